/* vulcan_b200.h — C ABI of the B200-native TSDF-fusion hot path.
 *
 * This is the drop-in boundary underneath the reference's public C++ classes
 * (vulcan::Volume, *Integrator, Tracer, Frame).  Every entry point cites the
 * reference interface it replaces (paths relative to /root/reference).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no C++ or torch types.
 *   - return value: 0 = ok, >0 = cudaError_t, <0 = VK_E_* below.  Nothing throws.
 *   - all work is stream-ordered on the `stream` argument (a cudaStream_t passed
 *     as void*; NULL = legacy default stream, which is what the reference uses for
 *     all of its launches, include/vulcan/device.h:40-52).
 *   - no process-global device state (the reference keeps its allocator counters
 *     in translation-unit __device__ globals, src/volume.cu:17-21; here they are
 *     per volume).
 *   - struct layouts below that mirror reference value types are byte images of
 *     those types so that the C++ wrappers can pass `&member` without repacking.
 */
#ifndef VULCAN_B200_H_
#define VULCAN_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VK_E_INVALID   (-1)  /* bad argument                                        */
#define VK_E_NODEVICE  (-2)  /* no CUDA device / kernels not built for this device  */
#define VK_E_OVERFLOW  (-3)  /* voxel pool or excess list exhausted (non fatal;     */
                             /* the reference drops such requests silently,         */
                             /* src/volume.cu:338-356)                              */
#define VK_E_IO        (-4)  /* snapshot file missing, truncated or of another version */
#define VK_E_TIMEOUT   (-5)  /* sharded volume: a peer did not arrive at a handshake within ~2 s */
#define VK_E_TOO_LARGE (-6)  /* vk_extract_mesh_static: the volume exceeds VK_MESH_STATIC_MAX_* */

/* vkVolume::math_mode: where float multiply-adds round.                               */
#define VK_MATH_REFERENCE 0  /* fused exactly where nvcc's default -fmad=true fuses them  */
                             /* in the reference's release build (bit-exact with it)      */
#define VK_MATH_STRICT    1  /* one rounding per source operation (reference built with   */
                             /* -fmad=false; the CPU oracle's strict mode)                */

#define VK_BLOCK_RES     8      /* Block::resolution, include/vulcan/block.h:13    */
#define VK_BLOCK_VOXELS  512    /* Block::voxel_count, include/vulcan/block.h:15   */
#define VK_BOUNDS_W      80     /* hard coded in src/tracer.cpp:56                 */
#define VK_BOUNDS_H      60     /* hard coded in src/tracer.cpp:57                 */
#define VK_SPACE_WORDS   65536  /* 128^3 bits, see vkVolume::space_bits            */
#define VK_MAX_SHARDS    8      /* ranks of one sharded volume (one NVSwitch box)  */

/* byte image of vulcan::Voxel (include/vulcan/voxel.h:41-49), 20 bytes */
typedef struct vkVoxel {
  float   distance;
  float   color[3];
  int16_t distance_weight;
  int16_t color_weight;
} vkVoxel;

/* byte image of vulcan::HashEntry (include/vulcan/hash.h:49-55), 16 bytes */
typedef struct vkHashEntry {
  int16_t block[3];
  int16_t pad;
  int32_t data;   /* voxel-block id, -1 = not allocated */
  int32_t next;   /* entry index of next chain element, -1 = end */
} vkHashEntry;

/* byte image of vulcan::Patch (include/vulcan/tracer.h:13-22), 16 bytes */
typedef struct vkPatch {
  int16_t origin[2];
  int16_t size[2];
  float   bounds[2];
} vkPatch;

/* byte image of vulcan::Transform (include/vulcan/transform.h:166-170):
 * column-major 4x4 matrix followed by its (cached) inverse, 128 bytes */
typedef struct vkTransform {
  float m[16];
  float inv[16];
} vkTransform;

/* byte image of vulcan::Projection (include/vulcan/projection.h:109-113) */
typedef struct vkProjection {
  float fx, fy, cx, cy;
} vkProjection;

/* byte image of vulcan::Light (include/vulcan/light.h:62-66) */
typedef struct vkLight {
  float intensity;
  float position[3];
} vkLight;

/* vulcan::Visibility (include/vulcan/types.h:6-11) */
enum { VK_VIS_UNKNOWN = 0, VK_VIS_FALSE = 1, VK_VIS_TRUE = 2 };

/* device-resident counters of one volume (ints) */
enum {
  VK_CNT_VISIBLE0   = 0,  /* visible count, slot 0                                  */
  VK_CNT_VISIBLE1   = 1,  /* visible count, slot 1                                  */
  VK_CNT_VOXEL_PTR  = 2,  /* top of free voxel-block stack (voxel_pointer)          */
  VK_CNT_EXCESS_PTR = 3,  /* next excess entry (excess_pointer)                     */
  VK_CNT_OVERFLOW   = 4,  /* number of dropped allocations                          */
  VK_CNT_NEW_BLOCKS = 5,  /* reserved                                               */
  VK_CNT_ALLOCATED  = 6,  /* blocks allocated in total                              */
  VK_CNT_PATCHES    = 7,  /* reserved                                               */
  VK_CNT_UPDATED    = 8,  /* 64-bit (slots 8,9): voxels written by the integrators  */
  VK_CNT_COUNT      = 16
};

/* One voxel-hash volume.  Replaces the data members of vulcan::Volume
 * (include/vulcan/volume.h:86-116).  All pointers are device pointers owned by
 * the library between vk_volume_create and vk_volume_destroy. */
struct vkShard;
typedef struct vkVolume {
  vkVoxel*     voxels;        /* [(M+E)*512], 256-byte aligned                      */
  vkHashEntry* entries;       /* [M+E]: [0,M) bucket heads, [M,M+E) excess          */
  int32_t*     free_list;     /* [M+E] free voxel-block ids (stack)                 */
  uint8_t*     visibility;    /* [M+E] vulcan::Visibility per entry                 */
  int32_t*     visible[2];    /* [M+E] each; visible[visible_slot] is current       */
  int32_t*     stamps;        /* [M+E] id of the last SetView that touched an entry */
  int32_t*     locks;         /* [M]   per-bucket insertion locks                   */
  uint64_t*    lock_tags;     /* [M]   block the bucket's inserter works on (lets rays of the same block move on) */
  uint32_t*    bucket_bits;   /* [ceil(M/32)] bit h set <=> bucket head h allocated */
  uint32_t*    space_bits;    /* [VK_SPACE_WORDS] bit of (x,y,z) mod 128 set <=> some block with    */
                              /* those low coordinate bits is allocated (free-space filter)       */
  float*       sdf;           /* [(M+E)*512] mirror of Voxel::distance alone, block-major (2 KB per  */
                              /* block, index z*64+y*8+x): what the ray marcher samples.  Kept in    */
                              /* step by the integrators, vk_volume_reset and vk_volume_load; after  */
                              /* writing `voxels` directly call vk_volume_sync_sdf.                  */
  int32_t*     counters;      /* [VK_CNT_COUNT] device                              */
  int32_t*     host_counters; /* [VK_CNT_COUNT] pinned host mirror, filled by vk_volume_sync_counters only: no   */
                              /* kernel stores to host memory (such a store queues behind the PCIe traffic of   */
                              /* concurrent downloads and delayed the kernel's completion by ~14 us per frame)   */
  int32_t      main_blocks;   /* M (also the hash modulus K)                        */
  int32_t      excess_blocks; /* E                                                  */
  float        voxel_length;
  float        trunc_length;
  float        min_depth;     /* Volume::depth_range_ (gates allocation only)       */
  float        max_depth;
  int32_t      shard_rank;    /* 0 when unsharded                                   */
  int32_t      shard_count;   /* 1 when unsharded                                   */
  int32_t      shard_shift;   /* log2 of the brick edge in blocks                   */
  int32_t      shard_halo;    /* 0: a rank stores the blocks it owns (exact peer-memory ray cast);  */
                              /* 1: also their 26 neighbours (for the key-merge ray cast)           */
  int32_t      visible_slot;  /* which of visible[] / VK_CNT_VISIBLE* is current    */
  int32_t      view_id;       /* number of vk_set_view calls so far                 */
  int32_t      math_mode;     /* VK_MATH_REFERENCE (default) or VK_MATH_STRICT       */
  int32_t      has_color;     /* a colour / light integrator has run (or voxels were written from outside):  */
                              /* while 0 every Voxel::color_weight is 0, so the ray marcher writes the colour  */
                              /* that implies - (0,0,0) - without gathering the 8 colour taps per ray          */
  struct vkShard* shard;      /* replicated lookup structures of a block-sharded volume (vk_shard_create), or NULL */
  /* Library-owned side stream + events (created by vk_volume_create): vk_trace computes the ray bounds of a view -  */
  /* which depend on the visible list of vk_set_view only, not on the voxels - on aux_stream while the caller's      */
  /* stream is still integrating, and joins before the ray march.  Results and stream-order semantics are unchanged. */
  void*        aux_stream;
  void*        ev_view;       /* recorded on the caller's stream at the end of vk_set_view                     */
  void*        ev_bounds;     /* recorded on aux_stream after the bounds kernel                                 */
  void*        ev_trace;      /* recorded on the caller's stream at the end of every ray cast (bounds scratch free) */
  int32_t      aux_valid;     /* ev_view belongs to the current view                                            */
  int32_t      reserved0;
} vkVolume;

/* One RGB-D frame on the device.  Replaces vulcan::Frame
 * (include/vulcan/frame.h:11-32); images are row-major, colour/normal images are
 * packed 3 x float per pixel like vulcan::ColorImage (include/vulcan/image.h:268-272). */
typedef struct vkFrame {
  const float* depth;         /* [depth_w*depth_h] metric z depth, 0 = invalid      */
  const float* color;         /* [color_w*color_h*3] or NULL                        */
  const float* normal;        /* [depth_w*depth_h*3] or NULL (LightIntegrator)      */
  int32_t      depth_w, depth_h;
  int32_t      color_w, color_h;
  vkProjection depth_proj;
  vkProjection color_proj;
  vkTransform  depth_to_world;
  vkTransform  depth_to_color;
} vkFrame;

/* ---- volume lifetime --------------------------------------------------------- */

/* Volume::Volume + Volume::Initialize (src/volume.cu:370-380,552-627).  Allocates
 * every pool with cudaMalloc and fills it like the reference: voxels = Voxel::Empty(),
 * entries invalid, free list = 0..B-1, voxel pointer B-1, excess pointer M,
 * visibility FALSE.  Defaults: trunc 0.04, voxel 0.008, depth range (0.1, 5). */
int vk_volume_create(vkVolume* v, int main_blocks, int excess_blocks, void* stream);

/* Re-runs the fills of vk_volume_create on the existing pools. */
int vk_volume_reset(vkVolume* v, void* stream);

/* ~Buffer<T> of each member (include/vulcan/buffer.h:28-31). */
int vk_volume_destroy(vkVolume* v);

/* ---- Volume::SetView (src/volume.cu:430-437) --------------------------------- */

/* ResetBlockVisibility + CreateAllocationRequests + HandleAllocationRequests +
 * UpdateBlockVisibility, iterated to the fixed point the reference reaches by
 * calling SetView repeatedly (apps/vulcan/vulcan.cu:316-318,
 * tests/tracer_test.cu:298-303) — in one call, with no host synchronisation.
 * The visible count stays on the device (counters[VK_CNT_VISIBLE0 + visible_slot]). */
int vk_set_view(vkVolume* v, const vkFrame* f, void* stream);

/* Rebuilds vkVolume::sdf from `voxels` for every allocated block (stream-ordered).  Only needed after
 * the voxel pool was written from outside the library (the reference hands out a mutable
 * Buffer<Voxel>&, include/vulcan/volume.h:48-50; the C++ / Python layers call this before the next
 * trace when that accessor was used). */
int vk_volume_sync_sdf(vkVolume* v, void* stream);

/* Copies the counters to host_counters and synchronises the stream.  Returns the
 * current visible-block count in *visible (Volume::GetBufferSize, src/volume.cu:537-543);
 * returns VK_E_OVERFLOW if any allocation has been dropped. */
int vk_volume_sync_counters(vkVolume* v, int* visible, void* stream);

/* Volume snapshots (no reference counterpart: the reference has no volume I/O).  vk_volume_save
 * writes the complete state of a volume - hash entries, free list, visibility, visible lists,
 * occupancy bitmaps, counters and the voxel blocks in use - to `path`; vk_volume_load restores it
 * into a volume created with the same block counts, which then continues exactly as the saved one
 * would have.  Both synchronise `stream`.  Little-endian, versioned; VK_E_IO on file errors. */
int vk_volume_save(const vkVolume* v, const char* path, void* stream);
int vk_volume_load(vkVolume* v, const char* path, void* stream);

/* ---- Integrators ------------------------------------------------------------- */

/* DepthIntegrator::Integrate (src/depth_integrator.cu:17-115). */
int vk_integrate_depth(vkVolume* v, const vkFrame* f, float min_depth, float max_depth,
                       float max_distance_weight, void* stream);

/* ColorIntegrator::Integrate (src/color_integrator.cu:17-204): depth update followed
 * by the colour update of the same voxel, in one pass over the voxel tiles. */
int vk_integrate_color(vkVolume* v, const vkFrame* f, float min_depth, float max_depth,
                       float max_distance_weight, float max_color_weight, void* stream);

/* LightIntegrator::ComputeFrameMask (src/light_integrator.cu:17-103,277-293). */
int vk_compute_frame_mask(const vkFrame* f, float depth_threshold, float* mask, void* stream);

/* LightIntegrator::Integrate (src/light_integrator.cu:270-354); `mask` is scratch of
 * depth_w*depth_h floats (LightIntegrator::frame_mask_). */
int vk_integrate_light(vkVolume* v, const vkFrame* f, float min_depth, float max_depth,
                       float max_distance_weight, float max_color_weight,
                       const vkLight* light, float depth_threshold, float* mask, void* stream);

/* ---- Tracer ------------------------------------------------------------------ */

/* Tracer::Trace (src/tracer.cpp:41-47): ray bounds from the current visible list,
 * per-pixel ray march, normals.  `bounds` is scratch of VK_BOUNDS_W*VK_BOUNDS_H*2 floats
 * (Tracer::bounds_).  out_depth/out_color/out_normal are the traced frame's images
 * (out_normal may be NULL).  `view` supplies pose, intrinsics and image size only. */
int vk_trace(const vkVolume* v, const vkFrame* view, float min_depth, float max_depth,
             float* bounds, float* out_depth, float* out_color, float* out_normal,
             void* stream);

/* Same, but additionally writes per-pixel merge keys (depth_bits << 32 | rgb8 << 8 | rank)
 * for the sharded nearest-hit merge; a pixel without an owned hit gets the key of +inf. */
int vk_trace_keys(const vkVolume* v, const vkFrame* view, float min_depth, float max_depth,
                  float* bounds, unsigned long long* out_keys, void* stream);

/* Decode merged keys into depth + colour images and compute normals from the merged depth. */
int vk_resolve_keys(const unsigned long long* keys, const vkFrame* view, float* out_depth,
                    float* out_color, float* out_normal, void* stream);

/* Step-wise entry points with the reference's free-function signatures
 * (include/vulcan/tracer.cuh:14-32; used directly by tests/tracer_test.cu). */
int vk_compute_patches(const int* indices, const vkHashEntry* entries, const vkTransform* Tcw,
                       const vkProjection* proj, float block_length, float min_depth,
                       float max_depth, int block_count, int image_w, int image_h,
                       int bounds_w, int bounds_h, vkPatch* patches, int* patch_count,
                       void* stream);
int vk_compute_bounds(const vkPatch* patches, float* bounds, int bounds_w, int patch_count,
                      void* stream);
int vk_reset_bounds(float* bounds, int count, void* stream);
int vk_compute_points(const vkHashEntry* entries, const vkVoxel* voxels, const float* bounds,
                      int block_count, float block_length, float voxel_length,
                      float trunc_length, const vkTransform* Twc, const vkProjection* proj,
                      float* depths, float* colors, int image_w, int image_h,
                      int bounds_w, int bounds_h, void* stream);

/* Frame::ComputeNormals / vulcan::ComputeNormals (src/frame.cu:7-122,183-192). */
int vk_compute_normals(const float* depths, const vkProjection* proj, float* normals,
                       int image_w, int image_h, void* stream);

/* Frame::FilterDepths / vulcan::FilterDepths (src/frame.cu:124-181,194-203). */
int vk_filter_depths(int image_w, int image_h, const float* src, float* dst, void* stream);

/* Input-side pre-pass of a sensor frame in one kernel: dst_depth = FilterDepths(src) and
 * dst_normals = ComputeNormals(dst_depth), bit-identical to vk_filter_depths followed by
 * vk_compute_normals (src/frame.cpp:8-36).  src and dst_depth must not alias.
 * Note: in the reference's release build Frame::FilterDepths filters into a temporary and the
 * copy back into the frame sits in CUDA_DEBUG(...), which NDEBUG removes (src/frame.cpp:18), so
 * the frame stays unfiltered there; the free function vulcan::FilterDepths is what this mirrors. */
int vk_prepare_frame(int image_w, int image_h, const float* src, const vkProjection* proj,
                     float* dst_depth, float* dst_normals, void* stream);

/* ---- sharding (no reference counterpart) -------------------------------------- */

/* Tables of all ranks of a sharded volume as seen from one rank: its own pointers in slot
 * `rank`, the others mapped from their processes (vk_ipc_open) over NVLink / NVSwitch. */
typedef struct vkPeerTable {
  int32_t            count, rank;
  const vkHashEntry* entries[VK_MAX_SHARDS];
  const vkVoxel*     voxels[VK_MAX_SHARDS];
  const float*       sdf[VK_MAX_SHARDS];      /* vkVolume::sdf of every rank */
} vkPeerTable;

/* CUDA IPC handles of a volume's hash entries, voxel pool and distance mirror (3 x cudaIpcMemHandle_t). */
typedef struct vkIpcHandles {
  unsigned char entries[64];
  unsigned char voxels[64];
  unsigned char sdf[64];
} vkIpcHandles;

/* Exact ray casting of a block-sharded volume.  Every rank holds the blocks it owns (plus a
 * one-block halo if shard_halo is set) and renders a band of image rows; a block of another rank is
 * read from its owner's tables through peer memory, so each ray takes exactly the steps it
 * takes in an unsharded volume.  Per traced view and rank:
 *   vk_trace_bounds   ray bounds of the LOCAL visible blocks into `bounds` (80 x 60 x 2 floats),
 *                     each cell as (min depth, -max depth), +FLT_MAX where empty: the caller
 *                     combines the ranks' grids with ONE element-wise MIN all-reduce, which is
 *                     also the barrier that orders every rank's integration before any rank's
 *                     peer reads;
 *   vk_trace_rows     marches rows [row_begin - 2, row_end + 2) of the view into the full-size
 *                     images and computes the normals of rows [row_begin, row_end); the caller
 *                     all-gathers the bands.  `bounds` is in the (min, -max) format of
 *                     vk_trace_bounds.  peers == NULL or count 1: plain unsharded trace. */
int vk_trace_bounds(const vkVolume* v, const vkFrame* view, float min_depth, float max_depth,
                    float* bounds, void* stream);
int vk_trace_rows(const vkVolume* v, const vkPeerTable* peers, const vkFrame* view, const float* bounds,
                  int row_begin, int row_end, float* out_depth, float* out_color, float* out_normal,
                  void* stream);
int vk_ipc_export(const vkVolume* v, vkIpcHandles* out);
int vk_ipc_open(const vkIpcHandles* handles, const vkHashEntry** entries, const vkVoxel** voxels, const float** sdf);
int vk_ipc_close(const vkHashEntry* entries, const vkVoxel* voxels, const float* sdf);

/* ---- block-sharded volume with replicated lookup structures (the multi-GPU mode of bench.py) --------
 *
 * A rank stores, allocates and integrates exactly the blocks it owns (10 KB of voxels each).  What the
 * ray marcher needs from a block - its hash entry (16 B) and its distances (vkVolume::sdf, 2 KB) - is
 * REPLICATED: every rank keeps a local copy of every other rank's hash table and distance mirror, and
 * the owner's allocation / integration kernels store each entry / each changed 16-byte piece of
 * distances into all copies as they write their own (plain stores into peer memory over NVLink /
 * NVSwitch: posted, nobody waits for them).  Ray casting is then purely LOCAL on every rank - every ray
 * takes exactly the steps of the unsharded march, so the images are bit-identical - and each rank
 * renders a band of rows.  Per traced view there are two flag handshakes over peer memory and no
 * library collective:
 *   1. ray bounds: the last CTA of the bounds kernel stores the rank's 80 x 60 grid into every peer's
 *      buffer, announces it, waits for the others, leaves the element-wise minimum in place and splits
 *      the image rows into `count` bands of equal active-cell counts.  This is also the barrier that
 *      orders every rank's replica stores (earlier in its stream) before any rank's ray cast.
 *   2. bands: each rank marches its band (+2 rows either side for the normals) into its own output
 *      images, stores the band into every peer's output images and announces it; after the handshake
 *      every rank holds the complete depth / colour / normal images.  With gather = 0 step 2 is left
 *      out (the consumer is the host: each rank downloads only its band, vk_shard_band).
 * Only the colour of a hit (8 taps of Voxel::color / color_weight, once per ray) is read from the
 * owner's voxel pool through peer memory, and only if the volume has colour (vkVolume::has_color).
 *
 * Setup (host): vk_shard_create on every rank BEFORE the volume allocates anything, exchange the
 * vkShardHandles (any byte transport: torch.distributed, MPI, a file), vk_shard_connect.  Ranks that
 * share a process (one thread per GPU with peer access enabled, or the one-GPU emulation of the tests)
 * call vk_shard_connect_local instead.  All volumes of a sharded set must have the same block counts. */
typedef struct vkShard {
  int32_t count, rank;
  int32_t width, height;                      /* size of the output images                                   */
  int32_t epoch;                              /* views traced so far                                         */
  int32_t main_blocks, excess_blocks;
  int32_t connected;                          /* 0 no, 1 IPC mappings (closed by vk_shard_destroy), 2 local  */
  vkHashEntry*   entries[VK_MAX_SHARDS];      /* hash table of rank r as THIS rank reads it: the volume's    */
  float*         sdf[VK_MAX_SHARDS];          /* own arrays for r == rank, local replicas otherwise          */
  const vkVoxel* voxels[VK_MAX_SHARDS];       /* voxel pools: own / peer-mapped (colour gather only)         */
  vkHashEntry*   push_entries[VK_MAX_SHARDS]; /* this rank's replica ON rank r (peer memory, write-only)     */
  float*         push_sdf[VK_MAX_SHARDS];
  void*          exchange[VK_MAX_SHARDS];     /* exchange + output buffer of every rank (own one included)   */
} vkShard;

/* What a rank publishes to its peers: CUDA IPC handles (cudaIpcMemHandle_t, 64 bytes each). */
typedef struct vkShardHandles {
  unsigned char voxels[64];
  unsigned char exchange[64];
  unsigned char rep_entries[VK_MAX_SHARDS][64];   /* [r]: this rank's replica of rank r's table (unused for r == rank) */
  unsigned char rep_sdf[VK_MAX_SHARDS][64];
} vkShardHandles;

int vk_shard_create(vkShard* s, vkVolume* v, int rank, int count, int brick_shift, int width, int height,
                    void* stream);
int vk_shard_export(const vkShard* s, const vkVolume* v, vkShardHandles* out);
int vk_shard_connect(vkShard* s, const vkShardHandles* all_ranks /* [count], own slot ignored */);
int vk_shard_connect_local(vkShard* const* shards, vkVolume* const* volumes, int count);
int vk_shard_destroy(vkShard* s, vkVolume* v);
/* Tracer::Trace of a sharded volume (collective: every rank calls it for the same view).  *out_* receive
 * this rank's output images (device memory inside its exchange buffer, valid until the second-next
 * call); with gather != 0 they are complete on return of the stream, with gather == 0 only rows
 * [band_begin, band_end) of vk_shard_band are. */
int vk_trace_sharded(vkVolume* v, const vkFrame* view, float min_depth, float max_depth, int gather,
                     float** out_depth, float** out_color, float** out_normal, void* stream);
/* The same view in three parts, for ranks that SHARE a GPU (the one-GPU emulation of the tests): kernels
 * of one device must never wait for one another, so every rank runs part 1 (local ray bounds, stored into
 * the peers and announced), then every rank part 2 (combine the bounds - all announcements are in - split
 * the rows, march the band, store it into the peers, announce), then every rank part 3 (nothing left to
 * wait for).  stage = 1, 2, 3; same arguments for all three. */
int vk_trace_sharded_staged(vkVolume* v, const vkFrame* view, float min_depth, float max_depth, int gather,
                            int stage, float** out_depth, float** out_color, float** out_normal, void* stream);
/* Rows this rank rendered in the last vk_trace_sharded call (synchronises the stream: the split is
 * computed on the device from the merged ray bounds).  VK_E_TIMEOUT if a handshake of this rank was ever
 * given up (a peer did not arrive within ~2 s; the images of that view are incomplete). */
int vk_shard_band(const vkVolume* v, int* band_begin, int* band_end, void* stream);
/* The same without synchronising, for streaming hosts: enqueues on `stream` the copy of the LAST traced view's band
 * table into host_bands[0..16) (page-locked; rank r renders rows [host_bands[r], host_bands[r+1])) and of the count of
 * handshakes given up into host_bands[16] (non-zero = VK_E_TIMEOUT).  The table of a view stays valid until the
 * second-next vk_trace_sharded of the volume. */
int vk_shard_band_async(const vkVolume* v, int* host_bands /* [17] */, void* stream);

/* Rank that owns block (bx, by, bz) when a volume is split over `count` ranks: a pure function
 * of the brick (bx >> shift, by >> shift, bz >> shift).  Host function; the kernels use the
 * same arithmetic.  With vkVolume::shard_halo = 1 a rank also stores the 26 neighbours of every
 * block it owns, so that the interpolation of the key-merge ray cast (vk_trace_keys) never
 * leaves the shard; only owned hits enter the merge keys. */
int vk_shard_owner(int bx, int by, int bz, int shift, int count);

/* ---- image pyramid + depth tracker (the caller side of Tracer::Trace) ---- */

#define VK_TRACK_STATE_FLOATS 80  /* solver state at the head of a tracker workspace */

/* Frame::Downsample (src/frame.cpp:38-57; kernels src/image.cu:101-165,183-274) in one launch:
 * depth and normals by nearest sample, colour by 2x2 average, into the device buffers dst already
 * points to (dst->depth / color / normal are written through although vkFrame declares them
 * const); sizes, halved intrinsics and the transforms are filled in.  Images that are NULL in src
 * are skipped.  Odd image sizes -> VK_E_INVALID ("even image dimensions required"). */
int vk_downsample_frame(const vkFrame* src, vkFrame* dst, void* stream);
/* Image::Downsample / ColorImage::Downsample (src/image.cu:183-274): channels 1 or 3 */
int vk_downsample_image(const float* src, int src_w, int src_h, int channels, int nearest, float* dst,
                        void* stream);

/* DepthTracker::ComputeResiduals / ComputeJacobian (src/depth_tracker.cu:279-344): one float /
 * six floats per frame pixel (Jacobian columns 3..5 are 0 with translation disabled). */
int vk_track_residuals(const vkFrame* keyframe, const vkFrame* frame, float* residuals, void* stream);
int vk_track_jacobian(const vkFrame* keyframe, const vkFrame* frame, int translation_enabled,
                      float* jacobian, void* stream);
/* floats a tracker workspace (device memory) needs on the current device */
int vk_track_workspace_floats(void);
/* DepthTracker::ComputeSystem (src/depth_tracker.cu:346-378): packed lower triangle of J^T J
 * (21 or 6 floats, row r column c <= r at r(r+1)/2+c) and J^T r (6 or 3 floats), device buffers.
 * Sums are deterministic (fixed order), unlike the reference's atomicAdd of CTA sums. */
int vk_track_system(const vkFrame* keyframe, const vkFrame* frame, int translation_enabled,
                    float* workspace, float* hessian, float* gradient, void* stream);
/* Tracker::Track (src/tracker.cpp:57-67) for the DepthTracker: up to max_iterations Gauss-Newton
 * steps, each ONE kernel that evaluates the system, solves it and applies
 * DepthTracker::ApplyUpdate (src/depth_tracker.cpp:23-88) to the pose in device memory; stops
 * like Tracker::ComputeUpdate when |update| < 1e-6 (src/tracker.cpp:152).  On return (the call
 * waits for the stream) frame->depth_to_world holds the tracked pose; *iterations = systems
 * solved. */
int vk_track_depth(const vkFrame* keyframe, vkFrame* frame, int translation_enabled, int max_iterations,
                   float* workspace, int* iterations, float* last_update_norm, void* stream);

/* ---- surface points (first slice; no finished reference counterpart) ------- */

/* Marching-cubes vertices of the whole volume: the zero crossings of the TSDF on the three lattice
 * edges every voxel owns (+x, +y, +z neighbour - edges 0, 3, 8 of its cube, the edge set of
 * src/extractor.cu:52-117), over all allocated blocks and across block faces, both end points
 * observed (distance_weight > 0) and of different sign; position and colour interpolated linearly.
 * points / colors: device float[3 * capacity] (colors may be NULL); *count (device int) receives the
 * number of crossings found, which may exceed capacity (then only the first `capacity` appended
 * are stored: call again with a larger buffer).  Order is arbitrary. */
int vk_extract_points(const vkVolume* v, float* points, float* colors, int capacity, int* count,
                      void* stream);

/* The triangle mesh of the whole volume: the vertices of vk_extract_points plus marching-cubes faces over
 * them (the reference's ExtractVertexIndicesKernel / ExtractFacesKernel are empty, src/extractor.cu:392-413).
 * One cube per voxel (the voxel and its +x / +y / +z neighbours, across block faces), meshed when all eight
 * corners are observed; case bit c set where !(distance > 0); triangle table generated by
 * tools/gen_mc_tables.py (segments on a cube face depend only on that face's corner signs, so neighbouring
 * cubes agree and the mesh is watertight wherever the surface is surrounded by observed voxels; normals point
 * towards positive distance).  faces: device int[3 * face_capacity], vertex indices into points; *point_count
 * and *face_count (device ints) receive the numbers found, which may exceed the capacities (then call again
 * with larger buffers).  Synchronises the stream (sizes its workspace from the allocated-block counter). */
int vk_extract_mesh(const vkVolume* v, float* points, float* colors, int point_capacity, int* point_count,
                    int* faces, int face_capacity, int* face_count, void* stream);

/* vk_extract_mesh without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same arguments, same result as vk_extract_mesh (the same two kernels);
 * the workspace vk_extract_mesh allocates per call lives in one file-scope `static __device__` array
 * instead (vk_extract.cu: g_mesh_static, 64 + VK_MESH_STATIC_MAX_POOL + 1536 * VK_MESH_STATIC_MAX_BLOCKS
 * ints = 12.5 MiB, held by every GPU that loads the library until the process exits).  Limits: main_blocks +
 * excess_blocks <= VK_MESH_STATIC_MAX_POOL and at most VK_MESH_STATIC_MAX_BLOCKS allocated blocks; beyond
 * either the call writes nothing and returns VK_E_TOO_LARGE (use vk_extract_mesh).  Calls no cudaMalloc /
 * cudaFree / allocating library routine.  Synchronises the stream once (reads the allocated-block counter)
 * and returns with the kernels enqueued.  Every call shares the array: calls that overlap on the GPU
 * (different streams or host threads of one device) race - serialise them. */
#define VK_MESH_STATIC_MAX_POOL   131072  /* main_blocks + excess_blocks */
#define VK_MESH_STATIC_MAX_BLOCKS 2048    /* allocated blocks            */
int vk_extract_mesh_static(const vkVolume* v, float* points, float* colors, int point_capacity, int* point_count,
                           int* faces, int face_capacity, int* face_count, void* stream);

/* ---- misc -------------------------------------------------------------------- */

/* Library version and the number of kernels launched by this library since load
 * (used by bench.py's gpu_launches). */
int vk_version(void);
unsigned long long vk_launch_count(void);

/* Math mode used by the entry points that take no vkVolume (vk_compute_normals,
 * vk_filter_depths, vk_compute_frame_mask, the step-wise tracer functions) and copied into
 * vkVolume::math_mode by vk_volume_create.  Plain host configuration, no device state. */
int vk_set_default_math_mode(int mode);
int vk_get_default_math_mode(void);

/* Division by loop-invariant divisors (block length, voxel length, truncation length, image
 * size) uses a reciprocal + one FMA residual step instead of the IEEE division sequence, after
 * the library has checked on the device, for that divisor, that the result equals __fdiv_rn for
 * every dividend the fast path is used on (all 2^24 mantissas / signs of one binade - inside the
 * guarded exponent range both sequences scale exactly with the dividend's exponent; ~20 us, cached
 * per value).  The first use of a divisor value synchronises `stream`, so callers that capture CUDA
 * graphs or stream frames call this once per length when a volume is configured (the C++ and Python
 * layers do so in Set*Length).  Returns 1 if the fast path is proven for d, 0 if the library will
 * use plain IEEE division for it. */
int vk_div_prepare(float d, void* stream);

#ifdef __cplusplus
}
#endif

#endif /* VULCAN_B200_H_ */
