"""Host logic of the block-sharded multi-GPU mode on CPU: the partition function agrees with the
library, covers every block exactly once, and the nearest-hit merge (MIN all-reduce of packed keys)
run over two gloo ranks reproduces the per-pixel nearest hit with its colour and owner."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import harness as H  # noqa: F401  (puts the repo on sys.path)
from vulcan_b200 import sharding as S


def test_partition_matches_library_and_is_a_partition():
    import vulcan_b200 as vb
    rng = np.random.default_rng(7)
    b = rng.integers(-300, 300, size=(4000, 3))
    for count in (1, 2, 4, 8):
        for shift in (0, 2, 3):
            own = S.shard_owner(b[:, 0], b[:, 1], b[:, 2], shift, count)
            assert own.min() >= 0 and own.max() < count
            for i in range(0, len(b), 97):
                assert vb.lib.vk_shard_owner(int(b[i, 0]), int(b[i, 1]), int(b[i, 2]), shift, count) == own[i]
            # ownership is per brick: all blocks of a brick share the owner
            brick = b >> shift
            _, inv = np.unique(brick, axis=0, return_inverse=True)
            for g in np.unique(inv)[:50]:
                assert len(set(own[inv.reshape(-1) == g])) == 1
    # every rank gets a fair share of bricks
    own = S.shard_owner(b[:, 0], b[:, 1], b[:, 2], 0, 8)
    assert np.bincount(own, minlength=8).min() > len(b) / 8 * 0.8
    assert vb.lib.vk_shard_owner(0, 0, 0, 3, 0) < 0


def test_key_packing_round_trip_and_order():
    rng = np.random.default_rng(3)
    depth = rng.uniform(0.1, 5.0, size=(6, 8)).astype(np.float32)
    depth[0, :3] = 0
    rgb = rng.uniform(0, 1, size=(6, 8, 3)).astype(np.float32)
    keys = S.pack_keys(depth, rgb, 5)
    d, c, r = S.unpack_keys(keys)
    assert np.array_equal(d, depth)
    assert np.abs(c - rgb)[depth > 0].max() <= 0.5 / 255 + 1e-6
    assert np.all(r[depth > 0] == 5) and np.all(r[depth == 0] == -1)
    # key order == depth order, misses last
    flat = keys.reshape(-1)
    order = np.argsort(flat, kind="stable")
    dd = depth.reshape(-1)[order]
    hits = dd[dd > 0]
    assert np.all(np.diff(hits) >= 0) and np.all(dd[len(hits):] == 0)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _merge_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(11)
    h, w = 24, 32
    # the same "true" per-rank hit images on every rank (seeded), each rank contributes its own
    depths = rng.uniform(0.5, 4.0, size=(world, h, w)).astype(np.float32)
    depths[rng.uniform(size=depths.shape) < 0.4] = 0      # pixels this rank does not own a hit for
    rgbs = rng.uniform(0, 1, size=(world, h, w, 3)).astype(np.float32)
    keys = torch.from_numpy(S.pack_keys(depths[rank], rgbs[rank], rank).copy())
    S.merge_keys(keys)
    d, c, r = S.unpack_keys(keys.numpy())
    big = np.where(depths > 0, depths, np.float32(np.inf))
    want_rank = np.argmin(big, axis=0)
    want_depth = np.min(big, axis=0)
    hit = np.isfinite(want_depth)
    assert np.array_equal(d[hit], want_depth[hit]) and np.all(d[~hit] == 0)
    assert np.array_equal(r[hit], want_rank[hit]) and np.all(r[~hit] == -1)
    want_rgb = np.take_along_axis(rgbs, want_rank[None, ..., None], axis=0)[0]
    assert np.abs(c - want_rgb)[hit].max() <= 0.5 / 255 + 1e-6
    with open(os.path.join(out_dir, "ok%d" % rank), "w") as f:
        f.write("ok")
    dist.destroy_process_group()


def test_nearest_hit_merge_over_two_gloo_ranks(tmp_path):
    world = 2
    mp.spawn(_merge_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all(os.path.exists(os.path.join(str(tmp_path), "ok%d" % r)) for r in range(world))
