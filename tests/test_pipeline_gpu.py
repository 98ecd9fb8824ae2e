"""FramePipeline (vulcan_b200/pipeline.py) must give, frame by frame, exactly the images the
synchronous calls give: overlapping the copies may not change a bit of the result."""
import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu

W, Hh = 320, 240


def _frames(n):
    K = np.array(H.synth.intrinsics(W, Hh), dtype=np.float32)
    out = []
    for k in range(n):
        R, t = H.synth.orbit_pose(k)
        out.append((H.synth.render_sphere_plane(W, Hh, K, R, t), H.raw32(R, t)))
    return K, out


def _meta(vb, K, T):
    f = vb.Frame()
    f.depth_projection = vb.Projection(*K)
    f.depth_to_world_transform = vb.Transform(T[:16].reshape(4, 4).T.copy(), T[16:].reshape(4, 4).T.copy())
    return f


@pytest.mark.parametrize("depth", [1, 2, 3])
def test_pipeline_matches_synchronous_calls(depth):
    import torch
    import vulcan_b200 as vb
    from vulcan_b200.pipeline import FramePipeline

    K, frames = _frames(7)

    def volume():
        vol = vb.Volume(32768, 8192)
        vol.SetVoxelLength(0.008)
        vol.SetTruncationLength(0.04)
        return vol

    # synchronous reference
    vol = volume()
    integ, tracer = vb.DepthIntegrator(vol), vb.Tracer(vol)
    want = []
    for d, T in frames:
        f = _meta(vb, K, T)
        f.depth_image = torch.from_numpy(d).cuda()
        vol.SetView(f)
        integ.Integrate(f)
        o = _meta(vb, K, T)
        o.depth_image = torch.zeros((Hh, W), dtype=torch.float32, device="cuda")
        tracer.Trace(o)
        want.append((o.depth_image.cpu().numpy(), o.color_image.cpu().numpy(), o.normal_image.cpu().numpy()))
    torch.cuda.synchronize()

    vol2 = volume()
    pipe = FramePipeline(vol2, vb.DepthIntegrator(vol2), vb.Tracer(vol2), W, Hh, depth=depth)
    pinned = [torch.from_numpy(d).pin_memory() for d, _ in frames]
    got = [None] * len(frames)
    for k, (d, T) in enumerate(frames):
        pipe.submit(pinned[k], _meta(vb, K, T))
        if k > 0 and depth > 1:
            got[k - 1] = tuple(x.numpy().copy() for x in pipe.result(k - 1))
        elif depth == 1:
            got[k] = tuple(x.numpy().copy() for x in pipe.result(k))
    got[-1] = tuple(x.numpy().copy() for x in pipe.result(len(frames) - 1))
    pipe.drain()
    for k in range(len(frames)):
        for a, b, name in zip(want[k], got[k], ("depth", "color", "normal")):
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (k, name)
    assert want[-1][0].max() > 0  # something was hit
    with pytest.raises(vb.VulcanError):
        pipe.result(0 if depth < len(frames) else 10_000)


def test_sharded_pipeline_single_rank_matches_synchronous_calls():
    """sharding.ShardedPipeline with ONE rank (its band is the whole image, the handshakes are trivial): the
    streamed bands equal the synchronous unsharded images bit for bit, frame by frame, with results taken two
    frames late.  (Several ranks need one GPU each: tools/sharded_check.py and bench.py --gpus N.)"""
    import torch
    import vulcan_b200 as vb
    from vulcan_b200 import sharding

    K, frames = _frames(8)

    def volume():
        vol = vb.Volume(32768, 8192)
        vol.SetVoxelLength(0.008)
        vol.SetTruncationLength(0.04)
        return vol

    vol = volume()
    integ, tracer = vb.DepthIntegrator(vol), vb.Tracer(vol)
    want = []
    for d, T in frames:
        f = _meta(vb, K, T)
        f.depth_image = torch.from_numpy(d).cuda()
        vol.SetView(f)
        integ.Integrate(f)
        o = _meta(vb, K, T)
        o.depth_image = torch.zeros((Hh, W), dtype=torch.float32, device="cuda")
        tracer.Trace(o)
        want.append((o.depth_image.cpu().numpy(), o.color_image.cpu().numpy(), o.normal_image.cpu().numpy()))
    torch.cuda.synchronize()

    vol2 = volume()
    shard = sharding.Shard(vol2, 0, 1, W, Hh)
    pipe = sharding.ShardedPipeline(shard, vb.DepthIntegrator(vol2), W, Hh, depth=3)
    pinned = [torch.from_numpy(d).pin_memory() for d, _ in frames]
    got = {}

    def take(j):
        b0, b1, hd, hc, hn = pipe.result(j)
        assert (b0, b1) == (0, Hh)
        got[j] = (hd.numpy().copy(), hc.numpy().copy(), hn.numpy().copy())

    for k, (d, T) in enumerate(frames):
        pipe.submit(pinned[k], _meta(vb, K, T))
        if k >= 2:
            take(k - 2)
    take(len(frames) - 2)
    take(len(frames) - 1)
    pipe.drain()
    for k in range(len(frames)):
        for a, b, name in zip(want[k], got[k], ("depth", "color", "normal")):
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (k, name)
    with pytest.raises(vb.VulcanError):
        pipe.result(0)
    shard.close()
    with pytest.raises(vb.VulcanError):
        sharding.ShardedPipeline(shard, vb.DepthIntegrator(vol2), W, Hh, depth=2)
