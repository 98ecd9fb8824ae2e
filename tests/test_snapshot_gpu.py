"""Volume snapshots (vk_volume_save / vk_volume_load): a volume restored from a snapshot must hold
exactly the saved state and continue exactly like the volume that was saved."""
import os

import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu


def _frame(vb, torch, scn, fr):
    K = scn["K"]
    f = vb.Frame()
    f.depth_projection = vb.Projection(*K)
    T = fr["T"]
    f.depth_to_world_transform = vb.Transform(T[:16].reshape(4, 4).T.copy(), T[16:].reshape(4, 4).T.copy())
    f.depth_image = torch.from_numpy(fr["depth"]).cuda()
    return f


def _state(vb, torch, scn, vol, frame):
    out = vb.Frame()
    out.depth_projection, out.depth_to_world_transform = frame.depth_projection, frame.depth_to_world_transform
    out.depth_image = torch.zeros((scn["h"], scn["w"]), dtype=torch.float32, device="cuda")
    vb.Tracer(vol).Trace(out)
    n_vis, counters = vol.SyncCounters()
    entries = vol.GetHashEntries().cpu().numpy()
    keys, data, idx = H.allocated_table(entries, scn["M"])
    vox = vol.GetVoxels().view(-1, 512, 5)[torch.as_tensor(data, device="cuda").long()].cpu().numpy()
    return dict(keys=keys, voxels=vox, vis_flags=vol.GetBlockVisibility().cpu().numpy()[idx],
                visible_keys=H.visible_keys(entries, vol.GetVisibleBlocks().cpu().numpy()),
                depth=out.depth_image.cpu().numpy(), color=out.color_image.cpu().numpy(),
                normal=out.normal_image.cpu().numpy(), counters=list(counters), entries=entries)


def test_snapshot_restores_and_continues(tmp_path):
    import torch
    import vulcan_b200 as vb
    scn = H.make_scenario("small", (0, 5, 10, 15, 20), "depth")
    path = str(tmp_path / "volume.vkvol")

    def volume():
        v = vb.Volume(scn["M"], scn["E"])
        v.SetVoxelLength(scn["voxel"])
        v.SetTruncationLength(scn["trunc"])
        return v

    a = volume()
    ia = vb.DepthIntegrator(a)
    f = None
    for fr in scn["frames"][:3]:
        f = _frame(vb, torch, scn, fr)
        a.SetView(f)
        ia.Integrate(f)
    a.Save(path)
    saved = _state(vb, torch, scn, a, f)
    assert os.path.getsize(path) < 0.2 * (scn["M"] + scn["E"]) * 10240, "only the blocks in use are stored"

    b = vb.Volume(scn["M"], scn["E"])   # default lengths: Load restores them
    b.Load(path)
    assert abs(b.GetVoxelLength() - scn["voxel"]) < 1e-9 and abs(b.GetTruncationLength() - scn["trunc"]) < 1e-9
    restored = _state(vb, torch, scn, b, f)
    # the restored volume is the saved one, entry for entry
    assert np.array_equal(saved["entries"], restored["entries"])
    assert saved["counters"][:8] == restored["counters"][:8]
    for k in ("keys", "voxels", "vis_flags", "visible_keys", "depth", "color", "normal"):
        assert np.array_equal(np.asarray(saved[k]).view(np.uint8), np.asarray(restored[k]).view(np.uint8)), k

    # both continue with the same frames: same blocks, voxels and images (block ids may differ)
    ib = vb.DepthIntegrator(b)
    for fr in scn["frames"][3:]:
        f = _frame(vb, torch, scn, fr)
        for vol, integ in ((a, ia), (b, ib)):
            vol.SetView(f)
            integ.Integrate(f)
    rep = H.compare(_state(vb, torch, scn, a, f), _state(vb, torch, scn, b, f), scn["trunc"])
    assert rep["alloc_equal"] and rep["visible_equal"] and rep["vis_flags_equal"], rep
    assert rep["dw_mismatch"] == 0 and rep["dist_bits_mismatch"] == 0, rep
    assert rep["depth_bits_mismatch"] == 0 and rep["normal_bits_mismatch"] == 0 and rep["color_img_max_abs"] == 0.0, rep


def test_snapshot_errors(tmp_path):
    import vulcan_b200 as vb
    v = vb.Volume(4096, 1024)
    with pytest.raises(vb.VulcanError):
        v.Load(str(tmp_path / "missing.vkvol"))
    p = tmp_path / "garbage.vkvol"
    p.write_bytes(b"not a snapshot" * 100)
    with pytest.raises(vb.VulcanError):
        v.Load(str(p))
    good = str(tmp_path / "good.vkvol")
    v.Save(good)
    other = vb.Volume(2048, 512)
    with pytest.raises(vb.VulcanError):
        other.Load(good)   # block counts differ
