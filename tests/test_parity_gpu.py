"""Parity of the CUDA path (through the C ABI) against the CPU oracle and against the
reference's own CUDA sources (oracle/_ref), on identical synthetic inputs.

Bars (BASELINE.json north_star): allocated block set and visible list equal as sorted
sets; integer weights equal; TSDF within 1e-4 (already normalised by the truncation
length); ray depth within 0.5 mm on pixels both hit; hit-mask mismatch <= 0.1 %.
Against the contraction-free targets (oracle, reference built with -fmad=false) the bar
is bit-exact.
"""
import json
import os

import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu

OUT = os.path.join(H.ROOT, "gpurun_out")


def _dump(name, rep):
    os.makedirs(OUT, exist_ok=True)
    with open(os.path.join(OUT, "parity_%s.json" % name), "w") as f:
        json.dump(rep, f, indent=1)
    print(name, json.dumps(rep))


def _assert_exact(rep, what):
    assert rep["alloc_equal"], (what, rep)
    assert rep["visible_equal"], (what, rep)
    assert rep["vis_flags_equal"], (what, rep)
    assert rep["dw_mismatch"] == 0 and rep["cw_mismatch"] == 0, (what, rep)
    assert rep["dist_bits_mismatch"] == 0, (what, rep)
    assert rep["color_bits_mismatch"] == 0, (what, rep)
    assert rep["depth_bits_mismatch"] == 0, (what, rep)
    assert rep["normal_bits_mismatch"] == 0, (what, rep)
    assert rep["color_img_max_abs"] == 0.0, (what, rep)


def _ref_extras(rep):
    """Blocks only the reference allocated (torn request writes, Q24: they differ between two
    runs of the reference itself) - never more than a handful."""
    extras = rep["alloc_only_b"] + rep["alloc_dup_b"]
    assert rep["alloc_only_a"] == 0 and rep["alloc_dup_a"] == 0, rep
    assert extras <= max(4, rep["alloc_counts"][0] // 500), rep
    assert rep["visible_only_a"] == 0 and rep["visible_only_b_known"] == 0, rep
    return extras


def _assert_exact_ref(rep, what):
    """Bit-exact on everything the reference computes deterministically."""
    extras = _ref_extras(rep)
    assert rep["vis_flags_equal"], (what, rep)
    assert rep["dw_mismatch"] == 0 and rep["cw_mismatch"] == 0, (what, rep)
    assert rep["dist_bits_mismatch"] == 0 and rep["color_bits_mismatch"] == 0, (what, rep)
    if extras == 0:
        assert rep["alloc_equal"] and rep["visible_equal"], (what, rep)
        assert rep["depth_bits_mismatch"] == 0 and rep["normal_bits_mismatch"] == 0, (what, rep)
        assert rep["color_img_max_abs"] == 0.0, (what, rep)
    else:  # rays crossing a phantom block of the reference step differently
        assert rep["hit_mismatch_frac"] <= 1e-3 and rep["depth_gt_half_mm_frac"] <= 1e-3, (what, rep)
        assert rep["depth_bits_mismatch"] <= 64 * extras + 0.002 * rep["pixels"], (what, rep)


def _assert_tolerance(rep, what):
    _ref_extras(rep)
    # knife-edge voxels may flip under FMA contraction; weights of all others are equal
    assert rep["dw_mismatch"] <= 1e-4 * rep["voxels"], (what, rep)
    assert rep["dist_max_abs"] <= 1e-4, (what, rep)
    assert rep["hit_mismatch_frac"] <= 1e-3, (what, rep)
    assert rep["depth_gt_half_mm_frac"] <= 1e-3, (what, rep)


SCENARIOS = [
    ("tiny", (0,), "depth", 1.2),
    ("small", (0,), "depth", 1.2),
    ("small", (0, 10, 20), "depth", 1.2),    # orbit: carried-over / dropped visibility
    ("small", (0, 5), "color", 1.2),
    ("small", (0, 5), "light", 1.2),
    ("small", (0, 1, 2, 3), "depth", 25.0),  # large view changes
]


@pytest.mark.parametrize("name,frames,mode,step", SCENARIOS)
def test_ours_vs_oracle_exact(name, frames, mode, step):
    scn = H.make_scenario(name, frames, mode, step)
    ours = H.run_ours(scn)
    orc = H.run_oracle(scn)
    rep = H.compare(ours, orc, scn["trunc"])
    _dump("oracle_%s_%s_%d_%g" % (name, mode, len(frames), step), rep)
    _assert_exact(rep, "ours vs oracle")


@pytest.mark.parametrize("name,frames,mode,step", SCENARIOS)
def test_ours_vs_reference_nofmad_exact(name, frames, mode, step):
    if not H.have_ref(nofmad=True):
        pytest.fail("oracle/_ref/libvulcan_ref_nofmad.so missing: run oracle/build_ref.sh")
    scn = H.make_scenario(name, frames, mode, step)
    ours = H.run_ours(scn)
    ref = H.run_ref(scn, nofmad=True)
    rep = H.compare(ours, ref, scn["trunc"])
    _dump("refnofmad_%s_%s_%d_%g" % (name, mode, len(frames), step), rep)
    _assert_exact_ref(rep, "ours vs reference (-fmad=false)")


@pytest.mark.parametrize("name,frames,mode,step", SCENARIOS)
def test_ours_vs_reference_release(name, frames, mode, step):
    if not H.have_ref():
        pytest.fail("oracle/_ref/libvulcan_ref.so missing: run oracle/build_ref.sh")
    scn = H.make_scenario(name, frames, mode, step)
    ours = H.run_ours(scn)
    ref = H.run_ref(scn)
    rep = H.compare(ours, ref, scn["trunc"])
    _dump("ref_%s_%s_%d_%g" % (name, mode, len(frames), step), rep)
    _assert_tolerance(rep, "ours vs reference (release)")


def test_c1_full_size_vs_reference():
    """BASELINE config 1 at full size: 640x480, 4 mm, Volume(262144, 65536)."""
    scn = H.make_scenario("c1", (0,), "depth")
    ours = H.run_ours(scn)
    ref = H.run_ref(scn, nofmad=True)
    rep = H.compare(ours, ref, scn["trunc"])
    _dump("refnofmad_c1", rep)
    _assert_exact_ref(rep, "c1 ours vs reference (-fmad=false)")
    ref2 = H.run_ref(scn)
    rep2 = H.compare(ours, ref2, scn["trunc"])
    _dump("ref_c1", rep2)
    _assert_tolerance(rep2, "c1 ours vs reference (release)")
