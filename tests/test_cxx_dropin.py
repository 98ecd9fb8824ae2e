"""The C++ drop-in layer (include/vulcan/*.h + vulcan_b200/cxx/vulcan.cpp -> libvulcan.so): the
reference's public classes over the C ABI.  CPU: it builds with a plain C++ compiler, exports the
reference's class surface and the headers keep the reference's value-type layouts.  GPU: the
reference's own known-answer scenarios (tests/cxx/test_dropin.cpp) pass."""
import os
import subprocess

import pytest

import harness as H

BUILD = os.path.join(H.ROOT, "vulcan_b200", "cxx", "build.sh")
LIB = os.path.join(H.ROOT, "vulcan_b200", "libvulcan.so")
BIN = os.path.join(H.ROOT, "tests", "cxx", "_bin", "test_dropin")


def _build():
    import vulcan_b200  # noqa: F401  (libvulcan_b200.so must exist: libvulcan.so links against it)
    subprocess.check_call(["bash", BUILD])


def test_cxx_library_builds_and_exports_reference_classes():
    _build()
    syms = subprocess.check_output(["nm", "-DC", "--defined-only", LIB], text=True)
    for needle in ("vulcan::Volume::Volume(int, int)", "vulcan::Volume::SetView(vulcan::Frame const&)",
                   "vulcan::Volume::GetVisibleBlocks() const", "vulcan::Volume::GetVoxels()",
                   "vulcan::Volume::GetHashEntries() const", "vulcan::Volume::GetAllocatedBlocks() const",
                   "vulcan::Volume::SetTruncationLength(float)", "vulcan::Volume::SetVoxelLength(float)",
                   "vulcan::Volume::SetDepthRange(float, float)",
                   "vulcan::Integrator::SetMaxDistanceWeight(float)", "vulcan::Integrator::SetMaxColorWeight(float)",
                   "vulcan::DepthIntegrator::Integrate(vulcan::Frame const&)",
                   "vulcan::ColorIntegrator::Integrate(vulcan::Frame const&)",
                   "vulcan::LightIntegrator::Integrate(vulcan::Frame const&)",
                   "vulcan::LightIntegrator::SetLight(vulcan::Light const&)",
                   "vulcan::Tracer::Trace(vulcan::Frame&)", "vulcan::Tracer::SetDepthRange(float, float)",
                   "vulcan::Frame::ComputeNormals()", "vulcan::Frame::FilterDepths()",
                   "vulcan::ComputePatches(", "vulcan::ComputeBounds(", "vulcan::ComputePoints(",
                   "vulcan::ComputeNormals(", "vulcan::ResetBoundsBuffer(",
                   "vulcan::FramePipeline::Submit(", "vulcan::FramePipeline::Result(int)",
                   "vulcan::ShardedTracer::Trace(", "vulcan::ShardedTracer::ConnectLocal(",
                   "vulcan::ShardedPipeline::Submit(", "vulcan::ShardedPipeline::Result(int)",
                   "vulcan_pipeline_create", "vulcan_pipeline_run", "vulcan_pipeline_submit", "vulcan_pipeline_result",
                   "vulcan_shard_handles_bytes", "vulcan_sharded_pipeline_create", "vulcan_sharded_pipeline_export",
                   "vulcan_sharded_pipeline_connect", "vulcan_sharded_pipeline_run", "vulcan_sharded_pipeline_destroy"):
        assert needle in syms, "libvulcan.so does not define " + needle
    assert os.path.exists(BIN)


def test_cxx_headers_keep_reference_layouts(tmp_path):
    """sizeof of the value types as Appendix B of SURVEY.md lists them (reference headers)."""
    src = tmp_path / "layout.cpp"
    src.write_text('#include <cstdio>\n#include <vulcan/vulcan.h>\nusing namespace vulcan;\n'
                   'int main() { printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(Voxel), sizeof(HashEntry), '
                   'sizeof(Block), sizeof(Patch), sizeof(Transform), sizeof(Projection), sizeof(Light), '
                   'sizeof(Vector3f)); Matrix4f m = Transform::Translate(1, 2, 3).GetMatrix(); '
                   'return (m(0, 3) == 1 && m.GetData()[12] == 1 && Transform::Translate(1, 2, 3).Inverse()'
                   '.GetMatrix()(2, 3) == -3) ? 0 : 1; }\n')
    exe = tmp_path / "layout"
    subprocess.check_call(["g++", "-std=c++14", "-I", os.path.join(H.ROOT, "include"), "-I", "/usr/local/cuda/include",
                           "-o", str(exe), str(src), "-L/usr/local/cuda/lib64", "-lcudart",
                           "-Wl,-rpath,/usr/local/cuda/lib64"])
    out = subprocess.check_output([str(exe)], text=True).split()
    assert [int(v) for v in out] == [20, 16, 8, 16, 128, 16, 16, 12]


@pytest.mark.gpu
def test_cxx_reference_scenarios_on_gpu():
    _build()
    res = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    print(res.stdout[-4000:], res.stderr[-2000:])
    assert res.returncode == 0, res.stdout[-4000:]
