"""The C++ facade (include/pbr/*.hpp) keeps the reference's pbr:: interfaces over the C ABI.  Host-only
behaviour (scene traversal order, MeshBuilder offsets, camera / model matrices -- pure CPU code in the
reference: Scene.cpp, MeshBuilder.cpp, CameraData.hpp, ModelPushConstant.hpp) is tested here without a GPU."""
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
CPP = ROOT / "tests" / "cpp"


@pytest.fixture(scope="module")
def built():
    subprocess.run(["make", "-C", str(CPP)], check=True, capture_output=True)
    return CPP / "build"


def test_host_facade_behaviour(built):
    r = subprocess.run([str(built / "facade_host_test")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    for case in ("sceneTraversalIsPostOrder", "findCameraReturnsFirstInPostOrder", "meshBuilderConcatenatesWithRunningOffsets",
                 "cameraDataMatchesClosedForm", "modelPushConstantNormalMatrix", "uniformRoundTrips"):
        assert f"ok {case}" in r.stdout


def test_facade_headers_keep_reference_names():
    """Every class / method the reference's client calls on this path (App.cpp:258-274, 387-392) exists."""
    text = "".join(p.read_text() for p in (ROOT / "include" / "pbr").rglob("*.hpp"))
    for name in ("class PbrRenderSystem", "struct PbrRenderSystemCreateInfo", "allocateGBuffer(IAllocator& allocator, vk::Extent2D extent)",
                 "auto render(vk::CommandBuffer cmdBuffer, Scene const& scene, GBuffer const& gBuffer, Image2D const& renderTarget",
                 "LIGHTING_PASS_OUTPUT_FORMAT = vk::Format::eR16G16B16A16Sfloat", "class TonemapperSystem",
                 "allocateHdrImage(IAllocator& allocator, vk::Extent2D extent)", "auto run(vk::CommandBuffer cmdBuffer, HdrImage const& hdrImage)",
                 "updateOutputTexture(vk::Image image, vk::ImageView", "getPositions()", "getNormals()", "getAlbedo()", "getDepth()",
                 "POSITIONS_FORMAT = vk::Format::eR16G16B16A16Sfloat", "ALBEDO_FORMAT = vk::Format::eR16G16B16A16Unorm",
                 "DEPTH_FORMAT = vk::Format::eD32Sfloat", "findCamera()", "iterateAllNodes()", "makeCameraData(", "makeModelPushConstant(",
                 "class MeshBuilder", "struct MaterialData", "class AsyncSubmitter", "class IAllocator", "makeGpuHandle("):
        assert name in text, name


def test_facade_fails_loudly_without_gpu(built):
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("GPU present")
    except ImportError:
        pass
    r = subprocess.run([str(built / "facade_frame"), "blinnphong", "4", "4", "0", "0", "5", "/tmp"], capture_output=True, text=True)
    assert r.returncode == 1 and "GpuHandle" in r.stderr  # std::runtime_error, no CPU fallback
