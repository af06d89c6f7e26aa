"""GPU (-m gpu): the drop-in, proven from the REFERENCE's side.

oracle/_ref/libref_gpu.so is the reference's own, unmodified Scene.cpp (+ its scene classes) in which the GPU branch of
CScene::LancerRayons -- `if (CVar::g_ComputerShadersON)`, Scene.cpp:692, formerly the GLSL dispatch -- runs the binding of
INTEGRATION.md (oracle/ref_gpu_binding.inc: the reference's getters -> rtb200_set_scene / rtb200_render -> m_InfoPixel),
linked with -lrtb200.  The test loads a bundled scene with the reference's parser, calls the reference's LancerRayons()
once per branch and compares what each branch hands to glTexImage2D (Scene.cpp:1562): the CUDA path behind the
reference's own switch against the reference's own CPU loop (Scene.cpp:1535-1563)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, bits_equal

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref_gpu():
    from oracle import refbind
    if not refbind.available("gpu"):
        pytest.skip("oracle/_ref/libref_gpu.so not built (needs /root/reference at build time)")
    r = refbind.RefRenderer("gpu")
    yield r
    r.free()


@pytest.mark.parametrize("scene,depth,W,H", [(1, 2, 640, 480), (5, 4, 320, 240), (9, 4, 400, 300), (7, 20, 256, 192), (2, 2, 200, 150)])
def test_reference_lancer_rayons_gpu_branch_equals_its_cpu_branch(ref_gpu, scene, depth, W, H):
    path = os.path.join(GOLDEN, "scenes", f"scene{scene}.dat")
    ref_gpu.load(path, max_bounces=depth)
    ref_gpu.use_gpu_branch(True)
    gpu, _ = ref_gpu.render_full(W, H)              # CScene::LancerRayons(): Scene.cpp:692 -> rtb200 -> glTexImage2D(m_InfoPixel)
    ref_gpu.load(path, max_bounces=depth)           # (a fresh scene: Pretraitement is not idempotent, Scene.cpp:140-147)
    ref_gpu.use_gpu_branch(False)
    cpu, _ = ref_gpu.render_full(W, H)              # the reference's own CPU loop, Scene.cpp:1535-1563
    assert gpu.shape == cpu.shape == (H, W, 3)
    assert np.isfinite(cpu).all() and cpu.max() > 0.1
    assert bits_equal(gpu, cpu), f"{int((gpu.view(np.uint32) != cpu.view(np.uint32)).sum())} differing floats, max |d| = {np.abs(gpu - cpu).max()}"


def test_other_flavours_refuse_the_gpu_branch():
    from oracle import refbind
    if not refbind.available("rec"):
        pytest.skip("oracle/_ref not built")
    r = refbind.RefRenderer("rec")
    with pytest.raises(RuntimeError):
        r.use_gpu_branch(True)
