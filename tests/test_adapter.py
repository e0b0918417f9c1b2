"""The C++ drop-in (adapter/CudaRenderWorld : RenderWorld).

CPU: it must compile against the reference's own headers (only where /root/reference exists).
GPU: oracle/_ref/adapter_check links the reference's classes, the adapter and liborion_vis.so into one
program and compares CudaRenderWorld with SimpleRenderWorld through the RenderWorld interface after
transform / box / removal / light mutations, main view and every shadow view (built in the container that
has the reference; the binary travels to the GPU box)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
CHECK = os.path.join(ROOT, "oracle", "_ref", "adapter_check")


def test_adapter_compiles_against_reference_headers(tmp_path):
    if not os.path.isdir(os.path.join(REF, "engine", "runtime")):
        pytest.skip("reference tree not present")
    inc = tmp_path / "inc"
    inc.mkdir()
    for c in ("core", "engine", "gpu", "input", "render_core", "graphics", "physics"):
        os.symlink(f"{REF}/engine/runtime/{c}/include", inc / c)
    os.symlink(f"{REF}/engine/3rdparty/glm/repo/glm", inc / "glm")
    # the adapter header is installed next to the reference's render headers
    (inc / "render").mkdir()
    for f in os.listdir(f"{REF}/engine/runtime/render/include"):
        os.symlink(f"{REF}/engine/runtime/render/include/{f}", inc / "render" / f)
    os.symlink(os.path.join(ROOT, "adapter", "cuda_render_world.h"), inc / "render" / "cuda_render_world.h")
    cmd = ["g++", "-std=c++14", "-fsyntax-only", "-Wall", "-Wno-pragmas", "-Wno-unused-parameter",
           "-DORION_PLATFORM_LINUX=1", f"-I{inc}", f"-I{REF}/engine/3rdparty/glm/include", f"-I{ROOT}/include",
           os.path.join(ROOT, "adapter", "cuda_render_world.cc")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]


@pytest.mark.gpu
def test_cuda_render_world_equals_simple_render_world():
    if not os.path.exists(CHECK):
        pytest.skip("oracle/_ref/adapter_check not built (needs /root/reference at build time)")
    r = subprocess.run([CHECK, "30000"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    assert "ADAPTER OK" in r.stdout
    assert r.stdout.count("compared") == 7


@pytest.mark.gpu
@pytest.mark.parametrize("n,seed", [(1, 11), (7, 12), (300, 13), (5000, 14), (5000, 15), (60000, 16)])
def test_cuda_render_world_on_reseeded_scenarios(n, seed):
    """The same end-to-end comparison on other draws of the scenario, tiny worlds included, with light ranges down to
    centimetres (shadow views whose far plane lies before their near plane)."""
    if not os.path.exists(CHECK):
        pytest.skip("oracle/_ref/adapter_check not built (needs /root/reference at build time)")
    r = subprocess.run([CHECK, str(n), str(seed)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-3000:]
    assert "ADAPTER OK" in r.stdout
