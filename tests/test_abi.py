"""CPU-side checks of the drop-in boundary: the shared library loads, exports every symbol
include/orion_vis.h declares, its host-side view/light set-up equals the oracle, and it refuses to
run without a CUDA device instead of falling back to the CPU."""
import os
import re

import numpy as np
import pytest

from orion_b200 import lib, scenes
import vis_testutil as util
from vis_testutil import same


def test_library_exports_every_declared_symbol():
    names = lib.declared_symbols()
    assert len(names) >= 30
    so = lib.load()
    missing = [n for n in names if not hasattr(so, n)]
    assert not missing, missing
    assert b"sm_100a" in so.ovis_version()


def test_struct_layouts_match_header():
    assert lib.LIGHT_DTYPE.itemsize == 56 and lib.VIEW_DTYPE.itemsize == 48


def test_host_view_math_equals_oracle(port):
    rng = np.random.default_rng(21)
    for v in list(util.random_views(rng, 200)) + [scenes.default_camera(), scenes.cubes_scene()["view"]]:
        v = np.asarray(v)
        assert same(lib.view_planes(v), port.view_planes(v))
        m, c = lib.view_matrices(v)
        pm, pc = port.view_matrices(v)
        assert same(m, pm) and same(c, pc)


def test_host_light_math_equals_oracle(port):
    rng = np.random.default_rng(22)
    lights = util.random_lights(rng, 300)
    for l in lights:
        l = np.asarray(l)
        if l["type"] == scenes.LIGHT_SPOT:
            assert same(lib.light_spot_aabb(l), port.light_spot_aabb(l))
        assert lib.light_shadow_views(l).tobytes() == port.light_shadow_views(l).tobytes()


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(lib.OvisError) as e:
        lib.World(0)
    assert e.value.code == lib.OVIS_ERR_CUDA
    assert "no CPU path" in str(e.value)


def test_product_package_never_touches_the_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(lib.__file__)))
    for dirpath, _, files in os.walk(os.path.join(root, "orion_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cc", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "pyoracle" not in text and "ovis_oracle" not in text and "liborion_ref" not in text, f


def test_header_is_plain_c(tmp_path):
    """include/orion_vis.h is the drop-in boundary: it must compile as C99 (no C++ or CUDA types in a signature), and its
    structs must have the sizes the bindings assume."""
    import shutil
    import subprocess
    cc = shutil.which("gcc") or shutil.which("cc")
    if not cc:
        pytest.skip("no C compiler")
    src = tmp_path / "abi.c"
    src.write_text('#include "orion_vis.h"\n#include <stdio.h>\n'
                   'int main(void) { printf("%zu %zu %zu %zu %zu\\n", sizeof(OvisView), sizeof(OvisLight), sizeof(OvisMaterial), '
                   'sizeof(OvisSceneInfo), sizeof(OvisGBufferTarget)); return 0; }\n')
    exe = tmp_path / "abi"
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([cc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", f"-I{root}/include", str(src), "-o", str(exe)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()]
    assert sizes[:3] == [lib.VIEW_DTYPE.itemsize, lib.LIGHT_DTYPE.itemsize, lib.MATERIAL_DTYPE.itemsize]
    assert sizes[3] == C_sizeof_scene_info() and sizes[4] == 40


def C_sizeof_scene_info():
    import ctypes
    return ctypes.sizeof(lib.SceneInfo)
