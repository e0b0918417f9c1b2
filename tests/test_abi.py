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
