"""A fixed handful of cases of the three randomised differential tools (tools/fuzz_multiview.py, fuzz_world.py,
fuzz_raster.py), the ones that found faults included, so that those regimes stay covered by the suite; the tools themselves
run hundreds of cases in a few GPU minutes."""
import os
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


@pytest.fixture(scope="module")
def oracle():
    from oracle import pyoracle
    return pyoracle.load("best")


@pytest.fixture(autouse=True)
def _clean_env():
    yield
    for k in ("OVIS_MV_DENSE", "OVIS_MV_ROUND_VIEWS"):      # the many-view tool sets these per case
        os.environ.pop(k, None)


def test_many_view_cases(oracle):
    import fuzz_multiview
    # 10009: an inverted box inside large light cubes with the grid forced (was filed under an empty cell range)
    # 120114: a light whose range is ~1e6 x the shadow near plane (the far planes keep three bits: the cube follows the planes)
    for seed in [10009, 120114] + list(range(70_000, 70_024)):
        desc, err = fuzz_multiview.run_case(oracle, seed, 4e7)
        assert err is None, (seed, desc, err)


def test_whole_frame_cases(oracle):
    import fuzz_world
    for seed in range(80_000, 80_040):
        desc, err = fuzz_world.run_case(oracle, seed)
        assert err is None, (seed, desc, err)


def test_rasteriser_cases(oracle):
    import fuzz_raster
    # 10008 ... 10106: hundreds of instances far smaller than a pixel (the bounding box is part of the definition)
    for seed in [10008, 10019, 10053, 10072, 10106] + list(range(90_000, 90_030)):
        desc, err = fuzz_raster.run_case(oracle, seed)
        assert err is None, (seed, desc, err)
    for seed in range(91_000, 91_015):                      # textured and normal-mapped materials
        desc, err = fuzz_raster.run_textured_case(oracle, seed)
        assert err is None, (seed, desc, err)


def test_single_view_cull_cases(oracle):
    import fuzz_cull
    others = []
    for seed in range(60_000, 60_040):
        desc, err = fuzz_cull.run_case(oracle, seed, others)
        assert err is None, (seed, desc, err)
    for w in others:
        w.close()


def test_sharded_world_cases(oracle):
    import fuzz_sharded
    for seed in range(50_000, 50_030):
        desc, err = fuzz_sharded.run_case(oracle, seed)
        assert err is None, (seed, desc, err)


def test_light_pass_cases(oracle):
    import fuzz_shading
    for seed in range(40_000, 40_030):
        desc, err = fuzz_shading.run_case(oracle, seed)
        assert err is None, (seed, desc, err)
