"""orion_b200 - B200-native scene-visibility path for the Orion engine.

``orion_b200.lib`` binds the C ABI of ``liborion_vis.so`` (CUDA, sm_100a); ``orion_b200.scenes`` builds
the synthetic BASELINE.json scenes. Nothing here computes visibility on the CPU.
"""
from . import scenes  # noqa: F401
