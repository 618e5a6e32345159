"""B200-native drop-in for the HT-3DGS rasterizer hot path.

The directory name starts with a digit, so import it with
``importlib.import_module("3dgs_hierarchical_training_b200")`` or, for the
drop-in plugin packages themselves, put this directory on ``sys.path``:

    import sys, importlib
    pkg = importlib.import_module("3dgs_hierarchical_training_b200")
    pkg.install()          # sys.path.insert(0, <this dir>)
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from simple_knn._C import distCUDA2

which is exactly what scene/gaussian_model_ht.py:20,39-42 and
gaussian_renderer/__init__.py:12 of the reference import.
"""
import os
import sys

PLUGIN_DIR = os.path.dirname(os.path.abspath(__file__))


def install():
    """Make `diff_gaussian_rasterization`, `simple_knn` and `htgs_native` importable."""
    if PLUGIN_DIR not in sys.path:
        sys.path.insert(0, PLUGIN_DIR)
    return PLUGIN_DIR
