"""Loader for the reference's own caller modules (not a test module).

tools/stage_reference_callers.py copies gaussian_renderer/, scene/, utils/, arguments/ of the reference, unmodified,
into baseline/_ref/htgs_callers/.  `load()` puts that directory on sys.path next to this repository's plugin packages,
installs import stubs for the third-party packages the image does not have, and returns the imported modules.

The stubs only satisfy `import` statements of code paths the tests never execute (SE3 pose parameters, PLY I/O through
plyfile, matplotlib plotting, ...): a stubbed module hands out MagicMock attributes.  Nothing of the rasterizer path
is stubbed -- `diff_gaussian_rasterization` and `simple_knn._C` resolve to the packages of this repository."""
import importlib
import importlib.abc
import importlib.machinery
import os
import sys
import types
from unittest import mock

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STAGED = os.path.join(ROOT, "baseline", "_ref", "htgs_callers")
# packages imported somewhere in the staged tree that this image lacks
STUBBED = ("lietorch", "plyfile", "matplotlib", "mpl_toolkits", "open3d", "kornia", "pytorch3d", "imageio", "timm", "evo",
           "roma", "trimesh", "mediapy", "lpips", "lpipsPyTorch", "configargparse", "torchmetrics", "skimage")


def staged():
    return os.path.isdir(os.path.join(STAGED, "gaussian_renderer"))


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        m = mock.MagicMock(name=f"{self.__name__}.{name}")
        setattr(self, name, m)
        return m


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    """Serves `import X[.y.z]` for every X in STUBBED that is really absent."""

    def __init__(self, names):
        self.names = set(names)

    def find_spec(self, fullname, path=None, target=None):
        if fullname.split(".")[0] in self.names:
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        m = _StubModule(spec.name)
        m.__path__ = []
        return m

    def exec_module(self, module):
        pass


_loaded = None


def load():
    """Returns a namespace with `gaussian_renderer`, `gaussian_model_ht`, `cameras`, `graphics_utils`, `stubbed`."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not staged():
        raise RuntimeError("reference callers not staged: run tools/stage_reference_callers.py where /root/reference exists")
    importlib.import_module("3dgs_hierarchical_training_b200").install()
    missing = []
    for name in STUBBED:
        try:
            if importlib.util.find_spec(name) is None:
                missing.append(name)
        except (ImportError, ValueError):
            missing.append(name)
    sys.meta_path.append(_StubFinder(missing))  # appended: real packages always win
    if STAGED not in sys.path:
        sys.path.insert(1, STAGED)  # behind this repository's plugin directory, ahead of everything else
    for pkg in ("scene", "utils", "gaussian_renderer", "arguments"):
        mod = sys.modules.get(pkg)
        if mod is not None and not getattr(mod, "__file__", "").startswith(STAGED):
            raise RuntimeError(f"a different top-level package `{pkg}` is already imported: {getattr(mod, '__file__', None)}")
    ns = types.SimpleNamespace()
    ns.gaussian_renderer = importlib.import_module("gaussian_renderer")
    ns.gaussian_model_ht = importlib.import_module("scene.gaussian_model_ht")
    ns.cameras = importlib.import_module("scene.cameras")
    ns.graphics_utils = importlib.import_module("utils.graphics_utils")
    ns.sh_utils = importlib.import_module("utils.sh_utils")
    ns.stubbed = missing
    _loaded = ns
    return ns


_trainer = None


def load_trainer():
    """The reference's trainer module (trainer/ht3dgs_trainer.py), unchanged.  Its base module asks torch.hub for the MiDaS
    help text at import time (trainer/trainer.py:40, needs the network): that one call is made a no-op."""
    global _trainer
    if _trainer is None:
        load()
        import torch
        torch.hub.help = lambda *a, **k: None
        _trainer = importlib.import_module("trainer.ht3dgs_trainer")
    return _trainer
