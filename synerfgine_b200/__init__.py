"""synerfgine_b200 — B200-native (sm_100a) implementation of the SyNeRFgine render hot path behind a C ABI.

    from synerfgine_b200 import Testbed, Engine, SyntheticModel

Layout: ``csrc/`` CUDA kernels + C ABI (libsnb.so), ``_lib.py`` ctypes binding, ``testbed.py`` the mirror of the
reference's pyngp.Testbed / sng::Engine surface, ``synthetic.py`` seeded models of the base.json architecture.
"""
from .synthetic import SyntheticModel  # noqa: F401


def __getattr__(name):  # torch-dependent classes are imported on first use
    if name == "Scene":
        from .scene import Scene
        return Scene
    if name in ("Testbed", "Engine"):
        from . import testbed
        return getattr(testbed, name)
    raise AttributeError(name)
