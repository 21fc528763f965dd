"""corsair_b200: B200-native correspondence-and-registration hot path of CORSAIR.

Drop-in modules with the reference's names and signatures (utils/find_nn.py, utils/eval_pose.py,
utils/preprocess.py, utils/symmetry.py, utils/retrieval.py, utils/pc_dist.py,
utils/_transform_estimation.py); every kernel runs in the sm_100a C-ABI library
(include/corsair_b200.h).  ``corsair_b200.install()`` aliases them over ``utils.*`` so the
reference's evaluation scripts pick them up unmodified.
"""
from ._capi import Context, CorsairError, get_context, load_library  # noqa: F401

__version__ = "0.1.0"


def install(package="utils"):
    """Alias the drop-in modules as ``<package>.find_nn`` etc. in ``sys.modules``."""
    import importlib
    import sys
    import types

    pkg = sys.modules.get(package)
    if pkg is None:
        pkg = types.ModuleType(package)
        pkg.__path__ = []
        sys.modules[package] = pkg
    for name in ("find_nn", "eval_pose", "preprocess", "symmetry", "retrieval", "pc_dist",
                 "_transform_estimation"):
        mod = importlib.import_module("corsair_b200." + name)
        sys.modules[package + "." + name] = mod
        setattr(pkg, name, mod)
    return pkg
