"""Import the reference's own ``utils.*`` modules from /root/reference (this container only).

TEST INFRASTRUCTURE.  The reference imports ``open3d``, ``transforms3d``, ``MinkowskiEngine``,
``matplotlib`` and ``vedo`` at module scope; none is installed.  Empty stub modules are
registered so the hot-path modules import; the only hot-path call that actually *needs* one
of them is ``registration_based_on_corr`` (Open3D RANSAC, utils/eval_pose.py:95), which the
caller patches with ``oracle.ransac_np``.  ``transforms3d.euler.euler2mat`` gets a small
restatement (rotation about +y) so ``eval_pose`` (utils/eval_pose.py:103-128) runs.

Never used on the GPU box: /root/reference does not exist there.
"""
import importlib
import os
import sys
import types

import numpy as np

REF_ROOT = "/root/reference"


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "utils"))


def _euler2mat(ai, aj, ak, axes="sxyz"):
    # static xyz: R = Rz(ak) @ Ry(aj) @ Rx(ai)
    ci, si = np.cos(ai), np.sin(ai)
    cj, sj = np.cos(aj), np.sin(aj)
    ck, sk = np.cos(ak), np.sin(ak)
    rx = np.array([[1, 0, 0], [0, ci, -si], [0, si, ci]])
    ry = np.array([[cj, 0, sj], [0, 1, 0], [-sj, 0, cj]])
    rz = np.array([[ck, -sk, 0], [sk, ck, 0], [0, 0, 1]])
    return rz @ ry @ rx


def _install_stubs():
    def stub(name):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__path__ = []
            sys.modules[name] = m
        return sys.modules[name]

    for name in ("open3d", "MinkowskiEngine", "vedo", "matplotlib", "matplotlib.cm",
                 "matplotlib.pyplot", "transforms3d", "transforms3d.euler"):
        stub(name)
    sys.modules["matplotlib"].cm = sys.modules["matplotlib.cm"]
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["transforms3d"].euler = sys.modules["transforms3d.euler"]
    sys.modules["transforms3d.euler"].euler2mat = _euler2mat


_cache = {}


def load(name):
    """Return the reference module ``utils.<name>`` (e.g. ``load('find_nn')``)."""
    if not available():
        raise RuntimeError("reference tree not present (expected on the GPU box)")
    if name in _cache:
        return _cache[name]
    _install_stubs()
    saved = sys.modules.get("utils")
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    mod = importlib.import_module("utils." + name)
    _cache[name] = mod
    return mod
