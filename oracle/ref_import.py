"""TEST INFRASTRUCTURE (container-only): import the UNMODIFIED reference from /root/reference.

Only `oracle/make_golden.py` and `tests/test_oracle_vs_reference.py` use this module, and only
when `/root/reference` exists (it does not exist on the GPU box).  Nothing is copied: the
reference sources are imported from where they lie.

The reference needs `yacs` (voxelnet/config.py:1), which is not installed in this image; a
minimal stand-in for `yacs.config.CfgNode` (attribute-access dict with `clone()`) is injected
into `sys.modules` before the import.  `np.random.shuffle` (voxelnet/utils.py:35) is replaced by
a no-op while `pcl_to_voxels` runs so that point order - and with it the first-T point selection
(voxelnet/utils.py:78-84) - is deterministic.
"""
from __future__ import annotations

import contextlib
import copy
import os
import sys
import types

REFERENCE_ROOT = "/root/reference"
REFERENCE_PKG = os.path.join(REFERENCE_ROOT, "voxelnet")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_PKG, "utils.py"))


class _CfgNode(dict):
    """Just enough of yacs.config.CfgNode for voxelnet/config.py."""

    def __getattr__(self, name):
        try:
            return self[name]
        except KeyError as exc:  # pragma: no cover
            raise AttributeError(name) from exc

    def __setattr__(self, name, value):
        self[name] = value

    def clone(self):
        return copy.deepcopy(self)

    def freeze(self):  # pragma: no cover
        pass

    def merge_from_file(self, path):  # pragma: no cover
        raise NotImplementedError


def _install_yacs_shim() -> None:
    if "yacs.config" in sys.modules:
        return
    yacs = types.ModuleType("yacs")
    yacs_config = types.ModuleType("yacs.config")
    yacs_config.CfgNode = _CfgNode
    yacs.config = yacs_config
    sys.modules["yacs"] = yacs
    sys.modules["yacs.config"] = yacs_config


_REF_MODULE_NAMES = ("config", "utils", "dataset", "model", "loss")


def load_reference():
    """Return a namespace with the reference's `utils`, `dataset`, `model` modules.

    The reference uses flat absolute imports (`from config import ...`), so its directory is
    put on sys.path for the duration of the import and the modules are then re-registered
    under private names so that they cannot shadow anything of ours called `utils`/`model`.
    """
    if not reference_available():
        raise RuntimeError("reference not present at %s" % REFERENCE_PKG)
    cached = sys.modules.get("_vnref_ns")
    if cached is not None:
        return cached
    _install_yacs_shim()
    saved = {n: sys.modules.pop(n) for n in _REF_MODULE_NAMES if n in sys.modules}
    sys.path.insert(0, REFERENCE_PKG)
    try:
        import importlib

        mods = {n: importlib.import_module(n) for n in ("config", "utils", "dataset", "model")}
    finally:
        sys.path.remove(REFERENCE_PKG)
        for n in _REF_MODULE_NAMES:
            m = sys.modules.pop(n, None)
            if m is not None:
                sys.modules["_vnref_" + n] = m
        sys.modules.update(saved)
    ns = types.SimpleNamespace(**mods)
    sys.modules["_vnref_ns"] = ns
    return ns


@contextlib.contextmanager
def no_shuffle():
    """Disable the in-place shuffle of voxelnet/utils.py:35."""
    import numpy as np

    orig = np.random.shuffle
    np.random.shuffle = lambda x: None
    try:
        yield
    finally:
        np.random.shuffle = orig


def ref_pcl_to_voxels(pcl, target: str):
    """Run the reference voxelizer on a copy of `pcl` with its shuffle disabled."""
    ref = load_reference()
    with no_shuffle():
        return ref.utils.pcl_to_voxels(pcl.copy(), target, False)


@contextlib.contextmanager
def ref_model_cfg(target: str):
    """Patch the model module's cfg clone for Ped/Cyc (read at call time: model.py:78,103)."""
    ref = load_reference()
    cfg = ref.model.cfg
    saved = (cfg.OBJECT.POINTS_PER_VOXEL, cfg.OBJECT.HEIGHT, cfg.OBJECT.WIDTH)
    if target != "Car":
        cfg.OBJECT.POINTS_PER_VOXEL, cfg.OBJECT.HEIGHT, cfg.OBJECT.WIDTH = 45, 200, 240
    try:
        yield ref
    finally:
        cfg.OBJECT.POINTS_PER_VOXEL, cfg.OBJECT.HEIGHT, cfg.OBJECT.WIDTH = saved


def kitti_frame_path(i: int) -> str:
    return os.path.join(
        REFERENCE_ROOT,
        "data/2011_09_26/2011_09_26_drive_0001_sync/velodyne_points/data/%010d.bin" % i,
    )
