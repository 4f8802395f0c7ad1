"""TEST INFRASTRUCTURE: import the UNMODIFIED reference.

Where from: `$VN_REFERENCE_ROOT` if set, else `/root/reference` (this container), else `baseline/_ref/` - the
byte-for-byte, git-ignored copy `oracle/install_ref.py` makes so that the reference travels to the GPU box
(`/root/reference` does not exist there).  Users: `oracle/make_golden*.py`, `tests/test_oracle_vs_reference.py`,
and `bench.py`'s `--impl reference` arm / `cpu_baseline` leg (the reference itself is what is timed, kind
"reference").  The sources are imported from where they lie; nothing here edits them.

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

_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _find_root() -> str:
    cands = [os.environ.get("VN_REFERENCE_ROOT"), "/root/reference", os.path.join(_REPO, "baseline", "_ref")]
    for c in cands:
        if c and os.path.isfile(os.path.join(c, "voxelnet", "utils.py")):
            return c
    return "/root/reference"


REFERENCE_ROOT = _find_root()
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
