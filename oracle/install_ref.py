"""TEST / BASELINE INFRASTRUCTURE (container only): place the UNMODIFIED reference where the GPU box can see it.

    python -m oracle.install_ref [--frames N]

`/root/reference` does not exist on the GPU box, but `gpurun` ships everything under `/root/repo` that is not
listed in `.gpurunignore`.  This recipe copies - byte for byte, nothing edited - the five modules of the reference's
hot path (voxelnet/{config,utils,dataset,model,loss}.py), its LICENSE, and some frames of its bundled KITTI drive into
`baseline/_ref/` (git-ignored: the copy never enters this repository's history; SURVEY.md section 8(c) "Shipping the
oracle to the GPU box").  On the box

  * `bench.py --impl reference` and the `cpu_baseline` leg time THAT code (`cpu_baseline.kind = "reference"`),
  * `tests/test_oracle_vs_reference.py` pins the oracle against it again on the box's numpy / torch / host CPU,
  * `bench.py --data kitti` reads real frames from `baseline/_ref/data`.

`oracle/ref_import.py` finds it (`/root/reference` first, then `baseline/_ref`, or `$VN_REFERENCE_ROOT`).
`__graft_entry__.build()` runs this recipe whenever `/root/reference` is present.
"""
from __future__ import annotations

import filecmp
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference"
DST = os.path.join(ROOT, "baseline", "_ref")
MODULES = ("config.py", "utils.py", "dataset.py", "model.py", "loss.py")
KITTI_REL = "data/2011_09_26/2011_09_26_drive_0001_sync/velodyne_points/data"
# frames the container tests read (0-3, 57, 107) + two batches of 16 for `bench.py --data kitti`
DEFAULT_FRAMES = tuple(range(32)) + (57, 107)


def _copy(src: str, dst: str) -> bool:
    if os.path.exists(dst) and filecmp.cmp(src, dst, shallow=False):
        return False
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    shutil.copyfile(src, dst)
    return True


def install(frames=DEFAULT_FRAMES, verbose: bool = False) -> str:
    """Returns the destination root.  Raises when /root/reference is absent (nothing to install from)."""
    if not os.path.isfile(os.path.join(SRC, "voxelnet", "utils.py")):
        raise RuntimeError("reference not present at %s" % SRC)
    n = 0
    for m in MODULES:
        n += _copy(os.path.join(SRC, "voxelnet", m), os.path.join(DST, "voxelnet", m))
    n += _copy(os.path.join(SRC, "LICENSE"), os.path.join(DST, "LICENSE"))
    for i in frames:
        name = "%010d.bin" % i
        n += _copy(os.path.join(SRC, KITTI_REL, name), os.path.join(DST, KITTI_REL, name))
    with open(os.path.join(DST, "README"), "w") as f:
        f.write("Unmodified copy of johanngerberding/voxelnet-pytorch files made by oracle/install_ref.py.\n"
                "Git-ignored; shipped to the GPU box by gpurun so that the reference itself is timed there.\n")
    if verbose:
        print("baseline/_ref: %d file(s) written, %d KITTI frames present" % (n, len(frames)))
    return DST


if __name__ == "__main__":
    k = None
    if "--frames" in sys.argv:
        k = int(sys.argv[sys.argv.index("--frames") + 1])
    install(DEFAULT_FRAMES if k is None else tuple(range(k)), verbose=True)
