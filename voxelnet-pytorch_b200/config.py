"""Grid and model constants of the hot path, restated as plain values.

Reference: voxelnet/config.py:35-76 (cfg.OBJECT.* for Car and Pedestrian/Cyclist) and the
hard-coded per-class constants of the voxelizer, voxelnet/utils.py:24-33.  The reference picks
the class at import time from a constant (`_C.OBJECT.NAME = "Car"`, config.py:36); here the
class is an explicit argument everywhere.
"""
from __future__ import annotations

from dataclasses import dataclass


@dataclass(frozen=True)
class GridConfig:
    name: str
    depth: int    # cfg.OBJECT.DEPTH   (z cells)
    height: int   # cfg.OBJECT.HEIGHT  (y cells)
    width: int    # cfg.OBJECT.WIDTH   (x cells)
    points_per_voxel: int  # cfg.OBJECT.POINTS_PER_VOXEL (T)
    x_range: tuple
    y_range: tuple
    z_range: tuple
    voxel_size: tuple = (0.2, 0.2, 0.4)  # (x, y, z)

    @property
    def cells(self) -> int:
        return self.depth * self.height * self.width


CAR = GridConfig("Car", 10, 400, 352, 35, (0.0, 70.4), (-40.0, 40.0), (-3.0, 1.0))
PED_CYC = GridConfig("Pedestrian", 10, 200, 240, 45, (0.0, 48.0), (-20.0, 20.0), (-3.0, 1.0))


def grid_config(target: str) -> GridConfig:
    """utils.py:24-33: 'Car' -> Car grid; any other class name -> Pedestrian/Cyclist grid."""
    return CAR if target == "Car" else PED_CYC


# The three ConvMD(3, cin, cout, 3, stride, padding) of model.py:207-209
MIDDLE_LAYERS = (
    dict(cin=128, cout=64, stride=(2, 1, 1), padding=(1, 1, 1)),
    dict(cin=64, cout=64, stride=(1, 1, 1), padding=(0, 1, 1)),
    dict(cin=64, cout=64, stride=(2, 1, 1), padding=(1, 1, 1)),
)
