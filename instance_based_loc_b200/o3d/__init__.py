"""Device-resident stand-in for the slice of ``open3d`` the reference's localisation path uses
(SURVEY.md section 8(b).2).  Same names and call signatures as open3d 0.17 so the reference's
modules and drivers run unedited, but every cloud lives in HBM as torch CUDA float64 tensors and
every operation is a libibl.so kernel (include/ibl.h).  No CPU fallback: importing is cheap,
but any geometric call raises if libibl.so or a GPU is missing.
"""
from __future__ import annotations

import types as _types

import numpy as _np
import torch as _torch

from .. import kernels as _k
from . import registration as _reg
from .pointcloud import (KDTreeSearchParamHybrid, KDTreeSearchParamKNN, KDTreeSearchParamRadius, PointCloud,
                         Vector2iVector, Vector3dVector)
from . import ply as _ply

__version__ = "0.17.0+ibl_b200"

geometry = _types.SimpleNamespace(PointCloud=PointCloud, KDTreeSearchParamHybrid=KDTreeSearchParamHybrid,
                                  KDTreeSearchParamKNN=KDTreeSearchParamKNN,
                                  KDTreeSearchParamRadius=KDTreeSearchParamRadius)
utility = _types.SimpleNamespace(Vector3dVector=Vector3dVector, Vector2iVector=Vector2iVector)
pipelines = _types.SimpleNamespace(registration=_reg)
io = _types.SimpleNamespace(write_point_cloud=_ply.write_point_cloud, read_point_cloud=_ply.read_point_cloud)
