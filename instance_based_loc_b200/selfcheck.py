"""Helpers shared by bench.py, the tools and the tests: an ObjectMemory built from a synthetic scene, and small pose
utilities.  Nothing here touches oracle/ -- the comparison against the CPU oracle lives in tests/ibl_selfcheck.py."""
from __future__ import annotations

import numpy as np
import torch


def build_memory(scene, spacing=0.01, voxel=0.01, bug_compatible=True, f=300.0):
    """-> (ObjectMemory on device, oracle-side memory list) built from the same host arrays."""
    from . import synth
    from .o3d.pointcloud import PointCloud
    from .object_memory.object_memory import ObjectMemory
    from .utils.depth_utils import voxel_down_sample_with_colors
    om = ObjectMemory("cuda", None, None, f, f, get_embeddings_func=None, log_enabled=False, bug_compatible=bug_compatible)
    host = []
    for j in range(scene.m):
        p, c = synth.memory_cloud(scene, j, spacing)
        pc = PointCloud(p)
        pc.colors = c
        pc = voxel_down_sample_with_colors(pc, voxel)                 # object_info.py:87-89 at memory-build time
        om.add_object(f"obj{j}", scene.memory_embs[j][0], pc, embeddings=list(scene.memory_embs[j]))
        host.append({"points": np.asarray(pc.points), "colors": np.asarray(pc.colors),
                     "embeddings": scene.memory_embs[j]})
    return om, host


def rot_deg(Ra, Rb):
    return float(np.degrees(np.arccos(np.clip((np.trace(Ra.T @ Rb) - 1) / 2, -1, 1))))


def pose_matrix(pose):
    from scipy.spatial.transform import Rotation
    T = np.eye(4)
    T[:3, :3] = Rotation.from_quat(pose[3:]).as_matrix()
    T[:3, 3] = pose[:3]
    return T
