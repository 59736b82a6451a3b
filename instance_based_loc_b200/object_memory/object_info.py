"""Mirror of the reference's object_memory/object_info.py ``ObjectInfo`` (:7-118): the per-object record of the
memory (names, stored embeddings, point cloud).  The cloud is a device-resident PointCloud; ``pcd`` / ``pcd_colors``
(the reference's 3xN numpy views, :72-74) are materialised lazily because nothing on the localise path reads them.
"""
from __future__ import annotations

import os
import pickle

import numpy as np

from ..o3d import io as _io
from ..o3d.pointcloud import PointCloud
from ..utils.depth_utils import voxel_down_sample_with_colors


class ObjectInfo:
    # Pickles name the reference's class path (object_memory/object_info.py:7), so a memory bank written here loads in the
    # unmodified reference and vice versa (ObjectMemory.read_pkl maps the path back to this class).
    __module__ = "object_memory.object_info"

    def __init__(self, id, name, emb, pointcloud, max_embeddings_num):
        self._version = 0                       # bumped by every mutation; ObjectMemory's caches key on it
        self.id = id
        self.names = [name]
        self.embeddings = [emb]
        self.pointcloud = pointcloud
        self.max_embeddings_num = max_embeddings_num
        self.mean_emb = None
        self.centroid = None
        self._compute_means()

    def __repr__(self):
        n = 0 if self.pointcloud is None else self.pointcloud.n()
        return f"ObjectInfo == ID: {self.id}, Names: {self.names}, Mean_Emb: {self.mean_emb.shape}, Num. Points: (3, {n})"

    def __setstate__(self, state):
        """Unpickling (ObjectMemory.load): records written by the reference carry its eager ``pcd`` / ``pcd_colors``
        host copies (empty for a save_to_pkl record) -- views here, so they are dropped."""
        state = dict(state)
        state.pop("pcd", None)
        state.pop("pcd_colors", None)
        self.__dict__.update(state)
        self.__dict__.setdefault("_version", 0)

    def __getstate__(self):
        """What the reference's ObjectInfo holds (:9-19, :72-74): public attributes plus the eager 3xN host copies."""
        state = {k: v for k, v in self.__dict__.items() if not k.startswith("_")}
        empty = np.zeros((3, 0))
        state["pcd"] = empty if self.pointcloud is None else self.pcd
        state["pcd_colors"] = empty if self.pointcloud is None else self.pcd_colors
        return state

    def __setattr__(self, key, value):
        if key in ("pointcloud", "embeddings"):
            self.__dict__["_version"] = self.__dict__.get("_version", 0) + 1
        object.__setattr__(self, key, value)

    def _touch(self):
        self._version = self.__dict__.get("_version", 0) + 1

    # the reference keeps eager 3xN host copies; here they are views on demand
    @property
    def pcd(self):
        return np.asarray(self.pointcloud.points).T

    @property
    def pcd_colors(self):
        return np.asarray(self.pointcloud.colors).T

    def _process_pointcloud(self):
        pass

    def _add_name(self, new_name):
        if new_name not in self.names:
            self.names.append(new_name)

    def _add_names(self, new_names):
        for new_name in new_names:
            self._add_name(new_name)

    def _add_embedding(self, new_emb):
        """:38-57 (memory build only)."""
        self._touch()
        if len(self.embeddings) < self.max_embeddings_num:
            self.embeddings.append(new_emb)
            return
        E = np.array(self.embeddings)
        d = np.linalg.norm(E - np.asarray(new_emb).reshape(1, -1), axis=1)
        order = np.argsort(d, kind="stable")
        j = int(order[1])                       # "the second closest (least similar)"
        dj = np.linalg.norm(E - E[j][None], axis=1)
        if np.min(dj) < d[j]:
            self.embeddings[j] = new_emb

    def _add_embeddings(self, new_embs):
        self.embeddings += new_embs          # (an assignment: __setattr__ bumps the version)

    def _add_pointcloud(self, new_pointcloud):
        merged = PointCloud()
        merged += PointCloud.from_tensors(self.pointcloud._p, self.pointcloud._c)
        merged += PointCloud.from_tensors(new_pointcloud._p, new_pointcloud._c)
        self.pointcloud = merged

    def _compute_means(self):
        """:76-78."""
        self.mean_emb = np.mean(np.array(self.embeddings), axis=0).squeeze()
        self.centroid = self.pointcloud.get_center() if self.pointcloud is not None else None

    def __add__(self, new_obj_info):
        self._add_names(new_names=new_obj_info.names)
        self._add_embeddings(new_embs=new_obj_info.embeddings)
        self._add_pointcloud(new_pointcloud=new_obj_info.pointcloud)
        return self

    def downsample(self, voxel_size):
        """:87-89."""
        self.pointcloud = voxel_down_sample_with_colors(self.pointcloud, voxel_size)

    def add_info(self, new_name, new_emb, new_pointcloud, align=False, max_iteration=30, max_correspondence_distance=0.01):
        if align:
            raise NotImplementedError("Aligning is a To-Do")
        self._add_name(new_name)
        self._add_embedding(new_emb)
        self._add_pointcloud(new_pointcloud)
        self._compute_means()

    def update_pointcloud_with_mask(self, mask):
        import torch
        keep = torch.as_tensor(np.asarray(mask).astype(np.uint8)).to(self.pointcloud._p.device)
        self.pointcloud = self.pointcloud._select_mask(keep)[0]

    def save(self, save_directory):
        """:109-118: pointcloud.ply (double xyz + uchar rgb) + info.pkl."""
        os.makedirs(save_directory, exist_ok=True)
        _io.write_point_cloud(os.path.join(save_directory, "pointcloud.ply"), self.pointcloud)
        with open(os.path.join(save_directory, "info.pkl"), "wb") as f:
            pickle.dump({"names": self.names, "embeddings": self.embeddings,
                         "max_embeddings_num": self.max_embeddings_num}, f)
