"""Mirror of the reference's ``ObjectMemory`` query path (object_memory/object_memory.py:60-161, 779-846,
852-1169): same constructor / ``localise`` / ``load`` / ``save_to_pkl`` signatures and the same result-affecting
behaviour (including the two bugs SURVEY.md section 0 lists, behind ``bug_compatible=True``), with the whole per-frame
computation on libibl.so kernels.

What is new relative to the reference (none of it changes results):
  * ``localise_arrays`` -- the same computation from in-memory rgb / depth / masks / embeddings (the detector and
    encoder are out of scope; benchmarks and tests inject synthetic masks and embeddings here);
  * memory-side caches: the normalised embedding bank, the union of all memory clouds with its search grid, and per
    (memory-object tuple) the centred union cloud with its normals, FPFH features, grids and colour gradients.  The
    reference recomputes all of these for every frame and every assignment.
"""
from __future__ import annotations

import contextlib
import copy
import os
import pickle
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch
from scipy.spatial.transform import Rotation

from .. import kernels as K
from ..o3d import registration as _reg
from ..o3d.pointcloud import PointCloud, remove_radius_outlier_many
from ..utils.depth_utils import DEFAULT_OUTLIER_REMOVAL_CONFIG, get_mask_coloured_pointclouds_from_depth, transform_pointcloud
from ..utils.fpfh_register import downsample_and_compute_fpfh, evaluate_transform, register_point_clouds
from ..utils.similarity_volume import SimVolume
from .object_info import ObjectInfo


def default_load_rgb(path):
    """object_memory.py:31-32 (imageio there; PIL here -- imageio is not in this image)."""
    from PIL import Image
    return np.asarray(Image.open(path))


def default_load_depth(path):
    """object_memory.py:34-40."""
    if path.split('.')[-1] == 'npy':
        return np.load(path)
    from PIL import Image
    return np.asarray(Image.open(path))


def check_if_floor(texts):
    """object_memory/object_finder_phrases.py:19-35 -- membership of a whole name in the list, not a substring test."""
    for word in ("floor", "ground", "earth"):
        if word in texts:
            return True
    return False


@contextlib.contextmanager
def _reference_class_path():
    """ObjectInfo pickles under the reference's path ``object_memory.object_info`` (so the unmodified reference can load
    the file); pickle insists that the path resolves while dumping, so it is aliased for the duration of the dump unless
    the caller already installed the drop-in aliases."""
    from . import object_info as _oi
    import sys
    added = []
    for name, mod in (("object_memory", sys.modules[__package__]), ("object_memory.object_info", _oi)):
        if name not in sys.modules:
            sys.modules[name] = mod
            added.append(name)
    try:
        if getattr(sys.modules["object_memory.object_info"], "ObjectInfo", None) is not ObjectInfo:
            raise pickle.PicklingError("another object_memory.object_info is imported in this process")
        yield
    finally:
        for name in added:
            sys.modules.pop(name, None)


class _MemoryUnpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if name == "ObjectInfo" and module.split(".")[-1] == "object_info":
            return ObjectInfo
        return super().find_class(module, name)


class ObjectMemory():
    def __init__(self, device, ram_pretrained_path, sam_checkpoint_path, camera_focal_lenth_x, camera_focal_lenth_y,
                 get_embeddings_func, log_enabled=True, mem_formation_bounding_box_threshold=0.3,
                 mem_formation_occlusion_overlap_threshold=0.9, object_info_max_embeddings_num=1000000,
                 load_rgb_image_func=default_load_rgb, load_depth_image_func=default_load_depth,
                 dataset_floor_thickness=0.1, lora_path=None, object_finder=None, bug_compatible=True):
        self.device = device
        self.ram_pretrained_path = ram_pretrained_path
        self.sam_checkpoint_path = sam_checkpoint_path
        self.camera_focal_lenth_x = camera_focal_lenth_x
        self.camera_focal_lenth_y = camera_focal_lenth_y
        self.log_enabled = log_enabled
        self.mem_formation_bounding_box_threshold = mem_formation_bounding_box_threshold
        self.mem_formation_occlusion_overlap_threshold = mem_formation_occlusion_overlap_threshold
        self.object_info_max_embeddings_num = object_info_max_embeddings_num
        self.get_embeddings_func = get_embeddings_func
        self.load_rgb_image_func = load_rgb_image_func
        self.load_depth_image_func = load_depth_image_func
        self.dataset_floor_thickness = dataset_floor_thickness
        # The detector (RAM + GroundingDINO + SAM, object_memory/object_finder.py) is out of scope: any object with
        # find(rgb_path, consider_floor) -> (crops, boxes, masks[n,1,H,W], phrases) can be injected.
        self.object_finder = object_finder
        self.bug_compatible = bug_compatible
        if self.get_embeddings_func is None and object_finder is not None:
            raise NotImplementedError("Need to pass in get_embeddings_func")
        self.memory: list[ObjectInfo] = []
        self.floors = None
        self._cache_key = None
        self._cache = {}
        self._lock = threading.RLock()
        self._tls = threading.local()
        self._pool, self._pool_workers = None, 0
        self.last_localise_info = {}
        self.cache_stats = {"lookups": 0, "misses": 0}   # _chosen_memory: per-assignment-tuple memory-side features

    def _log(self, statement):
        if self.log_enabled:
            print(f"\033[34m{statement}\033[0m")

    def __repr__(self):
        r = "".join(f"\t{o}\n" for o in self.memory)
        return r if r else "\tNo objects in memory yet."

    # ------------------------------------------------------------------------------------------ memory I/O
    def add_object(self, name, emb, pointcloud, embeddings=None):
        """Append one ObjectInfo (what process_image does after transform_pointcloud, :239-256)."""
        info = ObjectInfo(len(self.memory), name, np.asarray(emb), pointcloud, self.object_info_max_embeddings_num)
        if embeddings is not None:
            info.embeddings = [np.asarray(e) for e in embeddings]
            info._compute_means()
        self.memory.append(info)
        return info

    def downsample_all_objects(self, voxel_size):
        """:258-263."""
        for o in self.memory:
            o.downsample(voxel_size)
        if self.floors is not None:
            self.floors.downsample(voxel_size)

    # ------------------------------------------------------------------------------------------ memory build (8(f) N2)
    def process_image(self, rgb_image_path, depth_image_path, pose, consider_floor,
                      outlier_removal_config=DEFAULT_OUTLIER_REMOVAL_CONFIG, add_noise=False,
                      pose_noise={'trans': 0.0005, 'rot': 0.0005}, depth_noise=0.003, min_points=500,
                      will_cluster_later=True, depth_factor=1.):
        """:163-256 -- detect, back-project and outlier-filter every mask (kernels), move the clouds into the world frame
        (``transform_pointcloud``) and append one ObjectInfo per object with at least ``min_points`` points; floor-named
        objects are merged into ``self.floors``.  The noise option draws from numpy's global generator in the reference's
        order (pose translation, pose rotation, then one draw per cloud), so a seeded run reproduces it."""
        obj_phrases, embs, obj_pointclouds = self._get_object_info(rgb_image_path, depth_image_path, consider_floor,
                                                                   outlier_removal_config, depth_factor=depth_factor)
        if obj_phrases is None:
            self._log("ObjectMemory.process_image did NOT find any objects")
            return
        self._log(f"ObjectMemory.process_image found: {obj_phrases}")
        if add_noise:
            pose[:3] = pose[:3] + np.random.normal(0, pose_noise['trans'], pose[:3].shape)
            q = pose[3:] + np.random.normal(0, pose_noise['rot'], pose[3:].shape)
            nq = np.linalg.norm(q)
            pose[3:] = q if nq == 0 else q / nq
            noisy = []
            for pc in obj_pointclouds:
                noise = np.random.normal(0, depth_noise, (pc.n(), 3))
                noisy.append(PointCloud.from_tensors(pc._p + K.to_dev(noise), pc._c) if pc.n() else pc)
            obj_pointclouds = noisy
        transformed = [transform_pointcloud(pc, pose) for pc in obj_pointclouds]
        for obj_phrase, obj_emb, obj_pointcloud in zip(obj_phrases, embs, transformed):
            if obj_pointcloud.n() < min_points:
                self._log(f"\t\tSkipping as number of points {obj_pointcloud.n()} < min_points = {min_points}.")
                continue
            if not will_cluster_later:
                raise NotImplementedError("Only final clustering available currently")
            new_obj_info = ObjectInfo(len(self.memory), obj_phrase, obj_emb, obj_pointcloud,
                                      self.object_info_max_embeddings_num)
            if check_if_floor(new_obj_info.names):
                self.floors = new_obj_info if self.floors is None else self.floors + new_obj_info
                self._log(f"\tFloor Added: {new_obj_info}")
            else:
                self.memory.append(new_obj_info)
                self._log(f"\tObject Added: {new_obj_info}")

    def remove_points_below_floor(self):
        """:265-291 -- the lowest y over all objects (min/max kernel per cloud) plus ``dataset_floor_thickness`` is the cut;
        points at or below it are dropped, objects left empty are removed (while iterating, as in the reference)."""
        floor_height = float('inf')
        for info in self.memory:
            floor_height = min(float(K.bounds3(info.pointcloud._p)[0][1]), floor_height)
        for info in self.memory:
            keep = (info.pointcloud._p[:, 1] > floor_height + self.dataset_floor_thickness).to(torch.uint8)
            info.pointcloud = info.pointcloud._select_mask(keep.contiguous())[0]
            if info.pointcloud.n() == 0:
                self.memory.remove(info)

    def _dbscan_merge(self, objects, eps, min_points_per_cluster):
        """The block the reference repeats at :306-362, :498-547 and :619-668: DBSCAN over the union of the objects'
        clouds (``ibl_dbscan``), every object goes to the cluster of its FIRST point, the objects of one cluster are merged
        in list order, objects whose first point is noise are dropped.  The reference finds that cluster by testing, per
        label in ascending order, whether the first point's coordinates occur among the cluster's points and keeps the LAST
        label that matches (its "multiple labels" warning); the same answer read off the labels directly is the largest
        label over the points equal to the object's first point."""
        clouds = [o.pointcloud._p for o in objects]
        union = PointCloud.from_tensors(torch.cat(clouds, dim=0).contiguous())
        labels, n_clusters = union.cluster_dbscan_device(eps, min_points_per_cluster)
        assignments = np.full(len(objects), -1)
        for index, c in enumerate(clouds):
            same = (union._p == c[0:1]).all(dim=1)
            assignments[index] = int(labels[same].max().item())
        merged = []
        for label in range(n_clusters):
            to_combine = [objects[i] for i in range(len(objects)) if assignments[i] == label]
            if not to_combine:
                continue
            combined = to_combine[0]
            for obj_info in to_combine[1:]:
                combined = combined + obj_info
            merged.append(combined)
        return merged

    def _renumber(self):
        for i, obj_info in enumerate(self.memory):
            obj_info.id = i

    def recluster_objects_with_dbscan(self, eps=0.2, min_points_per_cluster=300, visualize=False):
        """:296-372."""
        self._log("Clustering using DBSCAN")
        self.memory = self._dbscan_merge(self.memory, eps, min_points_per_cluster)
        self._log(f"Updated memory size: {len(self.memory)}")
        self._renumber()

    def _embedding_labels(self, embedding_distance_threshold, one_minus):
        """The m x m part of :379-427 / :444-478: cosine matrix of the objects' mean embeddings (``1 -`` it in the
        combined variants), shifted and scaled to [0, 1], reversed, then scikit-learn's average-linkage agglomerative
        clustering on it -- the library and call the reference itself uses; m is the number of objects, host work."""
        from sklearn.cluster import AgglomerativeClustering
        all_obj_embs = np.array([obj.mean_emb for obj in self.memory])
        norms = np.linalg.norm(all_obj_embs, axis=1, keepdims=True)
        normalized_embeddings = all_obj_embs / norms
        distance_matrix = np.dot(normalized_embeddings, normalized_embeddings.T)
        if one_minus:
            distance_matrix = 1 - distance_matrix
        distance_matrix -= np.min(distance_matrix)
        distance_matrix /= np.max(distance_matrix)
        distance_matrix = 1 - distance_matrix
        agg = AgglomerativeClustering(n_clusters=None, distance_threshold=embedding_distance_threshold,
                                      metric='precomputed', linkage='average')
        agg.fit(distance_matrix)
        return agg.labels_

    def _merge_by_labels(self, labels):
        """:419-432 -- the first object of a label absorbs the later ones, result ordered by label."""
        new_memory = [None for _ in set(labels)]
        for new_label, old_obj_info in zip(labels, self.memory):
            if new_memory[new_label] is None:
                new_memory[new_label] = old_obj_info
            else:
                new_memory[new_label] = new_memory[new_label] + old_obj_info
        self.memory = new_memory
        self._renumber()

    def recluster_via_agglomerative_clustering(self, distance_func=None, embedding_distance_threshold=0.4,
                                               distance_threshold=0.1):
        """:379-432.  Only the default distance function exists in the reference (a custom one is never called there:
        the matrix always comes from the local ``df``), so ``distance_func`` is accepted and ignored."""
        self._merge_by_labels(self._embedding_labels(embedding_distance_threshold, one_minus=False))

    def recluster_via_combined(self, distance_func=None, embedding_distance_threshold=0.4, eps=0.4,
                               min_points_per_cluster=150):
        """:444-556 -- embedding clusters first, then DBSCAN inside every embedding cluster."""
        labels = self._embedding_labels(embedding_distance_threshold, one_minus=True)
        objects_post_dbscan = []
        for u in set(labels):
            objs_to_consider = [obj for i, obj in enumerate(self.memory) if labels[i] == u]
            objects_post_dbscan = objects_post_dbscan + self._dbscan_merge(objs_to_consider, eps, min_points_per_cluster)
        self.memory = objects_post_dbscan
        self._renumber()

    def _recluster_IoU(self, IoU_threshold=0.6, iou_func=None):
        """:710-747 -- average-linkage clustering on 1 - IoU of every pair of objects.  The reference's pair measure is
        ``calculate_obj_aligned_3d_IoU`` (Qhull oriented boxes + Objectron), which is not built here and raises
        (utils/IoU_ops.py); ``iou_func(pointcloud_i, pointcloud_j)`` selects another measure, e.g. the axis-aligned
        ``utils.IoU_ops.calculate_3d_IoU`` -- an explicit choice by the caller, never a silent substitute."""
        from sklearn.cluster import AgglomerativeClustering
        from ..utils.IoU_ops import calculate_obj_aligned_3d_IoU
        f = iou_func or calculate_obj_aligned_3d_IoU
        m = len(self.memory)
        IoUs = np.zeros((m, m))
        for i in range(m):
            for j in range(i, m):
                if i == j:
                    IoUs[i][j] = 1
                    continue
                IoUs[i][j] = 1 - f(self.memory[i].pointcloud, self.memory[j].pointcloud)
                IoUs[j][i] = IoUs[i][j]
        agg = AgglomerativeClustering(n_clusters=None, distance_threshold=1 - IoU_threshold, metric='precomputed',
                                      linkage='average')
        agg.fit(IoUs)
        self._merge_by_labels(agg.labels_)
        for o in self.memory:
            o._compute_means()

    def recluster_via_clustering_and_IoU(self, distance_func=None, embedding_distance_threshold=0.4, eps=0.4,
                                         min_points_per_cluster=150, IoU_threshold=0.25, iou_func=None):
        """:562-708 -- ``_recluster_IoU`` first, then exactly ``recluster_via_combined`` (the IoU clustering after the
        DBSCAN step is commented out in the reference)."""
        self._recluster_IoU(IoU_threshold, iou_func=iou_func)
        self.recluster_via_combined(distance_func, embedding_distance_threshold, eps, min_points_per_cluster)

    def save(self, save_directory):
        """:750-777 -- memory.txt (the repr), combined_pointcloud.ply, combined_pointcloud_with_floor.ply and one directory
        per object / for the floors (ObjectInfo.save: pointcloud.ply + info.pkl).  Like the reference it needs floors."""
        from ..o3d import io as _io
        from ..utils.depth_utils import combine_point_clouds
        os.makedirs(save_directory, exist_ok=True)
        individual_obj_save_dir = os.path.join(save_directory, "objects")
        os.makedirs(individual_obj_save_dir, exist_ok=True)
        individual_floor_save_dir = os.path.join(save_directory, "floors")
        os.makedirs(individual_floor_save_dir, exist_ok=True)
        with open(os.path.join(save_directory, "memory.txt"), "w") as f:
            f.write(self.__repr__())
        _io.write_point_cloud(os.path.join(save_directory, "combined_pointcloud.ply"),
                              combine_point_clouds(obj.pointcloud for obj in self.memory))
        _io.write_point_cloud(os.path.join(save_directory, "combined_pointcloud_with_floor.ply"),
                              combine_point_clouds([obj.pointcloud for obj in self.memory] + [self.floors.pointcloud]))
        for obj in self.memory:
            obj.save(os.path.join(individual_obj_save_dir, f"{obj.id}"))
        self.floors.save(os.path.join(individual_floor_save_dir, "all_floors"))
        self._log(f"Saved memory to {save_directory}")

    def save_to_pkl(self, save_directory):
        """:779-826 -- [(blank ObjectInfo, points f64[N,3], colors f64[N,3]), ...] plus a floors tuple.  As in the
        reference every record is a FRESH blank ObjectInfo (id = the builtin ``id``, :790; names / embeddings /
        means copied, no cloud), and the floors tuple is a separate record built from the LAST object (the reference's
        loop variable leaks, :804-824) -- not the floors and not the same Python object as the last memory record."""
        def blank(o):
            b = ObjectInfo(0, "", o.embeddings[0], None, 1e5)
            b.id = id
            b.names = o.names
            b.embeddings = o.embeddings
            b.max_embeddings_num = o.max_embeddings_num
            b.mean_emb = o.mean_emb
            b.centroid = o.centroid
            b.pointcloud = None
            return (b, np.asarray(o.pointcloud.points), np.asarray(o.pointcloud.colors))
        mem = [blank(o) for o in self.memory]
        floors = blank(self.memory[-1]) if self.memory else None
        with open(save_directory, "wb") as f, _reference_class_path():
            pickle.dump((mem, floors), f)

    @staticmethod
    def read_pkl(load_directory):
        """The raw content of a ``save_to_pkl`` file: ([(ObjectInfo-without-cloud, points, colors), ...], floors tuple).
        Files written by the REFERENCE name its own class path (``object_memory.object_info.ObjectInfo``); that path
        is mapped to this package's ObjectInfo, so a memory bank built with the reference loads without its tree."""
        with open(load_directory, "rb") as f:
            return _MemoryUnpickler(f).load()

    def load(self, load_directory):
        """:831-846."""
        pklable_memory, pklable_floors = self.read_pkl(load_directory)

        def conv(info_tuple):
            if info_tuple is None:
                return None
            blank_info, pcd_points, pcd_colors = info_tuple
            pcd = PointCloud(pcd_points)
            pcd.colors = pcd_colors
            blank_info.pointcloud = pcd
            return blank_info
        self.memory = [conv(it) for it in pklable_memory]
        self.floors = conv(pklable_floors)

    # ------------------------------------------------------------------------------------------ detection hand-off
    def _get_object_info(self, rgb_image_path, depth_image_path, consider_floor, outlier_removal_config, depth_factor=1.):
        """:125-161."""
        if self.object_finder is None:
            raise RuntimeError("no object_finder injected: the RAM/GroundingDINO/SAM detector is out of scope here; "
                               "use localise_arrays(rgb, depth, masks, embeddings) or pass object_finder=")
        imgs, boxes, masks, phrases = self.object_finder.find(rgb_image_path, consider_floor)
        if imgs is None:
            return None, None, None
        embs = np.stack([np.array(torch.as_tensor(self.get_embeddings_func(
            current_obj_grounded_img=imgs[i], current_obj_bounding_box=boxes[i], current_obj_mask=masks[i],
            current_obj_phrase=phrases[i], full_rgb_image=self.load_rgb_image_func(rgb_image_path),
            full_depth_image=self.load_depth_image_func(depth_image_path), consider_floor=consider_floor,
            device=self.device)).cpu()) for i in range(len(imgs))])
        clouds = get_mask_coloured_pointclouds_from_depth(
            depth_image=self.load_depth_image_func(depth_image_path) / depth_factor,
            rgb_image=self.load_rgb_image_func(rgb_image_path), masks=masks,
            focal_length_x=self.camera_focal_lenth_x, focal_length_y=self.camera_focal_lenth_y,
            outlier_removal_config=outlier_removal_config)
        assert len(imgs) == len(boxes) == len(masks) == len(phrases) == len(embs)
        return phrases, embs, clouds

    # ------------------------------------------------------------------------------------------ caches
    # Memory-side state is identical for every frame, so it is built once (under a lock: localise_batch runs frames
    # on several host threads / CUDA streams) and only read afterwards.  An entry is published after the building
    # stream has been synchronised, so readers on other streams never see half-built device data.
    def _memory_cache(self):
        # identity AND version of every record: ObjectInfo bumps ``_version`` on every mutation (cloud replaced or merged,
        # embedding appended or swapped in place at capacity), so neither a recycled id() nor an unchanged length can
        # leave stale device state behind
        key = (len(self.memory),) + tuple((id(m), m._version, id(m.pointcloud._p)) for m in self.memory)
        with self._lock:
            if key != self._cache_key:
                self._cache_key = key
                self._cache = {"assn": {}}
            return self._cache

    def cache_info(self):
        """-> entries / bytes of the per-assignment memory-side cache and its lookup statistics."""
        with self._lock:
            c = self._memory_cache()["assn"]
            return dict(self.cache_stats, entries=len(c), bytes=int(sum(e[3] for e in c.values())),
                        budget_bytes=int(self.assignment_cache_bytes))

    def clear_assignment_cache(self):
        """Drop the per-assignment-tuple memory-side state (centred union cloud, normals, FPFH, NN fields, colour
        gradients); per-object clouds, the embedding bank and the all-memory search grid stay."""
        with self._lock:
            torch.cuda.synchronize()
            self._memory_cache()["assn"].clear()

    def _embedding_bank(self):
        with self._lock:
            c = self._memory_cache()
            if "bank" not in c:
                rows = [np.asarray(e, np.float32).reshape(1, -1) for m in self.memory for e in m.embeddings]
                seg = np.cumsum([0] + [len(m.embeddings) for m in self.memory]).astype(np.int32)
                bank = (K.to_dev(np.concatenate(rows), torch.float32), K.to_dev(seg, torch.int32))
                for m in self.memory:
                    m._compute_means()                      # :913-914 (host-side means; nothing reads them here)
                torch.cuda.current_stream().synchronize()
                c["bank"] = bank
            return c["bank"]

    def _all_memory_pcd(self, threshold=0.02):
        with self._lock:
            c = self._memory_cache()
            if "all" not in c:
                allp = PointCloud()
                for obj in self.memory:
                    allp = allp + obj.pointcloud
                _reg._nn_grid(allp, threshold)              # evaluate_transform's search structure (:1104)
                torch.cuda.current_stream().synchronize()
                c["all"] = allp
            return c["all"]

    # Per-assignment memory-side state is kept under a BYTE budget with least-recently-used eviction: an entry is ~1 KB per
    # memory point (union cloud with normals / colours / gradients, FPFH rows, the matcher's packed operands) plus its two
    # exact-NN candidate fields -- 50-150 MB for a 50 k-point union.  A frame stream of one environment keeps meeting the same
    # tuples of co-visible objects; how many of them stay resident is a deployment choice (HBM is 180 GB).
    assignment_cache_bytes = 96 << 30

    @staticmethod
    def _assn_entry_bytes(pcd, feats):
        n = pcd.n()
        b = n * 24 * 2                                         # union points + colours
        if feats is not None:
            tgt, f = feats
            b += n * (24 * 3 + 264 + 448 + 2 * 148 + 4 + 16 + 40)      # normals, gradients, a spare row; FPFH; packs; fp32 rows + norms in both orders, permutation; grid rows
            for g in tgt._grids.values():
                if isinstance(g, K.Grid):
                    b += g.nnfield_bytes
        return int(b)

    def _chosen_memory(self, m_indices, fpfh_voxel_size, global_dist_factor, local_dist_factor):
        """centred union of the chosen memory clouds with everything register_point_clouds derives from it:
        normals + FPFH, the RANSAC / ICP search structures and the coloured-ICP gradients."""
        key = (tuple(int(i) for i in m_indices), float(fpfh_voxel_size), float(global_dist_factor),
               float(local_dist_factor))
        with self._lock:
            c = self._memory_cache()["assn"]
            self.cache_stats["lookups"] += 1
            hit = c.get(key)
            if hit is not None:
                c[key] = c.pop(key)                            # most recently used last (dicts keep insertion order)
                return hit[:3]
            self.cache_stats["misses"] += 1
            pcd = PointCloud()
            for m_index in m_indices:
                pcd = pcd + self.memory[m_index].pointcloud
            mean = pcd.get_center()
            pcd.translate(-mean)
            try:
                feats = downsample_and_compute_fpfh(pcd, fpfh_voxel_size)
                feats[1].pack()                            # matcher operands of the memory side, packed once
                tgt = feats[0]
                g_r = _reg._nn_grid(tgt, fpfh_voxel_size * global_dist_factor)
                g_i = _reg._nn_grid(tgt, fpfh_voxel_size * local_dist_factor)
                if tgt.has_colors():
                    _reg.ensure_color_gradient(tgt, fpfh_voxel_size * local_dist_factor)
                # what later frames need from the search structures are the two NN-field grids only: the k-NN grid (up to
                # 2^24 cells) and the shared neighbour lists (~1.2 KB per point) served the features and go
                tgt._grids = {k: g for k, g in tgt._grids.items() if g is g_r or g is g_i}
                pcd._p, pcd._c = tgt._p, tgt._c            # the copy the features were computed on holds the same values
            except K.IblError:
                raise                                      # a device / ABI error is not Open3D's "no normals" case
            except Exception:
                feats = None                               # register_point_clouds then takes its own fallback branch
            torch.cuda.current_stream().synchronize()
            nbytes = self._assn_entry_bytes(pcd, feats)
            used = sum(e[3] for e in c.values())
            while c and used + nbytes > self.assignment_cache_bytes:
                used -= c.pop(next(iter(c)))[3]            # least recently used first
                self.cache_stats["evictions"] = self.cache_stats.get("evictions", 0) + 1
            c[key] = (pcd, mean, feats, nbytes)
            # The entry's tensors were carved out of the block this stream had reserved for its transients.  Reserve again
            # NOW (the cold path: a no-op when enough of the pool is free, otherwise the driver maps memory here): a warm frame
            # that finds the pool full makes the stream-ordered allocator grow in the middle of a step, which was measured to
            # stall all four frame streams for ~650 ms (pool 80 192 -> 80 224 MiB, step 150 -> 855 ms).
            if self.scratch_bytes_per_stream and not os.environ.get("IBL_NO_RESERVE"):
                K.reserve_scratch(self.scratch_bytes_per_stream)
            return c[key][:3]

    # ------------------------------------------------------------------------------------------ localise
    def closest_similarities(self, detected_embs):
        """:913-936: S[i][j] = max over stored embeddings e of object j of <e/|e|, d_i/|d_i|> (float32)."""
        mem, seg = self._embedding_bank()
        det = detected_embs if isinstance(detected_embs, torch.Tensor) else torch.from_numpy(
            np.ascontiguousarray(np.asarray(detected_embs, np.float32)))
        return K.cosine_segmax(det.to(device=mem.device, dtype=torch.float32).contiguous(), mem, seg)

    def localise(self, image_path, depth_image_path, testname="", subtest_name="", save_point_clouds=False,
                 outlier_removal_config=None, fpfh_global_dist_factor=2, fpfh_local_dist_factor=0.4,
                 fpfh_voxel_size=0.05, topK=5, useLora=True, save_localised_pcd_path=None, consider_floor=False,
                 perform_semantic_icp=True, depth_factor=1., max_detected_object_num=7):
        """:852-1169."""
        if outlier_removal_config is None:
            outlier_removal_config = {"radius_nb_points": 8, "radius": 0.05}
        consider_floor = False
        detected_phrases, detected_embs, detected_pointclouds = self._get_object_info(
            image_path, depth_image_path, consider_floor=consider_floor, outlier_removal_config=outlier_removal_config,
            depth_factor=depth_factor)
        return self._localise_core(detected_phrases, detected_embs, detected_pointclouds, outlier_removal_config,
                                   fpfh_global_dist_factor, fpfh_local_dist_factor, fpfh_voxel_size,
                                   perform_semantic_icp, max_detected_object_num)

    def localise_arrays(self, rgb_image, depth_image, masks, embeddings, phrases=None, outlier_removal_config=None,
                        fpfh_global_dist_factor=2, fpfh_local_dist_factor=0.4, fpfh_voxel_size=0.05,
                        perform_semantic_icp=True, depth_factor=1., max_detected_object_num=7):
        """localise() from arrays: rgb uint8 [H,W,3], depth [H,W] (float32 metres or raw / depth_factor), masks
        bool [n,1,H,W] or [n,H,W], embeddings [n,D] -- numpy or CUDA tensors.  Same computation as localise after the
        detector / encoder hand-off (object_memory.py:147-154 onwards)."""
        if outlier_removal_config is None:
            outlier_removal_config = {"radius_nb_points": 8, "radius": 0.05}
        if masks is None or len(masks) == 0:
            return self._localise_core(None, None, None, outlier_removal_config, fpfh_global_dist_factor,
                                       fpfh_local_dist_factor, fpfh_voxel_size, perform_semantic_icp,
                                       max_detected_object_num)
        depth = depth_image
        if depth_factor != 1.:
            depth = depth_image / depth_factor
        clouds = get_mask_coloured_pointclouds_from_depth(depth, rgb_image, masks, self.camera_focal_lenth_x,
                                                          self.camera_focal_lenth_y, outlier_removal_config)
        embs = embeddings.detach() if isinstance(embeddings, torch.Tensor) else np.array(embeddings)
        phrases = phrases if phrases is not None else [f"object_{i}" for i in range(len(clouds))]
        return self._localise_core(phrases, embs, clouds, outlier_removal_config, fpfh_global_dist_factor,
                                   fpfh_local_dist_factor, fpfh_voxel_size, perform_semantic_icp,
                                   max_detected_object_num)

    # per frame stream: matcher candidate pool (~150 MB per call) + grids + k-NN lists, AND whatever memory-side caches
    # (per-assignment features, grids, NN fields: ~5 GB for the 40-object bank) happen to be built on that stream
    scratch_bytes_per_stream = 4 << 30

    def localise_batch(self, frames, workers=4, batched_registration=True, pipeline_chunks=2, **kw):
        """localise_arrays over a batch of frames (dicts with rgb / depth / masks / embs).  Frames are independent
        (SURVEY.md 8(e)); the work is organised in two stages:

          1. per frame, ``workers`` frames in flight on their own host threads / CUDA streams: back-projection, outlier
             filters, similarities, SimVolume assignments, and per assignment the centred detected union with its normals,
             FPFH features and feature correspondences (``_plan`` + ``_prepare_registrations``);
          2. on the caller's stream, ALL (frame, assignment) registration problems of the batch together: one batched
             RANSAC (``ibl_ransac_run_batch``), one batched coloured ICP (``ibl_icp_colored_batch``) and one batched
             evaluate_transform (``ibl_nn_score_batch``) -- a handful of launches and host round trips per stage for the
             whole batch instead of hundreds per frame.

        With several workers the two stages overlap: the batch is cut into ``pipeline_chunks`` groups and a group is
        registered (stage 2) while the workers are still in stage 1 of the next one.
        ``batched_registration=False`` runs stage 2 frame by frame inside the workers (the round-1 organisation).
        Results are identical to calling localise_arrays frame by frame either way.
        -> list of (pose, [assn, None], info)."""
        frames = list(frames)
        dev_index = torch.cuda.current_device()
        batched = bool(batched_registration) and _reg.ransac_sampler is None

        def one(fr):
            if not batched:
                pose, assn = self.localise_arrays(fr["rgb"], fr["depth"], fr["masks"], fr["embs"], **kw)
                return pose, assn, self._tls.info
            t0 = time.perf_counter()
            plan = self._plan_arrays(fr["rgb"], fr["depth"], fr["masks"], fr["embs"], **kw)
            t1 = time.perf_counter()
            if plan.get("registrations") is not None:
                self._prepare_registrations(plan)
            plan["info"]["stage1_ms"] = ((t1 - t0) * 1e3, (time.perf_counter() - t1) * 1e3)   # (plan, features + matching)
            return plan

        if workers <= 1 or len(frames) <= 1:
            out = [one(fr) for fr in frames]
        else:
            ready = torch.cuda.Event()
            ready.record()                                   # inputs produced on the caller's stream (e.g. H2D copies)

            def run(fr):
                torch.cuda.set_device(dev_index)
                st = getattr(self._tls, "stream", None)
                if st is None:
                    st = self._tls.stream = torch.cuda.Stream()
                    if self.scratch_bytes_per_stream and not os.environ.get("IBL_NO_RESERVE"):
                        with torch.cuda.stream(st):      # this frame stream's working set, mapped once (see ibl.h)
                            K.reserve_scratch(self.scratch_bytes_per_stream)
                st.wait_event(ready)
                with torch.cuda.stream(st):
                    r = one(fr)
                    st.synchronize()
                return r

            if self._pool is None or self._pool_workers != workers:
                if self._pool is not None:
                    self._pool.shutdown(wait=True)
                self._pool, self._pool_workers = ThreadPoolExecutor(max_workers=workers), workers
            # longest first: a frame's cost grows with its detected clouds, i.e. with its mask pixels; handing the heavy
            # frames out first keeps the last worker from finishing long after the others (results keep the input order).
            # Mask sizes come from the arrays' shapes and dtypes only -- no device read-back here.
            order = list(range(len(frames)))
            if not os.environ.get("IBL_NO_LPT"):
                order.sort(key=lambda i: -int(np.prod(np.shape(frames[i]["masks"]))))
            out = [None] * len(frames)
            futs = [self._pool.submit(run, frames[i]) for i in order]
            if not batched:
                for i, f in zip(order, futs):
                    out[i] = f.result()
                return out
            # Stage 2 of the frames that are ready runs on this thread WHILE the workers are still in stage 1 of the rest:
            # the batch is cut into ``pipeline_chunks`` groups in submission order, each registered as one batch.
            nchunks = max(1, min(int(pipeline_chunks), len(frames) // 4))
            cuts = [round(k * len(frames) / nchunks) for k in range(nchunks + 1)]
            timing = {"stage1_wait_s": 0.0, "stage2_s": 0.0}
            for a, b in zip(cuts[:-1], cuts[1:]):
                t0 = time.perf_counter()
                plans = [f.result() for f in futs[a:b]]
                t1 = time.perf_counter()
                for i, p in zip(order[a:b], plans):
                    out[i] = p
                self._register_batch([p for p in plans if p.get("registrations") is not None])
                timing["stage1_wait_s"] += t1 - t0
                timing["stage2_s"] += time.perf_counter() - t1
            self.last_batch_timing = timing
            return [self._finish(plan) + (plan["info"],) for plan in out]
        if not batched:
            return out
        t0 = time.perf_counter()
        self._register_batch([p for p in out if p.get("registrations") is not None])
        self.last_batch_timing = {"stage1_wait_s": 0.0, "stage2_s": time.perf_counter() - t0}
        return [self._finish(plan) + (plan["info"],) for plan in out]

    # ------------------------------------------------------------------------------------------ plan / register / finish
    def _plan_arrays(self, rgb_image, depth_image, masks, embeddings, phrases=None, outlier_removal_config=None,
                     fpfh_global_dist_factor=2, fpfh_local_dist_factor=0.4, fpfh_voxel_size=0.05,
                     perform_semantic_icp=True, depth_factor=1., max_detected_object_num=7):
        """localise_arrays up to (not including) the registrations."""
        if outlier_removal_config is None:
            outlier_removal_config = {"radius_nb_points": 8, "radius": 0.05}
        if masks is None or len(masks) == 0:
            return self._plan(None, None, None, outlier_removal_config, fpfh_global_dist_factor, fpfh_local_dist_factor,
                              fpfh_voxel_size, perform_semantic_icp, max_detected_object_num)
        depth = depth_image
        if depth_factor != 1.:
            depth = depth_image / depth_factor
        clouds = get_mask_coloured_pointclouds_from_depth(depth, rgb_image, masks, self.camera_focal_lenth_x,
                                                          self.camera_focal_lenth_y, outlier_removal_config)
        embs = embeddings.detach() if isinstance(embeddings, torch.Tensor) else np.array(embeddings)
        phrases = phrases if phrases is not None else [f"object_{i}" for i in range(len(clouds))]
        return self._plan(phrases, embs, clouds, outlier_removal_config, fpfh_global_dist_factor, fpfh_local_dist_factor,
                          fpfh_voxel_size, perform_semantic_icp, max_detected_object_num)

    def _plan(self, detected_phrases, detected_embs, detected_pointclouds, outlier_removal_config,
              fpfh_global_dist_factor, fpfh_local_dist_factor, fpfh_voxel_size, perform_semantic_icp,
              max_detected_object_num):
        """:895-1034 -- everything of localise() before register_point_clouds: top-7 selection, closest similarities,
        SimVolume assignments, second outlier filter, and per assignment the centred detected / memory unions.
        -> plan dict; plan["registrations"] is None when nothing was detected (sentinel pose, :895-896)."""
        info = {"assignments": [], "registrations": []}
        plan = {"info": info, "registrations": None, "voxel": fpfh_voxel_size, "gfac": fpfh_global_dist_factor,
                "lfac": fpfh_local_dist_factor}
        if detected_embs is None:
            return plan

        # take the largest clouds (:900-908; python's sort is stable, reverse keeps ties in input order)
        if len(detected_pointclouds) > max_detected_object_num:
            order = sorted(range(len(detected_pointclouds)), key=lambda i: detected_pointclouds[i].n(), reverse=True)
            order = order[:max_detected_object_num]
            detected_phrases = [detected_phrases[i] for i in order]
            detected_embs = detected_embs[order]
            detected_pointclouds = [detected_pointclouds[i] for i in order]

        if len(detected_embs) > len(self.memory):                              # :918-920
            detected_embs = detected_embs[:len(self.memory)]

        closest = self.closest_similarities(detected_embs)                     # :922-936 (device, float32)
        info["closest_similarities"] = closest

        sv = SimVolume(closest)                                                # :974-984
        subvolume_size = min(len(detected_pointclouds), 3)
        sv.fast_construct_volume(subvolume_size)
        assn_data = list(sv.get_top_indices_from_subvolumes(num_per_length=4))
        info["assignments"] = copy.deepcopy(assn_data)

        # second outlier removal on every detected cloud (:992-998)
        detected_pointclouds = remove_radius_outlier_many(detected_pointclouds, outlier_removal_config["radius_nb_points"],
                                                          outlier_removal_config["radius"])
        all_detected_pcd = PointCloud()
        for pcd in detected_pointclouds:                                       # :1010-1016
            all_detected_pcd = all_detected_pcd + pcd

        if perform_semantic_icp and len(assn_data):
            raise NotImplementedError                                          # :1036-1037, as in the reference

        regs = []
        unions = []
        for assn in assn_data:                                                 # :1020
            chosen_detected_pcd = PointCloud()
            for d_index, _ in assn:
                chosen_detected_pcd = chosen_detected_pcd + detected_pointclouds[d_index]
            unions.append(chosen_detected_pcd)
        # the means of all unions of the frame: one launch, one read-back (each equals get_center() of that union)
        means = K.mean3_many([u._p for u in unions]) if unions else []
        for assn, chosen_detected_pcd, detected_mean in zip(assn_data, unions, means):
            chosen_memory_pcd, memory_mean, target_feats = self._chosen_memory(
                [m for _, m in assn], fpfh_voxel_size, fpfh_global_dist_factor, fpfh_local_dist_factor)
            chosen_detected_pcd.translate(-detected_mean)                      # :1030-1034
            regs.append({"assn": assn, "source": chosen_detected_pcd, "target": chosen_memory_pcd,
                         "target_feats": target_feats, "detected_mean": detected_mean, "memory_mean": memory_mean})
        plan["registrations"] = regs
        plan["all_detected"] = all_detected_pcd
        return plan

    def _register_sequential(self, plan):
        """:1036-1108 one assignment after the other: register_point_clouds, then evaluate_transform of the composed
        global-frame transform over all detected points against the whole memory."""
        all_memory_pcd = self._all_memory_pcd()                                # :1001-1007
        for g in plan["registrations"]:
            rinfo = _reg.last_info()
            rinfo.clear()
            transform, rmse, fitness = register_point_clouds(
                g["source"], g["target"], voxel_size=plan["voxel"], global_dist_factor=plan["gfac"],
                local_dist_factor=plan["lfac"], target_features=g["target_feats"])
            G = self._global_transform(transform, g)
            full_rmse, full_fitness = evaluate_transform(plan["all_detected"], all_memory_pcd, trans_init=G)   # :1104
            g.update(transform=transform, rmse=rmse, fitness=fitness, full_rmse=full_rmse, full_fitness=full_fitness,
                     ransac=dict(rinfo.get("ransac", {})), icp=dict(rinfo.get("icp", {})))

    @staticmethod
    def _global_transform(transform, g):
        """:1096-1101 -- the centred registration expressed in the memory / camera frames."""
        G = np.array(transform, dtype=np.float64)
        G[:3, 3] = transform[:3, 3] + g["memory_mean"] - transform[:3, :3] @ g["detected_mean"]
        return G

    def _prepare_registrations(self, plan):
        """The per-problem front half of register_point_clouds (fpfh_register.py:100-112) for every assignment of a
        frame: source normals + FPFH and the mutual feature correspondences against the cached target features.  A failure
        Open3D would raise (no normals, empty cloud) marks the problem for the reference's fallback branch."""
        for g in plan["registrations"]:
            g["fallback"] = None
            try:
                src_down, src_fpfh = downsample_and_compute_fpfh(g["source"], plan["voxel"])
                if g["target_feats"] is not None:
                    tgt_down, tgt_fpfh = g["target_feats"]
                else:
                    tgt_down, tgt_fpfh = downsample_and_compute_fpfh(g["target"], plan["voxel"])
                g["src_down"], g["tgt_down"] = src_down, tgt_down
                g["corres"] = _reg.match_features(src_fpfh, tgt_fpfh, True, 3)
            except K.IblError:
                raise
            except Exception as e:                             # the reference's bare except (fpfh_register.py:137)
                g["fallback"] = repr(e)

    def _register_batch(self, plans):
        """Stage 2 of localise_batch: RANSAC, coloured ICP and evaluate_transform of every (frame, assignment) problem of
        the batch, each as ONE batched library call (fpfh_register.py:106-143,145-150 per problem)."""
        probs = [(plan, g) for plan in plans for g in plan["registrations"]]
        if not probs:
            return
        if not getattr(self._tls, "stage2_reserved", False) and self.scratch_bytes_per_stream and \
                not os.environ.get("IBL_NO_RESERVE"):
            # the batched stage's temporaries (entry lists of up to 2^20 hypotheses per problem, moved clouds, partials)
            # come from the stream-ordered pool: map this stream's share once instead of growing it step by step
            K.reserve_scratch(self.scratch_bytes_per_stream)
            self._tls.stage2_reserved = True
        all_memory_pcd = self._all_memory_pcd()
        by_cfg = {}
        for plan, g in probs:
            by_cfg.setdefault((plan["voxel"], plan["gfac"], plan["lfac"]), []).append((plan, g))
        for (voxel, gfac, lfac), group in by_cfg.items():
            max_d_r, max_d_i = voxel * gfac, voxel * lfac
            live = [(plan, g) for plan, g in group if g["fallback"] is None]
            # ---- RANSAC (fpfh_register.py:106-119)
            specs = []
            for plan, g in live:
                sd, td = g["src_down"], g["tgt_down"]
                colored = td.has_normals() and sd.has_colors() and td.has_colors()
                g["colored"] = colored
                grad = _reg.ensure_color_gradient(td, max_d_i)[1] if colored else None
                specs.append({"src": sd._p, "src_col": sd._c if colored else None, "tgt": td._p,
                              "tgt_nrm": td._n if colored else None, "tgt_col": td._c if colored else None, "tgt_grad": grad,
                              "grid_ransac": _reg._nn_grid(td, max_d_r), "grid_icp": _reg._nn_grid(td, max_d_i),
                              "corres": g["corres"], "seed": _reg.ransac_seed})
            if live:
                arr = K.reg_problems(specs)
                max_it = 4000000 if _reg.ransac_max_iteration is None else int(_reg.ransac_max_iteration)   # :119
                rr = K.ransac_run_batch(arr, max_it, 0.99, max_d_r, edge_thr=0.9, use_checkers=True,
                                        early_exit=bool(_reg.ransac_early_exit), block=_reg.ransac_batch_block)
                for (plan, g), r in zip(live, rr):
                    g["ransac"] = {k: r[k] for k in ("drawn", "validated", "best_hyp", "n_corres", "n_inliers", "err2")}
                    g["ransac_T"] = r["T"]
                # ---- coloured ICP from the RANSAC pose (:132-135); problems without colours / normals raise in Open3D
                ci = [k for k, (plan, g) in enumerate(live) if g["colored"]]
                if ci:
                    sub = K.reg_problems([specs[k] for k in ci])
                    ri = K.icp_colored_batch(sub, np.stack([live[k][1]["ransac_T"] for k in ci]), max_d_i)
                    for k, r in zip(ci, ri):
                        g = live[k][1]
                        g.update(transform=r["T"], rmse=r["inlier_rmse"], fitness=r["fitness"],
                                 icp={"mode": "colored", "iterations": r["iterations"], "n_inliers": r["n_inliers"]})
                for plan, g in live:
                    if not g["colored"]:
                        g["fallback"] = "ColoredICP requires colored pointclouds with target normals."
            # ---- the reference's except branch (:137-141): identity-initialised point-to-point ICP on the raw clouds
            for plan, g in group:
                if g["fallback"] is not None:
                    r = _reg.registration_icp(g["source"], g["target"], max_d_i, np.eye(4),
                                              _reg.TransformationEstimationPointToPoint())
                    g.update(transform=r.transformation, rmse=r.inlier_rmse, fitness=r.fitness,
                             icp=dict(_reg.last_info().get("icp", {})))
                    g.setdefault("ransac", {})
        # ---- evaluate_transform of every candidate over all detected points against the whole memory (:1104)
        Gs = [self._global_transform(g["transform"], g) for plan, g in probs]
        grid = _reg._nn_grid(all_memory_pcd, 0.02)
        scores = K.nn_score_batch(grid, [plan["all_detected"]._p if plan["all_detected"].n() else
                                         torch.zeros((0, 3), dtype=torch.float64, device=all_memory_pcd._p.device)
                                         for plan, g in probs], Gs, 0.02)
        for (plan, g), (n_in, err2) in zip(probs, scores):
            n = plan["all_detected"].n()
            g["full_fitness"] = (n_in / n) if n else 0.0
            g["full_rmse"] = float(np.sqrt(err2 / n_in)) if n_in else 0.0

    def _finish(self, plan):
        """:1111-1169 -- best assignment by full-memory fitness, pose assembly."""
        info = plan["info"]
        self.last_localise_info = info
        self._tls.info = info
        if plan["registrations"] is None:
            return np.array([0., 0., 0., 0., 0., 0., 1.]), [[], []]           # :895-896
        assn_data = []
        detected_mean = memory_mean = None
        assn = None
        for g in plan["registrations"]:
            assn = g["assn"]
            detected_mean, memory_mean = g["detected_mean"], g["memory_mean"]
            assn_data.append([assn, g["transform"], g["rmse"], g["fitness"], g["full_rmse"], g["full_fitness"],
                              detected_mean, memory_mean])
            info["registrations"].append({"assn": assn, "transform": g["transform"], "rmse": g["rmse"], "fitness": g["fitness"],
                                          "full_rmse": g["full_rmse"], "full_fitness": g["full_fitness"],
                                          "detected_mean": detected_mean, "memory_mean": memory_mean,
                                          "n_source": g["source"].n(), "n_target": g["target"].n(),
                                          "ransac": dict(g.get("ransac", {})), "icp": dict(g.get("icp", {}))})
        best = sorted(assn_data, key=lambda x: x[5], reverse=True)             # :1111 (stable, by full fitness)
        best_transform, best_assn = best[0][1], best[0][0]
        if not self.bug_compatible:
            detected_mean, memory_mean, assn = best[0][6], best[0][7], best_assn
        # bug_compatible: means of the LAST loop iteration (:1127 vs :1030-1031) and the last assn (:1169)
        R = np.array(best_transform[:3, :3])
        t = np.array(best_transform[:3, 3])
        tAvg = t + memory_mean - R @ detected_mean
        qAvg = Rotation.from_matrix(R).as_quat()
        localised_pose = np.concatenate((tAvg, qAvg))
        info["best_assn"] = best_assn
        info["best_transform"] = best_transform
        # the pose composed with the BEST assignment's own means (what :1127 intends); equals localised_pose whenever
        # the best assignment is also the last one, or with bug_compatible=False
        info["pose_best_means"] = np.concatenate((t + best[0][7] - R @ best[0][6], qAvg))
        return localised_pose, [assn, None]

    def _localise_core(self, detected_phrases, detected_embs, detected_pointclouds, outlier_removal_config,
                       fpfh_global_dist_factor, fpfh_local_dist_factor, fpfh_voxel_size, perform_semantic_icp,
                       max_detected_object_num):
        plan = self._plan(detected_phrases, detected_embs, detected_pointclouds, outlier_removal_config,
                          fpfh_global_dist_factor, fpfh_local_dist_factor, fpfh_voxel_size, perform_semantic_icp,
                          max_detected_object_num)
        if plan["registrations"] is not None:
            self._register_sequential(plan)
        return self._finish(plan)
