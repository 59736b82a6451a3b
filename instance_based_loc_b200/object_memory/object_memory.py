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

import copy
import os
import pickle
import threading
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch
from scipy.spatial.transform import Rotation

from .. import kernels as K
from ..o3d import registration as _reg
from ..o3d.pointcloud import PointCloud
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
        """:779-826 -- [(ObjectInfo-without-cloud, points f64[N,3], colors f64[N,3]), ...], floors tuple.  The
        reference builds the floors tuple from the LAST object (bug, :804-824); reproduced."""
        def strip(o):
            b = copy.copy(o)
            b.pointcloud = None
            return (b, np.asarray(o.pointcloud.points), np.asarray(o.pointcloud.colors))
        mem = [strip(o) for o in self.memory]
        with open(save_directory, "wb") as f:
            pickle.dump((mem, mem[-1] if mem else None), f)

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
        key = (len(self.memory),) + tuple(id(m.pointcloud._p) for m in self.memory) + \
              tuple(len(m.embeddings) for m in self.memory)
        with self._lock:
            if key != self._cache_key:
                self._cache_key = key
                self._cache = {"assn": {}}
            return self._cache

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

    def _chosen_memory(self, m_indices, fpfh_voxel_size, global_dist_factor, local_dist_factor):
        """centred union of the chosen memory clouds with everything register_point_clouds derives from it:
        normals + FPFH, the RANSAC / ICP search structures and the coloured-ICP gradients."""
        key = (tuple(int(i) for i in m_indices), float(fpfh_voxel_size), float(global_dist_factor),
               float(local_dist_factor))
        with self._lock:
            c = self._memory_cache()["assn"]
            if key not in c:
                pcd = PointCloud()
                for m_index in m_indices:
                    pcd = pcd + self.memory[m_index].pointcloud
                mean = pcd.get_center()
                pcd.translate(-mean)
                try:
                    feats = downsample_and_compute_fpfh(pcd, fpfh_voxel_size)
                    tgt = feats[0]
                    _reg._nn_grid(tgt, fpfh_voxel_size * global_dist_factor)
                    _reg._nn_grid(tgt, fpfh_voxel_size * local_dist_factor)
                    if tgt.has_colors():
                        _reg.ensure_color_gradient(tgt, fpfh_voxel_size * local_dist_factor)
                except Exception:
                    feats = None
                torch.cuda.current_stream().synchronize()
                if len(c) > 512:
                    c.pop(next(iter(c)))
                c[key] = (pcd, mean, feats)
            return c[key]

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

    def localise_batch(self, frames, workers=4, **kw):
        """localise_arrays over a batch of frames (dicts with rgb / depth / masks / embs), ``workers`` frames in flight
        at a time, each on its own host thread and CUDA stream: frames are independent, and the
        path is a long chain of small kernels with host decisions in between, so overlapping frames is what keeps
        the SMs busy.  Results are identical to calling localise_arrays frame by frame.
        -> list of (pose, [assn, None], info)."""
        frames = list(frames)
        dev_index = torch.cuda.current_device()

        def one(fr, stream=None):
            pose, assn = self.localise_arrays(fr["rgb"], fr["depth"], fr["masks"], fr["embs"], **kw)
            return pose, assn, self._tls.info

        if workers <= 1 or len(frames) <= 1:
            return [one(fr) for fr in frames]
        ready = torch.cuda.Event()
        ready.record()                                       # inputs produced on the caller's stream (e.g. H2D copies)

        def run(fr):
            torch.cuda.set_device(dev_index)
            st = getattr(self._tls, "stream", None)
            if st is None:
                st = self._tls.stream = torch.cuda.Stream()
                if self.scratch_bytes_per_stream and not os.environ.get("IBL_NO_RESERVE"):
                    with torch.cuda.stream(st):          # this frame stream's working set, mapped once (see ibl.h)
                        K.reserve_scratch(self.scratch_bytes_per_stream)
            st.wait_event(ready)
            with torch.cuda.stream(st):
                out = one(fr)
                st.synchronize()
            return out

        if self._pool is None or self._pool_workers != workers:
            if self._pool is not None:
                self._pool.shutdown(wait=True)
            self._pool, self._pool_workers = ThreadPoolExecutor(max_workers=workers), workers
        # longest first: a frame's cost grows with its detected clouds, i.e. with its mask pixels; handing the heavy
        # frames out first keeps the last worker from finishing long after the others (results keep the input order)
        order = list(range(len(frames)))
        try:
            if not os.environ.get("IBL_NO_LPT"):
                px = torch.stack([torch.as_tensor(fr["masks"]).sum() for fr in frames]).cpu().tolist()
                order.sort(key=lambda i: -px[i])
        except Exception:
            pass
        out = [None] * len(frames)
        for i, r in zip(order, self._pool.map(run, [frames[i] for i in order])):
            out[i] = r
        return out

    def _localise_core(self, detected_phrases, detected_embs, detected_pointclouds, outlier_removal_config,
                       fpfh_global_dist_factor, fpfh_local_dist_factor, fpfh_voxel_size, perform_semantic_icp,
                       max_detected_object_num):
        info = self.last_localise_info = {"assignments": [], "registrations": []}
        self._tls.info = info
        if detected_embs is None:
            return np.array([0., 0., 0., 0., 0., 0., 1.]), [[], []]           # :895-896

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
        detected_pointclouds = [obj.remove_radius_outlier(nb_points=outlier_removal_config["radius_nb_points"],
                                                          radius=outlier_removal_config["radius"])[0]
                                for obj in detected_pointclouds]
        all_memory_pcd = self._all_memory_pcd()                                # :1001-1007
        all_detected_pcd = PointCloud()
        for pcd in detected_pointclouds:                                       # :1010-1016
            all_detected_pcd = all_detected_pcd + pcd

        if perform_semantic_icp and len(assn_data):
            raise NotImplementedError                                          # :1036-1037, as in the reference

        detected_mean = memory_mean = None
        assn = None
        for assn_num, assn in enumerate(assn_data):                            # :1020
            chosen_detected_pcd = PointCloud()
            for d_index, _ in assn:
                chosen_detected_pcd = chosen_detected_pcd + detected_pointclouds[d_index]
            chosen_memory_pcd, memory_mean, target_feats = self._chosen_memory(
                [m for _, m in assn], fpfh_voxel_size, fpfh_global_dist_factor, fpfh_local_dist_factor)
            detected_mean = chosen_detected_pcd.get_center()
            chosen_detected_pcd.translate(-detected_mean)                      # :1030-1034
            transform, rmse, fitness = register_point_clouds(
                chosen_detected_pcd, chosen_memory_pcd, voxel_size=fpfh_voxel_size,
                global_dist_factor=fpfh_global_dist_factor, local_dist_factor=fpfh_local_dist_factor,
                target_features=target_feats)
            global_frame_transform = np.array(transform, dtype=np.float64)     # :1096-1101
            R = transform[:3, :3]
            global_frame_transform[:3, 3] = transform[:3, 3] + memory_mean - R @ detected_mean
            full_rmse, full_fitness = evaluate_transform(all_detected_pcd, all_memory_pcd,
                                                         trans_init=global_frame_transform)      # :1104
            assn_data[assn_num] = [assn, transform, rmse, fitness, full_rmse, full_fitness, detected_mean, memory_mean]
            info["registrations"].append({"assn": assn, "transform": transform, "rmse": rmse, "fitness": fitness,
                                          "full_rmse": full_rmse, "full_fitness": full_fitness})

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
        return localised_pose, [assn, None]
