#!/usr/bin/env python
"""Emit Open3D-produced golden vectors for the rows the reference delegates to Open3D.

THIS SCRIPT CANNOT RUN IN THE BUILD CONTAINER (open3d is not installable there, no network).  It is the one-file
recipe that turns "parity unpinned" into a pinned oracle on any machine that has the reference's pinned dependency:

    pip install open3d==0.17.0 numpy            # environment.yml:229
    python tests/golden/make_open3d_golden.py --reference-root /path/to/instance-based-loc \
           --out tests/golden/open3d_golden.npz

It needs nothing from this repository: inputs are the object clouds bundled with the reference
(out/360_trial_with_floor/objects/{0,1,2}/pointcloud.ply: double xyz + uchar rgb), cut down with fixed strides, and
seeded rigid transforms.  The .npz holds the INPUT arrays as well, so tests/test_open3d_golden_cpu.py needs neither
Open3D nor the reference tree; while the file is absent that test reports "parity unpinned".

Reference call sites covered (the arguments are the reference's own):
  utils/depth_utils.py:41,87, object_memory/object_memory.py:994   remove_radius_outlier(nb_points, radius)
  utils/fpfh_register.py:91-92      estimate_normals(KDTreeSearchParamHybrid(2 vs, 30))
  utils/fpfh_register.py:95-97      compute_fpfh_feature(KDTreeSearchParamHybrid(5 vs, 100))
  utils/fpfh_register.py:110-119    registration_ransac_based_on_feature_matching -> its deterministic pieces:
                                    correspondences_from_features (mutual filter), the 3-point
                                    TransformationEstimationPointToPoint.compute_transformation, evaluate_registration
                                    of the hypothesis; plus one seeded whole-RANSAC run as a sanity band
  utils/fpfh_register.py:132-135    registration_colored_icp(TransformationEstimationForColoredICP())
  utils/fpfh_register.py:138-141    registration_icp(TransformationEstimationPointToPoint())
  utils/fpfh_register.py:146-148    evaluate_registration(threshold 0.02)
  object_memory/object_memory.py:315,511,632   cluster_dbscan(eps, min_points)
  utils/IoU_ops.py:107-112          OrientedBoundingBox.create_from_points
"""
import argparse
import os

import numpy as np

VS = 0.05                      # fpfh_voxel_size (the drivers' CLI default)


def rigid(rng, max_angle, max_shift):
    a = rng.uniform(-max_angle, max_angle, 3)
    cx, sx, cy, sy, cz, sz = np.cos(a[0]), np.sin(a[0]), np.cos(a[1]), np.sin(a[1]), np.cos(a[2]), np.sin(a[2])
    T = np.eye(4)
    T[:3, :3] = [[cz * cy, cz * sy * sx - sz * cx, cz * sy * cx + sz * sx],
                 [sz * cy, sz * sy * sx + cz * cx, sz * sy * cx - cz * sx],
                 [-sy, cy * sx, cy * cx]]
    T[:3, 3] = rng.uniform(-max_shift, max_shift, 3)
    return T


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference-root", required=True)
    ap.add_argument("--out", default=os.path.join(os.path.dirname(os.path.abspath(__file__)), "open3d_golden.npz"))
    args = ap.parse_args()
    import open3d as o3d
    reg = o3d.pipelines.registration
    out = {"open3d_version": np.array(o3d.__version__)}
    assert o3d.__version__.startswith("0.17"), f"the reference pins open3d==0.17.0, found {o3d.__version__}"

    def load(k, stride):
        pc = o3d.io.read_point_cloud(os.path.join(args.reference_root, "out", "360_trial_with_floor", "objects", str(k),
                                                  "pointcloud.ply"))
        return np.asarray(pc.points)[::stride].copy(), np.asarray(pc.colors)[::stride].copy()

    def cloud(p, c=None, n=None):
        pc = o3d.geometry.PointCloud()
        pc.points = o3d.utility.Vector3dVector(p)
        if c is not None:
            pc.colors = o3d.utility.Vector3dVector(c)
        if n is not None:
            pc.normals = o3d.utility.Vector3dVector(n)
        return pc

    rng = np.random.default_rng(20240517)
    pA, cA = load(2, 2)            # ~4.9 k points
    pB, cB = load(0, 3)            # ~3.7 k points
    out.update(pA=pA, cA=cA, pB=pB, cB=cB)

    # ---- a5: remove_radius_outlier, the two settings of the reference (memory build 12 / localise 8)
    for nb in (8, 12):
        _, idx = cloud(pA).remove_radius_outlier(nb_points=nb, radius=0.05)
        out[f"outlier_idx_nb{nb}"] = np.asarray(idx, np.int64)

    # ---- a8: normals + FPFH (fpfh_register.py:86-98)
    feats = {}
    for name, p, c in (("A", pA, cA), ("B", pB, cB)):
        pc = cloud(p, c)
        pc.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius=VS * 2, max_nn=30))
        f = reg.compute_fpfh_feature(pc, o3d.geometry.KDTreeSearchParamHybrid(radius=VS * 5, max_nn=100))
        out[f"normals_{name}"] = np.asarray(pc.normals).copy()
        out[f"fpfh_{name}"] = np.asarray(f.data).copy()          # [33, N]
        feats[name] = (pc, f)
    # hybrid search itself (tie order, truncation): neighbour lists of the first 64 points of A
    tree = o3d.geometry.KDTreeFlann(cloud(pA))
    nb_idx = np.full((64, 30), -1, np.int64)
    nb_d2 = np.full((64, 30), np.nan)
    for i in range(64):
        k, idx, d2 = tree.search_hybrid_vector_3d(pA[i], VS * 2, 30)
        nb_idx[i, :k], nb_d2[i, :k] = np.asarray(idx), np.asarray(d2)
    out.update(hybrid_idx=nb_idx, hybrid_d2=nb_d2)

    # ---- a9: a registration problem with a known answer: source = a moved, cropped copy of A; target = A
    T_gt = rigid(rng, 0.6, 0.3)
    keep = pA[:, 0] > np.quantile(pA[:, 0], 0.35)
    pS = (pA[keep] - T_gt[:3, 3]) @ T_gt[:3, :3]                  # T_gt maps source -> target
    cS = cA[keep]
    out.update(pS=pS, cS=cS, T_gt=T_gt)
    src = cloud(pS, cS)
    src.estimate_normals(o3d.geometry.KDTreeSearchParamHybrid(radius=VS * 2, max_nn=30))
    fS = reg.compute_fpfh_feature(src, o3d.geometry.KDTreeSearchParamHybrid(radius=VS * 5, max_nn=100))
    tgt, fT = feats["A"]
    out["fpfh_S"] = np.asarray(fS.data).copy()
    if hasattr(reg, "correspondences_from_features"):
        out["corres_mutual"] = np.asarray(reg.correspondences_from_features(fS, fT, True), np.int64)
        out["corres_oneway"] = np.asarray(reg.correspondences_from_features(fS, fT, False), np.int64)
    # the per-hypothesis arithmetic: 3-point rigid estimate, then the score of that hypothesis over the whole source
    est = reg.TransformationEstimationPointToPoint(False)
    triples = rng.integers(0, len(pS), (32, 3))
    nnA = o3d.geometry.KDTreeFlann(tgt)
    T3, fit3, rmse3, ncor3 = [], [], [], []
    for tri in triples:
        # correspondences (i, nearest target of the TRUE position of i): geometrically consistent triples
        cor = []
        for i in tri:
            q = T_gt[:3, :3] @ pS[i] + T_gt[:3, 3]
            _, j, _ = nnA.search_knn_vector_3d(q, 1)
            cor.append([int(i), int(j[0])])
        T = est.compute_transformation(src, tgt, o3d.utility.Vector2iVector(np.asarray(cor, np.int32)))
        r = reg.evaluate_registration(src, tgt, VS * 1.5, T)
        T3.append(np.asarray(T)); fit3.append(r.fitness); rmse3.append(r.inlier_rmse); ncor3.append(len(r.correspondence_set))
        out.setdefault("triples_corres", []).append(cor)
    out.update(triples_corres=np.asarray(out["triples_corres"], np.int64), triples_T=np.asarray(T3),
               triples_fitness=np.asarray(fit3), triples_rmse=np.asarray(rmse3), triples_ncorres=np.asarray(ncor3, np.int64))
    # one whole RANSAC (thread-local RNG: only a band can be checked -- fitness within a few % and the pose near T_gt)
    o3d.utility.random.seed(7)
    rr = reg.registration_ransac_based_on_feature_matching(
        src, tgt, fS, fT, True, VS * 1.5, reg.TransformationEstimationPointToPoint(False), 3,
        [reg.CorrespondenceCheckerBasedOnEdgeLength(0.9), reg.CorrespondenceCheckerBasedOnDistance(VS * 1.5)],
        reg.RANSACConvergenceCriteria(4000000, 0.99))
    out.update(ransac_T=np.asarray(rr.transformation), ransac_fitness=rr.fitness, ransac_rmse=rr.inlier_rmse)

    # ---- a10: coloured ICP and point-to-point ICP from a perturbed ground truth (deterministic)
    T_init = rigid(rng, 0.03, 0.01) @ T_gt
    out["T_init"] = T_init
    for tag, max_d in (("tight", VS * 0.4), ("loose", VS * 1.5)):       # signature default / the drivers' CLI value
        r = reg.registration_colored_icp(src, tgt, max_d, T_init, reg.TransformationEstimationForColoredICP())
        out.update({f"cicp_{tag}_T": np.asarray(r.transformation), f"cicp_{tag}_fitness": r.fitness,
                    f"cicp_{tag}_rmse": r.inlier_rmse, f"cicp_{tag}_ncorres": len(r.correspondence_set)})
        r = reg.registration_icp(cloud(pS), cloud(pA), max_d, T_init, reg.TransformationEstimationPointToPoint())
        out.update({f"p2p_{tag}_T": np.asarray(r.transformation), f"p2p_{tag}_fitness": r.fitness,
                    f"p2p_{tag}_rmse": r.inlier_rmse, f"p2p_{tag}_ncorres": len(r.correspondence_set)})
    # one single coloured-ICP iteration (max_iteration = 1): pins the normal equations + the 6-vector -> matrix convention
    r = reg.registration_colored_icp(src, tgt, VS * 1.5, T_init, reg.TransformationEstimationForColoredICP(),
                                     reg.ICPConvergenceCriteria(1e-6, 1e-6, 1))
    out.update(cicp_one_T=np.asarray(r.transformation), cicp_one_fitness=r.fitness, cicp_one_rmse=r.inlier_rmse)

    # ---- a11: evaluate_registration at the reference's threshold
    r = reg.evaluate_registration(cloud(pS), cloud(pA), 0.02, T_gt)
    out.update(eval_fitness=r.fitness, eval_rmse=r.inlier_rmse, eval_corres=np.asarray(r.correspondence_set, np.int64))
    r = reg.evaluate_registration(cloud(pB), cloud(pA), 0.02, np.eye(4))
    out.update(eval_id_fitness=r.fitness, eval_id_rmse=r.inlier_rmse, eval_id_corres=np.asarray(r.correspondence_set, np.int64))

    # ---- N2: DBSCAN at the reference's settings, oriented bounding boxes
    pU = np.concatenate([pA, pB + np.array([0.0, 0.0, 0.0])])
    out["pU"] = pU
    for eps, mp in ((0.2, 300), (0.4, 150), (0.05, 12)):
        out[f"dbscan_{eps}_{mp}"] = np.asarray(cloud(pU).cluster_dbscan(eps=eps, min_points=mp, print_progress=False), np.int64)
    for name, p in (("A", pA), ("B", pB)):
        ob = o3d.geometry.OrientedBoundingBox.create_from_points(o3d.utility.Vector3dVector(p))
        out.update({f"obb_{name}_center": np.asarray(ob.center), f"obb_{name}_R": np.asarray(ob.R),
                    f"obb_{name}_extent": np.asarray(ob.extent), f"obb_{name}_corners": np.asarray(ob.get_box_points())})

    np.savez_compressed(args.out, **out)
    print(f"wrote {args.out}: {len(out)} arrays, open3d {o3d.__version__}")


if __name__ == "__main__":
    main()
