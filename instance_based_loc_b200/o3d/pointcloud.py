"""Device-resident PointCloud with the open3d 0.17 surface the reference's localisation path uses
(call sites utils/depth_utils.py:36-44,80-90, utils/fpfh_register.py:86-98,
object_memory/object_memory.py:992-1034, object_memory/object_info.py:63-76).

Points / colours / normals are CUDA float64 [N,3] tensors.  Every geometric method is a libibl.so
kernel (include/ibl.h); numpy only appears where the caller asks for host data (np.asarray).
"""
from __future__ import annotations

import numpy as np
import torch

from .. import kernels as K

F64 = torch.float64


class KDTreeSearchParamHybrid:
    def __init__(self, radius, max_nn):
        self.radius = float(radius)
        self.max_nn = int(max_nn)


class KDTreeSearchParamKNN:
    def __init__(self, knn=30):
        self.radius = float("inf")
        self.max_nn = int(knn)


class KDTreeSearchParamRadius:
    def __init__(self, radius):
        self.radius = float(radius)
        self.max_nn = 0


class DeviceVec:
    """Read view of a device [N,c] tensor that quacks like open3d.utility.Vector3dVector:
    np.asarray(v) (one D2H copy, cached), len(v), v[i]."""

    __slots__ = ("t", "_host")

    def __init__(self, t):
        self.t = t
        self._host = None

    def _h(self):
        if self._host is None:
            self._host = self.t.detach().cpu().numpy()
        return self._host

    def __array__(self, dtype=None, copy=None):
        h = self._h()
        return h if dtype is None else h.astype(dtype, copy=False)

    def __len__(self):
        return int(self.t.shape[0])

    def __getitem__(self, i):
        return self._h()[i]

    def __iter__(self):
        return iter(self._h())

    @property
    def shape(self):
        return tuple(self.t.shape)

    def __repr__(self):
        return f"DeviceVec({tuple(self.t.shape)}, {self.t.dtype}, {self.t.device})"


def Vector3dVector(arr=None):
    """Host staging object: assignment to PointCloud.points/colors/normals uploads it once."""
    if arr is None:
        return np.zeros((0, 3), np.float64)
    if isinstance(arr, (DeviceVec, torch.Tensor)):
        return arr
    return np.array(arr, dtype=np.float64).reshape(-1, 3)


def Vector2iVector(arr=None):
    if arr is None:
        return np.zeros((0, 2), np.int32)
    return np.array(arr, dtype=np.int32).reshape(-1, 2)


def _as_dev(v):
    if isinstance(v, DeviceVec):
        return v.t
    if isinstance(v, torch.Tensor):
        return K.to_dev(v).reshape(-1, 3)
    return K.to_dev(np.asarray(v, dtype=np.float64).reshape(-1, 3))


def remove_radius_outlier_many(clouds, nb_points, radius):
    """``[c.remove_radius_outlier(nb_points, radius)[0] for c in clouds]`` (A.2) with ONE grid build, ONE count kernel and
    ONE compaction for all clouds: the clouds are concatenated under a segmented grid whose radius query only counts
    neighbours of the query's own cloud, so every keep flag is the one the cloud's own grid gives.  The per-mask filters of
    a frame (utils/depth_utils.py:85-88; object_memory.py:992-998) go from two host round trips per cloud to three per
    frame."""
    if nb_points < 1 or radius <= 0:
        raise RuntimeError("Illegal input parameters, number of points and radius must be positive")
    live = [i for i, c in enumerate(clouds) if c.n() > 0]
    out = [PointCloud() for _ in clouds]
    if not live:
        return out
    if len(live) == 1:
        out[live[0]] = clouds[live[0]].remove_radius_outlier(nb_points, radius)[0]
        return out
    dev = clouds[live[0]]._p.device
    sizes = [clouds[i].n() for i in live]
    pts = torch.cat([clouds[i]._p for i in live]).contiguous()
    colored = all(clouds[i].has_colors() for i in live)
    cols = torch.cat([clouds[i]._c for i in live]).contiguous() if colored else None
    seg = torch.repeat_interleave(torch.arange(len(live), dtype=torch.int32, device=dev),
                                  torch.as_tensor(sizes, device=dev), output_size=int(sum(sizes)))
    grid = K.Grid(pts, radius * 1.001, seg=seg)
    keep = K.radius_outlier_mask(grid, radius, nb_points)
    ends = torch.as_tensor(np.cumsum(sizes) - 1, device=dev)
    kept = torch.cumsum(keep, dim=0, dtype=torch.int64)[ends].cpu().numpy()          # survivors up to the end of every cloud
    p2, c2, _, _ = K.select_rows(keep, pts, cols, None)
    a = 0
    for i, b in zip(live, kept.tolist()):
        if colored:
            out[i] = PointCloud.from_tensors(p2[a:b], c2[a:b], spacing=clouds[i]._spacing)
        else:                                                   # mixed: colours stay with the clouds that have them
            out[i] = PointCloud.from_tensors(p2[a:b], None, spacing=clouds[i]._spacing)
        a = b
    if not colored and any(clouds[i].has_colors() for i in live):
        for i in live:                                          # rare mixed case: fall back to the per-cloud filter
            out[i] = clouds[i].remove_radius_outlier(nb_points, radius)[0]
    return out


class PointCloud:
    def __init__(self, points=None):
        self._p = None
        self._c = None
        self._n = None
        self._grids = {}
        self._color_grad = None
        self._spacing = None          # mean point spacing, a PERFORMANCE hint for sizing search grids (never affects results)
        self._knn_hint = None         # (radius, max_nn) of a larger search that will follow on this cloud (see knn)
        if points is not None:
            self.points = points

    # ------------------------------------------------------------------ storage
    @staticmethod
    def from_tensors(p, c=None, n=None, spacing=None):
        r = PointCloud()
        r._p, r._c, r._n = p, c, n
        r._spacing = spacing
        return r


    def _empty(self):
        return torch.zeros((0, 3), dtype=F64, device=K._dev())

    def _invalidate(self):
        self._grids = {}
        self._color_grad = None

    @property
    def points(self):
        return DeviceVec(self._p if self._p is not None else self._empty())

    @points.setter
    def points(self, v):
        self._p = _as_dev(v)
        self._invalidate()

    @property
    def colors(self):
        return DeviceVec(self._c if self._c is not None else self._empty())

    @colors.setter
    def colors(self, v):
        self._c = _as_dev(v)
        self._color_grad = None

    @property
    def normals(self):
        return DeviceVec(self._n if self._n is not None else self._empty())

    @normals.setter
    def normals(self, v):
        self._n = _as_dev(v)
        self._color_grad = None

    def n(self):
        return 0 if self._p is None else int(self._p.shape[0])

    def has_points(self):
        return self.n() > 0

    def has_colors(self):
        return self.n() > 0 and self._c is not None and self._c.shape[0] == self.n()

    def has_normals(self):
        return self.n() > 0 and self._n is not None and self._n.shape[0] == self.n()

    def is_empty(self):
        return self.n() == 0

    def __repr__(self):
        return f"PointCloud with {self.n()} points."

    # ------------------------------------------------------------------ open3d A.10
    def __iadd__(self, o):
        if o.is_empty():
            return self
        # spacing hint of the union: areas add (area ~ n * spacing^2)
        if self.is_empty():
            self._spacing = o._spacing
        elif self._spacing and o._spacing:
            n1, n2 = self.n(), o.n()
            self._spacing = float(((n1 * self._spacing ** 2 + n2 * o._spacing ** 2) / (n1 + n2)) ** 0.5)
        else:
            self._spacing = None
        if (self.is_empty() or self.has_normals()) and o.has_normals():
            self._n = o._n if self.is_empty() else torch.cat([self._n, o._n])
        else:
            self._n = None
        if (self.is_empty() or self.has_colors()) and o.has_colors():
            self._c = o._c if self.is_empty() else torch.cat([self._c, o._c])
        else:
            self._c = None
        self._p = o._p if self.is_empty() else torch.cat([self._p, o._p])
        self._invalidate()
        return self

    def __add__(self, o):
        r = PointCloud()
        r += self
        r += o
        return r

    def translate(self, v, relative=True):
        if self.is_empty():
            return self
        v = np.asarray(v, np.float64).reshape(3)
        if not relative:
            v = v - self.get_center()
        self._p = self._p + torch.from_numpy(v).to(self._p.device)   # one rounded f64 add per coordinate
        self._invalidate()
        return self

    def transform(self, T):
        if self.is_empty():
            return self
        T = np.asarray(T, np.float64)
        self._p = K.transform_points(T, self._p)
        if self.has_normals():
            self._n = K.transform_points(T, self._n, rotate_only=True)
        self._invalidate()
        return self

    def get_center(self):
        return K.mean3(self._p) if self.n() else np.zeros(3)

    def select_by_index(self, idx, invert=False):
        n = self.n()
        keep = torch.zeros(n, dtype=torch.uint8, device=K._dev())
        it = idx.t.long() if isinstance(idx, DeviceVec) else torch.as_tensor(np.asarray(idx, np.int64)).to(keep.device)
        keep[it] = 1
        if invert:
            keep = 1 - keep
        return self._select_mask(keep)[0]

    def _select_mask(self, keep, want_index=False):
        p, c, nn, oi = K.select_rows(keep, self._p, self._c if self.has_colors() else None,
                                     self._n if self.has_normals() else None, want_index=want_index)
        return PointCloud.from_tensors(p, c, nn, spacing=self._spacing), oi

    def remove_radius_outlier(self, nb_points, radius, print_progress=False):
        """A.2: keep i iff #{j: d2(i,j) < r^2} (self included) > nb_points; order preserved."""
        if nb_points < 1 or radius <= 0:
            raise RuntimeError("Illegal input parameters, number of points and radius must be positive")
        if self.is_empty():
            return PointCloud(), []
        # cell a hair larger than the radius: the search then needs 1 ring of cells (3^3) instead of the 2 rings (5^3) that
        # cell == radius costs once the cell-coordinate rounding guard is added (radius_launch's S)
        keep = K.radius_outlier_mask(self.grid(radius * 1.001), radius, nb_points)
        out, oi = self._select_mask(keep, want_index=True)
        return out, DeviceVec(oi)

    def estimate_normals(self, search_param=None, fast_normal_computation=True):
        sp = search_param or KDTreeSearchParamKNN(30)
        if self.is_empty():
            return
        idx, _, cnt = self.knn(sp.radius, sp.max_nn, want_d2=False)
        self._n = K.normals(self._p, idx, cnt)
        self._color_grad = None

    def paint_uniform_color(self, c):
        c = torch.as_tensor(np.asarray(c, np.float64).reshape(1, 3)).to(K._dev())
        self._c = c.expand(self.n(), 3).contiguous()
        return self

    def voxel_down_sample(self, voxel_size):
        """open3d's own voxel_down_sample (map build only; output order is unspecified upstream)."""
        from ..utils.depth_utils import voxel_down_sample_with_colors
        return voxel_down_sample_with_colors(self, voxel_size)

    def cluster_dbscan(self, eps, min_points, print_progress=False):
        """Open3D ClusterDBSCAN (memory build: object_memory.py:315,511,632): one int label per point, -1 = noise, clusters
        numbered in the order the index-order scan meets them.  Returned as a host int32 array (the reference wraps the
        IntVector in np.array at once); ``cluster_dbscan_device`` keeps the labels on the device."""
        return self.cluster_dbscan_device(eps, min_points)[0].cpu().numpy()

    def cluster_dbscan_device(self, eps, min_points):
        if eps <= 0:
            raise RuntimeError("eps must be positive")
        if self.is_empty():
            return torch.zeros(0, dtype=torch.int32, device=K._dev()), 0
        return K.dbscan(self.grid(eps * 1.001), eps, min_points)

    def get_min_bound(self):
        return K.bounds3(self._p)[0] if self.n() else np.zeros(3)

    def get_max_bound(self):
        return K.bounds3(self._p)[1] if self.n() else np.zeros(3)

    # ------------------------------------------------------------------ neighbour plumbing
    def grid(self, cell, ring=0, max_cells=0, nnfield=0.0, nnfield_voxel=0.0):
        """Cached cell-sorted grid (optionally with ring field / exact NN candidate field) for this cloud;
        invalidated on any mutation."""
        key = (float(cell), int(ring), int(max_cells), float(nnfield), float(nnfield_voxel))
        g = self._grids.get(key)
        if g is None:
            g = K.Grid(self._p, cell, max_cells=max_cells, ring=ring, nnfield=nnfield, nnfield_voxel=nnfield_voxel)
            self._grids[key] = g
        return g

    def knn(self, radius, max_nn, want_d2=True):
        """Hybrid search of every point among the cloud's own points -> (idx int32 [n, stride], d2 | None, cnt int32 [n]);
        row i holds cnt[i] neighbours in ascending (d2, index) order, ``stride = idx.shape[1] >= max_nn``.

        A search is a prefix of every LARGER search on the same cloud: the max_nn nearest within ``radius`` are the first
        entries of the k' >= max_nn nearest within r' >= radius that lie inside ``radius`` (same order, same tie rule).
        So when a larger search with distances is already cached -- or announced through ``_knn_hint``, as
        downsample_and_compute_fpfh does for its FPFH search before it estimates the normals -- the result is derived
        from it (one comparison kernel) instead of searched again: normals (0.1 / 30) and FPFH (0.25 / 100) share one
        search per cloud."""
        r = radius if np.isfinite(radius) else 1e30
        last = self._grids.get("knn_last")
        hint = self._knn_hint
        if last is None and hint is not None and hint[0] >= r and hint[1] >= max_nn and (hint[0], hint[1]) != (r, max_nn):
            self._knn_hint = None
            self.knn(hint[0], hint[1], want_d2=True)                   # the larger search, cached below
            last = self._grids.get("knn_last")
        if last is not None and last[0] >= r and last[1] >= max_nn and self.n() > 0:
            R, Kmax, idx, d2, cnt = last
            if (R, Kmax) == (r, max_nn):
                return idx, (d2 if want_d2 else None), cnt
            sub = (d2[:, :max_nn] < r * r).sum(dim=1).to(torch.int32)    # d2 ascending, +inf beyond cnt
            return idx, (d2 if want_d2 else None), sub
        if self.n() == 0:
            cell = r
        else:
            # spacing hint: carried from the back-projection (depth / focal length) through filters, unions and rigid
            # moves; without one it is estimated with a probe grid, kept with the grids (dropped on any mutation)
            sp = self._spacing or self._grids.get("spacing")
            if sp is None:
                sp = self._grids["spacing"] = K.surface_spacing(self._p, min(r / 4.0, 0.05))
            cell = K.knn_cell_from_spacing(sp, max_nn, r)
        out = K.knn_hybrid(self.grid(cell), radius, max_nn, want_d2=want_d2)
        if want_d2:
            self._grids["knn_last"] = (r, max_nn) + tuple(out)
        return out

    # ------------------------------------------------------------------ copy / pickle
    def __deepcopy__(self, memo):
        cl = lambda t: None if t is None else t.clone()  # noqa: E731
        return PointCloud.from_tensors(cl(self._p), cl(self._c), cl(self._n), spacing=self._spacing)

    def __getstate__(self):
        h = lambda t: None if t is None else t.cpu().numpy()  # noqa: E731
        return {"points": h(self._p), "colors": h(self._c), "normals": h(self._n)}

    def __setstate__(self, st):
        self._grids = {}
        self._color_grad = None
        d = lambda a: None if a is None else K.to_dev(a)  # noqa: E731
        self._p, self._c, self._n = d(st["points"]), d(st["colors"]), d(st["normals"])
