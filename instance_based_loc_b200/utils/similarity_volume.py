"""Drop-in for the reference's utils/similarity_volume.py ``SimVolume`` (:12-270) on the libibl.so streaming
top-k kernel (ibl_simvolume_topk): the (m+1)^s fp16 sub-volumes are never materialised -- every entry is
recomputed on the fly with the reference's two fp16 roundings and the k best entries per sub-volume come out in
the reference's repeated-argmax order (value desc, first flat index on ties).  SURVEY.md Appendix B.
"""
from __future__ import annotations

import itertools

import numpy as np
import torch

from .. import kernels as K


def assignments_from_topk(flat, val, chosen_objects, m, s, n_det):
    """The bookkeeping half of get_top_indices_from_subvolumes (:226-270) on the top-k entries of every sub-volume
    (flat int64 [n_comb,k] indices into (m+1)^s, val float32 [n_comb,k], both in the reference's argmax order).

    The reference walks the n_comb*k entries in order, drops the "unassigned" coordinates, keeps the FIRST occurrence of
    every distinct assignment with its value, then for each length L takes the max(1, L) best by value (python's
    stable sort) and removes duplicates.  Same result here without the quadratic ``filtered not in assns`` list scans
    (21 ms of interpreter time per 7-detection frame): assignments become integer keys, first occurrences come from
    np.unique, the per-length selection from a stable argsort."""
    flat = np.asarray(flat, np.int64)
    val = np.asarray(val)
    ncomb, k = flat.shape
    chosen = np.asarray(chosen_objects, np.int64).reshape(ncomb, s)
    coords = np.stack(np.unravel_index(flat.reshape(-1), (m + 1,) * s), axis=1)          # [ncomb*k, s]
    det = np.repeat(chosen, k, axis=0)
    assigned = coords != m
    code = np.where(assigned, det * (m + 1) + coords + 1, 0)                               # 0 = dropped coordinate
    # assigned pairs first, in their original (ascending detection) order
    order = np.argsort(~assigned, axis=1, kind="stable")
    code = np.take_along_axis(code, order, axis=1)
    base = int(n_det) * (m + 1) + 2
    key = np.zeros(len(code), np.int64)
    for j in range(s):
        key = key * base + code[:, j]
    length = assigned.sum(axis=1)
    rows = np.nonzero(length > 0)[0]
    _, first = np.unique(key[rows], return_index=True)
    first = rows[np.sort(first)]                                                           # first occurrences, walk order
    vals = val.reshape(-1)[first]
    out = []
    for L in range(1, n_det + 1):
        sel = first[length[first] == L]
        if len(sel) == 0:
            continue
        best = sel[np.argsort(-vals[length[first] == L], kind="stable")[:max(1, L)]]
        for r in best:
            out.append([[int(det[r, j]), int(coords[r, j])] for j in range(s) if assigned[r, j]])
    return out


def _outer_chain(rows):
    """einsum('i,j') followed by einsum('...i,j') per further row on float16 arrays (:47-53, :172-179): every product is
    rounded to fp16 once, left to right."""
    vol = np.multiply.outer(rows[0], rows[1])
    for r in rows[2:]:
        vol = np.multiply.outer(vol, r)
    return vol


def _validity_mask(m, d):
    """Which entries of an (m+1)^d volume the reference's mask loops leave finite (:56-90, :127-154): the assigned
    coordinates (!= m) are pairwise distinct, any subset may be unassigned -- except that the closing
    ``mask[[-1] * (d + 1)] = -inf`` is numpy LIST indexing on axis 0, i.e. the whole slab ``vol[m]``: the first detection
    can never be unassigned."""
    if d > m:          # the loops start from permutations of d DISTINCT objects out of m: none exist, nothing is allowed
        return np.zeros((m + 1,) * d, dtype=bool)
    idx = np.indices((m + 1,) * d)
    ok = np.ones((m + 1,) * d, dtype=bool)
    for a in range(d):
        for b in range(a + 1, d):
            ok &= ~((idx[a] == idx[b]) & (idx[a] != m))
    ok[m] = False
    return ok


class SimVolume():
    def __init__(self, cosine_similarities) -> None:
        """:13-18.  Accepts a numpy array or a CUDA tensor [n_det, m] (float32)."""
        if isinstance(cosine_similarities, torch.Tensor):
            self._S = cosine_similarities.to(device=K._dev(), dtype=torch.float32).contiguous()
        else:
            # the reference assigns into a float16 array: whatever the input dtype, the value is rounded
            # ONCE to fp16.  float64 inputs are first narrowed to fp16 on the host to keep that single rounding.
            a = np.asarray(cosine_similarities)
            if a.dtype == np.float64:
                a = a.astype(np.float16)
            self._S = torch.from_numpy(np.ascontiguousarray(a.astype(np.float32))).to(K._dev())
        self.n_det, self.m = int(self._S.shape[0]), int(self._S.shape[1])
        self.chosen_objects = None
        self._subvolume_size = None

    @property
    def aug(self):
        """fp16 [n_det, m+1] augmented matrix, last column 1 ("unassigned") -- host copy for inspection."""
        aug = np.ones((self.n_det, self.m + 1), dtype=np.float16)
        aug[:, :-1] = self._S.cpu().numpy()
        return aug

    def fast_construct_volume(self, subvolume_size):
        """:102-164.  Only records which detection combinations exist; entries are produced on demand."""
        if self.n_det == 1:
            self.chosen_objects = [[0]]
            self._subvolume_size = 1
            return
        assert self.n_det >= subvolume_size
        if subvolume_size not in (2, 3):
            raise NotImplementedError("the localise path uses sub-volumes of min(n_det, 3) detections")
        if subvolume_size != min(self.n_det, 3):
            raise NotImplementedError("subvolume_size must be min(n_det, 3) (object_memory.py:980)")
        self._subvolume_size = subvolume_size
        self.chosen_objects = [i for i in itertools.combinations(range(self.n_det), subvolume_size)]

    @property
    def subvolumes(self):
        """:156-164 -- the masked fp16 sub-volumes themselves, one per entry of ``chosen_objects``.  The localise path
        never builds them (get_top_indices_from_subvolumes streams their top entries from the kernel); this
        materialises them on the host for callers of the legacy attribute."""
        assert self.chosen_objects is not None, "call fast_construct_volume first"
        aug = self.aug
        if self.n_det == 1:
            v = aug[0].copy()
            v[-1] = -np.inf                                    # :105-110
            return [v]
        allowed = _validity_mask(self.m, self._subvolume_size)
        out = []
        for chosen in self.chosen_objects:
            vol = _outer_chain([aug[c] for c in chosen])
            out.append(np.where(allowed, vol, np.float16(-np.inf)))
        return out

    def top_entries(self, k):
        """-> (flat int64 [n_comb,k], val float32 [n_comb,k]) host arrays, reference argmax order."""
        flat, val = K.simvolume_topk(self._S, k)
        return flat.cpu().numpy(), val.cpu().numpy()

    def get_top_indices_from_subvolumes(self, num_per_length=3):
        """:213-270."""
        assert self.chosen_objects is not None, "call fast_construct_volume first"
        k = num_per_length * self.n_det * 4
        flat, val = self.top_entries(k)
        return assignments_from_topk(flat, val, self.chosen_objects, self.m, self._subvolume_size, self.n_det)

    # ---- legacy API of the reference (not on the localise path; small dense host arrays) --------------------
    def construct_volume(self):
        """:30-96 -- the full (m+1)^n_det volume over ALL detections and its masked copy -> (volume, rep_volume).
        With fewer than two detections the reference prints and returns ``aug`` alone (:33-35)."""
        if self.n_det < 2:
            print("Too few detected embs")
            return self.aug
        volume = _outer_chain(list(self.aug))
        rep = np.where(_validity_mask(self.m, self.n_det), volume, np.float16(-np.inf))
        rep[np.isnan(rep)] = -np.inf
        return volume, rep

    def construct_volume_choose_e(self, chosen_e):
        """:169-180 -- unmasked product volume of the chosen detection rows."""
        assert len(chosen_e) <= self.n_det
        aug = self.aug
        return _outer_chain([aug[c] for c in chosen_e])

    def get_top_indices(self, vol, k):
        """:182-194 -- k rounds of argmax over the flattened volume (first flat index on ties); the volume is
        modified in place like the reference's."""
        top_k = []
        for _ in range(k):
            ind = np.unravel_index(np.argmax(vol, axis=None), vol.shape)
            top_k.append([ind, vol[ind]])
            vol[ind] = -np.inf
        return top_k

    def conv_coords_to_pairs(self, vol, coords):
        """:196-209 (pure bookkeeping, kept for API parity)."""
        assns = []
        unassigned_ind = vol.shape[0] - 1
        for c, cost in coords:
            assn = [[i, c_i] for i, c_i in enumerate(c) if c_i != unassigned_ind]
            if len(assn) == 0:
                continue
            assns.append([assn, cost])
        return assns
