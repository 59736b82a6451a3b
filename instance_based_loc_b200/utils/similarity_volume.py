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
        raise NotImplementedError("sub-volumes are not materialised on the device path; use "
                                  "get_top_indices_from_subvolumes (the only consumer in localise)")

    def top_entries(self, k):
        """-> (flat int64 [n_comb,k], val float32 [n_comb,k]) host arrays, reference argmax order."""
        flat, val = K.simvolume_topk(self._S, k)
        return flat.cpu().numpy(), val.cpu().numpy()

    def get_top_indices_from_subvolumes(self, num_per_length=3):
        """:213-270."""
        assert self.chosen_objects is not None, "call fast_construct_volume first"
        k = num_per_length * self.n_det * 4
        flat, val = self.top_entries(k)
        s = self._subvolume_size
        shape = (self.m + 1,) * s
        unassigned_ind = self.m
        all_filtered_topk = []
        assns = []
        for ci, chosen in enumerate(self.chosen_objects):
            coords = np.stack(np.unravel_index(flat[ci], shape), axis=1)
            for r in range(k):
                filtered = [[int(i), int(c)] for i, c in zip(chosen, coords[r]) if c != unassigned_ind]
                if len(filtered) == 0:
                    continue
                if filtered not in assns:
                    assns.append(filtered)
                    all_filtered_topk.append([filtered, val[ci, r]])
        filtered_topk = []
        for i in range(1, self.n_det + 1):
            to_pick = max(1, i)
            correct_length = [f for f in all_filtered_topk if len(f[0]) == i]
            correct_length = sorted(correct_length, key=lambda x: x[-1], reverse=True)[:to_pick]
            filtered_topk += correct_length
        assns = [a[0] for a in filtered_topk]
        no_dupe_assns = []
        for a in assns:
            if a not in no_dupe_assns:
                no_dupe_assns.append(a)
        return no_dupe_assns

    # ---- legacy API of the reference (not on the localise path) ---------------------------------------------
    def construct_volume(self):
        raise NotImplementedError("legacy full-volume path (similarity_volume.py:30-96) is not on the localise path")

    def construct_volume_choose_e(self, chosen_e):
        raise NotImplementedError("legacy path (similarity_volume.py:169-180) is not on the localise path")

    def get_top_indices(self, vol, k):
        raise NotImplementedError("legacy path (similarity_volume.py:182-194) is not on the localise path")

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
