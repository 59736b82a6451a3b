"""Driver-side evaluation of a localisation run, vectorised: what the reference's
``*_localisation_trial.py`` drivers do inline, one frame at a time, after ``ObjectMemory.localise`` -- pose errors
(tum_localisation_trial.py:231-232), success rule and cumulative error bins (:255-319), and the ``_results.txt`` report
(:253-344; the 8room / real / irl_rrc drivers carry the same block).  Poses are ``[x, y, z, qw, qx, qy, qz]``.

Host bookkeeping on N x 7 numbers: numpy, no device work.  The report text is byte-identical to the reference's for the
same error lists (running sums in frame order, the same float formatting, the same bin quirks: the last translation /
rotation bin's ``else`` makes "other" count everything NOT below 3.0 / 1.5, NaNs included).
"""
from __future__ import annotations

import io

import numpy as np

from .quaternion_ops import QuaternionOps

TRANSLATION_BINS = (0.1, 0.3, 0.6, 1.0, 1.5, 3.0)       # tum_localisation_trial.py:256
ROTATION_BINS = (0.1, 0.3, 0.6, 1.0, 1.5)               # :257
SUCCESS_TRANSLATION, SUCCESS_ROTATION = 0.6, 0.3        # :274


def pose_errors(target_poses, estimated_poses):
    """-> (translation_error [N], rotation_error [N]) for stacks of 7-vectors (or one pair -> two scalars):
    Euclidean distance of the positions and QuaternionOps.quaternion_error of the orientations (:231-232)."""
    t = np.asarray(target_poses, dtype=np.float64)
    e = np.asarray(estimated_poses, dtype=np.float64)
    if t.shape[-1] != 7 or e.shape != t.shape:
        raise ValueError(f"poses must be [..., 7] and of equal shape, got {t.shape} and {e.shape}")
    d = t[..., :3] - e[..., :3]
    trans = np.sqrt((d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2])
    return trans, QuaternionOps.quaternion_error(t[..., 3:], e[..., 3:])


def _cumulative_bins(err, edges):
    out = {repr(x): int((err < x).sum()) for x in edges}
    out["other"] = int(err.size - (err < edges[-1]).sum())
    return out


def bin_errors(trans_errors, rot_errors):
    """-> (d_tr, r_tr): the reference's cumulative bins, keys '0.1' ... 'other' (:256-257, :290-317)."""
    return (_cumulative_bins(np.asarray(trans_errors, dtype=np.float64).ravel(), TRANSLATION_BINS),
            _cumulative_bins(np.asarray(rot_errors, dtype=np.float64).ravel(), ROTATION_BINS))


def summarise(trans_errors, rot_errors):
    """-> dict(total, successes, success, avg_trans_error, avg_rot_error, median_trans_error, median_rot_error,
    success_rate, d_tr, r_tr).  Averages are the reference's running sums in frame order (``0 + e0 + e1 ...``, :287-288)
    divided by the count, so they round as its loop does."""
    te = np.asarray(trans_errors, dtype=np.float64).ravel()
    re_ = np.asarray(rot_errors, dtype=np.float64).ravel()
    if te.size != re_.size:
        raise ValueError("translation and rotation error lists differ in length")
    if te.size == 0:
        raise ZeroDivisionError("no evaluated frames")                         # the reference divides by total == 0
    ok = (te < SUCCESS_TRANSLATION) & (re_ < SUCCESS_ROTATION)
    d_tr, r_tr = bin_errors(te, re_)
    total = int(te.size)
    return {
        "total": total, "successes": int(ok.sum()), "success": ok,
        "avg_trans_error": np.add.accumulate(te)[-1] / total,                  # sequential, not pairwise
        "avg_rot_error": np.add.accumulate(re_)[-1] / total,
        "median_trans_error": np.median(te), "median_rot_error": np.median(re_),
        "success_rate": int(ok.sum()) / total * 100,
        "d_tr": d_tr, "r_tr": r_tr,
    }


def format_results(trans_errors, rot_errors, chosen_assignments, n_images=None):
    """The text of ``./out/<testname>_results.txt`` (:253-344).  ``chosen_assignments[i]`` is localise()'s second return
    value ``[assignment, moved_objects]``; ``n_images`` is the number the reference prints after "image" (the length
    of the evaluated index list), default the number of frames."""
    s = summarise(trans_errors, rot_errors)
    te = np.asarray(trans_errors, dtype=np.float64).ravel()
    re_ = np.asarray(rot_errors, dtype=np.float64).ravel()
    if len(chosen_assignments) != s["total"]:
        raise ValueError("one [assignment, moved_objects] entry per frame is required")
    n_images = s["total"] if n_images is None else n_images
    w = io.StringIO()
    for i in range(s["total"]):
        w.write(f"Pose {i + 1}, image {n_images}\n"
                f"Translation error: {te[i]}\n"
                f"Rotation errors: {re_[i]}\n"
                f"Assignment: {chosen_assignments[i][0]}\n"
                f"Moved objects: {chosen_assignments[i][1]}\n"
                f"{'SUCCESS' if s['success'][i] else 'MISALIGNED'}\n\n")
    w.write(f"Bagged results for {n_images} eval indices\n")
    for x in TRANSLATION_BINS:
        w.write(f"Translation error less than {x!r}: {s['d_tr'][repr(x)]}\n")
    w.write(f"Translation error greater than {TRANSLATION_BINS[-1]!r}: {s['d_tr']['other']}\n\n")
    for x in ROTATION_BINS:
        w.write(f"Rotation error less than {x!r}: {s['r_tr'][repr(x)]}\n")
    w.write(f"Rotation error greater than {ROTATION_BINS[-1]!r}: {s['r_tr']['other']}\n\n")
    w.write(f"Average Translation Error: {s['avg_trans_error']}\n"
            f"Average Rotation Error: {s['avg_rot_error']}\n"
            f"Median Translation Error: {s['median_trans_error']}\n"
            f"Median Rotation Error: {s['median_rot_error']}\n"
            f"Total Success Rate: {s['success_rate']}\n")
    return w.getvalue()


def write_results(path, trans_errors, rot_errors, chosen_assignments, n_images=None):
    """Write the report to ``path`` (the reference's ``./out/{testname}_results.txt``) and return the summary dict."""
    text = format_results(trans_errors, rot_errors, chosen_assignments, n_images)
    with open(path, "w") as f:
        f.write(text)
    return summarise(trans_errors, rot_errors)
