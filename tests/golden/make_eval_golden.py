"""Generates tests/golden/eval_golden.npz by running the UNMODIFIED reference code in the
build container:

 * ``utils/quaternion_ops.py`` is imported from /root/reference and called on seeded quaternion pairs (plus the edge
   cases: identical, antipodal -q, 180 degrees, unnormalised, zero);
 * the results-report block of ``tum_localisation_trial.py`` (from ``f = open(f"./out/...`` to ``f.close()``, lines
   253-344) is inline in the driver's main function, so it cannot be imported: its source lines are read from where
   they lie, dedented and exec'd on prepared error lists inside a temporary directory.  Nothing is copied into the
   repo; the report TEXT it writes is the fixture.

The reference tree does not exist on the GPU box; this script is the provenance.  Run: python tests/golden/make_eval_golden.py
"""
from __future__ import annotations

import contextlib
import io
import os
import sys
import tempfile
import textwrap
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402


def reference_report(trans_errors, rot_errors, chosen_assignments, n_indices, driver="tum_localisation_trial.py"):
    """Run the reference driver's own report block on the given lists -> the text of ./out/<testname>_results.txt."""
    src = open(os.path.join(ref_loader.REF, driver)).read().split("\n")
    first = next(i for i, l in enumerate(src) if l.strip().startswith('f = open(f"./out/{args.testname}_results.txt"'))
    last = next(i for i in range(first, len(src)) if src[i].strip() == "f.close()")
    block = textwrap.dedent("\n".join(src[first:last + 1]))
    ns = {
        "np": np, "tqdm": lambda it, **kw: it,
        "args": types.SimpleNamespace(testname="golden"),
        "eval_dataloader": types.SimpleNamespace(environment_indices=list(range(n_indices))),
        "trans_errors": list(trans_errors), "rot_errors": list(rot_errors),
        "chosen_assignments": chosen_assignments,
    }
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as tmp:
        os.makedirs(os.path.join(tmp, "out"))
        os.chdir(tmp)
        try:
            with contextlib.redirect_stdout(io.StringIO()):
                exec(compile(block, driver, "exec"), ns)
            return open(os.path.join(tmp, "out", "golden_results.txt")).read()
        finally:
            os.chdir(cwd)


def cases(rng):
    """Seeded pose lists: [N, 7] target / estimate pairs spanning every bin, with the quaternion edge cases."""
    n = 64
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    ang = np.concatenate([rng.uniform(0, 0.09, 16), rng.uniform(0.09, 0.7, 24), rng.uniform(0.7, np.pi, 24)])
    axis = rng.normal(size=(n, 3))
    axis /= np.linalg.norm(axis, axis=1, keepdims=True)
    dq = np.c_[np.cos(ang / 2), np.sin(ang / 2)[:, None] * axis]
    return q, dq, n


def main():
    qo = ref_loader.load("utils/quaternion_ops.py").QuaternionOps
    rng = np.random.default_rng(4041)
    out = {}
    q1, dq, n = cases(rng)
    q2 = np.stack([qo.quaternion_multiply(q1[i], dq[i]) for i in range(n)])
    q2[::3] *= -1.0                                                   # the other cover of the same rotation
    edge1 = np.array([[1., 0, 0, 0], [0.5, 0.5, 0.5, 0.5], [1., 0, 0, 0], [2., 0, 0, 0], [0., 0, 0, 0], [0.6, 0, 0.8, 0]])
    edge2 = np.array([[1., 0, 0, 0], [-0.5, -0.5, -0.5, -0.5], [0., 1, 0, 0], [0., 0, 3, 0], [1., 0, 0, 0], [0.6, 0, 0.8, 0]])
    q1, q2 = np.concatenate([q1, edge1]), np.concatenate([q2, edge2])
    out["q1"], out["q2"] = q1, q2
    out["q_mul"] = np.stack([qo.quaternion_multiply(a, b) for a, b in zip(q1, q2)])
    out["q_conj"] = np.stack([qo.quaternion_conjugate(a) for a in q1])
    out["q_err"] = np.array([qo.quaternion_error(a, b) for a, b in zip(q1, q2)])

    # pose lists as the driver builds them (:231-232), then its report
    m = len(q1)
    tgt_xyz = rng.uniform(-3, 3, (m, 3))
    step = np.concatenate([rng.uniform(0, 0.12, m // 3), rng.uniform(0.05, 1.2, m // 3), rng.uniform(0.5, 4.5, m - 2 * (m // 3))])
    dirs = rng.normal(size=(m, 3))
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    target = np.c_[tgt_xyz, q1]
    estimate = np.c_[tgt_xyz + dirs * step[:, None], q2]
    out["target_poses"], out["estimated_poses"] = target, estimate
    te = [np.linalg.norm(target[i, :3] - estimate[i, :3]) for i in range(m)]
    re_ = [qo.quaternion_error(target[i, 3:], estimate[i, 3:]) for i in range(m)]
    out["trans_errors"], out["rot_errors"] = np.array(te), np.array(re_)
    assn = [[[[int(a), int(b)] for a, b in zip(range(i % 4), rng.integers(0, 40, i % 4))], None] for i in range(m)]
    out["assn_len"] = np.array([len(a[0]) for a in assn], np.int32)
    out["assn_flat"] = np.array([x for a in assn for pair in a[0] for x in pair], np.int32)
    out["report"] = np.array(reference_report(te, re_, assn, m))
    # a second report with the bin-edge quirks: values exactly on the thresholds, a NaN and an inf
    te2 = [np.float64(x) for x in (0.1, 0.3, 0.6, 0.5999999999999999, 1.0, 1.5, 3.0, 2.9999, np.nan, np.inf, 0.0)]
    re2 = [np.float64(x) for x in (0.3, 0.29999999999999993, 0.1, 0.0, 1.5, 1.4999, np.nan, 0.6, 0.2, 0.1, 0.0)]
    assn2 = [[[], None]] * len(te2)
    out["edge_trans_errors"], out["edge_rot_errors"] = np.array(te2), np.array(re2)
    with np.errstate(all="ignore"):
        out["edge_report"] = np.array(reference_report(te2, re2, assn2, len(te2)))
    # the 8room driver carries the same block: its report for the first list must be the same text
    out["report_8room"] = np.array(reference_report(te, re_, assn, m, driver="8room_localisation_trial.py"))

    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "eval_golden.npz"), **out)
    print("wrote eval_golden.npz with", len(out), "arrays;", "8room == tum:", str(out["report_8room"]) == str(out["report"]))


if __name__ == "__main__":
    main()
