"""SURVEY.md 8(f) row N4 -- the evaluation step after ObjectMemory.localise: the vectorised QuaternionOps / pose errors /
cumulative bins / results report against golden vectors produced by the UNMODIFIED reference code
(tests/golden/make_eval_golden.py: utils/quaternion_ops.py imported, the drivers' report block exec'd from where it lies).
Bar: quaternion product and conjugate bit-exact; angles and distances <= 1e-14 (the reference takes its norms through
BLAS dot, whose rounding is the library's); the report text byte-identical for the same error lists."""
import os

import numpy as np
import pytest

from instance_based_loc_b200.utils import evaluation as ev
from instance_based_loc_b200.utils.quaternion_ops import QuaternionOps


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "eval_golden.npz"))


def _assignments(g):
    out, k = [], 0
    for n in g["assn_len"]:
        out.append([[[int(g["assn_flat"][k + 2 * i]), int(g["assn_flat"][k + 2 * i + 1])] for i in range(n)], None])
        k += 2 * n
    return out


def test_quaternion_ops_match_reference(g):
    q1, q2 = g["q1"], g["q2"]
    assert np.array_equal(QuaternionOps.quaternion_multiply(q1, q2), g["q_mul"])            # stacks, bit-exact
    assert np.array_equal(QuaternionOps.quaternion_conjugate(q1), g["q_conj"])
    for i in range(len(q1)):                                                                # the reference's call shape
        assert np.array_equal(QuaternionOps.quaternion_multiply(q1[i], q2[i]), g["q_mul"][i])
        assert np.array_equal(QuaternionOps.quaternion_conjugate(list(q1[i])), g["q_conj"][i])
        assert abs(QuaternionOps.quaternion_error(q1[i], q2[i]) - g["q_err"][i]) <= 1e-14
    np.testing.assert_allclose(QuaternionOps.quaternion_error(q1, q2), g["q_err"], rtol=0, atol=1e-14)
    with pytest.raises(ValueError):
        QuaternionOps.quaternion_multiply([1.0, 0, 0], [1.0, 0, 0, 0])


def test_quaternion_error_properties():
    rng = np.random.default_rng(5)
    q = rng.normal(size=(1000, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    r = rng.normal(size=(1000, 4))
    r /= np.linalg.norm(r, axis=1, keepdims=True)
    e = QuaternionOps.quaternion_error(q, r)
    assert np.array_equal(e, QuaternionOps.quaternion_error(q, -r))                          # double cover
    np.testing.assert_allclose(e, QuaternionOps.quaternion_error(r, q), rtol=0, atol=1e-14)  # symmetry
    assert np.all((e >= 0) & (e <= np.pi / 2 + 1e-15))                                       # half the rotation angle, folded
    assert np.all(QuaternionOps.quaternion_error(q, q) < 1e-7)
    # the value is half the relative rotation angle (folded to [0, pi/2]): check against rotation matrices
    def R(p):
        w, x, y, z = p.T
        return np.stack([np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
                         np.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
                         np.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], -2)
    tr = np.einsum("nij,nij->n", R(q), R(r))
    np.testing.assert_allclose(2 * e, np.arccos(np.clip((tr - 1) / 2, -1, 1)), atol=1e-6)


def test_pose_errors_match_reference(g):
    te, re_ = ev.pose_errors(g["target_poses"], g["estimated_poses"])
    np.testing.assert_allclose(te, g["trans_errors"], rtol=0, atol=1e-14)
    np.testing.assert_allclose(re_, g["rot_errors"], rtol=0, atol=1e-14)
    t1, r1 = ev.pose_errors(g["target_poses"][3], g["estimated_poses"][3])                   # one pair, as the driver calls it
    assert abs(t1 - g["trans_errors"][3]) <= 1e-14 and abs(r1 - g["rot_errors"][3]) <= 1e-14
    with pytest.raises(ValueError):
        ev.pose_errors(np.zeros((2, 6)), np.zeros((2, 6)))


def test_report_text_matches_reference(g, tmp_path):
    assn = _assignments(g)
    text = ev.format_results(g["trans_errors"], g["rot_errors"], assn)
    assert text == str(g["report"]) == str(g["report_8room"])
    s = ev.write_results(tmp_path / "r.txt", list(g["trans_errors"]), list(g["rot_errors"]), assn)
    assert (tmp_path / "r.txt").read_text() == str(g["report"])
    assert f"Total Success Rate: {s['success_rate']}\n" in text and s["total"] == len(assn)
    # bins are cumulative and consistent with the per-frame verdicts
    assert s["d_tr"]["0.1"] <= s["d_tr"]["0.3"] <= s["d_tr"]["0.6"] <= s["d_tr"]["3.0"] == s["total"] - s["d_tr"]["other"]
    assert text.count("SUCCESS\n") == s["successes"] and text.count("MISALIGNED\n") == s["total"] - s["successes"]


def test_report_edge_cases_match_reference(g):
    """values exactly on the thresholds (strict <), NaN / inf errors (land in 'other', never SUCCESS)."""
    te, re_ = g["edge_trans_errors"], g["edge_rot_errors"]
    with np.errstate(all="ignore"):
        text = ev.format_results(te, re_, [[[], None]] * len(te))
    assert text == str(g["edge_report"])
    d_tr, r_tr = ev.bin_errors(te, re_)
    assert d_tr["other"] == 3 and r_tr["other"] == 2                                         # 3.0, nan, inf / 1.5, nan
    with pytest.raises(ZeroDivisionError):
        ev.summarise([], [])
    with pytest.raises(ValueError):
        ev.format_results([0.1, 0.2], [0.1, 0.2], [[[], None]])
