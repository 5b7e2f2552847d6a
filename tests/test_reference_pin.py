"""The one part of the reference that compiles in this image - its vendored assignment solver
(third-party/hungarian-v2.0, used by src/segmentation_evaluation.cpp:139-165 through src/hungarian/hungarian.cpp) -
pins the evaluation scorer: golden vectors generated from it (tests/golden/hungarian_ref.json, always checked) and
the compiled solver itself (oracle/_ref/libhungarian_ref.so, when present)."""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

import orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libhungarian_ref.so")


def _labels_from_overlap(w):
    """Label arrays whose overlap table is w (ground-truth label 10+r, segmentation label 100+c: no label 0)."""
    r, c = np.nonzero(w)
    cnt = w[r, c]
    return np.repeat(r + 10, cnt).astype(np.uint32), np.repeat(c + 100, cnt).astype(np.uint32)


def _ref_lib():
    if os.path.isdir("/root/reference"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s", "_ref"])
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    return C.CDLL(REF_SO)


def test_scorers_agree_with_the_reference_solver_golden_vectors():
    """Every row is matched (rows <= columns), so #erroneous = #points - weight of the optimal matching, whichever of
    the tied optima a solver returns."""
    from scipy.optimize import linear_sum_assignment
    with open(os.path.join(ROOT, "tests", "golden", "hungarian_ref.json")) as f:
        cases = json.load(f)["cases"]
    assert len(cases) >= 50
    for case in cases:
        w = np.array(case["weights"], np.int64)
        assert int(w[np.arange(len(w)), case["match"]].sum()) == case["weight"]
        r, c = linear_sum_assignment(w, maximize=True)
        assert int(w[r, c].sum()) == case["weight"]
        gt, seg = _labels_from_overlap(w)
        if len(gt):
            assert orc.segmentation_error(gt, seg) == int(w.sum()) - case["weight"]


def test_evaluation_tool_agrees_with_the_reference_solver(tmp_path):
    """The product's scorer (include/tcs/evaluation.h through tcs_segmentation_evaluation) on overlap tables solved
    by the reference's own solver."""
    from tcs_b200 import pcd
    import test_tools
    test_tools._build()
    L = _ref_lib()
    rng = np.random.default_rng(3)
    for it in range(6):
        R = int(rng.integers(2, 12))
        Cc = int(rng.integers(R, 14))
        w = rng.integers(0, [4, 60, 900][it % 3], (R, Cc)).astype(np.int32)
        out = np.zeros(R, np.int32)
        assert L.ref_find_assignment(w.ctypes.data_as(C.c_void_p), R, Cc, out.ctypes.data_as(C.c_void_p)) == R
        want = int(w.sum()) - int(w[np.arange(R), out].sum())
        gt, seg = _labels_from_overlap(w)
        perm = rng.permutation(len(gt))
        gt, seg = gt[perm], seg[perm]
        pts, bad, _ = test_tools._run_eval(tmp_path, gt, seg, np.zeros((len(gt), 3), np.float32), "binary")
        assert pts == len(gt) and bad == want
        assert orc.segmentation_error(gt, seg) == want


def test_reference_triples_overload_gives_the_same_matching_weight():
    """findAssignment(triples) (hungarian.cpp:8-54, the overload the evaluation tool calls) numbers labels in first-
    encounter order and does not transpose; same optimum as the matrix overload."""
    L = _ref_lib()
    rng = np.random.default_rng(4)
    for _ in range(40):
        R = int(rng.integers(1, 10))
        Cc = int(rng.integers(R, 12))
        w = rng.integers(1, 50, (R, Cc)).astype(np.int32)
        w[rng.random((R, Cc)) < 0.3] = 0
        w[np.arange(R), rng.permutation(Cc)[:R]] += 1            # every row and at least R columns occur
        gi, si = np.nonzero(w)
        if len(set(gi.tolist())) > len(set(si.tolist())):
            continue
        l1, l2 = (gi * 3 + 1).astype(np.uint32), (si * 5 + 2).astype(np.uint32)
        cnt = w[gi, si].astype(np.uint64)
        o1, o2 = np.zeros(len(gi), np.uint32), np.zeros(len(gi), np.uint32)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        k = L.ref_find_assignment_triples(p(l1), p(l2), p(cnt), len(gi), p(o1), p(o2))
        got = sum(int(w[(a - 1) // 3, (b - 2) // 5]) for a, b in zip(o1[:k], o2[:k]))
        out = np.zeros(R, np.int32)
        L.ref_find_assignment(p(w), R, Cc, p(out))
        assert got == int(w[np.arange(R), out].sum())
