"""The label gate of the GPU parity tests (BASELINE.json north_star): labels equal the reference's except at
documented near-ties, |delta potential| < 1e-6.

A vertex where the GPU label differs from the oracle's is accepted only if the ORACLE's own float64 direct solve
has its two largest potentials closer than NEAR_TIE_GAP there; every such vertex is recorded (scene tag, count,
largest gap) in gpurun_out/label_parity.jsonl, from which profiles/r02_label_parity.md is written."""
import json
import os

import numpy as np

import orc

NEAR_TIE_GAP = 1e-6
_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LOG = os.path.join(_ROOT, "gpurun_out", "label_parity.jsonl")


def seed_colours(graph, seed_xyz, seed_label):
    """setSeeds + the seed painting of segment() (impl/random_walker_segmentation.hpp:82-101,149-152)."""
    labels = np.unique(seed_label)
    colours = np.zeros(graph.V, np.uint32)
    for s, v in enumerate(graph.nearest_vertices(seed_xyz)):
        if v >= 0:
            colours[v] = 1 + int(np.searchsorted(labels, seed_label[s]))
    return colours, len(labels)


def oracle_gaps(ref_graph, seed_xyz, seed_label, vertices):
    """top-2 gap of the oracle's float64 LDL^T potentials at `vertices` (re-solves on the oracle's final graph)."""
    colours, K = seed_colours(ref_graph, seed_xyz, seed_label)
    st, _, pot, _ = ref_graph.random_walker(colours, K, solver=orc.SOLVE_LDLT_F64, want_potentials=True)
    assert st == 0
    p = np.sort(pot[vertices].astype(np.float64), axis=1)
    if K == 1:
        return p[:, -1]
    return p[:, -1] - p[:, -2]


def check_labels(tag, gpu_labels, ref, seed_xyz, seed_label, bound=NEAR_TIE_GAP, info=None):
    """ref: result of orc.segment_cloud(..., solver=SOLVE_LDLT_F64).  Returns the number of mismatching vertices."""
    ref_labels = ref["vertex_colour"]
    assert len(gpu_labels) == len(ref_labels), (len(gpu_labels), len(ref_labels))
    diff = np.nonzero(np.asarray(gpu_labels) != ref_labels)[0]
    gaps = oracle_gaps(ref["graph"], seed_xyz, seed_label, diff) if len(diff) else np.zeros(0)
    rec = dict(tag=tag, V=int(len(ref_labels)), mismatches=int(len(diff)), max_gap=float(gaps.max()) if len(gaps) else 0.0,
               vertices=[int(v) for v in diff[:32]], gaps=[float(x) for x in gaps[:32]], bound=bound)
    if info:
        rec.update(iterations=int(info.get("iterations", 0)), max_rel_residual=float(max(info.get("rel_residual", [0.0]))))
    if os.path.isdir(os.path.dirname(_LOG)):
        with open(_LOG, "a") as f:
            f.write(json.dumps(rec) + "\n")
    assert len(gaps) == 0 or gaps.max() < bound, \
        "%s: %d label mismatches, largest oracle top-2 gap %.3e (bound %.1e)" % (tag, len(diff), gaps.max(), bound)
    return len(diff)


def check_labels_direct(tag, gpu_labels, ref_graph, colours, K, bound=NEAR_TIE_GAP):
    """Same gate for tests that drive tcs_random_walker / orc random_walker on an existing graph with given seed colours."""
    st, ref_labels, pot, _ = ref_graph.random_walker(colours, K, solver=orc.SOLVE_LDLT_F64, want_potentials=True)
    assert st == 0
    diff = np.nonzero(np.asarray(gpu_labels) != ref_labels)[0]
    gaps = np.zeros(0)
    if len(diff):
        p = np.sort(pot[diff].astype(np.float64), axis=1)
        gaps = p[:, -1] if K == 1 else p[:, -1] - p[:, -2]
    rec = dict(tag=tag, V=int(len(ref_labels)), mismatches=int(len(diff)), max_gap=float(gaps.max()) if len(gaps) else 0.0,
               vertices=[int(v) for v in diff[:32]], gaps=[float(x) for x in gaps[:32]], bound=bound)
    if os.path.isdir(os.path.dirname(_LOG)):
        with open(_LOG, "a") as f:
            f.write(json.dumps(rec) + "\n")
    assert len(gaps) == 0 or gaps.max() < bound, \
        "%s: %d label mismatches, largest oracle top-2 gap %.3e (bound %.1e)" % (tag, len(diff), gaps.max(), bound)
    return len(diff)
