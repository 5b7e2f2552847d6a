"""GPU parity of the nearest-neighbours graph builder (reference:
graph/impl/nearest_neighbors_graph_builder.hpp:49-119) against the oracle's brute-force restatement.

Integer outputs (vertex set, point->vertex map, edge list in insertion order, out-edge lists) are bit-exact; the
full-size case is checked through properties and a brute-force sample of queries.
"""
import numpy as np
import pytest

import orc
from tcs_b200 import capi
from tcs_b200.seeds import seeds_from_labels

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(0)
    yield c
    c.close()


def _subsample(sc, stride):
    """every stride-th row and column of the organised 640 x 480 scene"""
    idx = (np.arange(0, 480, stride)[:, None] * 640 + np.arange(0, 640, stride)[None, :]).ravel()
    return sc["xyz"][idx].copy(), sc["rgba"][idx].copy(), sc["label"][idx].copy()


def _clouds(scenes):
    rng = np.random.default_rng(5)
    out = {}
    out["scene_stride4"] = _subsample(scenes["test1"], 4)[:2]
    out["uniform_volume"] = (rng.random((6000, 3), dtype=np.float32), None)
    # integer lattice: most candidate distances tie exactly, so the (d^2, index) order decides
    gx, gy = np.meshgrid(np.arange(70, dtype=np.float32), np.arange(50, dtype=np.float32))
    lat = np.stack([gx.ravel(), gy.ravel(), np.zeros(gx.size, np.float32)], axis=1)
    out["lattice_ties"] = (lat[rng.permutation(len(lat))], None)
    # dense blobs plus far, isolated outliers (exhaustive fallback) and non-finite points
    blobs = np.concatenate([rng.normal(c, 0.02, (1500, 3)) for c in ((0, 0, 1), (0.5, 0.2, 1.3), (-0.4, 0.1, 0.9))])
    outl = rng.uniform(-40, 40, (25, 3))
    pts = np.concatenate([blobs, outl]).astype(np.float32)
    pts = pts[rng.permutation(len(pts))]
    pts[::97] = np.nan
    pts[5, 1] = np.inf
    out["blobs_outliers_nan"] = (pts, rng.integers(0, 2**32, len(pts), dtype=np.uint32))
    # exact duplicates
    base = rng.random((800, 3), dtype=np.float32)
    out["duplicates"] = (np.concatenate([base, base[:300], base[:50]]), None)
    out["line"] = (np.stack([np.linspace(0, 1, 500, dtype=np.float32), np.zeros(500, np.float32), np.ones(500, np.float32)], 1), None)
    out["single"] = (np.array([[0.1, 0.2, 0.3]], np.float32), None)
    out["pair"] = (np.array([[0.1, 0.2, 0.3], [0.3, 0.2, 0.1]], np.float32), None)
    out["identical"] = (np.tile(np.array([[1.0, 2.0, 3.0]], np.float32), (40, 1)), None)
    out["all_nan"] = (np.full((64, 3), np.nan, np.float32), None)
    out["empty"] = (np.zeros((0, 3), np.float32), None)
    return out


def _compare(g, o):
    assert (g.V, g.E) == (o.V, o.E)
    assert np.array_equal(g.point_to_vertex(), o.point_to_vertex())
    gx, gc, _, _ = g.vertices()
    ox, oc, _, _ = o.vertices()
    assert np.array_equal(gx.view(np.uint32), ox.view(np.uint32))
    assert np.array_equal(gc, oc)
    for a, b in zip(g.edges()[:2], o.edges()[:2]):
        assert np.array_equal(a, b)
    for a, b in zip(g.csr(), o.csr()):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("k", [1, 4, 10, 33, 64])
def test_knn_graph_bit_exact(ctx, scenes, k):
    for name, (xyz, rgba) in _clouds(scenes).items():
        ctx.upload_cloud(xyz, rgba)
        g = ctx.build_nn_graph(k)
        o = orc.Graph.from_cloud_nn(xyz, rgba, k)
        _compare(g, o)
        g.close()


@pytest.mark.parametrize("k,radius", [(10, 0.01), (10, 0.05), (3, 0.5), (40, 2.0), (10, 1e-6), (5, 1.0), (64, 1e3)])
def test_radius_graph_bit_exact(ctx, scenes, k, radius):
    for name, (xyz, rgba) in _clouds(scenes).items():
        ctx.upload_cloud(xyz, rgba)
        g = ctx.build_nn_graph(k, radius, use_radius=True)
        o = orc.Graph.from_cloud_nn(xyz, rgba, k, radius, use_radius=True)
        _compare(g, o)
        g.close()


def test_nn_graph_rejects_bad_arguments(ctx):
    ctx.upload_cloud(np.zeros((4, 3), np.float32), None)
    with pytest.raises(capi.TcsError):
        ctx.build_nn_graph(65)
    with pytest.raises(capi.TcsError):
        ctx.build_nn_graph(-1)
    with pytest.raises(capi.TcsError):
        ctx.build_nn_graph(10, 0.0, use_radius=True)
    g = ctx.build_nn_graph(0)
    assert (g.V, g.E) == (4, 0)


def test_knn_graph_full_scene_properties(ctx, scenes):
    """640 x 480 scene (about 240 k finite points), k = 10: structure properties and a brute-force sample."""
    sc = scenes["test13"]
    xyz = sc["xyz"]
    k = 10
    ctx.upload_cloud(xyz, sc["rgba"])
    g = ctx.build_nn_graph(k)
    finite = np.isfinite(xyz).all(axis=1)
    V = int(finite.sum())
    assert g.V == V
    p2v = g.point_to_vertex()
    assert np.array_equal(p2v[finite], np.arange(V, dtype=np.uint32)) and np.all(p2v[~finite] == 0xFFFFFFFF)
    src, tgt, _ = g.edges()
    assert np.all(src != tgt)
    und = np.minimum(src, tgt).astype(np.uint64) * V + np.maximum(src, tgt)
    assert len(np.unique(und)) == len(und)            # each undirected edge once
    assert np.all(np.diff(src.astype(np.int64)) >= 0)  # insertion order: by querying vertex
    deg = np.bincount(np.concatenate([src, tgt]), minlength=V)
    assert deg.min() >= k and V * k / 2 <= g.E <= V * k
    rp, col, _ = g.csr()
    pts = xyz[finite]
    rng = np.random.default_rng(3)
    for v in rng.integers(0, V, 60):
        d = pts[v] - pts
        d2 = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
        d2[v] = np.inf
        order = np.lexsort((np.arange(V), d2))[:k]
        assert set(order.tolist()) <= set(col[rp[v]:rp[v + 1]].tolist())


def test_random_walker_on_knn_graph(ctx, scenes):
    """The rest of the path runs unchanged on a kNN graph: features, weights and the solve against the oracle."""
    xyz, rgba, label = _subsample(scenes["test1"], 2)
    ctx.upload_cloud(xyz, rgba)
    g = ctx.build_nn_graph(8)
    src, tgt, _ = g.edges()
    gx, gc, _, _ = g.vertices()
    o = orc.Graph.from_edges(gx, src, tgt, rgba=gc)
    for h, mode, terms in ((g, capi.MODE_PARITY, capi.TERMS_CLASS), (o, orc.MODE_PARITY, orc.PRESET_CLASS)):
        h.normals(mode=mode)
        h.signed_curvatures()
        h.edge_weights(terms)
    assert np.array_equal(g.edges()[2].view(np.uint32), o.edges()[2].view(np.uint32))
    sx, sl, _ = seeds_from_labels(xyz, label, 0, 5)
    near = o.nearest_vertices(sx)
    assert np.array_equal(g.nearest_vertices(sx), near)
    labels = np.unique(sl)
    colours = np.zeros(o.V, np.uint32)
    for s, v in enumerate(near):
        colours[v] = 1 + int(np.searchsorted(labels, sl[s]))
    K = len(labels)
    st, cg, _, info = g.random_walker(colours, K)
    assert st == 0
    assert max(info.as_dict()["rel_residual"]) <= 1e-6
    import label_gate
    label_gate.check_labels_direct("test1 subsampled, 8-nearest-neighbour graph", cg, o, colours, K)
