"""tcs_nearest_vertices_noalloc (no counterpart in the reference): the
allocation-free seed look-up, called through the C ABI, against the existing tcs_nearest_vertices entry point.

Integer results: the bar is bit-exact, like every other index output of the path.  The file is named so that pytest
collects it last."""
import numpy as np
import pytest

from tcs_b200 import capi
from tcs_b200.seeds import seeds_from_labels

pytestmark = pytest.mark.gpu


def _seed_sets(sc):
    xyz = sc["xyz"]
    rng = np.random.default_rng(5)
    finite = np.nonzero(np.isfinite(xyz).all(axis=1))[0]
    # the largest allowed set: random finite scene points, jittered off the voxel centroids, plus a far outlier
    big = xyz[rng.choice(finite, capi.NOALLOC_MAX_SEEDS, replace=False)].astype(np.float32)
    big += rng.normal(0, 0.002, big.shape).astype(np.float32)
    big[7] = (50.0, -50.0, 80.0)
    # the recipe's own seeds (5 per label), with one NaN seed: no vertex is nearest -> -1 from both entry points
    small, _, _ = seeds_from_labels(xyz, sc["label"], sc["id"], 5)
    small = np.vstack([small, np.array([[np.nan, 0.0, 1.0]], np.float32)]).astype(np.float32)
    return big, small


def test_noalloc_matches_the_allocating_entry_point(scenes):
    sc = scenes["test1"]
    ctx = capi.Context(0)
    try:
        ctx.upload_cloud(sc["xyz"], sc["rgba"])
        g = ctx.build_voxel_graph(0.006)
        big, small = _seed_sets(sc)
        # ns = 0: set-up only
        assert len(g.nearest_vertices_noalloc(np.zeros((0, 3), np.float32))) == 0
        # two calls in a row on the shared scratch (the second, smaller one must not see the first one's leftovers),
        # then the first again
        for seeds in (big, small, big, small):
            want = g.nearest_vertices(seeds)
            got = g.nearest_vertices_noalloc(seeds)
            assert np.array_equal(got, want)
            assert got.min() >= -1 and got.max() < g.V
        assert g.nearest_vertices_noalloc(small)[-1] == -1
        # beyond the fixed maximum: a clear error, not a silent overflow of the scratch
        too_many = np.zeros((capi.NOALLOC_MAX_SEEDS + 1, 3), np.float32)
        with pytest.raises(capi.TcsError) as e:
            g.nearest_vertices_noalloc(too_many)
        assert e.value.code == capi.ERR_INPUT and "TCS_NOALLOC_MAX_SEEDS" in str(e.value)
        # the entry point still works after the rejected call
        assert np.array_equal(g.nearest_vertices_noalloc(small), g.nearest_vertices(small))
        g.close()
    finally:
        ctx.close()


def test_noalloc_on_a_second_graph_and_context(scenes):
    """Another context and a differently sized graph reuse the same array (same device): results stay exact."""
    sc = scenes["test47"]
    ctx = capi.Context(0)
    try:
        ctx.upload_cloud(sc["xyz"], sc["rgba"])
        g = ctx.build_voxel_graph(0.01)
        big, small = _seed_sets(sc)
        for seeds in (small, big):
            assert np.array_equal(g.nearest_vertices_noalloc(seeds), g.nearest_vertices(seeds))
        g.close()
    finally:
        ctx.close()
