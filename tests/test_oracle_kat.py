"""Known-answer tests pinning the CPU oracle from first principles (the reference ships no tests or golden vectors,
SURVEY.md section 4): hand-solvable graphs for the random walker, analytic cases for every feature stage, and a scipy
float64 cross-check of the linear solve.  CPU only."""
import numpy as np
import pytest

import orc


def path_graph(n, w=None):
    xyz = np.zeros((n, 3), np.float32)
    xyz[:, 0] = np.arange(n) * 0.01
    src = np.arange(1, n, dtype=np.uint32)      # add_edge(v, v - 1): like the voxel builder, towards visited vertices
    tgt = np.arange(0, n - 1, dtype=np.uint32)
    return orc.Graph.from_edges(xyz, src, tgt, weight=np.ones(n - 1, np.float32) if w is None else w)


@pytest.mark.parametrize("solver", [orc.SOLVE_LDLT_F32, orc.SOLVE_LDLT_F64, orc.SOLVE_PCG_F64])
def test_path_graph_linear_potentials(solver):
    """Unit-weight path with seeds at both ends: the potential of the left seed falls linearly from 1 to 0."""
    n = 11
    g = path_graph(n)
    colours = np.zeros(n, np.uint32)
    colours[0], colours[-1] = 1, 2
    st, out, pot, res = g.random_walker(colours, 2, solver=solver, want_potentials=True)
    assert st == 0
    expect = 1.0 - np.arange(n) / (n - 1.0)
    np.testing.assert_allclose(pot[:, 0], expect, atol=2e-6)
    np.testing.assert_allclose(pot[:, 1], 1.0 - expect, atol=2e-6)
    assert list(out[:5]) == [1] * 5 and list(out[6:]) == [2] * 5
    assert out[5] == 1          # exact tie at the midpoint: first maximal column wins (Eigen maxCoeff)
    assert res.max() < 1e-5


def test_weighted_path_is_a_resistor_chain():
    """Potentials on a path are a voltage divider: drop across edge i proportional to 1 / w_i."""
    w = np.array([1.0, 0.5, 0.25, 2.0], np.float32)
    g = path_graph(5, w)
    colours = np.array([1, 0, 0, 0, 2], np.uint32)
    st, out, pot, _ = g.random_walker(colours, 2, solver=orc.SOLVE_LDLT_F64, want_potentials=True)
    r = 1.0 / w
    expect = 1.0 - np.concatenate([[0], np.cumsum(r)]) / r.sum()
    np.testing.assert_allclose(pot[:, 0], expect, atol=1e-6)


def test_star_graph_weighted_average():
    """Centre of a star = weighted mean of the seeded leaves."""
    xyz = np.zeros((5, 3), np.float32)
    src = np.array([1, 2, 3, 4], np.uint32)
    tgt = np.zeros(4, np.uint32)
    w = np.array([1.0, 2.0, 3.0, 4.0], np.float32)
    g = orc.Graph.from_edges(xyz, src, tgt, weight=w)
    colours = np.array([0, 1, 1, 2, 3], np.uint32)
    st, out, pot, _ = g.random_walker(colours, 3, solver=orc.SOLVE_LDLT_F64, want_potentials=True)
    np.testing.assert_allclose(pot[0], [0.3, 0.3, 0.4], atol=1e-6)
    assert out[0] == 3


def test_seedless_component_stays_unlabelled_and_single_colour():
    """Two components: one with a single seed colour (everything takes it), one without seeds (stays 0,
    impl/random_walker.hpp:191-192)."""
    xyz = np.zeros((7, 3), np.float32)
    src = np.array([1, 2, 3, 5, 6], np.uint32)
    tgt = np.array([0, 1, 2, 4, 5], np.uint32)
    g = orc.Graph.from_edges(xyz, src, tgt, weight=np.ones(5, np.float32))
    nc, comp = g.components()
    assert nc == 2 and list(comp) == [0, 0, 0, 0, 1, 1, 1]
    colours = np.array([0, 2, 0, 0, 0, 0, 0], np.uint32)
    st, out, pot, _ = g.random_walker(colours, 2, want_potentials=True)
    assert st == 0
    assert list(out) == [2, 2, 2, 2, 0, 0, 0]
    assert np.all(pot[4:] == 0)


def test_isolated_vertex_is_excluded():
    xyz = np.zeros((4, 3), np.float32)
    g = orc.Graph.from_edges(xyz, np.array([1, 2], np.uint32), np.array([0, 1], np.uint32), weight=np.ones(2, np.float32))
    colours = np.array([1, 0, 2, 0], np.uint32)
    st, out, _, _ = g.random_walker(colours, 2)
    assert st == 0 and out[3] == 0 and out[1] in (1, 2)


def test_solver_against_scipy(scenes):
    """Potentials of the oracle's LDL^T on a real scene graph vs scipy's sparse LU in float64."""
    sp = pytest.importorskip("scipy.sparse")
    spl = pytest.importorskip("scipy.sparse.linalg")
    sc = scenes["test13"]
    g = orc.Graph.from_cloud(sc["xyz"], sc["rgba"], 0.01)
    g.normals(); g.signed_curvatures(); g.edge_weights()
    s, t, w = g.edges()
    V = g.V
    rng = np.random.default_rng(3)
    colours = np.zeros(V, np.uint32)
    seeds = rng.choice(V, 30, replace=False)
    colours[seeds] = 1 + (np.arange(30) % 3)
    st, out, pot, res = g.random_walker(colours, 3, solver=orc.SOLVE_LDLT_F64, want_potentials=True)
    assert st == 0 and res.max() < 1e-5   # potentials are returned rounded to float32
    W = sp.coo_matrix((np.r_[w, w].astype(np.float64), (np.r_[s, t], np.r_[t, s])), shape=(V, V)).tocsr()
    # the reference's diagonal is the float32 degree accumulated in edge order (impl/random_walker.hpp:103-114)
    deg = np.zeros(V, np.float32)
    for e in range(len(w)):
        deg[s[e]] += w[e]
        deg[t[e]] += w[e]
    deg = deg.astype(np.float64)
    L = sp.diags(deg) - W
    nc, comp = g.components()
    seeded = np.zeros(nc, bool)
    seeded[comp[seeds]] = True
    free = np.nonzero((colours == 0) & (deg >= np.finfo(np.float32).eps) & seeded[comp])[0]
    Lff = L[free][:, free].tocsc()
    for k in range(3):
        b = W[free][:, np.nonzero(colours == k + 1)[0]].sum(axis=1)
        x = spl.spsolve(Lff, np.asarray(b).ravel())
        np.testing.assert_allclose(pot[free, k], x, atol=2e-4)


def test_voxel_builder_hand_case():
    """Five points in three voxels (x keys 0, 1, 3 after the octree recentres its box by half the oversize):
    centroid = float mean in index order, rgba = truncated channel means, vertex order = descending Morton order
    (PCL <= 1.8 leaf iterator), edges only between 26-adjacent voxels, from the later to the earlier vertex."""
    z = 1.0
    xyz = np.array([[0.000, 0.0, z], [0.001, 0.0, z], [0.0200, 0.0, z], [0.0195, 0.0, z], [0.005, 0.0, z],
                    [np.nan, 0, 1]], np.float32)
    rgba = np.array([0xFF102030, 0xFF112233, 0xFF000000, 0xFF000003, 0xFF0A0B0C, 0], np.uint32)
    g = orc.Graph.from_cloud(xyz, rgba, 0.006)
    assert g.status == 0 and g.depth == 2 and g.V == 3
    p2v = g.point_to_vertex()
    assert list(p2v) == [2, 2, 0, 0, 1, 0xFFFFFFFF]
    assert [list(k) for k in g.keys()] == [[3, 1, 1], [1, 1, 1], [0, 1, 1]]
    vx, vc, _, _ = g.vertices()
    np.testing.assert_array_equal(vx[2], (xyz[0] + xyz[1]) / np.float32(2))
    assert vc[2] == 0xFF102131 and vc[0] == 0xFF000001 and vc[1] == 0xFF0A0B0C
    s, t, _ = g.edges()
    assert (list(s), list(t)) == ([2], [1])
    rp, col, eid = g.csr()
    assert list(rp) == [0, 0, 1, 2] and list(col) == [2, 1] and list(eid) == [0, 0]


def test_normals_of_a_plane():
    """Vertices on the plane z = 1 + 0.2 x: normal = +-(0.2, 0, -1)/|.| oriented towards the origin, curvature ~ 0."""
    u = np.arange(-10, 11) * 0.006
    X, Y = np.meshgrid(u, u)
    xyz = np.stack([X.ravel(), Y.ravel(), 1.0 + 0.2 * X.ravel()], 1).astype(np.float32)
    n = len(u)
    idx = np.arange(n * n).reshape(n, n)
    src = np.r_[idx[:, 1:].ravel(), idx[1:, :].ravel()].astype(np.uint32)
    tgt = np.r_[idx[:, :-1].ravel(), idx[:-1, :].ravel()].astype(np.uint32)
    g = orc.Graph.from_edges(xyz, src, tgt)
    g.normals()
    _, _, n4, cv = g.vertices()
    expect = np.array([0.2, 0.0, -1.0]) / np.sqrt(1.04)
    inner = idx[3:-3, 3:-3].ravel()
    np.testing.assert_allclose(n4[inner, :3], np.tile(expect, (len(inner), 1)), atol=6e-3)  # float32 single-pass covariance
    assert np.all(cv[inner] < 2e-2) and np.all(cv >= 0) and np.median(cv[inner]) < 2e-3
    # second pass keeps the orientation of the first (common.hpp:87-93)
    g.normals()
    _, _, n4b, _ = g.vertices()
    assert np.all(np.sum(n4b[inner, :3] * n4[inner, :3], axis=1) > 0.99)


def test_edge_weight_formula():
    """One edge, no normalisation: w = exp(-|dp|^2 / s_xyz) * exp(-|dn|^2 / 2 / s_n) ; small weights are coerced."""
    xyz = np.array([[0, 0, 1], [0.01, 0, 1]], np.float32)
    n4 = np.array([[0, 0, -1, 0], [0.6, 0, -0.8, 0]], np.float32)
    g = orc.Graph.from_edges(xyz, [1], [0], normals4=n4, curvature=np.array([-0.1, 0.2], np.float32))
    terms = [(orc.TERM_XYZ, 0.001, 1.0, orc.NORM_NONE), (orc.TERM_NORMAL, 0.5, 0.0, orc.NORM_NONE)]
    g.edge_weights(terms, threshold=0.0, policy=orc.SMALL_IGNORE)
    w = g.edges()[2][0]
    dn2 = 0.6 ** 2 + 0.2 ** 2
    np.testing.assert_allclose(w, np.exp(-1e-4 / 1e-3) * np.exp(-0.5 * dn2 / 0.5), rtol=1e-6)
    g.edge_weights(terms, threshold=0.9, policy=orc.SMALL_COERCE)
    assert g.edges()[2][0] == np.float32(0.9)
    g.edge_weights(terms, threshold=0.9, policy=orc.SMALL_REMOVE)
    assert g.E == 0


def test_global_normalisation_divides_by_the_mean():
    xyz = np.array([[0, 0, 1], [0.01, 0, 1], [0.03, 0, 1]], np.float32)
    g = orc.Graph.from_edges(xyz, [1, 2], [0, 1])
    g.edge_weights([(orc.TERM_XYZ, 2.0, 1.0, orc.NORM_GLOBAL)], threshold=0.0, policy=orc.SMALL_IGNORE)
    d = np.array([1e-4, 4e-4])
    np.testing.assert_allclose(g.edges()[2], np.exp(-(d / d.mean()) / 2.0), rtol=1e-5)


def test_signed_curvature_convex_vs_concave():
    """A ridge seen from above is convex (positive), a valley concave (negative) - common.hpp:105-132."""
    def run(sign):
        xyz = np.array([[-0.01, 0, 1.0 + sign * 0.005], [0, 0, 1.0], [0.01, 0, 1.0 + sign * 0.005]], np.float32)
        nl = np.array([-0.447 * sign, 0, -0.894, 0], np.float32)
        nr = np.array([0.447 * sign, 0, -0.894, 0], np.float32)
        n4 = np.stack([nl, [0, 0, -1, 0], nr]).astype(np.float32)
        g = orc.Graph.from_edges(xyz, [1, 2], [0, 1], normals4=n4, curvature=np.full(3, 0.05, np.float32))
        g.signed_curvatures()
        return g.vertices()[3]
    assert np.all(run(+1) > 0) and np.all(run(-1) < 0)


def test_segmentation_error_is_assignment_based():
    gt = np.array([0, 0, 0, 1, 1, 2, 2, 2], np.uint32)
    seg = np.array([5, 5, 5, 7, 7, 9, 9, 7], np.uint32)
    assert orc.segmentation_error(gt, seg) == 1
    assert orc.segmentation_error(gt, gt + 10) == 0


def test_scene_sizes_match_the_survey(scenes):
    """Integer facts of the inputs measured in the survey."""
    expect = {"test1": (41546, 252811), "test13": (33275, 203032), "test47": (39766, 232411),
              "test56": (38359, 224896), "test60": (36132, 211459)}
    for name, (V, E) in expect.items():
        g = orc.Graph.from_cloud(scenes[name]["xyz"], scenes[name]["rgba"], 0.006)
        assert (g.V, g.E, g.depth) == (V, E, 8), name
    g = orc.Graph.from_cloud(scenes["test1"]["xyz"], scenes["test1"]["rgba"], 0.01)
    assert (g.V, g.E) == (16235, 104525)


def test_nn_graph_known_answers():
    """nearest_neighbors_graph_builder.hpp:49-119 on inputs whose answer is known by hand."""
    nan = np.nan
    # points on a line at 0, 1, 3, 7 (+ one non-finite point that gets no vertex)
    xyz = np.array([[0, 0, 0], [nan, 0, 0], [1, 0, 0], [3, 0, 0], [7, 0, 0]], np.float32)
    g = orc.Graph.from_cloud_nn(xyz, None, 1)
    assert g.V == 4 and list(g.point_to_vertex()) == [0, 0xFFFFFFFF, 1, 2, 3]
    s, t, _ = g.edges()
    # v0 -> v1; v1 -> v0 (exists); v2 -> v1; v3 -> v2
    assert list(zip(s, t)) == [(0, 1), (2, 1), (3, 2)]
    g = orc.Graph.from_cloud_nn(xyz, None, 2)
    s, t, _ = g.edges()
    # v0: v1, v2; v1: v0 (exists), v2; v2: v1 (exists), v0 (exists); v3: v2, v1
    assert list(zip(s, t)) == [(0, 1), (0, 2), (1, 2), (3, 2), (3, 1)]
    # radius: strictly closer than the radius, at most num_neighbors of them
    g = orc.Graph.from_cloud_nn(xyz, None, 10, 2.0, use_radius=True)
    s, t, _ = g.edges()
    assert list(zip(s, t)) == [(0, 1)]          # |v1 - v2| = 2 is not < 2
    g = orc.Graph.from_cloud_nn(xyz, None, 1, 2.5, use_radius=True)
    s, t, _ = g.edges()
    assert list(zip(s, t)) == [(0, 1), (2, 1)]
    # equidistant candidates: lower index first
    sq = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [-1, 0, 0], [0, -1, 0]], np.float32)
    g = orc.Graph.from_cloud_nn(sq, None, 2)
    rp, col, _ = g.csr()
    assert list(col[rp[0]:rp[1]][:2]) == [1, 2]
    # more neighbours requested than points: complete graph
    g = orc.Graph.from_cloud_nn(sq, None, 10)
    assert g.E == 10
    assert orc.Graph.from_cloud_nn(np.zeros((0, 3), np.float32), None, 10).V == 0
