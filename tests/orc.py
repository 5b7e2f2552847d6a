"""ctypes binding of the CPU oracle (oracle/liborc.so).  TEST INFRASTRUCTURE: imported only by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs - never by tcs_b200/."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = os.path.join(_ROOT, "oracle", "liborc.so")

TERM_XYZ, TERM_NORMAL, TERM_CURVATURE, TERM_RGB = 0, 1, 2, 3
NORM_NONE, NORM_GLOBAL, NORM_LOCAL = 0, 1, 2
SMALL_IGNORE, SMALL_COERCE, SMALL_REMOVE = 0, 1, 2
MODE_PARITY, MODE_FAST = 0, 1
SOLVE_LDLT_F32, SOLVE_LDLT_F64, SOLVE_PCG_F64 = 0, 1, 2

# weight presets
PRESET_CLASS = [(TERM_XYZ, 3.0, 1.0, NORM_GLOBAL), (TERM_NORMAL, 0.01, 0.0, NORM_NONE),
                (TERM_CURVATURE, 0.0001, 0.0, NORM_NONE), (TERM_RGB, 3.0, 1.0, NORM_GLOBAL)]
PRESET_CLI = [(TERM_XYZ, 3.0, 1.0, NORM_GLOBAL), (TERM_NORMAL, 0.01, 0.01, NORM_NONE),
              (TERM_CURVATURE, 0.0001, 0.01, NORM_NONE), (TERM_RGB, 3.0, 1.0, NORM_LOCAL)]


class Term(C.Structure):
    _fields_ = [("kind", C.c_int), ("influence", C.c_float), ("convex_mult", C.c_float), ("normalization", C.c_int)]


def build_lib(force: bool = False) -> str:
    src_dir = os.path.join(_ROOT, "oracle")
    srcs = [os.path.join(src_dir, f) for f in os.listdir(src_dir) if f.endswith((".cpp", ".h", ".hpp"))]
    if force or not os.path.exists(_LIB) or any(os.path.getmtime(s) > os.path.getmtime(_LIB) for s in srcs):
        subprocess.check_call(["make", "-C", src_dir, "-s", "-B", "liborc.so"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build_lib())
        vp, sz = C.c_void_p, C.c_size_t
        L = _lib
        L.orc_build_voxel_graph.restype = vp
        L.orc_build_voxel_graph.argtypes = [vp, vp, sz, C.c_float, C.c_int, C.POINTER(C.c_int)]
        L.orc_build_nn_graph.restype = vp
        L.orc_build_nn_graph.argtypes = [vp, vp, sz, C.c_int, C.c_double, C.c_int, C.POINTER(C.c_int)]
        L.orc_graph_from_edges.restype = vp
        L.orc_graph_from_edges.argtypes = [sz, vp, vp, vp, vp, sz, vp, vp, vp]
        L.orc_graph_free.argtypes = [vp]
        for f in ("orc_num_vertices", "orc_num_edges", "orc_num_points"):
            getattr(L, f).restype = sz
            getattr(L, f).argtypes = [vp]
        L.orc_depth.argtypes = [vp]
        L.orc_get_bbox_min.argtypes = [vp, vp]
        L.orc_get_vertices.argtypes = [vp, vp, vp, vp, vp]
        L.orc_get_keys.argtypes = [vp, vp]
        L.orc_get_edges.argtypes = [vp, vp, vp, vp]
        L.orc_get_csr.argtypes = [vp, vp, vp, vp]
        L.orc_get_point_to_vertex.argtypes = [vp, vp]
        L.orc_set_xyz.argtypes = [vp, vp]
        L.orc_set_normals.argtypes = [vp, vp, vp]
        L.orc_set_weights.argtypes = [vp, vp]
        L.orc_normals.argtypes = [vp, C.c_int, C.c_int]
        L.orc_smoothen.argtypes = [vp, C.c_float, C.c_float]
        L.orc_signed_curvatures.argtypes = [vp]
        L.orc_edge_weights.argtypes = [vp, vp, C.c_int, C.c_float, C.c_int, C.c_int]
        L.orc_edge_weights_ex.argtypes = [vp, vp, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_components.argtypes = [vp, vp]
        L.orc_nearest_vertices.argtypes = [vp, vp, sz, vp]
        L.orc_random_walker.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp]
        L.orc_segment_cloud.restype = vp
        L.orc_segment_cloud.argtypes = [vp, vp, sz, C.c_float, C.c_int, C.c_int, vp, vp, sz, C.c_int, vp, vp,
                                        C.POINTER(C.c_int), vp, C.POINTER(C.c_int)]
        L.orc_segmentation_error.restype = C.c_long
        L.orc_segmentation_error.argtypes = [vp, vp, sz]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _u32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.uint32)


class Graph:
    """Owning handle of an orc_graph."""

    def __init__(self, handle):
        self.h = handle

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_graph_free(self.h)
            self.h = None

    # ---- construction
    @staticmethod
    def from_cloud(xyz, rgba, resolution, leaf_order_desc=1):
        xyz, rgba = _f32(xyz), _u32(rgba)
        st = C.c_int(0)
        h = lib().orc_build_voxel_graph(_p(xyz), _p(rgba), len(xyz), resolution, leaf_order_desc, C.byref(st))
        g = Graph(h)
        g.status = st.value
        return g

    @staticmethod
    def from_cloud_nn(xyz, rgba, num_neighbors=10, radius=0.0, use_radius=False):
        xyz, rgba = _f32(xyz), _u32(rgba)
        st = C.c_int(0)
        h = lib().orc_build_nn_graph(_p(xyz), _p(rgba), len(xyz), num_neighbors, radius, int(use_radius), C.byref(st))
        g = Graph(h)
        g.status = st.value
        return g

    @staticmethod
    def from_edges(xyz, src, tgt, weight=None, rgba=None, normals4=None, curvature=None):
        xyz, rgba, normals4, curvature = _f32(xyz), _u32(rgba), _f32(normals4), _f32(curvature)
        src, tgt, weight = _u32(src), _u32(tgt), _f32(weight)
        h = lib().orc_graph_from_edges(len(xyz), _p(xyz), _p(rgba), _p(normals4), _p(curvature), len(src), _p(src),
                                       _p(tgt), _p(weight))
        g = Graph(h)
        g.status = 0
        return g

    # ---- accessors
    @property
    def V(self):
        return lib().orc_num_vertices(self.h)

    @property
    def E(self):
        return lib().orc_num_edges(self.h)

    @property
    def depth(self):
        return lib().orc_depth(self.h)

    def bbox_min(self):
        m = np.zeros(3, np.float64)
        lib().orc_get_bbox_min(self.h, _p(m))
        return m

    def vertices(self):
        V = self.V
        xyz, rgba = np.zeros((V, 3), np.float32), np.zeros(V, np.uint32)
        n4, cv = np.zeros((V, 4), np.float32), np.zeros(V, np.float32)
        lib().orc_get_vertices(self.h, _p(xyz), _p(rgba), _p(n4), _p(cv))
        return xyz, rgba, n4, cv

    def keys(self):
        k = np.zeros((self.V, 3), np.uint32)
        lib().orc_get_keys(self.h, _p(k))
        return k

    def edges(self):
        E = self.E
        s, t, w = np.zeros(E, np.uint32), np.zeros(E, np.uint32), np.zeros(E, np.float32)
        lib().orc_get_edges(self.h, _p(s), _p(t), _p(w))
        return s, t, w

    def csr(self):
        rp, col, eid = np.zeros(self.V + 1, np.uint32), np.zeros(2 * self.E, np.uint32), np.zeros(2 * self.E, np.uint32)
        lib().orc_get_csr(self.h, _p(rp), _p(col), _p(eid))
        return rp, col, eid

    def point_to_vertex(self):
        m = np.zeros(lib().orc_num_points(self.h), np.uint32)
        lib().orc_get_point_to_vertex(self.h, _p(m))
        return m

    def set_xyz(self, xyz):
        xyz = _f32(xyz)
        lib().orc_set_xyz(self.h, _p(xyz))

    def set_normals(self, n4, curvature=None):
        n4, curvature = _f32(n4), _f32(curvature)
        lib().orc_set_normals(self.h, _p(n4), _p(curvature))

    def set_weights(self, w):
        w = _f32(w)
        lib().orc_set_weights(self.h, _p(w))

    # ---- stages
    def normals(self, one_ring=False, mode=MODE_PARITY):
        lib().orc_normals(self.h, int(one_ring), mode)

    def smoothen(self, spatial_sigma=0.012, influence_sigma=0.0012):
        lib().orc_smoothen(self.h, spatial_sigma, influence_sigma)

    def signed_curvatures(self):
        lib().orc_signed_curvatures(self.h)

    def edge_weights(self, terms=PRESET_CLASS, threshold=1e-5, policy=SMALL_COERCE, mode=MODE_PARITY, balancing=0):
        arr = (Term * len(terms))(*[Term(*t) for t in terms])
        return lib().orc_edge_weights_ex(self.h, arr, len(terms), threshold, policy, mode, balancing)

    def components(self):
        comp = np.zeros(self.V, np.int32)
        nc = lib().orc_components(self.h, _p(comp))
        return nc, comp

    def nearest_vertices(self, seed_xyz):
        seed_xyz = _f32(seed_xyz)
        out = np.zeros(len(seed_xyz), np.int32)
        lib().orc_nearest_vertices(self.h, _p(seed_xyz), len(seed_xyz), _p(out))
        return out

    def random_walker(self, colours, ncolours, solver=SOLVE_LDLT_F32, want_potentials=False):
        colours = np.array(colours, dtype=np.uint32)
        pot = np.zeros((ncolours, self.V), np.float32) if want_potentials else None  # column-major V x K
        res = np.zeros(ncolours, np.float64)
        st = lib().orc_random_walker(self.h, _p(colours), ncolours, solver, _p(pot), _p(res))
        return st, colours, (pot.T if pot is not None else None), res


def segment_cloud(xyz, rgba, resolution, seed_xyz, seed_label, preset=0, mode=MODE_PARITY, solver=SOLVE_LDLT_F32):
    """Whole class path.  Returns dict(graph, point_cluster, vertex_colour, K, stage_ms, status)."""
    xyz, rgba, seed_xyz, seed_label = _f32(xyz), _u32(rgba), _f32(seed_xyz), _u32(seed_label)
    n = len(xyz)
    pc = np.zeros(n, np.int32)
    vc = np.zeros(n, np.uint32)  # V <= n
    K, st = C.c_int(0), C.c_int(0)
    ms = np.zeros(10, np.float64)
    h = lib().orc_segment_cloud(_p(xyz), _p(rgba), n, resolution, preset, mode, _p(seed_xyz), _p(seed_label),
                                len(seed_label), solver, _p(pc), _p(vc), C.byref(K), _p(ms), C.byref(st))
    g = Graph(h)
    return dict(graph=g, point_cluster=pc, vertex_colour=vc[:g.V].copy(), K=K.value, stage_ms=ms, status=st.value)


def segmentation_error(gt, seg):
    gt, seg = _u32(gt), _u32(seg)
    return lib().orc_segmentation_error(_p(gt), _p(seg), len(gt))
