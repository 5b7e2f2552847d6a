"""ctypes binding of the C ABI (include/tcs_b200.h) of tcs_b200/libtcs_b200.so.

This is the thin Python view of the product library used by tests/ and bench.py.  There is no CPU fallback: if the
shared library is missing or no CUDA device is present, loading / context creation raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtcs_b200.so")

TERM_XYZ, TERM_NORMAL, TERM_CURVATURE, TERM_RGB = 0, 1, 2, 3
NORM_NONE, NORM_GLOBAL, NORM_LOCAL = 0, 1, 2
SMALL_IGNORE, SMALL_COERCE, SMALL_REMOVE = 0, 1, 2
MODE_PARITY, MODE_FAST = 0, 1
PRESET_CLASS, PRESET_CLI = 0, 1
BALANCE_GAUSSIAN, BALANCE_INVERSE, BALANCE_GAUSSIAN_SQUARED, BALANCE_LINEAR = 0, 1, 2, 3
OK, ERR_CUDA, ERR_INPUT, ERR_DEPTH, ERR_NOT_CONVERGED, ERR_STATE = 0, -1, -2, -3, -4, -5
MAX_LABELS = 64
NOALLOC_MAX_SEEDS = 1024
NIL = 0xFFFFFFFF

# weight presets
TERMS_CLASS = [(TERM_XYZ, 3.0, 1.0, NORM_GLOBAL), (TERM_NORMAL, 0.01, 0.0, NORM_NONE),
               (TERM_CURVATURE, 0.0001, 0.0, NORM_NONE), (TERM_RGB, 3.0, 1.0, NORM_GLOBAL)]
TERMS_CLI = [(TERM_XYZ, 3.0, 1.0, NORM_GLOBAL), (TERM_NORMAL, 0.01, 0.01, NORM_NONE),
             (TERM_CURVATURE, 0.0001, 0.01, NORM_NONE), (TERM_RGB, 3.0, 1.0, NORM_LOCAL)]

EXPORTS = [
    "tcs_ctx_create", "tcs_ctx_create_on_stream", "tcs_ctx_destroy", "tcs_last_error", "tcs_ctx_synchronize",
    "tcs_ctx_stream", "tcs_ctx_launch_count", "tcs_version", "tcs_cloud_upload", "tcs_cloud_set_device",
    "tcs_build_voxel_graph", "tcs_graph_create", "tcs_graph_destroy", "tcs_graph_sizes", "tcs_graph_bbox_min",
    "tcs_graph_download_vertices", "tcs_graph_download_keys", "tcs_graph_download_edges", "tcs_graph_download_csr",
    "tcs_graph_download_point_to_vertex", "tcs_graph_upload_xyz", "tcs_graph_upload_normals",
    "tcs_graph_upload_weights", "tcs_normals", "tcs_smoothen", "tcs_signed_curvatures", "tcs_edge_weights",
    "tcs_components", "tcs_solver_opts_default", "tcs_nearest_vertices", "tcs_random_walker", "tcs_segment_cloud",
    "tcs_graph_download_potentials", "tcs_graph_download_colours", "tcs_graph_d_colours",
    "tcs_graph_d_point_cluster", "tcs_bench_spmm", "tcs_check_spmm", "tcs_graph_download_best",
    "tcs_graph_d_best_potential", "tcs_host_alloc", "tcs_host_free", "tcs_build_nn_graph", "tcs_segment_batch", "tcs_nearest_vertices_noalloc", "tcs_edge_weights_ex",
]


class Term(C.Structure):
    _fields_ = [("kind", C.c_int), ("influence", C.c_float), ("convex_mult", C.c_float), ("normalization", C.c_int)]


class SolverOpts(C.Structure):
    _fields_ = [("tolerance", C.c_double), ("max_iterations", C.c_int), ("check_every", C.c_int),
                ("store_potentials", C.c_int), ("time_spmm", C.c_int), ("column_begin", C.c_int), ("column_count", C.c_int)]


class SolveInfo(C.Structure):
    _fields_ = [("iterations", C.c_int), ("converged", C.c_int), ("rows", C.c_int), ("nnz", C.c_longlong),
                ("ncolours", C.c_int), ("ncomponents", C.c_int), ("near_ties", C.c_longlong),
                ("rel_residual", C.c_double * MAX_LABELS), ("spmm_ms", C.c_float), ("spmm_launches", C.c_int),
                ("solve_ms", C.c_float), ("stage_ms", C.c_float * 12), ("spmm_kernel", C.c_int)]

    def as_dict(self):
        K = self.ncolours
        return dict(iterations=self.iterations, converged=bool(self.converged), rows=self.rows, nnz=self.nnz,
                    ncolours=K, ncomponents=self.ncomponents, near_ties=self.near_ties,
                    rel_residual=[self.rel_residual[k] for k in range(K)], spmm_ms=self.spmm_ms,
                    spmm_launches=self.spmm_launches, solve_ms=self.solve_ms, stage_ms=list(self.stage_ms),
                    spmm_kernel=self.spmm_kernel)


class Scene(C.Structure):
    """tcs_scene: one cloud + seeds of a batch (tcs_segment_batch)."""
    _fields_ = [("xyz", C.c_void_p), ("xyz_stride", C.c_size_t), ("rgba", C.c_void_p), ("rgba_stride", C.c_size_t),
                ("n", C.c_size_t), ("d_xyzc", C.c_void_p), ("d_has_color", C.c_int), ("seed_xyz", C.c_void_p), ("seed_label", C.c_void_p), ("ns", C.c_size_t),
                ("point_cluster", C.c_void_p), ("vertex_colour", C.c_void_p), ("V", C.c_size_t), ("E", C.c_size_t),
                ("ncolours", C.c_int), ("status", C.c_int)]


class TcsError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("tcs_b200 error %d: %s" % (code, msg))
        self.code = code


_lib = None


def lib():
    """Load libtcs_b200.so (raises if it has not been built: there is no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("%s not found: build it with `make -C tcs_b200/csrc` (no CPU fallback exists)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, sz, ci, cf = C.c_void_p, C.c_size_t, C.c_int, C.c_float
    PP = C.POINTER(vp)
    L.tcs_ctx_create.argtypes = [ci, PP]
    L.tcs_ctx_create_on_stream.argtypes = [ci, vp, PP]
    L.tcs_ctx_destroy.argtypes = [vp]
    L.tcs_ctx_destroy.restype = None
    L.tcs_last_error.argtypes = [vp]
    L.tcs_last_error.restype = C.c_char_p
    L.tcs_ctx_synchronize.argtypes = [vp]
    L.tcs_ctx_stream.argtypes = [vp]
    L.tcs_ctx_stream.restype = vp
    L.tcs_ctx_launch_count.argtypes = [vp]
    L.tcs_ctx_launch_count.restype = C.c_longlong
    L.tcs_version.restype = C.c_char_p
    L.tcs_cloud_upload.argtypes = [vp, vp, sz, vp, sz, sz]
    L.tcs_cloud_set_device.argtypes = [vp, vp, sz, ci]
    L.tcs_build_voxel_graph.argtypes = [vp, cf, ci, PP]
    L.tcs_build_nn_graph.argtypes = [vp, ci, C.c_double, ci, PP]
    L.tcs_graph_create.argtypes = [vp, sz, vp, vp, vp, vp, sz, vp, vp, vp, PP]
    L.tcs_graph_destroy.argtypes = [vp]
    L.tcs_graph_destroy.restype = None
    L.tcs_graph_sizes.argtypes = [vp, C.POINTER(sz), C.POINTER(sz), C.POINTER(sz), C.POINTER(ci)]
    L.tcs_graph_bbox_min.argtypes = [vp, vp]
    L.tcs_graph_download_vertices.argtypes = [vp, vp, vp, vp, vp]
    L.tcs_graph_download_keys.argtypes = [vp, vp]
    L.tcs_graph_download_edges.argtypes = [vp, vp, vp, vp]
    L.tcs_graph_download_csr.argtypes = [vp, vp, vp, vp]
    L.tcs_graph_download_point_to_vertex.argtypes = [vp, vp]
    L.tcs_graph_upload_xyz.argtypes = [vp, vp]
    L.tcs_graph_upload_normals.argtypes = [vp, vp, vp]
    L.tcs_graph_upload_weights.argtypes = [vp, vp]
    L.tcs_normals.argtypes = [vp, ci, ci]
    L.tcs_smoothen.argtypes = [vp, cf, cf]
    L.tcs_signed_curvatures.argtypes = [vp]
    L.tcs_edge_weights.argtypes = [vp, vp, ci, cf, ci, ci]
    L.tcs_edge_weights_ex.argtypes = [vp, vp, ci, cf, ci, ci, ci]
    L.tcs_components.argtypes = [vp, C.POINTER(ci), vp]
    L.tcs_solver_opts_default.argtypes = [C.POINTER(SolverOpts)]
    L.tcs_solver_opts_default.restype = None
    L.tcs_nearest_vertices.argtypes = [vp, vp, sz, vp]
    L.tcs_nearest_vertices_noalloc.argtypes = [vp, vp, sz, vp]
    L.tcs_random_walker.argtypes = [vp, vp, ci, C.POINTER(SolverOpts), vp, C.POINTER(SolveInfo)]
    L.tcs_segment_cloud.argtypes = [vp, cf, ci, ci, vp, vp, sz, C.POINTER(SolverOpts), vp, vp, C.POINTER(SolveInfo), PP]
    L.tcs_segment_batch.argtypes = [vp, C.POINTER(Scene), ci, cf, ci, ci, C.POINTER(SolverOpts), C.POINTER(SolveInfo)]
    L.tcs_graph_download_potentials.argtypes = [vp, vp]
    L.tcs_graph_download_colours.argtypes = [vp, vp]
    L.tcs_graph_d_colours.argtypes = [vp]
    L.tcs_graph_d_colours.restype = vp
    L.tcs_graph_d_point_cluster.argtypes = [vp]
    L.tcs_graph_d_point_cluster.restype = vp
    L.tcs_graph_d_best_potential.argtypes = [vp]
    L.tcs_graph_d_best_potential.restype = vp
    L.tcs_bench_spmm.argtypes = [vp, ci, C.POINTER(cf), C.POINTER(C.c_longlong)]
    L.tcs_graph_download_best.argtypes = [vp, vp, vp]
    L.tcs_check_spmm.argtypes = [vp, C.c_uint, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                 C.POINTER(ci)]
    _lib = L
    return L


class _DeviceSpan:
    """a borrowed device buffer for torch.as_tensor (CUDA array interface v2)"""

    def __init__(self, ptr, count, typestr):
        if not ptr:
            raise TcsError(ERR_STATE, "no device buffer (no solve on this graph yet)")
        self.__cuda_array_interface__ = dict(shape=(int(count),), typestr=typestr, data=(int(ptr), False), version=2,
                                             strides=None)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _u32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.uint32)


def default_opts(**kw) -> SolverOpts:
    o = SolverOpts()
    lib().tcs_solver_opts_default(C.byref(o))
    for k, v in kw.items():
        setattr(o, k, v)
    return o


class Context:
    """One device + one stream (tcs_ctx)."""

    def __init__(self, device: int = 0, stream: int | None = None):
        h = C.c_void_p()
        st = lib().tcs_ctx_create_on_stream(device, C.c_void_p(stream) if stream else None, C.byref(h))
        if st != OK:
            raise TcsError(st, (lib().tcs_last_error(None) or b"").decode())
        self.h, self.device = h, int(device)

    def close(self):
        if getattr(self, "h", None):
            lib().tcs_ctx_destroy(self.h)
            self.h = None

    __del__ = close

    def check(self, st, allow=()):
        if st != OK and st not in allow:
            raise TcsError(st, (lib().tcs_last_error(self.h) or b"").decode())
        return st

    def synchronize(self):
        self.check(lib().tcs_ctx_synchronize(self.h))

    @property
    def launches(self) -> int:
        return lib().tcs_ctx_launch_count(self.h)

    @property
    def stream(self) -> int:
        return lib().tcs_ctx_stream(self.h) or 0

    # ---- input cloud
    def upload_cloud(self, xyz, rgba=None):
        xyz, rgba = _f32(xyz), _u32(rgba)
        self._keep = (xyz, rgba)
        self.n_points_hint = len(xyz)
        self.check(lib().tcs_cloud_upload(self.h, _p(xyz), 12, _p(rgba), 4, len(xyz)))
        self.synchronize()

    def upload_cloud_aos(self, records: np.ndarray, xyz_offset: int, rgba_offset: int | None):
        """records: (n, stride) uint8 array of interleaved point structs (e.g. 32-byte pcl::PointXYZRGB)."""
        records = np.ascontiguousarray(records)
        n, stride = records.shape
        base = records.ctypes.data
        self._keep = records
        self.n_points_hint = n
        self.check(lib().tcs_cloud_upload(self.h, C.c_void_p(base + xyz_offset), stride,
                                          C.c_void_p(base + rgba_offset) if rgba_offset is not None else None, stride, n))
        self.synchronize()

    def set_device_cloud(self, device_ptr: int, n: int, has_color: bool = True):
        self.check(lib().tcs_cloud_set_device(self.h, C.c_void_p(device_ptr), n, int(has_color)))
        self.n_points_hint = n

    # ---- graph construction
    def build_voxel_graph(self, resolution: float, leaf_order_desc: int = 1) -> "Graph":
        h = C.c_void_p()
        self.check(lib().tcs_build_voxel_graph(self.h, resolution, leaf_order_desc, C.byref(h)))
        return Graph(self, h)

    def build_nn_graph(self, num_neighbors: int = 10, radius: float = 0.0, use_radius: bool = False) -> "Graph":
        """NearestNeighborsGraphBuilder (kNN / radius) on the uploaded cloud."""
        h = C.c_void_p()
        self.check(lib().tcs_build_nn_graph(self.h, num_neighbors, radius, int(use_radius), C.byref(h)))
        return Graph(self, h)

    def graph_from_edges(self, xyz, src, tgt, weight=None, rgba=None, normals4=None, curvature=None) -> "Graph":
        xyz, rgba, normals4, curvature = _f32(xyz), _u32(rgba), _f32(normals4), _f32(curvature)
        src, tgt, weight = _u32(src), _u32(tgt), _f32(weight)
        h = C.c_void_p()
        self.check(lib().tcs_graph_create(self.h, len(xyz), _p(xyz), _p(rgba), _p(normals4), _p(curvature), len(src),
                                          _p(src), _p(tgt), _p(weight), C.byref(h)))
        return Graph(self, h)

    def segment_cloud(self, resolution, seed_xyz, seed_label, preset=PRESET_CLASS, mode=MODE_PARITY, opts=None,
                      want_host=True):
        """Whole class path on the uploaded cloud.  Returns dict(graph, point_cluster, vertex_colour, info, status)."""
        seed_xyz, seed_label = _f32(seed_xyz), _u32(seed_label)
        opts = opts or default_opts()
        info = SolveInfo()
        h = C.c_void_p()
        n = self.n_points_hint
        pc = np.zeros(n, np.int32) if want_host else None
        vc = np.zeros(n, np.uint32) if want_host else None
        st = self.check(lib().tcs_segment_cloud(self.h, resolution, preset, mode, _p(seed_xyz), _p(seed_label),
                                                len(seed_label), C.byref(opts), _p(pc), _p(vc), C.byref(info),
                                                C.byref(h)), allow=(ERR_NOT_CONVERGED,))
        g = Graph(self, h)
        return dict(graph=g, point_cluster=pc, vertex_colour=None if vc is None else vc[:g.V].copy(), info=info,
                    status=st)

    def segment_batch(self, scenes, resolution, preset=PRESET_CLASS, mode=MODE_PARITY, opts=None):
        """Batched class path (tcs_segment_batch).  scenes: list of (xyz (n,3) f32, rgba (n,) u32 or None, seed_xyz,
        seed_label).  Returns (list of dict(point_cluster, vertex_colour, V, E, K), SolveInfo of the joint solve, status)."""
        arr = (Scene * len(scenes))()
        keep = []
        for i, (xyz, rgba, sx, sl) in enumerate(scenes):
            xyz, rgba, sx, sl = _f32(xyz), _u32(rgba), _f32(sx), _u32(sl)
            pc, vc = np.zeros(len(xyz), np.int32), np.zeros(len(xyz), np.uint32)
            keep.append((xyz, rgba, sx, sl, pc, vc))
            a = arr[i]
            a.xyz, a.xyz_stride, a.n = xyz.ctypes.data, 12, len(xyz)
            a.rgba, a.rgba_stride = (rgba.ctypes.data if rgba is not None else None), 4
            a.seed_xyz, a.seed_label, a.ns = sx.ctypes.data, sl.ctypes.data, len(sl)
            a.point_cluster, a.vertex_colour = pc.ctypes.data, vc.ctypes.data
        opts = opts or default_opts()
        info = SolveInfo()
        st = self.check(lib().tcs_segment_batch(self.h, arr, len(scenes), resolution, preset, mode, C.byref(opts),
                                                C.byref(info)), allow=(ERR_NOT_CONVERGED,))
        out = [dict(point_cluster=k[4], vertex_colour=k[5][:arr[i].V].copy(), V=arr[i].V, E=arr[i].E, K=arr[i].ncolours)
               for i, k in enumerate(keep)]
        return out, info, st

    n_points_hint = 0


class Graph:
    """Owning handle of a tcs_graph (device-resident point-cloud graph)."""

    def __init__(self, ctx: Context, handle):
        self.ctx, self.h = ctx, handle

    def close(self):
        if getattr(self, "h", None) and getattr(self.ctx, "h", None):
            lib().tcs_graph_destroy(self.h)
        self.h = None

    __del__ = close

    def _sizes(self):
        V, E, N, d = C.c_size_t(), C.c_size_t(), C.c_size_t(), C.c_int()
        lib().tcs_graph_sizes(self.h, C.byref(V), C.byref(E), C.byref(N), C.byref(d))
        return V.value, E.value, N.value, d.value

    V = property(lambda s: s._sizes()[0])
    E = property(lambda s: s._sizes()[1])
    N = property(lambda s: s._sizes()[2])
    depth = property(lambda s: s._sizes()[3])

    def bbox_min(self):
        m = np.zeros(3, np.float64)
        lib().tcs_graph_bbox_min(self.h, _p(m))
        return m

    def vertices(self):
        V = self.V
        xyz, rgba = np.zeros((V, 3), np.float32), np.zeros(V, np.uint32)
        n4, cv = np.zeros((V, 4), np.float32), np.zeros(V, np.float32)
        self.ctx.check(lib().tcs_graph_download_vertices(self.h, _p(xyz), _p(rgba), _p(n4), _p(cv)))
        return xyz, rgba, n4, cv

    def keys(self):
        k = np.zeros((self.V, 3), np.uint32)
        self.ctx.check(lib().tcs_graph_download_keys(self.h, _p(k)))
        return k

    def edges(self):
        E = self.E
        s, t, w = np.zeros(E, np.uint32), np.zeros(E, np.uint32), np.zeros(E, np.float32)
        self.ctx.check(lib().tcs_graph_download_edges(self.h, _p(s), _p(t), _p(w)))
        return s, t, w

    def csr(self):
        V, E = self.V, self.E
        rp, col, eid = np.zeros(V + 1, np.uint32), np.zeros(2 * E, np.uint32), np.zeros(2 * E, np.uint32)
        self.ctx.check(lib().tcs_graph_download_csr(self.h, _p(rp), _p(col), _p(eid)))
        return rp, col, eid

    def point_to_vertex(self):
        m = np.zeros(self.N, np.uint32)
        self.ctx.check(lib().tcs_graph_download_point_to_vertex(self.h, _p(m)))
        return m

    def set_xyz(self, xyz):
        xyz = _f32(xyz)
        self.ctx.check(lib().tcs_graph_upload_xyz(self.h, _p(xyz)))

    def set_normals(self, n4, curvature=None):
        n4, curvature = _f32(n4), _f32(curvature)
        self.ctx.check(lib().tcs_graph_upload_normals(self.h, _p(n4), _p(curvature)))

    def set_weights(self, w):
        w = _f32(w)
        self.ctx.check(lib().tcs_graph_upload_weights(self.h, _p(w)))

    # ---- stages
    def normals(self, one_ring=False, mode=MODE_PARITY):
        self.ctx.check(lib().tcs_normals(self.h, int(one_ring), mode))

    def smoothen(self, spatial_sigma=0.012, influence_sigma=0.0012):
        self.ctx.check(lib().tcs_smoothen(self.h, spatial_sigma, influence_sigma))

    def signed_curvatures(self):
        self.ctx.check(lib().tcs_signed_curvatures(self.h))

    def edge_weights(self, terms=TERMS_CLASS, threshold=1e-5, policy=SMALL_COERCE, mode=MODE_PARITY, balancing=BALANCE_GAUSSIAN):
        arr = (Term * len(terms))(*[Term(*t) for t in terms])
        self.ctx.check(lib().tcs_edge_weights_ex(self.h, arr, len(terms), threshold, policy, mode, balancing))

    def components(self):
        comp = np.zeros(self.V, np.int32)
        nc = C.c_int()
        self.ctx.check(lib().tcs_components(self.h, C.byref(nc), _p(comp)))
        return nc.value, comp

    def nearest_vertices(self, seed_xyz):
        seed_xyz = _f32(seed_xyz)
        out = np.zeros(len(seed_xyz), np.int32)
        self.ctx.check(lib().tcs_nearest_vertices(self.h, _p(seed_xyz), len(seed_xyz), _p(out)))
        return out

    def nearest_vertices_noalloc(self, seed_xyz):
        """tcs_nearest_vertices_noalloc: same result from the fixed device scratch (at most NOALLOC_MAX_SEEDS seeds)."""
        seed_xyz = _f32(seed_xyz)
        out = np.zeros(len(seed_xyz), np.int32)
        self.ctx.check(lib().tcs_nearest_vertices_noalloc(self.h, _p(seed_xyz), len(seed_xyz), _p(out)))
        return out

    def random_walker(self, colours, ncolours, opts=None, want_potentials=False):
        """Returns (status, colours, potentials (V, K) or None, SolveInfo)."""
        colours = np.array(colours, dtype=np.uint32)
        pot = np.zeros((ncolours, self.V), np.float32) if want_potentials else None  # column-major V x K
        info = SolveInfo()
        opts = opts or default_opts()
        st = self.ctx.check(lib().tcs_random_walker(self.h, _p(colours), ncolours, C.byref(opts), _p(pot),
                                                    C.byref(info)), allow=(ERR_NOT_CONVERGED,))
        return st, colours, (pot.T if pot is not None else None), info

    def colours(self):
        c = np.zeros(self.V, np.uint32)
        self.ctx.check(lib().tcs_graph_download_colours(self.h, _p(c)))
        return c

    def best(self):
        """(best potential, its colour) per vertex among the columns this solve covered."""
        val, col = np.zeros(self.V, np.float32), np.zeros(self.V, np.uint32)
        self.ctx.check(lib().tcs_graph_download_best(self.h, _p(val), _p(col)))
        return val, col

    def best_device(self):
        """(best potential f32, its colour as int32) as torch CUDA tensors aliasing the graph's device buffers - no copy,
        no host staging; valid while the graph lives.  Synchronizes the context stream first (torch reads on its own)."""
        import torch
        self.ctx.synchronize()
        V, dev = self.V, self.ctx.device
        val = torch.as_tensor(_DeviceSpan(lib().tcs_graph_d_best_potential(self.h), V, "<f4"), device="cuda:%d" % dev)
        col = torch.as_tensor(_DeviceSpan(lib().tcs_graph_d_colours(self.h), V, "<i4"), device="cuda:%d" % dev)
        return val, col

    def check_spmm(self, seed=1):
        d, r, dd, k = C.c_double(), C.c_double(), C.c_double(), C.c_int()
        self.ctx.check(lib().tcs_check_spmm(self.h, seed, C.byref(d), C.byref(r), C.byref(dd), C.byref(k)))
        return d.value, r.value, dd.value, k.value

    def bench_spmm(self, reps=20):
        ms, nbytes = C.c_float(), C.c_longlong()
        self.ctx.check(lib().tcs_bench_spmm(self.h, reps, C.byref(ms), C.byref(nbytes)))
        return ms.value, nbytes.value
