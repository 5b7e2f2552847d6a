"""The C++ drop-in API (include/tcs/*.h: pcl::segmentation::RandomWalkerSegmentation, pcl::graph::*), driven through
tests/cpp/api_check.cpp the way a caller of the reference would use it, checked against the CPU oracle."""
import os
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tcs_b200", "cpp", "api_check")


def _build():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "tcs_b200", "cpp"), "-s"])


def _write_input(path, xyz, rgba, sx, sl):
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(xyz)))
        f.write(np.ascontiguousarray(xyz, np.float32).tobytes())
        f.write(np.ascontiguousarray(rgba, np.uint32).tobytes())
        f.write(struct.pack("<Q", len(sl)))
        f.write(np.ascontiguousarray(sx, np.float32).tobytes())
        f.write(np.ascontiguousarray(sl, np.uint32).tobytes())


def _read_output(path, m):
    raw = open(path, "rb").read()
    cl = np.frombuffer(raw, np.int32, m)
    off = 4 * m
    V = struct.unpack_from("<Q", raw, off)[0]; off += 8
    col = np.frombuffer(raw, np.uint32, V, off); off += 4 * V
    K = struct.unpack_from("<Q", raw, off)[0]; off += 8
    pot = np.frombuffer(raw, np.float32, V * K, off).reshape(K, V).T
    return cl, col, pot


def test_cpp_api_compiles_and_fails_loudly_without_gpu(tmp_path):
    """Header-only API builds with plain g++ against the C ABI; with no CUDA device it throws, it does not fall back."""
    import torch
    _build()
    assert os.path.exists(EXE)
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    inp, out = str(tmp_path / "i.bin"), str(tmp_path / "o.bin")
    _write_input(inp, np.array([[0, 0, 1]], np.float32), np.array([0xFF000000], np.uint32),
                 np.array([[0, 0, 1]], np.float32), np.array([1], np.uint32))
    r = subprocess.run([EXE, "class", inp, out], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["class", "cli"])
@pytest.mark.parametrize("name", ["test1", "test47"])
def test_cpp_api_matches_oracle(scenes, tmp_path, mode, name):
    import orc
    from tcs_b200.seeds import seeds_from_labels
    _build()
    sc = scenes[name]
    sx, sl, _ = seeds_from_labels(sc["xyz"], sc["label"], sc["id"], 5)
    inp, out = str(tmp_path / "i.bin"), str(tmp_path / "o.bin")
    _write_input(inp, sc["xyz"], sc["rgba"], sx, sl)
    r = subprocess.run([EXE, mode, inp, out], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    ref = orc.segment_cloud(sc["xyz"], sc["rgba"], 0.006, sx, sl, preset=0 if mode == "class" else 1,
                            solver=orc.SOLVE_LDLT_F64)
    V = ref["graph"].V
    m = len(sc["xyz"]) if mode == "class" else V
    cl, col, pot = _read_output(out, m)
    assert len(col) == V and pot.shape == (V, ref["K"])
    import label_gate
    bad = label_gate.check_labels("%s C++ class API (%s)" % (name, mode), col, ref, sx, sl)
    if mode == "class":
        assert np.count_nonzero(cl != ref["point_cluster"]) <= 10 * bad
    else:   # graph input: clusters index vertices (impl/random_walker_segmentation.hpp:213-221)
        expect = np.where(ref["vertex_colour"] == 0, ref["K"], ref["vertex_colour"].astype(np.int64) - 1)
        assert np.count_nonzero(cl != expect) <= bad
    # potentials: rows sum to 1 on labelled vertices, seeds are one-hot
    lab = col > 0
    # (residual 1e-6 on these ill-conditioned systems leaves errors of ~1e-2 in the potentials)
    np.testing.assert_allclose(pot[lab].sum(axis=1), 1.0, atol=5e-2)


GIO = os.path.join(ROOT, "tcs_b200", "cpp", "graph_io_check")


def test_graph_io_and_subgraph_helpers_on_the_host(tmp_path):
    """saveGraph / loadGraph (include/impl/io.hpp:13-101) and createSubgraphsFromColorMap / Indices
    (graph/impl/common.hpp:151-239) need no device for a host-built graph."""
    _build()
    r = subprocess.run([GIO, str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0 and "graph_io_check ok" in r.stdout, r.stdout + r.stderr
    text = open(tmp_path / "graph.pcd").read()
    head, tail = text.split("# Edges\n")
    assert head.startswith("# .PCD v0.7") and "FIELDS x y z rgba normal_x normal_y normal_z curvature" in head
    assert tail.splitlines()[:3] == ["0 1 1", "1 2 0.5", "2 3 1e-05"]
    # the PCD part is a regular PCD file for any other reader
    from tcs_b200 import pcd
    cols, hdr = pcd.read_pcd(str(tmp_path / "graph.pcd"))
    assert hdr["n"] == 7 and cols["rgba"][1] == 0xFF102030


@pytest.mark.gpu
def test_graph_io_round_trip_of_a_device_graph(tmp_path):
    _build()
    r = subprocess.run([GIO, str(tmp_path), "device"], capture_output=True, text=True)
    assert r.returncode == 0 and "graph_io_check ok" in r.stdout, r.stdout + r.stderr


REFCALL = os.path.join(ROOT, "tcs_b200", "cpp", "ref_caller_check")


def test_reference_caller_code_compiles():
    """tests/cpp/ref_caller_check.cpp holds the reference callers' idioms verbatim (the saveGraph edge loop over
    as_range (boost::edges (graph)), boost::tie, the 3- and 5-argument randomWalker with property maps, the three point
    types of src/random_walker_segmentation.cpp:44): it must compile against include/tcs with plain g++."""
    _build()
    assert os.path.exists(REFCALL)


@pytest.mark.gpu
def test_reference_caller_code_runs():
    _build()
    r = subprocess.run([REFCALL], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ref_caller_check ok" in r.stdout, r.stdout + r.stderr
    # sparse colour ids, external maps, > 64 colours failing cleanly, the NaN seed, PointXYZ / PointXYZRGBA
    assert "RandomWalkerSegmentation<PointXYZ>" in r.stdout and "NaN seed" in r.stdout
    assert "distinct seed colours" in r.stderr
