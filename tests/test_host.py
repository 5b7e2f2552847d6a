"""CPU tests of the host-side logic: C-ABI surface, PCD reader/writer, seed recipe, synthetic scenes, sharding."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_abi_exports_every_declared_symbol():
    """include/tcs_b200.h is the contract: every function it declares must be exported by libtcs_b200.so."""
    from tcs_b200 import capi
    hdr = open(os.path.join(ROOT, "include", "tcs_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = sorted(set(re.findall(r"\b(tcs_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 35
    lib = capi.lib()
    missing = [f for f in declared if not hasattr(lib, f)]
    assert not missing, missing
    assert sorted(capi.EXPORTS) == declared
    assert b"sm_100a" in lib.tcs_version()


def test_no_cpu_fallback_without_device():
    """Without a CUDA device context creation fails loudly (TCS_ERR_CUDA) - there is no CPU path behind the ABI."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from tcs_b200 import capi
    with pytest.raises(capi.TcsError) as e:
        capi.Context(0)
    assert e.value.code == capi.ERR_CUDA


def test_product_does_not_touch_the_oracle():
    """The oracle is test infrastructure: nothing under tcs_b200/ or include/ may reference it."""
    bad = []
    for base in ("tcs_b200", "include"):
        for d, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", "Makefile")):
                    txt = open(os.path.join(d, f), errors="ignore").read()
                    if re.search(r"liborc|import orc|orc_[a-z]+\(|oracle/", txt):
                        bad.append(os.path.join(d, f))
    assert not bad, bad


def test_pcd_roundtrip_all_encodings(tmp_path):
    from tcs_b200 import pcd
    rng = np.random.default_rng(0)
    n = 1000
    fields = {"x": rng.standard_normal(n).astype(np.float32), "y": rng.standard_normal(n).astype(np.float32),
              "z": rng.standard_normal(n).astype(np.float32), "rgba": rng.integers(0, 2 ** 32, n, dtype=np.uint32),
              "label": rng.integers(0, 5, n).astype(np.uint32)}
    fields["x"][::7] = np.nan
    for enc in ("ascii", "binary", "binary_compressed"):
        p = str(tmp_path / ("t_%s.pcd" % enc))
        pcd.write_pcd(p, fields, mode=enc)
        got, hdr = pcd.read_pcd(p)
        for k, v in fields.items():
            if enc == "ascii" and v.dtype == np.float32:
                np.testing.assert_allclose(got[k], v, rtol=1e-6, equal_nan=True)
            else:
                assert np.array_equal(got[k].view(np.uint32), v.view(np.uint32)), (enc, k)


def test_lzf_decoder_backreferences():
    from tcs_b200 import pcd
    # literal "abcabcabcabc" encoded with a back reference: 3 literals, then copy 9 bytes from offset 3
    stream = bytes([2]) + b"abc" + bytes([(7 << 5) | 0, 0, 2])
    assert pcd.lzf_decompress(stream, 12) == b"abcabcabcabc"


def test_seed_recipe_is_deterministic(scenes):
    from tcs_b200.seeds import SplitMix64, seeds_from_labels
    assert SplitMix64(1).next() == 0x910A2DEC89025CC1          # published splitmix64 test vector
    sc = scenes["test1"]
    sx, sl, si = seeds_from_labels(sc["xyz"], sc["label"], 0, 5)
    assert len(sl) == 15 and sorted(set(sl.tolist())) == [1, 2, 3]
    assert np.all(np.isfinite(sx)) and np.array_equal(sc["label"][si] + 1, sl)
    sx2, sl2, si2 = seeds_from_labels(sc["xyz"], sc["label"], 0, 5)
    assert np.array_equal(si, si2)


def test_synthetic_scene_shape_and_determinism():
    from tcs_b200 import synth
    xyz, rgba, label = synth.make_c5(3)
    assert xyz.shape == (640 * 480, 3) and xyz.dtype == np.float32 and rgba.dtype == np.uint32
    fin = np.isfinite(xyz).all(axis=1)
    assert 0.70 < fin.mean() < 0.85                           # ~22 % invalid pixels
    assert sorted(np.unique(label[fin]).tolist()) == list(range(8))   # 7 objects + background, all visible
    assert np.all(rgba >> 24 == 255)
    xyz2, rgba2, label2 = synth.make_c5(3)
    assert np.array_equal(xyz.view(np.uint32), xyz2.view(np.uint32)) and np.array_equal(rgba, rgba2)
    # organised pinhole geometry: x / z is the pixel ray
    z = xyz[fin, 2]
    u = (np.nonzero(fin)[0] % 640 - 320.0) / synth.FX
    np.testing.assert_allclose(xyz[fin, 0] / z, u, atol=1e-5)


def test_sharding_round_robin():
    from tcs_b200 import shard
    assert shard.scenes_for_rank(10, 1, 4) == [1, 5, 9]
    got = sorted(sum((shard.scenes_for_rank(1024, r, 8) for r in range(8)), []))
    assert got == list(range(1024))
    assert shard.throughput([10, 20], [1000.0, 2000.0]) == 15.0
    with pytest.raises(ValueError):
        shard.scenes_for_rank(4, 4, 4)


_WORKER = r"""
import os, sys
sys.path.insert(0, sys.argv[1])
import torch.distributed as dist
from tcs_b200 import shard
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % sys.argv[2], rank=int(sys.argv[3]), world_size=2)
rank = dist.get_rank()
mine = shard.scenes_for_rank(7, rank, 2)
units, ms = shard.combine(100.0 * len(mine), 10.0 * (rank + 1), 2)
assert units == 700.0 and ms == 20.0, (units, ms)
dist.barrier()
dist.destroy_process_group()
print("ok", rank, mine)
"""


def test_combine_world_size_2_gloo(tmp_path):
    """The N > 1 bookkeeping (sum of vertices, max of time) over torch.distributed, gloo, two processes."""
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    port = str(29500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "ok 0 [0, 2, 4, 6]" in outs[0] and "ok 1 [1, 3, 5]" in outs[1]


def test_column_shard_blocks_and_merge():
    from tcs_b200 import shard
    assert [shard.columns_for_rank(16, r, 8) for r in range(8)] == [(2 * r, 2) for r in range(8)]
    blocks = [shard.columns_for_rank(18, r, 4) for r in range(4)]
    assert blocks == [(0, 5), (5, 5), (10, 4), (14, 4)]
    assert shard.columns_for_rank(3, 3, 4) == (3, 0)
    # merge: largest potential wins, lowest colour on ties, unlabelled stays 0
    val = np.array([[0.6, 0.5, 0.0, 0.2], [0.3, 0.5, 0.0, 0.7]], np.float32)
    col = np.array([[1, 2, 0, 2], [3, 4, 0, 3]], np.uint32)
    assert list(shard.merge_best(val, col)) == [1, 2, 0, 3]
    assert list(shard.merge_best(val[::-1], col[::-1])) == [1, 2, 0, 3]
    # a vertex whose best potential is <= 0 (tiny negative PCG values) is still labelled when a shard labelled it, like on
    # one GPU (k_argmax labels every row with a non-zero potential); a shard that left it unlabelled does not compete
    val = np.array([[-1e-9, 0.0, -2e-9], [0.0, 0.0, -1e-9]], np.float32)
    col = np.array([[2, 0, 1], [0, 0, 3]], np.uint32)
    assert list(shard.merge_best(val, col)) == [2, 0, 3]


def test_all_gather_best_world_size_2_gloo(tmp_path):
    """The one exchange of the label-column shard, over gloo on CPU (NCCL on the GPU box)."""
    script = tmp_path / "w.py"
    script.write_text('''
import os, sys
sys.path.insert(0, %r)
import numpy as np, torch.distributed as dist
from tcs_b200 import shard
dist.init_process_group("gloo")
rank = dist.get_rank()
val = np.array([0.9, 0.1, 0.5], np.float32) if rank == 0 else np.array([0.2, 0.8, 0.5], np.float32)
col = np.array([1, 1, 2], np.uint32) if rank == 0 else np.array([3, 4, 3], np.uint32)
gv, gc = shard.all_gather_best(val, col, 2)
assert gv.shape == (2, 3) and list(shard.merge_best(gv, gc)) == [1, 4, 2]
dist.destroy_process_group()
''' % ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29541", str(script)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr


def test_bench_reference_arm_prints_its_line():
    """`bench.py --impl reference` (the oracle timed on the host cores) must run without a GPU and print the contract's
    JSON line; ranks other than 0 exit without work."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--small", "--steps", "1",
                        "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "rw_segmentation_vertices_per_s" and line["value"] > 0
    # the reference is single-threaded per scene: the arm uses the host cores scene-parallel and says how many
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert "worker processes" in line["cpu_baseline"]["sample"]
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--config", "c5", "--cpu-procs", "2",
                        "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    c5 = json.loads(r.stdout.strip().splitlines()[-1])
    assert c5["impl"] == "reference" and c5["cpu_baseline"]["cores"] == 2 and c5["config"]["workload"].startswith("C5")
    assert line["e2e"] == {"value": line["value"], "unit": "vertices/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "2", "--small",
                        "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_balanced_batches_give_every_worker_the_same_number_of_batches():
    from tcs_b200 import shard
    assert shard.balanced_batch(1024, 3, 32) == 32          # 11 rounds of 3 batches: 1024 / 33 -> 32
    assert shard.balanced_batch(128, 3, 32) == 22           # 2 rounds: 6 batches of <= 22
    assert shard.balanced_batch(5, 3, 32) == 2 and shard.balanced_batch(0, 3, 32) == 1
    for n in (1, 7, 31, 96, 97, 128, 500, 1024):
        for w in (1, 2, 3, 5):
            for b in (1, 4, 32):
                step = shard.balanced_batch(n, w, b)
                nb = -(-n // step)
                assert 1 <= step <= b and nb <= w * max(1, -(-n // (w * b))), (n, w, b, step)
    with pytest.raises(ValueError):
        shard.balanced_batch(10, 0, 32)
