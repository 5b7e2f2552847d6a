"""Many scenes per call: tcs_segment_batch (block-diagonal joint solve) and the C++ multi-GPU driver above it
(include/tcs/batch_segmenter.h, tcs_b200/cpp/tcs_batch_bench) - BASELINE.json configs[4] (C5), SURVEY.md 8e row 1.
Each scene of a batch has to come out as the reference's single-scene segment() would label it
(impl/random_walker_segmentation.hpp:138-224): the oracle's float64 direct solve is the checker, through the 1e-6
near-tie gate of tests/label_gate.py."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "scripts"))
TOOL = os.path.join(ROOT, "tcs_b200", "cpp", "tcs_batch_bench")

from write_scenes import read_labels, write_scenes  # noqa: E402


def _build():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "tcs_b200", "cpp"), "-s", "tcs_batch_bench"])


def _small_scenes(count, w=160, h=120):
    """ragged little scenes: different sizes, label counts and seed counts"""
    from tcs_b200 import synth
    from tcs_b200.seeds import seeds_from_labels
    out = []
    for i in range(count):
        xyz, rgba, label = synth.make_scene(w + 16 * i, h + 8 * (i % 3), 3 + i % 4, 977 + i)
        sx, sl, _ = seeds_from_labels(xyz, label, i, 2 + i % 3)
        out.append((xyz, rgba, sx, sl))
    return out


def test_batch_tool_builds_and_fails_loudly_without_gpu(tmp_path):
    """plain g++ above the C ABI; on a box without a CUDA device the driver throws (exit 3), it never computes on the CPU"""
    import torch
    _build()
    path = str(tmp_path / "scenes.bin")
    write_scenes(path, _small_scenes(2))
    r = subprocess.run([TOOL, path, "--repeat", "0"], capture_output=True, text=True, timeout=300)
    if torch.cuda.is_available():
        assert r.returncode == 0, r.stderr
        assert json.loads(r.stdout.strip().splitlines()[-1])["scenes"] == 2
    else:
        assert r.returncode == 3 and "no CPU fallback" in r.stderr, (r.returncode, r.stderr)
    bad = str(tmp_path / "bad.bin")
    open(bad, "wb").write(b"\x05\x00\x00\x00\x00\x00\x00\x00\xff\xff")
    r = subprocess.run([TOOL, bad], capture_output=True, text=True, timeout=60)
    assert r.returncode == 2 and "scene" in r.stderr


@pytest.fixture(scope="module")
def ctx():
    from tcs_b200 import capi
    c = capi.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def ragged():
    import orc
    scenes = _small_scenes(5)
    refs = {}
    for mode in (orc.MODE_PARITY, orc.MODE_FAST):
        refs[mode] = [orc.segment_cloud(x, c, 0.006, sx, sl, preset=0, mode=mode, solver=orc.SOLVE_LDLT_F64)
                      for x, c, sx, sl in scenes]
    return scenes, refs


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["parity", "fast"])
def test_segment_batch_matches_oracle_per_scene(ctx, ragged, mode):
    import label_gate
    import orc
    from tcs_b200 import capi
    scenes, refs = ragged
    gm, om = (capi.MODE_PARITY, orc.MODE_PARITY) if mode == "parity" else (capi.MODE_FAST, orc.MODE_FAST)
    out, info, st = ctx.segment_batch(scenes, 0.006, preset=capi.PRESET_CLASS, mode=gm)
    assert st == capi.OK and info.as_dict()["converged"]
    for i, (o, ref, sc) in enumerate(zip(out, refs[om], scenes)):
        assert (o["V"], o["E"]) == (ref["graph"].V, ref["graph"].E)
        assert o["K"] == len(np.unique(sc[3]))
        nd = label_gate.check_labels("batch[%d] %s" % (i, mode), o["vertex_colour"], ref, sc[2], sc[3])
        assert np.count_nonzero(o["point_cluster"] != ref["point_cluster"]) <= 10 * nd


@pytest.mark.gpu
def test_segment_batch_equals_single_scene_calls(ctx, ragged):
    """a batch of one is the single-scene path; a joint solve may only differ from it at near-ties"""
    from tcs_b200 import capi
    scenes, _ = ragged
    joint, _, _ = ctx.segment_batch(scenes, 0.006, mode=capi.MODE_FAST)
    for i, sc in enumerate(scenes):
        one, _, st = ctx.segment_batch([sc], 0.006, mode=capi.MODE_FAST)
        assert st == capi.OK
        assert np.count_nonzero(one[0]["vertex_colour"] != joint[i]["vertex_colour"]) <= 2


@pytest.mark.gpu
def test_cpp_batch_segmenter_matches_oracle(ragged, tmp_path):
    """the C++ driver: two worker contexts on the device, batches of 2 (so the last batch is ragged), e2e from host buffers"""
    import label_gate
    import orc
    _build()
    scenes, refs = ragged
    path, lab = str(tmp_path / "scenes.bin"), str(tmp_path / "labels.bin")
    write_scenes(path, scenes)
    r = subprocess.run([TOOL, path, "--batch", "2", "--contexts", "2", "--repeat", "1", "--labels", lab],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["scenes"] == len(scenes) and line["not_converged"] == 0 and line["scenes_per_s"] > 0
    for i, (got, ref, sc) in enumerate(zip(read_labels(lab, len(scenes)), refs[orc.MODE_PARITY], scenes)):
        label_gate.check_labels("BatchSegmenter[%d]" % i, got, ref, sc[2], sc[3])


@pytest.mark.gpu
def test_cpp_batch_segmenter_cycles_a_job_over_every_visible_gpu(tmp_path):
    import torch
    _build()
    path = str(tmp_path / "scenes.bin")
    write_scenes(path, _small_scenes(3))
    g = torch.cuda.device_count()
    r = subprocess.run([TOOL, path, "--gpus", str(g), "--batch", "4", "--total", str(8 * g + 3), "--fast", "--repeat", "1"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["n_gpus"] == g and line["scenes"] == 8 * g + 3 and line["not_converged"] == 0
