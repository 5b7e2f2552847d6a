#!/usr/bin/env python
"""bench.py - headline benchmark of the tcs random-walker segmentation hot path on B200.

Metric (BASELINE.json): RW segmentation vertices/s end-to-end (graph build + random-walker solve), with the CG SpMM's
achieved HBM GB/s against the measured B200 peak in `roofline`.

--config c4 (default; BASELINE.json configs[3]): synthetic organised RGB-D cloud 4480x2880 px (~10 M finite points),
  5 mm voxels (~2.85 M vertices, ~19.6 M edges), 16 labels x 10 seeds, class weight preset, FAST arithmetic mode,
  multilevel-preconditioned PCG to a float64 relative residual of 1e-6 per label column.  One step = one whole
  segmentation of one scene.  At N > 1 every rank segments its own scene (independent scenes are the path's natural
  shard; no data-path collective) => weak scaling.  The same S = 2 scenes ROTATE over ranks and steps at every N,
  N = 1 included (rank r runs scene (r + step) mod S), so with an even number of steps every rank of every N does the
  same total work and the per-N values are comparable.
--config c5 (BASELINE.json configs[4]): 1024 synthetic 640x480 scenes (~47 k vertices, 8 labels x 5 seeds each, 6 mm
  voxels) round-robin over the ranks, segmented in batches through tcs_segment_batch (block-diagonal joint solve) from three
  worker contexts per GPU.
  One step = the whole job (1024 scenes); value = vertices/s (scenes/s in config).

  value : clouds already resident in HBM -> labels resident in HBM (CUDA events on the library's stream)
  e2e   : host (pinned) clouds -> upload -> segmentation -> labels back on the host, all inside the timer

`--impl reference` times the CPU restatement of the reference pipeline (oracle/; the reference itself needs PCL / Boost /
Eigen and cannot be built in this image) on the host cores: the reference is single-threaded per scene, so all cores are
used scene-parallel (one process per core, each running the whole pipeline on a bounded sample of the workload).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

C4_W, C4_H, C4_OBJECTS, C4_RES, C4_SEEDS = 4480, 2880, 15, 0.005, 10
C4_SEED0 = 20261104
CROP_W, CROP_H = 1024, 768     # bounded CPU sample: centre crop of the same scene
C5_RES, C5_SEEDS = 0.006, 5
NOMINAL_HBM_GBS = 8000.0


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c4", choices=["c4", "c5"])
    ap.add_argument("--small", action="store_true", help="debug: 1280x960 scene instead of the full C4 shape")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--scenes", type=int, default=1024, help="c5: scenes of the whole job")
    ap.add_argument("--distinct", type=int, default=128, help="c5: distinct generated scenes (cycled to fill the job)")
    ap.add_argument("--batch", type=int, default=32, help="c5: scenes per tcs_segment_batch call")
    ap.add_argument("--contexts", type=int, default=3, help="c5: worker contexts (streams) per GPU, batches round-robin over them")
    ap.add_argument("--cpu-procs", type=int, default=0, help="reference arm: worker processes (0 = all host cores, at most 32)")
    ap.add_argument("--shard", default="scenes", choices=["scenes", "columns"],
                    help="c4, N > 1: 'scenes' = one scene per GPU (weak scaling, no collective, the default); 'columns' = ONE "
                         "scene, its label columns split over the GPUs, one NCCL all-gather of (best potential, colour) "
                         "per vertex (strong scaling)")
    return ap.parse_args()


C4_ROTATION = 2


def c4_scene(args, index):
    """Scene `index` of the C4 rotation set: (w, h, xyz, rgba, label, seed_xyz, seed_label)."""
    from tcs_b200 import synth
    from tcs_b200.seeds import seeds_from_labels
    w, h = (1280, 960) if args.small else (C4_W, C4_H)
    xyz, rgba, label = synth.make_scene(w, h, C4_OBJECTS, C4_SEED0 + index)
    sx, sl, _ = seeds_from_labels(xyz, label, index, C4_SEEDS)
    return w, h, xyz, rgba, label, sx, sl


def crop_scene(w, h, xyz, rgba, label):
    from tcs_b200.seeds import seeds_from_labels
    cw, ch = min(CROP_W, w), min(CROP_H, h)
    r0, c0 = (h - ch) // 2, (w - cw) // 2
    X = xyz.reshape(h, w, 3)[r0:r0 + ch, c0:c0 + cw].reshape(-1, 3).copy()
    C = rgba.reshape(h, w)[r0:r0 + ch, c0:c0 + cw].reshape(-1).copy()
    L = label.reshape(h, w)[r0:r0 + ch, c0:c0 + cw].reshape(-1).copy()
    sx, sl, _ = seeds_from_labels(X, L, 0, C4_SEEDS)
    return cw, ch, X, C, sx, sl


def c5_scene(index):
    from tcs_b200 import synth
    from tcs_b200.seeds import seeds_from_labels
    xyz, rgba, label = synth.make_c5(index)
    sx, sl, _ = seeds_from_labels(xyz, label, index, C5_SEEDS)
    return xyz, rgba, sx, sl


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons of this rank's GPU, sampled DURING the timed region (B200_PROFILING.md's clocks
    line).  Reads NVML in-process (nvidia_ml_py): a spawned `nvidia-smi` takes a driver lock long enough to stall the
    host-synchronised CG loop for ~100 ms per query.  Falls back to nvidia-smi (sparse) only if NVML cannot be loaded.
    The thread idles until arm() and records nothing after disarm()."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt, self._armed = index, [], threading.Event(), threading.Event()
        self.call_ms = []
        self.nv, self.h, self.source = None, None, "nvidia-smi"
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            try:
                uuid = "GPU-" + str(torch.cuda.get_device_properties(index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
                ids = [v for v in vis.split(",") if v.strip()]
                self.h = pynvml.nvmlDeviceGetHandleByIndex(int(ids[index]) if ids and ids[index].strip().isdigit() else index)
            self.nv, self.source = pynvml, "nvml"
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            for _ in range(3):          # the first call of an NVML query initialises state in the driver (tens of ms, and it
                self._sample_nvml()     # stalls the GPU work of this process): pay that here, not inside the timed region
        except Exception:
            self.nv = None

    def arm(self):
        self._armed.set()

    def disarm(self):
        self._armed.clear()

    def _sample_nvml(self):
        nv = self.nv
        mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
        try:
            mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
        except Exception:
            mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
        flags = [(nv.nvmlClocksThrottleReasonHwSlowdown, "hw_slowdown"), (nv.nvmlClocksThrottleReasonHwThermalSlowdown, "hw_thermal_slowdown"),
                 (nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_thermal_slowdown"), (nv.nvmlClocksThrottleReasonSwPowerCap, "sw_power_cap")]
        return [str(mhz), str(self.max_mhz)] + ["Active" if mask & bit else "Not Active" for bit, _ in flags]

    def run(self):
        while not self._halt.is_set():
            if not self._armed.is_set():
                self._halt.wait(0.005)
                continue
            try:
                if self.nv is not None:
                    t0 = time.perf_counter()
                    self.rows.append(self._sample_nvml())
                    self.call_ms.append(1e3 * (time.perf_counter() - t0))
                else:
                    out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                    self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self._halt.wait(0.2 if self.nv is not None else 1.0)

    def stop(self):
        self._halt.set()
        if self.is_alive():
            self.join(timeout=6)
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for n, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_min_mhz": min(sm) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm), "source": self.source + ", sampled inside the timed region only",
                "query_ms_max": max(self.call_ms) if self.call_ms else None}


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6650.0, "fallback 6650 GB/s (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------------- CPU side (oracle)
# Only the cpu_baseline leg and --impl reference execute anything under oracle/ (through tests/orc.py), and only as the
# timed baseline / the checker - never as part of the product path.
_CPU_SAMPLE = {}      # filled before the worker pool forks: the workers inherit it


def _orc():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orc
    return orc


def _cpu_worker(i):
    """One pass of the CPU restatement (reference algorithm incl. the float32 sparse LDL^T) on sample i.  (V, seconds)."""
    orc = _orc()
    X, C, sx, sl, res = _CPU_SAMPLE["scenes"][i % len(_CPU_SAMPLE["scenes"])]
    t = time.perf_counter()
    r = orc.segment_cloud(X, C, res, sx, sl, preset=0, mode=orc.MODE_PARITY, solver=orc.SOLVE_LDLT_F32)
    return r["graph"].V, time.perf_counter() - t


def host_cores(args):
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        n = os.cpu_count() or 1
    return max(1, min(args.cpu_procs or n, 32, n))


def cpu_sample(args):
    """The bounded CPU sample of the workload: list of (xyz, rgba, seed_xyz, seed_label, resolution) + a description."""
    P = host_cores(args)
    if args.config == "c5":
        scenes = [c5_scene(i) + (C5_RES,) for i in range(P)]
        return scenes, "%d of the C5 scenes (640x480 px, ~47 k vertices each), one per worker process" % P
    w, h, xyz, rgba, label, _, _ = c4_scene(args, 0)
    cw, ch, X, C, sx, sl = crop_scene(w, h, xyz, rgba, label)
    return [(X, C, sx, sl, C4_RES)], "centre %dx%d px crop of C4 scene 0 (V = 186 k at full size), the same crop in every worker process" % (cw, ch)


def cpu_throughput(args, steps, warmup):
    """Scene-parallel CPU pipeline on all host cores: returns (vertices/s, seconds per step, cores, sample text, V per step)."""
    import multiprocessing as mp
    scenes, text = cpu_sample(args)
    _CPU_SAMPLE["scenes"] = scenes
    _orc().lib()                      # build / load the oracle before forking
    P = host_cores(args)
    with mp.get_context("fork").Pool(P) as pool:
        for _ in range(warmup):
            pool.map(_cpu_worker, range(P))
        t0 = time.perf_counter()
        V = 0
        for _ in range(steps):
            V += sum(v for v, _dt in pool.map(_cpu_worker, range(P)))
        dt = time.perf_counter() - t0
    return V / dt, dt / steps, P, text, V // max(1, steps)


def workload_config(args, world):
    if args.config == "c5":
        return {"workload": "C5: %d synthetic 640x480 RGB-D scenes (%d distinct, cycled; ~47 k vertices each), %g m voxels, "
                            "8 labels x %d seeds, class weight preset, FAST mode, PCG tol 1e-6; scenes round-robin over the "
                            "ranks, batches of at most %d through tcs_segment_batch" % (args.scenes, args.distinct, C5_RES, C5_SEEDS, args.batch),
                "scenes_per_step": args.scenes,
                "l2": "a batch's merged system (1.5 M rows x 8 columns: vectors 48 MB each, matrix 170 MB, hierarchy) exceeds L2"}
    w, h = (1280, 960) if args.small else (C4_W, C4_H)
    return {"workload": "C4: synthetic %dx%d RGB-D cloud (~%.1f M finite points), %g m voxels, %d labels x %d seeds, "
                        "class weight preset, FAST mode, PCG tol 1e-6; one scene per GPU" %
                        (w, h, w * h * 0.78 / 1e6, C4_RES, C4_OBJECTS + 1, C4_SEEDS),
            "scenes_per_step": world, "l2": "inputs and solver vectors are larger than L2 (cloud 160 MB, each n x K block 182 MB)"}


def run_reference(args, rank, world):
    if rank != 0:
        return                      # the CPU arm does not shard over GPUs: rank 0 alone runs and prints it
    warm = min(args.warmup, 1)
    val, s_per_step, P, text, v_step = cpu_throughput(args, args.steps, warm)
    sample = "%s; %d worker processes x whole pipeline incl. float32 sparse LDL^T per step (%d vertices per step)" % (text, P, v_step)
    print(json.dumps({
        "impl": "reference", "metric": "rw_segmentation_vertices_per_s", "value": val, "unit": "vertices/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": warm, "ms_per_step": 1e3 * s_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {"value": val, "unit": "vertices/s", "cores": P, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "vertices/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


# ---------------------------------------------------------------------------------------------------- GPU side
def _dist_setup(local_rank, world):
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    return torch, dist, barrier


def _reduce(torch, dist, world, maxes, sums):
    """(max over ranks of `maxes`, sum over ranks of `sums`) - bookkeeping only, no data-path collective."""
    tm = torch.tensor(maxes, dtype=torch.float64, device="cuda")
    ts = torch.tensor(sums, dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(ts, op=dist.ReduceOp.SUM)
    return tm.tolist(), ts.tolist()


def _pinned_records(torch, xyz, rgba):
    """Host side of the e2e arm: pinned interleaved records {x, y, z, rgba} (what a caller holding pcl::PointXYZRGB
    data passes)."""
    n = len(xyz)
    host = torch.empty((n, 4), dtype=torch.float32).pin_memory()
    host[:, :3] = torch.from_numpy(xyz)
    host[:, 3] = torch.from_numpy(rgba.view(np.float32))
    return host


def parity_and_same_input(args, ctx, capi, crop, cpu_labels_of):
    """GPU on the very crop the CPU arm times: label parity with the CPU run (oracle float64 top-2 gap at every
    mismatching vertex) and the GPU's time on that input (device time, cloud resident, FAST and PARITY modes)."""
    import torch
    cw, ch, X, C, csx, csl = crop
    orc = _orc()
    import label_gate
    ctx.upload_cloud(X, C)
    out = {}
    ref = cpu_labels_of()                      # oracle: MODE_PARITY, float32 LDL^T (the timed CPU run's labels)
    for mode_name, mode in (("parity", capi.MODE_PARITY), ("fast", capi.MODE_FAST)):
        best = None
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s = torch.cuda.ExternalStream(ctx.stream)
            e0.record(s)
            r = ctx.segment_cloud(C4_RES, csx, csl, preset=capi.PRESET_CLASS, mode=mode)
            e1.record(s)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
            labels = r["vertex_colour"]
            r["graph"].close()
        out[mode_name + "_ms"] = best
        if mode == capi.MODE_PARITY:
            diff = np.nonzero(labels != ref["vertex_colour"])[0]
            gaps = label_gate.oracle_gaps(ref["graph"], csx, csl, diff) if len(diff) else np.zeros(0)
            out["parity"] = {"input": "%dx%d crop, PARITY mode, GPU default solver vs oracle float32 LDL^T" % (cw, ch),
                             "vertices": int(len(labels)), "mismatches": int(len(diff)),
                             "near_ties": int(np.count_nonzero(gaps < label_gate.NEAR_TIE_GAP)),
                             "max_oracle_top2_gap": float(gaps.max()) if len(gaps) else 0.0, "bound": label_gate.NEAR_TIE_GAP}
    return out


def run_c4(args, rank, local_rank, world):
    torch, dist, barrier = _dist_setup(local_rank, world)
    from tcs_b200 import capi
    columns = args.shard == "columns" and world > 1
    # distinct scenes of the rotation: the same set at every N (N = 1 included), so that every rank of every N does the
    # same mix of work per pair of steps and the per-N values are comparable (scene 0: 136 CG iterations, scene 1: 200)
    S = 1 if columns else C4_ROTATION
    mine = sorted({(rank + s) % S for s in range(S)}) if not columns else [0]
    scenes = {k: c4_scene(args, k) for k in mine}
    stream = torch.cuda.Stream()
    ctx = capi.Context(local_rank, stream=stream.cuda_stream)
    host, dev, outs = {}, {}, {}
    for k, (w, h, xyz, rgba, label, sx, sl) in scenes.items():
        host[k] = _pinned_records(torch, xyz, rgba)
        dev[k] = host[k].to("cuda", non_blocking=False)
    n = len(next(iter(scenes.values()))[2])
    pc_host = torch.empty(n, dtype=torch.int32).pin_memory()
    vc_host = torch.empty(n, dtype=torch.int32).pin_memory()
    opts = capi.default_opts(time_spmm=1)
    if columns:     # ONE scene: this rank solves its block of label columns
        from tcs_b200 import shard
        opts.column_begin, opts.column_count = shard.columns_for_rank(len(np.unique(scenes[0][6])), rank, world)
        if opts.column_count == 0:
            raise SystemExit("bench.py: more ranks than label columns")

    def merge(graph):
        """The column shard's one exchange: all-gather of (best potential, colour) per vertex over NCCL."""
        if columns:
            tv, tc = graph.best_device()          # device-resident: no host staging before the all-gather
            shard._merge_best_nccl(tv, tc, world)

    def scene_of(step):
        return 0 if columns else (rank + step) % S

    def step_resident(step):
        k = scene_of(step)
        ctx.set_device_cloud(dev[k].data_ptr(), n, True)
        r = ctx.segment_cloud(C4_RES, scenes[k][5], scenes[k][6], preset=capi.PRESET_CLASS, mode=capi.MODE_FAST, opts=opts,
                              want_host=False)
        merge(r["graph"])
        return r

    def step_e2e(step):
        import ctypes as C
        k = scene_of(step)
        L = capi.lib()
        base = host[k].data_ptr()
        ctx.check(L.tcs_cloud_upload(ctx.h, C.c_void_p(base), 16, C.c_void_p(base + 12), 16, n))
        ctx.n_points_hint = n
        info = capi.SolveInfo()
        gh = C.c_void_p()
        sxx, sll = np.ascontiguousarray(scenes[k][5], np.float32), np.ascontiguousarray(scenes[k][6], np.uint32)
        st = L.tcs_segment_cloud(ctx.h, C4_RES, capi.PRESET_CLASS, capi.MODE_FAST, sxx.ctypes.data, sll.ctypes.data,
                                 len(sll), C.byref(opts), C.c_void_p(pc_host.data_ptr()),
                                 C.c_void_p(vc_host.data_ptr()), C.byref(info), C.byref(gh))
        ctx.check(st)
        g = capi.Graph(ctx, gh)
        merge(g)
        return dict(graph=g, info=info)

    def timed(step_fn, sampler=None):
        """W warm-up steps (every scene of the rotation at least once), then K steps between two events."""
        for s in range(max(args.warmup, S)):
            step_fn(s)["graph"].close()
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        infos, V, E = [], 0, 0
        launches0 = ctx.launches
        if sampler:
            sampler.arm()
        with torch.cuda.stream(stream):
            ev0.record(stream)
            for s in range(args.steps):
                t_host = time.perf_counter()
                r = step_fn(s)
                infos.append(r["info"].as_dict())
                if os.environ.get("TCS_BENCH_DEBUG"):
                    print("[rank %d] step %d: host %.1f ms, iterations %d, stage_ms %s" % (
                        rank, s, 1e3 * (time.perf_counter() - t_host), infos[-1]["iterations"],
                        [round(x, 1) for x in infos[-1]["stage_ms"][:12]]), file=sys.stderr, flush=True)
                V += r["graph"].V
                E = r["graph"].E
                r["graph"].close()
            ev1.record(stream)
        barrier()
        if sampler:
            sampler.disarm()
        return ev0.elapsed_time(ev1), infos, V, E, ctx.launches - launches0

    sampler = ClockSampler(local_rank)
    if not os.environ.get("TCS_BENCH_NO_CLOCKS"):
        sampler.start()
    ms, infos, V, E, launches = timed(step_resident, sampler)
    clocks = sampler.stop()
    ms_e2e, infos_e, Ve, _, launches_e = timed(step_e2e)
    its = [i["iterations"] for i in infos]
    (ms_max, ms_e2e_max, it_min_neg, it_max), (V_all, Ve_all, launches_all) = _reduce(
        torch, dist, world, [ms, ms_e2e, -float(sum(its)), float(sum(its))], [float(V), float(Ve), float(launches + launches_e)])
    if columns:     # one scene in total: its vertices count once
        V_all, Ve_all = float(V), float(Ve)
    if rank == 0:
        info = infos[-1]
        nrows, nnz, K_total = info["rows"], info["nnz"], info["ncolours"]
        K = opts.column_count if columns else K_total       # the SpMM of a column shard carries its own columns only
        b_spmm = 8 * nnz + 4 * (nrows + 1) + 8 * nrows * K      # SURVEY.md 8(d) / DESIGN.md
        spmm_ms = sum(i["spmm_ms"] for i in infos)
        spmm_n = sum(i["spmm_launches"] for i in infos)
        avg_ms = spmm_ms / max(1, spmm_n)
        peak, peak_src = peaks()
        achieved = b_spmm / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
        kname = {1: "k_spmm", 2: "k_spmm_tiled", 3: "k_spmm_ring"}.get(info.get("spmm_kernel"), "k_spmm")
        traffic = None    # DRAM bytes per launch of that kernel from the committed ncu --set full capture
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "spmm_traffic.json")))
            if tr.get("kernel") == kname and not args.small and not columns:
                traffic = int(tr["dram_bytes_read"]) + int(tr["dram_bytes_write"])
        except Exception:
            pass
        balanced = args.steps % S == 0
        out = {
            "metric": "rw_segmentation_vertices_per_s", "value": V_all / (ms_max * 1e-3), "unit": "vertices/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, S), "ms_per_step": ms_max / args.steps,
            "higher_is_better": True, "scaling": "strong" if columns else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": dict(workload_config(args, 1 if columns else world),
                           shard="label columns of one scene, one NCCL all-gather per step" if columns else
                           "independent scenes, no collective; %d distinct scenes rotate over ranks and steps (rank r runs "
                           "scene (r + step) mod %d)%s" % (S, S, "" if balanced else "; steps is not a multiple of it: ranks do unequal work"),
                           vertices=int(V // args.steps), edges=int(E), rows=nrows, nnz=nnz, labels=K_total,
                           pcg_iterations_per_step=sum(its) // args.steps, pcg_iterations_rank0=its,
                           pcg_iterations_total_min_max_over_ranks=[int(-it_min_neg), int(it_max)],
                           converged=all(i["converged"] for i in infos), max_rel_residual=max(max(i["rel_residual"]) for i in infos),
                           preconditioner="Jacobi" if os.environ.get("TCS_AMG") == "0" else
                           "Jacobi + V(1,1) cycle on a device-built aggregation hierarchy (amg.cuh)"),
            "e2e": {"value": Ve_all / (ms_e2e_max * 1e-3), "unit": "vertices/s", "h2d_bytes_per_step": 16 * n + 16 * len(scenes[mine[0]][6]),
                    "d2h_bytes_per_step": 4 * n + 4 * int(Ve // args.steps), "ms_per_step": ms_e2e_max / args.steps,
                    "warmup": max(args.warmup, S)},
            "gpu_launches": int(launches_all),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": kname + " (CG SpMM over all label columns)", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak if peak else None, "traffic": traffic,
                         "frac_dram": (traffic / (avg_ms * 1e-3) / 1e9 / peak) if traffic and avg_ms > 0 else None,
                         "frac_nominal": achieved / NOMINAL_HBM_GBS,
                         "algorithmic_bytes_per_launch": b_spmm, "avg_launch_ms": avg_ms,
                         "launches_per_step": spmm_n // args.steps, "spmm_share_of_step": spmm_ms / ms if ms else None,
                         "peak_source": peak_src},
            "stage_ms": {k: round(v, 3) for k, v in zip(
                ["build", "normals", "smoothen", "normals2", "signs", "weights", "components", "seeds", "assemble",
                 "solve", "argmax", "point_labels"], info["stage_ms"])},
        }
        if world == 1 and not args.no_cpu_baseline:
            w, h, xyz, rgba, label = scenes[0][:5]
            crop = crop_scene(w, h, xyz, rgba, label)
            keep = {}

            def cpu_once():
                if "r" not in keep:
                    orc = _orc()
                    t = time.perf_counter()
                    keep["r"] = orc.segment_cloud(crop[2], crop[3], C4_RES, crop[4], crop[5], preset=0, mode=orc.MODE_PARITY,
                                                  solver=orc.SOLVE_LDLT_F32)
                    keep["dt"] = time.perf_counter() - t
                return keep["r"]
            same = parity_and_same_input(args, ctx, capi, crop, cpu_once)
            v, dt = keep["r"]["graph"].V, keep["dt"]
            out["cpu_baseline"] = {"value": v / dt, "unit": "vertices/s", "cores": 1, "kind": "port",
                                   "sample": "centre %dx%d px crop of scene 0 (V=%d), whole pipeline incl. float32 sparse "
                                             "LDL^T, one pass on one core (the reference is single-threaded), %.1f s" % (crop[0], crop[1], v, dt)}
            out["parity"] = same["parity"]
            out["vs_cpu_same_input"] = {"input": "the cpu_baseline crop (V=%d)" % v, "cpu_ms": 1e3 * dt, "gpu_ms_parity_mode": same["parity_ms"],
                                        "gpu_ms_fast_mode": same["fast_ms"], "speedup_parity_mode": 1e3 * dt / same["parity_ms"]}
        print(json.dumps(out), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def _cpu_worker_labels(i):
    orc = _orc()
    X, C, sx, sl, res = _CPU_SAMPLE["scenes"][i]
    t = time.perf_counter()
    r = orc.segment_cloud(X, C, res, sx, sl, preset=0, mode=orc.MODE_PARITY, solver=orc.SOLVE_LDLT_F32)
    return r["graph"].V, time.perf_counter() - t, r["vertex_colour"]


def run_c5(args, rank, local_rank, world):
    import ctypes as C
    torch, dist, barrier = _dist_setup(local_rank, world)
    from tcs_b200 import capi, shard
    L = capi.lib()
    mine = shard.scenes_for_rank(args.scenes, rank, world)          # scene i -> rank i mod world
    D = max(1, min(args.distinct, args.scenes))
    distinct = sorted({j % D for j in mine})
    data = {k: c5_scene(k) for k in distinct}
    n = 640 * 480
    # W worker contexts (one stream each) share the GPU: while one batch is in its bandwidth-bound solve, another
    # worker's batch runs its launch-bound graph / hierarchy stages and its copies - the same scheme as the C++ driver
    # (include/tcs/batch_segmenter.h).  Batch i belongs to worker i mod W.
    W = max(1, args.contexts)
    streams = [torch.cuda.Stream() for _ in range(W)]
    fence = torch.cuda.Stream()
    ctxs = [capi.Context(local_rank, stream=st.cuda_stream) for st in streams]
    ctx = ctxs[0]
    host = {k: _pinned_records(torch, data[k][0], data[k][1]) for k in distinct}
    dev = {k: host[k].to("cuda", non_blocking=False) for k in distinct}
    seeds = {k: (np.ascontiguousarray(data[k][2], np.float32), np.ascontiguousarray(data[k][3], np.uint32)) for k in distinct}
    # batches of at most --batch scenes, sized so that every worker context gets the same number of batches (128 scenes
    # on 3 contexts: 6 batches of 21-22 instead of 4 of 32, which would leave two contexts idle in the second round)
    B = shard.balanced_batch(len(mine), W, max(1, args.batch))
    pc_host = [torch.empty((B, n), dtype=torch.int32).pin_memory() for _ in range(W)]
    vc_host = [torch.empty((B, n), dtype=torch.int32).pin_memory() for _ in range(W)]
    opts = capi.default_opts(time_spmm=0)
    opts_timed = capi.default_opts(time_spmm=1)

    def make_batches(resident):
        out = []
        for i, b0 in enumerate(range(0, len(mine), B)):
            ids = mine[b0:b0 + B]
            arr = (capi.Scene * len(ids))()
            for slot, j in enumerate(ids):
                k = j % D
                a = arr[slot]
                a.n, a.ns = n, len(seeds[k][1])
                a.seed_xyz, a.seed_label = seeds[k][0].ctypes.data, seeds[k][1].ctypes.data
                if resident:
                    a.d_xyzc, a.d_has_color = dev[k].data_ptr(), 1
                else:
                    base = host[k].data_ptr()
                    a.xyz, a.xyz_stride, a.rgba, a.rgba_stride = base, 16, base + 12, 16
                    a.point_cluster = pc_host[i % W].data_ptr() + slot * n * 4
                    a.vertex_colour = vc_host[i % W].data_ptr() + slot * n * 4
            out.append(arr)
        return out

    def run_batch(c, arr, o):
        info = capi.SolveInfo()
        c.check(L.tcs_segment_batch(c.h, arr, len(arr), C5_RES, capi.PRESET_CLASS, capi.MODE_FAST, C.byref(o), C.byref(info)))
        return sum(a.V for a in arr), info.as_dict()

    def job(batches):
        """One step = this rank's share of the job.  Returns (vertices, per-batch infos)."""
        res, errs = [[] for _ in range(W)], []

        def worker(w):
            try:
                torch.cuda.set_device(local_rank)
                for arr in batches[w::W]:
                    res[w].append(run_batch(ctxs[w], arr, opts))
            except Exception as e:       # surfaced below: a failing batch must fail the bench
                errs.append(e)

        th = [threading.Thread(target=worker, args=(w,)) for w in range(W)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        if errs:
            raise errs[0]
        flat = [r for w in range(W) for r in res[w]]
        return sum(r[0] for r in flat), [r[1] for r in flat]

    def timed(batches, sampler=None):
        for _ in range(args.warmup):
            job(batches)
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        V, infos = 0, []
        launches0 = sum(c.launches for c in ctxs)
        if sampler:
            sampler.arm()
        ev0.record(fence)
        for st in streams:
            st.wait_event(ev0)
        for _ in range(args.steps):
            v, inf = job(batches)
            V += v
            infos += inf
        for st in streams:                 # the step ends when the slowest worker stream has drained
            e = torch.cuda.Event()
            e.record(st)
            fence.wait_event(e)
        ev1.record(fence)
        barrier()
        if sampler:
            sampler.disarm()
        return ev0.elapsed_time(ev1), V, infos, sum(c.launches for c in ctxs) - launches0

    sampler = ClockSampler(local_rank)
    sampler.start()
    ms, V, infos, launches = timed(make_batches(True), sampler)
    clocks = sampler.stop()
    ms_e2e, Ve, infos_e, launches_e = timed(make_batches(False))
    # the SpMM of one batch timed ALONE (one context, nothing else on the GPU): the roofline's kernel time
    alone = [run_batch(ctx, arr, opts_timed)[1] for arr in make_batches(True)[:4]]
    (ms_max, ms_e2e_max), (V_all, Ve_all, launches_all, scenes_all) = _reduce(
        torch, dist, world, [ms, ms_e2e], [float(V), float(Ve), float(launches + launches_e), float(len(mine))])
    if rank == 0:
        full = [i for i in alone if i["rows"] == max(j["rows"] for j in alone)] or alone
        info = full[0]
        nrows, nnz, K = info["rows"], info["nnz"], info["ncolours"]
        b_spmm = 8 * nnz + 4 * (nrows + 1) + 8 * nrows * K
        spmm_ms = sum(i["spmm_ms"] for i in full)
        spmm_n = sum(i["spmm_launches"] for i in full)
        avg_ms = spmm_ms / max(1, spmm_n)
        peak, peak_src = peaks()
        achieved = b_spmm / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
        kname = {1: "k_spmm", 2: "k_spmm_tiled", 3: "k_spmm_ring"}.get(info.get("spmm_kernel"), "k_spmm")
        its = [i["iterations"] for i in infos[:len(infos) // max(1, args.steps)]]
        out = {
            "metric": "rw_segmentation_vertices_per_s", "value": V_all / (ms_max * 1e-3), "unit": "vertices/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(args, world), shard="independent scenes round-robin over the ranks, no collective",
                           scenes_per_s=scenes_all * args.steps / (ms_max * 1e-3),
                           scenes_per_s_e2e=scenes_all * args.steps / (ms_e2e_max * 1e-3),
                           vertices_per_scene=int(V // args.steps // max(1, len(mine))), batch_rows=nrows, batch_nnz=nnz, labels=K,
                           contexts_per_gpu=W, scenes_per_batch_rank0=B, pcg_iterations_per_batch_rank0=its, converged=all(i["converged"] for i in infos),
                           max_rel_residual=max(max(i["rel_residual"]) for i in infos),
                           preconditioner="Jacobi + V(1,1) cycle on a device-built aggregation hierarchy of the merged "
                                          "block-diagonal system; one CG recurrence per (scene, column)"),
            "e2e": {"value": Ve_all / (ms_e2e_max * 1e-3), "unit": "vertices/s",
                    "h2d_bytes_per_step": int(scenes_all) * (16 * n + 16 * 40), "d2h_bytes_per_step": int(scenes_all) * 4 * n + 4 * int(Ve_all // args.steps),
                    "ms_per_step": ms_e2e_max / args.steps, "warmup": args.warmup},
            "gpu_launches": int(launches_all),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": kname + " (CG SpMM of a batch's merged system, all label columns)",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak if peak else None,
                         "traffic": None, "frac_nominal": achieved / NOMINAL_HBM_GBS, "algorithmic_bytes_per_launch": b_spmm,
                         "avg_launch_ms": avg_ms, "launches_per_step": sum(i["spmm_launches"] for i in infos) // args.steps,
                         "spmm_share_of_solve": (avg_ms * full[0]["iterations"] / full[0]["stage_ms"][9]) if full[0]["stage_ms"][9] else None,
                         "timed": "one batch alone on the GPU, outside the timed region (the step overlaps %d contexts)" % W,
                         "peak_source": peak_src},
            "stage_ms_first_batch": {"graphs_features_systems": round(infos[0]["stage_ms"][0], 3), "merge": round(infos[0]["stage_ms"][8], 3),
                                     "solve": round(infos[0]["stage_ms"][9], 3), "labels_and_copies": round(infos[0]["stage_ms"][10], 3)},
        }
        if world == 1 and not args.no_cpu_baseline:
            import multiprocessing as mp
            P = host_cores(args)
            ids = list(range(min(P, D)))
            sample = [c5_scene(k) for k in ids]
            _CPU_SAMPLE["scenes"] = [s + (C5_RES,) for s in sample]
            _orc().lib()
            t0 = time.perf_counter()
            with mp.get_context("fork").Pool(len(ids)) as pool:
                res = pool.map(_cpu_worker_labels, range(len(ids)))
            dt = time.perf_counter() - t0
            vs = sum(r[0] for r in res)
            out["cpu_baseline"] = {"value": vs / dt, "unit": "vertices/s", "cores": len(ids), "kind": "port",
                                   "sample": "C5 scenes 0..%d, one per worker process, whole pipeline incl. float32 sparse LDL^T, "
                                             "%.1f s wall (%.2f s per scene on one core)" % (len(ids) - 1, dt, float(np.mean([r[1] for r in res])))}
            # parity: the same scenes through tcs_segment_batch in PARITY mode vs the CPU labels
            got, _info, _st = ctx.segment_batch([(s[0], s[1], s[2], s[3]) for s in sample], C5_RES, preset=capi.PRESET_CLASS, mode=capi.MODE_PARITY)
            import label_gate
            orc = _orc()
            mism, near, worst = 0, 0, 0.0
            for s, g, r in zip(sample, got, res):
                diff = np.nonzero(g["vertex_colour"] != r[2])[0]
                if len(diff):
                    ref = orc.segment_cloud(s[0], s[1], C5_RES, s[2], s[3], preset=0, mode=orc.MODE_PARITY, solver=orc.SOLVE_LDLT_F64)
                    gaps = label_gate.oracle_gaps(ref["graph"], s[2], s[3], diff)
                    mism += len(diff); near += int(np.count_nonzero(gaps < label_gate.NEAR_TIE_GAP)); worst = max(worst, float(gaps.max()))
            out["parity"] = {"input": "C5 scenes 0..%d, PARITY mode, tcs_segment_batch vs oracle float32 LDL^T" % (len(ids) - 1),
                             "vertices": int(vs), "mismatches": mism, "near_ties": near, "max_oracle_top2_gap": worst,
                             "bound": label_gate.NEAR_TIE_GAP}
        print(json.dumps(out), flush=True)
    for c in ctxs:
        c.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    elif args.config == "c5":
        run_c5(args, rank, local_rank, world)
    else:
        run_c4(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
