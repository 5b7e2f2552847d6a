#!/usr/bin/env python
"""bench.py - headline benchmark of the tcs random-walker segmentation hot path on B200.

Metric (BASELINE.json): RW segmentation vertices/s end-to-end (graph build + random-walker solve), with the CG SpMM's
achieved HBM GB/s against the measured B200 peak in `roofline`.

Workload (BASELINE.json configs[3], "C4"): synthetic organised RGB-D cloud 4480x2880 px (~10 M finite points), 5 mm
voxels (~2.8 M vertices, ~19.6 M edges), 16 labels x 10 seeds, class weight preset, FAST arithmetic mode, PCG to a
float64 relative residual of 1e-6 per label column.  One step = one whole segmentation of that scene.
At N > 1 every rank segments its own scene of the same shape (independent scenes are the path's natural shard; no
data-path collective) => weak scaling; value = vertices of all ranks / max-over-ranks time.

  value : cloud already resident in HBM -> labels resident in HBM (CUDA events on the library's stream)
  e2e   : host (pinned) cloud -> tcs_cloud_upload -> tcs_segment_cloud -> labels back on the host, all inside the timer

`--impl reference` times the CPU restatement of the reference pipeline (oracle/, single thread like the reference;
the reference itself needs PCL/Boost/Eigen and cannot be built in this image) on a bounded crop of the same scene.
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
CROP_W, CROP_H = 1024, 768     # bounded CPU sample: centre crop of the same scene


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--small", action="store_true", help="debug: 1280x960 scene instead of the full C4 shape")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--shard", default="scenes", choices=["scenes", "columns"],
                    help="N > 1: 'scenes' = one scene per GPU (weak scaling, no collective, the default); 'columns' = ONE "
                         "scene, its label columns split over the GPUs, one NCCL all-gather of (best potential, colour) "
                         "per vertex (strong scaling)")
    return ap.parse_args()


def make_scene(args, rank):
    from tcs_b200 import synth
    from tcs_b200.seeds import seeds_from_labels
    w, h = (1280, 960) if args.small else (C4_W, C4_H)
    xyz, rgba, label = synth.make_scene(w, h, C4_OBJECTS, 20261104 + rank)
    sx, sl, _ = seeds_from_labels(xyz, label, rank, C4_SEEDS)
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


def cpu_pipeline(X, C, sx, sl):
    """One pass of the CPU restatement (oracle) - reference algorithm incl. the sparse LDL^T solve.  Returns (V, s)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orc
    t = time.perf_counter()
    r = orc.segment_cloud(X, C, C4_RES, sx, sl, preset=0, mode=orc.MODE_PARITY, solver=orc.SOLVE_LDLT_F32)
    dt = time.perf_counter() - t
    return r["graph"].V, dt


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._halt = index, [], threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self._halt.wait(1.0)   # nvidia-smi queries perturb the GPU: sample sparsely

    def stop(self):
        self._halt.set()
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
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def run_reference(args, rank, world):
    if rank != 0:
        return
    w, h, xyz, rgba, label, _, _ = make_scene(args, 0)
    cw, ch, X, C, sx, sl = crop_scene(w, h, xyz, rgba, label)
    for _ in range(min(args.warmup, 1)):
        cpu_pipeline(X, C, sx, sl)
    t0 = time.perf_counter()
    V = 0
    for _ in range(args.steps):
        v, _dt = cpu_pipeline(X, C, sx, sl)
        V += v
    dt = time.perf_counter() - t0
    val = V / dt
    sample = "centre %dx%d px crop of the C4 scene (V=%d per step), whole pipeline incl. float32 sparse LDL^T" % (cw, ch, V // max(1, args.steps))
    print(json.dumps({
        "impl": "reference", "metric": "rw_segmentation_vertices_per_s", "value": val, "unit": "vertices/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {"value": val, "unit": "vertices/s", "cores": 1, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "vertices/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def workload_config(args, world):
    w, h = (1280, 960) if args.small else (C4_W, C4_H)
    return {"workload": "C4: synthetic %dx%d RGB-D cloud (~%.1f M finite points), %g m voxels, %d labels x %d seeds, "
                        "class weight preset, FAST mode, PCG tol 1e-6; one scene per GPU" %
                        (w, h, w * h * 0.78 / 1e6, C4_RES, C4_OBJECTS + 1, C4_SEEDS),
            "scenes_per_step": world, "l2": "inputs and solver vectors are larger than L2 (cloud 160 MB, each n x K block 182 MB)"}


def run_ours(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from tcs_b200 import capi
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    columns = args.shard == "columns" and world > 1
    w, h, xyz, rgba, label, sx, sl = make_scene(args, 0 if columns else rank)
    n = len(xyz)
    stream = torch.cuda.Stream()
    ctx = capi.Context(local_rank, stream=stream.cuda_stream)
    # host side: pinned interleaved records {x, y, z, rgba} (what a caller holding pcl::PointXYZRGB data passes)
    host = torch.empty((n, 4), dtype=torch.float32).pin_memory()
    host[:, :3] = torch.from_numpy(xyz)
    host[:, 3] = torch.from_numpy(rgba.view(np.float32))
    dev_cloud = host.to("cuda", non_blocking=False)
    host_np = host.numpy().view(np.uint8).reshape(n, 16)
    pc_host = torch.empty(n, dtype=torch.int32).pin_memory()
    vc_host = torch.empty(n, dtype=torch.int32).pin_memory()
    opts = capi.default_opts(time_spmm=1)
    if columns:     # ONE scene: this rank solves its block of label columns
        from tcs_b200 import shard
        opts.column_begin, opts.column_count = shard.columns_for_rank(len(np.unique(sl)), rank, world)
        if opts.column_count == 0:
            raise SystemExit("bench.py: more ranks than label columns")

    def merge(graph):
        """The shard's one exchange: all-gather of (best potential, colour) per vertex over NCCL, merged on device."""
        if columns:
            val, col = graph.best()
            shard._merge_best_nccl(val, col, world)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident():
        ctx.set_device_cloud(dev_cloud.data_ptr(), n, True)
        r = ctx.segment_cloud(C4_RES, sx, sl, preset=capi.PRESET_CLASS, mode=capi.MODE_FAST, opts=opts, want_host=False)
        merge(r["graph"])
        return r

    def step_e2e():
        import ctypes as C
        L = capi.lib()
        base = host_np.ctypes.data
        ctx.check(L.tcs_cloud_upload(ctx.h, C.c_void_p(base), 16, C.c_void_p(base + 12), 16, n))
        ctx.n_points_hint = n
        info = capi.SolveInfo()
        gh = C.c_void_p()
        sxx, sll = np.ascontiguousarray(sx, np.float32), np.ascontiguousarray(sl, np.uint32)
        st = L.tcs_segment_cloud(ctx.h, C4_RES, capi.PRESET_CLASS, capi.MODE_FAST, sxx.ctypes.data, sll.ctypes.data,
                                 len(sll), C.byref(opts), C.c_void_p(pc_host.data_ptr()),
                                 C.c_void_p(vc_host.data_ptr()), C.byref(info), C.byref(gh))
        ctx.check(st)
        g = capi.Graph(ctx, gh)
        merge(g)
        return dict(graph=g, info=info)

    # ---- device-resident arm
    last = None
    for _ in range(args.warmup):
        last = step_resident()
        last["graph"].close()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = ctx.launches
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    spmm_ms, spmm_n, V, iters = 0.0, 0, 0, 0
    infos = []
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for _ in range(args.steps):
            r = step_resident()
            info = r["info"].as_dict()
            infos.append(info)
            V += r["graph"].V
            spmm_ms += info["spmm_ms"]
            spmm_n += info["spmm_launches"]
            iters += info["iterations"]
            if _ < args.steps - 1:
                r["graph"].close()
        ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = ctx.launches - launches0
    clocks = sampler.stop()
    graph = r["graph"]
    info = infos[-1]
    E = graph.E
    converged = all(i["converged"] for i in infos)
    graph.close()

    # ---- end-to-end arm (host buffers, copies inside the timed region)
    r = step_e2e()
    r["graph"].close()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    Ve = 0
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(args.steps):
            r = step_e2e()
            Ve += r["graph"].V
            r["graph"].close()
        e1.record(stream)
    barrier()
    ms_e2e = e0.elapsed_time(e1)

    t = torch.tensor([ms, ms_e2e, -float(V), -float(Ve), float(launches)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)         # max time; V summed below
        tv = torch.tensor([float(V), float(Ve), float(launches)], dtype=torch.float64, device="cuda")
        dist.all_reduce(tv, op=dist.ReduceOp.SUM)
        V_all, Ve_all, launches_all = tv.tolist()
        if columns:     # one scene in total: its vertices count once
            V_all, Ve_all = float(V), float(Ve)
    else:
        V_all, Ve_all, launches_all = float(V), float(Ve), float(launches)
    ms_max, ms_e2e_max = t[0].item(), t[1].item()

    if rank == 0:
        nrows, nnz, K = info["rows"], info["nnz"], info["ncolours"]
        K_total = K
        if columns:
            K = opts.column_count       # the SpMM of this rank carries its own columns only
        b_spmm = 8 * nnz + 4 * (nrows + 1) + 8 * nrows * K          # SURVEY.md 8(d) / DESIGN.md
        avg_ms = spmm_ms / max(1, spmm_n)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        achieved = b_spmm / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
        kname = {1: "k_spmm", 2: "k_spmm_tiled", 3: "k_spmm_ring"}.get(info.get("spmm_kernel"), "k_spmm")
        traffic = None    # DRAM bytes per launch of that kernel from the committed ncu --set full capture
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "spmm_traffic.json")))
            if tr.get("kernel") == kname and not args.small and not columns:
                traffic = int(tr["dram_bytes_read"]) + int(tr["dram_bytes_write"])
        except Exception:
            pass
        out = {
            "metric": "rw_segmentation_vertices_per_s", "value": V_all / (ms_max * 1e-3), "unit": "vertices/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
            "higher_is_better": True, "scaling": "strong" if columns else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(args, 1 if columns else world), shard="label columns of one scene, one NCCL "
                           "all-gather per step" if columns else "independent scenes, no collective", vertices=int(V // args.steps), edges=int(E), rows=nrows, nnz=nnz,
                           labels=K_total, pcg_iterations_per_step=iters // args.steps, converged=converged,
                           max_rel_residual=max(info["rel_residual"]),
                           preconditioner="Jacobi" if os.environ.get("TCS_COARSE") == "0" else
                           "Jacobi + strong-edge aggregate coarse level (additive two-level)"),
            "e2e": {"value": Ve_all / (ms_e2e_max * 1e-3), "unit": "vertices/s", "h2d_bytes_per_step": 16 * n + 16 * len(sl),
                    "d2h_bytes_per_step": 4 * n + 4 * int(Ve // args.steps), "ms_per_step": ms_e2e_max / args.steps},
            "gpu_launches": int(launches_all),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": kname + " (CG SpMM over all label columns)", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak if peak else None, "traffic": traffic,
                         "algorithmic_bytes_per_launch": b_spmm, "avg_launch_ms": avg_ms,
                         "launches_per_step": spmm_n // args.steps, "spmm_share_of_step": spmm_ms / ms if ms else None,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s"},
            "stage_ms": {k: round(v, 3) for k, v in zip(
                ["build", "normals", "smoothen", "normals2", "signs", "weights", "components", "seeds", "assemble",
                 "solve", "argmax", "point_labels"], info["stage_ms"])},
        }
        if world == 1 and not args.no_cpu_baseline:
            cw, ch, X, C, csx, csl = crop_scene(w, h, xyz, rgba, label)
            v, dt = cpu_pipeline(X, C, csx, csl)
            out["cpu_baseline"] = {"value": v / dt, "unit": "vertices/s", "cores": 1, "kind": "port",
                                   "sample": "centre %dx%d px crop of the same scene (V=%d), whole pipeline incl. float32 "
                                             "sparse LDL^T, %.1f s" % (cw, ch, v, dt)}
        print(json.dumps(out), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
