#!/usr/bin/env python
"""bench.py - throughput of the polygonization hot path (BlobTree -> cull -> field -> marching cubes).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one full polygonization of the workload scene. Metric (BASELINE.json): MC cells/s = grid cells of
the job / device time of the step (field evals/s, triangles/s and ms per polygonization ride along).

  N = 1     BASELINE configs[1]: synthetic 1k point/line/cylinder blob scene at 512^3 cells
  N = 8     BASELINE configs[3]: 10k-primitive blob scene at 1024^3, MPU lattice sharded in z-slabs
  N = 2, 4  the same family in between (1250*N primitives, extent 13*N^(1/3), 512*N^(1/3) cells): weak scaling

GPU arm: the scene is linearized on the host, uploaded once, and every step runs psgpu_polygonize (device
time from CUDA events on the library's own stream, `value`) - then, for `e2e`, re-linearize-free upload of the
program from pinned host memory + psgpu_polygonize_host into pinned host buffers, timed by wall clock.
Both arms run the SAME semantics (PS_SimdPoly: infinite lines, inside = f >= iso, 4-sample + lerp roots), so
the meshes are comparable; `native` adds the numbers of the canonical mode (segments, bracketed root search).

Reference arm (--impl reference): the reference's own TBB/SIMD polygonizer (oracle/_ref, built from
/root/reference sources by oracle/Makefile; the C restatement if that build is absent) on all host cores, each
step a bounded z-slab sample of the same lattice.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def workload(n_gpus: int) -> dict:
    if n_gpus <= 1:
        return dict(name="blob1k@512^3 (BASELINE configs[1])", prims=1000, extent=12.0, seed=2, cells=512)
    if n_gpus == 8:
        return dict(name="blob10k@1024^3 z-slab sharded (BASELINE configs[3])", prims=10000, extent=26.0, seed=4, cells=1024)
    s = n_gpus ** (1.0 / 3.0)
    return dict(name="blob%dk@%d^3 z-slab sharded (between configs[1] and configs[3])" % (1250 * n_gpus // 1000, round(512 * s)),
                prims=1250 * n_gpus, extent=13.0 * s, seed=4, cells=int(round(512 * s)))


def build_scene(w):
    from parsip_b200 import blobtree as bt, scenes
    root = scenes.blobs(w["prims"], w["extent"], w["seed"])
    return root, bt


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)), "measured"
        except Exception:
            pass
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx = float(r[2])
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------------
# CPU legs (the only places that execute oracle/)
# ----------------------------------------------------------------------------------------------------------
def cpu_polygonizer(root, bt, n_prims):
    """The reference's own SIMD polygonizer when its build is present (kind 'reference'), else the C restatement."""
    from oracle import pyoracle
    soa = bt.to_binary_soa(bt.binarize(root))
    n = max(soa["n_prims"], soa["n_ops"])
    if pyoracle.ref_available(n):
        sc = pyoracle.RefScene(soa)
        sc.prepare_bboxes(0.05)
        return "reference", sc, (lambda cs, b, e: sc.polygonize_stream(cs, b, e, keep_mesh=False)[1])
    sc = pyoracle.SimdOracle(soa)
    sc.prepare_bboxes(0.05)
    return "port", sc, (lambda cs, b, e: sc.polygonize(cs, 0, b, e, keep_mesh=False)[1])


def cpu_sample(run, cs, dims, budget_s):
    """Times a z-slab sample of the lattice sized for about budget_s seconds. Returns (seconds, mpu layers, stats)."""
    nz = dims[2]
    mid = nz // 2
    t = time.perf_counter()
    st = run(cs, [0, 0, mid], [dims[0], dims[1], mid + 1])
    t1 = time.perf_counter() - t
    layers = int(max(1, min(nz, budget_s / max(t1, 1e-6))))
    if layers == 1:
        return t1, 1, st
    b = max(0, mid - layers // 2)
    e = min(nz, b + layers)
    t = time.perf_counter()
    st = run(cs, [0, 0, b], [dims[0], dims[1], e])
    return time.perf_counter() - t, e - b, st


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    w = workload(args.gpus)
    root, bt = build_scene(w)
    kind, sc, run = cpu_polygonizer(root, bt, w["prims"])
    from parsip_b200 import scenes
    lo, hi = sc.bbox()
    cs = scenes.cellsize_for(lo, hi, w["cells"])
    dims = [int(math.ceil(math.ceil(float(hi[a] - lo[a]) / cs) / 7.0)) for a in range(3)]
    cores = os.cpu_count() or 1
    per_step = max(1.0, min(20.0, 150.0 / max(1, args.steps + args.warmup)))
    # calibrate the sample once, then time K steps of it
    t0, layers, st = cpu_sample(run, cs, dims, per_step)
    mid = dims[2] // 2
    b = max(0, mid - layers // 2)
    e = min(dims[2], b + layers)
    for _ in range(max(0, args.warmup - 1)):
        run(cs, [0, 0, b], [dims[0], dims[1], e])
    times = []
    for _ in range(args.steps):
        t = time.perf_counter()
        st = run(cs, [0, 0, b], [dims[0], dims[1], e])
        times.append(time.perf_counter() - t)
    sec = float(np.sum(times))
    cells = float(dims[0] * dims[1] * (e - b)) * 343.0 * args.steps
    value = cells / sec
    sample = "z-slab of %d/%d MPU layers (%d MPUs, %d with surface) of %s" % (e - b, dims[2], st["mpus"], st["crossed"], w["name"])
    out = {
        "impl": "reference", "metric": "MC cells/s", "value": value, "unit": "cells/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sec / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": w["name"], "semantics": "PS_SimdPoly (infinite lines, f>=iso, 4-sample roots)", "cells": w["cells"],
                   "primitives": w["prims"], "cellsize": cs, "lattice_mpus": dims},
        "field_evals_per_s": st["simd_evals"] * 4.0 * args.steps / sec,
        "triangles_per_s": st["triangles"] * args.steps / sec,
        "cpu_baseline": {"value": value, "unit": "cells/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out))
    return 0


# ----------------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------------
def run_gpu(args):
    import torch
    import torch.distributed as dist

    import parsip_b200 as ps
    from parsip_b200 import scenes, sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("WORLD_SIZE %d != --gpus %d" % (world, args.gpus))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    w = workload(world)
    root, bt = build_scene(w)
    flat = bt.flatten(bt.binarize(root))
    lo, hi = ps.prepare_bboxes(flat)
    cs = scenes.cellsize_for(lo, hi, w["cells"])
    dims = ps.lattice_dims(cs, lo, hi)
    zb, ze = sharding.slab_bounds(dims[2], world, rank)
    ctx = ps.Context(local)
    prog = ps.linearize(flat, ps.SEM_SIMD, 0)
    ctx.upload(prog)
    mk = lambda **kw: ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_FOURSAMPLE, coord_mode=ps.COORD_GLOBAL, output=ps.OUT_INDEXED,
                                     cull=ps.CULL_INTERVAL, z_begin=zb, z_end=ze, **kw)
    prm = mk()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def exchange(nv, nt):
        """per-rank vertex / triangle counts -> global offsets (NCCL all-gather over NVLink when world > 1)"""
        if world == 1:
            return (0, 0), (nv, nt)
        return sharding.exchange_counts(nv, nt, torch.device("cuda", local))

    # ---- device-resident steps: `value`
    for _ in range(args.warmup):
        m = ctx.polygonize(prm, download=False)
        exchange(m["n_vertices"], m["n_triangles"])
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    dev_ms, launches, stats_sum = 0.0, 0, None
    field_ms = emit_ms = cull_ms = scan_ms = tri_ms = vertex_ms = color_ms = 0.0
    for _ in range(args.steps):
        flush.fill_(1)  # L2 flush between steps (untimed)
        torch.cuda.synchronize()
        t = time.perf_counter()
        m = ctx.polygonize(prm, download=False)
        offs, tot = exchange(m["n_vertices"], m["n_triangles"])
        wall = (time.perf_counter() - t) * 1e3
        st = ctx.stats()
        # device time of the step = CUDA events on the library's launching stream (cull .. emit); the count exchange
        # (16 B per rank) is added from its wall time when sharded
        dev_ms += st["ms_total"] + (max(0.0, wall - st["ms_total"]) if world > 1 else 0.0)
        field_ms += st["ms_field"]; emit_ms += st["ms_emit"]; cull_ms += st["ms_cull"]; scan_ms += st["ms_scan"]
        tri_ms += st["ms_tri"]; vertex_ms += st["ms_vertex"]; color_ms += st["ms_color"]
        launches += st["gpu_launches"]
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    stc = ctx.polygonize(mk(collect_stats=1), download=False)
    st = ctx.stats()

    # ---- end to end through the C ABI with HOST buffers: upload (pinned) + polygonize + mesh download (pinned)
    nv, nt = m["n_vertices"], m["n_triangles"]
    pin = lambda n: torch.empty((max(n, 1), 3), dtype=torch.float32, pin_memory=True).numpy()
    pos, nrm, col = pin(nv), pin(nv), pin(nv)
    tri = torch.empty((max(nt, 1), 3), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
    ctx.upload(prog); ctx.polygonize_host(prm, pos, nrm, col, tri)
    barrier()
    e2e_s = 0.0
    for _ in range(args.steps):
        flush.fill_(1)
        torch.cuda.synchronize()
        t = time.perf_counter()
        ctx.upload(prog)
        mh = ctx.polygonize_host(prm, pos, nrm, col, tri)
        exchange(mh["n_vertices"], mh["n_triangles"])
        e2e_s += time.perf_counter() - t
    barrier()

    # ---- canonical native mode (segments, bracketed root search), for information
    progN = ps.linearize(bt.flatten(root), ps.SEM_SCALAR, ps.LIN_REASSOCIATE)
    ctx.upload(progN)
    prmN = ps.Params.make(cs, lo, hi, z_begin=zb, z_end=ze)
    ctx.polygonize(prmN, download=False)
    nat_ms = 0.0
    for _ in range(max(1, min(args.steps, 5))):
        mn = ctx.polygonize(prmN, download=False)
        nat_ms += ctx.stats()["ms_total"]
    nat_ms /= max(1, min(args.steps, 5))
    stn = ctx.stats()

    # ---- reduce over ranks: max time, summed work
    def allmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); return float(t[0])

    def allsum(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.SUM); return float(t[0])

    dev_ms_max = allmax(dev_ms)
    e2e_s_max = allmax(e2e_s)
    cells_total = float(dims[0] * dims[1] * dims[2]) * 343.0  # grid cells of the whole job
    evals = allsum(float(st["field_evals"]))
    tris = allsum(float(nt)); verts = allsum(float(nv))
    launches_all = allsum(float(launches))
    h2d = prog.n_quads * 16 + prog.n_prims * 12
    d2h = nv * 36 + nt * 12
    flops = float(st["algorithmic_flops"])
    nat_ms_max = allmax(nat_ms)
    if world > 1:
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peaks, peak_src = read_peaks()
    fp32_tflops, _ = ctx.fp32_peak()
    sec = dev_ms_max * 1e-3
    value = cells_total * args.steps / sec
    # ---- rooflines, rank 0's launches (DESIGN.md section 5 states the per-unit figures):
    #   field_kernel / vertex_kernel / color_kernel  FP32 pipe: SURVEY 8(d) op counts of every MPU's pruned program x points evaluated
    #   tri_kernel                                   HBM: per surface MPU 2 KB field cache + 288 B edge masks read, 12 B per triangle written
    K = float(args.steps)
    traffic = {}
    tp = os.path.join(ROOT, "profiles", "traffic.json")  # dram__bytes_read+write per launch from the committed ncu --set full capture
    if os.path.exists(tp) and world == 1:
        try:
            traffic = json.load(open(tp)).get(w["name"], {})
        except Exception:
            traffic = {}
    fp32_src = "FFMA-chain microbenchmark run in this process (MEASURED_PEAKS.json has no fp32 figure)"

    def roof_fp32(name, fl, ms):
        ach = fl / max(ms, 1e-9) / 1e9
        return {"bound": "fp32", "kernel": name, "achieved": ach, "peak": fp32_tflops, "unit": "TFLOP/s", "frac": ach / fp32_tflops,
                "traffic": traffic.get(name), "peak_source": fp32_src, "algorithmic_flops_per_launch": fl, "ms_per_launch": ms}

    def roof_hbm(name, by, ms):
        ach = by / max(ms, 1e-9) / 1e6
        return {"bound": "hbm", "kernel": name, "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach / peaks["hbm_gbs"],
                "traffic": traffic.get(name), "peak_source": peak_src, "algorithmic_bytes_per_launch": by, "ms_per_launch": ms}

    vflops = float(st["vertex_eval_flops"])
    tri_bytes = float(st["mpus_crossed"]) * (2048 + 288 + 8) + nt * 12.0
    roofs = [roof_fp32("field_kernel", flops, field_ms / K), roof_fp32("vertex_kernel", 8.0 * vflops, vertex_ms / K),
             roof_fp32("color_kernel", vflops, color_ms / K), roof_hbm("tri_kernel", tri_bytes, tri_ms / K)]
    roofs.sort(key=lambda r: -r["ms_per_launch"])
    dominant, other = roofs[0], roofs[1:]
    field_ms_r0, emit_ms_r0 = field_ms / K, emit_ms / K

    out = {
        "metric": "MC cells/s", "value": value, "unit": "cells/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": w["name"], "semantics": "PS_SimdPoly (infinite lines, f>=iso, 4-sample roots)", "cells": w["cells"],
                   "primitives": w["prims"], "cellsize": cs, "lattice_mpus": list(dims), "sharding": "z-slabs of MPU layers, %d ranks" % world,
                   "l2": "flushed between steps (256 MiB fill, untimed); a step's field caches + mesh also exceed L2",
                   "cull": "interval", "output": "indexed mesh (global edge-id dedup)"},
        "field_evals_per_s": evals * args.steps / sec / args.steps * 1.0 / 1.0,
        "triangles_per_s": tris * args.steps / sec / args.steps * 1.0,
        "ms_per_polygonization": dev_ms_max / args.steps,
        "mesh": {"vertices": int(verts), "triangles": int(tris)},
        "stage_ms_rank0": {"cull": cull_ms / args.steps, "field": field_ms_r0, "scan": scan_ms / args.steps, "emit": emit_ms_r0,
                           "tri": tri_ms / K, "vertex": vertex_ms / K, "color": color_ms / K},
        "mpus_rank0": {"total": st["mpus_total"], "evaluated": st["mpus_evaluated"], "crossed": st["mpus_crossed"]},
        "e2e": {"value": cells_total * args.steps / e2e_s_max, "unit": "cells/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": 1e3 * e2e_s_max / args.steps},
        "gpu_launches": int(launches_all),
        "roofline": dominant, "roofline_other": other,
        "native": {"semantics": "CompactBlobTree (segments, f>iso), bracketed false-position roots", "ms_per_polygonization": nat_ms_max,
                   "cells_per_s": cells_total / (nat_ms_max * 1e-3), "vertices": mn["n_vertices"], "triangles": mn["n_triangles"],
                   "field_evals": stn["field_evals"]},
        "clocks": clocks,
    }
    out["field_evals_per_s"] = evals / (dev_ms_max / args.steps * 1e-3)
    out["triangles_per_s"] = tris / (dev_ms_max / args.steps * 1e-3)

    # ---- CPU baseline beside it (N = 1 only): bounded sample of the same workload on the host cores
    if world == 1 and not args.no_cpu_baseline:
        kind, sc, run = cpu_polygonizer(root, bt, w["prims"])
        rlo, rhi = sc.bbox()
        rcs = scenes.cellsize_for(rlo, rhi, w["cells"])
        rdims = [int(math.ceil(math.ceil(float(rhi[a] - rlo[a]) / rcs) / 7.0)) for a in range(3)]
        tsec, layers, rst = cpu_sample(run, rcs, rdims, 12.0)
        out["cpu_baseline"] = {"value": rdims[0] * rdims[1] * layers * 343.0 / tsec, "unit": "cells/s", "cores": os.cpu_count() or 1, "kind": kind,
                               "sample": "z-slab of %d/%d MPU layers (%d MPUs, %d with surface), %.1f s" % (layers, rdims[2], rst["mpus"], rst["crossed"], tsec),
                               "field_evals_per_s": rst["simd_evals"] * 4.0 / tsec, "triangles_per_s": rst["triangles"] / tsec}
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gpu", choices=["gpu", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "gpu" else max(args.warmup, 1)
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
