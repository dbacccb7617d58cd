#!/usr/bin/env python
"""bench.py -- FOF + unbinding throughput of the B200 structure-finding path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--nside 256]

A "step" is one pass of the hot path (SearchFullSet DM path: 3DFOF -> AdjustStructureForPeriod -> [6DFOF] ->
CheckUnboundGroups) over one synthetic clustered periodic box.  Default workload = BASELINE.json configs[2], the
configuration the metric is quoted on (1/2/4/8 B200) and the largest that fits one GPU: 3DFOF + 6DFOF + tree-potential
unbinding on a 512^3 box with the reference's stock thresholds.  `--config 2` = configs[1] (256^3, 3DFOF + direct-sum
unbinding); --nside / --fofbgtype / --pp-threshold override single settings.

  value     Mparticles/s with the particles already resident in HBM (vrgpu_search_full_set_resident), timed with
            CUDA events on the library's stream, max over ranks.
  e2e       the same metric through the reference-facing call vrgpu_search_full_set on HOST particles exactly as the
            adaptor hands them over (host/vrgpu_adaptor.cxx: pageable vector<Particle> records, a fresh pageable
            `new Int_t[n]` result): host->device unpack, the path, pfof + rewritten coordinates back.  The page-locked
            variant (caller pins its buffers) is the secondary key e2e_pinned.
  result    n_particles, ngroups, an order-independent 64-bit hash of the (pid, group id) membership and the number of
            marginal particles (|E| <= 1e-6 |phi|) the unbinding stage saw -- identical at every N.
  roofline  dominant kernel group (by measured device time), algorithmic bytes or flops from DESIGN.md.
  cpu_baseline  the reference's own OpenMP SearchFullSet (oracle/_ref/libvrref_omp.so) on a bounded sample.

`--impl reference` times only the reference CPU implementation (all host threads, OpenMP build) on a bounded sample of
the SAME workload: same generator, same stage set and thresholds, a 192^3 box (7.1e6 particles, above the 6e6 from which
the reference's OpenMP domain FOF engages, search.cxx:83-84).
Nothing here reads /root/reference; oracle/ is used only for the cpu_baseline / reference legs.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "velociraptor-stf_b200"))

import numpy as np  # noqa: E402

_REAL_STDOUT = None


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _REAL_STDOUT is not None:
        os.write(_REAL_STDOUT, data)
    else:
        sys.stdout.write(data.decode())
        sys.stdout.flush()


METRIC = "fof_unbind_throughput"
UNIT = "Mparticles/s"

# algorithmic work per particle / pair (DESIGN.md "Roofline accounting")
ALG_BYTES = {
    "fof3d.link_cells": 32.0,     # sorted position read once (24 B) + label written (8 B)
    "fof3d.radix_sort": 64.0,     # 4 digit passes x (8 B read + 8 B write) of (key32, index32)
    "fof3d.gather_pos": 48.0,     # 24 B read + 24 B write
    "fof3d.cell_keys": 28.0,      # 24 B read + 4 B key written
}
FLOPS_PER_ORDERED_PAIR = 8.0      # F_pp = 8 n (n-1) per group (16 flops per unordered pair, unbind.cxx:404-409)
FP64_PEAK_TFLOPS_NOMINAL = 40.0   # B200 vector FP64 (vendor figure; not in MEASURED_PEAKS.json)


def workload_name(nside, fofbgtype: int, pp: int) -> str:
    stages = "3DFOF" + ("+6DFOF" if fofbgtype <= 4 else "") + ("+direct-sum unbinding" if pp < 0 else "+tree/direct unbinding")
    if ZOOM_NPART:
        return (f"{stages} on a zoom-style single halo of {ZOOM_NPART} particles (NFW c=7 + 10% in subclumps, one 3D group, "
                f"L=100, b=0.2 of the mean spacing, seed 105)")
    gen = ", slab-wise generator" if nside >= SLAB_GENERATOR_FROM else ""
    return f"{stages} on a {nside}^3 synthetic clustered periodic DM box (L=100, b=0.2, seed {seed_for(nside)}{gen})"


ZOOM_NPART = 0   # --single-halo N: BASELINE.json configs[4] (set in main)
SLAB_GENERATOR_FROM = 512   # boxes from this side on come from vrsynth.make_slab: every rank draws only its own x-slab


def make_input(nside):
    import vrsynth
    if ZOOM_NPART:
        return vrsynth.make_box(ZOOM_NPART, 105, single_halo=True)
    if nside >= SLAB_GENERATOR_FROM:
        return vrsynth.make_slab(nside ** 3, seed_for(nside), 0.0, vrsynth.BOX)
    return vrsynth.make_box(nside ** 3, seed_for(nside))


def seed_for(nside: int) -> int:
    return {128: 101, 256: 102, 512: 103, 1024: 104}.get(nside, 100 + nside)


def measured_peaks() -> tuple[dict, str]:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi sampling DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.12)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm),
                       power_w_max=float(max(pw)))
        return out


def make_search_opts(vrgpu, nside: int, fofbgtype: int, pp: int):
    import vrsynth
    ell = vrsynth.BOX / nside
    param1 = (ell * ell) * (0.2 * 0.2) * (1.0 * 1.0)          # search.cxx:77
    ub = vrgpu.default_unbind_opts(pp_threshold=pp)
    return vrgpu.search_opts(vrsynth.BOX, param1, halo_min_size=20, fofbgtype=fofbgtype, bound_halos=1, unbind=ub)


def workload_config(args) -> dict:
    """The part of the JSON line that names the workload: identical in our arm and in the reference arm."""
    return {"workload": workload_name(args.nside, args.fofbgtype, args.pp_threshold),
            "stages": ["3DFOF", "AdjustStructureForPeriod"] + (["6DFOF (FoF_Field_search_type=%d)" % args.fofbgtype] if args.fofbgtype <= 4 else [])
                      + ["CheckUnboundGroups (Bound_halos=1)"],
            "thresholds": {"b": 0.2, "halo_min_size": 20, "ellhalo6dxfac": 1.0, "ellhalo6dvfac": 1.25,
                           "pp_threshold": args.pp_threshold, "large_threshold": 1000, "theta": 0.5, "bucket": 8,
                           "Eratio": 0.95, "minEfrac": 0.65, "unbindtype": 0},
            "l2": "inputs (GBs of columns) larger than the 126 MB L2; no explicit flush",
            "parallelism": "1 GPU" if args.gpus == 1 else f"{args.gpus} x-slabs, ghost exchange + boundary union-find + group placement over NCCL"}


_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def membership_hash(pid: np.ndarray, pfof: np.ndarray) -> int:
    """Order-independent 64-bit hash of the catalogue: sum over grouped particles of mix64(pid * K + gid) mod 2^64.  Group ids
    are canonical (size descending, smallest pid breaks ties), so equal hashes mean equal (pid -> group) maps."""
    tot = 0
    step = 1 << 24
    with np.errstate(over="ignore"):
        for a in range(0, pfof.shape[0], step):
            g = pfof[a:a + step]
            m = g > 0
            x = pid[a:a + step][m].astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15) + g[m].astype(np.uint64)
            x ^= x >> np.uint64(30); x *= np.uint64(0xBF58476D1CE4E5B9)
            x ^= x >> np.uint64(27); x *= np.uint64(0x94D049BB133111EB)
            x ^= x >> np.uint64(31)
            tot = (tot + int(np.add.reduce(x, dtype=np.uint64))) & 0xFFFFFFFFFFFFFFFF
    return tot


# ------------------------------------------------------------------------------------------------------
# reference / cpu baseline leg (oracle/_ref, test infrastructure -- never on the product path)
# ------------------------------------------------------------------------------------------------------
CPU_SAMPLE_NSIDE = 192   # 7.08e6 particles: ceil(N / 2e6) = 4 OpenMP regions, so OpenMPLocalSearch / OpenMPLinkAcross engage


def reference_step(nside: int, fofbgtype: int, seed: int | None = None, data: dict | None = None) -> dict:
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import vroracle
    import vrsynth
    kind = "reference"
    variant = "omp" if vroracle.Ref.available("omp") else ("serial" if vroracle.Ref.available("serial") else None)
    d = data if data is not None else vrsynth.make_box(nside ** 3, seed if seed is not None else seed_for(nside))
    cores = os.cpu_count() or 1
    if variant is not None:
        if variant == "omp":
            os.environ.setdefault("OMP_NUM_THREADS", str(cores))
        ref = vroracle.Ref(variant)
        o = vroracle.default_opts(period=vrsynth.BOX, ellxscale=vrsynth.BOX / nside, fofbgtype=fofbgtype, bound_halos=1,
                                  omp_fof=1 if variant == "omp" else 0)
        r = ref.search_full_set(d["pos"], d["vel"], d["mass"], d["pid"], o)
        sec, ng, pf = r["seconds"], r["ngroups"], r["pfof"]
        if variant != "omp":
            cores = 1
    else:  # compiled reference absent: our scalar C port
        kind, cores = "port", 1
        orc = vroracle.Oracle()
        t0 = time.perf_counter()
        b = 0.2 * vrsynth.BOX / nside
        pf, ng = orc.fof3d(d["pos"], b, vrsynth.BOX, 20, d["pid"])
        pos = orc.adjust_for_period(d["pos"], vrsynth.BOX, pf, ng)
        uo = vroracle.unbind_opts_from(vroracle.default_opts(), pp_threshold=150)
        u = orc.unbind(pos, d["vel"], d["mass"], pf, ng, uo, d["pid"], 1e-6)
        ng, pf = u["ngroups"], u["pfof"]
        sec = time.perf_counter() - t0
    return {"seconds": sec, "n": nside ** 3, "ngroups": ng, "kind": kind, "cores": cores, "n_grouped": int(np.count_nonzero(pf)),
            "sample": f"{nside}^3 box from the same generator (seed {seed if seed is not None else seed_for(nside)}, L=100, b=0.2), the reference's own "
                      f"SearchFullSet ({'OpenMP build, OMP_run_fof=1' if variant == 'omp' else 'serial build'}, -O2) with the same stage set "
                      f"(FoF_Field_search_type={fofbgtype}, Bound_halos=1) and the stock thresholds (UNBINDNUM 150, ompunbindnum 1000, theta 0.5, "
                      f"bucket 8); whole call timed, input generation and marshalling excluded"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import vrsynth
    ns = args.cpu_nside or CPU_SAMPLE_NSIDE
    d = vrsynth.make_box(ns ** 3, seed_for(ns))          # the driver copies the arrays into its own vector<Particle> every call
    # a CPU step of this sample takes tens of seconds: one untimed call is enough to fault the libraries and the arrays in
    warm = min(args.warmup, 1)
    for _ in range(warm):
        reference_step(ns, args.fofbgtype, data=d)
    secs = []
    r = None
    for _ in range(args.steps):
        r = reference_step(ns, args.fofbgtype, data=d)
        secs.append(r["seconds"])
    t = float(np.sum(secs))
    value = args.steps * r["n"] / t / 1e6
    cfg = workload_config(args)
    if args.pp_threshold < 0:
        cfg["reference_note"] = "the reference's UNBINDNUM is a compile-time 150: its arm uses kd-tree potentials above 150 members"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": warm, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
            "result": {"n_particles": r["n"], "ngroups": r["ngroups"], "n_grouped": r["n_grouped"]},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------
def traffic_table(args) -> tuple[dict, str | None]:
    """DRAM bytes per launch of the heaviest kernels from the committed `ncu --set full` capture of THIS configuration."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(path) or ZOOM_NPART:
        return {}, None
    tab = json.load(open(path))
    key = f"{args.nside}/{args.fofbgtype}/{args.pp_threshold}"
    e = tab.get(key)
    return (e.get("dram_bytes_per_launch", {}), e.get("source")) if e else ({}, None)


def e2e_arm(ctx, vrgpu, pristine, so, n, steps, warm, pinned: bool):
    """SearchFullSet through the C ABI on host records: pageable records + a fresh pageable result per call (what the
    adaptor passes), or page-locked records + a reused page-locked result."""
    import torch
    parts = vrgpu.particles_from_arrays(pristine["pos"], pristine["vel"], pristine["mass"], pristine["pid"], pinned=True) if pinned \
        else pristine.copy()
    out_buf = vrgpu.pinned_i64(n) if pinned else None
    ms, h2d, d2h, pf, ng = [], 0, 0, None, 0
    ctx.kernel_times_reset()
    for it in range(warm + steps):
        parts[:] = pristine                    # the call rewrites coordinates in place (AdjustStructureForPeriod, Q3): restore, untimed
        pf = None
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pf, ng = ctx.search_full_set(parts, so, out=out_buf)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) * 1e3
        if it >= warm:
            ms.append(dt)
            s = ctx.stats()
            h2d, d2h = s["h2d_bytes"], s["d2h_bytes"]
    io = {k: v["ms"] / (warm + steps) for k, v in ctx.kernel_times().items() if k.startswith("io.")}
    mean = float(np.mean(ms))
    return {"value": n / (mean * 1e-3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": mean,
            "ngroups": ng, "io_ms_per_step": io,
            "buffers": "page-locked particle records, page-locked reused pfof_out" if pinned
                       else "pageable particle records (vector<Particle>), fresh pageable pfof_out per call (new Int_t[n])"}, pf


def run_ours(args):
    import torch
    import vrgpu
    import vrsynth
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        return run_ours_multi(args, rank, world, local)

    nside = args.nside
    d = make_input(nside)
    n = d["n"]
    ctx = vrgpu.Context(local)
    so = make_search_opts(vrgpu, nside, args.fofbgtype, args.pp_threshold)
    pristine = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"])
    pid = np.ascontiguousarray(d["pid"], dtype=np.int64)
    del d
    stream = torch.cuda.ExternalStream(ctx.stream(), device=local)

    # ---- resident arm (`value`) ----
    ctx.particles_upload(pristine)
    ng = 0
    for _ in range(args.warmup):
        ng = ctx.search_full_set_resident(so)
    ctx.kernel_times_reset()
    sampler = ClockSampler(local)
    sampler.start()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    stage_ms = {}
    ev0.record(stream)
    for _ in range(args.steps):
        ng = ctx.search_full_set_resident(so)
        st = ctx.stats()
        launches += st["kernel_launches"]
        for k in ("sort_ms", "link_ms", "label_ms", "fof6d_ms", "potential_ms", "unbind_ms"):
            stage_ms[k] = stage_ms.get(k, 0.0) + st[k]
    ev1.record(stream)
    torch.cuda.synchronize()
    clocks = sampler.stop()
    ms_total = ev0.elapsed_time(ev1)
    ktimes = ctx.kernel_times()
    st_last = ctx.stats()
    ms_step = ms_total / args.steps
    value = n / (ms_step * 1e-3) / 1e6
    pf_res = ctx.resident_pfof(n)
    result = {"n_particles": n, "ngroups": ng, "n_grouped": int(np.count_nonzero(pf_res)), "membership_hash": "%016x" % membership_hash(pid, pf_res),
              "n_marginal": int(st_last["n_marginal"]), "n_groups3d": int(st_last["n_groups3d"])}
    del pf_res
    ctx.particles_release()

    # ---- e2e arm: the reference-facing call on host particles ----
    e2e, pf = e2e_arm(ctx, vrgpu, pristine, so, n, args.steps, max(1, args.warmup), pinned=(args.e2e_buffers == "pinned"))
    e2e["membership_hash"] = "%016x" % membership_hash(pid, pf)
    del pf
    e2e_other = None
    if not args.no_e2e_secondary:
        k2 = max(1, min(args.steps, 3))
        e2e_other, pf = e2e_arm(ctx, vrgpu, pristine, so, n, k2, 1, pinned=(args.e2e_buffers != "pinned"))
        e2e_other["steps"] = k2
        del pf

    # ---- roofline of the dominant kernel group ----
    peaks, peaks_kind = measured_peaks()
    fp64_peak = ctx.measure_fp64_tflops()   # DFMA microbenchmark on this GPU (MEASURED_PEAKS.json has no fp64 figure)
    traffic_tab, traffic_src = traffic_table(args)
    groups = {k: v for k, v in ktimes.items() if k != "copy_pristine"}
    # a timer that brackets other timers (a stage total) is not a kernel
    leaf = {k: v for k, v in groups.items() if k not in ("potential.tree", "potential.direct") and not k.endswith("_total")}
    dom = max(leaf, key=lambda k: leaf[k]["ms"]) if leaf else None
    roof_all = {}
    for k, v in groups.items():
        ms = v["ms"] / args.steps
        if ms <= 0:
            continue
        if k in ALG_BYTES:
            a = ALG_BYTES[k] * n / (ms * 1e-3) / 1e9
            roof_all[k] = {"bound": "hbm", "achieved": a, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": a / peaks["hbm_gbs"], "ms": ms}
        elif k in ("potential.direct", "potential.direct.sym") and args.pp_threshold < 0:
            a = FLOPS_PER_ORDERED_PAIR * st_last["n_pairs_potential"] / (ms * 1e-3) / 1e12
            roof_all[k] = {"bound": "fp64", "achieved": a, "peak": fp64_peak, "unit": "TFLOP/s", "frac": a / fp64_peak, "ms": ms,
                           "peak_kind": "measured DFMA microbenchmark (vendor vector figure 40)",
                           "note": "algorithmic 8 flops per ordered pair (all direct-sum groups of the step)"}
        elif k == "ptree.walk" and st_last["n_tree_interactions"] > 0:
            # F_tree = 12 flops per accepted cell / visited leaf particle, counted by the kernel
            a = 12.0 * st_last["n_tree_interactions"] / (ms * 1e-3) / 1e12
            roof_all[k] = {"bound": "fp64", "achieved": a, "peak": fp64_peak, "unit": "TFLOP/s", "frac": a / fp64_peak, "ms": ms,
                           "interactions": int(st_last["n_tree_interactions"]), "tree_nodes": int(st_last["n_tree_nodes"]),
                           "ginteractions_per_s": st_last["n_tree_interactions"] / (ms * 1e-3) / 1e9,
                           "peak_kind": "measured DFMA microbenchmark (vendor vector figure 40)"}
        else:
            roof_all[k] = {"ms": ms}
        if k in traffic_tab:
            roof_all[k]["traffic"] = traffic_tab[k]
    roofline = dict(roof_all.get(dom, {}))
    roofline.update({"kernel": dom, "traffic": traffic_tab.get(dom), "traffic_source": traffic_src if dom in traffic_tab else None,
                     "peak_kind": roofline.get("peak_kind", peaks_kind)})

    # ---- cpu baseline (bounded sample: the reference arm's box and settings) ----
    cpu = None
    if not args.no_cpu_baseline:
        r = reference_step(args.cpu_nside or CPU_SAMPLE_NSIDE, args.fofbgtype)
        cpu = {"value": r["n"] / r["seconds"] / 1e6, "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
               "seconds": r["seconds"]}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args), "result": result, "e2e": e2e,
            ("e2e_pinned" if args.e2e_buffers != "pinned" else "e2e_pageable"): e2e_other,
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "roofline_all": roof_all,
            "stage_ms_per_step": {k: v / args.steps for k, v in stage_ms.items()},
            "cpu_baseline": cpu}
    emit(line)
    ctx.close()


def run_ours_multi(args, rank, world, local):
    """N > 1: the SAME box, split into x-slabs (one per rank / GPU); ghost exchange + boundary merge + group placement
    over NCCL inside the library (vrgpu_dist_*).  Strong scaling: total work is the whole box whatever N."""
    import torch
    import torch.distributed as dist
    import vrgpu
    import vrsynth
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nside = args.nside
    n = nside ** 3
    L = vrsynth.BOX
    lo, hi = L * rank / world, L * (rank + 1) / world
    pinned = args.e2e_buffers == "pinned"
    if nside >= SLAB_GENERATOR_FROM:
        d = vrsynth.make_slab(n, seed_for(nside), lo, hi)
        sel = slice(None)
    else:
        d = vrsynth.make_box(n, seed_for(nside))
        x = d["pos"][:, 0]
        sel = np.flatnonzero((x >= lo) & (x < hi))
        del x
    parts = vrgpu.particles_from_arrays(d["pos"][sel], d["vel"][sel], d["mass"][sel], d["pid"][sel], pinned=pinned)
    pid = np.ascontiguousarray(d["pid"][sel], dtype=np.int64)
    n_loc = parts.shape[0]
    del d
    ctx = vrgpu.Context(local)
    uid = [vrgpu.dist_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ctx.dist_init(world, rank, uid[0])
    so = make_search_opts(vrgpu, nside, args.fofbgtype, args.pp_threshold)
    stream = torch.cuda.ExternalStream(ctx.stream(), device=local)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    # ---- resident arm ----
    ctx.particles_upload(parts)
    ng = 0
    for _ in range(args.warmup):
        ng = ctx.dist_search_full_set_resident(lo, hi, so)
    ctx.kernel_times_reset()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    ev0.record(stream)
    for _ in range(args.steps):
        ng = ctx.dist_search_full_set_resident(lo, hi, so)
        launches += ctx.stats()["kernel_launches"]
    ev1.record(stream)
    barrier()
    clocks = sampler.stop()
    t = torch.tensor([ev0.elapsed_time(ev1)], device=f"cuda:{local}", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    ktimes = ctx.kernel_times()
    st_last = ctx.stats()
    pf_res = ctx.resident_pfof(n_loc)
    loc = [membership_hash(pid, pf_res), int(np.count_nonzero(pf_res)), int(st_last["n_marginal"])]
    del pf_res
    ctx.particles_release()
    # ---- e2e arm: host particles of the slab in, global ids out ----
    e2e = []
    e2e_warm = max(1, args.warmup)
    out_buf = vrgpu.pinned_i64(n_loc) if pinned else None
    pf = None
    pristine = parts.copy()                    # pageable copy of the slab's records
    for it in range(e2e_warm + args.steps):
        pf = None
        parts[:] = pristine                    # the call rewrites coordinates in place (AdjustStructureForPeriod, Q3): restore, untimed
        barrier()
        t0 = time.perf_counter()
        pf, ng_e = ctx.dist_search_full_set(parts, lo, hi, so, out=out_buf)
        torch.cuda.synchronize()
        dt = torch.tensor([(time.perf_counter() - t0) * 1e3], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        if it >= e2e_warm:
            e2e.append(float(dt.item()))
    loc.append(membership_hash(pid, pf))
    s = ctx.stats()
    tot = torch.tensor([float(launches), float(s["h2d_bytes"]), float(s["d2h_bytes"])], device=f"cuda:{local}", dtype=torch.float64)
    dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    kms = {k: v["ms"] / args.steps for k, v in ktimes.items()}
    gathered = [None] * world
    dist.all_gather_object(gathered, (loc, kms))
    if rank == 0:
        peaks, peaks_kind = measured_peaks()
        value = n / (ms_step * 1e-3) / 1e6
        link_ms = kms.get("fof3d.link_cells", 0.0)
        roof = {"kernel": "fof3d.link_cells", "bound": "hbm", "unit": "GB/s", "peak": peaks["hbm_gbs"], "peak_kind": peaks_kind, "traffic": None,
                "ms": link_ms, "note": "rank 0's slab (owned + ghosts)"}
        if link_ms > 0:
            roof["achieved"] = ALG_BYTES["fof3d.link_cells"] * n_loc / (link_ms * 1e-3) / 1e9
            roof["frac"] = roof["achieved"] / peaks["hbm_gbs"]
        hres = sum(g[0][0] for g in gathered) & 0xFFFFFFFFFFFFFFFF
        he2e = sum(g[0][3] for g in gathered) & 0xFFFFFFFFFFFFFFFF
        # per-rank device time of the placed per-group stages: the imbalance the placement leaves
        heavy = ("potential.tree", "potential.direct", "fof6d.link", "fof6d.dispersion", "mg.consolidate_total", "mg.fof3d_total")
        per_rank = {k: [round(g[1].get(k, 0.0), 3) for g in gathered] for k in heavy if any(g[1].get(k, 0.0) > 0 for g in gathered)}
        # every timer: the slowest rank and its time (what the step waits for at the next collective)
        names = sorted({k for g in gathered for k in g[1]})
        slowest = {k: [int(np.argmax([g[1].get(k, 0.0) for g in gathered])), round(max(g[1].get(k, 0.0) for g in gathered), 3)] for k in names}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": workload_config(args),
                "result": {"n_particles": n, "ngroups": ng, "n_grouped": sum(g[0][1] for g in gathered), "membership_hash": "%016x" % hres,
                           "n_marginal": sum(g[0][2] for g in gathered)},
                "e2e": {"value": n / (float(np.mean(e2e)) * 1e-3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(tot[1].item()),
                        "d2h_bytes_per_step": int(tot[2].item()), "ms_per_step": float(np.mean(e2e)), "ngroups": ng_e,
                        "membership_hash": "%016x" % he2e,
                        "buffers": "page-locked records, reused page-locked pfof_out" if pinned else "pageable records, fresh pageable pfof_out per call"},
                "gpu_launches": int(tot[0].item()), "clocks": clocks, "roofline": roof, "kernel_ms_rank0": kms, "kernel_ms_per_rank": per_rank, "kernel_ms_slowest_rank": slowest,
                "cpu_baseline": None}
        emit(line)
    ctx.close()
    dist.destroy_process_group()


CONFIGS = {2: {"nside": 256, "fofbgtype": 5, "pp_threshold": -1},      # BASELINE.json configs[1]
           3: {"nside": 512, "fofbgtype": 3, "pp_threshold": 150}}     # BASELINE.json configs[2] (default)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=3, choices=sorted(CONFIGS), help="BASELINE.json configuration (1-based as in SURVEY 8: C2, C3)")
    ap.add_argument("--nside", type=int, default=0)
    ap.add_argument("--fofbgtype", type=int, default=0, help="5 = 3DFOF, 3 = adaptive 6DFOF inside each halo")
    ap.add_argument("--pp-threshold", type=int, default=None, help="<0: direct sum for every group (configs[1]); 150 = reference UNBINDNUM")
    ap.add_argument("--cpu-nside", type=int, default=0, help=f"box side of the bounded CPU sample (default {CPU_SAMPLE_NSIDE})")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-buffers", default="pageable", choices=["pageable", "pinned"],
                    help="host buffers of the headline end-to-end arm: pageable records + fresh pageable result (the adaptor's contract), or page-locked")
    ap.add_argument("--no-e2e-secondary", action="store_true", help="skip the short run of the other buffer kind")
    ap.add_argument("--single-halo", type=int, default=0, metavar="NPART",
                    help="zoom-style input (configs[4]): ONE halo of NPART particles instead of the clustered box (1 GPU)")
    args = ap.parse_args()
    c = CONFIGS[args.config]
    args.nside = args.nside or c["nside"]
    args.fofbgtype = args.fofbgtype or c["fofbgtype"]
    args.pp_threshold = c["pp_threshold"] if args.pp_threshold is None else args.pp_threshold
    if args.single_halo:
        global ZOOM_NPART
        ZOOM_NPART = args.single_halo
        args.nside = float(args.single_halo) ** (1.0 / 3.0)   # only its mean spacing L/nside enters (the linking length)
    # exactly ONE line on stdout: libraries (NCCL prints its version banner) write to fd 1 too, so everything except the
    # final JSON goes to stderr
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
