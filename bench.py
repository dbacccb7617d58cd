#!/usr/bin/env python
"""bench.py -- FOF + unbinding throughput of the B200 structure-finding path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--nside 256]

A "step" is one pass of the hot path (SearchFullSet DM path: 3DFOF -> AdjustStructureForPeriod -> [6DFOF] ->
CheckUnboundGroups) over one synthetic clustered periodic box.  Default workload = BASELINE.json configs[1]:
3DFOF + direct-sum potential unbinding on a 256^3 box on one B200.

  value     Mparticles/s with the particles already resident in HBM (vrgpu_search_full_set_resident), timed with
            CUDA events on the library's stream, max over ranks.
  e2e       the same metric through the reference-facing call vrgpu_search_full_set on pinned HOST particles
            (96-byte NBody::Particle AoS): host->device unpack, the path, pfof + unwrapped positions back.
  roofline  dominant kernel group (by measured device time), algorithmic bytes or flops from DESIGN.md.
  cpu_baseline  the reference's own OpenMP SearchFullSet (oracle/_ref/libvrref_omp.so) on a bounded sample.

`--impl reference` times only the reference CPU implementation (all host threads) on bounded samples.
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


# ------------------------------------------------------------------------------------------------------
# reference / cpu baseline leg (oracle/_ref, test infrastructure -- never on the product path)
# ------------------------------------------------------------------------------------------------------
def reference_step(nside: int, fofbgtype: int, seed: int | None = None) -> dict:
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import vroracle
    import vrsynth
    kind = "reference"
    variant = "omp" if vroracle.Ref.available("omp") else ("serial" if vroracle.Ref.available("serial") else None)
    d = vrsynth.make_box(nside ** 3, seed if seed is not None else seed_for(nside))
    cores = os.cpu_count() or 1
    if variant is not None:
        if variant == "omp":
            os.environ.setdefault("OMP_NUM_THREADS", str(cores))
        ref = vroracle.Ref(variant)
        o = vroracle.default_opts(period=vrsynth.BOX, ellxscale=vrsynth.BOX / nside, fofbgtype=fofbgtype, bound_halos=1,
                                  omp_fof=1 if variant == "omp" else 0)
        r = ref.search_full_set(d["pos"], d["vel"], d["mass"], d["pid"], o)
        sec, ng = r["seconds"], r["ngroups"]
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
        ng = orc.unbind(pos, d["vel"], d["mass"], pf, ng, uo, d["pid"], 1e-6)["ngroups"]
        sec = time.perf_counter() - t0
    return {"seconds": sec, "n": nside ** 3, "ngroups": ng, "kind": kind, "cores": cores,
            "sample": f"{nside}^3 box from the same generator (seed {seed if seed is not None else seed_for(nside)}), the reference's own SearchFullSet "
                      f"({'OpenMP' if variant == 'omp' else 'serial'} build, -O2), whole call timed"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # bounded sample: ~1.5e-5 s per particle on a handful of cores -> pick the box so K+W steps stay within minutes
    total_steps = args.steps + args.warmup
    ns = args.cpu_nside or (128 if total_steps <= 3 else (96 if total_steps <= 8 else 64))
    for _ in range(args.warmup):
        reference_step(ns, args.fofbgtype)
    secs = []
    r = None
    for _ in range(args.steps):
        r = reference_step(ns, args.fofbgtype)
        secs.append(r["seconds"])
    t = float(np.sum(secs))
    value = args.steps * r["n"] / t / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.nside, args.fofbgtype, args.pp_threshold), "cpu_sample_nside": ns},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------
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
    parts = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"], pinned=True)
    del d
    stream = torch.cuda.ExternalStream(ctx.stream(), device=local)

    # ---- resident arm (`value`) ----
    ctx.particles_upload(parts)
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
    ctx.particles_release()

    # ---- e2e arm: reference-facing call on pinned host particles ----
    e2e_ms = []
    h2d = d2h = 0
    ctx.kernel_times_reset()
    e2e_warm = max(1, args.warmup)
    # the caller's result buffer: page-locked and reused (default), or a fresh pageable array per call as the adaptor's
    # `new Int_t[n]` (main.cxx:427 delete[]s it) -- the latter costs ~9 ms more at 256^3 (profiles/r01f)
    out_buf = vrgpu.pinned_i64(n) if args.e2e_out == "pinned" else None
    for it in range(e2e_warm + args.steps):   # untimed warm-up calls first, like the resident arm
        parts[:] = pristine                    # the call rewrites positions in place (AdjustStructureForPeriod): restore, untimed
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pf, ng_e = ctx.search_full_set(parts, so, out=out_buf)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) * 1e3
        if it >= e2e_warm:
            e2e_ms.append(dt)
            s = ctx.stats()
            h2d, d2h = s["h2d_bytes"], s["d2h_bytes"]
    e2e_value = n / (float(np.mean(e2e_ms)) * 1e-3) / 1e6
    e2e_io = {k: v["ms"] / (e2e_warm + args.steps) for k, v in ctx.kernel_times().items() if k.startswith("io.")}

    # ---- roofline of the dominant kernel group ----
    peaks, peaks_kind = measured_peaks()
    fp64_peak = ctx.measure_fp64_tflops()   # DFMA microbenchmark on this GPU (MEASURED_PEAKS.json has no fp64 figure)
    traffic_tab = {}
    tpath = os.path.join(ROOT, "profiles", "r01f_traffic.json")
    if os.path.exists(tpath) and nside == 256 and not ZOOM_NPART and args.fofbgtype == 5 and args.pp_threshold < 0:
        traffic_tab = json.load(open(tpath)).get("dram_bytes_per_launch", {})
    kernel_of = {"potential.direct": "k_potential_sym", "fof3d.link_cells": "k_link_thread"}
    groups = {k: v for k, v in ktimes.items() if k != "copy_pristine"}
    dom = max(groups, key=lambda k: groups[k]["ms"]) if groups else None
    roof_all = {}
    for k, v in groups.items():
        ms = v["ms"] / args.steps
        if ms <= 0:
            continue
        if k in ALG_BYTES:
            a = ALG_BYTES[k] * n / (ms * 1e-3) / 1e9
            roof_all[k] = {"bound": "hbm", "achieved": a, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": a / peaks["hbm_gbs"], "ms": ms}
        elif k == "potential.direct":
            a = FLOPS_PER_ORDERED_PAIR * st_last["n_pairs_potential"] / (ms * 1e-3) / 1e12
            roof_all[k] = {"bound": "fp64", "achieved": a, "peak": fp64_peak, "unit": "TFLOP/s", "frac": a / fp64_peak, "ms": ms,
                           "peak_kind": "measured DFMA microbenchmark (vendor vector figure 40)",
                           "note": "algorithmic 8 flops per ordered pair; the symmetric kernel issues 11.25 FP64-pipe instructions per unordered pair"}
        elif k == "ptree.walk" and st_last["n_tree_interactions"] > 0:
            # F_tree = 12 flops per accepted cell / visited leaf particle, counted by the kernel
            a = 12.0 * st_last["n_tree_interactions"] / (ms * 1e-3) / 1e12
            roof_all[k] = {"bound": "fp64", "achieved": a, "peak": fp64_peak, "unit": "TFLOP/s", "frac": a / fp64_peak, "ms": ms,
                           "interactions": int(st_last["n_tree_interactions"]), "tree_nodes": int(st_last["n_tree_nodes"]),
                           "ginteractions_per_s": st_last["n_tree_interactions"] / (ms * 1e-3) / 1e9,
                           "peak_kind": "measured DFMA microbenchmark (vendor vector figure 40)"}
        else:
            roof_all[k] = {"ms": ms}
    if dom == "potential.tree" and "ptree.walk" in roof_all:
        dom = "ptree.walk" if roof_all["ptree.walk"]["ms"] >= roof_all.get("ptree.build", {"ms": 0})["ms"] else "ptree.build"
    roofline = dict(roof_all.get(dom, {}))
    roofline.update({"kernel": dom, "traffic": traffic_tab.get(kernel_of.get(dom, ""), None),
                     "traffic_source": "profiles/r01f_summary.md (ncu --set full, per launch)" if traffic_tab else None,
                     "peak_kind": roofline.get("peak_kind", peaks_kind)})
    for k, name in kernel_of.items():
        if k in roof_all and name in traffic_tab:
            roof_all[k]["traffic"] = traffic_tab[name]

    # ---- cpu baseline (bounded sample) ----
    cpu = None
    if not args.no_cpu_baseline:
        r = reference_step(args.cpu_nside or 128, args.fofbgtype)
        cpu = {"value": r["n"] / r["seconds"] / 1e6, "unit": UNIT, "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
               "seconds": r["seconds"]}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": workload_name(nside, args.fofbgtype, args.pp_threshold), "n_particles": n, "ngroups": ng,
                       "l2": "inputs (1.07 GB of columns at 256^3) larger than the 126 MB L2; no explicit flush",
                       "e2e_buffers": f"particle records page-locked; pfof_out {'page-locked, reused' if args.e2e_out == 'pinned' else 'fresh pageable array per call'}",
                       "parallelism": "1 GPU"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": float(np.mean(e2e_ms)), "ngroups": ng_e, "io_ms_per_step": e2e_io},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "roofline_all": roof_all,
            "stage_ms_per_step": {k: v / args.steps for k, v in stage_ms.items()},
            "cpu_baseline": cpu}
    emit(line)
    ctx.close()


def run_ours_multi(args, rank, world, local):
    """N > 1: the SAME box, split into x-slabs (one per rank / GPU); ghost exchange + boundary merge + group consolidation
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
    if nside >= SLAB_GENERATOR_FROM:
        d = vrsynth.make_slab(n, seed_for(nside), lo, hi)
        parts = vrgpu.particles_from_arrays(d["pos"], d["vel"], d["mass"], d["pid"], pinned=True)
        n_loc = d["n"]
        del d
    else:
        d = vrsynth.make_box(n, seed_for(nside))
        x = d["pos"][:, 0]
        sel = np.flatnonzero((x >= lo) & (x < hi))
        parts = vrgpu.particles_from_arrays(d["pos"][sel], d["vel"][sel], d["mass"][sel], d["pid"][sel], pinned=True)
        n_loc = len(sel)
        del d, x
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
    ctx.particles_release()
    # ---- e2e arm: host particles of the slab in, global ids out ----
    e2e = []
    e2e_warm = max(1, args.warmup)
    out_buf = vrgpu.pinned_i64(parts.shape[0]) if args.e2e_out == "pinned" else None
    for it in range(e2e_warm + args.steps):
        barrier()
        t0 = time.perf_counter()
        pf, ng_e = ctx.dist_search_full_set(parts, lo, hi, so, out=out_buf)
        torch.cuda.synchronize()
        dt = torch.tensor([(time.perf_counter() - t0) * 1e3], device=f"cuda:{local}", dtype=torch.float64)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        if it >= e2e_warm:
            e2e.append(float(dt.item()))
    s = ctx.stats()
    tot = torch.tensor([float(launches), float(s["h2d_bytes"]), float(s["d2h_bytes"])], device=f"cuda:{local}", dtype=torch.float64)
    dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    if rank == 0:
        peaks, peaks_kind = measured_peaks()
        value = n / (ms_step * 1e-3) / 1e6
        kms = {k: v["ms"] / args.steps for k, v in ktimes.items()}
        link_ms = kms.get("fof3d.link_cells", 0.0)
        roof = {"kernel": "fof3d.link_cells", "bound": "hbm", "unit": "GB/s", "peak": peaks["hbm_gbs"], "peak_kind": peaks_kind, "traffic": None,
                "ms": link_ms, "note": "rank 0's slab (owned + ghosts)"}
        if link_ms > 0:
            roof["achieved"] = ALG_BYTES["fof3d.link_cells"] * n_loc / (link_ms * 1e-3) / 1e9
            roof["frac"] = roof["achieved"] / peaks["hbm_gbs"]
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": workload_name(nside, args.fofbgtype, args.pp_threshold), "n_particles": n, "ngroups": ng,
                           "l2": "inputs larger than the 126 MB L2; no explicit flush",
                           "e2e_buffers": f"particle records page-locked; pfof_out {'page-locked, reused' if args.e2e_out == 'pinned' else 'fresh pageable array per call'}",
                           "parallelism": f"{world} x-slabs, ghost exchange + boundary union-find + group consolidation over NCCL"},
                "e2e": {"value": n / (float(np.mean(e2e)) * 1e-3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(tot[1].item()),
                        "d2h_bytes_per_step": int(tot[2].item()), "ms_per_step": float(np.mean(e2e)), "ngroups": ng_e},
                "gpu_launches": int(tot[0].item()), "clocks": clocks, "roofline": roof, "kernel_ms_rank0": kms, "cpu_baseline": None}
        emit(line)
    ctx.close()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--nside", type=int, default=256)
    ap.add_argument("--fofbgtype", type=int, default=5, help="5 = 3DFOF, 3 = adaptive 6DFOF inside each halo")
    ap.add_argument("--pp-threshold", type=int, default=-1, help="<0: direct sum for every group (configs[1]); 150 = reference UNBINDNUM")
    ap.add_argument("--cpu-nside", type=int, default=0, help="box side of the bounded CPU sample (default 128)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-out", default="pinned", choices=["pinned", "pageable"],
                    help="result buffer of the end-to-end arm: reused page-locked array, or a fresh pageable one per call")
    ap.add_argument("--single-halo", type=int, default=0, metavar="NPART",
                    help="zoom-style input (configs[4]): ONE halo of NPART particles instead of the clustered box (1 GPU)")
    args = ap.parse_args()
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
