#!/bin/bash
# First GPU call of the next measuring session: parity of the experimental rotation-step variant of k_potential_sym
# (VRGPU_SYM_VARIANT=1, csrc/potential.cu) and its time against the measured default, plus the tests that were written
# after round 1's GPU minutes ran out.  Also: the candidate-queue 3D linker (VRGPU_LINK_VARIANT=1|2, csrc/link2.cuh).  Usage: gpurun --timeout 900 -- 'bash tools/probe_sym_variant.sh'
set -u
mkdir -p gpurun_out
{
  echo "== fp64 pipes: does DMMA (mma.sync m8n8k4 f64) run beside the vector DFMA pipe?  (TFLOP/s)"
  timeout 120 python -c "import sys; sys.path.insert(0, 'velociraptor-stf_b200'); import vrgpu; print(vrgpu.Context(0).probe_fp64_pipes())" 2>&1 | tail -1
  echo "== gated tests (first GPU run)"
  VRGPU_RUN_UNVERIFIED=1 timeout 600 python -m pytest tests/test_gpu_energy.py tests/test_gpu_unbind.py tests/test_gpu_zoom.py tests/test_gpu_fof3d.py -q -m gpu 2>&1 | tail -5
  for v in 0 1 2; do
    echo "== VRGPU_SYM_VARIANT=$v: parity"
    VRGPU_SYM_VARIANT=$v timeout 600 python -m pytest tests/test_gpu_unbind.py tests/test_gpu_zoom.py tests/test_gpu_dropin.py -q -m gpu 2>&1 | tail -3
    echo "== VRGPU_SYM_VARIANT=$v: bench (default config)"
    VRGPU_SYM_VARIANT=$v timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/probe_sym_v$v.json
    python tools/bench_summary.py gpurun_out/probe_sym_v$v.json | head -8
  done
  for w in 0 1; do
    echo "== VRGPU_WALK_PREFETCH=$w: tree parity + bench (256^3, 6DFOF + tree unbinding)"
    VRGPU_WALK_PREFETCH=$w timeout 600 python -m pytest tests/test_gpu_tree.py -q -m gpu 2>&1 | tail -2
    VRGPU_WALK_PREFETCH=$w timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --fofbgtype 3 --pp-threshold 150 2>/dev/null | tail -1 > gpurun_out/probe_walk_p$w.json
    python tools/bench_summary.py gpurun_out/probe_walk_p$w.json | head -8
  done
  for v in 0 1 2; do
    echo "== VRGPU_LINK_VARIANT=$v: 3DFOF parity (single GPU, drop-in, 2 ranks) + bench (default config; look at fof3d.link_cells)"
    VRGPU_LINK_VARIANT=$v timeout 600 python -m pytest tests/test_gpu_fof3d.py tests/test_gpu_fof6d.py tests/test_gpu_dropin.py tests/test_gpu_multi.py -q -m gpu 2>&1 | tail -3
    VRGPU_LINK_VARIANT=$v timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/probe_link_v$v.json
    python tools/bench_summary.py gpurun_out/probe_link_v$v.json | head -12
  done
  for m in 0 262144; do
    echo "== VRGPU_DISP_MULTI_MIN=$m: zoom-style single halo of 2e7 particles (look at fof6d.dispersion)"
    VRGPU_DISP_MULTI_MIN=$m timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --single-halo 20000000 --fofbgtype 3 --pp-threshold 150 2>/dev/null | tail -1 > gpurun_out/probe_disp_m$m.json
    python tools/bench_summary.py gpurun_out/probe_disp_m$m.json | head -8
  done
  for a in 0 1; do
    echo "== VRGPU_ASYNC_WRITEBACK=$a: 6DFOF parity incl. pinned write-back + e2e of 256^3 6DFOF + tree unbinding (look at e2e and io.writeback)"
    VRGPU_ASYNC_WRITEBACK=$a timeout 600 python -m pytest tests/test_gpu_fof6d.py tests/test_gpu_zoom.py tests/test_gpu_dropin.py -q -m gpu 2>&1 | tail -2
    VRGPU_ASYNC_WRITEBACK=$a timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --fofbgtype 3 --pp-threshold 150 2>/dev/null | tail -1 > gpurun_out/probe_wb_a$a.json
    python tools/bench_summary.py gpurun_out/probe_wb_a$a.json | head -4
  done
} > gpurun_out/probe_sym_variant.log 2>&1
tail -40 gpurun_out/probe_sym_variant.log
