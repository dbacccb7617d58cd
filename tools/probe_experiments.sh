#!/bin/bash
# First GPU call of the next measuring session: every experiment that was prepared without a GPU (each off by default), its
# parity and its time against the measured default, plus the tests written after round 1's GPU minutes ran out.
# Usage: gpurun --timeout 2400 -- 'bash tools/probe_experiments.sh'            (everything, about half an hour)
#        gpurun --timeout 900  -- 'bash tools/probe_experiments.sh pipes gated sym'   (chosen sections)
# Sections: pipes gated sym walk link disp wb ties veldens.  Log: gpurun_out/probe_experiments.log, bench lines: gpurun_out/probe_*.json
set -u
mkdir -p gpurun_out
SECTIONS="${*:-pipes gated ties veldens sym walk link disp wb}"
bench() { # bench <out.json> <args...>: one bench line, CPU baseline skipped
  local out=$1; shift
  timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline "$@" 2>/dev/null | tail -1 > "$out"
  python tools/bench_summary.py "$out" | head -10
}
pipes() {
  echo "== fp64 pipes: does DMMA (mma.sync m8n8k4 f64) run beside the vector DFMA pipe?  (TFLOP/s)"
  timeout 120 python -c "import sys; sys.path.insert(0, 'velociraptor-stf_b200'); import vrgpu; print(vrgpu.Context(0).probe_fp64_pipes())" 2>&1 | tail -1
}
gated() {
  echo "== gated tests (first GPU run): binding-energy sort adaptor, pglist, Keep_background_potential=0, 3DFOF edge cases, many-window dispersion, exact tree ties, local velocity density"
  VRGPU_RUN_UNVERIFIED=1 timeout 1200 python -m pytest tests -q -m gpu --tb=short -p no:cacheprovider > gpurun_out/gated_pytest.log 2>&1
  tail -40 gpurun_out/gated_pytest.log
}
sym() {
  for v in 0 1 2; do
    echo "== VRGPU_SYM_VARIANT=$v: parity, then bench (default config; look at potential.direct.sym)"
    VRGPU_SYM_VARIANT=$v timeout 600 python -m pytest tests/test_gpu_unbind.py tests/test_gpu_zoom.py tests/test_gpu_dropin.py -q -m gpu 2>&1 | tail -3
    VRGPU_SYM_VARIANT=$v bench gpurun_out/probe_sym_v$v.json
  done
}
walk() {
  for w in 0 1; do
    echo "== VRGPU_WALK_PREFETCH=$w: tree parity + bench (256^3, 6DFOF + tree unbinding; look at ptree.walk)"
    VRGPU_WALK_PREFETCH=$w timeout 600 python -m pytest tests/test_gpu_tree.py -q -m gpu 2>&1 | tail -2
    VRGPU_WALK_PREFETCH=$w bench gpurun_out/probe_walk_p$w.json --fofbgtype 3 --pp-threshold 150
  done
}
link() {
  for v in 0 1 2; do
    echo "== VRGPU_LINK_VARIANT=$v: 3DFOF parity (single GPU, drop-in, 2 ranks) + bench (default config; look at fof3d.link_cells)"
    VRGPU_LINK_VARIANT=$v VRGPU_RUN_UNVERIFIED=1 timeout 600 python -m pytest tests/test_gpu_fof3d.py tests/test_gpu_fof6d.py tests/test_gpu_dropin.py tests/test_gpu_multi.py -q -m gpu 2>&1 | tail -3
    VRGPU_LINK_VARIANT=$v bench gpurun_out/probe_link_v$v.json
  done
}
disp() {
  for m in 0 262144; do
    echo "== VRGPU_DISP_MULTI_MIN=$m: zoom-style single halo of 2e7 particles (look at fof6d.dispersion)"
    VRGPU_DISP_MULTI_MIN=$m bench gpurun_out/probe_disp_m$m.json --single-halo 20000000 --fofbgtype 3 --pp-threshold 150
  done
}
wb() {
  for a in 0 1; do
    echo "== VRGPU_ASYNC_WRITEBACK=$a: 6DFOF parity incl. pinned write-back + 256^3 6DFOF + tree unbinding (look at e2e and io.writeback)"
    VRGPU_ASYNC_WRITEBACK=$a timeout 600 python -m pytest tests/test_gpu_fof6d.py tests/test_gpu_zoom.py tests/test_gpu_dropin.py -q -m gpu 2>&1 | tail -2
    VRGPU_ASYNC_WRITEBACK=$a bench gpurun_out/probe_wb_a$a.json --fofbgtype 3 --pp-threshold 150
  done
}
veldens() {
  echo "== row f4: vrgpu_velocity_density on one structure of 1e6 particles (per-leaf search), with the reference's OpenMP build on a sample"
  timeout 900 python tools/bench_veldens.py 1000000 3 --cpu 2>&1 | tail -1 | tee gpurun_out/probe_veldens.json
  echo "== the same with the sort-based tree (separates a failure of the search from one of the exact-ties build; densities then agree to ~1e-9, not 1e-12)"
  VRGPU_VD_SORT_TREE=1 timeout 900 python tools/bench_veldens.py 200000 1 --cpu 2>&1 | tail -1
}
ties() {
  echo "== VRGPU_TREE_TIES=1: the reference's quickselect arrangement (qsel.cuh): tied-coordinate parity at 1e-10, then the whole tree + unbinding suites under it"
  VRGPU_TREE_TIES=1 VRGPU_RUN_UNVERIFIED=1 timeout 900 python -m pytest tests/test_gpu_tree.py tests/test_gpu_unbind.py tests/test_gpu_zoom.py tests/test_gpu_dropin.py tests/test_gpu_energy.py -q -m gpu --tb=short -p no:cacheprovider > gpurun_out/ties_pytest.log 2>&1
  tail -30 gpurun_out/ties_pytest.log
  for t in 0 1; do
    echo "== VRGPU_TREE_TIES=$t bench (256^3, 6DFOF + tree unbinding; look at ptree.build), then the 2e7 single halo"
    VRGPU_TREE_TIES=$t bench gpurun_out/probe_ties_t$t.json --fofbgtype 3 --pp-threshold 150
    VRGPU_TREE_TIES=$t bench gpurun_out/probe_ties_zoom_t$t.json --single-halo 20000000 --fofbgtype 3 --pp-threshold 150
  done
}
{
  for s in $SECTIONS; do
    case $s in pipes|gated|sym|walk|link|disp|wb|ties|veldens) $s ;; *) echo "unknown section $s" ;; esac
  done
} > gpurun_out/probe_experiments.log 2>&1
tail -60 gpurun_out/probe_experiments.log
