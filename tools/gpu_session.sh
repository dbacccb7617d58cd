#!/bin/bash
# One measuring session on the GPU box: the parity suites named on the command line (default: all), then bench lines.
# Usage: gpurun --timeout 1200 -- 'bash tools/gpu_session.sh tree rest b256'
set -u
mkdir -p gpurun_out
for s in "$@"; do
  case $s in
    tree) timeout 600 python -m pytest tests/test_gpu_tree.py -q -m gpu --tb=short -p no:cacheprovider > gpurun_out/tree_pytest.log 2>&1; tail -30 gpurun_out/tree_pytest.log ;;
    rest) timeout 900 python -m pytest tests/test_gpu_unbind.py tests/test_gpu_zoom.py tests/test_gpu_dropin.py tests/test_gpu_veldens.py tests/test_gpu_energy.py -q -m gpu --tb=short -p no:cacheprovider > gpurun_out/rest_pytest.log 2>&1; tail -30 gpurun_out/rest_pytest.log ;;
    all) timeout 1500 python -m pytest tests -q -m gpu --tb=short -p no:cacheprovider > gpurun_out/all_pytest.log 2>&1; tail -30 gpurun_out/all_pytest.log ;;
    fof) timeout 900 python -m pytest tests/test_gpu_fof3d.py tests/test_gpu_fof6d.py tests/test_gpu_dropin.py -q -m gpu --tb=short -p no:cacheprovider > gpurun_out/fof_pytest.log 2>&1; tail -30 gpurun_out/fof_pytest.log ;;
    base) timeout 1500 python -m pytest tests/test_gpu_baseline_sizes.py -q -m gpu --tb=short -s -p no:cacheprovider > gpurun_out/base_pytest.log 2>&1; tail -12 gpurun_out/base_pytest.log ;;
    b256) timeout 600 python bench.py --nside 256 --steps 3 --warmup 2 --no-cpu-baseline --no-e2e-secondary > gpurun_out/b256_c3.json 2> gpurun_out/b256_c3.err; python tools/bench_summary.py gpurun_out/b256_c3.json 2>/dev/null | head -30; tail -3 gpurun_out/b256_c3.err ;;
    b256c2) timeout 600 python bench.py --config 2 --steps 3 --warmup 2 --no-cpu-baseline --no-e2e-secondary > gpurun_out/b256_c2.json 2> gpurun_out/b256_c2.err; python tools/bench_summary.py gpurun_out/b256_c2.json 2>/dev/null | head -30; tail -3 gpurun_out/b256_c2.err ;;
    b512) timeout 1500 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-e2e-secondary > gpurun_out/b512_c3.json 2> gpurun_out/b512_c3.err; python tools/bench_summary.py gpurun_out/b512_c3.json 2>/dev/null | head -40; tail -3 gpurun_out/b512_c3.err ;;
    zoom20) timeout 900 python bench.py --single-halo 20000000 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e-secondary > gpurun_out/zoom20.json 2> gpurun_out/zoom20.err; python tools/bench_summary.py gpurun_out/zoom20.json 2>/dev/null | head -24; tail -3 gpurun_out/zoom20.err ;;
    l512) bash tools/ncu_launches.sh r02j_launches_512c3 ;;
    l256) bash tools/ncu_launches.sh r02g_launches_256c3 --nside 256 ;;
    t256) bash tools/ncu_traffic.sh r02g_traffic_256c3 --nside 256 ;;
    t512) bash tools/ncu_traffic.sh r02j_traffic_512c3 ;;
    t256c2) bash tools/ncu_traffic.sh r02g_traffic_256c2 --config 2 ;;
    zz) timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu --tb=short -p no:cacheprovider > gpurun_out/zz_pytest.log 2>&1; tail -15 gpurun_out/zz_pytest.log ;;
    new) timeout 1200 python -m pytest tests/test_gpu_dropin.py tests/test_gpu_stf_files.py tests/test_zz_no_alloc.py -q -m gpu --tb=short -s -p no:cacheprovider > gpurun_out/new_pytest.log 2>&1; tail -40 gpurun_out/new_pytest.log ;;
    mg) for w in ${MG_WORLDS:-2 4 8}; do n=$(nvidia-smi -L | wc -l); [ $w -le $n ] || continue; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$w --master-addr 127.0.0.1 --master-port 29541 tests/dist/mg_check.py ${MG_NSIDE:-128} 2>&1 | grep -v "^W\|^\*" | tee gpurun_out/r02_mg${w}_check.log | tail -8; done ;;
    final) timeout 900 python bench.py > gpurun_out/r02l_bench_default.json 2> gpurun_out/r02l_bench_default.err; python tools/bench_summary.py gpurun_out/r02l_bench_default.json 2>/dev/null | head -30; tail -3 gpurun_out/r02l_bench_default.err ;;
    refarm) timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r02l_bench_reference.json 2> gpurun_out/r02l_bench_reference.err; cut -c1-400 gpurun_out/r02l_bench_reference.json; tail -3 gpurun_out/r02l_bench_reference.err ;;
    ncuwalk) bash tools/ncu_kernel.sh r02j_walk k_pt_walk 1 --nside 256 ;;
    *) echo "unknown section $s" ;;
  esac
done
