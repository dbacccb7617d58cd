mkdir -p gpurun_out
free -g | head -2; nproc
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 $TR --nproc-per-node 8 --master-port 29581 tests/dist/mg_check.py 64 > gpurun_out/mg8_check.log 2>&1; echo "mg8 rc=$?"; grep -E "OK|FAIL|mismatch" gpurun_out/mg8_check.log | tail -6
for N in 8 4; do
timeout 300 $TR --nproc-per-node $N --master-port 2958$N bench.py --gpus $N --steps 3 --warmup 2 > gpurun_out/bench_c2_n$N.json 2> gpurun_out/bench_c2_n$N.err; echo "c2 n$N rc=$?"
timeout 600 $TR --nproc-per-node $N --master-port 2959$N bench.py --gpus $N --nside 512 --fofbgtype 3 --pp-threshold 150 --steps 2 --warmup 1 > gpurun_out/bench_c3_n$N.json 2> gpurun_out/bench_c3_n$N.err; echo "c3 n$N rc=$?"
done
AV=$(free -g | awk '/Mem:/{print $7}')
if [ "$AV" -gt 600 ]; then
timeout 900 $TR --nproc-per-node 8 --master-port 29599 bench.py --gpus 8 --nside 1024 --fofbgtype 3 --pp-threshold 150 --steps 1 --warmup 1 > gpurun_out/bench_c4_n8.json 2> gpurun_out/bench_c4_n8.err; echo "c4 n8 rc=$?"; tail -3 gpurun_out/bench_c4_n8.err
else echo "C4 skipped: only $AV GB host memory available"; fi
for f in gpurun_out/bench_c2_n8.json gpurun_out/bench_c2_n4.json gpurun_out/bench_c3_n8.json gpurun_out/bench_c3_n4.json gpurun_out/bench_c4_n8.json; do python -c "
import json,sys
try:
    j=json.load(open('$f')); print('$f', round(j['value'],1), round(j['ms_per_step'],1), 'e2e', round(j['e2e']['value'],1), j['config']['ngroups'])
except Exception as e: print('$f', 'unreadable', e)
"; done
