mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for N in 8 4 2; do
timeout 300 $TR --nproc-per-node $N --master-port 2960$N bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_c2b_n$N.json 2> gpurun_out/bench_c2b_n$N.err; echo "c2 n$N rc=$?"
python -c "
import json
j=json.load(open('gpurun_out/bench_c2b_n$N.json')); print($N, round(j['value'],1), round(j['ms_per_step'],2), 'e2e', round(j['e2e']['value'],1), round(j['e2e']['ms_per_step'],1), j['config']['ngroups'])
k=j['kernel_ms_rank0']; print({a:round(k[a],2) for a in k if 'potential' in a or a.startswith('mg.')})"
done
