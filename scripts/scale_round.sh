#!/bin/bash
# Run on an 8-GPU box (gpurun --gpus 8): the strong-scaling bench at N = 8, 4, 2, 1 (both arms at N = 1), the configs[4] sweep at its
# stated size on 8 GPUs and the tomographic Fisher driver on 2 and 1 GPUs.   usage: scripts/scale_round.sh <tag>
TAG=${1:-r2s}
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for N in 8 2; do
  timeout 240 $TR --nproc-per-node $N --master-port $((29700 + N)) scripts/xchg_check.py > $O/${TAG}_xchg_check_n$N.log 2>&1; echo "xchg check N=$N rc=$?"; tail -1 $O/${TAG}_xchg_check_n$N.log
done
for N in 8 4 2; do
  timeout 600 $TR --nproc-per-node $N --master-port $((29500 + N)) bench.py --gpus $N --steps 10 --warmup 3 > $O/${TAG}_bench_n$N.json 2> $O/${TAG}_bench_n$N.err; echo "bench N=$N rc=$?"
done
timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 > $O/${TAG}_bench_n1.json 2> $O/${TAG}_bench_n1.err; echo "bench N=1 rc=$?"
timeout 900 $TR --nproc-per-node 8 --master-port 29611 scripts/sweep_bench.py --full > $O/${TAG}_sweep_full_n8.json 2> $O/${TAG}_sweep_full_n8.err; echo "sweep rc=$?"
timeout 300 $TR --nproc-per-node 2 --master-port 29622 scripts/tomography_fisher_bench.py > $O/${TAG}_tomography_fisher_n2.json 2> $O/${TAG}_tomography_fisher_n2.err; echo "tomo2 rc=$?"
timeout 300 python scripts/tomography_fisher_bench.py > $O/${TAG}_tomography_fisher_n1.json 2> $O/${TAG}_tomography_fisher_n1.err; echo "tomo1 rc=$?"
for N in 8 4 2 1; do python - <<PY
import json
b = json.loads(open('$O/${TAG}_bench_n$N.json').read().strip().splitlines()[-1])
print($N, 'value', round(b['value']), 'ms', round(b['ms_per_step'], 3), 'e2e', round(b['e2e']['value']), 'e2e ms', round(b['e2e']['ms_per_step'], 3), b['rates'].get('exchange_ms_per_step'), b.get('exchange'))
PY
done
tail -c 700 $O/${TAG}_sweep_full_n8.json; tail -c 400 $O/${TAG}_tomography_fisher_n2.json | head -c 600; echo; head -c 500 $O/${TAG}_tomography_fisher_n1.json
