#!/bin/bash
# Run on the GPU box (under gpurun): tests, smoke, bench (both arms), ncu launch list, one full ncu capture of the top kernel.
# usage: scripts/gpu_round.sh <tag> [quick]
TAG=${1:-r2}
QUICK=$2
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q --maxfail=20 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest.log
timeout 300 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/${TAG}_smoke.log
timeout 600 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
if [ -z "$QUICK" ]; then
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_ref.json 2> $O/${TAG}_bench_ref.err; echo "bench ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:n0_rows5 -s 1 -c 1 -o $O/${TAG}_n0_rows5 -f \
    python scripts/n0_profile_case.py > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
if [ -f $O/${TAG}_n0_rows5.ncu-rep ]; then
  ncu -i $O/${TAG}_n0_rows5.ncu-rep --page raw --csv > $O/${TAG}_n0_rows5_raw.csv 2>/dev/null
  python scripts/ncu_raw_summary.py $O/${TAG}_n0_rows5_raw.csv > $O/${TAG}_n0_rows5_ncu_full_summary.txt 2>&1
fi
fi
tail -25 $O/${TAG}_pytest.log; tail -2 $O/${TAG}_smoke.log; tail -5 $O/${TAG}_bench.err; cat $O/${TAG}_bench.json | cut -c 1-1500; cat $O/${TAG}_bench_ref.json 2>/dev/null | cut -c 1-600
