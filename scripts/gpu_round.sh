#!/bin/bash
# One GPU-box visit: parity tests, bench (both arms), ncu launch list, one full capture of the fkcc kernel.
# usage (from the repo root, under gpurun): bash scripts/gpu_round.sh [tag]
tag=${1:-r2}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/${tag}_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $out/${tag}_pytest.log
timeout 600 python __graft_entry__.py smoke > $out/${tag}_smoke.log 2>&1; echo "smoke exit $?" >> $out/${tag}_smoke.log
timeout 900 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?" >> $out/${tag}_bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err
timeout 600 python scripts/probe_throughput.py > $out/${tag}_probe.txt 2>&1
if [ -s $out/${tag}_bench.json ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_bench_launches.csv \
      python bench.py --steps 2 --warmup 3 --configs 20000000 --edges 262144 --no-robots > $out/${tag}_ncu_bench.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:fkcc_kernel -s 3 -c 1 -f -o $out/${tag}_panda_fkcc_mbm \
      python bench.py --steps 1 --warmup 3 --configs 8000000 --edges 4096 --no-robots > $out/${tag}_ncu_full.log 2>&1
fi
tail -3 $out/${tag}_pytest.log; tail -2 $out/${tag}_smoke.log; cat $out/${tag}_bench.json; cat $out/${tag}_bench_ref.json; cat $out/${tag}_probe.txt
