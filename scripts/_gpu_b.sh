o=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $o/r1n_pytest.log 2>&1; tail -6 $o/r1n_pytest.log
for r in panda ur5; do
  timeout 300 python scripts/probe_fkcc.py $r 2>&1 | tail -1
done | tee $o/r1n_ab.txt
