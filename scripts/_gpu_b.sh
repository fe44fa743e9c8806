o=gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > $o/r2a_pytest.log 2>&1; tail -3 $o/r2a_pytest.log
for r in panda ur5 fetch baxter; do timeout 300 python scripts/probe_fkcc.py $r 2>&1 | tail -1; done | tee $o/r2a_ab.txt
