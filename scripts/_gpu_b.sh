o=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $o/r1c_pytest.log 2>&1; tail -3 $o/r1c_pytest.log
for r in panda ur5 fetch baxter; do
  timeout 300 python scripts/probe_fkcc.py $r 2>&1 | tail -1 | sed 's/^/new /'
done | tee $o/r1c_ab.txt
for v in old v1 v2 v3 v4 v5; do
  VAMP_B200_LIB=build/libvamp_$v.so timeout 300 python scripts/probe_fkcc.py panda 2>&1 | tail -1 | sed "s/^/$v /"
done | tee -a $o/r1c_ab.txt
timeout 300 python scripts/probe_throughput.py 2>&1 | tee $o/r1c_probe.txt
