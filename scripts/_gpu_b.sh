o=gpurun_out
VAMP_B200_LIB=build/libvamp_fetchR.so timeout 300 python scripts/probe_fkcc.py fetch 2>&1 | tail -1 | sed "s/^/fetchR /" | tee $o/r1u_ab.txt
for r in fetch panda; do timeout 300 python scripts/probe_fkcc.py $r 2>&1 | tail -1; done | tee -a $o/r1u_ab.txt
