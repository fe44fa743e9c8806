o=gpurun_out
VAMP_B200_LIB=build/libvamp_mb.so timeout 120 python scripts/probe_fkcc.py panda 2>&1 | tail -1 | sed "s/^/mb /" | tee $o/r1r_ab.txt
timeout 120 python scripts/probe_fkcc.py panda 2>&1 | tail -1 | sed "s/^/head /" | tee -a $o/r1r_ab.txt
