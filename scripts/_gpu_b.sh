o=gpurun_out
timeout 900 python -m pytest tests -m gpu -q -k "motion or edges or m1 or m3 or path or dropin" > $o/r1q_pytest.log 2>&1; tail -4 $o/r1q_pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 --configs 20000000 > $o/r1q_bench.json 2> $o/r1q_bench.err; python -c "
import json;d=json.loads(open('$o/r1q_bench.json').read().strip().splitlines()[-1]);print('value',d['value'],'edges/s',d['extra']['edges_per_s'],'cpu edges',d['cpu_baseline']['edges_per_s'],d['cpu_baseline']['edge_verdicts_match_gpu'])"
