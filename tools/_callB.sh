timeout 600 python bench.py --steps 20 > gpurun_out/bench1.log 2>gpurun_out/bench1.err; tail -1 gpurun_out/bench1.log | python -c '
import sys, json
d = json.loads(sys.stdin.read())
print("ms", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], d["clocks"])
print(json.dumps(d["consumers"], indent=1))'
tail -3 gpurun_out/bench1.err
