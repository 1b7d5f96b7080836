#!/bin/bash
# one GPU-box pass at the end of a round: new graph cases, the bench line, the reference arm (development aid)
timeout 400 python -m pytest tests/test_runner_gpu.py -x -q -k "custom" 2>&1 | tail -4
timeout 600 python bench.py > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_reference.json 2>/dev/null
python - <<'P'
import json
d=json.loads(open("gpurun_out/r2_bench_final.json").read().strip().splitlines()[-1]); r=d["roofline"]
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], r["frac"], r["standalone_ntt_gbs"], d["cpu_baseline"], d["clocks"])
print({k:(round(v,1) if isinstance(v,float) else v) for k,v in d["secondary"].items() if k!="note"})
e=json.loads(open("gpurun_out/r2_bench_reference.json").read().strip().splitlines()[-1]); print(e["value"], e["cpu_baseline"])
P
