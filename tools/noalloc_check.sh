#!/bin/bash
# GPU-box pass for lsb_mod_raise_static: alone in a fresh process, then the whole GPU suite (it runs last), then the bench line
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -5
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
timeout 600 python bench.py > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err
tail -c 400 gpurun_out/r2_bench_final.err
python - <<'P'
import json
d=json.loads(open("gpurun_out/r2_bench_final.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], d["cpu_baseline"]["value"], d["clocks"])
P
