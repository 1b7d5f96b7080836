#!/bin/bash
# scratch-budget / NTT-variant sweep (development aid; results land in gpurun_out/)
for tw in 0 1; do
for mb in 96 512 2048 8192; do
  echo "== LSB_NTT_TWOPASS=$tw LSB_SCRATCH_MB=$mb"
  LSB_NTT_TWOPASS=$tw LSB_SCRATCH_MB=$mb python bench.py --steps 3 --warmup 3 --no-cpu 2>&1 | python -c "
import sys,json
for line in sys.stdin:
    try: d=json.loads(line)
    except Exception: print(line.strip()[:300]); continue
    r=d['roofline']; print('value %.0f e2e %.0f launches %d ntt_instep %.0f GB/s standalone %.0f GB/s share %.2f' % (d['value'], d['e2e']['value'], d['gpu_launches'], r['achieved'], r['standalone_ntt_gbs'], r['share_of_step'])); print(d['kernel_ms_per_step'])
"
done; done
