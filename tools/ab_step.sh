#!/bin/bash
# rebuild lsb_ntt_cluster.cu with different -D flags on the GPU box and time the cfg2 step (development aid)
cd lattisense_b200/csrc
for spec in "$@"; do
  echo "== $spec"
  envs=""; flags=""
  for w in $spec; do case "$w" in -D*) flags="$flags $w";; *) envs="$envs $w";; esac; done
  rm -f ../_build/lsb_ntt_cluster.o
  make -s NVFLAGS_EXTRA="$flags" >/dev/null 2>&1 || { echo build failed; continue; }
  (cd ../..; env $envs timeout 200 python bench.py --no-cpu --no-api --no-secondary --steps 10 2>&1 | python -c "
import sys,json
for line in sys.stdin:
    try: d=json.loads(line)
    except Exception: continue
    r=d['roofline']; print('step %.3f ms  ntt_instep %.0f GB/s frac %.3f standalone %.0f' % (d['ms_per_step'], r['achieved'], r['frac'], r['standalone_ntt_gbs'])); print(d['kernel_ms_per_step'])
")
done
