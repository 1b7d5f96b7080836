#!/bin/bash
# rebuild lsb_ntt_cluster.cu with different -D flags on the GPU box and time the NTT (development aid)
cd lattisense_b200/csrc
for flags in "$@"; do
  echo "== $flags"
  rm -f ../_build/lsb_ntt_cluster.o
  make -s NVFLAGS_EXTRA="$flags" >/dev/null 2>&1 || { echo build failed; continue; }
  (cd ../..; python tools/prof_ntt.py 14 8192 20 t; python tools/prof_ntt.py 16 2048 20 t; PROF_MOD0=0 PROF_NMOD=1 python tools/prof_ntt.py 16 2048 20 t) 2>&1 | grep -v "^ok"
done
