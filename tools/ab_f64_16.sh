#!/bin/bash
# rebuild lsb_ntt_cluster.cu with different -D flags on the GPU box and time the FP64-policy NTT at N=2^16 / 2^15 (development aid)
cd lattisense_b200/csrc
for spec in "$@"; do
  echo "== $spec"
  rm -f ../_build/lsb_ntt_cluster.o
  make -s NVFLAGS_EXTRA="$spec" >/dev/null 2>&1 || { echo build failed; continue; }
  (cd ../..; PROF_MOD0=1 PROF_NMOD=1 timeout 100 python tools/prof_ntt.py 16 2048 20 t | grep -v "^ok"; PROF_PARAMS=CKKS_BTP_N15 PROF_MOD0=1 PROF_NMOD=1 timeout 100 python tools/prof_ntt.py 15 4096 20 t | grep -v "^ok")
done
