#!/bin/bash
# rebuild lsb_ntt_cluster.cu with different -D flags on the GPU box and time the N=2^16 key-switch (development aid)
# usage: ab_ks16.sh "ENV=.. -Dflags" ...
cd lattisense_b200/csrc
for spec in "$@"; do
  echo "== $spec"
  envs=""; flags=""
  for w in $spec; do case "$w" in -D*) flags="$flags $w";; *) envs="$envs $w";; esac; done
  rm -f ../_build/lsb_ntt_cluster.o
  make -s NVFLAGS_EXTRA="$flags" >/dev/null 2>&1 || { echo build failed; continue; }
  (cd ../..; env $envs timeout 200 python tools/prof_ks16.py ks16 16 | head -3; env $envs timeout 100 python tools/prof_ks16.py ks16 1 | head -1; env $envs PROF_MOD0=0 PROF_NMOD=1 timeout 100 python tools/prof_ntt.py 16 2048 20 t | grep -v "^ok")
done
