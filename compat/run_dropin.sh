#!/bin/bash
# Runs the reference's own (unmodified) driver programs, linked against libb200attn.so, on the GPU box.
# 04_ring runs as 1 rank by default; set NRANKS=2 (needs 2 GPUs) for the real ring.
cd "$(dirname "$0")/_dropin" || exit 1
rc=0
for p in 01_fa1 02_fwd 02_bwd 03_1gpu; do
  echo "=== $p"; timeout 120 ./$p > $p.log 2>&1; echo "exit $?"; grep -E "PASS|FAIL|Pass|pass|match|Max diff|max_diff|Max difference" $p.log | tail -12
done
N=${NRANKS:-1}
rm -rf /tmp/b200_mpi; echo "=== 04_ring ($N ranks)"
for r in $(seq 0 $((N-1))); do RANK=$r WORLD_SIZE=$N timeout 300 ./04_ring > 04_ring.$r.log 2>&1 & done
wait
grep -E "PASS|FAIL|match|Total significant" 04_ring.0.log | tail -5
# the reference's communication diagnostics (ring pass-around, compute/transfer overlap) need >= 2 ranks
if [ "$N" -ge 2 ]; then
  for p in 01_nccl_verify 02_overlap; do
    rm -rf /tmp/b200_mpi; echo "=== $p ($N ranks)"
    for r in $(seq 0 $((N-1))); do RANK=$r WORLD_SIZE=$N timeout 120 ./$p > $p.$r.log 2>&1 & done
    wait
    grep -E "Received|completed|Total time" $p.0.log | tail -6
  done
fi
