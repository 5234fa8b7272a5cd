#!/bin/bash
# the reference's own unmodified drivers (01, 02 fwd / bwd, 03 1-GPU, 04 ring on 2 ranks, its nccl / overlap diagnostics) against libb200attn.so
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NRANKS=2 bash compat/run_dropin.sh > gpurun_out/dropin_2ranks.txt 2>&1
cat gpurun_out/dropin_2ranks.txt
for f in compat/_dropin/02_bwd.log compat/_dropin/04_ring.0.log; do echo "---- $f"; tail -15 $f; done >> gpurun_out/dropin_2ranks.txt
