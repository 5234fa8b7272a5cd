#!/bin/bash
# Runs every UMMA/TMA primitive probe in its own process (a bad descriptor can poison a context).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
n=$(./build/umma_probe)
: > gpurun_out/probe.log
for i in $(seq ${1:-0} $((n-1))); do
  timeout 30 ./build/umma_probe $i >> gpurun_out/probe.log 2>&1 || echo "[probe $i] exit code $?" >> gpurun_out/probe.log
done
cat gpurun_out/probe.log
