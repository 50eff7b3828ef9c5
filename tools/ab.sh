#!/usr/bin/env bash
# A/B of library variants on one GPU box: tools/ab.sh <bench_stages args...>   (variants: scratch/variants/*.so)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:${LD_LIBRARY_PATH:-}
: > gpurun_out/ab.jsonl
for rep in 1 2; do
for so in scratch/variants/*.so; do
  MSK144_B200_LIB=$PWD/$so timeout 600 python tools/bench_stages.py "$@" >> gpurun_out/ab.jsonl 2>> gpurun_out/ab.err
  MSK144_LDPC_NO_REFILL=1 MSK144_B200_LIB=$PWD/$so timeout 600 python tools/bench_stages.py noise >> gpurun_out/ab.jsonl 2>> gpurun_out/ab.err
done
done
cut -c1-330 gpurun_out/ab.jsonl
tail -5 gpurun_out/ab.err
