#!/usr/bin/env bash
# Helper for gpurun calls: tools/gpu_run.sh <step>...   (outputs under gpurun_out/)
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export LD_LIBRARY_PATH=/usr/local/cuda/lib64:${LD_LIBRARY_PATH:-}
for step in "$@"; do
  case "$step" in
    info)   nvidia-smi > gpurun_out/nvidia-smi.txt 2>&1; nproc >> gpurun_out/nvidia-smi.txt ;;
    smoke)  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" ;;
    tests)  timeout 1500 python -m pytest tests -m gpu -q -x --timeout=900 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/pytest_gpu.log ;;
    testsall) timeout 1500 python -m pytest tests -m gpu -q --timeout=900 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -60 gpurun_out/pytest_gpu.log ;;
    golden) timeout 900 python tools/make_golden.py gpurun_out/golden > gpurun_out/golden.log 2>&1; echo "golden rc=$?"; tail -12 gpurun_out/golden.log ;;
    bench)  timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.log; tail -5 gpurun_out/bench.err ;;
    benchref) timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "benchref rc=$?"; tail -3 gpurun_out/bench_ref.log; tail -5 gpurun_out/bench_ref.err ;;
    *) echo "unknown step $step" ;;
  esac
done
