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
    testsall) timeout 2400 python -m pytest tests -m gpu -q --timeout=900 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -60 gpurun_out/pytest_gpu.log ;;
    golden60) timeout 900 python tools/make_golden.py gpurun_out/golden audio_deep_60s iq_deep_60s > gpurun_out/golden60.log 2>&1; echo "golden60 rc=$?"; tail -4 gpurun_out/golden60.log; cp gpurun_out/golden/*_60s.npz tests/golden/ ;;
    golden) timeout 900 python tools/make_golden.py gpurun_out/golden > gpurun_out/golden.log 2>&1; echo "golden rc=$?"; tail -12 gpurun_out/golden.log ;;
    bench)  timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.log; tail -5 gpurun_out/bench.err ;;
    benchref) timeout 1200 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; echo "benchref rc=$?"; tail -3 gpurun_out/bench_ref.log; tail -5 gpurun_out/bench_ref.err ;;
    configs) timeout 1200 python tools/bench_configs.py > gpurun_out/configs.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"; cut -c1-600 gpurun_out/configs.jsonl; tail -5 gpurun_out/configs.err ;;
    fullsize) timeout 1500 python -m pytest tests/test_gpu_fullsize.py -m gpu -q --timeout=900 > gpurun_out/pytest_fullsize.log 2>&1; echo "fullsize rc=$?"; tail -40 gpurun_out/pytest_fullsize.log ;;
    stages) timeout 600 python tools/bench_stages.py > gpurun_out/stages.log 2>&1; echo "stages rc=$?"; cat gpurun_out/stages.log | cut -c1-400 ;;
    quick)  timeout 900 python -m pytest tests/test_golden.py tests/test_gpu_parity.py -m gpu -q -x --timeout=600 -k "ldpc or cli or stream_lines" > gpurun_out/pytest_quick.log 2>&1; echo "quick rc=$?"; tail -5 gpurun_out/pytest_quick.log ;;
    ncu_list)
      CMD="python bench.py --steps 1 --warmup 1 --seconds 60 --no-cpu-baseline"
      $CMD > gpurun_out/ncu_plain.log 2>&1 &&
      ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
      echo "ncu_list rc=$?"; tail -2 gpurun_out/ncu_plain.log; wc -l gpurun_out/launches.csv ;;
    ncu_full)
      CMD="python bench.py --steps 1 --warmup 1 --seconds 30 --no-cpu-baseline"
      $CMD > gpurun_out/ncu_plain2.log 2>&1 &&
      ncu --set full --clock-control none --import-source on -k regex:'scan2?_kernel|ldpc_kernel|frontend_kernel' -s 0 -c 3 -f -o gpurun_out/prof_full $CMD > gpurun_out/ncu_full.log 2>&1
      echo "ncu_full rc=$?"; tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/*.ncu-rep ;;
    ncu_scan)
      CMD="python bench.py --steps 1 --warmup 1 --seconds 30 --no-cpu-baseline"
      ncu --set full --clock-control none --import-source on -k regex:'scan2?_kernel' -s 0 -c 1 -f -o gpurun_out/prof_scan $CMD > gpurun_out/ncu_scan.log 2>&1
      echo "ncu_scan rc=$?"; tail -2 gpurun_out/ncu_scan.log; ls -la gpurun_out/prof_scan.ncu-rep ;;
    ncu_ldpc)
      CMD="python bench.py --steps 1 --warmup 1 --seconds 30 --no-cpu-baseline"
      ncu --set full --clock-control none --import-source on -k regex:'ldpc_kernel' -s 0 -c 1 -f -o gpurun_out/prof_ldpc $CMD > gpurun_out/ncu_ldpc.log 2>&1
      echo "ncu_ldpc rc=$?"; tail -2 gpurun_out/ncu_ldpc.log; ls -la gpurun_out/prof_ldpc.ncu-rep ;;
    ncu_front)
      for v in cur cur_fft; do
        MSK144_B200_LIB=$PWD/scratch/variants/$v.so ncu --set full --clock-control none --import-source on -k regex:'frontend_kernel' -s 4 -c 2 -f -o gpurun_out/prof_front_$v python tools/bench_stages.py frontend > gpurun_out/ncu_front_$v.log 2>&1
        echo "ncu_front $v rc=$?"
      done; ls -la gpurun_out/prof_front_*.ncu-rep ;;
    ab_scan)
      for f in 2 1 2 1; do MSK144_SCAN_FORM=$f timeout 300 python tools/bench_stages.py scan 2>&1 | sed "s/^/form$f /" ; done > gpurun_out/ab_scan.log 2>&1; cut -c1-330 gpurun_out/ab_scan.log ;;
    parity12)
      for f in 1 2; do MSK144_SCAN_FORM=$f timeout 900 python -m pytest tests/test_golden.py tests/test_gpu_parity.py -m gpu -q -x --timeout=600 > gpurun_out/pytest_form$f.log 2>&1; echo "form $f rc=$?"; tail -3 gpurun_out/pytest_form$f.log; done ;;
    scale8)   # on an 8-GPU box: in-process pool 1/2/4/8, then bench.py under torchrun at N = 8 and N = 4 (cfg2 replicas + cfg5 gather)
      nvidia-smi -L > gpurun_out/gpus.txt 2>&1
      timeout 600 python tools/bench_pool.py 256 > gpurun_out/pool_scale.jsonl 2> gpurun_out/pool_scale.err; echo "pool rc=$?"; cut -c1-420 gpurun_out/pool_scale.jsonl; tail -3 gpurun_out/pool_scale.err
      for n in 8 4; do
        timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 3 --warmup 3 --no-cli --no-cpu-baseline > gpurun_out/bench_n$n.log 2> gpurun_out/bench_n$n.err
        echo "bench n=$n rc=$?"; tail -1 gpurun_out/bench_n$n.log | cut -c1-300; tail -3 gpurun_out/bench_n$n.err
      done ;;
    multi) timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout=600 > gpurun_out/pytest_multi.log 2>&1; echo "multi rc=$?"; tail -5 gpurun_out/pytest_multi.log
           timeout 600 python tools/bench_pool.py 128 > gpurun_out/pool_1gpu.jsonl 2> gpurun_out/pool_1gpu.err; cut -c1-330 gpurun_out/pool_1gpu.jsonl; tail -3 gpurun_out/pool_1gpu.err ;;
    pool8) nvidia-smi -L > gpurun_out/gpus.txt 2>&1
           timeout 600 python tools/bench_pool.py 256 > gpurun_out/pool_scale.jsonl 2> gpurun_out/pool_scale.err; echo "pool rc=$?"; cut -c1-420 gpurun_out/pool_scale.jsonl; tail -3 gpurun_out/pool_scale.err ;;
    share)   # two decoders sharing GPU 0: does a scan CTA beside an LDPC CTA on the same SM fill the idle issue slots?
      run() { env "$@" timeout 300 python tools/bench_pool.py 64 60 "$DEVS" 2>&1 | grep cfg5_pool | cut -c1-330; }
      DEVS=0 run MSK144_SCAN_FORM=1; DEVS=0 run MSK144_SCAN_FORM=2
      DEVS=0,0 run MSK144_SCAN_FORM=1; DEVS=0,0 run MSK144_SCAN_FORM=2
      DEVS=0,0 run MSK144_SCAN_FORM=2 MSK144_LDPC_CTAS=1; DEVS=0,0,0 run MSK144_SCAN_FORM=2 MSK144_LDPC_CTAS=1 ;;
    ncu_fft)
      timeout 300 python tools/bench_stages.py frontend > gpurun_out/stages_frontend.jsonl 2>&1; echo "frontend plain rc=$?"; cut -c1-300 gpurun_out/stages_frontend.jsonl
      # launches of bench_stages.py frontend: (128-block groups) method 2: 3 per run x 7 runs, then method 1: 21, then the 4096-block groups: 2 per run x 7 runs each
      timeout 600 ncu --set full --clock-control none --import-source on -k regex:'frontend_kernel' -s 18 -c 1 -f -o gpurun_out/prof_front_m2_b128 python tools/bench_stages.py frontend > gpurun_out/ncu_front.log 2>&1
      timeout 600 ncu --set full --clock-control none --import-source on -k regex:'frontend_kernel' -s 39 -c 1 -f -o gpurun_out/prof_front_m1_b128 python tools/bench_stages.py frontend >> gpurun_out/ncu_front.log 2>&1
      timeout 600 ncu --set full --clock-control none --import-source on -k regex:'frontend_kernel' -s 52 -c 1 -f -o gpurun_out/prof_front_m2_b4096 python tools/bench_stages.py frontend >> gpurun_out/ncu_front.log 2>&1
      timeout 600 ncu --set full --clock-control none --import-source on -k regex:'frontend_kernel' -s 66 -c 1 -f -o gpurun_out/prof_front_m1_b4096 python tools/bench_stages.py frontend >> gpurun_out/ncu_front.log 2>&1
      echo "ncu_fft rc=$?"
      for r in m2_b128 m1_b128 m2_b4096 m1_b4096; do echo "##### $r"; python tools/ncu_summary.py gpurun_out/prof_front_$r.ncu-rep; done > gpurun_out/ncu_front_summary.txt 2>&1
      ls -la gpurun_out/prof_front_*.ncu-rep; rm -f gpurun_out/prof_front_*.ncu-rep; head -60 gpurun_out/ncu_front_summary.txt ;;
    summarize)   # text summaries of gpurun_out/prof_full.ncu-rep on the box (the report itself may be too large to travel)
      python tools/ncu_summary.py gpurun_out/prof_full.ncu-rep > gpurun_out/ncu_full_summary.txt 2>&1
      for k in scan_kernel ldpc_kernel; do
        python tools/ncu_lines.py gpurun_out/prof_full.ncu-rep $k 30 > gpurun_out/ncu_lines_$k.txt 2>&1
        ncu -i gpurun_out/prof_full.ncu-rep --page source --csv --kernel-name regex:$k > gpurun_out/src_$k.csv 2>/dev/null
        python tools/sass_hist.py gpurun_out/src_$k.csv 30 > gpurun_out/sass_hist_$k.txt 2>&1; rm -f gpurun_out/src_$k.csv
      done
      ncu -i gpurun_out/prof_full.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum > gpurun_out/ncu_full_dram.csv 2>&1
      du -sh gpurun_out ;;
    chz) timeout 900 python -m pytest tests/test_channelizer.py -m gpu -q -s --timeout=600 > gpurun_out/pytest_chz.log 2>&1; echo "chz rc=$?"; grep -E "differ|bin |passed|failed|Error|error" gpurun_out/pytest_chz.log | tail -30
         timeout 600 python tools/bench_channelizer.py > gpurun_out/chz_bench.jsonl 2> gpurun_out/chz_bench.err; cut -c1-500 gpurun_out/chz_bench.jsonl; tail -3 gpurun_out/chz_bench.err ;;
    ncu_chz)
      for n in 1 160; do
        ncu --set full --clock-control none --import-source on -k regex:'direct_kernel|fft_kernel' -s 5 -c 1 -f -o gpurun_out/prof_chz_$n python tools/bench_channelizer.py 120 $n > gpurun_out/ncu_chz_$n.log 2>&1
        echo "ncu_chz $n rc=$?"
      done; ls -la gpurun_out/prof_chz_*.ncu-rep ;;
    warpops) ./scratch/warpops > gpurun_out/warpops.txt 2>&1; cat gpurun_out/warpops.txt ;;
    scanprof) for so in scratch/variants_prof/*.so; do for f in 1 2; do echo "== $so form $f"; MSK144_SCAN_FORM=$f MSK144_B200_LIB=$PWD/$so timeout 300 python tools/scan_prof.py 2>&1 | tail -20; done; done > gpurun_out/scanprof.log 2>&1; cat gpurun_out/scanprof.log ;;
    ab_scan_variants)
      for rep in 1 2; do for so in scratch/variants/*.so; do for f in 1 2; do
        MSK144_SCAN_FORM=$f MSK144_B200_LIB=$PWD/$so timeout 300 python tools/bench_stages.py scan 2>&1 | grep stages_deep | sed "s/^/form$f /"
      done; done; done > gpurun_out/ab_scan_variants.log 2>&1; cut -c1-260 gpurun_out/ab_scan_variants.log ;;
    golden_multi) timeout 1500 python -m pytest tests/test_golden.py tests/test_gpu_multi.py -m gpu -q --timeout=900 -s > gpurun_out/pytest_gm.log 2>&1; echo "golden_multi rc=$?"; grep -E "passed|failed|_60s:|Error" gpurun_out/pytest_gm.log | tail -20 ;;
    *) echo "unknown step $step" ;;
  esac
done
