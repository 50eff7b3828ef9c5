#!/usr/bin/env python3
"""bench.py -- MSK144 deep-scan stream decode on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[1]): synthetic 15-minute 12 kHz s16 audio stream with injected MSK144
pings at SNR -4..+10 dB, deep scan --search-step=1 --search-width=500 --scan-depth=6
--nbadsync-threshold=3 (NF = 501 frequency hypotheses, 24,048 candidates per block, 4,165 blocks).
One step = one pass of the whole hot path (front end -> scan/softbits/gate -> LDPC) over the stream.
With N > 1 every rank decodes its own 15-minute stream (independent channels, no collective on the
data path): weak scaling, value = total audio seconds / max-over-ranks device time.

  value     audio seconds decoded per second, input resident in HBM, timed with CUDA events
  e2e       same through msk144_decode_host: pinned host input, H2D + kernels + D2H + host text decode
  roofline  FP32-issue roof of the dominant kernel (the path is not HBM-bound); HBM as a sub-object
  cli       the same stream through the drop-in executable (bytes on stdin, lines on stdout), start-up separated
  cfg5      BASELINE.json configs[4]: 1024 channels x 60 s, sharded over the ranks by msk144cudecoder_b200.sharding,
            host-side gather of the decode lines inside the timed region
  cpu_baseline, lines_identical_to_reference: see DESIGN.md section 6

--impl reference runs the UNMODIFIED reference (oracle/_ref/msk144cudecoder_ref: its own CUDA kernels recompiled
for sm_100a, its own CLI; the reference has no CPU implementation of the path) on a bounded sample of the same
stream -- the first 60 s -- once per step, plus one pass over the whole stream outside the step loop; its fixed
process start-up is measured on a one-block input and reported.  If that binary is not usable the arm falls back
to the CPU oracle port on all host cores.
"""
import argparse
import hashlib
import json
import multiprocessing as mp
import os
import re
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLAGS = dict(center=1500.0, width=500.0, step=1.0, depth=6, thr=3)
CLI_FLAGS = ["--search-step=1", "--search-width=500", "--scan-depth=6", "--nbadsync-threshold=3"]
STREAM_SECONDS = 900.0
SAMPLE_SECONDS = 60.0     # the reference arm's bounded sample: first 276 blocks == tests/golden/audio_deep_60s.npz
METRIC = "real_time_factor_deep_scan"
UNIT = "audio_s_per_s"
WORKLOAD = ("cfg2: synthetic 15-min 12 kHz s16 audio, pings at SNR -4..+10 dB; --search-step=1 --search-width=500 "
            "--scan-depth=6 --nbadsync-threshold=3 (NF=501, 24048 candidates/block, 4165 blocks)")
REF_EXE = os.path.join(ROOT, "oracle", "_ref", "msk144cudecoder_ref")
GOLDEN_60S = os.path.join(ROOT, "tests", "golden", "audio_deep_60s.npz")


def workload_config(world, seconds=STREAM_SECONDS):
    """The `config` object: identical in both arms (it names the workload, not the implementation)."""
    nblocks = (int(seconds * 12000) - 5184) // 2592 + 1
    return {"workload": WORKLOAD if seconds == STREAM_SECONDS else WORKLOAD + f" [shortened to {seconds:.0f} s]",
            "streams": world, "blocks_per_stream": nblocks, "flags": " ".join(CLI_FLAGS), "seed": 144,
            "decode_cache_quirk": "reproduced (MSK144_FLAG_COMPAT_DECODE_CACHE: a block's first text is reused, main.cu:437-445)",
            "l2": "ours: 256 MiB memset between steps, per-step intermediates (1.7 GB LLR queue) exceed L2; reference: one process per step",
            "parallelism": f"{world} independent stream(s), one process per GPU, no data-path collective"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("sm_max_mhz", 1965.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1965.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except (ValueError, IndexError):
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def algorithmic_flops(nf, depth, nblocks):
    """per-block work of scan + soft bits in the direct (reformulated, no early-out) form."""
    scan = nf * (1_798_848 + 31_104 * depth)
    soft = nf * depth * 8 * 15_300
    return nblocks * scan, nblocks * soft


# ---------------------------------------------------------------------------------------- CLI harness (both arms)
def mask_date(text):
    return [re.sub(r"date=\d{14}", "date=" + "x" * 14, l) for l in text.splitlines()]


def untie(lines):
    """Lines without f0 / pattern_idx: they name the winner among candidates tied in (num_avg, nbadsync), which the
    reference does not reproduce from one run to the next (its LLR race decides which marginal candidates decode, its
    unstable sort then picks)."""
    return [re.sub(r"f0=[^;]*; ", "", re.sub(r"pattern_idx=\d+; ", "", l)) for l in lines]


def compare_lines(a, b):
    """{identical: same text/snr/num_avg/nbadsync/order in every line, strict: lines equal in every field, of: lines}"""
    same = untie(a) == untie(b)
    return {"identical": same, "strict": sum(x == y for x, y in zip(a, b)) - 1 if same else None, "of": len(b) - 1}


def cli_once(cmd, data, timeout=1800):
    """One process: bytes on stdin, stdout captured.  Returns (wall seconds, masked stdout lines)."""
    env = dict(os.environ, LD_LIBRARY_PATH="/usr/local/cuda/lib64:" + os.environ.get("LD_LIBRARY_PATH", ""))
    t0 = time.perf_counter()
    p = subprocess.run(cmd, input=data, capture_output=True, env=env, timeout=timeout)
    dt = time.perf_counter() - t0
    out = p.stdout.decode(errors="replace")
    if p.returncode != 0 or not out.rstrip().endswith("Done"):
        raise RuntimeError(f"rc={p.returncode}: " + p.stderr.decode(errors="replace")[-300:])
    return dt, mask_date(out)


def cli_figures(cmd, sample, full, steps, warmup, audio_s_sample, audio_s_full):
    """Wall-clock figures of a stdin->stdout executable: fixed start-up on a one-block input, the 60 s sample per
    step (median), one pass over the whole stream.  Steady-state = start-up subtracted."""
    one_block = sample[: 5184 * 2]
    startup = float(np.median([cli_once(cmd, one_block)[0] for _ in range(5)]))
    times, lines, strict_same = [], None, 0
    for i in range(warmup + steps):
        dt, ls = cli_once(cmd, sample)
        if lines is not None:
            if untie(ls) != untie(lines):
                raise RuntimeError("output differs between two runs on the same input beyond f0 / pattern_idx tie-breaks")
            strict_same += ls == lines
        else:
            lines = ls
        if i >= warmup:
            times.append(dt)
    dt_med = float(np.median(times))
    # the steady-state figure of the sample means something only when the sample outlasts the start-up noise
    steady = audio_s_sample / (dt_med - startup) if dt_med - startup > 0.25 * dt_med and dt_med - startup > 0.2 else None
    out = {"startup_s": startup, "sample_s_median": dt_med, "sample_s_all": [round(t, 4) for t in times],
           "runs_with_strictly_identical_output": f"{strict_same} of {warmup + steps - 1} repeats (all identical apart from f0 / pattern_idx tie-breaks)",
           "raw_value": audio_s_sample / dt_med, "value": steady, "unit": UNIT,
           "sample": f"first {int(round(audio_s_sample / 0.216))} blocks ({audio_s_sample:.1f} s of audio) per process, "
                     "bytes on stdin, lines on stdout; value = steady state (start-up on a one-block input subtracted), "
                     "raw_value = whole process wall clock"}
    if full is not None:
        dt_full, full_lines = cli_once(cmd, full)
        full_all = [dt_full]
        if dt_full < 3.0:   # a process this short is within a few start-up jitters of its start-up: median of five
            full_all += [cli_once(cmd, full)[0] for _ in range(4)]
            dt_full = float(np.median(full_all))
        out.update(full_stream_s=dt_full, full_stream_s_all=[round(t, 4) for t in full_all], full_stream_raw_value=audio_s_full / dt_full,
                   full_stream_value=audio_s_full / max(1e-9, dt_full - startup), full_stream_lines=len(full_lines) - 1)
        if steady is None:   # a sample shorter than the start-up jitter: the whole stream is the figure
            out["value"] = out["full_stream_value"]
            out["sample"] += "; the sample runs shorter than the start-up jitter here, so value = the whole-stream steady state"
    return out, lines


def lines_digest(lines):
    return hashlib.sha256("\n".join(lines).encode()).hexdigest()[:16]


def golden_lines():
    if os.path.exists(GOLDEN_60S):
        return [str(s) for s in np.load(GOLDEN_60S)["cli_lines"]]
    return None


def cfg2_stream(seed=144, seconds=STREAM_SECONDS):
    from msk144cudecoder_b200 import synth
    return synth.cfg2_deep_audio(seconds=seconds, seed=seed)


# ---------------------------------------------------------------------------------------- reference arm
def run_reference(args, rank):
    """The reference arm: unmodified reference binary over its own CLI, bounded sample of the workload per step."""
    if rank != 0:
        return
    from msk144cudecoder_b200 import synth
    x, _ = cfg2_stream(seconds=args.seconds)
    ns = int(min(SAMPLE_SECONDS, args.seconds) * 12000)
    sample = np.ascontiguousarray(x[:ns])
    nblocks = synth.num_blocks(len(sample))
    audio_s = nblocks * 0.216
    audio_full = synth.num_blocks(len(x)) * 0.216
    kind, cores, note, extra = "reference", 1, "", {}
    usable = os.path.exists(REF_EXE)
    if usable:
        try:
            fig, lines = cli_figures([REF_EXE] + CLI_FLAGS, sample.tobytes(), None if args.no_full_stream else x.tobytes(),
                                     args.steps, args.warmup, audio_s, audio_full)
            val, dt = fig["value"] or fig["raw_value"], fig["sample_s_median"]
            gl = golden_lines()
            extra = {"cli": fig, "raw_value": fig["raw_value"], "startup_s": fig["startup_s"],
                     "full_stream_value": fig.get("full_stream_value"),
                     "lines_sha16": lines_digest(untie(lines)), "lines": len(lines) - 1,
                     "lines_vs_golden": compare_lines(lines, gl) if gl is not None and args.seconds >= SAMPLE_SECONDS else None}
            extra["lines_identical_to_golden"] = extra["lines_vs_golden"]["identical"] if extra["lines_vs_golden"] else None
            note = ("unmodified reference CLI (its CUDA kernels recompiled for sm_100a, no fast-math) on the same B200; one process "
                    "per step, median; value = steady state: the process start-up (CUDA context, cuFFT plan; measured on a one-block "
                    f"input: {fig['startup_s']:.2f} s) is subtracted, raw_value keeps it")
        except Exception as e:  # noqa: BLE001
            usable, note = False, f"reference binary unusable ({' '.join(str(e).split())[:120]}); CPU oracle port instead"
    if not usable:
        from oracle import orc
        kind, cores = "port", os.cpu_count() or 1
        nb_s = min(nblocks, int(os.environ.get("MSK144_BENCH_REF_BLOCKS", 8 * cores)))  # env: contract tests only
        audio_s = nb_s * 0.216
        cfg = orc.config(center=1500.0, width=500.0, step=1.0, depth=6, thr=3, scan_mode=1, nthreads=cores)
        times = []
        for i in range(min(1, args.warmup) + args.steps):
            t0 = time.perf_counter()
            orc.decode_stream(sample, cfg, max_blocks=nb_s)
            if i >= min(1, args.warmup):
                times.append(time.perf_counter() - t0)
        nblocks = nb_s
        dt = float(np.median(times))
        val = audio_s / dt
        note = note or "CPU oracle port (C++ restatement of the reference), all host cores"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args.gpus, args.seconds),
            "reference_sample": f"first {nblocks} blocks ({audio_s:.1f} s of audio) of the stream per step",
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": f"{nblocks} blocks ({audio_s:.1f} s of audio) per step; {note}"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    line.update(extra)
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------- cfg5 input (before CUDA)
def _cfg5_one(c):
    from msk144cudecoder_b200 import synth
    return synth.cfg5_channel(c, seconds=60.0)[0]


def cfg5_input(channels, world):
    """This rank's channels, synthesised on host cores in a fork pool (call before CUDA is initialised)."""
    if not channels:
        return np.zeros((0, 720000), np.int16)
    workers = max(1, min(16, (os.cpu_count() or 1) // max(1, world)))
    if workers == 1:
        return np.stack([_cfg5_one(c) for c in channels])
    with mp.get_context("fork").Pool(workers) as pool:
        return np.stack(pool.map(_cfg5_one, channels, chunksize=4))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--seconds", type=float, default=STREAM_SECONDS, help="stream length (default: the 15-min workload)")
    ap.add_argument("--batch-blocks", type=int, default=128)
    ap.add_argument("--cpu-blocks", type=int, default=0, help="blocks of the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-full-stream", action="store_true", help="skip the one whole-stream pass of the CLI legs")
    ap.add_argument("--no-cli", action="store_true", help="skip the CLI / reference-check legs of our arm")
    ap.add_argument("--cfg5-channels", type=int, default=1024, help="channels of the multi-channel leg (0 = skip)")
    ap.add_argument("--no-one-stream", action="store_true", help="skip the one-stream-over-all-ranks leg")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank)
        return

    from msk144cudecoder_b200 import sharding, synth

    # ---- host-side synthesis first (fork pool: before CUDA exists in this process) ---------------------
    x, pings = cfg2_stream(seed=144 + rank, seconds=args.seconds)
    my_chans = sharding.channels_of_rank(args.cfg5_channels, rank, world) if args.cfg5_channels > 0 else []
    x5 = cfg5_input(my_chans, world) if args.cfg5_channels > 0 else None

    import torch
    import torch.distributed as dist

    import msk144cudecoder_b200 as m

    assert torch.cuda.is_available(), "bench.py needs a GPU: the decoder has no CPU path"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # ---- workload: one 15-min stream per rank (rank r = channel r) ------------------------------------
    nsamp = len(x)
    nblocks = synth.num_blocks(nsamp)
    audio_s = nblocks * 0.216   # every block advances the stream by one 2592-sample hop
    x_pinned = torch.from_numpy(x).pin_memory()
    x_dev = x_pinned.to(dev, non_blocking=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    dec = m.Decoder(device=local_rank, max_batch_blocks=args.batch_blocks, compat_cache=True, **FLAGS)
    stream = torch.cuda.Stream(dev)   # the kernels are launched on this stream; the events below are recorded on it
    torch.cuda.synchronize(dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def device_step():
        with torch.cuda.stream(stream):
            flush.zero_()   # L2 flush between steps (inside the timed region: ~0.05 ms of a ~0.5 s step)
        dec.run_device(x_dev.data_ptr(), nsamp, 1, stream.cuda_stream)

    # ---- resident-input throughput -----------------------------------------------------------------
    for _ in range(args.warmup):
        device_step()
    barrier()
    dec.stage_times()   # drop warm-up timings
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clk:
        ev0.record(stream)
        for _ in range(args.steps):
            device_step()
        ev1.record(stream)
        barrier()
    ms_total = ev0.elapsed_time(ev1)
    st = dec.stage_times()     # per-kernel event sums over exactly the timed region
    t = torch.tensor([ms_total], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = world * audio_s * args.steps / (ms_max / 1e3)

    # ---- end to end through the public API: pinned host samples in, decode lines out -----------------
    caps = dict(line_cap=1 << 15, result_cap=0)   # the call a user makes: decode lines out
    for _ in range(min(args.warmup, 2)):
        dec.decode_host(x_pinned.numpy(), **caps)
    barrier()
    b0 = dec.stage_times()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        lines, _ = dec.decode_host(x_pinned.numpy(), **caps)
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0
    b1 = dec.stage_times()
    t2 = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t2, op=dist.ReduceOp.MAX)
    e2e_val = world * audio_s * args.steps / float(t2.item())
    h2d = (b1["h2d_bytes"] - b0["h2d_bytes"]) // args.steps
    d2h = (b1["d2h_bytes"] - b0["d2h_bytes"]) // args.steps
    found_texts = {l["text"] for l in lines}
    sent_texts = {p.text for p in pings}

    # ---- cfg5: 1024 channels x 60 s over all ranks, host-side gather of the lines inside the timed region ----
    cfg5 = None
    if args.cfg5_channels > 0:
        x5p = torch.from_numpy(x5).pin_memory() if len(my_chans) else None

        def cfg5_pass():
            def decode_channels(chans):
                return dec.decode_host(x5p.numpy(), nchan=len(chans), line_cap=1 << 17, result_cap=0)[0]
            t_a = time.perf_counter()
            mine = [dict(l, channel=my_chans[l["channel"]]) for l in decode_channels(my_chans)] if my_chans else []
            t_b = time.perf_counter()
            merged = sharding.gather_lines(mine, rank, world)
            return merged, t_b - t_a, time.perf_counter() - t_b

        cfg5_pass()                       # warm-up (buffers grow to the multi-channel shape)
        barrier()
        dec.stage_times()
        t0 = time.perf_counter()
        merged, t_dec, t_gather = cfg5_pass()
        barrier()
        t_pass = time.perf_counter() - t0
        s5 = dec.stage_times()
        tt = torch.tensor([t_pass, t_dec, t_gather], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_pass, t_dec, t_gather = (float(v) for v in tt.tolist())
        if rank == 0:
            key = hashlib.sha256(repr([(l["channel"], l["block"], l["text"], l["num_avg"], l["nbadsync"], l["snr"])
                                       for l in merged]).encode()).hexdigest()[:16]
            cfg5 = {"workload": f"cfg5: {args.cfg5_channels} independent synthetic 12 kHz channels x 60 s (seed 1000 + c, 20 pings each), deep-scan flags",
                    "sharding": f"contiguous channel ranges over {world} rank(s) (msk144cudecoder_b200.sharding), one decode_host call per rank "
                                "(pinned host samples in), gather_object of the decode lines to rank 0 + ordered merge, all inside the timed pass",
                    "channels": args.cfg5_channels, "seconds_per_channel": 60.0, "pass_s": t_pass, "decode_s_max_rank": t_dec,
                    "gather_merge_s": t_gather, "realtime_channels": args.cfg5_channels * 60.0 / t_pass,
                    "audio_s_per_s": args.cfg5_channels * 276 * 0.216 / t_pass, "merged_lines": len(merged), "merged_lines_sha16": key,
                    "rank0_kernel_ms": {"frontend": s5["frontend_ms"], "scan": s5["scan_ms"], "ldpc": s5["ldpc_ms"]},
                    "timing": "wall clock between barriers, max over ranks (includes host text decode and the gather); 1 warm-up + 1 timed pass"}
        del x5p

    # ---- ONE stream over all ranks: rank 0's 15-min stream cut into contiguous block ranges (SURVEY.md 8e
    # "Partitioning"); every rank runs the GPU stages of its range from pinned host samples, rank 0 gathers (powers, LDPC
    # successes) and replays SNR + text once in block order; all of it inside the timed pass ----
    one_stream = None
    if not args.no_one_stream:
        x1 = x_dev.clone()
        if world > 1:
            dist.broadcast(x1.view(torch.uint8), src=0)      # (NCCL has no int16) outside the timed region: every rank holds the capture, as it would a file
        x1p = x1.cpu().pin_memory()
        del x1
        x1n = x1p.numpy()

        phase = {"decode_s": 0.0, "replay_s": 0.0}

        def timed_chunk(b0, b1):
            t_a = time.perf_counter()
            out = dec.decode_chunk_host(x1n[2592 * b0: 2592 * (b1 - 1) + 5184], b0, result_cap=1 << 17)
            phase["decode_s"] = time.perf_counter() - t_a
            return out

        def timed_replay(pw, rs):
            t_a = time.perf_counter()
            out = m.replay_channel(pw, rs, compat_cache=True)
            phase["replay_s"] = time.perf_counter() - t_a
            return out

        def one_pass():
            t_a = time.perf_counter()
            out = sharding.decode_time_sharded(timed_chunk, timed_replay, nblocks, rank, world)
            return out, time.perf_counter() - t_a

        one_pass()
        barrier()
        n1 = max(1, min(args.steps, 3))
        t0 = time.perf_counter()
        for _ in range(n1):
            got1, _ = one_pass()
            barrier()
        t1 = torch.tensor([(time.perf_counter() - t0) / n1], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t1, op=dist.ReduceOp.MAX)
        if rank == 0:
            b0r, b1r = m.chunk_ranges(nblocks, world)[0]
            one_stream = {"workload": "the single cfg2 stream of rank 0, blocks cut into contiguous ranges over the ranks "
                                      "(msk144cudecoder_b200.sharding.decode_time_sharded: msk144_decode_chunk_host per rank, "
                                      "gather_object of (powers, LDPC successes) to rank 0, msk144_replay_channel there)",
                          "ranks": world, "blocks": nblocks, "blocks_rank0": b1r - b0r, "passes": n1, "pass_s": float(t1.item()),
                          "audio_s_per_s": audio_s / float(t1.item()), "lines": len(got1),
                          "identical_to_undivided_decode": got1 == lines,
                          "rank0_last_pass": {"decode_chunk_s": phase["decode_s"], "replay_s": phase["replay_s"],
                                              "gather_and_rest_s": max(0.0, float(t1.item()) - phase["decode_s"] - phase["replay_s"])},
                          "timing": "wall clock between barriers, max over ranks, mean of the passes; pinned host samples in, "
                                    "decode lines out on rank 0"}
        del x1p, x1n

    if rank == 0:
        hbm_peak, sm_mhz, peak_src = peaks()
        # dominant kernel by measured device time over the timed region
        per = {"frontend_kernel": st["frontend_ms"], "scan_kernel": st["scan_ms"], "ldpc_kernel": st["ldpc_ms"]}
        dom = max(per, key=per.get)
        groups = max(1, st["groups"])
        launches_per_step = max(1, groups // args.steps)
        ncw = b1["llr_codewords"]          # gated candidates of one pass over the stream (last e2e pass)
        nsucc = b1["ldpc_successes"]
        scan_flops, soft_flops = algorithmic_flops(dec.num_freqs, dec.depth, nblocks)
        # LDPC work actually executed: a codeword that does not converge runs 10 parity checks and the 9 message
        # updates between them; the (few) successes stop earlier and are not credited at all
        ldpc_updates = max(0, ncw - nsucc) * 9
        fp32_nominal = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12
        try:
            fp32_meas = dec.measure_fp32_peak()
        except Exception:  # noqa: BLE001
            fp32_meas = None
        dur = {k: v / 1e3 / args.steps for k, v in per.items()}   # seconds per pass
        # algorithmic work and bytes per pass of each kernel
        flops_alg = {"frontend_kernel": nblocks * 5184 * (26 * 4 + 12), "scan_kernel": scan_flops + soft_flops,
                     "ldpc_kernel": ldpc_updates * 6_200}
        bytes_alg = {"frontend_kernel": nblocks * (5184 * 2 + 5184 * 8 + 32),
                     "scan_kernel": nblocks * (5184 * 8 + dec.items_per_block * 16) + ncw * 544,
                     "ldpc_kernel": ncw * 544 + nsucc * 48}
        ach_tf = {k: flops_alg[k] / dur[k] / 1e12 if dur[k] > 0 else 0.0 for k in per}
        ach_gbs = {k: bytes_alg[k] / dur[k] / 1e9 if dur[k] > 0 else 0.0 for k in per}
        peak_tf = fp32_meas or fp32_nominal
        # dram__bytes_read + dram__bytes_write of the dominant kernel from the committed `ncu --set full` capture
        # (one 128-block launch group of this workload), scaled to this run's average launch
        traffic, ncu_pipes = None, None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
            traffic = tj["dram_bytes_per_launch"][dom] * (nblocks / launches_per_step) / tj["blocks_per_launch"]
            ncu_pipes = {"issue_active_pct": tj.get("issue_active_pct", {}).get(dom), **tj.get("pipe_pct", {}).get(dom, {}),
                         "source": tj.get("source")}
        except Exception:  # noqa: BLE001
            pass
        whole_tf = sum(flops_alg.values()) / (st["total_ms"] / 1e3 / args.steps) / 1e12
        roofline = {"bound": "fp32", "kernel": dom, "achieved": ach_tf[dom], "peak": peak_tf, "unit": "TFLOP/s",
                    "frac": ach_tf[dom] / peak_tf, "traffic": traffic,
                    "work": ("ldpc_kernel: (codewords - successes) x 9 message updates executed x 6,200 flop (SURVEY.md 8d), "
                             f"= {ldpc_updates} updates per step; scan_kernel: 8d's direct form per block"),
                    "peak_source": ("measured on this GPU by msk144_measure_fp32_peak (dependency-free FFMA stream, CUDA events)"
                                    if fp32_meas else "nominal (probe failed)"),
                    "peak_nominal": fp32_nominal, "peak_nominal_source": f"148 SM x 128 lanes x 2 x {sm_mhz:.0f} MHz",
                    "frac_of_nominal": ach_tf[dom] / fp32_nominal,
                    "launches_per_step": launches_per_step, "avg_launch_ms": per[dom] / groups,
                    "algorithmic_flop_per_launch": flops_alg[dom] / launches_per_step,
                    "hbm": {"achieved": ach_gbs[dom], "peak": hbm_peak, "unit": "GB/s", "frac": ach_gbs[dom] / hbm_peak,
                            "algorithmic_bytes_per_launch": bytes_alg[dom] / launches_per_step, "traffic": traffic,
                            "traffic_source": "profiles/ncu_traffic.json", "peak_source": peak_src},
                    "per_kernel": {k: {"ms_per_step": per[k] / args.steps, "share_of_step": per[k] / max(1e-9, st["total_ms"]),
                                       "fp32_tflops": ach_tf[k], "fp32_frac": ach_tf[k] / peak_tf,
                                       "hbm_gbs": ach_gbs[k], "hbm_frac": ach_gbs[k] / hbm_peak} for k in per},
                    "whole_step": {"fp32_tflops": whole_tf, "fp32_frac": whole_tf / peak_tf},
                    "binding": "instruction issue of a packed-FP32 / ALU / MUFU mix (not HBM: algorithmic HBM traffic is ~0.5 KB per "
                               "codeword / 41 KB per block); see DESIGN.md section 4 and profiles/",
                    "ncu_pipes_pct_of_peak": ncu_pipes}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": workload_config(world, args.seconds),
                "batch_blocks": args.batch_blocks, "build": m.build_info(),
                "clocks": clk.summary(),
                "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                        "decode_lines_per_step": len(lines), "messages_found": len(found_texts & sent_texts),
                        "messages_sent": len(sent_texts), "false_texts": len(found_texts - sent_texts)},
                "gpu_launches": int(st["groups"] * 3),
                "ldpc_codewords_per_s": ncw / dur["ldpc_kernel"] if dur["ldpc_kernel"] > 0 else None,
                "ldpc_codewords_per_step": int(ncw)}
        # the sharded legs before the (long) roofline object, so that a truncated tail of the line still carries them
        if one_stream is not None:
            line["one_stream"] = one_stream
        if cfg5 is not None:
            line["cfg5"] = cfg5
        line["roofline"] = roofline
        if world == 1 and not args.no_cli:
            # ---- the drop-in executable on the same bytes, and its output against the reference's ----------
            try:
                ns = int(min(SAMPLE_SECONDS, args.seconds) * 12000)
                sample = np.ascontiguousarray(x[:ns]).tobytes()
                fig, cli_lines = cli_figures([m.CLI_PATH, f"--device={local_rank}"] + CLI_FLAGS, sample,
                                             None if args.no_full_stream else x.tobytes(), min(args.steps, 5), 1,
                                             synth.num_blocks(ns) * 0.216, audio_s)
                line["cli"] = fig
                check = {"sample": f"first {synth.num_blocks(ns)} blocks of the stream through both executables, stdout compared "
                                   "line by line (date masked)", "ours_lines": len(cli_lines) - 1, "ours_sha16": lines_digest(untie(cli_lines))}
                gl = golden_lines()
                check["criterion"] = ("identical = every line equal in text, snr, num_avg, nbadsync and order; f0 / pattern_idx are a "
                                      "tie-break the reference itself does not reproduce between two runs (strict = lines equal in "
                                      "those two fields as well)")
                if gl is not None and args.seconds >= SAMPLE_SECONDS:
                    check["golden"] = compare_lines(cli_lines, gl)   # committed output of the reference (tests/golden)
                if os.path.exists(REF_EXE):
                    try:
                        _, ref_lines = cli_once([REF_EXE] + CLI_FLAGS, sample)   # the checker, outside every timed region
                        check["live"] = compare_lines(cli_lines, ref_lines)
                        _, ref_again = cli_once([REF_EXE] + CLI_FLAGS, sample)
                        check["reference_vs_itself"] = compare_lines(ref_again, ref_lines)
                        check["reference_lines"], check["reference_sha16"] = len(ref_lines) - 1, lines_digest(untie(ref_lines))
                    except Exception as e:  # noqa: BLE001
                        check["live_error"] = " ".join(str(e).split())[:160]
                verdicts = [check[k]["identical"] for k in ("golden", "live") if k in check]
                line["lines_identical_to_reference"] = all(verdicts) if verdicts else None
                line["reference_check"] = check
            except Exception as e:  # noqa: BLE001
                line["cli"] = {"error": " ".join(str(e).split())[:200]}
        if world == 1 and not args.no_cpu_baseline:
            from oracle import orc
            cores = os.cpu_count() or 1
            nb_s = args.cpu_blocks or max(1, min(nblocks, 8 * cores))
            cfg = orc.config(center=1500.0, width=500.0, step=1.0, depth=6, thr=3, scan_mode=1, nthreads=cores)
            t0 = time.perf_counter()
            ref_lines, nb_done = orc.decode_stream(x, cfg, max_blocks=nb_s)
            dt = time.perf_counter() - t0
            ours = sorted((l["block"], l["text"], l["num_avg"], l["nbadsync"], l["snr"]) for l in lines if l["block"] < nb_done)
            theirs = sorted((l["block"], l["text"], l["num_avg"], l["nbadsync"], l["snr"]) for l in ref_lines)
            line["cpu_baseline"] = {"value": nb_done * 0.216 / dt, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"first {nb_done} blocks ({nb_done * 0.216:.1f} s of audio) of the same stream, "
                                              f"C++ oracle port on {cores} threads, {dt:.1f} s; WSJT-X Fortran decoder "
                                              "unavailable on this box (empty submodule, no gfortran)",
                                    "lines_identical_to_gpu": ours == theirs}
            # LDPC stage alone on the host cores (SURVEY.md 8d "CPU baseline"): the oracle's restatement of
            # ldpc_kernel.cuh:100-249 and of WSJT-X's bpdecode128_90 control flow, all threads
            try:
                ncw_cpu = 1 << 16
                rng = np.random.default_rng(7)
                noise = (5.5 * rng.standard_normal((ncw_cpu, 128))).astype(np.float32)   # like gate-passing candidates: none converges
                sig, _ = synth.cfg4_llrs(ncw_cpu, 3.0, seed=149)
                ld = {}
                for name, fn, arr in (("noise_reference_semantics", orc.ldpc_decode_batch, noise),
                                      ("noise_wsjtx_semantics", orc.ldpc_decode_wsjtx_batch, noise),
                                      ("cfg4_ebn0_3db_reference_semantics", orc.ldpc_decode_batch, sig)):
                    fn(arr[:1024], nthreads=cores)
                    t0 = time.perf_counter()
                    fn(arr, nthreads=cores)
                    ld[name] = ncw_cpu / (time.perf_counter() - t0)
                line["cpu_baseline"]["ldpc_cw_per_s"] = dict(ld, codewords=ncw_cpu, cores=cores,
                                                             gpu_same_metric=line["ldpc_codewords_per_s"])
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"]["ldpc_cw_per_s"] = {"error": " ".join(str(e).split())[:200]}
        print(json.dumps(line), flush=True)
    dec.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
