#!/usr/bin/env python3
"""The other BASELINE.json configurations on one GPU, one JSON line each (bench.py is configs[1]):

  cfg1  60 s default-scan audio (stand-in for demo/0001.wav): width 100, depth 3
  cfg3  15-min s8 IQ stream, deep scan
  cfg4  10^7 LDPC codewords, Eb/N0 = 0..7 dB (1.25 M per point), codewords/s per point and overall,
        with the HBM roofline of the high-SNR end (528 B per codeword)
  cfg5  1024 channels x 60 s, deep scan, ONE decoder call (host samples in, lines out)

    python tools/bench_configs.py [cfg1 cfg3 cfg4 cfg5]
Timing: CUDA events / wall clock around the public calls, one warm-up pass, inputs larger than L2 except cfg1.
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import msk144cudecoder_b200 as m  # noqa: E402
from msk144cudecoder_b200 import synth  # noqa: E402

DEEP = dict(width=500.0, step=1.0, depth=6, thr=3)


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:  # noqa: BLE001
        return 6650.0


def stream_bench(name, x, pings, per_sample, reps=3, **kw):
    dev = torch.device("cuda", 0)
    nsamp = len(x) // per_sample
    nb = synth.num_blocks(nsamp)
    xp = torch.from_numpy(x).pin_memory()
    xd = xp.to(dev)
    with m.Decoder(compat_cache=True, max_batch_blocks=128, **kw) as d:
        s = torch.cuda.Stream(dev)
        d.run_device(xd.data_ptr(), nsamp, 1, s.cuda_stream)
        torch.cuda.synchronize()
        d.stage_times()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(reps):
            d.run_device(xd.data_ptr(), nsamp, 1, s.cuda_stream)
        e1.record(s)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        st = d.stage_times()
        d.decode_host(xp.numpy(), line_cap=1 << 15, result_cap=0)
        t0 = time.perf_counter()
        for _ in range(reps):
            lines, _ = d.decode_host(xp.numpy(), line_cap=1 << 15, result_cap=0)
        e2e = (time.perf_counter() - t0) / reps
    sent, got = {p.text for p in pings}, {l["text"] for l in lines}
    print(json.dumps({"bench": name, "blocks": nb, "audio_s": nb * 0.216, "rtf_resident": nb * 0.216 / (ms * 1e-3),
                      "rtf_e2e": nb * 0.216 / e2e, "ms": ms,
                      "share": {k: st[k] / max(1e-9, st["total_ms"]) for k in ("frontend_ms", "scan_ms", "ldpc_ms")},
                      "lines": len(lines), "messages_found": len(sent & got), "messages_sent": len(sent),
                      "false_texts": len(got - sent)}), flush=True)


def encode_gpu(msgs):
    n, dev = msgs.shape[0], msgs.device
    rem = torch.zeros(n, dtype=torch.int64, device=dev)
    ext = torch.cat([msgs.to(torch.int64), torch.zeros((n, 19), dtype=torch.int64, device=dev)], dim=1)
    for k in range(96):
        rem = (rem << 1) | ext[:, k]
        rem = torch.where((rem & 0x2000) != 0, rem ^ (0x2000 | 0x15D7), rem)
    rem &= 0x1FFF
    crc = torch.stack([(rem >> (12 - i)) & 1 for i in range(13)], dim=1)
    m90 = torch.cat([msgs.to(torch.int64), crc], dim=1).to(torch.float32)
    gen = torch.from_numpy(synth._GEN.astype(np.float32)).to(dev)
    return torch.cat([m90, torch.remainder(m90 @ gen.T, 2.0)], dim=1)


def cfg4():
    dev = torch.device("cuda", 0)
    per_point, reps = 1_250_000, 5
    g = torch.Generator(device=dev).manual_seed(146)
    peak = hbm_peak()
    tot_ms, rows = 0.0, []
    with m.Decoder(compat_cache=True, width=20, depth=1) as d:
        s = torch.cuda.Stream(dev)
        out = torch.empty((per_point, 4), dtype=torch.int32, device=dev)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        for ebn0 in range(8):
            msgs = torch.randint(0, 2, (per_point, 77), device=dev, generator=g, dtype=torch.uint8)
            cws = encode_gpu(msgs)
            sigma = float(np.sqrt(1.0 / (2.0 * (90.0 / 128.0) * 10 ** (ebn0 / 10))))
            llr = (2.0 * ((2.0 * cws - 1.0) + sigma * torch.randn((per_point, 128), device=dev, generator=g)) / sigma ** 2).contiguous()
            torch.cuda.synchronize()
            d.ldpc_decode_device(llr.data_ptr(), per_point, out.data_ptr(), s.cuda_stream)
            torch.cuda.synchronize()
            ms = 0.0
            for _ in range(reps):           # 640 MB of LLRs per pass > 126 MB L2; flushed anyway
                with torch.cuda.stream(s):
                    flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(s)
                d.ldpc_decode_device(llr.data_ptr(), per_point, out.data_ptr(), s.cuda_stream)
                e1.record(s)
                torch.cuda.synchronize()
                ms += e0.elapsed_time(e1) / reps
            found = out[:, 2] < 0
            iters = (out[:, 3] & 0xFF).float()
            idx = torch.arange(77, device=dev)
            w = out[:, :3].to(torch.int64) & 0xFFFFFFFF
            dec = ((w[:, idx >> 5] >> (idx & 31)) & 1).to(torch.uint8)
            good = found & (dec == msgs).all(dim=1)
            row = {"bench": "cfg4_ldpc", "ebn0_db": ebn0, "codewords": per_point, "ms": ms, "cw_per_s": per_point / ms * 1e3,
                   "decoded_frac": float(found.float().mean()), "correct_frac": float(good.float().mean()),
                   "undetected_errors": int((found & ~good).sum()), "mean_iters": float(iters.mean()),
                   "hbm_gbs": per_point * 528 / ms / 1e6, "hbm_frac": per_point * 528 / ms / 1e6 / peak}
            rows.append(row)
            tot_ms += ms
            print(json.dumps(row), flush=True)
    n = per_point * 8
    print(json.dumps({"bench": "cfg4_ldpc_total", "codewords": n, "ms": tot_ms, "cw_per_s": n / tot_ms * 1e3,
                      "hbm_peak_gbs": peak}), flush=True)


def cfg5(nchan=1024, seconds=60.0):
    t0 = time.perf_counter()
    chans, pings = [], []
    for c in range(nchan):
        x, p = synth.cfg5_channel(c, seconds=seconds)
        chans.append(x)
        pings.append({q.text for q in p})
    x = torch.from_numpy(np.concatenate(chans)).pin_memory()
    gen_s = time.perf_counter() - t0
    nb = synth.num_blocks(len(chans[0])) * nchan
    with m.Decoder(compat_cache=True, center=1500.0, max_batch_blocks=128, **DEEP) as d:
        d.decode_host(x.numpy()[: len(chans[0]) * 8], nchan=8, line_cap=1 << 12, result_cap=0)   # warm-up
        t0 = time.perf_counter()
        lines, _ = d.decode_host(x.numpy(), nchan=nchan, line_cap=1 << 19, result_cap=0)
        dt = time.perf_counter() - t0
    found = [set() for _ in range(nchan)]
    for l in lines:
        found[l["channel"]].add(l["text"])
    print(json.dumps({"bench": "cfg5_channels", "channels": nchan, "seconds_per_channel": seconds, "blocks": nb,
                      "audio_s": nb * 0.216, "wall_s": dt, "rtf_e2e": nb * 0.216 / dt, "channels_realtime": nb * 0.216 / dt,
                      "lines": len(lines), "messages_found": sum(len(f & s) for f, s in zip(found, pings)),
                      "messages_sent": sum(len(s) for s in pings), "false_texts": sum(len(f - s) for f, s in zip(found, pings)),
                      "h2d_bytes": int(x.numel() * 2), "synth_s": gen_s}), flush=True)


def deinterleave(nchan=1024, nsamp=720000, reps=20):
    """The ingest adapter for interleaved multi-channel captures: a pure transpose, bound by HBM (2 B in + 2 B out per sample)."""
    dev = torch.device("cuda", 0)
    x = torch.randint(-30000, 30000, (nsamp, nchan), dtype=torch.int16, device=dev)
    out = torch.empty((nchan, nsamp), dtype=torch.int16, device=dev)
    with m.Decoder(compat_cache=True, width=20, depth=1) as d:
        s = torch.cuda.Stream(dev)
        torch.cuda.synchronize()
        d.deinterleave_device(x.data_ptr(), out.data_ptr(), nsamp, nchan, s.cuda_stream)
        torch.cuda.synchronize()
        ok = bool(torch.equal(out, x.T))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(reps):      # 1.47 GB in + 1.47 GB out per pass: far larger than L2
            d.deinterleave_device(x.data_ptr(), out.data_ptr(), nsamp, nchan, s.cuda_stream)
        e1.record(s)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
    gbs = 2 * x.numel() * 2 / ms / 1e6
    print(json.dumps({"bench": "deinterleave", "channels": nchan, "samples_per_channel": nsamp, "ms": ms, "hbm_gbs": gbs,
                      "hbm_peak_gbs": hbm_peak(), "hbm_frac": gbs / hbm_peak(), "correct": ok}), flush=True)


def latency(seconds=30.0):
    """Real-time use: one 2592-sample hop pushed at a time (what arrives every 216 ms), wall clock push -> lines polled."""
    x, _ = synth.cfg2_deep_audio(seconds=seconds, seed=144)
    raw = x.tobytes()
    hop = 2592 * 2
    for name, kw in (("deep", dict(center=1500.0, **DEEP)), ("default", dict(width=100.0, step=2.0, depth=3, thr=1))):
        with m.Decoder(compat_cache=True, max_batch_blocks=8, **kw) as d:
            d.push(raw[: 5184 * 2])
            d.poll()
            ts = []
            for i in range(5184 * 2, len(raw) - hop + 1, hop):
                t0 = time.perf_counter()
                d.push(raw[i:i + hop])
                d.poll()
                ts.append(time.perf_counter() - t0)
        ts = np.array(ts[3:]) * 1e3
        print(json.dumps({"bench": "latency_per_block_" + name, "blocks": len(ts), "ms_median": float(np.median(ts)),
                          "ms_p99": float(np.percentile(ts, 99)), "ms_max": float(ts.max()), "budget_ms": 216.0}), flush=True)


def main():
    which = sys.argv[1:] or ["cfg1", "cfg3", "cfg4", "cfg5", "deinterleave", "latency"]
    if "latency" in which:
        latency()
    if "deinterleave" in which:
        deinterleave()
    if "cfg1" in which:
        x, p = synth.cfg1_demo()
        stream_bench("cfg1_demo_standin", x, p, 1, reps=10, center=1500.0, width=100.0, step=2.0, depth=3, thr=1)
    if "cfg3" in which:
        x, p = synth.cfg3_deep_iq()
        stream_bench("cfg3_iq_deep", x, p, 2, center=0.0, read_mode=2, **DEEP)
    if "cfg4" in which:
        cfg4()
    if "cfg5" in which:
        cfg5()


if __name__ == "__main__":
    main()
