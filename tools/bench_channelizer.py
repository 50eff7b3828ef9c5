#!/usr/bin/env python3
"""Channeliser throughput on one GPU: resident u8 capture at 1.92 Msps (the README's rtl_sdr rate), CUDA events on the
launching stream, input larger than L2.  One JSON line per channel count with the HBM and FP32 rooflines:
  bytes  = capture bytes read once + channel bytes written once            (algorithmic)
  flops  = frames x (32 M [branch filters] + 8 M per channel [direct kernel] or 5 M log2 M [FFT kernel, > 4 channels])"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import msk144cudecoder_b200 as m  # noqa: E402


def main():
    seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
    M = 160
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm_peak = float(peaks.get("hbm_gbs", 6540.0))
    with m.Decoder(compat_cache=True, width=20, depth=1) as d:
        fp32_peak = d.measure_fp32_peak()
    nframes = int(seconds * 12000)
    g = torch.Generator(device="cuda").manual_seed(1)
    cap = torch.randint(96, 160, (nframes * M * 2,), dtype=torch.uint8, device="cuda", generator=g)
    stride = (nframes + 7) // 8 * 8
    only = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else None
    for bins in ([0], [0, 5], [0, 5, -17, 40], list(range(8)), list(range(-16, 16)), list(range(-80, 80))):
        if only and len(bins) not in only:
            continue
        nchan = len(bins)
        out = torch.zeros((nchan, stride, 2), dtype=torch.int8, device="cuda")
        with m.Channelizer(M, bins, gain=8.0) as z:
            s = torch.cuda.Stream()
            for _ in range(3):
                z.channelize_device(cap.data_ptr(), nframes, out.data_ptr(), stride, 1, 0, s.cuda_stream)
            torch.cuda.synchronize()
            reps = 10
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s)
            for _ in range(reps):
                z.channelize_device(cap.data_ptr(), nframes, out.data_ptr(), stride, 1, 0, s.cuda_stream)
            e1.record(s)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / reps
        fft = nchan > 4
        nch = 4 if nchan > 2 else nchan
        groups = 1 if fft else (nchan + nch - 1) // nch
        bytes_alg = cap.numel() + nchan * nframes * 2
        # branch filters 16 FMA per sample; direct: 4 FMA per (sample, channel); FFT 8 x 4 x 5: ~5 M log2 M real operations
        flops = nframes * (2 * 16 * M + (5 * M * np.log2(M) if fft else 8 * M * nchan))
        print(json.dumps({"bench": "channelizer", "lib": os.path.basename(m.LIB_PATH), "decimation": M, "input": "u8 I/Q 1.92 Msps",
                          "seconds_of_capture": seconds, "channels": nchan, "ms": ms, "x_realtime": seconds / (ms * 1e-3),
                          "input_msps": nframes * M / ms / 1e3,
                          "roofline": {"bound": "hbm", "achieved": bytes_alg / ms / 1e6, "peak": hbm_peak, "unit": "GB/s",
                                       "frac": bytes_alg / ms / 1e6 / hbm_peak},
                          "kernel": "fft" if fft else "direct", "fp32": {"achieved": flops / ms / 1e9, "peak": fp32_peak, "unit": "TFLOP/s", "frac": flops / ms / 1e9 / fp32_peak},
                          "checksum": int(out.to(torch.int32).sum().item())}), flush=True)


if __name__ == "__main__":
    main()
