#!/usr/bin/env python3
"""Stage micro-benchmarks on one GPU (CUDA events, inputs resident):
   ldpc: codewords/s of ldpc_kernel on BASELINE config 4 (Eb/N0 sweep) and on gate-passing noise
   scan: ms per block of frontend/scan/ldpc on a 60 s deep-scan stream
Prints one JSON line per measurement."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import msk144cudecoder_b200 as m  # noqa: E402
from msk144cudecoder_b200 import synth  # noqa: E402


def time_ldpc(dec, llr_dev, out_dev, n, reps=5):
    s = torch.cuda.Stream()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dec.ldpc_decode_device(llr_dev.data_ptr(), n, out_dev.data_ptr(), s.cuda_stream)
    torch.cuda.synchronize()
    e0.record(s)
    for _ in range(reps):
        dec.ldpc_decode_device(llr_dev.data_ptr(), n, out_dev.data_ptr(), s.cuda_stream)
    e1.record(s)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    which = sys.argv[1:] or ["ldpc", "scan"]
    dev = torch.device("cuda", 0)
    if "ldpc" in which or "noise" in which:
        dec = m.Decoder(compat_cache=True, width=20, depth=1)
        n = 1 << 21
        out = torch.empty((n, 4), dtype=torch.int32, device=dev)
        g = torch.Generator(device=dev).manual_seed(1)
        noise = torch.randn((n, 128), device=dev, generator=g) * 5.5     # LLR rms like the soft-bit stage
        ms = time_ldpc(dec, noise, out, n)
        print(json.dumps({"bench": "ldpc_noise", "lib": os.path.basename(m.LIB_PATH), "form": "pair" if os.environ.get("MSK144_LDPC_NO_REFILL") else "stream",
                          "codewords": n, "ms": ms, "cw_per_s": n / ms * 1e3,
                          "found": int((out[:, 2] < 0).sum().item()),
                          "out_sha": __import__("hashlib").sha256(out.cpu().numpy().tobytes()).hexdigest()[:12]}), flush=True)
    if "ldpc" in which:
        base, msgs = synth.cfg4_llrs(1 << 15, 0.0, seed=146)
        for ebn0 in (0, 1, 2, 3, 4, 5, 6, 7):
            llr_h, _ = synth.cfg4_llrs(1 << 15, float(ebn0), seed=146 + ebn0)
            llr = torch.from_numpy(llr_h).to(dev).repeat(n >> 15, 1).contiguous()
            ms = time_ldpc(dec, llr, out, n)
            found = int((out[:, 2] < 0).sum().item())
            iters = (out[:, 3] & 0xff).float().mean().item()
            print(json.dumps({"bench": "ldpc_cfg4", "ebn0_db": ebn0, "codewords": n, "ms": ms, "cw_per_s": n / ms * 1e3,
                              "decoded_frac": found / n, "mean_iters": iters,
                              "hbm_gbs": n * 528 / ms / 1e6}), flush=True)
        dec.close()
    if "frontend" in which:
        # front-end kernels alone (CUDA events around the kernel inside the library): method 2 (frontend_kernel) and
        # method 1 (fft_frontend_kernel), launch groups of 128 blocks (the stream pipeline's) and of 4096 (cfg5's scale)
        x1, _ = synth.cfg2_deep_audio(seconds=60.0, seed=144)
        for nchan, batch in ((1, 128), (16, 4096)):
            xs = np.ascontiguousarray(np.tile(x1, nchan))
            xd = torch.from_numpy(xs).to(dev)
            for method in (2, 1):
                dec = m.Decoder(compat_cache=True, max_batch_blocks=batch, width=20.0, step=2.0, depth=1, thr=1, method=method)
                s = torch.cuda.Stream()
                for _ in range(2):
                    nb = dec.run_device(xd.data_ptr(), len(x1), nchan, s.cuda_stream)
                torch.cuda.synchronize()
                dec.stage_times()
                for _ in range(5):
                    dec.run_device(xd.data_ptr(), len(x1), nchan, s.cuda_stream)
                st = dec.stage_times()
                us = st["frontend_ms"] * 1e3 / st["blocks"]
                print(json.dumps({"bench": "frontend", "lib": os.path.basename(m.LIB_PATH), "method": method, "blocks_per_launch": min(batch, nb),
                                  "us_per_block": us, "algorithmic_gbs": (5184 * 2 + 5184 * 8 + 32) / us / 1e3,
                                  "launch_ms": st["frontend_ms"] / st["groups"]}), flush=True)
                dec.close()
    if "scan" in which:
        x, _ = synth.cfg2_deep_audio(seconds=60.0, seed=144)
        xd = torch.from_numpy(x).to(dev)
        for name, kw in [("deep", dict(center=1500.0, width=500.0, step=1.0, depth=6, thr=3)),
                         ("cfg1", dict(center=1500.0, width=100.0, step=2.0, depth=3, thr=1))]:
            dec = m.Decoder(compat_cache=True, max_batch_blocks=128, **kw)
            s = torch.cuda.Stream()
            for _ in range(2):
                nb = dec.run_device(xd.data_ptr(), len(x), 1, s.cuda_stream)
            torch.cuda.synchronize()
            dec.stage_times()
            for _ in range(3):
                dec.run_device(xd.data_ptr(), len(x), 1, s.cuda_stream)
            st = dec.stage_times()
            print(json.dumps({"bench": "stages_" + name, "lib": os.path.basename(m.LIB_PATH), "blocks": nb,
                              "us_per_block": {k: st[k] * 1e3 / st["blocks"] for k in ("frontend_ms", "scan_ms", "ldpc_ms", "total_ms")},
                              "rtf": 0.216 / (st["total_ms"] * 1e-3 / st["blocks"])}), flush=True)
            dec.close()


if __name__ == "__main__":
    main()
