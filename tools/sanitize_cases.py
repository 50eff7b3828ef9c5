#!/usr/bin/env python3
"""Small invocations of every kernel of libmsk144b200.so, meant to run under compute-sanitizer:

    compute-sanitizer --tool memcheck|racecheck|synccheck|initcheck python tools/sanitize_cases.py [case ...]

Sizes are a few blocks / a few hundred codewords: racecheck slows the shared-memory kernels down by two orders of
magnitude.  Each case prints one line; the sanitizer's own summary ("ERROR SUMMARY: n errors") is what counts.
Cases: pipe3 (frontend_kernel, scan_kernel<3>, ldpc_kernel<false>), deep8 (fft_frontend_kernel, scan_kernel<8> at
step 1.5), iq6 (IQ front end, scan_kernel<6>), rms (rms_kernel: more blocks than SMs), stage (msk144_frontend_host,
msk144_scan_host without twin sharing, ldpc_stream_kernel on a mixed-SNR batch), wsjtx (ldpc_kernel<true> with an
a-priori mask), deint (deinterleave_kernel, vector and ragged shapes), chz (channeliser: direct kernel for 1, 2, 3
channels, FFT kernel), noalloc (msk144_frontend_device_noalloc)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import msk144cudecoder_b200 as m  # noqa: E402
from msk144cudecoder_b200 import synth  # noqa: E402

HOP, N = 2592, 5184


def pipe3():
    x, _ = synth.cfg1_demo(seconds=2.0, seed=1)
    with m.Decoder(compat_cache=True, width=40, step=2, depth=3, thr=2) as d:
        lines, res = d.decode_host(x, result_cap=4096)
    return f"lines={len(lines)} results={len(res)}"


def deep8():
    x, _ = synth.cfg2_deep_audio(seconds=1.0, seed=144, npings=2)
    with m.Decoder(compat_cache=True, width=9, step=1.5, depth=8, thr=3, method=1) as d:
        lines, res = d.decode_host(x, result_cap=4096)
    return f"lines={len(lines)} results={len(res)}"


def iq6():
    x, _ = synth.cfg3_deep_iq(seconds=1.0, seed=145, npings=2)
    with m.Decoder(compat_cache=True, read_mode=2, center=0.0, width=12, step=2, depth=6, thr=3) as d:
        lines, res = d.decode_host(x, result_cap=4096)
    return f"lines={len(lines)} results={len(res)}"


def rms():
    x, _ = synth.cfg1_demo(seconds=40.0, seed=2)           # 184 blocks > 148 SMs in one launch group
    with m.Decoder(compat_cache=True, width=4, step=2, depth=1, thr=1, max_batch_blocks=256) as d:
        lines, _ = d.decode_host(x)
    with m.Decoder(compat_cache=True, width=4, step=2, depth=1, thr=1, max_batch_blocks=256, method=1) as d:
        lines1, _ = d.decode_host(x[1:])                   # unaligned rows for the scalar load path
    return f"lines={len(lines)}/{len(lines1)}"


def stage():
    x, _ = synth.cfg1_demo(seconds=1.0, seed=3)
    with m.Decoder(width=20, step=2, depth=4, thr=2) as d:
        blocks, powers = d.frontend(x)
        cands, llr, item = d.scan(blocks)
        noise = np.random.default_rng(1).standard_normal((300, 128)).astype(np.float32) * 5.5
        hi, _ = synth.cfg4_llrs(201, 7.0, seed=1)
        lo, _ = synth.cfg4_llrs(200, 2.0, seed=2)
        batch = np.concatenate([hi, noise, lo, llr[:99]])
        found, msg, iters, nhard = d.ldpc_decode(batch)
    return f"blocks={blocks.shape[0]} gated={llr.shape[0]} ldpc_found={int(found.sum())}/{batch.shape[0]}"


def wsjtx():
    hi, _ = synth.cfg4_llrs(150, 4.0, seed=3)
    lo, _ = synth.cfg4_llrs(151, 1.0, seed=4)
    with m.Decoder(width=20, depth=1, wsjtx_bp=True) as d:
        mask = np.zeros(128, np.uint8)
        mask[:8] = 1
        d.set_ap_mask(mask)
        found, *_ = d.ldpc_decode(np.concatenate([hi, lo]))
        x, _ = synth.cfg1_demo(seconds=1.0, seed=4)
        d.set_ap_mask(None)
        lines, _ = d.decode_host(x)
    return f"found={int(found.sum())}/301 lines={len(lines)}"


def deint():
    import torch
    out = []
    with m.Decoder(width=20, depth=1) as d:
        for nchan, spc in ((4, 5184), (3, 5190), (1, 777), (33, 65), (64, 64)):
            x = np.random.default_rng(nchan).integers(-3000, 3000, (spc, nchan)).astype(np.int16)
            d_in = torch.from_numpy(x).cuda()
            d_out = torch.zeros((nchan, spc), dtype=torch.int16, device="cuda")
            d.deinterleave_device(d_in.data_ptr(), d_out.data_ptr(), spc, nchan, torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            out.append(bool(np.array_equal(d_out.cpu().numpy(), x.T)))
    return f"ok={out}"


def chz():
    res = []
    for bins, M, fmt in (([0], 16, m.IQ_U8), ([1, -2], 16, m.IQ_S8), ([0, 3, -5], 20, m.IQ_S16), (list(range(-6, 7)), 16, m.IQ_U8),
                         (list(range(-40, 41, 2)), 160, m.IQ_U8)):
        cap, _ = synth.wideband_capture(bins[:2], decimation=M, seconds=0.05, seed=3, fmt=fmt)
        with m.Channelizer(M, bins, input_format=fmt) as z:
            y = z.channelize_host(cap)
            z.reset()
            raw = cap.tobytes()
            y2 = [z.push(raw[:len(raw) // 3 + 1]), z.push(raw[len(raw) // 3 + 1:])]
        res.append(f"{len(bins)}ch:{y.shape[1]}+{sum(v.shape[0] for v in y2)}")
    return " ".join(res)


def noalloc():
    import torch
    x, _ = synth.cfg1_demo(seconds=3.0, seed=5)
    nb = (x.size - N) // HOP + 1
    d_in = torch.from_numpy(x).cuda()
    d_blocks = torch.zeros((nb, N, 2), dtype=torch.float32, device="cuda")
    d_pow = torch.zeros((nb, 8), dtype=torch.float32, device="cuda")
    for _ in range(2):
        got = m.frontend_device_noalloc(d_in.data_ptr(), x.size, 1, d_blocks.data_ptr(), d_pow.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
    with m.Decoder(width=20, depth=1) as d:
        ref, _ = d.frontend(x)
    same = bool(np.array_equal(np.ascontiguousarray(d_blocks.cpu().numpy()).view(np.complex64)[..., 0], ref))
    return f"blocks={got} same={same}"


CASES = dict(pipe3=pipe3, deep8=deep8, iq6=iq6, rms=rms, stage=stage, wsjtx=wsjtx, deint=deint, chz=chz, noalloc=noalloc)

if __name__ == "__main__":
    m.lib()
    for name in (sys.argv[1:] or list(CASES)):
        print(f"case {name}: {CASES[name]()}", flush=True)
