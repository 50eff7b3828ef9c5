#!/usr/bin/env python3
"""Yardstick for the channeliser's tensor-core decision: how long would the DFT over the M = 160
polyphase branches take as a GEMM on the tensor cores?  cuBLAS (torch.matmul) stands in for the best tcgen05 kernel one
could write: frames x 2M (re, im of every branch) times 2M x 2M, FP16 and BF16 operands with FP32 accumulation, and TF32.
One such product carries 11 (FP16) / 8 (BF16) significant bits of the twiddles and of the branch outputs; parity of the s8
channel samples with the FP32 path needs the hi / lo split of both operands, i.e. 3 products (hi*hi, hi*lo, lo*hi).
Not part of the product: the shipped channeliser is the FP32 Stockham FFT (csrc/channelizer.cu, chz_fft.h)."""
import json

import torch

M, SECONDS = 160, 120.0
frames = int(SECONDS * 12000)
dev = torch.device("cuda", 0)
for name, dt in (("fp16", torch.float16), ("bf16", torch.bfloat16), ("tf32", torch.float32)):
    torch.backends.cuda.matmul.allow_tf32 = True
    a = torch.randn(frames, 2 * M, device=dev, dtype=dt)
    w = torch.randn(2 * M, 2 * M, device=dev, dtype=dt)
    for _ in range(3):
        c = a @ w
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        c = a @ w
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    flop = 2.0 * frames * (2 * M) * (2 * M)
    print(json.dumps({"bench": "dft_as_gemm_bound", "operands": name, "frames": frames, "k": 2 * M, "n": 2 * M, "ms_one_product": ms,
                      "tflops": flop / ms / 1e9, "ms_three_products_hi_lo_split": 3 * ms,
                      "bytes_in_out_gb": (a.numel() * a.element_size() + c.numel() * c.element_size()) / 1e9,
                      "compare": "fft_kernel, all 160 channels of the same 120 s capture, branch filters and s8 stores included: 1.36 ms "
                                 "(profiles/r02_channelizer_bench.jsonl)"}), flush=True)
    del a, w, c
