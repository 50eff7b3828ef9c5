#!/usr/bin/env python3
"""Per-phase cycle breakdown of ldpc_kernel on gate-passing noise (library built with -DMSK_LDPC_PROF)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import msk144cudecoder_b200 as m  # noqa: E402

dec = m.Decoder(compat_cache=True, width=20, depth=1)
n = 1 << 21
dev = torch.device("cuda", 0)
out = torch.empty((n, 4), dtype=torch.int32, device=dev)
g = torch.Generator(device=dev).manual_seed(1)
noise = torch.randn((n, 128), device=dev, generator=g) * 5.5
s = torch.cuda.Stream()
L = m.lib()
prof = (C.c_ulonglong * 8)()
dec.ldpc_decode_device(noise.data_ptr(), n, out.data_ptr(), s.cuda_stream)
L.msk144_debug_ldpc_prof(prof)
dec.ldpc_decode_device(noise.data_ptr(), n, out.data_ptr(), s.cuda_stream)
L.msk144_debug_ldpc_prof(prof)
names = ["pair setup (LLR loads)", "zn + ballots", "parities + tanh + scatter (+ check)", "round 1", "round 2", "gather", "output", "-"]
tot = sum(prof)
pairs = n // 2
for nm, v in zip(names, prof):
    print(f"{nm:38s} {v / pairs:9.0f} cycles per codeword pair  {100 * v / tot:5.1f}%")
print("total per pair", tot / pairs, " (per iteration: /9 message updates)")
