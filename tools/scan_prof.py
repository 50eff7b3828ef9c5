#!/usr/bin/env python3
"""Per-phase cycle breakdown of scan_kernel (needs a library built with -DMSK_SCAN_PROF, MSK144_B200_LIB=...)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import msk144cudecoder_b200 as m  # noqa: E402
from msk144cudecoder_b200 import synth  # noqa: E402

x, _ = synth.cfg2_deep_audio(seconds=60.0, seed=144)
xd = torch.from_numpy(x).to("cuda:0")
dec = m.Decoder(compat_cache=True, max_batch_blocks=128, center=1500.0, width=500.0, step=1.0, depth=6, thr=3)
s = torch.cuda.Stream()
L = m.lib()
out = (C.c_ulonglong * 16)()
dec.run_device(xd.data_ptr(), len(x), 1, s.cuda_stream)
L.msk144_debug_scan_prof(out)
nb = dec.run_device(xd.data_ptr(), len(x), 1, s.cuda_stream)
L.msk144_debug_scan_prof(out)
names = ["prologue S1+S2a", "S2b", "S3", "S4 phase: barrier wait", "S2a next (thread 0)", "S5 (warp 0)", "barrier wait after S5",
         "S4: slice maxima", "S4: hypot", "S4: top-8 replay", "S4: shuffles, atan2, sincos", "S4: soft bit 0", "-", "-", "-", "-"]
tot = sum(out)
items = nb * 501
for n, v in zip(names, out):
    print(f"{n:24s} {v / items:9.0f} cycles per (block, frequency)  {100 * v / tot:5.1f}%")
print("total", tot / items)
