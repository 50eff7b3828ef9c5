#!/usr/bin/env python3
"""cfg5 through the IN-PROCESS multi-GPU pool (msk144_pool_*, csrc/host/pool.cpp): one process, one host thread and one
set of CUDA streams per GPU, channels split into contiguous ranges, decode lines merged on the host in single-decoder
order.  This is the "per-GPU CUDA streams and a host-side gather" form of the north star; bench.py's cfg5 leg is the
one-process-per-GPU form of the same partition.

    python tools/bench_pool.py [channels] [seconds]        (default 256 channels x 60 s, deep-scan flags)

For every N in 1, 2, 4, 8 that the box has GPUs for: one warm-up pass (buffers grow), one timed pass (wall clock around
msk144_pool_decode_host: host samples in, merged lines out, text decode included).  Strong scaling: the channels are
fixed, N grows.  One JSON line per N; the merged lines of every N must be identical to N = 1's."""
import hashlib
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from msk144cudecoder_b200 import synth  # noqa: E402

SECONDS = float(sys.argv[2]) if len(sys.argv) > 2 else 60.0


def _one(c):
    return synth.cfg5_channel(c, seconds=SECONDS)[0]


def main():
    nchan = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    with mp.get_context("fork").Pool(min(os.cpu_count() or 1, 32)) as pool:      # before CUDA exists in this process
        x = np.stack(pool.map(_one, range(nchan), chunksize=4))
    import torch
    import msk144cudecoder_b200 as m
    ngpu = torch.cuda.device_count()
    flags = dict(center=1500.0, width=500.0, step=1.0, depth=6, thr=3)
    base = None
    # optional third argument: an explicit member list such as 0,0 (two decoders sharing GPU 0) instead of the 1/2/4/8 sweep
    lists = [[int(v) for v in sys.argv[3].split(",")]] if len(sys.argv) > 3 else [list(range(n)) for n in (1, 2, 4, 8) if n <= ngpu]
    for devs in lists:
        n = len(devs)
        with m.Pool(devs, compat_cache=True, **flags) as p:
            p.decode_host(x, nchan=nchan, line_cap=1 << 17)
            t0 = time.perf_counter()
            lines, _ = p.decode_host(x, nchan=nchan, line_cap=1 << 17)
            dt = time.perf_counter() - t0
        key = hashlib.sha256(repr([(l["channel"], l["block"], l["text"], l["snr"], l["num_avg"], l["nbadsync"])
                                   for l in lines]).encode()).hexdigest()[:16]
        if base is None:
            base = (dt, key)
        print(json.dumps({"bench": "cfg5_pool", "n_gpus": len(set(devs)), "members": devs,
                          "env": {k: os.environ[k] for k in ("MSK144_SCAN_FORM", "MSK144_LDPC_CTAS") if k in os.environ}, "channels": nchan, "seconds_per_channel": SECONDS,
                          "pass_s": dt, "realtime_channels": nchan * SECONDS / dt,
                          "audio_s_per_s": nchan * int(SECONDS * 12000 / 2592 - 1) * 0.216 / dt,
                          "speedup_vs_1": base[0] / dt, "efficiency": base[0] / dt / n, "merged_lines": len(lines),
                          "merged_lines_sha16": key, "identical_to_n1": key == base[1],
                          "timing": "wall clock around msk144_pool_decode_host (pinned staging, H2D, kernels, D2H, text "
                                    "decode and the ordered merge inside); 1 warm-up + 1 timed pass"}), flush=True)


if __name__ == "__main__":
    main()
