import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    """The CPU oracle (test infrastructure)."""
    from oracle import orc as o
    o.build()
    o.lib()
    return o


@pytest.fixture(scope="session")
def product():
    """The product package with its CUDA library built in-tree."""
    import msk144cudecoder_b200 as m
    if not os.path.exists(m.LIB_PATH):
        m.build()
    m.lib()
    return m


@pytest.fixture(scope="session")
def synth():
    from msk144cudecoder_b200 import synth as s
    return s


def one_ping_audio(synth, f0=1512.0, snr_db=6.0, start=3000, nframes=6, text="K1ABC W9XYZ EN37", nsamples=5184 * 2,
                   seed=3, phase=1.0):
    pings = [synth.Ping(start=start, nframes=nframes, f0=f0, snr_db=snr_db, text=text, phase=phase)]
    return synth.audio_stream(pings, nsamples, seed=seed), pings


def lines_key(lines, with_snr=True):
    """Comparable view of decode lines; f0 / pattern_idx are left out of the strict key because ties in
    (num_avg, nbadsync) are broken by an unstable sort in the reference (result_filter.cpp:61-69)."""
    out = []
    for l in lines:
        k = (l["block"], l["text"], l["num_avg"], l["nbadsync"])
        if with_snr:
            k = k + (l["snr"],)
        out.append(k)
    return sorted(out)


def candidate_set_agreement(pos_a, pos_b, depth):
    """Fraction of (frequency, pattern) groups whose 8 kept positions agree as SETS.

    Slot order inside a group depends on the order in which near-equal maxima evict each other
    (scan_kernel.cuh:276-353).  With circular frame sums the all-ones pattern (index 5) gives the same
    metric at p, p+864, ..., and 100100 (index 6) at p, p+2592, up to FP32 summation order, so those
    groups are compared modulo the symmetry."""
    a = np.asarray(pos_a).reshape(-1, depth, 8).astype(np.int64) % 5184
    b = np.asarray(pos_b).reshape(-1, depth, 8).astype(np.int64) % 5184
    good = total = 0
    for f in range(a.shape[0]):
        for p in range(depth):
            mod = 864 if p == 5 else (2592 if p == 6 else 5184)
            total += 1
            good += sorted((a[f, p] % mod).tolist()) == sorted((b[f, p] % mod).tolist())
    return good / total
