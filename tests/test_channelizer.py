"""The channeliser in front of the decoder (msk144_channelizer_*): wideband I/Q -> channels of 12 ksps s8 I/Q.

CPU: the numpy oracle's two independent forms agree, the prototype filter meets its stated response, streaming history.
CPU: the Stockham stages of csrc/chz_fft.h against a direct DFT for every decimation that has a plan (host harness).
GPU: the CUDA kernels against the oracle (s8 output: equal except for roundings that sit on a .5 boundary within FP32
accuracy -- counted, printed and bounded), both kernels (direct: 1/2/4 channels per group; FFT across the branches with
every radix), all three input formats, both output layouts, unaligned input, streaming == one shot, and capture ->
channeliser -> decoder end to end."""
import numpy as np
import pytest

from oracle import channelizer as oc

LSB_DIFF_MAX = 2e-3   # fraction of output bytes allowed to differ by one LSB (observed: printed by the tests)


def random_capture(rng, nframes, M, fmt):
    n = 2 * nframes * M
    if fmt == 1:
        return rng.integers(0, 256, n, dtype=np.uint8)
    if fmt == 2:
        return rng.integers(-128, 128, n, dtype=np.int8)
    return rng.integers(-32768, 32768, n, dtype=np.int16)


def test_oracle_forms_agree_and_stream():
    rng = np.random.default_rng(1)
    for M, bins in ((2, [0, 1]), (20, [0, 3, -2, 19]), (33, [5, -16])):
        x = rng.standard_normal(M * 150) + 1j * rng.standard_normal(M * 150)
        a = oc.channelize_direct(x, M, bins)
        b = oc.channelize_polyphase(x, M, bins)
        assert np.abs(a - b).max() < 1e-12
        # streaming: second half with the last 7 frames of the first as history == the second half of the whole
        cut = 70 * M
        c = oc.channelize_polyphase(x[cut:], M, bins, history=x[cut - 7 * M:cut])
        assert np.abs(c - b[:, 70:]).max() < 1e-12


def test_prototype_response():
    M = 160
    h = oc.prototype(M)
    assert len(h) == 8 * M and abs(h.sum() - 1.0) < 1e-12 and np.allclose(h, h[::-1])
    H = np.abs(np.fft.fft(h, 1 << 18))
    f = np.abs(np.fft.fftfreq(1 << 18, 1.0 / (M * 12000)))
    db = lambda sel: 20 * np.log10(np.maximum(H[sel], 1e-12))
    assert db(f <= 2000).min() > -0.01            # MSK144 (+-1.3 kHz + search width) passes untouched
    assert db(f <= 3400).min() > -0.2
    assert abs(db((f > 5990) & (f < 6010)).mean() + 6.02) < 0.1
    assert db(f >= 10000).max() < -79.0           # everything that folds onto |f| < 2 kHz


def test_fft_stages_against_direct_dft(tmp_path):
    """csrc/chz_fft.h compiles as plain C++: every stage is walked thread index by thread index on the host."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
    exe = str(tmp_path / "chz_fft_host")
    subprocess.run(["g++", "-O2", "-std=c++17", f"-I{cuda_inc}", os.path.join(root, "tests", "chz_fft_host.cpp"), "-o", exe], check=True)
    r = subprocess.run([exe], capture_output=True, text=True)
    rows = [l.split() for l in r.stdout.splitlines()]
    assert r.returncode == 0, r.stdout[-400:]
    assert len(rows) >= 100 and {"160", "256", "200", "33", "91"} <= {row[0] for row in rows}
    assert max(float(row[2]) for row in rows) < 5e-7


def lsb_report(name, got, want):
    d = np.abs(got.astype(np.int32) - want.astype(np.int32))
    frac = float((d > 0).mean())
    print(f"{name}: {int((d > 0).sum())} of {d.size} bytes differ by one LSB ({frac:.2e}); max diff {int(d.max())}")
    assert d.max() <= 1 and frac <= LSB_DIFF_MAX
    return frac


@pytest.mark.gpu
@pytest.mark.parametrize("M,fmt,bins,gain", [
    (160, 1, [0, 5, -17, 40], 8.0),               # the README's rate; direct kernel, 4 channels
    (160, 1, [3], 8.0),                           # one channel (the README's chain)
    (160, 1, list(range(-80, 80)), 8.0),          # every channel of the capture: FFT across the branches, 8 x 4 x 5
    (2, 2, [0, 1], 1.0),                          # smallest decimation
    (20, 3, [0, -3, 7], 0.9),                     # 240 ksps s16
    (80, 2, [1, 2], 2.0),
    (100, 1, [0, 1, 2, 3, 4, 5, 6, 7, -1], 8.0),  # FFT 4 x 5 x 5
    (200, 2, [0, -99, 64, 13], 6.0),              # 2.4 Msps
    (256, 1, list(range(-8, 8)), 12.0),           # 3.072 Msps, FFT 8 x 8 x 4
    (256, 3, list(range(-128, 128)), 12.0),       # s16, every channel
    (33, 3, [0, 16, -16], 0.9),                   # a decimation that is not a multiple of anything
    (33, 2, [0, 16, -16, 1, 2, 3, -5], 0.9),      # FFT 3 x 11 (dense radix 11)
    (91, 2, list(range(10)), 2.0),                # FFT 7 x 13
    (34, 1, [0, 1, -1, 16, -16, 5], 4.0),         # 2 x 17: no FFT plan, two groups of the direct kernel
])
def test_gpu_channelizer_matches_oracle(product, M, fmt, bins, gain):
    rng = np.random.default_rng(M * 7 + fmt)
    nframes = 1237                               # not a multiple of 8: the tail takes the byte path
    cap = random_capture(rng, nframes, M, fmt)
    with product.Channelizer(M, bins, input_format=fmt, gain=gain) as z:
        taps = z.prototype()
        got = z.channelize_host(cap)
        assert z.launches() == 1
    assert np.abs(taps - oc.prototype(M)).max() < 1e-8
    want = oc.quantize(oc.channelize_polyphase(oc.to_complex(cap, fmt), M, bins), gain)
    assert got.shape == want.shape == (len(bins), nframes, 2)
    sat = float((np.abs(want) == 127).mean())
    assert sat < 0.2                              # saturation is exercised (M = 160: ~1 % of bytes), but rarely
    lsb_report(f"M={M} fmt={fmt} nchan={len(bins)}", got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("bins", [[0, 5, -17], [0, 5, -17, 40, 79, -80]])   # direct kernel / FFT kernel
def test_gpu_channelizer_streaming_and_layouts(product, bins):
    import torch
    M, fmt = 160, 1
    rng = np.random.default_rng(5)
    nframes = 4000
    cap = random_capture(rng, nframes, M, fmt)
    raw = cap.tobytes()
    with product.Channelizer(M, bins, input_format=fmt, gain=8.0) as z:
        whole = z.channelize_host(cap)                                   # [c][f][2]
        parts, off = [], 0
        for size in [1, 319, 320, 7 * 320 + 11, 100000, 3, 640000 - 77] + [len(raw)]:
            parts.append(z.push(raw[off:off + size]))
            off += size
            if off >= len(raw):
                break
        stream = np.concatenate(parts)                                   # [f][c][2]
        assert stream.shape == (nframes, len(bins), 2)
        assert np.array_equal(stream.transpose(1, 0, 2), whole)          # bit-identical: same arithmetic per frame
        z.reset()
        assert np.array_equal(z.push(raw[:320 * 50]).transpose(1, 0, 2), whole[:, :50])
        # device entry point, both layouts, with a history that the caller vouches for
        d_in = torch.from_numpy(cap).cuda()
        frame_bytes = 2 * M
        stride = (nframes + 7) // 8 * 8
        d_cm = torch.zeros((len(bins), stride, 2), dtype=torch.int8, device="cuda")
        d_tm = torch.zeros((nframes, len(bins), 2), dtype=torch.int8, device="cuda")
        s = torch.cuda.Stream()
        z.channelize_device(d_in.data_ptr(), nframes, d_cm.data_ptr(), stride, 1, 0, s.cuda_stream)
        z.channelize_device(d_in.data_ptr(), nframes, d_tm.data_ptr(), 1, len(bins), 0, s.cuda_stream)
        d_late = torch.zeros((len(bins), stride, 2), dtype=torch.int8, device="cuda")
        z.channelize_device(d_in.data_ptr() + 1000 * frame_bytes, nframes - 1000, d_late.data_ptr(), stride, 1, 7, s.cuda_stream)
        torch.cuda.synchronize()
        assert np.array_equal(d_cm.cpu().numpy()[:, :nframes], whole)
        assert np.array_equal(d_tm.cpu().numpy().transpose(1, 0, 2), whole)
        assert np.array_equal(d_late.cpu().numpy()[:, :nframes - 1000], whole[:, 1000:])
        # an input pointer that is not 16-byte aligned (the staging copies are): same bytes, same answer
        d_off = torch.zeros(cap.size + 64, dtype=torch.uint8, device="cuda")
        for shift in (2, 6, 14):
            d_off[shift:shift + cap.size] = d_in
            d_cm.zero_()
            z.channelize_device(d_off.data_ptr() + shift, nframes, d_cm.data_ptr(), stride, 1, 0, s.cuda_stream)
            torch.cuda.synchronize()
            assert np.array_equal(d_cm.cpu().numpy()[:, :nframes], whole), shift
        with pytest.raises(product.Msk144Error):
            z.channelize_device(d_in.data_ptr(), nframes, d_cm.data_ptr(), stride, 1, 8, s.cuda_stream)
    for bad in (dict(decimation=1, bins=[0]), dict(decimation=257, bins=[0]), dict(decimation=160, bins=[160]),
                dict(decimation=160, bins=[0], gain=0.0), dict(decimation=160, bins=[0], input_format=4)):
        with pytest.raises(product.Msk144Error):
            product.Channelizer(**bad)


@pytest.mark.gpu
def test_gpu_capture_to_text(product, synth):
    """rtl_sdr-style capture (u8, 1.92 Msps) with pings in three channels -> channeliser -> decoder (--read-mode=2)."""
    M, bins = 160, [0, 5, -17]
    cap, truth = synth.wideband_capture(bins, M, seconds=3.0, seed=7, snr_lo=5.0, snr_hi=10.0)
    gain = 72.0                                   # noise floor of a channel at ~16 LSB, like synth.iq_stream
    with product.Channelizer(M, bins, gain=gain) as z:
        chans = z.channelize_host(cap)
    want = oc.quantize(oc.channelize_polyphase(oc.to_complex(cap, 1), M, bins), gain)
    lsb_report("capture", chans, want)
    kw = dict(read_mode=2, center=0.0, width=100.0, step=2.0, depth=3, thr=2, compat_cache=True)
    with product.Decoder(**kw) as d:
        lines, _ = d.decode_host(chans, nchan=len(bins))
        lines_oracle_input, _ = d.decode_host(want, nchan=len(bins))
    for c, k in enumerate(bins):
        texts = {l["text"] for l in lines if l["channel"] == c}
        sent = {p.text for p in truth[k]}
        print(f"bin {k}: sent {sorted(sent)} decoded {sorted(texts)}")
        assert sent <= texts and texts <= set(synth.STANDARD_MESSAGES)
        # a message is heard in its own channel only
        others = {p.text for kk in bins if kk != k for p in truth[kk]}
        assert not (texts & (others - sent))
    key = lambda ls: sorted((l["channel"], l["block"], l["text"]) for l in ls)
    assert key(lines) == key(lines_oracle_input)
    # streaming: capture bytes -> Channelizer.push -> Decoder.push (time-major, one 2-byte sample per channel and step)
    with product.Channelizer(M, bins, gain=gain) as z, product.Decoder(**kw) as d:
        raw, got = cap.tobytes(), []
        for off in range(0, len(raw), 1 << 20):
            tm = z.push(raw[off:off + (1 << 20)])
            if len(tm):
                d.push(tm.tobytes(), nchan=len(bins))
                got += d.poll()
    assert key(got) == key(lines)
