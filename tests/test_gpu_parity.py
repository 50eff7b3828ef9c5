"""GPU: the CUDA path, called through the C ABI, against the CPU oracle on the same seeded inputs.

Bars: payload bits / decode flags / decode lines exact; LLRs of matched
candidates max|d| / rms <= 1e-4; xb relative <= 1e-5.  The oracle uses glibc's libm where the
GPU uses libdevice (sincosf, tanhf, hypotf, atan2f differ in the last ulp), so a knife-edge candidate
can fall differently; such cases are counted and bounded, not hidden.
"""
import os
import subprocess

import numpy as np
import pytest

from conftest import candidate_set_agreement, lines_key, one_ping_audio

pytestmark = pytest.mark.gpu

LLR_TOL = 1e-4   # max |llr_gpu - llr_oracle| / rms(llr_oracle)
XB_TOL = 1e-5    # relative
# Bounds on the discrete differences between the GPU and the CPU oracle (glibc vs libdevice sincosf / hypotf / tanhf in
# the last ulp can move a near-tie).  Each test prints what it observed; the bounds are the observed values plus margin
# (profiles/r02_notes.md records the observations).
SET_AGREE_MIN = 0.98     # fraction of (frequency, pattern) groups with the same 8 positions
SAME_SLOT_MIN = 0.99     # fraction of asymmetric-pattern candidates in the same slot
NBAD_DIFFER_MAX = 1      # candidates whose nbadsync differs (a sync soft bit within rounding of zero)
FLAG_MISMATCH_MAX = 2    # LDPC verdicts / iteration counts / hard-error counts out of ~17,000 codewords


@pytest.fixture(scope="module")
def gpu(product):
    import torch
    assert torch.cuda.is_available(), "these tests need the B200"
    return product


def test_frontend_method2_matches_oracle(gpu, orc, synth):
    x, _ = synth.cfg1_demo(seconds=3.0, seed=2)
    with gpu.Decoder(compat_cache=True, width=20, depth=1) as d:
        blocks, powers = d.frontend(x)
    assert blocks.shape[0] == synth.num_blocks(len(x))
    for k in range(blocks.shape[0]):
        ref = orc.frontend_audio(x[k * 2592:k * 2592 + 5184], 2)
        err = np.max(np.abs(blocks[k] - ref)) / np.sqrt(np.mean(np.abs(ref) ** 2))
        assert err < 2e-6, (k, err)
        pref = orc.snr_powers(ref)
        assert np.allclose(powers[k], pref, rtol=2e-6)


def test_frontend_iq_matches_oracle(gpu, orc, synth):
    iq, _ = synth.cfg3_deep_iq(seconds=2.0, seed=5)
    with gpu.Decoder(compat_cache=True, read_mode=2, width=20, depth=1) as d:
        blocks, powers = d.frontend(iq)
    for k in range(blocks.shape[0]):
        ref = orc.frontend_iq(iq[2 * k * 2592:2 * (k * 2592 + 5184)])
        err = np.max(np.abs(blocks[k] - ref)) / np.sqrt(np.mean(np.abs(ref) ** 2))
        assert err < 2e-6, (k, err)


def test_frontend_fft_matches_oracle(gpu, orc, synth):
    x, _ = synth.cfg1_demo(seconds=2.0, seed=3)
    with gpu.Decoder(compat_cache=True, method=1, width=20, depth=1) as d:
        blocks, _ = d.frontend(x)
    for k in range(blocks.shape[0]):
        ref = orc.frontend_audio(x[k * 2592:k * 2592 + 5184], 1)
        err = np.max(np.abs(blocks[k] - ref)) / np.sqrt(np.mean(np.abs(ref) ** 2))
        assert err < 1e-5, (k, err)   # two different FP32 FFT factorizations


@pytest.mark.parametrize("depth,thr", [(3, 1), (8, 3)])
def test_scan_softbits_match_oracle(gpu, orc, synth, depth, thr):
    x, _ = one_ping_audio(synth, snr_db=4.0)
    blk = orc.frontend_audio(x[:5184])
    cfg = orc.config(width=60, step=2, depth=depth, thr=thr, scan_mode=1, nthreads=8)
    ref = orc.decode_block(blk, cfg)
    with gpu.Decoder(compat_cache=True, width=60, step=2, depth=depth, thr=thr) as d:
        cands, llr, llr_item = d.scan(blk[None, :])
    cands = cands[0]
    assert len(cands) == len(ref)
    agree = candidate_set_agreement(cands["pos"], ref["pos"], depth)
    same = cands["pos"] == ref["pos"]
    pat = ref["pattern_idx"]
    same_slot = same[(pat != 5) & (pat != 6)].mean()
    nbad_differ = int((cands["nbadsync"][same] != ref["nbadsync"][same]).sum())
    print(f"depth {depth}: candidate sets agree in {agree:.4f} of the (frequency, pattern) groups, same slot in {same_slot:.4f} of the "
          f"asymmetric-pattern candidates, nbadsync differs in {nbad_differ} of {int(same.sum())}")
    assert agree >= SET_AGREE_MIN and same_slot >= SAME_SLOT_MIN
    assert np.max(np.abs(cands["xb"][same] - ref["xb"][same]) / ref["xb"][same]) < XB_TOL
    # sync gate: identical where the position is identical (a sync soft bit within rounding of 0 could flip)
    assert nbad_differ <= NBAD_DIFFER_MAX
    # LLRs of gated candidates
    gated = np.where(same & (cands["llr_slot"] >= 0) & (ref["nbadsync"] <= thr))[0]
    assert len(gated) > 20
    row_of = {int(it): r for r, it in enumerate(llr_item)}
    worst = 0.0
    for i in gated:
        g = llr[row_of[int(i)]]
        r = ref["llr"][i]
        worst = max(worst, float(np.max(np.abs(g - r)) / np.sqrt(np.mean(r ** 2))))
    print(f"depth {depth}: worst |dLLR| / rms over {len(gated)} gated candidates {worst:.2e} (bound {LLR_TOL:.0e})")
    assert worst < LLR_TOL, worst
    # every candidate that passed the gate on the GPU has exactly one queue row
    assert sorted(row_of) == sorted(np.where(cands["llr_slot"] >= 0)[0].tolist())
    assert ((cands["llr_slot"] >= 0) == (cands["nbadsync"] <= thr)).all()


def test_ldpc_matches_oracle(gpu, orc, synth):
    llrs, msgs = [], []
    for ebn0 in (0.0, 2.0, 3.0, 4.0, 6.0):
        l, m = synth.cfg4_llrs(3000, ebn0, seed=146 + int(ebn0))
        llrs.append(l), msgs.append(m)
    llr = np.concatenate(llrs)
    msg = np.concatenate(msgs)
    # edge cases: zeros, exact zeros mixed in, huge magnitudes, pure noise
    rng = np.random.default_rng(9)
    extra = [np.zeros((4, 128), np.float32), (rng.standard_normal((500, 128)) * 2).astype(np.float32),
             (llr[:200] * 50).astype(np.float32)]
    z = llr[200:400].copy()
    z[:, ::7] = 0.0
    extra.append(z)
    llr_all = np.concatenate([llr] + extra)
    fo, mo, io, no = orc.ldpc_decode_batch(llr_all, nthreads=8)
    with gpu.Decoder(compat_cache=True, width=20, depth=1) as d:
        fg, mg, ig, ng = d.ldpc_decode(llr_all)
    flag_mismatch = int((fo != fg).sum())
    both = fo & fg
    it_mismatch, nh_mismatch = int((ig[both] != io[both]).sum()), int((ng[both] != no[both]).sum())
    print(f"GPU vs CPU oracle over {len(llr_all)} codewords: {flag_mismatch} decode flags, {it_mismatch} iteration counts, "
          f"{nh_mismatch} hard-error counts differ (glibc vs libdevice tanhf, last ulp)")
    assert flag_mismatch <= FLAG_MISMATCH_MAX, flag_mismatch
    assert (mo[both] == mg[both]).all()
    assert it_mismatch <= FLAG_MISMATCH_MAX and nh_mismatch <= FLAG_MISMATCH_MAX
    assert (ig[~fg] == 10).all() and (ng[~fg] == -1).all()
    # whatever decodes is a real codeword of the transmitted payload
    n = len(llr)
    assert (mg[:n][fg[:n]] == msg[fg[:n]]).all()
    assert fg[:n].mean() > 0.5


def test_stream_lines_match_oracle(gpu, orc, synth):
    x, pings = synth.cfg1_demo(seconds=20.0, seed=1)
    cfg = orc.config(width=100, step=2, depth=3, thr=1, nthreads=8)
    ref, nb = orc.decode_stream(x, cfg)
    with gpu.Decoder(compat_cache=True, width=100, step=2, depth=3, thr=1, max_batch_blocks=16) as d:
        lines, results = d.decode_host(x)
        st = d.stage_times()
    assert st["blocks"] == nb and st["kernel_launches"] >= 3 * ((nb + 15) // 16)
    assert lines_key(lines) == lines_key(ref)
    assert {l["text"] for l in lines} <= {p.text for p in pings}
    for r in results:
        assert gpu.decode_message(r["message"][:77]) in {p.text for p in pings}
    # batch size must not change anything (launch grouping is invisible)
    with gpu.Decoder(compat_cache=True, width=100, step=2, depth=3, thr=1, max_batch_blocks=5) as d:
        lines2, results2 = d.decode_host(x)
    assert lines_key(lines2) == lines_key(lines)
    assert np.array_equal(np.sort(results2, order=["block_index", "item_index"]),
                          np.sort(results, order=["block_index", "item_index"]))


def test_iq_deep_stream_matches_oracle(gpu, orc, synth):
    iq, pings = synth.cfg3_deep_iq(seconds=5.0, seed=145, npings=4)
    cfg = orc.config(center=0.0, width=500, step=1, depth=6, thr=3, read_mode=2, nthreads=8)
    ref, nb = orc.decode_stream(iq, cfg, max_blocks=6)
    with gpu.Decoder(compat_cache=True, center=0.0, width=500, step=1, depth=6, thr=3, read_mode=2) as d:
        lines, _ = d.decode_host(iq[: 2 * (5 * 2592 + 5184)])
    assert lines_key(lines) == lines_key(ref)


def test_push_poll_equals_batch(gpu, synth):
    x, _ = synth.cfg1_demo(seconds=10.0, seed=4)
    with gpu.Decoder(compat_cache=True, width=100, step=2, depth=3, thr=1) as d:
        batch, _ = d.decode_host(x)
        raw = x.tobytes()
        got = []
        rng = np.random.default_rng(0)
        i = 0
        while i < len(raw):
            n = int(rng.integers(1, 20000))       # ragged chunks, odd byte counts included
            d.push(raw[i:i + n])
            got += d.poll()
            i += n
    assert lines_key(got) == lines_key(batch)


def test_multichannel_equals_per_channel(gpu, synth):
    chans = [synth.cfg5_channel(c, seconds=4.0)[0] for c in range(3)]
    with gpu.Decoder(compat_cache=True, width=100, step=2, depth=3, thr=2) as d:
        per = []
        for c, x in enumerate(chans):
            l, _ = d.decode_host(x)
            per += [dict(li, channel=c) for li in l]
        multi, _ = d.decode_host(np.concatenate(chans), nchan=3)
    key = lambda ls: sorted((l["channel"],) + k for l, k in zip(ls, [lines_key([l])[0] for l in ls]))
    assert key(multi) == key(per)


def test_cli_matches_library(gpu, synth, tmp_path):
    x, _ = synth.cfg1_demo(seconds=8.0, seed=1)
    with gpu.Decoder(compat_cache=True, width=100, step=2, depth=3, thr=1) as d:
        lines, _ = d.decode_host(x)
    p = subprocess.run([gpu.CLI_PATH, "--search-width=100", "--scan-depth=3"], input=x.tobytes(), capture_output=True)
    assert p.returncode == 0
    out = p.stdout.decode().splitlines()
    assert out[-1] == "Done"
    assert b"Incomplete read error. rc=" in p.stderr and b"Actual parameters:" in p.stderr
    import re
    got = sorted(re.sub(r"date=\d{14}", "date=" + "x" * 14, l) for l in out[:-1])
    want = sorted(gpu.format_line(l, "x" * 14) for l in lines)
    assert got == want


def test_empty_and_short_inputs(gpu):
    with gpu.Decoder(compat_cache=True, width=20, depth=1) as d:
        assert d.decode_host(np.zeros(5183, np.int16))[0] == []
        lines, res = d.decode_host(np.zeros(5184, np.int16))     # all-zero block: rms 0 -> NaNs, no decode, no crash
        assert lines == [] and len(res) == 0


def test_cli_read_error_prints_epilogue_and_exit_code(gpu):
    """A failing read() ends the loop like the reference's short fread does (main.cu:274-278): the epilogue is printed
    and the decoder destroyed; the exit code reports the error.  stdin = a directory makes read() fail with EISDIR."""
    fd = os.open("/", os.O_RDONLY)
    try:
        r = subprocess.run([gpu.CLI_PATH, "--search-width=100", "--scan-depth=3"], stdin=fd, capture_output=True, timeout=120)
    finally:
        os.close(fd)
    assert r.returncode == 2
    assert b"read: Is a directory" in r.stderr and b"Incomplete read error. rc=0" in r.stderr
    assert r.stdout.decode().splitlines()[-1] == "Done"


def test_cli_wav_json_dedupe(gpu, synth):
    """Extensions either side of the path: a RIFF header is skipped with --wav (and decoded as 22 samples
    without it, as the reference does); --json carries the same fields; --dedupe prints one line per ping."""
    import json
    import re
    import struct
    x, pings = synth.cfg1_demo(seconds=12.0, seed=1)
    raw = x.tobytes()
    fmt = struct.pack("<HHIIHH", 1, 1, 12000, 24000, 2, 16)
    wav = b"RIFF" + struct.pack("<I", 36 + len(raw)) + b"WAVE" + b"fmt " + struct.pack("<I", 16) + fmt + b"data" + \
        struct.pack("<I", len(raw)) + raw
    base = [gpu.CLI_PATH, "--search-width=100", "--scan-depth=3"]
    run = lambda extra, data: subprocess.run(base + extra, input=data, capture_output=True).stdout.decode().splitlines()
    mask = lambda ls: [re.sub(r"date=\d{14}", "date=x", l) for l in ls]
    plain = run([], raw)
    assert plain[-1] == "Done" and len(plain) > 3
    assert mask(run(["--wav"], wav)) == mask(plain)          # header skipped: identical to the raw stream
    assert mask(run(["--wav"], raw)) == mask(plain)          # --wav on headerless input changes nothing
    js = [json.loads(l) for l in run(["--json"], raw)[:-1]]
    assert len(js) == len(plain) - 1
    for j, l in zip(js, plain):
        m = re.search(r"snr=\s*(-?\d+); f0=\s*([-\d.e+]+); num_avg=(\d+); nbadsync=(\d+); pattern_idx=(\d+);.*msg='(.*)';", l)
        assert (j["snr"], j["num_avg"], j["nbadsync"], j["pattern_idx"], j["msg"]) == \
            (int(m.group(1)), int(m.group(2 + 1)), int(m.group(4)), int(m.group(5)), m.group(6))
        assert abs(j["f0"] - float(m.group(2))) < 1e-3 and len(j["payload"]) == 77
    dd = run(["--dedupe", "--json"], raw)
    assert dd[-1] == "Done"
    ev = [json.loads(l) for l in dd[:-1]]
    # one event per ping (cfg1_demo: 8 pings of distinct start times; a text may be sent by more than one ping)
    assert len(ev) < len(js) and {e["msg"] for e in ev} == {j["msg"] for j in js}
    for e in ev:
        assert any(all(j[k] == e[k] for k in ("block", "msg", "num_avg", "nbadsync", "snr")) for j in js)
        near = [j for j in js if j["msg"] == e["msg"] and abs(j["block"] - e["block"]) <= 2]   # same event for sure
        assert (e["num_avg"], e["nbadsync"]) == min((j["num_avg"], j["nbadsync"]) for j in near)


def test_push_skips_wav_header_when_asked(gpu, synth):
    import struct
    x, _ = synth.cfg1_demo(seconds=6.0, seed=1)
    raw = x.tobytes()
    hdr = b"RIFF" + struct.pack("<I", 36 + len(raw)) + b"WAVE" + b"fmt " + struct.pack("<I", 16) + \
        struct.pack("<HHIIHH", 1, 1, 12000, 24000, 2, 16) + b"data" + struct.pack("<I", len(raw))
    with gpu.Decoder(compat_cache=True, width=100, step=2, depth=3, thr=1) as d:
        want, _ = d.decode_host(x)
    with gpu.Decoder(compat_cache=True, width=100, step=2, depth=3, thr=1, flags=gpu.FLAG_SKIP_WAV_HEADER) as d:
        got = []
        data = hdr + raw
        for i in range(0, 70, 7):               # the header arrives in fragments
            d.push(data[i:i + 7])
        d.push(data[70:])
        got += d.poll(4096)
    assert lines_key(got) == lines_key(want)


def test_twin_sharing_is_invisible(gpu, synth, monkeypatch):
    """A candidate at pos >= 5184 whose twin pos - 5184 sits in the same list has bit-identical soft bits; the pipeline
    decodes one of the two and reports both.  Every LDPC success record must be the same with the sharing off."""
    x, _ = synth.cfg2_deep_audio(seconds=12.0, seed=144, npings=24)
    kw = dict(center=1500.0, width=500.0, step=1.0, depth=6, thr=3)
    with gpu.Decoder(compat_cache=True, **kw) as d:
        lines, res = d.decode_host(x, result_cap=1 << 18)
        ncw = d.stage_times()["llr_codewords"]
    monkeypatch.setenv("MSK144_NO_TWINS", "1")
    with gpu.Decoder(compat_cache=True, **kw) as d:
        lines0, res0 = d.decode_host(x, result_cap=1 << 18)
        ncw0 = d.stage_times()["llr_codewords"]
    order = ["block_index", "item_index"]
    assert len(res) > 0 and np.array_equal(np.sort(res, order=order), np.sort(res0, order=order))
    assert lines_key(lines) == lines_key(lines0)
    assert 0.9 * ncw0 < ncw < ncw0                      # a few per cent fewer codewords reach the LDPC stage
    assert (res["pos"] >= 5184).any()                   # and twins are among the successes


@pytest.mark.parametrize("nchan,nsamp", [(1, 5184), (3, 7777), (8, 20000), (64, 12000), (130, 6402), (1024, 5184)])
def test_deinterleave_is_a_transpose(gpu, nchan, nsamp):
    """[t][c] -> [c][t] for vector-friendly and ragged shapes (both code paths of deinterleave_kernel)."""
    import torch
    rng = np.random.default_rng(nchan * 7 + nsamp)
    x = rng.integers(-32768, 32767, (nsamp, nchan), dtype=np.int16)
    xin = torch.from_numpy(x).cuda()
    out = torch.full((nchan, nsamp), 12345, dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    with gpu.Decoder(compat_cache=True, width=20, depth=1) as d:
        d.deinterleave_device(xin.data_ptr(), out.data_ptr(), nsamp, nchan)
        torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), x.T)


def test_interleaved_input_equals_channel_major(gpu, synth):
    chans = [synth.cfg5_channel(c, seconds=4.0)[0] for c in range(5)]
    cm = np.concatenate(chans)
    il = np.stack(chans, axis=1).copy()                    # [t][c]
    with gpu.Decoder(compat_cache=True, width=100, step=2, depth=3, thr=2) as d:
        a, ra = d.decode_host(cm, nchan=5)
        b, rb = d.decode_host(il, nchan=5, interleaved=True)
    assert len(a) > 0 and a == b
    assert np.array_equal(ra, rb)


def test_multichannel_streaming_equals_single_channel_streams(gpu, synth):
    """msk144_push_interleaved: N channels interleaved per time step, ragged pushes; every channel's lines (snr
    included: one recurrence per channel is carried) equal that channel pushed alone.  CLI --channels likewise."""
    nch = 4
    chans = [synth.cfg5_channel(c, seconds=6.0)[0] for c in range(nch)]
    il = np.stack(chans, axis=1).copy().tobytes()
    kw = dict(width=100, step=2, depth=3, thr=2)
    want = []
    for c in range(nch):
        with gpu.Decoder(compat_cache=True, **kw) as d:
            d.push(chans[c].tobytes())
            want += [dict(l, channel=c) for l in d.poll(4096)]
    with gpu.Decoder(compat_cache=True, **kw) as d:
        got, i, rng = [], 0, np.random.default_rng(1)
        while i < len(il):
            n = int(rng.integers(1, 60000))
            d.push(il[i:i + n], nchan=nch)
            got += d.poll(4096)
            i += n
        with pytest.raises(gpu.Msk144Error):
            d.push(il[:64], nchan=nch + 1)               # the channel count is fixed in mid-stream
    key = lambda ls: sorted((l["channel"],) + lines_key([l])[0] for l in ls)
    assert len(want) > 0 and key(got) == key(want)
    # time order (block, then channel), and the same sequence whatever the chunking: one push of everything
    order = [(l["block"], l["channel"]) for l in got]
    assert order == sorted(order)
    with gpu.Decoder(compat_cache=True, **kw) as d:
        d.push(il, nchan=nch)
        assert d.poll(4096) == got
    p = subprocess.run([gpu.CLI_PATH, "--search-width=100", "--scan-depth=3", "--nbadsync-threshold=2", f"--channels={nch}"],
                       input=il, capture_output=True)
    out = p.stdout.decode().splitlines()
    assert p.returncode == 0 and out[-1] == "Done"
    import re
    cli = sorted((int(re.search(r"channel=(\d+); $", l).group(1)), re.search(r"msg='(.*)';", l).group(1)) for l in out[:-1])
    assert cli == sorted((l["channel"], l["text"]) for l in want)


@pytest.mark.gpu
def test_wsjtx_bp_kernel_matches_oracle(gpu, orc, synth):
    """MSK144_FLAG_WSJTX_BP: ldpc_kernel<true> against the oracle's restatement of bpdecode128_90's control flow over the
    cfg4 sweep, pure noise, degenerate inputs, and with an a-priori mask.  (bpdecode128_90.f90 itself is not on this box.)"""
    llrs = [synth.cfg4_llrs(2500, float(e), seed=160 + e)[0] for e in range(8)]
    rng = np.random.default_rng(11)
    llrs += [(rng.standard_normal((6000, 128)) * 5.5).astype(np.float32), np.zeros((3, 128), np.float32)]
    z = llrs[3][:300].copy()
    z[:, ::5] = 0.0                       # exact zeros: nharderror must not count them
    llr = np.concatenate(llrs + [z])
    for ap in (None, (np.arange(128) < 28).astype(np.uint8)):
        fo, mo, io, no = orc.ldpc_decode_wsjtx_batch(llr, ap, 8)
        with gpu.Decoder(width=20, depth=1, wsjtx_bp=True) as d:
            d.set_ap_mask(ap)
            fg, mg, ig, ng = d.ldpc_decode(llr)
        mism = int((fo != fg).sum())
        both = fo & fg
        it_mism, nh_mism = int((ig != io).sum()), int((ng[both] != no[both]).sum())
        print(f"WSJT-X BP, apmask {'on' if ap is not None else 'off'}: {len(llr)} codewords, {int(fg.sum())} decoded, {mism} verdicts, "
              f"{it_mism} iteration counts (incl. early stops), {nh_mism} hard-error counts differ from the CPU oracle")
        assert mism <= FLAG_MISMATCH_MAX and it_mism <= 2 * FLAG_MISMATCH_MAX and nh_mism <= FLAG_MISMATCH_MAX
        assert (mo[both] == mg[both]).all()
        assert ((ig[~fg] >= 5) | (ng[~fg] >= 18)).all() and (ig <= 11).all()
    # the flag changes nothing for a decoder that does not set it, and a mask needs the flag
    with gpu.Decoder(width=20, depth=1) as d:
        with pytest.raises(gpu.Msk144Error):
            d.set_ap_mask(np.zeros(128, np.uint8))


@pytest.mark.gpu
def test_wsjtx_bp_stream_matches_oracle(gpu, orc, synth):
    """The whole pipeline with the flag: decode lines equal the oracle's with wsjtx_bp = 1."""
    x, pings = synth.cfg1_demo(seconds=6.0, seed=5)
    kw = dict(width=100, step=2, depth=3, thr=2)
    ref, nb = orc.decode_stream(x, orc.config(compat_cache=1, wsjtx_bp=1, nthreads=8, **kw))
    with gpu.Decoder(compat_cache=True, wsjtx_bp=True, **kw) as d:
        lines, _ = d.decode_host(x)
    assert len(ref) > 0 and lines_key(lines) == lines_key(ref)


def test_repeated_runs_are_bit_identical(gpu, synth):
    """Stand-in for racecheck (compute-sanitizer is closed on this pool): a data race in the shared-memory pipelines of
    scan_kernel / ldpc_kernel, or in the lock-free queue slots, would make outputs differ between runs of the same
    input.  Deep-scan settings (depth 8, step 1.5, thr 3), repeated in one decoder and across fresh decoders: every
    LDPC success, every decode line, every candidate record and every soft bit must be the same bits each time (queue
    slots are handed out by atomicAdd, so LLR rows are compared by the item they belong to)."""
    P = synth.Ping
    x = synth.audio_stream([P(start=5000, nframes=7, f0=1490.5, snr_db=6.0, text="CQ K1ABC FN42"),
                            P(start=30000, nframes=10, f0=1512.0, snr_db=1.0, text="K1ABC W9XYZ EN37"),
                            P(start=52000, nframes=4, f0=1478.0, snr_db=9.0, text="W9XYZ K1ABC R-03")], 6 * 12000, seed=77)
    kw = dict(compat_cache=True, width=60, step=1.5, depth=8, thr=3)

    def stage_view(d):
        blocks, powers = d.frontend(x)
        cands, llr, item = d.scan(blocks[:6])
        order = np.argsort(item, kind="stable")
        return (blocks.tobytes(), powers.tobytes(), cands["pos"].tobytes(), cands["xb"].tobytes(), cands["nbadsync"].tobytes(),
                item[order].tobytes(), llr[order].tobytes())

    runs, stages = [], []
    with gpu.Decoder(**kw) as d:
        for _ in range(4):
            lines, res = d.decode_host(x, result_cap=1 << 16)
            runs.append((lines, res.tobytes()))
        stages += [stage_view(d), stage_view(d)]
    for _ in range(2):
        with gpu.Decoder(**kw) as d:
            lines, res = d.decode_host(x, result_cap=1 << 16)
            runs.append((lines, res.tobytes()))
            stages.append(stage_view(d))
    assert len(runs[0][1]) > 0 and len(runs[0][0]) > 0
    assert all(r == runs[0] for r in runs[1:])
    assert all(s == stages[0] for s in stages[1:])
