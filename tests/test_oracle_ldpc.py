"""CPU: the oracle's LDPC(128,90) / CRC-13 against the code's invariants and the independent numpy encoder."""
import numpy as np


def test_parity_matrix_invariants(orc, synth):
    H = orc.parity_matrix()
    assert H.shape == (38, 128)
    assert (H == synth.parity_matrix()).all()
    assert (H.sum(0) == 3).all()                      # every bit in exactly three checks
    rw = H.sum(1)
    assert sorted(np.where(rw == 11)[0].tolist()) == [2, 4, 5, 26]   # ldpc_context.cuh:160-163
    assert ((rw == 10) | (rw == 11)).all() and rw.sum() == 384
    # no two checks share more than one bit (no 4-cycles)
    ov = H.astype(int) @ H.astype(int).T
    np.fill_diagonal(ov, 0)
    assert ov.max() <= 1


def test_crc13_kat(orc, synth):
    b = np.random.default_rng(0).integers(0, 2, 77).astype(np.uint8)
    assert orc.crc13(b) == 0x1515 == synth.crc13(b)
    assert orc.crc13(np.zeros(77, np.uint8)) == 0


def test_encoder_matches_numpy_and_satisfies_H(orc, synth):
    rng = np.random.default_rng(7)
    H = orc.parity_matrix().astype(int)
    for _ in range(50):
        m = rng.integers(0, 2, 77).astype(np.uint8)
        cw = orc.ldpc_encode(m)
        assert (cw == synth.ldpc_encode(m)).all()
        assert ((H @ cw) % 2 == 0).all()
        assert (cw[:77] == m).all()


def test_decode_noiseless_and_noisy(orc, synth):
    llr, msgs = synth.cfg4_llrs(400, 7.0, seed=11)
    found, dec, iters, nhard = orc.ldpc_decode_batch(llr, nthreads=4)
    assert found.mean() > 0.99
    assert (dec[found] == msgs[found]).all()
    assert (iters[found] <= 9).all() and (nhard[found] < 18).all()
    # clean codewords decode at iteration 0 with no hard errors
    clean = (2.0 * np.stack([synth.ldpc_encode(m) for m in msgs[:20]]).astype(np.float32) - 1.0) * 5.0
    f, d, it, nh = orc.ldpc_decode_batch(clean)
    assert f.all() and (it == 0).all() and (nh == 0).all() and (d == msgs[:20]).all()
    # pure noise never decodes and reports 10 iterations
    noise = np.random.default_rng(3).standard_normal((200, 128)).astype(np.float32) * 3
    f, _, it, nh = orc.ldpc_decode_batch(noise)
    assert not f.any() and (it == 10).all()


def test_hard_error_gate(orc, synth):
    """success needs nhard < 18 (ldpc_kernel.cuh:205): flip 20 weak bits of a strong codeword."""
    m = np.random.default_rng(5).integers(0, 2, 77).astype(np.uint8)
    cw = synth.ldpc_encode(m)
    llr = (2.0 * cw.astype(np.float32) - 1.0) * 8.0
    idx = np.random.default_rng(6).choice(128, 20, replace=False)
    llr2 = llr.copy()
    llr2[idx] = -0.05 * np.sign(llr[idx])
    f, d, it, nh = orc.ldpc_decode(llr2)
    # BP repairs the 20 weak bits but the decode is refused because 20 >= 18 hard disagreements
    assert not f
    llr3 = llr.copy()
    llr3[idx[:10]] = -0.05 * np.sign(llr[idx[:10]])
    f, d, it, nh = orc.ldpc_decode(llr3)
    assert f and nh == 10 and (d == m).all()


def test_single_and_batch_agree(orc, synth):
    llr, _ = synth.cfg4_llrs(64, 3.0, seed=21)
    fb, db, ib, nb = orc.ldpc_decode_batch(llr)
    for i in range(64):
        f, d, it, nh = orc.ldpc_decode(llr[i])
        assert f == fb[i] and it == ib[i] and nh == nb[i]
        if f:
            assert (d == db[i]).all()


# ---- WSJT-X control flow: oracle semantics; the kernel is pinned to it in test_gpu_parity.py
def test_wsjtx_bp_semantics(orc, synth):
    llr, msgs = synth.cfg4_llrs(3000, 3.0, seed=150)
    f, m, it, nh = orc.ldpc_decode_batch(llr, 8)
    fw, mw, itw, nhw = orc.ldpc_decode_wsjtx_batch(llr, None, 8)
    # same arithmetic: whatever both decode is the same payload, and it is the transmitted one
    both = f & fw
    assert both.sum() > 2000 and (m[both] == mw[both]).all() and (mw[fw] == msgs[fw]).all()
    # a decode found at check k < 10 by the reference kernel is found at the same check here (identical iterations up to
    # there), unless the early stop fired first -- which needs iter >= 5
    early = both & (it < 5)
    assert early.sum() > 1000 and (it[early] == itw[early]).all() and (nh[early] == nhw[early]).all()
    # the 11th check exists: failures that ran to the end report 11, early stops report 5..10
    fail = ~fw
    assert set(np.unique(itw[fail])) <= set(range(5, 12)) and (itw[fail] == 11).any() and (itw[fail] < 11).any()
    # decodes at the 11th check (iters == 10) are out of the reference kernel's reach
    assert (itw[fw] <= 10).all() and (it[f] <= 9).all()


def test_wsjtx_bp_noise_stops_early(orc):
    noise = (np.random.default_rng(3).standard_normal((4000, 128)) * 5.5).astype(np.float32)
    fw, _, itw, nhw = orc.ldpc_decode_wsjtx_batch(noise, None, 8)
    assert fw.sum() == 0 and (nhw == -1).all()
    frac_early = (itw < 11).mean()
    assert 0.2 < frac_early < 0.6           # about a third of pure-noise codewords is abandoned before the 11th check


def test_wsjtx_bp_apmask_freezes_bits(orc, synth):
    llr, msgs = synth.cfg4_llrs(500, 2.0, seed=151)
    cw = np.stack([orc.ldpc_encode(m) for m in msgs])
    ap = np.zeros(128, np.uint8)
    ap[:40] = 1
    strong = llr.copy()
    strong[:, :40] = (2.0 * cw[:, :40] - 1.0) * 20.0        # the caller's a-priori knowledge goes into the LLRs
    f0, _, _, _ = orc.ldpc_decode_wsjtx_batch(llr, None, 8)
    f1, m1, _, _ = orc.ldpc_decode_wsjtx_batch(strong, ap, 8)
    assert f1.sum() > f0.sum() and (m1[f1] == msgs[f1]).all()
    wrong = llr.copy()
    wrong[:, :40] = -(2.0 * cw[:, :40] - 1.0) * 20.0         # frozen to the wrong values: nothing may decode correctly
    f2, m2, _, _ = orc.ldpc_decode_wsjtx_batch(wrong, ap, 8)
    assert not (f2 & (m2 == msgs).all(axis=1)).any()
