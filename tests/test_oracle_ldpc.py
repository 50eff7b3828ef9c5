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
