"""CPU: unpack77 against 77-bit strings assembled BY HAND from the published WSJT-X field layout (the 77-bit message
protocol of FT8 / FT4 / MSK144: Franke, Somerville, Taylor, "The FT4 and FT8 Communication Protocols", QEX 2020, and
the `77-bit` section of the WSJT-X user guide).  WSJT-X's packjt77.f90 is not on this box and every other text check in
the tree goes through the builder's own `synth.pack77`, so these vectors are the one place where `unpack77`
(csrc/host/unpack77.cpp, the f_interop.cpp:61-71 seam) is held against field values worked out independently; the
arithmetic is in the comments.  A second assertion then holds `synth.pack77` to the same bits.

Layout used (most significant bit first):
  i3 = 1, 2 : n28a(28) ipa(1) n28b(28) ipb(1) ir(1) igrid4(15) i3(3)
  i3 = 4    : n12(12) n58(58) iflip(1) nrpt(2) icq(1) i3(3)
  i3 = 0    : 71 payload bits, n3(3), i3(3)    (n3 = 0 free text, n3 = 5 telemetry)
Standard call sign -> n28 = 2063592 (NTOKENS) + 4194304 (MAX22) + n, with the call right-aligned to 6 characters so that
its digit is the third one and n = ((((i1*36 + i2)*10 + i3)*27 + i4)*27 + i5)*27 + i6 over the alphabets
" 0-9A-Z" (37), "0-9A-Z" (36), "0-9" (10), " A-Z" (27) x 3.  CQ -> n28 = 2.
Grid "FN42" -> igrid4 = 1800*(F=5) + 100*(N=13) + 42; reports -> 32400 + 35 + report; RRR / RR73 / 73 -> 32402 / 3 / 4.
"""
import numpy as np
import pytest

NT = 2063592 + 4194304      # 6257896


def bits(*fields):
    out = []
    for value, width in fields:
        assert 0 <= value < (1 << width)
        out += [(value >> (width - 1 - k)) & 1 for k in range(width)]
    assert len(out) == 77
    return np.array(out, np.uint8)


# " K1ABC": (((( 0*36 + 20)*10 + 1)*27 + 1)*27 + 2)*27 + 3 = 200+1 = 201; *27+1 = 5428; *27+2 = 146558; *27+3 = 3957069
K1ABC = NT + 3957069
# " W9XYZ": W = 10 + 22 = 32: 32*10 + 9 = 329; *27 + 24 = 8907; *27 + 25 = 240514; *27 + 26 = 6493904
W9XYZ = NT + 6493904
# " G4ABC": G = 16: 16*10 + 4 = 164; *27 + 1 = 4429; *27 + 2 = 119585; *27 + 3 = 3228798
G4ABC = NT + 3228798

A38 = " 0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ/"
A42 = " 0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ+-./?"


def base(text, alphabet):
    v = 0
    for ch in text:
        v = v * len(alphabet) + alphabet.index(ch)
    return v


VECTORS = [
    # i3 = 1: CQ (n28 = 2), K1ABC, grid FN42 = 5*1800 + 13*100 + 42 = 10342
    ("CQ K1ABC FN42", bits((2, 28), (0, 1), (K1ABC, 28), (0, 1), (0, 1), (10342, 15), (1, 3))),
    # i3 = 1: R-03 -> ir = 1, igrid4 = 32400 + 35 - 3 = 32432
    ("K1ABC W9XYZ R-03", bits((K1ABC, 28), (0, 1), (W9XYZ, 28), (0, 1), (1, 1), (32432, 15), (1, 3))),
    # i3 = 1: RR73 -> 32403; 73 -> 32404
    ("W9XYZ K1ABC RR73", bits((W9XYZ, 28), (0, 1), (K1ABC, 28), (0, 1), (0, 1), (32403, 15), (1, 3))),
    ("K1ABC W9XYZ 73", bits((K1ABC, 28), (0, 1), (W9XYZ, 28), (0, 1), (0, 1), (32404, 15), (1, 3))),
    # i3 = 1 with /R on the second call (ipb = 1), grid EN37 = 4*1800 + 13*100 + 37 = 8537
    ("K1ABC W9XYZ/R EN37", bits((K1ABC, 28), (0, 1), (W9XYZ, 28), (1, 1), (0, 1), (8537, 15), (1, 3))),
    # i3 = 2 (EU VHF form of type 1): /P flag, grid IO91 = 8*1800 + 14*100 + 91 = 15891
    ("CQ G4ABC/P IO91", bits((2, 28), (0, 1), (G4ABC, 28), (1, 1), (0, 1), (15891, 15), (2, 3))),
    # i3 = 4: a non-standard call in full, 11 characters right-aligned, base 38; icq = 1 prints "CQ <call>"
    ("CQ PJ4/K1ABC", bits((0, 12), (base("  PJ4/K1ABC", A38), 58), (0, 1), (0, 2), (1, 1), (4, 3))),
    # i3 = 0, n3 = 0: free text, 13 characters, base 42, in 71 bits
    ("TNX BOB 73 GL", bits((base("TNX BOB 73 GL", A42), 71), (0, 3), (0, 3))),
    # i3 = 0, n3 = 5: telemetry, 71 bits printed as 18 hex digits
    ("123456789ABCDEF012", bits((0x123456789ABCDEF012, 71), (5, 3), (0, 3))),
]


@pytest.mark.parametrize("text,b", VECTORS, ids=[v[0] for v in VECTORS])
def test_unpack77_on_hand_assembled_bits(product, text, b):
    product.reset_hash_tables()
    assert product.decode_message(b) == text


@pytest.mark.parametrize("text,b", VECTORS, ids=[v[0] for v in VECTORS])
def test_builders_pack77_agrees_with_the_hand_assembled_bits(synth, text, b):
    if text == "123456789ABCDEF012":
        pytest.skip("synth.pack77 has no telemetry packer; the unpack side is covered above")
    got = np.asarray(synth.pack77(text), np.uint8)[:77]
    if text == "CQ PJ4/K1ABC":
        got, b = got[12:], b[12:]    # n12 (the sender's own hash) is not printed when icq = 1; the packer fills it in
    assert got.tolist() == b.tolist()


def test_hash_learnt_from_a_hand_assembled_type4_message(product):
    """i3 = 4 with icq = 0: n12 is the 12-bit hash of the OTHER call; ihashcall = (47055833459 * n) >> (64 - 12) over
    the call left-aligned in 11 characters, base 38 (published hash of the 77-bit protocol)."""
    product.reset_hash_tables()
    assert product.decode_message(VECTORS[6][1]) == "CQ PJ4/K1ABC"       # teaches PJ4/K1ABC
    n12 = ((47055833459 * base("PJ4/K1ABC  ", A38)) & (2 ** 64 - 1)) >> 52
    b = bits((n12, 12), (base("      W9XYZ", A38), 58), (0, 1), (2, 2), (0, 1), (4, 3))      # nrpt = 2 -> RR73
    assert product.decode_message(b) == "<PJ4/K1ABC> W9XYZ RR73"
    product.reset_hash_tables()
    assert product.decode_message(b) == "<...> W9XYZ RR73"
