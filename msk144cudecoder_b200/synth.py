"""Synthetic MSK144 stream generator (numpy only).

The reference ships a decoder only -- no encoder, modulator or test vector --
so every input used by tests/ and bench.py is made here, seeded and reproducible:

    text -> pack77 (standard i3=1 messages, free text, telemetry)
         -> CRC-13 (poly 0x15D7, ldpc_kernel.cuh:32-63)
         -> systematic LDPC(128,90) encode (generator derived from the parity-check structure of
            ldpc_context.cuh:10-139)
         -> 144-bit frame  s8 | cw[0:48] | s8 | cw[48:128]  (softbits_kernel.cuh:204-211)
         -> OQPSK / half-sine waveform, 864 samples at 12 kHz (msk_context.cuh:137-197)
         -> real s16 audio at an audio offset, or complex s8 IQ, + AWGN at a stated SNR
            (signal power / noise power in 2500 Hz).

This is an independent statement of the protocol: tests cross-check it against the oracle's encoder
and the oracle's / product's decoders.
"""
from __future__ import annotations

import dataclasses
from typing import List, Optional, Sequence, Tuple

import numpy as np

FS = 12000
NSPF = 864          # samples per 72 ms frame
NBLOCK = 5184       # one processing block = 6 frames (common.h:15)
HOP = NBLOCK // 2   # main.cu:284-288
S8 = np.array([0, 1, 1, 1, 0, 0, 1, 0], np.uint8)
NTOKENS = 2063592
MAX22 = 4194304
MAXGRID4 = 32400

# check -> member bits of the (128,90) code (same structure as oracle/ldpc_128_90_nm.h)
NM = [
    [1, 14, 26, 39, 52, 64, 76, 90, 93, 114], [2, 5, 27, 40, 53, 65, 77, 91, 93, 119],
    [3, 15, 28, 41, 54, 66, 76, 89, 92, 105, 117], [4, 16, 29, 42, 51, 63, 78, 91, 101, 118],
    [5, 17, 30, 43, 55, 67, 79, 88, 94, 107, 124], [6, 13, 31, 44, 56, 67, 78, 89, 95, 115, 120],
    [3, 18, 32, 42, 57, 68, 80, 96, 106, 120], [1, 19, 29, 38, 53, 69, 79, 97, 106, 127],
    [2, 20, 33, 45, 58, 66, 78, 98, 106, 122], [7, 14, 28, 46, 55, 70, 81, 99, 110, 127],
    [8, 21, 33, 43, 51, 71, 82, 100, 102, 125], [9, 16, 25, 47, 59, 72, 83, 90, 109, 120],
    [6, 22, 34, 37, 54, 72, 81, 100, 108, 119], [10, 18, 35, 48, 52, 69, 84, 101, 103, 125],
    [9, 19, 36, 45, 57, 70, 84, 104, 108, 121], [4, 22, 36, 46, 56, 73, 85, 92, 109, 124],
    [11, 12, 26, 40, 60, 67, 86, 96, 108, 112], [10, 15, 37, 44, 57, 71, 77, 98, 107, 114],
    [3, 21, 30, 40, 59, 73, 81, 104, 111, 114], [11, 23, 31, 38, 61, 62, 87, 98, 101, 117],
    [0, 18, 24, 44, 61, 74, 76, 99, 111, 118], [5, 24, 32, 49, 58, 70, 82, 97, 116, 117],
    [9, 15, 38, 50, 52, 65, 82, 94, 110, 123], [8, 23, 34, 41, 58, 75, 88, 93, 113, 121],
    [6, 16, 36, 49, 53, 74, 87, 110, 113, 122], [10, 24, 34, 47, 60, 64, 87, 104, 115, 124],
    [8, 20, 31, 48, 54, 68, 83, 85, 94, 112, 118], [1, 17, 27, 46, 62, 72, 86, 95, 105, 125],
    [11, 25, 30, 48, 63, 71, 80, 99, 105, 119], [12, 14, 27, 47, 50, 75, 84, 92, 102, 122],
    [7, 19, 35, 43, 56, 74, 77, 90, 112, 116], [4, 20, 28, 39, 50, 69, 80, 95, 116, 121],
    [7, 25, 33, 39, 61, 73, 79, 91, 115, 126], [0, 12, 13, 35, 41, 63, 65, 83, 100, 107],
    [13, 21, 26, 49, 62, 68, 88, 103, 109, 127], [0, 17, 29, 45, 59, 64, 89, 96, 113, 126],
    [2, 22, 32, 51, 55, 75, 86, 103, 111, 123], [23, 37, 42, 60, 66, 85, 97, 102, 123, 126],
]


def parity_matrix() -> np.ndarray:
    h = np.zeros((38, 128), np.uint8)
    for c, row in enumerate(NM):
        h[c, row] = 1
    return h


def _gf2_generator() -> np.ndarray:
    """parity(38) = G(38x90) @ msg90 over GF(2), from H = [A | B]: G = B^-1 A."""
    m = parity_matrix().copy()
    for col in range(38):
        piv = next(r for r in range(col, 38) if m[r, 90 + col])
        if piv != col:
            m[[piv, col]] = m[[col, piv]]
        for r in range(38):
            if r != col and m[r, 90 + col]:
                m[r] ^= m[col]
    return m[:, :90].copy()


_GEN = _gf2_generator()


def crc13(bits77: Sequence[int]) -> int:
    """Remainder of (77 message bits followed by 19 zero bits) modulo x^13 + 0x15D7."""
    rem = 0
    for b in list(bits77) + [0] * 19:
        rem = (rem << 1) | int(b)
        if rem & 0x2000:
            rem ^= 0x2000 | 0x15D7
    return rem & 0x1FFF


def ldpc_encode(bits77: Sequence[int]) -> np.ndarray:
    cw = np.zeros(128, np.uint8)
    cw[:77] = np.asarray(bits77, np.uint8)
    c = crc13(cw[:77])
    cw[77:90] = [(c >> (12 - i)) & 1 for i in range(13)]
    cw[90:] = (_GEN.astype(np.int64) @ cw[:90].astype(np.int64)) & 1
    return cw


# ------------------------------------------------------------------------------------------ pack77
_A1 = " 0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ"
_A2 = "0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ"
_A3 = "0123456789"
_A4 = " ABCDEFGHIJKLMNOPQRSTUVWXYZ"
_FREE = " 0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ+-./?"


_A38 = " 0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ/"


def ihashcall(call: str, m: int) -> int:
    """m-bit hash of a call sign (m = 10, 12, 22): base-38 value of the 11-character left-justified call times a
    fixed multiplier, top m bits of the 64-bit product."""
    c = call.strip("<>").ljust(11)[:11]
    n8 = 0
    for ch in c:
        n8 = 38 * n8 + _A38.index(ch)
    return ((47055833459 * n8) & 0xFFFFFFFFFFFFFFFF) >> (64 - m)


def is_standard_call(c: str) -> bool:
    """One or two leading characters (at most one of them a digit in front position), a digit, up to three letters."""
    if len(c) >= 3 and c[2].isdigit():
        c6 = c
    elif len(c) >= 2 and c[1].isdigit():
        c6 = " " + c
    else:
        return False
    if len(c6) > 6:
        return False
    c6 = c6.ljust(6)
    if not (c6[0] in _A1 and c6[1] in _A2 and c6[2] in _A3 and all(ch in _A4 for ch in c6[3:])):
        return False
    return " " not in c6.strip() and c6[1] != " "


def pack28(call: str) -> int:
    """28-bit field: tokens DE / QRZ / CQ / `CQ nnn` / `CQ xxxx` (written with an underscore or a blank), a hashed call
    `<CALL>` (22-bit hash), or a standard call (with the 3DA0 / 3X work-arounds of the protocol)."""
    if call == "DE":
        return 0
    if call == "QRZ":
        return 1
    if call == "CQ":
        return 2
    c = call.upper()
    if c.startswith("CQ_") or c.startswith("CQ "):
        t = c[3:]
        if t.isdigit() and len(t) == 3:
            return 3 + int(t)
        if t.isalpha() and 1 <= len(t) <= 4:
            v = 0
            for ch in t.rjust(4):
                v = v * 27 + _A4.index(ch)
            return 1003 + v
        raise ValueError(f"bad CQ token: {call}")
    if c.startswith("<") and c.endswith(">"):
        return NTOKENS + ihashcall(c, 22)
    if c.startswith("3DA0"):
        c = "3D0" + c[4:]
    elif c.startswith("3X") and len(c) > 2 and c[2].isalpha():
        c = "Q" + c[2:]
    if not is_standard_call(c):
        raise ValueError(f"not a standard call: {call}")
    c6 = (c if c[2].isdigit() and len(c) >= 3 else " " + c).ljust(6)
    i = [_A1.index(c6[0]), _A2.index(c6[1]), _A3.index(c6[2]), _A4.index(c6[3]), _A4.index(c6[4]),
         _A4.index(c6[5])]
    n = ((((i[0] * 36 + i[1]) * 10 + i[2]) * 27 + i[3]) * 27 + i[4]) * 27 + i[5]
    return NTOKENS + MAX22 + n


def _to_bits(v: int, n: int) -> List[int]:
    return [(v >> (n - 1 - k)) & 1 for k in range(n)]


def _grid4(t: str) -> Optional[int]:
    if len(t) == 4 and t[:2].isalpha() and t[2:].isdigit() and t[0] <= "R" and t[1] <= "R":
        return ((ord(t[0]) - 65) * 18 + (ord(t[1]) - 65)) * 100 + int(t[2:])
    return None


def _pack_type4(w: List[str]) -> Optional[List[int]]:
    """`<HASHED> NONSTD [RRR|RR73|73]`, `NONSTD <HASHED> [...]`, `CQ NONSTD`: one call in full (up to 11 characters of
    [0-9A-Z/]), the other as a 12-bit hash."""
    rpt = {"": 0, "RRR": 1, "RR73": 2, "73": 3}
    if len(w) == 2 and w[0] == "CQ" and 3 <= len(w[1]) <= 11 and all(ch in _A38[1:] for ch in w[1]):
        full, n12, iflip, nrpt, icq = w[1], 0, 0, 0, 1
    elif len(w) in (2, 3) and (len(w) == 2 or w[2] in rpt):
        a, b = w[0], w[1]
        if a.startswith("<") and not b.startswith("<"):
            full, n12, iflip = b, ihashcall(a, 12), 0
        elif b.startswith("<") and not a.startswith("<"):
            full, n12, iflip = a, ihashcall(b, 12), 1
        else:
            return None
        if not (3 <= len(full) <= 11 and all(ch in _A38[1:] for ch in full)):
            return None
        nrpt, icq = rpt[w[2] if len(w) == 3 else ""], 0
    else:
        return None
    n58 = 0
    for ch in full.rjust(11):
        n58 = n58 * 38 + _A38.index(ch)
    return _to_bits(n12, 12) + _to_bits(n58, 58) + [iflip] + _to_bits(nrpt, 2) + [icq] + _to_bits(4, 3)


def _pack_type5(w: List[str]) -> Optional[List[int]]:
    """EU VHF contest: `<CALL1> <CALL2> [R] RRSSSS GRID6` (report 52..59, serial 0..2047)."""
    if len(w) not in (4, 5) or not (w[0].startswith("<") and w[1].startswith("<")):
        return None
    ir = 1 if w[2] == "R" else 0
    if len(w) != 4 + ir:
        return None
    ex, g = w[2 + ir], w[3 + ir]
    if not (len(ex) == 6 and ex.isdigit() and len(g) == 6 and g[:2].isalpha() and g[2:4].isdigit() and g[4:].isalpha()):
        return None
    irpt, iserial = int(ex[:2]) - 52, int(ex[2:])
    if not (0 <= irpt <= 7 and iserial <= 2047 and g[0] <= "R" and g[1] <= "R" and g[4] <= "X" and g[5] <= "X"):
        return None
    ig = ((((ord(g[0]) - 65) * 18 + (ord(g[1]) - 65)) * 10 + int(g[2])) * 10 + int(g[3])) * 24 * 24 + (ord(g[4]) - 65) * 24 + (ord(g[5]) - 65)
    return (_to_bits(ihashcall(w[0], 12), 12) + _to_bits(ihashcall(w[1], 22), 22) + [ir] + _to_bits(irpt, 3)
            + _to_bits(iserial, 11) + _to_bits(ig, 25) + _to_bits(5, 3))


def pack77(msg: str) -> np.ndarray:
    """Pack a message into 77 bits: every type the reference's gate admits (decode_softbits.cpp:30).
      i3 = 1 / 2  `C1 C2 [R] GRID4 | [R]+-NN | RRR | RR73 | 73 | (nothing)`; C = DE, QRZ, CQ, `CQ nnn`, `CQ XXXX`,
                  a standard call with optional /R (i3 = 1) or /P (i3 = 2), or a hashed call `<CALL>`
      i3 = 4      one non-standard call in full + the other as `<CALL>` (12-bit hash), `CQ NONSTD`
      i3 = 5      EU VHF contest, both calls hashed
      0.5         telemetry, `TLM:<hex>` (71 bits);   0.0  free text of up to 13 characters (anything else)."""
    msg = msg.strip().upper()
    if msg.startswith("TLM:"):
        v = int(msg[4:], 16)
        if v >= 1 << 71:
            raise ValueError("telemetry exceeds 71 bits")
        return np.array(_to_bits(v, 71) + _to_bits(5, 3) + _to_bits(0, 3), np.uint8)
    w = msg.split()
    if len(w) >= 2 and w[0] == "CQ" and (w[1].isdigit() and len(w[1]) == 3 or w[1].isalpha() and len(w[1]) <= 4 and len(w) >= 3
                                         and not is_standard_call(w[1])):
        w = ["CQ_" + w[1]] + w[2:]          # directed CQ: the modifier belongs to the first field
    b5 = _pack_type5(w)
    if b5 is not None:
        return np.array(b5, np.uint8)
    try:
        if len(w) in (2, 3, 4):
            i3, ip = 1, [0, 0]
            calls = []
            for k in (0, 1):
                c = w[k]
                if c.endswith("/R") or c.endswith("/P"):
                    if is_standard_call(c[:-2]):
                        if c.endswith("/P"):
                            i3 = 2
                        ip[k], c = 1, c[:-2]
                calls.append(c)
            if "/P" in msg and "/R" in msg:
                raise ValueError
            n28a, n28b = pack28(calls[0]), pack28(calls[1])
            ir, rest = 0, w[2:]
            if len(rest) == 2:
                if rest[0] != "R" or _grid4(rest[1]) is None:
                    raise ValueError
                ir, ig = 1, _grid4(rest[1])
            elif len(rest) == 0:
                ig = MAXGRID4 + 1
            else:
                t = rest[0]
                if t == "RRR":
                    ig = MAXGRID4 + 2
                elif t == "RR73":
                    ig = MAXGRID4 + 3
                elif t == "73":
                    ig = MAXGRID4 + 4
                elif _grid4(t) is not None:
                    ig = _grid4(t)
                else:
                    if t[0] == "R":
                        ir, t = 1, t[1:]
                    snr = int(t)
                    if not (t[0] in "+-" and len(t) == 3 and -30 <= snr <= 49):
                        raise ValueError
                    ig = MAXGRID4 + 35 + snr
            bits = (_to_bits(n28a, 28) + [ip[0]] + _to_bits(n28b, 28) + [ip[1]] + [ir] + _to_bits(ig, 15)
                    + _to_bits(i3, 3))
            return np.array(bits, np.uint8)
    except (ValueError, IndexError):
        pass
    b4 = _pack_type4(w)
    if b4 is not None:
        return np.array(b4, np.uint8)
    if len(msg) > 13 or any(ch not in _FREE for ch in msg):
        raise ValueError(f"cannot pack: {msg!r}")
    v = 0
    for ch in msg.rjust(13):
        v = v * 42 + _FREE.index(ch)
    return np.array(_to_bits(v, 71) + _to_bits(0, 3) + _to_bits(0, 3), np.uint8)


# ------------------------------------------------------------------------------------------ waveform
def frame_bits(cw128: np.ndarray) -> np.ndarray:
    return np.concatenate([S8, cw128[:48], S8, cw128[48:]]).astype(np.uint8)


def frame_waveform(bits144: np.ndarray) -> np.ndarray:
    """864 complex samples, unit envelope, circular (the frame repeats seamlessly)."""
    pp = np.sin(np.arange(12) * np.pi / 12.0)
    b = 2.0 * bits144.astype(np.float64) - 1.0
    re = np.zeros(NSPF)
    im = np.zeros(NSPF)
    for i in range(72):
        re[12 * i:12 * i + 12] += b[2 * i + 1] * pp
        idx = (12 * i - 6 + np.arange(12)) % NSPF
        im[idx] += b[2 * i] * pp
    return re + 1j * im


def message_waveform(text: str) -> Tuple[np.ndarray, np.ndarray]:
    bits77 = pack77(text)
    return frame_waveform(frame_bits(ldpc_encode(bits77))), bits77


# ------------------------------------------------------------------------------------------ streams
@dataclasses.dataclass
class Ping:
    start: int          # first sample of the ping
    nframes: int        # consecutive repeats of the 72 ms frame
    f0: float           # Hz: audio frequency (audio mode) or baseband offset (IQ mode)
    snr_db: float       # signal power / noise power in 2500 Hz
    text: str
    phase: float = 0.0
    bits77: Optional[np.ndarray] = None


STANDARD_MESSAGES = [
    "CQ K1ABC FN42", "K1ABC W9XYZ EN37", "W9XYZ K1ABC -11", "K1ABC W9XYZ R-09", "W9XYZ K1ABC RRR",
    "K1ABC W9XYZ 73", "CQ RA9YER MO26", "RA9YER UA3ABC KO85", "UA3ABC RA9YER +05", "RA9YER UA3ABC R+02",
    "UA3ABC RA9YER RR73", "CQ G4ABC IO91", "G4ABC PA3XYZ JO22", "PA3XYZ G4ABC -03", "G4ABC PA3XYZ R-07",
    "PA3XYZ G4ABC 73", "CQ JA1XYZ PM95", "JA1XYZ VK2ABC QF56", "VK2ABC JA1XYZ +10", "JA1XYZ VK2ABC R+08",
    "VK2ABC JA1XYZ RRR", "CQ DL1ABC JO62", "DL1ABC OH2XYZ KP20", "OH2XYZ DL1ABC -15", "DL1ABC OH2XYZ R-12",
    "OH2XYZ DL1ABC RR73", "CQ W1AW FN31", "W1AW N0ABC DM79", "N0ABC W1AW +00", "W1AW N0ABC R-01",
    "N0ABC W1AW 73", "QRZ K1JT FN20",
]


def random_pings(rng: np.random.Generator, nsamples: int, npings: int, f_lo: float, f_hi: float,
                 snr_lo: float, snr_hi: float, frames_lo: int = 2, frames_hi: int = 12,
                 messages: Sequence[str] = STANDARD_MESSAGES) -> List[Ping]:
    pings = []
    for _ in range(npings):
        nfr = int(rng.integers(frames_lo, frames_hi + 1))
        start = int(rng.integers(0, max(1, nsamples - nfr * NSPF)))
        pings.append(Ping(start=start, nframes=nfr, f0=float(rng.uniform(f_lo, f_hi)),
                          snr_db=float(rng.uniform(snr_lo, snr_hi)),
                          text=str(messages[int(rng.integers(0, len(messages)))]),
                          phase=float(rng.uniform(0, 2 * np.pi))))
    pings.sort(key=lambda p: p.start)
    return pings


def _render(pings: Sequence[Ping], nsamples: int, amp_of) -> np.ndarray:
    sig = np.zeros(nsamples, np.complex128)
    cache = {}
    for p in pings:
        if p.text not in cache:
            cache[p.text] = message_waveform(p.text)
        wf, bits = cache[p.text]
        p.bits77 = bits
        n = p.nframes * NSPF
        end = min(nsamples, p.start + n)
        if end <= p.start:
            continue
        t = np.arange(p.start, end)
        base = np.tile(wf, p.nframes)[: end - p.start]
        sig[p.start:end] += amp_of(p) * base * np.exp(1j * (2 * np.pi * p.f0 * t / FS + p.phase))
    return sig


def audio_stream(pings: Sequence[Ping], nsamples: int, noise_rms: float = 1000.0,
                 seed: int = 0) -> np.ndarray:
    """Real s16 mono audio at 12 kHz (--read-mode=1).  SNR reference bandwidth 2500 Hz."""
    rng = np.random.default_rng(seed)
    amp = lambda p: noise_rms * np.sqrt(2.0 * 10 ** (p.snr_db / 10) * 2500.0 / (FS / 2))
    x = _render(pings, nsamples, amp).real + noise_rms * rng.standard_normal(nsamples)
    return np.clip(np.rint(x), -32768, 32767).astype(np.int16)


def iq_stream(pings: Sequence[Ping], nsamples: int, noise_rms: float = 16.0, seed: int = 0) -> np.ndarray:
    """Complex baseband, interleaved s8 I,Q at 12 ksps (--read-mode=2).  noise_rms is per component."""
    rng = np.random.default_rng(seed)
    amp = lambda p: noise_rms * np.sqrt(2.0 * 10 ** (p.snr_db / 10) * 2500.0 / FS)
    x = _render(pings, nsamples, amp)
    x = x + noise_rms * (rng.standard_normal(nsamples) + 1j * rng.standard_normal(nsamples))
    out = np.empty(2 * nsamples, np.int8)
    out[0::2] = np.clip(np.rint(x.real), -128, 127)
    out[1::2] = np.clip(np.rint(x.imag), -128, 127)
    return out


def num_blocks(nsamples: int) -> int:
    return (nsamples - NBLOCK) // HOP + 1 if nsamples >= NBLOCK else 0


# BASELINE.md section 3 configurations -------------------------------------------------------------
def cfg1_demo(seconds: float = 60.0, seed: int = 1):
    """Stand-in for demo/0001.wav (blob stripped from the reference tree)."""
    rng = np.random.default_rng(seed)
    n = int(seconds * FS)
    pings = random_pings(rng, n, 8, 1460.0, 1540.0, 3.0, 10.0, 5, 5)
    return audio_stream(pings, n, 1000.0, seed), pings


def cfg2_deep_audio(seconds: float = 900.0, seed: int = 144, npings: Optional[int] = None):
    rng = np.random.default_rng(seed)
    n = int(seconds * FS)
    if npings is None:
        npings = max(1, int(round(300 * seconds / 900.0)))
    pings = random_pings(rng, n, npings, 1260.0, 1740.0, -4.0, 10.0)
    return audio_stream(pings, n, 1000.0, seed), pings


def cfg3_deep_iq(seconds: float = 900.0, seed: int = 145, npings: Optional[int] = None):
    rng = np.random.default_rng(seed)
    n = int(seconds * FS)
    if npings is None:
        npings = max(1, int(round(300 * seconds / 900.0)))
    pings = random_pings(rng, n, npings, -249.0, 249.0, -4.0, 10.0)
    return iq_stream(pings, n, 16.0, seed), pings


def cfg4_llrs(ncw: int, ebn0_db: float, seed: int = 146):
    """Random payloads -> codewords -> BPSK + AWGN -> llr = 2y/sigma^2 (rate 90/128)."""
    rng = np.random.default_rng(seed)
    msgs = rng.integers(0, 2, (ncw, 77), dtype=np.uint8)
    cws = np.stack([ldpc_encode(m) for m in msgs]) if ncw <= 4096 else _encode_many(msgs)
    sigma = np.sqrt(1.0 / (2.0 * (90.0 / 128.0) * 10 ** (ebn0_db / 10)))
    y = (2.0 * cws.astype(np.float32) - 1.0) + sigma * rng.standard_normal(cws.shape).astype(np.float32)
    return (2.0 * y / sigma ** 2).astype(np.float32), msgs


def _encode_many(msgs: np.ndarray) -> np.ndarray:
    n = msgs.shape[0]
    cw = np.zeros((n, 128), np.uint8)
    cw[:, :77] = msgs
    # CRC of many messages at once: bitwise long division, vectorised over the batch
    rem = np.zeros(n, np.int64)
    ext = np.concatenate([msgs, np.zeros((n, 19), np.uint8)], axis=1)
    for k in range(96):
        rem = (rem << 1) | ext[:, k]
        hi = (rem & 0x2000) != 0
        rem[hi] ^= 0x2000 | 0x15D7
    rem &= 0x1FFF
    for i in range(13):
        cw[:, 77 + i] = (rem >> (12 - i)) & 1
    cw[:, 90:] = (cw[:, :90].astype(np.int64) @ _GEN.T.astype(np.int64)) & 1
    return cw


def cfg5_channel(channel: int, seconds: float = 60.0):
    rng = np.random.default_rng(1000 + channel)
    n = int(seconds * FS)
    pings = random_pings(rng, n, 20, 1260.0, 1740.0, -4.0, 10.0)
    return audio_stream(pings, n, 1000.0, 1000 + channel), pings


def wideband_capture(bins: Sequence[int], decimation: int = 160, seconds: float = 3.0, seed: int = 7, fmt: int = 1,
                     noise_lsb: float = 3.0, pings_per_channel: int = 3, snr_lo: float = 3.0, snr_hi: float = 10.0):
    """A wideband I/Q capture at decimation x 12 ksps (rtl_sdr -s 1920000 for 160) with MSK144 pings in the channels
    `bins` (offset bins[c] x 12 kHz from the centre), white noise of `noise_lsb` per component, quantised to
    fmt 1 = u8 (rtl_sdr), 2 = s8, 3 = s16 interleaved I,Q.  Returns (capture, {bin: pings})."""
    from scipy.signal import resample_poly
    rng = np.random.default_rng(seed)
    n12 = int(seconds * FS)
    M = decimation
    fs_in = M * FS
    x = noise_lsb * (rng.standard_normal(n12 * M) + 1j * rng.standard_normal(n12 * M))
    m = np.arange(n12 * M)
    truth = {}
    for k in bins:
        pings = random_pings(rng, n12, pings_per_channel, -40.0, 40.0, snr_lo, snr_hi, frames_lo=4, frames_hi=8)
        truth[k] = pings
        amp = lambda p: noise_lsb * np.sqrt(2.0 * 10 ** (p.snr_db / 10) * 2500.0 / fs_in)
        base = _render(pings, n12, amp)
        up = resample_poly(base, M, 1)                      # 12 ksps -> M x 12 ksps, unit pass-band gain
        x += up * np.exp(2j * np.pi * ((k * m) % M) / M)
    inter = np.empty(2 * len(x))
    inter[0::2], inter[1::2] = x.real, x.imag
    if fmt == 1:
        return np.clip(np.rint(inter + 127.5), 0, 255).astype(np.uint8), truth
    if fmt == 2:
        return np.clip(np.rint(inter), -128, 127).astype(np.int8), truth
    return np.clip(np.rint(inter * 64.0), -32768, 32767).astype(np.int16), truth
