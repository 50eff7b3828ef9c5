"""Host-side adapters either side of the hot path: WAV header recognition, JSON lines,
cross-block aggregation, and the pluggable unpack77 seam (f_interop.cpp:61-71).  No GPU needed."""
import json
import os
import struct
import subprocess

import numpy as np
import pytest


def wav_header(nbytes, channels=1, rate=12000, bits=16, extra_chunk=False, fmt_tag=1):
    fmt = struct.pack("<HHIIHH", fmt_tag, channels, rate, rate * channels * bits // 8, channels * bits // 8, bits)
    chunks = b"fmt " + struct.pack("<I", len(fmt)) + fmt
    if extra_chunk:
        chunks += b"LIST" + struct.pack("<I", 5) + b"abcde" + b"\0"     # odd length: padded to a word
    chunks += b"data" + struct.pack("<I", nbytes)
    return b"RIFF" + struct.pack("<I", 4 + len(chunks) + nbytes) + b"WAVE" + chunks


def test_wav_header_recognition(product):
    h = wav_header(1000)
    assert len(h) == 44
    assert product.wav_data_offset(h + b"\0" * 10) == (44, 1, 12000, 16)
    h2 = wav_header(1000, channels=2, rate=48000, bits=8, extra_chunk=True)
    assert product.wav_data_offset(h2)[0] == len(h2) == 44 + 8 + 6
    assert product.wav_data_offset(h2)[1:] == (2, 48000, 8)
    # raw samples: not RIFF
    raw = np.arange(100, dtype=np.int16).tobytes()
    assert product.wav_data_offset(raw)[0] == 0
    assert product.wav_data_offset(b"RI")[0] == -6 and product.wav_data_offset(b"RX")[0] == 0   # undecided / raw
    for k in range(4, 44):
        assert product.wav_data_offset(h[:k])[0] == -6                                            # incomplete header
    assert product.wav_data_offset(wav_header(10, fmt_tag=3))[0] == -1                            # IEEE float: rejected
    assert product.wav_data_offset(b"RIFF\0\0\0\0AVI " + b"\0" * 40)[0] == 0                      # RIFF but not WAVE


def _line(block, text, num_avg=1, nbadsync=0, snr=5, channel=0, f0=1500.0, pattern_idx=0):
    return dict(block=block, channel=channel, snr=snr, f0=f0, num_avg=num_avg, nbadsync=nbadsync,
                pattern_idx=pattern_idx, text=text, payload=bytes([1, 0] * 38 + [1]))


def test_json_line(product):
    l = _line(7, "CQ K1ABC FN42", num_avg=3, nbadsync=2, snr=-3, channel=4, f0=1498.5, pattern_idx=2)
    d = json.loads(product.format_line_json(l, "20260101120000"))
    assert d == {"channel": 4, "block": 7, "t": 1.512, "snr": -3, "f0": 1498.5, "num_avg": 3, "nbadsync": 2,
                 "pattern_idx": 2, "date": "20260101120000", "msg": "CQ K1ABC FN42", "payload": "10" * 38 + "1"}
    # characters that need escaping survive a round trip
    q = _line(0, 'A"B\\C')
    assert json.loads(product.format_line_json(q, "x"))["msg"] == 'A"B\\C'


def test_aggregator_merges_one_ping(product):
    a = product.Aggregator(max_gap=2)
    # one ping seen in blocks 10..13 (improving, then fading), a second message interleaved, a later repeat of the
    # first message (gap 5 > 2) that must become its own event
    a.add([_line(10, "K1ABC W9XYZ EN37", num_avg=3, nbadsync=2)])
    a.add([_line(11, "K1ABC W9XYZ EN37", num_avg=1, nbadsync=1, snr=4), _line(11, "CQ DL1ABC JO62", num_avg=2)])
    a.add([_line(12, "K1ABC W9XYZ EN37", num_avg=1, nbadsync=0, snr=6, f0=1501.0)])
    a.add([_line(13, "K1ABC W9XYZ EN37", num_avg=2, nbadsync=0)])
    assert a.poll(next_block=13) == []                               # everything can still grow
    ev = a.poll(next_block=14)                                       # 11 + 2 < 14: the CQ is finished, the ping is not
    assert [(e["text"], e["first_block"], e["last_block"], e["nlines"]) for e in ev] == [("CQ DL1ABC JO62", 11, 11, 1)]
    a.add([_line(19, "K1ABC W9XYZ EN37", num_avg=4, nbadsync=3)])    # closes the first ping by itself (13 + 2 < 19)
    ev = a.poll(next_block=20)
    assert len(ev) == 1
    e = ev[0]
    assert (e["first_block"], e["last_block"], e["nlines"]) == (10, 13, 4)
    assert (e["block"], e["num_avg"], e["nbadsync"], e["snr"], e["f0"]) == (12, 1, 0, 6, 1501.0)   # the best line
    ev = a.poll(flush=True)
    assert [(e["first_block"], e["nlines"], e["num_avg"]) for e in ev] == [(19, 1, 4)]
    assert a.poll(flush=True) == []
    a.close()


def test_aggregator_keeps_channels_apart(product):
    a = product.Aggregator(max_gap=1)
    a.add([_line(b, "CQ K1ABC FN42", channel=c) for c in (0, 1) for b in (3, 4)])
    ev = a.poll(flush=True)
    assert sorted((e["channel"], e["first_block"], e["last_block"], e["nlines"]) for e in ev) == [(0, 3, 4, 2), (1, 3, 4, 2)]
    # a block-order restart on another channel must not close the first channel's event early
    a.add([_line(100, "CQ K1ABC FN42", channel=0), _line(2, "CQ K1ABC FN42", channel=1), _line(101, "CQ K1ABC FN42", channel=0)])
    ev = a.poll(flush=True)
    assert sorted((e["channel"], e["nlines"]) for e in ev) == [(0, 2), (1, 1)]
    a.close()


STUB = r"""
#include <string.h>
#include <stddef.h>
/* a stand-in for a gfortran-built WSJT-X: same module-procedure ABI (hidden character lengths last) */
void __packjt77_MOD_unpack77(char* c77, const int* nrx, char* msg, int* ok, size_t l77, size_t l37) {
    memset(msg, ' ', l37);
    if (l77 != 77 || l37 != 37 || *nrx != 1) { *ok = 0; return; }
    memcpy(msg, "EXT:", 4);
    memcpy(msg + 4, c77, 20);          /* echo the first payload bits so the caller can check the plumbing */
    *ok = (c77[0] == '1');
}
"""


def test_external_unpack77_library(product, synth, tmp_path):
    src = tmp_path / "stub.c"
    src.write_text(STUB)
    so = tmp_path / "libstub77.so"
    subprocess.run(["gcc", "-shared", "-fPIC", "-o", str(so), str(src)], check=True)
    bits = synth.pack77("K1ABC W9XYZ EN37")
    assert product.decode_message(bits) == "K1ABC W9XYZ EN37"
    try:
        assert not product.set_unpack77_library(str(tmp_path / "missing.so"))
        assert product.lib().msk144_unpack77_is_external() == 0
        assert product.decode_message(bits) == "K1ABC W9XYZ EN37"          # built-in stays after a failed load
        assert product.set_unpack77_library(str(so))
        assert product.lib().msk144_unpack77_is_external() == 1
        b1 = bits.copy()
        b1[0] = 1
        assert product.decode_message(b1) == "EXT:" + "".join(str(int(v)) for v in b1[:20])
        b0 = bits.copy()
        b0[0] = 0
        assert product.decode_message(b0) is None                           # the library's ok flag is honoured
        # the i3/n3 gate of decode_softbits.cpp:30 still runs in front of the external library
        bad = b1.copy()
        bad[74:77] = [0, 1, 1]                                              # i3 = 3
        assert product.decode_message(bad) is None
    finally:
        assert product.set_unpack77_library(None)
    assert product.decode_message(bits) == "K1ABC W9XYZ EN37"


def test_cli_help_lists_extensions(product):
    out = subprocess.run([product.CLI_PATH, "--help"], capture_output=True, text=True).stdout
    for opt in ("--search-width", "--scan-depth", "--nbadsync-threshold", "--read-mode", "--wav", "--json", "--dedupe",
                "--unpack77-lib"):
        assert opt in out
