"""CPU: the C-ABI library loads and exports every symbol include/msk144_b200.h declares; with no GPU the
product fails loudly instead of falling back to anything."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "msk144_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(msk144_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(product):
    names = declared_functions()
    assert len(names) >= 18
    lib = ctypes.CDLL(product.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/msk144_b200.h but not exported"
    assert set(product.EXPORTED_SYMBOLS) <= set(names)
    # the gfortran-ABI alias the reference binds in f_interop.cpp:10-29
    assert hasattr(lib, "__packjt77_MOD_unpack77")


def test_library_has_sm100a_code_only(product):
    out = subprocess.run(["cuobjdump", "-lelf", product.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    assert not re.search(r"sm_(5|6|7|8|9)\d", out)


def test_config_defaults_match_reference_code(product):
    cfg = product.Config()
    product.lib().msk144_config_defaults(ctypes.byref(cfg))
    # main.cu:124-133: the CODE defaults, not the help text
    assert (cfg.search_step_hz, cfg.search_width_hz, cfg.scan_depth) == (2.0, 200.0, 4)
    assert (cfg.analytic_method, cfg.nbadsync_threshold, cfg.read_mode) == (2, 1, 1)
    assert cfg.center_frequency_hz == 1500.0


def test_bad_config_rejected(product):
    import torch
    for kw in (dict(step=0.0), dict(read_mode=3), dict(method=5)):
        with pytest.raises(product.Msk144Error):
            product.Decoder(**kw)
    if not torch.cuda.is_available():
        with pytest.raises(product.Msk144Error) as e:   # no CPU path: creation must fail loudly
            product.Decoder()
        assert "no CUDA device" in str(e.value) or "CUDA" in str(e.value)


def test_host_text_functions_need_no_gpu(product, synth):
    product.reset_hash_tables()
    for m in synth.STANDARD_MESSAGES + ["HELLO WORLD", "TNX 73 GL"]:
        assert product.decode_message(synth.pack77(m)) == m
    assert product.decode_message(synth.pack77("TLM:0123456789ABCDEF01")) == "123456789ABCDEF01"
    line = dict(snr=5, f0=1512.0, num_avg=1, nbadsync=0, pattern_idx=0, text="CQ K1ABC FN42")
    assert product.format_line(line, "20240101120000") == \
        "***  snr= 5; f0=  1512; num_avg=1; nbadsync=0; pattern_idx=0; date=20240101120000; msg='CQ K1ABC FN42'; "
    line.update(snr=-3, f0=1437.5)
    assert "snr=-3; f0=1437.5; " in product.format_line(line, "x" * 14)


def test_i3n3_gate(product, synth):
    """decode_softbits.cpp:30: types 0.1, 0.3, 0.4, 0.6+, 3 and >5 are dropped before unpack77."""
    import numpy as np
    base = synth.pack77("CQ K1ABC FN42")
    def with_type(i3, n3):
        b = base.copy()
        b[71:74] = [(n3 >> 2) & 1, (n3 >> 1) & 1, n3 & 1]
        b[74:77] = [(i3 >> 2) & 1, (i3 >> 1) & 1, i3 & 1]
        return b
    for i3, n3 in [(0, 1), (0, 3), (0, 4), (0, 6), (0, 7), (3, 0), (6, 0), (7, 5)]:
        assert product.decode_message(with_type(i3, n3)) is None
    assert product.decode_message(np.zeros(77, np.uint8)) is None    # blank free text fails in unpack77


def test_hashed_calls_persist(product, synth):
    """A type-4 message teaches the 12-bit hash of a non-standard call; a later message resolves it."""
    import numpy as np
    product.reset_hash_tables()
    A = " 0123456789ABCDEFGHIJKLMNOPQRSTUVWXYZ/"
    call = "PJ4/K1ABC".rjust(11)
    n58 = 0
    for ch in call:
        n58 = n58 * 38 + A.index(ch)
    def t4(n12, iflip, nrpt, icq):
        bits = [(n12 >> (11 - k)) & 1 for k in range(12)] + [(n58 >> (57 - k)) & 1 for k in range(58)]
        bits += [iflip, (nrpt >> 1) & 1, nrpt & 1, icq, 1, 0, 0]
        return np.array(bits, np.uint8)
    assert product.decode_message(t4(1234, 0, 0, 1)) == "CQ PJ4/K1ABC"
    # now somebody calls PJ4/K1ABC by hash: compute its 12-bit hash the way ihashcall does
    n8 = 0
    for ch in "PJ4/K1ABC".ljust(11):
        n8 = n8 * 38 + A.index(ch)
    n12 = ((47055833459 * n8) & (2 ** 64 - 1)) >> 52
    w9 = "W9XYZ".rjust(11)
    n58b = 0
    for ch in w9:
        n58b = n58b * 38 + A.index(ch)
    bits = [(n12 >> (11 - k)) & 1 for k in range(12)] + [(n58b >> (57 - k)) & 1 for k in range(58)] + [0, 1, 0, 0, 1, 0, 0]
    assert product.decode_message(np.array(bits, np.uint8)) == "<PJ4/K1ABC> W9XYZ RR73"
    product.reset_hash_tables()
    assert product.decode_message(np.array(bits, np.uint8)) == "<...> W9XYZ RR73"
