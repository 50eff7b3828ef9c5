"""ctypes binding of the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  Nothing under msk144cudecoder_b200/ imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

NSAMP, NLLR, NMSG = 5184, 128, 77


class Item(C.Structure):
    _fields_ = [("freq_idx", C.c_uint32), ("pattern_idx", C.c_uint32), ("pos", C.c_uint32), ("f0", C.c_float),
                ("nbadsync", C.c_int32), ("xb", C.c_float), ("num_avg", C.c_int32), ("llr", C.c_float * 128),
                ("is_message_present", C.c_int32), ("ldpc_iters", C.c_int32), ("ldpc_nhard", C.c_int32),
                ("message", C.c_uint8 * 80)]


class Line(C.Structure):
    _fields_ = [("block_index", C.c_int64), ("snr", C.c_int32), ("f0", C.c_float), ("num_avg", C.c_int32),
                ("nbadsync", C.c_int32), ("pattern_idx", C.c_int32), ("payload", C.c_uint8 * 80),
                ("text", C.c_char * 40)]


class Config(C.Structure):
    _fields_ = [("center_hz", C.c_float), ("width_hz", C.c_float), ("step_hz", C.c_float),
                ("scan_depth", C.c_int), ("nbadsync_threshold", C.c_int), ("read_mode", C.c_int),
                ("analytic_method", C.c_int), ("scan_mode", C.c_int), ("compat_cache", C.c_int),
                ("nthreads", C.c_int), ("wsjtx_bp", C.c_int)]


ITEM_DTYPE = np.dtype([("freq_idx", "<u4"), ("pattern_idx", "<u4"), ("pos", "<u4"), ("f0", "<f4"),
                       ("nbadsync", "<i4"), ("xb", "<f4"), ("num_avg", "<i4"), ("llr", "<f4", (128,)),
                       ("is_message_present", "<i4"), ("ldpc_iters", "<i4"), ("ldpc_nhard", "<i4"),
                       ("message", "u1", (80,))])
assert ITEM_DTYPE.itemsize == C.sizeof(Item), (ITEM_DTYPE.itemsize, C.sizeof(Item))


def build() -> str:
    subprocess.run(["make", "-s", "-C", _HERE, "liboracle.so"], check=True)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        f32p, u8p, i32p = C.POINTER(C.c_float), C.POINTER(C.c_uint8), C.POINTER(C.c_int32)
        L.orc_num_freqs.restype = C.c_int
        L.orc_num_freqs.argtypes = [C.c_float, C.c_float]
        L.orc_freq.restype = C.c_float
        L.orc_freq.argtypes = [C.c_float, C.c_float, C.c_float, C.c_int]
        L.orc_effective_depth.restype = C.c_int
        L.orc_effective_depth.argtypes = [C.c_int]
        L.orc_sync_waveform.argtypes = [f32p, f32p]
        L.orc_frontend_audio.argtypes = [C.c_void_p, C.c_int, f32p]
        L.orc_frontend_iq.argtypes = [C.c_void_p, f32p]
        L.orc_snr_powers.argtypes = [f32p, f32p]
        L.orc_snr_update.restype = C.c_float
        L.orc_snr_update.argtypes = [f32p, f32p, i32p]
        L.orc_scan.argtypes = [f32p, C.POINTER(Config), C.c_void_p]
        L.orc_softbits.argtypes = [f32p, C.c_float, C.c_int, C.c_uint32, f32p, i32p, f32p]
        L.orc_ldpc_decode.restype = C.c_int
        L.orc_ldpc_decode.argtypes = [f32p, u8p, i32p, i32p]
        L.orc_ldpc_decode_batch.argtypes = [f32p, C.c_size_t, u8p, u8p, i32p, i32p, C.c_int]
        L.orc_crc13.restype = C.c_uint32
        L.orc_crc13.argtypes = [u8p]
        L.orc_ldpc_decode_wsjtx_batch.argtypes = [f32p, C.c_size_t, u8p, u8p, u8p, i32p, i32p, C.c_int]
        L.orc_ldpc_encode.argtypes = [u8p, u8p]
        L.orc_ldpc_parity_matrix.argtypes = [u8p]
        L.orc_decode_block.argtypes = [f32p, C.POINTER(Config), C.c_void_p]
        L.orc_block_lines.restype = C.c_int
        L.orc_block_lines.argtypes = [C.c_void_p, C.c_int, C.POINTER(Config), C.c_int, C.c_int64,
                                      C.POINTER(Line), C.c_int]
        L.orc_decode_stream.restype = C.c_int64
        L.orc_decode_stream.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Config), C.c_int64, C.c_int64,
                                        C.POINTER(Line), C.c_int64, C.POINTER(C.c_int64)]
        L.orc_decode_message.restype = C.c_int
        L.orc_decode_message.argtypes = [u8p, C.c_char_p]
        L.orc_format_line.restype = C.c_int
        L.orc_format_line.argtypes = [C.POINTER(Line), C.c_char_p, C.c_char_p, C.c_int]
        L.msk144_unpack77_reset.argtypes = []
        _lib = L
    return _lib


def _f32(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _u8(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint8))


def _i32(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def config(center=1500.0, width=200.0, step=2.0, depth=4, thr=1, read_mode=1, method=2, scan_mode=1,
           compat_cache=1, nthreads=0, wsjtx_bp=0) -> Config:
    return Config(center, width, step, depth, thr, read_mode, method, scan_mode, compat_cache, nthreads, wsjtx_bp)


def num_freqs(width, step) -> int:
    return lib().orc_num_freqs(width, step)


def freq(center, width, step, idx) -> float:
    return lib().orc_freq(center, width, step, idx)


def sync_waveform() -> Tuple[np.ndarray, np.ndarray]:
    cb = np.zeros(84, np.float32)
    pp = np.zeros(12, np.float32)
    lib().orc_sync_waveform(_f32(cb), _f32(pp))
    return cb.view(np.complex64), pp


def frontend_audio(s16: np.ndarray, method: int = 2) -> np.ndarray:
    s16 = np.ascontiguousarray(s16, np.int16)
    assert s16.size == NSAMP
    out = np.zeros(2 * NSAMP, np.float32)
    lib().orc_frontend_audio(s16.ctypes.data, method, _f32(out))
    return out.view(np.complex64)


def frontend_iq(iq: np.ndarray) -> np.ndarray:
    iq = np.ascontiguousarray(iq, np.int8)
    assert iq.size == 2 * NSAMP
    out = np.zeros(2 * NSAMP, np.float32)
    lib().orc_frontend_iq(iq.ctypes.data, _f32(out))
    return out.view(np.complex64)


def snr_powers(blk: np.ndarray) -> np.ndarray:
    blk = np.ascontiguousarray(blk, np.complex64)
    p = np.zeros(8, np.float32)
    lib().orc_snr_powers(_f32(blk.view(np.float32)), _f32(p))
    return p


class SnrTracker:
    def __init__(self):
        self.state = np.zeros(1, np.float32)

    def update(self, powers8: np.ndarray) -> Tuple[float, int]:
        p = np.ascontiguousarray(powers8, np.float32)
        snri = np.zeros(1, np.int32)
        v = lib().orc_snr_update(_f32(self.state), _f32(p), _i32(snri))
        return float(v), int(snri[0])


def n_items(cfg: Config) -> int:
    return num_freqs(cfg.width_hz, cfg.step_hz) * lib().orc_effective_depth(cfg.scan_depth) * 8


def scan(blk: np.ndarray, cfg: Config) -> np.ndarray:
    blk = np.ascontiguousarray(blk, np.complex64)
    items = np.zeros(n_items(cfg), ITEM_DTYPE)
    lib().orc_scan(_f32(blk.view(np.float32)), C.byref(cfg), items.ctypes.data)
    return items


def softbits(blk: np.ndarray, freq_hz: float, pattern_idx: int, pos: int):
    blk = np.ascontiguousarray(blk, np.complex64)
    llr = np.zeros(128, np.float32)
    sb = np.zeros(144, np.float32)
    nb = np.zeros(1, np.int32)
    lib().orc_softbits(_f32(blk.view(np.float32)), freq_hz, pattern_idx, pos, _f32(llr), _i32(nb), _f32(sb))
    return llr, int(nb[0]), sb


def ldpc_decode(llr: np.ndarray):
    llr = np.ascontiguousarray(llr, np.float32)
    msg = np.zeros(80, np.uint8)
    it = np.zeros(1, np.int32)
    nh = np.zeros(1, np.int32)
    f = lib().orc_ldpc_decode(_f32(llr), _u8(msg), _i32(it), _i32(nh))
    return bool(f), msg[:77].copy(), int(it[0]), int(nh[0])


def ldpc_decode_batch(llr: np.ndarray, nthreads: int = 0):
    llr = np.ascontiguousarray(llr, np.float32).reshape(-1, 128)
    n = llr.shape[0]
    found = np.zeros(n, np.uint8)
    msg = np.zeros((n, 77), np.uint8)
    it = np.zeros(n, np.int32)
    nh = np.zeros(n, np.int32)
    lib().orc_ldpc_decode_batch(_f32(llr), n, _u8(found), _u8(msg), _i32(it), _i32(nh), nthreads)
    return found.astype(bool), msg, it, nh


def ldpc_decode_wsjtx_batch(llr: np.ndarray, apmask: np.ndarray = None, nthreads: int = 0):
    """WSJT-X bpdecode128_90 control flow (11 checks, early stop, optional a-priori mask of 128 0/1 bytes)."""
    llr = np.ascontiguousarray(llr, np.float32).reshape(-1, 128)
    n = llr.shape[0]
    found = np.zeros(n, np.uint8)
    msg = np.zeros((n, 77), np.uint8)
    it = np.zeros(n, np.int32)
    nh = np.zeros(n, np.int32)
    ap = None if apmask is None else np.ascontiguousarray(apmask, np.uint8)
    lib().orc_ldpc_decode_wsjtx_batch(_f32(llr), n, _u8(ap) if ap is not None else None, _u8(found), _u8(msg), _i32(it), _i32(nh),
                                      nthreads)
    return found.astype(bool), msg, it, nh


def crc13(bits77: np.ndarray) -> int:
    b = np.ascontiguousarray(bits77, np.uint8)
    return int(lib().orc_crc13(_u8(b)))


def ldpc_encode(bits77: np.ndarray) -> np.ndarray:
    b = np.ascontiguousarray(bits77, np.uint8)
    cw = np.zeros(128, np.uint8)
    lib().orc_ldpc_encode(_u8(b), _u8(cw))
    return cw


def parity_matrix() -> np.ndarray:
    h = np.zeros(38 * 128, np.uint8)
    lib().orc_ldpc_parity_matrix(_u8(h))
    return h.reshape(38, 128)


def decode_block(blk: np.ndarray, cfg: Config) -> np.ndarray:
    blk = np.ascontiguousarray(blk, np.complex64)
    items = np.zeros(n_items(cfg), ITEM_DTYPE)
    lib().orc_decode_block(_f32(blk.view(np.float32)), C.byref(cfg), items.ctypes.data)
    return items


def block_lines(items: np.ndarray, cfg: Config, snr: int, block_index: int = 0) -> List[dict]:
    cap = 256
    lines = (Line * cap)()
    n = lib().orc_block_lines(items.ctypes.data, len(items), C.byref(cfg), snr, block_index, lines, cap)
    return [_line_dict(lines[i]) for i in range(n)]


def _line_dict(l: Line) -> dict:
    return dict(block=int(l.block_index), snr=int(l.snr), f0=float(l.f0), num_avg=int(l.num_avg),
                nbadsync=int(l.nbadsync), pattern_idx=int(l.pattern_idx),
                payload=bytes(bytearray(l.payload)[:77]), text=l.text.decode("ascii", "replace"))


def decode_stream(raw: np.ndarray, cfg: Config, first_block: int = 0, max_blocks: int = 0,
                  cap: int = 1 << 16) -> Tuple[List[dict], int]:
    raw = np.ascontiguousarray(raw)
    lines = (Line * cap)()
    nb = C.c_int64(0)
    n = lib().orc_decode_stream(raw.ctypes.data, raw.nbytes, C.byref(cfg), first_block, max_blocks, lines, cap,
                                C.byref(nb))
    return [_line_dict(lines[i]) for i in range(min(n, cap))], int(nb.value)


def decode_message(bits77: np.ndarray) -> Optional[str]:
    b = np.ascontiguousarray(bits77, np.uint8)
    buf = C.create_string_buffer(64)
    ok = lib().orc_decode_message(_u8(b), buf)
    return buf.value.decode("ascii") if ok else None


def format_line(d: dict, date: Optional[str] = None) -> str:
    l = Line()
    l.block_index, l.snr, l.f0, l.num_avg = d["block"], d["snr"], d["f0"], d["num_avg"]
    l.nbadsync, l.pattern_idx = d["nbadsync"], d["pattern_idx"]
    l.text = d["text"].encode("ascii")
    buf = C.create_string_buffer(256)
    lib().orc_format_line(C.byref(l), date.encode() if date else None, buf, 256)
    return buf.value.decode("ascii")


def reset_hash_tables():
    lib().msk144_unpack77_reset()
