"""msk144cudecoder_b200 -- B200-native MSK144 stream decoder (Python face of libmsk144b200.so).

The product is the C-ABI shared library built from csrc/ (include/msk144_b200.h); this package is a
thin ctypes binding plus a host-side mirror of the reference's operator interface:

    MSK144SearchContext(center, width, step, depth, nbadsync_threshold)   # msk_context.cuh:23
    do_decode(ctx, samples) -> decode lines                                # main.cu:428-532

There is no CPU path: if the library is missing, or no sm_100-class GPU is usable, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MSK144_B200_LIB") or os.path.join(_HERE, "lib", "libmsk144b200.so")  # override: A/B builds
CLI_PATH = os.path.join(_HERE, "bin", "msk144cudecoder")

BLOCK_SAMPLES, HOP_SAMPLES = 5184, 2592
FLAG_COMPAT_DECODE_CACHE = 1
FLAG_SKIP_WAV_HEADER = 2
FLAG_WSJTX_BP = 4
FLAG_BLOCKING_SYNC = 8

EXPORTED_SYMBOLS = [
    "msk144_last_error", "msk144_config_defaults", "msk144_create", "msk144_destroy", "msk144_get_info",
    "msk144_get_stage_times", "msk144_push", "msk144_poll", "msk144_reset_stream", "msk144_decode_host",
    "msk144_decode_device", "msk144_run_device", "msk144_frontend_host", "msk144_scan_host",
    "msk144_ldpc_decode_host", "msk144_ldpc_decode_device", "msk144_decode_message", "msk144_format_line",
    "msk144_set_unpack77_library", "msk144_unpack77_is_external", "msk144_wav_data_offset", "msk144_format_line_json",
    "msk144_deinterleave_device", "msk144_decode_host_interleaved", "msk144_push_interleaved",
    "msk144_aggregator_create", "msk144_aggregator_destroy", "msk144_aggregator_add", "msk144_aggregator_poll",
    "msk144_pool_last_error", "msk144_pool_create", "msk144_pool_destroy", "msk144_pool_size", "msk144_pool_member",
    "msk144_pool_decode_host", "msk144_pool_push_interleaved", "msk144_pool_poll", "msk144_pool_reset_stream",
    "msk144_measure_fp32_peak", "msk144_build_info", "msk144_selftest_ldpc_math", "msk144_set_ap_mask",
    "msk144_channelizer_create", "msk144_channelizer_destroy", "msk144_channelizer_prototype", "msk144_channelizer_launches",
    "msk144_channelize_device", "msk144_channelize_host", "msk144_channelizer_push", "msk144_channelizer_reset",
    "msk144_frontend_device_noalloc", "msk144_decode_chunk_host", "msk144_replay_channel",
]
NOALLOC_MAX_BLOCKS = 8192  # MSK144_NOALLOC_MAX_BLOCKS
IQ_U8, IQ_S8, IQ_S16 = 1, 2, 3


class Msk144Error(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [("center_frequency_hz", C.c_float), ("search_step_hz", C.c_float), ("search_width_hz", C.c_float),
                ("scan_depth", C.c_int32), ("read_mode", C.c_int32), ("analytic_method", C.c_int32),
                ("nbadsync_threshold", C.c_int32), ("device", C.c_int32), ("max_batch_blocks", C.c_int32),
                ("flags", C.c_int32), ("reserved", C.c_int32 * 6)]


class ChannelizerConfig(C.Structure):
    _fields_ = [("input_format", C.c_int32), ("decimation", C.c_int32), ("gain", C.c_float), ("device", C.c_int32),
                ("reserved", C.c_int32 * 4)]


class Line(C.Structure):
    _fields_ = [("block_index", C.c_int64), ("channel", C.c_int32), ("snr", C.c_int32), ("f0", C.c_float),
                ("num_avg", C.c_int32), ("nbadsync", C.c_int32), ("pattern_idx", C.c_int32),
                ("payload", C.c_uint8 * 80), ("text", C.c_char * 40)]


class Event(C.Structure):
    _fields_ = [("first_block", C.c_int64), ("last_block", C.c_int64), ("nlines", C.c_int32), ("reserved", C.c_int32),
                ("best", Line)]


class Info(C.Structure):
    _fields_ = [("num_freqs", C.c_int32), ("scan_depth", C.c_int32), ("items_per_block", C.c_int32),
                ("sm_count", C.c_int32), ("left_hz", C.c_float), ("right_hz", C.c_float),
                ("max_batch_blocks", C.c_int32), ("reserved", C.c_int32 * 9)]


class StageTimes(C.Structure):
    _fields_ = [("frontend_ms", C.c_float), ("scan_ms", C.c_float), ("ldpc_ms", C.c_float), ("total_ms", C.c_float),
                ("blocks", C.c_int64), ("llr_codewords", C.c_int64), ("ldpc_successes", C.c_int64),
                ("kernel_launches", C.c_int64), ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64),
                ("groups", C.c_int64)]


RESULT_DTYPE = np.dtype([("block_index", "<i8"), ("channel", "<i4"), ("item_index", "<u4"), ("freq_idx", "<u4"),
                         ("pattern_idx", "<u4"), ("slot", "<u4"), ("pos", "<u4"), ("f0", "<f4"), ("xb", "<f4"),
                         ("num_avg", "<i4"), ("nbadsync", "<i4"), ("ldpc_iters", "<i4"), ("ldpc_nhard", "<i4"),
                         ("message", "u1", (80,))], align=True)
CAND_DTYPE = np.dtype([("pos", "<u4"), ("xb", "<f4"), ("nbadsync", "<i4"), ("llr_slot", "<i4")])
assert RESULT_DTYPE.itemsize == 136, RESULT_DTYPE.itemsize

_lib = None


def build(verbose: bool = False) -> str:
    """Compile the CUDA library and the CLI in-tree (nvcc -gencode arch=compute_100a,code=sm_100a)."""
    r = subprocess.run(["make", "-C", os.path.join(_HERE, "csrc")], capture_output=not verbose, text=True)
    if r.returncode != 0:
        raise Msk144Error("build failed:\n" + (r.stdout or "") + (r.stderr or ""))
    return LIB_PATH


def lib():
    """Load libmsk144b200.so.  Fails loudly when it has not been built: there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise Msk144Error(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(the decoder has no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp, sz = C.c_void_p, C.c_size_t
        L.msk144_last_error.restype = C.c_char_p
        L.msk144_config_defaults.argtypes = [C.POINTER(Config)]
        L.msk144_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
        L.msk144_destroy.argtypes = [vp]
        L.msk144_destroy.restype = None
        L.msk144_get_info.argtypes = [vp, C.POINTER(Info)]
        L.msk144_get_stage_times.argtypes = [vp, C.POINTER(StageTimes)]
        L.msk144_push.argtypes = [vp, vp, sz]
        L.msk144_push_interleaved.argtypes = [vp, vp, sz, C.c_int]
        L.msk144_poll.argtypes = [vp, C.POINTER(Line), sz, C.POINTER(sz)]
        L.msk144_reset_stream.argtypes = [vp]
        L.msk144_decode_host.argtypes = [vp, vp, sz, C.c_int, C.POINTER(Line), sz, C.POINTER(sz), vp, sz, C.POINTER(sz)]
        L.msk144_decode_device.argtypes = [vp, vp, sz, C.c_int, vp, C.POINTER(Line), sz, C.POINTER(sz), vp, sz,
                                           C.POINTER(sz)]
        L.msk144_decode_host_interleaved.argtypes = L.msk144_decode_host.argtypes
        L.msk144_deinterleave_device.argtypes = [vp, vp, vp, sz, C.c_int, vp]
        L.msk144_run_device.argtypes = [vp, vp, sz, C.c_int, vp, C.POINTER(C.c_int64)]
        L.msk144_frontend_host.argtypes = [vp, vp, sz, vp, vp, C.POINTER(C.c_int64)]
        L.msk144_frontend_device_noalloc.argtypes = [vp, sz, C.c_int, vp, vp, vp, C.POINTER(C.c_int64)]
        L.msk144_decode_chunk_host.argtypes = [vp, vp, sz, C.c_int64, vp, sz, C.POINTER(sz), vp, sz, C.POINTER(C.c_int64)]
        L.msk144_replay_channel.argtypes = [C.c_int32, C.c_int32, C.c_int64, C.c_int64, vp, vp, sz, C.POINTER(Line), sz, C.POINTER(sz)]
        L.msk144_scan_host.argtypes = [vp, vp, C.c_int64, vp, vp, vp, sz, C.POINTER(sz)]
        L.msk144_ldpc_decode_host.argtypes = [vp, vp, sz, vp, vp, vp, vp]
        L.msk144_ldpc_decode_device.argtypes = [vp, vp, sz, vp, vp]
        L.msk144_selftest_ldpc_math.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.msk144_decode_message.argtypes = [vp, C.c_char_p]
        L.msk144_format_line.argtypes = [C.POINTER(Line), C.c_char_p, C.c_char_p, sz]
        L.msk144_unpack77_reset.argtypes = []
        L.msk144_set_unpack77_library.argtypes = [C.c_char_p]
        L.msk144_wav_data_offset.argtypes = [vp, sz, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.msk144_wav_data_offset.restype = C.c_int64
        L.msk144_format_line_json.argtypes = [C.POINTER(Line), C.c_char_p, C.c_char_p, sz]
        L.msk144_aggregator_create.argtypes = [C.c_int32, C.POINTER(vp)]
        L.msk144_aggregator_destroy.argtypes = [vp]
        L.msk144_aggregator_destroy.restype = None
        L.msk144_aggregator_add.argtypes = [vp, C.POINTER(Line), sz]
        L.msk144_aggregator_poll.argtypes = [vp, C.c_int64, C.c_int, C.POINTER(Event), sz, C.POINTER(sz)]
        L.msk144_pool_last_error.restype = C.c_char_p
        L.msk144_pool_create.argtypes = [C.POINTER(Config), C.POINTER(C.c_int32), C.c_int32, C.POINTER(vp)]
        L.msk144_pool_destroy.argtypes = [vp]
        L.msk144_pool_destroy.restype = None
        L.msk144_pool_size.argtypes = [vp]
        L.msk144_pool_member.argtypes = [vp, C.c_int32]
        L.msk144_pool_member.restype = vp
        L.msk144_pool_decode_host.argtypes = L.msk144_decode_host.argtypes
        L.msk144_last_output.argtypes = [vp, C.POINTER(Line), sz, C.POINTER(sz), vp, sz, C.POINTER(sz)]
        L.msk144_pool_last_output.argtypes = L.msk144_last_output.argtypes
        L.msk144_pool_push_interleaved.argtypes = [vp, vp, sz, C.c_int]
        L.msk144_pool_poll.argtypes = [vp, C.POINTER(Line), sz, C.POINTER(sz)]
        L.msk144_pool_reset_stream.argtypes = [vp]
        L.msk144_measure_fp32_peak.argtypes = [vp, C.POINTER(C.c_double)]
        L.msk144_build_info.restype = C.c_char_p
        L.msk144_set_ap_mask.argtypes = [vp, vp]
        L.msk144_channelizer_create.argtypes = [C.POINTER(ChannelizerConfig), C.POINTER(C.c_int32), C.c_int32, C.POINTER(vp)]
        L.msk144_channelizer_destroy.argtypes = [vp]
        L.msk144_channelizer_destroy.restype = None
        L.msk144_channelizer_prototype.argtypes = [vp, vp, sz, C.POINTER(sz)]
        L.msk144_channelizer_launches.argtypes = [vp]
        L.msk144_channelizer_launches.restype = C.c_int64
        L.msk144_channelize_device.argtypes = [vp, vp, sz, C.c_int32, vp, sz, sz, vp]
        L.msk144_channelize_host.argtypes = [vp, vp, sz, vp, sz, C.POINTER(sz)]
        L.msk144_channelizer_push.argtypes = [vp, vp, sz, vp, sz, C.POINTER(sz)]
        L.msk144_channelizer_reset.argtypes = [vp]
        _lib = L
    return _lib


def build_info() -> str:
    """Toolkit the library was compiled with (the LDPC kernel's tanh reproduces that toolkit's libdevice)."""
    return lib().msk144_build_info().decode("ascii")


def _check(rc: int):
    if rc != 0:
        raise Msk144Error(f"msk144 error {rc}: {lib().msk144_last_error().decode('utf-8', 'replace')}")


def _line_dict(l: Line) -> dict:
    return dict(block=int(l.block_index), channel=int(l.channel), snr=int(l.snr), f0=float(l.f0),
                num_avg=int(l.num_avg), nbadsync=int(l.nbadsync), pattern_idx=int(l.pattern_idx),
                payload=bytes(bytearray(l.payload)[:77]), text=l.text.decode("ascii", "replace"))


class Decoder:
    """One GPU decoder instance (msk144_create .. msk144_destroy)."""

    def __init__(self, center: Optional[float] = None, width: float = 200.0, step: float = 2.0, depth: int = 4,
                 thr: int = 1, read_mode: int = 1, method: int = 2, device: int = 0, max_batch_blocks: int = 128,
                 compat_cache: bool = False, flags: int = 0, wsjtx_bp: bool = False):
        L = lib()
        flags |= FLAG_WSJTX_BP if wsjtx_bp else 0
        cfg = _make_config(center, width, step, depth, thr, read_mode, method, device, max_batch_blocks, compat_cache, flags)
        self.cfg = cfg
        self._h = C.c_void_p()
        _check(L.msk144_create(C.byref(cfg), C.byref(self._h)))
        info = Info()
        _check(L.msk144_get_info(self._h, C.byref(info)))
        self.num_freqs, self.depth, self.items_per_block = info.num_freqs, info.scan_depth, info.items_per_block
        self.sm_count = info.sm_count
        self.left_hz, self.right_hz = info.left_hz, info.right_hz

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().msk144_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- batch ---------------------------------------------------------------------------------
    def _decode(self, fn, ptr, spc, nchan, extra, line_cap, result_cap):
        lines = (Line * line_cap)()
        results = np.zeros(result_cap, RESULT_DTYPE)
        nl, nr = C.c_size_t(0), C.c_size_t(0)
        rc = fn(self._h, ptr, spc, nchan, *extra, lines, line_cap, C.byref(nl),
                results.ctypes.data if result_cap else None, result_cap, C.byref(nr) if result_cap else None)
        if rc == -5 and (nl.value > line_cap or nr.value > result_cap):
            # the decoder keeps its output: fetch it into arrays of the reported sizes (no second decode)
            line_cap, result_cap = max(line_cap, nl.value), (max(result_cap, nr.value) if result_cap else 0)
            lines = (Line * line_cap)()
            results = np.zeros(result_cap, RESULT_DTYPE)
            rc = lib().msk144_last_output(self._h, lines, line_cap, C.byref(nl), results.ctypes.data if result_cap else None,
                                          result_cap, C.byref(nr) if result_cap else None)
        _check(rc)
        return [_line_dict(lines[i]) for i in range(nl.value)], results[:nr.value].copy()

    def decode_host(self, samples: np.ndarray, nchan: int = 1, line_cap: int = 4096, result_cap: int = 1 << 16,
                    interleaved: bool = False):
        """samples: int16 (audio) or int8 I,Q pairs, `nchan` equal-length channels back to back, or -- with
        interleaved=True -- one sample of every channel per time step ([t][c]).
        result_cap = 0: decode lines only (the per-candidate LDPC successes are not returned)."""
        samples = np.ascontiguousarray(samples)
        spc = samples.nbytes // 2 // nchan
        fn = lib().msk144_decode_host_interleaved if interleaved else lib().msk144_decode_host
        return self._decode(fn, samples.ctypes.data, spc, nchan, (), line_cap, result_cap)

    def decode_chunk_host(self, samples: np.ndarray, block_base: int = 0, result_cap: int = 1 << 16):
        """GPU stages of a contiguous piece of ONE channel that starts at block `block_base` (msk144_decode_chunk_host):
        returns (results, powers[nblocks, 8]); no SNR, no text -- `replay_channel` does both over the pieces in order."""
        samples = np.ascontiguousarray(samples)
        ns = samples.nbytes // 2
        nb = (ns - BLOCK_SAMPLES) // HOP_SAMPLES + 1 if ns >= BLOCK_SAMPLES else 0
        results = np.zeros(max(result_cap, 1), RESULT_DTYPE)
        powers = np.zeros((nb, 8), np.float32)
        nr, got = C.c_size_t(0), C.c_int64(0)
        rc = lib().msk144_decode_chunk_host(self._h, samples.ctypes.data, ns, block_base, results.ctypes.data, results.size,
                                            C.byref(nr), powers.ctypes.data, nb, C.byref(got))
        if rc == -5:
            results = np.zeros(nr.value, RESULT_DTYPE)
            rc = lib().msk144_last_output(self._h, None, 0, None, results.ctypes.data, results.size, C.byref(nr))
        _check(rc)
        return results[:nr.value].copy(), powers

    def deinterleave_device(self, d_in: int, d_out: int, samples_per_channel: int, nchan: int, stream: int = 0):
        _check(lib().msk144_deinterleave_device(self._h, d_in, d_out, samples_per_channel, nchan, C.c_void_p(stream)))

    def decode_device(self, dev_ptr: int, samples_per_channel: int, nchan: int = 1, stream: int = 0,
                      line_cap: int = 4096, result_cap: int = 1 << 16):
        return self._decode(lib().msk144_decode_device, dev_ptr, samples_per_channel, nchan, (C.c_void_p(stream),),
                            line_cap, result_cap)

    def run_device(self, dev_ptr: int, samples_per_channel: int, nchan: int = 1, stream: int = 0) -> int:
        nb = C.c_int64(0)
        _check(lib().msk144_run_device(self._h, dev_ptr, samples_per_channel, nchan, C.c_void_p(stream), C.byref(nb)))
        return int(nb.value)

    def stage_times(self) -> dict:
        t = StageTimes()
        _check(lib().msk144_get_stage_times(self._h, C.byref(t)))
        return {k: getattr(t, k) for k, _ in StageTimes._fields_}

    # -- streaming -------------------------------------------------------------------------------
    def push(self, data: bytes, nchan: int = 1):
        """Raw stdin bytes of one channel, or (nchan > 1) of `nchan` channels interleaved per time step."""
        if nchan > 1:
            _check(lib().msk144_push_interleaved(self._h, data, len(data), nchan))
        else:
            _check(lib().msk144_push(self._h, data, len(data)))

    def poll(self, cap: int = 1024) -> List[dict]:
        lines = (Line * cap)()
        n = C.c_size_t(0)
        _check(lib().msk144_poll(self._h, lines, cap, C.byref(n)))
        return [_line_dict(lines[i]) for i in range(n.value)]

    def reset_stream(self):
        _check(lib().msk144_reset_stream(self._h))

    # -- stages ----------------------------------------------------------------------------------
    def frontend(self, samples: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        samples = np.ascontiguousarray(samples)
        ns = samples.nbytes // 2
        nb = (ns - BLOCK_SAMPLES) // HOP_SAMPLES + 1 if ns >= BLOCK_SAMPLES else 0
        blocks = np.zeros((nb, BLOCK_SAMPLES), np.complex64)
        powers = np.zeros((nb, 8), np.float32)
        n = C.c_int64(0)
        _check(lib().msk144_frontend_host(self._h, samples.ctypes.data, ns, blocks.ctypes.data, powers.ctypes.data,
                                          C.byref(n)))
        return blocks, powers

    def scan(self, blocks: np.ndarray, want_llr: bool = True):
        """blocks: complex64 [nblocks, 5184].  Returns (cands[nblocks, items], llr[n,128], llr_item[n]);
        want_llr=False returns only the candidates (no 512 B per gated candidate on the host)."""
        blocks = np.ascontiguousarray(blocks, np.complex64).reshape(-1, BLOCK_SAMPLES)
        nb = blocks.shape[0]
        cands = np.zeros((nb, self.items_per_block), CAND_DTYPE)
        if not want_llr:
            _check(lib().msk144_scan_host(self._h, blocks.ctypes.data, nb, cands.ctypes.data, None, None, 0, None))
            return cands
        cap = nb * self.items_per_block
        llr = np.zeros((cap, 128), np.float32)
        item = np.zeros(cap, np.uint32)
        n = C.c_size_t(0)
        _check(lib().msk144_scan_host(self._h, blocks.ctypes.data, nb, cands.ctypes.data, llr.ctypes.data,
                                      item.ctypes.data, cap, C.byref(n)))
        return cands, llr[:n.value].copy(), item[:n.value].copy()

    def ldpc_decode(self, llr: np.ndarray):
        llr = np.ascontiguousarray(llr, np.float32).reshape(-1, 128)
        n = llr.shape[0]
        found = np.zeros(n, np.uint8)
        msg = np.zeros((n, 77), np.uint8)
        iters = np.zeros(n, np.int32)
        nhard = np.zeros(n, np.int32)
        _check(lib().msk144_ldpc_decode_host(self._h, llr.ctypes.data, n, found.ctypes.data, msg.ctypes.data,
                                             iters.ctypes.data, nhard.ctypes.data))
        return found.astype(bool), msg, iters, nhard

    def selftest_ldpc_math(self) -> Tuple[int, int]:
        """All 2^32 floats through the LDPC kernel's packed tanh / platanh vs their scalar definitions:
        (tanh mismatches, platanh mismatches)."""
        a, b = C.c_uint64(0), C.c_uint64(0)
        _check(lib().msk144_selftest_ldpc_math(self._h, C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)

    def ldpc_decode_device(self, d_llr: int, ncw: int, d_out: int, stream: int = 0):
        _check(lib().msk144_ldpc_decode_device(self._h, d_llr, ncw, d_out, C.c_void_p(stream)))

    def set_ap_mask(self, apmask128: Optional[np.ndarray]):
        """wsjtx_bp decoders only: bits whose entry is non-zero keep zn = llr (bpdecode128_90's apmask); None clears."""
        if apmask128 is None:
            _check(lib().msk144_set_ap_mask(self._h, None))
        else:
            a = np.ascontiguousarray(apmask128, np.uint8)
            assert a.size == 128
            _check(lib().msk144_set_ap_mask(self._h, a.ctypes.data))

    def measure_fp32_peak(self) -> float:
        """Sustained FP32 FMA rate of this decoder's GPU in TFLOP/s (dependency-free FFMA stream, CUDA events)."""
        v = C.c_double(0.0)
        _check(lib().msk144_measure_fp32_peak(self._h, C.byref(v)))
        return float(v.value)


def _make_config(center, width, step, depth, thr, read_mode, method, device, max_batch_blocks, compat_cache, flags) -> Config:
    cfg = Config()
    lib().msk144_config_defaults(C.byref(cfg))
    cfg.center_frequency_hz = (1500.0 if read_mode == 1 else 0.0) if center is None else center
    cfg.search_width_hz, cfg.search_step_hz, cfg.scan_depth = width, step, depth
    cfg.nbadsync_threshold, cfg.read_mode, cfg.analytic_method = thr, read_mode, method
    cfg.device, cfg.max_batch_blocks = device, max_batch_blocks
    cfg.flags = (FLAG_COMPAT_DECODE_CACHE if compat_cache else 0) | flags
    return cfg


class Pool:
    """One decoder per listed GPU behind one handle (msk144_pool_*): channels are split into contiguous ranges, one
    host thread and one set of CUDA streams per GPU, lines merged on the host in single-decoder order."""

    def __init__(self, devices: List[int], center: Optional[float] = None, width: float = 200.0, step: float = 2.0,
                 depth: int = 4, thr: int = 1, read_mode: int = 1, method: int = 2, max_batch_blocks: int = 128,
                 compat_cache: bool = False, flags: int = 0):
        cfg = _make_config(center, width, step, depth, thr, read_mode, method, 0, max_batch_blocks, compat_cache, flags)
        devs = (C.c_int32 * len(devices))(*devices)
        self._h = C.c_void_p()
        rc = lib().msk144_pool_create(C.byref(cfg), devs, len(devices), C.byref(self._h))
        if rc != 0:
            raise Msk144Error(f"msk144 pool error {rc}: {lib().msk144_pool_last_error().decode('utf-8', 'replace')}")
        self.devices = list(devices)

    def _check(self, rc):
        if rc != 0:
            raise Msk144Error(f"msk144 pool error {rc}: {lib().msk144_pool_last_error().decode('utf-8', 'replace')}")

    def decode_host(self, samples: np.ndarray, nchan: int, line_cap: int = 4096, result_cap: int = 0):
        samples = np.ascontiguousarray(samples)
        spc = samples.nbytes // 2 // nchan
        lines = (Line * line_cap)()
        results = np.zeros(result_cap, RESULT_DTYPE)
        nl, nr = C.c_size_t(0), C.c_size_t(0)
        rc = lib().msk144_pool_decode_host(self._h, samples.ctypes.data, spc, nchan, lines, line_cap, C.byref(nl),
                                           results.ctypes.data if result_cap else None, result_cap,
                                           C.byref(nr) if result_cap else None)
        if rc == -5 and (nl.value > line_cap or nr.value > result_cap):  # the pool keeps its output: no second decode
            line_cap, result_cap = max(line_cap, nl.value), (max(result_cap, nr.value) if result_cap else 0)
            lines = (Line * line_cap)()
            results = np.zeros(result_cap, RESULT_DTYPE)
            rc = lib().msk144_pool_last_output(self._h, lines, line_cap, C.byref(nl),
                                               results.ctypes.data if result_cap else None, result_cap,
                                               C.byref(nr) if result_cap else None)
        self._check(rc)
        return [_line_dict(lines[i]) for i in range(nl.value)], results[:nr.value].copy()

    def push(self, data: bytes, nchan: int):
        self._check(lib().msk144_pool_push_interleaved(self._h, data, len(data), nchan))

    def poll(self, cap: int = 1024) -> List[dict]:
        out = []
        while True:
            lines = (Line * cap)()
            n = C.c_size_t(0)
            self._check(lib().msk144_pool_poll(self._h, lines, cap, C.byref(n)))
            out += [_line_dict(lines[i]) for i in range(n.value)]
            if n.value < cap:
                return out

    def reset_stream(self):
        self._check(lib().msk144_pool_reset_stream(self._h))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().msk144_pool_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


class Channelizer:
    """Wideband I/Q capture -> channels of 12 ksps s8 I/Q (msk144_channelizer_*): the GPU version of the README's
    `csdr fir_decimate_cc 8 | 5 | 4 | gain_ff 100 | convert_f_s8` chain (README.md:75-78), for many channels at once.
    `bins[c]` x 12 kHz is the offset of channel c from the centre of the capture."""

    def __init__(self, decimation: int, bins: List[int], input_format: int = IQ_U8, gain: float = 100.0, device: int = 0):
        cfg = ChannelizerConfig(input_format=input_format, decimation=decimation, gain=gain, device=device)
        arr = (C.c_int32 * len(bins))(*bins)
        self._h = C.c_void_p()
        _check(lib().msk144_channelizer_create(C.byref(cfg), arr, len(bins), C.byref(self._h)))
        self.bins, self.decimation, self.input_format = list(bins), decimation, input_format
        self.sample_bytes = 4 if input_format == IQ_S16 else 2

    def prototype(self) -> np.ndarray:
        n = C.c_size_t(0)
        _check(lib().msk144_channelizer_prototype(self._h, None, 0, C.byref(n)))
        taps = np.zeros(n.value, np.float32)
        _check(lib().msk144_channelizer_prototype(self._h, taps.ctypes.data, n.value, C.byref(n)))
        return taps

    def launches(self) -> int:
        return int(lib().msk144_channelizer_launches(self._h))

    def channelize_host(self, capture: np.ndarray) -> np.ndarray:
        """capture: interleaved I,Q (u8 / s8 / s16 as configured) -> int8 [nchan][nframes][2]"""
        capture = np.ascontiguousarray(capture)
        nf = capture.nbytes // (self.decimation * self.sample_bytes)
        out = np.zeros((len(self.bins), nf, 2), np.int8)
        n = C.c_size_t(0)
        _check(lib().msk144_channelize_host(self._h, capture.ctypes.data, capture.nbytes, out.ctypes.data, nf, C.byref(n)))
        assert n.value == nf
        return out

    def channelize_device(self, d_in: int, nframes: int, d_out: int, chan_stride: int, frame_stride: int = 1,
                          history_frames: int = 0, stream: int = 0):
        _check(lib().msk144_channelize_device(self._h, d_in, nframes, history_frames, d_out, chan_stride, frame_stride, stream))

    def push(self, data: bytes) -> np.ndarray:
        """-> int8 [nframes][nchan][2], the time-major layout Decoder.push(..., nchan) takes"""
        cap = (len(data) // (self.decimation * self.sample_bytes)) + 1
        out = np.zeros((cap, len(self.bins), 2), np.int8)
        n = C.c_size_t(0)
        _check(lib().msk144_channelizer_push(self._h, data, len(data), out.ctypes.data, cap, C.byref(n)))
        return out[:n.value]

    def reset(self):
        _check(lib().msk144_channelizer_reset(self._h))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().msk144_channelizer_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


# ---- mirror of the reference's operator interface ---------------------------------------------------
class MSK144SearchContext(Decoder):
    """MSK144SearchContext(center_freq, search_width, search_step, scan_depth, nbadsync_threshold)
    -- same argument order and meaning as msk_context.cuh:23."""

    def __init__(self, center_freq: float, search_width: float, search_step: float, scan_depth: int,
                 nbadsync_threshold: int, **kw):
        super().__init__(center=center_freq, width=search_width, step=search_step, depth=scan_depth,
                         thr=nbadsync_threshold, **kw)

    def scanDepth(self) -> int:  # msk_context.cuh:160
        return self.depth

    def leftBound(self) -> float:  # msk_context.cuh:127
        return self.left_hz

    def rightBound(self) -> float:
        return self.right_hz

    def getNBadSyncThreshold(self) -> int:
        return int(self.cfg.nbadsync_threshold)


def do_decode(ctx: MSK144SearchContext, samples: np.ndarray, nchan: int = 1):
    """main.cu:428-532 for every block of `samples` (raw stdin samples, not a pre-filtered block: the
    front end runs on the GPU too).  Returns (lines, results)."""
    return ctx.decode_host(samples, nchan)


def frontend_device_noalloc(d_samples: int, nsamples: int, read_mode: int, d_blocks: int, d_powers: int,
                            stream: int = 0) -> int:
    """msk144_frontend_device_noalloc: the method-2 / IQ front end over device buffers the caller already owns, for
    callers that may not allocate (see include/msk144_b200.h).  All pointers are device addresses on the current CUDA
    device; asynchronous on `stream`.  Returns the number of blocks written.  At most NOALLOC_MAX_BLOCKS blocks per
    call; calls that can overlap on the GPU share one scratch array and must be serialised by the caller."""
    n = C.c_int64(0)
    _check(lib().msk144_frontend_device_noalloc(d_samples, nsamples, read_mode, d_blocks, d_powers, stream, C.byref(n)))
    return n.value


def chunk_ranges(nblocks: int, parts: int) -> List[Tuple[int, int]]:
    """Contiguous, balanced block ranges [b0, b1) of one channel for `parts` decoders (the first nblocks % parts get one
    more); the piece of blocks [b0, b1) is samples [2592 b0, 2592 (b1 - 1) + 5184)."""
    base, extra = divmod(nblocks, parts)
    out, b = [], 0
    for j in range(parts):
        e = b + base + (1 if j < extra else 0)
        out.append((b, e))
        b = e
    return out


def replay_channel(powers: np.ndarray, results: np.ndarray, channel: int = 0, first_block: int = 0,
                   compat_cache: bool = False) -> List[dict]:
    """The ordered host pass over the output of `Decoder.decode_chunk_host` pieces (msk144_replay_channel, host only):
    powers [nblocks, 8] and results concatenated in block order -> the decode lines of the undivided channel."""
    powers = np.ascontiguousarray(powers, np.float32).reshape(-1, 8)
    results = np.ascontiguousarray(results, RESULT_DTYPE)
    flags = FLAG_COMPAT_DECODE_CACHE if compat_cache else 0
    cap = 4096
    while True:
        lines = (Line * cap)()
        n = C.c_size_t(0)
        rc = lib().msk144_replay_channel(flags, channel, first_block, powers.shape[0], powers.ctypes.data,
                                         results.ctypes.data if results.size else None, results.size, lines, cap, C.byref(n))
        if rc == -5 and n.value > cap:
            cap = n.value
            continue
        _check(rc)
        return [_line_dict(lines[i]) for i in range(n.value)]


def decode_message(bits77: np.ndarray) -> Optional[str]:
    """decode_softbits.cpp:23-55."""
    b = np.ascontiguousarray(bits77, np.uint8)
    buf = C.create_string_buffer(64)
    ok = lib().msk144_decode_message(b.ctypes.data, buf)
    return buf.value.decode("ascii") if ok else None


def format_line(d: dict, date: Optional[str] = None) -> str:
    """main.cu:406-419."""
    l = Line()
    l.snr, l.f0, l.num_avg, l.nbadsync, l.pattern_idx = d["snr"], d["f0"], d["num_avg"], d["nbadsync"], d["pattern_idx"]
    l.text = d["text"].encode("ascii")
    buf = C.create_string_buffer(256)
    lib().msk144_format_line(C.byref(l), date.encode() if date else None, buf, 256)
    return buf.value.decode("ascii")


def _line_struct(d: dict) -> Line:
    l = Line()
    l.block_index, l.channel = d.get("block", 0), d.get("channel", 0)
    l.snr, l.f0, l.num_avg, l.nbadsync, l.pattern_idx = d["snr"], d["f0"], d["num_avg"], d["nbadsync"], d["pattern_idx"]
    l.text = d["text"].encode("ascii")
    pl = d.get("payload")
    if pl is not None:
        C.memmove(l.payload, bytes(pl), min(80, len(pl)))
    return l


def format_line_json(d: dict, date: Optional[str] = None) -> str:
    """Machine-readable twin of format_line (one JSON object per line)."""
    buf = C.create_string_buffer(512)
    lib().msk144_format_line_json(C.byref(_line_struct(d)), date.encode() if date else None, buf, 512)
    return buf.value.decode("ascii")


def wav_data_offset(head: bytes):
    """(offset, channels, sample_rate, bits): offset > 0 = first sample byte of a RIFF/WAVE file, 0 = raw samples,
    -6 = need more bytes, -1 = RIFF but not PCM."""
    ch, sr, bits = C.c_int32(0), C.c_int32(0), C.c_int32(0)
    off = lib().msk144_wav_data_offset(head, len(head), C.byref(ch), C.byref(sr), C.byref(bits))
    return int(off), ch.value, sr.value, bits.value


def set_unpack77_library(path: Optional[str]) -> bool:
    """Route text decoding through a library exporting __packjt77_MOD_unpack77 (None = built-in)."""
    return lib().msk144_set_unpack77_library(path.encode() if path else None) == 0


class Aggregator:
    """One event per ping: merges the per-block repeats of a (channel, text) at most `max_gap` blocks apart."""

    def __init__(self, max_gap: int = 2):
        self._h = C.c_void_p()
        _check(lib().msk144_aggregator_create(max_gap, C.byref(self._h)))

    def add(self, lines: List[dict]):
        arr = (Line * max(1, len(lines)))(*[_line_struct(l) for l in lines])
        _check(lib().msk144_aggregator_add(self._h, arr, len(lines)))

    def poll(self, next_block: int = 0, flush: bool = False) -> List[dict]:
        out, n, ev = [], C.c_size_t(0), (Event * 256)()
        while True:
            _check(lib().msk144_aggregator_poll(self._h, next_block, 1 if flush else 0, ev, 256, C.byref(n)))
            out += [dict(_line_dict(ev[i].best), first_block=int(ev[i].first_block), last_block=int(ev[i].last_block),
                         nlines=int(ev[i].nlines)) for i in range(n.value)]
            if n.value < 256:
                return out

    def close(self):
        if self._h:
            lib().msk144_aggregator_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def reset_hash_tables():
    lib().msk144_unpack77_reset()
