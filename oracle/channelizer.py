"""TEST INFRASTRUCTURE -- numpy/float64 statement of the channeliser (include/msk144_b200.h, msk144_channelizer_*).

Not used by the product: only tests/ and tools/bench_channelizer.py's checker import this.  There is no reference code to
follow for this step: the reference's README documents an external csdr chain for one channel
(/root/reference/README.md:75-78: 1.92 Msps u8 I/Q -> fir_decimate_cc 8, 5, 4 -> gain_ff 100 -> convert_f_s8) and csdr
is not on this box, so parity is against the definition in the header, stated twice here in independent forms:

  channelize_direct     mix to baseband, convolve with the prototype, keep every M-th sample (the definition)
  channelize_polyphase  branch filters + one DFT row per channel (fast; equals the direct form to 1e-12)
"""
import numpy as np

T = 8
BETA = 7.857
FULL_SCALE = {1: 127.5, 2: 128.0, 3: 32768.0}


def prototype(M: int) -> np.ndarray:
    L = T * M
    n = np.arange(L) - 0.5 * (L - 1)
    h = (1.0 / M) * np.sinc(n / M) * np.kaiser(L, BETA)
    return h / h.sum()


def to_complex(capture: np.ndarray, fmt: int) -> np.ndarray:
    if fmt == 1:
        v = capture.view(np.uint8).astype(np.float64) - 127.5
    elif fmt == 2:
        v = capture.view(np.int8).astype(np.float64)
    else:
        v = capture.view(np.int16).astype(np.float64)
    return (v[0::2] + 1j * v[1::2]) / FULL_SCALE[fmt]


def quantize(y: np.ndarray, gain: float) -> np.ndarray:
    q = 127.0 * gain * y
    out = np.empty(y.shape + (2,), np.int8)
    out[..., 0] = np.clip(np.rint(q.real), -127, 127)
    out[..., 1] = np.clip(np.rint(q.imag), -127, 127)
    return out


def channelize_direct(x: np.ndarray, M: int, bins, h=None) -> np.ndarray:
    """x: complex samples at M x 12 ksps -> [nchan][nframes] complex, y_k[f] = sum_n h[n] v_k[fM + M-1 - n]"""
    h = prototype(M) if h is None else h
    nf = len(x) // M
    m = np.arange(nf * M)
    out = np.empty((len(bins), nf), np.complex128)
    for c, k in enumerate(bins):
        v = x[:nf * M] * np.exp(-2j * np.pi * ((k * m) % M) / M)
        out[c] = np.convolve(v, h)[M - 1::M][:nf]
    return out


def channelize_polyphase(x: np.ndarray, M: int, bins, h=None, history: np.ndarray = None) -> np.ndarray:
    """Same numbers through u_f[q] = sum_t h[tM + M-1-q] x[(f-t)M + q], y_k[f] = sum_q e^{-2 pi j k q / M} u_f[q].
    `history`: up to 7 frames of samples preceding x (streaming)."""
    h = prototype(M) if h is None else h
    nf = len(x) // M
    hist = np.zeros((T - 1) * M, np.complex128)
    if history is not None and len(history):
        hist[len(hist) - len(history):] = history
    frames = np.concatenate([hist, x[:nf * M]]).reshape(nf + T - 1, M)
    g = h.reshape(T, M)[:, ::-1]                       # g[t][q] = h[tM + M-1-q]
    u = np.zeros((nf, M), np.complex128)
    for t in range(T):
        u += g[t][None, :] * frames[T - 1 - t:T - 1 - t + nf]
    q = np.arange(M)
    w = np.exp(-2j * np.pi * ((np.asarray(bins)[:, None] * q[None, :]) % M) / M)
    return (u @ w.T).T.copy()
