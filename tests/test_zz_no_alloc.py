"""msk144_frontend_device_noalloc: the front end for callers that are not allowed to allocate GPU memory.

Who that is: a plugin inside a host program that owns the device's memory budget (here torch's caching allocator
plays that host: the test hands the library tensors it allocated and the library must not take a byte more), a live
receive loop where cudaMalloc / cudaFree would stall every stream of the device, a process that gave all free HBM
away.  The entry point keeps its only scratch (1 / rms per block) in a fixed array inside the library's device image,
so these tests check three things: same bits as the allocating entry point (msk144_frontend_host), on calls that
follow each other and so reuse the array; the stated size limit is enforced with a readable error; free device memory
does not drop across a call.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a failure here should not hide the
parity tests.  The GPU tests call the new entry point BEFORE creating any decoder, so that run alone in a fresh
process they also cover its own per-device set-up.
"""
import numpy as np
import pytest

N, HOP = 5184, 2592


def _device_run(product, torch, raw, read_mode, byte_offset=0):
    """raw: int16 [nsamples] (audio) or int8 [2 * nsamples] (IQ).  Returns (blocks complex64 [nb, N], powers [nb, 8])
    computed by the no-allocation entry point from device buffers this test owns."""
    nsamples = raw.nbytes // 2
    nb = (nsamples - N) // HOP + 1 if nsamples >= N else 0
    d_in = torch.zeros(raw.nbytes + 64, dtype=torch.uint8, device="cuda")
    d_in[byte_offset:byte_offset + raw.nbytes] = torch.from_numpy(raw.view(np.uint8)).cuda()
    d_blocks = torch.full((max(nb, 1), N, 2), float("nan"), dtype=torch.float32, device="cuda")
    d_powers = torch.full((max(nb, 1), 8), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    s = torch.cuda.current_stream()
    free_before = torch.cuda.mem_get_info()[0]
    got = product.frontend_device_noalloc(d_in.data_ptr() + byte_offset, nsamples, read_mode, d_blocks.data_ptr(),
                                          d_powers.data_ptr(), s.cuda_stream)
    assert got == nb
    torch.cuda.synchronize()
    _device_run.freed = torch.cuda.mem_get_info()[0] - free_before  # bytes of free device memory gained over the call
    blocks = np.ascontiguousarray(d_blocks.cpu().numpy()[:nb])
    return blocks.view(np.complex64)[..., 0], d_powers.cpu().numpy()[:nb]


def _same_bits(a, b):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))


@pytest.mark.gpu
def test_noalloc_frontend_equals_allocating_entry_point_on_calls_in_a_row(product, synth):
    import torch
    assert torch.cuda.is_available(), "needs the B200"
    torch.cuda.set_device(0)
    # first call: 300 blocks (more than the SM count); second call: other data, fewer blocks, straight after, so the
    # scratch array still holds the first call's values; third: an input pointer that is only 2-byte aligned
    a1 = synth.audio_stream([synth.Ping(start=9000, nframes=6, f0=1490.0, snr_db=5.0, text="CQ K1ABC FN42")], HOP * 301, seed=11)
    a2 = synth.audio_stream([synth.Ping(start=4000, nframes=4, f0=1530.0, snr_db=2.0, text="K1ABC W9XYZ R-03")], HOP * 41 + 777,
                            noise_rms=300.0, seed=12)
    a3 = a1[1000:1000 + HOP * 9]
    out1 = _device_run(product, torch, a1, 1)
    # (the first launch of a kernel in a process may make the driver load its code; from the second call on, free
    # device memory must not drop across the call: the buffers are this test's, the scratch is in the library image)
    out2 = _device_run(product, torch, a2, 1)
    assert _device_run.freed >= 0
    out3 = _device_run(product, torch, a3, 1, byte_offset=2)
    assert _device_run.freed >= 0
    with product.Decoder(method=2) as dec:
        for (blocks, powers), raw in ((out1, a1), (out2, a2), (out3, a3)):
            ref_blocks, ref_powers = dec.frontend(raw)
            assert blocks.shape[0] == ref_blocks.shape[0] > 0
            # same kernels, same FP32 order of operations: the tolerance is zero
            assert _same_bits(blocks, ref_blocks)
            assert _same_bits(powers, ref_powers)


@pytest.mark.gpu
def test_noalloc_frontend_iq_and_largest_call(product, synth):
    import torch
    torch.cuda.set_device(0)
    iq = synth.iq_stream([synth.Ping(start=3000, nframes=6, f0=20.0, snr_db=4.0, text="CQ K1ABC FN42")], HOP * 7, seed=5)
    blocks, powers = _device_run(product, torch, iq, 2)
    with product.Decoder(read_mode=2, center=0.0) as dec:
        ref_blocks, ref_powers = dec.frontend(iq)
    assert _same_bits(blocks, ref_blocks) and _same_bits(powers, ref_powers)
    # the largest call the scratch array allows; a block depends only on its own 5184 samples, so the last two
    # blocks are checked against a two-block call of the allocating entry point
    nb = product.NOALLOC_MAX_BLOCKS
    rng = np.random.default_rng(77)
    big = np.clip(np.rint(900.0 * rng.standard_normal(HOP * (nb + 1))), -32768, 32767).astype(np.int16)
    blocks, powers = _device_run(product, torch, big, 1)
    assert blocks.shape[0] == nb and np.isfinite(powers).all()
    with product.Decoder(method=2) as dec:
        ref_blocks, ref_powers = dec.frontend(big[HOP * (nb - 2):])
    assert _same_bits(blocks[-2:], ref_blocks) and _same_bits(powers[-2:], ref_powers)
    # short input: no block, nothing launched, nothing written
    assert _device_run(product, torch, big[:N - 1], 1)[0].shape[0] == 0


def test_noalloc_frontend_refuses_what_its_scratch_cannot_hold(product):
    """Argument checks come before any CUDA call, so this runs without a GPU."""
    nb = product.NOALLOC_MAX_BLOCKS
    fake = 0x1000  # never dereferenced: the size check refuses the call first
    with pytest.raises(product.Msk144Error, match=r"8193 blocks.*holds 8192"):
        product.frontend_device_noalloc(fake, HOP * (nb + 2), 1, fake, fake)
    with pytest.raises(product.Msk144Error, match="null device pointer"):
        product.frontend_device_noalloc(0, N, 1, fake, fake)
    with pytest.raises(product.Msk144Error, match="read_mode"):
        product.frontend_device_noalloc(fake, N, 3, fake, fake)
    with pytest.raises(product.Msk144Error, match="misaligned"):
        product.frontend_device_noalloc(fake, N, 1, fake + 4, fake)
