"""Bit-exactness obligations of the LDPC kernel's packed math (kernels.cuh: tanh2, platanh2).

CPU: the three-instruction constant division used by platanh2 equals the IEEE quotient of the reference's
platanh (ldpc_kernel.cuh:65-93) for every float of every segment at or above the kernel's LDPC_TINY guard
(exhaustive, tests/ldpc_divproof.c).
GPU: all 2^32 float inputs through the f32x2 tanh and platanh against libdevice tanhf and the reference-form
platanh with __fdiv_rn, on the device itself (msk144_selftest_ldpc_math)."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def kernel_tiny():
    src = open(os.path.join(ROOT, "msk144cudecoder_b200", "csrc", "kernels.cuh")).read()
    return float(re.search(r"LDPC_TINY\s*=\s*([0-9.e+-]+)f", src).group(1))


def test_constant_division_is_ieee_exact(tmp_path):
    exe = str(tmp_path / "divproof")
    subprocess.run(["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-o", exe,
                    os.path.join(ROOT, "tests", "ldpc_divproof.c"), "-lm"], check=True)
    r = subprocess.run([exe, repr(kernel_tiny())], capture_output=True, text=True, timeout=600)
    print(r.stdout)
    assert r.returncode == 0, r.stdout
    # segments 1..3 (z > 0.664) must be exact everywhere
    for k in (1, 2, 3):
        assert re.search(rf"segment {k} .*mismatches=0 ", r.stdout), r.stdout


@pytest.mark.gpu
def test_packed_math_matches_scalar_on_device(product):
    with product.Decoder() as d:
        bad_tanh, bad_platanh = d.selftest_ldpc_math()
    assert bad_tanh == 0
    assert bad_platanh == 0
