#!/usr/bin/env python3
"""Derive the LDPC(128,90) parity-check structure as a check->bits table (the WSJT-X "Nm" form)
from the bit->(slot,check) map the reference ships (ldpc_context.cuh:10-139).

Dev-time tool: /root/reference is only read here, never at test/bench time.  Prints a C table
`{bit indices of check c, ascending, -1 padded to 11}` for c = 0..37, plus invariants.
"""
import re, sys
import numpy as np

src = open("/root/reference/src/ldpc_context.cuh").read()
body = src[src.index("ldpc_reverse_map[128][3][2]"):src.index("};")]
nums = [int(x) for x in re.findall(r"-?\d+", body)][3:]  # skip the 128,3,2 dims
assert len(nums) == 128 * 6, len(nums)
mp = np.array(nums).reshape(128, 3, 2)   # [bit][k] = (slot, check)

rows = [[-1] * 11 for _ in range(38)]
for b in range(128):
    for k in range(3):
        slot, chk = mp[b, k]
        assert rows[chk][slot] == -1
        rows[chk][slot] = b
H = np.zeros((38, 128), np.uint8)
for c, r in enumerate(rows):
    bits = [x for x in r if x >= 0]
    assert bits == sorted(bits), "bits ascend with slot"
    assert r[:len(bits)] == bits, "slots contiguous"
    H[c, bits] = 1
assert (H.sum(0) == 3).all()
full = [c for c in range(38) if rows[c][10] >= 0]
assert full == [2, 4, 5, 26], full
# k-order of each bit's edges must be recoverable: record it
print("// check -> member bits (ascending == slot order); -1 = unused 11th slot")
for c, r in enumerate(rows):
    print("    {" + ", ".join(f"{x:3d}" for x in r) + "},")
print("// bit -> its three checks in the reference's edge order k=0,1,2")
for b in range(128):
    print("    {" + ", ".join(f"{mp[b,k,1]:2d}" for k in range(3)) + "},")
