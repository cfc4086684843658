"""Is N = 104 (a multiple of 8, not of 16) accepted with M = 128 / M = 64 by tcgen05.mma kind::tf32?"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
src = open(os.path.join(HERE, "umma_probe.py")).read().split("# ---- T0")[0].replace("__file__", repr(os.path.join(HERE, "umma_probe.py")))
exec(src)
for M in (64, 128):
    A = ints((M, 32)); B = ints((104, 32))
    D = run(b_kmajor_sw128(A), b_kmajor_sw128(B), desc(), desc(), idesc(M, 104), 4, 2, 2, 128)
    want = A @ B.T
    if M == 128:
        report("N=104 M=128", D, want)
    else:
        lanes = [[l for l in range(128) if np.array_equal(D[l, :104], want[m])] for m in range(64)]
        print("N=104 M=64 row->lane", [x[0] if x else -1 for x in lanes])
