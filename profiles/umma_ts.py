"""A operand from tensor memory (tcgen05.mma, [d], [a_tmem], b_desc): D = A · B^T with A = [128 rows][K] in TMEM columns."""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
src = open(os.path.join(HERE, "umma_probe.py")).read().split("# ---- T0")[0].replace("__file__", repr(os.path.join(HERE, "umma_probe.py")))
exec(src)
N, K = 104, 32
A = ints((128, K)); B = ints((N, K))
for a_col in (0, 300):
    out = np.empty((128, 256), np.float32)
    a_img = np.ascontiguousarray(A).tobytes()
    b_img = b_kmajor_sw128(B)
    st = lib.qp_debug_umma(0, a_img, len(a_img), b_img, len(b_img), C.c_uint64(0), C.c_uint64(desc()), idesc(128, N), 4, 8, 2,
                           128, 0, 256, C.c_float(-5.0), out.ctypes.data_as(C.c_void_p), K, a_col)
    assert st == 0, _capi.last_error()
    report(f"A from TMEM (columns {a_col}..), M=128 N=104 K=32, D at column 128", out[:, 128:128 + N], A @ B.T)
