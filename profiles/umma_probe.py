"""Exploration of tcgen05.mma operand layouts through qp_debug_umma (csrc/qp_probe.cuh): prints, for each layout
hypothesis the cluster kernel would rest on, whether the tensor core's result equals the numpy product.
Run on a B200:  python profiles/umma_probe.py
"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scala_akka_tensorflow_sharetrade_helper_b200 import _capi  # noqa: E402

lib = _capi.lib()


def idesc(M, N, a_mn=0, b_mn=0):
    return (1 << 4) | (2 << 7) | (2 << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)


def desc(start16=0, lbo=16, sbo=1024, layout=2):
    return (start16 & 0x3FFF) | ((lbo >> 4) << 16) | ((sbo >> 4) << 32) | (1 << 46) | (layout << 61)


def b_kmajor_sw128(B):
    """B [N, 32] fp32 -> K-major SWIZZLE_128B tile image (row r: 128 bytes, 16-byte chunk j at j ^ (r & 7))."""
    N = B.shape[0]
    img = np.zeros((N, 8, 4), np.float32)
    for r in range(N):
        for j in range(8):
            img[r, j ^ (r & 7)] = B[r, 4 * j:4 * j + 4]
    return img.tobytes()


def a_mn_noswizzle(A, rows, quad_pitch, k0=0):
    """A [M, K] fp32 (M = episodes) -> image [quad][row][4]: element (m, k) at quad (m/4), row k0 + k."""
    M, K = A.shape
    img = np.full((M // 4, quad_pitch // 16, 4), 7777.0, np.float32)
    for m in range(M):
        img[m // 4, k0:k0 + K, m % 4] = A[m]
    return img.tobytes()


def run(a_img, b_img, da, db, ide, n_mma, a_step, b_step, n_cols, d_col=0, d_lane=0, fill=-5.0):
    out = np.empty((128, n_cols), np.float32)
    st = lib.qp_debug_umma(0, a_img, len(a_img), b_img, len(b_img), C.c_uint64(da), C.c_uint64(db), ide, n_mma, a_step, b_step,
                           d_col, d_lane, n_cols, C.c_float(fill), out.ctypes.data_as(C.c_void_p), 0, 0)
    if st != 0:
        raise RuntimeError(_capi.last_error())
    return out


def report(name, D, want, rows=None):
    rows = range(want.shape[0]) if rows is None else rows
    got = D[list(rows), :want.shape[1]]
    ok = np.array_equal(got, want)
    print(f"{name}: {'OK' if ok else 'MISMATCH'}", flush=True)
    if not ok:
        bad = np.argwhere(got != want)
        print("   first mismatches:", [(int(i), int(j), float(got[i, j]), float(want[i, j])) for i, j in bad[:6]], " n_bad", len(bad))
    return ok


rng = np.random.default_rng(0)


def ints(shape, lo=-8, hi=9):
    return rng.integers(lo, hi, shape).astype(np.float32)


# ---- T0: sanity, the proven form: A and B K-major SWIZZLE_128B, M = 128, N = 112
N = 112
A = ints((128, 32)); B = ints((N, 32))
D = run(b_kmajor_sw128(A), b_kmajor_sw128(B), desc(), desc(), idesc(128, N), 4, 2, 2, 128)
report("T0  K-major SW128 A and B, M=128 N=112", D, A @ B.T)

# ---- T1/T2: A MN-major, no swizzle, [quad][row][4]; start row offset k0; which of LBO / SBO is the quad pitch
QP = 4096
for M in (128, 64):
    A = ints((M, 32)); B = ints((N, 32))
    want = A @ B.T
    for k0 in (0, 5):
        a_img = a_mn_noswizzle(A, 32, QP, k0)
        for nm, lbo, sbo in (("LBO=128,SBO=quad", 128, QP), ("LBO=quad,SBO=128", QP, 128)):
            D = run(a_img, b_kmajor_sw128(B), desc(k0, lbo, sbo, 0), desc(), idesc(M, N, a_mn=1), 4, 8, 2, 128)
            if M == 128:
                report(f"T1  A MN-major no-swizzle M=128 k0={k0} {nm}", D, want)
            else:
                # where do the 64 rows land?
                lanes = []
                for m in range(64):
                    hit = [l for l in range(128) if np.array_equal(D[l, :N], want[m])]
                    lanes.append(hit[0] if len(hit) == 1 else -1)
                print(f"T2  A MN-major no-swizzle M=64 k0={k0} {nm}: row->lane {lanes[:20]}... complete={all(x >= 0 for x in lanes)}")
                untouched = [l for l in range(128) if (D[l, :N] == -5.0).all()]
                print("     untouched lanes:", untouched[:8], "...", len(untouched))

# ---- T3: M=64 accumulator at lane offset 16
A = ints((64, 32)); B = ints((N, 32))
want = A @ B.T
try:
    D = run(a_mn_noswizzle(A, 32, QP, 0), b_kmajor_sw128(B), desc(0, 128, QP, 0), desc(), idesc(64, N, a_mn=1), 4, 8, 2, 128, d_lane=16)
    lanes = []
    for m in range(64):
        hit = [l for l in range(128) if np.array_equal(D[l, :N], want[m])]
        lanes.append(hit[0] if len(hit) == 1 else -1)
    print("T3  M=64 with lane offset 16: row->lane", lanes[:20], "... complete", all(x >= 0 for x in lanes))
except RuntimeError as e:
    print("T3  M=64 lane offset 16 failed:", e)

# ---- T4: A K-major, no swizzle: element (m, kk) at m*16 + (kk/4)*LBO + (kk%4)*4 (the SAME [quad][row][4] image read
# with M = rows, K = episodes), SBO = 128, start row offset r0, M = 128, K = 32 episodes (8 quads)
R = 200
X = ints((32, R))                  # X[e][row]
img = np.zeros((8, QP // 16, 4), np.float32)
for e in range(32):
    img[e // 4, :R, e % 4] = X[e]
Bz = ints((N, 32))                 # dZ1[n][e]
for r0 in (0, 3):
    want = X[:, r0:r0 + 128].T @ Bz.T          # D[m][n] = sum_e X[e][r0+m] * Bz[n][e]
    for nm, lbo, sbo in (("LBO=quad,SBO=128", QP, 128), ("LBO=128,SBO=quad", 128, QP)):
        D = run(img.tobytes(), b_kmajor_sw128(Bz), desc(r0, lbo, sbo, 0), desc(), idesc(128, N), 4, 2 * QP // 16, 2, 128)
        report(f"T4  A K-major no-swizzle (ring as backward operand) r0={r0} {nm}", D, want)

# ---- T6: precision of kind::tf32: are the low 13 mantissa bits of an operand truncated or rounded?
A = np.zeros((128, 32), np.float32); B = np.zeros((N, 32), np.float32)
A[:, 0] = 1.0
vals = np.float32(1.0) + np.arange(N, dtype=np.float32) * np.float32(2.0 ** -14)      # steps of 1/8 tf32 ulp
B[:, 0] = vals
D = run(b_kmajor_sw128(A), b_kmajor_sw128(B), desc(), desc(), idesc(128, N), 1, 0, 0, 128)
print("T6  tf32 operand handling, B = 1 + j*2^-14 -> D[0][j] - 1 in tf32 ulps (2^-10):", ((D[0, :24] - 1.0) * 1024).tolist())

# ---- T5: N = 104 with M = 128 and with M = 64
for M in (64, 128):      # (a rejected shape may kill the context: last)
    A = ints((M, 32)); B = ints((104, 32))
    try:
        D = run(b_kmajor_sw128(A), b_kmajor_sw128(B), desc(), desc(), idesc(M, 104), 4, 2, 2, 128)
        if M == 128:
            report("T5  N=104 M=128", D, A @ B.T)
        else:
            want = A @ B.T
            rows = [l for m in range(64) for l in range(128) if np.array_equal(D[l, :104], want[m])]
            print("T5  N=104 M=64: rows found", len(rows))
    except RuntimeError as e:
        print(f"T5  N=104 M={M} failed:", e)

