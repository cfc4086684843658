"""The tcgen05.mma (kind::tf32) operand-layout facts the persistent kernels rest on, each pinned against a numpy product
through qp_debug_umma (csrc/qp_probe.cuh): one CTA copies a byte image into shared memory, issues the MMAs with the
given matrix / instruction descriptors and hands back all 128 lanes of the accumulator.  Small integers, so every
product and sum is exact in tf32 and the comparison is bit-exact.

  * K-major SWIZZLE_128B A and B at M = 128 (the form every kernel here uses);
  * N = 104 — a multiple of 8, not of 16 — is a legal N at M = 128 (the pair kernel's N per CTA for 200 hidden units is
    104 with the padding column block);
  * the A operand may start at ANY 128-byte row of a SWIZZLE_128B buffer: the swizzle comes from the absolute address
    bits, so the backward contraction reads the price-window buffer from the row of window element 0 without a copy;
  * an M = 64 accumulator lies in lanes 0-15, 32-47, 64-79, 96-111;
  * A from tensor memory (the .ts form) gives the same product.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def idesc(M, N):
    # f32 accumulator, tf32 A and B, both K-major, N >> 3 at bit 17, M >> 4 at bit 24
    return (1 << 4) | (2 << 7) | (2 << 10) | ((N >> 3) << 17) | ((M >> 4) << 24)


def desc(start16=0, lbo=16, sbo=1024, layout=2):
    return (start16 & 0x3FFF) | ((lbo >> 4) << 16) | ((sbo >> 4) << 32) | (1 << 46) | (layout << 61)


def kmajor_sw128(X, rows=None, row0=0):
    """X [R, 32] fp32 -> image of `rows` 128-byte rows, X's row r at image row row0 + r, 16-byte chunk j at j ^ (row & 7)."""
    R = X.shape[0]
    rows = R if rows is None else rows
    img = np.full((rows, 8, 4), 7777.0, np.float32)
    for r in range(R):
        for j in range(8):
            img[row0 + r, j ^ ((row0 + r) & 7)] = X[r, 4 * j:4 * j + 4]
    return img.tobytes()


def run(a_img, b_img, da, db, ide, n_mma, a_step, b_step, n_cols, a_tmem_k=0, a_col=0, fill=-5.0):
    from scala_akka_tensorflow_sharetrade_helper_b200 import _capi
    out = np.empty((128, n_cols), np.float32)
    st = _capi.lib().qp_debug_umma(0, a_img, len(a_img), b_img, len(b_img), C.c_uint64(da), C.c_uint64(db), ide, n_mma, a_step, b_step,
                                   0, 0, n_cols, C.c_float(fill), out.ctypes.data_as(C.c_void_p), a_tmem_k, a_col)
    assert st == 0, _capi.last_error()
    return out


def ints(rng, shape):
    return rng.integers(-8, 9, shape).astype(np.float32)


@pytest.mark.parametrize("N", [112, 104, 56, 208])
def test_kmajor_sw128_m128(N):
    rng = np.random.default_rng(N)
    A, B = ints(rng, (128, 32)), ints(rng, (N, 32))
    n_cols = (N + 15) // 16 * 16
    D = run(kmajor_sw128(A), kmajor_sw128(B), desc(), desc(), idesc(128, N), 4, 2, 2, n_cols)
    assert np.array_equal(D[:, :N], A @ B.T)
    assert np.all(D[:, N:] == -5.0), "columns past N must not be written"


@pytest.mark.parametrize("row0", [1, 3, 8, 13, 77])
def test_kmajor_sw128_a_operand_may_start_at_any_row(row0):
    # the window buffer: 256 rows of 128 bytes, the operand = 128 rows from row0 on (descriptor start = row0 * 128 bytes)
    rng = np.random.default_rng(row0)
    A, B = ints(rng, (128, 32)), ints(rng, (104, 32))
    D = run(kmajor_sw128(A, rows=256, row0=row0), kmajor_sw128(B), desc(start16=row0 * 8), desc(), idesc(128, 104), 4, 2, 2, 112)
    assert np.array_equal(D[:, :104], A @ B.T)


def test_m64_accumulator_lanes():
    rng = np.random.default_rng(64)
    A, B = ints(rng, (64, 32)), ints(rng, (104, 32))
    D = run(kmajor_sw128(A), kmajor_sw128(B), desc(), desc(), idesc(64, 104), 4, 2, 2, 112)
    want = A @ B.T
    lanes = [16 * (m // 16) * 2 + m % 16 for m in range(64)]          # 0-15, 32-47, 64-79, 96-111
    assert np.array_equal(D[lanes, :104], want)
    others = [l for l in range(128) if l not in lanes]
    assert np.all(D[others, :104] == -5.0), "an M = 64 MMA must leave the other lanes alone"


def test_accumulate_over_two_k_blocks():
    # two 32-wide K blocks, each its own SW128 tile 16 KB (A) / N * 128 B (B) further on
    rng = np.random.default_rng(2)
    N = 104
    A, B = ints(rng, (128, 64)), ints(rng, (N, 64))
    a_img = kmajor_sw128(A[:, :32]) + kmajor_sw128(A[:, 32:])
    b_img = kmajor_sw128(B[:, :32], rows=128) + kmajor_sw128(B[:, 32:], rows=128)
    D0 = run(a_img, b_img, desc(), desc(), idesc(128, N), 4, 2, 2, 112)
    assert np.array_equal(D0[:, :N], A[:, :32] @ B[:, :32].T)
    # eight MMAs do not fit one (a_step, b_step) pair across the tile boundary: second block through the start fields
    D1 = run(a_img, b_img, desc(start16=16384 >> 4), desc(start16=16384 >> 4), idesc(128, N), 4, 2, 2, 112)
    assert np.array_equal(D1[:, :N], A[:, 32:] @ B[:, 32:].T)


def test_a_from_tensor_memory():
    rng = np.random.default_rng(5)
    N = 104
    A, B = ints(rng, (128, 32)), ints(rng, (N, 32))
    D = run(A.tobytes(), kmajor_sw128(B), 0, desc(), idesc(128, N), 4, 8, 2, 112, a_tmem_k=32, a_col=256)
    assert np.array_equal(D[:, :N], A @ B.T)
