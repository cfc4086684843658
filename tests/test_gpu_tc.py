"""The tcgen05 tensor-core GEMM of the GEMM-organised path on its own (qp_debug_tc_gemm), against
numpy in fp64.  Stated tolerances: 3xTF32 (hi/lo operand split, three MMAs per k-block) must be
fp32-grade, |dC| <= 2e-6 * sum_k |a||b|; plain TF32 gets its own, looser bound 2e-3 * sum_k |a||b|."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def tc_gemm(A, B, passes, splits=1):
    from scala_akka_tensorflow_sharetrade_helper_b200 import _capi
    lib = _capi.lib()
    M, K = A.shape
    N = B.shape[0]
    out = np.full((M, N), np.nan, np.float32)
    ms = C.c_float()
    st = lib.qp_debug_tc_gemm(0, M, N, K, A.ctypes.data_as(C.c_void_p), B.ctypes.data_as(C.c_void_p),
                              out.ctypes.data_as(C.c_void_p), passes, splits, C.byref(ms))
    assert st == 0, _capi.last_error()
    return out, ms.value


@pytest.mark.parametrize("M,N,K,splits", [(128, 128, 32, 1), (128, 128, 128, 1), (300, 260, 96, 1), (1024, 512, 512, 2),
                                          (256, 384, 1152, 3)])
def test_tc_gemm_3xtf32_is_fp32_grade(M, N, K, splits):
    rng = np.random.default_rng(M + N + K)
    A = (rng.standard_normal((M, K)) * rng.uniform(0.1, 100, (M, 1))).astype(np.float32)
    B = rng.standard_normal((N, K)).astype(np.float32)
    ref = A.astype(np.float64) @ B.astype(np.float64).T
    bound = np.abs(A).astype(np.float64) @ np.abs(B).astype(np.float64).T
    got, _ = tc_gemm(A, B, 3, splits)
    assert np.isfinite(got).all()
    assert (np.abs(got - ref) <= 2e-6 * bound + 1e-30).all(), float((np.abs(got - ref) / bound).max())
    got1, _ = tc_gemm(A, B, 1, splits)
    err1 = np.abs(got1 - ref) / bound
    assert err1.max() <= 2e-3
    assert err1.max() > 1e-5, "plain TF32 should be visibly less accurate than 3xTF32"


def test_tc_gemm_exact_on_small_integers():
    """Integers below 2^10 are exact in TF32 and their products/sums exact in fp32: bit-exact result,
    which pins the operand layout (swizzle, descriptor advance, TMEM lane mapping) precisely."""
    rng = np.random.default_rng(7)
    A = rng.integers(-8, 9, (256, 160)).astype(np.float32)
    B = rng.integers(-8, 9, (384, 160)).astype(np.float32)
    got, _ = tc_gemm(A, B, 3)
    np.testing.assert_array_equal(got, (A.astype(np.int64) @ B.astype(np.int64).T).astype(np.float32))
    got1, _ = tc_gemm(A, B, 1)
    np.testing.assert_array_equal(got1, got)
