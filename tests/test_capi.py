"""CPU tests of the drop-in boundary: the shared library loads, exports every symbol that
include/qpolicy.h declares, the ctypes stub covers each of them, and — without a GPU — the product
path fails loudly instead of falling back to anything."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "qpolicy.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(qp_[a-z0-9_]+)\s*\(", src)))


def test_header_compiles_as_plain_c(tmp_path):
    c = tmp_path / "t.c"
    c.write_text('#include "qpolicy.h"\nint main(void){ qp_config c; (void)c; return QP_OK; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(c),
                    "-o", str(tmp_path / "t.o")], check=True)


def test_library_exports_every_declared_symbol(built_lib):
    from scala_akka_tensorflow_sharetrade_helper_b200 import _capi
    lib = C.CDLL(built_lib)
    names = declared_functions()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in qpolicy.h but not exported by libqpolicy.so"
        assert n in _capi.SIGNATURES, f"{n} has no ctypes signature in _capi.SIGNATURES"
    assert set(_capi.SIGNATURES) <= set(names), "ctypes stub binds symbols the header does not declare"


def test_library_is_sm100a_and_links_no_torch(built_lib):
    out = subprocess.run(["cuobjdump", "-lelf", built_lib], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    ldd = subprocess.run(["ldd", built_lib], capture_output=True, text=True).stdout
    assert "torch" not in ldd and "libcudart" not in ldd  # static cudart, no framework dependency


def test_config_default_matches_reference_constants(built_lib):
    from scala_akka_tensorflow_sharetrade_helper_b200 import _capi
    cfg = _capi.QpConfig()
    assert _capi.lib().qp_config_default(C.byref(cfg)) == 0
    # QDecisionPolicyActor.scala:17-22, :50
    assert (cfg.window, cfg.hidden) == (201, 200)
    assert abs(cfg.epsilon - 0.9) < 1e-7 and abs(cfg.gamma - 0.001) < 1e-9 and abs(cfg.lr - 0.01) < 1e-9
    assert abs(cfg.bias - 0.1) < 1e-8 and cfg.step_div == 1000.0 and cfg.bookkeeping == 0


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="only meaningful on a box without a GPU")
def test_no_gpu_means_loud_failure_not_fallback(built_lib):
    import scala_akka_tensorflow_sharetrade_helper_b200 as q
    with pytest.raises(q.QPolicyError, match="no CUDA device"):
        q.QDecisionPolicy()


def test_invalid_config_is_rejected_before_touching_the_device(built_lib):
    from scala_akka_tensorflow_sharetrade_helper_b200 import _capi
    L = _capi.lib()
    cfg = _capi.QpConfig()
    L.qp_config_default(C.byref(cfg))
    cfg.hidden = 202
    h = C.c_void_p()
    assert L.qp_create(C.byref(cfg), C.byref(h)) == _capi.QP_EINVAL
    assert b"multiple of 4" in L.qp_last_error()
    assert L.qp_create(None, C.byref(h)) == _capi.QP_EINVAL


def test_every_option_key_is_documented_in_the_header():
    """qp_set_option / qp_get_option keys (strcmp(key, "...") in qp_api.cu) against the option list of include/qpolicy.h."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = open(os.path.join(root, "scala_akka_tensorflow_sharetrade_helper_b200", "csrc", "qp_api.cu")).read()
    hdr = open(os.path.join(root, "include", "qpolicy.h")).read()
    keys = set(re.findall(r'strcmp\(key, "([a-z0-9_]+)"\)', src))
    assert len(keys) >= 20
    missing = sorted(k for k in keys if f'"{k}"' not in hdr)
    assert not missing, f"options not documented in include/qpolicy.h: {missing}"
