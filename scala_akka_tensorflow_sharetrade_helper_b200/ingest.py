"""Price ingest: the reference's bundled `price, yyyy-MM-dd` text format, parsed by the library
(qp_parse_prices) exactly as SharePriceGetter.queryData does (SharePriceGetter.scala:89-101, :104-108),
then sorted by date and narrowed to fp32 (TrainerChildActor.scala:66).  Host-only: needs no GPU."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi


def parse_prices(text) -> np.ndarray:
    raw = text.encode() if isinstance(text, str) else bytes(text)
    lib = _capi.lib()
    n = C.c_int64()
    st = lib.qp_parse_prices(raw, len(raw), None, 0, C.byref(n))
    if st != _capi.QP_OK:
        raise RuntimeError(_capi.last_error())
    out = np.empty(n.value, np.float32)
    st = lib.qp_parse_prices(raw, len(raw), out.ctypes.data_as(C.c_void_p), n.value, C.byref(n))
    if st != _capi.QP_OK:
        raise RuntimeError(_capi.last_error())
    return out


def load_price_file(path: str) -> np.ndarray:
    with open(path, "rb") as f:
        return parse_prices(f.read())
