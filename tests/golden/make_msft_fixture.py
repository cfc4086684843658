"""Turns the reference's bundled price file into a committed fixture.

    python tests/golden/make_msft_fixture.py       # in the build container, where /root/reference exists

src/main/resources/MSFT-stock-prices-revised.txt is the data SharePriceGetter.queryData reads
(SharePriceGetter.scala:83-102) and ShareTradeHelper trains on (budget 2400, 0 shares, ten children on the one
series: ShareTradeHelper.scala:20-21, TrainerRouterActor.scala:36).  /root/reference does not exist on the GPU box,
so the file's text travels as a fixture: the raw bytes (so that qp_parse_prices itself runs in the GPU test), the
prices the oracle's restatement of the parser extracts from them, and a digest of the source.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ingest_oracle  # noqa: E402

SRC = "/root/reference/src/main/resources/MSFT-stock-prices-revised.txt"
HERE = os.path.dirname(os.path.abspath(__file__))

if __name__ == "__main__":
    raw = open(SRC, "rb").read()
    prices = ingest_oracle.parse_prices(raw.decode("utf-8"))
    assert prices.size == 6047 and prices.dtype == np.float32
    np.savez_compressed(os.path.join(HERE, "msft_prices.npz"), text=np.frombuffer(raw, np.uint8), prices=prices,
                        sha256=hashlib.sha256(raw).hexdigest(), n_rows=prices.size, n_steps=prices.size - 201)
    print("msft_prices.npz", os.path.getsize(os.path.join(HERE, "msft_prices.npz")), "bytes;", prices.size, "prices")
