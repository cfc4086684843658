"""CPU tests of the price ingest: qp_parse_prices (C ABI, host-only) against the
oracle's restatement of SharePriceGetter.queryData, on synthetic text with the malformed rows the
reference's parser drops, and on the reference's bundled file when it is present."""
import os

import numpy as np
import pytest

import scala_akka_tensorflow_sharetrade_helper_b200 as q
from oracle import ingest_oracle

MSFT = "/root/reference/src/main/resources/MSFT-stock-prices-revised.txt"

CASES = [
    "56.080002, 1992-07-22\n55.98, 1992-07-23\n56.150002, 1992-07-24\n",
    "2.5, 2001-01-03\n1.5, 2001-01-01\n",                              # sorted by date, not by file order
    "1.0, 2001-01-01\n9.0, 2001-01-01\n",                              # repeated date: last row wins (HashMap)
    "1.0,2001-01-01,ignored,more\n",                                   # Seq(price, date, _*)
    "only-one-field\n\n,\n1.0\n",                                      # fewer than two fields
    "abc, 2001-01-01\n1.0, 2001-1-01\n1.0, 01-01-2001\n1.0, 2001-13-01\n1.0, 2001-00-10\n1.0, 2001-01-32\n",
    "  7.25  ,\t2001-05-05  \r\n8.5d, 2001-05-06\r\n1e2, 2001-05-07\n.5, 2001-05-08\n-3., 2001-05-09\n",
    "4.5, 2001-02-31\n4.0, 2001-02-28\n1.0, 2000-02-30\n",           # SMART resolver clamps to month end
    "0x10, 2001-01-01\n1,5, 2001-01-01\n1.0e, 2001-01-02\n",
    "",
]


@pytest.mark.parametrize("text", CASES)
def test_parse_matches_restatement(text):
    got = q.parse_prices(text)
    want = ingest_oracle.parse_prices(text)
    assert got.dtype == np.float32
    np.testing.assert_array_equal(got, want)


def test_known_rows():
    np.testing.assert_array_equal(q.parse_prices(CASES[1]), np.float32([1.5, 2.5]))
    np.testing.assert_array_equal(q.parse_prices(CASES[2]), np.float32([9.0]))
    np.testing.assert_array_equal(q.parse_prices(CASES[4]), np.float32([]))
    np.testing.assert_array_equal(q.parse_prices(CASES[7]), np.float32([1.0, 4.0]))   # 2001-02-31 -> 02-28, overwritten by the later 4.0


@pytest.mark.skipif(not os.path.exists(MSFT), reason="reference data file only exists in the build container")
def test_reference_price_file():
    """6 047 rows, 1992-07-22 .. 2011-11-14, 15.2 .. 178.9375 -> 5 846 env-steps."""
    p = q.load_price_file(MSFT)
    assert p.size == 6047 and p[0] == np.float32(56.080002)
    assert float(p.min()) == np.float32(15.2) and float(p.max()) == 178.9375
    np.testing.assert_array_equal(p, ingest_oracle.parse_prices(open(MSFT).read()))
    assert p.size - 201 == 5846        # TrainerChildActor.scala:67
