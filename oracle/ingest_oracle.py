"""CPU restatement of the reference's price ingest — TEST INFRASTRUCTURE ONLY (see qp_oracle.c).

Follows SharePriceGetter.queryData (src/main/scala/SharePriceGetter.scala:89-101) and its parsers
(:104-108), then TrainerChildActor.train's ordering and narrowing (TrainerChildActor.scala:66):
    for each line: split(",").map(_.trim) ; Seq(price, date, _*) -> (parseDate(date), parseDouble(price))
    rows where either fails are dropped; HashMap(parsed: _*) keeps the LAST row of a repeated date;
    sharePrices.toIndexedSeq sorted by date, .toFloat.
"""
from __future__ import annotations

import datetime
import re

import numpy as np

_DATE = re.compile(r"^(\d{4})-(\d{2})-(\d{2})$")
_NUM = re.compile(r"^[+-]?(NaN|Infinity|((\d+\.?\d*|\.\d+)([eE][+-]?\d+)?))[fFdD]?$")


def _parse_date(s):
    m = _DATE.match(s)
    if not m:
        return None
    y, mo, d = (int(g) for g in m.groups())
    if y < 1 or not 1 <= mo <= 12 or not 1 <= d <= 31:
        return None
    # ResolverStyle.SMART: a day-of-month of 29..31 past the month's end resolves to its last day
    last = (datetime.date(y + (mo == 12), mo % 12 + 1, 1) - datetime.timedelta(days=1)).day
    return datetime.date(y, mo, min(d, last))


def _parse_double(s):
    if not _NUM.match(s):
        return None
    t = s.rstrip("fFdD") if not s.endswith(("NaN", "Infinity")) else s
    return float(t.replace("Infinity", "inf").replace("NaN", "nan"))


def parse_prices(text: str) -> np.ndarray:
    by_date = {}
    for line in text.split("\n"):
        line = line.rstrip("\r")
        fields = [f.strip(" \t\r\n\x0b\x0c") for f in line.split(",")]
        while fields and fields[-1] == "":
            fields.pop()
        if len(fields) < 2:
            continue
        date, price = _parse_date(fields[1]), _parse_double(fields[0])
        if date is None or price is None:
            continue
        by_date[date] = price
    return np.asarray([by_date[k] for k in sorted(by_date)], np.float64).astype(np.float32)
