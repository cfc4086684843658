"""Host-side pieces of bench.py that need no GPU: the clock-sample summary and the workload arithmetic."""
import datetime
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _line(ts, idx, sm, mx, power, flags):
    t = datetime.datetime.fromtimestamp(ts).strftime("%Y/%m/%d %H:%M:%S.%f")[:-3]
    return f"{t}, {idx}, {sm}, {mx}, {power}, " + ", ".join(flags)


def test_clock_summary_keeps_only_the_timed_region():
    b = _bench()
    na = ["Not Active"] * 4
    t0 = 1_700_000_000.0
    lines = [
        _line(t0 - 1.0, 0, 345, 1965, 80.0, na),                                   # idle before: outside
        _line(t0 + 0.1, 0, 1965, 1965, 700.0, na),
        _line(t0 + 0.1, 1, 1950, 1965, 690.0, ["Not Active", "Not Active", "Not Active", "Active"]),
        _line(t0 + 0.2, 0, 1965, 1965, 700.0, na),
        _line(t0 + 5.0, 0, 345, 1965, 80.0, ["Active", "Not Active", "Not Active", "Not Active"]),   # after: outside
        "garbage line",
    ]
    c = b.ClockSampler.summarise(lines, t0, t0 + 1.0, 2)
    assert c["samples"] == 3 and c["gpus"] == 2 and c["window"] == "timed region"
    assert c["sm_mhz"] == 1965 and c["sm_min_mhz"] == 1950 and c["sm_max_mhz"] == 1965
    assert c["reasons"] == ["sw_power_cap"]                 # the hw_slowdown sample lies outside the region


def test_clock_summary_falls_back_and_survives_no_samples():
    b = _bench()
    na = ["Not Active"] * 4
    t0 = 1_700_000_000.0
    c = b.ClockSampler.summarise([_line(t0 - 0.5, 0, 1800, 1965, 500.0, na)], t0, t0 + 0.01, 1)
    assert c["samples"] == 1 and c["sm_mhz"] == 1800 and c["window"].startswith("nearest samples")
    c = b.ClockSampler.summarise([], t0, t0 + 1.0, 1)
    assert c == dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)


def test_workload_arithmetic_matches_the_survey():
    b = _bench()
    assert b.D == 203 and b.P == 41200 and b.FLOP_PER_ENV_STEP == 246800
    assert abs(b.bytes_per_env_step(4096) - 1089.40625) < 1e-9
    assert b.SCALING == "weak"
