"""grackle_b200/csrc/fastmath.cuh (the solve kernel's lean log / exp / exp10 / divide) compiled
for the host and checked against libm in long double over the argument ranges of the path.
The device build uses the same source with MUFU.RCP64H as the reciprocal seed."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fastmath_matches_libm(tmp_path):
    exe = str(tmp_path / "fastmath_host")
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-o", exe,
                    os.path.join(ROOT, "tests", "fastmath_host.cpp")], check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    rows = {ln.split()[0]: ln.split()[1:] for ln in out.strip().splitlines()}
    assert rows["special"] == ["ok"], out          # NaN / Inf / 0 / negative / out-of-range -> libm result
    # max error in ulp of the exact result
    # (div: the short-dependency-chain form a r0 (1 + e)(1 + e^2) carries three roundings, <= 1.6 ulp)
    limits = {"log_wide": 3.0, "log_near1": 3.2, "log10": 4.0, "exp": 1.5, "exp_small": 1.5,
              "exp10": 2.0, "exp10_small": 2.0, "rcp": 1.0, "div": 1.6}
    for name, lim in limits.items():
        assert float(rows[name][0]) <= lim, (name, rows[name])
    # table-driven versions (the ones the kernels call): log in absolute terms near x = 1
    assert float(rows["tlog_wide"][0]) <= 4e-16 and float(rows["tlog_near1"][0]) <= 4e-16
    for name in ("tlog_far", "texp", "texp_small", "texp10"):
        assert float(rows[name][0]) <= 2.0, (name, rows[name])
    assert float(rows["tpow"][1]) <= 3e-14, rows["tpow"]
    # x**y = exp(y ln x) amplifies the log error by |y ln x| (here <= 38): relative bound instead
    assert float(rows["pow"][1]) <= 2e-14, rows["pow"]
