#!/usr/bin/env python
"""One calculate_* call per kind on device-resident fields (config C5 = C4 parameters),
for ncu captures and quick timing.   python tools/calc_probe.py [nk] [what ...]"""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from grackle_b200 import capi, synth_fields as sf, synth_tables  # noqa: E402

nk = int(sys.argv[1]) if len(sys.argv) > 1 else 16
kinds = sys.argv[2:] or ["pressure", "temperature", "gamma", "cooling_time"]
dims = [1024, 1024, nk]
n = dims[0] * dims[1] * dims[2]
lib = capi.load_product()
dev = torch.device("cuda", 0)
params, dt, _ = sf.config("C4", synth_tables.ensure())
ch = capi.Chemistry(lib, **params).initialize()
idx = torch.arange(n, dtype=torch.int64, device=dev)
fields = {k: v.contiguous() for k, v in sf.make_fields(idx, 3, params, ch.temperature_units, "B").items()}
g = capi.Grid(dims)
for k, v in fields.items():
    g.set(k, int(v.data_ptr()))
out = torch.empty(n, dtype=torch.float64, device=dev)
for what in kinds:
    ms = []
    for it in range(4):
        assert capi.calculate(ch, g, what, int(out.data_ptr())) == 1
        ms.append(lib.grb200_last_kernel_ms(C.byref(ch.rates)))
    t = min(ms[1:]) * 1e-3
    print("%-13s %.3f ms  %.2f Gcell/s" % (what, t * 1e3, n / t / 1e9))
