"""grb200_calculate_dust_temperature_noalloc (include/grackle.h): the allocation-free variant of
calculate_dust_temperature, whose gas-temperature scratch is a static device array.  It has no
counterpart in the reference, so it is checked
against the existing entry point, local_calculate_dust_temperature, which test_gpu_parity.py holds
against the oracle.  Two calls in a row (the scratch keeps the previous call's contents), the size
limit, and the host-array rejection.  Collected last on purpose (pytest -x).
"""
import ctypes as C

import numpy as np
import pytest

from grackle_b200 import capi, synth_fields as sf

pytestmark = pytest.mark.gpu
MAX_CELLS = 1 << 20          # GRB200_NOALLOC_MAX_CELLS


def _declare(lib):
    fn = lib.grb200_calculate_dust_temperature_noalloc
    fn.argtypes = [C.POINTER(capi.ChemistryData), C.POINTER(capi.ChemistryStorage), C.POINTER(capi.CodeUnits),
                   C.POINTER(capi.FieldData), C.c_void_p]
    fn.restype = C.c_int
    return fn


def _device_grid(torch, fields, dims):
    dev = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in fields.items()}
    g = capi.Grid(dims)
    for k, t in dev.items():
        g.set(k, int(t.data_ptr()))
    return g, dev


def test_noalloc_dust_temperature_matches_existing_entry_point(product_lib, synth_table):
    torch = pytest.importorskip("torch")
    assert product_lib.grb200_device_available() == 1
    noalloc = _declare(product_lib)
    params, dt, _ = sf.config("C3", synth_table)            # pc = 2, h2_on_dust: dust temperature is defined
    ch = capi.Chemistry(product_lib, **params).initialize()
    try:
        for call, (dims, seed0) in enumerate((([48, 32, 20], 0), ([31, 17, 9], 777777))):   # second call: other size, other cells
            n = int(np.prod(dims))
            idx = np.arange(seed0, seed0 + n, dtype=np.int64)
            fields = sf.make_fields(idx, 2, params, ch.temperature_units, "B", fh=ch.chem.HydrogenFractionByMass)
            g, dev = _device_grid(torch, fields, dims)
            ref = torch.full((n,), -7.0, dtype=torch.float64, device="cuda")
            out = torch.full((n,), -7.0, dtype=torch.float64, device="cuda")
            torch.cuda.synchronize()
            assert capi.calculate(ch, g, "dust_temperature", int(ref.data_ptr())) == 1
            before = product_lib.grb200_kernel_launch_count()
            rc = noalloc(C.byref(ch.chem), C.byref(ch.rates), C.byref(ch.units), C.byref(g.fd), C.c_void_p(int(out.data_ptr())))
            assert rc == 1, "call %d" % call
            assert product_lib.grb200_kernel_launch_count() - before == 2
            a, b = out.cpu().numpy(), ref.cpu().numpy()
            assert np.all(np.isfinite(a)) and np.all(a > 0)
            # same kernels, same inputs: the tolerance of the calculate_* tests (1e-9) - in fact bitwise
            assert float(np.max(np.abs(a - b) / np.abs(b))) <= 1e-9
            assert np.array_equal(a, b)
            for k, t in dev.items():                          # inputs untouched
                assert np.array_equal(t.cpu().numpy(), np.ascontiguousarray(fields[k])), k
    finally:
        ch.free()


def test_noalloc_rejects_what_it_cannot_serve(product_lib, synth_table, capfd):
    torch = pytest.importorskip("torch")
    noalloc = _declare(product_lib)
    params, dt, _ = sf.config("C3", synth_table)
    ch = capi.Chemistry(product_lib, **params).initialize()
    try:
        # a grid beyond the static scratch: FAIL with a message, output untouched
        dims = [MAX_CELLS + 1]
        g = capi.Grid(dims)
        buf = torch.ones(MAX_CELLS + 1, dtype=torch.float64, device="cuda")
        for k in sf.fields_for(2):
            g.set(k, int(buf.data_ptr()))
        out = torch.full((MAX_CELLS + 1,), -7.0, dtype=torch.float64, device="cuda")
        rc = noalloc(C.byref(ch.chem), C.byref(ch.rates), C.byref(ch.units), C.byref(g.fd), C.c_void_p(int(out.data_ptr())))
        assert rc == 0
        assert "GRB200_NOALLOC_MAX_CELLS" in capfd.readouterr().err
        assert bool((out == -7.0).all().item())
        # host arrays: rejected (staging them would allocate)
        n = 256
        idx = np.arange(n, dtype=np.int64)
        fields = sf.make_fields(idx, 2, params, ch.temperature_units, "B", fh=ch.chem.HydrogenFractionByMass)
        gh = capi.Grid([n])
        keep = {k: np.ascontiguousarray(v) for k, v in fields.items()}
        for k, v in keep.items():
            gh.set(k, v)
        outh = np.full(n, -7.0)
        rc = noalloc(C.byref(ch.chem), C.byref(ch.rates), C.byref(ch.units), C.byref(gh.fd), C.c_void_p(outh.ctypes.data))
        assert rc == 0 and np.all(outh == -7.0)
        assert "device-resident" in capfd.readouterr().err
    finally:
        ch.free()
