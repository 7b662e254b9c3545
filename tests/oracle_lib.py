"""The CPU ORACLE's loader - test infrastructure.  Only tests/, __graft_entry__.smoke() and
bench.py's CPU-baseline legs may import this module; nothing under grackle_b200/ does.

oracle/_ref/libgrackle_oracle.so exports the same C ABI as the product (the reference's own C
sources compiled in place + the C++ restatement of its Fortran), so the ctypes struct
definitions of grackle_b200/capi.py serve both.  Builds (oracle/Makefile):
  plain   -O2 -ffp-contract=off                      the checker
  fma     -O2 -ffp-contract=fast -mfma                the same sources, contraction on: finds the
                                                      cells whose reference answer is not
                                                      reproducible across compiler flags (parity.py)
  count   operation-counting build (flop counts for the roofline)
"""
import ctypes as C
import os

import numpy as np

from grackle_b200 import capi

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(REPO, "oracle", "_ref")
_NAMES = {"plain": "libgrackle_oracle.so", "fma": "libgrackle_oracle_fma.so", "count": "libgrackle_oracle_count.so"}


def load_oracle(kind="plain"):
    so = os.path.join(ORACLE_DIR, _NAMES[kind])
    if not os.path.exists(so):
        raise capi.ExtensionMissing("oracle library %s is missing (make -C oracle)" % so)
    lib = capi._declare(C.CDLL(so, mode=C.RTLD_LOCAL))
    if kind == "count":
        lib.oracle_flop_counts.argtypes = [C.POINTER(C.c_longlong), C.c_int]
        lib.oracle_flop_counts.restype = None
    lib.oracle_inject_cloudy.argtypes = [C.c_int, C.c_int, C.c_longlong, C.POINTER(C.c_longlong),
                                         C.POINTER(capi.c_dp), capi.c_dp, capi.c_dp, capi.c_dp]
    lib.oracle_inject_cloudy.restype = None
    lib.oracle_inject_uvb.argtypes = [C.c_longlong, C.POINTER(capi.c_dp)]
    lib.oracle_inject_uvb.restype = None
    lib.oracle_clear_injected.restype = None
    lib.oracle_last_subcycles_sum.restype = C.c_longlong
    lib.oracle_set_roundoff_noise.argtypes = [C.c_double, C.c_ulonglong]
    lib.oracle_set_roundoff_noise.restype = None
    lib.oracle_last_subcycles_max.restype = C.c_longlong
    return lib


class OracleChemistry(capi.Chemistry):
    """chemistry_data + storage + units bound to an oracle library"""

    # read the data file in Python and hand the raw arrays over (the reference's own loaders need libhdf5)
    def _before_initialize(self):
        from grackle_b200 import h5lite
        self.lib.oracle_clear_injected()
        need_file = (self.chem.primordial_chemistry == 0 or self.chem.metal_cooling == 1 or
                     (self.chem.UVbackground == 1 and self.chem.primordial_chemistry > 0))
        if not need_file:
            return
        f = h5lite.read_file(self._file.decode())
        old_style = 1 if "old_style" in f.attrs else 0
        keep = self._keep = []
        for which, group in enumerate(("Primordial", "Metals")):
            path = "CoolingRates/%s/Cooling" % group
            if path not in f:
                continue
            node = f[path]
            rank = int(node.attrs["Rank"])
            dims = np.atleast_1d(node.attrs["Dimension"]).astype(np.int64)
            pars = []
            for q in range(rank):
                name = "Temperature" if q == rank - 1 else "Parameter%d" % (q + 1)
                pars.append(np.ascontiguousarray(node.attrs[name], dtype=np.float64))
            cool = np.ascontiguousarray(node.data, dtype=np.float64).ravel()
            hpath, mpath = "CoolingRates/%s/Heating" % group, "CoolingRates/%s/MMW" % group
            heat = np.ascontiguousarray(f[hpath].data, dtype=np.float64).ravel() if hpath in f else None
            mmw = np.ascontiguousarray(f[mpath].data, dtype=np.float64).ravel() if mpath in f else None
            parr = (capi.c_dp * rank)(*[capi._dptr(p) for p in pars])
            darr = (C.c_longlong * rank)(*[int(d) for d in dims])
            keep += [pars, cool, heat, mmw]
            self.lib.oracle_inject_cloudy(which, old_style, rank, darr, parr, capi._dptr(cool), capi._dptr(heat), capi._dptr(mmw))
        if "UVBRates/z" in f:
            names = ["z", "Chemistry/k24", "Chemistry/k25", "Chemistry/k26", "Chemistry/k27",
                     "Chemistry/k28", "Chemistry/k29", "Chemistry/k30", "Chemistry/k31",
                     "Photoheating/piHI", "Photoheating/piHeI", "Photoheating/piHeII",
                     "CrossSections/hi_avg_crs", "CrossSections/hei_avg_crs", "CrossSections/heii_avg_crs"]
            vecs = []
            for n in names:
                p = "UVBRates/" + n
                vecs.append(np.ascontiguousarray(f[p].data, dtype=np.float64).ravel() if p in f else None)
            keep.append(vecs)
            varr = (capi.c_dp * 15)(*[capi._dptr(v) for v in vecs])
            self.lib.oracle_inject_uvb(len(vecs[0]), varr)

