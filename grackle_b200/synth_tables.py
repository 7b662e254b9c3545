"""Deterministic synthetic Grackle data file with the HM2012 schema.

The reference's `CloudyData_UVB=HM2012.h5` is not in the reference snapshot
(.MISSING_LARGE_BLOBS), yet four of the five BASELINE configs name it.  This writes a file
with the SAME on-disk schema (initialize_cloudy_data.c:92-305,
initialize_UVbackground_data.c:50-330): rank-3 (log n_H, z, T) Cloudy grids for
/CoolingRates/{Primordial,Metals}/{Cooling,Heating,MMW} and the /UVBRates vectors, filled
with smooth, physically plausible closed-form functions (no random numbers), so the
oracle and the CUDA path are compared on identical tables.  Every report states that the
table is synthetic.  If the real HM2012 file is supplied it is used unchanged.

Grid sizes follow the real file as described in SURVEY.md (29 n_H x 26 z x 161 T).
"""
import os

import numpy as np

from . import h5lite

N_NH, N_Z, N_T = 29, 26, 161
N_UVB_Z = 60


def _axes():
    log_nh = np.linspace(-10.0, 4.0, N_NH)
    # HM2012 tabulates z densely at low z, sparsely above; strictly increasing
    z = np.concatenate([np.linspace(0.0, 5.0, 18), np.linspace(5.6, 15.13, N_Z - 18)])
    # the real files store the temperature axis rounded to 7 significant digits (CloudyData_noUVB.h5:
    # 10, 11.22018, 12.58925, ...), i.e. only nominally uniform in log10 T: reproduce that, because
    # the interpolation weights must come from the STORED axis values (interpolators_g.F:80-98)
    T = np.array([float("%.7g" % v) for v in 10.0 ** np.linspace(1.0, 9.0, N_T)])
    return log_nh, z, T


def _uvb_shape(z):
    """relative UVB intensity vs z: peaks near z~2, falls steeply beyond z~6"""
    return 10.0 ** (-0.12 * (z - 2.0) ** 2 / (1.0 + 0.08 * z)) * (0.35 + 0.65 / (1.0 + (z / 7.0) ** 6))


def build_tables():
    log_nh, z, T = _axes()
    L = np.log10(T)
    nh = log_nh[:, None, None]
    zz = z[None, :, None]
    lt = L[None, None, :]
    uv = _uvb_shape(zz)

    # metal cooling per n_H^2 at solar metallicity [erg cm^3 s^-1]
    peak = -21.6 - 0.45 * ((lt - 5.3) / 0.9) ** 2 / (1.0 + 0.25 * np.abs(lt - 5.3))
    fine = -25.8 + 0.9 * np.tanh((lt - 2.0) / 0.8) - 0.03 * (nh + 3.0)
    brem = -23.4 + 0.5 * (lt - 7.5)
    metal_cool = np.log10(10.0 ** peak + 10.0 ** fine + 10.0 ** brem)
    metal_cool = metal_cool - 0.15 * uv * np.exp(-((lt - 4.6) / 0.7) ** 2) / (1.0 + 10.0 ** (nh + 2.0))
    metal_cool = metal_cool + 0.02 * np.sin(1.7 * lt + 0.3 * nh) * np.cos(0.9 * zz)
    metal_heat = -24.6 + np.log10(uv) - 0.55 * np.log10(1.0 + 10.0 ** (nh + 3.5)) - 0.08 * (lt - 4.0) ** 2 / (1 + 0.1 * lt)

    # primordial (H+He) cooling/heating and mean molecular weight
    hpeak = -21.9 - 1.6 * ((lt - 4.25) / 0.35) ** 2 / (1.0 + 2.0 * np.abs(lt - 4.25))
    hepeak = -22.4 - 1.2 * ((lt - 4.95) / 0.35) ** 2 / (1.0 + 2.0 * np.abs(lt - 4.95))
    ff = -23.6 + 0.5 * (lt - 7.0)
    low = -29.5 + 1.5 * (lt - 1.0)
    pri_cool = np.log10(10.0 ** hpeak + 10.0 ** hepeak + 10.0 ** ff + 10.0 ** np.minimum(low, -23.5))
    pri_cool = pri_cool - 0.3 * uv / (1.0 + 10.0 ** (nh + 2.5)) * np.exp(-((lt - 4.3) / 0.5) ** 2)
    pri_heat = -23.9 + np.log10(uv) - 0.8 * np.log10(1.0 + 10.0 ** (nh + 2.8)) - 0.05 * (lt - 4.0) ** 2
    ion = 1.0 / (1.0 + np.exp(-(lt - 4.15 + 0.12 * np.log10(1.0 + 10.0 ** (nh + 2.0)) - 0.25 * uv / (1 + 10.0 ** (nh + 2.0))) / 0.08))
    mmw = 1.2195 - (1.2195 - 0.5882) * ion + 0.0 * zz

    tables = {
        "Metals": {"Cooling": 10.0 ** metal_cool, "Heating": 10.0 ** metal_heat},
        "Primordial": {"Cooling": 10.0 ** pri_cool, "Heating": 10.0 ** pri_heat, "MMW": mmw},
    }
    axes = {"Parameter1": log_nh, "Parameter2": z, "Temperature": T}

    zu = np.linspace(0.0, 15.13, N_UVB_Z)
    s = _uvb_shape(zu)
    uvb = {
        "z": zu,
        "Chemistry/k24": 9.0e-13 * s,
        "Chemistry/k25": 1.1e-14 * s * 10.0 ** (-0.25 * np.maximum(zu - 2.5, 0.0)),
        "Chemistry/k26": 5.5e-13 * s,
        "Chemistry/k27": 1.1e-9 * s ** 0.6,
        "Chemistry/k28": 3.2e-10 * s ** 0.6,
        "Chemistry/k29": 2.6e-13 * s,
        "Chemistry/k30": 1.9e-12 * s,
        "Chemistry/k31": 8.0e-10 * s ** 0.7,
        "Photoheating/piHI": 3.6e-12 * s,           # eV/s
        "Photoheating/piHeI": 4.4e-12 * s,
        "Photoheating/piHeII": 2.1e-13 * s * 10.0 ** (-0.25 * np.maximum(zu - 2.5, 0.0)),
        "CrossSections/hi_avg_crs": 2.6e-18 * (1.0 + 0.05 * np.tanh(zu - 3.0)),
        "CrossSections/hei_avg_crs": 1.9e-18 * (1.0 + 0.04 * np.tanh(zu - 3.0)),
        "CrossSections/heii_avg_crs": 6.8e-19 * (1.0 + 0.03 * np.tanh(zu - 3.0)),
    }
    return tables, axes, uvb


def write(path):
    tables, axes, uvb = build_tables()
    dims = np.array([N_NH, N_Z, N_T], dtype=np.int64)
    root = h5lite.Group()
    cr = h5lite.Group()
    for grp, sets in tables.items():
        g = h5lite.Group()
        for name, arr in sets.items():
            attrs = {"Rank": np.int64(3), "Dimension": dims, "Parameter1": axes["Parameter1"],
                     "Parameter2": axes["Parameter2"], "Temperature": axes["Temperature"],
                     "Parameter1_Name": "hden", "Parameter2_Name": "redshift"}
            g[name] = h5lite.Dataset(np.ascontiguousarray(arr, dtype=np.float64), attrs)
        cr[grp] = g
    root["CoolingRates"] = cr
    ub = h5lite.Group()
    ub["Info"] = h5lite.Dataset("SYNTHETIC HM2012-schema UV background (grackle-b200 synth_tables.py)")
    ub["z"] = h5lite.Dataset(uvb["z"])
    for sub in ("Chemistry", "Photoheating", "CrossSections"):
        g = h5lite.Group()
        for k, v in uvb.items():
            if k.startswith(sub + "/"):
                g[k.split("/")[1]] = h5lite.Dataset(np.ascontiguousarray(v, dtype=np.float64))
        ub[sub] = g
    root["UVBRates"] = ub
    h5lite.write_file(path, root)
    return path


def default_path():
    """<repo>/gpurun_out is scratch; keep the generated table next to the package"""
    d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")
    os.makedirs(d, exist_ok=True)
    return os.path.join(d, "CloudyData_UVB=SYNTHETIC_HM2012_v2.h5")


def ensure(path=None):
    path = path or default_path()
    if not os.path.exists(path):
        write(path)
    return path


# ---------------------------------------------------------------------------------------
# Old-style metal tables of rank 4 and 5 (cool1d_cloudy_old_tables_g.F; the reference's
# `cloudy_metals_2008_{4,5}D.h5` family is not in the snapshot, only the rank-3 file is).
# Same on-disk schema as cloudy_metals_2008_3D.h5: root attribute `old_style`,
# /CoolingRates/Metals/{Cooling,Heating} with Rank / Dimension / Parameter1..Parameter(R-1) /
# Temperature attributes.  Axes: log10 n_H, log10 Z/Zsun, log10 electron fraction,
# [redshift], T.  /UVBRates is carried along so that UVbackground = 1 (which is what makes
# the loader read Heating) can be used with these files.
# ---------------------------------------------------------------------------------------
OLD_DIMS = {4: (8, 5, 6, 33), 5: (8, 5, 6, 7, 33)}


def build_old_style(rank):
    assert rank in (4, 5)
    dims = OLD_DIMS[rank]
    log_nh = np.linspace(-6.0, 4.0, dims[0])
    log_z = np.linspace(-4.0, 1.0, dims[1])
    log_ef = np.linspace(-5.0, 0.0, dims[2])
    T = 10.0 ** np.linspace(1.0, 8.0, dims[-1])
    axes = [log_nh, log_z, log_ef]
    if rank == 5:
        axes.append(np.linspace(0.0, 9.0, dims[3]))
    axes.append(T)
    shape = [1] * rank
    g = []
    for q, ax in enumerate(axes):
        sh = list(shape)
        sh[q] = len(ax)
        g.append((np.log10(ax) if q == rank - 1 else ax).reshape(sh))
    nh, lz, ef, lt = g[0], g[1], g[2], g[-1]
    zz = g[3] if rank == 5 else 0.0 * nh
    uv = _uvb_shape(zz)
    peak = -21.9 - 0.5 * ((lt - 5.2) / 0.9) ** 2 / (1.0 + 0.25 * np.abs(lt - 5.2))
    fine = -26.1 + 0.8 * np.tanh((lt - 2.0) / 0.8) - 0.04 * (nh + 2.0)
    cool = np.log10(10.0 ** peak + 10.0 ** fine) + 1.0 * lz - 0.35 * ef * (1.0 + 0.05 * lt)
    cool = cool + 0.03 * np.sin(1.3 * lt + 0.4 * nh + 0.7 * lz) * np.cos(0.5 * ef + 0.6 * zz)
    cool = cool - 0.1 * uv / (1.0 + 10.0 ** (nh + 2.0))
    heat = -25.0 + np.log10(uv) + 1.0 * lz - 0.5 * np.log10(1.0 + 10.0 ** (nh + 3.0)) \
        - 0.06 * (lt - 4.0) ** 2 + 0.2 * ef
    cool = np.broadcast_to(cool, dims)
    heat = np.broadcast_to(heat, dims)
    return axes, 10.0 ** cool, 10.0 ** heat


def write_old_style(path, rank):
    axes, cool, heat = build_old_style(rank)
    _, _, uvb = build_tables()
    dims = np.array(cool.shape, dtype=np.int64)
    attrs = {"Rank": np.int64(rank), "Dimension": dims}
    for q, ax in enumerate(axes):
        attrs["Temperature" if q == rank - 1 else "Parameter%d" % (q + 1)] = np.ascontiguousarray(ax)
    g = h5lite.Group()
    g["Cooling"] = h5lite.Dataset(np.ascontiguousarray(cool, dtype=np.float64), dict(attrs))
    g["Heating"] = h5lite.Dataset(np.ascontiguousarray(heat, dtype=np.float64), dict(attrs))
    cr = h5lite.Group()
    cr["Metals"] = g
    root = h5lite.Group(attrs={"old_style": np.int64(1)})
    root["CoolingRates"] = cr
    ub = h5lite.Group()
    ub["Info"] = h5lite.Dataset("SYNTHETIC UV background (grackle-b200 synth_tables.py)")
    ub["z"] = h5lite.Dataset(uvb["z"])
    for sub in ("Chemistry", "Photoheating", "CrossSections"):
        sg = h5lite.Group()
        for k, v in uvb.items():
            if k.startswith(sub + "/"):
                sg[k.split("/")[1]] = h5lite.Dataset(np.ascontiguousarray(v, dtype=np.float64))
        ub[sub] = sg
    root["UVBRates"] = ub
    h5lite.write_file(path, root)
    return path


def ensure_old_style(rank):
    d = os.path.dirname(default_path())
    path = os.path.join(d, "cloudy_metals_SYNTHETIC_%dD.h5" % rank)
    if not os.path.exists(path):
        write_old_style(path, rank)
    return path
