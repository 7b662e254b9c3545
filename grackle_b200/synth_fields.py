"""Deterministic synthetic input fields and the BASELINE configurations.

Every value is a pure function of (seed, global cell index, stream id) through a
counter-based generator (splitmix64 -> uniform [0,1)), so a GPU can fill its own slab and
the CPU oracle can fill any sub-block identically.  Works on numpy arrays or torch
tensors (same integer arithmetic; torch is used only to place the data in HBM).

Set H: the distribution of the reference's own timing harness
       (src/example/cxx_omp_example.C:111-114, 295-303, 830-859).
Set B: broad log-uniform n_H, T, Z, ionisation and H2 fraction (the headline set).
"""
import numpy as np

MH = 1.67262171e-24
KBOLTZ = 1.3806504e-16
SEED_H = 123
SEED_B = 20261104
Z_SOLAR = 0.01295

_M64 = (1 << 64) - 1


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _const(xp_like, v):
    v &= _M64
    if v >= 1 << 63:
        v -= 1 << 64
    if _is_torch(xp_like):
        import torch
        return torch.tensor(v, dtype=torch.int64, device=xp_like.device)
    return np.int64(v)


def _lsr(x, s):
    """logical shift right of int64 (both numpy and torch shift arithmetically)"""
    return (x >> s) & _const(x, (1 << (64 - s)) - 1)


def uniform(index, seed, stream):
    """index: int64 array/tensor of global cell ids -> float64 uniform [0,1)."""
    with np.errstate(over="ignore"):
        z = index * _const(index, 0x9E3779B97F4A7C15) + _const(index, (seed * 0xD1342543DE82EF95 + stream * 0xBF58476D1CE4E5B9) & _M64)
        z = (z ^ _lsr(z, 30)) * _const(index, 0xBF58476D1CE4E5B9)
        z = (z ^ _lsr(z, 27)) * _const(index, 0x94D049BB133111EB)
        z = z ^ _lsr(z, 31)
    top = _lsr(z, 11)
    if _is_torch(top):
        import torch
        return top.to(torch.float64) * (1.0 / 9007199254740992.0)
    return top.astype(np.float64) * (1.0 / 9007199254740992.0)


def _pow10(x):
    if _is_torch(x):
        import torch
        return torch.pow(torch.tensor(10.0, dtype=torch.float64, device=x.device), x)
    return np.power(10.0, x)


def cosmology_units(z=3.0, z_init=99.0, h=0.704, omega_m=0.268, box=1.0):
    """pygrackle/utilities/units.py:16-57"""
    a_units = 1.0 / (1.0 + z_init)
    return dict(comoving_coordinates=1, a_units=a_units, a_value=1.0 / (1.0 + z) / a_units,
                density_units=1.8788e-29 * omega_m * h ** 2 * (1 + z) ** 3,
                length_units=3.085678e24 * box / h / (1.0 + z),
                time_units=2.519445e17 / np.sqrt(omega_m) / h / (1 + z_init) ** 1.5)


def proper_units():
    """cxx_omp_example.C:186-193 / c_example.c"""
    return dict(comoving_coordinates=0, a_units=1.0, a_value=1.0, density_units=1.67e-24,
                length_units=1.0, time_units=1.0e12)


# NB C4: SURVEY 8d proposed dt = 10 Myr; with Set B's dense cells (n_H up to 100 cm^-3) that
# makes ~10 % of the cells exceed max_iterations=10000 in the reference algorithm itself
# (limit cycle around thermal equilibrium, sub-step ~ const), i.e. an error condition, so
# the benchmark uses 1 Myr like C1-C3 (no cell hits the limit; max ~ 400 subcycles).
CONFIGS = {
    # name: (chemistry parameters, units, dt in seconds, default grid)
    "C1": (dict(primordial_chemistry=0, metal_cooling=1, UVbackground=1), proper_units(), 3.15e13, (64, 64, 64)),
    "C2": (dict(primordial_chemistry=1, metal_cooling=1, UVbackground=1), proper_units(), 3.15e13, (128, 128, 128)),
    "C3": (dict(primordial_chemistry=2, metal_cooling=1, UVbackground=1, h2_on_dust=1,
                self_shielding_method=3), proper_units(), 3.15e13, (256, 256, 256)),
    "C4": (dict(primordial_chemistry=3, metal_cooling=1, UVbackground=1), cosmology_units(), 3.15e13, (1024, 1024, 1024)),
    # the parameter set of the reference's own examples and timing harness (c_example.c:52-70,
    # cxx_omp_example.C:160-193): pc = 3 + dust_chemistry (which switches on H2 on dust, grain
    # recombination cooling and photo-electric heating 2) + metals + UVB, proper units, input Set H
    "HARNESS": (dict(primordial_chemistry=3, dust_chemistry=1, metal_cooling=1, UVbackground=1), proper_units(),
                3.15e13, (128, 128, 128)),
}


def config(name, data_file):
    chem, units, dt_s, grid = CONFIGS[name]
    p = dict(use_grackle=1, with_radiative_cooling=1, grackle_data_file=data_file)
    p.update(chem)
    p.update(units)
    return p, dt_s / units["time_units"], grid


FIELD_ORDER = ["density", "internal_energy", "metal_density", "HI_density", "HII_density",
               "HeI_density", "HeII_density", "HeIII_density", "e_density", "HM_density",
               "H2I_density", "H2II_density", "DI_density", "DII_density", "HDI_density"]


def fields_for(pc):
    n = {0: 3, 1: 9, 2: 12, 3: 15}[pc]
    return FIELD_ORDER[:n]


def make_fields(index, pc, units, temperature_units, set_name="B", fh=0.76, gamma=5.0 / 3.0,
                dtoh=6.8e-5, set_c_geometry=(1024 * 1024, 1024)):
    """index: int64 global cell ids.  Returns dict field-name -> float64 array (code units).
    set_c_geometry = (cells per k-plane, k-planes of the GLOBAL grid), used by Set C only."""
    xp_log10 = (lambda t: t.log10()) if _is_torch(index) else np.log10
    out = {}
    if set_name == "H":
        u = lambda s: 0.9 + 0.2 * uniform(index, SEED_H, s)
        rho = 1.0 * u(0)
        out["density"] = rho
        out["metal_density"] = Z_SOLAR * rho * u(1)
        out["internal_energy"] = (1000.0 / temperature_units) * u(2)
        tiny = 1e-20
        if pc > 0:
            out["HI_density"] = fh * rho
            out["HII_density"] = tiny * rho
            out["HeI_density"] = (1.0 - fh) * rho
            out["HeII_density"] = tiny * rho
            out["HeIII_density"] = tiny * rho
            out["e_density"] = tiny * rho
        if pc > 1:
            out["HM_density"] = tiny * rho
            out["H2I_density"] = tiny * rho
            out["H2II_density"] = tiny * rho
        if pc > 2:
            out["DI_density"] = dtoh * rho
            out["DII_density"] = tiny * rho
            out["HDI_density"] = tiny * rho
        return {k: out[k] for k in fields_for(pc)}

    u = lambda s: uniform(index, SEED_B, s)
    if set_name == "C":
        # Set C = Set B with SPATIAL STRUCTURE along k (the slab axis of the multi-GPU decomposition):
        # the density rises by five decades from the first to the last k-plane of the global grid, with
        # one decade of cell-to-cell scatter, so subcycle counts - hence k-slab run times - differ
        # several-fold between slabs (Set B is i.i.d.: every slab costs the same by construction)
        plane, nk = set_c_geometry
        s_k = (index // plane) % nk
        s_k = (s_k.to(u(0).dtype) if _is_torch(index) else s_k.astype(np.float64)) / float(nk)
        n_h = _pow10(-4.0 + 5.0 * s_k + 1.0 * u(0))
    else:
        n_h = _pow10(-4.0 + 6.0 * u(0))             # proper cm^-3
    T = _pow10(2.0 + 5.0 * u(1))                    # K
    zmet = _pow10(-3.0 + 3.0 * u(2))                # Z / Zsun
    xe = _pow10(-4.0 + 4.0 * u(3))                  # ionised fraction of H
    fH2 = _pow10(-8.0 + 6.0 * u(4))                 # H2 mass fraction of H
    rho = n_h * MH / (fh * units["density_units"])  # code density (total, incl. metals)
    metal = Z_SOLAR * zmet * rho
    gas = rho - metal
    rhoH = fh * gas
    rhoHe = (1.0 - fh) * gas
    out["density"] = rho
    out["metal_density"] = metal
    if pc == 0:
        mu = 1.22 - 0.62 * xe
    else:
        H2I = fH2 * rhoH if pc > 1 else 0.0 * rhoH
        HM = 1e-10 * xe * rhoH if pc > 1 else 0.0 * rhoH
        H2II = 1e-3 * xe * H2I if pc > 1 else 0.0 * rhoH
        rest = rhoH - H2I - HM - H2II
        HII = xe * rest
        HI = rest - HII
        HeII = 0.5 * xe * rhoHe
        HeIII = 0.5 * xe * xe * rhoHe
        HeI = rhoHe - HeII - HeIII
        de = HII + HeII / 4.0 + HeIII / 2.0
        if pc > 1:
            de = de - HM + H2II / 2.0
        out.update(HI_density=HI, HII_density=HII, HeI_density=HeI, HeII_density=HeII,
                   HeIII_density=HeIII, e_density=de)
        if pc > 1:
            out.update(HM_density=HM, H2I_density=H2I, H2II_density=H2II)
        if pc > 2:
            out.update(DI_density=dtoh * HI, DII_density=dtoh * HII, HDI_density=1.5 * dtoh * H2I)
        nden = (HeI + HeII + HeIII) / 4.0 + HI + HII + de
        if pc > 1:
            nden = nden + HM + (H2I + H2II) / 2.0
        nden = nden + metal / 16.0
        mu = rho / nden
    out["internal_energy"] = T / (mu * (gamma - 1.0) * temperature_units)
    return {k: out[k] for k in fields_for(pc)}
