"""Seeded synthetic inputs in the reference's own file formats.

Writers for the TLI v6 line list (layout as read by readlineinfo.c:87-244,416-537 and
written by pylineread.py:199-427), the atmosphere text file (readatm.c:276-620, example
transit/examples/demo/HD209458b_demo.atm), CS/CIA tables (crosssec.c:86-233) and the
``-c`` configuration file (procopt.c:271-398).  Used by the golden-fixture generator
(tests/golden/make_golden.py), the parity tests and bench.py.  Pure numpy: nothing here
touches the GPU or the oracle.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field

import numpy as np

TLI_MAGIC = b"\xff\xb6\xb3\xab"  # pylineread.py:196-201 (little-endian)

# (name, mass amu, isotopic ratio) -- pylineread/inputs/isotopologues.dat:14-49
ISOTOPOLOGUES = {
    "H2O": [("161", 18.0105646, 9.973e-01), ("181", 20.014811, 1.999e-03),
            ("171", 19.014781, 3.719e-04), ("162", 19.016841, 3.107e-04),
            ("182", 21.021088, 6.230e-07), ("172", 20.021058, 1.158e-07)],
    "CO2": [("626", 43.98982920, 9.842e-01), ("636", 44.99318400, 1.106e-02),
            ("628", 45.99407560, 3.947e-03), ("627", 44.99404630, 7.340e-04),
            ("638", 46.99743043, 4.434e-05), ("637", 45.99740113, 8.246e-06),
            ("828", 47.99832200, 3.957e-06), ("728", 46.99829270, 1.472e-06),
            ("727", 45.99826370, 1.368e-07), ("838", 49.00167680, 4.446e-08)],
    "CO": [("26", 27.99491460, 9.865e-01), ("36", 28.99826943, 1.108e-02),
           ("28", 29.99916100, 1.978e-03), ("27", 28.99913170, 3.679e-04),
           ("38", 31.00251583, 2.223e-05), ("37", 30.00248653, 4.133e-06)],
    "CH4": [("211", 16.031300, 9.883e-01), ("311", 17.034655, 1.110e-02),
            ("212", 17.037577, 6.158e-04), ("312", 18.040932, 6.918e-06)],
}


@dataclass
class Database:
    """One TLI database = one molecule's line list."""
    name: str
    molecule: str
    temps: np.ndarray                      # [nT]
    iso_names: list
    iso_mass: np.ndarray                   # [nIso]
    iso_ratio: np.ndarray                  # [nIso]
    Z: np.ndarray                          # [nIso][nT]
    # per-isotope line arrays (wavelength um ascending), filled by make_lines
    wl: list = field(default_factory=list)
    elow: list = field(default_factory=list)
    gf: list = field(default_factory=list)


def make_database(molecule: str, niso: int | None = None, tmin=70.0, tmax=3000.0,
                  tstep=10.0, dbname: str | None = None) -> Database:
    isos = ISOTOPOLOGUES[molecule][:niso]
    temps = np.arange(tmin, tmax + 0.5 * tstep, tstep)
    # smooth, increasing partition function ~ T^1.5 with an isotope-dependent scale
    Z = np.stack([(1.0 + 0.07 * k) * (temps / 296.0) ** 1.5 * (50.0 + 20.0 * len(molecule))
                  * (1.0 + 1e-4 * temps) for k in range(len(isos))])
    return Database(dbname or f"HITRAN {molecule}", molecule, temps,
                    [i[0] for i in isos], np.array([i[1] for i in isos]),
                    np.array([i[2] for i in isos]), Z)


def make_lines(db: Database, nlines: int, wn_lo: float, wn_hi: float,
               rng: np.random.Generator, bands: int = 6, flat: bool = False,
               loggf=(-12.0, -6.0), elow_max=6000.0, dense_pairs: float = 0.05):
    """Fill ``db`` with ``nlines`` lines inside [wn_lo, wn_hi] cm-1.

    ``flat`` draws wavenumbers uniformly (ExoMol-like), otherwise from a band mixture.
    ``dense_pairs`` is the fraction of lines placed within a fine-grid step of another
    line so that the co-adding branch (extinction.c:449-462) is exercised.
    """
    niso = len(db.iso_names)
    # split lines over isotopes roughly like abundance^0.25 so minor ones get some
    w = db.iso_ratio ** 0.25
    counts = np.floor(nlines * w / w.sum()).astype(np.int64)
    counts[0] += nlines - counts.sum()
    db.wl, db.elow, db.gf = [], [], []
    for k in range(niso):
        n = int(counts[k])
        if flat:
            wn = rng.uniform(wn_lo, wn_hi, n)
        else:
            centers = rng.uniform(wn_lo, wn_hi, bands)
            widths = rng.uniform(0.02, 0.08, bands) * (wn_hi - wn_lo)
            which = rng.integers(0, bands, n)
            wn = centers[which] + widths[which] * rng.standard_normal(n)
            bad = (wn < wn_lo) | (wn > wn_hi)
            wn[bad] = rng.uniform(wn_lo, wn_hi, int(bad.sum()))
        npair = int(dense_pairs * n)
        if npair > 0 and n > 2 * npair:
            src = rng.integers(0, n, npair)
            dst = rng.choice(n, npair, replace=False)
            wn[dst] = np.clip(wn[src] + rng.uniform(-3e-4, 3e-4, npair), wn_lo, wn_hi)
        wl = 1.0 / (wn * 1e-4)             # microns
        order = np.argsort(wl, kind="stable")
        db.wl.append(wl[order])
        db.elow.append(rng.uniform(0.0, elow_max, n)[order])
        db.gf.append(10.0 ** rng.uniform(loggf[0], loggf[1], n)[order])
    return db


def write_tli(path: str, dbs: list, wl_lo_um: float, wl_hi_um: float) -> int:
    """TLI v6.  Returns the number of transitions written."""
    with open(path, "wb") as f:
        f.write(TLI_MAGIC)
        f.write(struct.pack("<3H", 6, 6, 3))
        f.write(struct.pack("<2d", wl_lo_um, wl_hi_um))
        f.write(struct.pack("<H", len(dbs)))
        for db in dbs:
            for s in (db.name, db.molecule):
                b = s.encode()
                f.write(struct.pack("<H", len(b)) + b)
            f.write(struct.pack("<2H", len(db.temps), len(db.iso_names)))
            f.write(np.asarray(db.temps, "<f8").tobytes())
            for k, nm in enumerate(db.iso_names):
                b = nm.encode()
                f.write(struct.pack("<H", len(b)) + b)
                f.write(struct.pack("<2d", db.iso_mass[k], db.iso_ratio[k]))
                f.write(np.asarray(db.Z[k], "<f8").tobytes())
        wl, iso, el, gf, counts = [], [], [], [], []
        base = 0
        for db in dbs:
            for k in range(len(db.iso_names)):
                wl.append(db.wl[k]); el.append(db.elow[k]); gf.append(db.gf[k])
                iso.append(np.full(len(db.wl[k]), base + k, "<i2"))
                counts.append(len(db.wl[k]))
            base += len(db.iso_names)
        n = int(sum(counts))
        f.write(struct.pack("<Q", n))
        f.write(struct.pack("<i", len(counts)))
        f.write(np.asarray(counts, "<u8").tobytes())
        f.write(np.concatenate(wl).astype("<f8").tobytes())
        f.write(np.concatenate(iso).astype("<i2").tobytes())
        f.write(np.concatenate(el).astype("<f8").tobytes())
        f.write(np.concatenate(gf).astype("<f8").tobytes())
    return n


def make_atmosphere(nlayers: int, species: dict, p_top_bar=1e-8, p_bot_bar=1e2,
                    t_top=700.0, t_bot=1100.0, r_bot_km=2.0e4, g_cgs=1500.0,
                    rng: np.random.Generator | None = None):
    """Layer table, bottom (high p) first as in HD209458b_demo.atm.

    ``species`` maps name -> constant mole fraction.  Radii are hydrostatic for the mean
    molecular mass so that they are strictly increasing (readatm.c:585-601).
    Returns dict(radius_km, p_bar, T, q[name]).
    """
    p = np.logspace(np.log10(p_bot_bar), np.log10(p_top_bar), nlayers)
    x = np.linspace(0.0, 1.0, nlayers)
    T = t_bot + (t_top - t_bot) * (3 * x**2 - 2 * x**3)
    if rng is not None:
        T = T + rng.uniform(-5.0, 5.0, nlayers)
    masses = {"H2": 2.01588, "He": 4.002602, "H2O": 18.01528, "CH4": 16.0425,
              "CO": 28.0101, "CO2": 44.0095, "H": 1.00794, "N2": 28.0134}
    mm = sum(q * masses[s] for s, q in species.items())
    kb, amu = 1.380658e-16, 1.66053886e-24
    r = np.empty(nlayers)
    r[0] = r_bot_km
    for i in range(1, nlayers):
        Tm = 0.5 * (T[i] + T[i - 1])
        H = kb * Tm / (mm * amu * g_cgs) / 1e5      # km
        r[i] = r[i - 1] + H * np.log(p[i - 1] / p[i])
    return dict(radius_km=r, p_bar=p, T=T, q={s: np.full(nlayers, q) for s, q in species.items()})


def write_atmosphere(path: str, atm: dict):
    names = list(atm["q"].keys())
    with open(path, "w") as f:
        f.write("# synthetic atmosphere (transit_b200.synth)\n\n")
        f.write("ur 1e5\nup 1e6\nq number\n\n#SPECIES\n" + " ".join(names) + "\n\n#TEADATA\n")
        f.write("#Radius    Pressure   Temp       " + " ".join(f"{n:10s}" for n in names) + "\n")
        for i in range(len(atm["T"])):
            f.write(f" {atm['radius_km'][i]:.4f} {atm['p_bar'][i]:.4e} {atm['T'][i]:.2f} "
                    + " ".join(f"{atm['q'][n][i]:.4e}" for n in names) + " \n")


def write_cia(path: str, sp1: str, sp2: str, wn: np.ndarray, temps: np.ndarray,
              scale=1e-7, seed=0):
    """Smooth positive synthetic CIA table, cm-1 amagat-2 (crosssec.c:98-233)."""
    rng = np.random.default_rng(seed)
    ph = rng.uniform(0, 2 * np.pi, 3)
    w = wn[:, None]
    t = temps[None, :]
    tab = scale * (1.0 + 0.5 * np.sin(w / 900.0 + ph[0]) + 0.3 * np.cos(w / 3100.0 + ph[1])) \
        * np.exp(-((w - 4000.0) / 9000.0) ** 2) * (t / 1000.0) ** 0.5 * (1 + 0.1 * np.sin(t / 700.0 + ph[2]))
    with open(path, "w") as f:
        f.write(f"# synthetic CIA table\n\ni {sp1} {sp2}\nt " + " ".join(f"{x:10.1f}" for x in temps) + "\n\n")
        f.write("# Wavenumber in cm-1, CIA coefficients in cm-1 amagat-2:\n")
        for i in range(len(wn)):
            f.write(f" {wn[i]:9.2f}    " + " ".join(f"{v:.4e}" for v in tab[i]) + "\n")


def write_cfg(path: str, **keys):
    """``key value`` lines (procopt.c:271-398).  Keys are the reference's long options."""
    with open(path, "w") as f:
        for k, v in keys.items():
            if v is None:
                continue
            if isinstance(v, (list, tuple)):
                v = " ".join(str(x) for x in v)
            f.write(f"{k} {v}\n")
