"""Synthetic problems of the shapes BASELINE.json names, as plain arrays.

``make_problem`` returns a ``Problem`` = the state ``transit_init`` leaves behind (grids,
species, line list, sampled atmosphere, CIA tables, geometry), built directly in memory with
seeded numpy.  ``write_reference_inputs`` writes the same problem in the reference's file
formats so that the unmodified reference binary can be timed on it (bench.py cpu_baseline).
No GPU, no oracle here.
"""
from __future__ import annotations

import os
from types import SimpleNamespace

import numpy as np

from . import synth

AMU, KB = 1.66053886e-24, 1.380658e-16

# molecules.dat rows used here: (ID, mass, radius A, polarizability) -- inputs/molecules.dat
MOLDAT = {"H2": (105, 2.01588, 2.89, 0.787), "He": (2, 4.0026020, 2.0, 0.208),
          "H2O": (101, 18.01528, 3.2, 1.501), "CH4": (102, 16.0425, 4.0, 2.448),
          "CO": (103, 28.0101, 2.8, 1.953), "CO2": (104, 44.0095, 2.8, 2.507)}

CONFIGS = {
    # C1: transit_demo.cfg shape (CH4, 2-4 um, eclipse, 5 angles)
    "C1": dict(wn_lo=2500.0, wn_hi=5000.0, wndelt=1.0, wnosamp=2160, nlayers=100, nlines=200_000,
               molecules={"CH4": (4, 1.0)}, species={"H2": 0.85, "He": 0.1496, "CH4": 4e-4},
               eclipse=True, angles=[0, 20, 40, 60, 80], toomuch=10.0, ethresh=1e-50, t_top=1000.0,
               t_bot=1650.0, cia=[("H2", "H2")], seed=1001, starrad=1.125),
    # C2: HAT-P-11b-shape transmission 0.3-5 um, 4 line-list molecules + 2 CIA tables
    "C2": dict(wn_lo=2000.0, wn_hi=33333.0, wndelt=1.0, wnosamp=2160, nlayers=100, nlines=1_000_000,
               molecules={"H2O": (6, 0.5), "CH4": (4, 0.2), "CO": (6, 0.05), "CO2": (10, 0.25)},
               species={"H2": 0.85, "He": 0.1485, "H2O": 5e-4, "CH4": 3e-4, "CO": 4e-4, "CO2": 3e-4},
               eclipse=False, toomuch=10.0, ethresh=1e-6, t_top=700.0, t_bot=1100.0,
               cia=[("H2", "H2"), ("H2", "He")], seed=1002, starrad=0.75),
}


class Problem(SimpleNamespace):
    pass


def grid_count(wn_i, wn_f, d, o):
    """makesample1's sample count (makesample.c:78-95)."""
    n = int(((1.0 + 1e-8) * wn_f - wn_i) / d + 1)
    return (abs(n) - 1) * o + 1


def partition_function(mol, k, T):
    """The analytic Z(T) synth.make_database tabulates (so tables and direct values agree)."""
    return (1.0 + 0.07 * k) * (T / 296.0) ** 1.5 * (50.0 + 20.0 * len(mol)) * (1.0 + 1e-4 * T)


def make_problem(name="C2", scale=1.0, nlayers=None, wn_hi=None, wnosamp=None, seed=None):
    """Build configuration ``name``; ``scale`` multiplies the line count (parity tests use <1)."""
    cfg = dict(CONFIGS[name])
    if nlayers: cfg["nlayers"] = nlayers
    if wn_hi: cfg["wn_hi"] = wn_hi
    if wnosamp: cfg["wnosamp"] = wnosamp
    if seed: cfg["seed"] = seed
    rng = np.random.default_rng(cfg["seed"])
    wn_i, d, o = cfg["wn_lo"], cfg["wndelt"], cfg["wnosamp"]
    nwn = grid_count(wn_i, cfg["wn_hi"], d, 1)
    onwn = (nwn - 1) * o + 1
    names = list(cfg["species"].keys())
    dbs = []
    nl = int(cfg["nlines"] * scale)
    for mol, (niso, frac) in cfg["molecules"].items():
        db = synth.make_database(mol, niso)
        synth.make_lines(db, max(int(nl * frac), niso), wn_i + 0.3, wn_i + (nwn - 1) * d - 0.3, rng,
                         bands=8, flat=(name != "C1"))
        dbs.append(db)
    wl, el, gf, iso, iso_mass, iso_ratio, iso_imol, iso_mol, iso_k = [], [], [], [], [], [], [], [], []
    base = 0
    for db in dbs:
        for k in range(len(db.iso_names)):
            wl.append(db.wl[k]); el.append(db.elow[k]); gf.append(db.gf[k])
            iso.append(np.full(len(db.wl[k]), base + k, np.int16))
            iso_mass.append(db.iso_mass[k]); iso_ratio.append(db.iso_ratio[k])
            iso_imol.append(names.index(db.molecule)); iso_mol.append(db.molecule); iso_k.append(k)
        base += len(db.iso_names)
    atm = synth.make_atmosphere(cfg["nlayers"], cfg["species"], t_top=cfg["t_top"], t_bot=cfg["t_bot"], rng=rng)
    T = np.round(atm["T"], 2)                      # what a %.2f atmosphere file carries
    p_bar = np.array([float(f"{x:.4e}") for x in atm["p_bar"]])
    rad_km = np.round(atm["radius_km"], 4)
    q = np.stack([np.array([float(f"{x:.4e}") for x in atm["q"][s]]) for s in names])
    mol_mass = np.array([MOLDAT[s][1] for s in names])
    # stateeqnford (transit.h:58-69), number abundances
    dens = np.stack([AMU * q[j] * (p_bar * 1e6) / KB / T * mol_mass[j] for j in range(len(names))])
    Z = np.stack([partition_function(iso_mol[i], iso_k[i], T) for i in range(len(iso_mass))])
    cia = []
    cwn = np.arange(20.0, 34000.1, 20.0)
    ctemps = np.array([200., 400., 600., 800., 1000., 1500., 2000., 3000.])
    for kk, (a, b) in enumerate(cfg["cia"]):
        rngc = np.random.default_rng(kk + 1)
        ph = rngc.uniform(0, 2 * np.pi, 3)
        w = cwn[:, None]; t = ctemps[None, :]
        tab = (1e-7 / (kk + 1)) * (1.0 + 0.5 * np.sin(w / 900.0 + ph[0]) + 0.3 * np.cos(w / 3100.0 + ph[1])) \
            * np.exp(-((w - 4000.0) / 9000.0) ** 2) * (t / 1000.0) ** 0.5 * (1 + 0.1 * np.sin(t / 700.0 + ph[2]))
        tab = np.array([[float(f"{v:.4e}") for v in row] for row in tab])
        cia.append(dict(wn=cwn.copy(), temp=ctemps.copy(), tab=tab, mol=[names.index(a), names.index(b)], names=(a, b)))
    return Problem(
        name=name, cfg=cfg, dbs=dbs, species_names=names,
        wn_i=wn_i, wn_d=d, wnosamp=o, nwn=nwn, onwn=onwn, wn_fct=1.0, wn=wn_i + np.arange(nwn) * d,
        mol_mass=mol_mass, mol_radius=np.array([MOLDAT[s][2] for s in names]) / 2.0 * 1e-8,   # readatm.c:686,706
        mol_pol=np.array([MOLDAT[s][3] for s in names]), mol_ID=np.array([MOLDAT[s][0] for s in names], np.int32),
        iso_mass=np.array(iso_mass), iso_ratio=np.array(iso_ratio), iso_imol=np.array(iso_imol, np.int32),
        iso_z=Z, mol_d=dens, mol_q=q, nmol=len(names), niso=len(iso_mass), nrad=cfg["nlayers"],
        wl=np.concatenate(wl), elow=np.concatenate(el), gf=np.concatenate(gf), isoid=np.concatenate(iso),
        wfct=1e-4, efct=1.0, rad=rad_km, rad_fct=1e5, p=p_bar, p_fct=1e6, t=T, t_fct=1.0,
        ethresh=cfg["ethresh"], toomuch=cfg["toomuch"], nwidth=20.0, eclipse=cfg["eclipse"],
        angles=np.array(cfg.get("angles", [0.0]), float), starrad=cfg["starrad"] * 6.957e10, transp=0, cia=cia,
        atm=atm)


def perturbed(prob, seed, dT=60.0, dex=0.3):
    """Another atmosphere on the same static problem (what BART feeds run_transit): a smooth
    temperature offset of at most ``dT`` K and log-abundance offsets of at most ``dex`` for the
    line-list molecules; densities and partition functions follow (stateeqnford, Z(T))."""
    import copy
    rng = np.random.default_rng(seed)
    p = copy.copy(prob)
    x = np.linspace(0.0, 1.0, prob.nrad)
    T = prob.t + dT * (rng.uniform(-1, 1) * (1 - x) + rng.uniform(-1, 1) * x) * 0.5
    T = np.round(np.clip(T, 250.0, 2900.0), 2)
    q = prob.mol_q.copy()
    for j, s in enumerate(prob.species_names):
        if s in prob.cfg["molecules"]:
            q[j] = q[j] * 10.0 ** rng.uniform(-dex, dex)
    p.t, p.mol_q = T, q
    p.mol_d = np.stack([AMU * q[j] * (prob.p * 1e6) / KB / T * prob.mol_mass[j] for j in range(prob.nmol)])
    iso_mol, iso_k, k = [], [], 0
    for db in prob.dbs:
        for kk in range(len(db.iso_names)):
            iso_mol.append(db.molecule); iso_k.append(kk)
    p.iso_z = np.stack([partition_function(iso_mol[i], iso_k[i], T) for i in range(prob.niso)])
    return p


def write_reference_inputs(prob, outdir, line_fraction=1.0, molfile=None, band=None, tag="run"):
    """The same problem as files for the unmodified reference binary (bench cpu_baseline)."""
    os.makedirs(outdir, exist_ok=True)
    if molfile is None:
        # our own molecules.dat in the layout getmoldata parses (readatm.c:625-717):
        # ID name mass diameter(A) radius-source polarizability description
        molfile = os.path.join(outdir, "molecules.dat")
        with open(molfile, "w") as f:
            f.write("# synthetic molecular info (transit_b200.workloads.MOLDAT)\n")
            for nm, (mid, mass, diam, pol) in MOLDAT.items():
                f.write(f" {mid:3d}    {nm:8s} {mass:10.6f}   {diam:5.2f}      01   {pol:8.3f}   {nm}\n")
    dbs = prob.dbs
    if line_fraction < 1.0:
        import copy
        dbs = copy.deepcopy(prob.dbs)
        for db in dbs:
            for k in range(len(db.wl)):
                n = max(1, int(len(db.wl[k]) * line_fraction))
                idx = np.sort(np.random.default_rng(7).choice(len(db.wl[k]), n, replace=False))
                db.wl[k], db.elow[k], db.gf[k] = db.wl[k][idx], db.elow[k][idx], db.gf[k][idx]
    wn_f = wn_full = prob.wn_i + (prob.nwn - 1) * prob.wn_d
    wn_lo = prob.wn_i
    if band is not None:          # a contiguous band of the coarse grid [j0, j1] (inclusive ends)
        wn_lo, wn_f = prob.wn_i + band[0] * prob.wn_d, prob.wn_i + band[1] * prob.wn_d
    n = synth.write_tli(os.path.join(outdir, "lines.tli"), dbs, 1e4 / (wn_full + 50), 1e4 / max(prob.wn_i - 50, 1.0))
    atm = dict(radius_km=prob.rad, p_bar=prob.p, T=prob.t, q={s: prob.mol_q[j] for j, s in enumerate(prob.species_names)})
    synth.write_atmosphere(os.path.join(outdir, "atm.dat"), atm)
    cs = []
    for k, tb in enumerate(prob.cia):
        fn = os.path.join(outdir, f"cia{k}.dat")
        with open(fn, "w") as f:
            f.write(f"# synthetic CIA table\n\ni {tb['names'][0]} {tb['names'][1]}\nt " +
                    " ".join(f"{x:10.1f}" for x in tb["temp"]) + "\n\n")
            for i in range(len(tb["wn"])):
                f.write(f" {tb['wn'][i]:9.2f}    " + " ".join(f"{v:.4e}" for v in tb["tab"][i]) + "\n")
        cs.append(fn)
    cfg = dict(atm=os.path.join(outdir, "atm.dat"), linedb=os.path.join(outdir, "lines.tli"), molfile=molfile,
               csfile=",".join(cs), wnlow=wn_lo, wnhigh=wn_f, wndelt=prob.wn_d, wnosamp=prob.wnosamp,
               wnfct=1.0, wlfct=1e-4, solution="eclipse" if prob.eclipse else "transit",
               toomuch=prob.toomuch, ethreshold=prob.ethresh, starrad=prob.cfg["starrad"], nwidth=20, verb=0,
               outspec=os.path.join(outdir, f"spec_{tag}.dat"))
    if prob.eclipse:
        cfg["raygrid"] = " ".join(str(int(a)) for a in prob.angles)
    synth.write_cfg(os.path.join(outdir, f"{tag}.cfg"), **cfg)
    return os.path.join(outdir, f"{tag}.cfg"), n
