"""ctypes bindings for oracle/liboracle.so -- the CPU checker (test infrastructure only).

Nothing under transit_b200/ imports this module; it is used by tests/, by
__graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None

f64p = C.POINTER(C.c_double)
i64p = C.POINTER(C.c_int64)
i32p = C.POINTER(C.c_int32)
i16p = C.POINTER(C.c_int16)
i8p = C.POINTER(C.c_int8)
f32p = C.POINTER(C.c_float)


class Grid(C.Structure):
    _fields_ = [("wn_i", C.c_double), ("wn_d", C.c_double), ("o", C.c_int),
                ("nwn", C.c_int64), ("onwn", C.c_int64)]


class Species(C.Structure):
    _fields_ = [("nmol", C.c_int), ("mass", f64p), ("radius", f64p), ("ID", i32p),
                ("niso", C.c_int), ("iso_mass", f64p), ("iso_ratio", f64p), ("iso_imol", i32p)]


class Lines(C.Structure):
    _fields_ = [("n", C.c_int64), ("wl", f64p), ("elow", f64p), ("gf", f64p), ("isoid", i16p),
                ("wfct", C.c_double), ("efct", C.c_double)]


class VoigtTable(C.Structure):
    _fields_ = [("nDop", C.c_int), ("nLor", C.c_int), ("aDop", f64p), ("aLor", f64p),
                ("psize", i64p), ("off", i64p), ("data", f32p), ("ndata", C.c_int64)]


class Trace(C.Structure):
    _fields_ = [("iown", i32p), ("idwn", i32p), ("idop", i32p), ("status", i8p),
                ("nsamples", C.c_int64), ("nadd", C.c_int64), ("nskip", C.c_int64),
                ("neval", C.c_int64)]


class Contrib(C.Structure):
    _fields_ = [("sc_flag", C.c_int), ("sc_logext", C.c_double), ("cl_flag", C.c_int),
                ("cl_ext", C.c_double), ("cl_top", C.c_double), ("cl_bot", C.c_double),
                ("cl_gamma", C.c_double), ("cl_Q", C.c_double), ("cl_r", C.c_double),
                ("cl_sig", C.c_double), ("cl_refwn", C.c_double), ("nH", f64p),
                ("mean_dens", f64p), ("nmol", C.c_int), ("pol", f64p), ("mol_d", f64p),
                ("mol_mass", f64p)]


def lib():
    """Load (building if needed) oracle/liboracle.so."""
    global _LIB
    if _LIB is None:
        so = os.path.join(ROOT, "oracle", "liboracle.so")
        src = os.path.join(ROOT, "oracle", "transit_oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True,
                           stdout=subprocess.DEVNULL)
        L = C.CDLL(so)
        L.to_voigt_build.restype = C.POINTER(VoigtTable)
        L.to_voigt_build.argtypes = [C.POINTER(Grid), C.c_int, C.c_int, C.c_double, C.c_double,
                                     C.c_double, C.c_double, C.c_float]
        L.to_voigt_free.argtypes = [C.POINTER(VoigtTable)]
        L.to_voigt_profile.restype = C.c_int64
        L.to_voigt_profile.argtypes = [f32p, C.c_double, C.c_float, C.c_float, C.c_float, C.c_int64]
        L.to_voigt_psize.restype = C.c_int64
        L.to_voigt_psize.argtypes = [C.c_double, C.c_float, C.c_float, C.c_float, C.c_int64]
        L.to_grid_count.restype = C.c_int64
        L.to_grid_count.argtypes = [C.c_double, C.c_double, C.c_double, C.c_int]
        L.to_computemolext.argtypes = [C.POINTER(Grid), C.POINTER(Lines), C.POINTER(Species),
                                       C.POINTER(VoigtTable), C.c_double, f64p, f64p, C.c_double,
                                       C.c_int, C.c_int, i32p, f64p, C.POINTER(Trace)]
        L.to_interpolmolext.argtypes = [f64p, C.c_int64, C.c_int64, C.c_int64, f64p, C.c_int64,
                                        C.c_double, f64p, f64p]
        L.to_interpcs_table.argtypes = [f64p, f64p, C.c_int, f64p, C.c_int, f64p, C.c_int64, f64p,
                                        C.c_int, f64p, f64p]
        L.to_tau.argtypes = [C.c_int, f64p, f64p, f64p, C.c_int64, C.c_double, f64p, C.c_int,
                             C.c_double, f64p, C.c_int, C.c_double, f64p, f64p,
                             C.POINTER(Contrib), C.c_double, C.c_int, f64p, i64p, i16p]
        L.to_totaltau1.restype = C.c_double
        L.to_totaltau1.argtypes = [C.c_double, f64p, C.c_double, f64p, C.c_long]
        L.to_eclipsetau.restype = C.c_double
        L.to_eclipsetau.argtypes = [C.c_double, f64p, f64p, C.c_long]
        L.to_modulationm1.restype = C.c_double
        L.to_modulationm1.argtypes = [f64p, C.c_long, C.c_double, f64p, C.c_long, C.c_double, C.c_double]
        L.to_modulation1.restype = C.c_double
        L.to_modulation1.argtypes = [f64p, C.c_long, C.c_double, f64p, C.c_long, C.c_double,
                                     C.c_double, C.c_int]
        L.to_eclipse_intens.restype = C.c_double
        L.to_eclipse_intens.argtypes = [f64p, C.c_double, C.c_double, C.c_long, f64p, C.c_long,
                                        C.c_double]
        L.to_flux.argtypes = [f64p, C.c_int, f64p, C.c_int64, f64p]
        L.to_spline_init.argtypes = [f64p, f64p, f64p, C.c_long]
        L.to_splinterp_pt.restype = C.c_double
        L.to_splinterp_pt.argtypes = [f64p, C.c_long, f64p, f64p, C.c_double]
        L.to_splinterp.argtypes = [C.c_long, f64p, f64p, C.c_long, f64p, f64p]
        L.to_radpress.argtypes = [C.c_double, C.c_double, C.c_double, f64p, f64p, f64p, f64p,
                                  C.c_int, C.c_double]
        L.to_binsearchapprox.argtypes = [f64p, C.c_double, C.c_int, C.c_int]
        _LIB = L
    return _LIB


def p(a, t=f64p):
    return a.ctypes.data_as(t)


def c64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class Problem:
    """Plain-array view of one configured problem (what the C-ABI and the oracle consume).

    Built either from a golden fixture (``from_golden``) or synthetically by bench/tests.
    """

    def __init__(self, **kw):
        self.__dict__.update(kw)
        self._keep = []

    # -- construction ---------------------------------------------------------------
    @classmethod
    def from_golden(cls, d, tli_lines=None):
        g = lambda k: d[k]
        s = lambda k: d[k].item()
        kw = dict(
            wn_i=s("wns_i"), wn_d=s("wns_d"), wnosamp=int(s("owns_o")), nwn=int(s("wns_n")),
            onwn=int(s("owns_n")), wn=c64(g("wns_v")), wn_fct=s("wns_fct"),
            mol_mass=c64(g("mol_mass")), mol_radius=c64(g("mol_radius")), mol_pol=c64(g("mol_pol")),
            mol_ID=np.ascontiguousarray(g("mol_ID"), np.int32),
            iso_mass=c64(g("iso_mass")), iso_ratio=c64(g("iso_ratio")),
            iso_imol=np.ascontiguousarray(g("iso_imol"), np.int32),
            iso_z=c64(g("iso_z")), mol_d=c64(g("mol_d")), mol_q=c64(g("mol_q")),
            rad=c64(g("rad_v")), rad_fct=s("rad_fct"), p=c64(g("atm_p")), p_fct=s("atm_pfct"),
            t=c64(g("atm_t")), t_fct=s("atm_tfct"), ethresh=s("ethresh"), toomuch=s("toomuch"),
            nwidth=float(s("timesalpha")), eclipse=bool(s("is_eclipse")),
            starrad=s("starrad_cgs"), transp=int(s("transpplanet")),
        )
        if "line_wl" in d:
            kw.update(wl=c64(g("line_wl")), elow=c64(g("line_elow")), gf=c64(g("line_gf")),
                      isoid=np.ascontiguousarray(g("line_isoid"), np.int16),
                      wfct=s("line_wfct"), efct=s("line_efct"))
        elif tli_lines is not None:
            kw.update(tli_lines)
        if "angles" in d:
            kw["angles"] = c64(g("angles"))
        kw["cia"] = [dict(wn=c64(d[f"cs{i}_wn"]), temp=c64(d[f"cs{i}_temp"]),
                          tab=c64(d[f"cs{i}_tab"]), mol=[int(x) for x in d[f"cs{i}_mol"]])
                     for i in range(int(s("cs_nfiles")))]
        kw["nmol"] = len(kw["mol_mass"]); kw["niso"] = len(kw["iso_mass"]); kw["nrad"] = len(kw["rad"])
        return cls(**kw)

    # -- ctypes views ---------------------------------------------------------------
    def grid(self):
        return Grid(self.wn_i, self.wn_d, self.wnosamp, self.nwn, self.onwn)

    def species(self):
        return Species(self.nmol, p(self.mol_mass), p(self.mol_radius), p(self.mol_ID, i32p),
                       self.niso, p(self.iso_mass), p(self.iso_ratio), p(self.iso_imol, i32p))

    def lines(self):
        return Lines(len(self.wl), p(self.wl), p(self.elow), p(self.gf), p(self.isoid, i16p),
                     self.wfct, self.efct)

    def cia_densfac(self, k, mol_d=None):
        """prod_k rho_k/(AMU m_k AMAGAT), crosssec.c:321-326"""
        AMU, AMAGAT = 1.66053886e-24, 2.68678e19
        mol_d = self.mol_d if mol_d is None else mol_d
        f = np.ones(self.nrad)
        for im in self.cia[k]["mol"]:
            f = f * (mol_d[im] / (AMU * self.mol_mass[im] * AMAGAT))
        return f

    def grid_mol_rows(self, molID):
        """iso -> row of a per-molecule output (valueinarray(op->molID, mol->ID[imol]))"""
        molID = list(molID)
        return np.array([molID.index(self.mol_ID[self.iso_imol[i]]) for i in range(self.niso)],
                        np.int32)

    def line_molIDs(self):
        """opacity.c:349-361: molecule IDs in first-appearance order of the isotopes"""
        out = []
        for i in range(self.niso):
            mid = int(self.mol_ID[self.iso_imol[i]])
            if mid not in out:
                out.append(mid)
        return np.array(out, np.int32)


class OracleVoigt:
    def __init__(self, prob, nDop, nLor, dmin, dmax, lmin, lmax):
        L = lib()
        self.g = prob.grid()
        self.ptr = L.to_voigt_build(C.byref(self.g), nDop, nLor, dmin, dmax, lmin, lmax,
                                    C.c_float(prob.nwidth))
        t = self.ptr.contents
        self.nDop, self.nLor = nDop, nLor
        self.aDop = np.ctypeslib.as_array(t.aDop, (nDop,)).copy()
        self.aLor = np.ctypeslib.as_array(t.aLor, (nLor,)).copy()
        self.psize = np.ctypeslib.as_array(t.psize, (nDop, nLor)).copy()
        self.off = np.ctypeslib.as_array(t.off, (nDop, nLor)).copy()
        self.data = np.ctypeslib.as_array(t.data, (t.ndata,)).copy()

    def __del__(self):
        try:
            lib().to_voigt_free(self.ptr)
        except Exception:
            pass


def voigt_from_arrays(aDop, aLor, psize, off, data):
    """VoigtTable struct over numpy arrays (e.g. the reference's own table from a fixture)."""
    aDop, aLor = c64(aDop), c64(aLor)
    psize = np.ascontiguousarray(psize, np.int64); off = np.ascontiguousarray(off, np.int64)
    data = np.ascontiguousarray(data, np.float32)
    vt = VoigtTable(len(aDop), len(aLor), p(aDop), p(aLor), p(psize, i64p), p(off, i64p),
                    p(data, f32p), data.size)
    vt._keep = (aDop, aLor, psize, off, data)
    return vt


def computemolext(prob, vt, temp, density, Z, ethresh, permol=0, molID=None, trace=False):
    L = lib()
    g, ln, sp = prob.grid(), prob.lines(), prob.species()
    density, Z = c64(density), c64(Z)
    if permol:
        rows = prob.grid_mol_rows(molID); nout = len(molID)
    else:
        rows = np.zeros(prob.niso, np.int32); nout = 1
    out = np.zeros((nout, prob.nwn))
    tr = None
    if trace:
        n = len(prob.wl)
        arrs = dict(iown=np.zeros(n, np.int32), idwn=np.zeros(n, np.int32),
                    idop=np.zeros(n, np.int32), status=np.zeros(n, np.int8))
        tr = Trace(p(arrs["iown"], i32p), p(arrs["idwn"], i32p), p(arrs["idop"], i32p),
                   p(arrs["status"], i8p), 0, 0, 0, 0)
    vtp = vt.ptr if isinstance(vt, OracleVoigt) else C.pointer(vt)
    L.to_computemolext(C.byref(g), C.byref(ln), C.byref(sp), vtp, temp, p(density), p(Z),
                       ethresh, permol, nout, p(rows, i32p), p(out),
                       C.byref(tr) if tr else None)
    if trace:
        arrs.update(nsamples=tr.nsamples, nadd=tr.nadd, nskip=tr.nskip, neval=tr.neval)
        return out, arrs
    return out


def interpcs(prob, temp=None, mol_d=None):
    L = lib()
    temp = c64(prob.t * prob.t_fct if temp is None else temp)
    wn = c64(prob.wn * prob.wn_fct)
    e_cs = np.zeros((prob.nwn, prob.nrad))
    for k, tb in enumerate(prob.cia):
        df = c64(prob.cia_densfac(k, mol_d))
        L.to_interpcs_table(p(tb["tab"]), p(tb["wn"]), len(tb["wn"]), p(tb["temp"]),
                            len(tb["temp"]), p(wn), prob.nwn, p(temp), prob.nrad, p(df), p(e_cs))
    return e_cs


def cloud_aux(prob, mol_names):
    """tau.c:194-214: mean_dens[r] = (sum_j d_j/m_j*q_j) * (sum_j m_j*q_j) and, for the species named "H2",
    nH[r] = d/m*q*N_A (the reference accumulates mean_dens into an uninitialised stack array; from zero here)."""
    NAVOGADRO = 6.02214076e23
    md = np.zeros(prob.nrad); nH = np.zeros(prob.nrad); mm = np.zeros(prob.nrad)
    for j, nm in enumerate(mol_names):
        md = md + prob.mol_d[j] / prob.mol_mass[j] * prob.mol_q[j]
        if nm == "H2":
            nH = prob.mol_d[j] / prob.mol_mass[j] * prob.mol_q[j] * NAVOGADRO
        mm = mm + prob.mol_mass[j] * prob.mol_q[j]
    return nH, md * mm


def contrib(prob, sc_flag=0, sc_logext=0.0, cl_flag=0, cl_ext=0.0, cl_top=0.0, cl_bot=0.0,
            cl_gamma=0.0, cl_Q=0.0, cl_r=0.0, cl_sig=0.0, cl_refwn=0.0, nH=None, md=None):
    nH = np.zeros(prob.nrad) if nH is None else c64(nH).copy()
    md = np.zeros(prob.nrad) if md is None else c64(md).copy()
    cb = Contrib(sc_flag, sc_logext, cl_flag, cl_ext, cl_top, cl_bot, cl_gamma, cl_Q, cl_r,
                 cl_sig, cl_refwn, p(nH), p(md), prob.nmol, p(prob.mol_pol), p(prob.mol_d),
                 p(prob.mol_mass))
    cb._keep = (nH, md)
    return cb


def tau(prob, e, e_cs, cb=None, lazy=True, rad=None, toomuch=None, h=None):
    L = lib()
    cb = cb or contrib(prob)
    toomuch = prob.toomuch if toomuch is None else toomuch
    rad = c64(prob.rad if rad is None else rad).copy()
    nrad = prob.nrad
    h = c64(rad[::-1]) if h is None else c64(h)   # makeipsample raddelt=-1 / tau.c:92-98, or an explicit sample
    nh = len(h)
    t = np.zeros((prob.nwn, nh)); last = np.zeros(prob.nwn, np.int64)
    comp = np.zeros(nrad, np.int16)
    e = c64(e); e_cs = c64(e_cs)
    L.to_tau(int(prob.eclipse), p(e), p(e_cs), p(prob.wn), prob.nwn, prob.wn_fct, p(rad), nrad,
             prob.rad_fct, p(h), nh, prob.rad_fct, p(prob.p), p(prob.t), C.byref(cb),
             toomuch, int(lazy), p(t), p(last, i64p), p(comp, i16p))
    return t, last, comp


def modulation(prob, t, last, rad=None, transp=None, modlevel=1, toomuch=None, ip=None):
    L = lib()
    rad = prob.rad if rad is None else rad
    ip = c64(rad[::-1]) if ip is None else c64(ip)
    out = np.zeros(prob.nwn)
    transp = prob.transp if transp is None else transp
    toomuch = prob.toomuch if toomuch is None else toomuch
    for w in range(prob.nwn):
        row = c64(t[w])
        if modlevel == -1:
            out[w] = L.to_modulationm1(p(row), int(last[w]), toomuch, p(ip), len(ip), prob.rad_fct, prob.starrad)
        else:
            out[w] = L.to_modulation1(p(row), int(last[w]), toomuch, p(ip), len(ip),
                                      prob.rad_fct, prob.starrad, int(transp))
    return out


def intensity_flux(prob, t, last, temp=None):
    L = lib()
    temp = c64(prob.t if temp is None else temp)
    nang = len(prob.angles)
    intens = np.zeros((nang, prob.nwn))
    for a in range(nang):
        for w in range(prob.nwn):
            row = c64(t[w])
            intens[a, w] = L.to_eclipse_intens(p(row), prob.wn[w], prob.wn_fct, int(last[w]),
                                               p(temp), prob.nrad, prob.angles[a])
    out = np.zeros(prob.nwn)
    L.to_flux(p(intens), nang, p(prob.angles), prob.nwn, p(out))
    return intens, out


def parse_tli(raw: bytes, wn_lo: float, wn_hi: float):
    """Minimal TLI v6 reader (readlineinfo.c:87-244,416-537): per isotope, keeps the lines
    with wavelength inside [1/wn_hi, 1/wn_lo] exactly as datafileBS + fread do."""
    import struct
    pos = 4 + 6
    wi, wf = struct.unpack_from("<2d", raw, pos); pos += 16
    (ndb,) = struct.unpack_from("<H", raw, pos); pos += 2
    dbs = []
    for _ in range(ndb):
        (n,) = struct.unpack_from("<H", raw, pos); pos += 2 + n
        (n,) = struct.unpack_from("<H", raw, pos); pos += 2
        mol = raw[pos:pos + n].decode(); pos += n
        nT, nI = struct.unpack_from("<2H", raw, pos); pos += 4
        T = np.frombuffer(raw, "<f8", nT, pos); pos += 8 * nT
        isos = []
        for _ in range(nI):
            (n,) = struct.unpack_from("<H", raw, pos); pos += 2 + n
            m, r = struct.unpack_from("<2d", raw, pos); pos += 16
            Z = np.frombuffer(raw, "<f8", nT, pos); pos += 8 * nT
            isos.append((m, r, Z))
        dbs.append((mol, T, isos))
    (nl,) = struct.unpack_from("<Q", raw, pos); pos += 8
    (ni,) = struct.unpack_from("<i", raw, pos); pos += 4
    cnt = np.frombuffer(raw, "<u8", ni, pos).astype(np.int64); pos += 8 * ni
    wl = np.frombuffer(raw, "<f8", nl, pos); pos += 8 * nl
    iso = np.frombuffer(raw, "<i2", nl, pos); pos += 2 * nl
    el = np.frombuffer(raw, "<f8", nl, pos); pos += 8 * nl
    gf = np.frombuffer(raw, "<f8", nl, pos)
    iniw, finw = 1.0 / (wn_hi * 1.0) / 1e-4, 1.0 / (wn_lo * 1.0) / 1e-4
    sel = []
    start = 0
    for c in cnt:
        seg = wl[start:start + c]
        a = int(np.searchsorted(seg, iniw, "left")); b = int(np.searchsorted(seg, finw, "right"))
        sel.append(np.arange(start + a, start + b))
        start += c
    sel = np.concatenate(sel)
    return dict(wl=c64(wl[sel]), elow=c64(el[sel]), gf=c64(gf[sel]),
                isoid=np.ascontiguousarray(iso[sel], np.int16), wfct=1e-4, efct=1.0), dbs


def cloud_variant(d, name):
    """Parameters of one variant of tests/golden/tiny_cloud.npz as keyword arguments of contrib()."""
    g = lambda k: d[f"{name}__{k}"].item()
    return dict(sc_flag=int(g("sc_flag")), sc_logext=g("sc_logext"), cl_flag=int(g("cl_flag")), cl_ext=g("cl_ext"),
                cl_top=g("cl_top"), cl_bot=g("cl_bot"), cl_gamma=g("cl_gamma"), cl_Q=g("cl_Q"), cl_r=g("cl_r"),
                cl_sig=g("cl_sig"), cl_refwn=g("cl_refwn"))


def golden_mol_names(d):
    return d["mol_names"].tobytes().decode().strip("\x00").split()
