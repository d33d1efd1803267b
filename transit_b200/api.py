"""Host-side mirror of the reference's library interface over the C ABI (include/transit_b200.h).

``Transit`` follows the call sequence BART uses on the reference's SWIG module
(transit/src/transit.i:11-19,95-104; transit/src/transit.c:25-228):

    transit_init  -> Transit.init(problem)           (grids, species, lines, Voigt table, CIA)
    run_transit   -> Transit.run_transit(atmosphere) (one spectrum: CIA + extinction + tau + spectrum)
    get_no_samples / get_waveno_arr / set_radius / set_cloudtop / set_scattering / free_memory

There is NO CPU path here: every call goes to libtransit_b200.so (hand-written sm_100a
kernels) and raises ``TransitError`` when the library or a GPU is missing.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "transit_b200", "libtransit_b200.so")
GEOM_TRANSIT, GEOM_ECLIPSE = 0, 1
AMU, KB = 1.66053886e-24, 1.380658e-16          # constants_tr.h:21,25 (cgs)

f64p = C.POINTER(C.c_double)
f32p = C.POINTER(C.c_float)
i64p = C.POINTER(C.c_int64)
i32p = C.POINTER(C.c_int32)
i16p = C.POINTER(C.c_int16)
i8p = C.POINTER(C.c_int8)

# every symbol include/transit_b200.h declares: (restype, argtypes)
ABI = {
    "tb_last_error": (C.c_char_p, []),
    "tb_version": (C.c_int, []),
    "tb_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "tb_destroy": (None, [C.c_void_p]),
    "tb_sync": (C.c_int, [C.c_void_p]),
    "tb_grid_count": (C.c_int64, [C.c_double, C.c_double, C.c_double, C.c_int]),
    "tb_set_wn_grid": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_int, C.c_int64, C.c_int64, C.c_double]),
    "tb_set_species": (C.c_int, [C.c_void_p, C.c_int, f64p, f64p, f64p, i32p, C.c_int, f64p, f64p, i32p]),
    "tb_upload_lines": (C.c_int, [C.c_void_p, f64p, f64p, f64p, i16p, C.c_int64, C.c_double, C.c_double]),
    "tb_build_profiles": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float]),
    "tb_get_profile_meta": (C.c_int, [C.c_void_p, f64p, f64p, i64p, i64p, i64p]),
    "tb_get_profile_data": (C.c_int, [C.c_void_p, f32p, C.c_int64]),
    "tb_set_atmosphere": (C.c_int, [C.c_void_p, C.c_int, f64p, C.c_double, f64p, C.c_double, f64p, C.c_double, f64p, f64p]),
    "tb_extinction_lbl": (C.c_int, [C.c_void_p, C.c_double]),
    "tb_molext_states": (C.c_int, [C.c_void_p, C.c_int, f64p, f64p, f64p, C.c_double, C.c_int, f64p]),
    "tb_get_line_trace": (C.c_int, [C.c_void_p, C.c_int, i32p, i32p, i32p, i8p]),
    "tb_get_sample_counts": (C.c_int, [C.c_void_p, i64p, C.c_int]),
    "tb_get_state_ilor": (C.c_int, [C.c_void_p, C.c_int, i32p]),
    "tb_set_fused": (C.c_int, [C.c_void_p, C.c_int]),
    "tb_set_fused_reach": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "tb_set_narrow_reach": (C.c_int, [C.c_void_p, C.c_int]),
    "tb_set_fused_variant": (C.c_int, [C.c_void_p, C.c_int]),
    "tb_set_rays": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_double]),
    "tb_opacity_share": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tb_tau_transit_static": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_double, C.c_int, C.c_void_p, C.c_double,
                                        C.c_int64, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.c_int]),
    "tb_set_trace": (C.c_int, [C.c_void_p, C.c_int]),
    "tb_get_fused_stats": (C.c_int, [C.c_void_p, i64p, i64p]),
    "tb_get_line_molIDs": (C.c_int, [C.c_void_p, i32p, i32p]),
    "tb_get_line_counts": (C.c_int, [C.c_void_p, i64p, i64p, i64p]),
    "tb_opacity_build": (C.c_int, [C.c_void_p, C.c_int, f64p, f64p, f64p, f64p, C.c_int, C.c_double, f64p]),
    "tb_opacity_load": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, i32p, f64p, f64p, f64p]),
    "tb_extinction_interp": (C.c_int, [C.c_void_p]),
    "tb_cia_clear": (C.c_int, [C.c_void_p]),
    "tb_cia_add_table": (C.c_int, [C.c_void_p, f64p, C.c_int, f64p, C.c_int, f64p, i32p, C.c_int]),
    "tb_cia_interp": (C.c_int, [C.c_void_p]),
    "tb_set_scattering": (C.c_int, [C.c_void_p, C.c_int, C.c_double]),
    "tb_set_cloud": (C.c_int, [C.c_void_p, C.c_int] + [C.c_double] * 8 + [f64p, f64p]),
    "tb_tau": (C.c_int, [C.c_void_p, C.c_int, C.c_double, f64p, C.c_int, C.c_double]),
    "tb_set_modlevel": (C.c_int, [C.c_void_p, C.c_int]),
    "tb_set_tau_streaming": (C.c_int, [C.c_void_p, C.c_int]),
    "tb_modulation": (C.c_int, [C.c_void_p, C.c_double, C.c_int, f64p]),
    "tb_intensity_flux": (C.c_int, [C.c_void_p, f64p, C.c_int, f64p]),
    "tb_spectrum": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, f64p, C.c_int, f64p]),
    "tb_spectrum_batch": (C.c_int, [C.c_void_p, C.c_int, C.c_int, f64p, C.c_double, f64p, C.c_double, f64p, C.c_double,
                                    f64p, f64p, C.c_int, C.c_double, C.c_double, C.c_double, C.c_int, f64p, C.c_int, f64p]),
    "tb_set_atmosphere_model": (C.c_int, [C.c_void_p, C.c_int, f64p, C.c_double, C.c_double, C.c_double, C.c_int,
                                          C.c_double, C.c_double, C.c_double]),
    "tb_run_transit_batch": (C.c_int, [C.c_void_p, C.c_int, f64p, C.c_double, C.c_double, C.c_int, f64p]),
    "tb_get_extinction": (C.c_int, [C.c_void_p, f64p]),
    "tb_get_cia": (C.c_int, [C.c_void_p, f64p]),
    "tb_get_tau": (C.c_int, [C.c_void_p, f64p, i64p]),
    "tb_get_intensity": (C.c_int, [C.c_void_p, f64p]),
    "tb_get_computed": (C.c_int, [C.c_void_p, i16p]),
    "tb_get_kernel_ms": (C.c_int, [C.c_void_p, C.c_char_p, f32p]),
    "tb_timer_start": (C.c_int, [C.c_void_p, C.c_char_p]),
    "tb_timer_stop": (C.c_int, [C.c_void_p, C.c_char_p]),
    "tb_launch_count": (C.c_int64, [C.c_void_p]),
    "tb_stream": (C.c_void_p, [C.c_void_p]),
    "tb_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "tb_device_spectrum": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), i64p]),
    "tb_microbench": (C.c_int, [C.c_void_p, f64p, f64p, f64p]),
    "tb_comm_unique_id": (C.c_int, [C.c_char_p]),
    "tb_comm_init": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_char_p]),
    "tb_comm_destroy": (C.c_int, [C.c_void_p]),
    "tb_comm_barrier": (C.c_int, [C.c_void_p]),
    "tb_balanced_bands": (C.c_int, [C.c_void_p, C.c_int, i64p]),
    "tb_set_bands": (C.c_int, [C.c_void_p, C.c_int, i64p, C.c_int]),
    "tb_rebalance_bands": (C.c_int, [C.c_void_p]),
    "tb_get_bands": (C.c_int, [C.c_void_p, i64p]),
    "tb_set_shard": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
}


EXCHANGE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int64)


class TransitError(RuntimeError):
    pass


_lib = None


def build_library(force: bool = False) -> str:
    """Compile the CUDA sources for sm_100a into transit_b200/libtransit_b200.so (in-tree)."""
    src = os.path.join(ROOT, "transit_b200", "csrc")
    if force:
        subprocess.run(["make", "-C", src, "clean"], check=True, stdout=subprocess.DEVNULL)
    subprocess.run(["make", "-C", src, "-j8"], check=True, stdout=subprocess.DEVNULL)
    return LIB_PATH


def load_library():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise TransitError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in ABI.items():
            fn = getattr(lib, name)          # AttributeError = header and library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _p(a, t=f64p):
    return None if a is None else a.ctypes.data_as(t)


def _c64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class Context:
    """Thin, typed wrapper of one ``tb_ctx``: one method per ABI entry point."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.tb_create(device, C.byref(h))
        if rc != 0:
            raise TransitError(f"tb_create({device}) failed [{rc}]: {self.lib.tb_last_error().decode()}")
        self.h = h
        self.nwn = self.nrad = self.nang = 0

    @classmethod
    def adopt(cls, handle, nwn, nrad, nmol, niso, nlines=0, nang=0):
        """Wrap a tb_ctx that somebody else owns (the C host's, through th_context()): taps only, never closed here."""
        self = cls.__new__(cls)
        self.lib = load_library()
        self.h = C.c_void_p(handle)
        self.nwn, self.nrad, self.nmol, self.niso, self.nlines, self.nang = nwn, nrad, nmol, niso, nlines, nang
        self.nh = nrad
        self._borrowed = True
        return self

    def close(self):
        if getattr(self, "_borrowed", False):
            self.h = None
            return
        if getattr(self, "h", None):
            self.lib.tb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc != 0:
            raise TransitError(f"{what} failed [{rc}]: {self.lib.tb_last_error().decode()}")

    # -- static problem -----------------------------------------------------------------
    def set_wn_grid(self, wn_i, wn_d, wnosamp, nwn, onwn, wn_fct=1.0):
        self._ck(self.lib.tb_set_wn_grid(self.h, wn_i, wn_d, int(wnosamp), int(nwn), int(onwn), wn_fct), "tb_set_wn_grid")
        self.nwn = int(nwn)
        self.wn = wn_i + np.arange(nwn) * wn_d

    def set_species(self, mol_mass, mol_radius, mol_pol, mol_ID, iso_mass, iso_ratio, iso_imol):
        a = [_c64(mol_mass), _c64(mol_radius), _c64(mol_pol), np.ascontiguousarray(mol_ID, np.int32),
             _c64(iso_mass), _c64(iso_ratio), np.ascontiguousarray(iso_imol, np.int32)]
        self.nmol, self.niso = len(a[0]), len(a[4])
        self._ck(self.lib.tb_set_species(self.h, self.nmol, _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3], i32p), self.niso,
                                         _p(a[4]), _p(a[5]), _p(a[6], i32p)), "tb_set_species")

    def upload_lines(self, wl, elow, gf, isoid, wfct=1e-4, efct=1.0):
        wl, elow, gf = _c64(wl), _c64(elow), _c64(gf)
        isoid = np.ascontiguousarray(isoid, np.int16)
        self.nlines = len(wl)
        self._ck(self.lib.tb_upload_lines(self.h, _p(wl), _p(elow), _p(gf), _p(isoid, i16p), len(wl), wfct, efct),
                 "tb_upload_lines")

    def build_profiles(self, nDop=60, nLor=60, dmin=1e-3, dmax=0.25, lmin=1e-4, lmax=10.0, nwidth=20.0):
        self.nDop, self.nLor = nDop, nLor
        self._ck(self.lib.tb_build_profiles(self.h, nDop, nLor, dmin, dmax, lmin, lmax, nwidth), "tb_build_profiles")

    def get_profiles(self):
        aD, aL = np.zeros(self.nDop), np.zeros(self.nLor)
        ps = np.zeros((self.nDop, self.nLor), np.int64); off = np.zeros_like(ps)
        nd = C.c_int64()
        self._ck(self.lib.tb_get_profile_meta(self.h, _p(aD), _p(aL), _p(ps, i64p), _p(off, i64p), C.byref(nd)),
                 "tb_get_profile_meta")
        data = np.zeros(nd.value, np.float32)
        self._ck(self.lib.tb_get_profile_data(self.h, _p(data, f32p), nd.value), "tb_get_profile_data")
        return dict(aDop=aD, aLor=aL, psize=ps, off=off, data=data)

    def set_atmosphere(self, rad, rad_fct, p, p_fct, t, t_fct, density, Z):
        rad, p, t, density = _c64(rad), _c64(p), _c64(t), _c64(density)
        Z = None if Z is None else _c64(Z)
        self.nrad = len(rad)
        self._ck(self.lib.tb_set_atmosphere(self.h, self.nrad, _p(rad), rad_fct, _p(p), p_fct, _p(t), t_fct,
                                            _p(density), _p(Z)), "tb_set_atmosphere")

    def set_scattering(self, flag, logext):
        self._ck(self.lib.tb_set_scattering(self.h, int(flag), float(logext)), "tb_set_scattering")

    def set_cloud(self, flag, ext=0.0, top=0.0, bot=0.0, gamma=0.0, Q=0.0, r=0.0, sig=0.0, refwn=0.0,
                  nH=None, mean_dens=None):
        nH = None if nH is None else _c64(nH)
        md = None if mean_dens is None else _c64(mean_dens)
        self._ck(self.lib.tb_set_cloud(self.h, int(flag), ext, top, bot, gamma, Q, r, sig, refwn, _p(nH), _p(md)),
                 "tb_set_cloud")

    # -- line sum -----------------------------------------------------------------------
    def extinction_lbl(self, ethresh):
        self._ck(self.lib.tb_extinction_lbl(self.h, ethresh), "tb_extinction_lbl")

    def molext_states(self, T, density, Z, ethresh, permol=0):
        T, density, Z = _c64(T), _c64(density), _c64(Z)
        ns = len(T)
        nout = self.line_molIDs().size if permol else 1
        out = np.zeros((ns, nout, self.nwn))
        self._ck(self.lib.tb_molext_states(self.h, ns, _p(T), _p(density), _p(Z), ethresh, int(permol), _p(out)),
                 "tb_molext_states")
        return out

    def line_trace(self, s=0):
        n = self.nlines
        a = [np.zeros(n, np.int32) for _ in range(3)]
        st = np.zeros(n, np.int8)
        self._ck(self.lib.tb_get_line_trace(self.h, s, _p(a[0], i32p), _p(a[1], i32p), _p(a[2], i32p), _p(st, i8p)),
                 "tb_get_line_trace")
        return dict(iown=a[0], idwn=a[1], idop=a[2], status=st)

    def set_fused(self, on=1):
        """0: two-kernel path; 1/True: fused where the grid has >= 96 tiles; 2: fused on small grids too."""
        self._ck(self.lib.tb_set_fused(self.h, int(on)), "tb_set_fused")

    def tau_transit_static(self, nrad, d_rad, rad_fct, nh, d_h, hfct, nwn, d_e, d_ecs, toomuch, d_tau, d_last, check=True):
        """tb_tau_transit_static: every d_* is a DEVICE address (int) of a buffer the caller owns; nothing is allocated."""
        self._ck(self.lib.tb_tau_transit_static(self.h, int(nrad), C.c_void_p(d_rad), float(rad_fct), int(nh),
                                                C.c_void_p(d_h), float(hfct), int(nwn), C.c_void_p(d_e), C.c_void_p(d_ecs),
                                                float(toomuch), C.c_void_p(d_tau), C.c_void_p(d_last), int(bool(check))),
                 "tb_tau_transit_static")

    def opacity_share(self, src):
        """Read the opacity grid of another context of the same GPU in place (no copy)."""
        self._ck(self.lib.tb_opacity_share(self.h, src.h), "tb_opacity_share")

    def set_rays(self, h, hfct):
        """Impact parameters tb_spectrum integrates along (None: the reversed radii, makeipsample with raddelt -1)."""
        self.user_nh = None if h is None else len(h)
        if h is None:
            self._ck(self.lib.tb_set_rays(self.h, None, 0, 1.0), "tb_set_rays")
            return
        h = np.ascontiguousarray(h, dtype=np.float64)
        self._ck(self.lib.tb_set_rays(self.h, h.ctypes.data_as(C.c_void_p), len(h), float(hfct)), "tb_set_rays")

    def set_fused_variant(self, v):
        self._ck(self.lib.tb_set_fused_variant(self.h, int(v)), "tb_set_fused_variant")

    def set_narrow_reach(self, r):
        self._ck(self.lib.tb_set_narrow_reach(self.h, int(r)), "tb_set_narrow_reach")

    def set_fused_reach(self, rmax, carried=False):
        self._ck(self.lib.tb_set_fused_reach(self.h, int(rmax), int(bool(carried))), "tb_set_fused_reach")

    def set_trace(self, on=True):
        self._ck(self.lib.tb_set_trace(self.h, int(bool(on))), "tb_set_trace")

    def fused_stats(self):
        a, b = C.c_int64(), C.c_int64()
        self._ck(self.lib.tb_get_fused_stats(self.h, C.byref(a), C.byref(b)), "tb_get_fused_stats")
        return dict(fused_states=a.value, fused_samples=b.value)

    def state_ilor(self, s=0):
        out = np.zeros(max(1, self.niso), np.int32)
        self._ck(self.lib.tb_get_state_ilor(self.h, s, _p(out, i32p)), "tb_get_state_ilor")
        return out[:self.niso]

    def sample_counts(self, ns):
        out = np.zeros(ns, np.int64)
        self._ck(self.lib.tb_get_sample_counts(self.h, _p(out, i64p), ns), "tb_get_sample_counts")
        return out

    def line_molIDs(self):
        n = C.c_int32()
        ids = np.zeros(256, np.int32)
        self._ck(self.lib.tb_get_line_molIDs(self.h, _p(ids, i32p), C.byref(n)), "tb_get_line_molIDs")
        return ids[:n.value].copy()

    def line_counts(self):
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        self._ck(self.lib.tb_get_line_counts(self.h, C.byref(a), C.byref(b), C.byref(c)), "tb_get_line_counts")
        return a.value, b.value, c.value

    def ngroups(self):
        return self.line_counts()[1]

    # -- opacity grid -------------------------------------------------------------------
    def opacity_build(self, temp, ziso, q, mm, bymass, ethresh):
        temp, ziso, q, mm = _c64(temp), _c64(ziso), _c64(q), _c64(mm)
        nrows = self.line_molIDs().size
        grid = np.zeros((self.nrad, len(temp), nrows, self.nwn))
        self._ck(self.lib.tb_opacity_build(self.h, len(temp), _p(temp), _p(ziso), _p(q), _p(mm), int(bymass), ethresh,
                                           _p(grid)), "tb_opacity_build")
        return grid

    def opacity_load(self, molID, temp, press, grid):
        grid = _c64(grid); temp = _c64(temp); press = _c64(press)
        molID = np.ascontiguousarray(molID, np.int32)
        Nl, Nt, Nm, Nw = grid.shape
        self._ck(self.lib.tb_opacity_load(self.h, Nm, Nt, Nl, Nw, _p(molID, i32p), _p(temp), _p(press), _p(grid)),
                 "tb_opacity_load")

    def extinction_interp(self):
        self._ck(self.lib.tb_extinction_interp(self.h), "tb_extinction_interp")

    # -- CIA, tau, spectra --------------------------------------------------------------
    def cia_clear(self):
        self._ck(self.lib.tb_cia_clear(self.h), "tb_cia_clear")

    def cia_add_table(self, wn, temp, tab, mol):
        wn, temp, tab = _c64(wn), _c64(temp), _c64(tab)
        mol = np.ascontiguousarray(mol, np.int32)
        self._ck(self.lib.tb_cia_add_table(self.h, _p(wn), len(wn), _p(temp), len(temp), _p(tab), _p(mol, i32p), len(mol)),
                 "tb_cia_add_table")

    def cia_interp(self):
        self._ck(self.lib.tb_cia_interp(self.h), "tb_cia_interp")

    def tau(self, geometry, toomuch, h=None, hfct=1.0):
        hp = None if h is None else _c64(h)
        self._ck(self.lib.tb_tau(self.h, geometry, toomuch, _p(hp), 0 if h is None else len(hp), hfct), "tb_tau")
        self.nh = self.nrad if h is None else len(hp)

    def set_tau_streaming(self, on=True):
        self._ck(self.lib.tb_set_tau_streaming(self.h, int(bool(on))), "tb_set_tau_streaming")

    def set_modlevel(self, level):
        self._ck(self.lib.tb_set_modlevel(self.h, int(level)), "tb_set_modlevel")

    def modulation(self, starrad_cm, transparent=0):
        out = np.zeros(self.nwn)
        self._ck(self.lib.tb_modulation(self.h, starrad_cm, int(transparent), _p(out)), "tb_modulation")
        return out

    def intensity_flux(self, angles_deg):
        a = _c64(angles_deg); out = np.zeros(self.nwn)
        self.nang = len(a)
        self._ck(self.lib.tb_intensity_flux(self.h, _p(a), len(a), _p(out)), "tb_intensity_flux")
        return out

    def spectrum(self, geometry, toomuch, ethresh, starrad_cm=1.0, transparent=0, angles_deg=None, out=None):
        a = None if angles_deg is None else _c64(angles_deg)
        out = np.zeros(self.nwn) if out is None else out
        if a is not None:
            self.nang = len(a)
        self._ck(self.lib.tb_spectrum(self.h, geometry, toomuch, ethresh, starrad_cm, int(transparent), _p(a),
                                      0 if a is None else len(a), _p(out)), "tb_spectrum")
        self.nh = getattr(self, "user_nh", None) or self.nrad
        return out

    # -- taps ---------------------------------------------------------------------------
    def spectrum_batch(self, rad, rad_fct, p, p_fct, t, t_fct, density, Z, geometry, toomuch, ethresh,
                       starrad_cm=1.0, transparent=0, angles_deg=None, out=None):
        """natm atmospheres in one call: rad/p/t [natm][nrad], density [natm][nmol][nrad], Z [natm][niso][nrad]."""
        rad, p, t, density = _c64(rad), _c64(p), _c64(t), _c64(density)
        Z = None if Z is None else _c64(Z)
        natm, nrad = rad.shape
        assert p.shape == (natm, nrad) and t.shape == (natm, nrad) and density.shape == (natm, self.nmol, nrad)
        assert Z is None or Z.shape == (natm, self.niso, nrad)
        a = None if angles_deg is None else _c64(angles_deg)
        out = np.zeros((natm, self.nwn)) if out is None else out
        assert out.shape == (natm, self.nwn) and out.flags.c_contiguous
        if a is not None:
            self.nang = len(a)
        self._ck(self.lib.tb_spectrum_batch(self.h, natm, nrad, _p(rad), rad_fct, _p(p), p_fct, _p(t), t_fct,
                                            _p(density), _p(Z), geometry, toomuch, ethresh, starrad_cm,
                                            int(transparent), _p(a), 0 if a is None else len(a), _p(out)),
                 "tb_spectrum_batch")
        self.nrad = nrad
        self.nh = nrad
        return out

    def set_atmosphere_model(self, p, p_fct, t_fct, rad_fct, bymass, p0, r0, gsurf):
        p = _c64(p)
        self._ck(self.lib.tb_set_atmosphere_model(self.h, len(p), _p(p), p_fct, t_fct, rad_fct, int(bymass), p0, r0, gsurf),
                 "tb_set_atmosphere_model")

    def run_transit_batch(self, inputs, toomuch, starrad_cm, transparent=0, out=None):
        """inputs[natm][nlayers*(1+nmol)] as BART hands them to run_transit; out[natm][nwn]."""
        x = _c64(inputs)
        natm = x.shape[0]
        out = np.zeros((natm, self.nwn)) if out is None else out
        self._ck(self.lib.tb_run_transit_batch(self.h, natm, _p(x), toomuch, starrad_cm, int(transparent), _p(out)),
                 "tb_run_transit_batch")
        return out

    def get_extinction(self):
        e = np.zeros((self.nrad, self.nwn))
        self._ck(self.lib.tb_get_extinction(self.h, _p(e)), "tb_get_extinction")
        return e

    def get_cia(self):
        e = np.zeros((self.nwn, self.nrad))
        self._ck(self.lib.tb_get_cia(self.h, _p(e)), "tb_get_cia")
        return e

    def get_tau(self):
        t = np.zeros((self.nwn, self.nh)); last = np.zeros(self.nwn, np.int64)
        self._ck(self.lib.tb_get_tau(self.h, _p(t), _p(last, i64p)), "tb_get_tau")
        return t, last

    def get_intensity(self):
        a = np.zeros((self.nang, self.nwn))
        self._ck(self.lib.tb_get_intensity(self.h, _p(a)), "tb_get_intensity")
        return a

    def get_computed(self):
        c = np.zeros(self.nrad, np.int16)
        self._ck(self.lib.tb_get_computed(self.h, _p(c, i16p)), "tb_get_computed")
        return c

    def kernel_ms(self, name):
        ms = C.c_float()
        self._ck(self.lib.tb_get_kernel_ms(self.h, name.encode(), C.byref(ms)), "tb_get_kernel_ms")
        return ms.value

    def timer_start(self, name):
        self._ck(self.lib.tb_timer_start(self.h, name.encode()), "tb_timer_start")

    def timer_stop(self, name):
        self._ck(self.lib.tb_timer_stop(self.h, name.encode()), "tb_timer_stop")

    def launch_count(self):
        return int(self.lib.tb_launch_count(self.h))

    def set_stream(self, stream_ptr):
        """Run on a caller-owned cudaStream_t (integer handle).  Handle 0 (torch's DEFAULT stream) means
        "back to the context's own stream" to the library: pass a real stream (sharding.private_stream())."""
        if not stream_ptr:
            raise TransitError("set_stream(0): the default stream cannot be shared with the context; "
                               "use transit_b200.sharding.private_stream()")
        self._ck(self.lib.tb_set_stream(self.h, C.c_void_p(stream_ptr)), "tb_set_stream")

    def device_spectrum(self):
        """(device pointer, n) of the last spectrum, for device-side consumers (NCCL gather)."""
        p, n = C.c_void_p(), C.c_int64()
        self._ck(self.lib.tb_device_spectrum(self.h, C.byref(p), C.byref(n)), "tb_device_spectrum")
        return p.value, n.value

    def set_shard(self, j_lo, j_hi, rank=0, world=1, exchange=None):
        """Wavenumber-band shard: line sum only for coarse bins [j_lo, j_hi).  ``exchange(what, ptr, count)``
        must all-reduce MAX in place on the context's stream (what 0: float64, 1: int32)."""
        cb = None
        if exchange is not None:
            def _cb(user, what, ptr, count):
                try:
                    exchange(int(what), int(ptr), int(count))
                    return 0
                except Exception as e:       # never let an exception cross the C boundary
                    print(f"transit_b200: exchange callback failed: {e!r}")
                    return 1
            cb = EXCHANGE_FN(_cb)
        self._exchange_cb = cb               # keep the trampoline alive as long as the context uses it
        self._ck(self.lib.tb_set_shard(self.h, int(j_lo), int(j_hi), int(rank), int(world),
                                       C.cast(cb, C.c_void_p) if cb else None, None), "tb_set_shard")

    # -- in-library NCCL exchange of a band-sharded run ------------------------------------
    @staticmethod
    def comm_unique_id():
        buf = C.create_string_buffer(128)
        lib = load_library()
        if lib.tb_comm_unique_id(buf) != 0:
            raise TransitError(f"tb_comm_unique_id failed: {lib.tb_last_error().decode()}")
        return buf.raw

    def comm_init(self, rank, world, id128):
        self._ck(self.lib.tb_comm_init(self.h, int(rank), int(world), C.c_char_p(bytes(id128))), "tb_comm_init")

    def balanced_bands(self, world):
        e = np.zeros(world + 1, np.int64)
        self._ck(self.lib.tb_balanced_bands(self.h, int(world), _p(e, i64p)), "tb_balanced_bands")
        return e

    def rebalance_bands(self, world):
        self._ck(self.lib.tb_rebalance_bands(self.h), "tb_rebalance_bands")
        e = np.zeros(world + 1, np.int64)
        self._ck(self.lib.tb_get_bands(self.h, _p(e, i64p)), "tb_get_bands")
        return e

    def set_bands(self, edges, rank):
        e = np.ascontiguousarray(edges, np.int64)
        self._ck(self.lib.tb_set_bands(self.h, len(e) - 1, _p(e, i64p), int(rank)), "tb_set_bands")

    def sync(self):
        self._ck(self.lib.tb_sync(self.h), "tb_sync")

    def microbench(self):
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        self._ck(self.lib.tb_microbench(self.h, C.byref(a), C.byref(b), C.byref(c)), "tb_microbench")
        return dict(dfma_tflops=a.value, dexp_gops=b.value, copy_gbs=c.value)


class Transit:
    """The reference's library interface (transit.c:14-22) on top of ``Context``.

    ``problem`` is any object with the plain-array attributes of transit_b200.workloads.Problem
    (what transit_init leaves in ``struct transit`` after reading the cfg, atmosphere, TLI,
    molecules.dat and CS files).
    """

    def __init__(self, device=0):
        self.ctx = Context(device)
        self.prob = None

    def transit_init(self, prob, nDop=60, nLor=60, dmin=1e-3, dmax=0.25, lmin=1e-4, lmax=10.0, opacity=None):
        c = self.ctx
        self.prob = prob
        c.set_wn_grid(prob.wn_i, prob.wn_d, prob.wnosamp, prob.nwn, prob.onwn, prob.wn_fct)
        c.set_species(prob.mol_mass, prob.mol_radius, prob.mol_pol, prob.mol_ID, prob.iso_mass, prob.iso_ratio,
                      prob.iso_imol)
        if opacity is None:
            c.upload_lines(prob.wl, prob.elow, prob.gf, prob.isoid, prob.wfct, prob.efct)
            c.build_profiles(nDop, nLor, dmin, dmax, lmin, lmax, prob.nwidth)
        else:
            c.opacity_load(opacity["molID"], opacity["temp"], opacity["press"], opacity["grid"])
        c.cia_clear()
        for tb in prob.cia:
            c.cia_add_table(tb["wn"], tb["temp"], tb["tab"], tb["mol"])
        c.set_atmosphere(prob.rad, prob.rad_fct, prob.p, prob.p_fct, prob.t, prob.t_fct, prob.mol_d, prob.iso_z)

    def get_no_samples(self):
        return self.ctx.nwn

    def get_waveno_arr(self):
        return self.ctx.wn.copy()

    def set_scattering(self, flag, logext):
        self.ctx.set_scattering(flag, logext)

    def cloud_aux(self):
        """tau.c:194-214: per layer the H2 number density nH = d/m*q*N_A of the species NAMED "H2" and the mean
        mass density (sum_j d_j/m_j*q_j)*(sum_j m_j*q_j), what cloud models 2-5 scale with."""
        p = self.prob
        names = list(getattr(p, "species_names", []))
        md = np.zeros(p.nrad); nH = np.zeros(p.nrad); mm = np.zeros(p.nrad)
        for j in range(p.nmol):
            md = md + p.mol_d[j] / p.mol_mass[j] * p.mol_q[j]
            if j < len(names) and names[j] == "H2":
                nH = p.mol_d[j] / p.mol_mass[j] * p.mol_q[j] * 6.02214076e23
            mm = mm + p.mol_mass[j] * p.mol_q[j]
        return nH, md * mm

    def set_cloud(self, flag, ext, top, bot, gamma=1.0, Q=1.0, r=1.0, sig=1.0, refwn=1.0):
        """--cloud type,ext,top,bot[,gamma[,Q,r | sig,refwn]] (argum.c:636-715; extinction.c:629-693)."""
        nH, md = self.cloud_aux()
        self.ctx.set_cloud(flag, ext, top, bot, gamma, Q, r, sig, refwn, nH, md)

    def set_cloudtop(self, cloudtop):
        """transit.c:103-109: constant-extinction cloud deck between cloudtop and cloudtop+10 (log p)."""
        self.ctx.set_cloud(1, 100.0, cloudtop, cloudtop + 10)

    def do_transit(self, out=None):
        p = self.prob
        geom = GEOM_ECLIPSE if p.eclipse else GEOM_TRANSIT
        return self.ctx.spectrum(geom, p.toomuch, p.ethresh, p.starrad, p.transp,
                                 getattr(p, "angles", None) if p.eclipse else None, out)

    def run_batch(self, atms, out=None):
        """do_transit for a list of sampled atmospheres (same layers and units as the configured one) through
        one tb_spectrum_batch call; out[natm][nwn]."""
        p = self.prob
        geom = GEOM_ECLIPSE if p.eclipse else GEOM_TRANSIT
        st = lambda k: np.stack([np.asarray(getattr(x, k), np.float64) for x in atms])
        return self.ctx.spectrum_batch(st("rad"), p.rad_fct, st("p"), p.p_fct, st("t"), p.t_fct, st("mol_d"), st("iso_z"),
                                       geom, p.toomuch, p.ethresh, p.starrad, p.transp,
                                       getattr(p, "angles", None) if p.eclipse else None, out)

    def free_memory(self):
        self.ctx.close()
