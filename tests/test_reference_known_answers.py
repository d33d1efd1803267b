"""Known answers the reference's own test material holds for the path, applied to the
CPU oracle (CPU tests) and to the CUDA path through the C ABI (gpu tests):

* analytic optical depth through a sphere for constant / outward- / inward-increasing extinction
  (transit/test/test_slantpath.c:66-76 `calcex`, :177-198 `const_ex_anal`, `incout_ex_anal`,
  `incin_ex_anal`, sweep :522-539) and analytic modulation for constant / increasing / decreasing
  tau (:231-306 `mod_ctau`, `mod_itau`, `mod_dtau`, sweep :491-517); the reference's pass criterion
  is a relative error below `maxerr = 1e-4` (:62);
* the archived printouts of those programs (tests/golden/ref_known_answers.npz, made by
  tests/golden/make_known_answers.py from transit/test/slantpath.080604.dat and voigt.080904.dat).

The 2004 build behind the printouts integrated with GSL; today's `totaltau1`/`modulation1` use the
reference's own non-uniform Simpson (numerical.c:390-525), so the archived *observed* digits pin the
quadrature only where it is exact (constant tau / constant extinction); elsewhere the archived
*expected* digits pin the analytic formulas restated here and the oracle and the kernels must
converge onto them: within the reference's own 1e-4 at 1 000 layers (optical depth) and from 100
layers up (modulation), within the stated coarser bounds below that (today's two-point and
parabolic first-interval constructions, slantpath.c:44-74, are coarser at 10 layers than GSL was).
"""
import ctypes as C

import numpy as np
import pytest

import oraclelib as O

MAXERR = 1e-4                       # test_slantpath.c:62
TAU_BOUND = {10: 3e-2, 100: 3e-3, 1000: MAXERR}
MOD_BOUND = {10: 3e-3, 100: MAXERR, 1000: MAXERR, 10000: MAXERR}
H_, LS_, KB_ = 6.6260755e-27, 2.99792458e10, 1.380658e-16   # constants_tr.h (cgs)


@pytest.fixture(scope="module")
def known(golden):
    return golden("ref_known_answers")


def calcex(alpha, rm, r):
    r = np.asarray(r, dtype=np.float64)
    if alpha == 0:
        return np.ones_like(r)
    return -alpha * (rm - r) if alpha < 0 else alpha * r


def tau_analytic(alpha, rm, ip):
    rat = rm / ip
    s = np.sqrt(rat * rat - 1)
    if alpha == 0:
        return 2 * np.sqrt(rm * rm - ip * ip)
    if alpha > 0:
        return alpha * (rm * ip * s + ip * ip * np.log(s + rat))
    return -alpha * (rm * ip * s - ip * ip * np.log(s + rat))


def mod_tau(kind, prm, ipmax, first, nip):
    k = np.arange(nip, dtype=np.float64)
    if kind == 0:
        return np.full(nip, prm)
    delt = (1 - first) / (nip - 1)
    return prm * ipmax * (1 - k * delt) if kind == 1 else prm * ipmax * k * delt


def mod_analytic(kind, prm, star, ipmax, first, toomuch):
    rath, ratl = ipmax / star, first * ipmax / star
    if kind == 0:
        return -np.exp(-prm) * (rath * rath - ratl * ratl) - np.exp(-toomuch) * ratl * ratl + rath * rath
    if kind == 1:
        r = -2 * (np.exp(-prm * ipmax * first) * (first * ipmax + 1 / prm)
                  - np.exp(-prm * ipmax) * (ipmax + 1 / prm)) / star / star / prm
    else:
        r = -2 * ((ipmax - 1 / prm) - np.exp(-prm * ipmax * (1 - first)) * (ipmax * first - 1 / prm)) / star / star / prm
    return r - np.exp(-toomuch) * ratl * ratl + rath * rath


def mod_rays(ipmax, first, nip):
    delt = ipmax * (1 - first) / (nip - 1)
    return np.array([first * ipmax + (nip - 1 - i) * delt for i in range(nip)])


def oracle_totaltau1(ip, rad, ex):
    return O.lib().to_totaltau1(float(ip), O.p(O.c64(rad).copy()), 1.0, O.p(O.c64(ex).copy()), len(rad))


# ------------------------------------------------------------------ oracle (CPU)

def test_oracle_totaltau1_analytic_sweep(known):
    """test_tau (test_slantpath.c:522-539): 3 extinction laws x 3 planet radii x 4 impact parameters x
    3 layer counts = 108 cases."""
    cases = known["tau_cases"]
    assert len(cases) == 108
    for alpha, rm, ip, nrad, expected, observed, err2004 in cases:
        nrad = int(nrad)
        a = tau_analytic(alpha, rm, ip)
        assert abs(a - expected) <= 1e-5 * abs(expected)          # the archived 6-digit expectation
        rad = (np.arange(nrad) + 1) * (rm / nrad)                 # tau_dens, :104-127
        r = oracle_totaltau1(ip, rad, calcex(alpha, rm, rad))
        err = abs(1 - r / a)
        if alpha == 0:
            # Simpson in s = sqrt(r^2 - r0^2) integrates a constant exactly; the 2004 build did too
            assert err < 1e-13 and abs(1 - r / observed) < 6e-10
        else:
            assert err < TAU_BOUND[nrad], (alpha, rm, ip, nrad, err)


def test_oracle_modulation1_analytic_sweep(known):
    """test_mod (test_slantpath.c:491-517): 3 tau laws x star radii x planet radii (not above the star)
    x 4 inner edges x 4 layer counts; transparent-planet term on, toomuch 30."""
    cases = known["mod_cases"]
    assert len(cases) == 288
    L = O.lib()
    for kind, prm, star, ipmax, first, nip, expected, observed, err2004 in cases:
        kind, nip = int(kind), int(nip)
        a = mod_analytic(kind, prm, star, ipmax, first, 30.0)
        assert abs(a - expected) <= 2e-6 * abs(expected)
        tau = mod_tau(kind, prm, ipmax, first, nip)
        ip = mod_rays(ipmax, first, nip)
        r = L.to_modulation1(O.p(tau), nip - 1, 30.0, O.p(ip), nip, 1.0, star, 1)
        err = abs(1 - r / a)
        if kind == 0:
            assert err < 1e-12 and abs(1 - r / observed) < 6e-10   # archived 10 digits reproduced
        else:
            assert err < MOD_BOUND[nip], (kind, star, ipmax, first, nip, err)


def voigt_archive_check(prof, dw, known):
    """prof: bin means on 2*psize+1 bins of width dw centred on the line, bin k covering points k, k+1
    (voigt.c:464-478 with one interval per bin).  The archive has 10 000 bins over +-75 cm-1 (spacing
    150/9999), values at the same half-bin offset, 4 significant digits."""
    x = known["voigt_x"]
    v = known["voigt_v0"].astype(np.float64)
    n = len(v)
    step = 150.0 / (n - 1)
    xa = -75.0 + step * np.arange(n) + step / 2                   # where the archived bins sit
    psz = (len(prof) - 1) // 2
    xo = (np.arange(len(prof)) - psz) * dw + dw / 2
    assert np.all(np.abs(x - (xa - step / 2)) < 6e-4)            # printed abscissa, 5 digits
    o = np.interp(xa, xo, prof.astype(np.float64))
    digit = 10.0 ** (np.floor(np.log10(v * (1 + 1e-6))) - 3)    # one unit of the 4th printed digit
    return float(np.max(np.abs(o - v) / digit))


def test_oracle_voigt_archived_table(known):
    """pu/test/test_voigt.c printout: Lorentz 1.5, Doppler 1 (half widths, cm-1), +-75 cm-1."""
    assert known["voigt_lor"] == 1.5 and known["voigt_dop"] == 1.0 and known["voigt_range"] == 75.0
    dw = 0.015
    L = O.lib()
    psz = L.to_voigt_psize(dw, 1.0, 1.5, 50.0, 10 ** 7)
    assert psz == 5000
    prof = np.zeros(2 * psz + 1, np.float32)
    L.to_voigt_profile(O.p(prof, O.f32p), dw, 1.0, 1.5, 50.0, 10 ** 7)
    assert voigt_archive_check(prof, dw, known) < 0.75           # half a digit of rounding + 3e-5 of interpolation


def isothermal_flux(wn, T):
    return np.pi * 2.0 * H_ * wn ** 3 * LS_ * LS_ / (np.exp(H_ * wn * LS_ / (KB_ * T)) - 1.0)


def _eclipse_problem(nrad=40, nwn=16, T=1200.0):
    rad = 7.0e4 + np.linspace(0.0, 3000.0, nrad)                  # km
    wn = 2500.0 + 10.0 * np.arange(nwn)
    e = np.outer(np.geomspace(3e-8, 1e-12, nrad), 1.0 + 0.3 * np.sin(wn))   # cm-1, tau(bottom) ~ 3
    return O.Problem(nwn=nwn, wn=wn, wn_i=wn[0], wn_d=10.0, wn_fct=1.0, rad=rad, nrad=nrad, rad_fct=1e5,
                     p=np.geomspace(1e2, 1e-6, nrad), p_fct=1e6, t=np.full(nrad, T), t_fct=1.0,
                     toomuch=10.0, eclipse=True, nmol=1, mol_pol=np.zeros(1), mol_d=np.ones((1, nrad)),
                     mol_mass=np.array([2.0]), angles=np.array([0.0, 20.0, 40.0, 60.0, 80.0])), e


def test_oracle_isothermal_eclipse_flux_is_pi_planck():
    """An isothermal atmosphere emits pi*B_nu(T) whatever its opacity: with B constant,
    eclipse_intens (eclipse.c:158) is B*dtau[last] - B*(dtau[last] - dtau[0]) = B, and flux (eclipse.c:262-279)
    sums pi*B over the annuli sin^2(90) - sin^2(0) = 1."""
    prob, e = _eclipse_problem()
    t, last, _ = O.tau(prob, e, np.zeros((prob.nwn, prob.nrad)), lazy=False)
    assert np.all(last >= 1)
    intens, flux = O.intensity_flux(prob, t, last)
    assert np.max(np.abs(flux / isothermal_flux(prob.wn, 1200.0) - 1)) < 1e-12


# ------------------------------------------------------------------ CUDA path through the C ABI

def _ctx_with_extinction(rad, rad_fct, e_cols, temp=None, wn_i=2500.0, wn_d=10.0):
    """A context whose extinction e[r][w] is the given array: a one-molecule opacity grid holding the
    values at both grid temperatures, unit density (interpolmolext, extinction.c:568-578)."""
    from transit_b200 import api
    nrad, nwn = e_cols.shape
    c = api.Context(0)
    c.set_wn_grid(wn_i, wn_d, 1, nwn, nwn, 1.0)
    c.set_species([2.0], [1.0], [0.0], [105], [2.0], [1.0], [0])
    temp = np.full(nrad, 1000.0) if temp is None else temp
    c.set_atmosphere(rad, rad_fct, np.geomspace(1e2, 1e-6, nrad), 1e6, temp, 1.0, np.ones((1, nrad)),
                     np.ones((1, nrad)))
    gtemp = np.array([500.0, 1500.0])
    grid = np.empty((nrad, 2, 1, nwn))
    grid[:, 0, 0, :] = e_cols
    grid[:, 1, 0, :] = e_cols
    c.opacity_load(np.array([105], np.int32), gtemp, np.geomspace(1e2, 1e-6, nrad) * 1e6, grid)
    c.extinction_interp()
    return c


@pytest.mark.gpu
@pytest.mark.parametrize("nrad", [10, 100])
def test_gpu_totaltau1_analytic_sweep(known, nrad):
    """The test_tau sweep through tb_tau: the three extinction laws are three wavenumber columns, the
    four impact parameters four rays (not on layer radii for 0.75 at 10 layers: interp_parab and the
    replaced first point are exercised off the grid)."""
    from transit_b200 import api
    for rm in (10.0, 100.0, 1000.0):
        rad = (np.arange(nrad) + 1) * (rm / nrad)
        alphas = (1.0, -1.0, 0.0)
        e = np.stack([calcex(a, rm, rad) for a in alphas], axis=1)
        c = _ctx_with_extinction(rad, 1.0, e)
        eg = c.get_extinction()
        assert np.max(np.abs(eg - e)) <= 4e-16 * np.max(np.abs(e))
        rays = np.array([0.9, 0.75, 0.5, 0.1]) * rm                # top first
        c.tau(api.GEOM_TRANSIT, 1e300, h=rays, hfct=1.0)
        t, last = c.get_tau()
        assert t.shape == (3, 4) and np.all(last == 3)
        for w, a in enumerate(alphas):
            for i, b in enumerate(rays):
                want = tau_analytic(a, rm, b)
                bound = 1e-13 if a == 0 else TAU_BOUND[nrad]
                assert abs(1 - t[w, i] / want) < bound, (a, rm, b, nrad)
                o = oracle_totaltau1(b, rad, eg[:, w])
                assert abs(t[w, i] - o) <= 1e-12 * abs(o), (a, rm, b, nrad, t[w, i], o)
        c.close()


@pytest.mark.gpu
@pytest.mark.parametrize("nrad", [100, 150])
def test_gpu_constant_extinction_modulation_closed_form(nrad):
    """tb_tau + tb_modulation against a closed form: constant extinction k in a shell b0..R gives
    tau(b) = 2k sqrt(R^2-b^2) and, with u = sqrt(R^2-b^2),
    M R*^2 = R^2 - 2 [1 - exp(-2k u0)(1 + 2k u0)]/(2k)^2 - exp(-toomuch) b0^2,  u0 = sqrt(R^2-b0^2)."""
    from transit_b200 import api
    R, b0, k, star, toomuch = 100.0, 50.0, 0.01, 400.0, 30.0
    rad = np.linspace(b0, R, nrad)
    e = np.full((nrad, 4), k)
    c = _ctx_with_extinction(rad, 1.0, e)
    c.tau(api.GEOM_TRANSIT, toomuch)
    t, last = c.get_tau()
    assert np.all(last == nrad - 1)
    rays = rad[::-1]
    want_tau = 2 * k * np.sqrt(np.maximum(R * R - rays * rays, 0.0))
    assert np.max(np.abs(t[0] - want_tau)) < 1e-12 * want_tau.max()
    u0 = np.sqrt(R * R - b0 * b0)
    integ = (1 - np.exp(-2 * k * u0) * (1 + 2 * k * u0)) / (2 * k) ** 2
    for transp in (0, 1):
        m = c.modulation(star, transp)
        want = (R * R - 2 * integ - transp * np.exp(-toomuch) * b0 * b0) / star ** 2
        o = O.lib().to_modulation1(O.p(O.c64(t[0])), nrad - 1, toomuch, O.p(O.c64(rays)), nrad, 1.0, star, transp)
        assert np.all(np.abs(m - o) <= 1e-12 * abs(o))
        # exp(-tau(b)) b has an infinite slope at b = R: the quadrature converges as h^1.5
        assert abs(1 - m[0] / want) < 5e-4
    c.close()


@pytest.mark.gpu
def test_gpu_isothermal_eclipse_flux_is_pi_planck():
    from transit_b200 import api
    prob, e = _eclipse_problem()
    c = _ctx_with_extinction(prob.rad, prob.rad_fct, e, temp=prob.t, wn_i=prob.wn_i, wn_d=prob.wn_d)
    c.tau(api.GEOM_ECLIPSE, prob.toomuch)
    t, last = c.get_tau()
    t0, last0, _ = O.tau(prob, c.get_extinction(), np.zeros((prob.nwn, prob.nrad)), lazy=False)
    assert np.array_equal(last, last0)
    assert np.max(np.abs(t - t0)) <= 1e-9 * np.max(np.abs(t0))
    flux = c.intensity_flux(prob.angles)
    assert np.max(np.abs(flux / isothermal_flux(prob.wn, 1200.0) - 1)) < 1e-12
    c.close()


@pytest.mark.gpu
def test_gpu_voigt_archived_table(known):
    """tb_build_profiles on a grid whose fine spacing is 0.015 cm-1, widths bracketing the archive's
    (Doppler 1, Lorentz 1.5): the (0, 0) profile of the table against the archived printout."""
    from transit_b200 import api
    c = api.Context(0)
    nwn = 2001
    c.set_wn_grid(1000.0, 0.15, 10, nwn, (nwn - 1) * 10 + 1, 1.0)
    c.build_profiles(2, 2, 1.0, 1.25, 1.5, 2.0, 50.0)
    t = c.get_profiles()
    assert abs(t["aDop"][0] - 1.0) < 1e-14 and abs(t["aLor"][0] - 1.5) < 1e-14 and t["psize"][0, 0] == 5000
    prof = t["data"][t["off"][0, 0]: t["off"][0, 0] + 2 * 5000 + 1]
    ref = np.zeros(10001, np.float32)
    O.lib().to_voigt_profile(O.p(ref, O.f32p), 0.15 / 10, 1.0, 1.5, 50.0, (nwn - 1) * 10 + 1)
    assert np.array_equal(prof, ref)                               # bit-exact against the oracle
    assert voigt_archive_check(prof, 0.015, known) < 0.75
    c.close()
