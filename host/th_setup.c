/* th_setup.c -- host-side set-up around the device path:
 *   makewnsample / makesample1     transit/src/makesample.c:27-120,308-400
 *   checkaddmm, stateeqnford       readatm.c:122-159, transit.h:58-69
 *   radpress                       readatm.c:787-865
 *   makeradsample (raddelt = -1)   makesample.c:408-549 (natural splines, spline.c:12-48,131-206)
 *   opacity() driver + file I/O    opacity.c:11-215,281-503
 */
#include <fcntl.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#include "th_internal.h"

/* multi-GPU run: this rank's wavenumber band, once the grid (and, line by line, the line index) is on the device */
static void set_bands(th_state *s) {
  if (th_world <= 1) return;
  int64_t edges[65];
  if (TH_DEV(tb_balanced_bands(s->ctx, th_world, edges)) || TH_DEV(tb_set_bands(s->ctx, th_world, edges, th_rank)))
    th_die("%s\n", tb_last_error());
  s->band_lo = edges[th_rank]; s->band_hi = edges[th_rank + 1];
}

/* ---- natural cubic spline, pu/src/spline.c:12-48 (tri + spline_init) and :131-183 */
void th_spline_init(double *z, const double *x, const double *y, long n) {
  double *u = (double *)calloc(n, sizeof(double)), *v = (double *)calloc(n, sizeof(double));
  for (long i = 1; i < n - 1; i++) {
    const double hi = x[i + 1] - x[i], hm = x[i] - x[i - 1];
    const double bi = (y[i + 1] - y[i]) / hi, bm = (y[i] - y[i - 1]) / hm;
    if (i == 1) { u[i] = 2 * (hi + hm); v[i] = 6 * (bi - bm); }
    else { u[i] = 2 * (hi + hm) - hm * hm / u[i - 1]; v[i] = 6 * (bi - bm) - v[i - 1] * hm / u[i - 1]; }
  }
  z[0] = z[n - 1] = 0;
  for (long i = n - 2; i > 0; i--) z[i] = (v[i] - (x[i + 1] - x[i]) * z[i + 1]) / u[i];
  free(u); free(v);
}

static long binsearchapprox(const double *a, double v, long lo, long hi) {   /* iomisc.c:1088-1108 */
  while (hi - lo > 1) {
    const long mid = (hi + lo) / 2;
    if (a[mid] > v) hi = mid; else lo = mid;
  }
  if (hi - lo == 1) return fabs(a[hi] - v) < fabs(a[lo] - v) ? hi : lo;
  return lo;
}

double th_splinterp_pt(const double *z, long n, const double *x, const double *y, double xout) {
  long idx = binsearchapprox(x, xout, 0, n - 1);
  if (idx == n - 1 || xout < x[idx]) idx--;
  if (idx < 0) idx = 0;
  const double h = x[idx + 1] - x[idx], dy = y[idx + 1] - y[idx];
  if (x[idx] == xout) return y[idx];
  if (h > 0) {
    const double dx = xout - x[idx];
    const double a = (z[idx + 1] - z[idx]) / (6 * h), b = 0.5 * z[idx];
    const double c = dy / h - h / 6 * (z[idx + 1] + 2 * z[idx]);
    return y[idx] + dx * (c + dx * (b + dx * a));
  }
  return 0;
}

static void splinterp(long n, const double *x, const double *y, long nout, const double *xout, double *yout) {
  if (n < 2) { for (long i = 0; i < nout; i++) yout[i] = y[0]; return; }
  double *z = (double *)calloc(n, sizeof(double));
  th_spline_init(z, x, y, n);
  for (long i = 0; i < nout; i++) yout[i] = th_splinterp_pt(z, n, x, y, xout[i]);
  free(z);
}

/* The same evaluation with the second derivatives z computed once per table, and the interval
 * found by walking from the previous point's interval.  binsearchapprox + the step back of
 * th_splinterp_pt select the largest idx with x[idx] <= xout, kept in [0, n-2]; the walk below
 * lands on that same idx, so every output is the same double. */
static void splinterp_cached(long n, const double *x, const double *y, const double *z, long nout,
                             const double *xout, double *yout) {
  if (n < 2) { for (long i = 0; i < nout; i++) yout[i] = y[0]; return; }
  long idx = 0;
  for (long i = 0; i < nout; i++) {
    const double xo = xout[i];
    while (idx < n - 2 && x[idx + 1] <= xo) idx++;
    while (idx > 0 && x[idx] > xo) idx--;
    const double h = x[idx + 1] - x[idx], dy = y[idx + 1] - y[idx];
    if (x[idx] == xo) { yout[i] = y[idx]; continue; }
    if (h > 0) {
      const double dx = xo - x[idx];
      const double a = (z[idx + 1] - z[idx]) / (6 * h), b = 0.5 * z[idx];
      const double c = dy / h - h / 6 * (z[idx + 1] + 2 * z[idx]);
      yout[i] = y[idx] + dx * (c + dx * (b + dx * a));
    } else
      yout[i] = 0;
  }
}

/* ---- wavenumber sampling */
int th_make_wn(th_state *s) {
  const th_hints *h = &s->h;
  double wi = 0, wf = 0;
  if (h->wn_i > 0) {
    if (h->wn_fct <= 0) th_die("User specified wavenumber factor is negative (%g).\n", h->wn_fct);
    wi = h->wn_i * h->wn_fct;
  } else if (h->wl_f > 0) {
    if (h->wl_fct <= 0) th_die("User specified wavelength factor is negative (%g).\n", h->wl_fct);
    wi = 1.0 / (h->wl_f * h->wl_fct);
  } else
    th_die("Initial wavenumber (nor final wavelength) were correctly provided by the user.\n");
  if (h->wn_f > 0) {
    if (h->wn_fct < 0) th_die("User specified wavenumber factor is negative (%g).\n", h->wn_fct);
    wf = h->wn_f * h->wn_fct;
  } else if (h->wl_i > 0) {
    if (h->wl_fct < 0) th_die("User specified wavelength factor is negative (%g).\n", h->wl_fct);
    wf = 1.0 / (h->wl_i * h->wl_fct);
  } else
    th_die("Final wavenumber (nor initial wavelength) were correctly provided by the user.\n");
  if (h->wn_d <= 0) th_die("Incorrect wavenumber spacing (%g), it must be positive.\n", h->wn_d);
  if (h->wn_o <= 0) th_die("makesample() :: Invalid oversampling (%d).\n", h->wn_o);
  if (wf < wi) th_die("makewnsample() :: final wavenumber (%g) is smaller than the initial one (%g).\n", wf, wi);
  s->wn_i = wi; s->wn_f = wf; s->wn_d = h->wn_d; s->wn_o = h->wn_o;
  s->nwn = tb_grid_count(wi, wf, h->wn_d, 1);
  s->onwn = tb_grid_count(wi, wf, h->wn_d, h->wn_o);
  if (s->nwn < 2) th_die("The wavenumber sampling has less than 2 points.\n");
  return 0;
}

/* ---- mean molecular mass and densities for every layer of the atmosphere file */
int th_densities(th_state *s) {
  const int n = s->nlayer, nmol = s->nmol;
  for (int r = 0; r < n; r++) {
    double mm = 0, sumq = 0;
    for (int i = 0; i < nmol; i++) {
      const double q = s->q[(size_t)i * n + r];
      mm += s->bymass ? q / s->molmass[i] : q * s->molmass[i];
      sumq += q;
    }
    if (s->bymass) mm = 1.0 / mm;
    if (sumq > 1.001) th_die("Sum of abundances of isotopes adds up to more than 1: %g\n", sumq);
    if (fabs(sumq - 1.0) > s->h.allowq && s->h.verb > 1)
      fprintf(stderr, "In radius %g (%i: %g in file), abundances don't add up to 1: %.9g\n", s->rad[r], r,
              s->rad[r] - s->zerorad, sumq);
    s->mm[r] = mm;
    const double p = s->p[r] * s->p_fct, t = s->t[r] * s->t_fct;
    for (int i = 0; i < nmol; i++) {
      const double rho = TH_AMU * s->q[(size_t)i * n + r] * p / TH_KB / t;
      s->d[(size_t)i * n + r] = s->bymass ? rho * mm : rho * s->molmass[i];
    }
  }
  return 0;
}

/* ---- hydrostatic radii from the reference level (p0, r0, gsurf) */
int th_radpress(th_state *s) {
  const int n = s->nlayer;
  const double g0 = s->h.gsurf, p0 = s->h.p0, r0 = s->h.r0, rfct = s->rad_fct;
  const double *temp = s->t, *mu = s->mm, *pr = s->p;
  double *radius = s->rad;
  if (p0 == 0 || r0 == 0 || g0 == 0)
    th_die("Surface gravity (%g cm s-2), reference pressure (%g) or reference radius (%g) were not defined.\n", g0, p0, r0);
  double minimum = 1e37, temp0, mu0, g;
  int i0 = -1;
  for (int i = 0; i < n; i++)
    if (fabs(pr[i] - p0) < minimum) { i0 = i; minimum = fabs(pr[i] - p0); }
  if ((pr[i0] > p0 && i0 + 1 >= n) || (pr[i0] <= p0 && i0 - 1 < 0))
    th_die("The reference pressure (%g) lies outside the pressure range of the atmosphere file.\n", p0);
  const double kb_amu = TH_KB / TH_AMU;
  if (pr[i0] > p0) {
    temp0 = temp[i0] + ((temp[i0 + 1] - temp[i0]) / log(pr[i0 + 1] / pr[i0])) * log(p0 / pr[i0]);
    mu0 = mu[i0] + ((mu[i0 + 1] - mu[i0]) / log(pr[i0 + 1] / pr[i0])) * log(p0 / pr[i0]);
    radius[i0] = r0 + 0.5 * (temp[i0] / mu[i0] + temp0 / mu0) * (kb_amu * log(p0 / pr[i0]) / g0) / rfct;
  } else {
    temp0 = temp[i0] + ((temp[i0 - 1] - temp[i0]) / log(pr[i0 - 1] / pr[i0])) * log(p0 / pr[i0]);
    mu0 = mu[i0] + ((mu[i0 - 1] - mu[i0]) / log(pr[i0 - 1] / pr[i0])) * log(p0 / pr[i0]);
    radius[i0] = r0 - 0.5 * (temp[i0] / mu[i0] + temp0 / mu0) * (kb_amu * log(pr[i0] / p0) / g0) / rfct;
  }
  g = g0 * pow(r0 / radius[i0], 2);
  for (int i = i0 - 1; i >= 0; i--) {
    radius[i] = radius[i + 1] - 0.5 * (temp[i] / mu[i] + temp[i + 1] / mu[i + 1]) * (kb_amu * log(pr[i] / pr[i + 1]) / g) / rfct;
    g = g * pow(radius[i + 1] / radius[i], 2);
  }
  g = g0 * pow(r0 / radius[i0], 2);
  for (int i = i0 + 1; i < n; i++) {
    radius[i] = radius[i - 1] + 0.5 * (temp[i] / mu[i] + temp[i - 1] / mu[i - 1]) * (kb_amu * log(pr[i - 1] / pr[i]) / g) / rfct;
    g = g * pow(radius[i - 1] / radius[i], 2);
  }
  return 0;
}

/* ---- makeradsample (makesample.c:408-549).  raddelt = -1 keeps the atmosphere file's own layers; a positive
 * raddelt builds an equidistant radius sample between radlow and radhigh (the file's first / last radius where
 * those are 0) by makesample (makesample.c:145-300) and carries temperature, pressure, mean molecular mass and
 * every species' density and abundance over by natural cubic splines in radius (makesample.c:508-530).  In both
 * cases the partition function of every isotope is splined in temperature at the layer temperatures and the
 * sampled atmosphere goes to the device; the resampled case also makes the impact-parameter sample of
 * makeipsample's second branch (makesample.c:577-596): from the top radius downwards in steps of raddelt. */
void th_file_layers(th_state *s) {
  if (!s->resampled) return;
  /* current = resampled (capacity other.cap), other = file */
  th_layers samp = {s->nlayer, s->other.cap, s->rad, s->p, s->t, s->mm, s->q, s->d};
  const th_layers f = s->other;
  s->nlayer = f.n; s->rad = f.rad; s->p = f.p; s->t = f.t; s->mm = f.mm; s->q = f.q; s->d = f.d;
  s->other = samp;
  s->resampled = 0;
}

static void resample_layers(th_state *s) {
  const th_hints *h = &s->h;
  const int fn = s->nlayer, nmol = s->nmol;
  if (h->rad_d <= 0)
    th_die("raddelt %g: the radius spacing has to be positive (or -1 to keep the atmosphere file's layers); "
           "makesample.c:229-233, :585-596.\n", h->rad_d);
  /* makesample(rad, hint = {radlow, radhigh, raddelt, radfct}, ref = atmosphere radii, TRH_RAD) */
  if (h->rad_fct > 0) s->rad_fct = h->rad_fct;             /* makesample.c:157-160 (values are NOT converted there either) */
  const double ri = h->rad_i <= 0 ? s->rad[0] : h->rad_i, rf = h->rad_f <= 0 ? s->rad[fn - 1] : h->rad_f;
  const double dl = h->rad_d;
  if (rf <= ri)
    th_die("Initial accepted sampling value (%g) is greater or equal than final accepted sample value (%g). "
           "radius was being hinted.\n", ri, rf);
  long n = (long)(((1.0 + 1e-8) * rf - ri) / dl + 1);      /* makesample.c:248 */
  if (n < 0) n = -n;
  if (n < 1 || n > 100000) th_die("raddelt %g gives %ld radius samples.\n", dl, n);
  th_layers *o = &s->other;                                 /* the spare set of buffers */
  if (o->cap < n) {
    o->cap = (int)n;
    o->rad = (double *)realloc(o->rad, n * sizeof(double)); o->p = (double *)realloc(o->p, n * sizeof(double));
    o->t = (double *)realloc(o->t, n * sizeof(double)); o->mm = (double *)realloc(o->mm, n * sizeof(double));
    o->q = (double *)realloc(o->q, (size_t)n * nmol * sizeof(double));
    o->d = (double *)realloc(o->d, (size_t)n * nmol * sizeof(double));
  }
  o->n = (int)n;
  o->rad[0] = ri;
  for (long k = 1; k < n; k++) o->rad[k] = ri + k * dl;     /* makesample.c:273-279, oversampling 1 */
  splinterp(fn, s->rad, s->t, n, o->rad, o->t);             /* makesample.c:508-510 */
  splinterp(fn, s->rad, s->p, n, o->rad, o->p);
  splinterp(fn, s->rad, s->mm, n, o->rad, o->mm);
  for (int i = 0; i < nmol; i++) {                          /* makesample.c:527-530 */
    splinterp(fn, s->rad, s->d + (size_t)i * fn, n, o->rad, o->d + (size_t)i * n);
    splinterp(fn, s->rad, s->q + (size_t)i * fn, n, o->rad, o->q + (size_t)i * n);
  }
  s->samp_f = rf;
  /* current <- resampled, other <- file */
  th_layers file = {fn, fn, s->rad, s->p, s->t, s->mm, s->q, s->d};
  const th_layers samp = *o;
  s->nlayer = samp.n; s->rad = samp.rad; s->p = samp.p; s->t = samp.t; s->mm = samp.mm; s->q = samp.q; s->d = samp.d;
  file.cap = samp.cap;                                      /* `other.cap` always describes the RESAMPLED buffers */
  s->other = file;
  s->resampled = 1;
  /* makeipsample: makesample(ips, hint = {0, -0, 0, 0}, ref = {d = -raddelt, i = top radius, f = bottom radius, o = 1}) */
  const double ii = s->rad[n - 1], ff = s->rad[0], dd = -dl;
  long nip = (long)(((1.0 - 1e-8) * ff - ii) / dd + 1);
  if (nip < 0) nip = -nip;
  s->ips = (double *)realloc(s->ips, (nip > 0 ? nip : 1) * sizeof(double));
  s->nip = (int)nip;
  for (long k = 0; k < nip; k++) s->ips[k] = k ? ii + k * dd : ii;
}

int th_make_rad(th_state *s) {
  if (s->h.rad_d != -1) resample_layers(s);
  const int n = s->nlayer, niso = s->niso;
  for (int i = 0; i < n; i++) {
    if (s->t[i] < s->tli_tmin) th_die("The layer %d in the atmospheric model has a lower temperature (%.1f K) than the lowest allowed TLI temperature (%.1f K).\n", i, s->t[i], s->tli_tmin);
    if (s->t[i] > s->tli_tmax) th_die("The layer %d in the atmospheric model has a higher temperature (%.1f K) than the highest allowed TLI temperature (%.1f K).\n", i, s->t[i], s->tli_tmax);
  }
  if (s->zcap < (int64_t)(niso ? niso : 1) * n) {
    s->zcap = (int64_t)(niso ? niso : 1) * n;
    s->Zlayer = (double *)realloc(s->Zlayer, (size_t)s->zcap * sizeof(double));
  }
  if (!s->isoZ2 && niso) {        /* run_transit comes here once per spectrum: the tridiagonal solves are done once */
    int64_t zn = 0;
    for (int i = 0; i < niso; i++)
      if (s->isoZoff[i] + s->db[s->isodb[i]].nT > zn) zn = s->isoZoff[i] + s->db[s->isodb[i]].nT;
    s->isoZ2 = (double *)calloc((size_t)zn, sizeof(double));
    for (int i = 0; i < niso; i++) {
      const th_db *db = s->db + s->isodb[i];
      if (db->nT >= 2) th_spline_init(s->isoZ2 + s->isoZoff[i], db->T, s->isoZ + s->isoZoff[i], db->nT);
    }
  }
  for (int i = 0; i < niso; i++) {
    const th_db *db = s->db + s->isodb[i];
    splinterp_cached(db->nT, db->T, s->isoZ + s->isoZoff[i], s->isoZ2 + s->isoZoff[i], n, s->t,
                     s->Zlayer + (size_t)i * n);                                             /* makesample.c:534-544 */
  }
  if (TH_DEV(tb_set_atmosphere(s->ctx, n, s->rad, s->rad_fct, s->p, s->p_fct, s->t, s->t_fct, s->d, niso ? s->Zlayer : NULL)))
    th_die("%s\n", tb_last_error());
  /* transit geometry integrates along the impact-parameter sample; eclipse along the reversed radii (tau.c:92-104) */
  if (TH_DEV(tb_set_rays(s->ctx, s->resampled && !s->eclipse ? s->ips : NULL, s->resampled && !s->eclipse ? s->nip : 0, s->rad_fct)))
    th_die("%s\n", tb_last_error());
  return 0;
}

/* ---- opacity(): Voigt table, and the grid: read it, or build + write it, or none */
static int file_exists(const char *p) {
  FILE *f = p ? fopen(p, "rb") : NULL;
  if (!f) return 0;
  fclose(f);
  return 1;
}

int th_opacity(th_state *s) {
  const th_hints *h = &s->h;
  s->have_grid_file = 0;
  if (h->f_opa && file_exists(h->f_opa)) {                       /* readopacity, opacity.c:432-503 */
    FILE *fp = fopen(h->f_opa, "rb");
    long dims[4];                                                /* Nmol, Ntemp, Nlayer, Nwave */
    if (fread(dims, sizeof(long), 4, fp) != 4) th_die("Opacity file '%s' is truncated.\n", h->f_opa);
    const long Nmol = dims[0], Nt = dims[1], Nl = dims[2], Nw = dims[3];
    if (Nw != s->nwn || Nl != s->nlayer)
      th_die("Opacity file '%s' holds %ld layers x %ld wavenumbers; the run needs %d x %lld.\n", h->f_opa, Nl, Nw,
             s->nlayer, (long long)s->nwn);
    int *molID = (int *)malloc(Nmol * sizeof(int));
    double *temp = (double *)malloc(Nt * sizeof(double)), *press = (double *)malloc(Nl * sizeof(double));
    double *wns = (double *)malloc(Nw * sizeof(double));
    const size_t ng = (size_t)Nl * Nt * Nmol * Nw;
    double *grid = (double *)malloc(ng * sizeof(double));
    if (!grid) th_die("Not enough host memory for the opacity grid (%zu bytes).\n", ng * sizeof(double));
    if (fread(molID, sizeof(int), Nmol, fp) != (size_t)Nmol || fread(temp, 8, Nt, fp) != (size_t)Nt ||
        fread(press, 8, Nl, fp) != (size_t)Nl || fread(wns, 8, Nw, fp) != (size_t)Nw || fread(grid, 8, ng, fp) != ng)
      th_die("Opacity file '%s' is truncated.\n", h->f_opa);
    fclose(fp);
    if (TH_DEV(tb_opacity_load(s->ctx, Nmol, Nt, Nl, Nw, molID, temp, press, grid))) th_die("%s\n", tb_last_error());
    free(molID); free(temp); free(press); free(wns); free(grid);
    s->have_grid_file = 1;
    set_bands(s);
    return 0;
  }
  /* line-by-line: line index + Voigt table (calcprofiles, opacity.c:218-277) */
  if (TH_DEV(tb_upload_lines(s->ctx, s->wl, s->elow, s->gf, s->isoid, s->nlines, 1e-4, 1.0))) th_die("%s\n", tb_last_error());
  if (TH_DEV(tb_build_profiles(s->ctx, h->nDop, h->nLor, h->dmin, h->dmax, h->lmin, h->lmax, h->nwidth))) th_die("%s\n", tb_last_error());
  set_bands(s);
  if (!h->f_opa) return 0;
  /* calcopacity (opacity.c:281-427): temperature grid, Z(T) per isotope, grid, file */
  const long Nt = (long)tb_grid_count(h->tlow, h->thigh, h->tempdelt, 1);   /* maketempsample, makesample.c:614-636 */
  if (Nt < 2) th_die("The opacity temperature grid needs at least 2 samples (tlow %g, thigh %g, tempdelt %g).\n", h->tlow, h->thigh, h->tempdelt);
  const int n = s->nlayer, niso = s->niso, nmol = s->nmol;
  double *temp = (double *)malloc(Nt * sizeof(double)), *ziso = (double *)malloc((size_t)(niso ? niso : 1) * Nt * sizeof(double));
  for (long t = 0; t < Nt; t++) temp[t] = h->tlow + t * h->tempdelt;
  if (temp[0] < s->tli_tmin || temp[Nt - 1] > s->tli_tmax)
    th_die("The opacity temperature grid [%.1f, %.1f] K exceeds the TLI partition-function range [%.1f, %.1f] K.\n",
           temp[0], temp[Nt - 1], s->tli_tmin, s->tli_tmax);
  for (int i = 0; i < niso; i++) {
    const th_db *db = s->db + s->isodb[i];
    splinterp(db->nT, db->T, s->isoZ + s->isoZoff[i], Nt, temp, ziso + (size_t)i * Nt);
  }
  int molID[256], nrows = 0;
  if (tb_get_line_molIDs(s->ctx, molID, &nrows)) th_die("%s\n", tb_last_error());
  const size_t ng = (size_t)n * Nt * nrows * s->nwn;
  double *grid = (double *)malloc(ng * sizeof(double));
  if (!grid) th_die("Not enough host memory for the opacity grid (%zu bytes).\n", ng * sizeof(double));
  if (TH_DEV(tb_opacity_build(s->ctx, (int)Nt, temp, ziso, s->q, s->mm, s->bymass, h->ethresh, grid))) th_die("%s\n", tb_last_error());
  /* the file (opacity.c:405-423).  A band-sharded build computed only this rank's columns of every row
   * o[r][t][m][.]: rank 0 writes the header, every rank writes its own column slices in place (file order has
   * the wavenumber innermost), and nobody goes on before all have written. */
  const long dims[4] = {nrows, Nt, n, (long)s->nwn};
  const off_t hdr = 4 * sizeof(long) + (off_t)nrows * sizeof(int) + 8 * ((off_t)Nt + n + s->nwn);
  if (th_rank == 0) {
    FILE *fp = fopen(h->f_opa, "wb");
    if (!fp) th_die("Opacity filename '%s' cannot be opened for writing.\n", h->f_opa);
    fwrite(dims, sizeof(long), 4, fp);
    fwrite(molID, sizeof(int), nrows, fp);
    fwrite(temp, 8, Nt, fp);
    for (int r = 0; r < n; r++) { const double pr = s->p[r] * s->p_fct; fwrite(&pr, 8, 1, fp); }
    for (int64_t w = 0; w < s->nwn; w++) { const double wn = s->wn_i + (double)w * s->wn_d; fwrite(&wn, 8, 1, fp); }
    if (th_world == 1) fwrite(grid, 8, ng, fp);
    fclose(fp);
  }
  if (th_world > 1) {
    if (TH_DEV(tb_comm_barrier(s->ctx))) th_die("%s\n", tb_last_error());      /* the header is there */
    const int fd = open(h->f_opa, O_WRONLY);
    if (fd < 0) th_die("Opacity filename '%s' cannot be opened for writing.\n", h->f_opa);
    const size_t nrow = (size_t)n * Nt * nrows, bw = (size_t)(s->band_hi - s->band_lo);
    for (size_t k = 0; k < nrow && bw; k++) {
      const off_t at = hdr + 8 * ((off_t)k * s->nwn + s->band_lo);
      if (pwrite(fd, grid + k * s->nwn + s->band_lo, 8 * bw, at) != (ssize_t)(8 * bw))
        th_die("Short write to opacity file '%s'.\n", h->f_opa);
    }
    close(fd);
    if (TH_DEV(tb_comm_barrier(s->ctx))) th_die("%s\n", tb_last_error());      /* the file is complete */
  }
  free(temp); free(ziso); free(grid);
  (void)nmol;
  s->have_grid_file = 1;
  return 0;
}
