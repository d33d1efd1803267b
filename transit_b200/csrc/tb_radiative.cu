// tb_radiative.cu -- CIA interpolation (crosssec.c:271-428), optical depth (tau.c:59-356 with
// slantpath.c:18-108 / eclipse.c:28-105), modulation (slantpath.c:274-436) and emergent
// intensity + flux (eclipse.c:117-287).
//
// Device layouts are wavenumber-innermost ([layer][wn], [ray][wn]) so that a warp of
// consecutive wavenumbers reads and writes coalesced; the taps in tb_context.cu transpose
// to the reference's [wn][layer] where it differs.  Ray geometry (the Simpson weights of
// numerical.c:390-525) depends only on the radii, so it is computed once per call by one
// thread per ray and shared by every wavenumber.
#include "tb_internal.cuh"

static inline int nblk(int64_t n, int t) { return (int)((n + t - 1) / t); }

// =============================================================== natural cubic splines

// tri + spline_init (pu/src/spline.c:12-48,186-206); scratch u,v of length n
__device__ void tb_spline_init(double *z, const double *x, const double *y, int n, double *u, double *v) {
  if (n > 2) {
    double h0 = x[1] - x[0], h1 = x[2] - x[1];
    double b0 = (y[1] - y[0]) / h0, b1 = (y[2] - y[1]) / h1;
    u[1] = 2 * (h1 + h0);
    v[1] = 6 * (b1 - b0);
    for (int i = 2; i < n - 1; i++) {
      double hi = x[i + 1] - x[i], hm = x[i] - x[i - 1];
      double bi = (y[i + 1] - y[i]) / hi, bm = (y[i] - y[i - 1]) / hm;
      u[i] = 2 * (hi + hm) - hm * hm / u[i - 1];
      v[i] = 6 * (bi - bm) - v[i - 1] * hm / u[i - 1];
    }
  }
  z[0] = z[n - 1] = 0;
  for (int i = n - 2; i > 0; i--) z[i] = (v[i] - (x[i + 1] - x[i]) * z[i + 1]) / u[i];
}

// splinterp_pt (spline.c:131-183)
__device__ double tb_splinterp_pt(const double *z, int N, const double *x, const double *y, double xout) {
  int idx = tb_binsearchapprox(x, xout, 0, N - 1);
  if (idx == N - 1 || xout < x[idx]) idx--;
  double x_lo = x[idx], x_hi = x[idx + 1], y_lo = y[idx], y_hi = y[idx + 1];
  double h = x_hi - x_lo, dy = y_hi - y_lo;
  if (x[idx] == xout) return y[idx];
  if (h > 0) {
    double dx = xout - x_lo;
    double a = (z[idx + 1] - z[idx]) / (6 * h), b = 0.5 * z[idx];
    double cc = dy / h - h / 6 * (z[idx + 1] + 2 * z[idx]);
    return y_lo + dx * (cc + dx * (b + dx * a));
  }
  return 0;
}

// =============================================================== CIA (a9)
//
// bicubicinterpolate (crosssec.c:353-428) is two families of natural cubic splines: along
// temperature for every table row, then along wavenumber for every layer.  What depends only
// on the table (second derivatives along T, the tridiagonal pivots u[] of the wavenumber
// spline, the table interval of every grid wavenumber) is computed once per table/grid; per
// spectrum there remain one evaluation per (row, layer), two first-order recurrences per
// layer and one cubic per (wavenumber, layer), all tables summed in one pass.

#define CIA_MAXT 64
#define CIA_MAXTAB 8

// per-table arguments of the two per-spectrum spline passes: every table goes in ONE launch
// (blockIdx.y = table) so that their latency-bound recurrences run side by side
struct CiaSplineTab {
  const double *tab, *zt, *ttemp, *h, *rh, *r, *ru;
  double *f2, *v, *z2;
  int ntw, ntt, fj, lj;
};
struct CiaSplinePack {
  CiaSplineTab t[CIA_MAXTAB];
};

// pass 1 (crosssec.c:407-412): f2[layer][row] = spline along T of table row `row` at the layer T
__global__ void k_cia_temp(CiaSplinePack P, const double *__restrict__ ltemp, int nrad) {
  const CiaSplineTab &T = P.t[blockIdx.y];
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T.ntw * nrad) return;
  const int j = t / T.ntw, i = t % T.ntw;
  if (j < T.fj || j >= T.lj) return;
  T.f2[t] = tb_splinterp_pt(T.zt + (size_t)i * T.ntt, T.ntt, T.ttemp, T.tab + (size_t)i * T.ntt, ltemp[j]);
}

// pass 2a (crosssec.c:415-416, spline.c:12-48): second derivatives along the table wavenumbers.
// Both sweeps of the tridiagonal solve are first-order linear recurrences
//     v[i] = c[i] - r[i] v[i-1],     z[i] = (v[i] - h[i] z[i+1]) ru[i]
// (r = h[i-1]/u[i-1], ru = 1/u: table-only, precomputed).  One warp per layer: every lane
// composes the affine map of its contiguous segment of rows, the 32 maps are chained through
// shuffles, and the lane then replays its segment from the right starting value.
__global__ void k_cia_wnspline(CiaSplinePack P) {
  const CiaSplineTab &T = P.t[blockIdx.y];
  const double *__restrict__ h = T.h, *__restrict__ rh = T.rh, *__restrict__ r = T.r, *__restrict__ ru = T.ru;
  const int ntw = T.ntw;
  const int lane = threadIdx.x & 31;
  const int j = T.fj + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  if (j >= T.lj) return;                                 // whole warp
  const double *__restrict__ y = T.f2 + (size_t)j * ntw;
  double *__restrict__ vv = T.v + (size_t)j * ntw, *__restrict__ zz = T.z2 + (size_t)j * ntw;
  const int L = (ntw - 2 + 31) / 32;                      // rows 1 .. ntw-2
  const int a = 1 + lane * L;
  const int b = a + L < ntw - 1 ? a + L : ntw - 1;
  const unsigned full = 0xffffffffu;
  // the loops below are first-order recurrences fed by loads that do not depend on them: unrolled
  // by 8 the loads of eight rows are in flight together instead of one L2 round trip per row
  // ---- forward sweep
  double A = 1.0, B = 0.0;
#pragma unroll 8
  for (int i = a; i < b; i++) {
    const double c = 6 * ((y[i + 1] - y[i]) * rh[i] - (y[i] - y[i - 1]) * rh[i - 1]);
    B = c - r[i] * B;
    A = -r[i] * A;
  }
  double carry = 0.0, mine = 0.0;
  for (int l = 0; l < 32; l++) {
    const double Al = __shfl_sync(full, A, l), Bl = __shfl_sync(full, B, l);
    if (lane == l) mine = carry;
    carry = Al * carry + Bl;
  }
  double vp = mine;
#pragma unroll 8
  for (int i = a; i < b; i++) {
    const double c = 6 * ((y[i + 1] - y[i]) * rh[i] - (y[i] - y[i - 1]) * rh[i - 1]);
    vp = c - r[i] * vp;
    vv[i] = vp;
  }
  // ---- backward sweep (each lane reads back only what it wrote)
  A = 1.0; B = 0.0;
#pragma unroll 8
  for (int i = b - 1; i >= a; i--) {
    B = (vv[i] - h[i] * B) * ru[i];
    A = -h[i] * ru[i] * A;
  }
  carry = 0.0; mine = 0.0;
  for (int l = 31; l >= 0; l--) {
    const double Al = __shfl_sync(full, A, l), Bl = __shfl_sync(full, B, l);
    if (lane == l) mine = carry;
    carry = Al * carry + Bl;
  }
  double zn = mine;
#pragma unroll 8
  for (int i = b - 1; i >= a; i--) {
    zn = (vv[i] - h[i] * zn) * ru[i];
    zz[i] = zn;
  }
  if (lane == 0) { zz[0] = 0; zz[ntw - 1] = 0; }
}

struct CiaDev {
  const double *wn, *f2, *z2, *densfac, *rh, *r6h, *h6;   // rh = 1/h, r6h = 1/(6h), h6 = h/6 per table interval
  const int32_t *widx;
  int fi, li, fj, lj, ntw;
};
struct CiaPack {
  CiaDev t[CIA_MAXTAB];
  int n;
};

// interval of every grid wavenumber in the table, as splinterp_pt picks it (spline.c:139-145)
__global__ void k_cia_widx(const double *__restrict__ twn, int ntw, double wn_i, double wn_d, double wn_fct,
                           int64_t nwn, int32_t *__restrict__ widx) {
  const int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (w >= nwn) return;
  const double x = wn_fct * __dadd_rn(wn_i, __dmul_rn((double)w, wn_d));
  int idx = tb_binsearchapprox(twn, x, 0, ntw - 1);
  if (idx == ntw - 1 || x < twn[idx]) idx--;
  widx[w] = idx < 0 ? 0 : idx;
}

// pass 2b (crosssec.c:417-419) + density scaling and positivity (crosssec.c:321-336), every
// table in file order; ecs[layer][wn]
__global__ void k_cia_eval(CiaPack P, int nrad, double wn_i, double wn_d, double wn_fct, int64_t nwn, int64_t w0,
                           int64_t w1, double *__restrict__ ecs) {
  const int64_t w = w0 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;      // columns [w0, w1): the band of a shard
  const int j = blockIdx.y;
  if (w >= w1) return;
  const double xout = wn_fct * __dadd_rn(wn_i, __dmul_rn((double)w, wn_d));
  double acc = 0.0;
  for (int n = 0; n < P.n; n++) {
    const CiaDev &T = P.t[n];
    if (w < T.fi || w >= T.li || j < T.fj || j >= T.lj) continue;
    const int idx = T.widx[w];
    const double x_lo = T.wn[idx], x_hi = T.wn[idx + 1];
    const double y_lo = T.f2[(size_t)j * T.ntw + idx], y_hi = T.f2[(size_t)j * T.ntw + idx + 1];
    const double z_lo = T.z2[(size_t)j * T.ntw + idx], z_hi = T.z2[(size_t)j * T.ntw + idx + 1];
    const double h = x_hi - x_lo, dy = y_hi - y_lo;
    double val = 0;
    if (x_lo == xout) val = y_lo;
    else if (h > 0) {                                   // splinterp_pt, spline.c:163-180
      // its three divisions have divisors that depend on the table interval only: 1/(6h), 1/h and h/6 are formed
      // once per table (correctly rounded), so the per-point cubic is multiplications -- each quotient within one
      // ulp of the reference's, the cross-section within a few 1e-16 (tolerance 1e-12)
      const double dx = xout - x_lo;
      const double a = (z_hi - z_lo) * T.r6h[idx], b = 0.5 * z_lo;
      const double cc = dy * T.rh[idx] - T.h6[idx] * (z_hi + 2 * z_lo);
      val = y_lo + dx * (cc + dx * (b + dx * a));
    }
    if (val > 0) acc += val * T.densfac[j];
  }
  ecs[(size_t)j * nwn + w] = acc;
}

extern "C" int tb_cia_clear(tb_ctx *c) {
  for (CiaTable *t : c->cia) {
    t->release();
    delete t;
  }
  c->cia.clear();
  c->have_ecs = false;
  return TB_OK;
}

extern "C" int tb_cia_add_table(tb_ctx *c, const double *wn, int nwn, const double *temp, int ntemp,
                                const double *tab, const int *mol, int nspec) {
  TB_REQUIRE(nwn >= 3 && ntemp >= 3 && ntemp <= CIA_MAXT, TB_ERR_ARG,
             "CIA table needs >=3 wavenumbers and 3..%d temperatures", CIA_MAXT);
  TB_REQUIRE(nspec >= 1 && nspec <= 2, TB_ERR_ARG, "CIA table has 1 or 2 species (crosssec.c:115-124)");
  TB_REQUIRE((int)c->cia.size() < CIA_MAXTAB, TB_ERR_ARG, "at most %d CIA tables", CIA_MAXTAB);
  for (int k = 0; k < nspec; k++)
    TB_REQUIRE(mol[k] >= 0 && mol[k] < c->nmol, TB_ERR_ARG, "CIA species index out of range");
  for (int i = 1; i < nwn; i++)
    TB_REQUIRE(wn[i] > wn[i - 1], TB_ERR_ARG, "CIA table wavenumbers must increase");
  TB_CUDA(cudaSetDevice(c->device));
  CiaTable *t = new CiaTable();
  t->nw = nwn; t->nt = ntemp; t->nspec = nspec;
  for (int k = 0; k < nspec; k++) t->mol[k] = mol[k];
  t->wn.assign(wn, wn + nwn);
  t->temp.assign(temp, temp + ntemp);
  t->tab.assign(tab, tab + (size_t)nwn * ntemp);
  // table-only parts of the two spline families, in the reference's operation order
  // (tri + spline_init, spline.c:12-48,186-206)
  std::vector<double> zt((size_t)nwn * ntemp, 0.0), h(nwn, 1.0), u(nwn, 1.0), uu(ntemp), vv(ntemp);
  std::vector<double> rh(nwn, 1.0), rr(nwn, 0.0), ru(nwn, 1.0);
  for (int i = 0; i < nwn; i++) {
    const double *y = tab + (size_t)i * ntemp;
    double *z = zt.data() + (size_t)i * ntemp;
    for (int k = 1; k < ntemp - 1; k++) {
      const double hk = temp[k + 1] - temp[k], hm = temp[k] - temp[k - 1];
      const double bk = (y[k + 1] - y[k]) / hk, bm = (y[k] - y[k - 1]) / hm;
      if (k == 1) { uu[k] = 2 * (hk + hm); vv[k] = 6 * (bk - bm); }
      else { uu[k] = 2 * (hk + hm) - hm * hm / uu[k - 1]; vv[k] = 6 * (bk - bm) - vv[k - 1] * hm / uu[k - 1]; }
    }
    z[0] = z[ntemp - 1] = 0;
    for (int k = ntemp - 2; k > 0; k--) z[k] = (vv[k] - (temp[k + 1] - temp[k]) * z[k + 1]) / uu[k];
  }
  for (int i = 0; i < nwn - 1; i++) h[i] = wn[i + 1] - wn[i];
  for (int i = 1; i < nwn - 1; i++)
    u[i] = (i == 1) ? 2 * (h[1] + h[0]) : 2 * (h[i] + h[i - 1]) - h[i - 1] * h[i - 1] / u[i - 1];
  std::vector<double> r6h(nwn, 0.0), h6(nwn, 0.0);
  for (int i = 0; i < nwn - 1; i++) { rh[i] = 1.0 / h[i]; r6h[i] = 1.0 / (6 * h[i]); h6[i] = h[i] / 6; }
  for (int i = 1; i < nwn - 1; i++) { ru[i] = 1.0 / u[i]; if (i > 1) rr[i] = h[i - 1] / u[i - 1]; }
  if (t->d_wn.ensure(nwn) || t->d_temp.ensure(ntemp) || t->d_tab.ensure((size_t)nwn * ntemp) ||
      t->d_zt.ensure((size_t)nwn * ntemp) || t->d_h.ensure(nwn) || t->d_u.ensure(nwn) || t->d_rh.ensure(nwn) ||
      t->d_r.ensure(nwn) || t->d_ru.ensure(nwn) || t->d_r6h.ensure(nwn) || t->d_h6.ensure(nwn)) {
    t->release();
    delete t;
    return TB_ERR_NOMEM;
  }
  cudaStream_t st = c->stream;
  TB_CUDA(cudaMemcpyAsync(t->d_wn.p, wn, nwn * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(t->d_temp.p, temp, ntemp * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(t->d_tab.p, tab, (size_t)nwn * ntemp * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(t->d_zt.p, zt.data(), zt.size() * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(t->d_h.p, h.data(), nwn * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(t->d_u.p, u.data(), nwn * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(t->d_rh.p, rh.data(), nwn * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(t->d_r.p, rr.data(), nwn * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(t->d_ru.p, ru.data(), nwn * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(t->d_r6h.p, r6h.data(), nwn * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(t->d_h6.p, h6.data(), nwn * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaStreamSynchronize(st));
  c->cia.push_back(t);
  c->have_ecs = false;
  return TB_OK;
}

// static per (table, grid): the table's wavenumber range on the grid and the table interval of every grid point
static int tb_cia_table_grid(tb_ctx *c, CiaTable *t) {
  if (t->grid_version == c->grid_version) return TB_OK;
  const int64_t nwn = c->nwn;
  auto gw = [&](int64_t k) { return c->wn_fct * (c->wn_i + (double)k * c->wn_d); };
  const double fx1 = t->wn[0], lx1 = t->wn[t->nw - 1];
  t->fi = 0; t->li = nwn;
  if (gw(0) > lx1 || gw(nwn - 1) < fx1) t->li = 0;
  else {
    while (t->fi < nwn && gw(t->fi) < fx1) t->fi++;
    for (int64_t i = 0; i < t->li; i++) if (gw(i) > lx1) t->li = i;
  }
  TB_TRY(t->d_widx.ensure(nwn));
  k_cia_widx<<<nblk(nwn, 256), 256, 0, c->stream>>>(t->d_wn.p, t->nw, c->wn_i, c->wn_d, c->wn_fct, nwn, t->d_widx.p);
  TB_LAUNCHED(c);
  t->grid_version = c->grid_version;
  return TB_OK;
}

int tb_cia_run(tb_ctx *c) {
  TB_REQUIRE(c->have_atm && c->have_grid, TB_ERR_STATE, "grid and atmosphere must be set before tb_cia_interp");
  TB_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int nrad = c->nrad, ntab = (int)c->cia.size();
  const int64_t nwn = c->nwn;
  TB_TRY(c->d_ecs.ensure((size_t)nrad * nwn));
  if (ntab == 0) {
    TB_CUDA(cudaMemsetAsync(c->d_ecs.p, 0, sizeof(double) * nrad * nwn, st));
    c->have_ecs = true;
    return TB_OK;
  }
  // host side, O(nrad * ntab): layer temperatures, density factors, layer ranges
  std::vector<double> stage((size_t)nrad * (1 + ntab));
  double *ltemp = stage.data();
  for (int i = 0; i < nrad; i++) ltemp[i] = c->t_fct * c->t[i];
  {  // interpcs refuses a layer outside the range common to all tables (crosssec.c:174-176,293-309);
     // a silently clipped layer would come back with zero CIA and a plausible, wrong spectrum
    double tmin = 0.0, tmax = 70000.0;
    for (int n = 0; n < ntab; n++) {
      tmin = fmax(tmin, c->cia[n]->temp[0]);
      tmax = fmin(tmax, c->cia[n]->temp[c->cia[n]->nt - 1]);
    }
    for (int i = 0; i < nrad; i++) {
      TB_REQUIRE(!(ltemp[i] < tmin), TB_ERR_DOMAIN,
                 "The layer %d in the atmospheric model has a lower temperature (%.1f K) than the lowest allowed "
                 "cross-section temperature (%.1f K).", i, ltemp[i], tmin);
      TB_REQUIRE(!(ltemp[i] > tmax), TB_ERR_DOMAIN,
                 "The layer %d in the atmospheric model has a higher temperature (%.1f K) than the highest allowed "
                 "cross-section temperature (%.1f K).", i, ltemp[i], tmax);
    }
  }
  TB_TRY(c->c_stage.ensure(stage.size()));
  CiaPack P;
  P.n = 0;
  auto gw = [&](int64_t k) { return c->wn_fct * (c->wn_i + (double)k * c->wn_d); };
  for (int n = 0; n < ntab; n++) {
    CiaTable *t = c->cia[n];
    // bicubicinterpolate's range logic (crosssec.c:366-393)
    const double fx1 = t->wn[0], lx1 = t->wn[t->nw - 1], fx2 = t->temp[0], lx2 = t->temp[t->nt - 1];
    TB_TRY(tb_cia_table_grid(c, t));
    if (t->fi >= t->li || ltemp[0] > lx2 || ltemp[nrad - 1] < fx2) continue;
    int fj = 0, lj = nrad;
    while (fj < nrad && ltemp[fj] < fx2) fj++;
    for (int j = 0; j < lj; j++) if (ltemp[j] > lx2) lj = j;
    if (fj >= lj) continue;
    double *df = stage.data() + (size_t)nrad * (1 + n);
    for (int i = 0; i < nrad; i++) {
      df[i] = 1.0;
      for (int k = 0; k < t->nspec; k++) {
        const int m = t->mol[k];
        df[i] *= c->dens[(size_t)m * nrad + i] / (TB_AMU * c->mol_mass[m] * TB_AMAGAT);
      }
    }
    TB_TRY(t->d_f2.ensure((size_t)nrad * t->nw)); TB_TRY(t->d_z2.ensure((size_t)nrad * t->nw));
    TB_TRY(t->d_v.ensure((size_t)nrad * t->nw));
    CiaDev &D = P.t[P.n++];
    D.wn = t->d_wn.p; D.f2 = t->d_f2.p; D.z2 = t->d_z2.p; D.widx = t->d_widx.p;
    D.rh = t->d_rh.p; D.r6h = t->d_r6h.p; D.h6 = t->d_h6.p;
    D.densfac = c->c_stage.p + (size_t)nrad * (1 + n);
    D.fi = (int)t->fi; D.li = (int)t->li; D.fj = fj; D.lj = lj; D.ntw = t->nw;
  }
  TB_TRY(tb_upload_small(c, c->c_stage.p, stage.data(), stage.size() * sizeof(double), st));
  tb_timer_begin(c, "cia");
  if (P.n > 0) {
    CiaSplinePack S;
    int k = 0, maxrows = 0, maxlayers = 0;
    for (int n = 0; n < ntab && k < P.n; n++) {
      CiaTable *t = c->cia[n];
      if (P.t[k].f2 != t->d_f2.p) continue;
      const CiaDev &D = P.t[k];
      CiaSplineTab &T = S.t[k++];
      T.tab = t->d_tab.p; T.zt = t->d_zt.p; T.ttemp = t->d_temp.p; T.h = t->d_h.p; T.rh = t->d_rh.p;
      T.r = t->d_r.p; T.ru = t->d_ru.p; T.f2 = t->d_f2.p; T.v = t->d_v.p; T.z2 = t->d_z2.p;
      T.ntw = t->nw; T.ntt = t->nt; T.fj = D.fj; T.lj = D.lj;
      if (t->nw > maxrows) maxrows = t->nw;
      if (D.lj - D.fj > maxlayers) maxlayers = D.lj - D.fj;
    }
    k_cia_temp<<<dim3(nblk((int64_t)maxrows * nrad, 256), k), 256, 0, st>>>(S, c->c_stage.p, nrad);
    k_cia_wnspline<<<dim3(nblk(maxlayers, 4), k), 128, 0, st>>>(S);
    c->launches += 2;
  }
  int64_t w0, w1;
  tb_band_columns(c, &w0, &w1);
  if (w1 > w0) {
    dim3 g(nblk(w1 - w0, 256), nrad);
    k_cia_eval<<<g, 256, 0, st>>>(P, nrad, c->wn_i, c->wn_d, c->wn_fct, nwn, w0, w1, c->d_ecs.p);
    TB_LAUNCHED(c);
  }
  tb_timer_end(c, "cia");
  TB_CUDA(cudaGetLastError());
  c->have_ecs = true;
  return TB_OK;
}

extern "C" int tb_cia_interp(tb_ctx *c) {
  TB_TRY(tb_cia_run(c));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  return TB_OK;
}

// =============================================================== ray geometry

#define TAU_RB 4     // rays per register block of k_tau_w
// meta[ray*4]: {rs, npts, mode, npairs}; mode 0: tau = 0, 1: normal, 2: two-point case, -1: error
// par[ray*4]:  {dx, x0, xr, trap_h0}
// coef[ray*stride + 4*i]: {2-hratio, hfactor, 2-1/hratio, hsum} (geth/simpson, numerical.c:390-525)
__global__ void k_ray_geometry(int eclipse, const double *__restrict__ rad, int nrad, double rfct,
                               const double *__restrict__ h, int nh, double hfct, int stride,
                               int32_t *__restrict__ meta, double *__restrict__ par,
                               double *__restrict__ coef, double *__restrict__ sbuf,
                               double *__restrict__ rcp, int32_t *__restrict__ err,
                               double *__restrict__ wg, double *__restrict__ w0) {
  // one block per ray: thread 0 places the ray, then the path coordinates s[i] (one sqrt each) and the
  // Simpson weights of the pairs (three divisions each) are independent and spread over the block
  // (one thread per ray ran them as a 27 us serial chain in front of every tau launch)
  // a batch of atmospheres (tb_batch.cu) launches natm*nh blocks: block a*nh + ri is ray ri of atmosphere a, whose
  // radii and impact parameters are the a-th slices; every output is indexed by the block
  const int ri = blockIdx.x;
  rad += (size_t)(ri / nh) * nrad;                             // h[natm][nh] is indexed by the block as it is
  __shared__ int sh_rs, sh_mode, sh_n;
  __shared__ double sh_r0;
  int32_t *mt = meta + ri * 4;
  double *pp = par + ri * 4;
  double *s = sbuf + (size_t)ri * (nrad + 3);
  if (threadIdx.x == 0) {
    // tau.c:274 as the reference is BUILT: gcc -ffast-math hoists 1/rfct out of the loop (objdump of
    // tau(): movsd 1.0; divsd rfct ahead of the wavenumber loop), so b = (h*hfct)*(1/rfct).  With an
    // IEEE division some impact parameters land one ulp below their own layer radius, the ray then
    // starts one layer lower and tau changes by ~20%: something the reference binary never shows.
    const double b = __dmul_rn(__dmul_rn(h[ri], hfct), __drcp_rn(rfct));
    int rs, mode = 1;
    if (eclipse) {
      rs = tb_binsearchapprox(rad, b, 0, nrad - 1);             // eclipse.c:39
      if (rs == nrad - 1) mode = 0;
    } else {
      // binsearchie (numerical.c:16-45)
      const double r0 = b;                                       // refraction index 1 (slantpath.c:32)
      if (rad[0] > r0) { mode = -1; rs = 0; }
      else if (rad[nrad - 1] < r0) { mode = 0; rs = 0; }         // -2: above the top -> 0
      else if (rad[nrad - 1] == r0) { mode = 0; rs = 0; }        // -5: outermost layer -> 0
      else {
        long i = 0, f = nrad - 1;
        while (f - i > 1) {
          long m = (f + i) >> 1;
          if (rad[m] > r0) f = m; else i = m;
        }
        rs = (int)i;
      }
    }
    int n = 0;
    double r0 = 0;
    if (mode <= 0) {
      if (mode < 0) atomicExch(err + 1, 1);
      mt[0] = rs; mt[1] = 0; mt[2] = mode; mt[3] = 0;
    } else {
      n = nrad - rs;
      r0 = eclipse ? rad[rs] : b;
      if (n == 2) {
        mode = 2;
        // interp_parab(rad-1, ex-1, r0): parabola through layers rs-1, rs, rs+1
        const double *x = rad + rs - 1;
        pp[0] = x[1] - x[0];
        pp[1] = x[0] / pp[0];
        pp[2] = r0;
        // three points: r0, midpoint, rad[rs+1] (slantpath.c:62-74, eclipse.c:68-80)
        double r3_0 = eclipse ? rad[rs] : r0, r3_2 = rad[rs + 1], r3_1 = (r3_0 + r3_2) / 2.0;
        s[0] = 0;
        if (eclipse) { s[1] = s[0] + (r3_1 - r3_0); s[2] = s[1] + (r3_2 - r3_1); }
        else { s[1] = sqrt(r3_1 * r3_1 - r0 * r0); s[2] = sqrt(r3_2 * r3_2 - r0 * r0); }
        n = 3;
      } else {
        const double *x = rad + rs;
        pp[0] = x[1] - x[0];
        pp[1] = x[0] / pp[0];
        pp[2] = r0;
        s[0] = 0;
        if (eclipse)                                             // a running sum: in order
          for (int i = 1; i < n; i++) s[i] = s[i - 1] + (rad[rs + i] - rad[rs + i - 1]);
      }
      rcp[ri * 2] = 1.0 / (2.0 * pp[0] * pp[0]);      // correctly rounded reciprocals for tb_div_r
      rcp[ri * 2 + 1] = 1.0 / pp[0];
    }
    sh_rs = rs; sh_mode = mode; sh_n = n; sh_r0 = r0;
  }
  __syncthreads();
  const int mode = sh_mode, n = sh_n, rs = sh_rs;
  // per-layer weights of the ray for k_tau_w (see there): rays in groups of four, wgr[layer][ray % 4]
  double *wgr = wg ? wg + ((size_t)(ri / TAU_RB) * nrad) * TAU_RB + (ri % TAU_RB) : nullptr;
  if (mode <= 0) {
    if (wgr) {
      for (int L = threadIdx.x; L < nrad; L += blockDim.x) wgr[(size_t)L * TAU_RB] = 0.0;
      if (threadIdx.x == 0) w0[ri] = 0.0;
    }
    return;
  }
  if (mode == 1 && !eclipse) {
    const double r0 = sh_r0;
    for (int i = 1 + threadIdx.x; i < n; i += blockDim.x) {
      double ri_ = rad[rs + i];
      s[i] = sqrt(__dsub_rn(__dmul_rn(ri_, ri_), __dmul_rn(r0, r0)));
    }
  }
  __syncthreads();
  const int even = (n % 2 == 0);
  const int np = (n - 1) / 2;
  double *cf = coef + (size_t)ri * stride;
  for (int i = threadIdx.x; i < np; i += blockDim.x) {
    int j = 2 * i + even;
    double h0 = s[j + 1] - s[j], h1 = s[j + 2] - s[j + 1];
    double hsum = h0 + h1, hratio = h1 / h0, hfactor = hsum * hsum / (h0 * h1);
    cf[4 * i + 0] = 2.0 - hratio;
    cf[4 * i + 1] = hfactor;
    cf[4 * i + 2] = 2.0 - 1.0 / hratio;
    cf[4 * i + 3] = hsum;
  }
  if (threadIdx.x == 0) {
    pp[3] = even ? (s[1] - s[0]) : 0.0;
    mt[0] = rs; mt[1] = n; mt[2] = mode; mt[3] = np;
  }
  if (!wgr) return;
  __syncthreads();                                   // cf[] of this block is complete
  // simpson() (numerical.c:454-525) is linear in the samples f_k = er[rs + k] (f_0 = the parabola value):
  //   sum_i (f_a c0_i + f_a+1 c1_i + f_a+2 c2_i) c3_i / 6  [+ h0 (f_0 + f_1)/2]  =  sum_k W_k f_k
  // W_k depends on the ray only, not on the wavenumber; layer L = rs + k carries W_k, every other layer zero.
  const double h0t = even ? (s[1] - s[0]) : 0.0;
  for (int L = threadIdx.x; L < nrad; L += blockDim.x) {
    double W = 0.0;
    const int k = L - rs;
    if (mode == 2) {                                 // two-point ray: samples y0, (y0 + yb)/2, yb = er[rs + 1]
      if (k == 0) W = (cf[0] + cf[1] / 2.0) * cf[3] / 6.0;
      else if (k == 1) W = (cf[2] + cf[1] / 2.0) * cf[3] / 6.0;
    } else if (k >= 0 && k < n) {
      if (k >= even) {
        const int m = k - even;
        if ((m & 1) == 0) {
          const int i = m >> 1;
          if (i < np) W += cf[4 * i] * cf[4 * i + 3];
          if (i >= 1) W += cf[4 * (i - 1) + 2] * cf[4 * (i - 1) + 3];
        } else {
          const int i = (m - 1) >> 1;
          W += cf[4 * i + 1] * cf[4 * i + 3];
        }
      }
      W /= 6.0;
      if (even && k <= 1) W += h0t / 2.0;
    }
    if (k == 0) { w0[ri] = W; W = 0.0; }             // the parabola value's weight is kept apart
    wgr[(size_t)L * TAU_RB] = W;
  }
}

// =============================================================== tau (a10-a13)

struct TauPar {
  int eclipse, nrad, nh, stride;
  int64_t nwn, w0, w1;                 // columns [w0, w1) are computed (the band of a shard, else the grid)
  double rfct, toomuch, wn_i, wn_d, wn_fct;
  int sc_flag, cl_flag;
  double sc_fac;                       // pow(10, logext) * E0H2
  double cl_ext, cl_top, cl_bot, cl_gamma, cl_Q, cl_r, cl_sig, cl_refwn;
};


// interp_parab (numerical.c:182-195) with the per-ray constants dx, x0 = x[0]/dx and the
// reciprocals rcp[0] = 1/(2 dx^2), rcp[1] = 1/dx (k_ray_geometry)
__device__ __forceinline__ double tb_parab(double dx, double x0, double xr, const double *__restrict__ rcp,
                                           double y0, double y1, double y2) {
  const double my = y0 + y2 - 2 * y1;
  const double a = tb_div_r(my, 2.0 * dx * dx, rcp[0]);
  const double b = tb_div_r(y2 - y1 - (x0 + 1.5) * my, dx, rcp[1]);
  const double cc = y0 + x0 * (y2 - 4 * y1 + 3 * y0 + x0 * my) / 2.0;
  return xr * xr * a + xr * b + cc;
}

// The production layout, both geometries: a block owns 32 wavenumbers (one per lane) and
// TAU_NW warps share the rays (ray ri -> warp ri % TAU_NW), all reading the block's extinction
// columns er[layer][lane] from shared memory.  totaltau1 restores what it modifies
// (slantpath.c:98-99), so transit rays are independent.  eclipsetau does NOT restore the
// extrapolated value it stores at its starting layer (eclipse.c:62-66): ray ri leaves
// er[rs] = parabola(er[rs], er[rs+1], er[rs+2]) behind for the rays below it.  That chain is
// O(layers) per wavenumber and touches each layer once, top down, so warp 0 runs it first
// (the two-point ray next to the top, which reads the not-yet-modified layer below it, is
// evaluated inside the chain); after that every ray's Simpson sum reads only final values and
// the rays are independent again.  The reference's early exit (tau.c:277-287) becomes: each warp
// stops a lane at its first ray above `toomuch`, the block takes the minimum over warps as
// last[w] and clears what was written beyond it (the reference leaves zeros there).
#define TAU_NW 8
__global__ void __launch_bounds__(32 * TAU_NW, 4)
k_tau_transit(TauPar P, const double *__restrict__ e, const double *__restrict__ ecs,
              const double *__restrict__ press, const double *__restrict__ temp,
              const double *__restrict__ aux, const double *__restrict__ mol_d,
              const double *__restrict__ mol_pol, const double *__restrict__ mol_mass, int nmol,
              const int32_t *__restrict__ meta, const double *__restrict__ par,
              const double *__restrict__ coef, const double *__restrict__ rcp, double *__restrict__ tau,
              int32_t *__restrict__ last) {
  extern __shared__ double sm[];
  __shared__ int sh_last[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  {  // a batch of atmospheres (tb_batch.cu): blockIdx.y selects the atmosphere's slices
    const size_t a = blockIdx.y;
    e += a * P.nrad * P.nwn; ecs += a * P.nrad * P.nwn; temp += a * P.nrad; aux += a * 2 * P.nrad;
    mol_d += a * (size_t)nmol * P.nrad; meta += a * P.nh * 4; par += a * P.nh * 4; coef += a * (size_t)P.nh * P.stride;
    rcp += a * P.nh * 2; tau += a * P.nh * P.nwn; last += a * P.nwn;
  }
  const int64_t w = P.w0 + blockIdx.x * 32LL + lane;
  const bool inb = w < P.w1;
  const int nrad = P.nrad;
  if (warp == 0) sh_last[lane] = P.nh;
  if (inb) {
    const double wn = P.wn_fct * __dadd_rn(P.wn_i, __dmul_rn((double)w, P.wn_d));
    double kBP = 0, kFH = 0, refw = 1;
    int ctop = -1, cbot = -1;   // cloud occupies layers (cbot, ctop]
    if (P.cl_flag && P.cl_ext != 0) {
      const double top = pow(10.0, P.cl_top), bot = pow(10.0, P.cl_bot);
      int i = nrad - 1;
      for (; i >= 0; i--) if (press[i] >= top) break;
      ctop = i;
      for (; i >= 0; i--) if (press[i] >= bot) break;
      cbot = i;
      const double x = 2 * TB_PI * P.cl_r * wn;
      refw = pow(P.cl_refwn, P.cl_gamma);
      kBP = P.cl_ext * pow(wn, P.cl_gamma);
      kFH = P.cl_ext / (P.cl_Q * pow(x, -1 * P.cl_gamma) + pow(x, 0.2));
    }
    const double wn4 = P.sc_flag == 1 ? pow(wn, 4.0) : 0.0;
    for (int r = warp; r < nrad; r += TAU_NW) {
      double es = 0, ec = 0;
      if (P.sc_flag == 1) es = P.sc_fac * press[r] / temp[r] * wn4;
      else if (P.sc_flag == 2) {
        for (int j = 0; j < nmol; j++)
          es += TB_PI * 8e-32 / 3. * pow(mol_pol[j], 2.0) * pow(2. * TB_PI * wn * TB_MICRON, 4.0) *
                mol_d[(size_t)j * nrad + r] / mol_mass[j] * TB_NAVOGADRO;
      }
      if (r <= ctop && r > cbot) {
        const double md = aux[nrad + r];
        switch (P.cl_flag) {
          case 1: ec = P.cl_ext; break;
          case 2: ec = P.cl_ext * md; break;
          case 3: ec = kBP * md; break;
          case 4: ec = kFH * md; break;
          case 5: ec = aux[r] * kBP * P.cl_sig / refw * md; break;
        }
      }
      // tau.c:231-232, same left-to-right sum
      sm[r * 32 + lane] = e[(size_t)r * P.nwn + w] + es + ec + ecs[(size_t)r * P.nwn + w];
    }
  }
  __shared__ double sh_t2[32];
  __shared__ volatile int sh_prog;          // eclipse: rays whose starting layer the chain has passed
  if (threadIdx.x == 0) sh_prog = 0;
  __syncthreads();
  // eclipse: warp 0 is the PRODUCER of the extrapolation chain, warps 1.. consume: ray ri may start as
  // soon as the chain has stored its starting layer (everything above it is final by then), so the
  // O(layers) dependent chain overlaps the Simpson sums instead of preceding them
  const int w0 = P.eclipse ? 1 : 0, nrw = TAU_NW - w0;
  if (P.eclipse && warp == 0) {
    double *ew = sm + lane;
    for (int ri = 0; ri < P.nh; ri++) {
      if (inb) {
        const int32_t *mt = meta + ri * 4;
        const int mode = mt[2], rs = mt[0];
        const double *pp = par + ri * 4;
        if (mode == 1) {
          ew[rs * 32] = tb_parab(pp[0], pp[1], pp[2], rcp + ri * 2, ew[rs * 32], ew[(rs + 1) * 32], ew[(rs + 2) * 32]);
        } else if (mode == 2) {
          const double *cf = coef + (size_t)ri * P.stride;
          const double y0 = tb_parab(pp[0], pp[1], pp[2], rcp + ri * 2, ew[(rs - 1) * 32], ew[rs * 32], ew[(rs + 1) * 32]);
          const double yb = ew[(rs + 1) * 32];
          const double ya = (y0 + yb) / 2.0;
          sh_t2[lane] = (y0 * cf[0] + ya * cf[1] + yb * cf[2]) * cf[3] / 6.0;
        }
      }
      __syncwarp();
      if (lane == 0) { __threadfence_block(); sh_prog = ri + 1; }
    }
  }
  const double *er = sm + lane;
  int myfirst = P.nh;                       // first ray of this warp above toomuch (none yet)
  // Simpson coefficients of a ray: one row of `coef`, the same for every wavenumber.  Read pair by
  // pair from global memory they cost an L2 round trip per pair (the rows of all rays do not fit
  // beside 25 KB of extinction columns in L1); instead the warp fetches the whole row of its NEXT ray
  // coalesced into registers while it integrates the current one out of shared memory.
  double *cw = sm + (size_t)nrad * 32 + (size_t)warp * P.stride;
  constexpr int CFR = 10;                   // registers per lane for one row (stride <= 320)
  double pre[CFR];
  auto fetch = [&](int ri) {
    if (ri < P.nh) {
      const double *cf = coef + (size_t)ri * P.stride;
      const int n4 = 4 * meta[ri * 4 + 3];
#pragma unroll
      for (int k = 0; k < CFR; k++) pre[k] = (lane + 32 * k < n4) ? cf[lane + 32 * k] : 0.0;
    }
  };
  if (warp >= w0) fetch(warp - w0);
  for (int ri = warp - w0; ri < P.nh && warp >= w0; ri += nrw) {     // whole warp: lanes outside the grid idle
    const int32_t *mt = meta + ri * 4;
    const int mode = mt[2];
    const int n4 = 4 * mt[3];
    __syncwarp();
#pragma unroll
    for (int k = 0; k < CFR; k++) if (lane + 32 * k < n4) cw[lane + 32 * k] = pre[k];
    __syncwarp();
    fetch(ri + nrw);
    if (P.eclipse) {                          // wait for the chain to have passed this ray's starting layer
      while (sh_prog <= ri) { }
      __threadfence_block();
    }
    if (inb && myfirst == P.nh) {
      double t = 0.0;
      if (mode > 0) {
        const int rs = mt[0], n = mt[1], np = mt[3];
        const double *pp = par + ri * 4;
        const double *cf = cw;
        if (mode == 2 && P.eclipse) {
          t = sh_t2[lane];                                   // evaluated inside the chain (see above)
        } else if (mode == 2) {
          const double y0 = tb_parab(pp[0], pp[1], pp[2], rcp + ri * 2, er[(rs - 1) * 32], er[rs * 32], er[(rs + 1) * 32]);
          const double yb = er[(rs + 1) * 32];
          const double ya = (y0 + yb) / 2.0;
          t = (y0 * cf[0] + ya * cf[1] + yb * cf[2]) * cf[3] / 6.0;
        } else {
          const double y0 = P.eclipse ? er[rs * 32]          // the chain left it there
                                      : tb_parab(pp[0], pp[1], pp[2], rcp + ri * 2, er[rs * 32], er[(rs + 1) * 32], er[(rs + 2) * 32]);
          const int even = (n % 2 == 0);
          double res = 0.0;
          // same terms in the same order as simpson() (numerical.c:454-525); the third sample of a pair
          // is the first of the next and stays in a register, the four weights come as two 128-bit reads
          double a0 = even ? er[(rs + 1) * 32] : y0;
          const double *ep = er + (rs + even + 1) * 32;
          const double2 *c2 = reinterpret_cast<const double2 *>(cf);
#pragma unroll 2
          for (int i = 0; i < np; i++) {
            const double2 c01 = c2[2 * i], c23 = c2[2 * i + 1];
            const double e1 = ep[0], e2 = ep[32];
            ep += 64;
            res += (a0 * c01.x + e1 * c01.y + e2 * c23.x) * c23.y;
            a0 = e2;
          }
          t = res / 6.0;
          if (even) t += pp[3] * (y0 + er[(rs + 1) * 32]) / 2;
        }
        t = P.rfct * (P.eclipse ? t : 2 * t);
      }
      tau[(size_t)ri * P.nwn + w] = t;
      if (t > P.toomuch) myfirst = ri;                   // tau.c:277-287: this lane is done
    }
    if (__all_sync(0xffffffffu, !inb || myfirst != P.nh)) break;
  }
  if (inb) {
    atomicMin(&sh_last[lane], myfirst);
  }
  __syncthreads();
  if (inb) {
    const int lst = sh_last[lane];                       // == nh: never above toomuch
    if (warp == 0) last[w] = lst < P.nh ? lst : P.nh - 1;
    // clear what this warp wrote beyond the block-wide first exceed
    const int upto = myfirst < P.nh ? myfirst : P.nh - 1;
    for (int ri = warp - w0; ri <= upto && warp >= w0; ri += nrw)
      if (ri > lst) tau[(size_t)ri * P.nwn + w] = 0.0;
  }
}


// ---- eclipse geometry in O(layers) per wavenumber --------------------------------------------------
// eclipsetau (eclipse.c:28-105) integrates along the VERTICAL: the path coordinate of ray ri is a running
// sum of layer spacings, whichever layer the ray starts from, so the Simpson weights of the pair of
// intervals (a, a+1, a+2) are the same for every ray that contains it (to rounding of the running sums),
// every ray's pairs end at the top layer, and the rays take their starting layers top down, one layer
// each.  The O(layers^2) work of the general kernel therefore collapses: walking the rays top down,
//   * the unrestored extrapolation chain stores er[rs] = parabola(er[rs], er[rs+1], er[rs+2]) (eclipse.c:62-66);
//   * a ray with an odd number of points adds ONE new pair -- its own first pair, with its own
//     coefficients (k_ray_geometry) -- to the running sum S of the pairs above it;
//   * a ray with an even number of points has exactly the pairs of the ray above it plus the trapezoid of
//     its first interval (numerical.c:454-525).
// One thread per wavenumber, no shared extinction columns, the inputs streamed once: 24 B per (layer,
// wavenumber), the HBM figure of SURVEY 8d.  The pairs of a ray are summed from the top down instead of
// the reference's bottom up: same terms, a different association (tau within 1e-9: tested).
struct EclRay {           // per ray, wavenumber-independent (from k_ray_geometry's tables)
  double dx, x0, xr, rcp0, rcp1;      // interp_parab constants of the ray's starting layer
  double c0, c1, c2, c3;              // its first pair of intervals
  double h0;                          // first interval (trapezoid of an even point count)
  int mode, n;
  int pad0, pad1;
};

__global__ void k_ecl_rays(int nh, int nrad, int stride, const int32_t *__restrict__ meta,
                           const double *__restrict__ par, const double *__restrict__ coef,
                           const double *__restrict__ rcp, EclRay *__restrict__ rays, int32_t *__restrict__ err) {
  const int ri = blockIdx.x * blockDim.x + threadIdx.x;
  if (ri >= nh) return;
  const int32_t *mt = meta + ri * 4;
  EclRay r;
  r.mode = mt[2]; r.n = mt[1];
  r.dx = par[ri * 4]; r.x0 = par[ri * 4 + 1]; r.xr = par[ri * 4 + 2]; r.h0 = par[ri * 4 + 3];
  r.rcp0 = rcp[ri * 2]; r.rcp1 = rcp[ri * 2 + 1];
  const double *cf = coef + (size_t)ri * stride;
  r.c0 = cf[0]; r.c1 = cf[1]; r.c2 = cf[2]; r.c3 = cf[3];
  r.pad0 = r.pad1 = 0;
  // the streaming kernel relies on ray ri starting at layer nrad-1-ri (the rays ARE the layers, tau.c:92-98)
  if (r.mode > 0 && mt[0] != nrad - 1 - ri) atomicExch(err, 1);
  if (r.mode == 0 && ri != 0) atomicExch(err, 1);
  rays[ri] = r;
}

__device__ __forceinline__ double tb_extra_ext(const TauPar &P, int r, double wn, double wn4, double kBP, double kFH,
                                               double refw, int ctop, int cbot, const double *__restrict__ press,
                                               const double *__restrict__ temp, const double *__restrict__ aux,
                                               const double *__restrict__ mol_d, const double *__restrict__ mol_pol,
                                               const double *__restrict__ mol_mass, int nmol, double &ec_out) {
  double es = 0, ec = 0;
  const int nrad = P.nrad;
  if (P.sc_flag == 1) es = P.sc_fac * press[r] / temp[r] * wn4;
  else if (P.sc_flag == 2) {
    for (int j = 0; j < nmol; j++)
      es += TB_PI * 8e-32 / 3. * pow(mol_pol[j], 2.0) * pow(2. * TB_PI * wn * TB_MICRON, 4.0) *
            mol_d[(size_t)j * nrad + r] / mol_mass[j] * TB_NAVOGADRO;
  }
  if (r <= ctop && r > cbot) {
    const double md = aux[nrad + r];
    switch (P.cl_flag) {
      case 1: ec = P.cl_ext; break;
      case 2: ec = P.cl_ext * md; break;
      case 3: ec = kBP * md; break;
      case 4: ec = kFH * md; break;
      case 5: ec = aux[r] * kBP * P.cl_sig / refw * md; break;
    }
  }
  ec_out = ec;
  return es;
}

__global__ void __launch_bounds__(128)
k_tau_eclipse(TauPar P, const double *__restrict__ e, const double *__restrict__ ecs,
              const double *__restrict__ press, const double *__restrict__ temp,
              const double *__restrict__ aux, const double *__restrict__ mol_d,
              const double *__restrict__ mol_pol, const double *__restrict__ mol_mass, int nmol,
              const EclRay *__restrict__ rays, double *__restrict__ tau, int32_t *__restrict__ last) {
  extern __shared__ __align__(16) unsigned char esm[];
  EclRay *sr = reinterpret_cast<EclRay *>(esm);
  for (int i = threadIdx.x; i < P.nh * (int)(sizeof(EclRay) / 8); i += blockDim.x)
    reinterpret_cast<double *>(sr)[i] = reinterpret_cast<const double *>(rays)[i];
  __syncthreads();
  const int64_t w = P.w0 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (w >= P.w1) return;
  const int nrad = P.nrad;
  const double wn = P.wn_fct * __dadd_rn(P.wn_i, __dmul_rn((double)w, P.wn_d));
  double kBP = 0, kFH = 0, refw = 1;
  int ctop = -1, cbot = -1;
  if (P.cl_flag && P.cl_ext != 0) {
    const double top = pow(10.0, P.cl_top), bot = pow(10.0, P.cl_bot);
    int i = nrad - 1;
    for (; i >= 0; i--) if (press[i] >= top) break;
    ctop = i;
    for (; i >= 0; i--) if (press[i] >= bot) break;
    cbot = i;
    const double x = 2 * TB_PI * P.cl_r * wn;
    refw = pow(P.cl_refwn, P.cl_gamma);
    kBP = P.cl_ext * pow(wn, P.cl_gamma);
    kFH = P.cl_ext / (P.cl_Q * pow(x, -1 * P.cl_gamma) + pow(x, 0.2));
  }
  const double wn4 = P.sc_flag == 1 ? pow(wn, 4.0) : 0.0;
  const bool extra = P.sc_flag != 0 || (P.cl_flag && P.cl_ext != 0);
  // tau.c:231-232, same left-to-right sum
  auto load = [&](int r) -> double {
    double es = 0, ec = 0;
    if (extra) es = tb_extra_ext(P, r, wn, wn4, kBP, kFH, refw, ctop, cbot, press, temp, aux, mol_d, mol_pol, mol_mass, nmol, ec);
    return __ldcs(e + (size_t)r * P.nwn + w) + es + ec + __ldcs(ecs + (size_t)r * P.nwn + w);
  };
  // ray 0 is the top layer itself: tau = 0 (eclipse.c:46-47)
  tau[w] = 0.0;
  int lst = P.nh - 1;
  if (0.0 > P.toomuch) { last[w] = 0; return; }
  if (P.nh == 1) { last[w] = 0; return; }
  // sliding window of the column, top down: e1 = er[rs+1], e2 = er[rs+2] (final values), er[rs] loaded ahead
  double e2 = 0.0, e1 = load(nrad - 1);
  double cur = load(nrad - 2);                     // er[rs] of ray 1
  double nxt = nrad >= 3 ? load(nrad - 3) : 0.0;   // the layer below, still unmodified: ray 1's parabola reads it
  // three more layers in flight: one coalesced 256-byte request per warp and array every iteration, issued
  // four rays before its value is needed
  double n2 = nrad >= 4 ? load(nrad - 4) : 0.0, n3 = nrad >= 5 ? load(nrad - 5) : 0.0, n4 = nrad >= 6 ? load(nrad - 6) : 0.0;
  double S = 0.0;                                  // running sum of the pairs above, (f0 c0 + f1 c1 + f2 c2) c3
  for (int ri = 1; ri < P.nh; ri++) {
    const EclRay &R = sr[ri];
    const int rs = nrad - 1 - ri;
    double t;
    if (R.mode == 2) {
      // two points: parabola through layers rs-1, rs, rs+1 at rad[rs], midpoint construction, value restored
      const double rc[2] = {R.rcp0, R.rcp1};
      const double y0 = tb_parab(R.dx, R.x0, R.xr, rc, nxt, cur, e1);
      const double yb = e1, ya = (y0 + yb) / 2.0;
      t = (y0 * R.c0 + ya * R.c1 + yb * R.c2) * R.c3 / 6.0;
    } else {
      const double rc[2] = {R.rcp0, R.rcp1};
      cur = tb_parab(R.dx, R.x0, R.xr, rc, cur, e1, e2);          // stays modified for the rays below
      if ((R.n & 1) != 0) {                                        // odd point count: one new pair, then the ones above
        S = (cur * R.c0 + e1 * R.c1 + e2 * R.c2) * R.c3 + S;
        t = S / 6.0;
      } else
        t = S / 6.0 + R.h0 * (cur + e1) / 2;
    }
    t = P.rfct * t;
    tau[(size_t)ri * P.nwn + w] = t;
    if (t > P.toomuch) { lst = ri; break; }                        // tau.c:277-287
    e2 = e1; e1 = cur; cur = nxt; nxt = n2; n2 = n3; n3 = n4;
    n4 = rs - 5 >= 0 ? load(rs - 5) : 0.0;
  }
  last[w] = lst;
}


// ---- transit geometry as t = W er ------------------------------------------------------------------
// totaltau1's Simpson sum is a fixed linear map of the extinction column: per ray one weight per
// layer (k_ray_geometry) plus the weight of the parabola value at the ray's closest approach, which stays the
// exact interp_parab expression (its conditioning is what the 1e-9 tolerance on tau comes from).  A warp
// integrates FOUR consecutive rays at once: per layer one shared read of the column (32 wavenumbers, one per
// lane), the four weights as two warp-uniform 128-bit reads, four FMAs -- 0.75 loads per FMA where the pair-wise
// kernel issues 4 loads per 5 floating-point instructions and 2.5 of those per layer.  Layers below a ray's
// start carry the weight zero, so the loop is the same for the four rays.  Early exit, last[] and the zero
// tail as in k_tau_transit.
#define TAU_WL 2      // wavenumbers per lane of k_tau_w (a block owns 32 * TAU_WL columns)
// MINB = 4 (64 registers, 8 bytes spilled) puts a grid of up to 592 blocks -- C2: 490 -- on the machine in ONE wave
// (0.090 -> 0.072 ms); a grid of many waves (C4, batches) runs the 80-register build at 3 blocks per SM.
template <int MINB>
__global__ void __launch_bounds__(32 * TAU_NW, MINB)
k_tau_w(TauPar P, const double *__restrict__ e, const double *__restrict__ ecs,
        const double *__restrict__ press, const double *__restrict__ temp,
        const double *__restrict__ aux, const double *__restrict__ mol_d,
        const double *__restrict__ mol_pol, const double *__restrict__ mol_mass, int nmol,
        const int32_t *__restrict__ meta, const double *__restrict__ par, const double *__restrict__ rcp,
        const double *__restrict__ wg, const double *__restrict__ w0, double *__restrict__ tau,
        int32_t *__restrict__ last) {
  extern __shared__ __align__(16) double sm[];          // er[layer][32 lanes][TAU_WL]
  __shared__ int sh_last[32 * TAU_WL];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nrad = P.nrad, ngrp = (P.nh + TAU_RB - 1) / TAU_RB;
  {  // a batch of atmospheres (tb_run_transit_batch): blockIdx.y selects the atmosphere's slices
    const size_t a = blockIdx.y;
    e += a * nrad * P.nwn; ecs += a * nrad * P.nwn; temp += a * nrad; aux += a * 2 * nrad;
    mol_d += a * (size_t)nmol * nrad; meta += a * P.nh * 4; par += a * P.nh * 4; rcp += a * P.nh * 2;
    wg += a * (size_t)ngrp * nrad * TAU_RB; w0 += a * P.nh; tau += a * P.nh * P.nwn; last += a * P.nwn;
  }
  const int64_t wbase = P.w0 + blockIdx.x * (32LL * TAU_WL) + lane * TAU_WL;    // this lane's first column
  bool inb[TAU_WL];
  bool any = false;
#pragma unroll
  for (int u = 0; u < TAU_WL; u++) { inb[u] = wbase + u < P.w1; any |= inb[u]; }
  if (warp == 0)
    for (int u = 0; u < TAU_WL; u++) sh_last[lane * TAU_WL + u] = P.nh;
  const bool extra = P.sc_flag != 0 || (P.cl_flag && P.cl_ext != 0);
#pragma unroll
  for (int u = 0; u < TAU_WL; u++) {
    const int64_t w = wbase + u;
    if (!inb[u]) {
      for (int r = warp; r < nrad; r += TAU_NW) sm[(r * 32 + lane) * TAU_WL + u] = 0.0;
      continue;
    }
    const double wn = P.wn_fct * __dadd_rn(P.wn_i, __dmul_rn((double)w, P.wn_d));
    double kBP = 0, kFH = 0, refw = 1, wn4 = 0;
    int ctop = -1, cbot = -1;
    if (extra) {
      if (P.cl_flag && P.cl_ext != 0) {
        const double top = pow(10.0, P.cl_top), bot = pow(10.0, P.cl_bot);
        int i = nrad - 1;
        for (; i >= 0; i--) if (press[i] >= top) break;
        ctop = i;
        for (; i >= 0; i--) if (press[i] >= bot) break;
        cbot = i;
        const double x = 2 * TB_PI * P.cl_r * wn;
        refw = pow(P.cl_refwn, P.cl_gamma);
        kBP = P.cl_ext * pow(wn, P.cl_gamma);
        kFH = P.cl_ext / (P.cl_Q * pow(x, -1 * P.cl_gamma) + pow(x, 0.2));
      }
      wn4 = P.sc_flag == 1 ? pow(wn, 4.0) : 0.0;
    }
    for (int r = warp; r < nrad; r += TAU_NW) {
      double es = 0, ec = 0;
      if (extra) es = tb_extra_ext(P, r, wn, wn4, kBP, kFH, refw, ctop, cbot, press, temp, aux, mol_d, mol_pol, mol_mass, nmol, ec);
      // tau.c:231-232, same left-to-right sum
      sm[(r * 32 + lane) * TAU_WL + u] = e[(size_t)r * P.nwn + w] + es + ec + ecs[(size_t)r * P.nwn + w];
    }
  }
  __syncthreads();
  const double2 *er = reinterpret_cast<const double2 *>(sm) + lane;       // er[L * 32] = this lane's two columns of layer L
  int myfirst[TAU_WL];                      // first ray of this warp above toomuch (none yet), per column
#pragma unroll
  for (int u = 0; u < TAU_WL; u++) myfirst[u] = P.nh;
  for (int g = warp; g < ngrp; g += TAU_NW) {
    const int ri0 = g * TAU_RB;
    int rsv[TAU_RB], mdv[TAU_RB], rs_min = nrad;
#pragma unroll
    for (int j = 0; j < TAU_RB; j++) {
      const int ri = ri0 + j;
      rsv[j] = nrad; mdv[j] = 0;
      if (ri < P.nh) { rsv[j] = meta[ri * 4]; mdv[j] = meta[ri * 4 + 2]; }
      if (mdv[j] > 0 && rsv[j] < rs_min) rs_min = rsv[j];
    }
    double ax[TAU_RB], ay[TAU_RB];
#pragma unroll
    for (int j = 0; j < TAU_RB; j++) {
      ax[j] = 0.0; ay[j] = 0.0;
      const int ri = ri0 + j;
      if (mdv[j] > 0 && any) {            // weight of the parabola value at the ray's closest approach
        const int b = rsv[j] - (mdv[j] == 2 ? 1 : 0);        // the two-point ray takes layers rs-1, rs, rs+1
        const double *pp = par + ri * 4;
        const double2 f0 = er[b * 32], f1 = er[(b + 1) * 32], f2 = er[(b + 2) * 32];
        const double wq = w0[ri];
        ax[j] = wq * tb_parab(pp[0], pp[1], pp[2], rcp + ri * 2, f0.x, f1.x, f2.x);
        ay[j] = wq * tb_parab(pp[0], pp[1], pp[2], rcp + ri * 2, f0.y, f1.y, f2.y);
      }
    }
    if (any && (myfirst[0] == P.nh || myfirst[1] == P.nh)) {
      const double2 *wp = reinterpret_cast<const double2 *>(wg + ((size_t)g * nrad) * TAU_RB);
#pragma unroll 4
      for (int L = rs_min + 1; L < nrad; L++) {
        const double2 ev = er[L * 32];
        const double2 wa = __ldg(wp + 2 * L), wb = __ldg(wp + 2 * L + 1);
        ax[0] = fma(wa.x, ev.x, ax[0]); ay[0] = fma(wa.x, ev.y, ay[0]);
        ax[1] = fma(wa.y, ev.x, ax[1]); ay[1] = fma(wa.y, ev.y, ay[1]);
        ax[2] = fma(wb.x, ev.x, ax[2]); ay[2] = fma(wb.x, ev.y, ay[2]);
        ax[3] = fma(wb.y, ev.x, ax[3]); ay[3] = fma(wb.y, ev.y, ay[3]);
      }
#pragma unroll
      for (int j = 0; j < TAU_RB; j++) {
        const int ri = ri0 + j;
        if (ri >= P.nh) continue;
#pragma unroll
        for (int u = 0; u < TAU_WL; u++) {
          if (inb[u] && myfirst[u] == P.nh) {
            const double t = mdv[j] > 0 ? P.rfct * (2 * (u == 0 ? ax[j] : ay[j])) : 0.0;
            tau[(size_t)ri * P.nwn + wbase + u] = t;
            if (t > P.toomuch) myfirst[u] = ri;              // tau.c:277-287: this column is done
          }
        }
      }
    }
    bool done = true;
#pragma unroll
    for (int u = 0; u < TAU_WL; u++) done &= (!inb[u] || myfirst[u] != P.nh);
    if (__all_sync(0xffffffffu, done)) break;
  }
#pragma unroll
  for (int u = 0; u < TAU_WL; u++)
    if (inb[u]) atomicMin(&sh_last[lane * TAU_WL + u], myfirst[u]);
  __syncthreads();
#pragma unroll
  for (int u = 0; u < TAU_WL; u++) {
    if (!inb[u]) continue;
    const int64_t w = wbase + u;
    const int lst = sh_last[lane * TAU_WL + u];          // == nh: never above toomuch
    if (warp == 0) last[w] = lst < P.nh ? lst : P.nh - 1;
    // clear what this warp wrote beyond the block-wide first exceed
    const int upto = myfirst[u] < P.nh ? myfirst[u] : P.nh - 1;
    for (int g = warp; g < ngrp; g += TAU_NW)
      for (int j = 0; j < TAU_RB; j++) {
        const int ri = g * TAU_RB + j;
        if (ri <= upto && ri > lst) tau[(size_t)ri * P.nwn + w] = 0.0;
      }
  }
}

static int tb_tau_run(tb_ctx *c, int geometry, double toomuch, const double *h, int nh, double hfct, bool sync) {
  TB_REQUIRE(c->have_e && c->have_atm, TB_ERR_STATE, "extinction must be computed before tb_tau");
  TB_REQUIRE(geometry == TB_GEOM_TRANSIT || geometry == TB_GEOM_ECLIPSE, TB_ERR_ARG, "bad geometry");
  TB_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int nrad = c->nrad;
  const int64_t nwn = c->nwn;
  if (!h) {                                   // makeipsample with raddelt=-1 / tau.c:92-98
    nh = nrad;
    hfct = c->rad_fct;
    c->h_h.resize(nrad);
    for (int i = 0; i < nrad; i++) c->h_h[i] = c->rad[nrad - 1 - i];
  } else {
    TB_REQUIRE(nh >= 1, TB_ERR_ARG, "need at least one ray");
    c->h_h.assign(h, h + nh);
  }
  if (geometry == TB_GEOM_ECLIPSE) TB_REQUIRE(nh == nrad, TB_ERR_ARG, "eclipse rays are the layers (tau.c:92-98)");
  c->nh = nh; c->hfct = hfct; c->geometry = geometry; c->toomuch = toomuch;
  if (!c->have_ecs) {                         // no CIA tables: zeros
    TB_TRY(c->d_ecs.ensure((size_t)nrad * nwn));
    TB_CUDA(cudaMemsetAsync(c->d_ecs.p, 0, sizeof(double) * nrad * nwn, st));
    c->have_ecs = true;
  }
  const int stride = 4 * ((nrad + 2) / 2 + 1);
  DevBuf<double> &d_h = c->r_h, &d_s = c->r_s;
  TB_TRY(d_h.ensure(nh)); TB_TRY(d_s.ensure((size_t)nh * (nrad + 3)));
  TB_TRY(c->r_meta.ensure((size_t)nh * 4)); TB_TRY(c->r_par.ensure((size_t)nh * 4));
  TB_TRY(c->r_coef.ensure((size_t)nh * stride)); TB_TRY(c->r_rcp.ensure((size_t)nh * 2));
  TB_TRY(c->d_tau.ensure((size_t)nh * nwn)); TB_TRY(c->d_last.ensure(nwn));
  TB_TRY(c->d_es_aux.ensure(2 * (size_t)nrad));
  std::vector<double> aux(2 * (size_t)nrad, 0.0);
  if ((int)c->nH.size() == nrad) for (int i = 0; i < nrad; i++) { aux[i] = c->nH[i]; aux[nrad + i] = c->mean_dens[i]; }
  TB_TRY(tb_upload_small(c, c->d_es_aux.p, aux.data(), aux.size() * sizeof(double), st));
  TB_TRY(tb_upload_small(c, d_h.p, c->h_h.data(), nh * sizeof(double), st));
  if (!c->err_keep) TB_CUDA(cudaMemsetAsync(c->d_err.p, 0, 4 * sizeof(int32_t), st));
  TB_CUDA(cudaMemsetAsync(c->d_tau.p, 0, sizeof(double) * nh * nwn, st));   // calloc, tau.c:34
  const bool wform = geometry == TB_GEOM_TRANSIT && c->tau_stream_ecl;      // t = W er (k_tau_w)
  const int ngrp = (nh + TAU_RB - 1) / TAU_RB;
  if (wform) { TB_TRY(c->r_wg.ensure((size_t)ngrp * nrad * TAU_RB)); TB_TRY(c->r_w0.ensure(nh)); }
  k_ray_geometry<<<nh, 64, 0, st>>>(geometry == TB_GEOM_ECLIPSE, c->d_rad.p, nrad, c->rad_fct, d_h.p, nh,
                                              hfct, stride, c->r_meta.p, c->r_par.p, c->r_coef.p, d_s.p, c->r_rcp.p, c->d_err.p,
                                              wform ? c->r_wg.p : nullptr, wform ? c->r_w0.p : nullptr);
  TauPar P;
  P.eclipse = geometry == TB_GEOM_ECLIPSE; P.nrad = nrad; P.nh = nh; P.stride = stride; P.nwn = nwn;
  tb_band_columns(c, &P.w0, &P.w1);
  const int64_t ncol = P.w1 - P.w0;
  P.rfct = c->rad_fct; P.toomuch = toomuch; P.wn_i = c->wn_i; P.wn_d = c->wn_d; P.wn_fct = c->wn_fct;
  P.sc_flag = c->sc_flag; P.cl_flag = c->cl_flag;
  P.sc_fac = pow(10.0, c->sc_logext) * TB_E0H2;
  P.cl_ext = c->cl_ext; P.cl_top = c->cl_top; P.cl_bot = c->cl_bot; P.cl_gamma = c->cl_gamma;
  P.cl_Q = c->cl_Q; P.cl_r = c->cl_r; P.cl_sig = c->cl_sig; P.cl_refwn = c->cl_refwn;
  tb_timer_begin(c, "tau");
  if (ncol <= 0) {
    // an empty band: nothing to integrate
  } else if (geometry == TB_GEOM_ECLIPSE && c->tau_stream_ecl && nh == nrad && nrad >= 3 &&
             (size_t)nh * sizeof(EclRay) <= 48 * 1024) {
    TB_TRY(c->r_ecl.ensure((size_t)nh * sizeof(EclRay)));
    k_ecl_rays<<<nblk(nh, 128), 128, 0, st>>>(nh, nrad, stride, c->r_meta.p, c->r_par.p, c->r_coef.p, c->r_rcp.p,
                                              reinterpret_cast<EclRay *>(c->r_ecl.p), c->d_err.p);
    k_tau_eclipse<<<nblk(ncol, 128), 128, (size_t)nh * sizeof(EclRay), st>>>(
        P, c->d_e.p, c->d_ecs.p, c->d_p.p, c->d_t.p, c->d_es_aux.p, c->d_dens.p, c->d_mol_pol.p, c->d_mol_mass.p,
        c->nmol, reinterpret_cast<const EclRay *>(c->r_ecl.p), c->d_tau.p, c->d_last.p);
    TB_LAUNCHED(c);
  } else if (wform) {
    const size_t shm = (size_t)nrad * 32 * TAU_WL * sizeof(double);
    TB_REQUIRE(shm <= 200 * 1024, TB_ERR_ARG, "too many layers (%d) for the tau kernel's shared memory", nrad);
    if (c->tauw_shm_set != (int)shm) {
      TB_CUDA(cudaFuncSetAttribute(k_tau_w<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm));
      TB_CUDA(cudaFuncSetAttribute(k_tau_w<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm));
      c->tauw_shm_set = (int)shm;
    }
    auto tauw = nblk(ncol, 32 * TAU_WL) <= 148 * 4 ? k_tau_w<4> : k_tau_w<3>;
    tauw<<<nblk(ncol, 32 * TAU_WL), 32 * TAU_NW, shm, st>>>(P, c->d_e.p, c->d_ecs.p, c->d_p.p, c->d_t.p, c->d_es_aux.p, c->d_dens.p,
                                                    c->d_mol_pol.p, c->d_mol_mass.p, c->nmol, c->r_meta.p, c->r_par.p,
                                                    c->r_rcp.p, c->r_wg.p, c->r_w0.p, c->d_tau.p, c->d_last.p);
  } else {
    const size_t shm = ((size_t)nrad * 32 + (size_t)TAU_NW * stride) * sizeof(double);
    TB_REQUIRE(stride <= 320, TB_ERR_ARG, "too many layers (%d) for the tau kernel's coefficient registers", nrad);
    TB_REQUIRE(shm <= 200 * 1024, TB_ERR_ARG, "too many layers (%d) for the tau kernel's shared memory", nrad);
    if (c->tau_shm_set != (int)shm) {
      TB_CUDA(cudaFuncSetAttribute(k_tau_transit, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm));
      c->tau_shm_set = (int)shm;
    }
    k_tau_transit<<<nblk(ncol, 32), 32 * TAU_NW, shm, st>>>(P, c->d_e.p, c->d_ecs.p, c->d_p.p, c->d_t.p, c->d_es_aux.p,
                                                          c->d_dens.p, c->d_mol_pol.p, c->d_mol_mass.p, c->nmol,
                                                          c->r_meta.p, c->r_par.p, c->r_coef.p, c->r_rcp.p, c->d_tau.p,
                                                          c->d_last.p);
  }
  tb_timer_end(c, "tau");
  c->launches += 2;
  TB_CUDA(cudaGetLastError());
  c->have_tau = true;
  if (sync) TB_TRY(tb_check_errors(c));
  return TB_OK;
}

// device-side error flags (set by k_ray_geometry / k_modulation), read back with one sync
int tb_check_errors(tb_ctx *c) {
  if (c->defer_check) return TB_OK;          // inside a batch: flags accumulate, read back once
  int32_t err[4];
  TB_CUDA(cudaMemcpyAsync(err, c->d_err.p, sizeof(err), cudaMemcpyDeviceToHost, c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  TB_REQUIRE(err[0] == 0, TB_ERR_STATE, "eclipse rays are not the layers taken top down (tau.c:92-98): internal error");
  TB_REQUIRE(err[1] == 0, TB_ERR_DOMAIN, "a ray's closest approach lies below the bottom layer (slantpath.c:40-44)");
  TB_REQUIRE(err[2] == 0, TB_ERR_DOMAIN, "less than 3 items for radial integration (slantpath.c:393-397)");
  TB_REQUIRE(err[3] == 0, TB_ERR_DOMAIN, "Optical depth didn't reach the limiting value at some wavenumber.  Cannot use "
             "critical radius technique (-1) (slantpath.c:308-316)");
  return TB_OK;
}

// 1 (default): eclipse tau by the O(layers) streaming kernel; 0: the general O(layers^2) kernel (comparison tap)
extern "C" int tb_set_tau_streaming(tb_ctx *c, int on) {
  c->tau_stream_ecl = on != 0;
  return TB_OK;
}

// The impact-parameter sample tb_spectrum integrates along (makeipsample, makesample.c:558-604).  With raddelt = -1
// it is the radius sample reversed (h = NULL, the default); a resampled radius grid comes with its own sample,
// built from the top radius downwards, which is NOT bit for bit the reversed radii.
extern "C" int tb_set_rays(tb_ctx *c, const double *h, int nh, double hfct) {
  if (!h || nh <= 0) { c->user_h.clear(); return TB_OK; }
  c->user_h.assign(h, h + nh);
  c->user_hfct = hfct;
  return TB_OK;
}

extern "C" int tb_tau(tb_ctx *c, int geometry, double toomuch, const double *h, int nh, double hfct) {
  return tb_tau_run(c, geometry, toomuch, h, nh, hfct, true);
}

// =============================================================== modulation (a14)

#define MOD_QL 4   // lanes per wavenumber in k_modulation
__global__ void k_modulation(int64_t nwn, int64_t w0, int64_t w1, int nh, const double *__restrict__ tau,
                             const int32_t *__restrict__ last, const double *__restrict__ ip, double ipfct,
                             double toomuch, double srad, int transp, double *__restrict__ out,
                             int32_t *__restrict__ err) {
  // The impact parameters and the non-uniform Simpson weights (geth/simpson, numerical.c:390-525)
  // are the same for every wavenumber: three FP64 divisions per pair of intervals.  A pair is
  // identified by its innermost ray r (points r, r-1, r-2 in ascending radius) whatever the
  // cut-off ray of a wavenumber is, so each block tabulates xs[r] = ip[r]*ipfct and
  // cf[r] = {2-hratio, hfactor, 2-1/hratio, hsum} once in shared memory.
  extern __shared__ __align__(16) double smod[];
  double *xs = smod, *cf = smod + ((nh + 1) & ~1);   // 16-byte aligned rows
  {  // a batch of atmospheres (tb_batch.cu): blockIdx.y selects the atmosphere's slices
    const size_t a = blockIdx.y;
    tau += a * nh * nwn; last += a * nwn; ip += a * nh; out += a * nwn;
  }
  for (int i = threadIdx.x; i < nh; i += blockDim.x) xs[i] = ip[i] * ipfct;
  __syncthreads();
  for (int r = 2 + threadIdx.x; r < nh; r += blockDim.x) {
    const double h0 = xs[r - 1] - xs[r], h1 = xs[r - 2] - xs[r - 1];
    const double hsum = h0 + h1, hratio = h1 / h0, hfactor = hsum * hsum / (h0 * h1);
    cf[4 * r + 0] = 2.0 - hratio;
    cf[4 * r + 1] = hfactor;
    cf[4 * r + 2] = 2.0 - 1.0 / hratio;
    cf[4 * r + 3] = hsum;
  }
  __syncthreads();
  // MOD_QL lanes share a wavenumber: each integrates a contiguous quarter of the pairs (a dependent
  // exp-and-add chain of ~50 pairs per wavenumber left the machine at 10 % occupancy with one lane
  // each); the partial sums are added in a fixed butterfly, lane 0 of the quad finishes.
  const int64_t wq = w0 + (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / MOD_QL;   // columns [w0, w1)
  const int ql = threadIdx.x & (MOD_QL - 1);
  const bool inb = wq < w1;
  const int64_t w = inb ? wq : w1 - 1;
  const int L0 = last[w];
  int L1 = L0 + 1;
  if (L1 > nh - 1) L1 = nh - 1;
  const int n = L1 + 1;                                 // points handed to simps (slantpath.c:392-408)
  const bool bad = n < 3;
  // point k (ascending radius) is ray i = n-1-k; integrand exp(-tau_i)*ip_i for i<=L0, else 0
  auto Y = [&](int i) { return (i <= L0) ? exp(-tau[(size_t)i * nwn + w]) * xs[i] : 0.0; };
  const int even = (n % 2 == 0);
  const int np = bad ? 0 : (n - 1) / 2;
  const int i0 = (np * ql) / MOD_QL, i1 = (np * (ql + 1)) / MOD_QL;
  double res = 0.0;
  if (i1 > i0) {
    int r = n - 1 - even - 2 * i0;                      // innermost ray of this lane's first pair
    double y0 = Y(r);
    for (int i = i0; i < i1; i++, r -= 2) {
      const double y1 = Y(r - 1), y2 = Y(r - 2);
      const double2 c01 = *reinterpret_cast<const double2 *>(cf + 4 * r);
      const double2 c23 = *reinterpret_cast<const double2 *>(cf + 4 * r + 2);
      res += (y0 * c01.x + y1 * c01.y + y2 * c23.x) * c23.y;
      y0 = y2;
    }
  }
#pragma unroll
  for (int d = 1; d < MOD_QL; d <<= 1) res += __shfl_xor_sync(0xffffffffu, res, d);
  if (!inb || ql != 0) return;
  if (bad) { atomicExch(err + 2, 1); out[w] = 0; return; }
  double integ = res / 6.0;
  if (even) integ += (xs[n - 2] - xs[n - 1]) * (Y(n - 1) + Y(n - 2)) / 2;
  const double rtop = xs[0];
  double rr = rtop * rtop - 2.0 * integ;
  if (transp) {
    const double tl = tau[(size_t)L0 * nwn + w];
    const double maxtau = tl > toomuch ? tl : toomuch;
    const double rb = xs[n - 1];
    rr -= exp(-maxtau) * rb * rb;
  }
  out[w] = rr * (1.0 / (srad * srad));
}

// modulationm1 (slantpath.c:446-473, `modlevel -1`): the planet as an opaque disc whose radius is where
// tau reaches toomuch, by a line through the two rays around the cut-off (interp_line, pu numerical.c:200-210).
// A wavenumber whose tau never reached toomuch is the reference's error -1 (slantpath.c:308-316).
__global__ void k_modulation_m1(int64_t nwn, int64_t w0, int64_t w1, int nh, const double *__restrict__ tau,
                                const int32_t *__restrict__ last, const double *__restrict__ ip, double ipfct,
                                double toomuch, double srad, double *__restrict__ out, int32_t *__restrict__ err) {
  const int64_t w = w0 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (w >= w1) return;
  const int L = last[w];
  if (tau[(size_t)L * nwn + w] < toomuch) { atomicExch(err + 3, 1); out[w] = -1; return; }
  const int ini = L - 1 < 0 ? 0 : L - 1;
  // ipv[1] is left unset by the reference when the very first ray is already opaque (ini = last = 0)
  const double y0 = ip[ini] * ipfct, y1 = (ini + 1 <= L) ? ip[ini + 1] * ipfct : 0.0;
  const double x0 = tau[(size_t)ini * nwn + w], x1 = ini + 1 < nh ? tau[(size_t)(ini + 1) * nwn + w] : 0.0;
  const double dx = x1 - x0, m = (y1 - y0) / dx;
  const double muchrad = y0 + (toomuch - x0) * m;
  out[w] = muchrad * muchrad / (srad * srad);
}

extern "C" int tb_set_modlevel(tb_ctx *c, int level) {
  TB_REQUIRE(level == 1 || level == -1, TB_ERR_ARG,
             "slantpath:: modulationperwn:: Level %i of detail has not been implemented to compute modulation.", level);
  c->modlevel = level;
  return TB_OK;
}

extern "C" int tb_modulation(tb_ctx *c, double starrad_cm, int transparent, double *out) {
  TB_REQUIRE(c->have_tau && c->geometry == TB_GEOM_TRANSIT, TB_ERR_STATE, "tb_tau (transit) must precede tb_modulation");
  TB_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  DevBuf<double> &d_ip = c->r_ip;
  TB_TRY(d_ip.ensure(c->nh));
  TB_TRY(c->d_out.ensure(c->nwn));
  TB_TRY(tb_upload_small(c, d_ip.p, c->h_h.data(), c->nh * sizeof(double), st));
  tb_timer_begin(c, "modulation");
  const size_t shm = (5 * (size_t)c->nh + 2) * sizeof(double);   // xs[nh] (padded to even) + cf[4*nh]
  TB_REQUIRE(shm <= 48 * 1024, TB_ERR_ARG, "too many rays (%d) for the modulation kernel's shared memory", c->nh);
  int64_t w0, w1;
  tb_band_columns(c, &w0, &w1);
  if (w1 <= w0) {
    // an empty band
  } else if (c->modlevel == -1)
    k_modulation_m1<<<nblk(w1 - w0, 128), 128, 0, st>>>(c->nwn, w0, w1, c->nh, c->d_tau.p, c->d_last.p, d_ip.p, c->hfct,
                                                        c->toomuch, starrad_cm, c->d_out.p, c->d_err.p);
  else
    k_modulation<<<nblk((w1 - w0) * MOD_QL, 128), 128, shm, st>>>(c->nwn, w0, w1, c->nh, c->d_tau.p, c->d_last.p, d_ip.p, c->hfct,
                                                  c->toomuch, starrad_cm, transparent, c->d_out.p, c->d_err.p);
  tb_timer_end(c, "modulation");
  TB_LAUNCHED(c);
  TB_TRY(tb_comm_gather_spectrum(c));         // band shard: every rank gets the whole spectrum (NCCL all-gather)
  if (out) TB_CUDA(cudaMemcpyAsync(out, c->d_out.p, c->nwn * sizeof(double), cudaMemcpyDeviceToHost, st));
  return tb_check_errors(c);
}

// =============================================================== intensity + flux (a15)

#define TB_MAXANG 16
// QL lanes share a wavenumber, each integrating a contiguous part of the layers (a grid of a few
// thousand wavenumbers leaves one lane per wavenumber a chain of ~600 exp() on a nearly empty
// machine); the trapezoid partial sums are added in a fixed butterfly, the lane holding the last
// layer finishes.  QL = 1 is the plain loop (large grids: enough wavenumbers to fill the machine).
template <int QL>
__global__ void k_intensity_flux(int64_t nwn, int64_t w0, int64_t w1, int nrad, const double *__restrict__ tau,
                                 const int32_t *__restrict__ last, const double *__restrict__ temp,
                                 double wn_i, double wn_d, double wn_fct, int nang,
                                 const double *__restrict__ ang, const double *__restrict__ area,
                                 double *__restrict__ intens, double *__restrict__ out) {
  extern __shared__ __align__(16) double sif[];
  double *skt = sif, *srkt = sif + nrad;                // k_B*T of point i (layer nrad-1-i) and its reciprocal
  for (int i = threadIdx.x; i < nrad; i += blockDim.x) {
    const double kt = TB_KB * temp[nrad - 1 - i];
    skt[i] = kt;
    srkt[i] = 1.0 / kt;
  }
  __syncthreads();
  const int64_t wq = w0 + (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / QL;        // columns [w0, w1)
  const int ql = threadIdx.x & (QL - 1);
  const bool inb = wq < w1;
  const int64_t w = inb ? wq : w1 - 1;
  const int L = last[w];
  const double wv = __dadd_rn(wn_i, __dmul_rn((double)w, wn_d));   // wns.v[w]
  const double ww = wv * wn_fct;
  const double num = 2.0 * TB_H * pow(ww, 3.0) * TB_LS * TB_LS;     // eclipse.c:154
  // the divisors of the two per-(layer, wavenumber) divisions -- cos(angle) and k_B*T(layer) -- do not
  // depend on the wavenumber: their reciprocals are formed once and every quotient is completed to the
  // correctly rounded one by tb_div_r (the value of the division for ~6 instructions instead of ~40)
  double icos[TB_MAXANG], rcos[TB_MAXANG], dprev[TB_MAXANG], tz[TB_MAXANG];
  for (int a = 0; a < nang; a++) { icos[a] = cos(ang[a] * TB_DEGREES); rcos[a] = 1.0 / icos[a]; tz[a] = 0; dprev[a] = 0; }
  double Bprev = 0;
  const double hcw = TB_H * ww * TB_LS;
  const int n = L + 1;
  const int i0 = (n * ql) / QL, i1 = (n * (ql + 1)) / QL;
  // a lane that does not start at the top first evaluates the point before its part
  for (int i = (i0 > 0 ? i0 - 1 : 0); i < i1; i++) {
    const double t = tau[(size_t)i * nwn + w];
    const double B = num / (exp(tb_div_r(hcw, skt[i], srkt[i])) - 1.0);
    for (int a = 0; a < nang; a++) {
      const double dt = exp(tb_div_r(-t, icos[a], rcos[a]));
      if (i >= i0 && i > 0) tz[a] += (dt - dprev[a]) * (B + Bprev);  // integ_trapz, numerical.c:168-170
      dprev[a] = dt;
    }
    Bprev = B;
  }
  if (QL > 1) {
    for (int a = 0; a < nang; a++) {
#pragma unroll
      for (int d = 1; d < QL; d <<= 1) tz[a] += __shfl_xor_sync(0xffffffffu, tz[a], d);
    }
  }
  if (!inb || ql != QL - 1) return;       // the last lane holds the last layer (i1 = n, n >= 1)
  double f = 0;
  for (int a = 0; a < nang; a++) {
    const double I = Bprev * dprev[a] - 0.5 * tz[a];                  // eclipse.c:158
    intens[(size_t)a * nwn + w] = I;
    f += TB_PI * I * area[a];                                         // eclipse.c:275-279
  }
  out[w] = f;
}

extern "C" int tb_intensity_flux(tb_ctx *c, const double *angles_deg, int nang, double *out) {
  TB_REQUIRE(c->have_tau && c->geometry == TB_GEOM_ECLIPSE, TB_ERR_STATE, "tb_tau (eclipse) must precede tb_intensity_flux");
  TB_REQUIRE(nang >= 1 && nang <= TB_MAXANG, TB_ERR_ARG, "1..%d angles supported", TB_MAXANG);
  TB_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  // area grid (eclipse.c:262-276)
  std::vector<double> grid(nang + 1), area(nang), ang(angles_deg, angles_deg + nang);
  grid[0] = 0.0 * TB_DEGREES;
  grid[nang] = 90.0 * TB_DEGREES;
  for (int i = 1; i < nang; i++) grid[i] = (ang[i - 1] + ang[i]) * TB_DEGREES / 2.0;
  for (int i = 0; i < nang; i++) area[i] = pow(sin(grid[i + 1]), 2.0) - pow(sin(grid[i]), 2.0);
  DevBuf<double> &d_a = c->r_ang;
  TB_TRY(d_a.ensure(2 * (size_t)nang));
  TB_TRY(c->d_intens.ensure((size_t)nang * c->nwn));
  TB_TRY(c->d_out.ensure(c->nwn));
  TB_TRY(tb_upload_small(c, d_a.p, ang.data(), nang * sizeof(double), st));
  TB_TRY(tb_upload_small(c, d_a.p + nang, area.data(), nang * sizeof(double), st));
  int64_t w0, w1;
  tb_band_columns(c, &w0, &w1);
  tb_timer_begin(c, "intensity");
  if (w1 <= w0) {
    // an empty band
  } else if (c->nwn < 100000)
    k_intensity_flux<8><<<nblk((w1 - w0) * 8, 128), 128, 2 * (size_t)c->nrad * sizeof(double), st>>>(c->nwn, w0, w1, c->nrad, c->d_tau.p, c->d_last.p, c->d_t.p, c->wn_i,
                                                            c->wn_d, c->wn_fct, nang, d_a.p, d_a.p + nang, c->d_intens.p,
                                                            c->d_out.p);
  else
    k_intensity_flux<1><<<nblk(w1 - w0, 128), 128, 2 * (size_t)c->nrad * sizeof(double), st>>>(c->nwn, w0, w1, c->nrad, c->d_tau.p, c->d_last.p, c->d_t.p, c->wn_i,
                                                        c->wn_d, c->wn_fct, nang, d_a.p, d_a.p + nang, c->d_intens.p,
                                                        c->d_out.p);
  tb_timer_end(c, "intensity");
  TB_LAUNCHED(c);
  TB_TRY(tb_comm_gather_spectrum(c));         // band shard: every rank gets the whole flux spectrum
  if (out) TB_CUDA(cudaMemcpyAsync(out, c->d_out.p, c->nwn * sizeof(double), cudaMemcpyDeviceToHost, st));
  c->nang = nang;
  return tb_check_errors(c);
}

// =============================================================== do_transit in one call

extern "C" int tb_spectrum(tb_ctx *c, int geometry, double toomuch, double ethresh, double starrad_cm,
                           int transparent, const double *angles_deg, int nang, double *out) {
  // interpcs (transit.c:147) does not depend on the molecular extinction: its small latency-bound
  // kernels run on a second stream beside the line sum and join before tau
  cudaStream_t main_st = c->stream;
  TB_CUDA(cudaSetDevice(c->device));
  TB_CUDA(cudaEventRecord(c->ev_fork, main_st));
  TB_CUDA(cudaStreamWaitEvent(c->aux_stream, c->ev_fork, 0));
  // the extinction is enqueued FIRST: its kernels start while the host is still enqueueing the CIA chain on the
  // second stream (a band shard's step is short enough for those ~40 us of host time to show)
  // (from a resident grid the CIA chain is the longer of the two -- C4: 2.0 against 1.2 ms -- and goes first)
  int rc_e = TB_OK, rc_cia = TB_OK;
  for (int pass = 0; pass < 2; pass++) {
    if ((pass == 0) == !c->have_grid_o) {
      if (c->have_grid_o) rc_e = tb_interp_run(c);         // interpolmolext (opacity file given)
      else rc_e = tb_extinction_lbl(c, ethresh);           // computemolext per layer
    } else {
      c->stream = c->aux_stream;
      rc_cia = tb_cia_run(c);
      c->stream = main_st;
    }
  }
  TB_CUDA(cudaEventRecord(c->ev_join, c->aux_stream));
  TB_CUDA(cudaStreamWaitEvent(main_st, c->ev_join, 0));    // also on error: never leave the fork dangling
  TB_TRY(rc_e);
  TB_TRY(rc_cia);
  if (c->user_h.empty()) TB_TRY(tb_tau_run(c, geometry, toomuch, nullptr, 0, 0, false));   // tau, transit.c:165,189
  else {
    const std::vector<double> h(c->user_h);
    TB_TRY(tb_tau_run(c, geometry, toomuch, h.data(), (int)h.size(), c->user_hfct, false));
  }
  if (geometry == TB_GEOM_ECLIPSE) return tb_intensity_flux(c, angles_deg, nang, out);
  return tb_modulation(c, starrad_cm, transparent, out);
}

// run_transit for a batch of atmospheres (BART hands a retrieval chain's samples one by one through
// transit.c:118-122; SURVEY 8b asks for `*_batched` entry points): atmosphere a is the a-th slice
// of each array, laid out as tb_set_atmosphere takes one.  The spectra are computed back to back on
// the context's stream -- no host synchronisation, error read-back or device->host copy between
// atmospheres -- and delivered to out[natm][nwn] with one copy.  Same kernels, same results as natm
// calls of tb_set_atmosphere + tb_spectrum.
extern "C" int tb_spectrum_batch(tb_ctx *c, int natm, int nrad, const double *rad, double rad_fct,
                                 const double *p, double p_fct, const double *t, double t_fct,
                                 const double *density, const double *Z, int geometry, double toomuch,
                                 double ethresh, double starrad_cm, int transparent,
                                 const double *angles_deg, int nang, double *out) {
  TB_REQUIRE(natm >= 1 && nrad >= 1 && rad && p && t && density && out, TB_ERR_ARG, "tb_spectrum_batch: bad arguments");
  TB_REQUIRE(c->have_species && c->have_grid, TB_ERR_STATE, "grid and species must be set before tb_spectrum_batch");
  TB_CUDA(cudaSetDevice(c->device));
  const int64_t nwn = c->nwn;
  TB_TRY(c->d_out_batch.ensure((size_t)natm * nwn));
  int rc = TB_OK;
  c->defer_check = true;
  for (int a = 0; a < natm && rc == TB_OK; a++) {
    c->err_keep = a > 0;
    rc = tb_set_atmosphere(c, nrad, rad + (size_t)a * nrad, rad_fct, p + (size_t)a * nrad, p_fct,
                           t + (size_t)a * nrad, t_fct, density + (size_t)a * c->nmol * nrad,
                           Z ? Z + (size_t)a * c->niso * nrad : nullptr);
    if (rc == TB_OK)
      rc = tb_spectrum(c, geometry, toomuch, ethresh, starrad_cm, transparent, angles_deg, nang, nullptr);
    if (rc == TB_OK && cudaMemcpyAsync(c->d_out_batch.p + (size_t)a * nwn, c->d_out.p, nwn * sizeof(double),
                                       cudaMemcpyDeviceToDevice, c->stream) != cudaSuccess)
      rc = TB_ERR_CUDA;
  }
  c->defer_check = false;
  c->err_keep = false;
  if (rc != TB_OK) { cudaStreamSynchronize(c->stream); return rc; }
  TB_CUDA(cudaMemcpyAsync(out, c->d_out_batch.p, (size_t)natm * nwn * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  return tb_check_errors(c);
}

// =============================================================== batched retrieval (a16 on the device, C5)
//
// BART hands run_transit one atmosphere at a time: in = [T(nlayers) | q_mol0(nlayers) | ...] (readatm.c:735-744).
// From the opacity grid a spectrum is a chain of small latency-bound kernels (~0.37 ms of which ~0.1 ms is
// arithmetic), so a batch of atmospheres goes through EVERY stage in one launch each:
//   k_atm_setup    reloadatm + checkaddmm + stateeqnford + radpress (readatm.c:122-159,722-865, transit.h:58-69)
//                  per atmosphere on the device: densities, mean molecular mass, hydrostatic radii, impact
//                  parameters, the grid's temperature brackets, the CIA density factors, cloud densities;
//   k_grid_interp  over natm*nlayers virtual layers (every grid plane a batch touches is read while it is in L2);
//   k_cia_*        the two spline families over natm*nlayers layers;
//   k_ray_geometry, k_tau_transit, k_modulation with the atmosphere as a grid dimension.
// One upload of the inputs, one read-back of the spectra and one error check per batch.
struct BatchPar {
  int nrad, nmol, bymass, ih2, Nt, Nm, ntab;
  double p_fct, t_fct, rad_fct, p0, r0, gsurf;
  double cia_tmin, cia_tmax;
  int tab_mol[CIA_MAXTAB][2], tab_nspec[CIA_MAXTAB];
};

#define ATM_MAXLAYERS 1024
__global__ void k_atm_setup(BatchPar B, int natm, const double *__restrict__ in, const double *__restrict__ p,
                            const double *__restrict__ mol_mass, const double *__restrict__ gtemp,
                            const int *__restrict__ gmol_idx, double *__restrict__ ltemp /*[natm][nrad] kelvin*/,
                            double *__restrict__ traw /*[natm][nrad] file units*/, double *__restrict__ dens /*[natm][nmol][nrad]*/,
                            double *__restrict__ rad /*[natm][nrad]*/, double *__restrict__ ip /*[natm][nrad]*/,
                            InterpLayer *__restrict__ L /*[natm*nrad]*/, double *__restrict__ rho_g /*[Nm][natm*nrad]*/,
                            double *__restrict__ densfac /*[ntab][natm*nrad]*/, double *__restrict__ aux /*[natm][2][nrad]*/,
                            int32_t *__restrict__ err) {
  __shared__ double sh_t[ATM_MAXLAYERS], sh_mm[ATM_MAXLAYERS], sh_rad[ATM_MAXLAYERS], sh_lp[ATM_MAXLAYERS];
  const int a = blockIdx.x, n = B.nrad, nmol = B.nmol, nvl = natm * n;
  const double *x = in + (size_t)a * n * (1 + nmol);
  for (int r = threadIdx.x; r + 1 < n; r += blockDim.x) sh_lp[r] = log(p[r] / p[r + 1]);   // radpress: log(pr[i]/pr[i+1])
  for (int r = threadIdx.x; r < n; r += blockDim.x) {
    const double T = x[r];
    double mm = 0, sumq = 0;
    for (int i = 0; i < nmol; i++) {                       // checkaddmm (readatm.c:122-159)
      const double q = x[(size_t)n * (i + 1) + r];
      mm += B.bymass ? q / mol_mass[i] : q * mol_mass[i];
      sumq += q;
    }
    if (B.bymass) mm = 1.0 / mm;
    if (sumq > 1.001) atomicExch(err + 4, 1);
    sh_t[r] = T; sh_mm[r] = mm;
    const double pr = p[r] * B.p_fct, tk = T * B.t_fct;
    double md = 0, nH = 0, mean_mm = 0;
    for (int i = 0; i < nmol; i++) {                       // stateeqnford (transit.h:58-69)
      const double q = x[(size_t)n * (i + 1) + r];
      const double rho = TB_AMU * q * pr / TB_KB / tk;
      const double d = B.bymass ? rho * mm : rho * mol_mass[i];
      dens[((size_t)a * nmol + i) * n + r] = d;
      md += d / mol_mass[i] * q;                           // tau.c:194-214
      if (i == B.ih2) nH = d / mol_mass[i] * q * TB_NAVOGADRO;
      mean_mm += mol_mass[i] * q;
    }
    aux[(size_t)a * 2 * n + r] = nH;
    aux[(size_t)a * 2 * n + n + r] = md * mean_mm;
    ltemp[(size_t)a * n + r] = tk;
    traw[(size_t)a * n + r] = T;
    if (tk < B.cia_tmin || tk > B.cia_tmax) atomicExch(err + 5, 1);   // interpcs exits (crosssec.c:293-309)
    // interpolmolext's temperature bracket (extinction.c:562-564)
    int lo = 0, hi = B.Nt;
    while (hi - lo > 1) {
      const int mid = (hi + lo) / 2;
      if (gtemp[mid] > tk) hi = mid; else lo = mid;
    }
    int it = (fabs(gtemp[hi] - tk) < fabs(gtemp[lo] - tk)) ? hi : lo;
    if (tk < gtemp[it]) it--;
    if (it < 0 || it + 1 >= B.Nt) { atomicExch(err + 6, 1); it = it < 0 ? 0 : B.Nt - 2; }
    InterpLayer l;
    l.it = it; l.t = tk; l.tlo = gtemp[it]; l.thi = gtemp[it + 1]; l.rdt = 1.0 / (l.thi - l.tlo);
    L[(size_t)a * n + r] = l;
    for (int m = 0; m < B.Nm; m++) rho_g[(size_t)m * nvl + (size_t)a * n + r] = dens[((size_t)a * nmol + gmol_idx[m]) * n + r];
    for (int k = 0; k < B.ntab; k++) {                      // crosssec.c:321-326
      double f = 1.0;
      for (int q2 = 0; q2 < B.tab_nspec[k]; q2++) {
        const int m = B.tab_mol[k][q2];
        f *= dens[((size_t)a * nmol + m) * n + r] / (TB_AMU * mol_mass[m] * TB_AMAGAT);
      }
      densfac[(size_t)k * nvl + (size_t)a * n + r] = f;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {                                  // radpress (readatm.c:787-865): a recurrence over the layers
    const double g0 = B.gsurf, p0 = B.p0, r0 = B.r0, rfct = B.rad_fct;
    const double *temp = sh_t, *mu = sh_mm, *pr = p;
    double *radius = sh_rad;
    double minimum = 1e37, temp0, mu0, g;
    int i0 = -1;
    for (int i = 0; i < n; i++)
      if (fabs(pr[i] - p0) < minimum) { i0 = i; minimum = fabs(pr[i] - p0); }
    const double kb_amu = TB_KB / TB_AMU;
    if ((pr[i0] > p0 && i0 + 1 >= n) || (pr[i0] <= p0 && i0 - 1 < 0)) { atomicExch(err + 7, 1); i0 = i0 + 1 >= n ? n - 2 : 1; }
    if (pr[i0] > p0) {
      temp0 = temp[i0] + ((temp[i0 + 1] - temp[i0]) / log(pr[i0 + 1] / pr[i0])) * log(p0 / pr[i0]);
      mu0 = mu[i0] + ((mu[i0 + 1] - mu[i0]) / log(pr[i0 + 1] / pr[i0])) * log(p0 / pr[i0]);
      radius[i0] = r0 + 0.5 * (temp[i0] / mu[i0] + temp0 / mu0) * (kb_amu * log(p0 / pr[i0]) / g0) / rfct;
    } else {
      temp0 = temp[i0] + ((temp[i0 - 1] - temp[i0]) / log(pr[i0 - 1] / pr[i0])) * log(p0 / pr[i0]);
      mu0 = mu[i0] + ((mu[i0 - 1] - mu[i0]) / log(pr[i0 - 1] / pr[i0])) * log(p0 / pr[i0]);
      radius[i0] = r0 - 0.5 * (temp[i0] / mu[i0] + temp0 / mu0) * (kb_amu * log(pr[i0] / p0) / g0) / rfct;
    }
    // (pow(x, 2) of a correctly rounded libm is the rounded product x*x; the logarithms of the layers' pressure
    // ratios do not depend on the atmosphere and were formed by the whole block: the chain below is ~10 operations
    // per layer instead of a log and a pow)
    { const double q = r0 / radius[i0]; g = g0 * (q * q); }
    for (int i = i0 - 1; i >= 0; i--) {
      radius[i] = radius[i + 1] - 0.5 * (temp[i] / mu[i] + temp[i + 1] / mu[i + 1]) * (kb_amu * sh_lp[i] / g) / rfct;
      const double q = radius[i + 1] / radius[i];
      g = g * (q * q);
    }
    { const double q = r0 / radius[i0]; g = g0 * (q * q); }
    for (int i = i0 + 1; i < n; i++) {
      radius[i] = radius[i - 1] + 0.5 * (temp[i] / mu[i] + temp[i - 1] / mu[i - 1]) * (kb_amu * sh_lp[i - 1] / g) / rfct;
      const double q = radius[i - 1] / radius[i];
      g = g * (q * q);
    }
  }
  __syncthreads();
  for (int r = threadIdx.x; r < n; r += blockDim.x) {
    rad[(size_t)a * n + r] = sh_rad[r];
    ip[(size_t)a * n + r] = sh_rad[n - 1 - r];              // makeipsample with raddelt = -1 (makesample.c:564-573)
    if (r > 0 && !(sh_rad[r] > sh_rad[r - 1])) atomicExch(err + 7, 1);
  }
}

extern "C" int tb_set_atmosphere_model(tb_ctx *c, int nrad, const double *p, double p_fct, double t_fct, double rad_fct,
                                       int bymass, double p0, double r0, double gsurf) {
  TB_REQUIRE(nrad >= 3 && nrad <= ATM_MAXLAYERS && p, TB_ERR_ARG, "tb_set_atmosphere_model: 3..%d layers", ATM_MAXLAYERS);
  TB_REQUIRE(p0 != 0 && r0 != 0 && gsurf != 0, TB_ERR_ARG,
             "Surface gravity (%g) or reference pressure (%g) or radius (%g) were not defined.", gsurf, p0, r0);
  TB_CUDA(cudaSetDevice(c->device));
  c->am_nrad = nrad; c->am_p.assign(p, p + nrad);
  c->am_pfct = p_fct; c->am_tfct = t_fct; c->am_rfct = rad_fct; c->am_bymass = bymass;
  c->am_p0 = p0; c->am_r0 = r0; c->am_gsurf = gsurf;
  c->am_ih2 = -1;                       // tau.c:207 looks H2 up by NAME; molecules.dat gives it the ID 105
  for (int i = 0; i < c->nmol; i++) if (c->mol_ID[i] == 105) c->am_ih2 = i;
  TB_TRY(c->b_p.ensure(nrad));
  TB_CUDA(cudaMemcpyAsync(c->b_p.p, p, nrad * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  return TB_OK;
}

extern "C" int tb_run_transit_batch(tb_ctx *c, int natm, const double *in, double toomuch, double starrad_cm,
                                    int transparent, double *out) {
  TB_REQUIRE(natm >= 1 && in && out, TB_ERR_ARG, "tb_run_transit_batch: bad arguments");
  TB_REQUIRE(c->am_nrad > 0, TB_ERR_STATE, "tb_set_atmosphere_model must precede tb_run_transit_batch");
  TB_REQUIRE(c->have_grid_o, TB_ERR_STATE, "tb_run_transit_batch serves spectra from a resident opacity grid (tb_opacity_load)");
  TB_REQUIRE(c->modlevel == 1, TB_ERR_STATE, "tb_run_transit_batch: modlevel 1 only");
  TB_REQUIRE(!(c->banded && c->shard_world > 1), TB_ERR_STATE, "a retrieval batch shards over atmospheres, not bands");
  const int nrad = c->am_nrad, nmol = c->nmol, nh = nrad, ntab = (int)c->cia.size();
  TB_REQUIRE(c->gNlayer == nrad, TB_ERR_ARG, "opacity grid has %lld layers, the atmosphere model %d", (long long)c->gNlayer, nrad);
  TB_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int64_t nwn = c->nwn;
  const int Bmax = natm < 32 ? natm : 32;
  const int Nt = (int)c->gNtemp, Nm = (int)c->gNmol;
  const int stride = 4 * ((nrad + 2) / 2 + 1);
  const size_t shm_tau = ((size_t)nrad * 32 + (size_t)TAU_NW * stride) * sizeof(double);
  TB_REQUIRE(stride <= 320 && shm_tau <= 200 * 1024, TB_ERR_ARG, "too many layers (%d) for the tau kernel", nrad);
  const size_t nin = (size_t)nrad * (1 + nmol), nvl_max = (size_t)Bmax * nrad;
  // ---- batch buffers
  TB_TRY(c->b_in.ensure(Bmax * nin)); TB_TRY(c->b_ltemp.ensure(nvl_max)); TB_TRY(c->b_traw.ensure(nvl_max));
  TB_TRY(c->b_dens.ensure(nvl_max * nmol)); TB_TRY(c->b_rad.ensure(nvl_max)); TB_TRY(c->b_ip.ensure(nvl_max));
  TB_TRY(c->b_L.ensure(nvl_max * sizeof(InterpLayer))); TB_TRY(c->b_rho.ensure(nvl_max * Nm));
  TB_TRY(c->b_densfac.ensure(nvl_max * (ntab ? ntab : 1))); TB_TRY(c->b_aux.ensure(2 * nvl_max));
  TB_TRY(c->b_e.ensure(nvl_max * nwn)); TB_TRY(c->b_ecs.ensure(nvl_max * nwn)); TB_TRY(c->b_tau.ensure(nvl_max * nwn));
  TB_TRY(c->b_last.ensure((size_t)Bmax * nwn)); TB_TRY(c->b_out.ensure((size_t)Bmax * nwn));
  TB_TRY(c->r_meta.ensure(nvl_max * 4)); TB_TRY(c->r_par.ensure(nvl_max * 4)); TB_TRY(c->r_coef.ensure(nvl_max * stride));
  TB_TRY(c->r_rcp.ensure(nvl_max * 2)); TB_TRY(c->r_s.ensure(nvl_max * (nrad + 3)));
  if (c->b_static_version != c->grid_version || c->b_gtemp.n < (size_t)Nt + 1) {
    TB_TRY(c->b_gtemp.ensure(Nt + 1)); TB_TRY(c->b_gmol.ensure(Nm));
    TB_CUDA(cudaMemcpyAsync(c->b_gtemp.p, c->gtemp.data(), (Nt + 1) * sizeof(double), cudaMemcpyHostToDevice, st));
    TB_CUDA(cudaMemcpyAsync(c->b_gmol.p, c->gmol_idx.data(), Nm * sizeof(int), cudaMemcpyHostToDevice, st));
    TB_CUDA(cudaStreamSynchronize(st));
    c->b_static_version = c->grid_version;
  }
  BatchPar B;
  B.nrad = nrad; B.nmol = nmol; B.bymass = c->am_bymass; B.Nt = Nt; B.Nm = Nm; B.ntab = ntab;
  B.p_fct = c->am_pfct; B.t_fct = c->am_tfct; B.rad_fct = c->am_rfct; B.p0 = c->am_p0; B.r0 = c->am_r0; B.gsurf = c->am_gsurf;
  B.ih2 = c->am_ih2;
  B.cia_tmin = 0.0; B.cia_tmax = 70000.0;
  for (int n = 0; n < ntab; n++) {
    CiaTable *t = c->cia[n];
    TB_TRY(tb_cia_table_grid(c, t));
    B.cia_tmin = fmax(B.cia_tmin, t->temp[0]); B.cia_tmax = fmin(B.cia_tmax, t->temp[t->nt - 1]);
    B.tab_nspec[n] = t->nspec; B.tab_mol[n][0] = t->mol[0]; B.tab_mol[n][1] = t->mol[1];
    TB_TRY(t->d_f2.ensure(nvl_max * t->nw)); TB_TRY(t->d_z2.ensure(nvl_max * t->nw)); TB_TRY(t->d_v.ensure(nvl_max * t->nw));
  }
  const size_t shm_tauw = (size_t)nrad * 32 * TAU_WL * sizeof(double);
  if (c->tauw_shm_set != (int)shm_tauw) {
    TB_CUDA(cudaFuncSetAttribute(k_tau_w<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm_tauw));
    TB_CUDA(cudaFuncSetAttribute(k_tau_w<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm_tauw));
    c->tauw_shm_set = (int)shm_tauw;
  }
  TB_REQUIRE(nh % TAU_RB == 0, TB_ERR_ARG, "tb_run_transit_batch: the layer count (%d) must be a multiple of %d", nh, TAU_RB);
  TB_TRY(c->r_wg.ensure(nvl_max * nrad)); TB_TRY(c->r_w0.ensure(nvl_max));
  TauPar P;
  P.eclipse = 0; P.nrad = nrad; P.nh = nh; P.stride = stride; P.nwn = nwn; P.w0 = 0; P.w1 = nwn;
  P.rfct = c->am_rfct; P.toomuch = toomuch; P.wn_i = c->wn_i; P.wn_d = c->wn_d; P.wn_fct = c->wn_fct;
  P.sc_flag = c->sc_flag; P.cl_flag = c->cl_flag;
  P.sc_fac = pow(10.0, c->sc_logext) * TB_E0H2;
  P.cl_ext = c->cl_ext; P.cl_top = c->cl_top; P.cl_bot = c->cl_bot; P.cl_gamma = c->cl_gamma;
  P.cl_Q = c->cl_Q; P.cl_r = c->cl_r; P.cl_sig = c->cl_sig; P.cl_refwn = c->cl_refwn;
  const size_t shm_mod = (5 * (size_t)nh + 2) * sizeof(double);
  TB_REQUIRE(shm_mod <= 48 * 1024, TB_ERR_ARG, "too many rays (%d) for the modulation kernel's shared memory", nh);
  TB_CUDA(cudaMemsetAsync(c->d_err.p, 0, 8 * sizeof(int32_t), st));
  c->toomuch = toomuch;
  tb_timer_begin(c, "batch");
  for (int a0 = 0; a0 < natm; a0 += Bmax) {
    const int nb = natm - a0 < Bmax ? natm - a0 : Bmax;
    const int nvl = nb * nrad;
    TB_CUDA(cudaMemcpyAsync(c->b_in.p, in + (size_t)a0 * nin, nb * nin * sizeof(double), cudaMemcpyHostToDevice, st));
    k_atm_setup<<<nb, 128, 0, st>>>(B, nb, c->b_in.p, c->b_p.p, c->d_mol_mass.p, c->b_gtemp.p, c->b_gmol.p, c->b_ltemp.p,
                                    c->b_traw.p, c->b_dens.p, c->b_rad.p, c->b_ip.p, reinterpret_cast<InterpLayer *>(c->b_L.p),
                                    c->b_rho.p, c->b_densfac.p, c->b_aux.p, c->d_err.p);
    TB_TRY(tb_interp_launch(c, nvl, nrad, reinterpret_cast<const InterpLayer *>(c->b_L.p), c->b_rho.p, c->b_e.p));
    // ---- CIA over nvl layers (every layer is inside every table's range: k_atm_setup flags the others)
    if (ntab > 0) {
      CiaPack CP;
      CiaSplinePack SP;
      int k = 0, maxrows = 0;
      for (int n = 0; n < ntab; n++) {
        CiaTable *t = c->cia[n];
        if (t->fi >= t->li) continue;
        CiaDev &D = CP.t[k];
        D.wn = t->d_wn.p; D.f2 = t->d_f2.p; D.z2 = t->d_z2.p; D.widx = t->d_widx.p;
        D.rh = t->d_rh.p; D.r6h = t->d_r6h.p; D.h6 = t->d_h6.p;
        D.densfac = c->b_densfac.p + (size_t)n * nvl;
        D.fi = (int)t->fi; D.li = (int)t->li; D.fj = 0; D.lj = nvl; D.ntw = t->nw;
        CiaSplineTab &T = SP.t[k];
        T.tab = t->d_tab.p; T.zt = t->d_zt.p; T.ttemp = t->d_temp.p; T.h = t->d_h.p; T.rh = t->d_rh.p;
        T.r = t->d_r.p; T.ru = t->d_ru.p; T.f2 = t->d_f2.p; T.v = t->d_v.p; T.z2 = t->d_z2.p;
        T.ntw = t->nw; T.ntt = t->nt; T.fj = 0; T.lj = nvl;
        if (t->nw > maxrows) maxrows = t->nw;
        k++;
      }
      CP.n = k;
      if (k > 0) {
        k_cia_temp<<<dim3(nblk((int64_t)maxrows * nvl, 256), k), 256, 0, st>>>(SP, c->b_ltemp.p, nvl);
        k_cia_wnspline<<<dim3(nblk(nvl, 4), k), 128, 0, st>>>(SP);
      }
      k_cia_eval<<<dim3(nblk(nwn, 256), nvl), 256, 0, st>>>(CP, nvl, c->wn_i, c->wn_d, c->wn_fct, nwn, 0, nwn, c->b_ecs.p);
      c->launches += 3;
    } else
      TB_CUDA(cudaMemsetAsync(c->b_ecs.p, 0, sizeof(double) * nvl * nwn, st));
    // ---- tau + modulation, the atmosphere as a grid dimension
    TB_CUDA(cudaMemsetAsync(c->b_tau.p, 0, sizeof(double) * nvl * nwn, st));
    k_ray_geometry<<<nb * nh, 64, 0, st>>>(0, c->b_rad.p, nrad, c->am_rfct, c->b_ip.p, nh, c->am_rfct, stride, c->r_meta.p,
                                           c->r_par.p, c->r_coef.p, c->r_s.p, c->r_rcp.p, c->d_err.p, c->r_wg.p, c->r_w0.p);
    auto tauw = nblk(nwn, 32 * TAU_WL) * nb <= 148 * 4 ? k_tau_w<4> : k_tau_w<3>;
    tauw<<<dim3(nblk(nwn, 32 * TAU_WL), nb), 32 * TAU_NW, shm_tauw, st>>>(P, c->b_e.p, c->b_ecs.p, c->b_p.p, c->b_traw.p, c->b_aux.p,
                                                                   c->b_dens.p, c->d_mol_pol.p, c->d_mol_mass.p, nmol,
                                                                   c->r_meta.p, c->r_par.p, c->r_rcp.p, c->r_wg.p, c->r_w0.p,
                                                                   c->b_tau.p, c->b_last.p);
    k_modulation<<<dim3(nblk(nwn * MOD_QL, 128), nb), 128, shm_mod, st>>>(nwn, 0, nwn, nh, c->b_tau.p, c->b_last.p, c->b_ip.p,
                                                                         c->am_rfct, toomuch, starrad_cm, transparent,
                                                                         c->b_out.p, c->d_err.p);
    c->launches += 5;
    TB_CUDA(cudaMemcpyAsync(out + (size_t)a0 * nwn, c->b_out.p, (size_t)nb * nwn * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (a0 + Bmax < natm) TB_CUDA(cudaStreamSynchronize(st));   // the batch buffers are reused
  }
  tb_timer_end(c, "batch");
  TB_CUDA(cudaGetLastError());
  int32_t err[8];
  TB_CUDA(cudaMemcpyAsync(err, c->d_err.p, sizeof(err), cudaMemcpyDeviceToHost, st));
  TB_CUDA(cudaStreamSynchronize(st));
  TB_REQUIRE(err[4] == 0, TB_ERR_DOMAIN, "Sum of abundances of isotopes adds up to more than 1 in some layer of the batch");
  TB_REQUIRE(err[5] == 0, TB_ERR_DOMAIN, "A layer of the batch has a temperature outside the cross-section tables' range (crosssec.c:293-309)");
  TB_REQUIRE(err[6] == 0, TB_ERR_DOMAIN, "A layer of the batch has a temperature outside the opacity grid");
  TB_REQUIRE(err[7] == 0, TB_ERR_DOMAIN, "The reference pressure lies outside the atmosphere, or the hydrostatic radii do not increase");
  TB_REQUIRE(err[1] == 0, TB_ERR_DOMAIN, "a ray's closest approach lies below the bottom layer (slantpath.c:40-44)");
  TB_REQUIRE(err[2] == 0, TB_ERR_DOMAIN, "less than 3 items for radial integration (slantpath.c:393-397)");
  return TB_OK;
}

// =============================================================== transit tau without any allocation
//
// tb_tau_transit_static: the transit optical depth t = W er (k_ray_geometry + k_tau_w, the kernels of tb_tau) for
// callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
// It has no counterpart in the reference.  The ray geometry --
// path coordinates, Simpson coefficients, the weight matrix W, the per-ray parabola constants, the error flags: the
// scratch every call of tb_tau keeps in cudaMalloc'ed context buffers -- lives here in ONE file-scope device array of
// fixed size: every GPU that loads the library holds TB_STATIC_SCRATCH_BYTES of it until the process exits.  Inputs
// and outputs are the caller's device buffers.  The array is shared by every call on a device and keeps the last
// call's contents: the part a call uses is cleared first, and calls that overlap on the GPU (two streams, two
// contexts of one device) race on it -- the caller serialises them.
#define TB_STATIC_MAXLAYERS 128            // most layers AND most rays of a call
#define TB_STATIC_STRIDE (4 * ((TB_STATIC_MAXLAYERS + 2) / 2 + 1))
#define TB_STATIC_NDBL ((size_t)TB_STATIC_MAXLAYERS * (TB_STATIC_MAXLAYERS + 3)      /* s     */ \
                        + (size_t)TB_STATIC_MAXLAYERS * 4                            /* par   */ \
                        + (size_t)TB_STATIC_MAXLAYERS * TB_STATIC_STRIDE             /* coef  */ \
                        + (size_t)TB_STATIC_MAXLAYERS * 2                            /* rcp   */ \
                        + (size_t)TB_STATIC_MAXLAYERS * TB_STATIC_MAXLAYERS          /* wg    */ \
                        + (size_t)TB_STATIC_MAXLAYERS)                               /* w0    */
#define TB_STATIC_NINT ((size_t)TB_STATIC_MAXLAYERS * 4 + 8)                         /* meta, err */
#define TB_STATIC_SCRATCH_BYTES (TB_STATIC_NDBL * 8 + TB_STATIC_NINT * 4)
static_assert(TB_STATIC_MAXLAYERS == TB_TAU_STATIC_MAXLAYERS && TB_STATIC_SCRATCH_BYTES == TB_TAU_STATIC_SCRATCH_BYTES,
              "include/transit_b200.h states the limits of tb_tau_transit_static");
static __device__ __align__(256) unsigned char tb_static_scratch[TB_STATIC_SCRATCH_BYTES];

// In host code `tb_static_scratch` is a host-side stand-in, not a device address: the address a kernel can use comes
// from cudaGetSymbolAddress, asked once per device with that device current.
#define TB_STATIC_MAXDEV 64
static unsigned char *tb_static_addr[TB_STATIC_MAXDEV];

extern "C" int tb_tau_transit_static(tb_ctx *c, int nrad, const double *d_rad, double rad_fct, int nh,
                                     const double *d_h, double hfct, int64_t nwn, const double *d_e,
                                     const double *d_ecs, double toomuch, double *d_tau, int32_t *d_last, int check) {
  TB_REQUIRE(c != nullptr, TB_ERR_ARG, "tb_tau_transit_static: no context");
  TB_REQUIRE(d_rad && d_h && d_e && d_ecs && d_tau && d_last, TB_ERR_ARG,
             "tb_tau_transit_static: rad, h, e, e_cs, tau and last are device buffers of the caller, none may be NULL");
  TB_REQUIRE(nrad >= 2 && nrad <= TB_STATIC_MAXLAYERS, TB_ERR_ARG,
             "tb_tau_transit_static: %d layers; the static scratch holds 2..%d (tb_tau has no such limit)", nrad,
             TB_STATIC_MAXLAYERS);
  TB_REQUIRE(nh >= 1 && nh <= TB_STATIC_MAXLAYERS, TB_ERR_ARG,
             "tb_tau_transit_static: %d rays; the static scratch holds 1..%d (tb_tau has no such limit)", nh,
             TB_STATIC_MAXLAYERS);
  TB_REQUIRE(nwn >= 1 && nwn < ((int64_t)1 << 31), TB_ERR_ARG, "tb_tau_transit_static: %lld wavenumbers", (long long)nwn);
  TB_REQUIRE(c->device >= 0 && c->device < TB_STATIC_MAXDEV, TB_ERR_ARG, "tb_tau_transit_static: device %d", c->device);
  TB_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  if (!tb_static_addr[c->device]) {            // once per device, that device current
    unsigned char *ptr = nullptr;
    TB_CUDA(cudaGetSymbolAddress((void **)&ptr, tb_static_scratch));
    TB_REQUIRE(ptr != nullptr, TB_ERR_CUDA, "tb_tau_transit_static: no device address for the static scratch");
    tb_static_addr[c->device] = ptr;
  }
  // ---- carve-up of the array (doubles first: every part stays 8-byte aligned)
  const int stride = 4 * ((nrad + 2) / 2 + 1);
  const int ngrp = (nh + TAU_RB - 1) / TAU_RB;
  double *dbl = reinterpret_cast<double *>(tb_static_addr[c->device]);
  double *s_s = dbl;                                  dbl += (size_t)nh * (nrad + 3);
  double *s_par = dbl;                                dbl += (size_t)nh * 4;
  double *s_coef = dbl;                               dbl += (size_t)nh * stride;
  double *s_rcp = dbl;                                dbl += (size_t)nh * 2;
  double *s_wg = dbl;                                 dbl += (size_t)ngrp * nrad * TAU_RB;
  double *s_w0 = dbl;                                 dbl += nh;
  int32_t *s_meta = reinterpret_cast<int32_t *>(dbl);
  int32_t *s_err = s_meta + (size_t)nh * 4;
  const size_t used = (size_t)(reinterpret_cast<unsigned char *>(s_err + 8) - tb_static_addr[c->device]);
  TB_REQUIRE(used <= TB_STATIC_SCRATCH_BYTES, TB_ERR_ARG, "tb_tau_transit_static: internal: scratch carve-up too large");
  // the array keeps the last call's contents: error flags, the weights of the rays a ragged last group of four
  // does not have, and everything else this call will read start from zero
  TB_CUDA(cudaMemsetAsync(tb_static_addr[c->device], 0, used, st));
  TB_CUDA(cudaMemsetAsync(d_tau, 0, sizeof(double) * (size_t)nh * nwn, st));   // calloc, tau.c:34: zero beyond last[]
  k_ray_geometry<<<nh, 64, 0, st>>>(0, d_rad, nrad, rad_fct, d_h, nh, hfct, stride, s_meta, s_par, s_coef, s_s, s_rcp,
                                    s_err, s_wg, s_w0);
  TauPar P;
  memset(&P, 0, sizeof(P));
  P.eclipse = 0; P.nrad = nrad; P.nh = nh; P.stride = stride; P.nwn = nwn; P.w0 = 0; P.w1 = nwn;
  P.rfct = rad_fct; P.toomuch = toomuch; P.wn_i = 0; P.wn_d = 0; P.wn_fct = 1;
  P.sc_flag = 0; P.cl_flag = 0;                      // no scattering, no cloud: the caller folds them into e
  const size_t shm = (size_t)nrad * 32 * TAU_WL * sizeof(double);
  const bool one_wave = nblk(nwn, 32 * TAU_WL) <= 148 * 4;
  auto tauw = one_wave ? k_tau_w<4> : k_tau_w<3>;
  {  // the dynamic shared memory limit of the kernel is per device and shared with tb_tau's launches: raise, never lower
    cudaFuncAttributes fa;
    TB_CUDA(cudaFuncGetAttributes(&fa, tauw));
    if ((size_t)fa.maxDynamicSharedSizeBytes < shm)
      TB_CUDA(cudaFuncSetAttribute(tauw, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)shm));
  }
  tauw<<<nblk(nwn, 32 * TAU_WL), 32 * TAU_NW, shm, st>>>(P, d_e, d_ecs, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                                                        0, s_meta, s_par, s_rcp, s_wg, s_w0, d_tau, d_last);
  c->launches += 2;
  TB_CUDA(cudaGetLastError());
  if (check) {                                       // the geometry's error flags, one small copy and a sync of the stream
    int32_t err[4];
    TB_CUDA(cudaMemcpyAsync(err, s_err, sizeof(err), cudaMemcpyDeviceToHost, st));
    TB_CUDA(cudaStreamSynchronize(st));
    TB_REQUIRE(err[1] == 0, TB_ERR_DOMAIN, "a ray's closest approach lies below the bottom layer (slantpath.c:40-44)");
  }
  return TB_OK;
}
