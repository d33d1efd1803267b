// tb_opacity.cu -- the opacity grid: calcopacity (transit/src/opacity.c:281-427), readopacity
// (opacity.c:432-503) and interpolmolext (extinction.c:534-581).  The grid keeps the file's
// order o[layer][temp][molecule][wavenumber] in HBM (wavenumber innermost = coalesced), and
// replaces the reference's SysV shared-memory copies (opacity.c:507-630) by residency.
#include "tb_internal.cuh"

static inline int nblk(int64_t n, int t) { return (int)((n + t - 1) / t); }

extern "C" int tb_opacity_load(tb_ctx *c, int64_t Nmol, int64_t Ntemp, int64_t Nlayer, int64_t Nwave,
                               const int *molID, const double *temp, const double *press,
                               const double *grid) {
  TB_REQUIRE(c->have_grid && c->have_species, TB_ERR_STATE, "grid and species must be set before tb_opacity_load");
  TB_REQUIRE(Nmol >= 1 && Ntemp >= 2 && Nlayer >= 1 && Nwave == c->nwn, TB_ERR_ARG,
             "opacity grid shape (%lld mol, %lld T, %lld layers, %lld wn) does not fit the problem (nwn=%lld)",
             (long long)Nmol, (long long)Ntemp, (long long)Nlayer, (long long)Nwave, (long long)c->nwn);
  TB_CUDA(cudaSetDevice(c->device));
  c->gNmol = Nmol; c->gNtemp = Ntemp; c->gNlayer = Nlayer; c->gNwave = Nwave;
  c->gtemp.assign(temp, temp + Ntemp);
  c->gtemp.push_back(INFINITY);   // binsearchapprox(gtemp, T, 0, Ntemp) looks one past the end
  c->gpress.assign(press, press + Nlayer);
  c->gmolID.assign(molID, molID + Nmol);
  c->gmol_idx.assign(Nmol, -1);
  for (int m = 0; m < Nmol; m++) {       // valueinarray(mol->ID, gmol[m], nmol), extinction.c:575
    for (int k = 0; k < c->nmol; k++)
      if (c->mol_ID[k] == molID[m]) { c->gmol_idx[m] = k; break; }
    TB_REQUIRE(c->gmol_idx[m] >= 0, TB_ERR_ARG, "opacity-grid molecule ID %d is not in the atmosphere", molID[m]);
  }
  const size_t n = (size_t)Nlayer * Ntemp * Nmol * Nwave;
  TB_TRY(c->d_grid.ensure(n));
  if (grid) {
    TB_CUDA(cudaMemcpyAsync(c->d_grid.p, grid, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    TB_CUDA(cudaStreamSynchronize(c->stream));
  }
  c->have_grid_o = true;
  return TB_OK;
}


// One grid in HBM for several contexts of a GPU (what the reference's SysV shared-memory scheme,
// opacity.c:89-201,507-630, does between processes): dst reads src's grid in place -- no second copy of a 2.6 GB
// grid per retrieval thread.  The grid is read-only after tb_opacity_load; src must outlive dst's use of it (a grid
// loaded again into src, or src destroyed, leaves dst with a dangling pointer: call tb_opacity_share again).
extern "C" int tb_opacity_share(tb_ctx *dst, tb_ctx *src) {
  TB_REQUIRE(dst && src && dst != src, TB_ERR_ARG, "tb_opacity_share: two different contexts");
  TB_REQUIRE(src->have_grid_o && src->d_grid.p, TB_ERR_STATE, "tb_opacity_share: the source context holds no opacity grid");
  TB_REQUIRE(dst->device == src->device, TB_ERR_ARG, "tb_opacity_share: contexts on devices %d and %d", dst->device, src->device);
  TB_REQUIRE(dst->have_grid && dst->have_species, TB_ERR_STATE, "grid and species must be set before tb_opacity_share");
  TB_REQUIRE(dst->nwn == src->gNwave, TB_ERR_ARG, "tb_opacity_share: the grid has %lld wavenumbers, the problem %lld",
             (long long)src->gNwave, (long long)dst->nwn);
  std::vector<int> idx(src->gNmol, -1);
  for (int m = 0; m < src->gNmol; m++) {
    for (int k = 0; k < dst->nmol; k++)
      if (dst->mol_ID[k] == src->gmolID[m]) { idx[m] = k; break; }
    TB_REQUIRE(idx[m] >= 0, TB_ERR_ARG, "opacity-grid molecule ID %d is not in the atmosphere", src->gmolID[m]);
  }
  dst->gNmol = src->gNmol; dst->gNtemp = src->gNtemp; dst->gNlayer = src->gNlayer; dst->gNwave = src->gNwave;
  dst->gtemp = src->gtemp; dst->gpress = src->gpress; dst->gmolID = src->gmolID;
  dst->gmol_idx = idx;
  dst->d_grid.borrow(src->d_grid.p, src->d_grid.n);
  dst->have_grid_o = true;
  return TB_OK;
}

// extinction.c:568-578: e[r][w] = sum_m rho_m[r] * (o_lo*(T+ - T) + o_hi*(T - T-))/(T+ - T-)
// A batch of atmospheres (tb_batch.cu) is nrad*natm "virtual layers": virtual layer v = a*ngrid + r reads the grid
// planes of layer r = v % ngrid with its own temperature bracket and densities; rho[Nmol][nrad], e[nrad][nwn].
__global__ void k_grid_interp(int64_t nwn, int nrad, int ngrid, int Ntemp, int Nmol, const double *__restrict__ grid,
                              const InterpLayer *__restrict__ L, const double *__restrict__ rho /*[Nmol][nrad]*/,
                              double *__restrict__ e) {
  int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int r = blockIdx.y + blockIdx.z * gridDim.y;
  if (w >= nwn || r >= nrad) return;
  // blocks are dealt layer-major, atmosphere fastest: the atmospheres of a batch that bracket a layer's temperature
  // alike read the same grid planes while those are in L2 (2 MB per (atmosphere, layer), 126 MB of L2)
  const int natm = nrad / ngrid;
  r = (r % natm) * ngrid + r / natm;
  const InterpLayer l = L[r];
  const double *lo = grid + (((size_t)(r % ngrid) * Ntemp + l.it) * Nmol) * nwn + w;
  const double *hi = lo + (size_t)Nmol * nwn;
  double acc = 0.0;
  // the 2 x 4 plane reads of four molecules are issued before the first division depends on one:
  // same operations in the same order as the loop over m, four times the bytes in flight
  for (int m0 = 0; m0 < Nmol; m0 += 4) {
    double vlo[4], vhi[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const bool in = m0 + k < Nmol;
      // one atmosphere: every plane is read once (streaming hint); a batch: the planes are shared through L2
      vlo[k] = !in ? 0.0 : (natm == 1 ? __ldcs(lo + (size_t)(m0 + k) * nwn) : __ldg(lo + (size_t)(m0 + k) * nwn));
      vhi[k] = !in ? 0.0 : (natm == 1 ? __ldcs(hi + (size_t)(m0 + k) * nwn) : __ldg(hi + (size_t)(m0 + k) * nwn));
    }
#pragma unroll
    for (int k = 0; k < 4; k++)
      if (m0 + k < Nmol) {
        // the divisor depends on the layer only: its reciprocal comes with the layer, the quotient is completed to
        // the correctly rounded one (tb_div_r)
        const double ext = tb_div_r(vlo[k] * (l.thi - l.t) + vhi[k] * (l.t - l.tlo), l.thi - l.tlo, l.rdt);
        acc += rho[(m0 + k) * nrad + r] * ext;
      }
  }
  e[(size_t)r * nwn + w] = acc;
}

int tb_interp_run(tb_ctx *c) {
  TB_REQUIRE(c->have_grid_o && c->have_atm, TB_ERR_STATE, "tb_opacity_load and tb_set_atmosphere must precede the interpolation");
  TB_REQUIRE(c->gNlayer == c->nrad, TB_ERR_ARG, "opacity grid has %lld layers, atmosphere %d", (long long)c->gNlayer, c->nrad);
  TB_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int nrad = c->nrad, Nt = (int)c->gNtemp, Nm = (int)c->gNmol;
  std::vector<InterpLayer> L(nrad);
  std::vector<double> rho((size_t)Nm * nrad);
  for (int r = 0; r < nrad; r++) {
    const double T = c->t[r] * c->t_fct;
    // binsearchapprox(gtemp, temp, 0, Ntemp) then step down (extinction.c:562-564)
    int lo = 0, hi = Nt;
    while (hi - lo > 1) {
      int mid = (hi + lo) / 2;
      if (c->gtemp[mid] > T) hi = mid; else lo = mid;
    }
    int it = (fabs(c->gtemp[hi] - T) < fabs(c->gtemp[lo] - T)) ? hi : lo;
    if (T < c->gtemp[it]) it--;
    TB_REQUIRE(it >= 0 && it + 1 < Nt, TB_ERR_DOMAIN,
               "layer %d temperature %.2f K is outside the opacity grid [%.1f, %.1f] K", r, T, c->gtemp[0], c->gtemp[Nt - 1]);
    L[r].it = it; L[r].t = T; L[r].tlo = c->gtemp[it]; L[r].thi = c->gtemp[it + 1];
    L[r].rdt = 1.0 / (L[r].thi - L[r].tlo);
    for (int m = 0; m < Nm; m++) rho[(size_t)m * nrad + r] = c->dens[(size_t)c->gmol_idx[m] * nrad + r];
  }
  // staging buffers live in the context: a retrieval calls this once per atmosphere
  TB_TRY(c->i_stage.ensure(nrad * sizeof(InterpLayer))); TB_TRY(c->i_rho.ensure(rho.size()));
  TB_TRY(c->d_e.ensure((size_t)nrad * c->nwn));
  InterpLayer *dLp = reinterpret_cast<InterpLayer *>(c->i_stage.p);
  double *drhop = c->i_rho.p;
  TB_TRY(tb_upload_small(c, dLp, L.data(), nrad * sizeof(InterpLayer), st));   // copied into the pinned ring:
  TB_TRY(tb_upload_small(c, drhop, rho.data(), rho.size() * sizeof(double), st)); // the locals may die
  dim3 g(nblk(c->nwn, 256), nrad);
  tb_timer_begin(c, "interp");
  k_grid_interp<<<g, 256, 0, st>>>(c->nwn, nrad, nrad, Nt, Nm, c->d_grid.p, dLp, drhop, c->d_e.p);
  tb_timer_end(c, "interp");
  TB_LAUNCHED(c);
  TB_CUDA(cudaGetLastError());
  c->have_e = true;
  return TB_OK;
}

// interpolmolext for a BATCH of atmospheres (tb_run_transit_batch): a block owns 256 wavenumbers of ONE grid layer
// and serves every atmosphere from it.  The atmospheres of a retrieval bracket a layer's temperature with a handful
// of neighbouring grid temperatures, so the block walks those brackets upwards -- the upper planes of one bracket
// stay in registers as the lower planes of the next -- and every grid plane is read ONCE per batch instead of once
// per atmosphere (2 MB per (atmosphere, layer) otherwise: 6.4 GB per 32 atmospheres of C5).  Per atmosphere the
// operations and their order are those of k_grid_interp: same bits.
#define GI_MAXMOL 8
__global__ void k_grid_interp_batch(int64_t nwn, int natm, int ngrid, int Ntemp, int Nmol, const double *__restrict__ grid,
                                    const InterpLayer *__restrict__ L /*[natm*ngrid]*/,
                                    const double *__restrict__ rho /*[Nmol][natm*ngrid]*/, double *__restrict__ e) {
  const int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int r = blockIdx.y, nvl = natm * ngrid;
  __shared__ int sh_it[64];
  int lo_it = Ntemp, hi_it = -1;
  for (int a0 = 0; a0 < natm; a0 += 64) {              // (warp-uniform: every thread scans the same brackets)
    __syncthreads();
    if (threadIdx.x < 64 && a0 + threadIdx.x < natm) sh_it[threadIdx.x] = L[(size_t)(a0 + threadIdx.x) * ngrid + r].it;
    __syncthreads();
    for (int k = 0; k < 64 && a0 + k < natm; k++) { lo_it = min(lo_it, sh_it[k]); hi_it = max(hi_it, sh_it[k]); }
  }
  if (w >= nwn) return;
  const double *pl = grid + (((size_t)r * Ntemp) * Nmol) * nwn + w;
  double vlo[GI_MAXMOL], vhi[GI_MAXMOL];
#pragma unroll
  for (int m = 0; m < GI_MAXMOL; m++) vlo[m] = m < Nmol ? __ldcs(pl + ((size_t)lo_it * Nmol + m) * nwn) : 0.0;
  for (int it = lo_it; it <= hi_it; it++) {
#pragma unroll
    for (int m = 0; m < GI_MAXMOL; m++) vhi[m] = m < Nmol ? __ldcs(pl + ((size_t)(it + 1) * Nmol + m) * nwn) : 0.0;
    for (int a = 0; a < natm; a++) {
      const int v = a * ngrid + r;
      const InterpLayer l = L[v];
      if (l.it != it) continue;
      double acc = 0.0;
#pragma unroll
      for (int m = 0; m < GI_MAXMOL; m++)
        if (m < Nmol) {
          const double ext = tb_div_r(vlo[m] * (l.thi - l.t) + vhi[m] * (l.t - l.tlo), l.thi - l.tlo, l.rdt);
          acc += rho[(size_t)m * nvl + v] * ext;
        }
      e[(size_t)v * nwn + w] = acc;
    }
#pragma unroll
    for (int m = 0; m < GI_MAXMOL; m++) vlo[m] = vhi[m];
  }
}

int tb_interp_launch(tb_ctx *c, int nvl, int ngrid, const InterpLayer *dL, const double *drho, double *e) {
  tb_timer_begin(c, "interp");
  if (c->gNmol <= GI_MAXMOL && nvl > ngrid) {
    k_grid_interp_batch<<<dim3(nblk(c->nwn, 256), ngrid), 256, 0, c->stream>>>(c->nwn, nvl / ngrid, ngrid, (int)c->gNtemp,
                                                                              (int)c->gNmol, c->d_grid.p, dL, drho, e);
  } else {       // one atmosphere, or more molecules than the batch kernel keeps in registers: a virtual layer per block row
    const int ny = nvl < 32768 ? nvl : 32768;
    dim3 g(nblk(c->nwn, 256), ny, (nvl + ny - 1) / ny);
    k_grid_interp<<<g, 256, 0, c->stream>>>(c->nwn, nvl, ngrid, (int)c->gNtemp, (int)c->gNmol, c->d_grid.p, dL, drho, e);
  }
  tb_timer_end(c, "interp");
  TB_LAUNCHED(c);
  return TB_OK;
}

extern "C" int tb_extinction_interp(tb_ctx *c) { return tb_interp_run(c); }

// calcopacity (opacity.c:387-403): every (layer, temperature) pair is one state of the line
// sum with permol=1; densities follow stateeqnford (transit.h:58-69).
extern "C" int tb_opacity_build(tb_ctx *c, int Ntemp, const double *temp, const double *ziso,
                                const double *q, const double *mm, int bymass, double ethresh, double *grid) {
  TB_REQUIRE(c->have_atm && c->have_lines && c->have_prof, TB_ERR_STATE,
             "atmosphere, lines and profiles must be set before tb_opacity_build");
  TB_REQUIRE(Ntemp >= 1 && temp && ziso && q && grid, TB_ERR_ARG, "tb_opacity_build: bad arguments");
  const int nrad = c->nrad, nmol = c->nmol, niso = c->niso;
  const int ns = nrad * Ntemp;
  std::vector<double> T(ns), dens((size_t)ns * nmol), Z((size_t)ns * (niso ? niso : 1));
  for (int r = 0; r < nrad; r++)
    for (int t = 0; t < Ntemp; t++) {
      const int s = r * Ntemp + t;
      T[s] = temp[t];
      const double press = c->p[r] * c->p_fct;            // opacity.c:345
      for (int j = 0; j < nmol; j++) {
        const double rho = TB_AMU * q[(size_t)j * nrad + r] * press / TB_KB / temp[t];
        dens[(size_t)s * nmol + j] = bymass ? rho * mm[r] : rho * c->mol_mass[j];
      }
      for (int i = 0; i < niso; i++) Z[(size_t)s * niso + i] = ziso[(size_t)i * Ntemp + t];
    }
  TB_TRY(tb_states_run(c, ns, T.data(), dens.data(), Z.data(), ethresh, 1, nullptr, grid));
  // keep the freshly built grid resident for interpolation, like a second run would read it
  std::vector<double> press(nrad);
  for (int r = 0; r < nrad; r++) press[r] = c->p[r] * c->p_fct;
  TB_TRY(tb_opacity_load(c, c->nrows, Ntemp, nrad, c->nwn, c->row_molID.data(), temp, press.data(), grid));
  return TB_OK;
}
