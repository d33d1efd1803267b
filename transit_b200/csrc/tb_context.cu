// tb_context.cu -- context life cycle, static problem description, taps, micro-benchmarks.
#include <cstring>
#include <stdarg.h>
#include "tb_internal.cuh"

static thread_local char g_err[1024] = "";

void tb_set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char *tb_last_error(void) { return g_err; }
extern "C" int tb_version(void) { return 100; }

void tb_timer_begin(tb_ctx *c, const char *name) {
  KernelTimer &t = c->timers[name];
  if (!t.a) {
    cudaEventCreate(&t.a);
    cudaEventCreate(&t.b);
  }
  cudaEventRecord(t.a, c->stream);
}
void tb_timer_end(tb_ctx *c, const char *name) {
  KernelTimer &t = c->timers[name];
  cudaEventRecord(t.b, c->stream);
  t.pending = true;
}

extern "C" int tb_get_kernel_ms(tb_ctx *c, const char *name, float *ms) {
  auto it = c->timers.find(name);
  TB_REQUIRE(it != c->timers.end(), TB_ERR_ARG, "no timer named '%s'", name);
  KernelTimer &t = it->second;
  if (t.pending) {
    TB_CUDA(cudaEventSynchronize(t.b));
    TB_CUDA(cudaEventElapsedTime(&t.ms, t.a, t.b));
    t.pending = false;
  }
  *ms = t.ms;
  return TB_OK;
}
extern "C" int tb_timer_start(tb_ctx *c, const char *name) {
  TB_CUDA(cudaSetDevice(c->device));
  tb_timer_begin(c, name);
  return TB_OK;
}
extern "C" int tb_timer_stop(tb_ctx *c, const char *name) {
  TB_CUDA(cudaSetDevice(c->device));
  tb_timer_end(c, name);
  return TB_OK;
}
extern "C" int64_t tb_launch_count(tb_ctx *c) { return c->launches; }
extern "C" void *tb_stream(tb_ctx *c) { return (void *)c->stream; }

extern "C" int tb_set_stream(tb_ctx *c, void *stream) {
  TB_CUDA(cudaSetDevice(c->device));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  c->stream = stream ? (cudaStream_t)stream : c->own_stream;
  return TB_OK;
}

extern "C" int tb_device_spectrum(tb_ctx *c, void **ptr, int64_t *n) {
  TB_REQUIRE(c->d_out.p && c->nwn > 0, TB_ERR_STATE, "no spectrum computed yet");
  *ptr = (void *)c->d_out.p;
  *n = c->nwn;
  return TB_OK;
}

static int tb_create_resources(tb_ctx *c) {
  TB_CUDA(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
  c->stream = c->own_stream;
  TB_CUDA(cudaStreamCreateWithFlags(&c->aux_stream, cudaStreamNonBlocking));
  TB_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
  TB_CUDA(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
  TB_TRY(c->d_err.ensure(8));
  TB_CUDA(cudaMemsetAsync(c->d_err.p, 0, 8 * sizeof(int32_t), c->stream));
  return TB_OK;
}

extern "C" int tb_create(int device, tb_ctx **out) {
  TB_REQUIRE(out, TB_ERR_ARG, "tb_create: out is NULL");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    tb_set_error("no CUDA device available (%s); this library has no CPU fallback",
                 e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    return TB_ERR_CUDA;
  }
  TB_REQUIRE(device >= 0 && device < ndev, TB_ERR_ARG, "device %d out of range (0..%d)", device,
             ndev - 1);
  TB_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  TB_CUDA(cudaGetDeviceProperties(&prop, device));
  TB_REQUIRE(prop.major >= 10, TB_ERR_CUDA,
             "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major,
             prop.minor);
  tb_ctx *c = new tb_ctx();
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  const int rc = tb_create_resources(c);
  if (rc != TB_OK) {           // whatever was created so far goes away with the context
    tb_destroy(c);
    return rc;
  }
  *out = c;
  return TB_OK;
}

extern "C" int tb_set_shard(tb_ctx *c, int64_t j_lo, int64_t j_hi, int rank, int world, tb_exchange_fn fn,
                            void *user) {
  TB_REQUIRE(c->have_grid, TB_ERR_STATE, "tb_set_wn_grid must precede tb_set_shard");
  TB_REQUIRE(world >= 1 && rank >= 0 && rank < world, TB_ERR_ARG, "tb_set_shard: rank %d of %d", rank, world);
  TB_REQUIRE(world == 1 || fn || (c->comm && c->comm_world == world && c->comm_rank == rank), TB_ERR_ARG,
             "tb_set_shard: world > 1 needs an exchange callback or a communicator of that size (tb_comm_init)");
  if (j_lo == 0 && j_hi == 0) j_hi = c->nwn;
  TB_REQUIRE(j_lo >= 0 && j_lo <= j_hi && j_hi <= c->nwn, TB_ERR_ARG, "tb_set_shard: band [%lld, %lld) outside the grid",
             (long long)j_lo, (long long)j_hi);
  TB_REQUIRE(j_lo % 128 == 0 && (j_hi % 128 == 0 || j_hi == c->nwn), TB_ERR_ARG,
             "tb_set_shard: band ends must be multiples of 128 (or the grid end)");
  c->band_lo = j_lo; c->band_hi = j_hi;
  c->banded = !(j_lo == 0 && j_hi == c->nwn);       // an EMPTY band (j_lo == j_hi > 0) is a valid shard
  c->shard_rank = rank; c->shard_world = world;
  c->exch = fn; c->exch_user = user;
  return TB_OK;
}

extern "C" int tb_sync(tb_ctx *c) {
  TB_CUDA(cudaSetDevice(c->device));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  return TB_OK;
}

extern "C" void tb_destroy(tb_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (c->aux_stream) cudaStreamSynchronize(c->aux_stream);
  tb_comm_destroy(c);
  // every DevBuf member frees itself in its destructor (delete c below)
  for (CiaTable *t : c->cia) delete t;
  c->cia.clear();
  for (auto &kv : c->timers) {
    if (kv.second.a) cudaEventDestroy(kv.second.a);
    if (kv.second.b) cudaEventDestroy(kv.second.b);
  }
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  if (c->aux_stream) cudaStreamDestroy(c->aux_stream);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  if (c->h_ring) cudaFreeHost(c->h_ring);
  delete c;
}

// makesample1 (makesample.c:78-95)
extern "C" int64_t tb_grid_count(double wn_i, double wn_f, double wn_d, int osamp) {
  int64_t n = (int64_t)(((1.0 + 1e-8) * wn_f - wn_i) / wn_d + 1);
  if (n < 0) n = -n;
  return (n - 1) * osamp + 1;
}

extern "C" int tb_set_wn_grid(tb_ctx *c, double wn_i, double wn_d, int wnosamp, int64_t nwn,
                              int64_t onwn, double wn_fct) {
  TB_REQUIRE(wn_d > 0 && wnosamp >= 1 && nwn >= 2, TB_ERR_ARG,
             "wavenumber grid needs spacing>0, oversampling>=1 and >=2 samples (extinction.c:176)");
  TB_REQUIRE(onwn == (nwn - 1) * wnosamp + 1, TB_ERR_ARG,
             "onwn (%lld) != (nwn-1)*wnosamp+1 (%lld)", (long long)onwn,
             (long long)((nwn - 1) * wnosamp + 1));
  TB_REQUIRE(onwn < 2000000000LL, TB_ERR_ARG, "fine grid too large for int indices (reference limit)");
  c->wn_i = wn_i;
  c->wn_d = wn_d;
  c->o = wnosamp;
  c->nwn = nwn;
  c->onwn = onwn;
  c->wn_fct = wn_fct;
  c->odwn = wn_d / (double)wnosamp;
  c->own_last = wn_i + (double)(onwn - 1) * c->odwn;  // owns.v[onwn-1], mul then add
  c->have_grid = true;
  c->grid_version++;
  c->have_lines = c->have_prof = false;  // both depend on the grid
  return TB_OK;
}

// host -> device copy of a small array through the context's pinned ring (see tb_ctx::h_ring)
int tb_upload_small(tb_ctx *c, void *dst, const void *src, size_t bytes, cudaStream_t st) {
  if (!bytes) return TB_OK;
  if (bytes > TB_RING_BYTES / 8) {                     // not small: the ordinary staged copy
    TB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st));
    return TB_OK;
  }
  if (!c->h_ring) TB_CUDA(cudaMallocHost((void **)&c->h_ring, TB_RING_BYTES));
  const size_t need = (bytes + 63) & ~size_t(63);
  if (c->ring_pos + need > TB_RING_BYTES) {            // wrap: whatever still reads the ring must have left
    TB_CUDA(cudaStreamSynchronize(c->stream));
    TB_CUDA(cudaStreamSynchronize(c->own_stream));
    TB_CUDA(cudaStreamSynchronize(c->aux_stream));
    if (st != c->stream) TB_CUDA(cudaStreamSynchronize(st));
    c->ring_pos = 0;
  }
  char *slot = c->h_ring + c->ring_pos;
  memcpy(slot, src, bytes);
  TB_CUDA(cudaMemcpyAsync(dst, slot, bytes, cudaMemcpyHostToDevice, st));
  c->ring_pos += need;
  return TB_OK;
}

template <typename T>
static int upload(tb_ctx *c, DevBuf<T> &b, const T *h, size_t n) {
  TB_TRY(b.ensure(n ? n : 1));
  return tb_upload_small(c, b.p, h, n * sizeof(T), c->stream);
}

extern "C" int tb_set_species(tb_ctx *c, int nmol, const double *mol_mass, const double *mol_radius,
                              const double *mol_pol, const int *mol_ID, int niso,
                              const double *iso_mass, const double *iso_ratio,
                              const int *iso_imol) {
  TB_REQUIRE(nmol > 0 && niso >= 0 && niso < 32000, TB_ERR_ARG, "bad species counts");
  TB_CUDA(cudaSetDevice(c->device));
  c->nmol = nmol;
  c->niso = niso;
  c->mol_mass.assign(mol_mass, mol_mass + nmol);
  c->mol_radius.assign(mol_radius, mol_radius + nmol);
  if (mol_pol) c->mol_pol.assign(mol_pol, mol_pol + nmol); else c->mol_pol.assign(nmol, 0.0);
  c->mol_ID.assign(mol_ID, mol_ID + nmol);
  c->iso_mass.assign(iso_mass, iso_mass + niso);
  c->iso_ratio.assign(iso_ratio, iso_ratio + niso);
  c->iso_imol.assign(iso_imol, iso_imol + niso);
  // molecules with line transitions in first-appearance order (opacity.c:349-361)
  c->row_molID.clear();
  c->iso_row.assign(niso, 0);
  for (int i = 0; i < niso; i++) {
    TB_REQUIRE(iso_imol[i] >= 0 && iso_imol[i] < nmol, TB_ERR_ARG, "iso_imol[%d] out of range", i);
    int id = mol_ID[iso_imol[i]], row = -1;
    for (size_t k = 0; k < c->row_molID.size(); k++)
      if (c->row_molID[k] == id) row = (int)k;
    if (row < 0) {
      row = (int)c->row_molID.size();
      c->row_molID.push_back(id);
    }
    c->iso_row[i] = row;
  }
  c->nrows = (int)c->row_molID.size();
  TB_TRY(upload(c, c->d_mol_mass, c->mol_mass.data(), nmol));
  TB_TRY(upload(c, c->d_mol_radius, c->mol_radius.data(), nmol));
  TB_TRY(upload(c, c->d_mol_pol, c->mol_pol.data(), nmol));
  TB_TRY(upload(c, c->d_iso_mass, c->iso_mass.data(), niso));
  TB_TRY(upload(c, c->d_iso_ratio, c->iso_ratio.data(), niso));
  TB_TRY(upload(c, c->d_iso_imol, c->iso_imol.data(), niso));
  TB_TRY(upload(c, c->d_iso_row, c->iso_row.data(), niso));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  c->have_species = true;
  c->have_lines = false;
  return TB_OK;
}

extern "C" int tb_get_line_molIDs(tb_ctx *c, int *molID, int *nmolrows) {
  TB_REQUIRE(c->have_species, TB_ERR_STATE, "tb_set_species not called");
  *nmolrows = c->nrows;
  if (molID) for (int i = 0; i < c->nrows; i++) molID[i] = c->row_molID[i];
  return TB_OK;
}

extern "C" int tb_get_line_counts(tb_ctx *c, int64_t *nlines, int64_t *ngroups, int64_t *nmembers) {
  TB_REQUIRE(c->have_lines, TB_ERR_STATE, "tb_upload_lines not called");
  if (nlines) *nlines = c->nlines;
  if (ngroups) *ngroups = c->G;
  if (nmembers) *nmembers = c->M;
  return TB_OK;
}

extern "C" int tb_set_atmosphere(tb_ctx *c, int nrad, const double *rad, double rad_fct,
                                 const double *p, double p_fct, const double *t, double t_fct,
                                 const double *density, const double *Z) {
  TB_REQUIRE(c->have_species, TB_ERR_STATE, "tb_set_species must precede tb_set_atmosphere");
  TB_REQUIRE(nrad >= 1, TB_ERR_ARG, "need at least one layer (extinction.c:168)");
  TB_CUDA(cudaSetDevice(c->device));
  c->nrad = nrad;
  c->rad_fct = rad_fct;
  c->p_fct = p_fct;
  c->t_fct = t_fct;
  c->rad.assign(rad, rad + nrad);
  c->p.assign(p, p + nrad);
  c->t.assign(t, t + nrad);
  c->dens.assign(density, density + (size_t)c->nmol * nrad);
  if (Z && c->niso) c->Z.assign(Z, Z + (size_t)c->niso * nrad); else c->Z.assign((size_t)c->niso * nrad, 1.0);
  for (int i = 1; i < nrad; i++)
    TB_REQUIRE(rad[i] > rad[i - 1], TB_ERR_ARG, "radii must be strictly increasing (readatm.c:585-601)");
  TB_TRY(upload(c, c->d_rad, c->rad.data(), nrad));
  TB_TRY(upload(c, c->d_p, c->p.data(), nrad));
  TB_TRY(upload(c, c->d_t, c->t.data(), nrad));
  TB_TRY(upload(c, c->d_dens, c->dens.data(), (size_t)c->nmol * nrad));
  c->have_atm = true;
  c->have_e = c->have_ecs = c->have_tau = false;
  return TB_OK;
}

extern "C" int tb_set_scattering(tb_ctx *c, int flag, double logext) {
  TB_REQUIRE(flag >= 0 && flag <= 2, TB_ERR_ARG, "scattering flag must be 0, 1 or 2");
  c->sc_flag = flag;
  c->sc_logext = logext;
  return TB_OK;
}

extern "C" int tb_set_cloud(tb_ctx *c, int flag, double cloudext, double cloudtop, double cloudbot,
                            double gamma, double Q, double r, double sig, double refwn,
                            const double *nH, const double *mean_dens) {
  TB_REQUIRE(flag >= 0 && flag <= 5, TB_ERR_ARG, "cloud flag must be 0..5");
  TB_REQUIRE(c->have_atm, TB_ERR_STATE, "tb_set_atmosphere must precede tb_set_cloud");
  c->cl_flag = flag; c->cl_ext = cloudext; c->cl_top = cloudtop; c->cl_bot = cloudbot;
  c->cl_gamma = gamma; c->cl_Q = Q; c->cl_r = r; c->cl_sig = sig; c->cl_refwn = refwn;
  c->nH.assign(c->nrad, 0.0);
  c->mean_dens.assign(c->nrad, 0.0);
  if (nH) c->nH.assign(nH, nH + c->nrad);
  if (mean_dens) c->mean_dens.assign(mean_dens, mean_dens + c->nrad);
  return TB_OK;
}

// ---- taps ------------------------------------------------------------------------------

extern "C" int tb_get_extinction(tb_ctx *c, double *e) {
  TB_REQUIRE(c->have_e, TB_ERR_STATE, "no extinction computed yet");
  TB_CUDA(cudaSetDevice(c->device));
  TB_CUDA(cudaMemcpyAsync(e, c->d_e.p, sizeof(double) * c->nrad * c->nwn, cudaMemcpyDeviceToHost,
                          c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  return TB_OK;
}

extern "C" int tb_get_cia(tb_ctx *c, double *e_cs) {
  TB_REQUIRE(c->have_ecs, TB_ERR_STATE, "tb_cia_interp not run");
  TB_CUDA(cudaSetDevice(c->device));
  std::vector<double> tmp((size_t)c->nrad * c->nwn);
  TB_CUDA(cudaMemcpyAsync(tmp.data(), c->d_ecs.p, sizeof(double) * tmp.size(),
                          cudaMemcpyDeviceToHost, c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  for (int r = 0; r < c->nrad; r++)       // device [rad][wn] -> reference [wn][rad]
    for (int64_t w = 0; w < c->nwn; w++) e_cs[w * c->nrad + r] = tmp[(size_t)r * c->nwn + w];
  return TB_OK;
}

extern "C" int tb_get_tau(tb_ctx *c, double *t, int64_t *last) {
  TB_REQUIRE(c->have_tau, TB_ERR_STATE, "tb_tau not run");
  TB_CUDA(cudaSetDevice(c->device));
  std::vector<double> tmp((size_t)c->nh * c->nwn);
  std::vector<int32_t> l32(c->nwn);
  TB_CUDA(cudaMemcpyAsync(tmp.data(), c->d_tau.p, sizeof(double) * tmp.size(),
                          cudaMemcpyDeviceToHost, c->stream));
  TB_CUDA(cudaMemcpyAsync(l32.data(), c->d_last.p, sizeof(int32_t) * c->nwn, cudaMemcpyDeviceToHost,
                          c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  if (t)
    for (int r = 0; r < c->nh; r++)       // device [ray][wn] -> reference [wn][ray]
      for (int64_t w = 0; w < c->nwn; w++) t[w * c->nh + r] = tmp[(size_t)r * c->nwn + w];
  if (last) for (int64_t w = 0; w < c->nwn; w++) last[w] = l32[w];
  return TB_OK;
}

extern "C" int tb_get_intensity(tb_ctx *c, double *intens) {
  TB_REQUIRE(c->nang > 0, TB_ERR_STATE, "tb_intensity_flux not run");
  TB_CUDA(cudaSetDevice(c->device));
  TB_CUDA(cudaMemcpyAsync(intens, c->d_intens.p, sizeof(double) * c->nang * c->nwn,
                          cudaMemcpyDeviceToHost, c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  return TB_OK;
}

// The reference evaluates layers lazily (tau.c:238-271): a layer is computed when some ray
// of some wavenumber dips below the deepest one computed so far.  We evaluate every layer;
// this tap reports which ones the reference would have, from tau.last[].
extern "C" int tb_get_computed(tb_ctx *c, int16_t *computed) {
  TB_REQUIRE(c->have_tau, TB_ERR_STATE, "tb_tau not run");
  TB_CUDA(cudaSetDevice(c->device));
  std::vector<int32_t> l32(c->nwn);
  TB_CUDA(cudaMemcpyAsync(l32.data(), c->d_last.p, sizeof(int32_t) * c->nwn, cudaMemcpyDeviceToHost,
                          c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  int deepest = 0;
  for (int64_t w = 0; w < c->nwn; w++) if (l32[w] > deepest) deepest = l32[w];
  int lastr = c->nrad - 1;
  for (int ri = 0; ri <= deepest; ri++)
    while (lastr > 0 && c->h_h[ri] * c->hfct < c->rad[lastr] * c->rad_fct) lastr--;
  for (int r = 0; r < c->nrad; r++) computed[r] = (r >= lastr) ? 1 : 0;
  return TB_OK;
}

// ---- micro-benchmarks for the roofline denominators -----------------------------------

__global__ void k_mb_dfma(double *out, int iters) {
  double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5,
         a6 = a0 + 6, a7 = a0 + 7;
  const double m = 0.9999999, b = 1e-7;
  for (int i = 0; i < iters; i++) {
    a0 = fma(a0, m, b); a1 = fma(a1, m, b); a2 = fma(a2, m, b); a3 = fma(a3, m, b);
    a4 = fma(a4, m, b); a5 = fma(a5, m, b); a6 = fma(a6, m, b); a7 = fma(a7, m, b);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

__global__ void k_mb_dexp(double *out, int iters) {
  double x = -1.0 - threadIdx.x * 1e-3, acc = 0;
  for (int i = 0; i < iters; i++) {
    acc += exp(x);
    x -= 1e-6;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

__global__ void k_mb_copy(const double4 *__restrict__ a, double4 *__restrict__ b, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    b[i] = a[i];
}

extern "C" int tb_microbench(tb_ctx *c, double *dfma_tflops, double *dexp_gops, double *copy_gbs) {
  TB_CUDA(cudaSetDevice(c->device));
  const int blocks = c->sm_count * 8, threads = 256;
  DevBuf<double> out;
  TB_TRY(out.ensure((size_t)blocks * threads));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float ms = 0, best;
  const int it_f = 20000, it_e = 2000;
  best = 1e30f;
  for (int rep = 0; rep < 4; rep++) {
    cudaEventRecord(e0, c->stream);
    k_mb_dfma<<<blocks, threads, 0, c->stream>>>(out.p, it_f);
    cudaEventRecord(e1, c->stream);
    TB_CUDA(cudaEventSynchronize(e1));
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep && ms < best) best = ms;
  }
  *dfma_tflops = 2.0 * 8 * it_f * (double)blocks * threads / (best * 1e-3) / 1e12;
  best = 1e30f;
  for (int rep = 0; rep < 4; rep++) {
    cudaEventRecord(e0, c->stream);
    k_mb_dexp<<<blocks, threads, 0, c->stream>>>(out.p, it_e);
    cudaEventRecord(e1, c->stream);
    TB_CUDA(cudaEventSynchronize(e1));
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep && ms < best) best = ms;
  }
  *dexp_gops = (double)it_e * blocks * threads / (best * 1e-3) / 1e9;
  size_t n4 = (size_t)1 << 26;  // 2 GiB per buffer
  DevBuf<double4> a, b;
  TB_TRY(a.ensure(n4));
  TB_TRY(b.ensure(n4));
  cudaMemsetAsync(a.p, 0, n4 * sizeof(double4), c->stream);
  best = 1e30f;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0, c->stream);
    k_mb_copy<<<c->sm_count * 16, 512, 0, c->stream>>>(a.p, b.p, n4);
    cudaEventRecord(e1, c->stream);
    TB_CUDA(cudaEventSynchronize(e1));
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep && ms < best) best = ms;
  }
  *copy_gbs = 2.0 * n4 * sizeof(double4) / (best * 1e-3) / 1e9;
  c->launches += 13;
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  out.release(); a.release(); b.release();
  return TB_OK;
}
