// tb_internal.cuh -- shared declarations of the sm_100a implementation (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <math.h>
#include <map>
#include <string>
#include <vector>
#include "../../include/transit_b200.h"

// ---- physical constants, transit/include/constants_tr.h:19-53 (cgs) -----------------
#define TB_PI 3.141592653589793
#define TB_AMU 1.66053886e-24
#define TB_EC 4.8032068e-10
#define TB_LS 2.99792458e10
#define TB_ME 9.1093897e-28
#define TB_KB 1.380658e-16
#define TB_H 6.6260755e-27
#define TB_SIGCTE (TB_PI * TB_EC * TB_EC / TB_LS / TB_LS / TB_ME / TB_AMU)
#define TB_EXPCTE (TB_H * TB_LS / TB_KB)
#define TB_NAVOGADRO 6.02214076e23
#define TB_SQRTLN2 0.83255461115769775635
#define TB_E0H2 4.911e-23
#define TB_MICRON 1e-4
#define TB_AMAGAT 2.68678e19
#define TB_DEGREES (TB_PI / 180.0)

void tb_set_error(const char *fmt, ...);

#define TB_CUDA(call)                                                                   \
  do {                                                                                  \
    cudaError_t e_ = (call);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      tb_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      return TB_ERR_CUDA;                                                               \
    }                                                                                   \
  } while (0)
#define TB_TRY(call)            \
  do {                          \
    int r_ = (call);            \
    if (r_ != TB_OK) return r_; \
  } while (0)
#define TB_REQUIRE(cond, code, ...) \
  do {                              \
    if (!(cond)) {                  \
      tb_set_error(__VA_ARGS__);    \
      return (code);                \
    }                               \
  } while (0)

// Device buffer that remembers its size.  It frees itself when its owner goes away (tb_destroy sets the
// device and deletes the context; locals leave scope), so no release list can drift from the struct.
template <typename T>
struct DevBuf {
  T *p = nullptr;
  size_t n = 0;
  DevBuf() = default;
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  ~DevBuf() { release(); }
  bool owned = true;           // false: p belongs to another context's buffer (borrow), never freed here
  void borrow(T *q, size_t m) {
    release();
    p = q; n = m; owned = false;
  }
  int ensure(size_t want) {
    if (want <= n && p && owned) return TB_OK;
    if (p && owned) cudaFree(p);
    owned = true;
    p = nullptr;
    n = 0;
    if (want == 0) return TB_OK;
    cudaError_t e = cudaMalloc((void **)&p, want * sizeof(T));
    if (e != cudaSuccess) {
      tb_set_error("cudaMalloc(%zu bytes) failed: %s", want * sizeof(T), cudaGetErrorString(e));
      return TB_ERR_NOMEM;
    }
    n = want;
    return TB_OK;
  }
  void release() {
    if (p && owned) cudaFree(p);
    p = nullptr;
    n = 0;
    owned = true;
  }
};

// One cell (idop, ilor) of the Voigt table as the line-sum kernels see it.  The profile is
// stored "polyphase": sample idx of the reference layout lives at off + (idx % o)*nk + idx/o,
// so the samples one line contributes to consecutive coarse bins (idx, idx+o, idx+2o, ...)
// are consecutive floats (one coalesced 128-byte read per warp instead of 32 sectors).
#define TB_PROF_PAD 128   // doubles of zero padding either side of the polyphase table
struct ProfEntry {
  int32_t off;      // offset of the profile in the polyphase float table
  int32_t psize;    // half size in fine samples
  int32_t nk;       // coarse samples per phase: 2*psize/o + 1
  int32_t pad;      // psize / o
};

// What k_strength leaves per (state, group) for k_accumulate: one 16-byte record, so that a
// visit is one 128-bit read (one sector per two records) instead of three scattered ones.
struct __align__(16) LineRec {
  double k;        // strength (0 for skipped groups)
  int32_t iown;    // fine bin of the group
  uint32_t pid;    // low 16 bits: profile cell, PID_SKIP, PID_NOREACH + cell, or PID_FB + isotope;
                   // high 16 bits: how many dead records follow this one inside its 256-group block
};

struct InterpLayer {   // per layer of an interpolated spectrum: lower grid-temperature index and the weights' ingredients
  int it;
  double t, tlo, thi, rdt;   // rdt = RN(1 / (thi - tlo))
};

struct CiaTable {
  int nw = 0, nt = 0, nspec = 0;
  int mol[2] = {0, 0};
  std::vector<double> wn, temp, tab;
  DevBuf<double> d_wn, d_temp, d_tab;
  DevBuf<double> d_zt, d_h, d_u, d_rh, d_r, d_ru, d_r6h, d_h6;   // table-only spline parts (see tb_radiative.cu)
  DevBuf<double> d_f2, d_z2, d_v;     // per-spectrum scratch, [row][layer]
  DevBuf<int32_t> d_widx;             // table interval of every grid wavenumber
  int64_t fi = 0, li = 0;             // grid range inside the table (no extrapolation)
  int grid_version = -1;
  void release() {
    d_wn.release(); d_temp.release(); d_tab.release(); d_zt.release(); d_h.release(); d_u.release(); d_rh.release(); d_r.release(); d_ru.release(); d_r6h.release(); d_h6.release();
    d_f2.release(); d_z2.release(); d_v.release(); d_widx.release();
  }
};

struct KernelTimer {
  cudaEvent_t a = nullptr, b = nullptr;
  bool pending = false;
  float ms = 0.f;
};

struct tb_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr, own_stream = nullptr;
  cudaStream_t aux_stream = nullptr;                 // CIA runs here, beside the line sum (tb_spectrum)
  // pinned host ring behind tb_upload_small: cudaMemcpyAsync from PAGEABLE memory first waits for the
  // stream's earlier work and then stages, so every small per-atmosphere upload used to be a hidden
  // stream synchronisation; from the ring it is a plain enqueue and the caller's buffer is free at once
  char *h_ring = nullptr;
  size_t ring_pos = 0;
  int tauw_shm_set = -1;
  int tau_shm_set = -1;                              // dynamic shared memory size k_tau_transit was last opted into
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  int64_t launches = 0;
  std::map<std::string, KernelTimer> timers;

  // ---- wavenumber grids (a2)
  bool have_grid = false;
  int grid_version = 0;
  double wn_i = 0, wn_d = 0, odwn = 0, own_last = 0, wn_fct = 1;
  int o = 1;
  int64_t nwn = 0, onwn = 0;

  // ---- species
  bool have_species = false;
  int nmol = 0, niso = 0, nrows = 0;  // nrows: line-list molecules (opacity.c:349-361)
  std::vector<double> mol_mass, mol_radius, mol_pol, iso_mass, iso_ratio;
  std::vector<int> mol_ID, iso_imol, iso_row, row_molID;
  DevBuf<double> d_mol_mass, d_mol_radius, d_mol_pol, d_iso_mass, d_iso_ratio;
  DevBuf<int> d_iso_imol, d_iso_row;

  // ---- line index (a1), see tb_lines.cu
  bool have_lines = false;
  int64_t nlines = 0, M = 0, G = 0;
  double wfct = 1e-4, efct = 1.0;
  // original order (kept for the trace tap)
  DevBuf<int32_t> l_iown, l_gpos;
  DevBuf<uint8_t> l_flag;  // bit0 in range, bit1 head
  // bin-sorted member lines
  DevBuf<double> m_wavn, m_elow, m_gf;
  DevBuf<int16_t> m_iso;
  DevBuf<uint8_t> m_inr;
  DevBuf<double> q_g, q_a, q_w;   // in-range member lines, strongest (g/m) first, for k_kmax
  DevBuf<int16_t> q_iso;
  int64_t Mq = 0;
  DevBuf<int32_t> x_line;      // member positions of the co-added (non-leading) lines
  // bin-sorted groups
  DevBuf<int32_t> g_start, g_iown, g_idwn, g_orig;
  DevBuf<int16_t> g_iso;
  DevBuf<double> g_wavn;       // leader wavenumber, sorted order
  DevBuf<double> g_a0, g_gf, g_lg;   // leader: -EXPCTE*efct*elow, gf, log(gf) (+inf for co-add groups)
  DevBuf<double> go_wavn;      // leader wavenumber, original group order
  DevBuf<int32_t> go_pos;      // sorted position of every original group
  DevBuf<int32_t> iso_og0, iso_og1;   // [niso] range of original groups of each isotope (one run each: og_contig)
  bool og_contig = true;
  DevBuf<int32_t> binstart;    // [nrows][nwn+1]
  DevBuf<int32_t> substart;    // [nrows][nsub]: finer start index for the narrow path (tb_lines.cu)
  int64_t nsub = 0;
  int nsube = 1;

  // ---- wavenumber-band shard (tb_set_shard): line sum only for coarse bins [band_lo, band_hi)
  bool banded = false;                // false = whole grid
  int64_t band_lo = 0, band_hi = 0;   // coarse bins [band_lo, band_hi); may be empty
  int shard_rank = 0, shard_world = 1;
  tb_exchange_fn exch = nullptr;
  void *exch_user = nullptr;
  std::vector<int32_t> h_binstart;    // host copy of binstart (group ranges of a band)
  std::vector<int64_t> h_xr_key, h_xr; // co-added member ranges of the band's groups per row (k_extra), keyed by (pa, pb)
  // in-library exchange of a band-sharded run (tb_comm.cu): NCCL communicator, band table, gather staging
  void *comm = nullptr;
  int comm_rank = 0, comm_world = 1;
  std::vector<int64_t> band_edges;    // [world + 1]
  std::vector<double> tile_weight;    // line groups per 128-bin tile (tb_rebalance_bands)
  DevBuf<double> g_send, g_recv;
  DevBuf<int64_t> g_edges;
  int g_edges_world = 0;

  // ---- Voigt table (a3/a4)
  bool have_prof = false;
  int nDop = 0, nLor = 0;
  float nwidth = 20.f;
  std::vector<double> aDop, aLor;
  double dop_lmin = 0, dop_rstep = 0;   // aDop[i] = 10^(dop_lmin + i/dop_rstep)
  std::vector<int64_t> psize, poff;
  int64_t ndata = 0;
  DevBuf<double> d_aDop, d_aLor;
  DevBuf<int32_t> d_psize;
  DevBuf<int64_t> d_poff;
  DevBuf<float> d_pdata;      // reference layout (test tap)
  DevBuf<double> d_pdata2;    // polyphase layout, widened to double (line sum)
  DevBuf<ProfEntry> d_ptab;   // [nDop*nLor]

  // ---- atmosphere
  bool have_atm = false;
  int nrad = 0;
  double rad_fct = 1, p_fct = 1, t_fct = 1;
  std::vector<double> rad, p, t, dens, Z;
  DevBuf<double> d_rad, d_p, d_t, d_dens;

  // ---- state batch scratch (tb_lines.cu)
  DevBuf<double> s_T, s_dens, s_Z, s_alphal, s_alphad, s_kmax;
  DevBuf<int32_t> s_ilor, s_idop0, s_lastq, s_fb, s_fbpid, s_wtile, s_reach, s_reachP, s_fbmax;
  DevBuf<double> s_facz, s_rtemp, s_ral, s_xterm, s_lcz;
  DevBuf<LineRec> s_rec;      // [state][group]
  DevBuf<int32_t> s_fclass, s_fP, s_lastq2;   // s_lastq2: carried-index source of the fused states (k_lastq_scan)
  bool kmax_block_skip = false;  // tb_upload_lines: no negative energies / gf, k_kmax may leave whole blocks
  int fuse_rmax = 8;
  int fuse_variant = 1;          // 0: k_linesum_fused, 1: k_linesum_fused_w
  int narrow_r = 2;              // NARROW_R of tb_lines.cu
  bool fuse_carried = false;     // fuse states whose lines partly keep a carried Doppler index (k_lastq_scan)   // per batch of 8 states: taken by k_linesum_fused? / widest half-size (fine bins)
  int fuse_enabled = 1;          // 0 off, 1 on where the grid has enough tiles, 2 always (tests)
  bool trace_enabled = false;
  int fuse_shm_set = -1, last_chunk_s0 = 0, last_chunk_cs = 0;
  DevBuf<uint8_t> s_hasfb;    // [state][tile]
  DevBuf<unsigned long long> s_nsamp;
  DevBuf<double> s_part;      // partial sums of wide-reach tiles shared by several blocks (small grids)
  DevBuf<unsigned> s_tcnt;    // ... and their arrival counters
  DevBuf<double> s_out;
  int last_ns = 0, last_permol = 0;
  bool nsamp_valid = false;

  // ---- results
  DevBuf<double> d_e, d_ecs, d_tau, d_intens, d_out;
  DevBuf<int32_t> d_last;
  DevBuf<int32_t> d_err;      // [8] device-side error flags
  bool defer_check = false, err_keep = false;   // tb_spectrum_batch: one error read-back per batch
  DevBuf<double> d_out_batch;
  bool have_e = false, have_ecs = false, have_tau = false;
  int nh = 0, geometry = 0, nang = 0;
  double toomuch = 0;
  std::vector<double> h_h;
  std::vector<double> user_h;    // tb_set_rays: impact parameters for tb_spectrum (empty: reversed radii)
  double user_hfct = 1;
  double hfct = 1;

  // ---- ray geometry scratch
  DevBuf<double> r_coef, r_par, r_h, r_s, r_ip, r_ang, r_rcp, r_wg, r_w0;
  DevBuf<uint8_t> r_ecl;               // per-ray constants of the streaming eclipse kernel (tb_radiative.cu)
  bool tau_stream_ecl = true;          // eclipse tau in O(layers) per wavenumber (false: the general kernel)
  DevBuf<int32_t> r_meta;

  // ---- opacity grid (a6-a8)
  bool have_grid_o = false;
  int64_t gNmol = 0, gNtemp = 0, gNlayer = 0, gNwave = 0;
  std::vector<double> gtemp, gpress;
  std::vector<int> gmolID, gmol_idx;
  DevBuf<double> d_grid, i_rho;
  DevBuf<uint8_t> i_stage;    // per-layer interpolation brackets (tb_opacity.cu)

  // ---- CIA
  std::vector<CiaTable *> cia;
  DevBuf<double> c_stage;     // [layer temperatures | density factor per table]

  // ---- scattering / clouds
  int modlevel = 1;                                  // modulation1 (1) or modulationm1 (-1), slantpath.c:487-503
  int sc_flag = 0;
  double sc_logext = 0;
  int cl_flag = 0;
  double cl_ext = 0, cl_top = 0, cl_bot = 0, cl_gamma = 0, cl_Q = 0, cl_r = 0, cl_sig = 0, cl_refwn = 0;
  std::vector<double> nH, mean_dens;
  DevBuf<double> d_es_aux;  // [nH | mean_dens]

  // ---- batched retrieval (tb_radiative.cu: tb_set_atmosphere_model / tb_run_transit_batch)
  int am_nrad = 0, am_bymass = 0, am_ih2 = -1, b_static_version = -1;
  double am_pfct = 1, am_tfct = 1, am_rfct = 1, am_p0 = 0, am_r0 = 0, am_gsurf = 0;
  std::vector<double> am_p;
  DevBuf<double> b_p, b_in, b_ltemp, b_traw, b_dens, b_rad, b_ip, b_rho, b_densfac, b_aux, b_e, b_ecs, b_tau, b_out, b_gtemp;
  DevBuf<uint8_t> b_L;
  DevBuf<int32_t> b_last;
  DevBuf<int> b_gmol;

  // CUB temp storage
  DevBuf<uint8_t> cub_tmp;
};

// timing helpers: record CUDA events around the dominant kernels on the context stream
void tb_timer_begin(tb_ctx *c, const char *name);
void tb_timer_end(tb_ctx *c, const char *name);
#define TB_LAUNCHED(c) ((c)->launches++)

// ---- device helpers shared by several translation units ------------------------------

// owns.v[k] exactly as makesample1 fills it (makesample.c:100-104): si + k*osd, as a
// separate multiply and add (no FMA contraction: bin assignment must be bit-exact).
__device__ __forceinline__ double tb_fine_wn(double wn_i, double odwn, long long k) {
  return __dadd_rn(wn_i, __dmul_rn((double)k, odwn));
}

// binsearchapprox (pu/src/iomisc.c:1088-1108): nearest element by bisection on [lo,hi]
__device__ __forceinline__ int tb_binsearchapprox(const double *__restrict__ a, double v, int lo,
                                                  int hi) {
  while (hi - lo > 1) {
    int mid = (hi + lo) / 2;
    if (a[mid] > v) hi = mid; else lo = mid;
  }
  if (hi - lo == 1) return (fabs(a[hi] - v) < fabs(a[lo] - v)) ? hi : lo;
  return lo;
}

// exp(x) for the Boltzmann and stimulated-emission factors (extinction.c:409-413,455-458):
// x <= 0 there.  Cody-Waite reduction x = k ln2 + r, |r| <= ln2/2, degree-13 Taylor polynomial
// (truncation < 5e-18 relative), scaling by an exponent-field add: < 1 ulp like the library
// exp() it replaces, in about a quarter of the instructions.  Outside [-690, 690] (results
// near the denormal range, or the bogus positive arguments of negative energies) it defers
// to the library.
__device__ __forceinline__ double tb_exp(double x) {
  if (!(x > -690.0 && x < 690.0)) return exp(x);
  const double kf = rint(x * 1.4426950408889634074);
  double r = fma(kf, -6.93147180369123816490e-01, x);       // ln2 high part
  r = fma(kf, -1.90821492927058770002e-10, r);              // ln2 low part
  double p = 1.6059043836821614599e-10;                     // 1/13!
  p = fma(p, r, 2.0876756987868098979e-9);
  p = fma(p, r, 2.5052108385441718775e-8);
  p = fma(p, r, 2.7557319223985890653e-7);
  p = fma(p, r, 2.7557319223985892511e-6);
  p = fma(p, r, 2.4801587301587301566e-5);
  p = fma(p, r, 1.9841269841269841253e-4);
  p = fma(p, r, 1.3888888888888889419e-3);
  p = fma(p, r, 8.3333333333333332177e-3);
  p = fma(p, r, 4.1666666666666664354e-2);
  p = fma(p, r, 1.6666666666666665741e-1);
  p = fma(p, r, 0.5);
  p = fma(p, r, 1.0);
  p = fma(p, r, 1.0);
  const int k = (int)kf;
  return p * __hiloint2double((k + 1023) << 20, 0);
}

// RN(x / y) given r = RN(1 / y): quotient estimate, exact remainder by FMA, one correction (Markstein):
// the same correctly rounded quotient as the division, without recomputing the reciprocal of a
// constant divisor for every element (an FP64 division is ~40 dependent instructions)
__device__ __forceinline__ double tb_div_r(double x, double y, double r) {
  const double q = x * r;
  return fma(fma(-q, y, x), r, q);
}

// internal entry points between translation units
int tb_states_run(tb_ctx *c, int ns, const double *T, const double *dens, const double *Z,
                  double ethresh, int permol, double *d_out_rows, double *h_out);
int tb_cia_run(tb_ctx *c);
int tb_check_errors(tb_ctx *c);
#define TB_RING_BYTES (size_t(4) << 20)
int tb_upload_small(tb_ctx *c, void *dst, const void *src, size_t bytes, cudaStream_t st);
int tb_interp_run(tb_ctx *c);
int tb_interp_launch(tb_ctx *c, int nvl, int ngrid, const InterpLayer *dL, const double *drho, double *e);
int tb_comm_allreduce_max(tb_ctx *c, int what, void *device_ptr, int64_t count);
int tb_comm_gather_spectrum(tb_ctx *c);
extern "C" int tb_comm_destroy(tb_ctx *c);
// columns [w0, w1) the per-wavenumber stages (CIA evaluation, tau, modulation, intensities) compute: the band of a
// sharded run that gathers its spectrum in the library, else the whole grid
static inline void tb_band_columns(const tb_ctx *c, int64_t *w0, int64_t *w1) {
  const bool cut = c->banded && c->comm && c->shard_world > 1 && !c->band_edges.empty();
  *w0 = cut ? c->band_lo : 0;
  *w1 = cut ? c->band_hi : c->nwn;
}
