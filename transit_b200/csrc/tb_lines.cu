// tb_lines.cu -- the extinction line sum (computemolext, transit/src/extinction.c:281-529)
// re-designed for the GPU.
//
// The reference walks the line list once per layer, sequentially, carrying state from line
// to line (co-added runs, the per-isotope Doppler index).  Here the layer-INDEPENDENT part of
// that walk is done once, on the device, when the lines are uploaded:
//   * fine/coarse bin of every line (extinction.c:445-447,476), bit-exact (no FMA contraction);
//   * the greedy co-add runs (extinction.c:449-462) as a forward-pointer list whose heads are
//     found by pointer doubling (deterministic, O(n log n), no sequential pass);
//   * groups sorted by (molecule row, coarse bin) so that every output bin's contributors are
//     one contiguous window of the group list.
// Per batch of layer states (T, densities, partition functions) four kernels run:
//   k_state_setup -> widths and width indices per isotope (extinction.c:364-395)
//   k_kmax        -> strength maximum over all in-range lines (extinction.c:399-427)
//   k_strength    -> group strengths, threshold, Doppler index (extinction.c:439-483)
//   k_accumulate  -> Voigt-profile gather per output bin, fixed summation order, no atomics
//                    on floating-point data (extinction.c:486-509)
#include <cub/cub.cuh>
#include <stdlib.h>

#define NSUB 8             // sub-cells per coarse bin in the narrow path's start index
#include "tb_internal.cuh"


// ------------------------------------------------------------------ upload-time index

// wavn, in-range flag, nearest fine bin (extinction.c:431,436,445-447)
__global__ void k_line_prep(const double *__restrict__ wl, int64_t n, double wfct, double wn_i,
                            double odwn, double own_last, double *__restrict__ wavn,
                            int32_t *__restrict__ iown, uint8_t *__restrict__ flag) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  double w = __ddiv_rn(1.0, __dmul_rn(wl[i], wfct));
  wavn[i] = w;
  bool inr = !((w < wn_i) || (w > own_last));
  int32_t io = 0;
  if (inr) {
    io = (int32_t)__ddiv_rn(__dsub_rn(w, wn_i), odwn);
    if (fabs(__dsub_rn(w, tb_fine_wn(wn_i, odwn, io + 1))) <
        fabs(__dsub_rn(w, tb_fine_wn(wn_i, odwn, io))))
      io++;
  }
  iown[i] = io;
  flag[i] = inr ? 1 : 0;
}

// next[i]: the line the reference's for-loop visits after i (extinction.c:430,449-462)
__global__ void k_line_next(const double *__restrict__ wavn, const int16_t *__restrict__ iso,
                            const int32_t *__restrict__ iown, const uint8_t *__restrict__ flag,
                            int64_t n, double wn_i, double odwn, int32_t *__restrict__ next) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int64_t j = i + 1;
  if (flag[i] & 1) {
    const double center = tb_fine_wn(wn_i, odwn, iown[i]);
    const int16_t is = iso[i];
    while (j < n && iso[j] == is && fabs(__dsub_rn(wavn[j], center)) < odwn) j++;
  }
  next[i] = (int32_t)j;
}

// one pointer-doubling round: heads reached so far mark the node 2^k visits ahead
__global__ void k_mark_round(const int32_t *__restrict__ jump, int32_t *__restrict__ jump_out,
                             uint8_t *__restrict__ flag, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int32_t j = jump[i];
  if (j < n) {
    if (flag[i] & 2) flag[j] |= 2;   // byte store of an idempotent bit: benign race
    jump_out[i] = jump[j];
  } else
    jump_out[i] = j;
}

__global__ void k_iota(int32_t *__restrict__ a, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) a[i] = (int32_t)i;
}

__global__ void k_leader_flags(const uint8_t *__restrict__ flag, int64_t n, uint8_t *__restrict__ isl) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) isl[i] = (flag[i] == 3) ? 1 : 0;
}

// per group (original order): sort key and size; also the sortedness check
__global__ void k_group_keys(const int32_t *__restrict__ lead, int64_t G, const double *__restrict__ wavn,
                             const int16_t *__restrict__ iso, const int32_t *__restrict__ next,
                             const int *__restrict__ iso_row, const int32_t *__restrict__ iown,
                             double wn_i, double dwn, int64_t onwn,
                             uint64_t *__restrict__ key, int32_t *__restrict__ idwn_out,
                             int32_t *__restrict__ size, double *__restrict__ go_wavn,
                             int32_t *__restrict__ err) {
  int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (g >= G) return;
  int32_t L = lead[g];
  double w = wavn[L];
  idwn_out[g] = (int32_t)__ddiv_rn(__dsub_rn(w, wn_i), dwn);    // extinction.c:476
  key[g] = (uint64_t)iso_row[iso[L]] * (uint64_t)onwn + (uint64_t)iown[L];
  size[g] = next[L] - L;
  go_wavn[g] = w;
  if (g + 1 < G) {
    int32_t L2 = lead[g + 1];
    if (iso[L2] == iso[L] && wavn[L2] > w) atomicExch(err, 1);
  }
}

// scatter groups and their member lines into bin-sorted order
__global__ void k_group_fill(const int32_t *__restrict__ order, const int32_t *__restrict__ idwn_orig,
                             const int32_t *__restrict__ gstart, int64_t G,
                             const int32_t *__restrict__ lead, const int32_t *__restrict__ next,
                             const double *__restrict__ wavn, const double *__restrict__ elow,
                             const double *__restrict__ gf, const int16_t *__restrict__ iso,
                             const int32_t *__restrict__ iown, const uint8_t *__restrict__ flag,
                             int64_t nwn, int32_t *__restrict__ g_iown, int32_t *__restrict__ g_idwn,
                             int32_t *__restrict__ g_orig, int16_t *__restrict__ g_iso,
                             double *__restrict__ g_wavn, double *__restrict__ m_wavn,
                             double *__restrict__ m_elow, double *__restrict__ m_gf,
                             int16_t *__restrict__ m_iso, uint8_t *__restrict__ m_inr,
                             int32_t *__restrict__ l_gpos, int32_t *__restrict__ x_line, double efct,
                             double *__restrict__ g_a0, double *__restrict__ g_gf, double *__restrict__ g_lg) {
  int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= G) return;
  int32_t g = order[p], L = lead[g], E = next[L], s = gstart[p];
  g_a0[p] = __dmul_rn(__dmul_rn(-TB_EXPCTE, efct), elow[L]);
  g_gf[p] = gf[L];
  // log of the leader's gf: lets k_strength bound a lone line's strength without exp();
  // co-add groups (and anything not a positive finite gf) always take the exact path
  g_lg[p] = (E - L == 1 && gf[L] > 0) ? log(gf[L]) : (gf[L] == 0 && E - L == 1 ? -INFINITY : INFINITY);
  g_iown[p] = iown[L];
  g_idwn[p] = idwn_orig[g];
  g_orig[p] = g;
  g_iso[p] = iso[L];
  g_wavn[p] = wavn[L];
  l_gpos[L] = (int32_t)p;
  for (int32_t l = L; l < E; l++, s++) {
    m_wavn[s] = wavn[l];
    m_elow[s] = elow[l];
    m_gf[s] = gf[l];
    m_iso[s] = iso[l];
    m_inr[s] = flag[l] & 1;
    if (l != L) x_line[s - p - 1] = s;     // co-added member: the (s-p-1)-th non-leading member
  }
}

// position of every group (original, i.e. file order) in the sorted order, and the contiguous range of original
// groups each isotope occupies (TLI order: by isotope, then ascending wavelength) -- what k_lastq_scan walks
__global__ void k_group_pos(int64_t G, const int32_t *__restrict__ g_orig, int32_t *__restrict__ go_pos) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p < G) go_pos[g_orig[p]] = (int32_t)p;
}
__global__ void k_iso_ranges(int64_t G, const int32_t *__restrict__ go_pos, const int16_t *__restrict__ g_iso,
                             int32_t *__restrict__ og0, int32_t *__restrict__ og1, int32_t *__restrict__ nruns) {
  const int64_t og = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (og >= G) return;
  const int iso = g_iso[go_pos[og]];
  if (og == 0 || g_iso[go_pos[og - 1]] != iso) { og0[iso] = (int32_t)og; atomicAdd(&nruns[iso], 1); }
  if (og == G - 1 || g_iso[go_pos[og + 1]] != iso) og1[iso] = (int32_t)og + 1;
}

// the strength-maximum pass wants the lines with the largest state-independent factor first
__global__ void k_q_keys(int64_t M, const double *__restrict__ m_gf, const int16_t *__restrict__ m_iso,
                         const uint8_t *__restrict__ m_inr, const double *__restrict__ iso_mass,
                         const double *__restrict__ iso_ratio, double *__restrict__ key,
                         int32_t *__restrict__ idx) {
  int64_t l = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (l >= M) return;
  const int iso = m_iso[l];
  key[l] = m_inr[l] ? __dmul_rn(__dmul_rn(iso_ratio[iso], TB_SIGCTE), m_gf[l]) / iso_mass[iso] : -1.0;
  idx[l] = (int32_t)l;
}
__global__ void k_q_fill(int64_t Mq, const int32_t *__restrict__ idx, const double *__restrict__ m_wavn,
                         const double *__restrict__ m_elow, const double *__restrict__ m_gf,
                         const int16_t *__restrict__ m_iso, const double *__restrict__ iso_ratio, double efct,
                         double *__restrict__ q_g, double *__restrict__ q_a, double *__restrict__ q_w,
                         int16_t *__restrict__ q_iso) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= Mq) return;
  const int l = idx[i], iso = m_iso[l];
  q_g[i] = __dmul_rn(__dmul_rn(iso_ratio[iso], TB_SIGCTE), m_gf[l]);          // extinction.c:419-421
  q_a[i] = __dmul_rn(__dmul_rn(-TB_EXPCTE, efct), m_elow[l]);
  q_w[i] = __dmul_rn(-TB_EXPCTE, m_wavn[l]);
  q_iso[i] = (int16_t)iso;
}
__global__ void k_count_pos(int64_t M, const double *__restrict__ key_sorted_desc, int32_t *__restrict__ n) {
  int64_t l = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (l >= M) return;
  // keys are sorted descending: the boundary between in-range (>= 0) and out-of-range (-1) lines
  if (key_sorted_desc[l] >= 0 && (l + 1 == M || key_sorted_desc[l + 1] < 0)) *n = (int32_t)(l + 1);
}

// binstart[row][j] = first sorted group of `row` whose fine bin is >= o*j (groups are sorted by
// (row, fine bin), so the groups centred in coarse cell j are binstart[row][j] .. [j+1])
__global__ void k_binstart(const uint64_t *__restrict__ key_sorted, int64_t G, int64_t nkeys, int64_t nwn,
                           int64_t onwn, int o, int32_t *__restrict__ binstart) {
  int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k > nkeys) return;
  const uint64_t want = (uint64_t)(k / nwn) * (uint64_t)onwn + (uint64_t)(k % nwn) * (uint64_t)o;
  int64_t lo = 0, hi = G;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (key_sorted[mid] < want) lo = mid + 1; else hi = mid;
  }
  binstart[k] = (int32_t)lo;
}

// substart[row][u] = first sorted group of `row` whose fine bin is >= ceil(u*o/nsube): the narrow
// path's start index (nsube sub-cells per coarse bin), so that a bin finds the first record of its
// window with one read instead of a bisection of dependent reads
__global__ void k_substart(const uint64_t *__restrict__ key_sorted, int64_t G, int nrows, int64_t nsub,
                           int64_t onwn, int o, int nsube, int32_t *__restrict__ substart) {
  int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (k >= (int64_t)nrows * nsub) return;
  const int64_t row = k / nsub, u = k % nsub;
  int64_t f = (u * o + nsube - 1) / nsube;
  if (f > onwn) f = onwn;
  const uint64_t want = (uint64_t)row * (uint64_t)onwn + (uint64_t)f;
  int64_t lo = 0, hi = G;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (key_sorted[mid] < want) lo = mid + 1; else hi = mid;
  }
  substart[k] = (int32_t)lo;
}

static inline int nblk(int64_t n, int t) { return (int)((n + t - 1) / t); }

extern "C" int tb_upload_lines(tb_ctx *c, const double *wl, const double *elow, const double *gf,
                               const int16_t *isoid, int64_t n, double wfct, double efct) {
  TB_REQUIRE(c->have_grid && c->have_species, TB_ERR_STATE,
             "tb_set_wn_grid and tb_set_species must precede tb_upload_lines");
  TB_REQUIRE(n >= 0 && n < 2000000000LL, TB_ERR_ARG, "line count out of range");
  TB_REQUIRE((int64_t)c->nrows * c->nwn < 2000000000LL, TB_ERR_ARG, "nrows*nwn too large");
  TB_CUDA(cudaSetDevice(c->device));
  {  // per isotope the wavelengths must not decrease over the WHOLE list (not just between neighbours:
     // an isotope split into two runs would get a wrong carried Doppler index, extinction.c:480-483)
    std::vector<double> lastwl(c->niso > 0 ? c->niso : 1, -INFINITY);
    for (int64_t i = 0; i < n; i++) {
      TB_REQUIRE(isoid[i] >= 0 && isoid[i] < c->niso, TB_ERR_ARG, "isoid[%lld]=%d out of range",
                 (long long)i, (int)isoid[i]);
      TB_REQUIRE(!(wl[i] < lastwl[isoid[i]]), TB_ERR_UNSORTED,
                 "line list is not sorted by ascending wavelength within isotope %d (line %lld; "
                 "pylineread.py:369-383 guarantees it; the Doppler-index carry relies on it)",
                 (int)isoid[i], (long long)i);
      lastwl[isoid[i]] = wl[i];
    }
  }
  {  // k_kmax leaves whole blocks of its strength-ordered list when their strongest line cannot raise a maximum: that
     // bound needs both exponential factors <= 1, i.e. no negative lower-state energy (and no negative gf or factor)
    bool plain = efct > 0;
    for (int64_t i = 0; i < n && plain; i++) plain = elow[i] >= 0 && gf[i] >= 0 && wl[i] > 0;
    c->kmax_block_skip = plain;
  }
  cudaStream_t st = c->stream;
  c->nlines = n; c->wfct = wfct; c->efct = efct;
  c->have_lines = false;
  c->h_binstart.clear();
  c->h_xr_key.clear();
  c->G = c->M = c->Mq = 0;
  const int64_t nkeys = (int64_t)c->nrows * c->nwn;
  TB_TRY(c->binstart.ensure(nkeys + 1));
  if (n == 0) {
    TB_CUDA(cudaMemsetAsync(c->binstart.p, 0, (nkeys + 1) * sizeof(int32_t), st));
    c->have_lines = true;
    return TB_OK;
  }
  DevBuf<double> d_wl, d_wavn, d_elow, d_gf;
  DevBuf<int16_t> d_iso;
  DevBuf<int32_t> nxt, jA, jB, lead, idw, ord, ord_s, size, size_s, ngrp;
  DevBuf<uint64_t> key, key_s;
  DevBuf<uint8_t> isl;
  TB_TRY(d_wl.ensure(n)); TB_TRY(d_wavn.ensure(n)); TB_TRY(d_elow.ensure(n)); TB_TRY(d_gf.ensure(n));
  TB_TRY(d_iso.ensure(n)); TB_TRY(nxt.ensure(n)); TB_TRY(jA.ensure(n)); TB_TRY(jB.ensure(n));
  TB_TRY(isl.ensure(n)); TB_TRY(lead.ensure(n)); TB_TRY(ngrp.ensure(1));
  TB_TRY(c->l_iown.ensure(n)); TB_TRY(c->l_gpos.ensure(n)); TB_TRY(c->l_flag.ensure(n));
  TB_CUDA(cudaMemcpyAsync(d_wl.p, wl, n * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(d_elow.p, elow, n * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(d_gf.p, gf, n * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(d_iso.p, isoid, n * sizeof(int16_t), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemsetAsync(c->l_gpos.p, 0xff, n * sizeof(int32_t), st));
  TB_CUDA(cudaMemsetAsync(c->d_err.p, 0, sizeof(int32_t), st));

  const int T = 256;
  k_line_prep<<<nblk(n, T), T, 0, st>>>(d_wl.p, n, wfct, c->wn_i, c->odwn, c->own_last, d_wavn.p,
                                        c->l_iown.p, c->l_flag.p);
  k_line_next<<<nblk(n, T), T, 0, st>>>(d_wavn.p, d_iso.p, c->l_iown.p, c->l_flag.p, n, c->wn_i,
                                        c->odwn, nxt.p);
  // heads of the visit list: root = line 0, then pointer doubling
  uint8_t one = 0;
  TB_CUDA(cudaMemcpyAsync(&one, c->l_flag.p, 1, cudaMemcpyDeviceToHost, st));
  TB_CUDA(cudaStreamSynchronize(st));
  one |= 2;
  TB_CUDA(cudaMemcpyAsync(c->l_flag.p, &one, 1, cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(jA.p, nxt.p, n * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  int rounds = 1;
  while ((1LL << rounds) < n + 1) rounds++;
  int32_t *ja = jA.p, *jb = jB.p;
  for (int r = 0; r <= rounds; r++) {
    k_mark_round<<<nblk(n, T), T, 0, st>>>(ja, jb, c->l_flag.p, n);
    int32_t *t = ja; ja = jb; jb = t;
  }
  c->launches += 3 + rounds;
  // leaders = heads that are in range, in original order
  k_leader_flags<<<nblk(n, T), T, 0, st>>>(c->l_flag.p, n, isl.p);
  size_t tmpb = 0;
  cub::CountingInputIterator<int32_t> cnt(0);
  cub::DeviceSelect::Flagged(nullptr, tmpb, cnt, isl.p, lead.p, ngrp.p, (int)n, st);
  TB_TRY(c->cub_tmp.ensure(tmpb));
  cub::DeviceSelect::Flagged(c->cub_tmp.p, tmpb, cnt, isl.p, lead.p, ngrp.p, (int)n, st);
  int32_t G32 = 0;
  TB_CUDA(cudaMemcpyAsync(&G32, ngrp.p, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  TB_CUDA(cudaStreamSynchronize(st));
  const int64_t G = G32;
  c->G = G;
  if (G == 0) {
    TB_CUDA(cudaMemsetAsync(c->binstart.p, 0, (nkeys + 1) * sizeof(int32_t), st));
    c->have_lines = true;
    c->og_contig = true;
    return TB_OK;
  }
  TB_TRY(key.ensure(G)); TB_TRY(key_s.ensure(G)); TB_TRY(ord.ensure(G)); TB_TRY(ord_s.ensure(G)); TB_TRY(idw.ensure(G));
  TB_TRY(size.ensure(G)); TB_TRY(size_s.ensure(G)); TB_TRY(c->go_wavn.ensure(G));
  k_group_keys<<<nblk(G, T), T, 0, st>>>(lead.p, G, d_wavn.p, d_iso.p, nxt.p, c->d_iso_row.p, c->l_iown.p,
                                         c->wn_i, c->wn_d, c->onwn, key.p, idw.p, size.p, c->go_wavn.p,
                                         c->d_err.p);
  // stable sort of (key, original group index): equal keys keep the reference's line order
  k_iota<<<nblk(G, T), T, 0, st>>>(ord.p, G);
  int end_bit = 1;
  while (end_bit < 63 && (1ULL << end_bit) < (uint64_t)c->nrows * (uint64_t)c->onwn + 1) end_bit++;
  tmpb = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmpb, key.p, key_s.p, ord.p, ord_s.p, (int)G, 0, end_bit, st);
  TB_TRY(c->cub_tmp.ensure(tmpb));
  cub::DeviceRadixSort::SortPairs(c->cub_tmp.p, tmpb, key.p, key_s.p, ord.p, ord_s.p, (int)G, 0, end_bit, st);
  // sizes in sorted order -> member offsets
  // (gather through a tiny transform iterator replacement: reuse SortPairs with sizes as values)
  tmpb = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmpb, key.p, key_s.p, size.p, size_s.p, (int)G, 0, end_bit, st);
  TB_TRY(c->cub_tmp.ensure(tmpb));
  cub::DeviceRadixSort::SortPairs(c->cub_tmp.p, tmpb, key.p, key_s.p, size.p, size_s.p, (int)G, 0, end_bit, st);
  TB_TRY(c->g_start.ensure(G + 1));
  tmpb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmpb, size_s.p, c->g_start.p, (int)G, st);
  TB_TRY(c->cub_tmp.ensure(tmpb));
  cub::DeviceScan::ExclusiveSum(c->cub_tmp.p, tmpb, size_s.p, c->g_start.p, (int)G, st);
  int32_t last_start = 0, last_size = 0, err = 0;
  TB_CUDA(cudaMemcpyAsync(&last_start, c->g_start.p + (G - 1), 4, cudaMemcpyDeviceToHost, st));
  TB_CUDA(cudaMemcpyAsync(&last_size, size_s.p + (G - 1), 4, cudaMemcpyDeviceToHost, st));
  TB_CUDA(cudaMemcpyAsync(&err, c->d_err.p, 4, cudaMemcpyDeviceToHost, st));
  TB_CUDA(cudaStreamSynchronize(st));
  TB_REQUIRE(err == 0, TB_ERR_UNSORTED,
             "line list is not sorted by ascending wavelength within an isotope "
             "(pylineread.py:369-383 guarantees it; the Doppler-index carry relies on it)");
  const int64_t M = (int64_t)last_start + last_size;
  c->M = M;
  TB_CUDA(cudaMemcpyAsync(c->g_start.p + G, &M, 4, cudaMemcpyHostToDevice, st));
  TB_TRY(c->m_wavn.ensure(M)); TB_TRY(c->m_elow.ensure(M)); TB_TRY(c->m_gf.ensure(M));
  TB_TRY(c->m_iso.ensure(M)); TB_TRY(c->m_inr.ensure(M)); TB_TRY(c->x_line.ensure(M - G + 1));
  TB_TRY(c->g_iown.ensure(G)); TB_TRY(c->g_idwn.ensure(G)); TB_TRY(c->g_orig.ensure(G));
  TB_TRY(c->g_iso.ensure(G)); TB_TRY(c->g_wavn.ensure(G));
  TB_TRY(c->g_a0.ensure(G)); TB_TRY(c->g_gf.ensure(G)); TB_TRY(c->g_lg.ensure(G));
  k_group_fill<<<nblk(G, T), T, 0, st>>>(ord_s.p, idw.p, c->g_start.p, G, lead.p, nxt.p, d_wavn.p,
                                         d_elow.p, d_gf.p, d_iso.p, c->l_iown.p, c->l_flag.p, c->nwn,
                                         c->g_iown.p, c->g_idwn.p, c->g_orig.p, c->g_iso.p,
                                         c->g_wavn.p, c->m_wavn.p, c->m_elow.p, c->m_gf.p, c->m_iso.p,
                                         c->m_inr.p, c->l_gpos.p, c->x_line.p, efct, c->g_a0.p, c->g_gf.p,
                                         c->g_lg.p);
  k_binstart<<<nblk(nkeys + 1, T), T, 0, st>>>(key_s.p, G, nkeys, c->nwn, c->onwn, c->o, c->binstart.p);
  c->nsube = c->o < NSUB ? c->o : NSUB;
  c->nsub = ((c->onwn - 1) * c->nsube) / c->o + 2;
  TB_REQUIRE(c->onwn * c->nsube < 4000000000LL, TB_ERR_ARG, "fine grid too large for the 32-bit start index");
  TB_TRY(c->substart.ensure((size_t)c->nrows * c->nsub));
  k_substart<<<nblk((int64_t)c->nrows * c->nsub, T), T, 0, st>>>(key_s.p, G, c->nrows, c->nsub, c->onwn, c->o,
                                                               c->nsube, c->substart.p);
  {  // original (file) order -> sorted position, isotope ranges in file order (k_lastq_scan)
    TB_TRY(c->go_pos.ensure(G)); TB_TRY(c->iso_og0.ensure(c->niso + 1)); TB_TRY(c->iso_og1.ensure(c->niso + 1));
    DevBuf<int32_t> nruns;
    TB_TRY(nruns.ensure(c->niso + 1));
    TB_CUDA(cudaMemsetAsync(c->iso_og0.p, 0, (c->niso + 1) * sizeof(int32_t), st));
    TB_CUDA(cudaMemsetAsync(c->iso_og1.p, 0, (c->niso + 1) * sizeof(int32_t), st));
    TB_CUDA(cudaMemsetAsync(nruns.p, 0, (c->niso + 1) * sizeof(int32_t), st));
    k_group_pos<<<nblk(G, T), T, 0, st>>>(G, c->g_orig.p, c->go_pos.p);
    k_iso_ranges<<<nblk(G, T), T, 0, st>>>(G, c->go_pos.p, c->g_iso.p, c->iso_og0.p, c->iso_og1.p, nruns.p);
    std::vector<int32_t> h(c->niso + 1, 0);
    TB_CUDA(cudaMemcpyAsync(h.data(), nruns.p, (c->niso + 1) * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    TB_CUDA(cudaStreamSynchronize(st));
    c->og_contig = true;                 // every isotope's lines form ONE run of the list (the TLI order)
    for (int i = 0; i < c->niso; i++) if (h[i] > 1) c->og_contig = false;
  }
  c->launches += 11;
  {  // strength-ordered copy of the in-range member lines for k_kmax
    DevBuf<double> qk, qk_s;
    DevBuf<int32_t> qi, qi_s;
    TB_TRY(qk.ensure(M)); TB_TRY(qk_s.ensure(M)); TB_TRY(qi.ensure(M)); TB_TRY(qi_s.ensure(M));
    k_q_keys<<<nblk(M, T), T, 0, st>>>(M, c->m_gf.p, c->m_iso.p, c->m_inr.p, c->d_iso_mass.p, c->d_iso_ratio.p,
                                       qk.p, qi.p);
    tmpb = 0;
    cub::DeviceRadixSort::SortPairsDescending(nullptr, tmpb, qk.p, qk_s.p, qi.p, qi_s.p, (int)M, 0, 64, st);
    TB_TRY(c->cub_tmp.ensure(tmpb));
    cub::DeviceRadixSort::SortPairsDescending(c->cub_tmp.p, tmpb, qk.p, qk_s.p, qi.p, qi_s.p, (int)M, 0, 64, st);
    TB_CUDA(cudaMemsetAsync(ngrp.p, 0, sizeof(int32_t), st));
    k_count_pos<<<nblk(M, T), T, 0, st>>>(M, qk_s.p, ngrp.p);
    int32_t Mq = 0;
    TB_CUDA(cudaMemcpyAsync(&Mq, ngrp.p, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    TB_CUDA(cudaStreamSynchronize(st));
    c->Mq = Mq;
    TB_TRY(c->q_g.ensure(Mq + 1)); TB_TRY(c->q_a.ensure(Mq + 1)); TB_TRY(c->q_w.ensure(Mq + 1));
    TB_TRY(c->q_iso.ensure(Mq + 1));
    if (Mq > 0)
      k_q_fill<<<nblk(Mq, T), T, 0, st>>>(Mq, qi_s.p, c->m_wavn.p, c->m_elow.p, c->m_gf.p, c->m_iso.p,
                                          c->d_iso_ratio.p, efct, c->q_g.p, c->q_a.p, c->q_w.p, c->q_iso.p);
    c->launches += 5;
    TB_CUDA(cudaStreamSynchronize(st));
    qk.release(); qk_s.release(); qi.release(); qi_s.release();
  }
  TB_CUDA(cudaStreamSynchronize(st));
  d_wl.release(); d_wavn.release(); d_elow.release(); d_gf.release(); d_iso.release();
  nxt.release(); jA.release(); jB.release(); lead.release(); key.release(); key_s.release();
  ord.release(); ord_s.release(); idw.release(); size.release(); size_s.release(); ngrp.release(); isl.release();
  c->have_lines = true;
  return TB_OK;
}

// ------------------------------------------------------------------ per-state kernels

#define PID_SKIP 0x7FFFu   // group below the strength threshold (extinction.c:467-470)
#define PID_FB   0x8000u   // + isotope: "use the isotope's carried Doppler index"
#define PID_NOREACH 0x4000u   // + cell: evaluated, but the profile covers no grid point (nothing to add)
#define PID_DEAD(pid) (((pid) & 0xC000u) == 0x4000u)   // PID_SKIP or PID_NOREACH: the line sum passes over it
#define ACC_TILE 128       // output bins per block of k_accumulate (4 warps x 32 bins)
#define FB 8               // states per batch: k_strength<8>'s block, the unit the fused / two-kernel choice is made for
#define MAXISO_SH 128      // isotopes handled with shared-memory maxima in k_strength

// exact x / d for 0 <= x < 2^32 by one 64x64->high multiply (d fixed per context)
struct FastDiv {
  uint64_t M;
  int d;
};
__device__ __forceinline__ int fd_div(int x, const FastDiv f) {
  return f.d == 1 ? x : (int)__umul64hi((uint64_t)(unsigned)x, f.M);
}

// extinction.c:364-395: Lorentz/Doppler widths and their table indices per (state, isotope)
__global__ void k_state_setup(int ns, int niso, int nmol, const double *__restrict__ T,
                              const double *__restrict__ dens, const double *__restrict__ mol_mass,
                              const double *__restrict__ mol_radius, const double *__restrict__ iso_mass,
                              const int *__restrict__ iso_imol, const double *__restrict__ aDop, int nDop,
                              const double *__restrict__ aLor, int nLor, double wn0,
                              double *__restrict__ alphal, double *__restrict__ alphad,
                              int32_t *__restrict__ ilor, int32_t *__restrict__ idop0,
                              int32_t *__restrict__ lastq, double *__restrict__ kmax, int nrows,
                              unsigned long long *__restrict__ nsamp, int32_t *__restrict__ wtile,
                              uint8_t *__restrict__ hasfb, int ntiles, const double *__restrict__ Z,
                              const double *__restrict__ iso_ratio, double *__restrict__ facz,
                              double *__restrict__ rtemp, double *__restrict__ ral) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < ns * nrows) kmax[t] = 0.0;
  if (t < ns) { nsamp[t] = 0ULL; rtemp[t] = 1.0 / T[t]; }   // nsamp already points at this chunk
  for (int i = t; i < ns * ntiles; i += gridDim.x * blockDim.x) { wtile[i] = -1; hasfb[i] = 0; }
  if (t >= ns * niso) return;
  int s = t / niso, i = t % niso;
  double temp = T[s];
  double fdoppler = sqrt(2 * TB_KB * temp / TB_AMU) * TB_SQRTLN2 / TB_LS;
  double florentz = sqrt(2 * TB_KB * temp / TB_PI / TB_AMU) / (TB_AMU * TB_LS);
  double al = 0.0;
  for (int j = 0; j < nmol; j++) {
    double cs = mol_radius[j] + mol_radius[iso_imol[i]];
    // separate multiply/add as the reference's x86-64 build has no FMA
    double term = __dmul_rn(__dmul_rn(__dmul_rn(dens[s * nmol + j] / mol_mass[j], cs), cs),
                            sqrt(__dadd_rn(1 / iso_mass[i], 1 / mol_mass[j])));
    al = __dadd_rn(al, term);
  }
  al *= florentz;
  double ad = fdoppler / sqrt(iso_mass[i]);
  alphal[t] = al;
  alphad[t] = ad;
  idop0[t] = tb_binsearchapprox(aDop, __dmul_rn(ad, wn0), 0, nDop);
  ilor[t] = tb_binsearchapprox(aLor, al, 0, nLor);
  lastq[t] = -1;
  // the state- and isotope-dependent factor of extinction.c:464-465
  facz[t] = __dmul_rn(TB_SIGCTE, iso_ratio[i]) / __dmul_rn(iso_mass[i], Z[t]);
  ral[t] = 1.0 / al;
}

__device__ __forceinline__ void atomicMaxPosDouble(double *addr, double v) {
  // positive doubles order like their bit patterns: exact, order-independent maximum
  atomicMax((unsigned long long *)addr, (unsigned long long)__double_as_longlong(v));
}

// extinction.c:399-427: per-line strength and its maximum per output row.  One thread per
// member line, SB states per thread so the 26 B line record is read once per SB states.  A line
// whose strength bound g/m/Z (both exponential factors are <= 1 and every rounding is
// monotone) cannot raise the maximum found so far skips its two exp(): the maximum stays
// exact, only the amount of work depends on the visiting order.
#define KMAX_ROWS 16   // output rows reduced in shared memory (more rows: straight to global)
#define KMAX_ISO 64    // isotopes whose skip thresholds are tabulated per block in shared memory
#define KMAX_LPT 8     // lines per thread
template <int SB, int LPT = KMAX_LPT>
__global__ void k_kmax(int64_t l0, int64_t l1, int ns, int niso, const double *__restrict__ q_g,
                       const double *__restrict__ q_a, const double *__restrict__ q_w,
                       const int16_t *__restrict__ q_iso, const double *__restrict__ rtemp,
                       const double *__restrict__ Z, const double *__restrict__ iso_mass,
                       const int *__restrict__ iso_row, int permol, int nrows, double *kmax, int block_skip) {
  // the running maxima are read once per block (not per warp: every warp of the chip asking
  // one L2 slice for the same sector serialises there) and merged back once per block
  __shared__ unsigned long long sh_cur[SB][KMAX_ROWS], sh_new[SB][KMAX_ROWS];
  // the skip bound (1-1e-10)*cur*mass*z depends on (state, isotope) only once the block has its
  // snapshot of the maxima: tabulated, the skip path of a (line, state) pair is one shared-memory
  // read and one comparison (same products in the same order as the inline form: same decisions)
  __shared__ double sh_thr[SB][KMAX_ISO];
  const int s0 = blockIdx.y * SB;
  const bool use_sh = nrows <= KMAX_ROWS;
  const bool use_thr = use_sh && niso <= KMAX_ISO;
  if (use_sh) {
    for (int i = threadIdx.x; i < SB * nrows; i += blockDim.x) {
      const int k = i / nrows, r = i % nrows, s = s0 + k;
      const unsigned long long v = s < ns ? __ldcg((const unsigned long long *)&kmax[s * nrows + r]) : 0ULL;
      sh_cur[k][r] = v;
      sh_new[k][r] = v;
    }
    __syncthreads();
    if (use_thr) {
      // The list is in descending order of key = g/mass, and a line's strength is at most key/z (both exponential
      // factors are <= 1, every rounding monotone).  If the key of the block's FIRST line is below
      // (1-1e-9)*cur*z for every (state, isotope) of the block, no line of the block -- nor of any later block --
      // can raise a maximum: the block leaves before reading its lines (1e-9: ten million times the rounding of
      // the two products and the division).  The maxima stay exact; only the work depends on it.
      __shared__ int sh_stay;
      const int64_t lf = l0 + (int64_t)blockIdx.x * LPT * blockDim.x;
      const double key0 = (block_skip && lf < l1) ? q_g[lf] / iso_mass[q_iso[lf]] : INFINITY;
      if (threadIdx.x == 0) sh_stay = 0;
      __syncthreads();
      for (int i = threadIdx.x; i < SB * niso; i += blockDim.x) {
        const int k = i / niso, iso = i % niso, s = s0 + k;
        const double cur = __longlong_as_double((long long)sh_cur[k][permol ? iso_row[iso] : 0]);
        sh_thr[k][iso] = s < ns ? 0.9999999999 * cur * iso_mass[iso] * Z[s * niso + iso] : 0.0;
        if (s < ns && !(key0 < 0.999999999 * cur * Z[s * niso + iso])) sh_stay = 1;
      }
      __syncthreads();
      if (!sh_stay) return;
    }
  }
#pragma unroll 1
  for (int rep = 0; rep < LPT; rep++) {
    const int64_t l = l0 + ((int64_t)blockIdx.x * LPT + rep) * blockDim.x + threadIdx.x;
    if (l >= l1) break;
    const int iso = q_iso[l];
    const double g = q_g[l];
    const double a = q_a[l];      // > 0 only for a (bogus) negative lower-state energy: never skipped
    double w = 0, mass = 0;
    int row = 0;
    bool loaded = false;
#pragma unroll
    for (int k = 0; k < SB; k++) {
      const int s = s0 + k;
      if (s >= ns) break;
      // g/mass/z <= cur  <=  g < (1-1e-10)*cur*mass*z: no division, no further loads on the skip path
      if (use_thr) { if (!(a > 0) && g < sh_thr[k][iso]) continue; }
      if (!loaded) { w = q_w[l]; mass = iso_mass[iso]; row = permol ? iso_row[iso] : 0; loaded = true; }
      const double cur = use_sh ? __longlong_as_double((long long)sh_cur[k][row]) : __ldcg(&kmax[s * nrows + row]);
      const double z = Z[s * niso + iso];
      if (a > 0 || !(g < 0.9999999999 * cur * mass * z)) {
        const double rT = rtemp[s];
        const double v = g * tb_exp(a * rT) * (1 - tb_exp(w * rT)) / mass / z;
        if (v > cur) {
          if (use_sh) atomicMax(&sh_new[k][row], (unsigned long long)__double_as_longlong(v));
          else atomicMaxPosDouble(&kmax[s * nrows + row], v);
        }
      }
    }
  }
  if (use_sh) {
    __syncthreads();
    for (int i = threadIdx.x; i < SB * nrows; i += blockDim.x) {
      const int k = i / nrows, r = i % nrows, s = s0 + k;
      if (s < ns && sh_new[k][r] > sh_cur[k][r])
        atomicMax((unsigned long long *)&kmax[s * nrows + r], sh_new[k][r]);
    }
  }
}

// The co-added (non-leading) members of a group (extinction.c:449-462) are few; their terms
// gf*exp(-E/T)*(1-exp(-w/T)) are evaluated here, one thread per (member, state), so that
// k_strength's member loop is loads and adds only (a divergent loop with two exp() in it
// costs every warp the full price for one busy lane).
// (one definition for k_extra and k_lastq_scan: the same instructions, the same bits)
__device__ __forceinline__ double tb_member_term(int l, double rT, double efct, const double *__restrict__ m_wavn,
                                                 const double *__restrict__ m_elow, const double *__restrict__ m_gf) {
  const double a = __dmul_rn(__dmul_rn(-TB_EXPCTE, efct), m_elow[l]);
  return __dmul_rn(__dmul_rn(m_gf[l], tb_exp(a * rT)), (1 - tb_exp(__dmul_rn(-TB_EXPCTE, m_wavn[l]) * rT)));
}
// members x0 <= x < x1 (a band shard: only the members of the groups that can reach the band)
__global__ void k_extra(int64_t X, int64_t x0, int64_t x1, int ns, const int32_t *__restrict__ x_line,
                        const double *__restrict__ m_wavn, const double *__restrict__ m_elow,
                        const double *__restrict__ m_gf, const double *__restrict__ rtemp, double efct,
                        double *__restrict__ xterm) {
  const int64_t x = x0 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int s = blockIdx.y;
  if (x >= x1) return;
  xterm[(size_t)s * X + x] = tb_member_term(x_line[x], rtemp[s], efct, m_wavn, m_elow, m_gf);
}

// log-domain threshold offsets: lcz[s][iso] = log(facz) - log(ethresh*kmax[s][row(iso)]).  A lone
// line with  a0/T + log(gf) + lcz < -1e-9  has  gf*exp(a0/T)*facz  (an upper bound of its strength:
// the stimulated-emission factor is <= 1) below the threshold by a relative 1e-9, ten thousand
// times the rounding noise of either side: it is below extinction.c:467's threshold whatever the
// last bits of exp() are.  Everything else takes the exact path.
__global__ void k_logthr(int ns, int niso, int nrows, int permol, double ethresh,
                         const int *__restrict__ iso_row, const double *__restrict__ facz,
                         const double *__restrict__ kmax, double *__restrict__ lcz) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ns * niso) return;
  const int s = t / niso, iso = t % niso;
  const double thr = __dmul_rn(ethresh, kmax[s * nrows + (permol ? iso_row[iso] : 0)]);
  const double f = facz[t];
  lcz[t] = (thr > 0 && f > 0 && isfinite(thr) && isfinite(f)) ? log(f) - log(thr) : INFINITY;
}

// extinction.c:439-483 for one (group, state): co-added strength in member order, scaling,
// threshold, density factor, Doppler-index decision.  Output per (state, group): the strength
// and a 16-bit profile id.  A block owns 256 consecutive groups and SB states:
//   phase 1  one thread per group: the log-domain bound above for each of the SB states (one FMA,
//            no exp); pairs that may pass are appended to a shared-memory list;
//   phase 2  the list is processed densely, one (group, state) pair per thread: both exp(), the
//            co-added members, threshold, Doppler cell.  More than half of all pairs are below
//            the threshold and interleaved with the survivors; without the list every warp would
//            pay the full price for its busiest lane;
//   phase 3  strengths and ids leave through shared memory: whole sectors, coalesced.
// The carried-index and tile-reach maxima are reduced in shared memory first: a handful of
// global integer atomics per block instead of one per group.
#define STR_NR 8   // group ranges one launch of k_strength can cover (a band shard: one range per molecule row)
struct StrRanges {
  long long p_off[STR_NR], p_end[STR_NR];
  int blk0[STR_NR + 1];    // first block of each range
  int n;
};
#define STR_NT 4   // consecutive accumulate tiles a block's reach maxima are kept for in shared memory
template <int SB>
__global__ void __launch_bounds__(256, 6) k_strength(int64_t G, StrRanges rg, int ns, int niso, int o,
                           FastDiv fd_tile, FastDiv fd_o,
                           const int32_t *__restrict__ g_start,
                           const int16_t *__restrict__ g_iso, const int32_t *__restrict__ g_orig,
                           const int32_t *__restrict__ g_iown,
                           const double *__restrict__ g_wavn, const double *__restrict__ g_a0,
                           const double *__restrict__ g_gf, const double *__restrict__ g_lg,
                           const double *__restrict__ rtemp, const double *__restrict__ facz,
                           const double *__restrict__ lcz,
                           const double *__restrict__ ral, const double *__restrict__ xterm, int64_t X,
                           float dop_lmin, float dop_rstep,
                           const double *__restrict__ dens, int nmol,
                           const int *__restrict__ iso_imol, const int *__restrict__ iso_row,
                           const double *__restrict__ alphal, const double *__restrict__ alphad,
                           const int32_t *__restrict__ ilor, const double *__restrict__ aDop, int nDop,
                           const ProfEntry *__restrict__ ptab, int nLor, double ethresh,
                           int permol, int nrows, const double *__restrict__ kmax,
                           LineRec *__restrict__ rec_out,
                           int32_t *__restrict__ lastq, int32_t *__restrict__ wtile,
                           uint8_t *__restrict__ hasfb, int ntiles, const int32_t *__restrict__ fclass) {
  static_assert(SB == 8, "k_strength and k_linesum_fused batch the states alike");
  if (fclass[blockIdx.y]) return;               // this batch of states belongs to k_linesum_fused
  __shared__ int32_t sh_lastq[SB][MAXISO_SH];
  __shared__ double sh_aDop[256];
  __shared__ double sh_k[SB][256];
  __shared__ uint16_t sh_pid[SB][256];
  __shared__ unsigned sh_dead[SB][8];
  __shared__ uint16_t sh_list[SB * 256];
  __shared__ int32_t sh_wt[SB][STR_NT];
  __shared__ int32_t sh_fb[SB][STR_NT];
  __shared__ int sh_cnt;
  const bool use_sh = niso <= MAXISO_SH;
  const int tid = threadIdx.x;
  if (use_sh)
    for (int i = tid; i < SB * MAXISO_SH; i += blockDim.x) (&sh_lastq[0][0])[i] = -1;
  for (int i = tid; i <= nDop; i += blockDim.x) sh_aDop[i] = aDop[i];   // [nDop] = +inf sentinel
  if (tid < SB * STR_NT) { (&sh_wt[0][0])[tid] = -1; (&sh_fb[0][0])[tid] = 0; }
  if (tid == 0) sh_cnt = 0;
  __syncthreads();
  int ir = 0;
  while (ir + 1 < rg.n && (int)blockIdx.x >= rg.blk0[ir + 1]) ir++;
  const int64_t p_end = rg.p_end[ir];
  const int64_t pbase = rg.p_off[ir] + (blockIdx.x - rg.blk0[ir]) * (int64_t)blockDim.x;   // groups [p_off, p_end) of range ir
  const int64_t p = pbase + tid;
  const int s0 = blockIdx.y * SB;
  const bool act = p < p_end;
  const unsigned lane = tid & 31;
  const int tile0 = fd_div(g_iown[pbase], fd_tile);        // pbase < p_end always
  // ---- phase 1
  {
    double a0 = 0, lg = 0;
    int iso = 0;
    if (act) { a0 = g_a0[p]; lg = g_lg[p]; iso = g_iso[p]; }
#pragma unroll
    for (int k = 0; k < SB; k++) {
      const int s = s0 + k;
      bool alive = false;
      if (act && s < ns) alive = !(fma(a0, rtemp[s], lg + lcz[s * niso + iso]) < -1e-9);
      sh_k[k][tid] = 0.0;
      sh_pid[k][tid] = (uint16_t)PID_SKIP;
      const unsigned m = __ballot_sync(0xffffffffu, alive);
      if (m) {
        int base = 0;
        if (lane == 0) base = atomicAdd(&sh_cnt, __popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (alive) sh_list[base + __popc(m & ((1u << lane) - 1))] = (uint16_t)(tid | (k << 8));
      }
    }
  }
  __syncthreads();
  // ---- phase 2
  const int cnt = sh_cnt;
  for (int it = tid; it < cnt; it += blockDim.x) {
    const int item = sh_list[it];
    const int pl = item & 255, k = item >> 8;
    const int64_t q = pbase + pl;
    const int s = s0 + k;
    const int iso = g_iso[q];
    const int b = g_start[q], e = g_start[q + 1];
    const double wlead = g_wavn[q];
    const double rT = rtemp[s];
    const int si = s * niso + iso;
    const double fz = facz[si], thr = __dmul_rn(ethresh, kmax[s * nrows + (permol ? iso_row[iso] : 0)]);
    const double ge1 = __dmul_rn(g_gf[q], tb_exp(g_a0[q] * rT));
    double kk = 0.0;
    // a lone line whose strength without the (<= 1) stimulated-emission factor is already
    // clearly below the threshold is skipped without the second exp()
    bool below = (e - b == 1) && ge1 * fz < 0.999999 * thr;
    if (!below) {
      double acc = __dmul_rn(ge1, (1 - tb_exp(__dmul_rn(-TB_EXPCTE, wlead) * rT)));
      for (int l = b + 1; l < e; l++)                        // co-added members (extinction.c:449-462)
        acc = __dadd_rn(acc, xterm[(size_t)s * X + (l - (int)q - 1)]);
      kk = __dmul_rn(acc, fz);
      below = kk < thr;                                      // extinction.c:467-470
    }
    if (below) continue;
    if (!permol) kk = __dmul_rn(kk, dens[s * nmol + iso_imol[iso]]);
    const double al = alphal[si], adw = __dmul_rn(alphad[si], wlead);
    // extinction.c:480-483: alphad*wavn/alphal >= 0.1, decided by the exact quotient whenever
    // the reciprocal product is within rounding distance of the boundary
    const double qa = adw * ral[si];
    const bool refresh = fabs(qa - 1e-1) > 1e-12 ? qa >= 1e-1 : adw / al >= 1e-1;
    const int iown = g_iown[q];
    const int tile = fd_div(iown, fd_tile) - tile0;           // fine bin / (o * ACC_TILE)
    const bool tsh = tile >= 0 && tile < STR_NT;
    unsigned pid;
    if (refresh) {
      // binsearchapprox (iomisc.c:1088-1108) on the log-spaced table: start from the
      // closed-form cell and settle with exact comparisons: lo = last entry <= adw
      int lo = (int)((__log2f((float)adw) * 0.30102999566f - dop_lmin) * dop_rstep);
      lo = lo < 0 ? 0 : (lo > nDop - 1 ? nDop - 1 : lo);
      while (lo > 0 && sh_aDop[lo] > adw) lo--;
      while (sh_aDop[lo + 1] <= adw) lo++;                    // [nDop] = +inf stops it
      const int hi = lo + 1;
      const int id = (fabs(sh_aDop[hi] - adw) < fabs(sh_aDop[lo] - adw)) ? hi : lo;
      pid = id * nLor + ilor[si];
      const int ps = ptab[pid].psize;
      {  // does the profile cover any grid point at all?  |iown - o*nearest| <= psize
        const int dist = iown - fd_div(iown + (o >> 1), fd_o) * o;
        if ((dist < 0 ? -dist : dist) > ps) pid |= PID_NOREACH;
      }
      {  // the maxima rarely move after a block's first records: look before the (conflicting) atomic
        const int og = g_orig[q];
        if (use_sh) { if (og > sh_lastq[k][iso]) atomicMax(&sh_lastq[k][iso], og); }
        else atomicMax(&lastq[si], og);
      }
      // widest profile per (state, tile of output bins) among those that reach a bin
      if (!(pid & PID_NOREACH)) {
        if (tsh) { if (ps > sh_wt[k][tile]) atomicMax(&sh_wt[k][tile], ps); }
        else atomicMax(&wtile[s * ntiles + tile + tile0], ps);
      }
    } else {
      pid = PID_FB + iso;
      if (tsh) sh_fb[k][tile] = 1;
      else hasfb[s * ntiles + tile + tile0] = 1;
    }
    sh_k[k][pl] = kk;
    sh_pid[k][pl] = (uint16_t)pid;
  }
  __syncthreads();
  // ---- phase 3
  // dead-record bitmaps per (state, warp): each record carries the length of the dead run after it
  // (within this block, up to 63 - lane), which the narrow walk of k_accumulate jumps over
#pragma unroll
  for (int k = 0; k < SB; k++) {
    const unsigned m = __ballot_sync(0xffffffffu, act && PID_DEAD((unsigned)sh_pid[k][tid]));
    if ((tid & 31) == 0) sh_dead[k][tid >> 5] = m;
  }
  __syncthreads();
  if (act) {
    const int io = g_iown[p];
    const int lane = tid & 31, w = tid >> 5;
#pragma unroll
    for (int k = 0; k < SB; k++) {
      const int s = s0 + k;
      if (s < ns) {
        // written for skipped groups too: whole sectors, no fill reads
        const double kk = sh_k[k][tid];
        const unsigned long long dm =
            (((unsigned long long)(w < 7 ? sh_dead[k][w + 1] : 0u) << 32) | sh_dead[k][w]) >> (lane + 1);
        const int skip = __ffsll((long long)~dm) - 1;
        int4 v;
        v.x = __double2loint(kk); v.y = __double2hiint(kk); v.z = io; v.w = (int)sh_pid[k][tid] | (skip << 16);
        *reinterpret_cast<int4 *>(rec_out + (size_t)s * G + p) = v;
      }
    }
  }
  if (tid < SB * STR_NT) {
    const int k = tid / STR_NT, t = tid % STR_NT, s = s0 + k;
    if (s < ns && tile0 + t < ntiles) {
      if (sh_wt[k][t] > __ldcg(&wtile[s * ntiles + tile0 + t])) atomicMax(&wtile[s * ntiles + tile0 + t], sh_wt[k][t]);
      if (sh_fb[k][t]) hasfb[s * ntiles + tile0 + t] = 1;
    }
  }
  if (use_sh) {
    for (int i = tid; i < SB * niso; i += blockDim.x) {
      const int k = i / niso, is = i % niso, s = s0 + k;
      const int v = sh_lastq[k][is];
      // most blocks cannot raise the maximum (sorted order runs against the original order):
      // look before paying for an atomic on an address every block of this state shares
      if (s < ns && v >= 0 && v > __ldcg(&lastq[s * niso + is])) atomicMax(&lastq[s * niso + is], v);
    }
  }
}

// carried Doppler index per (state, isotope): the index left by the last evaluated line that
// refreshed it, else the one computed at wn[0] (extinction.c:393,480-483).
__global__ void k_fallback(int ns, int niso, const int32_t *__restrict__ fclass, int which, const int32_t *__restrict__ lastq,
                           const int32_t *__restrict__ idop0, const double *__restrict__ go_wavn,
                           const double *__restrict__ alphad, const int32_t *__restrict__ ilor,
                           const double *__restrict__ aDop, int nDop, int nLor,
                           const ProfEntry *__restrict__ ptab, int32_t *__restrict__ fb,
                           int32_t *__restrict__ fbpid, int32_t *__restrict__ fbmax) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ns * niso) return;
  int s = t / niso;
  if ((fclass[s / FB] != 0) != (which != 0)) return;    // which = 1: the fused states (lastq from k_lastq_scan), 0: the others
  int q = lastq[t];
  int id = (q < 0) ? idop0[t] : tb_binsearchapprox(aDop, __dmul_rn(alphad[t], go_wavn[q]), 0, nDop);
  fb[t] = id;
  const int cell = id * nLor + ilor[t];
  fbpid[t] = cell;
  atomicMax(&fbmax[s], ptab[cell].psize);
}

// reach[s][t] / reachP[s][t]: how many coarse bins (and fine bins) to either side an output bin
// of tile t must look for contributing groups.  A group whose fine bin lies in coarse cell d and
// whose profile half-size is psize touches bins d - psize/o .. d + 1 + psize/o.
__global__ void k_reach(int ns, int ntiles, int o, int utiles, const int32_t *__restrict__ wtile,
                        const uint8_t *__restrict__ hasfb, const int32_t *__restrict__ fbmax,
                        int32_t *__restrict__ reach, int32_t *__restrict__ reachP) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ns * ntiles) return;
  const int s = i / ntiles, t = i % ntiles;
  int R = 0, P = -1;
  const int u0 = t - utiles < 0 ? 0 : t - utiles, u1 = t + utiles + 1 > ntiles ? ntiles : t + utiles + 1;
  for (int u = u0; u < u1; u++) {      // utiles: how far the widest profile of the table can reach
    int ps = wtile[s * ntiles + u];
    if (hasfb[s * ntiles + u]) ps = max(ps, fbmax[s]);
    if (ps < 0) continue;
    const int w = 1 + ps / o;
    const int du = u > t ? u - t : t - u;
    const int gap = du == 0 ? 0 : (du - 1) * ACC_TILE + 1;   // closest pair of bins of the two tiles
    if (w >= gap) { R = max(R, w); P = max(P, ps); }
  }
  reach[i] = R;
  reachP[i] = P;
}

// extinction.c:486-509 turned inside out: every output bin gathers its contributing groups
// from the windows of each molecule row, always in sorted (row, fine bin) order: a fixed
// summation order, no floating-point atomics.
//   contributes  <=>  0 <= o*j - (iown - psize) <= 2*psize   (same test as the reference loop)
// A block owns ACC_TILE consecutive bins of one state.  Two schedules, each with a fixed
// summation order:
//   * narrow reach (R <= NARROW_R bins): one thread per bin; it starts at the first record of the
//     sub-cell holding o*j - P (P = widest half-size that can reach the tile; substart[] has NSUB
//     sub-cells per bin, so no bisection) and walks the fine-bin-sorted records up to o*j + P;
//   * wide reach: the block walks the union window of its 128 bins once: the window's records are
//     split into four contiguous runs, one per warp; a warp fetches 32 records coalesced,
//     resolves them (table address, covered bins by an exact multiply-high division), compacts
//     the survivors into shared memory and broadcasts them; lane i adds the samples of bins
//     i, i+32, i+64, i+96 of the tile, each one contiguous 256-byte read of the polyphase
//     table.  The four partial sums per bin are added in warp order.
#define NARROW_R 2
struct AccRec {
  double k;
  int addr0;   // table address of bin 0's sample
  int lanes;   // lo | (hi - lo) << 8: bins (tile-relative, lo..hi) the profile covers
};
#define ACC_WW 16                          // warps of a wide-reach block (the window's records are split 16 ways)
struct WideShared {
  AccRec rec[ACC_WW][32];
  double part[ACC_WW][ACC_TILE];
  unsigned ticket;
};
#define ACC_SPLIT_MAX 8                    // blocks that may share one wide-reach tile on a small grid

// Two instances over the same (tile, state) grid: WIDE = false runs 128 threads and takes the tiles
// of narrow reach (and zero-fills the tiles nothing reaches), WIDE = true runs ACC_WW warps and takes
// the rest; a block whose tile belongs to the other instance leaves at once.  Sixteen warps per wide
// tile keep the machine full when a band shard (or a small batch of states) leaves few tiles.
template <bool WIDE>
__global__ void __launch_bounds__(WIDE ? 32 * ACC_WW : ACC_TILE)
k_accumulate(int64_t nwn, int64_t G, int ns, int niso, int o, FastDiv fd, int nrows, int permol, int ntiles,
             int tile_off, const int32_t *__restrict__ binstart, const int32_t *__restrict__ substart, int64_t nsub,
             int nsube, const LineRec *__restrict__ rec_in,
             const int32_t *__restrict__ fbpid, const ProfEntry *__restrict__ ptab,
             const int32_t *__restrict__ reach, const int32_t *__restrict__ reachP,
             const double *__restrict__ pdata, double *__restrict__ out,
             unsigned long long *__restrict__ nsamp, int S, double *__restrict__ part_g,
             unsigned *__restrict__ tilecnt, const int32_t *__restrict__ fclass, int narrow_r) {
  const int s = blockIdx.y;
  if (fclass[s >> 3]) return;                    // summed by k_linesum_fused
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // S > 1 (wide instance on a grid of few tiles): S blocks share a tile, block k of them takes the
  // segments k*ACC_WW .. k*ACC_WW+ACC_WW-1 of the window's S*ACC_WW record segments
  const int ksplit = S > 1 ? (int)(blockIdx.x % S) : 0;
  const int tile = (S > 1 ? (int)(blockIdx.x / S) : (int)blockIdx.x) + tile_off;   // a band shard launches only its own tiles
  const int64_t jt = (int64_t)tile * ACC_TILE;
  const int64_t j = jt + tid;
  const bool inb = j < nwn && tid < ACC_TILE;
  const int R = reach[s * ntiles + tile];
  const int P = reachP[s * ntiles + tile];
  const int4 *rs = reinterpret_cast<const int4 *>(rec_in + (size_t)s * G);
  const int32_t *fbs = fbpid + (size_t)s * niso;
  unsigned long long cnt = 0;
  if (WIDE != (P >= 0 && R > narrow_r)) return;  // the other instance's tile
  if (P < 0) {                                   // nothing reaches this tile
    for (int row = 0; row < (permol ? nrows : 1); row++)
      if (inb) out[((size_t)s * (permol ? nrows : 1) + row) * nwn + j] = 0.0;
    return;
  }
  if (!WIDE) {
    double acc = 0.0;
    if (inb) {
      const int64_t jhi = j + R + 1 > nwn ? nwn : j + R + 1;
      const int oj = (int)(o * j);
      const int lo_f = oj - P < 0 ? 0 : oj - P, hi_f = oj + P;
      const int u = fd_div(lo_f * nsube, fd);      // sub-cell of the window's first fine bin
      for (int row = 0; row < nrows; row++) {
        const int p1 = binstart[(size_t)row * nwn + jhi];
        // software-pipelined walk: the next record (one 128-bit read) is in flight while the
        // current one is resolved
        int p = substart[(size_t)row * nsub + u];
        int4 r_n = p < p1 ? __ldg(rs + p) : make_int4(0, 0, 0x7fffffff, PID_SKIP);
        while (p < p1 && r_n.z <= hi_f) {
          const int io = r_n.z;
          unsigned pid = (unsigned)r_n.w & 0xFFFFu;
          const double kk = __hiloint2double(r_n.y, r_n.x);
          p += 1 + (int)((unsigned)r_n.w >> 16);   // over the dead records that follow (k_strength counted them)
          if (p < p1) r_n = __ldg(rs + p);
          if (PID_DEAD(pid)) continue;
          if (pid >= PID_FB) pid = fbs[pid - PID_FB];
          const ProfEntry pe = ptab[pid];
          const int idx = oj - io + pe.psize;
          if (idx >= 0 && idx <= 2 * pe.psize) {
            int ad = idx;
            if (pe.nk != 1) {
              const int q = fd_div(idx, fd);
              ad = (idx - q * o) * pe.nk + q;
            }
            acc = fma(kk, pdata[pe.off + ad], acc);
            cnt++;
          }
        }
        if (permol) {
          out[((size_t)s * nrows + row) * nwn + j] = acc;
          acc = 0.0;
        }
      }
      if (!permol) out[(size_t)s * nwn + j] = acc;
    }
  } else {
    __shared__ __align__(16) WideShared w;
    const int maxbin = (int)(nwn - 1 - jt < ACC_TILE - 1 ? nwn - 1 - jt : ACC_TILE - 1);
    const int64_t ulo = jt - R < 0 ? 0 : jt - R;
    const int64_t uhi = jt + ACC_TILE + R > nwn ? nwn : jt + ACC_TILE + R;
    const int ojt = (int)(o * jt);
    double acc[ACC_TILE / 32];
#pragma unroll
    for (int q = 0; q < ACC_TILE / 32; q++) acc[q] = 0.0;
    for (int row = 0; row < nrows; row++) {
      const int32_t *bs = binstart + (size_t)row * nwn;
      const int p0 = bs[ulo], p1 = bs[uhi];
      // the window's records in chunks of 32, dealt round-robin to the ACC_WW * S warps that share the tile: the
      // alive records and their spans vary smoothly along the window, so every warp gets the same mix (contiguous
      // runs left a quarter of the stall samples at the block's last barrier); a fixed deal, hence a fixed order
      const int nseg = ACC_WW * S;
      const int pb = p1;
      for (int base = p0 + (ksplit * ACC_WW + warp) * 32; base < pb; base += nseg * 32) {
        const int p = base + lane;
        const int4 rr = p < pb ? __ldg(rs + p) : make_int4(0, 0, 0, PID_SKIP);
        unsigned pid = (unsigned)rr.w & 0xFFFFu;
        bool alive = !PID_DEAD(pid);
        AccRec rec;
        if (alive) {
          if (pid >= PID_FB) pid = fbs[pid - PID_FB];
          const ProfEntry pe = ptab[pid];
          const int idx0 = ojt - rr.z + pe.psize;                // sample index for bin 0 of the tile
          // bins with 0 <= idx0 + bin*o <= 2*psize
          const int lo = idx0 >= 0 ? 0 : fd_div(-idx0 + o - 1, fd);
          int hi = 2 * pe.psize - idx0;
          hi = hi < 0 ? -1 : fd_div(hi, fd);
          if (hi > maxbin) hi = maxbin;
          if (lo > hi) alive = false;
          else {
            const int first = idx0 + lo * o;                     // >= 0
            const int q = fd_div(first, fd);
            rec.addr0 = pe.off + (first - q * o) * pe.nk + q - lo;
            rec.lanes = lo | ((hi - lo) << 8);
            rec.k = __hiloint2double(rr.y, rr.x);
            cnt += hi - lo + 1;
          }
        }
        const unsigned m = __ballot_sync(0xffffffffu, alive);
        if (alive) w.rec[warp][__popc(m & ((1u << lane) - 1))] = rec;
        // the 32-bin quarters of the tile any record of the chunk touches (the chunk's records are neighbours in
        // fine bin: away from the deepest layers they all sit in one or two quarters)
        unsigned qm = 0;
        if (alive) {
          const int lo = rec.lanes & 255, hi = lo + (rec.lanes >> 8);
          qm = ((2u << (hi >> 5)) - 1u) & ~((1u << (lo >> 5)) - 1u);
        }
        qm = __reduce_or_sync(0xffffffffu, qm);
        __syncwarp();
        const int n = __popc(m);
        if (__popc(qm) > 2) {
          // record-major: four predicated (load, FMA) pairs per record
#pragma unroll 2
          for (int i = 0; i < n; i++) {
            const AccRec r = w.rec[warp][i];
            const int d = lane - (r.lanes & 255);                  // bin - lo for q = 0
            const unsigned span = (unsigned)(r.lanes >> 8);        // hi - lo
            const double *pp = pdata + r.addr0 + lane;
#pragma unroll
            for (int q = 0; q < ACC_TILE / 32; q++)
              if ((unsigned)(d + 32 * q) <= span) acc[q] = fma(r.k, __ldg(pp + 32 * q), acc[q]);
          }
        } else {
          // quarter-major: only the touched quarters walk the records (same records in the same order per bin)
#pragma unroll
          for (int q = 0; q < ACC_TILE / 32; q++) {
            if (!((qm >> q) & 1u)) continue;                       // warp-uniform
            const int bin = lane + 32 * q;
            const double *pp = pdata + bin;
#pragma unroll 4
            for (int i = 0; i < n; i++) {
              const AccRec r = w.rec[warp][i];
              if ((unsigned)(bin - (r.lanes & 255)) <= (unsigned)(r.lanes >> 8)) acc[q] = fma(r.k, __ldg(pp + r.addr0), acc[q]);
            }
          }
        }
        __syncwarp();
      }
      if (permol || row == nrows - 1) {
#pragma unroll
        for (int q = 0; q < ACC_TILE / 32; q++) { w.part[warp][lane + 32 * q] = acc[q]; acc[q] = 0.0; }
        __syncthreads();
        const size_t orow = (size_t)s * (permol ? nrows : 1) + (permol ? row : 0);
        double v = 0.0;
        if (tid < ACC_TILE) {
          v = w.part[0][tid];
#pragma unroll
          for (int g = 1; g < ACC_WW; g++) v += w.part[g][tid];
        }
        if (S == 1) {
          if (inb) out[orow * nwn + j] = v;
        } else {
          // the S partial sums of a tile meet in global memory; the block that arrives last adds
          // them in split order (a fixed order whichever block that is)
          double *pt = part_g + ((orow * ntiles + tile) * S) * ACC_TILE;
          if (tid < ACC_TILE) pt[(size_t)ksplit * ACC_TILE + tid] = v;
          __threadfence();
          __syncthreads();
          if (tid == 0) w.ticket = atomicAdd(&tilecnt[orow * ntiles + tile], 1u);
          __syncthreads();
          if (w.ticket == (unsigned)(S - 1)) {
            __threadfence();
            if (inb) {
              double t = __ldcg(pt + tid);
              for (int k2 = 1; k2 < S; k2++) t += __ldcg(pt + (size_t)k2 * ACC_TILE + tid);
              out[orow * nwn + j] = t;
            }
          }
        }
        __syncthreads();
      }
    }
  }
  // sample counter (the benchmark's unit of work): integer atomics only
  for (int d = 16; d > 0; d >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
  if (lane == 0 && cnt) atomicAdd(&nsamp[s], cnt);
}


// ------------------------------------------------------------------ fused line sum of the narrow layers
//
// For a state (layer) whose profiles are narrow, k_strength + k_accumulate spend their time on the
// 16-byte record every (group, state) pair leaves in HBM and on fetching it back bin by bin.  Where
//   (1) every isotope always refreshes its Doppler index (alphad*wn_i/alphal >= 0.1 with margin, so the
//       carried index of extinction.c:480-483 -- a quantity global over the line list -- is never used), and
//   (2) the widest profile any line of the state can take reaches at most FUSE_RMAX coarse bins,
// both halves run in ONE kernel and the records never leave the SM.  A block owns a tile of 128 output
// bins and FB = 8 consecutive states.  Per molecule row it walks the tile's window of groups (tile +-
// the widest profile of that row near the tile, found through binstart) in chunks of up to FUSE_CAPA alive groups:
//   phase 1  one thread per candidate, 256 at a time: the log-domain pretest of k_strength for the 8 states
//            (no exp), ordered compaction of the groups alive in any state (warp ballots + one block prefix);
//   phase 2  one thread per alive group, its constants in registers, a loop over its alive states:
//            both exp(), co-added members, threshold, Doppler cell, "covers no grid point" test -- the
//            decisions of k_strength, expression by expression -- into 16-byte shared-memory records
//            {k, table address of the tile's bin 0, first and last covered bin} ordered by fine bin;
//   phase 3  one thread per (bin, 4 states): bisection for the first record of the bin's window, then the
//            walk of k_accumulate's narrow schedule over the shared records: one 128-bit shared read, two
//            comparisons and, on a hit, one table read + FMA per (record, state).
// The sum of a bin runs over rows, chunks and records in sorted (row, fine bin) order: the same fixed
// order, the same FMA chain and therefore the same bits as the unfused narrow path; tiles are independent,
// so band shards stay bit-identical.  No floating-point atomics.
#define FBK 4                // states per block of the fused kernel
#define FPT (FBK / 2)        // states per thread in phase 3 (two threads per bin)
#define FUSE_BLOCKS 4        // resident blocks per SM the kernel is compiled for (64 registers)
#define FUSE_CAPA 512        // alive groups per chunk (two full passes of phase 2)
#define FUSE_NT 256
#define FUSE_MAXISO 64
#define FUSE_RMAX 8          // default widest reach, in coarse bins, of a fused state (tb_set_fused_reach)
#define FUSE_MIN_TILES 96     // grids with fewer 128-bin tiles keep the two-kernel path (tb_set_fused(c, 2) forces fusion)

// Which batches of FB states the fused kernel takes (decided on the device: no read-back on the way)
__global__ void k_fuse_class(int ns, int niso, double wn_i, double own_last, int o, int enable, int need_refresh, int rmax,
                             const double *__restrict__ alphad, const double *__restrict__ ral,
                             const int32_t *__restrict__ ilor, const double *__restrict__ aDop, int nDop, int nLor,
                             const ProfEntry *__restrict__ ptab, int32_t *__restrict__ fclass,
                             int32_t *__restrict__ fP) {
  __shared__ int sh_ok, sh_P;
  const int s0 = blockIdx.x * FB;
  if (threadIdx.x == 0) { sh_ok = 1; sh_P = 0; }
  __syncthreads();
  for (int i = threadIdx.x; i < FB * niso; i += blockDim.x) {
    const int k = i / niso, iso = i % niso, s = s0 + k;
    if (s >= ns) continue;
    const int si = s * niso + iso;
    const double ad = alphad[si];
    // every in-range line has wavn >= wn_i: alphad*wavn/alphal >= this, with a margin far above rounding
    if (need_refresh && !(__dmul_rn(ad, wn_i) * ral[si] >= 0.1000001)) sh_ok = 0;
    // ... and wavn <= own_last: no line takes a Doppler cell beyond this one (nearest-cell search is monotone)
    const int id = tb_binsearchapprox(aDop, __dmul_rn(ad, own_last), 0, nDop);
    atomicMax(&sh_P, ptab[id * nLor + ilor[si]].psize);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int P = sh_P;
    const long long lim = (long long)rmax * o < 0x3fffffff ? (long long)rmax * o : 0x3fffffff;
    fclass[blockIdx.x] = (enable && sh_ok && niso >= 1 && niso <= FUSE_MAXISO && P <= lim) ? 1 : 0;
    fP[blockIdx.x] = P;
  }
}

// fill the records of fused states with "below threshold" when the per-line trace is wanted (tb_set_trace)
__global__ void k_fuse_trace_fill(int64_t G, int ns, const int32_t *__restrict__ fclass,
                                  const int32_t *__restrict__ g_iown, LineRec *__restrict__ rec) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int s = blockIdx.y;
  if (p >= G || s >= ns || !fclass[s / FB]) return;
  LineRec r;
  r.k = 0.0; r.iown = g_iown[p]; r.pid = PID_SKIP;
  rec[(size_t)s * G + p] = r;
}


// extinction.c:480-483 as k_strength decides it: refresh the isotope's Doppler index at this line?
__device__ __forceinline__ bool tb_refresh(double adw, double al, double ral) {
  const double qa = adw * ral;
  return fabs(qa - 1e-1) > 1e-12 ? qa >= 1e-1 : adw / al >= 1e-1;
}

// Is group q evaluated at state s (extinction.c:439-470)?  The expressions of k_strength / k_linesum_fused.
__device__ __forceinline__ bool tb_group_evaluated(int q, size_t sX, double rT, double fz, double thr,
                                                   const int32_t *__restrict__ g_start, const double *__restrict__ g_gf,
                                                   const double *__restrict__ g_a0, const double *__restrict__ g_wavn,
                                                   const double *__restrict__ xterm) {
  const int b = g_start[q], e = g_start[q + 1];
  const double ge1 = __dmul_rn(g_gf[q], tb_exp(g_a0[q] * rT));
  if ((e - b == 1) && ge1 * fz < 0.999999 * thr) return false;
  double a = __dmul_rn(ge1, (1 - tb_exp(__dmul_rn(-TB_EXPCTE, g_wavn[q]) * rT)));
  for (int l = b + 1; l < e; l++) a = __dadd_rn(a, xterm[sX + (l - q - 1)]);
  return !(__dmul_rn(a, fz) < thr);
}
// ... with the co-added members' terms evaluated here (k_extra's expression) instead of read from xterm: k_lastq_scan
// looks at groups anywhere in the list, a band shard's k_extra only fills the members its own groups need
__device__ __forceinline__ bool tb_group_evaluated_x(int q, double rT, double fz, double thr, double efct,
                                                     const int32_t *__restrict__ g_start, const double *__restrict__ g_gf,
                                                     const double *__restrict__ g_a0, const double *__restrict__ g_wavn,
                                                     const int32_t *__restrict__ x_line, const double *__restrict__ m_wavn,
                                                     const double *__restrict__ m_elow, const double *__restrict__ m_gf) {
  const int b = g_start[q], e = g_start[q + 1];
  const double ge1 = __dmul_rn(g_gf[q], tb_exp(g_a0[q] * rT));
  if ((e - b == 1) && ge1 * fz < 0.999999 * thr) return false;
  double a = __dmul_rn(ge1, (1 - tb_exp(__dmul_rn(-TB_EXPCTE, g_wavn[q]) * rT)));
  for (int l = b + 1; l < e; l++) a = __dadd_rn(a, tb_member_term(x_line[l - q - 1], rT, efct, m_wavn, m_elow, m_gf));
  return !(__dmul_rn(a, fz) < thr);
}

// The carried Doppler index without walking the list (extinction.c:480-483): per (state, isotope) the reference
// leaves, for the lines that do not refresh it, the index of the LAST line in file order that did refresh it and
// was evaluated.  An isotope's lines come in one run of descending wavenumber and "refreshes" is monotone in the
// wavenumber, so the refreshing lines are a prefix of the run: its end is bisected, and the run is searched
// backwards from there for the first evaluated group -- in chunks of 256 groups dealt to LASTQ_NB = 8 blocks per
// (state, isotope), each chunk through the log-domain pretest first (an isotopologue whose lines are all below the
// threshold costs one cheap pass, not a serial walk), hits merged by an integer atomicMax, a chunk skipped as soon
// as a hit nearer the end of the prefix is known.  A pair whose lines ALL refresh never reads the carried index
// and is left alone.  Nothing here depends on a band shard: every rank finds the same index.
#define LASTQ_NB 8
__global__ void __launch_bounds__(256)
k_lastq_scan(int ns, int niso, int nrows, int permol, double ethresh, double wn_i,
             const int32_t *__restrict__ iso_og0, const int32_t *__restrict__ iso_og1,
             const int32_t *__restrict__ go_pos, const double *__restrict__ go_wavn,
             const int32_t *__restrict__ g_start, const double *__restrict__ g_gf,
             const double *__restrict__ g_a0, const double *__restrict__ g_lg, const double *__restrict__ g_wavn,
             const double *__restrict__ rtemp, const double *__restrict__ facz, const double *__restrict__ lcz,
             const double *__restrict__ kmax, const int *__restrict__ iso_row,
             const double *__restrict__ alphal, const double *__restrict__ alphad,
             const double *__restrict__ ral, const int32_t *__restrict__ x_line, const double *__restrict__ m_wavn,
             const double *__restrict__ m_elow, const double *__restrict__ m_gf, double efct,
             const int32_t *__restrict__ fclass, int all, int32_t *lastq /* preset to -1 */) {
  const int pair = blockIdx.x, s = pair / niso, iso = pair % niso;      // (x: up to 2^31 pairs; y: the chunk dealer)
  if (!all && !fclass[s / FB]) return;           // (else) k_strength finds it for the states it evaluates
  const int og0 = iso_og0[iso], og1 = iso_og1[iso];
  if (og1 <= og0) return;
  const double al = alphal[pair], ad = alphad[pair], rl = ral[pair];
  if (tb_refresh(__dmul_rn(ad, wn_i), al, rl)) return;      // every in-range line refreshes: the carried index is never read
  __shared__ int sh_best;
  int l = og0, h = og1;                          // first group of the run that does NOT refresh
  while (l < h) {
    const int mid = (l + h) >> 1;
    if (tb_refresh(__dmul_rn(ad, go_wavn[mid]), al, rl)) l = mid + 1; else h = mid;
  }
  const int ogb = l - 1;
  const double rT = rtemp[s], fz = facz[pair], lz = lcz[pair];
  const double thr = __dmul_rn(ethresh, kmax[s * nrows + (permol ? iso_row[iso] : 0)]);
  for (int j = blockIdx.y;; j += LASTQ_NB) {
    const int hi = ogb - 256 * j;
    if (hi < og0) break;
    if (__ldcg(lastq + pair) >= hi) break;       // a hit nearer the end of the prefix: nothing here can beat it
    if (threadIdx.x == 0) sh_best = -1;
    __syncthreads();
    const int og = hi - (int)threadIdx.x;
    if (og >= og0) {
      const int q = go_pos[og];
      if (!(fma(g_a0[q], rT, g_lg[q] + lz) < -1e-9) &&
          tb_group_evaluated_x(q, rT, fz, thr, efct, g_start, g_gf, g_a0, g_wavn, x_line, m_wavn, m_elow, m_gf))
        atomicMax(&sh_best, og);
    }
    __syncthreads();
    const int best = sh_best;
    __syncthreads();
    if (best >= 0) {
      if (threadIdx.x == 0) atomicMax(lastq + pair, best);
      break;
    }
  }
}

struct FuseTab {           // per (state of the block, isotope), in shared memory
  double lcz, facz, thr, densf, alphad, alphal, ral;
};
// What phase 2 leaves per (alive group, state) for phase 3: the strength and everything the gather needs
// as one 128-bit shared-memory read.  The profile sample bin j of the tile takes is pdata[A + j]:
//   idx = o*j - iown + psize = o*(j - q + m) + rho   (iown = o*q + r, psize - r = o*m + rho, 0 <= rho < o)
//   polyphase address = off + rho*nk + (j - q + m)   =>   A = off + rho*nk + m - q + jt  (j tile-relative)
struct __align__(16) FuseRec {
  double k;
  int A;
  short jl, jh;            // tile-relative bins the profile covers (jl > jh: none)
};

template <bool CARRIED>      // CARRIED: some lines keep the isotope's carried Doppler index (fbpid) instead of their own
__global__ void __launch_bounds__(FUSE_NT, FUSE_BLOCKS)
k_linesum_fused(int64_t nwn, int64_t onwn, int64_t G, int ns, int niso, int nmol, int o, FastDiv fd_o, FastDiv fd_tile,
                int nrows, int permol, int tile_off, double wn_i, double odwn,
                const int32_t *__restrict__ fclass, const int32_t *__restrict__ fP,
                const int32_t *__restrict__ binstart, const int32_t *__restrict__ g_start,
                const int16_t *__restrict__ g_iso, const int32_t *__restrict__ g_iown,
                const double *__restrict__ g_wavn, const double *__restrict__ g_a0,
                const double *__restrict__ g_gf, const double *__restrict__ g_lg,
                const double *__restrict__ rtemp, const double *__restrict__ facz, const double *__restrict__ lcz,
                const double *__restrict__ xterm, int64_t X, float dop_lmin, float dop_rstep,
                const double *__restrict__ dens, const int *__restrict__ iso_imol, const int *__restrict__ iso_row,
                const double *__restrict__ alphad, const int32_t *__restrict__ ilor,
                const double *__restrict__ alphal, const double *__restrict__ ral, const int32_t *__restrict__ fbpid,
                const double *__restrict__ aDop, int nDop, const ProfEntry *__restrict__ ptab, int nLor,
                double ethresh, const double *__restrict__ kmax, const double *__restrict__ pdata,
                double *__restrict__ out, unsigned long long *__restrict__ nsamp, LineRec *__restrict__ rec_out) {
  const int batch = blockIdx.y / (FB / FBK);       // classification is per batch of FB states (as k_strength)
  if (!fclass[batch]) return;                      // k_strength + k_accumulate take this batch
  extern __shared__ __align__(16) unsigned char fsm[];
  // ---- shared-memory carve-up
  FuseRec *sh_rec = reinterpret_cast<FuseRec *>(fsm);                            // [FBK][CAPA]
  int *sh_iown = reinterpret_cast<int *>(sh_rec + FBK * FUSE_CAPA);               // [CAPA]
  int *sh_gidx = sh_iown + FUSE_CAPA;                                            // [CAPA]
  double *sh_aDop = reinterpret_cast<double *>(sh_gidx + FUSE_CAPA);             // [256]
  double *sh_rT = sh_aDop + 256;                                                 // [FBK]
  FuseTab *sh_tab = reinterpret_cast<FuseTab *>(sh_rT + FBK);                    // [FBK][niso]
  int *sh_ilor = reinterpret_cast<int *>(sh_tab + FBK * niso);                   // [FBK][niso]
  int *sh_fbpid = sh_ilor + FBK * niso;                                          // [FBK][niso] carried cell
  unsigned char *sh_mask = reinterpret_cast<unsigned char *>(sh_fbpid + FBK * niso);  // [CAPA]
  __shared__ int sh_wcnt[2][FUSE_NT / 32];
  __shared__ int sh_Prow[16];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int s0 = blockIdx.y * FBK;
  const int tile = blockIdx.x + tile_off;
  const int64_t jt = (int64_t)tile * ACC_TILE;
  const int Pb = fP[batch];
  for (int i = tid; i <= nDop; i += FUSE_NT) sh_aDop[i] = aDop[i];               // [nDop] = +inf sentinel
  if (tid < FBK) sh_rT[tid] = rtemp[s0 + tid < ns ? s0 + tid : ns - 1];
  if (tid < 16) sh_Prow[tid] = 0;
  for (int i = tid; i < FBK * niso; i += FUSE_NT) {
    const int k = i / niso, iso = i % niso, s = s0 + k < ns ? s0 + k : ns - 1;
    const int si = s * niso + iso;
    FuseTab t;
    t.lcz = lcz[si]; t.facz = facz[si];
    t.thr = __dmul_rn(ethresh, kmax[s * nrows + (permol ? iso_row[iso] : 0)]);
    t.densf = dens[s * nmol + iso_imol[iso]];
    t.alphad = alphad[si]; t.alphal = alphal[si]; t.ral = ral[si];
    sh_tab[i] = t;
    sh_ilor[i] = ilor[si];
    sh_fbpid[i] = CARRIED ? fbpid[si] : 0;
  }
  __syncthreads();
  // ---- the widest profile a line of each molecule row can take in THIS tile: nothing that reaches the tile
  // has its centre above fine bin o*(jt+127) + Pb, the Doppler cell grows with the wavenumber and the half-size
  // with the cell, so the cell of that wavenumber bounds every line of (state, isotope) near the tile
  {
    int64_t ftop = (int64_t)o * (jt + ACC_TILE - 1) + Pb + 1;
    if (ftop > onwn - 1) ftop = onwn - 1;
    const double wtop = tb_fine_wn(wn_i, odwn, ftop);
    for (int i = tid; i < FBK * niso; i += FUSE_NT) {
      const int k = i / niso, iso = i % niso;
      if (s0 + k >= ns) continue;
      const int id = tb_binsearchapprox(sh_aDop, __dmul_rn(sh_tab[i].alphad, wtop), 0, nDop);
      const int row = nrows <= 16 ? iso_row[iso] : 0;
      // ... and the lines that do not refresh the Doppler index take the carried cell (extinction.c:480-483)
      const int pc = CARRIED ? ptab[sh_fbpid[i]].psize : 0, pr_ = ptab[id * nLor + sh_ilor[i]].psize;
      atomicMax(&sh_Prow[row], pc > pr_ ? pc : pr_);
    }
  }
  // ---- this thread's output bin and its four states
  const int jr = tid & (ACC_TILE - 1);
  const int64_t j = jt + jr;
  const int half = tid >> 7;                       // states FPT*half .. FPT*half+FPT-1
  const bool inb = j < nwn;
  const int oj = (int)(o * (inb ? j : nwn - 1));
  double acc[FPT];
  unsigned cnt[FPT];
#pragma unroll
  for (int kk = 0; kk < FPT; kk++) { acc[kk] = 0.0; cnt[kk] = 0u; }
  const unsigned nsmask = ns - s0 >= FBK ? (1u << FBK) - 1u : (ns - s0 > 0 ? (1u << (ns - s0)) - 1u : 0u);      // states that exist
  int buf = 0;
  __syncthreads();
  for (int row = 0; row < nrows; row++) {
    const int P = nrows <= 16 ? sh_Prow[row] : Pb;   // more than 16 line-list molecules: the batch-wide bound
    const int R = P / o + 1;
    const int64_t jlo = jt - R < 0 ? 0 : jt - R;
    const int64_t jhi = jt + ACC_TILE + R > nwn ? nwn : jt + ACC_TILE + R;
    const int pa = binstart[(size_t)row * nwn + jlo], pb = binstart[(size_t)row * nwn + jhi];
    const int lo_f = oj - P, hi_f = oj + P;
    int p = pa;
    while (p < pb) {
      // ---- phase 1: candidates 256 at a time through the log-domain pretest, the groups alive in any state
      // appended IN ORDER to the chunk's list until it is (nearly) full
      const int chunk0 = p;
      int nalive = 0;
      while (p < pb && nalive <= FUSE_CAPA - FUSE_NT) {
        const int q = p + tid;
        unsigned m8 = 0;
        if (q < pb) {
          const double a0 = g_a0[q], lg = g_lg[q];
          const FuseTab *tb = sh_tab + g_iso[q];
#pragma unroll
          for (int k = 0; k < FBK; k++)
            if (!(fma(a0, sh_rT[k], lg + tb[k * niso].lcz) < -1e-9)) m8 |= 1u << k;
          m8 &= nsmask;
        }
        const unsigned ball = __ballot_sync(0xffffffffu, m8 != 0);
        if (lane == 0) sh_wcnt[buf][warp] = __popc(ball);
        __syncthreads();
        int base = nalive, tot = 0;
#pragma unroll
        for (int w = 0; w < FUSE_NT / 32; w++) {
          const int cw = sh_wcnt[buf][w];
          if (w < warp) base += cw;
          tot += cw;
        }
        if (m8) {
          const int c = base + __popc(ball & ((1u << lane) - 1));
          sh_gidx[c] = q - chunk0;
          sh_mask[c] = (unsigned char)m8;
        }
        nalive += tot;
        p += FUSE_NT;
        buf ^= 1;
      }
      __syncthreads();
      // ---- phase 2: one thread per alive group, a loop over its alive states
      for (int c = tid; c < nalive; c += FUSE_NT) {
        const int q = chunk0 + sh_gidx[c];
        const unsigned m8 = sh_mask[c];
        const int iso = g_iso[q];
        const int b = g_start[q], e = g_start[q + 1];
        const double wlead = g_wavn[q], gf = g_gf[q], a0 = g_a0[q];
        const int iown = g_iown[q];
        const double wneg = __dmul_rn(-TB_EXPCTE, wlead);
        const bool home = rec_out && fd_div(iown, fd_tile) == tile;   // trace: each group is written by its own tile
        sh_iown[c] = iown;
        int lo = -1;                                               // Doppler cell of the previous state: the search resumes there
#pragma unroll 1
        for (int k = 0; k < FBK; k++) {
          FuseRec rr;
          rr.k = 0.0; rr.A = 0; rr.jl = 1; rr.jh = 0;                // dead unless proven otherwise
          if ((m8 >> k) & 1u) {
            const int s = s0 + k;
            const double rT = sh_rT[k];
            const FuseTab t = sh_tab[k * niso + iso];
            // the first co-added member's term is requested before the two exp(): its L2 round trip hides behind them
            const double *xt = xterm + (size_t)s * X + (b - q);     // member l of the group: xt[l - b - 1]
            const double x1 = e - b > 1 ? __ldg(xt) : 0.0;
            const double ge1 = __dmul_rn(gf, tb_exp(a0 * rT));
            double kk = 0.0;
            bool below = (e - b == 1) && ge1 * t.facz < 0.999999 * t.thr;
            if (!below) {
              double a = __dmul_rn(ge1, (1 - tb_exp(wneg * rT)));
              if (e - b > 1) {                                      // co-added members, in order (extinction.c:449-462)
                a = __dadd_rn(a, x1);
                for (int l = b + 2; l < e; l++) a = __dadd_rn(a, xt[l - b - 1]);
              }
              kk = __dmul_rn(a, t.facz);
              below = kk < t.thr;                                  // extinction.c:467-470
            }
            if (!below) {
              if (!permol) kk = __dmul_rn(kk, t.densf);
              const double adw = __dmul_rn(t.alphad, wlead);
              unsigned pid;
              if (!CARRIED || tb_refresh(adw, t.alphal, t.ral)) {   // extinction.c:480-483
                // binsearchapprox (iomisc.c:1088-1108) on the log-spaced table, as in k_strength
                if (lo < 0) {
                  lo = (int)((__log2f((float)adw) * 0.30102999566f - dop_lmin) * dop_rstep);
                  lo = lo < 0 ? 0 : (lo > nDop - 1 ? nDop - 1 : lo);
                }
                while (lo > 0 && sh_aDop[lo] > adw) lo--;
                while (sh_aDop[lo + 1] <= adw) lo++;
                const int id = (fabs(sh_aDop[lo + 1] - adw) < fabs(sh_aDop[lo] - adw)) ? lo + 1 : lo;
                pid = (unsigned)(id * nLor + sh_ilor[k * niso + iso]);
              } else
                pid = (unsigned)sh_fbpid[k * niso + iso];          // the index carried from the last line that refreshed it
              const ProfEntry pe = ptab[pid];
              const int ps = pe.psize;
              // bins with |o*j - iown| <= ps: ceil((iown-ps)/o) .. floor((iown+ps)/o), on the grid
              const int lof = iown - ps;
              int jl = lof <= 0 ? 0 : fd_div(lof + o - 1, fd_o);
              int jh = fd_div(iown + ps, fd_o);
              if (jh > (int)(nwn - 1)) jh = (int)(nwn - 1);
              const bool reach = jl <= jh;                          // else: covers no grid point (nothing to add)
              if (reach) {
                // sample index of bin jl: i0 = o*jl - iown + ps, in [0, o) unless jl was clamped to the grid start;
                // bin j then reads sample i0 + o*(j - jl): polyphase address off + (i0 % o)*nk + i0/o + (j - jl)
                int i0 = o * jl - lof, m = 0;
                if (i0 >= o) { m = fd_div(i0, fd_o); i0 -= m * o; }
                jl -= (int)jt; jh -= (int)jt;
                rr.k = kk;
                rr.A = pe.off + i0 * pe.nk + m - jl;                // + tile-relative bin
                rr.jl = (short)(jl < 0 ? 0 : jl);
                rr.jh = (short)(jh > ACC_TILE - 1 ? ACC_TILE - 1 : jh);
              }
              if (home) {
                LineRec r;
                r.k = kk; r.iown = iown; r.pid = reach ? pid : (pid | PID_NOREACH);
                rec_out[(size_t)s * G + q] = r;
              }
            }
          }
          sh_rec[k * FUSE_CAPA + c] = rr;
        }
      }
      __syncthreads();
      // ---- phase 3: the bin's window of the shared records, in fine-bin order
      if (inb && nalive > 0) {
        int lo = 0, hi = nalive;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (sh_iown[mid] < lo_f) lo = mid + 1; else hi = mid;
        }
        const FuseRec *rp = sh_rec + (size_t)(FPT * half) * FUSE_CAPA;
        for (int c = lo; c < nalive && sh_iown[c] <= hi_f; c++) {
#pragma unroll
          for (int kk = 0; kk < FPT; kk++) {
            const int4 v = *reinterpret_cast<const int4 *>(rp + kk * FUSE_CAPA + c);
            const int jl = (int)(short)(v.w & 0xFFFF), jh = v.w >> 16;
            if (jr >= jl && jr <= jh) {
              acc[kk] = fma(__hiloint2double(v.y, v.x), pdata[v.z + jr], acc[kk]);
              cnt[kk]++;
            }
          }
        }
      }
      // the next chunk's phase 1 passes at least one barrier before its phase 2 rewrites the records
    }
    if (permol) {
#pragma unroll
      for (int kk = 0; kk < FPT; kk++) {
        const int s = s0 + FPT * half + kk;
        if (inb && s < ns) out[((size_t)s * nrows + row) * nwn + j] = acc[kk];
        acc[kk] = 0.0;
      }
    }
  }
  if (!permol) {
#pragma unroll
    for (int kk = 0; kk < FPT; kk++) {
      const int s = s0 + FPT * half + kk;
      if (inb && s < ns) out[(size_t)s * nwn + j] = acc[kk];
    }
  }
  // sample counters (integer atomics): one add per warp and state
#pragma unroll
  for (int kk = 0; kk < FPT; kk++) {
    unsigned v = cnt[kk];
    for (int dlt = 16; dlt > 0; dlt >>= 1) v += __shfl_xor_sync(0xffffffffu, v, dlt);
    const int s = s0 + FPT * half + kk;
    if (lane == 0 && v && s < ns) atomicAdd(&nsamp[s], (unsigned long long)v);
  }
}

// ---- the same line sum with WARP-AUTONOMOUS phases.  k_linesum_fused spends a third of its stall samples at its
// two block barriers (profiles/r2_prof_fused_summary.txt): 256 threads wait for the slowest warp's candidate loads
// in phase 1 and for the ragged last pass of phase 2.  Here a warp owns 32 output bins x 4 states and runs the three
// phases on its own window of groups with __syncwarp only; a block is 8 such warps = one 128-bin tile x one batch
// of FB = 8 states (warps 0-3: states 0-3, warps 4-7: states 4-7), which share the per-(state, isotope) tables.
//   phase 1  32 candidates per step through the log-domain pretest; the survivors join a QUEUE (ordered);
//   phase 2  as soon as the queue holds 32 groups (or the window is exhausted) one lane per group evaluates them --
//            always a full pass, the rest of the queue waits for the next round -- into 32 shared records per state;
//   phase 3  lane = (bin mod 8, state): the round's 32 records span a bin or two, so the bins they can touch (first
//            fine bin - P .. last + P) are a handful; each lane walks the 32 records of its state for its bin(s) of
//            that range: 128-bit shared read, "covers my bin" test, table read + FMA.  Four accumulators per lane
//            (bins bq, bq+8, bq+16, bq+24 of the unit).
// The price is the halo: a 32-bin unit re-evaluates the groups within the widest profile of its edges (window
// 32 + 2R bins instead of 128 + 2R), a few per cent where the profiles reach +-1..2 bins.  Each bin still sums its
// records over rows and sorted records: the same FMA chain, the same bits as k_linesum_fused.
#define FW_Q 64              // queue entries per warp (a round consumes 32)
#define FW_NW 8              // warps per block
template <bool CARRIED>
__global__ void __launch_bounds__(32 * FW_NW, 4)
k_linesum_fused_w(int64_t nwn, int64_t onwn, int64_t G, int ns, int niso, int nmol, int o, FastDiv fd_o, FastDiv fd_tile,
                  int nrows, int permol, int tile_off, double wn_i, double odwn,
                  const int32_t *__restrict__ fclass, const int32_t *__restrict__ fP,
                  const int32_t *__restrict__ binstart, const int32_t *__restrict__ g_start,
                  const int16_t *__restrict__ g_iso, const int32_t *__restrict__ g_iown,
                  const double *__restrict__ g_wavn, const double *__restrict__ g_a0,
                  const double *__restrict__ g_gf, const double *__restrict__ g_lg,
                  const double *__restrict__ rtemp, const double *__restrict__ facz, const double *__restrict__ lcz,
                  const double *__restrict__ xterm, int64_t X, float dop_lmin, float dop_rstep,
                  const double *__restrict__ dens, const int *__restrict__ iso_imol, const int *__restrict__ iso_row,
                  const double *__restrict__ alphad, const int32_t *__restrict__ ilor,
                  const double *__restrict__ alphal, const double *__restrict__ ral, const int32_t *__restrict__ fbpid,
                  const double *__restrict__ aDop, int nDop, const ProfEntry *__restrict__ ptab, int nLor,
                  double ethresh, const double *__restrict__ kmax, const double *__restrict__ pdata,
                  double *__restrict__ out, unsigned long long *__restrict__ nsamp, LineRec *__restrict__ rec_out) {
  const int batch = blockIdx.y;
  if (!fclass[batch]) return;                      // k_strength + k_accumulate take this batch
  extern __shared__ __align__(16) unsigned char fsm[];
  FuseRec *sh_rec_all = reinterpret_cast<FuseRec *>(fsm);                          // [FW_NW][FBK][32]
  int *sh_q_all = reinterpret_cast<int *>(sh_rec_all + FW_NW * FBK * 32);           // [FW_NW][FW_Q] group index
  int *sh_io_all = sh_q_all + FW_NW * FW_Q;                                        // [FW_NW][32] fine bins of the round
  double *sh_aDop = reinterpret_cast<double *>(sh_io_all + FW_NW * 32);            // [256]
  double *sh_rT = sh_aDop + 256;                                                   // [FB]
  FuseTab *sh_tab = reinterpret_cast<FuseTab *>(sh_rT + FB);                       // [FB][niso]
  int *sh_ilor = reinterpret_cast<int *>(sh_tab + FB * niso);                      // [FB][niso]
  int *sh_fbpid = sh_ilor + FB * niso;                                             // [FB][niso] carried cell
  int *sh_lo0 = sh_fbpid + FB * niso;                                              // [FB][niso] Doppler cell below the tile's window
  unsigned char *sh_m_all = reinterpret_cast<unsigned char *>(sh_lo0 + FB * niso);   // [FW_NW][FW_Q] alive states
  __shared__ int sh_Prow[16];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int sb = batch * FB;                       // first state of the batch
  const int tile = blockIdx.x + tile_off;
  const int Pb = fP[batch];
  for (int i = tid; i <= nDop; i += 32 * FW_NW) sh_aDop[i] = aDop[i];              // [nDop] = +inf sentinel
  if (tid < FB) sh_rT[tid] = rtemp[sb + tid < ns ? sb + tid : ns - 1];
  if (tid < 16) sh_Prow[tid] = 0;
  for (int i = tid; i < FB * niso; i += 32 * FW_NW) {
    const int k = i / niso, iso = i % niso, s = sb + k < ns ? sb + k : ns - 1;
    const int si = s * niso + iso;
    FuseTab t;
    t.lcz = lcz[si]; t.facz = facz[si];
    t.thr = __dmul_rn(ethresh, kmax[s * nrows + (permol ? iso_row[iso] : 0)]);
    t.densf = dens[s * nmol + iso_imol[iso]];
    t.alphad = alphad[si]; t.alphal = alphal[si]; t.ral = ral[si];
    sh_tab[i] = t;
    sh_ilor[i] = ilor[si];
    sh_fbpid[i] = CARRIED ? fbpid[si] : 0;
  }
  __syncthreads();
  {  // the widest profile a line of each molecule row can take in this tile (see k_linesum_fused), over the batch
    int64_t ftop = (int64_t)o * ((int64_t)tile * ACC_TILE + ACC_TILE - 1) + Pb + 1;
    if (ftop > onwn - 1) ftop = onwn - 1;
    const double wtop = tb_fine_wn(wn_i, odwn, ftop);
    for (int i = tid; i < FB * niso; i += 32 * FW_NW) {
      const int k = i / niso, iso = i % niso;
      if (sb + k >= ns) continue;
      const int id = tb_binsearchapprox(sh_aDop, __dmul_rn(sh_tab[i].alphad, wtop), 0, nDop);
      const int row = nrows <= 16 ? iso_row[iso] : 0;
      const int pc = CARRIED ? ptab[sh_fbpid[i]].psize : 0, pr_ = ptab[id * nLor + sh_ilor[i]].psize;
      atomicMax(&sh_Prow[row], pc > pr_ ? pc : pr_);
      // ... and the last cell at or below the Doppler width of the LOWEST wavenumber any candidate of the tile can
      // have (tile start - the batch-wide reach - one bin): phase 2's search starts there and only walks up
      int64_t fbot = (int64_t)o * ((int64_t)tile * ACC_TILE - 1) - Pb - 1;
      if (fbot < 0) fbot = 0;
      const double abot = __dmul_rn(sh_tab[i].alphad, tb_fine_wn(wn_i, odwn, fbot));
      int l0 = 0, l1 = nDop - 1;                   // largest l in [0, nDop-1] with aDop[l] <= abot (0 if none)
      while (l0 < l1) {
        const int mid = (l0 + l1 + 1) >> 1;
        if (sh_aDop[mid] <= abot) l0 = mid; else l1 = mid - 1;
      }
      sh_lo0[i] = l0;
    }
  }
  __syncthreads();                                 // the last block-wide barrier
  // ---- this warp's unit: bins jt .. jt+31 of the tile, states s0 .. s0+3
  const int quad = warp >> 2;                      // 0: states 0-3 of the batch, 1: states 4-7
  const int s0 = sb + quad * FBK;
  const int64_t jt = (int64_t)tile * ACC_TILE + (warp & 3) * 32;
  if (jt >= nwn || s0 >= ns) return;
  FuseRec *sh_rec = sh_rec_all + (size_t)warp * FBK * 32;
  int *sh_q = sh_q_all + warp * FW_Q, *sh_io = sh_io_all + warp * 32;
  unsigned char *sh_m = sh_m_all + warp * FW_Q;
  const FuseTab *tab = sh_tab + (size_t)quad * FBK * niso;
  const int *qilor = sh_ilor + quad * FBK * niso, *qfbpid = sh_fbpid + quad * FBK * niso, *qlo0 = sh_lo0 + quad * FBK * niso;
  const double *qrT = sh_rT + quad * FBK;
  const int myk = lane & 3, bq = lane >> 2;        // phase 3: this lane's state and its bin modulo 8
  const FuseRec *myrec = sh_rec + myk * 32;
  const int maxbin = (int)(nwn - 1 - jt < 31 ? nwn - 1 - jt : 31);
  double acc[4];
  unsigned cnt = 0u;
#pragma unroll
  for (int i = 0; i < 4; i++) acc[i] = 0.0;
  const unsigned nsmask = ns - s0 >= FBK ? (1u << FBK) - 1u : (1u << (ns - s0)) - 1u;      // states that exist
  for (int row = 0; row < nrows; row++) {
    const int P = nrows <= 16 ? sh_Prow[row] : Pb;
    const int R = P / o + 1;
    const int64_t jlo = jt - R < 0 ? 0 : jt - R;
    const int64_t jhi = jt + 32 + R > nwn ? nwn : jt + 32 + R;
    const int pa = binstart[(size_t)row * nwn + jlo], pb = binstart[(size_t)row * nwn + jhi];
    int p = pa, nq = 0;
    while (true) {
      // ---- phase 1: fill the queue
      while (p < pb && nq < 32) {
        const int q = p + lane;
        unsigned m8 = 0;
        if (q < pb) {
          const double a0 = g_a0[q], lg = g_lg[q];
          const FuseTab *tb = tab + g_iso[q];
#pragma unroll
          for (int k = 0; k < FBK; k++)
            if (!(fma(a0, qrT[k], lg + tb[k * niso].lcz) < -1e-9)) m8 |= 1u << k;
          m8 &= nsmask;
        }
        const unsigned ball = __ballot_sync(0xffffffffu, m8 != 0);
        if (m8) {
          const int c = nq + __popc(ball & ((1u << lane) - 1));
          sh_q[c] = q;
          sh_m[c] = (unsigned char)m8;
        }
        nq += __popc(ball);
        p += 32;
      }
      if (nq == 0) break;
      __syncwarp();
      const int n = nq < 32 ? nq : 32;
      // ---- phase 2: one lane per queued group, a loop over its alive states (k_linesum_fused's, verbatim)
      int iown = 0;
      if (lane < n) {
        const int q = sh_q[lane];
        const unsigned m8 = sh_m[lane];
        const int iso = g_iso[q];
        const int b = g_start[q], e = g_start[q + 1];
        const double wlead = g_wavn[q], gf = g_gf[q], a0 = g_a0[q];
        iown = g_iown[q];
        const double wneg = __dmul_rn(-TB_EXPCTE, wlead);
        const int gq = fd_div(iown, fd_o), gr = iown - gq * o;      // iown = o*gq + gr: the bin ranges below need no division
        const bool home = rec_out && gq >= (int)jt && gq < (int)jt + 32;   // trace: written by the group's own unit
#pragma unroll 1
        for (int k = 0; k < FBK; k++) {
          FuseRec rr;
          rr.k = 0.0; rr.A = 0; rr.jl = 1; rr.jh = 0;                // dead unless proven otherwise
          if ((m8 >> k) & 1u) {
            const int s = s0 + k;
            const double rT = qrT[k];
            const FuseTab t = tab[k * niso + iso];
            const double *xt = xterm + (size_t)s * X + (b - q);
            const double x1 = e - b > 1 ? __ldg(xt) : 0.0;
            const double ge1 = __dmul_rn(gf, tb_exp(a0 * rT));
            double kk = 0.0;
            bool below = (e - b == 1) && ge1 * t.facz < 0.999999 * t.thr;
            if (!below) {
              double a = __dmul_rn(ge1, (1 - tb_exp(wneg * rT)));
              if (e - b > 1) {                                      // co-added members, in order (extinction.c:449-462)
                a = __dadd_rn(a, x1);
                for (int l = b + 2; l < e; l++) a = __dadd_rn(a, xt[l - b - 1]);
              }
              kk = __dmul_rn(a, t.facz);
              below = kk < t.thr;                                  // extinction.c:467-470
            }
            if (!below) {
              if (!permol) kk = __dmul_rn(kk, t.densf);
              const double adw = __dmul_rn(t.alphad, wlead);
              unsigned pid;
              if (!CARRIED || tb_refresh(adw, t.alphal, t.ral)) {   // extinction.c:480-483
                // binsearchapprox (iomisc.c:1088-1108): the last cell at or below adw, found upwards from the tile's
                // start cell (the Doppler width grows with the wavenumber), then the nearer of it and the next
                int lo = qlo0[k * niso + iso];
                while (sh_aDop[lo + 1] <= adw) lo++;
                const int id = (fabs(sh_aDop[lo + 1] - adw) < fabs(sh_aDop[lo] - adw)) ? lo + 1 : lo;
                pid = (unsigned)(id * nLor + qilor[k * niso + iso]);
              } else
                pid = (unsigned)qfbpid[k * niso + iso];
              const ProfEntry pe = ptab[pid];
              const int ps = pe.psize;
              const int lof = iown - ps;
              // ps = o*pe.pad + pr: ceil((iown - ps)/o) = gq - pad + (gr > pr), floor((iown + ps)/o) = gq + pad + (gr + pr >= o)
              const int pr = ps - pe.pad * o;
              int jl = lof <= 0 ? 0 : gq - pe.pad + (gr > pr ? 1 : 0);
              int jh = gq + pe.pad + (gr + pr >= o ? 1 : 0);
              if (jh > (int)(nwn - 1)) jh = (int)(nwn - 1);
              const bool reach = jl <= jh;
              if (reach) {
                int i0 = o * jl - lof, m = 0;
                if (i0 >= o) { m = fd_div(i0, fd_o); i0 -= m * o; }    // only where jl was clamped to the grid start
                jl -= (int)jt; jh -= (int)jt;
                rr.k = kk;
                rr.A = pe.off + i0 * pe.nk + m - jl;                // + unit-relative bin
                rr.jl = (short)(jl < 0 ? 0 : jl);
                rr.jh = (short)(jh > 31 ? 31 : jh);
              }
              if (home) {
                LineRec r;
                r.k = kk; r.iown = iown; r.pid = reach ? pid : (pid | PID_NOREACH);
                rec_out[(size_t)s * G + q] = r;
              }
            }
          }
          sh_rec[k * 32 + lane] = rr;
        }
      }
      sh_io[lane] = lane < n ? iown : 0x7fffffff;
      // bins the round can touch: first fine bin - P .. last fine bin + P (the queue is in fine-bin order)
      const int io_min = __shfl_sync(0xffffffffu, iown, 0), io_max = __shfl_sync(0xffffffffu, iown, n - 1);
      int b_lo = io_min - P <= 0 ? 0 : fd_div(io_min - P, fd_o) - (int)jt;
      int b_hi = fd_div(io_max + P, fd_o) - (int)jt;
      if (b_lo < 0) b_lo = 0;
      if (b_hi > maxbin) b_hi = maxbin;
      // the part of the queue that waits for the next round moves to the front
      int nxt_q = 0;
      unsigned char nxt_m = 0;
      if (lane + 32 < nq) { nxt_q = sh_q[lane + 32]; nxt_m = sh_m[lane + 32]; }
      __syncwarp();
      if (lane + 32 < nq) { sh_q[lane] = nxt_q; sh_m[lane] = nxt_m; }
      nq -= n;
      // ---- phase 3: lane = (bin mod 8, state) over the round's records
#pragma unroll
      for (int i = 0; i < 4; i++) {
        if (8 * i + 7 < b_lo || 8 * i > b_hi) continue;             // warp-uniform
        const int bin = 8 * i + bq;
        if (bin >= b_lo && bin <= b_hi) {
          const int ob = o * ((int)jt + bin), lo_f = ob - P, hi_f = ob + P;
          int c = 0;                                                // first record of the bin's window: 5 bisection steps
#pragma unroll
          for (int st = 16; st > 0; st >>= 1)
            if (sh_io[c + st - 1] < lo_f) c += st;
          int ce = 0;                                               // one past its last record
#pragma unroll
          for (int st = 16; st > 0; st >>= 1)
            if (sh_io[ce + st - 1] <= hi_f) ce += st;
          if (ce == 31 && sh_io[31] <= hi_f) ce = 32;
          // two records per step: both table reads are in flight before the first FMA waits for its sample
          for (; c < ce; c += 2) {
            const int4 v0 = *reinterpret_cast<const int4 *>(myrec + c);
            const int4 v1 = *reinterpret_cast<const int4 *>(myrec + (c + 1 < ce ? c + 1 : c));
            const bool h0 = bin >= (int)(short)(v0.w & 0xFFFF) && bin <= (v0.w >> 16);
            const bool h1 = c + 1 < ce && bin >= (int)(short)(v1.w & 0xFFFF) && bin <= (v1.w >> 16);
            const double p0 = h0 ? pdata[v0.z + bin] : 0.0, p1 = h1 ? pdata[v1.z + bin] : 0.0;
            if (h0) { acc[i] = fma(__hiloint2double(v0.y, v0.x), p0, acc[i]); cnt++; }
            if (h1) { acc[i] = fma(__hiloint2double(v1.y, v1.x), p1, acc[i]); cnt++; }
          }
        }
      }
      __syncwarp();                                // the next round rewrites the records
    }
    if (permol) {
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int s = s0 + myk;
        if (8 * i + bq <= maxbin && s < ns) out[((size_t)s * nrows + row) * nwn + jt + 8 * i + bq] = acc[i];
        acc[i] = 0.0;
      }
    }
  }
  if (!permol) {
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int s = s0 + myk;
      if (8 * i + bq <= maxbin && s < ns) out[(size_t)s * nwn + jt + 8 * i + bq] = acc[i];
    }
  }
  // sample counters (integer atomics): lanes of equal state add up, one add per warp and state
  for (int dlt = 4; dlt < 32; dlt <<= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, dlt);
  if (lane < 4 && cnt && s0 + lane < ns) atomicAdd(&nsamp[s0 + lane], (unsigned long long)cnt);
}

static size_t fusew_shared_bytes(int niso) {
  return (size_t)FW_NW * FBK * 32 * sizeof(FuseRec) + (size_t)FW_NW * FW_Q * 4 + (size_t)FW_NW * 32 * 4 + 256 * 8 + FB * 8 +
         (size_t)FB * niso * sizeof(FuseTab) + (size_t)FB * niso * 12 + (size_t)FW_NW * FW_Q + 16;
}

static size_t fuse_shared_bytes(int niso) {
  return (size_t)FBK * FUSE_CAPA * sizeof(FuseRec) + (size_t)FUSE_CAPA * 4 * 2 + 256 * 8 + FBK * 8 +
         (size_t)FBK * niso * sizeof(FuseTab) + (size_t)FBK * niso * 8 + FUSE_CAPA + 16;
}


// ------------------------------------------------------------------ host driver

int tb_states_run(tb_ctx *c, int ns, const double *T, const double *dens, const double *Z,
                  double ethresh, int permol, double *d_out_rows, double *h_out) {
  TB_REQUIRE(c->have_lines && c->have_prof, TB_ERR_STATE,
             "tb_upload_lines and tb_build_profiles must precede the line sum");
  TB_REQUIRE(c->niso < 0x7FFF, TB_ERR_ARG, "too many isotopes (%d)", c->niso);
  TB_REQUIRE(c->nDop * c->nLor < (int)PID_NOREACH, TB_ERR_ARG, "Voigt table has too many cells (%d)", c->nDop * c->nLor);
  TB_CUDA(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int niso = c->niso, nmol = c->nmol, nrows = c->nrows;
  const int nout = permol ? nrows : 1;
  const int64_t G = c->G, M = c->M, nwn = c->nwn;
  const int ntiles = (int)((nwn + ACC_TILE - 1) / ACC_TILE);
  c->last_ns = ns;
  c->last_permol = permol;
  // chunk the states so the per-(state, group) scratch stays below 48 GB of the 180 GB
  int chunk = ns;
  const size_t per_state = (size_t)G * 16 + (size_t)(M - G) * 8 + (size_t)nout * nwn * 8 + 64;
  const size_t budget = (size_t)48 << 30;
  if (per_state * chunk > budget) chunk = (int)(budget / per_state);
  if (chunk < 1) chunk = 1;
  if (chunk > 65535) chunk = 65535;            // gridDim.y
  TB_TRY(c->s_T.ensure(chunk)); TB_TRY(c->s_dens.ensure((size_t)chunk * nmol));
  TB_TRY(c->s_Z.ensure((size_t)chunk * (niso ? niso : 1)));
  TB_TRY(c->s_alphal.ensure((size_t)chunk * niso + 1)); TB_TRY(c->s_alphad.ensure((size_t)chunk * niso + 1));
  TB_TRY(c->s_ilor.ensure((size_t)chunk * niso + 1)); TB_TRY(c->s_idop0.ensure((size_t)chunk * niso + 1));
  TB_TRY(c->s_lastq.ensure((size_t)chunk * niso + 1)); TB_TRY(c->s_fb.ensure((size_t)chunk * niso + 1));
  TB_TRY(c->s_fbpid.ensure((size_t)chunk * niso + 1)); TB_TRY(c->s_fbmax.ensure(chunk));
  TB_TRY(c->s_wtile.ensure((size_t)chunk * ntiles)); TB_TRY(c->s_reach.ensure((size_t)chunk * ntiles));
  TB_TRY(c->s_reachP.ensure((size_t)chunk * ntiles));
  TB_TRY(c->s_facz.ensure((size_t)chunk * niso + 1)); TB_TRY(c->s_rtemp.ensure(chunk));
  TB_TRY(c->s_ral.ensure((size_t)chunk * niso + 1)); TB_TRY(c->s_lcz.ensure((size_t)chunk * niso + 1));
  const int64_t X = M - G;
  TB_TRY(c->s_xterm.ensure((size_t)chunk * X + 1));
  TB_TRY(c->s_hasfb.ensure((size_t)chunk * ntiles));
  TB_TRY(c->s_kmax.ensure((size_t)chunk * nrows + 1));
  TB_TRY(c->s_nsamp.ensure(ns));
  TB_TRY(c->s_rec.ensure((size_t)chunk * G + 1));
  const int nbatch_max = (chunk + FB - 1) / FB;
  TB_TRY(c->s_fclass.ensure(nbatch_max)); TB_TRY(c->s_fP.ensure(nbatch_max));
  const size_t fuse_shm = fuse_shared_bytes(niso), fusew_shm = fusew_shared_bytes(niso);
  // a tile of 128 bins x 4 states is the fused kernel's unit of parallelism: a grid of a few dozen tiles (C1: 20)
  // cannot fill 148 SMs with it (measured: C1 1.29 -> 2.9 ms per step), so small grids keep the two-kernel path
  const bool fuse_on = c->fuse_enabled && niso >= 1 && niso <= FUSE_MAXISO && G > 0 && fuse_shm <= 100 * 1024 && fusew_shm <= 100 * 1024 &&
                       (ntiles >= FUSE_MIN_TILES || c->fuse_enabled > 1);
  if (fuse_on && c->fuse_shm_set != (int)fuse_shm) {
    TB_CUDA(cudaFuncSetAttribute(k_linesum_fused<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fuse_shm));
    TB_CUDA(cudaFuncSetAttribute(k_linesum_fused<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fuse_shm));
    TB_CUDA(cudaFuncSetAttribute(k_linesum_fused_w<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fusew_shm));
    TB_CUDA(cudaFuncSetAttribute(k_linesum_fused_w<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fusew_shm));
    c->fuse_shm_set = (int)fuse_shm;
  }
  double *d_out = d_out_rows;
  if (!d_out) {
    TB_TRY(c->s_out.ensure((size_t)chunk * nout * nwn));
    d_out = c->s_out.p;
  }
  const int Tn = 256;
  constexpr int SB = 4;      // states per block of k_kmax's seed launch (few lines: more blocks)
  constexpr int SBB = 10;    // ... and of the bulk launch (line records read once per SBB states)
  constexpr int SBS = 8;     // states per block of k_strength (group records are read once per SBS states)
  auto fastdiv = [](int64_t d) {
    FastDiv f;
    f.d = (int)d;
    f.M = d == 1 ? 0 : (~0ULL) / (uint64_t)d + 1;
    return f;
  };
  const FastDiv fd_o = fastdiv(c->o), fd_tile = fastdiv((int64_t)c->o * ACC_TILE);
  TB_REQUIRE(c->nDop < 255, TB_ERR_ARG, "nDop must be < 255");
  // widest reach any profile of the table can have, in coarse bins and in accumulate tiles
  int64_t max_psize = 0;
  for (size_t i = 0; i < c->psize.size(); i++) max_psize = c->psize[i] > max_psize ? c->psize[i] : max_psize;
  const int64_t Rmax = 2 + max_psize / c->o;
  const int utiles = (int)((Rmax + ACC_TILE - 1) / ACC_TILE) + 1;
  // wavenumber-band shard (tb_set_shard): which tiles this context accumulates and which groups
  // can reach them
  const bool banded = c->banded;
  const int64_t b_lo = banded ? c->band_lo : 0, b_hi = banded ? c->band_hi : nwn;
  const int tile_lo = (int)(b_lo / ACC_TILE), tile_hi = (int)((b_hi + ACC_TILE - 1) / ACC_TILE);
  std::vector<int64_t> pa(nrows, 0), pb(nrows, 0);
  if (banded && G > 0) {
    if (c->h_binstart.empty()) {
      c->h_binstart.resize((size_t)nrows * nwn + 1);
      TB_CUDA(cudaMemcpyAsync(c->h_binstart.data(), c->binstart.p, c->h_binstart.size() * sizeof(int32_t),
                              cudaMemcpyDeviceToHost, st));
      TB_CUDA(cudaStreamSynchronize(st));
    }
    const int64_t jl = b_lo - Rmax < 0 ? 0 : b_lo - Rmax, jh = b_hi + Rmax > nwn ? nwn : b_hi + Rmax;
    for (int r = 0; r < nrows && b_hi > b_lo; r++) {      // an empty band needs no strengths at all
      pa[r] = c->h_binstart[(size_t)r * nwn + jl];
      pb[r] = c->h_binstart[(size_t)r * nwn + jh];
    }
  }
  TB_REQUIRE(c->shard_world == 1 || c->exch || c->comm, TB_ERR_STATE,
             "sharded context without an exchange callback or a communicator (tb_comm_init)");
  tb_timer_begin(c, "molext");
  for (int s0 = 0; s0 < ns; s0 += chunk) {
    const int cs = (ns - s0 < chunk) ? ns - s0 : chunk;
    TB_TRY(tb_upload_small(c, c->s_T.p, T + s0, cs * sizeof(double), st));
    TB_TRY(tb_upload_small(c, c->s_dens.p, dens + (size_t)s0 * nmol, (size_t)cs * nmol * sizeof(double), st));
    if (niso) TB_TRY(tb_upload_small(c, c->s_Z.p, Z + (size_t)s0 * niso, (size_t)cs * niso * sizeof(double), st));
    double *dst = d_out_rows ? d_out + (size_t)s0 * nout * nwn : d_out;
    int nset = cs * (niso > nrows ? niso : nrows);
    if (nset < cs) nset = cs;
    TB_CUDA(cudaMemsetAsync(c->s_fbmax.p, 0, cs * sizeof(int32_t), st));
    k_state_setup<<<nblk(nset, 128), 128, 0, st>>>(
        cs, niso, nmol, c->s_T.p, c->s_dens.p, c->d_mol_mass.p, c->d_mol_radius.p, c->d_iso_mass.p,
        c->d_iso_imol.p, c->d_aDop.p, c->nDop, c->d_aLor.p, c->nLor, c->wn_i, c->s_alphal.p,
        c->s_alphad.p, c->s_ilor.p, c->s_idop0.p, c->s_lastq.p, c->s_kmax.p, nrows,
        c->s_nsamp.p + s0, c->s_wtile.p, c->s_hasfb.p, ntiles, c->s_Z.p, c->d_iso_ratio.p, c->s_facz.p,
        c->s_rtemp.p, c->s_ral.p);
    TB_LAUNCHED(c);
    if (banded)    // bins outside the band come back as zero
      TB_CUDA(cudaMemsetAsync(dst, 0, (size_t)cs * nout * nwn * sizeof(double), st));
    const int nbatch = (cs + FB - 1) / FB;
    k_fuse_class<<<nbatch, 128, 0, st>>>(cs, niso, c->wn_i, c->own_last, c->o, fuse_on ? 1 : 0, (c->fuse_carried && c->og_contig) ? 0 : 1, c->fuse_rmax,
                                         c->s_alphad.p, c->s_ral.p,
                                         c->s_ilor.p, c->d_aDop.p, c->nDop, c->nLor, c->d_ptab.p, c->s_fclass.p, c->s_fP.p);
    TB_LAUNCHED(c);
    c->last_chunk_s0 = s0; c->last_chunk_cs = cs;
    if (G > 0) {
      // lines in descending order of their state-independent factor: a small seed launch finds
      // (nearly) the maxima, after which the bulk skips both exp() of almost every line
      const int64_t Mq = c->Mq, seed = Mq < 16384 ? Mq : 16384;
      tb_timer_begin(c, "kmax");
      if (seed > 0)
        // (one line per thread: the seed is 16 384 lines, eight per thread would leave it 200 blocks for 148 SMs)
        k_kmax<SB, 1><<<dim3(nblk(seed, Tn), (cs + SB - 1) / SB), Tn, 0, st>>>(
            0, seed, cs, niso, c->q_g.p, c->q_a.p, c->q_w.p, c->q_iso.p, c->s_rtemp.p, c->s_Z.p, c->d_iso_mass.p,
            c->d_iso_row.p, permol, nrows, c->s_kmax.p, c->kmax_block_skip ? 1 : 0);
      // the bulk is split over the ranks of a sharded run (every rank ran the seed, so every
      // rank skips alike); the exact maximum is completed by the all-reduce below
      // (with the block skip of k_kmax the whole bulk costs ~0.04 ms: a rank of a library-sharded run then walks it all
      // itself and needs no all-reduce -- one rendez-vous of the ranks less per spectrum, the same exact maxima)
      const bool kmax_local = c->shard_world > 1 && !c->exch && c->kmax_block_skip;
      const int kw = kmax_local ? 1 : c->shard_world, kr = kmax_local ? 0 : c->shard_rank;
      const int64_t per = (Mq - seed + kw - 1) / kw;
      const int64_t q0 = seed + per * kr, q1 = q0 + per < Mq ? q0 + per : Mq;
      if (q1 > q0)
        k_kmax<SBB><<<dim3(nblk(q1 - q0, Tn * KMAX_LPT), (cs + SBB - 1) / SBB), Tn, 0, st>>>(
            q0, q1, cs, niso, c->q_g.p, c->q_a.p, c->q_w.p, c->q_iso.p, c->s_rtemp.p, c->s_Z.p, c->d_iso_mass.p,
            c->d_iso_row.p, permol, nrows, c->s_kmax.p, c->kmax_block_skip ? 1 : 0);
      TB_LAUNCHED(c);
      if (c->shard_world > 1 && !kmax_local) {
        if (c->exch)
          TB_REQUIRE(c->exch(c->exch_user, TB_EXCH_KMAX, c->s_kmax.p, (int64_t)cs * nrows) == 0, TB_ERR_STATE,
                     "exchange callback failed (strength maxima)");
        else TB_TRY(tb_comm_allreduce_max(c, TB_EXCH_KMAX, c->s_kmax.p, (int64_t)cs * nrows));
      }
      tb_timer_end(c, "kmax");
      tb_timer_begin(c, "strength");
      if (X > 0 && !banded) {
        k_extra<<<dim3(nblk(X, Tn), cs), Tn, 0, st>>>(X, 0, X, cs, c->x_line.p, c->m_wavn.p, c->m_elow.p, c->m_gf.p,
                                                      c->s_rtemp.p, c->efct, c->s_xterm.p);
        TB_LAUNCHED(c);
      } else if (X > 0) {
        // a band shard: the members of the groups [pa, pb) of every row are x in [g_start[pa] - pa, g_start[pb] - pb)
        // (the four numbers per row are read back once per band table)
        std::vector<int64_t> key;
        for (int r = 0; r < nrows; r++) { key.push_back(pa[r]); key.push_back(pb[r]); }
        if (key != c->h_xr_key) {
          std::vector<int32_t> gs(2 * (size_t)nrows, 0);
          for (int r = 0; r < nrows; r++) {
            TB_CUDA(cudaMemcpyAsync(&gs[2 * r], c->g_start.p + pa[r], sizeof(int32_t), cudaMemcpyDeviceToHost, st));
            TB_CUDA(cudaMemcpyAsync(&gs[2 * r + 1], c->g_start.p + pb[r], sizeof(int32_t), cudaMemcpyDeviceToHost, st));
          }
          TB_CUDA(cudaStreamSynchronize(st));
          c->h_xr.assign(2 * (size_t)nrows, 0);
          for (int r = 0; r < nrows; r++) { c->h_xr[2 * r] = gs[2 * r] - pa[r]; c->h_xr[2 * r + 1] = gs[2 * r + 1] - pb[r]; }
          c->h_xr_key = key;
        }
        for (int r = 0; r < nrows; r++) {
          const int64_t x0 = c->h_xr[2 * r], x1 = c->h_xr[2 * r + 1];
          if (x1 <= x0) continue;
          k_extra<<<dim3(nblk(x1 - x0, Tn), cs), Tn, 0, st>>>(X, x0, x1, cs, c->x_line.p, c->m_wavn.p, c->m_elow.p, c->m_gf.p,
                                                            c->s_rtemp.p, c->efct, c->s_xterm.p);
          TB_LAUNCHED(c);
        }
      }
      k_logthr<<<nblk((int64_t)cs * niso, 128), 128, 0, st>>>(cs, niso, nrows, permol, ethresh, c->d_iso_row.p,
                                                              c->s_facz.p, c->s_kmax.p, c->s_lcz.p);
      TB_LAUNCHED(c);
      // carried Doppler index by bisection + backward scan of the file order (k_lastq_scan) instead of maxima over the
      // evaluated groups: for the fused states that keep one, and -- in a band-sharded run with the exchange in the
      // library -- for EVERY state: the scan reads the whole line index, so each rank finds the global index by
      // itself and the all-reduce of the carried index (one collective, one rendez-vous of the ranks) is not needed
      const bool scan_all = c->shard_world > 1 && !c->exch && c->og_contig;
      const bool scan_fused = fuse_on && c->fuse_carried && c->og_contig;
      if (scan_all || scan_fused) {
        TB_TRY(c->s_lastq2.ensure((size_t)chunk * niso + 1));
        TB_CUDA(cudaMemsetAsync(c->s_lastq2.p, 0xff, (size_t)cs * niso * sizeof(int32_t), st));
        k_lastq_scan<<<dim3(cs * niso, LASTQ_NB), 256, 0, st>>>(
            cs, niso, nrows, permol, ethresh, c->wn_i, c->iso_og0.p, c->iso_og1.p, c->go_pos.p, c->go_wavn.p, c->g_start.p,
            c->g_gf.p, c->g_a0.p, c->g_lg.p, c->g_wavn.p, c->s_rtemp.p, c->s_facz.p, c->s_lcz.p, c->s_kmax.p,
            c->d_iso_row.p, c->s_alphal.p, c->s_alphad.p, c->s_ral.p, c->x_line.p, c->m_wavn.p, c->m_elow.p, c->m_gf.p,
            c->efct, c->s_fclass.p, scan_all ? 1 : 0,
            c->s_lastq2.p);
        TB_LAUNCHED(c);
      }
      if (fuse_on && tile_hi > tile_lo) {
        tb_timer_end(c, "strength");
        tb_timer_begin(c, "fused");
        if (scan_fused) {      // (else only always-refreshing states are fused: their carried index is never read)
          k_fallback<<<nblk(cs * niso, 128), 128, 0, st>>>(
              cs, niso, c->s_fclass.p, 1, c->s_lastq2.p, c->s_idop0.p, c->go_wavn.p, c->s_alphad.p, c->s_ilor.p, c->d_aDop.p,
              c->nDop, c->nLor, c->d_ptab.p, c->s_fb.p, c->s_fbpid.p, c->s_fbmax.p);
          TB_LAUNCHED(c);
        }
        LineRec *trace_rec = nullptr;
        if (c->trace_enabled) {                 // per-line trace wanted: records of the fused states too
          k_fuse_trace_fill<<<dim3(nblk(G, Tn), cs), Tn, 0, st>>>(G, cs, c->s_fclass.p, c->g_iown.p, c->s_rec.p);
          trace_rec = c->s_rec.p;
        }
        const bool carried = scan_fused;
        auto fused_kernel = c->fuse_variant ? (carried ? k_linesum_fused_w<true> : k_linesum_fused_w<false>)
                                            : (carried ? k_linesum_fused<true> : k_linesum_fused<false>);
        fused_kernel<<<dim3(tile_hi - tile_lo, c->fuse_variant ? nbatch : nbatch * (FB / FBK)),
                       c->fuse_variant ? 32 * FW_NW : FUSE_NT, c->fuse_variant ? fusew_shm : fuse_shm, st>>>(
            nwn, c->onwn, G, cs, niso, nmol, c->o, fd_o, fd_tile, nrows, permol, tile_lo, c->wn_i, c->odwn,
            c->s_fclass.p, c->s_fP.p, c->binstart.p,
            c->g_start.p, c->g_iso.p, c->g_iown.p, c->g_wavn.p, c->g_a0.p, c->g_gf.p, c->g_lg.p, c->s_rtemp.p,
            c->s_facz.p, c->s_lcz.p, c->s_xterm.p, X, (float)c->dop_lmin, (float)c->dop_rstep, c->s_dens.p,
            c->d_iso_imol.p, c->d_iso_row.p, c->s_alphad.p, c->s_ilor.p, c->s_alphal.p, c->s_ral.p, c->s_fbpid.p,
            c->d_aDop.p, c->nDop, c->d_ptab.p, c->nLor,
            ethresh, c->s_kmax.p, c->d_pdata2.p, dst, c->s_nsamp.p + s0, trace_rec);
        TB_LAUNCHED(c);
        tb_timer_end(c, "fused");
        tb_timer_begin(c, "strength");
      }
      for (int r0 = 0; r0 < (banded ? nrows : 1); r0 += STR_NR) {
        StrRanges rg;
        rg.n = 0; rg.blk0[0] = 0;
        for (int r = r0; r < (banded ? nrows : 1) && rg.n < STR_NR; r++) {
          const int64_t p_off = banded ? pa[r] : 0, p_end = banded ? pb[r] : G;
          if (p_end <= p_off) continue;
          rg.p_off[rg.n] = p_off; rg.p_end[rg.n] = p_end;
          rg.blk0[rg.n + 1] = rg.blk0[rg.n] + nblk(p_end - p_off, Tn);
          rg.n++;
        }
        if (rg.n == 0) continue;
        dim3 g3(rg.blk0[rg.n], (cs + SBS - 1) / SBS);
      k_strength<SBS><<<g3, Tn, 0, st>>>(
          G, rg, cs, niso, c->o, fd_tile, fd_o, c->g_start.p, c->g_iso.p, c->g_orig.p, c->g_iown.p, c->g_wavn.p,
          c->g_a0.p, c->g_gf.p, c->g_lg.p, c->s_rtemp.p, c->s_facz.p, c->s_lcz.p, c->s_ral.p, c->s_xterm.p, X,
          (float)c->dop_lmin, (float)c->dop_rstep, c->s_dens.p, nmol,
          c->d_iso_imol.p, c->d_iso_row.p, c->s_alphal.p, c->s_alphad.p, c->s_ilor.p,
          c->d_aDop.p, c->nDop, c->d_ptab.p, c->nLor, ethresh, permol, nrows, c->s_kmax.p,
          c->s_rec.p, c->s_lastq.p, c->s_wtile.p, c->s_hasfb.p, ntiles, c->s_fclass.p);
      }
      if (c->shard_world > 1 && !scan_all) {
        if (c->exch)
          TB_REQUIRE(c->exch(c->exch_user, TB_EXCH_LASTQ, c->s_lastq.p, (int64_t)cs * niso) == 0, TB_ERR_STATE,
                     "exchange callback failed (carried Doppler index)");
        else TB_TRY(tb_comm_allreduce_max(c, TB_EXCH_LASTQ, c->s_lastq.p, (int64_t)cs * niso));
      }
      tb_timer_end(c, "strength");
      k_fallback<<<nblk(cs * niso, 128), 128, 0, st>>>(
          cs, niso, c->s_fclass.p, 0, scan_all ? c->s_lastq2.p : c->s_lastq.p, c->s_idop0.p, c->go_wavn.p, c->s_alphad.p, c->s_ilor.p, c->d_aDop.p,
          c->nDop, c->nLor, c->d_ptab.p, c->s_fb.p, c->s_fbpid.p, c->s_fbmax.p);
      c->launches += 3;
    }
    k_reach<<<nblk((int64_t)cs * ntiles, 128), 128, 0, st>>>(cs, ntiles, c->o, utiles, c->s_wtile.p, c->s_hasfb.p,
                                                             c->s_fbmax.p, c->s_reach.p, c->s_reachP.p);
    dim3 g4(tile_hi - tile_lo, cs);
    // A grid of few tiles leaves the wide-reach instance one block per (tile, deep layer): the deepest
    // layers of C1 (20 tiles) each ran 0.7 ms on 20 SMs.  There S blocks share a tile.  S depends on
    // the WHOLE grid, the chunk of states and the row count only, never on the band of a shard, so a
    // banded run adds in the same order as a single GPU.
    int S = 1;
    if (ntiles < 128) {
      S = (256 + ntiles - 1) / ntiles;
      if (S > ACC_SPLIT_MAX) S = ACC_SPLIT_MAX;
      if ((size_t)S * chunk * nout * ntiles * ACC_TILE * sizeof(double) > ((size_t)1 << 30)) S = 1;
    }
    if (S > 1) {
      TB_TRY(c->s_part.ensure((size_t)S * chunk * nout * ntiles * ACC_TILE));
      TB_TRY(c->s_tcnt.ensure((size_t)chunk * nout * ntiles));
      TB_CUDA(cudaMemsetAsync(c->s_tcnt.p, 0, (size_t)cs * nout * ntiles * sizeof(unsigned), st));
    }
    tb_timer_begin(c, "accumulate");
    if (tile_hi > tile_lo) {
      k_accumulate<true><<<dim3((tile_hi - tile_lo) * S, cs), 32 * ACC_WW, 0, st>>>(
                                          nwn, G, cs, niso, c->o, fd_o, nrows, permol, ntiles, tile_lo, c->binstart.p,
                                          c->substart.p, c->nsub, c->nsube, c->s_rec.p, c->s_fbpid.p, c->d_ptab.p,
                                          c->s_reach.p, c->s_reachP.p, c->d_pdata2.p, dst,
                                          c->s_nsamp.p + s0, S, c->s_part.p, c->s_tcnt.p, c->s_fclass.p, c->narrow_r);
      k_accumulate<false><<<g4, ACC_TILE, 0, st>>>(nwn, G, cs, niso, c->o, fd_o, nrows, permol, ntiles, tile_lo, c->binstart.p,
                                          c->substart.p, c->nsub, c->nsube, c->s_rec.p, c->s_fbpid.p, c->d_ptab.p,
                                          c->s_reach.p, c->s_reachP.p, c->d_pdata2.p, dst,
                                          c->s_nsamp.p + s0, 1, nullptr, nullptr, c->s_fclass.p, c->narrow_r);
    }
    tb_timer_end(c, "accumulate");
    c->launches += 3;
    TB_CUDA(cudaGetLastError());
    if (h_out)
      TB_CUDA(cudaMemcpyAsync(h_out + (size_t)s0 * nout * nwn, dst, (size_t)cs * nout * nwn * sizeof(double),
                              cudaMemcpyDeviceToHost, st));
    // the state scratch is reused by the next chunk: the host arrays staged above must be consumed
    if (s0 + cs < ns) TB_CUDA(cudaStreamSynchronize(st));
  }
  tb_timer_end(c, "molext");
  if (h_out) TB_CUDA(cudaStreamSynchronize(st));
  c->nsamp_valid = false;
  return TB_OK;
}

extern "C" int tb_molext_states(tb_ctx *c, int ns, const double *T, const double *density,
                                const double *Z, double ethresh, int permol, double *out) {
  TB_REQUIRE(ns > 0 && T && density && out, TB_ERR_ARG, "tb_molext_states: bad arguments");
  return tb_states_run(c, ns, T, density, Z, ethresh, permol, nullptr, out);
}

extern "C" int tb_extinction_lbl(tb_ctx *c, double ethresh) {
  TB_REQUIRE(c->have_atm, TB_ERR_STATE, "tb_set_atmosphere not called");
  const int nrad = c->nrad, nmol = c->nmol, niso = c->niso;
  std::vector<double> T(nrad), dens((size_t)nrad * nmol), Z((size_t)nrad * (niso ? niso : 1));
  for (int r = 0; r < nrad; r++) {
    T[r] = c->t[r] * c->t_fct;                               // tau.c:169,259
    for (int j = 0; j < nmol; j++) dens[(size_t)r * nmol + j] = c->dens[(size_t)j * nrad + r];
    for (int i = 0; i < niso; i++) Z[(size_t)r * niso + i] = c->Z[(size_t)i * nrad + r];
  }
  TB_TRY(c->d_e.ensure((size_t)nrad * c->nwn));
  TB_TRY(tb_states_run(c, nrad, T.data(), dens.data(), Z.data(), ethresh, 0, c->d_e.p, nullptr));
  c->have_e = true;
  return TB_OK;
}

extern "C" int tb_get_sample_counts(tb_ctx *c, int64_t *nsamples, int ns) {
  TB_REQUIRE(ns <= c->last_ns, TB_ERR_ARG, "only %d states in the last batch", c->last_ns);
  TB_CUDA(cudaSetDevice(c->device));
  std::vector<unsigned long long> h(ns);
  TB_CUDA(cudaMemcpyAsync(h.data(), c->s_nsamp.p, ns * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  for (int i = 0; i < ns; i++) nsamples[i] = (int64_t)h[i];
  return TB_OK;
}

// per-line trace in original order, for bit-exact index parity against the oracle
__global__ void k_trace(int64_t n, int64_t G, int s_local, int niso, const uint8_t *__restrict__ flag,
                        const int32_t *__restrict__ l_iown, const int32_t *__restrict__ l_gpos,
                        const int32_t *__restrict__ g_idwn, const int16_t *__restrict__ g_iso,
                        const LineRec *__restrict__ recs, int nLor,
                        const int32_t *__restrict__ fb, int32_t *__restrict__ iown,
                        int32_t *__restrict__ idwn, int32_t *__restrict__ idop, int8_t *__restrict__ status) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint8_t f = flag[i];
  int8_t st = 0;
  int32_t io = 0, id = 0, dp = -1;
  if (f == 3) {
    int32_t p = l_gpos[i];
    io = l_iown[i];
    id = g_idwn[p];
    const unsigned pid = recs[(size_t)s_local * G + p].pid & 0xFFFFu;
    if (pid == PID_SKIP) st = 3;
    else {
      st = 1;
      dp = pid >= PID_FB ? fb[s_local * niso + g_iso[p]] : (int)(pid & (PID_NOREACH - 1)) / nLor;
    }
  } else if (!(f & 2)) st = 2;     // visited only as a co-added member
  iown[i] = io; idwn[i] = id; idop[i] = dp; status[i] = st;
}

extern "C" int tb_get_line_trace(tb_ctx *c, int s, int32_t *iown, int32_t *idwn, int32_t *idop,
                                 int8_t *status) {
  TB_REQUIRE(c->have_lines && c->last_ns > 0, TB_ERR_STATE, "no line-sum batch has run");
  TB_REQUIRE(s >= 0 && s < c->last_ns, TB_ERR_ARG, "state %d out of range", s);
  TB_REQUIRE((size_t)c->last_ns * c->G + 1 <= c->s_rec.n, TB_ERR_STATE,
             "trace is only available when the batch ran as a single chunk");
  TB_CUDA(cudaSetDevice(c->device));
  const int64_t n = c->nlines;
  if (n == 0) return TB_OK;
  if (!c->trace_enabled && c->G > 0) {          // the fused kernel keeps its records on chip unless asked
    int32_t fc = 0;
    TB_CUDA(cudaMemcpyAsync(&fc, c->s_fclass.p + s / FB, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    TB_CUDA(cudaStreamSynchronize(c->stream));
    TB_REQUIRE(fc == 0, TB_ERR_STATE, "state %d went through the fused line sum, which writes per-line records only "
               "when asked: call tb_set_trace(c, 1) before the batch", s);
  }
  DevBuf<int32_t> a, b, d;
  DevBuf<int8_t> e;
  TB_TRY(a.ensure(n)); TB_TRY(b.ensure(n)); TB_TRY(d.ensure(n)); TB_TRY(e.ensure(n));
  k_trace<<<nblk(n, 256), 256, 0, c->stream>>>(n, c->G, s, c->niso, c->l_flag.p, c->l_iown.p,
                                               c->l_gpos.p, c->g_idwn.p, c->g_iso.p,
                                               c->s_rec.p, c->nLor, c->s_fb.p, a.p, b.p, d.p, e.p);
  TB_CUDA(cudaMemcpyAsync(iown, a.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
  TB_CUDA(cudaMemcpyAsync(idwn, b.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
  TB_CUDA(cudaMemcpyAsync(idop, d.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
  TB_CUDA(cudaMemcpyAsync(status, e.p, n, cudaMemcpyDeviceToHost, c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  a.release(); b.release(); d.release(); e.release();
  return TB_OK;
}

// Lorentz-width index of every isotope at state `s` of the last batch (extinction.c:394): with the trace's
// Doppler index it names the Voigt cell, hence profsize, minj and maxj of extinction.c:486-496
extern "C" int tb_get_state_ilor(tb_ctx *c, int s, int32_t *ilor) {
  TB_REQUIRE(c->have_lines && c->last_ns > 0, TB_ERR_STATE, "no line-sum batch has run");
  TB_REQUIRE(s >= 0 && s < c->last_ns, TB_ERR_ARG, "state %d out of range", s);
  TB_REQUIRE((size_t)c->last_ns * c->niso + 1 <= c->s_ilor.n, TB_ERR_STATE,
             "only available when the batch ran as a single chunk");
  TB_CUDA(cudaSetDevice(c->device));
  if (c->niso == 0) return TB_OK;
  TB_CUDA(cudaMemcpyAsync(ilor, c->s_ilor.p + (size_t)s * c->niso, c->niso * sizeof(int32_t), cudaMemcpyDeviceToHost,
                          c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  return TB_OK;
}

// Per-line records of the fused states (test tap; costs 16 B of HBM per (group, state) while on)
extern "C" int tb_set_trace(tb_ctx *c, int on) {
  c->trace_enabled = on != 0;
  return TB_OK;
}
// 0: every state through k_strength + k_accumulate (the two-kernel path); 1 (default): narrow states fused
extern "C" int tb_set_fused(tb_ctx *c, int on) {
  c->fuse_enabled = on < 0 ? 0 : (on > 2 ? 2 : on);     // 2: also on grids of fewer than FUSE_MIN_TILES tiles
  return TB_OK;
}
// how many states of the last batch the fused kernel took, and the profile updates it executed
extern "C" int tb_get_fused_stats(tb_ctx *c, int64_t *fused_states, int64_t *fused_samples) {
  TB_REQUIRE(c->last_ns > 0, TB_ERR_STATE, "no line-sum batch has run");
  TB_CUDA(cudaSetDevice(c->device));
  const int cs = c->last_chunk_cs, s0 = c->last_chunk_s0, nb = (cs + FB - 1) / FB;
  std::vector<int32_t> fc(nb > 0 ? nb : 1, 0);
  std::vector<unsigned long long> h(cs > 0 ? cs : 1, 0);
  if (nb > 0) TB_CUDA(cudaMemcpyAsync(fc.data(), c->s_fclass.p, nb * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
  if (cs > 0) TB_CUDA(cudaMemcpyAsync(h.data(), c->s_nsamp.p + s0, cs * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  int64_t nsf = 0, smp = 0;
  for (int s = 0; s < cs; s++)
    if (fc[s / FB]) { nsf++; smp += (int64_t)h[s]; }
  if (fused_states) *fused_states = nsf;
  if (fused_samples) *fused_samples = smp;
  return TB_OK;
}


// Which states the fused kernel takes (tuning / comparison tap): widest reach in coarse bins (default FUSE_RMAX) and
// whether states whose lines partly keep a CARRIED Doppler index are fused too (default no: measured on C2, the
// two-kernel path is as fast on those layers -- 25 us against 30 us per layer -- and k_lastq_scan costs 35 us).
// Which fused kernel: 0 = k_linesum_fused (a block per 128 bins x 4 states, block barriers between the phases),
// 1 = k_linesum_fused_w (a warp per 32 bins x 4 states, no block barrier).  Same sums, bit for bit.
extern "C" int tb_set_fused_variant(tb_ctx *c, int v) {
  TB_REQUIRE(v == 0 || v == 1, TB_ERR_ARG, "tb_set_fused_variant: 0 or 1");
  c->fuse_variant = v;
  return TB_OK;
}

// Widest reach, in coarse bins, of a tile summed by the one-thread-per-bin schedule of k_accumulate (default NARROW_R)
extern "C" int tb_set_narrow_reach(tb_ctx *c, int r) {
  TB_REQUIRE(r >= 0 && r <= 4096, TB_ERR_ARG, "tb_set_narrow_reach: 0..4096 coarse bins");
  c->narrow_r = r;
  return TB_OK;
}

extern "C" int tb_set_fused_reach(tb_ctx *c, int rmax, int carried) {
  TB_REQUIRE(rmax >= 1 && rmax <= 4096, TB_ERR_ARG, "tb_set_fused_reach: 1..4096 coarse bins");
  c->fuse_rmax = rmax;
  c->fuse_carried = carried != 0;
  return TB_OK;
}
