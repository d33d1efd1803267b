// tb_voigt.cu -- the Voigt-profile table: calcprofiles (transit/src/opacity.c:218-277) ->
// getprofile (extinction.c:8-57) -> voigtn / voigtxy / meanintegSimp|Trap
// (pu/src/voigt.c:132-200,369-554).
//
// One launch builds every unique profile: a thread owns one output bin, evaluates the
// Pierluissi approximation at the bin's fine sub-samples and reduces them in the
// reference's float operation order (odd points, doubled, even points, doubled, ends).
// voigtxy runs in FP64 here (the reference uses x87 long double and rounds to float).
#include "tb_internal.cuh"

struct ProfDesc {
  int64_t off;      // first output float
  int64_t blk0;     // first block of this profile
  int32_t nwn;      // 2*psize+1 output bins
  int32_t last;     // fine intervals per bin (ipo-1); 0 in point-sample (QUICK) mode
  int32_t mode;     // 0 point samples, 1 Simpson, 2 trapezoid
  int32_t pad;
  double dwn;       // half width handed to voigtn
  double dint;      // fine spacing
  double alphaD, alphaL;
};

__constant__ double c_ferf[32];  // 1/(n!(2n+1)), voigt.c:46-108 (n <= 29 is ever used)

static const double h_ferf[32] = {
    1.000000000000000000000,    0.333333333333333333333,    0.100000000000000000000,
    2.38095238095238095238e-2,  4.62962962962962962963e-3,  7.57575757575757575758e-4,
    1.06837606837606837607e-4,  1.32275132275132275132e-5,  1.45891690009337068161e-6,
    1.45038522231504687645e-7,  1.31225329638028050726e-8,  1.08922210371485733805e-9,
    8.35070279514723959168e-11, 5.94779401363763503681e-12, 3.95542951645852576340e-13,
    2.46682701026445692771e-14, 1.44832646435981372650e-15, 8.03273501241577360914e-17,
    4.22140728880708823303e-18, 2.10785519144213582486e-19, 1.00251649349077191670e-20,
    4.55184675892820028624e-22, 1.97706475387790517483e-23, 8.23014929921422135684e-25,
    3.28926034917575173275e-26, 1.26410789889891635220e-27, 4.67848351551848577373e-29,
    1.66976179341737202699e-30, 5.75419164398217177220e-32, 1.91694286210978253077e-33,
    6.18030758822279613746e-35, 1.93035720881510785656e-36};

// voigt.c:132-200
__device__ float tb_voigtxy(double x, double y, double alphaD) {
  const double SQRTLN2PI = 0.46971863934982566689, TWOOSQRTPI = 1.12837916709551257389;
  const double x2y2 = __dsub_rn(__dmul_rn(x, x), __dmul_rn(y, y));
  const double xy2 = __dmul_rn(__dmul_rn(2.0, x), y);
  if (x < 3 && y < 1.8) {  // Region I: series for w(z)
    double sn, cs;
    sincos(xy2, &sn, &cs);
    const int n = (x < 1 ? 15 : (int)(6.842 * x + 8.0)) + 1;
    double orr = y, oi = -x, ar = y, ai = -x;
    for (int i = 1; i <= n; i++) {
      double ni = orr * xy2 + oi * x2y2;
      double nr = orr * x2y2 - oi * xy2;
      ai += ni * c_ferf[i];
      ar += nr * c_ferf[i];
      oi = ni;
      orr = nr;
    }
    return (float)(SQRTLN2PI / alphaD * exp(-x2y2) *
                   (cs * (1 - ar * TWOOSQRTPI) - sn * ai * TWOOSQRTPI));
  } else if (x < 5 && y < 5) {  // Region II
    const double A1 = 0.46131350, A2 = 0.19016350, A3 = 0.09999216, A4 = 1.78449270,
                 A5 = 0.002883894, A6 = 5.52534370;
    double ar = xy2 * xy2, nr = xy2 * x, ni = x2y2 - A2, ai = x2y2 - A4, oi = x2y2 - A6;
    return (float)(SQRTLN2PI / alphaD *
                   (A1 * ((nr - ni * y) / (ni * ni + ar)) + A3 * ((nr - ai * y) / (ai * ai + ar)) +
                    A5 * ((nr - oi * y) / (oi * oi + ar))));
  } else {  // Region III
    const double B1 = 0.51242424, B2 = 0.27525510, B3 = 0.05176536, B4 = 2.72474500;
    double ar = xy2 * xy2, nr = xy2 * x, ni = x2y2 - B2, ai = x2y2 - B4;
    return (float)(SQRTLN2PI / alphaD *
                   (B1 * ((nr - ni * y) / (ni * ni + ar)) + B3 * ((nr - ai * y) / (ai * ai + ar))));
  }
}

__global__ void k_voigt_table(const ProfDesc *__restrict__ desc, int nprof, float *__restrict__ out) {
  // locate this block's profile (blk0 is ascending)
  int lo = 0, hi = nprof - 1;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (desc[mid].blk0 <= (int64_t)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const ProfDesc d = desc[lo];
  const int64_t k = ((int64_t)blockIdx.x - d.blk0) * blockDim.x + threadIdx.x;
  if (k >= d.nwn) return;
  const double y = TB_SQRTLN2 * d.alphaL / d.alphaD;
  auto sample = [&](int64_t i) -> float {   // voigt.c:448-451
    double x = TB_SQRTLN2 * fabs(__dsub_rn(__dmul_rn(d.dint, (double)i), d.dwn)) / d.alphaD;
    return tb_voigtxy(x, y, d.alphaD);
  };
  float res;
  if (d.mode == 0) {
    res = sample(k);                         // VOIGT_QUICK, voigt.c:456-459
  } else {
    const int last = d.last;
    const int64_t base = k * last;
    float acc = 0.f;
    if (d.mode == 1) {                       // meanintegSimp, voigt.c:489-521
      for (int i = 1; i < last; i += 2) acc = __fadd_rn(acc, sample(base + i));
      acc = __fmul_rn(acc, 2.f);
      for (int i = 2; i < last; i += 2) acc = __fadd_rn(acc, sample(base + i));
      acc = __fmul_rn(acc, 2.f);
      acc = __fadd_rn(acc, __fadd_rn(sample(base), sample(base + last)));
      res = (float)((double)acc / (last * 3.0));
    } else {                                 // meanintegTrap, voigt.c:527-554
      for (int i = 1; i < last; i++) acc = __fadd_rn(acc, sample(base + i));
      float ends = __fadd_rn(sample(base), sample(base + last));
      res = (float)(((double)acc + (double)ends / 2.0) / (double)last);
    }
  }
  out[d.off + k] = res;
}

// reference layout -> polyphase layout (see ProfEntry)
__global__ void k_polyphase(const ProfDesc *__restrict__ desc, const int32_t *__restrict__ off2,
                            int nprof, int o, const float *__restrict__ in, double *__restrict__ out) {
  int lo = 0, hi = nprof - 1;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (desc[mid].blk0 <= (int64_t)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const ProfDesc d = desc[lo];
  const int64_t k = ((int64_t)blockIdx.x - d.blk0) * blockDim.x + threadIdx.x;
  if (k >= d.nwn) return;
  const int nk = (d.nwn - 1) / o + 1;
  out[off2[lo] + (k % o) * (int64_t)nk + k / o] = (double)in[d.off + k];
}

// getprofile's size rule (extinction.c:21-35) with the reference's float narrowing of the widths
static int64_t prof_half_size(double dwn, float dop, float lor, float ta, int64_t nwave) {
  double bigalpha = dop;
  if (bigalpha < lor) bigalpha = lor;
  double wvgt = bigalpha * ta;
  int nvgt = 2 * (long)(wvgt / dwn + 0.5) + 1;
  if (nvgt < 2) nvgt = 3;
  if (nvgt > 2 * (int)nwave) nvgt = 2 * (int)nwave + 1;
  return nvgt / 2;
}

extern "C" int tb_build_profiles(tb_ctx *c, int nDop, int nLor, float dmin, float dmax, float lmin,
                                 float lmax, float nwidth) {
  TB_REQUIRE(c->have_grid, TB_ERR_STATE, "tb_set_wn_grid must precede tb_build_profiles");
  // the line sum packs a cell into 14 bits of a record and keeps aDop[] in 256 shared doubles: refuse here,
  // before a table of hundreds of MB is built, what tb_states_run could not use
  TB_REQUIRE(nDop >= 2 && nLor >= 2 && nDop < 255 && (int64_t)nDop * nLor < 0x4000, TB_ERR_ARG,
             "need 2 <= nDop < 255, nLor >= 2 and nDop*nLor < 16384 (got %d x %d)", nDop, nLor);
  TB_REQUIRE(dmin > 0 && dmax > dmin && lmin > 0 && lmax > lmin && nwidth > 0, TB_ERR_ARG,
             "bad width ranges");
  TB_CUDA(cudaSetDevice(c->device));
  c->nDop = nDop; c->nLor = nLor; c->nwidth = nwidth;
  // logspace (iomisc.c:1064-1083); one +inf sentinel past the end because binsearchapprox is
  // called with hi = n (extinction.c:393-394): a width beyond the table clamps to the last entry
  auto logspace = [](double mn, double mx, int n, std::vector<double> &v) {
    double lmn = log10(mn), lmx = log10(mx), step = (lmx - lmn) / (n - 1.0);
    v.resize(n + 1);
    for (int i = 0; i < n; i++) v[i] = pow(10, lmn + i * step);
    v[n] = INFINITY;
  };
  logspace(dmin, dmax, nDop, c->aDop);
  c->dop_lmin = log10((double)dmin);
  c->dop_rstep = (nDop - 1.0) / (log10((double)dmax) - log10((double)dmin));
  logspace(lmin, lmax, nLor, c->aLor);
  c->psize.assign((size_t)nDop * nLor, 0);
  c->poff.assign((size_t)nDop * nLor, 0);
  std::vector<ProfDesc> desc;
  const int TPB = 128;
  int64_t total = 0, blocks = 0;
  for (int i = 0; i < nDop; i++)
    for (int j = 0; j < nLor; j++) {
      size_t cell = (size_t)i * nLor + j;
      if (c->aDop[i] * 10.0 < c->aLor[j] && i != 0) {   // opacity.c:262-265: share row i-1
        c->psize[cell] = c->psize[cell - nLor];
        c->poff[cell] = c->poff[cell - nLor];
        continue;
      }
      const float dop = (float)c->aDop[i], lor = (float)c->aLor[j];
      int64_t ps = prof_half_size(c->odwn, dop, lor, nwidth, c->onwn);
      c->psize[cell] = ps;
      c->poff[cell] = total;
      ProfDesc d;
      d.off = total;
      d.blk0 = blocks;
      d.nwn = (int32_t)(2 * ps + 1);
      d.alphaD = dop;
      d.alphaL = lor;
      d.dwn = c->odwn * (long)(d.nwn / 2);
      // voigtn's choice of sub-sampling (voigt.c:399-433)
      const bool quick = d.nwn > 99999;
      double ddwn = 2.0 * d.dwn / (d.nwn - 1);
      int nint = 50;
      double dint = d.alphaD / (nint - 1);
      if (ddwn < dint || quick) {
        dint = ddwn;
        nint = d.nwn + 1;
      } else {
        nint = (int)(ddwn / dint) + 1;
        if (nint & 1) nint++;
        nint = d.nwn * nint + 1;
        dint = 2.0 * d.dwn / (nint - 1);
      }
      d.dint = dint;
      if (quick) {
        d.mode = 0;
        d.last = 0;
      } else {
        int ipo = (int)((float)(nint - 1) / d.nwn) + 1;
        d.last = ipo - 1;
        d.mode = (ipo & 1) ? 1 : 2;
      }
      d.pad = 0;
      desc.push_back(d);
      total += d.nwn;
      blocks += (d.nwn + TPB - 1) / TPB;
    }
  c->ndata = total;
  // polyphase copy: min(o, n) phases of nk = (n-1)/o + 1 samples each per unique profile
  std::vector<int32_t> off2(desc.size());
  std::vector<ProfEntry> ptab((size_t)nDop * nLor);
  int64_t total2 = TB_PROF_PAD;   // zero padding either side: the warp path reads lanes it then discards
  {
    std::map<int64_t, size_t> by_off;
    for (size_t k = 0; k < desc.size(); k++) {
      const int64_t n = desc[k].nwn, nk = (n - 1) / c->o + 1, nph = n < c->o ? n : c->o;
      TB_REQUIRE(total2 < 2000000000LL, TB_ERR_ARG, "Voigt table too large for 32-bit offsets");
      off2[k] = (int32_t)total2;
      by_off[desc[k].off] = k;
      total2 += nk * nph;
    }
    TB_REQUIRE(total2 < 2000000000LL, TB_ERR_ARG, "Voigt table too large for 32-bit offsets");
    for (size_t cell = 0; cell < ptab.size(); cell++) {
      const size_t k = by_off[c->poff[cell]];
      ptab[cell].off = off2[k];
      ptab[cell].psize = (int32_t)c->psize[cell];
      ptab[cell].nk = (int32_t)((2 * c->psize[cell]) / c->o + 1);
      ptab[cell].pad = (int32_t)(c->psize[cell] / c->o);      // psize = o*pad + remainder: k_linesum_fused_w's bin range without a division
    }
  }
  TB_REQUIRE(blocks < 2147483647LL, TB_ERR_ARG, "Voigt table too large for one launch");
  std::vector<int32_t> ps32(c->psize.begin(), c->psize.end());
  TB_TRY(c->d_aDop.ensure(nDop + 1)); TB_TRY(c->d_aLor.ensure(nLor + 1));
  TB_TRY(c->d_psize.ensure(ps32.size())); TB_TRY(c->d_poff.ensure(c->poff.size()));
  TB_TRY(c->d_pdata.ensure(total));
  TB_TRY(c->d_pdata2.ensure(total2 + TB_PROF_PAD));
  TB_TRY(c->d_ptab.ensure(ptab.size()));
  DevBuf<ProfDesc> d_desc;
  DevBuf<int32_t> d_off2;
  TB_TRY(d_desc.ensure(desc.size()));
  TB_TRY(d_off2.ensure(off2.size()));
  cudaStream_t st = c->stream;
  TB_CUDA(cudaMemcpyToSymbolAsync(c_ferf, h_ferf, sizeof(h_ferf), 0, cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(c->d_aDop.p, c->aDop.data(), (nDop + 1) * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(c->d_aLor.p, c->aLor.data(), (nLor + 1) * sizeof(double), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(c->d_psize.p, ps32.data(), ps32.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(c->d_poff.p, c->poff.data(), c->poff.size() * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(d_desc.p, desc.data(), desc.size() * sizeof(ProfDesc), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(d_off2.p, off2.data(), off2.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemcpyAsync(c->d_ptab.p, ptab.data(), ptab.size() * sizeof(ProfEntry), cudaMemcpyHostToDevice, st));
  TB_CUDA(cudaMemsetAsync(c->d_pdata2.p, 0, (total2 + TB_PROF_PAD) * sizeof(double), st));
  tb_timer_begin(c, "voigt");
  k_voigt_table<<<(unsigned)blocks, TPB, 0, st>>>(d_desc.p, (int)desc.size(), c->d_pdata.p);
  tb_timer_end(c, "voigt");
  k_polyphase<<<(unsigned)blocks, TPB, 0, st>>>(d_desc.p, d_off2.p, (int)desc.size(), c->o, c->d_pdata.p,
                                                c->d_pdata2.p);
  c->launches += 2;
  TB_CUDA(cudaGetLastError());
  TB_CUDA(cudaStreamSynchronize(st));
  d_desc.release();
  d_off2.release();
  c->aDop.resize(nDop);
  c->aLor.resize(nLor);
  c->have_prof = true;
  return TB_OK;
}

extern "C" int tb_get_profile_meta(tb_ctx *c, double *aDop, double *aLor, int64_t *psize, int64_t *off,
                                   int64_t *ndata) {
  TB_REQUIRE(c->have_prof, TB_ERR_STATE, "tb_build_profiles not called");
  if (aDop) memcpy(aDop, c->aDop.data(), c->nDop * sizeof(double));
  if (aLor) memcpy(aLor, c->aLor.data(), c->nLor * sizeof(double));
  if (psize) memcpy(psize, c->psize.data(), c->psize.size() * sizeof(int64_t));
  if (off) memcpy(off, c->poff.data(), c->poff.size() * sizeof(int64_t));
  if (ndata) *ndata = c->ndata;
  return TB_OK;
}

extern "C" int tb_get_profile_data(tb_ctx *c, float *data, int64_t ndata) {
  TB_REQUIRE(c->have_prof, TB_ERR_STATE, "tb_build_profiles not called");
  TB_REQUIRE(ndata == c->ndata, TB_ERR_ARG, "ndata mismatch (%lld vs %lld)", (long long)ndata,
             (long long)c->ndata);
  TB_CUDA(cudaSetDevice(c->device));
  TB_CUDA(cudaMemcpyAsync(data, c->d_pdata.p, ndata * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  return TB_OK;
}
