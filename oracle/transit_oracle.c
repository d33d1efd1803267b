/* oracle/transit_oracle.c -- TEST INFRASTRUCTURE ONLY (see transit_oracle.h).
 *
 * Plain-C restatement of the reference's hot path.  Not a copy: the reference works
 * on one global `struct transit`; every function here takes plain arrays and is
 * re-entrant.  Each function cites the reference lines whose behaviour it restates.
 * Compiled with strict IEEE arithmetic (no -ffast-math, no FMA contraction), so it is
 * also the "what does the arithmetic say without compiler licence" arm; agreement
 * with the reference binary (built WITH -ffast-math) is asserted by
 * tests/test_oracle_golden.py: indices exact, values to ~1e-13.
 */
#include "transit_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

/* constants_tr.h:21-53 */
#define PI_ 3.141592653589793
#define AMU_ 1.66053886e-24
#define EC_ 4.8032068e-10
#define LS_ 2.99792458e10
#define ME_ 9.1093897e-28
#define KB_ 1.380658e-16
#define H_ 6.6260755e-27
#define SIGCTE_ (PI_*EC_*EC_/LS_/LS_/ME_/AMU_)
#define EXPCTE_ (H_*LS_/KB_)
#define NAVOGADRO_ 6.02214076e23
#define SQRTLN2_ 0.83255461115769775635
#define E0H2_ 4.911e-23
#define MICRON_ 1e-4
#define AMAGAT_ 2.68678e19          /* crosssec.h / constants: amagat in cm-3 */
#define DEGREES_ (PI_/180.0)

/* ------------------------------------------------------------------ pu/iomisc.c */

/* iomisc.c:1088-1108: nearest element, recursive bisection on [lo,hi] */
int to_binsearchapprox(const double *a, double v, int lo, int hi){
  while(hi - lo > 1){
    int mid = (hi + lo)/2;
    if(a[mid] > v) hi = mid; else lo = mid;
  }
  if(hi - lo == 1)
    return (fabs(a[hi]-v) < fabs(a[lo]-v)) ? hi : lo;
  return lo;                         /* hi==lo: the reference exits; callers never do this */
}

/* iomisc.c:1064-1083 */
void to_logspace(double mn, double mx, int n, double *out){
  double lmin = log10(mn), lmax = log10(mx), step = (lmax-lmin)/(n-1.0);
  for(int i=0; i<n; i++) out[i] = pow(10, lmin + i*step);
}

/* makesample.c:78-95: n = trunc(((1+1e-8) f - i)/d + 1), then oversampled */
int64_t to_grid_count(double wn_i, double wn_f, double d, int o){
  int64_t n = (int64_t)(((1.0+1e-8)*wn_f - wn_i)/d + 1);
  if(n < 0) n = -n;
  return (n-1)*o + 1;
}

/* ------------------------------------------------------------------ pu/voigt.c */

static const double ferf[61] = {   /* 1/(n!(2n+1)), voigt.c:46-108 */
  1.000000000000000000000, 0.333333333333333333333, 0.100000000000000000000,
  2.38095238095238095238e-2, 4.62962962962962962963e-3, 7.57575757575757575758e-4,
  1.06837606837606837607e-4, 1.32275132275132275132e-5, 1.45891690009337068161e-6,
  1.45038522231504687645e-7, 1.31225329638028050726e-8, 1.08922210371485733805e-9,
  8.35070279514723959168e-11, 5.94779401363763503681e-12, 3.95542951645852576340e-13,
  2.46682701026445692771e-14, 1.44832646435981372650e-15, 8.03273501241577360914e-17,
  4.22140728880708823303e-18, 2.10785519144213582486e-19, 1.00251649349077191670e-20,
  4.55184675892820028624e-22, 1.97706475387790517483e-23, 8.23014929921422135684e-25,
  3.28926034917575173275e-26, 1.26410789889891635220e-27, 4.67848351551848577373e-29,
  1.66976179341737202699e-30, 5.75419164398217177220e-32, 1.91694286210978253077e-33,
  6.18030758822279613746e-35, 1.93035720881510785656e-36, 5.84675500746883629630e-38,
  1.71885606280178362397e-39, 4.90892396452342296700e-41, 1.36304126177913957635e-42,
  3.68249351546114573519e-44, 9.68728023887076175384e-46, 2.48306909745491159104e-47,
  6.20565791963739670594e-49, 1.51310794954121709805e-50, 3.60157930981012591661e-52,
  8.37341968387228154283e-54, 1.90254122728987952724e-55, 4.22678975419355257584e-57,
  9.18642950239868569596e-59, 1.95410258232417110410e-60, 4.07013527785325672298e-62,
  8.30461450592911058168e-64, 1.66058051345108993284e-65, 3.25539546201302778914e-67,
  6.25918411694871134025e-69, 1.18076183891157008800e-70, 2.18621042295388572103e-72,
  3.97425272266506578576e-74, 7.09571739181805357327e-76, 1.24466597738907071213e-77,
  2.14564844309633852739e-79, 3.63615636540051474579e-81, 6.05939744697137480783e-83,
  9.93207019544894768776e-85};

/* voigt.c:132-200: Pierluissi 3-region approximation, evaluated in long double as the
 * reference does, rounded to float on store. */
static float voigtxy(double x, double y, double alphaD){
  const double SQRTLN2PI = 0.46971863934982566689, TWOOSQRTPI = 1.12837916709551257389;
  long double or_, nr, oi, ni, ar, ai;
  /* the reference computes x*x-y*y and 2*x*y in double and widens the result */
  const long double X2Y2 = (long double)(x*x - y*y);
  const long double XY2  = (long double)(2*x*y);
  if(x < 3 && y < 1.8){                         /* Region I */
    const long double cosxy = cosl(XY2), sinxy = sinl(XY2);
    int n = (x < 1 ? 15 : (int)(6.842*x + 8.0)) + 1, i = 1;
    ar = or_ = y;  ai = oi = -x;
    while(1){
      ni = or_*XY2 + oi*X2Y2;
      nr = or_*X2Y2 - oi*XY2;
      if(i > n) break;
      ai += ni*ferf[i];
      ar += nr*ferf[i];
      oi = ni; or_ = nr; i++;
    }
    return (float)(SQRTLN2PI/alphaD*exp((double)(-X2Y2)) *
                   (cosxy*(1 - ar*TWOOSQRTPI) - sinxy*ai*TWOOSQRTPI));
  }
  else if(x < 5 && y < 5){                      /* Region II */
    const double A1=0.46131350, A2=0.19016350, A3=0.09999216, A4=1.78449270,
                 A5=0.002883894, A6=5.52534370;
    ar = XY2*XY2; nr = XY2*x; ni = X2Y2-A2; ai = X2Y2-A4; oi = X2Y2-A6;
    return (float)(SQRTLN2PI/alphaD*(A1*((nr-ni*y)/(ni*ni+ar)) +
                                     A3*((nr-ai*y)/(ai*ai+ar)) +
                                     A5*((nr-oi*y)/(oi*oi+ar))));
  }
  else{                                         /* Region III */
    const double B1=0.51242424, B2=0.27525510, B3=0.05176536, B4=2.72474500;
    ar = XY2*XY2; nr = XY2*x; ni = X2Y2-B2; ai = X2Y2-B4;
    return (float)(SQRTLN2PI/alphaD*(B1*((nr-ni*y)/(ni*ni+ar)) +
                                     B3*((nr-ai*y)/(ai*ai+ar))));
  }
}

/* extinction.c:21-35: profile half-size in fine-grid samples */
int64_t to_voigt_psize(double dwn, float dop, float lor, float ta, int64_t nwave){
  double bigalpha = dop;
  if(bigalpha < lor) bigalpha = lor;
  double wvgt = bigalpha*ta;
  int nvgt = 2*(long)(wvgt/dwn + 0.5) + 1;
  if(nvgt < 2) nvgt = 3;
  if(nvgt > 2*(int)nwave) nvgt = 2*(int)nwave + 1;
  return nvgt/2;
}

/* extinction.c:8-57 (getprofile) + voigt.c:369-483 (voigtn) + 489-554 (bin means).
 * out must hold 2*psize+1 floats.  Returns psize. */
int64_t to_voigt_profile(float *out, double dwn_fine, float dop, float lor, float ta,
                         int64_t nwave){
  int64_t psz = to_voigt_psize(dwn_fine, dop, lor, ta, nwave);
  int nwn = (int)(2*psz + 1);
  int quick = nwn > 99999;                          /* _voigt_maxelements, voigt.c:109 */
  double dwn = dwn_fine*(long)(nwn/2);              /* half width handed to voigtn    */
  double alphaL = lor, alphaD = dop;
  double y = SQRTLN2_*alphaL/alphaD;
  double ddwn = 2.0*dwn/(nwn-1);
  int nint = 50;
  double dint = alphaD/(nint-1);
  if(ddwn < dint || quick){ dint = ddwn; nint = nwn + 1; }
  else{
    nint = (int)(ddwn/dint) + 1;
    if(nint & 1) nint++;
    nint = nwn*nint + 1;
    dint = 2.0*dwn/(nint-1);
  }
  float *aint = (float *)calloc(nint, sizeof(float));
  for(int i=0; i<nint; i++){
    double x = SQRTLN2_*fabs(dint*i - dwn)/alphaD;
    aint[i] = voigtxy(x, y, alphaD);
  }
  if(quick){
    for(int i=0; i<nwn; i++) out[i] = aint[i];
  }
  else{
    int ipo = (int)((float)(nint-1)/nwn) + 1;       /* points per bin, voigt.c:464-465 */
    const float *in = aint;
    if(ipo & 1){                                    /* Simpson, voigt.c:489-521 (float ops) */
      int last = ipo - 1;
      for(int k=0; k<nwn; k++){
        float acc = 0;
        for(int i=1; i<last; i+=2) acc += in[i];
        acc *= 2;
        for(int i=2; i<last; i+=2) acc += in[i];
        acc *= 2;
        acc += in[0] + in[last];
        acc = (float)(acc/(last*3.0));
        out[k] = acc;
        in += last;
      }
    }
    else{                                           /* trapezoid, voigt.c:527-554 */
      int last = ipo - 1;
      for(int k=0; k<nwn; k++){
        float acc = 0;
        for(int i=1; i<last; i++) acc += in[i];
        acc = (float)((acc + (in[0]+in[last])/2.0)/(double)last);
        out[k] = acc;
        in += last;
      }
    }
  }
  free(aint);
  return psz;
}

/* opacity.c:218-277 */
to_voigt_table *to_voigt_build(const to_grid *g, int nDop, int nLor, double dmin,
                               double dmax, double lmin, double lmax, float ta){
  to_voigt_table *t = (to_voigt_table *)calloc(1, sizeof(*t));
  t->nDop = nDop; t->nLor = nLor;
  t->aDop = (double *)malloc(nDop*sizeof(double));
  t->aLor = (double *)malloc(nLor*sizeof(double));
  to_logspace(dmin, dmax, nDop, t->aDop);
  to_logspace(lmin, lmax, nLor, t->aLor);
  t->psize = (int64_t *)calloc((size_t)nDop*nLor, sizeof(int64_t));
  t->off   = (int64_t *)calloc((size_t)nDop*nLor, sizeof(int64_t));
  double odwn = g->wn_d/g->o;
  int64_t total = 0;
  for(int i=0; i<nDop; i++) for(int j=0; j<nLor; j++){
    if(t->aDop[i]*10.0 < t->aLor[j] && i != 0){     /* share the previous row's profile */
      t->psize[i*nLor+j] = t->psize[(i-1)*nLor+j];
      t->off[i*nLor+j]   = t->off[(i-1)*nLor+j];
    }
    else{
      t->psize[i*nLor+j] = to_voigt_psize(odwn, (float)t->aDop[i], (float)t->aLor[j], ta, g->onwn);
      t->off[i*nLor+j] = total;
      total += 2*t->psize[i*nLor+j] + 1;
    }
  }
  t->ndata = total;
  t->data = (float *)malloc((size_t)total*sizeof(float));
  for(int i=0; i<nDop; i++) for(int j=0; j<nLor; j++){
    if(t->aDop[i]*10.0 < t->aLor[j] && i != 0) continue;
    to_voigt_profile(t->data + t->off[i*nLor+j], odwn, (float)t->aDop[i],
                     (float)t->aLor[j], ta, g->onwn);
  }
  return t;
}

void to_voigt_free(to_voigt_table *t){
  if(!t) return;
  free(t->aDop); free(t->aLor); free(t->psize); free(t->off); free(t->data); free(t);
}

/* ------------------------------------------------------------ extinction.c */

/* owns.v[k] as makesample1 fills it (makesample.c:100-104): si + k*osd */
static inline double fine_wn(const to_grid *g, int64_t k){
  return g->wn_i + (double)k*(g->wn_d/(double)g->o);
}

/* extinction.c:281-529 */
int to_computemolext(const to_grid *g, const to_lines *ln, const to_species *sp,
                     const to_voigt_table *vt, double temp, const double *density,
                     const double *Z, double ethresh, int permol, int Nout,
                     const int *iso_out, double *out, to_trace *tr){
  const int niso = sp->niso, nmol = sp->nmol, nDop = vt->nDop, nLor = vt->nLor;
  const int ofactor = g->o;
  const int64_t nwn = g->nwn, onwn = g->onwn, nlines = ln->n;
  const double dwn = g->wn_d/1.0;                  /* wns.d/wns.o, wns.o == 1 */
  const double odwn = g->wn_d/g->o;                /* owns.d/owns.o           */
  const double wn0 = g->wn_i;                      /* wn[0]                   */
  const double own_last = fine_wn(g, onwn-1);
  double *alphal = (double *)calloc(niso, sizeof(double));
  double *alphad = (double *)calloc(niso, sizeof(double));
  int *idop = (int *)calloc(niso, sizeof(int)), *ilor = (int *)calloc(niso, sizeof(int));
  double *kmax = (double *)calloc(Nout > 0 ? Nout : 1, sizeof(double));
  int m = 0;
  int64_t nadd = 0, nskip = 0, neval = 0, nsamp = 0;

  for(int64_t k=0; k<(int64_t)Nout*nwn; k++) out[k] = 0.0;

  double fdoppler = sqrt(2*KB_*temp/AMU_)*SQRTLN2_/LS_;
  double florentz = sqrt(2*KB_*temp/PI_/AMU_)/(AMU_*LS_);

  for(int i=0; i<niso; i++){                       /* :368-395 */
    alphal[i] = 0.0;
    for(int j=0; j<nmol; j++){
      double cs = sp->radius[j] + sp->radius[sp->iso_imol[i]];
      alphal[i] += density[j]/sp->mass[j]*cs*cs*sqrt(1/sp->iso_mass[i] + 1/sp->mass[j]);
    }
    alphal[i] *= florentz;
    alphad[i] = fdoppler/sqrt(sp->iso_mass[i]);
    idop[i] = to_binsearchapprox(vt->aDop, alphad[i]*wn0, 0, nDop);
    ilor[i] = to_binsearchapprox(vt->aLor, alphal[i],     0, nLor);
  }

  for(int64_t l=0; l<nlines; l++){                 /* :399-427 strength maximum */
    double wavn = 1.0/(ln->wl[l]*ln->wfct);
    int i = ln->isoid[l];
    if(permol) m = iso_out[i];
    if(wavn < wn0 || wavn > own_last) continue;
    double k = sp->iso_ratio[i]*SIGCTE_*ln->gf[l]*exp(-EXPCTE_*ln->efct*ln->elow[l]/temp)*
               (1 - exp(-EXPCTE_*wavn/temp))/sp->iso_mass[i]/Z[i];
    if(kmax[m] == 0) kmax[m] = k; else kmax[m] = fmax(kmax[m], k);
  }
  if(tr) for(int64_t l=0; l<nlines; l++) tr->status[l] = 0;

  for(int64_t l=0; l<nlines; l++){                 /* :430-511 */
    double wavn = 1.0/(ln->wl[l]*ln->wfct);
    int i = ln->isoid[l];
    if(permol) m = iso_out[i];
    if(wavn < wn0 || wavn > own_last) continue;
    int64_t lead = l;
    double k = ln->gf[l]*exp(-EXPCTE_*ln->efct*ln->elow[l]/temp)*(1 - exp(-EXPCTE_*wavn/temp));
    int iown = (int)((wavn - wn0)/odwn);
    if(fabs(wavn - fine_wn(g, iown+1)) < fabs(wavn - fine_wn(g, iown))) iown++;
    while(l != nlines-1 && ln->isoid[l+1] == i){   /* :449-462 co-add */
      double next = 1.0/(ln->wl[l+1]*ln->wfct);
      if(fabs(next - fine_wn(g, iown)) < odwn){
        nadd++; l++;
        if(tr) tr->status[l] = 2;
        k += ln->gf[l]*exp(-EXPCTE_*ln->efct*ln->elow[l]/temp)*(1 - exp(-EXPCTE_*next/temp));
      }
      else break;
    }
    k *= SIGCTE_*sp->iso_ratio[i]/(sp->iso_mass[i]*Z[i]);
    int idwn = (int)((wavn - wn0)/dwn);
    if(tr){ tr->iown[lead] = iown; tr->idwn[lead] = idwn; tr->idop[lead] = -1; }
    if(k < ethresh*kmax[m]){ nskip++; if(tr) tr->status[lead] = 3; continue; }
    if(permol == 0) k *= density[sp->iso_imol[i]];
    if(alphad[i]*wavn/alphal[i] >= 1e-1)
      idop[i] = to_binsearchapprox(vt->aDop, alphad[i]*wavn, 0, nDop);
    int64_t psz = vt->psize[idop[i]*nLor + ilor[i]];
    const float *prof = vt->data + vt->off[idop[i]*nLor + ilor[i]];
    int64_t subw = iown - (int64_t)idwn*ofactor;
    int64_t offset = iown - psz;
    int64_t minj = idwn - (psz - subw)/ofactor;
    int64_t maxj = idwn + (psz + subw)/ofactor;
    if(minj < 0) minj = 0;
    if(maxj >= nwn) maxj = nwn - 1;
    int64_t bj = ofactor*minj - offset;
    double *row = out + (int64_t)m*nwn;
    for(int64_t j=minj; j<=maxj; j++){
      if(bj > 2*psz) break;
      if(bj >= 0){ row[j] += k*prof[bj]; nsamp++; }
      bj += ofactor;
    }
    neval++;
    if(tr){ tr->status[lead] = 1; tr->idop[lead] = idop[i]; }
  }
  if(tr){ tr->nsamples = nsamp; tr->nadd = nadd; tr->nskip = nskip; tr->neval = neval; }
  free(alphal); free(alphad); free(idop); free(ilor); free(kmax);
  return 0;
}

/* extinction.c:534-581.  dens_of_gridmol[m] = density at layer r of the grid's m-th
 * molecule (mol->molec[valueinarray(mol->ID, gmol[m])].d[r]).                       */
int to_interpolmolext(const double *grid, int64_t Ntemp, int64_t Nmol, int64_t Nwave,
                      const double *gtemp, int64_t r, double temp,
                      const double *dens_of_gridmol, double *e_row){
  int it = to_binsearchapprox(gtemp, temp, 0, (int)Ntemp);
  if(temp < gtemp[it]) it--;
  const double *lo = grid + ((r*Ntemp + it)*Nmol)*Nwave;
  const double *hi = grid + ((r*Ntemp + it + 1)*Nmol)*Nwave;
  for(int64_t i=0; i<Nwave; i++)
    for(int64_t m=0; m<Nmol; m++){
      double ext = (lo[m*Nwave+i]*(gtemp[it+1]-temp) + hi[m*Nwave+i]*(temp-gtemp[it]))/
                   (gtemp[it+1]-gtemp[it]);
      e_row[i] += dens_of_gridmol[m]*ext;
    }
  return 0;
}

/* ------------------------------------------------------------------ pu/spline.c */

/* spline.c:12-48 (tri) + 186-206 (spline_init) */
void to_spline_init(double *z, const double *x, const double *y, long n){
  double *h = (double *)calloc(n-1, sizeof(double)), *b = (double *)calloc(n-1, sizeof(double)),
         *u = (double *)calloc(n-1, sizeof(double)), *v = (double *)calloc(n-1, sizeof(double));
  for(long i=0; i<n-1; i++) h[i] = x[i+1]-x[i];
  for(long i=0; i<n-1; i++) b[i] = (y[i+1]-y[i])/h[i];
  if(n > 2){
    u[1] = 2*(h[1]+h[0]);
    v[1] = 6*(b[1]-b[0]);
  }
  for(long i=2; i<n-1; i++){
    u[i] = 2*(h[i]+h[i-1]) - h[i-1]*h[i-1]/u[i-1];
    v[i] = 6*(b[i]-b[i-1]) - v[i-1]*h[i-1]/u[i-1];
  }
  z[0] = z[n-1] = 0;
  for(long i=n-2; i>0; i--) z[i] = (v[i] - h[i]*z[i+1])/u[i];
  free(h); free(b); free(u); free(v);
}

/* spline.c:131-183 */
double to_splinterp_pt(const double *z, long N, const double *x, const double *y, double xout){
  int idx = to_binsearchapprox(x, xout, 0, (int)N-1);
  if(idx == N-1 || xout < x[idx]) idx--;
  double x_lo = x[idx], x_hi = x[idx+1], y_lo = y[idx], y_hi = y[idx+1];
  double h = x_hi-x_lo, dy = y_hi-y_lo;
  if(x[idx] == xout) return y[idx];
  if(h > 0){
    double dx = xout-x_lo;
    double a = (z[idx+1]-z[idx])/(6*h), b = 0.5*z[idx];
    double c = dy/h - h/6*(z[idx+1] + 2*z[idx]);
    return y_lo + dx*(c + dx*(b + dx*a));
  }
  return 0;
}

/* spline.c:55-128 (splinterp + spline3) */
void to_splinterp(long N, const double *xi, const double *yi, long nx, const double *x, double *y){
  double *z = (double *)calloc(N, sizeof(double)), *h = (double *)calloc(N-1, sizeof(double));
  for(long i=0; i<N-1; i++) h[i] = xi[i+1]-xi[i];
  to_spline_init(z, xi, yi, N);
  for(long n=0; n<nx; n++){
    int i = to_binsearchapprox(xi, x[n], 0, (int)N-1);
    if(i == N-1 || x[n] < xi[i]) i--;
    double B = (yi[i+1]-yi[i])/h[i] - h[i]/6*(z[i+1] + 2*z[i]);
    double d = x[n]-xi[i];
    y[n] = yi[i] + d*B + pow(d,2)*0.5*z[i] + pow(d,3)*(z[i+1]-z[i])/(6*h[i]);
  }
  free(z); free(h);
}

/* ------------------------------------------------------------------ crosssec.c */

/* crosssec.c:353-428 (bicubicinterpolate) + 314-336 (density scaling, positivity).
 * tab[ntw][ntt]; densfac[nrad] = prod_k rho_k/(AMU m_k AMAGAT).  Adds to e_cs[nwn][nrad]. */
int to_interpcs_table(const double *tab, const double *twn, int ntw, const double *ttemp,
                      int ntt, const double *wn, int64_t nwn, const double *temp,
                      int nrad, const double *densfac, double *e_cs){
  /* x1=table wn, x2=table T, t1=grid wn, t2=layer T */
  double fx1 = twn[0], lx1 = twn[ntw-1], fx2 = ttemp[0], lx2 = ttemp[ntt-1];
  int64_t li = nwn, fi = 0; long lj = nrad, fj = 0;
  if(wn[0] > lx1 || wn[nwn-1] < fx1 || temp[0] > lx2 || temp[nrad-1] < fx2) return 0;
  while(wn[fi++] < fx1); fi--;
  for(int64_t i=0; i<li; i++) if(wn[i] > lx1) li = i;
  while(temp[fj++] < fx2); fj--;
  for(long j=0; j<lj; j++) if(temp[j] > lx2) lj = j;
  double *z1 = (double *)calloc(ntt, sizeof(double)), *z2 = (double *)calloc(ntw, sizeof(double));
  double *f2 = (double *)calloc((size_t)nrad*ntw, sizeof(double));
  double *res = (double *)calloc((size_t)nwn*nrad, sizeof(double));
  for(int i=0; i<ntw; i++){
    to_spline_init(z1, ttemp, tab + (size_t)i*ntt, ntt);
    for(long j=fj; j<lj; j++)
      f2[j*ntw+i] = to_splinterp_pt(z1, ntt, ttemp, tab + (size_t)i*ntt, temp[j]);
  }
  for(long j=fj; j<lj; j++){
    to_spline_init(z2, twn, f2 + j*ntw, ntw);
    for(int64_t i=fi; i<li; i++)
      res[i*nrad+j] += to_splinterp_pt(z2, ntw, twn, f2 + j*ntw, wn[i]);
  }
  for(int i=0; i<nrad; i++)
    for(int64_t j=0; j<nwn; j++)
      if(res[j*nrad+i] > 0) e_cs[j*nrad+i] += res[j*nrad+i]*densfac[i];
  free(z1); free(z2); free(f2); free(res);
  return 0;
}

/* ------------------------------------------------------ scattering and clouds */

/* extinction.c:586-624 */
void to_extscat(double *e, int n, int flag, double logext, const double *press,
                const double *temp, int nmol, const double *pol, const double *mol_d,
                const double *mol_mass, double wn){
  if(flag == 1){
    for(int i=0; i<n; i++) e[i] = pow(10.0, logext)*E0H2_*press[i]/temp[i]*pow(wn, 4);
  }
  else if(flag == 2){
    for(int i=0; i<n; i++){
      e[i] = 0;
      for(int j=0; j<nmol; j++)
        e[i] += PI_*8e-32/3.*pow(pol[j], 2)*pow(2.*PI_*wn*MICRON_, 4)*mol_d[(size_t)j*n+i]/
                mol_mass[j]*NAVOGADRO_;
    }
  }
  else memset(e, 0, n*sizeof(double));
}

/* extinction.c:629-693 */
void to_extcloud(double *e, int n, int flag, double cloudext, double cloudtop,
                 double cloudbot, double gamma, double Q, double r, double sig,
                 double refwn, const double *nH, const double *press,
                 const double *density, double wn){
  double top = pow(10, cloudtop), bot = pow(10, cloudbot);
  double x = 2*PI_*r*wn, rw = pow(refwn, gamma), kBP = cloudext*pow(wn, gamma);
  double kFH = cloudext/(Q*pow(x, -1*gamma) + pow(x, 0.2));
  int i;
  if(!cloudext){ memset(e, 0, n*sizeof(double)); return; }
  for(i=n-1; i>=0; i--){ if(press[i] >= top) break; e[i] = 0.0; }
  for(; i>=0; i--){
    if(press[i] >= bot) break;
    switch(flag){
      case 1: e[i] = cloudext; break;
      case 2: e[i] = cloudext*density[i]; break;
      case 3: e[i] = kBP*density[i]; break;
      case 4: e[i] = kFH*density[i]; break;
      case 5: e[i] = nH[i]*kBP*sig/rw*density[i]; break;
    }
  }
  for(; i>=0; i--) e[i] = 0.0;
}

/* ------------------------------------------------------------ pu/numerical.c */

/* numerical.c:16-45 */
static int binsearchie(const double *arr, long i, long f, double val){
  if(arr[i] > val) return -1;
  if(arr[f] < val) return -2;
  if(arr[f] == val) return -5;
  if(i == f && arr[i] != val) return -3;
  while(f-i > 1){
    long m = (f+i) >> 1;
    if(arr[m] > val) f = m; else i = m;
  }
  return (int)i;
}

/* numerical.c:182-195 */
static double interp_parab(const double *x, const double *y, double xr){
  const double dx = x[1]-x[0], x0 = x[0]/dx, my = y[0]+y[2]-2*y[1];
  const double a = my/(2.0*dx*dx);
  const double b = (y[2]-y[1]-(x0+1.5)*my)/dx;
  const double c = y[0] + x0*(y[2]-4*y[1]+3*y[0]+x0*my)/2.0;
  return xr*xr*a + xr*b + c;
}

/* numerical.c:390-525 (makeh, geth, simps, simpson) folded into one routine:
 * integral of y over abscissae s[0..n-1] */
static double simps_s(const double *y, const double *s, int n){
  if(n == 1) return 0.0;
  if(n == 2) return (s[1]-s[0])*(y[0]+y[1])/2;
  int even = (n % 2 == 0);
  double res = 0.0;
  for(int i=0; i<(n-1)/2; i++){
    int j = 2*i + even;
    double h0 = s[j+1]-s[j], h1 = s[j+2]-s[j+1];
    double hsum = h0+h1, hratio = h1/h0, hfactor = hsum*hsum/(h0*h1);
    res += (y[j]*(2.0-hratio) + y[j+1]*hfactor + y[j+2]*(2.0-1.0/hratio))*hsum;
  }
  double integ = res/6.0;
  if(even) integ += (s[1]-s[0])*(y[0]+y[1])/2;
  return integ;
}

/* slantpath.c:18-108 */
double to_totaltau1(double b, double *rad, double refr, double *ex, long nrad){
  double r0 = b/refr, x3[3], r3[3];
  int rs = binsearchie(rad, 0, nrad-1, r0);
  if(rs == -5 || rs == -2) return 0;
  if(rs < 0){ fprintf(stderr, "to_totaltau1: closest approach out of range\n"); return NAN; }
  rad += rs; ex += rs; nrad -= rs;
  const double tmpex = *ex, tmprad = *rad;
  double *rad0 = rad, *ex0 = ex;
  if(nrad == 2) *ex = interp_parab(rad-1, ex-1, r0);
  else          *ex = interp_parab(rad,   ex,   r0);
  *rad = r0;
  if(nrad == 2){
    x3[0] = ex[0]; x3[2] = ex[1]; x3[1] = (ex[0]+ex[1])/2.0;
    r3[0] = rad[0]; r3[2] = rad[1]; r3[1] = (rad[0]+rad[1])/2.0;
    *rad = tmprad; *ex = tmpex;
    rad = r3; ex = x3; nrad++;
  }
  double *s = (double *)malloc(nrad*sizeof(double));
  s[0] = 0;
  for(long i=1; i<nrad; i++) s[i] = sqrt(rad[i]*rad[i] - r0*r0);
  double res = simps_s(ex, s, (int)nrad);
  free(s);
  *ex0 = tmpex; *rad0 = tmprad;
  return 2*res;
}

/* eclipse.c:28-105: ex[rs] keeps the parabola value unless only two points remain */
double to_eclipsetau(double height, double *rad, double *ex, long nrad_all){
  int rs = to_binsearchapprox(rad, height, 0, (int)nrad_all-1);
  if(rs == nrad_all-1) return 0.0;
  rad += rs; ex += rs;
  long nrad = nrad_all - rs;
  double x3[3], r3[3];
  const double tmpex = *ex;
  if(nrad == 2) *ex = interp_parab(rad-1, ex-1, rad[0]);
  else          *ex = interp_parab(rad,   ex,   rad[0]);
  if(nrad == 2){
    x3[0] = ex[0]; x3[2] = ex[1]; x3[1] = (ex[1]+ex[0])/2.0;
    r3[0] = rad[0]; r3[2] = rad[1]; r3[1] = (rad[0]+rad[1])/2.0;
    *ex = tmpex;
    rad = r3; ex = x3; nrad++;
  }
  double *s = (double *)malloc(nrad*sizeof(double));
  s[0] = 0.0;
  for(long i=1; i<nrad; i++) s[i] = s[i-1] + (rad[i]-rad[i-1]);
  double res = simps_s(ex, s, (int)nrad);
  free(s);
  return res;
}

/* ------------------------------------------------------------------ tau.c */

/* tau.c:59-356 */
int to_tau(int eclipse, const double *e, const double *e_cs, const double *wn,
           int64_t nwn, double wfct, double *rad, int nrad, double rfct,
           const double *h, int nh, double hfct, const double *press,
           const double *temp, const to_contrib *cb, double toomuch, int lazy,
           double *t, int64_t *last, int16_t *computed){
  double *er = (double *)malloc(nrad*sizeof(double)), *e_s = (double *)malloc(nrad*sizeof(double)),
         *e_c = (double *)malloc(nrad*sizeof(double));
  int16_t *comp = (int16_t *)calloc(nrad, sizeof(int16_t));
  int lastr = nrad-1;
  if(!lazy) for(int i=0; i<nrad; i++) comp[i] = 1;
  comp[nrad-1] = 1;
  memset(t, 0, (size_t)nwn*nh*sizeof(double));
  for(int64_t wi=0; wi<nwn; wi++){
    double *tw = t + wi*nh;
    double w = wn[wi]*wfct;
    to_extscat(e_s, nrad, cb->sc_flag, cb->sc_logext, press, temp, cb->nmol, cb->pol,
               cb->mol_d, cb->mol_mass, w);
    to_extcloud(e_c, nrad, cb->cl_flag, cb->cl_ext, cb->cl_top, cb->cl_bot, cb->cl_gamma,
                cb->cl_Q, cb->cl_r, cb->cl_sig, cb->cl_refwn, cb->nH, press, cb->mean_dens, w);
    for(int r=0; r<nrad; r++)
      er[r] = (comp[r] ? e[(int64_t)r*nwn+wi] : 0.0) + e_s[r] + e_c[r] + e_cs[wi*nrad+r];
    int ri;
    for(ri=0; ri<nh; ri++){
      if(h[ri]*hfct < rad[lastr]*rfct){
        do{
          if(!comp[--lastr]){
            comp[lastr] = 1;
            er[lastr] = e[(int64_t)lastr*nwn+wi] + e_s[lastr] + e_c[lastr] + e_cs[wi*nrad+lastr];
          }
        }while(h[ri]*hfct < rad[lastr]*rfct);
      }
      /* tau.c:274 as the reference is BUILT: -ffast-math hoists 1/rfct out of the loop (objdump of
         tau(): movsd 1.0; divsd rfct before the wavenumber loop), so b = (h*hfct)*(1/rfct).  With
         IEEE division some impact parameters land one ulp below their own layer radius and the ray
         picks the layer underneath: a 20% change in tau the reference binary never shows. */
      double b = h[ri]*hfct*(1.0/rfct);
      tw[ri] = rfct*(eclipse ? to_eclipsetau(b, rad, er, nrad)
                             : to_totaltau1(b, rad, 1.0, er, nrad));
      if(tw[ri] > toomuch){ last[wi] = ri; break; }
    }
    if(ri == nh) last[wi] = ri-1;
  }
  if(computed) memcpy(computed, comp, nrad*sizeof(int16_t));
  free(er); free(e_s); free(e_c); free(comp);
  return 0;
}

/* slantpath.c:350-436 */
double to_modulation1(const double *tau, long last, double toomuch, const double *ip,
                      long ipn, double ipfct, double srad, int transp){
  long ipn1 = ipn-1, i;
  const double maxtau = tau[last] > toomuch ? tau[last] : toomuch;
  double *rinteg = (double *)calloc(ipn, sizeof(double)), *ipv = (double *)calloc(ipn, sizeof(double));
  for(i=0; i<=last; i++){
    ipv[ipn1-i] = ip[i]*ipfct;
    rinteg[ipn1-i] = exp(-tau[i])*ipv[ipn1-i];
  }
  last += 1;
  if(last > ipn1) last = ipn1;
  for(; i<=last; i++){ ipv[ipn1-i] = ip[i]*ipfct; rinteg[ipn1-i] = 0; }
  last++;
  if(last < 3){ free(rinteg); free(ipv); return NAN; }
  double res = simps_s(rinteg+ipn-last, ipv+ipn-last, (int)last);
  res = ipv[ipn1]*ipv[ipn1] - 2.0*res;
  if(transp) res -= exp(-maxtau)*ipv[ipn-last]*ipv[ipn-last];
  res *= 1.0/(srad*srad);
  free(rinteg); free(ipv);
  return res;
}

/* modulationm1, slantpath.c:446-473 (interp_line: pu/src/numerical.c:200-210): the planet as an opaque disc
 * whose radius is where tau reaches toomuch.  Returns -1 if toomuch was not reached (slantpath.c:308-316).
 * When the first ray is already opaque the reference leaves ipv[1] unset; zero stands in for it here. */
double to_modulationm1(const double *tau, long last, double toomuch, const double *ip,
                       long ipn, double ipfct, double srad){
  long i, ini;
  double ipv[2] = {0.0, 0.0};
  (void)ipn;
  if(tau[last] < toomuch) return -1;
  ini = ++last - 2;
  if(ini < 0) ini = 0;
  for(i=ini; i<last; i++) ipv[i-ini] = ip[i]*ipfct;
  {
    const double *x = tau + ini;
    const double dx = x[1] - x[0];
    const double m = (ipv[1] - ipv[0]) / dx;
    const double muchrad = ipv[0] + (toomuch - x[0]) * m;
    return muchrad*muchrad/(srad*srad);
  }
}

/* eclipse.c:117-160 (+ integ_trapz numerical.c:154-172) */
double to_eclipse_intens(const double *tau, double w, double wfct, long last,
                         const double *temp, long rnn, double angle_deg){
  double angle = angle_deg*DEGREES_;
  double *B = (double *)malloc(rnn*sizeof(double)), *dtau = (double *)malloc(rnn*sizeof(double));
  for(long i=0; i<=last; i++){
    dtau[i] = exp(-tau[i]/cos(angle));
    B[i] = (2.0*H_*pow(w*wfct, 3.0)*LS_*LS_)/(exp(H_*w*wfct*LS_/(KB_*temp[rnn-1-i])) - 1.0);
  }
  double tz = 0;
  for(long i=0; i<last; i++) tz += (dtau[i+1]-dtau[i])*(B[i+1]+B[i]);
  double res = B[last]*dtau[last] - 0.5*tz;
  free(B); free(dtau);
  return res;
}

/* eclipse.c:242-287 */
void to_flux(const double *intens, int nang, const double *angles, int64_t nwn, double *out){
  double *grid = (double *)calloc(nang+1, sizeof(double));
  grid[0] = 0.0*DEGREES_; grid[nang] = 90.0*DEGREES_;
  for(int i=1; i<nang; i++) grid[i] = (angles[i-1]+angles[i])*DEGREES_/2.0;
  for(int64_t w=0; w<nwn; w++) out[w] = 0;
  for(int i=0; i<nang; i++){
    double area = pow(sin(grid[i+1]), 2.0) - pow(sin(grid[i]), 2.0);
    for(int64_t w=0; w<nwn; w++) out[w] += PI_*intens[(int64_t)i*nwn+w]*area;
  }
  free(grid);
}

/* readatm.c:787-865 */
int to_radpress(double g0, double p0, double r0, const double *temp, const double *mu,
                const double *pressure, double *radius, int nlayers, double rfct){
  double temp0, mu0, g, minimum = 1e37;
  int i, i0 = -1;
  for(i=0; i<nlayers; i++)
    if(fabs(pressure[i]-p0) < minimum){ i0 = i; minimum = fabs(pressure[i]-p0); }
  if(i0 < 0) return -1;
  if(pressure[i0] > p0){
    temp0 = temp[i0] + ((temp[i0+1]-temp[i0])/log(pressure[i0+1]/pressure[i0]))*log(p0/pressure[i0]);
    mu0   = mu[i0]   + ((mu[i0+1]-mu[i0])/log(pressure[i0+1]/pressure[i0]))*log(p0/pressure[i0]);
    radius[i0] = r0 + 0.5*(temp[i0]/mu[i0] + temp0/mu0)*(KB_/AMU_*log(p0/pressure[i0])/g0)/rfct;
  }
  else{
    temp0 = temp[i0] + ((temp[i0-1]-temp[i0])/log(pressure[i0-1]/pressure[i0]))*log(p0/pressure[i0]);
    mu0   = mu[i0]   + ((mu[i0-1]-mu[i0])/log(pressure[i0-1]/pressure[i0]))*log(p0/pressure[i0]);
    radius[i0] = r0 - 0.5*(temp[i0]/mu[i0] + temp0/mu0)*(KB_/AMU_*log(pressure[i0]/p0)/g0)/rfct;
  }
  g = g0*pow(r0/radius[i0], 2);
  for(i=i0-1; i>=0; i--){
    radius[i] = radius[i+1] - 0.5*(temp[i]/mu[i] + temp[i+1]/mu[i+1])*
                (KB_/AMU_*log(pressure[i]/pressure[i+1])/g)/rfct;
    g = g*pow(radius[i+1]/radius[i], 2);
  }
  g = g0*pow(r0/radius[i0], 2);
  for(i=i0+1; i<nlayers; i++){
    radius[i] = radius[i-1] + 0.5*(temp[i]/mu[i] + temp[i-1]/mu[i-1])*
                (KB_/AMU_*log(pressure[i-1]/pressure[i])/g)/rfct;
    g = g*pow(radius[i-1]/radius[i], 2);
  }
  return 1;
}
