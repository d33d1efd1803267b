/* oracle/transit_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C99, strict IEEE, single thread) of the reference's
 * line-by-line spectrum hot path, one function per reference function, on plain
 * arrays.  It is the checker for the CUDA path: only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  It is pinned
 * against the unmodified reference through tests/golden/*.npz
 * (tests/test_oracle_golden.py).  Citations are into /root/reference.
 */
#ifndef TRANSIT_ORACLE_H
#define TRANSIT_ORACLE_H
#include <stdint.h>

/* Voigt table, opacity.c:218-277 (calcprofiles) */
typedef struct {
  int nDop, nLor;
  double *aDop, *aLor;      /* logspace(), iomisc.c:1064-1083            */
  int64_t *psize;           /* [nDop*nLor] half-size, fine-grid samples  */
  int64_t *off;             /* [nDop*nLor] offset into data (shared rows repeat) */
  float *data;              /* concatenated profiles, 2*psize+1 each     */
  int64_t ndata;
} to_voigt_table;

/* wavenumber grids, makesample.c:27-120,308-400 */
typedef struct {
  double wn_i;              /* wns.i == owns.i                           */
  double wn_d;              /* coarse spacing wns.d                      */
  int    o;                 /* owns.o (wnosamp)                          */
  int64_t nwn;              /* coarse samples                            */
  int64_t onwn;             /* fine samples                              */
} to_grid;

typedef struct {
  int nmol; const double *mass, *radius; const int *ID;      /* molecules.dat  */
  int niso; const double *iso_mass, *iso_ratio; const int *iso_imol;
} to_species;

typedef struct {
  int64_t n; const double *wl, *elow, *gf; const int16_t *isoid;
  double wfct, efct;
} to_lines;

/* per-line trace of computemolext's decisions (for bit-exact index parity) */
typedef struct {
  int32_t *iown, *idwn, *idop;   /* [n] valid for status 1 and 3 (idop: status 1) */
  int8_t  *status;               /* 0 out of range, 1 evaluated leader, 2 co-added, 3 skipped (threshold) */
  int64_t nsamples;              /* executed kiso[m][j] += ... updates (extinction.c:507) */
  int64_t nadd, nskip, neval;
} to_trace;

int    to_binsearchapprox(const double *a, double v, int lo, int hi);
void   to_logspace(double mn, double mx, int n, double *out);
int64_t to_grid_count(double wn_i, double wn_f, double d, int o);   /* makesample1 n */

int64_t to_voigt_profile(float *out, double dwn, float dop, float lor, float ta,
                         int64_t nwave);                         /* extinction.c:8-57 */
int64_t to_voigt_psize(double dwn, float dop, float lor, float ta, int64_t nwave);
to_voigt_table *to_voigt_build(const to_grid *g, int nDop, int nLor, double dmin,
                               double dmax, double lmin, double lmax, float ta);
void   to_voigt_free(to_voigt_table *t);

/* extinction.c:281-529.  out[Nout][nwn] (Nout = 1 when !permol); iso_out[niso] maps an
 * isotope to its row of `out` when permol (valueinarray(op->molID, ...)).            */
int to_computemolext(const to_grid *g, const to_lines *ln, const to_species *sp,
                     const to_voigt_table *vt, double temp, const double *density,
                     const double *Z, double ethresh, int permol, int Nout,
                     const int *iso_out, double *out, to_trace *tr);

/* extinction.c:534-581.  grid[Nlayer][Ntemp][Nmol][Nwave]; adds onto e_row[Nwave]     */
int to_interpolmolext(const double *grid, int64_t Ntemp, int64_t Nmol, int64_t Nwave,
                      const double *gtemp, int64_t r, double temp,
                      const double *dens_of_gridmol, double *e_row);

/* crosssec.c:271-428 for one table; adds onto e_cs[nwn][nrad]                        */
int to_interpcs_table(const double *tab, const double *twn, int ntw, const double *ttemp,
                      int ntt, const double *wn, int64_t nwn, const double *temp,
                      int nrad, const double *densfac, double *e_cs);

/* extinction.c:586-693 */
void to_extscat(double *e, int n, int flag, double logext, const double *press,
                const double *temp, int nmol, const double *pol, const double *mol_d,
                const double *mol_mass, double wn);
void to_extcloud(double *e, int n, int flag, double cloudext, double cloudtop,
                 double cloudbot, double gamma, double Q, double r, double sig,
                 double refwn, const double *nH, const double *press,
                 const double *density, double wn);

/* slantpath.c:18-108 / eclipse.c:28-105 (ex is modified as the reference does)       */
double to_totaltau1(double b, double *rad, double refr, double *ex, long nrad);
double to_eclipsetau(double height, double *rad, double *ex, long nrad);

typedef struct {
  int sc_flag; double sc_logext;
  int cl_flag; double cl_ext, cl_top, cl_bot, cl_gamma, cl_Q, cl_r, cl_sig, cl_refwn;
  const double *nH, *mean_dens;          /* [nrad], tau.c:194-214 */
  int nmol; const double *pol, *mol_d, *mol_mass;  /* mol_d[nmol][nrad] */
} to_contrib;

/* tau.c:59-356.  e[nrad][nwn] holds EVERY layer; `lazy` replays the reference's lazy
 * evaluation order (uncomputed layers read as 0, computed[] is produced).            */
int to_tau(int eclipse, const double *e, const double *e_cs, const double *wn,
           int64_t nwn, double wfct, double *rad, int nrad, double rfct,
           const double *h, int nh, double hfct, const double *press,
           const double *temp, const to_contrib *cb, double toomuch, int lazy,
           double *t, int64_t *last, int16_t *computed);

/* slantpath.c:350-436 */
double to_modulationm1(const double *tau, long last, double toomuch, const double *ipv,
                       long ipn, double ipfct, double srad);
double to_modulation1(const double *tau, long last, double toomuch, const double *ipv,
                      long ipn, double ipfct, double srad, int transp);
/* eclipse.c:117-160 */
double to_eclipse_intens(const double *tau, double w, double wfct, long last,
                         const double *temp, long rnn, double angle_deg);
/* eclipse.c:242-287 */
void to_flux(const double *intens, int nang, const double *angles_deg, int64_t nwn,
             double *out);

/* spline.c */
void   to_spline_init(double *z, const double *x, const double *y, long n);
double to_splinterp_pt(const double *z, long n, const double *x, const double *y,
                       double xout);
void   to_splinterp(long N, const double *xi, const double *yi, long nx,
                    const double *xout, double *yout);

/* readatm.c:787-865 */
int to_radpress(double g, double p0, double r0, const double *temp, const double *mu,
                const double *pressure, double *radius, int nlayer, double rfct);
#endif
