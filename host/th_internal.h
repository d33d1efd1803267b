/* th_internal.h -- the C host of the B200 transit path.
 *
 * Keeps the reference's process-level face (transit [-c file] [--key value ...], cfg keys,
 * atmosphere / TLI / molecules.dat / CS / opacity file formats, spectrum output) and its
 * library-level face (transit_init, run_transit, ... transit/src/transit.i:11-19,95-104), and
 * hands every array of the hot path to libtransit_b200.so through include/transit_b200.h.
 * No physics of the hot path is computed here: only I/O, units, and the O(layers)
 * per-atmosphere setup (densities, hydrostatic radii, partition-function splines).
 */
#ifndef TH_INTERNAL_H
#define TH_INTERNAL_H
#include <stdint.h>
#include <stdio.h>
#include "../include/transit_b200.h"

#define TH_AMU 1.66053886e-24
#define TH_KB 1.380658e-16
#define TH_SUNRADIUS 6.957e10
#define TH_MAXNAME 64

/* transithint (structures_tr.h:300-370): what the option table of argum.c:112-320 fills */
typedef struct {
  char *f_atm, *f_line, *f_molfile, *f_outspec, *f_toomuch, *f_outsample, *f_outintens, *f_opa, *f_saveext;
  char *csfiles;                  /* comma-separated */
  char *solname;
  char *raygrid, *qmol, *qscale;
  double rad_d;                   /* raddelt (-1: keep the atmosphere file's layers)          */
  double rad_i, rad_f, rad_fct;   /* radlow, radhigh, radfct (0: take the atmosphere file's)  */
  double wl_i, wl_f, wl_fct;      /* wllow, wlhigh, wlfct                                     */
  double wn_i, wn_f, wn_d, wn_fct;
  int wn_o;
  int nDop, nLor;
  float dmin, dmax, lmin, lmax, nwidth;
  double ethresh, toomuch;
  double tlow, thigh, tempdelt;
  double starrad;
  double p0, r0, gsurf, allowq;
  int transparent, opabreak, verb, savefiles, modlevel;
  int cl_flag;
  double cl_ext, cl_top, cl_bot, cl_gamma, cl_Q, cl_r, cl_sig, cl_refwn;
  int sc_flag;
  double sc_logext;
} th_hints;

typedef struct {          /* one TLI database (readlineinfo.c:87-244) */
  char name[TH_MAXNAME], molname[TH_MAXNAME];
  int nT, niso, first;    /* first: index of its first isotope in the global list */
  double *T;
} th_db;

typedef struct {        /* one set of layers: the atmosphere file's, or the resampled ones (raddelt != -1) */
  int n, cap;
  double *rad, *p, *t, *mm, *q, *d;
} th_layers;

typedef struct {
  th_hints h;
  /* wavenumber sampling (makewnsample) */
  double wn_i, wn_f, wn_d;
  int wn_o;
  int64_t nwn, onwn;
  /* atmosphere file (getatm): bottom layer first */
  int nlayer, nmol, bymass;
  double rad_fct, p_fct, t_fct, zerorad;
  double *rad, *p, *t, *mm;       /* [nlayer] file units */
  double *q, *d;                  /* [nmol][nlayer] abundance, density g/cm3 */
  char (*molname)[TH_MAXNAME];
  int *molID;
  double *molmass, *molradius, *molpol;
  /* line database (readtli_bin, readdatarng) */
  int ndb, niso;
  th_db *db;
  double *isomass, *isoratio, *isoZ;   /* isoZ[iso][nT of its db] flattened via isoZoff */
  int64_t *isoZoff;
  double *isoZ2;                  /* second derivatives of each isoZ spline, same layout; filled on first use */
  int *isodb, *isoimol;
  double tli_tmin, tli_tmax;
  int64_t nlines;
  double *wl, *elow, *gf;
  int16_t *isoid;
  /* CS files */
  int ncs;
  /* sampled state handed to the device (makeradsample) */
  double *Zlayer;                 /* [niso][nlayer] */
  int64_t zcap;
  /* raddelt != -1 (makeradsample's makesample branch, makesample.c:484-486): nlayer/rad/p/t/mm/q/d above hold the
   * RESAMPLED atmosphere once th_make_rad has run, and `other` keeps the file's own layers (run_transit's input) */
  int resampled;
  th_layers other;
  double samp_f;                  /* the accepted final radius (what outsample prints; the last sample may fall short) */
  int nip;                        /* impact parameters of makeipsample's makesample branch, top first */
  double *ips;
  /* geometry */
  int eclipse, nang;
  double angles[16];
  /* device */
  tb_ctx *ctx;
  int have_grid_file, initialised, rebalanced;
  int64_t band_lo, band_hi;       /* this rank's wavenumber band (whole grid when world = 1) */
  double *out;
} th_state;

/* Host-logic test hook, compiled ONLY into the test build (-DTH_TEST_TAPS: host/transit_b200_taps and
 * host/libtransit_host_taps.so, used by tests/test_host_cli.py): with TRANSIT_B200_DRYRUN=<file> no
 * device call is made (th_dry != 0) and do_transit writes the sampled atmosphere to <file> instead of
 * computing a spectrum.  The shipped host has no such switch: every device call is made.     */
#ifdef TH_TEST_TAPS
extern int th_dry;
#define TH_DEV(call) (th_dry ? 0 : (call))
#else
#define TH_DEV(call) (call)
#endif

/* th_config.c */
void th_defaults(th_hints *h);
int  th_parse_args(th_hints *h, int argc, char **argv);
/* th_inputs.c */
void th_die(const char *fmt, ...);
int  th_read_atmosphere(th_state *s);
int  th_read_tli(th_state *s);
int  th_read_cs(th_state *s);
/* th_setup.c */
void th_spline_init(double *z, const double *x, const double *y, long n);
double th_splinterp_pt(const double *z, long n, const double *x, const double *y, double xout);
int  th_make_wn(th_state *s);
int  th_densities(th_state *s);
int  th_radpress(th_state *s);
int  th_make_rad(th_state *s);
void th_file_layers(th_state *s);   /* raddelt != -1: make the atmosphere file's layers current again (reloadatm works on them) */
int  th_opacity(th_state *s);
/* th_api.c: the reference's library interface */
void transit_init(int argc, char **argv);
int  get_no_samples(void);
void get_waveno_arr(double *waveno_arr, int waveno);
void set_radius(double refradius);
void set_cloudtop(double cloudtop);
void set_scattering(int flag, double scattering);
void run_transit(double *re_input, int transint, double *transit_out, int transit_out_size);
void do_transit(double *transit_out);
void run_transit_batch(double *re_input, int natm, double *transit_out);
void free_memory(void);
void th_write_spectrum(void);
/* test/diagnostic accessor: the device context behind the library face (NULL before transit_init) */
/* multi-GPU: rank / world of this process and the pipe ends the 128-byte communicator id travels through
 * (set by th_main.c before transit_init; world 1 = single GPU) */
extern int th_rank, th_world, th_idfd[64];
tb_ctx *th_context(void);
void th_sizes(int64_t *out);
#endif
