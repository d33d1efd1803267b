/* th_api.c -- the reference's library interface (transit/src/transit.c:14-228, transit.i:11-19)
 * on top of libtransit_b200.so: same function names, argument meaning and error behaviour
 * (messages on stderr and exit(EXIT_FAILURE); no return codes), one global state like the
 * reference's `struct transit transit` (transit.c:7).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "th_internal.h"

static th_state S;
#ifdef TH_TEST_TAPS
int th_dry = 0;
#endif

#define CK(call) do { if (TH_DEV(call) != TB_OK) th_die("%s\n", tb_last_error()); } while (0)

static void apply_cloud_scattering(void) {
  const th_hints *h = &S.h;
  CK(tb_set_scattering(S.ctx, h->sc_flag, h->sc_logext));
  /* tau.c:194-214, restated term by term (including the abundance factor the reference carries in
   * both sums): mean_dens[r] = (sum_j d_j/m_j*q_j) * (sum_j m_j*q_j), and the H2 number density
   * nH[r] = d_H2/m_H2*q_H2*N_A for the species NAMED "H2" only.  The reference accumulates mean_dens
   * into an uninitialised stack array (tau.c:129,203); the sum starts from zero here.           */
  const int n = S.nlayer;
  double *nH = (double *)calloc(n, sizeof(double)), *md = (double *)calloc(n, sizeof(double));
  for (int r = 0; r < n; r++) {
    double mean_mm = 0;
    for (int j = 0; j < S.nmol; j++) {
      const double d = S.d[(size_t)j * n + r], q = S.q[(size_t)j * n + r];
      md[r] += d / S.molmass[j] * q;
      if (!strcmp("H2", S.molname[j])) nH[r] = d / S.molmass[j] * q * 6.02214076e23;
      mean_mm += S.molmass[j] * q;
    }
    md[r] *= mean_mm;
  }
  CK(tb_set_cloud(S.ctx, h->cl_flag, h->cl_ext, h->cl_top, h->cl_bot, h->cl_gamma, h->cl_Q, h->cl_r, h->cl_sig,
                  h->cl_refwn, nH, md));
  free(nH); free(md);
}

void transit_init(int argc, char **argv) {
  memset(&S, 0, sizeof(S));
  th_defaults(&S.h);
  th_parse_args(&S.h, argc, argv);
  const th_hints *h = &S.h;
  if (h->nwidth < 1) th_die("Times of maximum width has to be greater than one: %g\n", h->nwidth);
  if (h->ethresh <= 0) th_die("Extinction threshold has to be positive: %g\n", h->ethresh);
  if (strcmp(h->solname, "eclipse") == 0) S.eclipse = 1;
  else if (strcmp(h->solname, "transit") == 0 || strcmp(h->solname, "slant path") == 0 ||
           strcmp(h->solname, "slantpath") == 0) S.eclipse = 0;
  else th_die("Solution kind '%s' is invalid! Currently accepted are: eclipse, transit.\n", h->solname);
  if (S.eclipse) {                                         /* raygrid: incident angles in degrees */
    char *g = strdup(h->raygrid), *e;
    for (char *p = g; *p && S.nang < 16;) {
      const double v = strtod(p, &e);
      if (e == p) break;
      S.angles[S.nang++] = v;
      p = e;
      while (*p == ' ' || *p == ',' || *p == '\t') p++;
    }
    free(g);
    if (S.nang < 1) th_die("raygrid '%s' holds no angle.\n", h->raygrid);
  }
  int dev = 0;
#ifdef TH_TEST_TAPS
  th_dry = getenv("TRANSIT_B200_DRYRUN") != NULL;
#endif
  if (getenv("TRANSIT_B200_DEVICE")) dev = atoi(getenv("TRANSIT_B200_DEVICE"));
  CK(tb_create(dev, &S.ctx));
  th_make_wn(&S);                                          /* makewnsample */
  th_read_atmosphere(&S);                                  /* getatm */
  th_read_tli(&S);                                         /* readlineinfo */
  CK(tb_set_wn_grid(S.ctx, S.wn_i, S.wn_d, S.wn_o, S.nwn, S.onwn, 1.0));
  CK(tb_set_species(S.ctx, S.nmol, S.molmass, S.molradius, S.molpol, S.molID, S.niso, S.isomass, S.isoratio,
                    S.isoimol));
  th_make_rad(&S);                                         /* makeradsample */
  th_opacity(&S);                                          /* opacity */
  th_read_cs(&S);                                          /* readcs */
  apply_cloud_scattering();
  S.out = (double *)calloc(S.nwn, sizeof(double));
  S.initialised = 1;
}

int get_no_samples(void) { return (int)S.nwn; }

void get_waveno_arr(double *waveno_arr, int waveno) {
  for (int i = 0; i < (int)S.nwn && i < waveno; i++)
    waveno_arr[i] = S.initialised ? S.wn_i + (double)i * S.wn_d : -1;
  if (!S.initialised) printf("Transit not initialized, please run init. Values set -1\n");
}

void set_radius(double refradius) { S.h.r0 = refradius; }

void set_cloudtop(double cloudtop) {                       /* transit.c:103-109 */
  S.h.cl_top = cloudtop; S.h.cl_bot = cloudtop + 10; S.h.cl_ext = 100; S.h.cl_flag = 1;
  if (S.initialised) apply_cloud_scattering();
}

void set_scattering(int flag, double scattering) {
  S.h.sc_flag = flag; S.h.sc_logext = scattering;
  if (S.initialised) CK(tb_set_scattering(S.ctx, flag, scattering));
}

void do_transit(double *transit_out) {
  if (!S.initialised) { printf("Transit init not run, please initialize transit.\n"); return; }
  if (S.h.opabreak) return;                                /* justOpacity */
  const th_hints *h = &S.h;
#ifdef TH_TEST_TAPS
  if (th_dry) {                                            /* host-logic tap: the sampled state as doubles */
    FILE *f = fopen(getenv("TRANSIT_B200_DRYRUN"), "wb");
    if (!f) th_die("cannot write the dry-run dump\n");
    const double hdr[8] = {(double)S.nlayer, (double)S.nmol, (double)S.niso, (double)S.nwn, (double)S.onwn,
                           (double)S.nlines, S.wn_i, S.wn_d};
    fwrite(hdr, 8, 8, f);
    fwrite(S.rad, 8, S.nlayer, f); fwrite(S.p, 8, S.nlayer, f); fwrite(S.t, 8, S.nlayer, f);
    fwrite(S.mm, 8, S.nlayer, f); fwrite(S.d, 8, (size_t)S.nmol * S.nlayer, f);
    fwrite(S.Zlayer, 8, (size_t)S.niso * S.nlayer, f);
    fwrite(S.molmass, 8, S.nmol, f); fwrite(S.molradius, 8, S.nmol, f);
    fwrite(S.isomass, 8, S.niso, f); fwrite(S.isoratio, 8, S.niso, f);
    if (S.nlines) { fwrite(S.wl, 8, S.nlines, f); fwrite(S.elow, 8, S.nlines, f); fwrite(S.gf, 8, S.nlines, f); }
    fclose(f);
    return;
  }
#endif
  CK(tb_spectrum(S.ctx, S.eclipse ? TB_GEOM_ECLIPSE : TB_GEOM_TRANSIT, h->toomuch, h->ethresh,
                 h->starrad * TH_SUNRADIUS, h->transparent, S.eclipse ? S.angles : NULL, S.nang, S.out));
  if (transit_out) memcpy(transit_out, S.out, S.nwn * sizeof(double));
}

/* reloadatm (readatm.c:722-784): in = [T(nlayers) | q_mol0(nlayers) | q_mol1(nlayers) | ...] */
void run_transit(double *re_input, int transint, double *transit_out, int transit_out_size) {
  const int n = S.nlayer;
  (void)transint; (void)transit_out_size;
  for (int i = 0; i < n; i++) S.t[i] = re_input[i];
  for (int j = 0; j < S.nmol; j++)
    for (int i = 0; i < n; i++) S.q[(size_t)j * n + i] = re_input[(size_t)n * (j + 1) + i];
  th_densities(&S);
  th_radpress(&S);
  th_make_rad(&S);
  apply_cloud_scattering();
  do_transit(transit_out);
}

void free_memory(void) {
  if (S.ctx) tb_destroy(S.ctx);
  S.ctx = NULL;
  S.initialised = 0;
}

/* printmod (slantpath.c:510-555) / printflux (eclipse.c:350-380) */
void th_write_spectrum(void) {
  if (S.h.opabreak) return;
  FILE *f = stdout;
  if (S.h.f_outspec && S.h.f_outspec[0] != '-') {
    f = fopen(S.h.f_outspec, "w");
    if (!f) th_die("Cannot open output file '%s'.\n", S.h.f_outspec);
  }
  if (S.eclipse) {
    fprintf(f, "#wvl [um]%*sFlux [erg/s/cm]\n", 6, " ");
    for (int64_t i = 0; i < S.nwn; i++)
      fprintf(f, "%-15.10g%-18.9g\n", 1e4 / ((S.wn_i + (double)i * S.wn_d) / 1.0), S.out[i]);
  } else {
    fprintf(f, "#wvl [um]        modulation\n");
    for (int64_t i = 0; i < S.nwn; i++)
      fprintf(f, "%-17.9g%-18.9g\n", 1 / ((S.wn_i + (double)i * S.wn_d) / 1.0 * 1e-4), S.out[i]);
  }
  if (f != stdout) fclose(f);
}
