/* th_api.c -- the reference's library interface (transit/src/transit.c:14-228, transit.i:11-19)
 * on top of libtransit_b200.so: same function names, argument meaning and error behaviour
 * (messages on stderr and exit(EXIT_FAILURE); no return codes), one global state like the
 * reference's `struct transit transit` (transit.c:7).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>
#include <unistd.h>
#include "th_internal.h"

static th_state S;
int th_rank = 0, th_world = 1, th_idfd[64];
#ifdef TH_TEST_TAPS
int th_dry = 0;
#endif

#define CK(call) do { if (TH_DEV(call) != TB_OK) th_die("%s\n", tb_last_error()); } while (0)

tb_ctx *th_context(void) { return S.ctx; }
/* sizes of the problem behind the library face: layers, molecules, isotopes, wavenumbers, lines, angles */
void th_sizes(int64_t *out) {
  out[0] = S.resampled ? S.other.n : S.nlayer;    /* layers of the atmosphere file: what run_transit's input holds */
  out[1] = S.nmol; out[2] = S.niso; out[3] = S.nwn; out[4] = S.nlines; out[5] = S.nang;
}

/* timecheck (transitstd.c:359-374): the "Check point" lines users grep, printed at verbosity >= 4
 * (TOUT_RESULT) with the reference's iteration / stage numbering (transit.c:33-66,143-199).  The first
 * line carries seconds since the epoch, as the reference's does (its t0 starts at 0).                 */
static double th_t0 = 0;
static long th_itr = 0;
static void timecheck(long index, const char *str) {
  struct timeval tv;
  gettimeofday(&tv, NULL);
  const double sec = tv.tv_sec + 1e-6 * tv.tv_usec;
  if (S.h.verb >= 4) printf("Check point: %02li - %02li %s:  dt = %.4f sec.\n\n", th_itr, index, str, sec - th_t0);
  th_t0 = sec;
}

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
  timecheck(0, "processparameters");
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
  int dev = th_rank;                                       /* one process per GPU: rank r drives device r */
#ifdef TH_TEST_TAPS
  th_dry = getenv("TRANSIT_B200_DRYRUN") != NULL;
#endif
  if (getenv("TRANSIT_B200_DEVICE")) dev = atoi(getenv("TRANSIT_B200_DEVICE"));
  CK(tb_create(dev, &S.ctx));
  th_make_wn(&S);                                          /* makewnsample */
  timecheck(1, "makewnsample");
  th_read_atmosphere(&S);                                  /* getatm */
  timecheck(2, "getatm");
  th_read_tli(&S);                                         /* readlineinfo */
  timecheck(3, "readlineinfo");
  CK(tb_set_wn_grid(S.ctx, S.wn_i, S.wn_d, S.wn_o, S.nwn, S.onwn, 1.0));
  CK(tb_set_species(S.ctx, S.nmol, S.molmass, S.molradius, S.molpol, S.molID, S.niso, S.isomass, S.isoratio,
                    S.isoimol));
  th_make_rad(&S);                                         /* makeradsample */
  timecheck(4, "makeradsample");
  S.band_lo = 0; S.band_hi = S.nwn;
  if (th_world > 1) {
    /* wavenumber-band sharding over the GPUs of the box: rank 0 makes the NCCL id and passes it
     * down the pipes th_main.c opened before forking; every rank joins the communicator */
    char id[128];
    if (th_rank == 0) {
      CK(tb_comm_unique_id(id));
      for (int r = 1; r < th_world; r++)
        if (write(th_idfd[r], id, 128) != 128) th_die("cannot hand the communicator id to rank %d\n", r);
    } else if (read(th_idfd[th_rank], id, 128) != 128)
      th_die("rank %d: no communicator id from rank 0\n", th_rank);
    CK(tb_comm_init(S.ctx, th_rank, th_world, id));
  }
  th_opacity(&S);                                          /* opacity */
  timecheck(5, "opacity");
  th_read_cs(&S);                                          /* readcs */
  timecheck(6, "readcs");
  if (h->modlevel != 1 && h->modlevel != -1)
    th_die("slantpath:: modulationperwn:: Level %i of detail has not been implemented to compute modulation.\n", h->modlevel);
  apply_cloud_scattering();
  if (S.have_grid_file && !S.resampled && h->p0 != 0 && h->r0 != 0 && h->gsurf != 0)      /* what run_transit_batch needs */
    CK(tb_set_atmosphere_model(S.ctx, S.nlayer, S.p, S.p_fct, S.t_fct, S.rad_fct, S.bymass, h->p0, h->r0, h->gsurf));
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

/* printsample / outsample (makesample.c:642-672,743-772), written only with `savefiles yes` as in the
 * reference (makesample.c:598); struct transit's `wavs` sample is never filled there, hence the zeros */
static void write_outsample(void) {
  FILE *out = stdout;
  if (strcmp(S.h.f_outsample, "-") != 0 && !(out = fopen(S.h.f_outsample, "w"))) {
    fprintf(stderr, "Cannot open file '%s' for writing sampling data.\n", S.h.f_outsample);
    return;
  }
  const int n = S.nlayer;
  const char *hdr = "############################\n   %-12s Sampling\n----------------------------\n";
  fprintf(out, hdr, "Wavenumber");
  fprintf(out, "Factor to cgs units: %g\nInitial value: %g\nFinal value: %g\nSpacing: %g\nOversample: %i\n"
               "Number of elements: %lli\n", 1.0, S.wn_i, S.wn_f, S.wn_d, 1, (long long)S.nwn);
  fprintf(out, hdr, "Wavelength");
  fprintf(out, "Factor to cgs units: %g\nInitial value: %g\nFinal value: %g\nSpacing: %g\nOversample: %i\n"
               "Number of elements: %lli\n", 0.0, 0.0, 0.0, 0.0, 0, 0LL);
  fprintf(out, hdr, "Radius");
  fprintf(out, "Factor to cgs units: %g\nInitial value: %g\nFinal value: %g\nSpacing: %g\n"
               "Number of elements: %lli\nValues: ", S.rad_fct, S.rad[0], S.resampled ? S.samp_f : S.rad[n - 1],
          S.resampled ? S.h.rad_d : 0.0, (long long)n);
  for (int i = 0; i < n; i++) fprintf(out, " %12.8g", S.rad[i]);
  fprintf(out, "\n");
  fprintf(out, hdr, "Impact parameter");
  fprintf(out, "Factor to cgs units: %g\nInitial value: %g\nFinal value: %g\nSpacing: %g\nOversample: %i\n"
               "Number of elements: %lli\nValues: ", S.rad_fct, S.rad[n - 1], S.rad[0], S.resampled ? -S.h.rad_d : 0.0,
          S.resampled ? 1 : 0, (long long)(S.resampled ? S.nip : n));
  for (int i = 0; i < (S.resampled ? S.nip : n); i++) fprintf(out, " %12.8g", S.resampled ? S.ips[i] : S.rad[n - 1 - i]);
  fprintf(out, "\n");
  if (out != stdout) fclose(out);
}

/* savefile_extinct (extinction.c:62-91): "@E@S@", e[rad][wn] as doubles, computed[rad] as shorts
 * (#define _Bool short, transit.h:128-130).  The reference evaluates layers lazily, so rows it never
 * reached hold zeros there; the same rows are zeroed here.  An existing file is overwritten: the
 * restore half (restfile_extinct, an optimisation of repeated runs) is not needed by a path that
 * recomputes every layer in a millisecond.                                                        */
static void write_saveext(void) {
  const int n = S.nlayer;
  double *e = (double *)malloc((size_t)n * S.nwn * sizeof(double));
  short *comp = (short *)calloc(n, sizeof(short));
  int16_t *c16 = (int16_t *)calloc(n, sizeof(int16_t));
  if (!e || !comp || !c16) th_die("Not enough host memory for the extinction savefile.\n");
  CK(tb_get_extinction(S.ctx, e));
  CK(tb_get_computed(S.ctx, c16));
  FILE *fp = fopen(S.h.f_saveext, "w");
  if (!fp) {
    fprintf(stderr, "Extinction savefile '%s' cannot be opened for writing.\nContinuing without saving\n", S.h.f_saveext);
  } else {
    for (int r = 0; r < n; r++) {
      comp[r] = c16[r];
      if (!c16[r]) memset(e + (size_t)r * S.nwn, 0, S.nwn * sizeof(double));
    }
    fwrite("@E@S@", 1, 5, fp);
    fwrite(e, sizeof(double), (size_t)n * S.nwn, fp);
    fwrite(comp, sizeof(short), n, fp);
    fclose(fp);
  }
  free(e); free(comp); free(c16);
}

/* printtoomuch (tau.c:611-640) */
static void write_toomuch(void) {
  const int n = S.resampled && !S.eclipse ? S.nip : S.nlayer;   /* rays: the layers, or the resampled impact parameters */
  double *t = (double *)malloc((size_t)S.nwn * n * sizeof(double));
  int64_t *last = (int64_t *)malloc(S.nwn * sizeof(int64_t));
  if (!t || !last) th_die("Not enough host memory for the toomuch file.\n");
  CK(tb_get_tau(S.ctx, t, last));
  FILE *out = stdout;
  if (S.h.f_toomuch[0] != '-') out = fopen(S.h.f_toomuch, "w");
  if (!out) {
    fprintf(stderr, "Cannot open '%s' for writing depth where the optical depth reached toomuch.\n", S.h.f_toomuch);
  } else {
    fprintf(out, "# Wavelength   Max Optical   Radius at the    Radius\n"
                 "   (microns)         depth   max depth (km)    index\n");
    for (int64_t w = 0; w < S.nwn; w++) {
      const double wn = S.wn_i + (double)w * S.wn_d;
      /* the rays are the impact parameters, top first: ips.v[i] = rads.v[n-1-i] (makesample.c:564-573) */
      fprintf(out, "%12.7f   %.5e     %12.4f     %04ld\n", 1.0 / wn * 1.0 * 1e4, t[(size_t)w * n + last[w]],
              (S.resampled ? S.ips[last[w]] : S.rad[n - 1 - last[w]]) * S.rad_fct / 1e5, (long)last[w]);
    }
    if (out != stdout) fclose(out);
  }
  free(t); free(last);
}

/* printintens (eclipse.c:293-348) */
static void write_intens(void) {
  double *a = (double *)malloc((size_t)S.nang * S.nwn * sizeof(double));
  if (!a) th_die("Not enough host memory for the intensity file.\n");
  CK(tb_get_intensity(S.ctx, a));
  FILE *outf = fopen(S.h.f_outintens, "w");
  if (!outf) th_die("Cannot open intensity file '%s'.\n", S.h.f_outintens);
  fprintf(outf, "#wvl %*s", 10, " ");
  for (int i = 0; i < S.nang; i++) fprintf(outf, "I[%4.1lf deg]%*s", S.angles[i], 7, " ");
  fprintf(outf, "\n#[um]%*s", 10, " ");
  for (int i = 0; i < S.nang; i++) fprintf(outf, "[erg/s/cm/sr]%*s", 5, " ");
  fprintf(outf, "\n");
  for (int64_t w = 0; w < S.nwn; w++) {
    fprintf(outf, "%-15.10g", 1e4 / ((S.wn_i + (double)w * S.wn_d) / 1.0));
    for (int i = 0; i < S.nang; i++) fprintf(outf, "%-18.9g", a[(size_t)i * S.nwn + w]);
    fprintf(outf, "\n");
  }
  fclose(outf);
  free(a);
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
  if (h->savefiles && h->f_outsample) write_outsample();   /* makeipsample, makesample.c:598-599 */
  timecheck(6, "makeipsample");
  CK(tb_set_modlevel(S.ctx, h->modlevel));
  if (th_world > 1 && !S.rebalanced) {     /* one spectrum to measure the bands, cut again by measured cost, then the real one */
    CK(tb_spectrum(S.ctx, S.eclipse ? TB_GEOM_ECLIPSE : TB_GEOM_TRANSIT, h->toomuch, h->ethresh,
                   h->starrad * TH_SUNRADIUS, h->transparent, S.eclipse ? S.angles : NULL, S.nang, S.out));
    if (!S.have_grid_file) CK(tb_rebalance_bands(S.ctx));
    S.rebalanced = 1;
  }
  CK(tb_spectrum(S.ctx, S.eclipse ? TB_GEOM_ECLIPSE : TB_GEOM_TRANSIT, h->toomuch, h->ethresh,
                 h->starrad * TH_SUNRADIUS, h->transparent, S.eclipse ? S.angles : NULL, S.nang, S.out));
  if (h->verb >= 4) {      /* one device call covers the reference's stages: their share of it, from the device timers */
    float ms;
    const char *stage[][2] = {{"cia", "interpcs"}, {"molext", "extwn + lazy extinction (tau.c:238-271)"},
                              {"interp", "interpolmolext"}, {"tau", S.eclipse ? "tau eclipse" : "tau transit"},
                              {"intensity", "emergent intensity + flux"}, {"modulation", "modulation"}};
    for (int k = 0; k < 6; k++)
      if (tb_get_kernel_ms(S.ctx, stage[k][0], &ms) == TB_OK && ms > 0)
        printf("  device: %-44s %.4f sec.\n", stage[k][1], 1e-3 * ms);
  }
  timecheck(9, "interpcs");
  timecheck(10, "idxrefrac");
  timecheck(11, "extwn");
  timecheck(12, S.eclipse ? "tau eclipse" : "tau transit");
  if (h->f_saveext) write_saveext();                       /* tau.c:340-341 */
  if (h->f_toomuch) write_toomuch();                       /* tau.c:344-345 */
  if (S.eclipse) {
    for (int a = 0; a < S.nang; a++) timecheck(13, "emergent intensity");
    if (h->f_outintens && h->f_outintens[0] != '-') write_intens();   /* eclipse.c:230-231 */
    timecheck(14, "flux");
  } else
    timecheck(13, "modulation");
  if (transit_out) memcpy(transit_out, S.out, S.nwn * sizeof(double));
  timecheck(14, "THE END");
  th_itr++;
}

/* reloadatm (readatm.c:722-784): in = [T(nlayers) | q_mol0(nlayers) | q_mol1(nlayers) | ...] */
void run_transit(double *re_input, int transint, double *transit_out, int transit_out_size) {
  th_file_layers(&S);                              /* raddelt != -1: reloadatm updates the atmosphere FILE's layers */
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

/* run_transit for natm atmospheres at once: re_input[natm][nlayers*(1+nmol)], transit_out[natm][nwn].  From a
 * resident opacity grid in transit geometry the whole batch goes through the device in one pass (reloadatm's
 * densities and hydrostatic radii included, tb_run_transit_batch); any other configuration is a loop of
 * run_transit.  An extension of the reference's interface (it has no batched call), same input layout. */
void run_transit_batch(double *re_input, int natm, double *transit_out) {
  const int nin = (S.resampled ? S.other.n : S.nlayer) * (1 + S.nmol);
  if (!S.initialised) { printf("Transit init not run, please initialize transit.\n"); return; }
  if (S.have_grid_file && !S.eclipse && !S.resampled && S.h.modlevel == 1 && S.h.p0 != 0 && S.h.r0 != 0 && S.h.gsurf != 0 && th_world == 1) {
    CK(tb_run_transit_batch(S.ctx, natm, re_input, S.h.toomuch, S.h.starrad * TH_SUNRADIUS, S.h.transparent, transit_out));
    return;
  }
  for (int a = 0; a < natm; a++) run_transit(re_input + (size_t)a * nin, nin, transit_out + (size_t)a * S.nwn, (int)S.nwn);
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
