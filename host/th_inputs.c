/* th_inputs.c -- readers of the reference's input formats:
 *   atmosphere file   transit/src/readatm.c:276-620 (keywords ur/up/ut, q, z, n, #SPECIES; rows
 *                     radius pressure temperature q1..qN; bottom-to-top or reversed)
 *   molecules.dat     readatm.c:625-717 (ID name mass diameter source polarizability ...)
 *   TLI v6            readlineinfo.c:87-244 (header) and :416-537 (per-isotope wavelength range)
 *   CS / CIA files    crosssec.c:86-233 ("i sp1 [sp2]", "t T1 T2 ...", rows "wn v(T1) ...")
 * Error behaviour follows the reference: a message and exit(EXIT_FAILURE).
 */
#include <ctype.h>
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include "th_internal.h"

void th_die(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  fprintf(stderr, "transit_b200: ");
  vfprintf(stderr, fmt, ap);
  va_end(ap);
  exit(EXIT_FAILURE);
}

static char *skipblank(char *p) { while (*p == ' ' || *p == '\t') p++; return p; }
static char *nextfield(char *p) { while (*p && !isspace((unsigned char)*p)) p++; return skipblank(p); }
static void getname(const char *p, char *out) {
  int n = 0;
  while (*p == ' ' || *p == '\t') p++;
  while (*p && !isspace((unsigned char)*p) && n < TH_MAXNAME - 1) out[n++] = *p++;
  out[n] = 0;
}
static int countfields(const char *p) {
  int n = 0;
  while (*p) {
    while (*p && isspace((unsigned char)*p)) p++;
    if (*p) n++;
    while (*p && !isspace((unsigned char)*p)) p++;
  }
  return n;
}

/* ------------------------------------------------------------------ molecules.dat */
static void read_moldata(th_state *s) {
  FILE *fp = fopen(s->h.f_molfile, "r");
  if (!fp) th_die("Molecular info file '%s' cannot be opened.\n", s->h.f_molfile);
  char line[1024], name[TH_MAXNAME];
  int *found = (int *)calloc(s->nmol, sizeof(int));
  int in_block = 0;
  while (fgets(line, sizeof line, fp)) {
    if (line[0] == '#' || line[0] == '\n' || line[0] == 0) { if (in_block) break; continue; }
    in_block = 1;                                    /* the first block of data lines (readatm.c:645-656) */
    char *lp = line;
    const int id = (int)strtol(lp, NULL, 10);
    lp = nextfield(skipblank(lp));
    getname(lp, name);
    lp = nextfield(lp);
    const double mass = strtod(lp, NULL);
    lp = nextfield(lp);
    const double radius = strtod(lp, NULL) / 2.0;
    lp = nextfield(lp);                              /* radius source */
    lp = nextfield(lp);
    const double pol = strtod(lp, NULL);
    for (int i = 0; i < s->nmol; i++)
      if (!found[i] && strcmp(s->molname[i], name) == 0) {
        s->molID[i] = id; s->molmass[i] = mass;
        s->molradius[i] = radius * 1e-8;             /* Angstrom -> cm (readatm.c:706) */
        s->molpol[i] = pol;
        found[i] = 1;
      }
  }
  fclose(fp);
  for (int i = 0; i < s->nmol; i++)
    if (!found[i]) th_die("Molecule '%s' from the atmosphere file is not in '%s'.\n", s->molname[i], s->h.f_molfile);
  free(found);
}

/* ------------------------------------------------------------------ atmosphere file */
int th_read_atmosphere(th_state *s) {
  if (!s->h.f_atm || strcmp(s->h.f_atm, "-") == 0) th_die("getatm() :: No atmospheric file specified.\n");
  FILE *fp = fopen(s->h.f_atm, "r");
  if (!fp) th_die("Atmospheric info file '%s' cannot be opened.\n", s->h.f_atm);
  char line[4096], key[TH_MAXNAME];
  s->rad_fct = s->p_fct = s->t_fct = 1.0;
  s->zerorad = 0; s->bymass = 0; s->nmol = 0;
  long datapos = 0;
  for (;;) {                                          /* keyword section (readatm.c:288-385) */
    datapos = ftell(fp);
    if (!fgets(line, sizeof line, fp))
      th_die("readatm :: EOF unexpectedly found in '%s' while no t,p data points have been read.\n", s->h.f_atm);
    const char c = line[0];
    if (c == '\n') continue;
    if (c == '#') {
      getname(line + 1, key);
      if (strcmp(key, "SPECIES") == 0) {
        if (!fgets(line, sizeof line, fp)) th_die("readatm :: missing species line in '%s'.\n", s->h.f_atm);
        s->nmol = countfields(line);
        s->molname = (char(*)[TH_MAXNAME])calloc(s->nmol, TH_MAXNAME);
        char *lp = skipblank(line);
        for (int i = 0; i < s->nmol; i++) { getname(lp, s->molname[i]); lp = nextfield(lp); }
      }
      continue;
    }
    if (c == 'q') {
      const char k = *skipblank(line + 1) | 0x20;
      if (k == 'n') s->bymass = 0; else if (k == 'm') s->bymass = 1;
      else fprintf(stderr, "'q' option in the atmosphere file can only be followed by 'm' or 'n'.\n");
      continue;
    }
    if (c == 'z') { s->zerorad = atof(line + 1); continue; }
    if (c == 'u') {
      if (line[1] == 'r') s->rad_fct = atof(line + 2);
      else if (line[1] == 'p') s->p_fct = atof(line + 2);
      else if (line[1] == 't') s->t_fct = atof(line + 2);
      else th_die("Invalid unit factor indication in atmosphere file.\n");
      continue;
    }
    if (c == 'n') continue;                           /* free-text identifier */
    break;                                            /* first data row */
  }
  if (!s->nmol) th_die("No species were found in the atmosphere file '%s'.\n", s->h.f_atm);
  s->molID = (int *)calloc(s->nmol, sizeof(int));
  s->molmass = (double *)calloc(s->nmol, sizeof(double));
  s->molradius = (double *)calloc(s->nmol, sizeof(double));
  s->molpol = (double *)calloc(s->nmol, sizeof(double));
  read_moldata(s);

  /* qmol / qscale (argum.c:881-890, readatm.c:394-406): log10 scale factors for the named species, applied to every
   * data row as it is read; H2 and He are then renormalised at their own ratio so that the row adds up again */
  const int nmol = s->nmol;
  int nq = 0, qidx[64];
  double qsc[64];
  if (s->h.qscale) {
    const char *lp = s->h.qscale;
    while (*lp == ' ') lp++;
    while (*lp && nq < 64) {
      qsc[nq++] = atof(lp);
      while (*lp && *lp != ' ') lp++;
      while (*lp == ' ') lp++;
    }
    int nn = 0;
    lp = s->h.qmol ? s->h.qmol : "";
    while (*lp == ' ') lp++;
    while (*lp) {
      char nm[TH_MAXNAME];
      int k = 0;
      while (*lp && *lp != ' ') { if (k < TH_MAXNAME - 1) nm[k++] = *lp; lp++; }
      nm[k] = 0;
      while (*lp == ' ') lp++;
      if (nn < 64) {
        qidx[nn] = -1;
        for (int j = 0; j < nmol; j++) if (strcmp(nm, s->molname[j]) == 0) { qidx[nn] = j; break; }
      }
      nn++;
    }
    if (nn != nq) th_die("qscale (%d) and qmol (%d) should have the same number of elements.\n", nq, nn);
  }
  int iH2 = -1, iHe = -1;                            /* valueinarray(mol->ID, 105 | 2), readatm.c:465-466 */
  for (int j = 0; j < nmol; j++) { if (s->molID[j] == 105 && iH2 < 0) iH2 = j; if (s->molID[j] == 2 && iHe < 0) iHe = j; }
  if (nq > 0 && (iH2 < 0 || iHe < 0)) th_die("qscale needs H2 and He in the atmosphere file (readatm.c:536-541).\n");

  /* data rows (readatm.c:498-534) */
  fseek(fp, datapos, SEEK_SET);
  int cap = 64, n = 0;
  double *rad = (double *)malloc(cap * sizeof(double)), *p = (double *)malloc(cap * sizeof(double));
  double *t = (double *)malloc(cap * sizeof(double)), *q = (double *)malloc((size_t)cap * nmol * sizeof(double));
  while (fgets(line, sizeof line, fp)) {
    if (line[0] == '#' || line[0] == '\n') continue;
    if (n == cap) {
      cap *= 2;
      rad = (double *)realloc(rad, cap * sizeof(double)); p = (double *)realloc(p, cap * sizeof(double));
      t = (double *)realloc(t, cap * sizeof(double)); q = (double *)realloc(q, (size_t)cap * nmol * sizeof(double));
    }
    char *lp = line, *e;
    rad[n] = strtod(lp, &e) + s->zerorad;
    if (e == lp) th_die("Line of file '%s': field 1 (radius) does not have a valid value:\n%s", s->h.f_atm, line);
    lp = e; p[n] = strtod(lp, &e);
    if (e == lp) th_die("Line of file '%s': field 2 (pressure) does not have a valid value:\n%s", s->h.f_atm, line);
    lp = e; t[n] = strtod(lp, &e);
    if (e == lp) th_die("Line of file '%s': field 3 (temperature) does not have a valid value:\n%s", s->h.f_atm, line);
    if (rad[n] < 0 || p[n] < 0 || t[n] < 0) th_die("A negative value was found in atmosphere file %s.\n", s->h.f_atm);
    for (int i = 0; i < nmol; i++) {
      lp = e;
      q[(size_t)n * nmol + i] = strtod(lp, &e);
      if (e == lp) th_die("The line %s of file %s contains %d abundance values, when there were %d expected.\n",
                          line, s->h.f_atm, i, nmol);
      for (int j = 0; j < nq; j++)
        if (qidx[j] == i) { q[(size_t)n * nmol + i] *= pow(10.0, qsc[j]); break; }     /* readatm.c:519-522 */
      if (q[(size_t)n * nmol + i] < 0) th_die("A negative abundance was found in atmosphere file %s.\n", s->h.f_atm);
    }
    if (nq > 0) {                                      /* readatm.c:536-541 */
      double sumq2 = 0;
      for (int i = 0; i < nmol; i++) if (i != iH2 && i != iHe) sumq2 += q[(size_t)n * nmol + i];
      const double ratio = q[(size_t)n * nmol + iH2] / q[(size_t)n * nmol + iHe];
      q[(size_t)n * nmol + iHe] = (1 - sumq2) / (1.0 + ratio);
      q[(size_t)n * nmol + iH2] = ratio * (1 - sumq2) / (1.0 + ratio);
    }
    n++;
  }
  fclose(fp);
  if (n < 1) th_die("No layers in atmosphere file '%s'.\n", s->h.f_atm);
  /* bottom-to-top order (readatm.c:585-615) */
  int sorted = 1, reversed = 1;
  for (int i = 0; i < n - 1; i++) {
    if (rad[i] >= rad[i + 1] || p[i] <= p[i + 1]) sorted = 0;
    if (rad[i] <= rad[i + 1] || p[i] >= p[i + 1]) reversed = 0;
  }
  if (!sorted && !reversed && n > 1)
    th_die("The atmospheric layers are neither sorted from the bottom up, nor from the top down.\n");
  s->nlayer = n;
  s->rad = (double *)malloc(n * sizeof(double)); s->p = (double *)malloc(n * sizeof(double));
  s->t = (double *)malloc(n * sizeof(double)); s->mm = (double *)malloc(n * sizeof(double));
  s->q = (double *)malloc((size_t)n * nmol * sizeof(double)); s->d = (double *)malloc((size_t)n * nmol * sizeof(double));
  for (int i = 0; i < n; i++) {
    const int src = (reversed && n > 1) ? n - 1 - i : i;
    s->rad[i] = rad[src]; s->p[i] = p[src]; s->t[i] = t[src];
    for (int j = 0; j < nmol; j++) s->q[(size_t)j * n + i] = q[(size_t)src * nmol + j];
  }
  free(rad); free(p); free(t); free(q);
  return th_densities(s);
}

static int access_ok(const char *path) {
  FILE *f = fopen(path, "rb");
  if (!f) return 0;
  fclose(f);
  return 1;
}

/* ------------------------------------------------------------------ TLI v6 */
static void rd(void *dst, size_t sz, size_t n, FILE *fp, const char *what) {
  if (fread(dst, sz, n, fp) != n) th_die("TLI file ended while reading %s.\n", what);
}

/* datafileBS (readlineinfo.c:16-77) on an ascending segment of the memory-mapped file; the TLI header
 * has no alignment, so the doubles are fetched with memcpy */
static inline double tli_at(const unsigned char *v, int64_t i) {
  double x;
  memcpy(&x, v + 8 * i, 8);
  return x;
}
static int64_t tli_bs(const unsigned char *v, int64_t nfields, double target, int up) {
  int64_t lo = 0, hi = nfields - 1, loc;
  do {
    loc = (hi + lo) / 2;
    if (target > tli_at(v, loc)) lo = loc; else hi = loc;
  } while (hi - lo > 1);
  if (up) {
    loc = lo;
    while (loc < nfields - 1) { if (tli_at(v, loc + 1) > target) break; loc++; }
  } else {
    loc = hi;
    while (loc > 0) { if (tli_at(v, loc - 1) < target) break; loc--; }
  }
  return loc;
}

int th_read_tli(th_state *s) {
  s->ndb = s->niso = 0; s->nlines = 0;
  s->tli_tmin = 0; s->tli_tmax = 1e30;
  if (!s->h.f_line) return 0;                          /* opacity file only */
  FILE *fp = fopen(s->h.f_line, "rb");
  if (!fp) th_die("Line info file '%s' is not available.\n", s->h.f_line);
  int32_t magic;
  uint16_t ver[3], ndb, rs, nT, nI;
  double wi, wf;
  rd(&magic, 4, 1, fp, "magic bytes");
  if (magic != (int32_t)0xabb3b6ff) th_die("The file '%s' is not a TLI file (wrong magic bytes).\n", s->h.f_line);
  rd(ver, 2, 3, fp, "version");
  if (ver[0] != 6) th_die("The TLI version of '%s' is %d; this program reads version 6 (transit.h:30).\n", s->h.f_line, ver[0]);
  rd(&wi, 8, 1, fp, "wavelength range"); rd(&wf, 8, 1, fp, "wavelength range");
  rd(&ndb, 2, 1, fp, "number of databases");
  s->ndb = ndb;
  s->db = (th_db *)calloc(ndb, sizeof(th_db));
  int niso = 0, cap = 16;
  int64_t zcap = 4096, zn = 0;
  s->isomass = (double *)malloc(cap * sizeof(double)); s->isoratio = (double *)malloc(cap * sizeof(double));
  s->isodb = (int *)malloc(cap * sizeof(int)); s->isoZoff = (int64_t *)malloc(cap * sizeof(int64_t));
  s->isoZ = (double *)malloc(zcap * sizeof(double));
  double tmin = 0, tmax = 1e30;
  for (int i = 0; i < ndb; i++) {
    th_db *db = s->db + i;
    rd(&rs, 2, 1, fp, "DB name"); if (rs >= TH_MAXNAME) th_die("TLI database name too long.\n");
    rd(db->name, 1, rs, fp, "DB name");
    rd(&rs, 2, 1, fp, "molecule name"); if (rs >= TH_MAXNAME) th_die("TLI molecule name too long.\n");
    rd(db->molname, 1, rs, fp, "molecule name");
    rd(&nT, 2, 1, fp, "temperature count"); rd(&nI, 2, 1, fp, "isotope count");
    db->nT = nT; db->niso = nI; db->first = niso;
    db->T = (double *)malloc(nT * sizeof(double));
    rd(db->T, 8, nT, fp, "temperatures");
    if (db->T[0] > tmin) tmin = db->T[0];               /* readlineinfo.c: li->tmin/tmax */
    if (db->T[nT - 1] < tmax) tmax = db->T[nT - 1];
    for (int j = 0; j < nI; j++) {
      if (niso == cap) {
        cap *= 2;
        s->isomass = (double *)realloc(s->isomass, cap * sizeof(double));
        s->isoratio = (double *)realloc(s->isoratio, cap * sizeof(double));
        s->isodb = (int *)realloc(s->isodb, cap * sizeof(int));
        s->isoZoff = (int64_t *)realloc(s->isoZoff, cap * sizeof(int64_t));
      }
      char nm[256];
      rd(&rs, 2, 1, fp, "isotope name"); if (rs >= sizeof nm) th_die("TLI isotope name too long.\n");
      rd(nm, 1, rs, fp, "isotope name");
      rd(s->isomass + niso, 8, 1, fp, "isotope mass"); rd(s->isoratio + niso, 8, 1, fp, "isotope ratio");
      while (zn + nT > zcap) { zcap *= 2; s->isoZ = (double *)realloc(s->isoZ, zcap * sizeof(double)); }
      rd(s->isoZ + zn, 8, nT, fp, "partition function");
      s->isoZoff[niso] = zn; zn += nT;
      s->isodb[niso] = i;
      niso++;
    }
  }
  s->niso = niso; s->tli_tmin = tmin; s->tli_tmax = tmax;
  /* setimol (readlineinfo.c:246-278): isotope -> molecule of the atmosphere file */
  s->isoimol = (int *)malloc((niso ? niso : 1) * sizeof(int));
  for (int i = 0; i < niso; i++) {
    s->isoimol[i] = -1;
    for (int j = 0; j < s->nmol; j++)
      if (strcmp(s->db[s->isodb[i]].molname, s->molname[j]) == 0) { s->isoimol[i] = j; break; }
    if (s->isoimol[i] < 0)
      th_die("The molecule '%s' of the line database is not in the atmosphere file.\n", s->db[s->isodb[i]].molname);
  }
  /* checkrange (readlineinfo.c:280-340), wavelengths in cm */
  const double wlmin = 1.0 / s->wn_f, wlmax = 1.0 / s->wn_i;
  if (wi * 1e-4 > wlmax) th_die("Database minimum wavelength (%g cm) is larger than the requested maximum (%g cm).\n", wi * 1e-4, wlmax);
  if (wf * 1e-4 < wlmin) th_die("Database maximum wavelength (%g cm) is smaller than the requested minimum (%g cm).\n", wf * 1e-4, wlmin);
  if (s->h.f_opa && access_ok(s->h.f_opa)) { fclose(fp); return 0; }   /* grid exists: lines are not needed */

  uint64_t nl;
  int32_t ni;
  rd(&nl, 8, 1, fp, "number of transitions"); rd(&ni, 4, 1, fp, "number of isotopes with lines");
  uint64_t *cnt = (uint64_t *)malloc((ni ? ni : 1) * sizeof(uint64_t));
  rd(cnt, 8, ni, fp, "transitions per isotope");
  /* readdatarng (readlineinfo.c:416-537): per isotope keep [1/wn_f, 1/wn_i] (microns).  The four
   * arrays (wavelength, isotope, E_low, gf; nl entries each) are NOT read: the file is mapped, the
   * two ends of every isotope's block are found by datafileBS's search in the mapped wavelengths, and
   * only the selected slices are copied -- an ExoMol-scale TLI (10^8 lines, 2.6 GB) costs what the
   * requested wavenumber range holds. */
  const long hdr_end = ftell(fp);
  fseek(fp, 0, SEEK_END);
  const long fsize = ftell(fp);
  if (hdr_end < 0 || fsize < 0 || (uint64_t)(fsize - hdr_end) < nl * 26)
    th_die("TLI file ended while reading the line transitions.\n");
  const unsigned char *map = NULL;
  if (nl) {
    map = (const unsigned char *)mmap(NULL, (size_t)fsize, PROT_READ, MAP_PRIVATE, fileno(fp), 0);
    if (map == MAP_FAILED) th_die("Cannot map the line info file '%s'.\n", s->h.f_line);
  }
  const unsigned char *m_wl = map + hdr_end, *m_id = m_wl + 8 * nl, *m_el = m_id + 2 * nl, *m_gf = m_el + 8 * nl;
  const double iniw = 1.0 / (s->wn_f * 1.0) / 1e-4, finw = 1.0 / (s->wn_i * 1.0) / 1e-4;
  int64_t *lo = (int64_t *)malloc((ni ? ni : 1) * sizeof(int64_t)), *cntsel = (int64_t *)calloc(ni ? ni : 1, sizeof(int64_t));
  int64_t off = 0, n = 0;
  for (int i = 0; i < ni; i++) {
    if (cnt[i] > 0) {
      const int64_t a = tli_bs(m_wl + 8 * off, (int64_t)cnt[i], iniw, 0) + off, b = tli_bs(m_wl + 8 * off, (int64_t)cnt[i], finw, 1) + off;
      if (b - a + 1 > 0) { lo[i] = a; cntsel[i] = b - a + 1; n += b - a + 1; }
    }
    off += (int64_t)cnt[i];
  }
  double *wl = (double *)malloc((n ? n : 1) * sizeof(double)), *el = (double *)malloc((n ? n : 1) * sizeof(double));
  double *gf = (double *)malloc((n ? n : 1) * sizeof(double));
  int16_t *id = (int16_t *)malloc((n ? n : 1) * sizeof(int16_t));
  int64_t k = 0;
  for (int i = 0; i < ni; i++) {
    const int64_t c = cntsel[i];
    if (!c) continue;
    memcpy(wl + k, m_wl + 8 * lo[i], 8 * c); memcpy(el + k, m_el + 8 * lo[i], 8 * c);
    memcpy(gf + k, m_gf + 8 * lo[i], 8 * c); memcpy(id + k, m_id + 2 * lo[i], 2 * c);
    k += c;
  }
  if (map) munmap((void *)map, (size_t)fsize);
  fclose(fp);
  free(lo); free(cntsel);
  free(cnt);
  s->nlines = n; s->wl = wl; s->elow = el; s->gf = gf; s->isoid = id;
  return 0;
}

/* ------------------------------------------------------------------ CS / CIA files */
int th_read_cs(th_state *s) {
  s->ncs = 0;
  if (TH_DEV(tb_cia_clear(s->ctx))) th_die("%s\n", tb_last_error());
  if (!s->h.csfiles || !*s->h.csfiles) return 0;
  char *list = strdup(s->h.csfiles);
  for (char *file = strtok(list, ","); file; file = strtok(NULL, ",")) {
    while (*file == ' ') file++;
    FILE *fp = fopen(file, "r");
    if (!fp) th_die("Cannot open Cross-section file '%s'.\n", file);
    char *line = NULL, name[TH_MAXNAME];
    size_t cap = 0;
    int nspec = 0, nt = 0, mol[2] = {-1, -1}, have_row = 0;
    double temps[256];
    while (getline(&line, &cap, fp) > 0) {               /* header (crosssec.c:98-160) */
      if (line[0] == '#' || line[0] == '\n') continue;
      if (line[0] == 'i') {
        char *lp = skipblank(line + 1);
        nspec = countfields(lp);
        if (nspec != 1 && nspec != 2) th_die("Wrong line in CS file '%s': species line must name 1 or 2 species.\n", file);
        for (int k = 0; k < nspec; k++) {
          getname(lp, name);
          for (int i = 0; i < s->nmol; i++) if (strcmp(s->molname[i], name) == 0) mol[k] = i;
          if (mol[k] < 0) th_die("CS species '%s' was not found in the atmosphere file.\n", name);
          lp = nextfield(lp);
        }
        continue;
      }
      if (line[0] == 't') {
        char *lp = skipblank(line + 1), *e;
        nt = countfields(lp);
        if (!nt || nt > 256) th_die("Wrong line in CS file '%s': bad temperature list.\n", file);
        for (int n = 0; n < nt; n++) {
          temps[n] = strtod(lp, &e);
          if (e == lp) th_die("Invalid temperature field %d in CS file '%s'.\n", n + 1, file);
          if ((*e | 0x20) == 'k') e++;
          lp = e;
        }
        continue;
      }
      have_row = 1;
      break;
    }
    if (!nspec || !nt || !have_row) th_die("CS file '%s' ended before any data row.\n", file);
    int wa = 1024, n = 0;
    double *wn = (double *)malloc(wa * sizeof(double)), *a = (double *)malloc((size_t)wa * nt * sizeof(double));
    do {                                                   /* data rows (crosssec.c:190-233) */
      if (line[0] == '#' || line[0] == '\n') continue;
      if (n == wa) {
        wa *= 2;
        wn = (double *)realloc(wn, wa * sizeof(double)); a = (double *)realloc(a, (size_t)wa * nt * sizeof(double));
      }
      char *lp = line, *e;
      wn[n] = strtod(lp, &e);
      if (e == lp) th_die("Invalid wavenumber in CS file '%s', data row %d.\n", file, n + 1);
      for (int i = 0; i < nt; i++) {
        lp = e;
        a[(size_t)n * nt + i] = strtod(lp, &e);
        if (e == lp) th_die("Invalid cross-section field %d in CS file '%s', data row %d.\n", i + 1, file, n + 1);
      }
      n++;
    } while (getline(&line, &cap, fp) > 0);
    fclose(fp);
    free(line);
    /* wavenumber boundaries check (crosssec.c:250-259) */
    const double w0 = s->wn_i, w1 = s->wn_i + (double)(s->nwn - 1) * s->wn_d;
    if (wn[0] > w0 || wn[n - 1] < w1)
      th_die("Wavenumber range [%.2f, %.2f] of CS file '%s' does not cover Transit's range [%.2f, %.2f].\n",
             wn[0], wn[n - 1], file, w0, w1);
    if (TH_DEV(tb_cia_add_table(s->ctx, wn, n, temps, nt, a, mol, nspec))) th_die("%s\n", tb_last_error());
    free(wn); free(a);
    s->ncs++;
  }
  free(list);
  return 0;
}
