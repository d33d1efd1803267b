/* th_config.c -- the reference's option table and cfg syntax (transit/src/argum.c:112-320,
 * pu/src/procopt.c:271-398,640-700):
 *   - defaults are delivered first, as if typed;
 *   - `-c file` / `--config_file file` reads `key<blanks>value` lines, '#' and empty lines skipped;
 *     the key is matched by PREFIX against the option names IN TABLE ORDER, first match wins
 *     (that is how the demo's `ethresh` selects `ethreshold`);
 *   - `--key value` on the command line follows getopt_long: exact name, else unique prefix.
 * Options of the reference that do not touch the spectrum path are accepted and ignored.
 */
#include <ctype.h>
#include <stdlib.h>
#include <string.h>
#include "th_internal.h"

enum { O_IGN = 1, O_VERB, O_QUIET, O_CFG, O_ATM, O_LINEDB, O_OUTTOOMUCH, O_OUTSAMPLE, O_OUTSPEC, O_OUTINTENS,
       O_MOLFILE, O_RADDELT, O_RADLOW, O_RADHIGH, O_RADFCT, O_ALLOWQ, O_RPRESS, O_RRADIUS, O_GSURF, O_QMOL, O_QSCALE, O_WLLOW, O_WLHIGH,
       O_WLFCT, O_WNLOW, O_WNHIGH, O_WNDELT, O_WNOSAMP, O_WNFCT, O_NDOP, O_NLOR, O_DMIN, O_DMAX, O_LMIN,
       O_LMAX, O_NWIDTH, O_ETHRESH, O_CLOUD, O_CLOUDTOP, O_SCATTERING, O_CSFILE, O_OPACITYFILE, O_TLOW,
       O_THIGH, O_TEMPDELT, O_OPABREAK, O_SOLUTION, O_TOOMUCH, O_STARRAD, O_TRANSPARENT, O_RAYGRID,
       O_VERSION, O_HELP, O_SAVEEXT, O_SAVEFILES, O_MODLEVEL };

typedef struct { const char *name; char shortc; int has_arg; const char *def; int id; } th_opt;

/* same names, same order, same defaults as argum.c:112-320 */
static const th_opt table[] = {
  {"version", 'V', 0, NULL, O_VERSION}, {"help", 'h', 0, NULL, O_HELP}, {"quiet", 'q', 0, NULL, O_QUIET},
  {"verb", 'v', 1, "2", O_VERB}, {"config_file", 'c', 1, NULL, O_CFG},
  {"atm", 0, 1, NULL, O_ATM}, {"linedb", 0, 1, NULL, O_LINEDB}, {"outtoomuch", 0, 1, NULL, O_OUTTOOMUCH},
  {"outsample", 0, 1, NULL, O_OUTSAMPLE}, {"outspec", 0, 1, "outspectrum", O_OUTSPEC},
  {"outintens", 0, 1, NULL, O_OUTINTENS}, {"molfile", 0, 1, "../inputs/molecules.dat", O_MOLFILE},
  {"savefiles", 0, 1, "no", O_SAVEFILES},
  {"raddelt", 0, 1, "-1", O_RADDELT}, {"radlow", 0, 1, "0", O_RADLOW}, {"radhigh", 0, 1, "0", O_RADHIGH},
  {"radfct", 0, 1, "0", O_RADFCT},
  {"allowq", 0, 1, "0.00001", O_ALLOWQ}, {"refpress", 0, 1, NULL, O_RPRESS}, {"refradius", 0, 1, NULL, O_RRADIUS},
  {"gsurf", 0, 1, NULL, O_GSURF}, {"qmol", 0, 1, NULL, O_QMOL}, {"qscale", 0, 1, NULL, O_QSCALE},
  {"wllow", 0, 1, NULL, O_WLLOW}, {"wlhigh", 0, 1, NULL, O_WLHIGH}, {"wlfct", 0, 1, "1e-4", O_WLFCT},
  {"wnlow", 0, 1, NULL, O_WNLOW}, {"wnhigh", 0, 1, NULL, O_WNHIGH}, {"wndelt", 0, 1, "0", O_WNDELT},
  {"wnosamp", 0, 1, "2160", O_WNOSAMP}, {"wnfct", 0, 1, "0", O_WNFCT},
  {"ndop", 0, 1, "60", O_NDOP}, {"nlor", 0, 1, "60", O_NLOR}, {"dmin", 0, 1, "1e-3", O_DMIN},
  {"dmax", 0, 1, "0.25", O_DMAX}, {"lmin", 0, 1, "1e-4", O_LMIN}, {"lmax", 0, 1, "10.0", O_LMAX},
  {"nwidth", 'a', 1, "20", O_NWIDTH},
  {"ethreshold", 0, 1, "1e-8", O_ETHRESH}, {"cloud", 0, 1, NULL, O_CLOUD}, {"cloudtop", 0, 1, NULL, O_CLOUDTOP},
  {"scattering", 0, 1, NULL, O_SCATTERING}, {"detailext", 0, 1, NULL, O_IGN}, {"detailcia", 0, 1, NULL, O_IGN},
  {"csfile", 0, 1, NULL, O_CSFILE}, {"saveext", 0, 1, NULL, O_SAVEEXT},
  {"opacityfile", 0, 1, NULL, O_OPACITYFILE}, {"tlow", 0, 1, "500", O_TLOW}, {"thigh", 0, 1, "3000", O_THIGH},
  {"tempdelt", 0, 1, "100.0", O_TEMPDELT}, {"justOpacity", 0, 0, NULL, O_OPABREAK},
  {"shareOpacity", 0, 0, NULL, O_IGN},
  {"solution", 's', 1, "eclipse", O_SOLUTION}, {"toomuch", 0, 1, "20", O_TOOMUCH}, {"taulevel", 0, 1, "1", O_IGN},
  {"modlevel", 0, 1, "1", O_MODLEVEL}, {"detailtau", 0, 1, NULL, O_IGN},
  {"starrad", 0, 1, "1.125", O_STARRAD}, {"gorbpar", 0, 1, NULL, O_IGN}, {"gorbparfct", 0, 1, NULL, O_IGN},
  {"transparent", 0, 0, NULL, O_TRANSPARENT}, {"raygrid", 0, 1, "0 20 40 60 80", O_RAYGRID},
  {NULL, 0, 0, NULL, 0}};

static char *dupstr(const char *v) {
  size_t n = strlen(v);
  while (n && isspace((unsigned char)v[n - 1])) n--;
  char *r = (char *)malloc(n + 1);
  memcpy(r, v, n);
  r[n] = 0;
  return r;
}
static void setstr(char **dst, const char *v) {
  free(*dst);
  *dst = dupstr(v);
}

static int process_file(th_hints *h, const char *path);

static void apply(th_hints *h, const th_opt *o, const char *v) {
  switch (o->id) {
    case O_IGN: break;
    case O_OUTTOOMUCH: setstr(&h->f_toomuch, v); break;
    case O_OUTSAMPLE: setstr(&h->f_outsample, v); break;
    case O_SAVEEXT: setstr(&h->f_saveext, v); break;
    case O_SAVEFILES:         /* argum.c:457-470: 'yes' or 'no' */
      if (strcmp(v, "yes") == 0) h->savefiles = 1; else if (strcmp(v, "no") == 0) h->savefiles = 0;
      else { printf("\nAllowed arguments for savefiles are: 'yes' or 'no'\n\n"); exit(EXIT_FAILURE); }
      break;
    case O_MODLEVEL: h->modlevel = atoi(v); break;
    case O_VERSION: printf("transit_b200 host (B200 line-by-line path), interface of transit v4.0\n"); exit(0);
    case O_HELP:
      printf("usage: transit_b200 [-c cfgfile] [--key value ...]   keys: ");
      for (const th_opt *p = table; p->name; p++) printf("%s ", p->name);
      printf("\n");
      exit(0);
    case O_QUIET: h->verb = 1; break;
    case O_VERB: h->verb = (int)strtol(v, NULL, 10); break;
    case O_CFG: { char *f = dupstr(v); process_file(h, f); free(f); break; }
    case O_ATM: setstr(&h->f_atm, v); break;
    case O_LINEDB: setstr(&h->f_line, v); break;
    case O_OUTSPEC: setstr(&h->f_outspec, v); break;
    case O_OUTINTENS: setstr(&h->f_outintens, v); break;
    case O_MOLFILE: setstr(&h->f_molfile, v); break;
    case O_RADDELT: h->rad_d = atof(v); break;
    case O_RADLOW: h->rad_i = atof(v); break;          /* argum.c:499-509 */
    case O_RADHIGH: h->rad_f = atof(v); break;
    case O_RADFCT: h->rad_fct = atof(v); break;
    case O_ALLOWQ: h->allowq = atof(v); break;
    case O_RPRESS: h->p0 = atof(v); break;
    case O_RRADIUS: h->r0 = atof(v); break;
    case O_GSURF: h->gsurf = atof(v); break;
    case O_QMOL: setstr(&h->qmol, v); break;
    case O_QSCALE: setstr(&h->qscale, v); break;
    case O_WLLOW: h->wl_i = atof(v); break;
    case O_WLHIGH: h->wl_f = atof(v); break;
    case O_WLFCT: h->wl_fct = atof(v); break;
    case O_WNLOW: h->wn_i = atof(v); break;
    case O_WNHIGH: h->wn_f = atof(v); break;
    case O_WNDELT: h->wn_d = atof(v); break;
    case O_WNOSAMP: h->wn_o = (int)atof(v); break;
    case O_WNFCT: h->wn_fct = atof(v); break;
    case O_NDOP: h->nDop = atoi(v); break;
    case O_NLOR: h->nLor = atoi(v); break;
    case O_DMIN: h->dmin = (float)atof(v); break;
    case O_DMAX: h->dmax = (float)atof(v); break;
    case O_LMIN: h->lmin = (float)atof(v); break;
    case O_LMAX: h->lmax = (float)atof(v); break;
    case O_NWIDTH: h->nwidth = (float)atof(v); break;
    case O_ETHRESH: h->ethresh = atof(v); break;
    case O_CLOUDTOP:          /* argum.c: constant-extinction deck below cloudtop (log bar) */
      h->cl_flag = 1; h->cl_top = atof(v); h->cl_bot = h->cl_top + 10; h->cl_ext = 100; break;
    case O_CLOUD: {           /* cloudtype,cloudext,cloudtop,cloudbot[,gamma[,Q,r | sig,refwn]] */
      double a[8] = {0, 0, 0, 1, 1, 1, 1, 1};
      if (strstr(v, "ext")) h->cl_flag = 1; else if (strstr(v, "opa")) h->cl_flag = 2;
      else if (strstr(v, "B17")) h->cl_flag = 3; else if (strstr(v, "F18")) h->cl_flag = 4;
      else if (strstr(v, "P19")) h->cl_flag = 5;
      else th_die("Syntax error in option '--cloud': unknown cloud type in '%s'.\n", v);
      const char *p = v + 3;
      int n = 0;
      while (*p == ',' && n < 8) { char *e; a[n++] = strtod(p + 1, &e); p = e; }
      if (n < 3) th_die("Syntax error in option '--cloud', parameters need to be given as "
                        "cloudtype,cloudext,cloudtop,cloudbot.\n");
      h->cl_ext = a[0]; h->cl_top = a[1]; h->cl_bot = a[2];
      if (h->cl_top > h->cl_bot) th_die("Syntax error in '--cloud', the cloud top (%g) needs to be less than "
                                        "the cloud bottom (%g) .\n", h->cl_top, h->cl_bot);
      h->cl_gamma = h->cl_flag >= 3 ? a[3] : 1; h->cl_Q = 1; h->cl_r = 1; h->cl_sig = 1; h->cl_refwn = 1;
      if (h->cl_flag == 4) { h->cl_Q = a[4]; h->cl_r = a[5]; }
      if (h->cl_flag == 5) { h->cl_sig = a[4]; h->cl_refwn = a[5]; }
      break;
    }
    case O_SCATTERING:
      if (strstr(v, "polar")) h->sc_flag = 2; else { h->sc_flag = 1; h->sc_logext = atof(v); }
      break;
    case O_CSFILE: setstr(&h->csfiles, v); break;
    case O_OPACITYFILE: setstr(&h->f_opa, v); break;
    case O_TLOW: h->tlow = atof(v); break;
    case O_THIGH: h->thigh = atof(v); break;
    case O_TEMPDELT: h->tempdelt = atof(v); break;
    case O_OPABREAK: h->opabreak = 1; break;
    case O_SOLUTION: setstr(&h->solname, v); break;
    case O_TOOMUCH: h->toomuch = atof(v); break;
    case O_STARRAD: h->starrad = atof(v); break;
    case O_TRANSPARENT: h->transparent = 1; break;
    case O_RAYGRID: setstr(&h->raygrid, v); break;
  }
}

/* getoptfrom (procopt.c:640-700): prefix match in table order, first wins */
static int process_file(th_hints *h, const char *path) {
  FILE *fp = fopen(path, "r");
  if (!fp) th_die("Unable to succesfully open parameter file '%s'\n", path);
  char *line = NULL;
  size_t cap = 0;
  while (getline(&line, &cap, fp) > 0) {
    size_t n = strlen(line);
    while (n && (line[n - 1] == '\n' || line[n - 1] == '\r')) line[--n] = 0;
    if (!*line || *line == '#') continue;
    char *v = line;
    while (*v && *v != ' ' && *v != '\t') v++;
    const size_t flen = (size_t)(v - line);
    while (*v == ' ' || *v == '\t') v++;
    const th_opt *o = table;
    for (; o->name; o++)
      if (strncmp(o->name, line, flen) == 0) break;
    if (!o->name) th_die("Unknown, unsupported, or missing parameter to option '%.*s' in '%s'\n", (int)flen, line, path);
    if (o->has_arg && !*v) th_die("Option '%s' in '%s' needs a value\n", o->name, path);
    apply(h, o, v);
  }
  free(line);
  fclose(fp);
  return 0;
}

void th_defaults(th_hints *h) {
  memset(h, 0, sizeof(*h));
  for (const th_opt *o = table; o->name; o++)
    if (o->def) apply(h, o, o->def);
}

int th_parse_args(th_hints *h, int argc, char **argv) {
  for (int i = 1; i < argc; i++) {
    const char *a = argv[i];
    const th_opt *hit = NULL;
    if (a[0] == '-' && a[1] == '-') {
      const char *name = a + 2;
      const char *eq = strchr(name, '=');
      const size_t len = eq ? (size_t)(eq - name) : strlen(name);
      int nmatch = 0;
      for (const th_opt *o = table; o->name; o++) {
        if (strlen(o->name) == len && strncmp(o->name, name, len) == 0) { hit = o; nmatch = 1; break; }
        if (strncmp(o->name, name, len) == 0) { if (!nmatch) hit = o; nmatch++; }
      }
      if (!hit || nmatch > 1) th_die("Unknown or ambiguous option '%s', use '-h' to see the available options.\n", a);
      const char *v = "";
      if (hit->has_arg) {
        if (eq) v = eq + 1;
        else if (i + 1 < argc) v = argv[++i];
        else th_die("Option '%s' needs a value\n", a);
      }
      apply(h, hit, v);
    } else if (a[0] == '-' && a[1] && !a[2]) {
      for (const th_opt *o = table; o->name; o++)
        if (o->shortc == a[1]) { hit = o; break; }
      if (!hit) th_die("Unknown option '%s'\n", a);
      const char *v = "";
      if (hit->has_arg) {
        if (i + 1 < argc) v = argv[++i]; else th_die("Option '%s' needs a value\n", a);
      }
      apply(h, hit, v);
    } else
      th_die("Unexpected argument '%s'\n", a);
  }
  return 0;
}
