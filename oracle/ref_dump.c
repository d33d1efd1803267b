/* oracle/ref_dump.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Our own driver, linked against the UNMODIFIED reference objects (oracle/Makefile,
 * target _ref/ref_dump).  It runs the reference stage by stage exactly as
 * transit_init()/do_transit() do (transit/src/transit.c:25-74,125-214) and writes
 * every array of the hot path at full double precision into one container file,
 * *before* do_transit's clean-up would free them.  tests/golden/make_golden.py turns
 * that container into the committed .npz fixtures.
 *
 * usage: ref_dump <out.bin> [--full-ext] [--rerun <in.f64>] [--idx r0,r1,..] -- <transit argv...>
 *   --full-ext : additionally call computemolext() for EVERY layer (the reference
 *                only evaluates the layers its rays reach, tau.c:238-271)
 *   --rerun f  : after the first spectrum, call run_transit() with the doubles in f
 *                (reloadatm path, transit.c:118-122) and dump its output too; a file holding k
 *                inputs back to back gives k calls (outputs rerun_out [k][nwn])
 *   --idx list : (ref_dump_idx only: the build with the patched copy of extinction.c, see
 *                oracle/Makefile) for every listed layer call computemolext() once more with the
 *                hooks armed and dump per line: status (0 not visited / out of range, 1 evaluated
 *                leader, 2 co-added member, 3 leader below threshold), iown, idwn, idop, ilor, minj, maxj
 *
 * container record: u32 namelen | name | u32 dtype(0=f64,1=i32,2=i64,3=f32,4=i16)
 *                   | u32 ndim | u64 dims[ndim] | raw little-endian data
 */
#include <transit.h>

extern struct transit transit;
void transit_init(int argc, char **argv);
void run_transit(double *re_input, int transint, double *transit_out, int n);
int  reloadatm(struct transit *tr, double *input);
int  computemolext(struct transit *tr, PREC_RES **kiso, PREC_ATM temp,
                   PREC_ATM *density, double *Z, int permol);

static FILE *out;

#ifdef REF_IDX
/* hooks called from the patched copy of extinction.c (oracle/ref_idx_hooks.h) */
long ref_idx_lead = 0;
static int idx_armed = 0;
static signed char *ix_status;
static int *ix_iown, *ix_idwn, *ix_idop, *ix_ilor, *ix_minj, *ix_maxj;
void ref_idx_hook(long lead, long last, long iown, long idwn, int idop, int ilor, long minj, long maxj){
  long l;
  if(!idx_armed) return;
  ix_status[lead] = 1;
  for(l=lead+1; l<=last; l++) ix_status[l] = 2;
  ix_iown[lead] = (int)iown; ix_idwn[lead] = (int)idwn; ix_idop[lead] = idop; ix_ilor[lead] = ilor;
  ix_minj[lead] = (int)minj; ix_maxj[lead] = (int)maxj;
}
void ref_idx_skip(long lead, long last, long iown){
  long l;
  if(!idx_armed) return;
  ix_status[lead] = 3;
  for(l=lead+1; l<=last; l++) ix_status[l] = 2;
  ix_iown[lead] = (int)iown;
}
#endif

static void rec(const char *name, unsigned dtype, unsigned ndim,
                const unsigned long long *dims, const void *data){
  static const size_t esz[] = {8, 4, 8, 4, 2};
  unsigned nl = (unsigned)strlen(name);
  size_t n = 1;
  for(unsigned i=0; i<ndim; i++) n *= dims[i];
  fwrite(&nl, 4, 1, out); fwrite(name, 1, nl, out);
  fwrite(&dtype, 4, 1, out); fwrite(&ndim, 4, 1, out);
  fwrite(dims, 8, ndim, out);
  if(n) fwrite(data, esz[dtype], n, out);
}
static void rec1(const char *name, unsigned dtype, unsigned long long n, const void *d){
  rec(name, dtype, 1, &n, d);
}
static void rec2(const char *name, unsigned dtype, unsigned long long a,
                 unsigned long long b, const void *d){
  unsigned long long dims[2] = {a, b};
  rec(name, dtype, 2, dims, d);
}
static void scal_f(const char *name, double v){ rec1(name, 0, 1, &v); }
static void scal_i(const char *name, long long v){ rec1(name, 2, 1, &v); }

static void dump_static(struct transit *tr){
  struct lineinfo *li = tr->ds.li;
  struct isotopes *iso = tr->ds.iso;
  struct molecules *mol = tr->ds.mol;
  struct opacity *op = tr->ds.op;
  struct cross *cr = tr->ds.cross;
  long i, j;
  char nm[64];

  scal_f("wns_i", tr->wns.i); scal_f("wns_d", tr->wns.d); scal_f("wns_f", tr->wns.f);
  scal_i("wns_o", tr->wns.o); scal_i("wns_n", tr->wns.n); scal_f("wns_fct", tr->wns.fct);
  rec1("wns_v", 0, tr->wns.n, tr->wns.v);
  scal_f("owns_i", tr->owns.i); scal_f("owns_d", tr->owns.d);
  scal_i("owns_o", tr->owns.o); scal_i("owns_n", tr->owns.n);
  scal_f("owns_last", tr->owns.v[tr->owns.n-1]);
  scal_f("timesalpha", tr->timesalpha);
  scal_f("ethresh", tr->ds.th->ethresh);
  scal_f("toomuch", tr->ds.th->toomuch);
  scal_i("is_eclipse", strcmp(tr->sol->name, "eclipse") == 0);

  /* line list (readlineinfo.c:416-537) */
  scal_i("nlines", li->n_l);
  if(tr->fp_opa == NULL || li->lt.wl){
    rec1("line_wl",   0, li->n_l, li->lt.wl);
    rec1("line_elow", 0, li->n_l, li->lt.elow);
    rec1("line_gf",   0, li->n_l, li->lt.gf);
    rec1("line_isoid",4, li->n_l, li->lt.isoid);
    scal_f("line_wfct", li->lt.wfct); scal_f("line_efct", li->lt.efct);
  }

  /* isotopes and molecules */
  long niso = iso->n_i, nmol = mol->nmol, nrad = tr->rads.n;
  scal_i("niso", niso); scal_i("nmol", nmol); scal_i("nrad", nrad);
  scal_i("iso_nmol", iso->nmol);
  double *tmp = calloc((niso>nmol?niso:nmol)*nrad + 8, sizeof(double));
  for(i=0; i<niso; i++) tmp[i] = iso->isof[i].m;
  rec1("iso_mass", 0, niso, tmp);
  rec1("iso_ratio", 0, niso, iso->isoratio);
  rec1("iso_imol", 1, niso, iso->imol);
  for(i=0; i<niso; i++) for(j=0; j<nrad; j++) tmp[i*nrad+j] = iso->isov[i].z[j];
  rec2("iso_z", 0, niso, nrad, tmp);
  {
    int dbidx[512]; for(i=0; i<niso; i++) dbidx[i] = iso->isof[i].d;
    rec1("iso_db", 1, niso, dbidx);
    scal_i("ndb", iso->n_db);
    for(i=0; i<iso->n_db; i++){
      sprintf(nm, "db%ld_T", i);  rec1(nm, 0, li->db[i].t, li->db[i].T);
      sprintf(nm, "db%ld_first", i); scal_i(nm, iso->db[i].s);
      sprintf(nm, "db%ld_niso", i);  scal_i(nm, iso->db[i].i);
      for(j=0; j<(long)iso->db[i].i; j++){
        sprintf(nm, "db%ld_iso%ld_Z", i, j);
        rec1(nm, 0, li->db[i].t, li->isov[iso->db[i].s+j].z);
      }
    }
  }
  rec1("mol_mass", 0, nmol, mol->mass);
  rec1("mol_radius", 0, nmol, mol->radius);
  rec1("mol_pol", 0, nmol, mol->pol);
  rec1("mol_ID", 1, nmol, mol->ID);
  for(i=0; i<nmol; i++) for(j=0; j<nrad; j++) tmp[i*nrad+j] = mol->molec[i].d[j];
  rec2("mol_d", 0, nmol, nrad, tmp);
  for(i=0; i<nmol; i++) for(j=0; j<nrad; j++) tmp[i*nrad+j] = mol->molec[i].q[j];
  rec2("mol_q", 0, nmol, nrad, tmp);
  free(tmp);
  {
    char names[4096]; names[0] = 0;
    for(i=0; i<nmol; i++){ strcat(names, mol->name[i]); strcat(names, " "); }
    rec1("mol_names", 4, (strlen(names)+1)/2, names);
  }

  /* sampled atmosphere (makesample.c:408-549) */
  rec1("rad_v", 0, nrad, tr->rads.v); scal_f("rad_fct", tr->rads.fct);
  rec1("atm_p", 0, nrad, tr->atm.p);  scal_f("atm_pfct", tr->atm.pfct);
  rec1("atm_t", 0, nrad, tr->atm.t);  scal_f("atm_tfct", tr->atm.tfct);
  rec1("atm_mm", 0, nrad, tr->atm.mm);
  scal_i("atm_bymass", tr->ds.at->mass);
  scal_f("p0", tr->p0); scal_f("r0", tr->r0); scal_f("gsurf", tr->gsurf);

  /* Voigt table (opacity.c:218-277) */
  if(op->profile){
    long nD = op->nDop, nL = op->nLor;
    scal_i("nDop", nD); scal_i("nLor", nL);
    rec1("aDop", 0, nD, op->aDop); rec1("aLor", 0, nL, op->aLor);
    rec2("profsize", 2, nD, nL, op->profsize[0]);
    /* concatenated profiles, one per (i,j) cell (shared ones repeated by offset) */
    long long *off = calloc(nD*nL, sizeof(long long));
    long long total = 0;
    for(i=0; i<nD; i++) for(j=0; j<nL; j++){
      if(i>0 && op->profile[i][j] == op->profile[i-1][j])
        off[i*nL+j] = off[(i-1)*nL+j];
      else{
        off[i*nL+j] = total;
        total += 2*op->profsize[i][j] + 1;
      }
    }
    float *buf = malloc(total*sizeof(float));
    for(i=0; i<nD; i++) for(j=0; j<nL; j++)
      if(!(i>0 && op->profile[i][j] == op->profile[i-1][j]))
        memcpy(buf+off[i*nL+j], op->profile[i][j],
               (2*op->profsize[i][j]+1)*sizeof(float));
    rec2("prof_off", 2, nD, nL, off);
    rec1("prof_data", 3, total, buf);
    free(off); free(buf);
  }
  /* opacity grid (opacity.c:405-423 / 432-503) */
  if(op->o){
    scal_i("op_Nmol", op->Nmol); scal_i("op_Ntemp", op->Ntemp);
    scal_i("op_Nlayer", op->Nlayer); scal_i("op_Nwave", op->Nwave);
    rec1("op_molID", 1, op->Nmol, op->molID);
    rec1("op_temp", 0, op->Ntemp, op->temp);
    rec1("op_press", 0, op->Nlayer, op->press);
    rec1("op_wns", 0, op->Nwave, op->wns);
    double *g = malloc(sizeof(double)*op->Nlayer*op->Ntemp*op->Nmol*op->Nwave);
    long r, t, m;
    for(r=0; r<op->Nlayer; r++) for(t=0; t<op->Ntemp; t++) for(m=0; m<op->Nmol; m++)
      memcpy(g + ((r*op->Ntemp+t)*op->Nmol+m)*op->Nwave, op->o[r][t][m],
             op->Nwave*sizeof(double));
    unsigned long long dims[4] = {op->Nlayer, op->Ntemp, op->Nmol, op->Nwave};
    rec("op_o", 0, 4, dims, g);
    free(g);
  }

  /* cross-section tables (crosssec.c:9-268) */
  scal_i("cs_nfiles", cr->nfiles);
  for(i=0; i<cr->nfiles; i++){
    sprintf(nm, "cs%ld_wn", i);   rec1(nm, 0, cr->nwave[i], cr->wn[i]);
    sprintf(nm, "cs%ld_temp", i); rec1(nm, 0, cr->ntemp[i], cr->temp[i]);
    double *t2 = malloc(sizeof(double)*cr->nwave[i]*cr->ntemp[i]);
    for(j=0; j<cr->nwave[i]; j++)
      memcpy(t2+j*cr->ntemp[i], cr->cs[i][j], cr->ntemp[i]*sizeof(double));
    sprintf(nm, "cs%ld_tab", i);  rec2(nm, 0, cr->nwave[i], cr->ntemp[i], t2);
    free(t2);
    sprintf(nm, "cs%ld_mol", i);  rec1(nm, 1, cr->nspec[i], cr->mol[i]);
  }

  /* geometry, clouds, scattering */
  scal_f("starrad_cgs", tr->ds.sg->starrad * tr->ds.sg->starradfct);
  scal_i("transpplanet", tr->ds.sg->transpplanet);
  scal_i("sc_flag", tr->ds.sc->flag); scal_f("sc_logext", tr->ds.sc->logext);
  scal_i("cl_flag", tr->ds.cl->flag); scal_f("cl_ext", tr->ds.cl->cloudext);
  scal_f("cl_top", tr->ds.cl->cloudtop); scal_f("cl_bot", tr->ds.cl->cloudbot);
  scal_f("cl_gamma", tr->ds.cl->gamma); scal_f("cl_Q", tr->ds.cl->Q);
  scal_f("cl_r", tr->ds.cl->r); scal_f("cl_sig", tr->ds.cl->sig);
  scal_f("cl_refwn", tr->ds.cl->refwn);
  if(tr->ann) rec1("angles", 0, tr->ann, tr->angles);
}

/* The stage sequence of do_transit (transit.c:138-199) with taps. */
static void spectrum_with_taps(struct transit *tr, const char *pfx, int full_ext){
  char nm[64];
  long nwn = tr->wns.n, nrad = tr->rads.n, i;
  int is_ecl = strcmp(tr->sol->name, "eclipse") == 0;

  if(makeipsample(tr) < 0) exit(3);
  if(interpcs(tr) != 0) exit(3);
  if(idxrefrac(tr) != 0) exit(3);
  if(extwn(tr) != 0) exit(3);
  if(init_optdepth(tr) != 0) exit(3);
  if(tau(tr) != 0) exit(3);

  sprintf(nm, "%sips_v", pfx);  rec1(nm, 0, tr->ips.n, tr->ips.v);
  sprintf(nm, "%sips_fct", pfx); scal_f(nm, tr->ips.fct);
  sprintf(nm, "%se_cs", pfx);   rec2(nm, 0, nwn, nrad, tr->ds.cross->e[0]);
  sprintf(nm, "%sext", pfx);    rec2(nm, 0, nrad, nwn, tr->ds.ex->e[0]);
  sprintf(nm, "%scomputed", pfx); rec1(nm, 4, nrad, tr->ds.ex->computed);
  long nh = is_ecl ? nrad : tr->ips.n;
  sprintf(nm, "%stau", pfx);    rec2(nm, 0, nwn, nh, tr->ds.tau->t[0]);
  sprintf(nm, "%slast", pfx);   rec1(nm, 2, nwn, tr->ds.tau->last);

  if(is_ecl){
    for(i=0; i<tr->ann; i++){
      tr->angleIndex = i;
      if(emergent_intens(tr) != 0) exit(3);
    }
    sprintf(nm, "%sintens", pfx); rec2(nm, 0, tr->ann, nwn, tr->ds.intens->a[0]);
    if(flux(tr) != 0) exit(3);
  }
  else if(modulation(tr) != 0) exit(3);
  sprintf(nm, "%sout", pfx); rec1(nm, 0, nwn, tr->ds.out->o);

  if(full_ext && tr->fp_opa == NULL && tr->f_line != NULL){
    /* every layer through the reference's own line sum (extinction.c:281-529) */
    double *full = calloc(nrad*nwn, sizeof(double));
    double *dens = calloc(tr->ds.mol->nmol, sizeof(double));
    double *Z    = calloc(tr->ds.iso->n_i, sizeof(double));
    long r; int k;
    for(r=0; r<nrad; r++){
      double *row = full + r*nwn;
      for(k=0; k<tr->ds.mol->nmol; k++) dens[k] = tr->ds.mol->molec[k].d[r];
      for(k=0; k<tr->ds.iso->n_i; k++)  Z[k] = tr->ds.iso->isov[k].z[r];
      computemolext(tr, &row, tr->atm.t[r]*tr->atm.tfct, dens, Z, 0);
    }
    sprintf(nm, "%sext_full", pfx); rec2(nm, 0, nrad, nwn, full);
    free(full);
  }
}

int main(int argc, char **argv){
  int full_ext = 0, a = 2, rerun_taps = 0;
  const char *rerun = NULL, *idxlist = NULL;
  if(argc < 4){ fprintf(stderr, "usage: ref_dump out.bin [--full-ext] [--rerun f] -- transit-args\n"); return 2; }
  for(; a < argc && strcmp(argv[a], "--"); a++){
    if(!strcmp(argv[a], "--full-ext")) full_ext = 1;
    else if(!strcmp(argv[a], "--rerun")) rerun = argv[++a];
    else if(!strcmp(argv[a], "--rerun-taps")) rerun_taps = 1;   /* taps of the re-run too (prefix rr_) */
    else if(!strcmp(argv[a], "--idx")) idxlist = argv[++a];
  }
  if(a >= argc) return 2;
  argv[a] = argv[0];                     /* transit's argv[0] */
  out = fopen(argv[1], "wb");
  if(!out){ perror("ref_dump"); return 2; }

  transit_init(argc - a, argv + a);
  dump_static(&transit);
  if(transit.opabreak){ fclose(out); return 0; }

  spectrum_with_taps(&transit, "", full_ext);

#ifdef REF_IDX
  if(idxlist && transit.fp_opa == NULL && transit.f_line != NULL){
    struct transit *tr = &transit;
    long nl = tr->ds.li->n_l, nwn = tr->wns.n;
    double *row = calloc(nwn, sizeof(double));
    double *dens = calloc(tr->ds.mol->nmol, sizeof(double));
    double *Z    = calloc(tr->ds.iso->n_i, sizeof(double));
    ix_status = malloc(nl); ix_iown = malloc(4*nl); ix_idwn = malloc(4*nl); ix_idop = malloc(4*nl);
    ix_ilor = malloc(4*nl); ix_minj = malloc(4*nl); ix_maxj = malloc(4*nl);
    char *lst = strdup(idxlist), *tok;
    for(tok = strtok(lst, ","); tok; tok = strtok(NULL, ",")){
      long r = atol(tok); int k; char nm[64];
      if(r < 0 || r >= tr->rads.n) continue;
      memset(ix_status, 0, nl); memset(ix_iown, 0, 4*nl); memset(ix_idwn, 0, 4*nl); memset(ix_idop, 0xff, 4*nl);
      memset(ix_ilor, 0xff, 4*nl); memset(ix_minj, 0, 4*nl); memset(ix_maxj, 0, 4*nl);
      for(k=0; k<tr->ds.mol->nmol; k++) dens[k] = tr->ds.mol->molec[k].d[r];
      for(k=0; k<tr->ds.iso->n_i; k++)  Z[k] = tr->ds.iso->isov[k].z[r];
      idx_armed = 1;
      computemolext(tr, &row, tr->atm.t[r]*tr->atm.tfct, dens, Z, 0);
      idx_armed = 0;
      /* dtype 4 = i16 is the smallest the container has: status goes out as i32 */
      { int *st32 = malloc(4*nl); long l; for(l=0; l<nl; l++) st32[l] = ix_status[l];
        sprintf(nm, "idx%ld_status", r); rec1(nm, 1, nl, st32); free(st32); }
      sprintf(nm, "idx%ld_iown", r); rec1(nm, 1, nl, ix_iown);
      sprintf(nm, "idx%ld_idwn", r); rec1(nm, 1, nl, ix_idwn);
      sprintf(nm, "idx%ld_idop", r); rec1(nm, 1, nl, ix_idop);
      sprintf(nm, "idx%ld_ilor", r); rec1(nm, 1, nl, ix_ilor);
      sprintf(nm, "idx%ld_minj", r); rec1(nm, 1, nl, ix_minj);
      sprintf(nm, "idx%ld_maxj", r); rec1(nm, 1, nl, ix_maxj);
      sprintf(nm, "idx%ld_ext", r);  rec1(nm, 0, nwn, row);
    }
    free(lst);
  }
#endif

  if(rerun){
    FILE *fi = fopen(rerun, "rb");
    long nin = transit.ds.at->rads.n * (1 + transit.ds.mol->nmol), nwn = transit.wns.n, nrr = 1, k;   /* reloadatm reads the atmosphere FILE's layers */
    if(fi){ fseek(fi, 0, SEEK_END); nrr = ftell(fi) / (8*nin); fseek(fi, 0, SEEK_SET); }
    if(nrr < 1) nrr = 1;
    double *in = malloc(nrr*nin*sizeof(double)), *o = malloc(nrr*nwn*sizeof(double));
    if(!fi || fread(in, 8, nrr*nin, fi) != (size_t)(nrr*nin)){ fprintf(stderr, "bad rerun file\n"); return 2; }
    fclose(fi);
    /* do_transit frees what spectrum_with_taps allocated only inside do_transit itself;
       our taps allocated fresh arrays, so release them the way do_transit does. */
    freemem_samp(&transit.ips);
    freemem_idexrefrac(transit.ds.ir, &transit.pi);
    freemem_extinction(transit.ds.ex, &transit.pi);
    freemem_tau(transit.ds.tau, &transit.pi);
    freemem_outputray(transit.ds.out, &transit.pi);
    if(rerun_taps){            /* run_transit = reloadatm + do_transit (transit.c:118-122) */
      if(reloadatm(&transit, in) < 0) exit(3);
      spectrum_with_taps(&transit, "rr_", full_ext);
      memcpy(o, transit.ds.out->o, nwn*sizeof(double));
    }
    else
      for(k=0; k<nrr; k++)
        run_transit(in + k*nin, (int)nin, o + k*nwn, (int)nwn);
    if(nrr > 1) rec2("rerun_out_all", 0, nrr, nwn, o);
    in += (nrr-1)*nin; o += (nrr-1)*nwn;       /* the single-call records describe the LAST call */
    rec1("rerun_in", 0, nin, in);
    rec1("rerun_out", 0, nwn, o);
    rec1("rerun_atm_t", 0, transit.rads.n, transit.atm.t);
    rec1("rerun_rad_v", 0, transit.rads.n, transit.rads.v);
    {
      long i, j, nmol = transit.ds.mol->nmol, nrad = transit.rads.n, niso = transit.ds.iso->n_i;
      double *tmp = malloc(sizeof(double)*(nmol>niso?nmol:niso)*nrad);
      for(i=0; i<nmol; i++) for(j=0; j<nrad; j++) tmp[i*nrad+j] = transit.ds.mol->molec[i].d[j];
      rec2("rerun_mol_d", 0, nmol, nrad, tmp);
      for(i=0; i<niso; i++) for(j=0; j<nrad; j++) tmp[i*nrad+j] = transit.ds.iso->isov[i].z[j];
      rec2("rerun_iso_z", 0, niso, nrad, tmp);
      free(tmp);
    }
  }
  fclose(out);
  return 0;
}
