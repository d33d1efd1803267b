/* transit_b200.h -- C ABI of the B200 (sm_100a) implementation of Transit's line-by-line
 * spectrum hot path.  This is the only boundary the host program (C, keeping the
 * reference's CLI / cfg keys / file formats) crosses: plain pointers and sizes, int
 * return codes (0 ok, <0 error, text from tb_last_error()), host owns every host
 * pointer, all device buffers live in the context, no global state, NO CPU FALLBACK:
 * every entry point fails with TB_ERR_CUDA when no sm_100 device is usable.
 *
 * Each entry point names the reference function(s) it replaces (paths into the
 * exosports/transit tree).  Array layouts at the boundary are the reference's.
 */
#ifndef TRANSIT_B200_H
#define TRANSIT_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct tb_ctx tb_ctx;

enum { TB_OK = 0, TB_ERR_ARG = -1, TB_ERR_CUDA = -2, TB_ERR_STATE = -3, TB_ERR_UNSORTED = -4,
       TB_ERR_NOMEM = -5, TB_ERR_DOMAIN = -6 };
enum { TB_GEOM_TRANSIT = 0, TB_GEOM_ECLIPSE = 1 };   /* ray_solution slantpath / eclipsepath */

const char *tb_last_error(void);
int  tb_version(void);

/* One context = one device (struct transit, transit/src/transit.c:7, made re-entrant). */
int  tb_create(int device, tb_ctx **out);
void tb_destroy(tb_ctx *c);
int  tb_sync(tb_ctx *c);

/* makewnsample / makesample1 (transit/src/makesample.c:27-120,308-400): coarse grid
 * wns.v[k] = wn_i + k*wn_d, fine grid owns.v[k] = wn_i + k*(wn_d/wnosamp).
 * nwn/onwn are the counts makesample1 computes (tb_grid_count).                       */
int64_t tb_grid_count(double wn_i, double wn_f, double wn_d, int osamp);
int  tb_set_wn_grid(tb_ctx *c, double wn_i, double wn_d, int wnosamp, int64_t nwn, int64_t onwn,
                    double wn_fct);

/* struct molecules / struct isotopes (transit/include/structures_tr.h:232-251),
 * filled by getmoldata (readatm.c:625-717), readtli_bin + setimol (readlineinfo.c:87-278). */
int  tb_set_species(tb_ctx *c, int nmol, const double *mol_mass, const double *mol_radius,
                    const double *mol_pol, const int *mol_ID, int niso, const double *iso_mass,
                    const double *iso_ratio, const int *iso_imol);

/* struct line_transition as readdatarng leaves it (readlineinfo.c:416-537): SoA, per
 * isotope ascending wavelength.  Builds the layer-independent line index on the device
 * (fine/coarse bins, co-add groups of extinction.c:449-462, bin-sorted order).
 * TB_ERR_UNSORTED if wavelengths are not non-decreasing within an isotope.            */
int  tb_upload_lines(tb_ctx *c, const double *wl, const double *elow, const double *gf,
                     const int16_t *isoid, int64_t n, double wfct, double efct);

/* calcprofiles (opacity.c:218-277) -> getprofile (extinction.c:8-57) -> voigtn
 * (pu/src/voigt.c:369-483).  dmin..lmax are float as in struct transithint
 * (structures_tr.h:333); nwidth = timesalpha.                                         */
int  tb_build_profiles(tb_ctx *c, int nDop, int nLor, float dmin, float dmax, float lmin,
                       float lmax, float nwidth);
/* test taps: sizes then data of the table (concatenated float profiles)              */
int  tb_get_profile_meta(tb_ctx *c, double *aDop, double *aLor, int64_t *psize, int64_t *off,
                         int64_t *ndata);
int  tb_get_profile_data(tb_ctx *c, float *data, int64_t ndata);

/* Sampled atmosphere as makeradsample leaves it (makesample.c:408-549):
 * rad[nrad] (bottom first) * rad_fct = cm; p[nrad]*p_fct = barye; t[nrad]*t_fct = K;
 * density[nmol][nrad] g/cm3; Z[niso][nrad] partition function at the layer T.         */
int  tb_set_atmosphere(tb_ctx *c, int nrad, const double *rad, double rad_fct, const double *p,
                       double p_fct, const double *t, double t_fct, const double *density,
                       const double *Z);

/* computemolext with permol=0 for every layer (extinction.c:281-529, driven lazily by
 * tau.c:159-177,238-271 in the reference; here all layers in one batch).             */
int  tb_extinction_lbl(tb_ctx *c, double ethresh);
/* generic batch: `ns` states (T, density[nmol], Z[niso]); out[ns][nout][nwn] on the host;
 * nout = 1 (permol=0) or the number of line-list molecules in first-appearance order. */
int  tb_molext_states(tb_ctx *c, int ns, const double *T, const double *density /*[ns][nmol]*/,
                      const double *Z /*[ns][niso]*/, double ethresh, int permol, double *out);
/* per-line trace of the last tb_molext_states/tb_extinction_lbl state `s`, in the ORIGINAL
 * line order: status 0 out of range, 1 evaluated, 2 co-added, 3 below threshold; iown/idwn
 * for 1 and 3, idop for 1.  nsamples[ns] = executed profile updates per state.        */
int  tb_get_line_trace(tb_ctx *c, int s, int32_t *iown, int32_t *idwn, int32_t *idop,
                       int8_t *status);
int  tb_get_sample_counts(tb_ctx *c, int64_t *nsamples, int ns);
/* The line sum of a state whose profiles are narrow runs as ONE kernel (strength + gather, records kept in
 * shared memory) on grids of at least 96 tiles of 128 bins; tb_set_fused(c, 0) sends every state through the
 * two-kernel path instead, tb_set_fused(c, 2) fuses on small grids too (same bits for the
 * narrow-reach states: same summation order).  tb_set_trace(c, 1) makes the fused kernel also leave the
 * per-line records tb_get_line_trace reads (off by default: 16 B of HBM per (group, state)).
 * tb_get_fused_stats: states of the last batch taken by the fused kernel and its executed profile updates. */
int  tb_set_fused(tb_ctx *c, int on);
/* which states are fused: widest reach in coarse bins (default 8) and whether states whose lines partly keep a
 * carried Doppler index (extinction.c:480-483; found by a bisection + backward scan instead of the list walk) are
 * fused as well (default 0: only states where every line refreshes its index) */
int  tb_set_fused_reach(tb_ctx *c, int rmax, int carried);
/* which fused kernel: 1 (default) = a warp per 32 bins x 4 states, no block barriers; 0 = a block per 128 bins x 4
 * states with block-wide phases.  Same sums, bit for bit (tested).                                          */
int  tb_set_fused_variant(tb_ctx *c, int v);
/* two-kernel path: tiles whose reach is at most r coarse bins (default 2) are summed one thread per bin, the
 * others by the 16-warp window walk of k_accumulate (tuning knob; any value gives the same sums to rounding) */
int  tb_set_narrow_reach(tb_ctx *c, int r);
int  tb_set_trace(tb_ctx *c, int on);
int  tb_get_fused_stats(tb_ctx *c, int64_t *fused_states, int64_t *fused_samples);
/* ilor[niso] of state `s` of the last batch (extinction.c:394); with the trace's idop it names the Voigt cell */
int  tb_get_state_ilor(tb_ctx *c, int s, int32_t *ilor);
int  tb_get_line_molIDs(tb_ctx *c, int *molID, int *nmolrows);   /* opacity.c:349-361 */
/* sizes of the device line index: lines uploaded, co-add groups (evaluated leaders), member lines */
int  tb_get_line_counts(tb_ctx *c, int64_t *nlines, int64_t *ngroups, int64_t *nmembers);

/* calcopacity (opacity.c:281-427): grid[Nlayer][Ntemp][Nmol][Nwave] written to `grid`
 * (host, file order).  ziso[niso][Ntemp]; q[nmol][nrad] number abundances; press cgs.  */
int  tb_opacity_build(tb_ctx *c, int Ntemp, const double *temp, const double *ziso,
                      const double *q, const double *mm, int bymass, double ethresh, double *grid);
/* readopacity (opacity.c:432-503): make a grid resident in HBM                        */
int  tb_opacity_load(tb_ctx *c, int64_t Nmol, int64_t Ntemp, int64_t Nlayer, int64_t Nwave,
                     const int *molID, const double *temp, const double *press,
                     const double *grid);
/* interpolmolext for every layer (extinction.c:534-581)                               */
/* one grid in HBM for several contexts of ONE GPU (replaces shareopacity / the SysV segment, opacity.c:89-201,507-630):
 * dst reads src's grid in place; src must keep its grid for as long as dst uses it                          */
int  tb_opacity_share(tb_ctx *dst, tb_ctx *src);
int  tb_extinction_interp(tb_ctx *c);

/* readcs tables (crosssec.c:9-268) and interpcs/bicubicinterpolate (crosssec.c:271-428) */
int  tb_cia_clear(tb_ctx *c);
int  tb_cia_add_table(tb_ctx *c, const double *wn, int nwn, const double *temp, int ntemp,
                      const double *tab /*[nwn][ntemp]*/, const int *mol, int nspec);
int  tb_cia_interp(tb_ctx *c);

/* computeextscat / computeextcloud parameters (extinction.c:586-693; struct extscat,
 * extcloud).  nH/mean_dens [nrad] as tau.c:194-214 builds them (may be NULL -> zeros). */
int  tb_set_scattering(tb_ctx *c, int flag, double logext);
int  tb_set_cloud(tb_ctx *c, int flag, double cloudext, double cloudtop, double cloudbot,
                  double gamma, double Q, double r, double sig, double refwn,
                  const double *nH, const double *mean_dens);

/* tau (tau.c:59-356) with totaltau1 (slantpath.c:18-108) or eclipsetau (eclipse.c:28-105);
 * rays h[nh]*hfct (impact parameters, top first; NULL = reversed radii as makeipsample
 * with raddelt=-1, makesample.c:564-573).                                             */
int  tb_tau(tb_ctx *c, int geometry, double toomuch, const double *h, int nh, double hfct);
/* the rays tb_spectrum uses (makeipsample, makesample.c:558-604): h[nh]*hfct, top first; NULL or nh = 0 goes back
 * to the reversed radii.  A resampled radius grid (raddelt != -1) brings its own impact-parameter sample.   */
int  tb_set_rays(tb_ctx *c, const double *h, int nh, double hfct);
/* eclipse geometry: 1 (default) = the O(layers)-per-wavenumber streaming kernel (every ray's Simpson pairs end
 * at the top layer, so ray ri adds one pair to the sum of the rays above it); 0 = the general kernel that
 * integrates every ray on its own.  Same terms, pairs associated top down instead of bottom up.            */
int  tb_set_tau_streaming(tb_ctx *c, int on);
/* modulationperwn's dispatch (slantpath.c:477-503, cfg key `modlevel`): 1 = modulation1 (default),
 * -1 = modulationm1 (slantpath.c:446-473, opaque disc at tau = toomuch); anything else is refused
 * with the reference's message.                                                          */
int  tb_set_modlevel(tb_ctx *c, int level);
/* tb_tau for transit geometry WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same kernels and results as tb_tau(TB_GEOM_TRANSIT) with scattering and clouds off; the scratch of the
 * ray geometry (path coordinates, Simpson coefficients, the weight matrix of t = W er, error flags) lives in a
 * file-scope `static __device__` array of TB_TAU_STATIC_SCRATCH_BYTES that every GPU loading the library holds until
 * exit.  All pointers are DEVICE buffers of the caller: rad[nrad], h[nh] (impact parameters, top first), e[nrad][nwn],
 * e_cs[nrad][nwn] (zeros if there is none), out tau[nh][nwn] (zero beyond last[], as tau.c:34 leaves it) and
 * last[nwn] (int32).  nrad and nh beyond TB_TAU_STATIC_MAXLAYERS are refused with TB_ERR_ARG.  Work is enqueued on
 * the context's stream; check != 0 reads the geometry's error flags back (one 16-byte copy + stream sync).  The
 * array is shared by all calls on a device and keeps the last call's contents (the part a call uses is cleared
 * first): calls that overlap on the GPU race on it -- serialise them.  Calls no cudaMalloc / cudaFree.        */
#define TB_TAU_STATIC_MAXLAYERS 128
#define TB_TAU_STATIC_SCRATCH_BYTES 544800
int  tb_tau_transit_static(tb_ctx *c, int nrad, const double *d_rad, double rad_fct, int nh, const double *d_h,
                           double hfct, int64_t nwn, const double *d_e, const double *d_ecs, double toomuch,
                           double *d_tau, int32_t *d_last, int check);
/* modulation/modulation1 (slantpath.c:274-436): out[nwn]                              */
int  tb_modulation(tb_ctx *c, double starrad_cm, int transparent, double *out);
/* eclipse_intens/emergent_intens/flux (eclipse.c:117-287): out[nwn] flux             */
int  tb_intensity_flux(tb_ctx *c, const double *angles_deg, int nang, double *out);

/* do_transit (transit.c:125-214) in one call: CIA + extinction (LBL if no grid is loaded,
 * else grid interpolation) + tau + modulation | intensity+flux.  out[nwn].            */
int  tb_spectrum(tb_ctx *c, int geometry, double toomuch, double ethresh, double starrad_cm,
                 int transparent, const double *angles_deg, int nang, double *out);

/* run_transit (transit.c:118-122) for a batch of atmospheres, as a retrieval hands them over one by
 * one: atmosphere a is the a-th slice of rad[natm][nrad], p, t, density[natm][nmol][nrad],
 * Z[natm][niso][nrad] (each slice as tb_set_atmosphere takes it).  The spectra are computed back to
 * back with no host synchronisation in between; out[natm][nwn].  Results are those of natm calls of
 * tb_set_atmosphere + tb_spectrum.                                                       */
int  tb_spectrum_batch(tb_ctx *c, int natm, int nrad, const double *rad, double rad_fct,
                       const double *p, double p_fct, const double *t, double t_fct,
                       const double *density, const double *Z, int geometry, double toomuch,
                       double ethresh, double starrad_cm, int transparent,
                       const double *angles_deg, int nang, double *out);

/* run_transit for a BATCH of atmospheres exactly as BART hands them over (readatm.c:735-744):
 * in[natm][nlayers*(1+nmol)] = [T(nlayers) | q_mol0(nlayers) | q_mol1(nlayers) ...] per atmosphere, spectra
 * out[natm][nwn].  The per-atmosphere set-up of reloadatm (readatm.c:722-784: checkaddmm, stateeqnford, radpress)
 * runs ON THE DEVICE, and every stage -- grid interpolation, CIA, tau, modulation -- is launched once per batch of
 * up to 32 atmospheres.  Needs a resident opacity grid (tb_opacity_load), transit geometry, modlevel 1, and the
 * static part of the atmosphere model: the file's pressures (file units), the unit factors, `q number|mass`, and the
 * hydrostatic reference level refpress / refradius / gsurf (argum.c).  H2 (for the cloud model P19) is the
 * molecule with ID 105, as in the shipped molecules.dat.                                                       */
int  tb_set_atmosphere_model(tb_ctx *c, int nrad, const double *p, double p_fct, double t_fct, double rad_fct,
                             int bymass, double p0, double r0, double gsurf);
int  tb_run_transit_batch(tb_ctx *c, int natm, const double *in, double toomuch, double starrad_cm,
                          int transparent, double *out);

/* test taps in the reference's layouts                                                */
int  tb_get_extinction(tb_ctx *c, double *e /*[nrad][nwn]*/);
int  tb_get_cia(tb_ctx *c, double *e_cs /*[nwn][nrad]*/);
int  tb_get_tau(tb_ctx *c, double *t /*[nwn][nh]*/, int64_t *last /*[nwn]*/);
int  tb_get_intensity(tb_ctx *c, double *intens /*[nang][nwn]*/);
int  tb_get_computed(tb_ctx *c, int16_t *computed /*[nrad]*/);   /* tau.c lazy-evaluation set */

/* timing taps: device milliseconds of the last call's dominant kernel (CUDA events on the
 * context stream) and launch counter                                                 */
int  tb_get_kernel_ms(tb_ctx *c, const char *name, float *ms);
/* user-named CUDA-event pair on the context stream (bench.py brackets whole steps with it) */
int  tb_timer_start(tb_ctx *c, const char *name);
int  tb_timer_stop(tb_ctx *c, const char *name);
int64_t tb_launch_count(tb_ctx *c);
void *tb_stream(tb_ctx *c);          /* cudaStream_t of the context */
/* run on a caller-owned stream (e.g. the host framework's current stream) from now on;
 * NULL returns to the context's own stream.  The caller keeps ownership of `stream`.   */
int  tb_set_stream(tb_ctx *c, void *stream);
/* device address of the last spectrum (out[nwn], FP64) for device-side consumers such as an
 * NCCL gather of per-rank spectra; valid until the next call that produces a spectrum.  */
int  tb_device_spectrum(tb_ctx *c, void **ptr, int64_t *n);

/* Wavenumber-band shard (the reference's only way to use more cores is N processes
 * over disjoint wavenumber bands, each re-reading the whole line list).  The context keeps the
 * whole line index but computes the line sum only for coarse bins [j_lo, j_hi) -- both multiples
 * of 128 or the grid ends; (0, 0) = the whole grid.  Two quantities of computemolext are global
 * over the line list: the strength maximum behind the threshold (extinction.c:399-427, 467) and
 * the Doppler index carried from the last evaluated line (extinction.c:480-483).  With world > 1
 * the library calls `fn` (or, when fn is NULL and tb_comm_init has run, ncclAllReduce itself) twice per batch
 * of states for an in-place all-reduce MAX over `count`
 * elements at `device_ptr` (doubles for TB_EXCH_KMAX, int32 for TB_EXCH_LASTQ), to be enqueued
 * on the context's stream (e.g. ncclAllReduce, or torch.distributed on the stream given to
 * tb_set_stream); both maxima are exact, so the banded results are bit-identical to one GPU.
 * Bins outside the band come back as zero.                                              */
enum { TB_EXCH_KMAX = 0, TB_EXCH_LASTQ = 1 };
typedef int (*tb_exchange_fn)(void *user, int what, void *device_ptr, int64_t count);
int  tb_set_shard(tb_ctx *c, int64_t j_lo, int64_t j_hi, int rank, int world, tb_exchange_fn fn,
                  void *user);

/* In-library exchange of a band-sharded run (one process per GPU; NCCL bound at run time with dlopen, so a
 * single-GPU user needs none).  Rank 0 makes the 128-byte id (tb_comm_unique_id = ncclGetUniqueId) and the host
 * hands it to the other ranks by whatever it has (a pipe between forked ranks, MPI, torch.distributed, a file);
 * every rank then calls tb_comm_init (ncclCommInitRank on the context's device).  tb_balanced_bands cuts the
 * coarse grid into `world` contiguous bands balanced by line groups per bin (ends on multiples of 128; the same
 * table on every rank); tb_set_bands makes this context rank `rank` of that table.  From then on tb_spectrum
 * evaluates only its band -- line sum, CIA evaluation, tau, modulation / intensities -- completes the two global
 * maxima of computemolext with ncclAllReduce(MAX) on the context's stream, and ends with an ncclAllGather of the
 * band slices, so that EVERY rank returns the whole spectrum, bit-identical to a single-GPU run.            */
int  tb_comm_unique_id(char *id128);
int  tb_comm_init(tb_ctx *c, int rank, int world, const char *id128);
int  tb_comm_destroy(tb_ctx *c);
int  tb_comm_barrier(tb_ctx *c);     /* all ranks have finished what they enqueued (tiny all-reduce + stream sync) */
int  tb_balanced_bands(tb_ctx *c, int world, int64_t *edges /*[world+1]*/);
int  tb_set_bands(tb_ctx *c, int world, const int64_t *edges /*[world+1]*/, int rank);
/* after a first spectrum: all ranks exchange the device time of their line sum, re-weight each band's line groups
 * by the cost per group measured there and cut the grid again (a collective call; same result on every rank) */
int  tb_rebalance_bands(tb_ctx *c);
int  tb_get_bands(tb_ctx *c, int64_t *edges /*[world+1]*/);

/* device micro-benchmarks used for the roofline denominators (FP64 FMA / exp / copy)   */
int  tb_microbench(tb_ctx *c, double *dfma_tflops, double *dexp_gops, double *copy_gbs);

#ifdef __cplusplus
}
#endif
#endif
