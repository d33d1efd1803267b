// tb_comm.cu -- the exchange of a wavenumber-band sharded run, inside the library.
//
// One process per GPU.  A band shard needs three small collectives per spectrum, all enqueued on the
// context's stream between the kernels they separate (no host synchronisation, no host callback):
//   * ncclAllReduce(MAX, double) of the strength maxima behind the threshold (extinction.c:399-427,467),
//   * ncclAllReduce(MAX, int32)  of the carried Doppler index (extinction.c:480-483)  -- both exact,
//   * ncclAllGather of the band slices of the spectrum, padded to the widest band and put back at the
//     band offsets by one small kernel.
// NCCL is bound at RUN time (dlopen of libnccl.so.2): a single-GPU user needs no NCCL at all, and inside a
// process that already carries NCCL (torch) the same copy is used.  The 128-byte unique id is created by
// rank 0 (tb_comm_unique_id) and handed to the other ranks by whatever the host has -- a pipe between the
// forked ranks of host/transit_b200, a torch.distributed broadcast in bench.py, a file.
#include <dlfcn.h>
#include <nccl.h>
#include "tb_internal.cuh"

struct TbNccl {
  void *h = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
static TbNccl g_nccl;

static int tb_nccl_load() {
  if (g_nccl.h) return TB_OK;
  const char *names[] = {"libnccl.so.2", "libnccl.so", nullptr};
  void *h = nullptr;
  for (int i = 0; names[i] && !h; i++) h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
  TB_REQUIRE(h, TB_ERR_STATE, "a band-sharded run needs NCCL, and libnccl.so.2 cannot be loaded: %s", dlerror());
#define SYM(field, name)                                                                          \
  g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, name));                        \
  TB_REQUIRE(g_nccl.field, TB_ERR_STATE, "libnccl.so.2 has no symbol %s", name)
  SYM(GetUniqueId, "ncclGetUniqueId");
  SYM(CommInitRank, "ncclCommInitRank");
  SYM(CommDestroy, "ncclCommDestroy");
  SYM(AllReduce, "ncclAllReduce");
  SYM(AllGather, "ncclAllGather");
  SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
  g_nccl.h = h;
  return TB_OK;
}

#define TB_NCCL(call)                                                                       \
  do {                                                                                      \
    ncclResult_t r_ = (call);                                                               \
    if (r_ != ncclSuccess) {                                                                \
      tb_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r_)); \
      return TB_ERR_CUDA;                                                                   \
    }                                                                                       \
  } while (0)

extern "C" int tb_comm_unique_id(char *id128) {
  TB_REQUIRE(id128, TB_ERR_ARG, "tb_comm_unique_id: NULL buffer");
  TB_TRY(tb_nccl_load());
  ncclUniqueId id;
  TB_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(id128, id.internal, NCCL_UNIQUE_ID_BYTES);
  return TB_OK;
}

extern "C" int tb_comm_init(tb_ctx *c, int rank, int world, const char *id128) {
  TB_REQUIRE(world >= 1 && rank >= 0 && rank < world && id128, TB_ERR_ARG, "tb_comm_init: rank %d of %d", rank, world);
  TB_REQUIRE(!c->comm, TB_ERR_STATE, "tb_comm_init: the context already has a communicator");
  TB_TRY(tb_nccl_load());
  TB_CUDA(cudaSetDevice(c->device));
  ncclUniqueId id;
  memcpy(id.internal, id128, NCCL_UNIQUE_ID_BYTES);
  ncclComm_t comm;
  TB_NCCL(g_nccl.CommInitRank(&comm, world, id, rank));
  c->comm = (void *)comm;
  c->comm_rank = rank;
  c->comm_world = world;
  return TB_OK;
}

extern "C" int tb_comm_destroy(tb_ctx *c) {
  if (c->comm && g_nccl.h) {
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    g_nccl.CommDestroy((ncclComm_t)c->comm);
  }
  c->comm = nullptr;
  c->comm_world = 1;
  c->comm_rank = 0;
  return TB_OK;
}

// in-place all-reduce MAX on the context's stream: what = TB_EXCH_KMAX (double) or TB_EXCH_LASTQ (int32)
int tb_comm_allreduce_max(tb_ctx *c, int what, void *device_ptr, int64_t count) {
  TB_REQUIRE(c->comm, TB_ERR_STATE, "sharded context without a communicator (tb_comm_init) or an exchange callback");
  TB_NCCL(g_nccl.AllReduce(device_ptr, device_ptr, (size_t)count, what == TB_EXCH_KMAX ? ncclDouble : ncclInt32, ncclMax,
                           (ncclComm_t)c->comm, c->stream));
  return TB_OK;
}

// Contiguous bands of the coarse grid, one per rank, balanced by the number of line groups centred in them
// (the line sum's cost per bin is groups x profile width, and the width profile is the same for every band);
// ends on multiples of the 128-bin accumulate tile, the last band ends at the grid end, bands may be empty.
// Every rank computes the same table from the same line index.
static int tb_cut_bands(tb_ctx *c, int world, const std::vector<double> *scale_per_tile, int64_t *edges);

extern "C" int tb_balanced_bands(tb_ctx *c, int world, int64_t *edges) {
  TB_REQUIRE(c->have_grid && world >= 1 && edges, TB_ERR_STATE, "tb_set_wn_grid must precede tb_balanced_bands");
  return tb_cut_bands(c, world, nullptr, edges);
}

// Measured rebalancing: the cost of a band is not proportional to its line groups (profiles widen with the
// wavenumber, the deep layers' wings cross band ends), so after a first spectrum every rank contributes the device
// time of its line sum (ncclAllGather of one double), each band's groups are re-weighted by the cost per group
// measured in it, and the grid is cut again.  The same table on every rank; results do not depend on the cut.
extern "C" int tb_rebalance_bands(tb_ctx *c) {
  TB_REQUIRE(c->comm && c->comm_world > 1 && (int)c->band_edges.size() == c->comm_world + 1, TB_ERR_STATE,
             "tb_rebalance_bands: tb_comm_init and tb_set_bands must come first");
  const int world = c->comm_world;
  float ms = 0.f, ms_k = 0.f;
  TB_TRY(tb_get_kernel_ms(c, "molext", &ms));
  // the strength-maximum stage ends with the all-reduce, i.e. with waiting for the slowest rank: not this band's cost
  if (tb_get_kernel_ms(c, "kmax", &ms_k) != TB_OK) ms_k = 0.f;
  TB_CUDA(cudaSetDevice(c->device));
  TB_TRY(c->g_send.ensure(1)); TB_TRY(c->g_recv.ensure(world));
  const double mine = ms - ms_k > 0.01f ? ms - ms_k : 0.01;
  TB_CUDA(cudaMemcpyAsync(c->g_send.p, &mine, sizeof(double), cudaMemcpyHostToDevice, c->stream));
  TB_NCCL(g_nccl.AllGather(c->g_send.p, c->g_recv.p, 1, ncclDouble, (ncclComm_t)c->comm, c->stream));
  std::vector<double> cost(world);
  TB_CUDA(cudaMemcpyAsync(cost.data(), c->g_recv.p, world * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  // cost per unit of weight inside each old band, as a scale per 128-bin tile
  const int64_t ntile = (c->nwn + 127) / 128;
  std::vector<double> scale(ntile, 1.0);
  std::vector<int64_t> one(world + 1);
  std::vector<double> unit(ntile, 1.0);
  {
    // weights of the old bands with unit scale
    std::vector<int64_t> tmp(world + 1);
    std::vector<double> wsum(world, 0.0);
    c->tile_weight.clear();
    TB_TRY(tb_cut_bands(c, world, &unit, tmp.data()));          // fills c->tile_weight
    for (int r = 0; r < world; r++)
      for (int64_t t = c->band_edges[r] / 128; t < (c->band_edges[r + 1] + 127) / 128 && t < ntile; t++) wsum[r] += c->tile_weight[t];
    for (int r = 0; r < world; r++)
      for (int64_t t = c->band_edges[r] / 128; t < (c->band_edges[r + 1] + 127) / 128 && t < ntile; t++)
        scale[t] = wsum[r] > 0 && cost[r] > 0 ? cost[r] / wsum[r] : 1.0;
  }
  std::vector<int64_t> edges(world + 1);
  TB_TRY(tb_cut_bands(c, world, &scale, edges.data()));
  return tb_set_bands(c, world, edges.data(), c->comm_rank);
}

static int tb_cut_bands(tb_ctx *c, int world, const std::vector<double> *scale_per_tile, int64_t *edges) {
  TB_CUDA(cudaSetDevice(c->device));
  const int64_t nwn = c->nwn;
  const int nrows = c->nrows;
  const bool lines = c->have_lines && c->G > 0;        // no line list (opacity-grid run): bands of equal width
  if (c->h_binstart.empty() && lines) {
    c->h_binstart.resize((size_t)nrows * nwn + 1);
    TB_CUDA(cudaMemcpyAsync(c->h_binstart.data(), c->binstart.p, c->h_binstart.size() * sizeof(int32_t),
                            cudaMemcpyDeviceToHost, c->stream));
    TB_CUDA(cudaStreamSynchronize(c->stream));
  }
  const int64_t ntile = (nwn + 127) / 128;
  std::vector<double> cum(ntile + 1, 0.0);
  for (int64_t t = 0; t < ntile; t++) {
    const int64_t j0 = t * 128, j1 = (t + 1) * 128 < nwn ? (t + 1) * 128 : nwn;
    double w = 1e-9 * (double)(j1 - j0);
    if (lines)
      for (int r = 0; r < nrows; r++) {
        // binstart[row][nwn] of the last row is G; within a row the entry past the last bin is the next row's first
        const size_t a = (size_t)r * nwn + j0, b = (size_t)r * nwn + j1;
        w += (double)(c->h_binstart[b] - c->h_binstart[a]);
      }
    if (scale_per_tile) {
      if ((int64_t)c->tile_weight.size() != ntile) c->tile_weight.assign(ntile, 0.0);
      c->tile_weight[t] = w;
      w *= (*scale_per_tile)[t];
    }
    cum[t + 1] = cum[t] + w;
  }
  int64_t prev = 0;
  edges[0] = 0;
  for (int r = 1; r < world; r++) {
    const double want = cum[ntile] * r / world;
    int64_t t = 0;
    while (t < ntile && cum[t] < want) t++;          // first tile boundary with cum >= want
    if (t < prev) t = prev;
    if (t > ntile) t = ntile;
    edges[r] = t * 128 < nwn ? t * 128 : nwn;
    prev = t;
  }
  edges[world] = nwn;
  return TB_OK;
}

// Band table of a sharded run: rank r owns coarse bins [edges[r], edges[r+1]).  Sets this context's shard
// (tb_set_shard) with the in-library exchange; the table is what the spectrum gather needs.
extern "C" int tb_set_bands(tb_ctx *c, int world, const int64_t *edges, int rank) {
  TB_REQUIRE(world >= 1 && edges && rank >= 0 && rank < world, TB_ERR_ARG, "tb_set_bands: rank %d of %d", rank, world);
  TB_REQUIRE(world == 1 || (c->comm && c->comm_world == world && c->comm_rank == rank), TB_ERR_STATE,
             "tb_set_bands: tb_comm_init(rank %d, world %d) must come first", rank, world);
  for (int r = 0; r < world; r++)
    TB_REQUIRE(edges[r] <= edges[r + 1], TB_ERR_ARG, "tb_set_bands: band edges must not decrease");
  TB_REQUIRE(edges[0] == 0 && edges[world] == c->nwn, TB_ERR_ARG, "tb_set_bands: the bands must cover the grid");
  c->band_edges.assign(edges, edges + world + 1);
  c->g_edges_world = 0;                 // the device copy of the table (spectrum gather) is stale
  c->exch = nullptr;
  c->exch_user = nullptr;
  if (world == 1) return tb_set_shard(c, 0, 0, 0, 1, nullptr, nullptr);
  // tb_set_shard wants a callback or a communicator for world > 1: the communicator is there
  return tb_set_shard(c, edges[rank], edges[rank + 1], rank, world, nullptr, nullptr);
}

__global__ void k_band_pack(const double *__restrict__ out, int64_t lo, int64_t n, int64_t wmax, double *__restrict__ buf) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < wmax) buf[i] = i < n ? out[lo + i] : 0.0;
}
__global__ void k_band_unpack(const double *__restrict__ all, const int64_t *__restrict__ edges, int world, int64_t wmax,
                              int64_t nwn, double *__restrict__ out) {
  const int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (w >= nwn) return;
  int r = 0;
  while (r + 1 < world && w >= edges[r + 1]) r++;
  out[w] = all[(size_t)r * wmax + (w - edges[r])];
}

// all-gather the band slices of the spectrum d_out[nwn] (this rank computed [band_lo, band_hi) of it) so that
// every rank holds the whole spectrum: pack -> ncclAllGather -> unpack, all on the context's stream
int tb_comm_gather_spectrum(tb_ctx *c) {
  if (!c->banded || c->shard_world <= 1 || !c->comm || c->band_edges.empty()) return TB_OK;
  const int world = c->shard_world;
  int64_t wmax = 0;
  for (int r = 0; r < world; r++) wmax = c->band_edges[r + 1] - c->band_edges[r] > wmax ? c->band_edges[r + 1] - c->band_edges[r] : wmax;
  if (wmax == 0) return TB_OK;
  TB_TRY(c->g_send.ensure(wmax)); TB_TRY(c->g_recv.ensure((size_t)world * wmax));
  if (c->g_edges_world != world) {
    TB_TRY(c->g_edges.ensure(world + 1));
    TB_CUDA(cudaMemcpyAsync(c->g_edges.p, c->band_edges.data(), (world + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, c->stream));
    TB_CUDA(cudaStreamSynchronize(c->stream));
    c->g_edges_world = world;
  }
  cudaStream_t st = c->stream;
  tb_timer_begin(c, "gather");
  k_band_pack<<<(unsigned)((wmax + 255) / 256), 256, 0, st>>>(c->d_out.p, c->band_lo, c->band_hi - c->band_lo, wmax, c->g_send.p);
  TB_NCCL(g_nccl.AllGather(c->g_send.p, c->g_recv.p, (size_t)wmax, ncclDouble, (ncclComm_t)c->comm, st));
  k_band_unpack<<<(unsigned)((c->nwn + 255) / 256), 256, 0, st>>>(c->g_recv.p, c->g_edges.p, world, wmax, c->nwn, c->d_out.p);
  tb_timer_end(c, "gather");
  c->launches += 2;
  TB_CUDA(cudaGetLastError());
  return TB_OK;
}

// every rank has enqueued and finished everything before it: a one-element all-reduce, then a stream sync
extern "C" int tb_comm_barrier(tb_ctx *c) {
  if (!c->comm || c->comm_world <= 1) return tb_sync(c);
  TB_CUDA(cudaSetDevice(c->device));
  TB_TRY(c->g_send.ensure(1));
  TB_CUDA(cudaMemsetAsync(c->g_send.p, 0, sizeof(double), c->stream));
  TB_NCCL(g_nccl.AllReduce(c->g_send.p, c->g_send.p, 1, ncclDouble, ncclMax, (ncclComm_t)c->comm, c->stream));
  TB_CUDA(cudaStreamSynchronize(c->stream));
  return TB_OK;
}

extern "C" int tb_get_bands(tb_ctx *c, int64_t *edges) {
  TB_REQUIRE(!c->band_edges.empty() && edges, TB_ERR_STATE, "tb_set_bands not called");
  memcpy(edges, c->band_edges.data(), c->band_edges.size() * sizeof(int64_t));
  return TB_OK;
}
