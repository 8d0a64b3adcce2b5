// gnc.cu -- C ABI of libgnc.so (include/gnc.h): contexts, launch logic, host-buffer path, gather.
//
// No torch types, no exceptions across the boundary.  Device code is in classify.cuh.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "classify.cuh"
#include "json_scan.cuh"

using gnc::Params;
using gnc::ScratchHeader;

namespace {

// ---- kernel variants: nodes per tile x pipeline depth x bytes per stage x CTAs per SM ------------
struct Variant {
  const char* name;
  int tile_nodes, stage_bytes, threads;
  size_t smem;
  const void* fn;
  void (*launch)(const Params&, int grid, size_t smem, cudaStream_t);
  const void* fn_fleet;  // same kernel with the fused multi-GPU gather compiled in
  void (*launch_fleet)(const Params&, int grid, size_t smem, cudaStream_t);
};

template <int TN, int ST, int SB, int MC, bool FLEET, int G>
void launch_variant(const Params& p, int grid, size_t smem, cudaStream_t stream) {
  gnc::classify_kernel<TN, ST, SB, MC, FLEET, G><<<grid, gnc::cta_threads<TN, G>(), smem, stream>>>(p);
}
template <int TN, int ST, int SB, int MC, int G = 1>
Variant make_variant(const char* name) {
  return Variant{name, TN, SB, gnc::cta_threads<TN, G>(), sizeof(gnc::Smem<TN, ST, SB, G>),
                 (const void*)gnc::classify_kernel<TN, ST, SB, MC, false, G>, launch_variant<TN, ST, SB, MC, false, G>,
                 (const void*)gnc::classify_kernel<TN, ST, SB, MC, true, G>, launch_variant<TN, ST, SB, MC, true, G>};
}

// Ordered by preference: the first of the first kPickable whose stage holds an average tile with 5 % to
// spare is used; the last of those reads tiles that do not fit straight from global memory.  The rest are the
// one-group kernels of the first round, kept selectable (GNC_VARIANT) for comparison and for the tests.
// name: t<nodes per tile>[s<stages>][g<walker groups per CTA>]
const Variant kVariants[] = {
    make_variant<256, 3, 65536, 1, 2>("t256g2"),   // rows up to ~243 B on average (ordinary nodes: ~207 B)
    make_variant<192, 3, 65536, 1, 2>("t192g2"),   // ~325 B
    make_variant<128, 3, 65536, 1, 2>("t128g2"),   // ~487 B
    make_variant<64, 3, 65536, 1, 2>("t64g2"),     // ~975 B
    make_variant<32, 2, 49152, 2>("t32"),
    make_variant<224, 2, 49792, 2>("t224"),
    make_variant<128, 3, 29696, 2>("t128s3"),
    make_variant<64, 2, 49152, 2>("t64"),
};
constexpr int kPickable = 5;
constexpr int kNumVariants = sizeof(kVariants) / sizeof(kVariants[0]);

}  // namespace

struct gnc_ctx {
  int device = 0, rank = 0, world = 1;
  int sm_count = 0;
  int max_digits = 4300;
  uint32_t epoch = 0;
  uint64_t launches = 0;
  int forced_variant = -1;
  int ctas_per_sm[kNumVariants] = {};
  // scratch: header + tile states
  void* scratch = nullptr;
  size_t scratch_tiles = 0;
  // host-buffer path staging
  struct HostStage {
    uint8_t* d_rows = nullptr;
    uint32_t* d_row_ptr = nullptr;
    uint8_t* d_flags = nullptr;
    uint32_t* d_idx = nullptr;
    int64_t* d_total = nullptr;
    int64_t* d_vals = nullptr;
    gnc_defer_rec* d_defer = nullptr;
    int64_t* d_counts = nullptr;
    int64_t* h_counts = nullptr;  // pinned
    size_t rows_cap = 0, nodes_cap = 0, defer_cap = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;    // counts are in h_counts
    cudaEvent_t kernel = nullptr;  // this stage's kernel has finished with the shared scratch
  } hs[2];
  unsigned long long* trace = nullptr;  // GNC_TRACE builds: set with gnc_debug_trace
  // NCCL (resolved at run time from the already-loaded libnccl.so.2)
  void* nccl_lib = nullptr;
  void* comm = nullptr;
  int64_t* d_all_counts = nullptr;  // [world][GNC_N_COUNTS]
  int64_t* h_all_counts = nullptr;  // pinned
  struct Fleet {                    // fused gather (gnc_fleet_enable)
    bool on = false;
    uint32_t cap = 0, slot_u32 = 0;
    size_t region_bytes = 0, bytes = 0;
    char* base[gnc::kFleetMaxWorld] = {};  // every rank's fleet buffer as mapped here (own: local allocation)
    uint32_t shard_nodes = 0;              // the largest shard a slot has room for
    unsigned long long epoch = 0;
    unsigned int* unpack_done = nullptr;
    // gnc_fleet_target: where the fleet-wide list goes, known before the classify launch
    uint32_t* target_idx = nullptr;
    int64_t* target_counts = nullptr;
    uint32_t target_cap = 0;
    unsigned long long unpacked_epoch = 0;  // the epoch whose classify launch did the rebuild itself
  } fleet;
  void* json_scratch = nullptr;     // gnc_json_to_pnl working arrays
  size_t json_scratch_bytes = 0;
  uint32_t* gather_buf = nullptr;   // [world][32 + gather_cap] u32: counts block + indices of every rank
  size_t gather_cap = 0;
  std::string err;
};

namespace {

int comm_init(gnc_ctx* c, const void* id, size_t id_len);
void comm_destroy(gnc_ctx* c);

int fail(gnc_ctx* c, int code, const std::string& msg) {
  if (c) c->err = msg;
  return code;
}
#define CUDA_TRY(c, expr)                                                                         \
  do {                                                                                            \
    cudaError_t e_ = (expr);                                                                      \
    if (e_ != cudaSuccess)                                                                        \
      return fail(c, GNC_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorName(e_) + ": " + cudaGetErrorString(e_)); \
  } while (0)

uint32_t tiles_for(uint32_t n_nodes, int tile_nodes) { return (n_nodes + tile_nodes - 1) / tile_nodes; }

int ensure_scratch(gnc_ctx* c, size_t tiles) {
  if (c->scratch && tiles <= c->scratch_tiles) return GNC_OK;
  size_t want = (std::max<size_t>(tiles, 4096) + gnc::kLookbackWindow - 1) / gnc::kLookbackWindow * gnc::kLookbackWindow;
  if (c->scratch) {
    CUDA_TRY(c, cudaDeviceSynchronize());
    CUDA_TRY(c, cudaFree(c->scratch));
    c->scratch = nullptr;
    c->scratch_tiles = 0;
  }
  size_t bytes = sizeof(ScratchHeader) + want * (sizeof(unsigned long long) + sizeof(uint32_t));
  CUDA_TRY(c, cudaMalloc(&c->scratch, bytes));
  CUDA_TRY(c, cudaMemset(c->scratch, 0, bytes));
  CUDA_TRY(c, cudaDeviceSynchronize());
  c->scratch_tiles = want;
  return GNC_OK;
}

int pick_variant(const gnc_ctx* c, uint64_t rows_bytes, uint32_t n_nodes) {
  if (c->forced_variant >= 0) return c->forced_variant;
  const double avg = n_nodes ? (double)rows_bytes / n_nodes : 0.0;
  for (int v = 0; v < kPickable; v++)
    if (avg * kVariants[v].tile_nodes * 1.05 + 32 <= kVariants[v].stage_bytes) return v;
  return kPickable - 1;  // tiles that do not fit are read straight from global memory
}

bool aligned16(const void* p) { return ((uintptr_t)p & 15u) == 0; }

int check_batch(gnc_ctx* c, const gnc_batch* in, const gnc_result* out) {
  if (!in || !out) return fail(c, GNC_ERR_ARG, "null batch or result");
  if (!out->counts) return fail(c, GNC_ERR_ARG, "result.counts is null");
  if (in->n_nodes) {
    if (!in->rows || !in->row_ptr) return fail(c, GNC_ERR_ARG, "batch.rows / batch.row_ptr is null");
    if (!aligned16(in->rows) || !aligned16(in->row_ptr))
      return fail(c, GNC_ERR_ARG, "batch.rows and batch.row_ptr must be 16-byte aligned");
    if (in->rows_bytes > 0xFFFFFFFFull - (1u << 20)) return fail(c, GNC_ERR_ARG, "rows_bytes exceeds the PNL-v2 limit (4 GiB - 1 MiB)");
    if (!out->flags || !out->gpu_idx || !out->gpu_total || !out->gpu_vals)
      return fail(c, GNC_ERR_ARG, "result.flags / gpu_idx / gpu_total / gpu_vals is null");
    if (!aligned16(out->gpu_vals)) return fail(c, GNC_ERR_ARG, "result.gpu_vals must be 16-byte aligned");
    if ((uint64_t)in->node_base + in->n_nodes > 0x100000000ull)
      return fail(c, GNC_ERR_ARG, "node_base + n_nodes exceeds 2^32");
    if (out->defer_cap && !out->defer) return fail(c, GNC_ERR_ARG, "result.defer is null but defer_cap > 0");
  }
  return GNC_OK;
}

// ---- fused gather: per-launch addresses inside every rank's fleet buffer -------------------------
// buffer layout: [parity 0: world slots][parity 1: world slots][flags: 2 x world u64][acks: world u64]
void fleet_params(gnc_ctx* c, Params* p) {
  gnc_ctx::Fleet& f = c->fleet;
  const unsigned long long e = ++f.epoch;
  const size_t parity = e & 1;
  p->fleet_world = (uint32_t)c->world;
  p->fleet_rank = (uint32_t)c->rank;
  p->fleet_cap = f.cap;
  p->fleet_slot_u32 = f.slot_u32;
  p->fleet_epoch = e;
  p->fleet_ack_need = e >= 3 ? e - 2 : 0;
  for (int q = 0; q < c->world; q++) {
    p->fleet_peer[q] = (uint32_t*)(f.base[q] + parity * f.region_bytes);
    p->fleet_flag[q] = (unsigned long long*)(f.base[q] + 2 * f.region_bytes) + parity * c->world;
  }
  p->fleet_ack = (const unsigned long long*)(f.base[c->rank] + 2 * f.region_bytes + 2 * (size_t)c->world * 8);
  p->fleet_out_idx = f.target_idx;
  p->fleet_out_counts = f.target_counts;
  p->fleet_out_cap = f.target_cap;
  p->fleet_unpack_done = f.unpack_done;
  for (int q = 0; q < c->world; q++)
    p->fleet_ack_peer[q] = (unsigned long long*)(f.base[q] + 2 * f.region_bytes + 2 * (size_t)c->world * 8);
}

__global__ void fleet_publish_empty_kernel(const Params p) {
  if (threadIdx.x == 0) {
    for (uint32_t q = 0; q < p.fleet_world; q++)
      for (int s = 0; s < GNC_N_COUNTS; s++)
        reinterpret_cast<long long*>(p.fleet_peer[q] + (size_t)p.fleet_rank * p.fleet_slot_u32)[s] = 0;
    for (uint32_t q = 0; q < p.fleet_world; q++) {
      uint32_t* h = p.fleet_peer[q] + (size_t)p.fleet_rank * p.fleet_slot_u32;
      h[gnc::FLEET_H_TILE_NODES] = 32, h[gnc::FLEET_H_TILES] = 0, h[gnc::FLEET_H_NODE_BASE] = p.node_base, h[gnc::FLEET_H_NODES] = 0;
    }
    __threadfence_system();
    for (uint32_t q = 0; q < p.fleet_world; q++) gnc::st_sys_u64(p.fleet_flag[q] + p.fleet_rank, p.fleet_epoch);
  }
}

struct AckPtrs {
  unsigned long long* p[gnc::kFleetMaxWorld];
};

// The rebuild as a launch of its own (gnc_gather when the target was not given to the classify launch,
// empty shards): see gnc::fleet_unpack in classify.cuh.
constexpr int kUnpackWarps = 8;
__global__ void __launch_bounds__(kUnpackWarps * 32) fleet_unpack_kernel(
    const uint32_t* slots, const unsigned long long* flags, unsigned long long epoch, int world, int rank, uint32_t slot_u32, uint32_t cap,
    uint32_t* out_idx, uint32_t out_cap, int64_t* out_counts, unsigned int* done, AckPtrs ack_peer) {
  __shared__ __align__(16) uint32_t buf[kUnpackWarps * 2 * gnc::kUnpackHalfWords];
  __shared__ unsigned long long sh[64];
  __shared__ uint32_t lut[256];
  gnc::FleetUnpack u;
  u.slots = slots;
  u.flags = flags;
  u.epoch = epoch;
  u.world = world;
  u.rank = rank;
  u.slot_u32 = slot_u32;
  u.cap = cap;
  u.out_idx = out_idx;
  u.out_cap = out_cap;
  u.out_counts = out_counts;
  u.done = done;
  u.ack_peer = ack_peer.p;
  u.timing = reinterpret_cast<unsigned long long*>(done) + 8;
  u.t_before = 0;
  gnc::fleet_unpack(u, buf, sh, lut);
}

}  // namespace

extern "C" {

int gnc_abi_version(void) { return GNC_ABI_VERSION; }

int gnc_create(gnc_ctx** out, int device, int rank, int world, const void* nccl_id, size_t nccl_id_len) {
  if (!out) return GNC_ERR_ARG;
  *out = nullptr;
  gnc_ctx* c = new (std::nothrow) gnc_ctx();
  if (!c) return GNC_ERR_NOMEM;
  *out = c;  // returned even on failure so the caller can read gnc_last_error, then gnc_destroy
  c->device = device;
  c->rank = rank;
  c->world = world < 1 ? 1 : world;
  if (rank < 0 || rank >= c->world) return fail(c, GNC_ERR_ARG, "rank outside [0, world)");
  CUDA_TRY(c, cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_TRY(c, cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10)
    return fail(c, GNC_ERR_STATE, std::string("libgnc is built for sm_100a (B200); device is ") + prop.name);
  c->sm_count = prop.multiProcessorCount;
  for (int v = 0; v < kNumVariants; v++) {
    CUDA_TRY(c, cudaFuncSetAttribute(kVariants[v].fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kVariants[v].smem));
    CUDA_TRY(c, cudaFuncSetAttribute(kVariants[v].fn_fleet, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kVariants[v].smem));
    int n = 0;
    CUDA_TRY(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kVariants[v].fn, kVariants[v].threads, kVariants[v].smem));
    if (n < 1) return fail(c, GNC_ERR_STATE, std::string("kernel variant does not fit an SM: ") + kVariants[v].name);
    c->ctas_per_sm[v] = n;
  }
  if (const char* f = getenv("GNC_VARIANT")) {
    for (int v = 0; v < kNumVariants; v++)
      if (!strcmp(f, kVariants[v].name)) c->forced_variant = v;
    if (c->forced_variant < 0) return fail(c, GNC_ERR_ARG, std::string("GNC_VARIANT names no kernel variant: ") + f);
  }
  int rc = ensure_scratch(c, 1 << 16);
  if (rc) return rc;
  if (c->world > 1) {
    rc = comm_init(c, nccl_id, nccl_id_len);
    if (rc) return rc;
  }
  return GNC_OK;
}

int gnc_reserve(gnc_ctx* c, uint32_t max_nodes) {
  if (!c) return GNC_ERR_ARG;
  CUDA_TRY(c, cudaSetDevice(c->device));
  return ensure_scratch(c, tiles_for(max_nodes, 32));  // the smallest tile any variant has
}

int gnc_set_max_str_digits(gnc_ctx* c, int max_digits) {
  if (!c || max_digits < 0) return GNC_ERR_ARG;
  c->max_digits = max_digits;
  return GNC_OK;
}

}  // extern "C"

namespace {
// gnc_classify with a lower bound for row offsets: the host-buffer path hands the kernel a staging
// buffer minus the chunk's first offset, so offsets below `rows_lo` must be contained, not followed.
int classify_impl(gnc_ctx* c, const gnc_batch* in, const gnc_result* out, void* cuda_stream, uint32_t rows_lo) {
  if (!c) return GNC_ERR_ARG;
  int rc = check_batch(c, in, out);
  if (rc) return rc;
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  CUDA_TRY(c, cudaSetDevice(c->device));
  if (in->n_nodes == 0) {
    CUDA_TRY(c, cudaMemsetAsync(out->counts, 0, GNC_N_COUNTS * sizeof(int64_t), stream));
    if (c->fleet.on) {  // an empty shard still has to tell the other ranks so
      Params p = {};
      fleet_params(c, &p);  // (the rebuild of this epoch is then gnc_gather's: unpacked_epoch stays behind)
      fleet_publish_empty_kernel<<<1, 32, 0, stream>>>(p);
      CUDA_TRY(c, cudaGetLastError());
      c->launches++;
    }
    return GNC_OK;
  }
  const int v = pick_variant(c, in->rows_bytes, in->n_nodes);
  const Variant& var = kVariants[v];
  const uint32_t tiles = tiles_for(in->n_nodes, var.tile_nodes);
  rc = ensure_scratch(c, tiles);
  if (rc) return rc;
  c->epoch = (c->epoch + 1) & gnc::kEpochMask;
  if (c->epoch == 0) {  // epoch wrapped: descriptors from 2^22 launches ago would look current
    CUDA_TRY(c, cudaMemsetAsync((char*)c->scratch + sizeof(ScratchHeader), 0,
                                c->scratch_tiles * (sizeof(unsigned long long) + sizeof(uint32_t)), stream));
    c->epoch = 1;
  }
  Params p = {};
  p.rows = in->rows;
  p.row_ptr = in->row_ptr;
  p.rows_bytes = in->rows_bytes;
  p.rows_lo = rows_lo;
  p.n_nodes = in->n_nodes;
  p.node_base = in->node_base;
  p.num_tiles = tiles;
  p.epoch = c->epoch;
  p.max_digits = c->max_digits;
  p.flags = out->flags;
  p.gpu_idx = out->gpu_idx;
  p.gpu_total = out->gpu_total;
  p.gpu_vals = out->gpu_vals;
  p.defer = out->defer;
  p.counts = out->counts;
  p.out_cap = out->out_cap;
  p.defer_cap = out->defer ? out->defer_cap : 0;
  p.scratch = (ScratchHeader*)c->scratch;
  p.inc = (unsigned long long*)((char*)c->scratch + sizeof(ScratchHeader));
  p.agg = (uint32_t*)(p.inc + c->scratch_tiles);
  p.trace = nullptr;
#ifdef GNC_TRACE
  p.trace = c->trace;
#endif
  if (c->fleet.on) {
    if (in->n_nodes > c->fleet.shard_nodes) return fail(c, GNC_ERR_ARG, "shard larger than the fleet buffers were sized for (gnc_fleet_enable)");
    fleet_params(c, &p);
  }
  const int grid = (int)std::min<uint32_t>(tiles, (uint32_t)(c->sm_count * c->ctas_per_sm[v]));
  if (c->fleet.on) {
    var.launch_fleet(p, grid, var.smem, stream);
    if (p.fleet_out_idx) c->fleet.unpacked_epoch = p.fleet_epoch;
  } else {
    var.launch(p, grid, var.smem, stream);
  }
  CUDA_TRY(c, cudaGetLastError());
  c->launches++;
  return GNC_OK;
}
}  // namespace

extern "C" {

int gnc_classify(gnc_ctx* c, const gnc_batch* in, const gnc_result* out, void* cuda_stream) {
  return classify_impl(c, in, out, cuda_stream, 0u);
}

uint64_t gnc_launch_count(const gnc_ctx* c) { return c ? c->launches : 0; }

const char* gnc_variant_for(uint64_t rows_bytes, uint32_t n_nodes) {
  gnc_ctx probe;  // forced_variant = -1: the choice made from the shard's shape alone
  return kVariants[pick_variant(&probe, rows_bytes, n_nodes)].name;
}

int gnc_unicode_tables(const uint32_t** nd_runs, int* n_runs, const uint32_t** spaces, int* n_spaces) {
  static const uint32_t runs[GNC_ND_RUNS] = {GNC_ND_RUN_LIST};
  static const uint32_t ws[GNC_WIDE_SPACES] = {GNC_WIDE_SPACE_LIST};
  if (nd_runs) *nd_runs = runs;
  if (n_runs) *n_runs = GNC_ND_RUNS;
  if (spaces) *spaces = ws;
  if (n_spaces) *n_spaces = GNC_WIDE_SPACES;
  return GNC_OK;
}

const char* gnc_last_error(const gnc_ctx* c) { return c ? c->err.c_str() : "null context"; }

#ifdef GNC_TRACE
// Debug builds only: device buffer of 296*64*8 u64 the next launches write phase timestamps into.
int gnc_debug_trace(gnc_ctx* c, void* device_buffer) {
  c->trace = (unsigned long long*)device_buffer;
  return GNC_OK;
}
#endif

}  // extern "C"

// =================================================================================================
// Host-buffer path: the shard lives in host memory; results come back to host memory.
// =================================================================================================
namespace {

template <class T>
int regrow(gnc_ctx* c, T** p, size_t count) {
  if (*p) CUDA_TRY(c, cudaFree(*p));
  *p = nullptr;
  CUDA_TRY(c, cudaMalloc((void**)p, count * sizeof(T)));
  return GNC_OK;
}

int ensure_host_stage(gnc_ctx* c, gnc_ctx::HostStage& h, size_t rows_bytes, size_t nodes, size_t defer_cap) {
  if (!h.stream) {
    CUDA_TRY(c, cudaStreamCreateWithFlags(&h.stream, cudaStreamNonBlocking));
    CUDA_TRY(c, cudaEventCreateWithFlags(&h.done, cudaEventDisableTiming));
    CUDA_TRY(c, cudaEventCreateWithFlags(&h.kernel, cudaEventDisableTiming));
    CUDA_TRY(c, cudaMalloc((void**)&h.d_counts, GNC_N_COUNTS * sizeof(int64_t)));
    CUDA_TRY(c, cudaMallocHost((void**)&h.h_counts, GNC_N_COUNTS * sizeof(int64_t)));
  }
  int rc;
  if (rows_bytes + 64 > h.rows_cap) {
    size_t cap = rows_bytes + rows_bytes / 8 + 4096;
    if ((rc = regrow(c, &h.d_rows, cap))) return rc;
    h.rows_cap = cap;
  }
  if (nodes > h.nodes_cap) {
    size_t cap = nodes + nodes / 8 + 1024;
    if ((rc = regrow(c, &h.d_row_ptr, cap + 8))) return rc;
    if ((rc = regrow(c, &h.d_flags, cap))) return rc;
    if ((rc = regrow(c, &h.d_idx, cap))) return rc;
    if ((rc = regrow(c, &h.d_total, cap))) return rc;
    if ((rc = regrow(c, &h.d_vals, cap * 4))) return rc;
    h.nodes_cap = cap;
  }
  if (defer_cap > h.defer_cap) {
    if ((rc = regrow(c, &h.d_defer, defer_cap))) return rc;
    h.defer_cap = defer_cap;
  }
  return GNC_OK;
}

}  // namespace

extern "C" int gnc_classify_host(gnc_ctx* c, const gnc_batch* in, const gnc_result* out) {
  if (!c) return GNC_ERR_ARG;
  if (!in || !out || !out->counts) return fail(c, GNC_ERR_ARG, "null batch, result or counts");
  CUDA_TRY(c, cudaSetDevice(c->device));
  memset(out->counts, 0, GNC_N_COUNTS * sizeof(int64_t));
  if (in->n_nodes == 0) return GNC_OK;
  if (!in->rows || !in->row_ptr || !out->flags || !out->gpu_idx || !out->gpu_total || !out->gpu_vals)
    return fail(c, GNC_ERR_ARG, "null host buffer");
  // The chunks below are classified on their own: the fused gather (one epoch per fleet-wide step)
  // stays out of it.
  struct FleetPause {
    gnc_ctx* c;
    bool was;
    explicit FleetPause(gnc_ctx* ctx) : c(ctx), was(ctx->fleet.on) { c->fleet.on = false; }
    ~FleetPause() { c->fleet.on = was; }
  } fleet_pause(c);
  // Chunks of ~1M nodes: chunk k+1's upload overlaps chunk k's kernel and download.
  uint32_t kChunk = 1u << 20;
  if (const char* e = getenv("GNC_HOST_CHUNK")) {  // tests: small chunks exercise the chunk seams
    const long v = atol(e);
    if (v >= 32 && v <= (1l << 24)) kChunk = (uint32_t)v;
  }
  const uint32_t n = in->n_nodes;
  uint64_t g_total = 0, d_total = 0;
  int64_t acc[GNC_N_COUNTS] = {0};
  unsigned __int128 key[4] = {0, 0, 0, 0};
  struct Pending {
    bool live = false;
    uint32_t first = 0, cnt = 0;
    uint64_t g_at = 0, d_at = 0;
  } pend[2];
  auto drain = [&](int b) -> int {
    gnc_ctx::HostStage& h = c->hs[b];
    Pending& q = pend[b];
    if (!q.live) return GNC_OK;
    CUDA_TRY(c, cudaEventSynchronize(h.done));
    const uint64_t g = (uint64_t)h.h_counts[GNC_C_GPU_NODES];
    const uint64_t d = std::min<uint64_t>((uint64_t)h.h_counts[GNC_C_DEFERRED], h.defer_cap);
    const uint64_t room = out->out_cap > g_total ? out->out_cap - g_total : 0, gk = std::min(g, room);
    CUDA_TRY(c, cudaMemcpyAsync(out->flags + q.first, h.d_flags, q.cnt, cudaMemcpyDeviceToHost, h.stream));
    if (gk) {
      CUDA_TRY(c, cudaMemcpyAsync(out->gpu_idx + g_total, h.d_idx, gk * 4, cudaMemcpyDeviceToHost, h.stream));
      CUDA_TRY(c, cudaMemcpyAsync(out->gpu_total + g_total, h.d_total, gk * 8, cudaMemcpyDeviceToHost, h.stream));
      CUDA_TRY(c, cudaMemcpyAsync(out->gpu_vals + 4 * g_total, h.d_vals, gk * 32, cudaMemcpyDeviceToHost, h.stream));
    }
    const uint64_t droom = out->defer && out->defer_cap > d_total ? out->defer_cap - d_total : 0, dk = std::min(d, droom);
    if (dk)
      CUDA_TRY(c, cudaMemcpyAsync(out->defer + d_total, h.d_defer, dk * sizeof(gnc_defer_rec), cudaMemcpyDeviceToHost, h.stream));
    CUDA_TRY(c, cudaStreamSynchronize(h.stream));
    for (int s : {GNC_C_GPU_NODES, GNC_C_READY_NODES, GNC_C_DEFERRED, GNC_C_SCANNED, GNC_C_MALFORMED})
      acc[s] += h.h_counts[s];
    acc[GNC_C_OUT_DROPPED] += h.h_counts[GNC_C_OUT_DROPPED] + (int64_t)(g - gk);
    acc[GNC_C_DEFER_DROPPED] += (int64_t)((uint64_t)h.h_counts[GNC_C_DEFERRED] - dk);  // the kernel's own drops are part of this difference
    for (int k = 0; k < 4; k++)
      key[k] += ((unsigned __int128)(uint64_t)h.h_counts[GNC_C_KEY_HI + k] << 64) | (uint64_t)h.h_counts[GNC_C_KEY_LO + k];
    g_total += gk;
    d_total += dk;
    q.live = false;
    return GNC_OK;
  };
  // One chunk: upload, classify, start the counts download.  Returns with the work enqueued.
  auto submit = [&](int b, uint32_t first, uint32_t defer_cap) -> int {
    const uint32_t cnt = std::min(kChunk, n - first);
    const uint64_t b0 = in->row_ptr[first], b1 = in->row_ptr[first + cnt];
    if (b1 < b0 || b1 > in->rows_bytes) return fail(c, GNC_ERR_ARG, "row_ptr is not monotone inside rows_bytes");
    gnc_ctx::HostStage& h = c->hs[b];
    int rc;
    if ((rc = ensure_host_stage(c, h, b1 - b0 + 16, cnt, defer_cap))) return rc;
    // rows are copied from a 16-byte aligned source offset so row_ptr can stay absolute minus a constant
    const uint64_t a0 = b0 & ~15ull;
    CUDA_TRY(c, cudaMemcpyAsync(h.d_rows, in->rows + a0, b1 - a0, cudaMemcpyHostToDevice, h.stream));
    CUDA_TRY(c, cudaMemsetAsync(h.d_rows + (b1 - a0), 0, 48, h.stream));
    CUDA_TRY(c, cudaMemcpyAsync(h.d_row_ptr, in->row_ptr + first, (cnt + 1) * 4ull, cudaMemcpyHostToDevice, h.stream));
    gnc_batch db;
    db.rows = h.d_rows - a0;  // device offsets == host offsets, so are defer offsets; the kernel treats any
    db.row_ptr = h.d_row_ptr; // offset below a0 (a corrupt row_ptr inside the chunk) as a malformed row
    db.rows_bytes = b1;
    db.n_nodes = cnt;
    db.node_base = in->node_base + first;
    gnc_result dr;
    dr.flags = h.d_flags;
    dr.gpu_idx = h.d_idx;
    dr.gpu_total = h.d_total;
    dr.gpu_vals = h.d_vals;
    dr.defer = h.d_defer;
    dr.counts = h.d_counts;
    dr.out_cap = cnt;
    dr.defer_cap = (uint32_t)h.defer_cap;
    // the tail of row_ptr the bulk copy may touch
    CUDA_TRY(c, cudaMemsetAsync(h.d_row_ptr + cnt + 1, 0xFF, 7 * 4, h.stream));
    // both stages' kernels share the context's scratch: this one starts after the other stage's has ended
    if (c->hs[b ^ 1].kernel && pend[b ^ 1].live) CUDA_TRY(c, cudaStreamWaitEvent(h.stream, c->hs[b ^ 1].kernel, 0));
    if ((rc = classify_impl(c, &db, &dr, h.stream, (uint32_t)a0))) return rc;
    CUDA_TRY(c, cudaEventRecord(h.kernel, h.stream));
    CUDA_TRY(c, cudaMemcpyAsync(h.h_counts, h.d_counts, GNC_N_COUNTS * sizeof(int64_t), cudaMemcpyDeviceToHost, h.stream));
    CUDA_TRY(c, cudaEventRecord(h.done, h.stream));
    pend[b].live = true;
    pend[b].first = first;
    pend[b].cnt = cnt;
    return GNC_OK;
  };
  // An error mid-pipeline leaves the other stage in flight: wait for both before the staging buffers
  // (owned by the context, reused by the next call) and the caller's host buffers are let go.
  auto bail = [&](int rc) -> int {
    const std::string keep = c->err;
    for (auto& h : c->hs)
      if (h.stream) cudaStreamSynchronize(h.stream);
    pend[0].live = pend[1].live = false;
    c->err = keep;
    return rc;
  };
  int b = 0;
  const uint32_t defer_cap = out->defer ? std::max<uint32_t>(out->defer_cap, 1) : 1;
  for (uint32_t first = 0; first < n; first += kChunk, b ^= 1) {
    int rc = drain(b);
    if (!rc) rc = submit(b, first, defer_cap);
    if (rc) return bail(rc);
  }
  for (int k = 0; k < 2; k++, b ^= 1) {
    int rc = drain(b);
    if (rc) return bail(rc);
  }
  memcpy(out->counts, acc, sizeof(acc));
  for (int k = 0; k < 4; k++) {
    out->counts[GNC_C_KEY_LO + k] = (int64_t)(uint64_t)key[k];
    out->counts[GNC_C_KEY_HI + k] = (int64_t)(uint64_t)(key[k] >> 64);
  }
  if (acc[GNC_C_OUT_DROPPED] || acc[GNC_C_DEFER_DROPPED])
    return fail(c, GNC_ERR_OVERFLOW, "out_cap or defer_cap too small; see counts[13], counts[14]");
  return GNC_OK;
}

extern "C" void gnc_destroy(gnc_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  comm_destroy(c);
  for (auto& h : c->hs) {
    if (h.stream) cudaStreamDestroy(h.stream);
    if (h.done) cudaEventDestroy(h.done);
    if (h.kernel) cudaEventDestroy(h.kernel);
    cudaFree(h.d_rows);
    cudaFree(h.d_row_ptr);
    cudaFree(h.d_flags);
    cudaFree(h.d_idx);
    cudaFree(h.d_total);
    cudaFree(h.d_vals);
    cudaFree(h.d_defer);
    cudaFree(h.d_counts);
    if (h.h_counts) cudaFreeHost(h.h_counts);
  }
  cudaFree(c->scratch);
  cudaFree(c->json_scratch);
  delete c;
}


// =================================================================================================
// Multi-GPU: NCCL, resolved at run time so the library loads (and the CPU tests run) without it and
// so a process that already holds a libnccl.so.2 (torch's) shares that one copy.
// =================================================================================================
namespace {

struct NcclId {
  char bytes[128];
};
struct NcclApi {
  int (*GetUniqueId)(NcclId*) = nullptr;
  int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
  int (*Broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  void* lib = nullptr;
  std::string err;
};
constexpr int kNcclUint8 = 1, kNcclUint32 = 3, kNcclInt64 = 4;

NcclApi* nccl_api() {
  static NcclApi api;
  if (api.lib || !api.err.empty()) return &api;
  const char* names[] = {getenv("GNC_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    if (!n) continue;
    api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (api.lib) break;
  }
  if (!api.lib) {
    api.err = std::string("cannot load libnccl: ") + dlerror();
    return &api;
  }
#define GNC_SYM(field, name)                                        \
  *(void**)(&api.field) = dlsym(api.lib, name);                      \
  if (!api.field) api.err = std::string("libnccl lacks ") + name;
  GNC_SYM(GetUniqueId, "ncclGetUniqueId")
  GNC_SYM(CommInitRank, "ncclCommInitRank")
  GNC_SYM(CommDestroy, "ncclCommDestroy")
  GNC_SYM(AllGather, "ncclAllGather")
  GNC_SYM(Broadcast, "ncclBroadcast")
  GNC_SYM(GroupStart, "ncclGroupStart")
  GNC_SYM(GroupEnd, "ncclGroupEnd")
  GNC_SYM(GetErrorString, "ncclGetErrorString")
#undef GNC_SYM
  return &api;
}

#define NCCL_TRY(c, expr)                                                                                 \
  do {                                                                                                    \
    int r_ = (expr);                                                                                      \
    if (r_ != 0) return fail(c, GNC_ERR_COMM, std::string(#expr) + ": " + nccl_api()->GetErrorString(r_)); \
  } while (0)

int comm_init(gnc_ctx* c, const void* id, size_t id_len) {
  NcclApi* n = nccl_api();
  if (!n->err.empty()) return fail(c, GNC_ERR_COMM, n->err);
  if (!id || id_len != sizeof(NcclId)) return fail(c, GNC_ERR_ARG, "world > 1 needs the 128-byte id from gnc_comm_id");
  NcclId nid;
  memcpy(nid.bytes, id, sizeof(nid));
  NCCL_TRY(c, n->CommInitRank(&c->comm, c->world, nid, c->rank));
  CUDA_TRY(c, cudaMalloc((void**)&c->d_all_counts, (size_t)c->world * GNC_N_COUNTS * sizeof(int64_t)));
  CUDA_TRY(c, cudaMallocHost((void**)&c->h_all_counts, ((size_t)c->world * GNC_N_COUNTS + 1) * sizeof(int64_t)));
  return GNC_OK;
}

void comm_destroy(gnc_ctx* c) {
  if (c->fleet.on) {
    for (int q = 0; q < c->world; q++)
      if (q != c->rank && c->fleet.base[q]) cudaIpcCloseMemHandle(c->fleet.base[q]);
    cudaFree(c->fleet.base[c->rank]);
    c->fleet.on = false;
  }
  if (c->comm) nccl_api()->CommDestroy(c->comm);
  c->comm = nullptr;
  cudaFree(c->d_all_counts);
  cudaFree(c->gather_buf);
  if (c->h_all_counts) cudaFreeHost(c->h_all_counts);
}

}  // namespace

namespace {

// One rank's slot of the gather buffer, in u32 units: [0, 32) its counts block, [32, 32 + cap) its indices.
constexpr uint32_t kSlotHeader = 2 * GNC_N_COUNTS;

__global__ void gather_pack_kernel(const int64_t* counts, const uint32_t* idx, uint32_t* slot, uint32_t cap) {
  const unsigned long long g = min((unsigned long long)counts[GNC_C_GPU_NODES], (unsigned long long)cap);
  if (blockIdx.x == 0 && threadIdx.x < kSlotHeader) slot[threadIdx.x] = reinterpret_cast<const uint32_t*>(counts)[threadIdx.x];
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < g; i += (unsigned long long)gridDim.x * blockDim.x)
    slot[kSlotHeader + i] = idx[i];
}

// Every rank's slot -> one dense, rank-ordered index list and the exact sum of the counts blocks.
__global__ void gather_unpack_kernel(const uint32_t* buf, uint32_t slot_u32, uint32_t cap, int world, uint32_t* out_idx,
                                     uint32_t out_cap, int64_t* out_counts) {
  __shared__ unsigned long long off[65];
  if (threadIdx.x == 0) {
    off[0] = 0;
    for (int r = 0; r < world; r++) off[r + 1] = off[r] + (unsigned long long)reinterpret_cast<const int64_t*>(buf + (size_t)r * slot_u32)[GNC_C_GPU_NODES];
  }
  __syncthreads();
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x, t0 = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  for (int r = 0; r < world; r++) {
    const uint32_t* src = buf + (size_t)r * slot_u32 + kSlotHeader;
    const unsigned long long cnt = min(off[r + 1] - off[r], (unsigned long long)cap);
    for (unsigned long long i = t0; i < cnt; i += stride)
      if (off[r] + i < out_cap) out_idx[off[r] + i] = src[i];
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    long long sum[GNC_N_COUNTS];
    unsigned __int128 key[4] = {0, 0, 0, 0};
    unsigned long long lost = 0;
    for (int s = 0; s < GNC_N_COUNTS; s++) sum[s] = 0;
    for (int r = 0; r < world; r++) {
      const long long* cr = reinterpret_cast<const long long*>(buf + (size_t)r * slot_u32);
      for (int s = 0; s < GNC_N_COUNTS; s++) sum[s] += cr[s];
      for (int k = 0; k < 4; k++)
        key[k] += ((unsigned __int128)(unsigned long long)cr[GNC_C_KEY_HI + k] << 64) | (unsigned long long)cr[GNC_C_KEY_LO + k];
      const unsigned long long cnt = off[r + 1] - off[r];
      if (cnt > cap) lost += cnt - cap;
    }
    if (off[world] > out_cap) lost = max(lost, off[world] - out_cap);
    for (int k = 0; k < 4; k++) {
      sum[GNC_C_KEY_LO + k] = (long long)(unsigned long long)key[k];
      sum[GNC_C_KEY_HI + k] = (long long)(unsigned long long)(key[k] >> 64);
    }
    sum[GNC_C_OUT_DROPPED] += (long long)lost;
    for (int s = 0; s < GNC_N_COUNTS; s++) out_counts[s] = sum[s];
  }
}

}  // namespace

extern "C" int gnc_comm_id(void* id, size_t cap, size_t* len) {
  NcclApi* n = nccl_api();
  if (!n->err.empty() || !id || cap < sizeof(NcclId)) return n->err.empty() ? GNC_ERR_ARG : GNC_ERR_COMM;
  NcclId nid;
  if (n->GetUniqueId(&nid) != 0) return GNC_ERR_COMM;
  memcpy(id, nid.bytes, sizeof(nid));
  if (len) *len = sizeof(nid);
  return GNC_OK;
}

extern "C" int gnc_fleet_enable(gnc_ctx* c, uint32_t fleet_nodes_cap, void* cuda_stream) {
  if (!c) return GNC_ERR_ARG;
  if (c->world == 1) return GNC_OK;
  if (c->world > gnc::kFleetMaxWorld) return fail(c, GNC_ERR_ARG, "the fused gather supports up to 8 ranks");
  if (!c->comm) return fail(c, GNC_ERR_STATE, "context has no communicator");
  if (c->fleet.on) return fail(c, GNC_ERR_STATE, "the fused gather is already enabled");
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  CUDA_TRY(c, cudaSetDevice(c->device));
  NcclApi* n = nccl_api();
  gnc_ctx::Fleet& f = c->fleet;
  // a shard of up to ceil(fleet_nodes_cap / world) nodes; tile records take at most 2 words per 32 nodes (the 32-node tile variant)
  f.shard_nodes = (uint32_t)(((size_t)fleet_nodes_cap + c->world - 1) / c->world);
  f.cap = (uint32_t)(((((size_t)f.shard_nodes + 31) / 32) * 2 + 8 + 3) & ~(size_t)3);
  f.slot_u32 = gnc::kFleetHeader + f.cap;
  f.region_bytes = ((size_t)c->world * f.slot_u32 * 4 + 255) & ~(size_t)255;
  f.bytes = 2 * f.region_bytes + 3 * (size_t)c->world * 8 + 256;
  // Every step below is collective.  A rank on which something fails still takes part in the two
  // all-gathers (the second carries a success flag from every rank), so no rank is left waiting in a
  // collective for one that has already returned; the fused gather is enabled only if all succeeded.
  char* local = nullptr;
  char* d_handles = nullptr;
  int opened[gnc::kFleetMaxWorld] = {};
  auto undo = [&]() {
    for (int q = 0; q < c->world; q++)
      if (opened[q]) cudaIpcCloseMemHandle(f.base[q]);
    for (int q = 0; q < c->world; q++) f.base[q] = nullptr;
    if (local) cudaFree(local);
    if (d_handles) cudaFree(d_handles);
  };
  // a record per rank: its IPC handle, then a 64-byte status block whose first byte says "failed"
  constexpr size_t kRec = sizeof(cudaIpcMemHandle_t) + 64;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "record layout");
  struct Rec {
    cudaIpcMemHandle_t h;
    unsigned char bad[64];
  };
  static_assert(sizeof(Rec) == kRec, "record layout");
  Rec mine = {};
  std::string why;
  auto note = [&](cudaError_t e, const char* what) {
    if (e != cudaSuccess && why.empty()) why = std::string(what) + ": " + cudaGetErrorString(e);
    return e == cudaSuccess;
  };
  bool ok = note(cudaMalloc((void**)&local, f.bytes), "cudaMalloc(fleet buffer)") && note(cudaMemset(local, 0, f.bytes), "cudaMemset") &&
            note(cudaIpcGetMemHandle(&mine.h, local), "cudaIpcGetMemHandle");
  mine.bad[0] = ok ? 0 : 1;
  std::vector<Rec> all(c->world);
  if (cudaMalloc((void**)&d_handles, kRec * c->world) != cudaSuccess) {
    // without this staging buffer the rank cannot even take part in the collective: nothing to do but fail here
    undo();
    return fail(c, GNC_ERR_NOMEM, "gnc_fleet_enable: cannot allocate the handle exchange buffer");
  }
  auto exchange = [&]() -> int {  // every rank's record to every rank (through the communicator that already exists)
    CUDA_TRY(c, cudaMemcpyAsync(d_handles + kRec * c->rank, &mine, kRec, cudaMemcpyHostToDevice, stream));
    NCCL_TRY(c, n->AllGather(d_handles + kRec * c->rank, d_handles, kRec, kNcclUint8, c->comm, stream));
    CUDA_TRY(c, cudaMemcpyAsync(all.data(), d_handles, kRec * c->world, cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(c, cudaStreamSynchronize(stream));
    return GNC_OK;
  };
  int rc = exchange();
  if (rc) {
    undo();
    return rc;
  }
  for (int q = 0; q < c->world; q++) ok = ok && !all[q].bad[0];
  if (ok) {
    for (int q = 0; q < c->world && ok; q++) {
      if (q == c->rank) {
        f.base[q] = local;
      } else {
        void* ptr = nullptr;
        ok = note(cudaIpcOpenMemHandle(&ptr, all[q].h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
        if (ok) {
          f.base[q] = (char*)ptr;
          opened[q] = 1;
        }
      }
    }
  }
  // nobody writes into a peer before every peer has zeroed its buffer and mapped the others: one more
  // trip round the ranks, carrying each rank's verdict
  mine.bad[0] = ok ? 0 : 1;
  rc = exchange();
  if (rc) {
    undo();
    return rc;
  }
  int first_bad = -1;
  for (int q = 0; q < c->world; q++)
    if (all[q].bad[0] && first_bad < 0) first_bad = q;
  if (first_bad >= 0) {
    undo();
    cudaGetLastError();
    return fail(c, GNC_ERR_CUDA, "gnc_fleet_enable: peer mapping unavailable (first failing rank " + std::to_string(first_bad) + ")" +
                                     (why.empty() ? "" : "; here: " + why));
  }
  CUDA_TRY(c, cudaFree(d_handles));
  d_handles = nullptr;
  f.unpack_done = (unsigned int*)(local + f.bytes - 256);
  f.epoch = 0;
  f.on = true;
  return GNC_OK;
}

extern "C" int gnc_fleet_target(gnc_ctx* c, const gnc_result* all) {
  if (!c) return GNC_ERR_ARG;
  gnc_ctx::Fleet& f = c->fleet;
  if (!all) {
    f.target_idx = nullptr;
    f.target_counts = nullptr;
    f.target_cap = 0;
    return GNC_OK;
  }
  if (!all->gpu_idx || !all->counts) return fail(c, GNC_ERR_ARG, "gnc_fleet_target: gpu_idx or counts is null");
  f.target_idx = all->gpu_idx;
  f.target_counts = all->counts;
  f.target_cap = all->out_cap;
  return GNC_OK;
}

extern "C" int gnc_fleet_timing(gnc_ctx* c, uint64_t out[8]) {
  if (!c || !out) return GNC_ERR_ARG;
  if (!c->fleet.on) return fail(c, GNC_ERR_STATE, "the fused gather is not enabled");
  CUDA_TRY(c, cudaSetDevice(c->device));
  unsigned long long* t = reinterpret_cast<unsigned long long*>(c->fleet.unpack_done) + 8;
  CUDA_TRY(c, cudaMemcpy(out, t, 8 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  CUDA_TRY(c, cudaMemset(t, 0, 8 * sizeof(uint64_t)));
  return GNC_OK;
}

extern "C" int gnc_gather(gnc_ctx* c, const gnc_result* local, const gnc_result* all, void* cuda_stream) {
  if (!c) return GNC_ERR_ARG;
  if (!local || !all || !local->counts || !all->counts) return fail(c, GNC_ERR_ARG, "null result or counts");
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  CUDA_TRY(c, cudaSetDevice(c->device));
  if (c->world == 1) {  // nothing to combine: `all` may alias `local`
    if (all->counts != local->counts)
      CUDA_TRY(c, cudaMemcpyAsync(all->counts, local->counts, GNC_N_COUNTS * sizeof(int64_t), cudaMemcpyDeviceToDevice, stream));
    return GNC_OK;
  }
  if (!c->comm) return fail(c, GNC_ERR_STATE, "context has no communicator");
  NcclApi* n = nccl_api();
  if (c->fleet.on && c->fleet.epoch && !all->gpu_total && !all->gpu_vals) {
    // Fused gather: the classify kernel already stored this rank's indices and counts into every
    // rank's fleet buffer over NVLink.  What is left is local: wait for the other ranks' flags, pack
    // the slots into the dense rank-ordered list, add the counts up, acknowledge.  No collective.
    if (!all->gpu_idx) return fail(c, GNC_ERR_ARG, "gpu_idx is null");
    gnc_ctx::Fleet& f = c->fleet;
    if (f.unpacked_epoch == f.epoch && f.target_idx == all->gpu_idx && f.target_counts == all->counts && f.target_cap == all->out_cap)
      return GNC_OK;  // the classify launch of this epoch has rebuilt the list into this very target already
    const size_t parity = f.epoch & 1;
    AckPtrs ap = {};
    for (int q = 0; q < c->world; q++) ap.p[q] = (unsigned long long*)(f.base[q] + 2 * f.region_bytes + 2 * (size_t)c->world * 8);
    fleet_unpack_kernel<<<2 * c->sm_count, kUnpackWarps * 32, 0, stream>>>(
        (const uint32_t*)(f.base[c->rank] + parity * f.region_bytes),
        (const unsigned long long*)(f.base[c->rank] + 2 * f.region_bytes) + parity * c->world, f.epoch, c->world, c->rank, f.slot_u32,
        f.cap, all->gpu_idx, all->out_cap, all->counts, f.unpack_done, ap);
    CUDA_TRY(c, cudaGetLastError());
    c->launches += 1;
    return GNC_OK;
  }
  if (!all->gpu_total && !all->gpu_vals) {
    // Indices and counts only (the north star's gather): ONE in-place all-gather of fixed-size slots
    // -- each rank's counts block followed by its indices -- then a kernel that packs the slots into
    // the dense rank-ordered list and adds the counts up.  No host round trip, nothing waits on the CPU.
    if (!all->gpu_idx || !local->gpu_idx) return fail(c, GNC_ERR_ARG, "gpu_idx is null");
    const size_t cap = (((size_t)all->out_cap + c->world - 1) / c->world + 3) & ~(size_t)3;   // same on every rank
    if (cap > c->gather_cap || !c->gather_buf) {
      if (c->gather_buf) {
        CUDA_TRY(c, cudaStreamSynchronize(stream));
        CUDA_TRY(c, cudaFree(c->gather_buf));
        c->gather_buf = nullptr;
      }
      CUDA_TRY(c, cudaMalloc((void**)&c->gather_buf, (size_t)c->world * (kSlotHeader + cap) * sizeof(uint32_t)));
      c->gather_cap = cap;
    }
    const uint32_t slot_u32 = (uint32_t)(kSlotHeader + c->gather_cap);
    uint32_t* mine = c->gather_buf + (size_t)c->rank * slot_u32;
    const int blocks = 4 * c->sm_count;
    gather_pack_kernel<<<blocks, 256, 0, stream>>>(local->counts, local->gpu_idx, mine, (uint32_t)c->gather_cap);
    CUDA_TRY(c, cudaGetLastError());
    NCCL_TRY(c, n->AllGather(mine, c->gather_buf, slot_u32, kNcclUint32, c->comm, stream));
    gather_unpack_kernel<<<blocks, 256, 0, stream>>>(c->gather_buf, slot_u32, (uint32_t)c->gather_cap, c->world, all->gpu_idx,
                                                     all->out_cap, all->counts);
    CUDA_TRY(c, cudaGetLastError());
    c->launches += 3;
    return GNC_OK;
  }
  // Records too: offsets are needed on the host, so this path synchronises the stream once.
  // 1. every rank's counts block to every rank; the sum (exact, 128-bit for the key totals) is the
  //    all-reduce of the counts, and the per-rank GPU-node counts give the gather offsets
  //    (slot 15 of the block is spare: it carries each rank's out_cap, so that "too small" is decided from
  //    numbers every rank has -- a rank that returned on its own would leave the others in the group below)
  constexpr int kBlk = GNC_N_COUNTS;
  int64_t* mine_blk = c->d_all_counts + (size_t)c->rank * kBlk;
  CUDA_TRY(c, cudaMemcpyAsync(mine_blk, local->counts, (GNC_N_COUNTS - 1) * sizeof(int64_t), cudaMemcpyDeviceToDevice, stream));
  c->h_all_counts[(size_t)c->world * kBlk] = (int64_t)all->out_cap;   // pinned staging word past the blocks
  CUDA_TRY(c, cudaMemcpyAsync(mine_blk + (GNC_N_COUNTS - 1), c->h_all_counts + (size_t)c->world * kBlk, sizeof(int64_t),
                              cudaMemcpyHostToDevice, stream));
  NCCL_TRY(c, n->AllGather(mine_blk, c->d_all_counts, kBlk, kNcclInt64, c->comm, stream));
  CUDA_TRY(c, cudaMemcpyAsync(c->h_all_counts, c->d_all_counts, (size_t)c->world * kBlk * sizeof(int64_t),
                              cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(c, cudaStreamSynchronize(stream));
  int64_t sum[GNC_N_COUNTS] = {0};
  unsigned __int128 key[4] = {0, 0, 0, 0};
  std::vector<uint64_t> off(c->world + 1, 0);
  uint64_t min_cap = ~0ull;
  for (int r = 0; r < c->world; r++) {
    const int64_t* cr = c->h_all_counts + (size_t)r * kBlk;
    for (int s = 0; s < GNC_N_COUNTS - 1; s++) sum[s] += cr[s];
    for (int k = 0; k < 4; k++)
      key[k] += ((unsigned __int128)(uint64_t)cr[GNC_C_KEY_HI + k] << 64) | (uint64_t)cr[GNC_C_KEY_LO + k];
    off[r + 1] = off[r] + (uint64_t)cr[GNC_C_GPU_NODES];
    min_cap = std::min<uint64_t>(min_cap, (uint64_t)cr[GNC_N_COUNTS - 1]);
  }
  for (int k = 0; k < 4; k++) {
    sum[GNC_C_KEY_LO + k] = (int64_t)(uint64_t)key[k];
    sum[GNC_C_KEY_HI + k] = (int64_t)(uint64_t)(key[k] >> 64);
  }
  if (off[c->world] > min_cap)   // the same verdict on every rank
    return fail(c, GNC_ERR_OVERFLOW, "some rank's all.out_cap is smaller than the fleet's GPU-node count");
  memcpy(c->h_all_counts, sum, sizeof(sum));  // reuse the pinned block as the upload source
  CUDA_TRY(c, cudaMemcpyAsync(all->counts, c->h_all_counts, sizeof(sum), cudaMemcpyHostToDevice, stream));
  // 2. all-gather-v of the compacted records: one broadcast per rank into its rank-ordered slice
  NCCL_TRY(c, n->GroupStart());
  for (int r = 0; r < c->world; r++) {
    const uint64_t cnt = off[r + 1] - off[r];
    if (!cnt) continue;
    NCCL_TRY(c, n->Broadcast(local->gpu_idx, all->gpu_idx + off[r], cnt, kNcclUint32, r, c->comm, stream));
    if (all->gpu_total && local->gpu_total)
      NCCL_TRY(c, n->Broadcast(local->gpu_total, all->gpu_total + off[r], cnt, kNcclInt64, r, c->comm, stream));
    if (all->gpu_vals && local->gpu_vals)
      NCCL_TRY(c, n->Broadcast(local->gpu_vals, all->gpu_vals + 4 * off[r], 4 * cnt, kNcclInt64, r, c->comm, stream));
  }
  NCCL_TRY(c, n->GroupEnd());
  c->launches += 2;
  return GNC_OK;
}


// =================================================================================================
// Classification without any allocation (gnc_classify_static, include/gnc.h).
//
// Some hosts of this library may not allocate device memory at all once they are running: an agent
// admitted to a node under a hard device-memory cap that is already spent on its buffers, or a
// caller on a latency path that cannot afford cudaMalloc's device-wide synchronisation (or its
// failure).  For them the kernel's scratch -- the tile ticket, the accumulators and the look-back
// descriptors that gnc_create otherwise cudaMallocs -- lives in the file-scope array below, which
// every GPU that loads libgnc.so carries from module load until the process exits.
// =================================================================================================
namespace {

constexpr size_t kStaticScratchBytes = sizeof(ScratchHeader) + (size_t)GNC_STATIC_MAX_TILES * (sizeof(unsigned long long) + sizeof(uint32_t));
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// Per-device host-side state for the static path.  `scratch` is the DEVICE address of
// g_static_scratch on that device: in host code the array's name is only a handle for the runtime,
// not an address a kernel could use, so the address is asked for once per device.
constexpr int kStaticMaxDevices = 64;
struct StaticState {
  bool ready;
  void* scratch;
  uint32_t epoch;
  int sm_count;
  int ctas_per_sm[kNumVariants];
};
StaticState g_static_state[kStaticMaxDevices];
// Host threads that call in for the same or different devices take turns through the set-up, the epoch
// bump and the launch; this orders the HOST side only -- kernels on different streams still share the array.
std::mutex g_static_mutex;

int static_fail(char* err, size_t cap, int code, const char* what, const char* detail) {
  if (err && cap) snprintf(err, cap, "%s%s%s", what, detail ? ": " : "", detail ? detail : "");
  return code;
}
#define STATIC_CUDA_TRY(expr)                                                              \
  do {                                                                                     \
    cudaError_t e_ = (expr);                                                               \
    if (e_ != cudaSuccess) return static_fail(err, err_cap, GNC_ERR_CUDA, #expr, cudaGetErrorString(e_)); \
  } while (0)

}  // namespace

extern "C" int gnc_classify_static(int device, const gnc_batch* in, const gnc_result* out, int max_str_digits,
                                   void* cuda_stream, char* err, size_t err_cap) {
  if (err && err_cap) err[0] = 0;
  if (device < 0 || device >= kStaticMaxDevices) return static_fail(err, err_cap, GNC_ERR_ARG, "device index outside [0, 64)", nullptr);
  if (!in || !out || !out->counts) return static_fail(err, err_cap, GNC_ERR_ARG, "null batch, result or counts", nullptr);
  if (max_str_digits < 0) return static_fail(err, err_cap, GNC_ERR_ARG, "max_str_digits is negative", nullptr);
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  std::lock_guard<std::mutex> guard(g_static_mutex);
  STATIC_CUDA_TRY(cudaSetDevice(device));
  StaticState& st = g_static_state[device];
  if (!st.ready) {
    // First call on this device (it is current): learn the device address of the scratch array and
    // the launch geometry.  No cudaMalloc here either, but this is where the CUDA runtime may create
    // the device context and load this library's module -- which is when the scratch array and the
    // kernel images take up their device memory -- so it can fail on a device whose memory is already
    // full, and it must not run inside a stream capture.
    void* ptr = nullptr;
    STATIC_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_static_scratch));
    cudaDeviceProp prop;
    STATIC_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return static_fail(err, err_cap, GNC_ERR_STATE, "libgnc is built for sm_100a (B200); device is", prop.name);
    for (int v = 0; v < kNumVariants; v++) {
      STATIC_CUDA_TRY(cudaFuncSetAttribute(kVariants[v].fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kVariants[v].smem));
      int n = 0;
      STATIC_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kVariants[v].fn, kVariants[v].threads, kVariants[v].smem));
      if (n < 1) return static_fail(err, err_cap, GNC_ERR_STATE, "kernel variant does not fit an SM", kVariants[v].name);
      st.ctas_per_sm[v] = n;
    }
    st.sm_count = prop.multiProcessorCount;
    st.scratch = ptr;
    st.epoch = 0;
    st.ready = true;
  }
  if (in->n_nodes == 0) {
    STATIC_CUDA_TRY(cudaMemsetAsync(out->counts, 0, GNC_N_COUNTS * sizeof(int64_t), stream));
    return GNC_OK;
  }
  if (!in->rows || !in->row_ptr || !out->flags || !out->gpu_idx || !out->gpu_total || !out->gpu_vals)
    return static_fail(err, err_cap, GNC_ERR_ARG, "null input or output buffer", nullptr);
  if (!aligned16(in->rows) || !aligned16(in->row_ptr) || !aligned16(out->gpu_vals))
    return static_fail(err, err_cap, GNC_ERR_ARG, "rows, row_ptr and gpu_vals must be 16-byte aligned", nullptr);
  if (in->rows_bytes > 0xFFFFFFFFull - (1u << 20)) return static_fail(err, err_cap, GNC_ERR_ARG, "rows_bytes exceeds the PNL-v2 limit (4 GiB - 1 MiB)", nullptr);
  if ((uint64_t)in->node_base + in->n_nodes > 0x100000000ull) return static_fail(err, err_cap, GNC_ERR_ARG, "node_base + n_nodes exceeds 2^32", nullptr);
  if (out->defer_cap && !out->defer) return static_fail(err, err_cap, GNC_ERR_ARG, "result.defer is null but defer_cap > 0", nullptr);

  // same tile-shape choice as gnc_classify
  gnc_ctx probe;
  const int v = pick_variant(&probe, in->rows_bytes, in->n_nodes);
  const Variant& var = kVariants[v];
  const uint32_t tiles = tiles_for(in->n_nodes, var.tile_nodes);
  if (tiles > GNC_STATIC_MAX_TILES) {
    char msg[160];
    snprintf(msg, sizeof msg, "%u nodes at %d nodes per tile need %u tiles; the static scratch holds %d (use gnc_classify, or split the shard)",
             in->n_nodes, var.tile_nodes, tiles, GNC_STATIC_MAX_TILES);
    return static_fail(err, err_cap, GNC_ERR_ARG, "shard too large for gnc_classify_static", msg);
  }
  // The array still holds whatever the previous call left.  The header (ticket, finished-CTA count,
  // accumulators) must start at zero; the descriptors are told apart by the epoch tag instead, and
  // are wiped only when the 22-bit epoch wraps.
  STATIC_CUDA_TRY(cudaMemsetAsync(st.scratch, 0, sizeof(ScratchHeader), stream));
  st.epoch = (st.epoch + 1) & gnc::kEpochMask;
  if (st.epoch == 0) {
    STATIC_CUDA_TRY(cudaMemsetAsync((char*)st.scratch + sizeof(ScratchHeader), 0, kStaticScratchBytes - sizeof(ScratchHeader), stream));
    st.epoch = 1;
  }
  Params p = {};
  p.rows = in->rows;
  p.row_ptr = in->row_ptr;
  p.rows_bytes = in->rows_bytes;
  p.n_nodes = in->n_nodes;
  p.node_base = in->node_base;
  p.num_tiles = tiles;
  p.epoch = st.epoch;
  p.max_digits = max_str_digits;
  p.flags = out->flags;
  p.gpu_idx = out->gpu_idx;
  p.gpu_total = out->gpu_total;
  p.gpu_vals = out->gpu_vals;
  p.defer = out->defer;
  p.counts = out->counts;
  p.out_cap = out->out_cap;
  p.defer_cap = out->defer ? out->defer_cap : 0;
  p.scratch = (ScratchHeader*)st.scratch;
  p.inc = (unsigned long long*)((char*)st.scratch + sizeof(ScratchHeader));
  p.agg = (uint32_t*)(p.inc + GNC_STATIC_MAX_TILES);
  p.trace = nullptr;
  const int grid = (int)std::min<uint32_t>(tiles, (uint32_t)(st.sm_count * st.ctas_per_sm[v]));
  var.launch(p, grid, var.smem, stream);
  STATIC_CUDA_TRY(cudaGetLastError());
  return GNC_OK;
}


// =================================================================================================
// Raw /api/v1/nodes JSON -> PNL-v2 on the device (include/gnc.h: gnc_json_to_pnl).
// =================================================================================================
namespace {

template <class T, class Op>
int device_exclusive_scan(gnc_ctx* c, const T* in, unsigned long long n, T* out, T* block_totals, T* grand, Op op, T ident,
                          cudaStream_t stream) {
  if (n == 0) {
    if (grand) CUDA_TRY(c, cudaMemsetAsync(grand, 0, sizeof(T), stream));
    return GNC_OK;
  }
  const unsigned long long nb = (n + 1023) / 1024;
  gnc::js::scan_block_totals<T, Op><<<(unsigned)nb, 1024, 0, stream>>>(in, n, block_totals, op, ident);
  gnc::js::scan_totals_block<T, Op><<<1, 1024, 0, stream>>>(block_totals, nb, op, ident, grand);
  gnc::js::scan_apply<T, Op><<<(unsigned)nb, 1024, 0, stream>>>(in, n, block_totals, out, op, ident);
  CUDA_TRY(c, cudaGetLastError());
  c->launches += 3;
  return GNC_OK;
}

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace

extern "C" int gnc_json_to_pnl(gnc_ctx* c, const uint8_t* json, uint64_t json_bytes, const gnc_json_out* out, void* cuda_stream) {
  using namespace gnc::js;
  if (!c) return GNC_ERR_ARG;
  if (!json || !out || !out->rows || !out->row_ptr || !out->info) return fail(c, GNC_ERR_ARG, "null json or output buffer");
  if (!aligned16(out->rows) || !aligned16(out->row_ptr)) return fail(c, GNC_ERR_ARG, "rows and row_ptr must be 16-byte aligned");
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  CUDA_TRY(c, cudaSetDevice(c->device));
  const unsigned long long n = json_bytes, n_chunks = (n + kChunk - 1) / kChunk, nodes_cap = out->nodes_cap;
  if (n_chunks >= (1ull << 32)) return fail(c, GNC_ERR_ARG, "json document too large (1 TiB limit)");
  // carve the working arrays out of one scratch allocation
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
  const size_t nbc = (n_chunks + 1023) / 1024 + 1, nbn = (nodes_cap + 1023) / 1024 + 1;
  const size_t mask_bytes = (size_t)n_chunks * kChunk / 2;  // the structural index: 16 bytes per 32 document bytes
  const size_t n_tiles = (n_chunks + kTileChunks - 1) / kTileChunks, tiles_pad = align_up(n_tiles + 1, kLookWindow);
  const size_t look_bytes = tiles_pad * 8 + 256;  // lexer look-back states, then the tile ticket
  // per chunk (256 document bytes)
  const size_t o_work = take(sizeof(Work));
  const size_t o_masks = take(mask_bytes);                 // the index
  const size_t o_depth = take(n_chunks * 4);               // start depth relative to the tile
  const size_t o_minrel = take(n_chunks * 4);              // lower bound of the depth inside the chunk
  const size_t o_evc = take(n_chunks * 4), o_evp = take(n_chunks * 4), o_btu = take(nbc * 4);  // events: counts, prefix, block totals
  const size_t o_evpack = take(n_chunks * 4);              // places of a chunk's first four events
  const size_t o_clist = take(n_chunks * 4);               // chunks the events kernel walks
  // per tile (128 chunks)
  const size_t o_look = take(look_bytes);
  const size_t o_tilebal = take((n_tiles + 1) * 4), o_tilestart = take((n_tiles + 1) * 4), o_bt32 = take((n_tiles / 1024 + 2) * 4);
  // per node
  const size_t o_elem = take((nodes_cap + 1) * 8);         // where each element of items starts
  const size_t o_sizes = take((nodes_cap + 1) * 4), o_plain = take((nodes_cap + 1) * 8), o_offs = take((nodes_cap + 1) * 8);
  const size_t o_bt64 = take(nbn * 8), o_total = take(8);
  const size_t o_plans = take((nodes_cap + 1) * sizeof(NodePlan));
  const size_t o_desc = take((nodes_cap + 1) * (size_t)kMaxDesc * sizeof(uint2));  // the measuring pass's notes
  if (off > c->json_scratch_bytes) {
    if (c->json_scratch) {
      CUDA_TRY(c, cudaStreamSynchronize(stream));
      CUDA_TRY(c, cudaFree(c->json_scratch));
      c->json_scratch = nullptr;
      c->json_scratch_bytes = 0;
    }
    CUDA_TRY(c, cudaMalloc(&c->json_scratch, off + off / 8));
    c->json_scratch_bytes = off + off / 8;
  }
  char* base = (char*)c->json_scratch;
  Work* work = (Work*)(base + o_work);
  int* depth0 = (int*)(base + o_depth);
  unsigned int *evc = (unsigned int*)(base + o_evc), *evp = (unsigned int*)(base + o_evp), *btu = (unsigned int*)(base + o_btu);
  unsigned long long *elem = (unsigned long long*)(base + o_elem), *plain = (unsigned long long*)(base + o_plain),
                     *offs = (unsigned long long*)(base + o_offs), *bt64 = (unsigned long long*)(base + o_bt64),
                     *total = (unsigned long long*)(base + o_total);
  unsigned int* sizes = (unsigned int*)(base + o_sizes);
  NodePlan* plans = (NodePlan*)(base + o_plans);
  Masks mk;
  mk.w = (uint4*)(base + o_masks);
  int* min_rel = (int*)(base + o_minrel);
  unsigned int* evpack = (unsigned int*)(base + o_evpack);
  unsigned int* chunk_list = (unsigned int*)(base + o_clist);
  unsigned long long* look = (unsigned long long*)(base + o_look);
  unsigned long long* lex_state = look;
  unsigned int* ticket = (unsigned int*)(look + tiles_pad);
  int *tile_bal = (int*)(base + o_tilebal), *tile_start = (int*)(base + o_tilestart), *bt32 = (int*)(base + o_bt32);
  const Depths dp{depth0, tile_start};
  uint2* desc = (uint2*)(base + o_desc);
  if (!aligned16(json)) return fail(c, GNC_ERR_ARG, "json must be 16-byte aligned");
  CUDA_TRY(c, cudaMemsetAsync(work, 0, sizeof(Work), stream));
  int rc;
  const unsigned cb = (unsigned)((n_chunks + 255) / 256);
  if (n_chunks) {
    const unsigned tb = (unsigned)((n_chunks + kTileChunks - 1) / kTileChunks);
    constexpr size_t kTileSmem = (size_t)kTileChunks * kRowStride;
    // one pass: masks, per-chunk start depth (lexer state and depth travel from tile to tile by look-back)
    CUDA_TRY(c, cudaMemsetAsync(look, 0, look_bytes, stream));
    index_kernel<<<tb, kTileChunks, kTileSmem, stream>>>(json, n, mk, depth0, tile_bal, min_rel, lex_state, ticket, n_chunks);
    if ((rc = device_exclusive_scan<int, OpAddI>(c, tile_bal, n_tiles, tile_start, bt32, nullptr, OpAddI(), 0, stream))) return rc;  // tile balances -> tile start depths
    events_select_kernel<<<cb, 256, 0, stream>>>(dp, min_rel, evc, chunk_list, work, n_chunks);
    events_kernel<<<8 * c->sm_count, 256, 0, stream>>>(json, mk, dp, chunk_list, evc, evpack, work);
    if ((rc = device_exclusive_scan<unsigned int, OpAddU>(c, evc, n_chunks, evp, btu, nullptr, OpAddU(), 0u, stream))) return rc;
    pick_items_kernel<<<1, 1, 0, stream>>>(json, n, mk, dp, evp, work, (unsigned int)std::min<unsigned long long>(nodes_cap, 0xFFFFFFFFull));
    element_starts_kernel<<<cb, 256, 0, stream>>>(json, mk, dp, evc, evpack, evp, work, elem, n_chunks);
    c->launches += 6;
  }
  const unsigned nb = (unsigned)std::max<unsigned long long>(1, (nodes_cap + 127) / 128);
  if (nodes_cap) {
    CUDA_TRY(c, cudaMemsetAsync(sizes, 0, (nodes_cap + 1) * 4, stream));
    measure_rows_kernel<<<nb, 128, 0, stream>>>(json, n, mk, elem, work, sizes, plans, desc);
    strip_flag_kernel<<<(unsigned)((nodes_cap + 255) / 256), 256, 0, stream>>>(sizes, plain, (unsigned int)nodes_cap);
    c->launches += 2;
  }
  if ((rc = device_exclusive_scan<unsigned long long, OpAddULL>(c, plain, nodes_cap, offs, bt64, total, OpAddULL(), 0ull, stream))) return rc;
  CUDA_TRY(c, cudaFuncSetAttribute(write_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWriteSmem));
  write_rows_kernel<<<nb, kWriteThreads, kWriteSmem, stream>>>(json, n, mk, plans, desc, elem, work, sizes, offs, total, out->rows, out->rows_cap, out->row_ptr, (long long*)out->info);
  CUDA_TRY(c, cudaGetLastError());
  c->launches += 1;
  return GNC_OK;
}
