/* gnc.h -- C ABI of libgnc.so, the B200 (sm_100a) node-classification library.
 *
 * The reference (ahaljh/k8s-gpu-node-checker) exposes no FFI: its seam for this path is the Python
 * function list_gpu_nodes(api) (check-gpu-node.py:215-226).  The entry points below are what a
 * binding for that function needs: one call that classifies a packed node list on the device and
 * one that combines the ranks of a sharded fleet.  Each comment names the reference lines the
 * call replaces.  No exceptions cross the boundary; every call returns 0 or a negative gnc_status
 * and leaves the text in gnc_last_error().  A context belongs to one GPU and one caller thread at a
 * time; distinct contexts are independent.
 *
 * Input format: PNL-v2 (DESIGN.md "Data layout"; packer in k8s-gpu-node-checker_b200/pnl.py).
 */
#ifndef GNC_H_
#define GNC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GNC_ABI_VERSION 2

typedef struct gnc_ctx gnc_ctx;

typedef enum gnc_status {
  GNC_OK = 0,
  GNC_ERR_ARG = -1,      /* null / misaligned pointer, size out of range */
  GNC_ERR_CUDA = -2,     /* a CUDA runtime call failed; text has the CUDA error */
  GNC_ERR_NOMEM = -3,
  GNC_ERR_STATE = -4,    /* call not valid for this context (e.g. gather without a communicator) */
  GNC_ERR_COMM = -5,     /* NCCL failure */
  GNC_ERR_OVERFLOW = -6  /* an output table was too small; counts say by how much */
} gnc_status;

/* Per-node flag bits written to gnc_result.flags. */
#define GNC_F_GPU 0x01u       /* gpus > 0: the node is in gpu_nodes          (check-gpu-node.py:222) */
#define GNC_F_READY 0x02u     /* is_ready(node)                                (check-gpu-node.py:172-178) */
#define GNC_F_DEFERRED 0x04u  /* a kept value needs > 18 digits: see gnc_defer_rec; GNC_F_GPU not decided */
#define GNC_F_KEY0 0x08u      /* bits 3..6: key k of GPU_RESOURCE_KEYS is in gpu_breakdown (:186-195) */
#define GNC_F_MALFORMED 0x80u /* the row's own lengths did not fit the row; the rest of it was ignored */

/* Slots of the counts block (int64[GNC_N_COUNTS]). */
enum {
  GNC_C_GPU_NODES = 0,   /* len(gpu_nodes)        before deferred nodes are added (check-gpu-node.py:275) */
  GNC_C_READY_NODES = 1, /* len(ready_gpu_nodes)  likewise                         (check-gpu-node.py:276) */
  GNC_C_DEFERRED = 2,    /* nodes with GNC_F_DEFERRED */
  GNC_C_SCANNED = 3,     /* nodes classified */
  GNC_C_KEY_LO = 4,      /* [4..7]  per-key GPU totals over the GPU nodes, low 64 bits            */
  GNC_C_KEY_HI = 8,      /* [8..11] high 64 bits (two's-complement 128-bit; exact for any fleet)  */
  GNC_C_MALFORMED = 12,
  GNC_C_DEFER_DROPPED = 13, /* deferred nodes that did not fit gnc_result.defer_cap */
  GNC_C_OUT_DROPPED = 14,   /* GPU nodes that did not fit gnc_result.out_cap */
  GNC_N_COUNTS = 16
};

/* One packed shard, DEVICE pointers.  rows and row_ptr must be 16-byte aligned, rows readable for
 * rows_bytes + 32 bytes and row_ptr for n_nodes + 1 + 4 entries (bulk copies move 16-byte units). */
typedef struct gnc_batch {
  const uint8_t* rows;     /* node rows, back to back */
  const uint32_t* row_ptr; /* n_nodes + 1 byte offsets into rows */
  uint64_t rows_bytes;     /* row_ptr[n_nodes] */
  uint32_t n_nodes;
  uint32_t node_base;      /* fleet index of this shard's node 0; added to every index written */
} gnc_batch;

/* A node whose value texts the caller must add up with wider integers: off/len locate the value of
 * key k inside gnc_batch.rows (off == 0xFFFFFFFF: key not kept). Texts already passed the grammar. */
typedef struct gnc_defer_rec {
  uint32_t node;
  uint32_t off[4];
  uint32_t len[4];
} gnc_defer_rec;

/* Caller-allocated DEVICE outputs.  gpu_* hold one record per GPU node, in fleet order. */
typedef struct gnc_result {
  uint8_t* flags;        /* [n_nodes] */
  uint32_t* gpu_idx;     /* [out_cap]     node index (node_base + i) of each GPU node, ascending */
  int64_t* gpu_total;    /* [out_cap]     "gpus"           (check-gpu-node.py:201) */
  int64_t* gpu_vals;     /* [out_cap][4]  "gpu_breakdown" values in GPU_RESOURCE_KEYS order, 0 if absent; 16-byte aligned */
  gnc_defer_rec* defer;  /* [defer_cap]   unordered */
  int64_t* counts;       /* [GNC_N_COUNTS] */
  uint32_t out_cap;      /* n_nodes is always enough */
  uint32_t defer_cap;
} gnc_result;

/* ABI version of the loaded library (compare with GNC_ABI_VERSION). */
int gnc_abi_version(void);

/* Create the context for CUDA device `device` as rank `rank` of `world`.  For world > 1 pass the
 * ncclUniqueId bytes rank 0 obtained from gnc_comm_id(); for world == 1 pass NULL/0.
 * Stands where the reference builds its client (check-gpu-node.py:253): one per process. */
int gnc_create(gnc_ctx** out, int device, int rank, int world, const void* nccl_id, size_t nccl_id_len);

/* Fill `id` (at least 128 bytes) with a fresh ncclUniqueId; returns its length in *len. */
int gnc_comm_id(void* id, size_t cap, size_t* len);

/* Size the library's scratch for shards of up to max_nodes nodes now, so that gnc_classify never
 * allocates (it otherwise grows the scratch on demand, which synchronises the device). */
int gnc_reserve(gnc_ctx*, uint32_t max_nodes);

/* CPython's int() refuses more than this many digit characters (sys.get_int_max_str_digits(),
 * default 4300; 0 = no limit).  Default 4300. */
int gnc_set_max_str_digits(gnc_ctx*, int max_digits);

/* Classify one shard: the loop of list_gpu_nodes (check-gpu-node.py:220-225) with gpu_capacity
 * (:181-196), is_ready (:172-178) and the gpus sum (:201) per node, the ordered gpu_nodes filter
 * (:222-223) and the counts one_shot derives (:275-276).  Work is enqueued on `cuda_stream`
 * (a cudaStream_t, NULL = default stream); the call does not wait for it. */
int gnc_classify(gnc_ctx*, const gnc_batch* in, const gnc_result* out, void* cuda_stream);

/* Combine the ranks of a sharded fleet (no-op for world == 1): all-reduce of counts, then an
 * all-gather of every rank's gpu_idx/gpu_total/gpu_vals into `all` at rank-ordered offsets, which
 * reproduces the reference's single ordered list (check-gpu-node.py:220-225).  `local` is this
 * rank's gnc_classify result; `all` must hold the fleet-wide GPU-node count. */
int gnc_gather(gnc_ctx*, const gnc_result* local, const gnc_result* all, void* cuda_stream);

/* Multi-GPU only, optional, collective (every rank calls it once, after gnc_create): switch the
 * context to the fused gather.  Afterwards gnc_classify also stores, per tile of nodes, which of them are
 * GPU nodes (one bit each) and how many GPU nodes precede the tile, and at the end the counts block, straight
 * into every rank's buffer over NVLink while the kernel runs; gnc_gather (indices + counts form) only waits
 * for the other ranks' flags and rebuilds the ordered index list locally -- no collective follows the
 * kernel.  `fleet_nodes_cap` bounds the fleet's node count (a shard: at most its share, rounded up).  Ranks must call
 * gnc_classify and gnc_gather in step (the same number of times); up to 8 ranks on one NVLink domain.
 * Returns an error (and leaves the NCCL path in place) where peer mapping is not available. */
int gnc_fleet_enable(gnc_ctx*, uint32_t fleet_nodes_cap, void* cuda_stream);

/* Optional, after gnc_fleet_enable: name the gather target (`all` of gnc_gather; indices + counts form) BEFORE
 * the classify launches.  gnc_classify then also rebuilds the fleet-wide ordered list into it once its own
 * tiles are done and every rank's records have arrived -- classification and gather are ONE kernel launch --
 * and gnc_gather with that same target has nothing left to do.  Every rank must do the same (all name a target
 * or none does: the ranks acknowledge each other's buffers in step).  NULL clears the target.  The buffers
 * must stay valid until the launches that use them have finished. */
int gnc_fleet_target(gnc_ctx*, const gnc_result* all);

/* Diagnostics of the fused gather since the last call (synchronous; resets them), from in-kernel timestamps in
 * ns: out[0] time CTAs spent, after their tiles, waiting for every rank's flag, summed over CTAs and launches;
 * out[1] likewise the time they spent rebuilding the index list; out[2] the longest single rebuild; out[3] how
 * many CTAs were added up (divide the sums by it); out[4] the time CTAs spent in the launch before that (their
 * share of the classification), summed likewise; out[5..7] reserved. */
int gnc_fleet_timing(gnc_ctx*, uint64_t out[8]);

/* Same classification from HOST memory: copies the shard to the device in pipelined chunks,
 * classifies, and copies flags / GPU-node records / counts back to the caller's HOST buffers
 * (the gnc_batch / gnc_result pointers are host pointers here).  Blocks until the results are in
 * host memory.  This is the call a binding without device buffers of its own uses. */
int gnc_classify_host(gnc_ctx*, const gnc_batch* host_in, const gnc_result* host_out);

/* The same classification as gnc_classify for callers that cannot allocate: no context, no
 * cudaMalloc or cudaFree, no allocating library call, on any call.  (No counterpart in the
 * reference; it exists for hosts that run under a device-memory cap they have already spent, or that
 * cannot take the device-wide synchronisation and possible failure of an allocation on their path.)
 * The kernel's scratch is a fixed array inside libgnc.so, GNC_STATIC_SCRATCH_BYTES of device memory
 * that every GPU which loads the library holds until the process exits.  It serves shards of up to
 * GNC_STATIC_MAX_TILES tiles -- about 16.7M nodes of ordinary size (256 per tile), fewer when rows
 * are large (down to 32 per tile, 2M nodes); a larger shard is refused with GNC_ERR_ARG and a text
 * saying so.  Inputs and outputs are the caller's DEVICE buffers, exactly as for gnc_classify;
 * max_str_digits is sys.get_int_max_str_digits() (4300 by default, 0 = no limit).
 *  - The first call on each device looks up the array's device address and the launch geometry.  If the
 *    process has not used that device yet, this is also when the CUDA runtime creates its context and
 *    loads this library's module, i.e. when the array and the kernel images take up their device memory:
 *    make that call early, before the memory budget is spent, and outside any stream capture.
 *  - Host threads may call concurrently (the set-up and launch are serialised by a lock inside).
 *  - All calls on a device share the one array, which keeps the previous call's contents; each call
 *    clears what has to start from zero, on `cuda_stream`.  Two calls whose kernels overlap on the
 *    GPU (different streams, different threads) race on it: order them.
 * `err` (may be NULL) receives the error text, at most err_cap bytes. */
#define GNC_STATIC_MAX_TILES 65536
#define GNC_STATIC_SCRATCH_BYTES (256 + GNC_STATIC_MAX_TILES * 12)
int gnc_classify_static(int device, const gnc_batch* in, const gnc_result* out, int max_str_digits,
                        void* cuda_stream, char* err, size_t err_cap);

/* Raw /api/v1/nodes JSON bytes (DEVICE memory) -> PNL-v2 columns (DEVICE memory), ready for gnc_classify.
 * Stands where the reference lets the kubernetes client fetch and deserialise the NodeList
 * (api.list_node().items, check-gpu-node.py:217): only what the path reads is extracted -- each item's
 * status.capacity entries (:183-187) and the type/status of its status.conditions (:173-177) -- with
 * json.loads semantics (last duplicate key wins, \uXXXX escapes, surrogate pairs).  Non-string scalars
 * follow the client's str() conversion of declared-'str' fields: an integer token is its own text (zero,
 * "0" or "-0", is "0": kept, value 0), floats / true / false / containers never survive int(str(.)), null
 * stays None; a condition's `status: true` is "True".  Asynchronous on `cuda_stream`.
 * info[0] nodes, info[1] rows_bytes (= row_ptr[nodes]), info[2] status bits (1 no root "items" array,
 * 2 nodes_cap too small, 4 rows_cap too small, 8 unbalanced JSON, 16 too many root-level arrays),
 * info[3]/[4] byte positions of the items array's brackets, info[5] how many nodes needed the general
 * byte-serial parser (escapes, duplicate keys or non-string values in the parts read; 0 on ordinary API
 * output, which is handled from a bit-mask index of the document).  `json` must be 16-byte aligned. */
typedef struct gnc_json_out {
  uint8_t* rows;      /* [rows_cap], 16-byte aligned */
  uint32_t* row_ptr;  /* [nodes_cap + 5], 16-byte aligned */
  int64_t* info;      /* [8] */
  uint64_t rows_cap;
  uint32_t nodes_cap;
} gnc_json_out;
int gnc_json_to_pnl(gnc_ctx*, const uint8_t* json, uint64_t json_bytes, const gnc_json_out* out, void* cuda_stream);

/* Kernels this context has launched since it was created (for benchmark accounting). */
uint64_t gnc_launch_count(const gnc_ctx*);

/* Name of the kernel variant (nodes per tile, e.g. "t224") gnc_classify picks for a shard of this shape when
 * GNC_VARIANT is not set; profiles and benchmark lines name it so that a capture is matched to the code it measured. */
const char* gnc_variant_for(uint64_t rows_bytes, uint32_t n_nodes);

/* Unicode tables the quantity parser was built with (decimal-digit run starts, wide spaces). */
int gnc_unicode_tables(const uint32_t** nd_runs, int* n_runs, const uint32_t** spaces, int* n_spaces);

const char* gnc_last_error(const gnc_ctx*);
void gnc_destroy(gnc_ctx*);

#ifdef __cplusplus
}
#endif
#endif /* GNC_H_ */
