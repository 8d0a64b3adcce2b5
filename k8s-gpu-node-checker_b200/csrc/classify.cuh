// classify.cuh -- the node-classification kernel (sm_100a).
//
// Persistent CTAs (two per SM), three warp roles, no CTA-wide barrier in the loop:
//   producer  takes tile tickets in fleet order two tiles ahead, prefetches each tile into L2
//             (cp.async.bulk.prefetch.L2) and later stages it in shared memory with TMA bulk copies
//             (cp.async.bulk + mbarrier complete_tx);
//   walkers   one thread per node: scan the row's length table without a branch per entry, look at
//             the text of the few candidate entries, publish flags / ranks / partial counts, and
//             write the tile's GPU-node records a few iterations later from a private stash;
//   scan      two warps turn per-warp counts into output positions with a 512-tile-wide decoupled
//             look-back over global descriptors and fold the tile's counts into registers.
// GPU nodes come out in fleet order (ballot ranks + prefix, no atomics decide positions).  With the
// fused gather enabled each index is also stored into every other rank's buffer over NVLink.
//
// Reference behaviour implemented (file:line in /root/reference/check-gpu-node.py):
//   key match + quantity parse  :186-195      Ready test  :172-178      gpus sum / filter  :201, :222-225
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/gnc.h"
#include "unicode_tables.inc"

namespace gnc {

#ifdef GNC_TRACE
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define GNC_T(slot, it_, val) do { if (p.trace && (it_) < 64) p.trace[((size_t)blockIdx.x * 64 + (it_)) * 8 + (slot)] = (val); } while (0)
// position markers for debugging a hang or fault: one word per warp in a (host-mapped) buffer past the timestamps
#define GNC_D(code) do { if (p.trace && (threadIdx.x & 31) == 0) ((volatile unsigned long long*)p.trace)[160000 + blockIdx.x * 32 + (threadIdx.x >> 5)] = (unsigned long long)(code); } while (0)
#else
#define GNC_T(slot, it_, val) do { } while (0)
#define GNC_D(code) do { } while (0)
#endif
#ifdef GNC_TRACE
// per-tile progress: 1 ticket taken, 2 load issued, 3 walkers have it, 4 aggregate published (debugging)
#define GNC_TILE(tile_, st_) do { if (p.trace && (uint32_t)(tile_) < 60000u) ((volatile unsigned long long*)p.trace)[200000 + (tile_)] = ((unsigned long long)blockIdx.x << 32) | ((unsigned long long)(threadIdx.x >> 5) << 8) | (st_); } while (0)
#else
#define GNC_TILE(tile_, st_) do { } while (0)
#endif

// ------------------------------------------------------------------------------------------------
// Library-owned scratch (global memory beyond the caller's inputs and outputs).
// ------------------------------------------------------------------------------------------------
struct ScratchHeader {
  uint32_t ticket;       // next tile to hand out
  uint32_t done;         // CTAs that have finished
  uint32_t defer_count;  // deferred nodes seen
  uint32_t pad;
  unsigned long long accum[GNC_N_COUNTS];
  unsigned long long reserved[14];
};
static_assert(sizeof(ScratchHeader) == 256, "tile states start 256 bytes in");

// Look-back descriptors, one pair per tile, valid only when their tag equals the launch's epoch
// (1 .. 2^22-1; zeroed memory is never valid):
//   agg[t] u32: [31:10] epoch, [9:0]  GPU nodes in tile t
//   inc[t] u64: [63:32] epoch, [31:0] GPU nodes in tiles 0..t
constexpr uint32_t kEpochMask = (1u << 22) - 1;
constexpr int kLookbackSpan = 16;                   // tiles per lane per round
constexpr int kLookbackWindow = 32 * kLookbackSpan; // tiles per round

constexpr int kFleetMaxWorld = 8;
// A rank's slot in a fleet buffer: the counts block (16 x i64), then {tile nodes, tiles, node_base, nodes}, then
// one record per tile: the tile's exclusive GPU-node prefix within the shard and its walker warps' GPU-node
// ballots (tile_nodes / 32 words).  Which nodes are GPU nodes travels as bit masks, 1/20 of the bytes of the
// index list; every rank rebuilds the ordered indices from them locally.
constexpr uint32_t kFleetHeader = 2 * GNC_N_COUNTS + 8;
enum : int { FLEET_H_TILE_NODES = 2 * GNC_N_COUNTS, FLEET_H_TILES, FLEET_H_NODE_BASE, FLEET_H_NODES };
constexpr unsigned long long kFleetTimeoutNs = 2000000000ull;

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned long long ld_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

struct Params {
  const uint8_t* rows;
  const uint32_t* row_ptr;
  uint64_t rows_bytes;
  uint32_t rows_lo;  // offsets below this are not readable (host-buffer path: `rows` is a staging buffer minus its first offset)
  uint32_t n_nodes;
  uint32_t node_base;
  uint32_t num_tiles;
  uint32_t epoch;
  int max_digits;
  uint8_t* flags;
  uint32_t* gpu_idx;
  int64_t* gpu_total;
  int64_t* gpu_vals;
  gnc_defer_rec* defer;
  int64_t* counts;
  uint32_t out_cap;
  uint32_t defer_cap;
  ScratchHeader* scratch;
  unsigned long long* inc;  // [tiles rounded up to the look-back window]
  uint32_t* agg;            // likewise
  unsigned long long* trace;  // debug builds only (GNC_TRACE): [cta][iteration][8] timestamps
  // Fused gather (multi-GPU, gnc_fleet_enable): every tile's GPU-node ballots and prefix are also stored
  // straight into every rank's fleet buffer over NVLink while the kernel runs, so that no collective follows it.
  uint32_t fleet_world;                      // 0: off
  uint32_t fleet_rank;
  uint32_t fleet_cap;                        // record words per slot
  uint32_t fleet_slot_u32;                   // u32 per slot: kFleetHeader + fleet_cap
  unsigned long long fleet_epoch;            // this launch's number
  unsigned long long fleet_ack_need;         // every peer must have consumed this epoch before its slots are rewritten
  uint32_t* fleet_peer[kFleetMaxWorld];      // peer-mapped slots [world][fleet_slot_u32] of this epoch's parity, per rank
  unsigned long long* fleet_flag[kFleetMaxWorld];  // peer-mapped arrival flags [world] of this parity, per rank
  const unsigned long long* fleet_ack;       // local: [world] epochs each peer has consumed
  // Gather target known before the launch (gnc_fleet_target): the kernel goes on to rebuild the fleet-wide
  // index list itself once its tiles are done -- classification and gather are then ONE launch.
  uint32_t* fleet_out_idx;                   // null: the rebuild is left to gnc_gather
  int64_t* fleet_out_counts;
  uint32_t fleet_out_cap;
  unsigned int* fleet_unpack_done;           // local: CTAs that have finished reading the slots
  unsigned long long* fleet_ack_peer[kFleetMaxWorld];  // every rank's acknowledgement array (peer-mapped)
};

// ------------------------------------------------------------------------------------------------
// Rebuild of the fleet-wide ordered index list from the tile records all ranks stored here (fused gather).
// Called by every thread of every CTA of a grid: by the classify kernel's own CTAs once their tiles are done,
// or by fleet_unpack_kernel when the gather target was not known at launch.
//   1. wait until every rank's slot of this epoch has arrived in the LOCAL fleet buffer (flags);
//   2. a half-warp takes a group of whole tiles of one rank, up to 16 ballot words: one word per lane, each
//      word's place from its tile's prefix and a segmented scan of the popcounts; the group's indices are one
//      contiguous piece of the output, assembled in shared memory and stored coalesced; the next group's words
//      are requested before this one's are expanded;
//   3. the counts blocks are added up (one thread per slot), and the last CTA tells every rank that its slot
//      here was consumed.
// `buf` is 2 * kUnpackHalfWords words of shared memory per warp (16-byte aligned), `sh` 64 words and `lut` 256
// words per CTA.
// ------------------------------------------------------------------------------------------------
// shared-memory words a half-warp's buffer takes: up to 512 indices + 3 of alignment shift, then a spare word for the
// bits that are clear; a multiple of 4 so that every half's buffer is 16-byte aligned
constexpr uint32_t kUnpackSpare = 516, kUnpackHalfWords = 520;

struct FleetUnpack {
  const uint32_t* slots;             // [world][slot_u32] this epoch's parity of the local fleet buffer
  const unsigned long long* flags;   // [world] arrival flags of this parity
  unsigned long long epoch;
  int world, rank;
  uint32_t slot_u32, cap;
  uint32_t* out_idx;
  uint32_t out_cap;
  int64_t* out_counts;
  unsigned int* done;
  unsigned long long* const* ack_peer;  // [world]
  unsigned long long* timing;           // [8] ns: sum of flag waits, sum of rebuild times, longest rebuild; CTAs added up; sum of t_before
  unsigned long long t_before;          // ns this CTA spent in the launch before it came here (its share of the classification)
};

// One group of the rebuild, by a half-warp (see fleet_unpack): a lane's ballot word `m` of a tile with prefix `pf`,
// whose first node is node0; `base` = GPU nodes of the ranks before this one.  Both halves of the warp call together.
__device__ __noinline__ void unpack_group(uint32_t* __restrict__ out_idx, uint32_t out_cap, uint32_t m, uint32_t pf, uint32_t node0, bool have,
                                          unsigned long long base, uint32_t* mybuf, const uint32_t* __restrict__ lut, int hl) {
  const uint32_t cnt = have ? __popc(m) : 0u;
  uint32_t incl = cnt;
#pragma unroll
  for (int d = 1; d < 16; d <<= 1) {
    const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d, 16);
    if (hl >= d) incl += t;
  }
  const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 15, 16);
  // tiles of a group follow one another in the output: the group's piece starts at its first tile's prefix
  const uint32_t pf0 = __shfl_sync(0xFFFFFFFFu, pf, 0, 16);
  const unsigned long long o0 = base + pf0;
  // The piece is laid out in the buffer as it will lie in memory modulo 16 bytes (buffer word i <-> output word
  // o0 - shift + i), so that its aligned middle goes out as 16-byte stores.  The lane's word is expanded bit by
  // bit with no branch and no loop-carried dependence: bit b goes to (bits below b that are set), a clear bit
  // to a spare word.
  const uint32_t shift = (uint32_t)((reinterpret_cast<uintptr_t>(out_idx + o0) >> 2) & 3u);  // (whatever the list's own alignment)
  // (population counts per bit would do, but POPC issues at a quarter of the rate of the integer pipe and 31 of
  // them per word were most of the rebuild's time: the table gives a byte's eight counts in one shared-memory load)
  const uint32_t mm = have ? m : 0u;
  const uint32_t l0 = lut[mm & 0xFFu], l1 = lut[(mm >> 8) & 0xFFu], l2 = lut[(mm >> 16) & 0xFFu], l3 = lut[mm >> 24];
  const uint32_t at0 = shift + incl - cnt, at1 = at0 + (l0 >> 24), at2 = at1 + (l1 >> 24), at3 = at2 + (l2 >> 24);
#pragma unroll
  for (int b = 0; b < 32; b++) {
    const uint32_t l = b < 8 ? l0 : b < 16 ? l1 : b < 24 ? l2 : l3, a = b < 8 ? at0 : b < 16 ? at1 : b < 24 ? at2 : at3;
    const bool on = (mm >> b) & 1u;
    mybuf[on ? a + ((l >> (3 * (b & 7))) & 7u) : kUnpackSpare] = node0 + b;
  }
  __syncwarp();
  const uint32_t end = shift + total;  // buffer words [shift, end) are the piece
  if (o0 + total <= out_cap) {
    uint4* dst4 = reinterpret_cast<uint4*>(out_idx + (o0 - shift));
    const uint4* src4 = reinterpret_cast<const uint4*>(mybuf);
    for (uint32_t v = hl; 4u * v < end; v += 16) {
      if (4u * v >= shift && 4u * v + 4u <= end) {
        dst4[v] = src4[v];
      } else {
        for (uint32_t k = max(4u * v, shift); k < min(4u * v + 4u, end); k++) out_idx[o0 - shift + k] = mybuf[k];
      }
    }
  } else {
    for (uint32_t k = shift + hl; k < end; k += 16)
      if (o0 - shift + k < out_cap) out_idx[o0 - shift + k] = mybuf[k];
  }
  __syncwarp();
}

__device__ __forceinline__ void fleet_unpack(const FleetUnpack& u, uint32_t* buf, unsigned long long* sh, uint32_t* lut) {
  unsigned long long* off = sh;                         // [world + 1] GPU nodes before each rank
  unsigned long long* goff = sh + kFleetMaxWorld + 1;   // [world + 1] groups before each rank
  int* timed_out = reinterpret_cast<int*>(sh + 2 * (kFleetMaxWorld + 1));
  uint32_t* hdr = reinterpret_cast<uint32_t*>(sh + 2 * (kFleetMaxWorld + 1) + 1);  // [world][4]: tile nodes, tiles, node base, tiles per group
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, n_warps_cta = blockDim.x >> 5;
  const unsigned long long t_enter = global_ns();
  if (tid == 0) *timed_out = 0;
  for (int v = tid; v < 256; v += blockDim.x) {  // byte -> its eight "set bits below bit k" counts, 3 bits each, and its popcount << 24
    uint32_t e = (uint32_t)__popc(v) << 24;
    for (int k = 1; k < 8; k++) e |= (uint32_t)__popc(v & ((1 << k) - 1)) << (3 * k);
    lut[v] = e;
  }
  __syncthreads();
  if (tid < u.world) {
    const unsigned long long t0 = global_ns();
    while (ld_sys_u64(u.flags + tid) < u.epoch)
      if (global_ns() - t0 > kFleetTimeoutNs) {
        *timed_out = 1;
        break;
      }
  }
  __syncthreads();
  __threadfence_system();
  const unsigned long long t_flags = global_ns();
  // A group is up to 16 ballot words of whole tiles of one rank, handled by a half-warp (a lane per word): finer
  // than a warp's 32 words, so that the ~10^4 groups of a 10M-node fleet spread evenly over the grid's warps.
  auto tiles_per_group = [](uint32_t wpt) { return wpt >= 16u ? 1u : 16u / max(wpt, 1u); };
  if (tid < u.world) {  // every rank's header words, read once
    const uint32_t* h = u.slots + (size_t)tid * u.slot_u32;
    const uint32_t tn = __ldcv(h + FLEET_H_TILE_NODES);
    hdr[4 * tid + 0] = tn;
    hdr[4 * tid + 1] = __ldcv(h + FLEET_H_TILES);
    hdr[4 * tid + 2] = __ldcv(h + FLEET_H_NODE_BASE);
    hdr[4 * tid + 3] = tiles_per_group(tn / 32);
    off[tid + 1] = (unsigned long long)__ldcv(reinterpret_cast<const long long*>(h) + GNC_C_GPU_NODES);
    if (tn / 32 > 16u) *timed_out = 1;  // tiles wider than a half-warp's 16 words: no kernel variant writes them
  }
  __syncthreads();
  if (tid == 0) {
    off[0] = goff[0] = 0;
    for (int r = 0; r < u.world; r++) {
      off[r + 1] += off[r];
      goff[r + 1] = goff[r] + (hdr[4 * r + 1] + hdr[4 * r + 3] - 1) / hdr[4 * r + 3];
    }
  }
  __syncthreads();
  const unsigned long long t_hdr = global_ns();
  const int hl = lane & 15, half = lane >> 4;                  // lane within the half-warp, which half
  uint32_t* mybuf = buf + (size_t)warp * (2 * kUnpackHalfWords) + half * kUnpackHalfWords;
  const unsigned long long n_halves = 2ull * gridDim.x * n_warps_cta, n_groups = goff[u.world];
  // what a lane needs for group `grp`: its ballot word and its tile's prefix
  struct Piece {
    uint32_t m, pf, node0;  // ballot word, tile prefix, first node of the word
    int r;
    bool have;
  };
  auto fetch = [&](unsigned long long grp) {
    Piece pc;
    pc.m = pc.pf = pc.node0 = 0;
    pc.r = 0;
    pc.have = false;
    if (grp >= n_groups) return pc;
    int r = 0;
    while (grp >= goff[r + 1]) r++;
    const uint32_t* h = u.slots + (size_t)r * u.slot_u32;
    const uint32_t tn = hdr[4 * r], wpt = tn / 32, tiles = hdr[4 * r + 1], tpg = hdr[4 * r + 3];
    const uint32_t tl = wpt ? hl / wpt : 0u, w = wpt ? hl % wpt : 0u;
    const unsigned long long tile = (grp - goff[r]) * tpg + tl;
    pc.r = r;
    pc.have = wpt != 0u && wpt <= 16u && tl < tpg && tile < tiles && (tile + 1) * (wpt + 1) <= u.cap;
    if (pc.have) {
      const uint32_t* rec = h + kFleetHeader + (size_t)tile * (wpt + 1);
      pc.m = __ldcv(rec + 1 + w);
      pc.pf = __ldcv(rec);
      pc.node0 = hdr[4 * r + 2] + (uint32_t)tile * tn + 32u * w;
    }
    return pc;
  };
  // Groups are taken kBatch at a time: their words are all requested before the first is expanded, so a half-warp
  // pays the trip to L2 once per batch, not once per group (the expansion itself is a fraction of that trip).
  constexpr int kBatch = 4;
  unsigned long long grp = 2ull * ((unsigned long long)blockIdx.x * n_warps_cta + warp) + half;
  while (__any_sync(0xFFFFFFFFu, grp < n_groups)) {  // the two halves of a warp run their groups side by side
    Piece pcs[kBatch];
#pragma unroll
    for (int i = 0; i < kBatch; i++) pcs[i] = fetch(grp + (unsigned long long)i * n_halves);
#pragma unroll
    for (int i = 0; i < kBatch; i++)  // (one copy of the code, called: unrolled four times it overflows the instruction cache)
      unpack_group(u.out_idx, u.out_cap, pcs[i].m, pcs[i].pf, pcs[i].node0, pcs[i].have, off[pcs[i].r], mybuf, lut, hl);
    grp += (unsigned long long)kBatch * n_halves;
  }
  const unsigned long long t_loop = global_ns();
  if (blockIdx.x == 0 && tid < GNC_N_COUNTS) {  // the all-reduce of the counts: exact sums, 128-bit for the key totals
    const int sl = tid;
    long long sum = 0;
    if (sl >= GNC_C_KEY_LO && sl < GNC_C_KEY_HI + 4) {
      const int k = (sl - GNC_C_KEY_LO) & 3;
      unsigned __int128 key = 0;
      for (int r = 0; r < u.world; r++) {
        const long long* cr = reinterpret_cast<const long long*>(u.slots + (size_t)r * u.slot_u32);
        key += ((unsigned __int128)(unsigned long long)__ldcv(cr + GNC_C_KEY_HI + k) << 64) | (unsigned long long)__ldcv(cr + GNC_C_KEY_LO + k);
      }
      sum = sl < GNC_C_KEY_HI ? (long long)(unsigned long long)key : (long long)(unsigned long long)(key >> 64);
    } else {
      for (int r = 0; r < u.world; r++) sum += __ldcv(reinterpret_cast<const long long*>(u.slots + (size_t)r * u.slot_u32) + sl);
      if (sl == GNC_C_OUT_DROPPED) {
        if (off[u.world] > u.out_cap) sum += (long long)(off[u.world] - u.out_cap);
        if (*timed_out) sum += 1ll << 41;
      }
    }
    u.out_counts[sl] = sum;
  }
  __syncthreads();
  if (tid == 0) {
    const unsigned long long t_done = global_ns();
    if (u.timing) {  // sums over CTAs and launches (the host divides), and how many CTAs were added up
      atomicAdd(u.timing + 0, t_flags - t_enter);
      atomicAdd(u.timing + 1, t_done - t_flags);
      atomicMax(u.timing + 2, t_done - t_flags);
      atomicAdd(u.timing + 3, 1ull);
      atomicAdd(u.timing + 4, u.t_before);
      atomicAdd(u.timing + 5, t_hdr - t_flags);
      atomicAdd(u.timing + 6, t_loop - t_hdr);   // (this thread's warp; the CTA's slowest warp shows in the total)
    }
    __threadfence();
    if (atomicAdd(u.done, 1u) == gridDim.x - 1) {  // every CTA has read the slots: peers may overwrite them
      *u.done = 0;
      __threadfence_system();
      for (int q = 0; q < u.world; q++) st_sys_u64(u.ack_peer[q] + u.rank, u.epoch);
    }
  }
}

__constant__ uint32_t c_nd_runs[GNC_ND_RUNS] = {GNC_ND_RUN_LIST};
__constant__ uint32_t c_wide_spaces[GNC_WIDE_SPACES] = {GNC_WIDE_SPACE_LIST};

// ------------------------------------------------------------------------------------------------
// mbarrier / TMA bulk-copy wrappers (PTX ISA: mbarrier, cp.async.bulk).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t addr = smem_addr(bar), ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
  } while (!ok);
}
// global -> shared bulk copy; src, dst and bytes are multiples of 16.
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_addr(dst)),
               "l"(src), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}
// ask for [src, src + bytes) to be brought into L2 (no destination, no completion to wait for)
__device__ __forceinline__ void bulk_prefetch_l2(const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int threads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 ld_volatile_v4(const uint32_t* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

// ------------------------------------------------------------------------------------------------
// Byte access.  Rows sit at arbitrary byte offsets; words are fetched as two aligned 32-bit loads and
// funnel-shifted (the buffers are padded so the second word is always readable).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld32u(const uint8_t* buf, uint32_t off) {
  const uint32_t* w = reinterpret_cast<const uint32_t*>(buf) + (off >> 2);
  return __funnelshift_r(w[0], w[1], (off & 3u) << 3);
}

__host__ __device__ constexpr uint32_t pack4(const char* s, int i, int n) {
  return (i < n ? (uint32_t)(uint8_t)s[i] : 0u) | (i + 1 < n ? (uint32_t)(uint8_t)s[i + 1] << 8 : 0u) |
         (i + 2 < n ? (uint32_t)(uint8_t)s[i + 2] << 16 : 0u) | (i + 3 < n ? (uint32_t)(uint8_t)s[i + 3] << 24 : 0u);
}

// ------------------------------------------------------------------------------------------------
// int(str) -- CPython's base-10 grammar over UTF-8 (see oracle/cgn_oracle.py:int_grammar for the
// spelled-out rules).  Generic-pointer, out of line: only values that are not a single ASCII digit
// come here.
// ------------------------------------------------------------------------------------------------
enum { Q_ABSENT = 0, Q_OK = 1, Q_BIG = 2 };
enum { C_WS = 10, C_US = 11, C_PLUS = 12, C_MINUS = 13, C_OTHER = 14 };

__device__ __forceinline__ int classify_cp(uint32_t cp) {
  if (cp < 127) {
    if (cp - '0' <= 9u) return (int)(cp - '0');
    if (cp - 9u <= 4u || cp == 32) return C_WS;
    if (cp == '_') return C_US;
    if (cp == '+') return C_PLUS;
    if (cp == '-') return C_MINUS;
    return C_OTHER;
  }
  for (int i = 0; i < GNC_WIDE_SPACES; i++)
    if (c_wide_spaces[i] == cp) return C_WS;
  int lo = 0, hi = GNC_ND_RUNS;  // last run starting at or below cp
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (c_nd_runs[mid] <= cp) lo = mid; else hi = mid;
  }
  uint32_t d = cp - c_nd_runs[lo];
  return d <= 9u ? (int)d : C_OTHER;
}

// Returns (value << 2) | status: nothing is passed by address, so callers keep their state in registers.
__device__ __noinline__ long long parse_quantity_general(const uint8_t* s, uint32_t n, int max_digits) {
  uint32_t i = 0, sig = 0;
  int phase = 0;  // 0 leading ws, 1 after sign, 2 in digits, 3 after '_', 4 trailing ws
  bool neg = false;
  unsigned long long v = 0, digits = 0;
  while (i < n) {
    uint32_t b0 = s[i], cp = b0, need = 0;
    if (b0 >= 0x80) {  // generalised UTF-8: surrogates pass, overlong / truncated / > 0x10FFFF do not
      if (b0 < 0xC2 || b0 > 0xF4) return Q_ABSENT;
      need = b0 < 0xE0 ? 1 : (b0 < 0xF0 ? 2 : 3);
      if (n - i <= need) return Q_ABSENT;
      cp = b0 & (0x3Fu >> need);
      for (uint32_t j = 1; j <= need; j++) {
        uint32_t b = s[i + j];
        if ((b & 0xC0u) != 0x80u) return Q_ABSENT;
        cp = (cp << 6) | (b & 0x3Fu);
      }
      if ((need == 2 && cp < 0x800u) || (need == 3 && (cp < 0x10000u || cp > 0x10FFFFu))) return Q_ABSENT;
    }
    i += 1 + need;
    int c = classify_cp(cp);
    if (c == C_OTHER) return Q_ABSENT;
    if (c <= 9) {
      if (phase == 4) return Q_ABSENT;
      phase = 2;
      digits++;
      if (sig | (uint32_t)c) {
        sig++;
        if (sig <= 18) v = v * 10 + (unsigned)c;
      }
    } else if (c == C_WS) {
      if (phase == 0 || phase == 4) continue;
      if (phase != 2) return Q_ABSENT;
      phase = 4;
    } else if (c == C_US) {
      if (phase != 2) return Q_ABSENT;
      phase = 3;
    } else {
      if (phase != 0) return Q_ABSENT;
      neg = (c == C_MINUS);
      phase = 1;
    }
  }
  if (phase != 2 && phase != 4) return Q_ABSENT;
  if (max_digits > 0 && digits > (unsigned long long)max_digits) return Q_ABSENT;
  if (sig > 18) return Q_BIG;
  return ((neg ? -(long long)v : (long long)v) << 2) | Q_OK;  // |v| < 10^18 < 2^60
}

// ------------------------------------------------------------------------------------------------
// Per-node result of a walk.  Values travel as four int16 (a quantity is almost always one digit);
// a node with a value outside int16 is "wide": its exact values are read again, out of line, where
// they are needed.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t R_PRESENT = 0xFu;         // bits 0..3: key k kept in gpu_breakdown
constexpr uint32_t R_BIG_SHIFT = 4;          // bits 4..7: key k kept but wider than 18 digits
constexpr uint32_t R_WIDE = 1u << 8;
constexpr uint32_t R_MALFORMED = 1u << 9;
constexpr uint32_t R_READY = 1u << 10;

struct NodeResult {
  uint32_t v01, v23;  // int16 x 4
  uint32_t bits;
};

__device__ __forceinline__ void set_value(NodeResult& r, int k, int st, long long val) {
  const uint32_t bit = 1u << k;
  r.bits = (r.bits & ~(bit | (bit << R_BIG_SHIFT))) | (st != Q_ABSENT ? bit : 0u) | (st == Q_BIG ? bit << R_BIG_SHIFT : 0u);
  if (st != Q_OK) val = 0;
  if ((unsigned long long)(val + 32768) >= 65536ull) {
    r.bits |= R_WIDE;
    val = 0;
  }
  const uint32_t sh = (k & 1) << 4, m = 0xFFFFu << sh, x = ((uint32_t)val & 0xFFFFu) << sh;
  if (k & 2) r.v23 = (r.v23 & ~m) | x;
  else r.v01 = (r.v01 & ~m) | x;
}

// Slow, general, out-of-line walk of one row from a generic pointer, byte by byte: both row forms,
// any alignment, exact 64-bit values, value-text positions.  Extended rows, rows of tiles too large
// to stage, wide and deferred nodes come here; the hot path below never does.
struct SlowRow {
  long long v[4];
  uint32_t off[4], len[4];
  uint32_t bits;  // R_PRESENT / big / R_MALFORMED / R_READY (never R_WIDE)
};

__constant__ char c_key_text[4][20] = {"nvidia.com/gpu", "amd.com/gpu", "gpu.intel.com/i915", "intel.com/gpu"};
__constant__ uint32_t c_key_len[4] = {14, 11, 18, 13};

__device__ __forceinline__ uint32_t ldb32(const uint8_t* p) {
  return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}

__device__ __noinline__ void slow_row(const uint8_t* buf, uint32_t s, uint32_t e, int max_digits, SlowRow* out) {
  for (int k = 0; k < 4; k++) {
    out->v[k] = 0;
    out->off[k] = 0xFFFFFFFFu;
    out->len[k] = 0;
  }
  out->bits = 0;
  if (e - s < 2) {
    if (e != s) out->bits = R_MALFORMED;
    return;
  }
  const bool ext = buf[s] == 0xFF;
  unsigned long long n_cap, n_cond, tab, stride;
  if (ext) {
    if (e - s < 12) {
      out->bits = R_MALFORMED;
      return;
    }
    n_cap = ldb32(buf + s + 4);
    n_cond = ldb32(buf + s + 8);
    tab = (unsigned long long)s + 12;
    stride = 8;
  } else {
    n_cap = buf[s];
    n_cond = buf[s + 1];
    tab = (unsigned long long)s + 2;
    stride = 2;
  }
  const unsigned long long n = n_cap + n_cond;
  unsigned long long off = tab + stride * n;
  if (off > e) {
    out->bits = R_MALFORMED;
    return;
  }
  uint32_t bits = 0;
  for (unsigned long long i = 0; i < n; i++) {
    const unsigned long long kl = ext ? ldb32(buf + tab + 8 * i) : buf[tab + 2 * i];
    const unsigned long long vl = ext ? ldb32(buf + tab + 8 * i + 4) : buf[tab + 2 * i + 1];
    if (off + kl + vl > e) {
      bits |= R_MALFORMED;
      break;
    }
    const uint8_t* key = buf + off;
    const uint8_t* val = key + kl;
    if (i < n_cap) {
      for (int k = 0; k < 4; k++) {
        if (kl != c_key_len[k]) continue;
        bool eq = true;
        for (uint32_t q = 0; q < c_key_len[k]; q++) eq = eq && key[q] == (uint8_t)c_key_text[k][q];
        if (!eq) continue;
        // a repeated key overwrites: dict / json.loads last-wins
        const long long pq = vl ? parse_quantity_general(val, (uint32_t)vl, max_digits) : (long long)Q_ABSENT;
        const int st = (int)(pq & 3);
        const long long x = pq >> 2;
        const uint32_t bit = 1u << k;
        bits = (bits & ~(bit | (bit << R_BIG_SHIFT))) | (st != Q_ABSENT ? bit : 0u) | (st == Q_BIG ? bit << R_BIG_SHIFT : 0u);
        out->v[k] = st == Q_OK ? x : 0;
        out->off[k] = st == Q_ABSENT ? 0xFFFFFFFFu : (uint32_t)(off + kl);
        out->len[k] = st == Q_ABSENT ? 0u : (uint32_t)vl;
      }
    } else if (kl == 5 && vl == 4 && key[0] == 'R' && key[1] == 'e' && key[2] == 'a' && key[3] == 'd' && key[4] == 'y' &&
               val[0] == 'T' && val[1] == 'r' && val[2] == 'u' && val[3] == 'e') {
      bits |= R_READY;  // first hit returns (check-gpu-node.py:176-177)
      break;
    }
    off += kl + vl;
  }
  out->bits = bits;
}

__device__ __noinline__ uint3 slow_row_packed(const uint8_t* buf, uint32_t s, uint32_t e, int max_digits) {
  SlowRow sr;
  slow_row(buf, s, e, max_digits, &sr);
  uint3 r = make_uint3(0u, 0u, sr.bits);  // v01, v23, bits
  for (int k = 0; k < 4; k++) {
    const bool fits = (unsigned long long)(sr.v[k] + 32768) < 65536ull;
    if (!fits) r.z |= R_WIDE;
    const uint32_t x = fits ? ((uint32_t)sr.v[k] & 0xFFFFu) : 0u;
    if (k & 2) r.y |= x << ((k & 1) << 4);
    else r.x |= x << ((k & 1) << 4);
  }
  return r;
}

__device__ __noinline__ void emit_defer(const Params& p, uint32_t node, uint32_t s, uint32_t e, uint32_t slot) {
  SlowRow sr;
  slow_row(p.rows, s, e, p.max_digits, &sr);
  gnc_defer_rec rec;
  rec.node = p.node_base + node;
  for (int k = 0; k < 4; k++) {
    rec.off[k] = sr.off[k];
    rec.len[k] = sr.len[k];
  }
  p.defer[slot] = rec;
}

// A real key (length and first four bytes match): compare the rest of the key against per-key
// constants picked by select -- no divergence between lanes holding different keys -- and parse the
// value.  `rec` is (row-relative text offset << 16) | vlen << 8 | klen.
__device__ __forceinline__ void resolve_key(const uint8_t* buf, uint32_t s, uint32_t rec, int max_digits, NodeResult& r) {
  const uint32_t kl = rec & 0xFFu, vl = (rec >> 8) & 0xFFu, kp = s + (rec >> 16);
  const int k = kl == 14u ? 0 : kl == 11u ? 1 : kl == 18u ? 2 : 3;
  const uint32_t e1 = k == 0 ? pack4("ia.c", 0, 4) : k == 1 ? pack4("com/", 0, 4) : k == 2 ? pack4("inte", 0, 4) : pack4("l.co", 0, 4);
  const uint32_t e2 = k == 0 ? pack4("om/g", 0, 4) : k == 1 ? pack4("gpu", 0, 3) : k == 2 ? pack4("l.co", 0, 4) : pack4("m/gp", 0, 4);
  const uint32_t m2 = k == 1 ? 0xFFFFFFu : 0xFFFFFFFFu;
  const uint32_t e3 = k == 0 ? pack4("pu", 0, 2) : k == 1 ? 0u : k == 2 ? pack4("m/i9", 0, 4) : pack4("u", 0, 1);
  const uint32_t m3 = k == 0 ? 0xFFFFu : k == 1 ? 0u : k == 2 ? 0xFFFFFFFFu : 0xFFu;
  const uint32_t m4 = k == 2 ? 0xFFFFu : 0u;
  const bool eq = ld32u(buf, kp + 4) == e1 && (ld32u(buf, kp + 8) & m2) == e2 && (ld32u(buf, kp + 12) & m3) == e3 &&
                  (ld32u(buf, kp + 16) & m4) == (pack4("15", 0, 2) & m4);
  if (!eq) return;
  const uint32_t vp = kp + kl;
  long long pq = Q_ABSENT;
  if (vl == 1) {
    const uint32_t d = (uint32_t)buf[vp] - '0';
    pq = d <= 9u ? (long long)((d << 2) | Q_OK) : parse_quantity_general(buf + vp, 1, max_digits);
  } else if (vl != 0) {
    pq = parse_quantity_general(buf + vp, vl, max_digits);
  }
  set_value(r, k, (int)(pq & 3), pq >> 2);  // a repeated key overwrites: dict / json.loads last-wins
}

// The hot walk: one node's row out of shared memory.
//  1. scan the length table, two entries per 32-bit word, with no branch per entry: running text
//     offset by integer adds; entries that could matter (a capacity key of length 11/13/14/18, a
//     condition of lengths 5 and 4) are pushed into small shift registers;
//  2. look at the candidates' text, oldest first.  Every lane has about the same number of
//     candidates, so the warp stays converged here whatever position the keys had in the rows.
// Returns false when the row needs the general walk (extended form, misaligned or >= 64 KB row, more
// candidates than the registers hold, texts that run past the row).
__device__ __forceinline__ bool walk_row(const uint8_t* buf, uint32_t s, uint32_t e, int max_digits, NodeResult& r) {
  r.v01 = r.v23 = 0;
  r.bits = 0;
  if (e - s < 2) {
    if (e != s) r.bits = R_MALFORMED;
    return true;
  }
  if ((s & 3u) || e - s >= 65536u) return false;
  const uint16_t* tab = reinterpret_cast<const uint16_t*>(buf + s);  // tab[0] header, tab[1 + i] = (klen, vlen) of entry i
  const uint32_t hdr = tab[0];
  const uint32_t n_cap = hdr & 0xFFu, n_cond = hdr >> 8;
  if (n_cap == 0xFFu) return false;
  const uint32_t n = n_cap + n_cond;
  uint32_t rel = 2u + 2u * n;  // row-relative offset of the text of the next entry
  if (s + rel > e) {
    r.bits = R_MALFORMED;
    return true;
  }
  uint32_t k0 = 0, k1 = 0, k2 = 0, k3 = 0, nk = 0;  // capacity candidates, k0 newest
  uint32_t q0 = 0, q1 = 0, nq = 0;                  // condition candidates, q0 newest
  uint32_t idx = 0;
  for (; idx < n_cap; ++idx) {
    const uint32_t hv = tab[1 + idx], kl = hv & 0xFFu;
    if (((0x46800u >> (kl & 31u)) & 1u) && kl < 32u) {
      k3 = k2, k2 = k1, k1 = k0, k0 = (rel << 16) | hv;
      ++nk;
    }
    rel += kl + (hv >> 8);
  }
  for (; idx < n; ++idx) {
    const uint32_t hv = tab[1 + idx];
    if (hv == (5u | (4u << 8))) {
      q1 = q0, q0 = (rel << 16) | hv;
      ++nq;
    }
    rel += (hv & 0xFFu) + (hv >> 8);
  }
  // A row whose texts do not all fit it (rel is where the last one would end) is malformed: the general walk
  // knows what counts in such a row.  Past this line every entry lies inside the row, so no candidate needs a
  // bounds check and the order of the candidates only matters for a key that appears twice.
  if (nk > 4u || nq > 2u || rel > e - s) return false;
  // capacity candidates, oldest first: the first four key bytes.  The few that pass (one per GPU node) are
  // resolved afterwards, so that lanes whose key sat at different candidate positions compare key tails and
  // parse values together.
  uint32_t real_a = 0, real_b = 0, n_real = 0;  // real_a older
#pragma unroll
  for (int i = 3; i >= 0; --i) {
    const uint32_t rec = i == 3 ? k3 : i == 2 ? k2 : i == 1 ? k1 : k0;
    if ((uint32_t)i < nk) {
      const uint32_t kl = rec & 0xFFu;
      const uint32_t e0 = kl == 14u ? pack4("nvid", 0, 4) : kl == 11u ? pack4("amd.", 0, 4) : kl == 18u ? pack4("gpu.", 0, 4) : pack4("inte", 0, 4);
      if (ld32u(buf, s + (rec >> 16)) == e0) {
        real_a = real_b;
        real_b = rec;
        ++n_real;
      }
    }
  }
  if (n_real > 2u) return false;
  if (n_real > 1u) resolve_key(buf, s, real_a, max_digits, r);
  if (n_real > 0u) resolve_key(buf, s, real_b, max_digits, r);
  // condition candidates: any (Ready, True) will do (check-gpu-node.py:175-177)
#pragma unroll
  for (int i = 1; i >= 0; --i) {
    const uint32_t rec = i == 1 ? q1 : q0;
    if ((uint32_t)i < nq) {
      const uint32_t kp = s + (rec >> 16);
      if (ld32u(buf, kp) == pack4("Read", 0, 4) && ld32u(buf, kp + 4) == pack4("yTru", 0, 4) && (ld32u(buf, kp + 8) & 0xFFu) == 'e')
        r.bits |= R_READY;
    }
  }
  return true;
}

// 128-bit accumulate of a signed 64-bit addend into (lo, hi); exact under concurrent callers.
__device__ __forceinline__ void acc128_add(unsigned long long* lo, unsigned long long* hi, long long v) {
  if (v == 0) return;
  unsigned long long uv = (unsigned long long)v;
  unsigned long long old = atomicAdd(lo, uv);
  unsigned long long carry = (old + uv < old) ? 1ull : 0ull;
  unsigned long long dh = (v < 0 ? ~0ull : 0ull) + carry;
  if (dh) atomicAdd(hi, dh);
}

// ------------------------------------------------------------------------------------------------
// Decoupled look-back, 512 tiles wide.  One warp; lane l reads the 16 aggregates of tiles
// [g + 16 l, g + 16 l + 16) of the aligned 512-tile group g below `t`, and the inclusive prefix at the
// end of that span.  The nearest valid inclusive prefix (highest lane) ends the search; the lanes
// above it add their aggregates.  Width matters more than depth here: the prefix front can advance
// by at most one window per global round trip, so a 32-wide look-back caps the whole kernel at
// ~20 tiles/us, far below what the memory system feeds.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t lookback(const Params& p, int t, uint32_t total, int lane) {
  const uint32_t tag = p.epoch & kEpochMask;  // agg[t] was published by the walkers the moment the tile was counted
  uint32_t excl = 0;
  int hi = t;  // tiles [0, hi) still to be accounted for
  while (hi > 0) {
    const int g0 = (hi - 1) & ~(kLookbackWindow - 1);
    const int base = g0 + kLookbackSpan * lane;
    const int need = min(max(hi - base, 0), kLookbackSpan);
    uint32_t sum, incv = 0;
    int top;  // highest lane holding a valid inclusive prefix, -1 if none
    bool good;
    do {
      sum = 0;
      bool ok = true;
      if (need > 0) {
#pragma unroll
        for (int q = 0; q < kLookbackSpan / 4; q++) {
          const uint4 a = ld_volatile_v4(p.agg + base + 4 * q);
          const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
          for (int k = 0; k < 4; k++) {
            if (4 * q + k < need) {
              ok = ok && (w[k] >> 10) == tag;
              sum += w[k] & 1023u;
            }
          }
        }
      }
      bool has_inc = false;
      if (need == kLookbackSpan) {
        const unsigned long long d = ld_volatile_u64(p.inc + base + kLookbackSpan - 1);
        has_inc = (uint32_t)(d >> 32) == tag;
        incv = (uint32_t)d;
      }
      const uint32_t m = __ballot_sync(0xFFFFFFFFu, has_inc);
      top = m ? 31 - __clz(m) : -1;
      good = lane > top ? ok : true;
#ifdef GNC_TRACE
      {
        const uint32_t badm = __ballot_sync(0xFFFFFFFFu, !good);
        if (badm) GNC_D(0x5000000ull | ((unsigned long long)(uint32_t)(g0 + kLookbackSpan * (__ffs(badm) - 1)) << 32) | (uint32_t)(t & 0xFFFFF));
      }
#endif
    } while (!__all_sync(0xFFFFFFFFu, good));
    excl += __reduce_add_sync(0xFFFFFFFFu, lane > top ? sum : (lane == top ? incv : 0u));
    if (top >= 0) break;
    hi = g0;
  }
  if (lane == 0) st_volatile_u64(p.inc + t, ((unsigned long long)tag << 32) | (excl + total));
  return excl;
}

// ------------------------------------------------------------------------------------------------
// The kernel.  TN nodes per tile, one thread per node.  Warps: TN/32 walkers, 1 producer, 2 scan.
// ------------------------------------------------------------------------------------------------
// Walker groups.  A CTA's walkers form GROUPS groups of TN threads; the CTA's q-th tile is walked by group
// q % GROUPS out of stage q % STAGES, so a CTA has GROUPS tiles being walked and STAGES - GROUPS being
// loaded.  (What bounds the kernel is how many walker warps an SM holds -- they run dependent integer
// chains, ~6 cycles per instruction -- and what bounds that is shared memory: with one group a CTA needs
// two whole stages per TN walkers, with two groups three stages serve 2 TN walkers.)
// Ring: a tile's records are written when its group is about to reuse the tile's ring slot, i.e.
// kRing / GROUPS - 1 of the group's iterations after it was walked.
template <int GROUPS>
__host__ __device__ constexpr int ring_slots() { return GROUPS == 1 ? 6 : 8; }
enum { P_GPU = 0, P_READY, P_DEFERRED, P_MALFORMED, P_KEY0, P_N = 8 };

template <int TN, int STAGES, int STAGE_BYTES, int GROUPS = 1>
struct Smem {
  static constexpr int kRing = ring_slots<GROUPS>();
  alignas(128) uint8_t stage[STAGES][STAGE_BYTES + 32];
  alignas(16) uint32_t rp[STAGES][TN + 8];
  alignas(8) uint64_t full[STAGES][GROUPS];  // producer/TMA -> the group that walks the tile: the stage holds it.  One barrier per
                                             // (stage, group): a waiter must see every phase of its barrier, or parities alias
  uint64_t pref_full[kRing];         // scan warp -> walkers: output positions are in warp_base[]
  unsigned long long acc[GNC_N_COUNTS];
  int tile[STAGES];
  uint32_t lo[STAGES];
  uint32_t hi[STAGES];      // rows of the tile must end at or before this byte
  uint32_t direct[STAGES];  // 1: tile too large to stage, rows are read from global memory
  alignas(16) int part[kRing][TN / 32][P_N];
  uint32_t warp_base[kRing][TN / 32];
  uint32_t ballot[kRing][TN / 32];  // the walker warps' GPU-node ballots (fused gather: the scan warp sends the tile's record)
  int scan_tile[kRing];
  uint32_t tile_cnt[kRing];  // [31:16] walker warps done with the tile, [15:0] GPU nodes counted so far
  alignas(8) uint2 stash_vals[kRing][TN];  // int16 x 4 per node
  uint8_t stash_meta[kRing][TN];           // bit 0 GPU node, bit 1 wide, bits 2..6 rank in warp
};

template <int TN, int GROUPS = 1>
__host__ __device__ constexpr int cta_threads() { return GROUPS * TN + 96; }

// FLEET: the fused multi-GPU gather (peer stores while emitting); compiled out of the single-GPU kernels.
template <int TN, int STAGES, int STAGE_BYTES, int MIN_CTAS, bool FLEET, int GROUPS = 1>
__global__ void __launch_bounds__(cta_threads<TN, GROUPS>(), MIN_CTAS) classify_kernel(const __grid_constant__ Params p) {
  using S = Smem<TN, STAGES, STAGE_BYTES, GROUPS>;
  extern __shared__ __align__(128) uint8_t smem_raw[];
  S& sm = *reinterpret_cast<S*>(smem_raw);
  constexpr int kWalkWarps = TN / 32;            // per group
  constexpr int kAllWalkWarps = GROUPS * kWalkWarps;
  constexpr int kRing = S::kRing;
  constexpr int kStagePeriod = (STAGES % GROUPS == 0) ? STAGES : STAGES * GROUPS;  // lcm, for 1 or 2 groups
  static_assert(TN % 32 == 0 && kWalkWarps <= 32 && TN < 1024, "one scan-warp lane per walker warp; 10-bit tile counts");
  static_assert(GROUPS == 1 || GROUPS == 2, "two scan warps: they alternate (one group) or serve a group each");
  static_assert(kRing % 2 == 0 && kRing % GROUPS == 0 && STAGES > GROUPS - 1 + 0, "ring slots and stages per group");
  // Hand-offs to warps that would otherwise do nothing but poll use hardware named barriers (the
  // waiting warp is descheduled, it issues nothing): stage s free -> producer, counts of ring slot b
  // ready -> the scan warp that owns slot b.  Both have TN walker threads arriving and one warp waiting.
  // (ring slot b always belongs to scan warp b & 1, and with two groups to group b & 1)
  constexpr int kBarStage = 1, kBarCnt = 1 + STAGES, kBarThreads = TN + 32;
  static_assert(kBarCnt + kRing <= 16, "named barrier ids");
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const unsigned long long t_cta = FLEET ? global_ns() : 0ull;

  if (tid == 0) {
    for (int s = 0; s < STAGES; s++)
      for (int g = 0; g < GROUPS; g++) mbar_init(&sm.full[s][g], 1);
    for (int b = 0; b < kRing; b++) {
      mbar_init(&sm.pref_full[b], 1);
    }
    mbar_init_fence();
  }
  if (tid == 0) GNC_T(7, 0, gtimer());  // CTA entry
  if (tid < GNC_N_COUNTS) sm.acc[tid] = 0;
  if (tid < kRing) sm.tile_cnt[tid] = 0;
  if (FLEET && tid < (int)p.fleet_world && p.fleet_ack_need) {
    // the slots about to be rewritten were last filled two launches ago: every peer must have read them
    const unsigned long long t0 = global_ns();
    while (ld_sys_u64(p.fleet_ack + tid) < p.fleet_ack_need)
      if (global_ns() - t0 > kFleetTimeoutNs) {
        atomicAdd(&sm.acc[GNC_C_OUT_DROPPED], 1ull << 40);  // poisons the counts: the host reports it
        break;
      }
  }
  __syncthreads();

  if (warp == kAllWalkWarps) {
    // ------------------------------ producer ------------------------------
    {
      // Lane 0 runs a small software pipeline, the other lanes only keep it company at the named barrier.  A
      // tile costs two dependent trips to L2 before its bytes can be asked for -- the ticket (atomicAdd, tiles
      // are handed out in fleet order), then the two row_ptr bounds -- ~1.4 us, which with one CTA per SM
      // and two walker groups is most of a tile period.  So the trips of different tiles overlap: in the
      // iteration that copies tile `it` to shared memory, the ticket of tile it + 2 and the bounds of tile
      // it + 1 are requested first and used last.  Tile it + 1 is prefetched into L2 as soon as its bounds
      // are known: HBM latency is then paid while its stage is still busy with an earlier tile, and the later
      // bulk copy is an L2 hit.  (More lead than that puts ticket order further from completion order and the
      // look-backs wait longer.)
      struct Plan {
        int tile;  // -1: no more tiles
        uint32_t lo_al, hi, bytes, cnt;
        bool direct;
      };
      auto finish_plan = [&](uint32_t tile, uint32_t lo, uint32_t hi) {
        Plan pl;
        const uint32_t first = tile * TN;
        pl.tile = (int)tile;
        pl.cnt = min((uint32_t)TN, p.n_nodes - first);
        pl.hi = hi;
        pl.lo_al = lo & ~15u;
        pl.bytes = ((pl.hi + 15u) & ~15u) - pl.lo_al;
        pl.direct = pl.hi < lo || pl.hi > p.rows_bytes || lo < p.rows_lo || pl.bytes > (uint32_t)STAGE_BYTES;
        if (!pl.direct && pl.bytes) bulk_prefetch_l2(p.rows + pl.lo_al, pl.bytes);
        bulk_prefetch_l2(p.row_ptr + first, ((pl.cnt + 1) * 4u + 15u) & ~15u);
        return pl;
      };
      Plan cur;
      cur.tile = -1;
      int next_tile = -1;  // ticket of tile it + 1
      bool dry = false;    // a ticket past the last tile has been drawn: no more are taken
      if (lane == 0) {
        const uint32_t t0 = atomicAdd(&p.scratch->ticket, 1u);
        if (t0 < p.num_tiles) {
          GNC_TILE(t0, 1);
          const uint32_t t1 = atomicAdd(&p.scratch->ticket, 1u);
          const uint32_t first = t0 * TN;
          cur = finish_plan(t0, __ldg(p.row_ptr + first), __ldg(p.row_ptr + first + min((uint32_t)TN, p.n_nodes - first)));
          if (t1 < p.num_tiles) next_tile = (int)t1;
          else dry = true;
        } else {
          dry = true;
        }
      }
      int ended = 0;
      for (uint32_t it = 0;; ++it) {
        const int s = it % STAGES;
        uint32_t ticket = 0, lo1 = 0, hi1 = 0;
        if (lane == 0) {  // requested now, looked at after this tile's copy has been issued
          if (!dry) ticket = atomicAdd(&p.scratch->ticket, 1u);
          if (next_tile >= 0) {
            const uint32_t first = (uint32_t)next_tile * TN;
            lo1 = __ldg(p.row_ptr + first);
            hi1 = __ldg(p.row_ptr + first + min((uint32_t)TN, p.n_nodes - first));
          }
        }
        const int cur_tile = __shfl_sync(0xFFFFFFFFu, cur.tile, 0);
        GNC_D(0x1000000ull | (it << 8) | 1);
        if (it >= (uint32_t)STAGES) named_bar_sync(kBarStage + s, kBarThreads);  // every walker has left stage s
        GNC_D(0x1000000ull | (it << 8) | 2);
        if (cur_tile < 0) {  // no more tiles: one end marker per group, in the stages their next tiles would have used
          if (lane == 0) {
            sm.tile[s] = -1;
            mbar_arrive(&sm.full[s][it % GROUPS]);
          }
          if (++ended == GROUPS) break;
          continue;
        }
        if (lane == 0) {
          const uint32_t first = (uint32_t)cur.tile * TN;
          const uint32_t rp_bytes = ((cur.cnt + 1) * 4u + 15u) & ~15u;
          sm.tile[s] = cur.tile;
          sm.lo[s] = cur.lo_al;
          sm.hi[s] = cur.direct ? (uint32_t)min(p.rows_bytes, (uint64_t)0xFFFFFFFFu) : cur.hi;
          sm.direct[s] = cur.direct ? 1u : 0u;
          const uint32_t stage_bytes = cur.direct ? 0u : cur.bytes;
          uint64_t* full = &sm.full[s][it % GROUPS];
          mbar_arrive_expect_tx(full, rp_bytes + stage_bytes);
          bulk_g2s(sm.rp[s], p.row_ptr + first, rp_bytes, full);
          if (stage_bytes) bulk_g2s(sm.stage[s], p.rows + cur.lo_al, stage_bytes, full);
          GNC_TILE(cur.tile, 2 | (it << 16));
          // the next tile's plan (and its L2 prefetch), the ticket after that
          cur.tile = -1;
          if (next_tile >= 0) cur = finish_plan((uint32_t)next_tile, lo1, hi1);
          next_tile = -1;
          if (!dry) {
            if (ticket < p.num_tiles) {
              next_tile = (int)ticket;
              GNC_TILE(ticket, 1);
            } else {
              dry = true;
            }
          }
        }
      }
    }
    __syncwarp();  // lanes 1..31 rejoin lane 0 before the CTA-wide barrier below
  } else if (warp > kAllWalkWarps) {
    // ------------------------------ scan warps: this CTA's tile `it` goes to warp it & 1 -------------
    long long a_gpu = 0, a_ready = 0, a_def = 0, a_bad = 0, a_key[4] = {0, 0, 0, 0};  // lane 0 only
    for (uint32_t it = warp - kAllWalkWarps - 1;; it += 2) {
      const int b = it % kRing;
      GNC_D(0x2000000ull | (it << 8) | 1);
      named_bar_sync(kBarCnt + b, kBarThreads);  // the walkers' partial counts of this tile are in part[b]
      const int tile = sm.scan_tile[b];
      GNC_D(0x2000000ull | (it << 8) | 2 | ((unsigned long long)(uint32_t)tile << 32));
      if (tile < 0) break;
      GNC_T(5, it, gtimer());
      int part[P_N];
#pragma unroll
      for (int q = 0; q < P_N; q++) part[q] = lane < kWalkWarps ? sm.part[b][lane][q] : 0;
      const uint32_t c = (uint32_t)part[P_GPU];
      uint32_t incl = c;
#pragma unroll
      for (int d = 1; d < kWalkWarps; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= d) incl += t;
      }
      const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, kWalkWarps - 1);
      const uint32_t excl = lookback(p, tile, total, lane);
      if (lane < kWalkWarps) sm.warp_base[b][lane] = excl + incl - c;
      if (FLEET && (uint32_t)(tile + 1) * (kWalkWarps + 1) <= p.fleet_cap) {
        // the tile's record -- its exclusive prefix, then its warps' ballots -- into every rank's buffer: a lane per
        // (rank, word), so a rank gets its 4 (kWalkWarps + 1) bytes as one contiguous run of stores over NVLink
        const uint32_t base = p.fleet_rank * p.fleet_slot_u32 + kFleetHeader + (uint32_t)tile * (kWalkWarps + 1);
        for (uint32_t idx = lane; idx < p.fleet_world * (kWalkWarps + 1); idx += 32) {
          const uint32_t q = idx / (kWalkWarps + 1), w = idx - q * (kWalkWarps + 1);
          p.fleet_peer[q][base + w] = w ? sm.ballot[b][w - 1] : excl;
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.pref_full[b]);
      GNC_T(6, it, gtimer());
      // the tile's share of the counts block
      a_gpu += total;
      a_ready += __reduce_add_sync(0xFFFFFFFFu, part[P_READY]);
      a_def += __reduce_add_sync(0xFFFFFFFFu, part[P_DEFERRED]);
      a_bad += __reduce_add_sync(0xFFFFFFFFu, part[P_MALFORMED]);
#pragma unroll
      for (int k = 0; k < 4; k++) a_key[k] += __reduce_add_sync(0xFFFFFFFFu, part[P_KEY0 + k]);
    }
    if (lane == 0) {
      atomicAdd(&sm.acc[GNC_C_GPU_NODES], (unsigned long long)a_gpu);
      atomicAdd(&sm.acc[GNC_C_READY_NODES], (unsigned long long)a_ready);
      atomicAdd(&sm.acc[GNC_C_DEFERRED], (unsigned long long)a_def);
      atomicAdd(&sm.acc[GNC_C_MALFORMED], (unsigned long long)a_bad);
      for (int k = 0; k < 4; k++) acc128_add(&sm.acc[GNC_C_KEY_LO + k], &sm.acc[GNC_C_KEY_HI + k], a_key[k]);
    }
    __syncwarp();
  } else {
    // ------------------------------ walkers: thread j owns node j of every tile -------------------
    // (of the tiles its group walks: the CTA's tiles grp, grp + GROUPS, ...; `it` below is the CTA's tile number)
    const int grp = warp / kWalkWarps, wg = warp - grp * kWalkWarps;
    const int j = wg * 32 + lane;
    constexpr int kLag = kRing / GROUPS - 1;  // a tile's records are written this many of the group's iterations after it was walked
    uint32_t lag_first[kLag];  // first node of the tiles walked 1 .. kLag iterations ago (registers: static indices only)
#pragma unroll
    for (int q = 0; q < kLag; q++) lag_first[q] = 0;

    // Write a stashed tile's GPU-node records now that its output positions are known.
    auto emit_stashed = [&](uint32_t it_e, uint32_t first_e) {
      const int b = it_e % kRing;
      mbar_wait(&sm.pref_full[b], (it_e / kRing) & 1u);
      GNC_D(0x3000000ull | (it_e << 8) | 7 | ((unsigned long long)sm.warp_base[b][wg] << 32));
      const uint32_t meta = sm.stash_meta[b][j];
      bool dropped = false;
      if (meta & 1u) {
        const uint32_t pos = sm.warp_base[b][wg] + (meta >> 2);
        if (pos < p.out_cap) {
          long long v0, v1, v2, v3;
          if (meta & 2u) {  // values did not fit 16 bits: read them again from the row
            SlowRow sr;
            slow_row(p.rows, __ldg(p.row_ptr + first_e + j), __ldg(p.row_ptr + first_e + j + 1), p.max_digits, &sr);
            v0 = sr.v[0], v1 = sr.v[1], v2 = sr.v[2], v3 = sr.v[3];
          } else {
            const uint2 sv = sm.stash_vals[b][j];
            v0 = (short)(sv.x & 0xFFFFu), v1 = (short)(sv.x >> 16), v2 = (short)(sv.y & 0xFFFFu), v3 = (short)(sv.y >> 16);
          }
          p.gpu_idx[pos] = p.node_base + first_e + j;
          p.gpu_total[pos] = v0 + v1 + v2 + v3;
          longlong2* vo = reinterpret_cast<longlong2*>(p.gpu_vals + 4ull * pos);
          vo[0] = make_longlong2(v0, v1);
          vo[1] = make_longlong2(v2, v3);
        } else {
          dropped = true;
        }
      }
      if (__any_sync(0xFFFFFFFFu, dropped)) {
        const uint32_t drop_b = __ballot_sync(0xFFFFFFFFu, dropped);
        if (lane == 0) atomicAdd(&sm.acc[GNC_C_OUT_DROPPED], (unsigned long long)__popc(drop_b));
      }
    };

    uint32_t it = grp;
    for (;; it += GROUPS) {
      const int s = it % STAGES;
      GNC_D(0x3000000ull | (it << 8) | 1);
      mbar_wait(&sm.full[s][grp], (it / kStagePeriod) & 1u);  // this group's tiles come to stage s every kStagePeriod tiles
      const int tile = sm.tile[s];
      GNC_D(0x3000000ull | (it << 8) | 2 | ((unsigned long long)(uint32_t)tile << 32));
      if (tile < 0) break;
      if (lane == 0 && wg == 0) GNC_TILE(tile, 3 | (it << 16));
      if (tid == 0) GNC_T(0, it / GROUPS, gtimer());
      const uint32_t first = (uint32_t)tile * TN;
      const uint32_t cnt = min((uint32_t)TN, p.n_nodes - first);
      const bool live = (uint32_t)j < cnt;
      const uint32_t lo_al = sm.lo[s];
      const bool direct = sm.direct[s] != 0;
      // a row must lie inside the tile's byte range (inside `rows` when read directly); one that does
      // not (a row_ptr that decreases or points past the data) is an empty, malformed row
      uint32_t rs = 0, re = 0;
      bool bad_ptr = false;
      if (live) {
        rs = sm.rp[s][j];
        re = sm.rp[s][j + 1];
        bad_ptr = re < rs || re > sm.hi[s] || rs < (direct ? p.rows_lo : sm.rp[s][0]);
        if (bad_ptr) rs = re = (direct ? p.rows_lo : sm.rp[s][0]);
      }
      GNC_D(0x3000000ull | (it << 8) | 8 | ((unsigned long long)(rs - lo_al) << 32) | ((unsigned long long)direct << 63) | ((unsigned long long)bad_ptr << 62));
      NodeResult r;
      r.v01 = r.v23 = r.bits = 0;
      if (live) {
        if (direct || !walk_row(sm.stage[s], rs - lo_al, re - lo_al, p.max_digits, r)) {
          const uint3 q = slow_row_packed(p.rows, rs, re, p.max_digits);
          r.v01 = q.x, r.v23 = q.y, r.bits = q.z;
        }
      }
      __syncwarp();
      GNC_D(0x3000000ull | (it << 8) | 3);
      named_bar_arrive(kBarStage + s, kBarThreads);  // this warp no longer reads the stage
      if (tid == 0) GNC_T(1, it / GROUPS, gtimer());

      // -------- flags, ranks, partial counts; the records go out kRing-1 iterations later --------
      long long wv[4] = {0, 0, 0, 0};  // exact values of a wide node (rare)
      int v0 = (short)(r.v01 & 0xFFFFu), v1 = (short)(r.v01 >> 16), v2 = (short)(r.v23 & 0xFFFFu), v3 = (short)(r.v23 >> 16);
      bool positive = v0 + v1 + v2 + v3 > 0;
      const bool wide = (r.bits & R_WIDE) != 0;
      if (wide) {
        SlowRow sr;
        slow_row(p.rows, rs, re, p.max_digits, &sr);
        for (int k = 0; k < 4; k++) wv[k] = sr.v[k];
        positive = wv[0] + wv[1] + wv[2] + wv[3] > 0;  // |v| < 10^18 each
      }
      const bool ready = (r.bits & R_READY) != 0;
      const bool malformed = bad_ptr || (r.bits & R_MALFORMED) != 0;
      const bool deferred = ((r.bits >> R_BIG_SHIFT) & 0xFu) != 0;
      const bool is_gpu = live && !deferred && positive;
      const uint32_t f = (is_gpu ? GNC_F_GPU : 0u) | (ready ? GNC_F_READY : 0u) | (deferred ? GNC_F_DEFERRED : 0u) |
                         ((r.bits & R_PRESENT) << 3) | (malformed ? GNC_F_MALFORMED : 0u);
      if (live) p.flags[first + j] = (uint8_t)f;
      const uint32_t gb = __ballot_sync(0xFFFFFFFFu, is_gpu);
      const uint32_t rank = __popc(gb & ((1u << lane) - 1u));
      const uint32_t n_ready = __popc(__ballot_sync(0xFFFFFFFFu, is_gpu && ready));
      const uint32_t n_def = __popc(__ballot_sync(0xFFFFFFFFu, live && deferred));
      const uint32_t n_bad = __popc(__ballot_sync(0xFFFFFFFFu, live && malformed));
      const bool count_narrow = is_gpu && !wide;
      const int s0 = __reduce_add_sync(0xFFFFFFFFu, count_narrow ? v0 : 0);
      const int s1 = __reduce_add_sync(0xFFFFFFFFu, count_narrow ? v1 : 0);
      const int s2 = __reduce_add_sync(0xFFFFFFFFu, count_narrow ? v2 : 0);
      const int s3 = __reduce_add_sync(0xFFFFFFFFu, count_narrow ? v3 : 0);
      const int b = it % kRing;
      if (lane == 0) {
        int4* pp = reinterpret_cast<int4*>(sm.part[b][wg]);
        pp[0] = make_int4(__popc(gb), (int)n_ready, (int)n_def, (int)n_bad);  // P_GPU, P_READY, P_DEFERRED, P_MALFORMED
        pp[1] = make_int4(s0, s1, s2, s3);                                     // P_KEY0 ..
        if (wg == 0) sm.scan_tile[b] = tile;
        if (FLEET) sm.ballot[b][wg] = gb;
        // the last warp to count its share publishes the tile's aggregate at once: other CTAs'
        // look-backs must never wait for this CTA's scan warp to get round to the tile
        const uint32_t old = atomicAdd(&sm.tile_cnt[b], (uint32_t)__popc(gb) | (1u << 16));
        if ((old >> 16) == (uint32_t)kWalkWarps - 1) {
          sm.tile_cnt[b] = 0;
          st_volatile_u32(p.agg + tile, ((p.epoch & kEpochMask) << 10) | ((old & 0xFFFFu) + (uint32_t)__popc(gb)));
          GNC_TILE(tile, 4 | (it << 16));
        }
      }
      __syncwarp();
      GNC_D(0x3000000ull | (it << 8) | 4);
      named_bar_arrive(kBarCnt + b, kBarThreads);  // the scan warp can start on this tile
      if (is_gpu && wide) {
#pragma unroll
        for (int k = 0; k < 4; k++) acc128_add(&sm.acc[GNC_C_KEY_LO + k], &sm.acc[GNC_C_KEY_HI + k], wv[k]);
      }
      if (live && deferred) {
        const uint32_t slot = atomicAdd(&p.scratch->defer_count, 1u);
        if (slot < p.defer_cap) emit_defer(p, first + j, rs, re, slot);
        else atomicAdd(&sm.acc[GNC_C_DEFER_DROPPED], 1ull);
      }
      if (tid == 0) GNC_T(2, it / GROUPS, gtimer());

      // the tile this group walked kLag iterations ago: its positions have had that long to arrive
      GNC_D(0x3000000ull | (it << 8) | 5);
      if (it >= (uint32_t)(kLag * GROUPS)) emit_stashed(it - kLag * GROUPS, lag_first[kLag - 1]);
      GNC_D(0x3000000ull | (it << 8) | 6);
      if (tid == 0) GNC_T(3, it / GROUPS, gtimer());
      sm.stash_vals[b][j] = make_uint2(r.v01, r.v23);
      sm.stash_meta[b][j] = (uint8_t)((is_gpu ? 1u : 0u) | (wide ? 2u : 0u) | (rank << 2));
#pragma unroll
      for (int q = kLag - 1; q > 0; q--) lag_first[q] = lag_first[q - 1];
      lag_first[0] = first;
      if (tid == 0) GNC_T(4, it / GROUPS, gtimer());
    }
    if (tid == 0) GNC_T(7, 1, gtimer());  // last tile walked
    // drain the ring, then release the scan warps that wait for this group (both of them when there is one
    // group, its own when there are two): their next phase completes with no tile
#pragma unroll
    for (int q = kLag - 1; q >= 0; q--)
      if (it >= (uint32_t)((q + 1) * GROUPS)) emit_stashed(it - (q + 1) * GROUPS, lag_first[q]);
    __syncwarp();
#pragma unroll
    for (int d = 0; d < 2; d += GROUPS) {
      if (j == 0) sm.scan_tile[(it + d) % kRing] = -1;
      __syncwarp();
      named_bar_arrive(kBarCnt + (int)((it + d) % kRing), kBarThreads);
    }
  }

  // ------------------------------ grid-wide finish ------------------------------
  GNC_D(0x4000000ull);
  __syncthreads();
  GNC_D(0x4000001ull);
  if (tid == 0) GNC_T(7, 2, gtimer());  // ring drained, every warp here
  // One thread per count: done by thread 0 alone, the sixteen dependent round trips to L2 here and in the
  // last CTA below were ~10 us at the very end of every launch.
  ScratchHeader* sc = p.scratch;
  if (tid < GNC_N_COUNTS) {
    const int c = tid;
    const unsigned long long v = sm.acc[c];
    if (c >= GNC_C_KEY_LO && c < GNC_C_KEY_HI) {  // lo with carry into hi
      if (v) {
        const unsigned long long old = atomicAdd(&sc->accum[c], v);
        if (old + v < old) atomicAdd(&sc->accum[c + 4], 1ull);
      }
    } else if (v) {
      atomicAdd(&sc->accum[c], v);
    }
  }
  __syncthreads();
  if (tid == 0) {
    if (FLEET) __threadfence_system();  // this CTA's peer stores are out before it counts as done
    else __threadfence();
    const uint32_t prev = atomicAdd(&sc->done, 1u);
    __threadfence();
    sm.tile_cnt[0] = prev == gridDim.x - 1 ? 1u : 0u;  // (the ring's counters are idle by now)
  }
  __syncthreads();
  if (sm.tile_cnt[0]) {  // last CTA: publish the counts block and re-arm the scratch
    if (tid < GNC_N_COUNTS) {
      const int c = tid;
      long long v = (long long)ld_volatile_u64(&sc->accum[c]);
      if (c == GNC_C_SCANNED) v = p.n_nodes;
      p.counts[c] = v;
      sc->accum[c] = 0;
      if (FLEET)
        for (uint32_t q = 0; q < p.fleet_world; q++)
          reinterpret_cast<long long*>(p.fleet_peer[q] + (size_t)p.fleet_rank * p.fleet_slot_u32)[c] = v;
    } else if (FLEET && tid < GNC_N_COUNTS + 4) {  // what the receivers need to read the tile records
      const int h = FLEET_H_TILE_NODES + tid - GNC_N_COUNTS;
      const uint32_t v = h == FLEET_H_TILE_NODES ? (uint32_t)TN : h == FLEET_H_TILES ? p.num_tiles : h == FLEET_H_NODE_BASE ? p.node_base : p.n_nodes;
      for (uint32_t q = 0; q < p.fleet_world; q++) p.fleet_peer[q][(size_t)p.fleet_rank * p.fleet_slot_u32 + h] = v;
    }
    __syncthreads();
    if (tid == 0) {
      if (FLEET) {  // every index and the counts are on their way: raise this rank's flag everywhere
        __threadfence_system();
        for (uint32_t q = 0; q < p.fleet_world; q++) st_sys_u64(p.fleet_flag[q] + p.fleet_rank, p.fleet_epoch);
      }
      sc->ticket = 0;
      sc->defer_count = 0;
      __threadfence();
      sc->done = 0;
    }
  }
  if (FLEET && p.fleet_out_idx) {
    // The gather, in the same launch: every rank's records for this epoch land in the local buffer while the
    // kernels run; once they are all here (flags), this grid rebuilds the fleet-wide list.  The stages are free
    // by now (every bulk copy was waited for by the walkers that used it): they serve as the warps' buffers.
    __syncthreads();
    static_assert((size_t)STAGES * (STAGE_BYTES + 32) >= (size_t)(cta_threads<TN, GROUPS>() / 32) * 2 * kUnpackHalfWords * 4 + 1024, "a buffer per warp and the byte table inside the stages");
    FleetUnpack u;
    u.slots = p.fleet_peer[p.fleet_rank];
    u.flags = p.fleet_flag[p.fleet_rank];
    u.epoch = p.fleet_epoch;
    u.world = (int)p.fleet_world;
    u.rank = (int)p.fleet_rank;
    u.slot_u32 = p.fleet_slot_u32;
    u.cap = p.fleet_cap;
    u.out_idx = p.fleet_out_idx;
    u.out_cap = p.fleet_out_cap;
    u.out_counts = p.fleet_out_counts;
    u.done = p.fleet_unpack_done;
    u.ack_peer = p.fleet_ack_peer;
    u.timing = reinterpret_cast<unsigned long long*>(p.fleet_unpack_done) + 8;  // the spare words after the counter
    u.t_before = global_ns() - t_cta;
    uint32_t* ubuf = reinterpret_cast<uint32_t*>(&sm.stage[0][0]);
    fleet_unpack(u, ubuf, reinterpret_cast<unsigned long long*>(&sm.rp[0][0]), ubuf + (size_t)(cta_threads<TN, GROUPS>() / 32) * 2 * kUnpackHalfWords);
  }
  if (tid == 0) GNC_T(7, 3, gtimer());  // CTA exit
}

}  // namespace gnc
