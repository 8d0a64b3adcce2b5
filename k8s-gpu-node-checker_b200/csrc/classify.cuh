// classify.cuh -- the node-classification kernel (sm_100a).
//
// One persistent CTA per tile stream.  A producer warp takes tile tickets in fleet order and stages
// each tile's bytes in shared memory with TMA bulk copies (cp.async.bulk + mbarrier); TN "capacity"
// threads and TN "condition" threads each walk one node's row out of shared memory; GPU nodes are
// compacted in fleet order with warp ballots, a CTA scan and a decoupled look-back across tiles.
//
// Reference behaviour implemented (file:line in /root/reference/check-gpu-node.py):
//   key match + quantity parse  :186-195      Ready test  :172-178      gpus sum / filter  :201, :222-225
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/gnc.h"
#include "unicode_tables.inc"

namespace gnc {

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

// Tile descriptor for the look-back: [63:62] status, [61:32] launch epoch, [31:0] GPU-node count.
constexpr uint64_t kStatusAggregate = 1ull << 62;
constexpr uint64_t kStatusInclusive = 2ull << 62;
constexpr uint32_t kEpochMask = (1u << 30) - 1;

struct Params {
  const uint8_t* rows;
  const uint32_t* row_ptr;
  uint64_t rows_bytes;
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
  unsigned long long* tile_state;
};

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
__host__ __device__ constexpr uint32_t mask4(int i, int n) {
  return (i < n ? 0xFFu : 0u) | (i + 1 < n ? 0xFF00u : 0u) | (i + 2 < n ? 0xFF0000u : 0u) | (i + 3 < n ? 0xFF000000u : 0u);
}

// bytes [2, N) of the key at `kp` against literal S (bytes 0..1 were checked from the entry header)
template <int N>
__device__ __forceinline__ bool key_tail_eq(const uint8_t* buf, uint32_t kp, const char (&S)[N + 1]) {
  bool eq = true;
#pragma unroll
  for (int i = 2; i < N; i += 4) eq = eq && ((ld32u(buf, kp + i) & mask4(i, N)) == pack4(S, i, N));
  return eq;
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

__device__ __noinline__ int parse_quantity_general(const uint8_t* s, uint32_t n, int max_digits, long long* out) {
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
  *out = neg ? -(long long)v : (long long)v;
  return Q_OK;
}

// Entry header at `pos` inside [pos, end): lengths, position of the key, first two key bytes.
struct Entry {
  uint32_t kl, vl, kp, k01;
  bool ok;
};
__device__ __forceinline__ Entry read_entry(const uint8_t* buf, uint32_t pos, uint32_t end) {
  Entry e;
  e.ok = false;
  e.kl = e.vl = e.kp = e.k01 = 0;
  if (end - pos < 2) return e;
  uint32_t w = ld32u(buf, pos);
  e.kl = w & 0xFFu;
  e.vl = (w >> 8) & 0xFFu;
  e.kp = pos + 2;
  e.k01 = w >> 16;
  if (e.kl == 255u || e.vl == 255u) {  // 32-bit length escapes
    if (e.kl == 255u) {
      if (end - e.kp < 4) return e;
      e.kl = ld32u(buf, e.kp);
      e.kp += 4;
    }
    if (e.vl == 255u) {
      if (end - e.kp < 4) return e;
      e.vl = ld32u(buf, e.kp);
      e.kp += 4;
    }
    e.k01 = ld32u(buf, e.kp) & 0xFFFFu;
    if ((unsigned long long)e.kl + e.vl > (unsigned long long)(end - e.kp)) return e;
  } else if (e.kl + e.vl > end - e.kp) {
    return e;
  }
  e.ok = true;
  return e;
}

// Row header: capacity section [*cap_b, *cap_e), conditions [*cap_e, e). False if the header is bad.
__device__ __forceinline__ bool row_sections(const uint8_t* buf, uint32_t s, uint32_t e, uint32_t* cap_b, uint32_t* cap_e) {
  *cap_b = *cap_e = e;
  if (e - s < 2) return e == s;
  uint32_t w = ld32u(buf, s), len = w & 0xFFFFu, p = s + 2;
  if (len == 0xFFFFu) {
    if (e - s < 6) return false;
    len = ld32u(buf, s + 2);
    p = s + 6;
  }
  if (len > e - p) return false;
  *cap_b = p;
  *cap_e = p + len;
  return true;
}

struct CapResult {
  long long v[4];
  uint32_t present;  // bit k: key k kept
  uint32_t big;      // bit k: key k kept but wider than 18 digits
  bool malformed;
};

__device__ __forceinline__ void walk_capacity(const uint8_t* buf, uint32_t s, uint32_t e, int max_digits, CapResult& r) {
  r.v[0] = r.v[1] = r.v[2] = r.v[3] = 0;
  r.present = r.big = 0;
  uint32_t pos, cap_e;
  r.malformed = !row_sections(buf, s, e, &pos, &cap_e);
  if (r.malformed) return;
  while (pos < cap_e) {
    Entry en = read_entry(buf, pos, cap_e);
    if (!en.ok) {
      r.malformed = true;
      break;
    }
    int k = -1;
    if (en.kl == 14) {
      if (en.k01 == pack4("nv", 0, 2) && key_tail_eq<14>(buf, en.kp, "nvidia.com/gpu")) k = 0;
    } else if (en.kl == 11) {
      if (en.k01 == pack4("am", 0, 2) && key_tail_eq<11>(buf, en.kp, "amd.com/gpu")) k = 1;
    } else if (en.kl == 18) {
      if (en.k01 == pack4("gp", 0, 2) && key_tail_eq<18>(buf, en.kp, "gpu.intel.com/i915")) k = 2;
    } else if (en.kl == 13) {
      if (en.k01 == pack4("in", 0, 2) && key_tail_eq<13>(buf, en.kp, "intel.com/gpu")) k = 3;
    }
    uint32_t vp = en.kp + en.kl;
    if (k >= 0) {  // a repeated key overwrites: dict / json.loads last-wins
      int st = Q_ABSENT;
      long long val = 0;
      if (en.vl == 1) {
        uint32_t d = (uint32_t)buf[vp] - '0';
        if (d <= 9u) {
          st = Q_OK;
          val = d;
        } else {
          st = parse_quantity_general(buf + vp, 1, max_digits, &val);
        }
      } else if (en.vl != 0) {
        st = parse_quantity_general(buf + vp, en.vl, max_digits, &val);
      }
      uint32_t bit = 1u << k;
      r.present = (r.present & ~bit) | (st != Q_ABSENT ? bit : 0u);
      r.big = (r.big & ~bit) | (st == Q_BIG ? bit : 0u);
      val = (st == Q_OK) ? val : 0;
      if (k == 0) r.v[0] = val;
      else if (k == 1) r.v[1] = val;
      else if (k == 2) r.v[2] = val;
      else r.v[3] = val;
    }
    pos = vp + en.vl;
  }
}

// bit 0: Ready, bit 1: malformed
__device__ __forceinline__ uint32_t walk_conditions(const uint8_t* buf, uint32_t s, uint32_t e) {
  uint32_t cap_b, pos;
  if (!row_sections(buf, s, e, &cap_b, &pos)) return 2u;
  while (pos < e) {
    Entry en = read_entry(buf, pos, e);
    if (!en.ok) return 2u;
    if (en.kl == 5 && en.vl == 4 && en.k01 == pack4("Re", 0, 2) && ld32u(buf, en.kp + 2) == pack4("adyT", 0, 4) &&
        (ld32u(buf, en.kp + 6) & 0xFFFFFFu) == pack4("rue", 0, 3))
      return 1u;  // first hit returns (check-gpu-node.py:176-177)
    pos = en.kp + en.kl + en.vl;
  }
  return 0u;
}

// Deferred node: walk its row again (from global memory) and record where each kept value's text is.
__device__ __noinline__ void emit_defer(const Params& p, uint32_t node, uint32_t s, uint32_t e, uint32_t slot) {
  gnc_defer_rec rec;
  rec.node = p.node_base + node;
  for (int k = 0; k < 4; k++) {
    rec.off[k] = 0xFFFFFFFFu;
    rec.len[k] = 0;
  }
  const uint8_t* buf = p.rows;
  uint32_t pos, cap_e;
  if (row_sections(buf, s, e, &pos, &cap_e)) {
    while (pos < cap_e) {
      Entry en = read_entry(buf, pos, cap_e);
      if (!en.ok) break;
      int k = -1;
      if (en.kl == 14 && en.k01 == pack4("nv", 0, 2) && key_tail_eq<14>(buf, en.kp, "nvidia.com/gpu")) k = 0;
      if (en.kl == 11 && en.k01 == pack4("am", 0, 2) && key_tail_eq<11>(buf, en.kp, "amd.com/gpu")) k = 1;
      if (en.kl == 18 && en.k01 == pack4("gp", 0, 2) && key_tail_eq<18>(buf, en.kp, "gpu.intel.com/i915")) k = 2;
      if (en.kl == 13 && en.k01 == pack4("in", 0, 2) && key_tail_eq<13>(buf, en.kp, "intel.com/gpu")) k = 3;
      uint32_t vp = en.kp + en.kl;
      if (k >= 0) {
        long long val;
        int st = en.vl ? parse_quantity_general(buf + vp, en.vl, p.max_digits, &val) : Q_ABSENT;
        rec.off[k] = st == Q_ABSENT ? 0xFFFFFFFFu : vp;
        rec.len[k] = st == Q_ABSENT ? 0u : en.vl;
      }
      pos = vp + en.vl;
    }
  }
  p.defer[slot] = rec;
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
// The kernel.  TN nodes per tile; 2*TN walker threads + one producer warp.
// ------------------------------------------------------------------------------------------------
template <int TN, int STAGES, int STAGE_BYTES>
struct Smem {
  alignas(128) uint8_t stage[STAGES][STAGE_BYTES + 16];
  alignas(16) uint32_t rp[STAGES][TN + 8];
  alignas(8) uint64_t full[STAGES];
  uint64_t empty[STAGES];
  unsigned long long acc[GNC_N_COUNTS];
  int tile[STAGES];
  uint32_t lo[STAGES];
  uint32_t hi[STAGES];          // rows of the tile must end at or before this byte
  uint32_t direct[STAGES];      // 1: tile too large to stage, walkers read global memory
  uint32_t cond_bits[2][TN / 32];
  uint32_t cond_bad[2][TN / 32];
  uint32_t warp_cnt[TN / 32];
  uint32_t tile_base;
};

template <int TN, int STAGES, int STAGE_BYTES, int MIN_CTAS>
__global__ void __launch_bounds__(2 * TN + 32, MIN_CTAS) classify_kernel(const __grid_constant__ Params p) {
  using S = Smem<TN, STAGES, STAGE_BYTES>;
  extern __shared__ __align__(128) uint8_t smem_raw[];
  S& sm = *reinterpret_cast<S*>(smem_raw);
  constexpr int kWalkers = 2 * TN;
  constexpr int kCapWarps = TN / 32;
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], 1);
    }
    mbar_init_fence();
  }
  if (tid < GNC_N_COUNTS) sm.acc[tid] = 0;
  __syncthreads();

  if (warp == kWalkers / 32) {
    // ------------------------------ producer ------------------------------
    if (lane == 0) {
      for (uint32_t it = 0;; ++it) {
        const int s = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1u;
        const uint32_t tile = atomicAdd(&p.scratch->ticket, 1u);
        if (tile >= p.num_tiles) {
          mbar_wait(&sm.empty[s], ph ^ 1u);
          sm.tile[s] = -1;
          mbar_arrive(&sm.full[s]);
          break;
        }
        const uint32_t first = tile * TN;
        const uint32_t cnt = min((uint32_t)TN, p.n_nodes - first);
        const uint32_t lo = __ldg(p.row_ptr + first), hi = __ldg(p.row_ptr + first + cnt);
        const uint32_t lo_al = lo & ~15u;
        const uint32_t bytes = ((hi + 15u) & ~15u) - lo_al;
        const uint32_t rp_bytes = ((cnt + 1) * 4u + 15u) & ~15u;
        const bool direct = hi < lo || hi > p.rows_bytes || bytes > (uint32_t)STAGE_BYTES;
        mbar_wait(&sm.empty[s], ph ^ 1u);
        sm.tile[s] = (int)tile;
        sm.lo[s] = lo_al;
        sm.hi[s] = direct ? (uint32_t)min(p.rows_bytes, (uint64_t)0xFFFFFFFFu) : hi;
        sm.direct[s] = direct ? 1u : 0u;
        const uint32_t stage_bytes = direct ? 0u : bytes;
        mbar_arrive_expect_tx(&sm.full[s], rp_bytes + stage_bytes);
        bulk_g2s(sm.rp[s], p.row_ptr + first, rp_bytes, &sm.full[s]);
        if (stage_bytes) bulk_g2s(sm.stage[s], p.rows + lo_al, stage_bytes, &sm.full[s]);
      }
    }
    __syncwarp();  // lanes 1..31 rejoin lane 0 before the CTA-wide barrier below
  } else {
    // ------------------------------ walkers -------------------------------
    const bool cap_role = warp < kCapWarps;
    const int j = cap_role ? tid : tid - TN;  // node within the tile
    for (uint32_t it = 0;; ++it) {
      const int s = it % STAGES;
      const uint32_t ph = (it / STAGES) & 1u;
      mbar_wait(&sm.full[s], ph);
      const int tile = sm.tile[s];
      if (tile < 0) break;
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
        bad_ptr = re < rs || re > sm.hi[s] || (!direct && rs < sm.rp[s][0]);
        if (bad_ptr) rs = re = (direct ? 0u : sm.rp[s][0]);
      }
      CapResult cr;
      cr.v[0] = cr.v[1] = cr.v[2] = cr.v[3] = 0;
      cr.present = cr.big = 0;
      cr.malformed = false;
      if (cap_role) {
        if (live) {
          if (!direct) walk_capacity(sm.stage[s], rs - lo_al, re - lo_al, p.max_digits, cr);
          else walk_capacity(p.rows, rs, re, p.max_digits, cr);
        }
      } else {
        uint32_t c = 0;
        if (live) c = direct ? walk_conditions(p.rows, rs, re) : walk_conditions(sm.stage[s], rs - lo_al, re - lo_al);
        const uint32_t rb = __ballot_sync(0xFFFFFFFFu, c & 1u), bb = __ballot_sync(0xFFFFFFFFu, c & 2u);
        if (lane == 0) {
          sm.cond_bits[it & 1][warp - kCapWarps] = rb;
          sm.cond_bad[it & 1][warp - kCapWarps] = bb;
        }
      }
      named_bar_sync(1, kWalkers);  // every walker is done with stage s; condition bits are visible
      if (tid == 0) mbar_arrive(&sm.empty[s]);
      if (!cap_role) continue;

      // -------- capacity threads: flags, ordered compaction, counts --------
      const bool ready = (sm.cond_bits[it & 1][warp] >> lane) & 1u;
      const bool malformed = cr.malformed || bad_ptr || ((sm.cond_bad[it & 1][warp] >> lane) & 1u);
      const long long total = cr.v[0] + cr.v[1] + cr.v[2] + cr.v[3];  // |v| < 10^18 each
      const bool deferred = cr.big != 0;
      const bool is_gpu = live && !deferred && total > 0;
      uint32_t f = (is_gpu ? GNC_F_GPU : 0u) | (ready ? GNC_F_READY : 0u) | (deferred ? GNC_F_DEFERRED : 0u) |
                   (cr.present << 3) | (malformed ? GNC_F_MALFORMED : 0u);
      const uint32_t gb = __ballot_sync(0xFFFFFFFFu, is_gpu);
      const uint32_t rank = __popc(gb & ((1u << lane) - 1u));
      if (lane == 0) sm.warp_cnt[warp] = __popc(gb);
      named_bar_sync(2, TN);
      if (warp == 0) {
        uint32_t c = lane < kCapWarps ? sm.warp_cnt[lane] : 0u, incl = c;
#pragma unroll
        for (int d = 1; d < kCapWarps; d <<= 1) {
          uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
          if (lane >= d) incl += t;
        }
        const uint32_t tile_total = __shfl_sync(0xFFFFFFFFu, incl, kCapWarps - 1);
        if (lane < kCapWarps) sm.warp_cnt[lane] = incl - c;  // exclusive offset of each warp
        // decoupled look-back over earlier tiles' descriptors
        const unsigned long long tag = (unsigned long long)(p.epoch & kEpochMask) << 32;
        uint32_t excl = 0;
        if (tile == 0) {
          if (lane == 0) st_volatile_u64(p.tile_state, kStatusInclusive | tag | tile_total);
        } else {
          if (lane == 0) st_volatile_u64(p.tile_state + tile, kStatusAggregate | tag | tile_total);
          int look = tile - 1;
          while (true) {
            const int t = look - lane;
            unsigned long long d = kStatusInclusive | tag;  // tiles before 0: inclusive prefix 0
            if (t >= 0) {
              do {
                d = ld_volatile_u64(p.tile_state + t);
              } while ((d >> 62) == 0 || ((d >> 32) & kEpochMask) != (p.epoch & kEpochMask));
            }
            const uint32_t incl_mask = __ballot_sync(0xFFFFFFFFu, (d >> 62) == 2);
            const int stop = incl_mask ? __ffs(incl_mask) - 1 : 31;  // nearest inclusive descriptor
            uint32_t v = lane <= stop ? (uint32_t)d : 0u;
            v = __reduce_add_sync(0xFFFFFFFFu, v);
            excl += v;
            if (incl_mask) break;
            look -= 32;
          }
          if (lane == 0) st_volatile_u64(p.tile_state + tile, kStatusInclusive | tag | (excl + tile_total));
        }
        if (lane == 0) {
          sm.tile_base = excl;
          sm.acc[GNC_C_GPU_NODES] += tile_total;
        }
      }
      named_bar_sync(2, TN);
      const uint32_t out_pos = sm.tile_base + sm.warp_cnt[warp] + rank;
      if (live) p.flags[first + j] = (uint8_t)f;
      bool dropped = false;
      if (is_gpu) {
        if (out_pos < p.out_cap) {
          p.gpu_idx[out_pos] = p.node_base + first + j;
          p.gpu_total[out_pos] = total;
          longlong2* vo = reinterpret_cast<longlong2*>(p.gpu_vals + 4ull * out_pos);
          vo[0] = make_longlong2(cr.v[0], cr.v[1]);
          vo[1] = make_longlong2(cr.v[2], cr.v[3]);
        } else {
          dropped = true;
        }
      }
      // counts: warp-reduced, then a handful of shared-memory atomics per warp
      const uint32_t n_ready = __popc(__ballot_sync(0xFFFFFFFFu, is_gpu && ready));
      const uint32_t def_b = __ballot_sync(0xFFFFFFFFu, live && deferred);
      const uint32_t bad_b = __ballot_sync(0xFFFFFFFFu, live && malformed);
      const uint32_t drop_b = __ballot_sync(0xFFFFFFFFu, dropped);
      // 32-bit warp sums are exact while every value is within +-2^26
      const bool wide = is_gpu && ((unsigned long long)(cr.v[0] + (1 << 26)) >> 27 | (unsigned long long)(cr.v[1] + (1 << 26)) >> 27 |
                                   (unsigned long long)(cr.v[2] + (1 << 26)) >> 27 | (unsigned long long)(cr.v[3] + (1 << 26)) >> 27) != 0;
      if (!__any_sync(0xFFFFFFFFu, wide)) {
        int s0 = __reduce_add_sync(0xFFFFFFFFu, is_gpu ? (int)cr.v[0] : 0);
        int s1 = __reduce_add_sync(0xFFFFFFFFu, is_gpu ? (int)cr.v[1] : 0);
        int s2 = __reduce_add_sync(0xFFFFFFFFu, is_gpu ? (int)cr.v[2] : 0);
        int s3 = __reduce_add_sync(0xFFFFFFFFu, is_gpu ? (int)cr.v[3] : 0);
        if (lane == 0) {
          acc128_add(&sm.acc[GNC_C_KEY_LO + 0], &sm.acc[GNC_C_KEY_HI + 0], s0);
          acc128_add(&sm.acc[GNC_C_KEY_LO + 1], &sm.acc[GNC_C_KEY_HI + 1], s1);
          acc128_add(&sm.acc[GNC_C_KEY_LO + 2], &sm.acc[GNC_C_KEY_HI + 2], s2);
          acc128_add(&sm.acc[GNC_C_KEY_LO + 3], &sm.acc[GNC_C_KEY_HI + 3], s3);
        }
      } else if (is_gpu) {
#pragma unroll
        for (int k = 0; k < 4; k++) acc128_add(&sm.acc[GNC_C_KEY_LO + k], &sm.acc[GNC_C_KEY_HI + k], cr.v[k]);
      }
      if (lane == 0) {
        if (n_ready) atomicAdd(&sm.acc[GNC_C_READY_NODES], (unsigned long long)n_ready);
        if (def_b) atomicAdd(&sm.acc[GNC_C_DEFERRED], (unsigned long long)__popc(def_b));
        if (bad_b) atomicAdd(&sm.acc[GNC_C_MALFORMED], (unsigned long long)__popc(bad_b));
        if (drop_b) atomicAdd(&sm.acc[GNC_C_OUT_DROPPED], (unsigned long long)__popc(drop_b));
      }
      if (live && deferred) {
        const uint32_t slot = atomicAdd(&p.scratch->defer_count, 1u);
        if (slot < p.defer_cap) emit_defer(p, first + j, rs, re, slot);
        else atomicAdd(&sm.acc[GNC_C_DEFER_DROPPED], 1ull);
      }
    }
  }

  // ------------------------------ grid-wide finish ------------------------------
  __syncthreads();
  if (tid == 0) {
    ScratchHeader* sc = p.scratch;
#pragma unroll
    for (int c = 0; c < GNC_N_COUNTS; c++) {
      if (c >= GNC_C_KEY_LO && c < GNC_C_KEY_HI) {  // lo with carry into hi
        unsigned long long v = sm.acc[c];
        if (v) {
          unsigned long long old = atomicAdd(&sc->accum[c], v);
          if (old + v < old) atomicAdd(&sc->accum[c + 4], 1ull);
        }
      } else if (sm.acc[c]) {
        atomicAdd(&sc->accum[c], sm.acc[c]);
      }
    }
    __threadfence();
    const uint32_t prev = atomicAdd(&sc->done, 1u);
    if (prev == gridDim.x - 1) {  // last CTA: publish the counts block and re-arm the scratch
      __threadfence();
      for (int c = 0; c < GNC_N_COUNTS; c++) {
        long long v = (long long)ld_volatile_u64(&sc->accum[c]);
        if (c == GNC_C_SCANNED) v = p.n_nodes;
        p.counts[c] = v;
        sc->accum[c] = 0;
      }
      sc->ticket = 0;
      sc->defer_count = 0;
      __threadfence();
      sc->done = 0;
    }
  }
}

}  // namespace gnc
