// json_scan.cuh -- raw /api/v1/nodes JSON bytes -> PNL-v2 columns, on the device (sm_100a).
//
// Replaces, for the classification path, the third-party deserialisation behind
// `api.list_node().items` (/root/reference/check-gpu-node.py:217): only the parts of each node the
// reference reads are extracted -- `status.capacity` entries (check-gpu-node.py:183-187) and the
// `type` / `status` of each element of `status.conditions` (:173-177) -- with json.loads semantics
// (duplicate object keys: last wins; \uXXXX escapes, surrogate pairs joined, lone surrogates kept).
//
// Structure: a structural index first (bit masks, one bit per byte), then everything walks the masks.
//   1. index:   one pass over the document (index_kernel).  32 KB tiles staged through shared memory, a
//               256-byte chunk per thread; per 32-byte group the quote and backslash masks by SWAR byte
//               compares; each chunk's effect on the lexer state {escape pending, inside string} as a
//               transition function, scanned inside the CTA and carried from tile to tile by decoupled
//               look-back.  With the start state known: escaped bytes, unescaped quotes, in-string mask
//               (prefix xor), and per group the kind of every structural byte -- string open, string
//               close, bracket open / close and comma outside strings -- as three bit planes, plus the
//               backslashes inside strings (16 bytes per 32 document bytes), plus every
//               chunk's start depth relative to its tile; a small kernel scans the tiles' bracket balances
//   2. events:  element separators of root-level arrays (the '[' itself and each ',' at depth 2) are
//               counted per chunk from the masks and scanned; root-level array brackets are listed; one
//               thread then picks the LAST root key "items" and the range of events that are its elements
//   3. nodes:   one thread per element, whole warps in step.  The fast walker steps from structural bit
//               to structural bit (find-first-set over the masks; containers that do not matter are
//               skipped a group at a time by popcount) and touches document bytes only to compare the few
//               keys that matter.  Anything unusual in the parts that matter -- an escape in a key or
//               value, a duplicate of a key that matters, a non-string quantity -- sends that node to the
//               general byte-serial parser below, which implements the full semantics.  Two passes:
//               measure (notes where every row entry's text is), scan of row sizes, write (warps gather
//               the noted texts together, rows are assembled in shared memory and copied out coalesced).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace gnc {
namespace js {

constexpr int kChunk = 256;            // bytes per thread in the chunk passes
constexpr int kGroup = 32;             // bytes per mask word
constexpr int kGroupsPerChunk = kChunk / kGroup;
constexpr int kTileChunks = 128;       // chunks per CTA in the byte passes (one per thread)
constexpr int kRowStride = kChunk + 16;  // shared-memory row pitch: 16-byte reads by 8 adjacent threads hit 8 different bank quads
constexpr int kMaxRootArrays = 64;

enum : int { INFO_NODES = 0, INFO_ROWS_BYTES = 1, INFO_STATUS = 2, INFO_ITEMS_BEGIN = 3, INFO_ITEMS_END = 4, INFO_GENERAL_NODES = 5, INFO_N = 8 };
enum : int { ST_NO_ITEMS = 1, ST_NODES_CAP = 2, ST_ROWS_CAP = 4, ST_BAD_JSON = 8, ST_TOO_MANY_ROOT_ARRAYS = 16 };

struct Work {                 // device scratch, one per call
  unsigned long long root_open[kMaxRootArrays];   // positions of '[' opening a root-level array
  unsigned long long root_close[kMaxRootArrays];  // positions of the ']' closing one
  unsigned int n_open, n_close;
  unsigned long long items_begin, items_end;      // byte positions of the chosen '[' and its ']'
  unsigned long long ev_base;                     // rank of the event at items_begin
  unsigned int n_nodes;
  unsigned int status;
  unsigned int general_nodes;                     // nodes that went through the byte-serial parser
  unsigned int n_listed;                          // chunks the events kernel has to walk
};

// The structural index: one 16-byte record per 32 document bytes (group g -> w[g]); bit i of a word = byte 32 g + i.
//   { p0, p1, p2, backslashes inside strings }: p2 p1 p0 at a bit spell the kind of structural byte there --
//   1 opening quote, 2 closing quote, 3 '{' '[', 4 '}' ']', 5 ','  (brackets and commas outside strings only); 0: none.
// Kinds as bit planes rather than one mask per kind: the walkers' "next structural byte" is one OR of three
// words and three bit tests instead of an OR of five and a five-way select, and the index is half the bytes.
struct Masks {
  uint4* w;
};
__device__ __forceinline__ uint32_t plane_open(const uint4 a) { return a.x & a.y; }          // kind 3
__device__ __forceinline__ uint32_t plane_close(const uint4 a) { return a.z & ~a.x; }        // kind 4
__device__ __forceinline__ uint32_t plane_comma(const uint4 a) { return a.z & a.x; }         // kind 5
__device__ __forceinline__ uint32_t plane_str_open(const uint4 a) { return a.x & ~a.y & ~a.z; }  // kind 1

// Depth (brackets open outside strings) at the start of a chunk = depth inside its tile + the tile's start depth.
struct Depths {
  const int* local;        // [chunks] relative to the start of the chunk's tile
  const int* tile_start;   // [tiles]
  __device__ __forceinline__ int at(unsigned long long c) const { return __ldg(tile_start + c / kTileChunks) + __ldg(local + c); }
};

// ---- lexer state ---------------------------------------------------------------------------------
// State = e | s << 1: e = the first byte of the chunk is escaped by a backslash before it, s = the chunk
// starts inside a string.  A chunk's transition function holds 2 bits per start state.  (Backslashes
// are taken as escapes wherever they stand; in valid JSON they only occur inside strings.)
__device__ __forceinline__ uint32_t compose(uint32_t f, uint32_t g) {  // first f, then g
  return ((g >> (2 * (f & 3))) & 3) | (((g >> (2 * ((f >> 2) & 3))) & 3) << 2) | (((g >> (2 * ((f >> 4) & 3))) & 3) << 4) |
         (((g >> (2 * ((f >> 6) & 3))) & 3) << 6);
}
constexpr uint32_t kIdentity = 0 | (1 << 2) | (2 << 4) | (3 << 6);

// ---- SWAR byte compares ----------------------------------------------------------------------------
// 0x80 in every byte of w that equals the byte replicated in k4 (exact: no borrow between bytes)
__device__ __forceinline__ uint32_t eq_bytes(uint32_t w, uint32_t k4) {
  const uint32_t x = w ^ k4;
  const uint32_t t = (x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;
  return ~(t | x) & 0x80808080u;
}
// the four 0x80 flags of z as a nibble, byte 0 in bit 0 (the partial products never collide)
__device__ __forceinline__ uint32_t flags_to_nibble(uint32_t z) { return (z * 0x00204081u) >> 28; }
// the same nibble left in the top four bits (the other bits cleared)
__device__ __forceinline__ uint32_t flags_to_top_nibble(uint32_t z) { return (z * 0x00204081u) & 0xF0000000u; }

struct RawMasks {
  uint32_t q, b, o, c, k;
};
// quotes and backslashes of one 32-byte group (all the lexer needs).  The word loops are kept rolled: fully
// unrolled, the kernel's code no longer fits the instruction cache (15 % of its samples were instruction fetch).
__device__ __forceinline__ void raw_quotes(const uint32_t* __restrict__ w, RawMasks& m) {
  uint32_t q = 0, b = 0;
#pragma unroll 2
  for (int i = 0; i < 8; i++) {
    const uint32_t x = w[i];
    q = (q >> 4) | flags_to_top_nibble(eq_bytes(x, 0x22222222u));  // after eight words the first word's nibble is at bit 0
    b = (b >> 4) | flags_to_top_nibble(eq_bytes(x, 0x5C5C5C5Cu));
  }
  m.q = q, m.b = b;
}
// brackets and commas of one group
__device__ __forceinline__ void raw_brackets(const uint32_t* __restrict__ w, RawMasks& m) {
  uint32_t o = 0, c = 0, k = 0;
#pragma unroll 2
  for (int i = 0; i < 8; i++) {
    const uint32_t x = w[i], folded = x & 0xDFDFDFDFu;  // '{' -> '[', '}' -> ']' (nothing else lands on them)
    o = (o >> 4) | flags_to_top_nibble(eq_bytes(folded, 0x5B5B5B5Bu));
    c = (c >> 4) | flags_to_top_nibble(eq_bytes(folded, 0x5D5D5D5Du));
    k = (k >> 4) | flags_to_top_nibble(eq_bytes(x, 0x2C2C2C2Cu));
  }
  m.o = o, m.c = c, m.k = k;
}

// Bytes escaped by a backslash, given the backslash mask of a group and whether its first byte is
// escaped from the group before.  An unescaped backslash escapes the byte after it, whatever that is;
// pe becomes whether the next group's first byte is escaped.  One iteration per escape sequence.
__device__ __forceinline__ uint32_t escaped_bytes(uint32_t backslash, uint32_t& pe) {
  if ((backslash | pe) == 0) return 0;
  uint32_t esc = pe, rem = backslash & ~pe;
  pe = 0;
  while (rem) {
    const int p = __ffs(rem) - 1;
    if (p == 31) {
      pe = 1;
      break;
    }
    esc |= 2u << p;
    rem &= ~(3u << p);
  }
  return esc;
}
__device__ __forceinline__ uint32_t prefix_xor(uint32_t x) {
  x ^= x << 1;
  x ^= x << 2;
  x ^= x << 4;
  x ^= x << 8;
  x ^= x << 16;
  return x;
}

// Stage one tile of the document (kTileChunks chunks of 256 B) into shared memory, coalesced; bytes past
// the end of the document read as spaces.
__device__ __forceinline__ void stage_tile(const uint8_t* __restrict__ json, unsigned long long n, unsigned long long tile_chunk0,
                                           uint8_t* sh) {
  const unsigned long long base = tile_chunk0 * kChunk;
  constexpr int kVecs = kTileChunks * kChunk / 16, kPer = kVecs / kTileChunks;  // 16-byte pieces: per tile, per thread
  if (base + (unsigned long long)kTileChunks * kChunk <= n) {
    // a whole tile: every load of the thread is in flight before the first store
    uint4 v[kPer];
#pragma unroll
    for (int r = 0; r < kPer; r++) v[r] = __ldg(reinterpret_cast<const uint4*>(json + base) + threadIdx.x + r * kTileChunks);
#pragma unroll
    for (int r = 0; r < kPer; r++) {
      const int i = threadIdx.x + r * kTileChunks;
      *reinterpret_cast<uint4*>(sh + (i >> 4) * kRowStride + (i & 15) * 16) = v[r];
    }
    return;
  }
  for (int i = threadIdx.x; i < kVecs; i += kTileChunks) {
    const unsigned long long at = base + 16ull * i;
    uint4 v = make_uint4(0x20202020u, 0x20202020u, 0x20202020u, 0x20202020u);
    if (at + 16 <= n) {
      v = __ldg(reinterpret_cast<const uint4*>(json + at));
    } else if (at < n) {
      uint8_t tmp[16];
      for (int k = 0; k < 16; k++) tmp[k] = at + k < n ? json[at + k] : (uint8_t)0x20;
      memcpy(&v, tmp, 16);
    }
    *reinterpret_cast<uint4*>(sh + (i >> 4) * kRowStride + (i & 15) * 16) = v;
  }
}

// ---- one pass over the document: the lexer state travels from tile to tile by decoupled look-back --------------
// Tiles are taken in ticket order, so every earlier tile is already running.  A tile publishes what it
// adds (its transition function) as soon as it knows it, then looks back over the tiles before it -- 128
// at a time, 4 per lane -- to the nearest one that has published its full prefix, and publishes its own.
// State word: flag << 32 | payload, flag 1 = this tile alone, 2 = all tiles up to and including this one.
// (The bracket depth was first carried the same way, by a second look-back after the masks: nothing was
// left to hide that wait behind, and a third of the kernel's samples sat at its barrier.  Depths are now
// written relative to the tile and a small kernel scans the tiles' balances.)
constexpr unsigned long long kFlagAgg = 1ull << 32, kFlagInc = 2ull << 32;
constexpr int kLookWindow = 128;

__device__ __forceinline__ void ld_volatile_2x_u64x2(const unsigned long long* p, unsigned long long (&w)[4]) {
  asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[0]), "=l"(w[1]) : "l"(p));
  asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[2]), "=l"(w[3]) : "l"(p + 2));
}
struct ComposeMonoid {  // payload: a transition function; order matters
  static __device__ __forceinline__ uint32_t identity() { return kIdentity; }
  static __device__ __forceinline__ uint32_t combine(uint32_t first, uint32_t then) { return compose(first, then); }
};
// What tile t adds, visible to the tiles after it at once (one thread).
__device__ __forceinline__ void tile_publish(unsigned long long* st, int t, uint32_t mine) {
  *reinterpret_cast<volatile unsigned long long*>(st + t) = kFlagAgg | mine;
}
// One full warp, some time after tile_publish (whatever the tile can do in between hides the wait).  Returns
// the combination of tiles [0, t); st[t] goes from "alone" to "prefix".
template <class M>
__device__ uint32_t tile_resolve(unsigned long long* st, int t, uint32_t mine, int lane) {
  uint32_t acc = M::identity();  // tiles [hi, t)
  int hi = t;
  while (hi > 0) {
    const int w0 = (hi - 1) & ~(kLookWindow - 1);  // the aligned window holding tile hi - 1
    const int base = w0 + 4 * lane;
    unsigned long long w[4];
    bool ok, has_inc;
    int top;
    while (true) {  // only flags are looked at while waiting; the payloads are combined once, afterwards
      w[0] = w[1] = w[2] = w[3] = 0;
      if (base < hi) ld_volatile_2x_u64x2(st + base, w);  // the state array is padded to a whole window
      ok = true, has_inc = false;
#pragma unroll
      for (int k = 3; k >= 0; --k) {  // nearest tile first
        if (base + k < hi && !has_inc) {
          const uint32_t flag = (uint32_t)(w[k] >> 32);
          ok = ok && flag != 0;
          has_inc = flag == 2u;
        }
      }
      const uint32_t m = __ballot_sync(0xFFFFFFFFu, has_inc && ok);
      top = m ? 31 - __clz(m) : -1;  // the nearest tile that knows its full prefix
      if (__all_sync(0xFFFFFFFFu, lane < top || ok)) break;
      __nanosleep(64);
    }
    uint32_t part = M::identity();
    {
      bool stop = false;
#pragma unroll
      for (int k = 3; k >= 0; --k) {
        if (base + k < hi && !stop) {
          part = M::combine((uint32_t)w[k], part);
          stop = (uint32_t)(w[k] >> 32) == 2u;
        }
      }
    }
    if (lane < top) part = M::identity();
    // ordered combination over the lanes: lower lanes hold earlier tiles
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t other = __shfl_down_sync(0xFFFFFFFFu, part, d);
      if (lane + d < 32) part = M::combine(part, other);
    }
    acc = M::combine(__shfl_sync(0xFFFFFFFFu, part, 0), acc);
    if (top >= 0) break;
    hi = w0;
  }
  if (lane == 0) *reinterpret_cast<volatile unsigned long long*>(st + t) = kFlagInc | M::combine(acc, mine);
  return acc;
}

// The index record of every group, every chunk's depth at its start (relative to the tile) and a lower bound of the depth inside it.
__global__ void __launch_bounds__(kTileChunks) index_kernel(const uint8_t* __restrict__ json, unsigned long long n, Masks mk,
                                                            int* __restrict__ depth_local, int* __restrict__ tile_bal,
                                                            int* __restrict__ min_rel, unsigned long long* lex_state,
                                                            unsigned int* ticket, unsigned long long n_chunks) {
  extern __shared__ __align__(16) uint8_t sh[];
  __shared__ uint32_t s_tile, s_warp[kTileChunks / 32], s_prefix;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
  __syncthreads();
  const int tile = (int)s_tile;
  const unsigned long long chunk0 = (unsigned long long)tile * kTileChunks;
  stage_tile(json, n, chunk0, sh);
  __syncthreads();
  const unsigned long long c = chunk0 + threadIdx.x;
  const bool have = c < n_chunks;
  const uint4* row = reinterpret_cast<const uint4*>(sh + threadIdx.x * kRowStride);
  RawMasks m[kGroupsPerChunk];
  uint32_t func = kIdentity;
  if (have) {
    uint32_t pe0 = 0, pe1 = 1, par0 = 0, par1 = 0;  // escape carry and quote parity, for e = 0 and e = 1 at the chunk start
#pragma unroll
    for (int g = 0; g < kGroupsPerChunk; g++) {
      raw_quotes(reinterpret_cast<const uint32_t*>(row + 2 * g), m[g]);
      const bool same = pe0 == pe1;
      const uint32_t q0 = m[g].q & ~escaped_bytes(m[g].b, pe0);
      par0 ^= __popc(q0) & 1;
      if (same) {
        pe1 = pe0;
        par1 ^= __popc(q0) & 1;
      } else {
        const uint32_t q1 = m[g].q & ~escaped_bytes(m[g].b, pe1);
        par1 ^= __popc(q1) & 1;
      }
    }
    // f(e | s << 1) = pe_e | (s ^ par_e) << 1
    const uint32_t f0 = pe0 | (par0 << 1), f1 = pe1 | (par1 << 1), f2 = pe0 | ((par0 ^ 1u) << 1), f3 = pe1 | ((par1 ^ 1u) << 1);
    func = f0 | (f1 << 2) | (f2 << 4) | (f3 << 6);
  }
  // ---- lexer state at the start of my chunk: ordered scan inside the CTA, look-back across tiles ----
  uint32_t inc = func;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t before = __shfl_up_sync(0xFFFFFFFFu, inc, d);
    if (lane >= d) inc = compose(before, inc);
  }
  uint32_t excl = __shfl_up_sync(0xFFFFFFFFu, inc, 1);
  if (lane == 0) excl = kIdentity;
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  uint32_t tile_func = kIdentity;
#pragma unroll
  for (int q = 0; q < kTileChunks / 32; q++) {
    if (q == warp) excl = compose(tile_func, excl);  // the warps before mine, then the lanes before me
    tile_func = compose(tile_func, s_warp[q]);
  }
  if (threadIdx.x == 0) tile_publish(lex_state, tile, tile_func);
  // the brackets and commas do not depend on the lexer state: found while the tiles before this one catch up
  if (have) {
#pragma unroll
    for (int g = 0; g < kGroupsPerChunk; g++) raw_brackets(reinterpret_cast<const uint32_t*>(row + 2 * g), m[g]);
  }
  if (warp == 0) {
    const uint32_t before = tile_resolve<ComposeMonoid>(lex_state, tile, tile_func, lane);
    if (lane == 0) s_prefix = before;
  }
  __syncthreads();
  const uint32_t st = compose(s_prefix, excl) & 3u;  // (every earlier chunk, in order)(state 0)
  // ---- the masks ----
  uint32_t pe = st & 1u, in_str = (st >> 1) ? 0xFFFFFFFFu : 0u;
  int bal = 0, low = 0;
  if (have) {
    uint4* out = mk.w + c * kGroupsPerChunk;
#pragma unroll
    for (int g = 0; g < kGroupsPerChunk; g++) {
      const uint32_t q = m[g].q & ~escaped_bytes(m[g].b, pe);
      const uint32_t s = prefix_xor(q) ^ in_str;  // bit i: byte i is inside a string (opening quote included, closing quote not)
      in_str = (s >> 31) ? 0xFFFFFFFFu : 0u;
      const uint32_t o = m[g].o & ~s, cl = m[g].c & ~s;
      const uint32_t so = q & s, sc = q & ~s, k = m[g].k & ~s;
      out[g] = make_uint4(so | o | k, sc | o, cl | k, m[g].b & s);
      low = min(low, bal - __popc(cl));  // as if every close of the group came first
      bal += __popc(o) - __popc(cl);
    }
  }
  // ---- depth at the start of my chunk, relative to the tile; the tiles' balances are scanned by a small kernel
  //      afterwards (a second look-back here would only add a wait at the end of every tile) ----
  int dinc = bal;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int before = __shfl_up_sync(0xFFFFFFFFu, dinc, d);
    if (lane >= d) dinc += before;
  }
  __syncthreads();  // s_warp is free again
  if (lane == 31) s_warp[warp] = (uint32_t)dinc;
  __syncthreads();
  int dexcl = dinc - bal, tile_total = 0;
#pragma unroll
  for (int q = 0; q < kTileChunks / 32; q++) {
    if (q == warp) dexcl += tile_total;
    tile_total += (int)s_warp[q];
  }
  if (threadIdx.x == 0) tile_bal[tile] = tile_total;
  if (have) {
    depth_local[c] = dexcl;
    min_rel[c] = low;
  }
}

// ---- generic exclusive scan over n items, 1024 per block: block totals, scan of totals, apply ---------
// In-order scan of one value per thread over a 1024-thread CTA: shuffles inside the warps, the 32 warp totals
// through shared memory (sh[32]).  Returns the inclusive value; `excl` gets everything before this thread.
template <class T, class Op>
__device__ __forceinline__ T block_scan_1024(T v, T* sh, Op op, T ident, T& excl) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const T t = __shfl_up_sync(0xFFFFFFFFu, v, d);
    if (lane >= d) v = op(t, v);
  }
  T before = __shfl_up_sync(0xFFFFFFFFu, v, 1);  // inclusive value of the lane before mine
  if (lane == 0) before = ident;
  if (lane == 31) sh[warp] = v;
  __syncthreads();
  if (warp == 0) {
    T w = sh[lane];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const T t = __shfl_up_sync(0xFFFFFFFFu, w, d);
      if (lane >= d) w = op(t, w);
    }
    sh[lane] = w;
  }
  __syncthreads();
  const T warps_before = warp ? sh[warp - 1] : ident;
  excl = op(warps_before, before);
  return op(warps_before, v);
}
template <class T, class Op>
__global__ void scan_block_totals(const T* __restrict__ in, unsigned long long n, T* __restrict__ totals, Op op, T ident) {
  __shared__ T sh[32];
  const unsigned long long i = blockIdx.x * 1024ull + threadIdx.x;
  T excl;
  const T incl = block_scan_1024<T, Op>(i < n ? in[i] : ident, sh, op, ident, excl);
  if (threadIdx.x == 1023) totals[blockIdx.x] = incl;
}
// one block: thread t owns a contiguous run of the totals (kept in order)
template <class T, class Op>
__global__ void scan_totals_block(T* totals, unsigned long long nb, Op op, T ident, T* grand) {
  __shared__ T sh[32];
  const unsigned long long per = (nb + 1023) / 1024, lo = min(nb, threadIdx.x * per), hi = min(nb, lo + per);
  T run = ident;
  for (unsigned long long b = lo; b < hi; b++) run = op(run, totals[b]);
  T before;
  const T incl = block_scan_1024<T, Op>(run, sh, op, ident, before);
  if (grand && threadIdx.x == 1023) *grand = incl;
  for (unsigned long long b = lo; b < hi; b++) {
    const T t = totals[b];
    totals[b] = before;
    before = op(before, t);
  }
}
template <class T, class Op>
__global__ void scan_apply(const T* __restrict__ in, unsigned long long n, const T* __restrict__ totals, T* __restrict__ out,
                           Op op, T ident) {
  __shared__ T sh[32];
  const unsigned long long i = blockIdx.x * 1024ull + threadIdx.x;
  T before;  // everything before me in the block
  block_scan_1024<T, Op>(i < n ? in[i] : ident, sh, op, ident, before);
  if (i < n) out[i] = op(totals[blockIdx.x], before);
}
struct OpAddI {
  __device__ int operator()(int a, int b) const { return a + b; }
};
struct OpAddU {
  __device__ unsigned int operator()(unsigned int a, unsigned int b) const { return a + b; }
};
struct OpAddULL {
  __device__ unsigned long long operator()(unsigned long long a, unsigned long long b) const { return a + b; }
};

// ---- events ------------------------------------------------------------------------------------------
// Visit, in order, the brackets and commas outside strings of chunk c that lie before `limit`:
// f(pos, kind, depth_before) with kind 0 open, 1 close, 2 comma.  Groups that stay deeper than the root
// levels are passed over by popcount.  Returns the depth after the last byte visited.
// (Stepping bracket to bracket and looking at commas only in stretches at depth 2 was slower: more
// branches for the same few bits.)
template <class F>
__device__ __forceinline__ int visit_structurals(const Masks& mk, unsigned long long c, int depth, unsigned long long limit, F&& f) {
  const uint4* pw = mk.w + c * kGroupsPerChunk;
  uint32_t ow[8], cw[8], kw[8];
#pragma unroll
  for (int g = 0; g < kGroupsPerChunk; g++) {
    const uint4 a = __ldg(pw + g);
    ow[g] = plane_open(a), cw[g] = plane_close(a), kw[g] = plane_comma(a);
  }
#pragma unroll
  for (int g = 0; g < kGroupsPerChunk; g++) {
    const unsigned long long base = (c * kGroupsPerChunk + g) * (unsigned long long)kGroup;
    if (base >= limit) break;
    uint32_t o = ow[g], cl = cw[g], k = kw[g];
    if (limit - base < 32) {
      const uint32_t keep = (1u << (unsigned)(limit - base)) - 1u;
      o &= keep, cl &= keep, k &= keep;
    }
    if (depth - __popc(cl) > 2) {
      depth += __popc(o) - __popc(cl);
      continue;
    }
    uint32_t any = o | cl | k;
    while (any) {
      const int bit = __ffs(any) - 1;
      const uint32_t bm = 1u << bit;
      any &= any - 1;
      if (o & bm) {
        f(base + bit, 0, depth);
        depth++;
      } else if (cl & bm) {
        f(base + bit, 1, depth);
        depth--;
      } else {
        f(base + bit, 2, depth);
      }
    }
  }
  return depth;
}

// events: the '[' of every root-level array and every ',' directly inside one (depth 2 outside strings).
// Two kernels: the first drops the chunks that lie deeper than the root levels throughout (four in five) and
// lists the others; the second walks the listed chunks, one per thread.  (One kernel doing both ran its
// walks at two active lanes per warp: the few chunks that matter were spread over all the warps.)
__global__ void events_select_kernel(Depths dp, const int* __restrict__ min_rel, unsigned int* __restrict__ ev_count,
                                     unsigned int* __restrict__ chunk_list, Work* w, unsigned long long n_chunks) {
  const unsigned long long c = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  const bool in = c < n_chunks;
  const bool keep = in && dp.at(c) + min_rel[c] <= 2;
  if (in && !keep) ev_count[c] = 0;
  // one atomic on the global counter per CTA (one per warp made 164k atomics on one address: 100 us)
  __shared__ unsigned int s_count, s_base;
  if (threadIdx.x == 0) s_count = 0;
  __syncthreads();
  const uint32_t m = __ballot_sync(0xFFFFFFFFu, keep);
  const int lane = threadIdx.x & 31;
  unsigned int at = 0;
  if (lane == 0 && m) at = atomicAdd(&s_count, (unsigned int)__popc(m));
  at = __shfl_sync(0xFFFFFFFFu, at, 0);
  __syncthreads();
  if (threadIdx.x == 0 && s_count) s_base = atomicAdd(&w->n_listed, s_count);
  __syncthreads();
  if (keep) chunk_list[s_base + at + __popc(m & ((1u << lane) - 1u))] = (unsigned int)c;
}

// Thread per listed chunk.  First every group's start depth by popcount (no branches), which also tells
// which groups can touch the root levels -- one or two of the eight; then a loop whose iteration is "my next
// such group", so that lanes whose interesting groups sit at different places in their chunks still walk them
// together (a loop over the eight groups with the bit walk inside ran each lane's walk on its own).
__global__ void events_kernel(const uint8_t* __restrict__ json, Masks mk, Depths dp, const unsigned int* __restrict__ chunk_list,
                              unsigned int* __restrict__ ev_count, unsigned int* __restrict__ ev_pack, Work* w) {
  const unsigned int n_listed = w->n_listed;
  for (unsigned int t0 = blockIdx.x * blockDim.x; t0 < n_listed; t0 += gridDim.x * blockDim.x) {  // whole warps stay in the loop
    const unsigned int t = t0 + threadIdx.x;
    const bool have = t < n_listed;
    const unsigned long long c = have ? chunk_list[t] : 0;
    const uint4* pw = mk.w + c * kGroupsPerChunk;
    const int d0 = have ? dp.at(c) : 0;
    // pass 1: depth at the start of each group, relative to d0, packed as eight int16; which groups come near the root
    uint32_t rel[4] = {0, 0, 0, 0}, near = 0;
    int bal = 0;
    if (have) {
#pragma unroll
      for (int g = 0; g < kGroupsPerChunk; g++) {
        const uint4 a = __ldg(pw + g);
        const int closes = __popc(plane_close(a));
        rel[g >> 1] |= ((uint32_t)bal & 0xFFFFu) << (16 * (g & 1));
        if (d0 + bal - closes <= 2) near |= 1u << g;
        bal += __popc(plane_open(a)) - closes;
      }
    }
    unsigned int cnt = 0, pack = 0;  // pack: where in the chunk the first four events are (element_starts_kernel reads it)
    const unsigned long long chunk_base = c * kChunk;
    while (__any_sync(0xFFFFFFFFu, near != 0)) {
      if (near == 0) continue;
      const int g = __ffs(near) - 1;
      near &= near - 1;
      const uint32_t r2 = g < 2 ? rel[0] : g < 4 ? rel[1] : g < 6 ? rel[2] : rel[3];
      int depth = d0 + (int)(short)((r2 >> (16 * (g & 1))) & 0xFFFFu);
      const uint4 a = __ldg(pw + g);
      const uint32_t o = plane_open(a), cl = plane_close(a), k = plane_comma(a);
      const unsigned long long base = chunk_base + (unsigned long long)g * kGroup;
      uint32_t any = o | cl | k;
      while (any) {
        const int bit = __ffs(any) - 1;
        const uint32_t bm = 1u << bit;
        const unsigned long long pos = base + bit;
        any &= any - 1;
        if (o & bm) {
          if (depth == 1 && json[pos] == '[') {
            if (cnt < 4) pack |= (unsigned int)(pos - chunk_base) << (8 * cnt);
            cnt++;
            const unsigned int q = atomicAdd(&w->n_open, 1u);
            if (q < kMaxRootArrays) w->root_open[q] = pos;
          }
          depth++;
        } else if (cl & bm) {
          if (depth == 2) {
            if (json[pos] == ']') {
              const unsigned int q = atomicAdd(&w->n_close, 1u);
              if (q < kMaxRootArrays) w->root_close[q] = pos;
            }
          } else if (depth <= 0) {
            atomicOr(&w->status, (unsigned int)ST_BAD_JSON);
          }
          depth--;
        } else if (depth == 2) {
          if (cnt < 4) pack |= (unsigned int)(pos - chunk_base) << (8 * cnt);
          cnt++;
        }
      }
    }
    if (have) {
      ev_count[c] = cnt;
      ev_pack[c] = pack;
    }
  }
}

// number of events strictly before byte `pos` (pos lies outside strings)
__device__ unsigned long long events_before(const uint8_t* json, const Masks& mk, const Depths& dp, const unsigned int* ev_prefix,
                                            unsigned long long pos) {
  const unsigned long long c = pos / kChunk;
  unsigned long long cnt = ev_prefix[c];
  visit_structurals(mk, c, dp.at(c), pos, [&](unsigned long long at, int kind, int d) {
    if ((kind == 0 && d == 1 && json[at] == '[') || (kind == 2 && d == 2)) cnt++;
  });
  return cnt;
}

// position of the byte after each element's separator
__global__ void element_starts_kernel(const uint8_t* __restrict__ json, Masks mk, Depths dp,
                                      const unsigned int* __restrict__ ev_count, const unsigned int* __restrict__ ev_pack,
                                      const unsigned int* __restrict__ ev_prefix, const Work* w,
                                      unsigned long long* __restrict__ elem_pos, unsigned long long n_chunks) {
  const unsigned long long c = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (c >= n_chunks) return;
  const unsigned int cnt = ev_count[c];
  if (cnt == 0) return;
  const unsigned long long lo = w->ev_base, hi = lo + w->n_nodes;
  unsigned long long rank = ev_prefix[c];
  if (rank >= hi || rank + cnt <= lo) return;  // no event of the items array here
  if (cnt <= 4) {  // the events kernel left their places behind
    const unsigned int pack = ev_pack[c];
    for (unsigned int e = 0; e < cnt; e++, rank++)
      if (rank >= lo && rank < hi) elem_pos[rank - lo] = c * kChunk + ((pack >> (8 * e)) & 0xFFu) + 1;
    return;
  }
  visit_structurals(mk, c, dp.at(c), ~0ull, [&](unsigned long long pos, int kind, int d) {
    if ((kind == 0 && d == 1 && json[pos] == '[') || (kind == 2 && d == 2)) {
      if (rank >= lo && rank < hi) elem_pos[rank - lo] = pos + 1;
      rank++;
    }
  });
}

// ---- general per-node parser (byte-serial; full json.loads semantics) ---------------------------------------------------------------------------------
struct Cur {
  const uint8_t* p;
  unsigned long long i, n;
  __device__ int peek() const { return i < n ? (int)p[i] : -1; }
  __device__ void ws() {
    while (i < n && (p[i] == ' ' || p[i] == '\t' || p[i] == '\n' || p[i] == '\r')) i++;
  }
};

__device__ __forceinline__ int hexv(int c) {
  if (c >= '0' && c <= '9') return c - '0';
  c |= 0x20;
  return (c >= 'a' && c <= 'f') ? c - 'a' + 10 : -1;
}

// Decoded bytes of a JSON string, one at a time. Cursor starts after the opening quote.
struct StrReader {
  Cur* c;
  uint32_t pend;   // up to 3 pending UTF-8 continuation bytes, low byte first
  int npend;
  bool done;
  __device__ StrReader(Cur* cur) : c(cur), pend(0), npend(0), done(false) {}
  __device__ int u4() {  // four hex digits at the cursor -> code unit, -1 if bad
    if (c->i + 4 > c->n) return -1;
    int v = 0;
    for (int k = 0; k < 4; k++) {
      const int h = hexv(c->p[c->i + k]);
      if (h < 0) return -1;
      v = (v << 4) | h;
    }
    c->i += 4;
    return v;
  }
  __device__ int next() {  // next decoded byte, -1 at the closing quote (consumed) or on a broken string
    if (npend) {
      const int b = pend & 0xFF;
      pend >>= 8;
      npend--;
      return b;
    }
    if (done || c->i >= c->n) {
      done = true;
      return -1;
    }
    int ch = c->p[c->i++];
    if (ch == '"') {
      done = true;
      return -1;
    }
    if (ch != '\\') return ch;
    if (c->i >= c->n) {
      done = true;
      return -1;
    }
    ch = c->p[c->i++];
    switch (ch) {
      case 'b': return 8;
      case 'f': return 12;
      case 'n': return 10;
      case 'r': return 13;
      case 't': return 9;
      case 'u': break;
      default: return ch;  // \" \\ \/
    }
    int cp = u4();
    if (cp < 0) {
      done = true;
      return -1;
    }
    if (cp >= 0xD800 && cp < 0xDC00 && c->i + 6 <= c->n && c->p[c->i] == '\\' && c->p[c->i + 1] == 'u') {
      const unsigned long long save = c->i;
      c->i += 2;
      const int lo = u4();
      if (lo >= 0xDC00 && lo < 0xE000) cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
      else c->i = save;  // not a pair: the high surrogate stands alone
    }
    if (cp < 0x80) return cp;
    if (cp < 0x800) {
      pend = 0x80 | (cp & 0x3F);
      npend = 1;
      return 0xC0 | (cp >> 6);
    }
    if (cp < 0x10000) {  // includes lone surrogates (written as 'surrogatepass' writes them)
      pend = (0x80 | ((cp >> 6) & 0x3F)) | ((0x80u | (cp & 0x3F)) << 8);
      npend = 2;
      return 0xE0 | (cp >> 12);
    }
    pend = (0x80 | ((cp >> 12) & 0x3F)) | ((0x80u | ((cp >> 6) & 0x3F)) << 8) | ((0x80u | (cp & 0x3F)) << 16);
    npend = 3;
    return 0xF0 | (cp >> 18);
  }
};

// cursor at the opening quote; skips the string
__device__ void skip_string(Cur& c) {
  c.i++;
  while (c.i < c.n) {
    const int ch = c.p[c.i++];
    if (ch == '\\') c.i++;
    else if (ch == '"') return;
  }
}
// cursor at the first byte of any value; leaves it after the value
__device__ void skip_value(Cur& c) {
  c.ws();
  int ch = c.peek();
  if (ch == '"') {
    skip_string(c);
    return;
  }
  if (ch == '{' || ch == '[') {
    int depth = 0;
    while (c.i < c.n) {
      ch = c.p[c.i];
      if (ch == '"') {
        skip_string(c);
        continue;
      }
      c.i++;
      if (ch == '{' || ch == '[') depth++;
      else if (ch == '}' || ch == ']') {
        if (--depth == 0) return;
      }
    }
    return;
  }
  while (c.i < c.n) {  // number / true / false / null
    ch = c.p[c.i];
    if (ch == ',' || ch == '}' || ch == ']' || ch == ' ' || ch == '\t' || ch == '\n' || ch == '\r') return;
    c.i++;
  }
}
// cursor at an opening quote: does the decoded string equal lit[0..len)? Cursor ends after the string.
__device__ bool string_equals(Cur& c, const char* lit, int len) {
  c.i++;
  StrReader r(&c);
  int k = 0;
  bool eq = true;
  for (int b = r.next(); b >= 0; b = r.next(), k++) eq = eq && k < len && b == (uint8_t)lit[k];
  return eq && k == len;
}
__device__ unsigned long long string_len(Cur c) {  // by value: measures without moving the caller
  c.i++;
  StrReader r(&c);
  unsigned long long k = 0;
  while (r.next() >= 0) k++;
  return k;
}
__device__ void string_copy(Cur& c, uint8_t* dst) {
  c.i++;
  StrReader r(&c);
  for (int b = r.next(); b >= 0; b = r.next()) *dst++ = (uint8_t)b;
}

// In the object at the cursor ('{'), the position of the value of the LAST member named lit, or ~0.
// Leaves the cursor after the object.
__device__ unsigned long long find_last_member(Cur& c, const char* lit, int len) {
  unsigned long long found = ~0ull;
  c.i++;  // '{'
  while (true) {
    c.ws();
    int ch = c.peek();
    if (ch == '}') {
      c.i++;
      break;
    }
    if (ch == ',') {
      c.i++;
      continue;
    }
    if (ch != '"') break;  // malformed: stop
    const bool hit = string_equals(c, lit, len);
    c.ws();
    if (c.peek() != ':') break;
    c.i++;
    c.ws();
    if (hit) found = c.i;
    skip_value(c);
  }
  return found;
}

// Same for two names in one pass over the object (each: the LAST member of that name, or ~0).
__device__ void find_last_members2(Cur& c, const char* lit_a, int len_a, const char* lit_b, int len_b,
                                   unsigned long long* pos_a, unsigned long long* pos_b) {
  *pos_a = *pos_b = ~0ull;
  c.i++;  // '{'
  while (true) {
    c.ws();
    int ch = c.peek();
    if (ch == '}') {
      c.i++;
      break;
    }
    if (ch == ',') {
      c.i++;
      continue;
    }
    if (ch != '"') break;
    // decode the key once, comparing against both names
    c.i++;
    StrReader r(&c);
    int k = 0;
    bool ea = true, eb = true;
    for (int b = r.next(); b >= 0; b = r.next(), k++) {
      ea = ea && k < len_a && b == (uint8_t)lit_a[k];
      eb = eb && k < len_b && b == (uint8_t)lit_b[k];
    }
    ea = ea && k == len_a;
    eb = eb && k == len_b;
    c.ws();
    if (c.peek() != ':') break;
    c.i++;
    c.ws();
    if (ea) *pos_a = c.i;
    if (eb) *pos_b = c.i;
    skip_value(c);
  }
}

// value at the cursor as a capacity quantity: how many bytes the PNL value has.  The client declares
// status.capacity as dict(str, str) and converts every non-null value with str(.) (ApiClient.__deserialize
// -> __deserialize_primitive: klass(data)); the reference then skips falsy strings and takes int(str(.))
// (check-gpu-node.py:188, :192).  So:
//   string -> decoded text;
//   integer token -> itself, except that a zero ("0", "-0": json.loads gives int 0, str gives "0") is the
//     one byte "0" -- truthy, kept, value 0, exactly like the string "0";
//   a token with a fraction or an exponent -> a float, whose str() never passes int(): nothing;
//   true / false ("True" / "False"), null (None: falsy), containers (their repr) -> nothing.
__device__ int number_class(const uint8_t* p, unsigned long long b, unsigned long long e) {  // 0 float, 1 integer zero, 2 other integer
  bool zero = true;
  for (unsigned long long i = b; i < e; i++) {
    const int ch = p[i];
    if (ch == '.' || ch == 'e' || ch == 'E') return 0;
    if (ch >= '1' && ch <= '9') zero = false;
  }
  return zero ? 1 : 2;
}
__device__ unsigned long long quantity_len(Cur c) {
  const int ch = c.peek();
  if (ch == '"') return string_len(c);
  if (ch == '-' || (ch >= '0' && ch <= '9')) {
    const unsigned long long b = c.i;
    skip_value(c);
    const int k = number_class(c.p, b, c.i);
    return k == 0 ? 0 : k == 1 ? 1 : c.i - b;
  }
  return 0;
}
__device__ void quantity_copy(Cur c, uint8_t* dst) {
  const int ch = c.peek();
  if (ch == '"') {
    string_copy(c, dst);
  } else if (ch == '-' || (ch >= '0' && ch <= '9')) {
    const unsigned long long b = c.i;
    skip_value(c);
    const int k = number_class(c.p, b, c.i);
    if (k == 1) *dst = '0';
    else if (k == 2)
      for (unsigned long long i = b; i < c.i; i++) *dst++ = c.p[i];
  }
}
// a condition's type / status: declared 'str' in V1NodeCondition, so the client applies str(.) as well:
// true -> "True" (a JSON true in `status` therefore counts as Ready), false -> "False"; numbers and
// containers become texts that can equal neither "Ready" nor "True" and are stored empty; null stays None.
__device__ unsigned long long cond_text_len(Cur c) {
  const int ch = c.peek();
  return ch == '"' ? string_len(c) : ch == 't' ? 4 : ch == 'f' ? 5 : 0;
}
__device__ void cond_text_copy(Cur c, uint8_t* dst) {
  const int ch = c.peek();
  if (ch == '"') string_copy(c, dst);
  else if (ch == 't') memcpy(dst, "True", 4);
  else if (ch == 'f') memcpy(dst, "False", 5);
}

// Row builder: pass 0 measures (entry counts, text bytes, widest length), pass 1 writes.
struct RowOut {
  uint8_t* row;        // pass 1: where the row goes
  bool ext;            // pass 1: extended form
  unsigned long long n_cap, n_cond, text, maxlen;  // pass 0 results
  unsigned long long k;                             // entries emitted so far (pass 1)
  unsigned long long text_off;                      // pass 1: next text byte, relative to row
  bool write;
  __device__ void table(unsigned long long kl, unsigned long long vl) {
    if (ext) {
      uint32_t a = (uint32_t)kl, b = (uint32_t)vl;
      memcpy(row + 12 + 8 * k, &a, 4);
      memcpy(row + 12 + 8 * k + 4, &b, 4);
    } else {
      row[2 + 2 * k] = (uint8_t)kl;
      row[2 + 2 * k + 1] = (uint8_t)vl;
    }
    k++;
  }
};

__device__ void emit_capacity(Cur c, RowOut& o) {  // cursor at '{' of the capacity object
  c.i++;
  while (true) {
    c.ws();
    int ch = c.peek();
    if (ch == '}' || ch < 0) break;
    if (ch == ',') {
      c.i++;
      continue;
    }
    if (ch != '"') break;
    Cur key = c;
    skip_string(c);
    c.ws();
    if (c.peek() != ':') break;
    c.i++;
    c.ws();
    const unsigned long long kl = string_len(key), vl = quantity_len(c);
    if (!o.write) {
      o.n_cap++;
      o.text += kl + vl;
      o.maxlen = max(o.maxlen, max(kl, vl));
    } else {
      o.table(kl, vl);
      string_copy(key, o.row + o.text_off);
      quantity_copy(c, o.row + o.text_off + kl);
      o.text_off += kl + vl;
    }
    skip_value(c);
  }
}

__device__ void emit_conditions(Cur c, RowOut& o) {  // cursor at '[' of the conditions array
  c.i++;
  while (true) {
    c.ws();
    int ch = c.peek();
    if (ch == ']' || ch < 0) break;
    if (ch == ',') {
      c.i++;
      continue;
    }
    if (ch != '{') {  // not an object: the client builds no V1NodeCondition from it
      skip_value(c);
      continue;
    }
    unsigned long long tpos, spos;
    find_last_members2(c, "type", 4, "status", 6, &tpos, &spos);  // leaves c after the object
    Cur t = c, s = c;
    unsigned long long tl = 0, sl = 0;
    if (tpos != ~0ull) {
      t.i = tpos;
      tl = cond_text_len(t);
    }
    if (spos != ~0ull) {
      s.i = spos;
      sl = cond_text_len(s);
    }
    if (!o.write) {
      o.n_cond++;
      o.text += tl + sl;
      o.maxlen = max(o.maxlen, max(tl, sl));
    } else {
      o.table(tl, sl);
      if (tl) cond_text_copy(t, o.row + o.text_off);
      if (sl) cond_text_copy(s, o.row + o.text_off + tl);
      o.text_off += tl + sl;
    }
  }
}

// What the measuring pass learns about a node and the writing pass reuses.
struct NodePlan {
  unsigned long long cap_pos, cond_pos;  // '{' of status.capacity and '[' of status.conditions, ~0 if none
  unsigned int n_cap, n_cond;
  unsigned int fast, pad_;               // fast: the structural-index walker handled this node
};

// Locate the element's status.capacity object and status.conditions array (last duplicates win).
__device__ void node_locate(const uint8_t* json, unsigned long long n, unsigned long long pos, NodePlan& pl) {
  pl.cap_pos = pl.cond_pos = ~0ull;
  Cur c{json, pos, n};
  c.ws();
  if (c.peek() != '{') return;                                  // not an object: an empty node
  Cur probe = c;
  const unsigned long long st = find_last_member(probe, "status", 6);
  if (st == ~0ull) return;
  c.i = st;
  if (c.peek() != '{') return;                                  // null / other: status falsy
  unsigned long long cap, cond;
  find_last_members2(c, "capacity", 8, "conditions", 10, &cap, &cond);
  if (cap != ~0ull && json[cap] == '{') pl.cap_pos = cap;
  if (cond != ~0ull && json[cond] == '[') pl.cond_pos = cond;
}
__device__ void node_emit(const uint8_t* json, unsigned long long n, const NodePlan& pl, RowOut& o) {
  if (pl.cap_pos != ~0ull) emit_capacity(Cur{json, pl.cap_pos, n}, o);
  if (pl.cond_pos != ~0ull) emit_conditions(Cur{json, pl.cond_pos, n}, o);
}

// One thread: which root-level array is the value of the LAST key "items", and which events are its elements.
__global__ void pick_items_kernel(const uint8_t* __restrict__ json, unsigned long long n, Masks mk, Depths dp,
                                  const unsigned int* __restrict__ ev_prefix, Work* w, unsigned int nodes_cap) {
  if (w->n_open > kMaxRootArrays || w->n_close > kMaxRootArrays) {
    w->status |= ST_TOO_MANY_ROOT_ARRAYS;
    w->n_nodes = 0;
    return;
  }
  long long best = -1;
  for (unsigned int k = 0; k < w->n_open; k++) {
    const unsigned long long pos = w->root_open[k];
    long long i = (long long)pos - 1;
    auto ws = [&](long long j) { return j >= 0 && (json[j] == ' ' || json[j] == '\t' || json[j] == '\n' || json[j] == '\r'); };
    while (ws(i)) i--;
    if (i < 0 || json[i] != ':') continue;
    i--;
    while (ws(i)) i--;
    if (i < 1 || json[i] != '"') continue;
    // i is the closing quote of the key; its opening quote is the last string-open bit before it.  The key is
    // compared decoded, so that an escaped spelling of "items" counts as json.loads counts it.
    long long open = -1;
    for (long long g = i >> 5; g >= 0 && open < 0; g--) {
      uint32_t so = plane_str_open(mk.w[g]);
      if (g == (i >> 5)) so &= (1u << (unsigned)(i & 31)) - 1u;
      if (so) open = (g << 5) + (31 - __clz(so));
    }
    if (open < 0) continue;
    Cur key{json, (unsigned long long)open, n};
    if (!string_equals(key, "items", 5)) continue;
    if ((long long)pos > best) best = (long long)pos;
  }
  if (best < 0) {
    w->status |= ST_NO_ITEMS;
    w->n_nodes = 0;
    return;
  }
  unsigned long long end = ~0ull;
  for (unsigned int k = 0; k < w->n_close; k++)
    if (w->root_close[k] > (unsigned long long)best && w->root_close[k] < end) end = w->root_close[k];
  if (end == ~0ull) {
    w->status |= ST_BAD_JSON;
    w->n_nodes = 0;
    return;
  }
  w->items_begin = (unsigned long long)best;
  w->items_end = end;
  w->ev_base = events_before(json, mk, dp, ev_prefix, (unsigned long long)best);
  unsigned long long cnt = events_before(json, mk, dp, ev_prefix, end) - w->ev_base;   // '[' + commas
  // an empty array has the '[' event but no element
  unsigned long long j = (unsigned long long)best + 1;
  while (j < end && (json[j] == ' ' || json[j] == '\t' || json[j] == '\n' || json[j] == '\r')) j++;
  if (j == end) cnt = 0;
  if (cnt > nodes_cap) {
    w->status |= ST_NODES_CAP;
    cnt = nodes_cap;
  }
  w->n_nodes = (unsigned int)cnt;
}

// ---- fast walker over the structural index --------------------------------------------------------------
// One flat state machine, one step per loop iteration: consume the next structural bit of the current
// group, or move to the next group, or (while skipping a container) pass a whole group by popcount.
// Every lane of a warp runs the same loop, so lanes at different places in their nodes stay converged;
// a recursive-descent walker here ran at 1.4 active lanes per warp.
enum : int { SK_STR_OPEN = 0, SK_STR_CLOSE = 1, SK_OPEN = 2, SK_CLOSE = 3, SK_COMMA = 4 };
enum : int { L_NODE = 0, L_STATUS = 1, L_CAP = 2, L_CONDS = 3, L_COND = 4, L_ITEM = 5 };
enum : int { KEY_OTHER = 0, KEY_A = 1, KEY_B = 2 };  // NODE: A status.  STATUS: A capacity, B conditions.  COND: A type, B status.

__host__ __device__ constexpr uint32_t le4(char a, char b = 0, char c = 0, char d = 0) {
  return (uint32_t)(uint8_t)a | (uint32_t)(uint8_t)b << 8 | (uint32_t)(uint8_t)c << 16 | (uint32_t)(uint8_t)d << 24;
}

// the twelve document bytes from `at` on, as three little-endian words (zeros past the end)
__device__ __forceinline__ void key_words(const uint8_t* __restrict__ json, unsigned long long n, unsigned long long at, uint32_t& k0,
                                          uint32_t& k1, uint32_t& k2) {
  const unsigned long long base = at & ~3ull;
  const uint32_t* w = reinterpret_cast<const uint32_t*>(json + base);
  const uint32_t a = base + 4 <= n ? __ldg(w) : 0u, b = base + 8 <= n ? __ldg(w + 1) : 0u, c = base + 12 <= n ? __ldg(w + 2) : 0u,
                 d = base + 16 <= n ? __ldg(w + 3) : 0u;
  const unsigned sh = (unsigned)(at & 3) * 8;
  k0 = __funnelshift_r(a, b, sh);
  k1 = __funnelshift_r(b, c, sh);
  k2 = __funnelshift_r(c, d, sh);
}
__device__ __forceinline__ void copy_text(uint8_t* dst, const uint8_t* __restrict__ src, unsigned long long len) {
  for (unsigned long long i = 0; i < len; i++) dst[i] = src[i];
}
// What the measuring pass notes down per row entry, so that the writing pass neither walks nor compares
// again: x = offset of the key (type) text from the start of the element, y = (value text - key text)
// as int16 | key length << 16 | value length << 24.
constexpr int kMaxDesc = 16;

// Walk from byte `pos` until the level `top` closes.  top = L_NODE: `pos` is the start of an element of
// `items`; finds status.capacity / status.conditions (pl), sizes the row and, when `desc` is given, notes
// the entries down (n_desc < 0: they did not fit the short form).  top = L_CAP or L_CONDS: `pos` is
// just inside the container (writing pass for rows without notes).  top < 0: a lane with no node,
// along for the votes.  False: something the fast walker does not do -- the general parser takes the node.
//
// Every lane of the warp must call this function together.  One iteration = one member of an object
// (key and value) or one element of an array, whatever the level: the same code for every lane, and
// nodes of the same shape stay in step.  The vote at the top brings the lanes back together once per
// iteration (left alone, lanes that took different paths never re-join and the warp runs at a handful
// of active lanes).  Positions inside are 32-bit, relative to the 32-byte group the walk starts in.
__device__ bool fast_walk(const uint8_t* __restrict__ json, unsigned long long n, const Masks& mk, unsigned long long pos, int top,
                          NodePlan& pl, RowOut& o, uint2* __restrict__ desc, int& n_desc) {
  const unsigned long long g_end = (n + kGroup - 1) / kGroup, g0 = pos >> 5;
  const bool idle = g0 >= g_end || top < 0;
  const unsigned long long base0 = g0 << 5;               // document position of relative position 0
  const uint32_t rel0 = (uint32_t)(pos & 31);             // relative position of the start of the walk
  const uint4* wp = mk.w + (idle ? 0 : g0);               // the current group's record
  uint32_t left = idle ? 0u : (uint32_t)min(g_end - g0 - 1, 0x03FFFFFFull);  // groups that may still follow (2 GiB per walk)
  uint32_t gr = 0;                                        // current group, relative
  uint32_t p0 = 0, p1 = 0, p2 = 0, bk = 0;                // kind planes and backslashes of the group, bits not yet consumed
  uint4 na = make_uint4(0, 0, 0, 0);                      // the next group, fetched ahead
  if (!idle) {
    const uint4 a = __ldg(wp);
    const uint32_t keep = ~0u << rel0;
    p0 = a.x & keep, p1 = a.y & keep, p2 = a.z & keep, bk = a.w & keep;
    if (left) na = __ldg(wp + 1);
  }
  bool ended = false;   // ran off the document
  uint32_t bacc = 0;    // backslashes passed since the last opening quote
  auto advance = [&]() {
    if (left == 0) {
      ended = true;
      p0 = p1 = p2 = bk = 0;
      return;
    }
    --left, ++gr, ++wp;
    bacc |= bk;
    p0 = na.x, p1 = na.y, p2 = na.z, bk = na.w;
    if (left) na = __ldg(wp + 1);
  };
  // the next structural byte: its kind, its relative position in `at`; -1 at the end of the document
  auto next = [&](uint32_t& at) -> int {
    uint32_t any;
    while ((any = p0 | p1 | p2) == 0) {
      advance();
      if (ended) return -1;
    }
    const int bit = __ffs(any) - 1;
    const uint32_t keep = ~((2u << bit) - 1u);
    at = (gr << 5) + bit;
    const int kind = (int)(((p0 >> bit) & 1u) | (((p1 >> bit) & 1u) << 1) | (((p2 >> bit) & 1u) << 2)) - 1;  // the planes spell kind + 1
    bacc = kind == SK_STR_OPEN ? 0u : (bacc | (bk & ~keep));  // at a closing quote: non-zero iff the string held a backslash
    p0 &= keep, p1 &= keep, p2 &= keep, bk &= keep;
    return kind;
  };
  // the next bracket or comma, strings passed over unseen (the planes hold nothing for the bytes inside them): how a
  // member whose key does not matter is left behind in one step instead of three
  auto next_delim = [&](uint32_t& at) -> int {
    uint32_t any;
    while ((any = (p0 & p1) | p2) == 0) {
      advance();
      if (ended) return -1;
    }
    const int bit = __ffs(any) - 1;
    const uint32_t keep = ~((2u << bit) - 1u);
    at = (gr << 5) + bit;
    const int kind = ((p2 >> bit) & 1u) ? (((p0 >> bit) & 1u) ? SK_COMMA : SK_CLOSE) : SK_OPEN;
    p0 &= keep, p1 &= keep, p2 &= keep, bk &= keep;
    return kind;
  };
  // the opening bracket has just been consumed: pass its container, whole groups at a time where it cannot close
  auto skip_container = [&]() -> bool {
    int depth = 1;
    while (true) {
      const uint32_t op = p0 & p1, cl = p2 & ~p0;  // kinds 3 and 4
      if (__popc(cl) < depth) {
        depth += __popc(op) - __popc(cl);
      } else {
        uint32_t oc = op | cl;
        while (oc) {
          const int bit = __ffs(oc) - 1;
          oc &= oc - 1;
          if ((op >> bit) & 1u) {
            depth++;
          } else if (--depth == 0) {
            const uint32_t keep = ~((2u << bit) - 1u);
            p0 &= keep, p1 &= keep, p2 &= keep, bk &= keep;
            return true;
          }
        }
      }
      advance();
      if (ended) return false;
    }
  };

  // level L_ITEM: the element itself (a value with no key, like the elements of L_CONDS)
  int level = top == L_NODE ? L_ITEM : top;
  bool run = !idle, ok = !idle;
  uint32_t tp = 0, tl = 0, sp = 0, sl = 0;                 // the current condition's type and status text (relative)
  uint32_t n_cap = 0, n_cond = 0, text = 0, maxlen = 0;    // what the measuring pass wants to know
  bool seen_status = false, seen_cap = false, seen_cond = false, have_t = false, have_s = false;
  // one row entry: key text (kt, kl) and value text (vt, vl), relative positions
  auto entry = [&](uint32_t kt, uint32_t kl, uint32_t vt, uint32_t vl, bool is_cap) {
    if (o.write) {
      o.table(kl, vl);
      copy_text(o.row + o.text_off, json + base0 + kt, kl);
      copy_text(o.row + o.text_off + kl, json + base0 + vt, vl);
      o.text_off += kl + vl;
      return;
    }
    if (desc && n_desc >= 0) {
      const uint32_t x = (kl ? kt : vl ? vt : rel0) - rel0;
      const int d = kl && vl ? (int)(vt - kt) : 0;  // with no key text the value text sits at x itself
      if (n_desc >= kMaxDesc || kl >= 255 || vl >= 255 || d < -32768 || d > 32767 || (is_cap && n_cond))
        n_desc = -1;  // too many, too long, too far apart, or conditions seen before capacity: walk again when writing
      else
        desc[n_desc++] = make_uint2(x, (uint32_t)(d & 0xFFFF) | kl << 16 | vl << 24);
    }
    if (is_cap) n_cap++;
    else n_cond++;
    text += kl + vl;
    maxlen = max(maxlen, max(kl, vl));
  };

  while (__any_sync(0xFFFFFFFFu, run)) {
    if (!run) continue;
    const bool keyed = level != L_ITEM && level != L_CONDS;
    uint32_t at = 0, p = 0, q = 0, q2 = 0, kl = 0;
    int kind = next(at);
    if (kind == SK_COMMA && level != L_ITEM) kind = next(at);  // the separator before this member (after the element itself a comma ends it)
    bool pop = kind == SK_CLOSE;
    int key = KEY_OTHER;
    if (!pop && keyed) {
      // ---- the key ----
      p = at;
      if (kind != SK_STR_OPEN || next(q) != SK_STR_CLOSE || bacc) {  // an escaped key could spell one that matters
        ok = run = false;
        continue;
      }
      kl = q - p - 1;
      const bool look = level == L_NODE ? kl == 6 : level == L_STATUS ? (kl == 8 || kl == 10) : level == L_COND ? (kl == 4 || kl == 6) : false;
      if (look) {
        uint32_t k0, k1, k2;
        key_words(json, n, base0 + p + 1, k0, k1, k2);
        const bool is_status = kl == 6 && k0 == le4('s', 't', 'a', 't') && (k1 & 0xFFFFu) == le4('u', 's');
        bool again = false;  // a second member of a name that matters: last one wins, general parser
        if (level == L_NODE) {
          if (is_status) key = KEY_A, again = seen_status, seen_status = true;
        } else if (level == L_STATUS) {
          if (kl == 8 && k0 == le4('c', 'a', 'p', 'a') && k1 == le4('c', 'i', 't', 'y')) key = KEY_A, again = seen_cap, seen_cap = true;
          else if (kl == 10 && k0 == le4('c', 'o', 'n', 'd') && k1 == le4('i', 't', 'i', 'o') && (k2 & 0xFFFFu) == le4('n', 's'))
            key = KEY_B, again = seen_cond, seen_cond = true;
        } else {
          if (kl == 4 && k0 == le4('t', 'y', 'p', 'e')) key = KEY_A, again = have_t;
          else if (is_status) key = KEY_B, again = have_s;
        }
        if (again) {
          ok = run = false;
          continue;
        }
      }
      // the first structural byte of the value; for a member that cannot matter, the delimiter that ends it
      kind = (level != L_CAP && key == KEY_OTHER) ? next_delim(at) : next(at);
    }
    if (!pop) {
      // ---- the value (or array element): `kind` at `at` is its first structural byte ----
      const bool need_string = level == L_CAP || (level == L_COND && key != KEY_OTHER);
      if (kind == SK_STR_OPEN) {
        if (next(q2) != SK_STR_CLOSE || (need_string && bacc)) {
          ok = run = false;
          continue;
        }
        const uint32_t vl = q2 - at - 1;
        if (level == L_CAP) {
          entry(p + 1, kl, at + 1, vl, true);
        } else if (level == L_COND && key != KEY_OTHER) {
          if (key == KEY_A) tp = at + 1, tl = vl, have_t = true;
          else sp = at + 1, sl = vl, have_s = true;
        } else if (level == L_ITEM) {
          run = false;  // a string element: an empty node
        }
      } else if (need_string || kind < 0) {
        ok = run = false;  // not a string where a string is read (the general parser knows the rules), or the end of the document
        continue;
      } else if (kind == SK_OPEN) {
        const bool wanted = level == L_ITEM || level == L_CONDS || (level == L_NODE && key == KEY_A) || (level == L_STATUS && key != KEY_OTHER);
        const int ch = wanted ? json[base0 + at] : 0;
        if (level == L_ITEM && ch == '{') {
          level = L_NODE;
        } else if (level == L_NODE && key == KEY_A && ch == '{') {
          level = L_STATUS, seen_cap = seen_cond = false;
        } else if (level == L_STATUS && key == KEY_A && ch == '{') {
          pl.cap_pos = base0 + at;
          level = L_CAP;
        } else if (level == L_STATUS && key == KEY_B && ch == '[') {
          pl.cond_pos = base0 + at;
          level = L_CONDS;
        } else if (level == L_CONDS && ch == '{') {
          level = L_COND;
          tp = tl = sp = sl = 0;
          have_t = have_s = false;
        } else if (!skip_container()) {  // anything else: passed over (an element that is no object is an empty node / no condition)
          ok = run = false;
          continue;
        } else if (level == L_ITEM) {
          run = false;  // the element was an array: an empty node
        }
      } else if (kind == SK_CLOSE) {  // a number / true / false / null, then the end of the container
        pop = true;
      } else if (level == L_ITEM) {
        run = false;  // a scalar element (its comma): an empty node
      }
    }
    if (pop) {
      if (level == L_COND) {
        entry(tp, tl, sp, sl, false);
        level = L_CONDS;
      } else if (level == top || level == L_ITEM) {
        run = false;  // done (L_ITEM: the items array ended where an element should start -- an empty node)
      } else if (level == L_STATUS) {
        level = L_NODE;
      } else {  // L_CAP, L_CONDS
        level = L_STATUS;
      }
    }
  }
  if (!o.write) o.n_cap += n_cap, o.n_cond += n_cond, o.text += text, o.maxlen = max(o.maxlen, (unsigned long long)maxlen);
  return ok;
}

// row size in bytes (4-byte padded); bit 31 set when the extended form is needed
__global__ void measure_rows_kernel(const uint8_t* __restrict__ json, unsigned long long n, Masks mk,
                                    const unsigned long long* __restrict__ elem_pos, Work* w, unsigned int* __restrict__ sizes,
                                    NodePlan* __restrict__ plans, uint2* __restrict__ desc) {
  const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool have = i < w->n_nodes;
  if (!__any_sync(0xFFFFFFFFu, have)) return;
  RowOut o{};
  o.write = false;
  NodePlan pl;
  pl.cap_pos = pl.cond_pos = ~0ull;
  int n_desc = 0;
  const bool fast_ok = fast_walk(json, n, mk, have ? elem_pos[i] : 0, have ? L_NODE : -1, pl, o, desc + (size_t)kMaxDesc * i, n_desc);  // whole warp
  if (!have) return;
  pl.fast = fast_ok ? (n_desc >= 0 ? 2u : 1u) : 0u;
  if (!pl.fast) {
    o = RowOut{};
    o.write = false;
    node_locate(json, min(n, w->items_end), elem_pos[i], pl);
    node_emit(json, min(n, w->items_end), pl, o);
    atomicAdd(&w->general_nodes, 1u);
  }
  pl.n_cap = (unsigned int)o.n_cap;
  pl.n_cond = (unsigned int)o.n_cond;
  plans[i] = pl;
  const bool ext = o.n_cap >= 255 || o.n_cond >= 255 || o.maxlen >= 255;
  unsigned long long bytes = ext ? 12 + 8 * (o.n_cap + o.n_cond) + o.text : 2 + 2 * (o.n_cap + o.n_cond) + o.text;
  bytes = (bytes + 3) & ~3ull;
  if (bytes >= (1ull << 31)) bytes = (1ull << 31) - 4;          // cannot fit PNL anyway: rows_cap check trips
  sizes[i] = (unsigned int)bytes | (ext ? 0x80000000u : 0u);
}
__global__ void strip_flag_kernel(const unsigned int* __restrict__ sizes, unsigned long long* __restrict__ plain, unsigned int n) {
  const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) plain[i] = sizes[i] & 0x7FFFFFFFu;
}

// `len` document bytes from position `at` to dst: aligned 32-bit loads, funnel-shifted; byte stores.
__device__ __forceinline__ void copy_from_doc(uint8_t* dst, const uint8_t* __restrict__ json, unsigned long long n, unsigned long long at,
                                              unsigned int len) {
  if (len == 0) return;
  unsigned long long base = at & ~3ull;
  const unsigned sh = (unsigned)(at & 3) * 8;
  auto word = [&](unsigned long long b) -> uint32_t {  // the aligned word at b; a word that straddles the end is read bytewise
    if (b + 4 <= n) return __ldg(reinterpret_cast<const uint32_t*>(json + b));
    uint32_t v = 0;
    for (unsigned k = 0; b + k < n; k++) v |= (uint32_t)json[b + k] << (8 * k);
    return v;
  };
  uint32_t prev = word(base);
  for (unsigned int j = 0; j < len; j += 4) {
    base += 4;
    const uint32_t nxt = (sh && j + 4 - (sh >> 3) < len) || (!sh && j + 4 < len) ? word(base) : 0u;
    const uint32_t v = __funnelshift_r(prev, nxt, sh);
    prev = nxt;
    const unsigned int m = min(4u, len - j);
    dst[j] = (uint8_t)v;
    if (m > 1) dst[j + 1] = (uint8_t)(v >> 8);
    if (m > 2) dst[j + 2] = (uint8_t)(v >> 16);
    if (m > 3) dst[j + 3] = (uint8_t)(v >> 24);
  }
}

constexpr int kWriteThreads = 128;
constexpr unsigned int kWriteStage = 32768;   // bytes of rows a CTA assembles in shared memory before one coalesced copy out
constexpr unsigned int kSpan = 1280;          // bytes of document one node's texts may span to be gathered by the whole warp
constexpr int kGatherBatch = 4;               // nodes whose spans a warp holds in shared memory at a time (8 lanes cut each node's row)
constexpr unsigned int kWriteSmem = kWriteStage + (kWriteThreads / 32) * kGatherBatch * kSpan + kWriteThreads * (8 + 4 + 4 + 4 + 8);  // 4 CTAs per SM

// Writing pass.  A CTA's rows are one contiguous piece of the output: they are assembled in shared memory and
// copied out coalesced.  The texts of a node (capacity keys and values, condition types and statuses)
// lie within about a kilobyte of the document; the measuring pass noted where.  A thread fetching its own
// node's texts one after another waits a memory round trip for each, so the warp fetches them together:
// for four nodes at a time all 32 lanes copy the nodes' spans into shared memory with independent 16-byte
// loads (all issued before the first is used), then eight lanes per node cut its row out of shared memory.
// The pass is latency-bound: shared memory is kept to 55.5 KB per CTA so that four CTAs fit an SM.
__global__ void __launch_bounds__(kWriteThreads) write_rows_kernel(
    const uint8_t* __restrict__ json, unsigned long long n, Masks mk, const NodePlan* __restrict__ plans, const uint2* __restrict__ desc,
    const unsigned long long* __restrict__ elem_pos, const Work* w, const unsigned int* __restrict__ sizes,
    const unsigned long long* __restrict__ offs, const unsigned long long* __restrict__ total, uint8_t* __restrict__ rows,
    unsigned long long rows_cap, uint32_t* __restrict__ row_ptr, long long* __restrict__ info) {
  extern __shared__ __align__(16) uint8_t sh[];
  uint8_t* gather = sh + kWriteStage;                                                    // [warps][kGatherBatch][kSpan]
  unsigned long long* s_lo = reinterpret_cast<unsigned long long*>(gather + (kWriteThreads / 32) * kGatherBatch * kSpan);  // [threads] span start
  uint8_t** s_rowp = reinterpret_cast<uint8_t**>(s_lo + kWriteThreads);                    // [threads] where the row is assembled
  unsigned int* s_len = reinterpret_cast<unsigned int*>(s_rowp + kWriteThreads);           // [threads] span length, 0: not gathered
  int* s_shift = reinterpret_cast<int*>(s_len + kWriteThreads);                            // [threads] node start - span start
  unsigned int* s_ne = reinterpret_cast<unsigned int*>(s_shift + kWriteThreads);           // [threads] entries
  const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned int nn = w->n_nodes;
  const bool fits = *total + 32 <= rows_cap && *total < 0xFFF00000ull;
  if (i == 0) {
    info[INFO_NODES] = nn;
    info[INFO_ROWS_BYTES] = fits ? (long long)*total : 0;
    info[INFO_STATUS] = w->status | (fits ? 0 : ST_ROWS_CAP);
    info[INFO_ITEMS_BEGIN] = (long long)w->items_begin;
    info[INFO_ITEMS_END] = (long long)w->items_end;
    info[INFO_GENERAL_NODES] = w->general_nodes;
    row_ptr[nn] = fits ? (uint32_t)*total : 0u;
    for (int k = 1; k <= 4; k++) row_ptr[nn + k] = row_ptr[nn];
  }
  const unsigned int first = blockIdx.x * blockDim.x;
  if (first >= nn) return;  // whole CTA
  const bool have = i < nn;
  if (!fits) {
    if (have) row_ptr[i] = 0;
    return;  // whole CTA
  }
  const unsigned int last = min(nn, first + blockDim.x);
  const unsigned long long blk_lo = offs[first], blk_hi = last < nn ? offs[last] : *total;
  const bool staged = blk_hi - blk_lo <= kWriteStage;
  if (staged) {
    for (unsigned int k = threadIdx.x * 4; k < (unsigned int)(blk_hi - blk_lo); k += blockDim.x * 4) *reinterpret_cast<uint32_t*>(sh + k) = 0u;
    __syncthreads();
  }
  const unsigned long long off = have ? offs[i] : 0;
  const unsigned int bytes = have ? sizes[i] & 0x7FFFFFFFu : 0u;
  uint8_t* row = staged ? sh + (off - blk_lo) : rows + off;
  NodePlan m;
  m.cap_pos = m.cond_pos = ~0ull;
  m.n_cap = m.n_cond = m.fast = 0;
  RowOut o{};
  o.write = true;
  o.row = row;
  unsigned int ne = 0;
  if (have) {
    row_ptr[i] = (uint32_t)off;
    if (!staged)
      for (unsigned int k = 0; k < bytes; k += 4) *reinterpret_cast<uint32_t*>(row + k) = 0u;   // rows are 4-byte aligned
    o.ext = (sizes[i] >> 31) != 0;
    m = plans[i];
    ne = m.n_cap + m.n_cond;
    if (o.ext) {
      row[0] = row[1] = 0xFF;
      uint32_t a = (uint32_t)m.n_cap, b = (uint32_t)m.n_cond;
      memcpy(row + 4, &a, 4);
      memcpy(row + 8, &b, 4);
      o.text_off = 12 + 8 * (unsigned long long)ne;
    } else {
      row[0] = (uint8_t)m.n_cap;
      row[1] = (uint8_t)m.n_cond;
      o.text_off = 2 + 2 * (unsigned long long)ne;
    }
  }
  // ---- rows with notes (fast == 2: at most kMaxDesc entries, short form) ----
  const bool noted = have && m.fast == 2;
  uint2 d[kMaxDesc];
  unsigned int lo = 0xFFFFFFFFu, hi = 0;  // the node's texts lie in [node_pos + lo, node_pos + hi)
#pragma unroll
  for (int e = 0; e < kMaxDesc; e++) {
    d[e] = make_uint2(0, 0);
    if (noted && (unsigned int)e < ne) {
      d[e] = __ldg(desc + (size_t)kMaxDesc * i + e);
      const unsigned int kl = (d[e].y >> 16) & 0xFFu, vl = d[e].y >> 24;
      const unsigned int vt = kl ? d[e].x + (int)(short)(d[e].y & 0xFFFFu) : d[e].x;
      if (kl) lo = min(lo, d[e].x), hi = max(hi, d[e].x + kl);
      if (vl) lo = min(lo, vt), hi = max(hi, vt + vl);
    }
  }
  const unsigned long long node_pos = have ? elem_pos[i] : 0;
  const unsigned long long span_lo = (node_pos + (lo <= hi ? lo : 0)) & ~15ull;          // 16-byte aligned document position
  const unsigned int span_len = lo < hi ? (unsigned int)(node_pos + hi - span_lo) : 0u;
  const bool gathered = noted && span_len <= kSpan;
  s_lo[threadIdx.x] = span_lo;
  s_len[threadIdx.x] = gathered ? span_len : 0u;
  s_shift[threadIdx.x] = (int)(long long)(node_pos - span_lo);
  s_rowp[threadIdx.x] = row;
  s_ne[threadIdx.x] = gathered ? ne : 0u;
  __syncwarp();
  uint8_t* wbuf = gather + warp * (kGatherBatch * kSpan);
  auto load16 = [&](unsigned long long at) -> uint4 {
    if (at + 16 <= n) return __ldg(reinterpret_cast<const uint4*>(json + at));
    uint8_t tmp[16];
    for (int t = 0; t < 16; t++) tmp[t] = at + t < n ? json[at + t] : (uint8_t)0;
    uint4 v;
    memcpy(&v, tmp, 16);
    return v;
  };
  for (int b0 = 0; b0 < 32; b0 += kGatherBatch) {
    // every load of the batch is issued before the first store waits for one
    uint4 v[kGatherBatch][2];
    unsigned int jlen[kGatherBatch];
#pragma unroll
    for (int j = 0; j < kGatherBatch; j++) {
      const unsigned long long jlo = s_lo[warp * 32 + b0 + j];
      jlen[j] = s_len[warp * 32 + b0 + j];
      const unsigned int k0 = lane * 16, k1 = k0 + 512;
      v[j][0] = v[j][1] = make_uint4(0, 0, 0, 0);
      if (k0 < jlen[j]) v[j][0] = load16(jlo + k0);
      if (k1 < jlen[j]) v[j][1] = load16(jlo + k1);
    }
#pragma unroll
    for (int j = 0; j < kGatherBatch; j++) {
      const unsigned int k0 = lane * 16, k1 = k0 + 512, k2 = k0 + 1024;
      if (k0 < jlen[j]) *reinterpret_cast<uint4*>(wbuf + j * kSpan + k0) = v[j][0];
      if (k1 < jlen[j]) *reinterpret_cast<uint4*>(wbuf + j * kSpan + k1) = v[j][1];
      if (k2 < jlen[j]) *reinterpret_cast<uint4*>(wbuf + j * kSpan + k2) = load16(s_lo[warp * 32 + b0 + j] + k2);  // spans past 1 KB: rare
    }
    __syncwarp();
    {
      // eight lanes per node: lane q of them writes entries q and q + 8.  The notes stay in the owning lane's
      // registers and are handed round by shuffles (a copy in shared memory cost a fourth CTA per SM).
      const int j = lane >> 3, q = lane & 7, node = warp * 32 + b0 + j;
      const unsigned int ne_j = s_ne[node];
      unsigned int run = 2 + 2 * ne_j, off_a = 0, off_b = 0;  // where my entries' texts go
      uint2 da = make_uint2(0, 0), db = make_uint2(0, 0);
#pragma unroll
      for (int e = 0; e < kMaxDesc; e++) {
        uint2 de;
        de.x = __shfl_sync(0xFFFFFFFFu, d[e].x, b0 + j);
        de.y = __shfl_sync(0xFFFFFFFFu, d[e].y, b0 + j);
        if ((unsigned int)e < ne_j) {
          if (e == q) da = de, off_a = run;
          if (e == q + 8) db = de, off_b = run;
          run += ((de.y >> 16) & 0xFFu) + (de.y >> 24);
        }
      }
      if (ne_j) {
        uint8_t* rowp = s_rowp[node];
        const uint8_t* src = wbuf + j * kSpan + s_shift[node];  // src[x] = document byte (node start + x)
        auto put = [&](unsigned int e, uint2 de, unsigned int text_off) {
          const unsigned int kl = (de.y >> 16) & 0xFFu, vl = de.y >> 24;
          const unsigned int vt = kl ? de.x + (int)(short)(de.y & 0xFFFFu) : de.x;
          rowp[2 + 2 * e] = (uint8_t)kl;
          rowp[3 + 2 * e] = (uint8_t)vl;
          for (unsigned int t = 0; t < kl; t++) rowp[text_off + t] = src[de.x + t];
          for (unsigned int t = 0; t < vl; t++) rowp[text_off + kl + t] = src[vt + t];
        };
        if ((unsigned int)q < ne_j) put(q, da, off_a);
        if ((unsigned int)q + 8 < ne_j) put(q + 8, db, off_b);
      }
    }
    __syncwarp();
  }
  if (noted && !gathered) {  // texts spread too far apart: fetched by the owning thread
    unsigned long long text_off = o.text_off;
    for (unsigned int e = 0; e < ne; e++) {
      const uint2 de = __ldg(desc + (size_t)kMaxDesc * i + e);
      const unsigned int kl = (de.y >> 16) & 0xFFu, vl = de.y >> 24;
      const long long dlt = (short)(de.y & 0xFFFFu);
      row[2 + 2 * e] = (uint8_t)kl;
      row[3 + 2 * e] = (uint8_t)vl;
      const unsigned long long kt = node_pos + de.x;
      copy_from_doc(row + text_off, json, n, kt, kl);
      copy_from_doc(row + text_off + kl, json, n, kl ? kt + dlt : kt, vl);
      text_off += kl + vl;
    }
  }
  if (have && m.fast == 0) node_emit(json, min(n, w->items_end), m, o);
  // rows without notes (too many or too long entries): the index walk again, capacity first, then
  // conditions -- the same walk as the measuring pass, so it succeeds again.  Whole warps call it.
  const bool rewalk = have && m.fast == 1;
  if (__any_sync(0xFFFFFFFFu, rewalk)) {
    NodePlan unused;
    int none = -1;
    const bool cap = rewalk && m.cap_pos != ~0ull, cond = rewalk && m.cond_pos != ~0ull;
    fast_walk(json, n, mk, cap ? m.cap_pos + 1 : 0, cap ? L_CAP : -1, unused, o, nullptr, none);
    fast_walk(json, n, mk, cond ? m.cond_pos + 1 : 0, cond ? L_CONDS : -1, unused, o, nullptr, none);
  }
  if (staged) {
    __syncthreads();
    uint32_t* dst = reinterpret_cast<uint32_t*>(rows + blk_lo);  // row offsets are multiples of 4
    for (unsigned int k = threadIdx.x; k < (unsigned int)((blk_hi - blk_lo) >> 2); k += blockDim.x) dst[k] = reinterpret_cast<const uint32_t*>(sh)[k];
  }
}

}  // namespace js
}  // namespace gnc
