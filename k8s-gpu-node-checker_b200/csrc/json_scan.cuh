// json_scan.cuh -- raw /api/v1/nodes JSON bytes -> PNL-v2 columns, on the device (sm_100a).
//
// Replaces, for the classification path, the third-party deserialisation behind
// `api.list_node().items` (/root/reference/check-gpu-node.py:217): only the parts of each node the
// reference reads are extracted -- `status.capacity` entries (check-gpu-node.py:183-187) and the
// `type` / `status` of each element of `status.conditions` (:173-177) -- with json.loads semantics
// (duplicate object keys: last wins; \uXXXX escapes, surrogate pairs joined, lone surrogates kept).
//
// Structure (all data-parallel until the per-node parse):
//   1. lexer:   256-byte chunks; each chunk's effect on the 3-state string lexer is a function
//               {outside, in string, after backslash} -> state; an exclusive scan of function
//               compositions gives every chunk its start state
//   2. depth:   per-chunk bracket balance outside strings, exclusive sum scan -> start depth
//   3. events:  element separators of root-level arrays (the '[' itself and each ',' at depth 2) are
//               counted per chunk and scanned; root-level array brackets are listed; one thread then
//               picks the LAST root key "items" and the range of events that are its elements
//   4. nodes:   one thread per element parses its node (a small recursive-descent walker over the
//               bytes), first to measure its PNL row, then -- after a scan of row sizes -- to write it
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace gnc {
namespace js {

constexpr int kChunk = 256;
constexpr int kMaxRootArrays = 64;

enum : int { INFO_NODES = 0, INFO_ROWS_BYTES = 1, INFO_STATUS = 2, INFO_ITEMS_BEGIN = 3, INFO_ITEMS_END = 4, INFO_N = 8 };
enum : int { ST_NO_ITEMS = 1, ST_NODES_CAP = 2, ST_ROWS_CAP = 4, ST_BAD_JSON = 8, ST_TOO_MANY_ROOT_ARRAYS = 16 };

struct Work {                 // device scratch, one per call
  unsigned long long root_open[kMaxRootArrays];   // positions of '[' opening a root-level array
  unsigned long long root_close[kMaxRootArrays];  // positions of the ']' closing one
  unsigned int n_open, n_close;
  unsigned long long items_begin, items_end;      // byte positions of the chosen '[' and its ']'
  unsigned long long ev_base;                     // rank of the event at items_begin
  unsigned int n_nodes;
  unsigned int status;
};

// ---- lexer -------------------------------------------------------------------------------------
__device__ __forceinline__ int lex_step(int st, uint32_t c) {
  if (st == 0) return c == '"' ? 1 : 0;
  if (st == 1) return c == '\\' ? 2 : (c == '"' ? 0 : 1);
  return 1;
}
// a chunk's transition function, 2 bits per start state
__device__ __forceinline__ uint32_t compose(uint32_t f, uint32_t g) {  // first f, then g
  return ((g >> (2 * (f & 3))) & 3) | (((g >> (2 * ((f >> 2) & 3))) & 3) << 2) | (((g >> (2 * ((f >> 4) & 3))) & 3) << 4);
}
constexpr uint32_t kIdentity = 0 | (1 << 2) | (2 << 4);

// Bytes of chunk c in order, fetched 16 at a time (the document is 16-byte aligned; the last,
// partial chunk is read bytewise).
template <class F>
__device__ __forceinline__ void for_chunk_bytes(const uint8_t* __restrict__ json, unsigned long long n, unsigned long long c, F&& f) {
  const unsigned long long b = c * kChunk, e = min(n, b + kChunk);
  if (e - b == kChunk) {
    const uint4* v = reinterpret_cast<const uint4*>(json + b);
#pragma unroll 4
    for (int q = 0; q < kChunk / 16; q++) {
      const uint4 x = __ldg(v + q);
      const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
      for (int k = 0; k < 16; k++) f(b + 16 * q + k, (w[k >> 2] >> (8 * (k & 3))) & 0xFFu);
    }
  } else {
    for (unsigned long long i = b; i < e; i++) f(i, (uint32_t)json[i]);
  }
}

__global__ void lex_funcs_kernel(const uint8_t* __restrict__ json, unsigned long long n, uint8_t* __restrict__ funcs,
                                 unsigned long long n_chunks) {
  const unsigned long long c = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (c >= n_chunks) return;
  int s0 = 0, s1 = 1, s2 = 2;
  for_chunk_bytes(json, n, c, [&](unsigned long long, uint32_t ch) {
    s0 = lex_step(s0, ch);
    s1 = lex_step(s1, ch);
    s2 = lex_step(s2, ch);
  });
  funcs[c] = (uint8_t)(s0 | (s1 << 2) | (s2 << 4));
}

// ---- generic 3-kernel exclusive scan over n items, 1024 per block ---------------------------------
template <class T, class Op>
__global__ void scan_block_totals(const T* __restrict__ in, unsigned long long n, T* __restrict__ totals, Op op, T ident) {
  __shared__ T sh[1024];
  const unsigned long long i = blockIdx.x * 1024ull + threadIdx.x;
  sh[threadIdx.x] = i < n ? in[i] : ident;
  __syncthreads();
  for (int d = 1; d < 1024; d <<= 1) {
    T v = threadIdx.x >= d ? op(sh[threadIdx.x - d], sh[threadIdx.x]) : sh[threadIdx.x];
    __syncthreads();
    sh[threadIdx.x] = v;
    __syncthreads();
  }
  if (threadIdx.x == 1023) totals[blockIdx.x] = sh[1023];
}
template <class T, class Op>
__global__ void scan_totals_serial(T* totals, unsigned long long nb, Op op, T ident, T* grand) {
  // one thread: nb is n/1024, a few thousand at most
  T run = ident;
  for (unsigned long long b = 0; b < nb; b++) {
    T t = totals[b];
    totals[b] = run;
    run = op(run, t);
  }
  if (grand) *grand = run;
}
template <class T, class Op>
__global__ void scan_apply(const T* __restrict__ in, unsigned long long n, const T* __restrict__ totals, T* __restrict__ out,
                           Op op, T ident) {
  __shared__ T sh[1024];
  const unsigned long long i = blockIdx.x * 1024ull + threadIdx.x;
  const T mine = i < n ? in[i] : ident;
  sh[threadIdx.x] = mine;
  __syncthreads();
  for (int d = 1; d < 1024; d <<= 1) {
    T v = threadIdx.x >= d ? op(sh[threadIdx.x - d], sh[threadIdx.x]) : sh[threadIdx.x];
    __syncthreads();
    sh[threadIdx.x] = v;
    __syncthreads();
  }
  // exclusive: everything before me in the block, after the block's own prefix
  const T before = threadIdx.x ? sh[threadIdx.x - 1] : ident;
  if (i < n) out[i] = op(totals[blockIdx.x], before);
}
struct OpCompose {
  __device__ uint8_t operator()(uint8_t a, uint8_t b) const { return (uint8_t)compose(a, b); }
};
struct OpAddI {
  __device__ int operator()(int a, int b) const { return a + b; }
};
struct OpAddU {
  __device__ unsigned int operator()(unsigned int a, unsigned int b) const { return a + b; }
};
struct OpAddULL {
  __device__ unsigned long long operator()(unsigned long long a, unsigned long long b) const { return a + b; }
};

// start state of a chunk = (composition of all earlier chunks)(state 0)
__device__ __forceinline__ int start_state(const uint8_t* prefix_funcs, unsigned long long c) { return prefix_funcs[c] & 3; }

// Visit the bytes of chunk c that lie outside strings: f(pos, byte, depth_before); brackets move depth.
template <class F>
__device__ __forceinline__ int visit_chunk(const uint8_t* __restrict__ json, unsigned long long n, unsigned long long c, int st,
                                           int depth, F&& f) {
  for_chunk_bytes(json, n, c, [&](unsigned long long i, uint32_t ch) {
    if (st == 0) {
      if (ch == '"') {
        st = 1;
      } else {
        f(i, ch, depth);
        if (ch == '{' || ch == '[') depth++;
        else if (ch == '}' || ch == ']') depth--;
      }
    } else {
      st = lex_step(st, ch);
    }
  });
  return depth;
}

__global__ void depth_delta_kernel(const uint8_t* __restrict__ json, unsigned long long n, const uint8_t* __restrict__ pf,
                                   int* __restrict__ delta, unsigned long long n_chunks) {
  const unsigned long long c = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (c >= n_chunks) return;
  delta[c] = visit_chunk(json, n, c, start_state(pf, c), 0, [](unsigned long long, uint32_t, int) {});
}

// events: the '[' of every root-level array and every ',' directly inside one (depth 2 outside strings)
__global__ void events_kernel(const uint8_t* __restrict__ json, unsigned long long n, const uint8_t* __restrict__ pf,
                              const int* __restrict__ depth0, unsigned int* __restrict__ ev_count, Work* w,
                              unsigned long long n_chunks) {
  const unsigned long long c = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (c >= n_chunks) return;
  unsigned int cnt = 0;
  visit_chunk(json, n, c, start_state(pf, c), depth0[c], [&](unsigned long long pos, uint32_t ch, int d) {
    if (ch == '[' && d == 1) {
      cnt++;
      const unsigned int k = atomicAdd(&w->n_open, 1u);
      if (k < kMaxRootArrays) w->root_open[k] = pos;
    } else if (ch == ',' && d == 2) {
      cnt++;
    } else if (ch == ']' && d == 2) {
      const unsigned int k = atomicAdd(&w->n_close, 1u);
      if (k < kMaxRootArrays) w->root_close[k] = pos;
    } else if ((ch == '}' || ch == ']') && d <= 0) {
      atomicOr(&w->status, (unsigned int)ST_BAD_JSON);
    }
  });
  ev_count[c] = cnt;
}

// number of events strictly before byte `pos` (pos lies outside strings)
__device__ unsigned long long events_before(const uint8_t* json, unsigned long long n, const uint8_t* pf, const int* depth0,
                                            const unsigned int* ev_prefix, unsigned long long pos) {
  const unsigned long long c = pos / kChunk;
  unsigned long long cnt = ev_prefix[c];
  visit_chunk(json, min(n, pos), c, start_state(pf, c), depth0[c], [&](unsigned long long, uint32_t ch, int d) {
    if ((ch == '[' && d == 1) || (ch == ',' && d == 2)) cnt++;
  });
  return cnt;
}

// One thread: which root-level array is the value of the LAST key "items", and which events are its elements.
__global__ void pick_items_kernel(const uint8_t* __restrict__ json, unsigned long long n, const uint8_t* __restrict__ pf,
                                  const int* __restrict__ depth0, const unsigned int* __restrict__ ev_prefix, Work* w,
                                  unsigned int nodes_cap) {
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
    if (i < 6 || json[i] != '"' || json[i - 1] != 's' || json[i - 2] != 'm' || json[i - 3] != 'e' || json[i - 4] != 't' ||
        json[i - 5] != 'i' || json[i - 6] != '"')
      continue;
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
  w->ev_base = events_before(json, n, pf, depth0, ev_prefix, (unsigned long long)best);
  unsigned long long cnt = events_before(json, n, pf, depth0, ev_prefix, end) - w->ev_base;   // '[' + commas
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

// position of the byte after each element's separator
__global__ void element_starts_kernel(const uint8_t* __restrict__ json, unsigned long long n, const uint8_t* __restrict__ pf,
                                      const int* __restrict__ depth0, const unsigned int* __restrict__ ev_prefix, const Work* w,
                                      unsigned long long* __restrict__ elem_pos, unsigned long long n_chunks) {
  const unsigned long long c = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
  if (c >= n_chunks) return;
  const unsigned long long lo = w->ev_base, hi = lo + w->n_nodes;
  unsigned long long rank = ev_prefix[c];
  if (rank >= hi) return;
  visit_chunk(json, n, c, start_state(pf, c), depth0[c], [&](unsigned long long pos, uint32_t ch, int d) {
    if ((ch == '[' && d == 1) || (ch == ',' && d == 2)) {
      if (rank >= lo && rank < hi) elem_pos[rank - lo] = pos + 1;
      rank++;
    }
  });
}

// ---- per-node parser ---------------------------------------------------------------------------------
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

// value at the cursor as a capacity quantity: how many bytes the PNL value has
//   string -> decoded text; number -> its token, or nothing when it is zero (a falsy value is
//   skipped, check-gpu-node.py:188); true/false/null/containers -> nothing (int(str(.)) fails or falsy)
__device__ bool number_is_zero(const uint8_t* p, unsigned long long b, unsigned long long e) {
  for (unsigned long long i = b; i < e; i++) {
    const int ch = p[i];
    if (ch == 'e' || ch == 'E') break;
    if (ch >= '1' && ch <= '9') return false;
  }
  return true;
}
__device__ unsigned long long quantity_len(Cur c) {
  const int ch = c.peek();
  if (ch == '"') return string_len(c);
  if (ch == '-' || (ch >= '0' && ch <= '9')) {
    const unsigned long long b = c.i;
    skip_value(c);
    return number_is_zero(c.p, b, c.i) ? 0 : c.i - b;
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
    if (!number_is_zero(c.p, b, c.i))
      for (unsigned long long i = b; i < c.i; i++) *dst++ = c.p[i];
  }
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
      if (t.peek() == '"') tl = string_len(t);
    }
    if (spos != ~0ull) {
      s.i = spos;
      if (s.peek() == '"') sl = string_len(s);
    }
    if (!o.write) {
      o.n_cond++;
      o.text += tl + sl;
      o.maxlen = max(o.maxlen, max(tl, sl));
    } else {
      o.table(tl, sl);
      if (tl) string_copy(t, o.row + o.text_off);
      if (sl) string_copy(s, o.row + o.text_off + tl);
      o.text_off += tl + sl;
    }
  }
}

// What the measuring pass learns about a node and the writing pass reuses.
struct NodePlan {
  unsigned long long cap_pos, cond_pos;  // '{' of status.capacity and '[' of status.conditions, ~0 if none
  unsigned int n_cap, n_cond;
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

// row size in bytes (4-byte padded); bit 31 set when the extended form is needed
__global__ void measure_rows_kernel(const uint8_t* __restrict__ json, unsigned long long n, const unsigned long long* __restrict__ elem_pos,
                                    const Work* w, unsigned int* __restrict__ sizes, NodePlan* __restrict__ plans) {
  const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= w->n_nodes) return;
  RowOut o{};
  o.write = false;
  NodePlan pl;
  node_locate(json, min(n, w->items_end), elem_pos[i], pl);
  node_emit(json, min(n, w->items_end), pl, o);
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

__global__ void write_rows_kernel(const uint8_t* __restrict__ json, unsigned long long n, const NodePlan* __restrict__ plans,
                                  const Work* w, const unsigned int* __restrict__ sizes, const unsigned long long* __restrict__ offs,
                                  const unsigned long long* __restrict__ total, uint8_t* __restrict__ rows, unsigned long long rows_cap,
                                  uint32_t* __restrict__ row_ptr, long long* __restrict__ info) {
  const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned int nn = w->n_nodes;
  const bool fits = *total + 32 <= rows_cap && *total < 0xFFF00000ull;
  if (i == 0) {
    info[INFO_NODES] = nn;
    info[INFO_ROWS_BYTES] = fits ? (long long)*total : 0;
    info[INFO_STATUS] = w->status | (fits ? 0 : ST_ROWS_CAP);
    info[INFO_ITEMS_BEGIN] = (long long)w->items_begin;
    info[INFO_ITEMS_END] = (long long)w->items_end;
    row_ptr[nn] = fits ? (uint32_t)*total : 0u;
    for (int k = 1; k <= 4; k++) row_ptr[nn + k] = row_ptr[nn];
  }
  if (i >= nn) return;
  if (!fits) {
    row_ptr[i] = 0;
    return;
  }
  const unsigned long long off = offs[i];
  row_ptr[i] = (uint32_t)off;
  const unsigned int bytes = sizes[i] & 0x7FFFFFFFu;
  uint8_t* row = rows + off;
  for (unsigned int k = 0; k < bytes; k += 4) *reinterpret_cast<uint32_t*>(row + k) = 0u;   // rows are 4-byte aligned
  RowOut o{};
  o.write = true;
  o.row = row;
  o.ext = (sizes[i] >> 31) != 0;
  const NodePlan m = plans[i];
  const unsigned long long ne = (unsigned long long)m.n_cap + m.n_cond;
  if (o.ext) {
    row[0] = row[1] = 0xFF;
    uint32_t a = (uint32_t)m.n_cap, b = (uint32_t)m.n_cond;
    memcpy(row + 4, &a, 4);
    memcpy(row + 8, &b, 4);
    o.text_off = 12 + 8 * ne;
  } else {
    row[0] = (uint8_t)m.n_cap;
    row[1] = (uint8_t)m.n_cond;
    o.text_off = 2 + 2 * ne;
  }
  node_emit(json, min(n, w->items_end), m, o);
}

}  // namespace js
}  // namespace gnc
