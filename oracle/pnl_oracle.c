/* CPU restatement of the classification path over PNL-v2 bytes -- TEST INFRASTRUCTURE ONLY.
 *
 * Checker for the CUDA library at sizes where the Python restatement (oracle/cgn_oracle.py) is
 * too slow, and the multi-threaded CPU baseline bench.py may report.  Never linked into, loaded
 * by or called from the product (k8s-gpu-node-checker_b200/).
 *
 * Follows /root/reference/check-gpu-node.py ("cgn.py"):
 *   key scan            cgn.py:39-44, :186-187   four fixed keys, capacity only
 *   falsy skip          cgn.py:188               empty value == absent
 *   int(str(val))       cgn.py:192-195           CPython base-10 int grammar, errors drop the key
 *   is_ready            cgn.py:172-178           any (type "Ready", status "True"), first hit wins
 *   gpus / filter       cgn.py:201, :222-225     sum of kept values, keep gpus > 0, keep order
 *   counts / exit code  cgn.py:274-276, :289-293
 * Parity: pinned.  tests/test_oracle_pnl.py holds it to tests/golden/ (outputs of the reference
 * itself) on every template and fleet there.
 *
 * Values needing more than 18 significant digits do not fit the int64 outputs: the node is
 * marked DEFERRED and a record with the value texts' positions is emitted; the caller finishes
 * those few nodes with arbitrary-precision integers (the product does the same).
 */
#include <stdint.h>
#include <string.h>

#include "pnl_unicode.h"

#define F_GPU 1u
#define F_READY 2u
#define F_DEFERRED 4u
#define F_KEY0 8u /* bits 3..6: key k kept in gpu_breakdown */
#define F_MALFORMED 128u

enum { Q_ABSENT = 0, Q_OK = 1, Q_BIG = 2 };

typedef struct {
  uint32_t node;
  uint32_t off[4];
  uint32_t len[4];
} pnl_defer;

static const uint32_t ND_RUN[PNLO_ND_RUNS] = {PNLO_ND_RUN_LIST};
static const uint32_t WIDE_SPACE[PNLO_WIDE_SPACES] = {PNLO_WIDE_SPACE_LIST};

static const char* const KEYS[4] = {"nvidia.com/gpu", "amd.com/gpu", "gpu.intel.com/i915", "intel.com/gpu"};
static const uint32_t KEY_LEN[4] = {14, 11, 18, 13};

/* Generalised UTF-8 (surrogates allowed, as 'surrogatepass' writes them). Returns bytes used, 0 if bad. */
static int decode_utf8(const uint8_t* p, uint32_t avail, uint32_t* cp) {
  uint8_t b0 = p[0];
  if (b0 < 0x80) { *cp = b0; return 1; }
  if (b0 < 0xC2) return 0;
  if (b0 < 0xE0) {
    if (avail < 2 || (p[1] & 0xC0) != 0x80) return 0;
    *cp = ((uint32_t)(b0 & 0x1F) << 6) | (p[1] & 0x3F);
    return 2;
  }
  if (b0 < 0xF0) {
    if (avail < 3 || (p[1] & 0xC0) != 0x80 || (p[2] & 0xC0) != 0x80) return 0;
    *cp = ((uint32_t)(b0 & 0x0F) << 12) | ((uint32_t)(p[1] & 0x3F) << 6) | (p[2] & 0x3F);
    return *cp >= 0x800 ? 3 : 0;
  }
  if (b0 < 0xF5) {
    if (avail < 4 || (p[1] & 0xC0) != 0x80 || (p[2] & 0xC0) != 0x80 || (p[3] & 0xC0) != 0x80) return 0;
    *cp = ((uint32_t)(b0 & 0x07) << 18) | ((uint32_t)(p[1] & 0x3F) << 12) | ((uint32_t)(p[2] & 0x3F) << 6) |
          (p[3] & 0x3F);
    return (*cp >= 0x10000 && *cp <= 0x10FFFF) ? 4 : 0;
  }
  return 0;
}

/* Symbol classes after CPython's decimal-and-space transform. */
enum { C_DIGIT0 = 0 /* 0..9 */, C_WS = 10, C_US = 11, C_PLUS = 12, C_MINUS = 13, C_OTHER = 14, C_BAD = 15 };

static int classify_cp(uint32_t cp) {
  if (cp < 127) {
    if (cp >= '0' && cp <= '9') return (int)(cp - '0');
    if ((cp >= 9 && cp <= 13) || cp == 32) return C_WS;
    if (cp == '_') return C_US;
    if (cp == '+') return C_PLUS;
    if (cp == '-') return C_MINUS;
    return C_OTHER;
  }
  for (int i = 0; i < PNLO_WIDE_SPACES; i++)
    if (WIDE_SPACE[i] == cp) return C_WS;
  for (int i = 0; i < PNLO_ND_RUNS; i++)
    if (cp >= ND_RUN[i] && cp < ND_RUN[i] + 10) return (int)(cp - ND_RUN[i]);
  return C_BAD;
}

/* int(text) over UTF-8 bytes. Q_OK: *out is the value (|v| < 10^18). Q_BIG: valid but wider. */
static int parse_quantity(const uint8_t* s, uint32_t n, int max_digits, int64_t* out) {
  if (n == 0) return Q_ABSENT;
  uint32_t i = 0;
  int phase = 0; /* 0 leading ws, 1 after sign (need digit), 2 in digits, 3 after '_', 4 trailing ws */
  int neg = 0;
  uint64_t v = 0;
  uint32_t sig = 0;
  uint64_t digits = 0;
  while (i < n) {
    uint32_t cp;
    int used = decode_utf8(s + i, n - i, &cp);
    if (!used) return Q_ABSENT;
    i += (uint32_t)used;
    int c = classify_cp(cp);
    if (c == C_BAD || c == C_OTHER) return Q_ABSENT;
    if (c <= 9) {
      if (phase == 4) return Q_ABSENT;
      phase = 2;
      digits++;
      if (sig || c) {
        sig++;
        if (sig <= 18) v = v * 10 + (uint64_t)c;
      }
    } else if (c == C_WS) {
      if (phase == 0 || phase == 4) continue;
      if (phase != 2) return Q_ABSENT; /* after a sign or '_' */
      phase = 4;
    } else if (c == C_US) {
      if (phase != 2) return Q_ABSENT;
      phase = 3;
    } else { /* sign */
      if (phase != 0) return Q_ABSENT;
      neg = (c == C_MINUS);
      phase = 1;
    }
  }
  if (phase != 2 && phase != 4) return Q_ABSENT;
  if (max_digits > 0 && digits > (uint64_t)max_digits) return Q_ABSENT;
  if (sig > 18) return Q_BIG;
  *out = neg ? -(int64_t)v : (int64_t)v;
  return Q_OK;
}

typedef struct {
  uint8_t flags;
  int64_t vals[4];
  int64_t total;
  uint32_t voff[4], vlen[4];
} node_out;

static uint32_t rd32(const uint8_t* p) {
  uint32_t t;
  memcpy(&t, p, 4);
  return t;
}

/* Row layout (PNL-v2, k8s-gpu-node-checker_b200/pnl.py): a length table, then the texts.
 *   row  := n_cap:u8 n_cond:u8 (klen:u8 vlen:u8){n_cap+n_cond} text pad
 *   xrow := 0xFF 0xFF 0 0 n_cap:u32 n_cond:u32 (klen:u32 vlen:u32){n_cap+n_cond} text pad
 * An entry whose text would end past the row makes the row malformed and ends the walk. */
static void classify_row(const uint8_t* rows, uint32_t s, uint32_t e, int max_digits, node_out* o) {
  int st[4] = {Q_ABSENT, Q_ABSENT, Q_ABSENT, Q_ABSENT};
  memset(o, 0, sizeof(*o));
  for (int k = 0; k < 4; k++) o->voff[k] = 0xFFFFFFFFu;
  uint32_t flags = 0;
  if (e - s < 2) {
    o->flags = (e != s) ? F_MALFORMED : 0;
    return;
  }
  const int ext = rows[s] == 0xFF;
  uint64_t n_cap, n_cond, tab, stride;
  if (ext) {
    if (e - s < 12) { o->flags = F_MALFORMED; return; }
    n_cap = rd32(rows + s + 4);
    n_cond = rd32(rows + s + 8);
    tab = (uint64_t)s + 12;
    stride = 8;
  } else {
    n_cap = rows[s];
    n_cond = rows[s + 1];
    tab = (uint64_t)s + 2;
    stride = 2;
  }
  const uint64_t n = n_cap + n_cond;
  uint64_t off = tab + stride * n;
  if (off > e) { o->flags = F_MALFORMED; return; }
  for (uint64_t i = 0; i < n; i++) {
    const uint64_t kl = ext ? rd32(rows + tab + 8 * i) : rows[tab + 2 * i];
    const uint64_t vl = ext ? rd32(rows + tab + 8 * i + 4) : rows[tab + 2 * i + 1];
    if (off + kl + vl > e) { flags |= F_MALFORMED; break; }
    const uint8_t* key = rows + off;
    const uint8_t* val = rows + off + kl;
    if (i < n_cap) {
      /* capacity: cgn.py:186-195. A repeated key overwrites (dict / json.loads last-wins). */
      for (int k = 0; k < 4; k++) {
        if (kl == KEY_LEN[k] && memcmp(key, KEYS[k], KEY_LEN[k]) == 0) {
          int64_t v = 0;
          st[k] = parse_quantity(val, (uint32_t)vl, max_digits, &v);
          o->vals[k] = st[k] == Q_OK ? v : 0;
          o->voff[k] = st[k] == Q_ABSENT ? 0xFFFFFFFFu : (uint32_t)(off + kl);
          o->vlen[k] = st[k] == Q_ABSENT ? 0 : (uint32_t)vl;
        }
      }
    } else if (kl == 5 && vl == 4 && memcmp(key, "Ready", 5) == 0 && memcmp(val, "True", 4) == 0) {
      flags |= F_READY; /* conditions: cgn.py:175-177, first hit returns */
      break;
    }
    off += kl + vl;
  }
  int big = 0;
  int64_t total = 0;
  for (int k = 0; k < 4; k++) {
    if (st[k] != Q_ABSENT) flags |= (F_KEY0 << k);
    if (st[k] == Q_BIG) big = 1;
    total += o->vals[k]; /* |v| < 10^18 each: no overflow */
  }
  if (big) flags |= F_DEFERRED;
  else if (total > 0) flags |= F_GPU; /* cgn.py:222 */
  o->total = total;
  o->flags = (uint8_t)flags;
}

/* counts[16]: 0 GPU nodes, 1 Ready GPU nodes, 2 deferred, 3 nodes scanned, 4..7 per-key totals (low
 * 64 bits), 8..11 (high 64 bits, two's complement 128-bit), 12 malformed rows, 13 defer records that
 * did not fit, 14..15 zero. Outputs sized for n nodes. Returns 0. */
int pnl_classify(const uint8_t* rows, const uint32_t* row_ptr, uint32_t n, uint32_t node_base, int max_digits,
                 uint8_t* flags, uint32_t* gpu_idx, int64_t* gpu_total, int64_t* gpu_vals, int64_t* counts,
                 pnl_defer* defer, uint32_t defer_cap) {
  uint64_t g = 0, r = 0, d = 0, bad = 0, dropped = 0;
  __int128 tot[4] = {0, 0, 0, 0};
  for (uint32_t i = 0; i < n; i++) {
    node_out o;
    classify_row(rows, row_ptr[i], row_ptr[i + 1], max_digits, &o);
    flags[i] = o.flags;
    if (o.flags & F_MALFORMED) bad++;
    if (o.flags & F_DEFERRED) {
      if (d < defer_cap) {
        defer[d].node = node_base + i;
        for (int k = 0; k < 4; k++) { defer[d].off[k] = o.voff[k]; defer[d].len[k] = o.vlen[k]; }
      } else {
        dropped++;
      }
      d++;
    }
    if (o.flags & F_GPU) {
      gpu_idx[g] = node_base + i;
      gpu_total[g] = o.total;
      for (int k = 0; k < 4; k++) { gpu_vals[g * 4 + k] = o.vals[k]; tot[k] += o.vals[k]; }
      g++;
      if (o.flags & F_READY) r++;
    }
  }
  memset(counts, 0, 16 * sizeof(int64_t));
  counts[0] = (int64_t)g;
  counts[1] = (int64_t)r;
  counts[2] = (int64_t)d;
  counts[3] = (int64_t)n;
  for (int k = 0; k < 4; k++) {
    counts[4 + k] = (int64_t)(uint64_t)tot[k];
    counts[8 + k] = (int64_t)(tot[k] >> 64);
  }
  counts[12] = (int64_t)bad;
  counts[13] = (int64_t)dropped;
  return 0;
}

/* Flags only, for timing the scan without the compaction writes. */
int pnl_flags_only(const uint8_t* rows, const uint32_t* row_ptr, uint32_t n, int max_digits, uint8_t* flags) {
  for (uint32_t i = 0; i < n; i++) {
    node_out o;
    classify_row(rows, row_ptr[i], row_ptr[i + 1], max_digits, &o);
    flags[i] = o.flags;
  }
  return 0;
}

int pnl_parse_quantity(const uint8_t* s, uint32_t n, int max_digits, int64_t* out) {
  return parse_quantity(s, n, max_digits, out);
}
