/* _fastpack -- CPython extension: V1Node objects -> PNL-v2 columns.
 *
 * The same marshalling as pnl.node_row / pnl.encode_row (the attribute reads of
 * check-gpu-node.py:173-176, :183-187), without a Python-level loop per entry:
 *     pack_nodes(nodes, condition_class) -> (rows: bytes, row_ptr: bytes /u32 little-endian/, n_nodes)
 * rows carries ROWS_PAD zero bytes and row_ptr ROWPTR_PAD repeats of the total after the last valid
 * element, as the device buffers require.  Type adaptation only; nothing is classified here.
 */
#define PY_SSIZE_T_CLEAN
#include <Python.h>
#include <stdint.h>
#include <string.h>

#define ROWS_PAD 32
#define ROWPTR_PAD 4

typedef struct {
  char* p;
  size_t n, cap;
} buf_t;

static int buf_reserve(buf_t* b, size_t extra) {
  if (b->n + extra <= b->cap) return 0;
  size_t cap = b->cap ? b->cap : 1 << 16;
  while (cap < b->n + extra) cap *= 2;
  char* q = (char*)realloc(b->p, cap);
  if (!q) {
    PyErr_NoMemory();
    return -1;
  }
  b->p = q;
  b->cap = cap;
  return 0;
}
static int buf_put(buf_t* b, const void* src, size_t len) {
  if (buf_reserve(b, len)) return -1;
  memcpy(b->p + b->n, src, len);
  b->n += len;
  return 0;
}

/* UTF-8 bytes of a str with 'surrogatepass' semantics; *owner must be DECREF'd when non-NULL. */
static int utf8_of(PyObject* s, const char** data, Py_ssize_t* len, PyObject** owner) {
  *owner = NULL;
  *data = PyUnicode_AsUTF8AndSize(s, len);
  if (*data) return 0;
  PyErr_Clear(); /* lone surrogates: the strict codec refuses them */
  *owner = PyUnicode_AsEncodedString(s, "utf-8", "surrogatepass");
  if (!*owner) return -1;
  *data = PyBytes_AS_STRING(*owner);
  *len = PyBytes_GET_SIZE(*owner);
  return 0;
}

typedef struct {
  const char* k;
  Py_ssize_t kl;
  const char* v;
  Py_ssize_t vl;
  PyObject* ko;
  PyObject* vo; /* owners (may be NULL) */
} entry_t;

typedef struct {
  entry_t* e;
  size_t n, cap;
} entries_t;

static int entries_push(entries_t* es, entry_t* en) {
  if (es->n == es->cap) {
    size_t cap = es->cap ? es->cap * 2 : 32;
    entry_t* q = (entry_t*)realloc(es->e, cap * sizeof(entry_t));
    if (!q) {
      PyErr_NoMemory();
      return -1;
    }
    es->e = q;
    es->cap = cap;
  }
  es->e[es->n++] = *en;
  return 0;
}
static void entries_clear(entries_t* es) {
  for (size_t i = 0; i < es->n; i++) {
    Py_XDECREF(es->e[i].ko);
    Py_XDECREF(es->e[i].vo);
  }
  es->n = 0;
}

/* A capacity value as the device will parse it: str as is; otherwise "" when falsy, str(value) when
 * truthy, "?" when either raises (the reference's try block swallows it, check-gpu-node.py:188-195). */
static int quantity_of(PyObject* v, entry_t* en) {
  if (PyUnicode_Check(v)) return utf8_of(v, &en->v, &en->vl, &en->vo);
  int truth = PyObject_IsTrue(v);
  if (truth == 0) {
    en->v = "";
    en->vl = 0;
    return 0;
  }
  if (truth > 0) {
    PyObject* s = PyObject_Str(v);
    if (s) {
      const char* d;
      Py_ssize_t l;
      PyObject* o;
      if (utf8_of(s, &d, &l, &o) == 0) {
        if (o) {
          Py_DECREF(s);
          en->vo = o;
        } else {
          en->vo = s; /* the str owns its cached UTF-8 */
        }
        en->v = d;
        en->vl = l;
        return 0;
      }
      Py_DECREF(s);
    }
  }
  PyErr_Clear();
  en->v = "?";
  en->vl = 1;
  return 0;
}

static PyObject *s_status, *s_capacity, *s_conditions, *s_type, *s_items;

static int collect(PyObject* node, PyObject* cond_cls, entries_t* caps, entries_t* conds) {
  PyObject* st = PyObject_GetAttr(node, s_status);
  if (!st) return -1;
  int t = PyObject_IsTrue(st);
  if (t <= 0) {
    Py_DECREF(st);
    return t;
  }
  PyObject* cap = PyObject_GetAttr(st, s_capacity);
  if (!cap) {
    Py_DECREF(st);
    return -1;
  }
  t = PyObject_IsTrue(cap);
  if (t < 0) goto fail_cap;
  if (t) {
    PyObject* items = PyDict_CheckExact(cap) ? NULL : PyObject_CallMethodNoArgs(cap, s_items);
    if (!PyDict_CheckExact(cap) && !items) goto fail_cap;
    PyObject *k, *v;
    if (!items) {
      Py_ssize_t pos = 0;
      while (PyDict_Next(cap, &pos, &k, &v)) {
        if (!PyUnicode_Check(k)) continue; /* cannot equal any of the four str keys */
        entry_t en = {0};
        if (utf8_of(k, &en.k, &en.kl, &en.ko) || quantity_of(v, &en) || entries_push(caps, &en)) goto fail_cap;
      }
    } else {
      PyObject* it = PyObject_GetIter(items);
      Py_DECREF(items);
      if (!it) goto fail_cap;
      PyObject* pair;
      while ((pair = PyIter_Next(it))) {
        if (!PyTuple_Check(pair) || PyTuple_GET_SIZE(pair) != 2) {
          Py_DECREF(pair);
          continue;
        }
        k = PyTuple_GET_ITEM(pair, 0);
        v = PyTuple_GET_ITEM(pair, 1);
        if (PyUnicode_Check(k)) {
          entry_t en = {0};
          /* keep the pair alive through the owners: take strong refs */
          if (utf8_of(k, &en.k, &en.kl, &en.ko) || quantity_of(v, &en)) {
            Py_DECREF(pair);
            Py_DECREF(it);
            goto fail_cap;
          }
          if (!en.ko) {
            Py_INCREF(k);
            en.ko = k;
          }
          if (!en.vo && PyUnicode_Check(v)) {
            Py_INCREF(v);
            en.vo = v;
          }
          if (entries_push(caps, &en)) {
            Py_DECREF(pair);
            Py_DECREF(it);
            goto fail_cap;
          }
        }
        Py_DECREF(pair);
      }
      Py_DECREF(it);
      if (PyErr_Occurred()) goto fail_cap;
    }
  }
  Py_DECREF(cap);
  {
    PyObject* cl = PyObject_GetAttr(st, s_conditions);
    if (!cl) {
      Py_DECREF(st);
      return -1;
    }
    t = PyObject_IsTrue(cl);
    if (t < 0) {
      Py_DECREF(cl);
      Py_DECREF(st);
      return -1;
    }
    if (t) {
      PyObject* it = PyObject_GetIter(cl);
      if (!it) {
        Py_DECREF(cl);
        Py_DECREF(st);
        return -1;
      }
      PyObject* c;
      while ((c = PyIter_Next(it))) {
        int is = PyObject_IsInstance(c, cond_cls); /* check-gpu-node.py:176 */
        if (is < 0) {
          Py_DECREF(c);
          Py_DECREF(it);
          Py_DECREF(cl);
          Py_DECREF(st);
          return -1;
        }
        if (is) {
          entry_t en = {0};
          PyObject* ty = PyObject_GetAttr(c, s_type);
          PyObject* su = ty ? PyObject_GetAttr(c, s_status) : NULL;
          int bad = !ty || !su;
          if (!bad) {
            if (PyUnicode_Check(ty)) {
              bad = utf8_of(ty, &en.k, &en.kl, &en.ko);
              if (!bad && !en.ko) {
                Py_INCREF(ty);
                en.ko = ty;
              }
            } else {
              en.k = "";
            }
            if (!bad && PyUnicode_Check(su)) {
              bad = utf8_of(su, &en.v, &en.vl, &en.vo);
              if (!bad && !en.vo) {
                Py_INCREF(su);
                en.vo = su;
              }
            } else if (!bad) {
              en.v = "";
            }
          }
          Py_XDECREF(ty);
          Py_XDECREF(su);
          if (bad || entries_push(conds, &en)) {
            Py_DECREF(c);
            Py_DECREF(it);
            Py_DECREF(cl);
            Py_DECREF(st);
            return -1;
          }
        }
        Py_DECREF(c);
      }
      Py_DECREF(it);
      if (PyErr_Occurred()) {
        Py_DECREF(cl);
        Py_DECREF(st);
        return -1;
      }
    }
    Py_DECREF(cl);
  }
  Py_DECREF(st);
  return 0;
fail_cap:
  Py_DECREF(cap);
  Py_DECREF(st);
  return -1;
}

static int emit_row(buf_t* rows, entries_t* caps, entries_t* conds) {
  const size_t n = caps->n + conds->n;
  int ext = caps->n >= 255 || conds->n >= 255;
  size_t text = 0;
  for (int part = 0; part < 2; part++) {
    entries_t* es = part ? conds : caps;
    for (size_t i = 0; i < es->n; i++) {
      if (es->e[i].kl >= 255 || es->e[i].vl >= 255) ext = 1;
      text += (size_t)es->e[i].kl + (size_t)es->e[i].vl;
    }
  }
  const size_t head = ext ? 12 + 8 * n : 2 + 2 * n;
  const size_t total = (head + text + 3) & ~(size_t)3;
  if (buf_reserve(rows, total)) return -1;
  unsigned char* row = (unsigned char*)rows->p + rows->n;
  memset(row, 0, total);
  if (ext) {
    uint32_t a = (uint32_t)caps->n, b = (uint32_t)conds->n;
    row[0] = row[1] = 0xFF;
    memcpy(row + 4, &a, 4);
    memcpy(row + 8, &b, 4);
  } else {
    row[0] = (unsigned char)caps->n;
    row[1] = (unsigned char)conds->n;
  }
  size_t k = 0, off = head;
  for (int part = 0; part < 2; part++) {
    entries_t* es = part ? conds : caps;
    for (size_t i = 0; i < es->n; i++, k++) {
      const entry_t* en = &es->e[i];
      if (ext) {
        uint32_t a = (uint32_t)en->kl, b = (uint32_t)en->vl;
        memcpy(row + 12 + 8 * k, &a, 4);
        memcpy(row + 12 + 8 * k + 4, &b, 4);
      } else {
        row[2 + 2 * k] = (unsigned char)en->kl;
        row[2 + 2 * k + 1] = (unsigned char)en->vl;
      }
      memcpy(row + off, en->k, (size_t)en->kl);
      memcpy(row + off + en->kl, en->v, (size_t)en->vl);
      off += (size_t)en->kl + (size_t)en->vl;
    }
  }
  rows->n += total;
  return 0;
}

static PyObject* pack_nodes(PyObject* self, PyObject* args) {
  PyObject *nodes, *cond_cls;
  if (!PyArg_ParseTuple(args, "OO", &nodes, &cond_cls)) return NULL;
  PyObject* seq = PySequence_Fast(nodes, "nodes must be a sequence");
  if (!seq) return NULL;
  const Py_ssize_t n = PySequence_Fast_GET_SIZE(seq);
  buf_t rows = {0}, ptr = {0};
  entries_t caps = {0}, conds = {0};
  PyObject* result = NULL;
  uint32_t zero = 0;
  if (buf_put(&ptr, &zero, 4)) goto done;
  for (Py_ssize_t i = 0; i < n; i++) {
    entries_clear(&caps);
    entries_clear(&conds);
    if (collect(PySequence_Fast_GET_ITEM(seq, i), cond_cls, &caps, &conds) < 0) goto done;
    if (emit_row(&rows, &caps, &conds)) goto done;
    if (rows.n > 0xFFF00000u) {
      PyErr_SetString(PyExc_ValueError, "PNL-v2 rows are limited to 4 GiB - 1 MiB per list; split the fleet");
      goto done;
    }
    uint32_t end = (uint32_t)rows.n;
    if (buf_put(&ptr, &end, 4)) goto done;
  }
  {
    uint32_t end = (uint32_t)rows.n;
    char pad[ROWS_PAD] = {0};
    if (buf_put(&rows, pad, ROWS_PAD)) goto done;
    for (int k = 0; k < ROWPTR_PAD; k++)
      if (buf_put(&ptr, &end, 4)) goto done;
    result = Py_BuildValue("(y#y#n)", rows.p ? rows.p : "", (Py_ssize_t)rows.n, ptr.p, (Py_ssize_t)ptr.n, n);
  }
done:
  entries_clear(&caps);
  entries_clear(&conds);
  free(caps.e);
  free(conds.e);
  free(rows.p);
  free(ptr.p);
  Py_DECREF(seq);
  return result;
}


/* ------------------------------------------------------------------------------------------------
 * emit_summary_json: the `--json` document of check-gpu-node.py:273-279, i.e. exactly the text of
 *   json.dumps({"total_nodes": G, "ready_nodes": R, "nodes": [...]}, ensure_ascii=False, indent=2)
 * written straight from the device's decisions and each GPU node's payload, without building the
 * list of dicts or a Python-level loop per node.
 *   emit_summary_json(nodes, gpu_idx: int64 buffer[G], flags: uint8 buffer[N], totals: int64 buffer[G],
 *                     vals: int64 buffer[G*4], wide: dict node -> (gpus:int, breakdown:dict) | None, node_base)
 * Returns str, or None when some payload value is of a type this writer does not format (the caller
 * then uses json.dumps). */
static const char* const KEY_NAMES[4] = {"nvidia.com/gpu", "amd.com/gpu", "gpu.intel.com/i915", "intel.com/gpu"};

static int put_s(buf_t* b, const char* s) { return buf_put(b, s, strlen(s)); }
/* json.dumps string rules with ensure_ascii=False */
static int put_json_str(buf_t* b, PyObject* s) {
  const char* d;
  Py_ssize_t n;
  PyObject* owner;
  if (utf8_of(s, &d, &n, &owner)) return -1;
  int rc = buf_put(b, "\"", 1);
  Py_ssize_t run = 0;
  for (Py_ssize_t i = 0; i < n && !rc; i++) {
    const unsigned char c = (unsigned char)d[i];
    const char* esc = NULL;
    char tmp[8];
    switch (c) {
      case '"': esc = "\\\""; break;
      case '\\': esc = "\\\\"; break;
      case '\n': esc = "\\n"; break;
      case '\r': esc = "\\r"; break;
      case '\t': esc = "\\t"; break;
      case '\b': esc = "\\b"; break;
      case '\f': esc = "\\f"; break;
      default:
        if (c < 0x20) {
          snprintf(tmp, sizeof tmp, "\\u%04x", c);
          esc = tmp;
        }
    }
    if (esc) {
      rc = buf_put(b, d + run, (size_t)(i - run)) || put_s(b, esc);
      run = i + 1;
    }
  }
  if (!rc) rc = buf_put(b, d + run, (size_t)(n - run)) || buf_put(b, "\"", 1);
  Py_XDECREF(owner);
  return rc;
}
/* 1 written, 0 unsupported type, -1 error */
static int put_scalar(buf_t* b, PyObject* v) {
  if (v == Py_None) return put_s(b, "null") ? -1 : 1;
  if (PyUnicode_Check(v)) return put_json_str(b, v) ? -1 : 1;
  if (v == Py_True) return put_s(b, "true") ? -1 : 1;
  if (v == Py_False) return put_s(b, "false") ? -1 : 1;
  if (PyLong_CheckExact(v)) {
    PyObject* t = PyObject_Str(v);
    if (!t) return -1;
    Py_ssize_t n;
    const char* d = PyUnicode_AsUTF8AndSize(t, &n);
    int rc = d ? buf_put(b, d, (size_t)n) : -1;
    Py_DECREF(t);
    return rc ? -1 : 1;
  }
  return 0;
}
static int put_i64(buf_t* b, long long v) {
  char tmp[32];
  snprintf(tmp, sizeof tmp, "%lld", v);
  return put_s(b, tmp);
}

static PyObject *s_metadata, *s_name, *s_labels, *s_spec, *s_taints, *s_key, *s_value, *s_effect;

static PyObject* emit_summary_json(PyObject* self, PyObject* args) {
  PyObject *nodes, *wide;
  Py_buffer idx, flags, totals, vals;
  long long node_base;
  if (!PyArg_ParseTuple(args, "Oy*y*y*y*OL", &nodes, &idx, &flags, &totals, &vals, &wide, &node_base)) return NULL;
  PyObject* seq = PySequence_Fast(nodes, "nodes must be a sequence");
  PyObject* result = NULL;
  buf_t b = {0};
  int unsupported = 0;
  if (!seq) goto done;
  {
    const Py_ssize_t g = idx.len / 8, n_nodes = PySequence_Fast_GET_SIZE(seq);
    const long long* gi = (const long long*)idx.buf;
    const long long* gt = (const long long*)totals.buf;
    const long long* gv = (const long long*)vals.buf;
    const unsigned char* fl = (const unsigned char*)flags.buf;
    long long ready = 0;
    for (Py_ssize_t i = 0; i < g; i++) ready += (fl[gi[i] - node_base] >> 1) & 1;
    if (put_s(&b, "{\n  \"total_nodes\": ") || put_i64(&b, (long long)g) || put_s(&b, ",\n  \"ready_nodes\": ") ||
        put_i64(&b, ready) || put_s(&b, ",\n  \"nodes\": ["))
      goto done;
    if (g == 0 && put_s(&b, "]")) goto done;
    for (Py_ssize_t i = 0; i < g && !unsupported; i++) {
      const long long node_i = gi[i];
      if (node_i - node_base < 0 || node_i - node_base >= n_nodes) {
        PyErr_SetString(PyExc_IndexError, "gpu_idx outside the node list");
        goto done;
      }
      PyObject* node = PySequence_Fast_GET_ITEM(seq, node_i - node_base);
      const unsigned f = fl[node_i - node_base];
      if (put_s(&b, i ? ",\n    {\n      \"name\": " : "\n    {\n      \"name\": ")) goto done;
      /* name / labels: check-gpu-node.py:203, :207 */
      PyObject* md = PyObject_GetAttr(node, s_metadata);
      if (!md) goto done;
      int md_true = PyObject_IsTrue(md);
      PyObject* labels = NULL;
      if (md_true > 0) {
        PyObject* nm = PyObject_GetAttr(md, s_name);
        labels = nm ? PyObject_GetAttr(md, s_labels) : NULL;
        int rc = (nm && labels) ? put_scalar(&b, nm) : -1;
        Py_XDECREF(nm);
        if (rc < 0) { Py_XDECREF(labels); Py_DECREF(md); goto done; }
        if (rc == 0) unsupported = 1;
      } else if (md_true == 0) {
        if (put_s(&b, "\"\"")) { Py_DECREF(md); goto done; }
      } else { Py_DECREF(md); goto done; }
      Py_DECREF(md);
      if (put_s(&b, ",\n      \"ready\": ") || put_s(&b, (f & 2) ? "true" : "false") || put_s(&b, ",\n      \"gpus\": ")) { Py_XDECREF(labels); goto done; }
      /* gpus / gpu_breakdown: from the device, or from the bignum table */
      PyObject* wrec = NULL;
      if (wide != Py_None && PyDict_Size(wide)) {
        PyObject* key = PyLong_FromLongLong(node_i);
        wrec = key ? PyDict_GetItemWithError(wide, key) : NULL; /* borrowed */
        Py_XDECREF(key);
        if (!wrec && PyErr_Occurred()) { Py_XDECREF(labels); goto done; }
      }
      if (wrec) {
        PyObject* gp = PyTuple_GetItem(wrec, 0);
        PyObject* bd = PyTuple_GetItem(wrec, 1);
        if (!gp || !bd || put_scalar(&b, gp) != 1 || put_s(&b, ",\n      \"gpu_breakdown\": {")) { Py_XDECREF(labels); goto done; }
        PyObject *k, *v;
        Py_ssize_t pos = 0, cnt = 0;
        while (PyDict_Next(bd, &pos, &k, &v)) {
          if (put_s(&b, cnt++ ? ",\n        " : "\n        ") || put_json_str(&b, k) || put_s(&b, ": ") || put_scalar(&b, v) != 1) { Py_XDECREF(labels); goto done; }
        }
        if (put_s(&b, cnt ? "\n      }" : "}")) { Py_XDECREF(labels); goto done; }
      } else {
        if (put_i64(&b, gt[i]) || put_s(&b, ",\n      \"gpu_breakdown\": {")) { Py_XDECREF(labels); goto done; }
        int cnt = 0;
        for (int k = 0; k < 4; k++)
          if (f & (8u << k)) {
            if (put_s(&b, cnt++ ? ",\n        \"" : "\n        \"") || put_s(&b, KEY_NAMES[k]) || put_s(&b, "\": ") || put_i64(&b, gv[4 * i + k])) { Py_XDECREF(labels); goto done; }
          }
        if (put_s(&b, cnt ? "\n      }" : "}")) { Py_XDECREF(labels); goto done; }
      }
      /* labels */
      if (put_s(&b, ",\n      \"labels\": {")) { Py_XDECREF(labels); goto done; }
      {
        Py_ssize_t cnt = 0;
        if (labels) {
          int lt = PyObject_IsTrue(labels);
          if (lt < 0) { Py_DECREF(labels); goto done; }
          if (lt) {
            if (!PyDict_Check(labels)) unsupported = 1;
            else {
              PyObject *k, *v;
              Py_ssize_t pos = 0;
              while (PyDict_Next(labels, &pos, &k, &v) && !unsupported) {
                if (!PyUnicode_Check(k)) { unsupported = 1; break; }
                if (put_s(&b, cnt++ ? ",\n        " : "\n        ") || put_json_str(&b, k) || put_s(&b, ": ")) { Py_DECREF(labels); goto done; }
                int rc = put_scalar(&b, v);
                if (rc < 0) { Py_DECREF(labels); goto done; }
                if (rc == 0) unsupported = 1;
              }
            }
          }
          Py_DECREF(labels);
        }
        if (put_s(&b, cnt ? "\n      }" : "}")) goto done;
      }
      /* taints: check-gpu-node.py:208-211 */
      if (put_s(&b, ",\n      \"taints\": [")) goto done;
      {
        Py_ssize_t cnt = 0;
        PyObject* spec = PyObject_GetAttr(node, s_spec);
        if (!spec) goto done;
        int st = PyObject_IsTrue(spec);
        if (st < 0) { Py_DECREF(spec); goto done; }
        PyObject* taints = NULL;
        if (st) {
          taints = PyObject_GetAttr(spec, s_taints);
          if (!taints) PyErr_Clear(); /* getattr(node.spec, "taints", None) */
        }
        Py_DECREF(spec);
        if (taints) {
          int tt = PyObject_IsTrue(taints);
          if (tt < 0) { Py_DECREF(taints); goto done; }
          if (tt) {
            PyObject* it = PyObject_GetIter(taints);
            if (!it) { Py_DECREF(taints); goto done; }
            PyObject* t;
            while ((t = PyIter_Next(it)) && !unsupported) {
              PyObject* parts[3] = {PyObject_GetAttr(t, s_key), NULL, NULL};
              parts[1] = parts[0] ? PyObject_GetAttr(t, s_value) : NULL;
              parts[2] = parts[1] ? PyObject_GetAttr(t, s_effect) : NULL;
              int bad = !parts[2];
              static const char* const names[3] = {"\"key\": ", "\"value\": ", "\"effect\": "};
              if (!bad) bad = put_s(&b, cnt++ ? ",\n        {" : "\n        {");
              for (int q = 0; q < 3 && !bad && !unsupported; q++) {
                bad = put_s(&b, q ? ",\n          " : "\n          ") || put_s(&b, names[q]);
                if (!bad) {
                  int rc = put_scalar(&b, parts[q]);
                  if (rc < 0) bad = 1;
                  if (rc == 0) unsupported = 1;
                }
              }
              if (!bad) bad = put_s(&b, "\n        }");
              for (int q = 0; q < 3; q++) Py_XDECREF(parts[q]);
              Py_DECREF(t);
              if (bad) { Py_DECREF(it); Py_DECREF(taints); goto done; }
            }
            Py_DECREF(it);
            if (PyErr_Occurred()) { Py_DECREF(taints); goto done; }
          }
          Py_DECREF(taints);
        }
        if (put_s(&b, cnt ? "\n      ]" : "]")) goto done;
      }
      if (put_s(&b, "\n    }")) goto done;
    }
    if (g && put_s(&b, "\n  ]")) goto done;
    if (put_s(&b, "\n}")) goto done;
  }
  if (unsupported) {
    result = Py_None;
    Py_INCREF(result);
  } else {
    result = PyUnicode_DecodeUTF8(b.p, (Py_ssize_t)b.n, "surrogatepass");
  }
done:
  free(b.p);
  Py_XDECREF(seq);
  PyBuffer_Release(&idx);
  PyBuffer_Release(&flags);
  PyBuffer_Release(&totals);
  PyBuffer_Release(&vals);
  return result;
}


/* ------------------------------------------------------------------------------------------------
 * build_records: (gpu_nodes, ready_gpu_nodes) exactly as check-gpu-node.py:199-226 returns them -- one
 * six-key dict per GPU node (name, ready, gpus, gpu_breakdown, labels, taints, in that order; :202-212),
 * GPU nodes in fleet order, the Ready list holding the same dict objects (:222-225) -- from the device's
 * decisions plus each node's pass-through payload, without a Python-level loop per node.
 *   build_records(nodes, gpu_idx: int64 buffer[G], flags: uint8 buffer[N], totals: int64 buffer[G],
 *                 vals: int64 buffer[G*4], wide: dict node -> (gpus:int, breakdown:dict) | None, node_base)
 */
static PyObject *s_ready, *s_gpus, *s_gpu_breakdown, *s_key_names[4];

static PyObject* build_records(PyObject* self, PyObject* args) {
  PyObject *nodes, *wide;
  Py_buffer idx, flags, totals, vals;
  long long node_base;
  if (!PyArg_ParseTuple(args, "Oy*y*y*y*OL", &nodes, &idx, &flags, &totals, &vals, &wide, &node_base)) return NULL;
  PyObject* seq = PySequence_Fast(nodes, "nodes must be a sequence");
  PyObject *gpu_list = NULL, *ready_list = NULL, *result = NULL;
  if (!seq) goto done;
  {
    const Py_ssize_t g = idx.len / 8, n_nodes = PySequence_Fast_GET_SIZE(seq);
    const long long* gi = (const long long*)idx.buf;
    const long long* gt = (const long long*)totals.buf;
    const long long* gv = (const long long*)vals.buf;
    const unsigned char* fl = (const unsigned char*)flags.buf;
    const int have_wide = wide != Py_None && PyDict_Size(wide) > 0;
    gpu_list = PyList_New(g);
    ready_list = PyList_New(0);
    if (!gpu_list || !ready_list) goto done;
    for (Py_ssize_t i = 0; i < g; i++) {
      const long long at = gi[i] - node_base;
      if (at < 0 || at >= n_nodes || at >= flags.len) {
        PyErr_SetString(PyExc_IndexError, "gpu_idx outside the node list");
        goto done;
      }
      PyObject* node = PySequence_Fast_GET_ITEM(seq, at);
      const unsigned f = fl[at];
      PyObject *name = NULL, *labels = NULL, *taints = NULL, *gpus = NULL, *bd = NULL, *rec = NULL;
      int ok = 0;
      /* name / labels: check-gpu-node.py:203, :207 */
      PyObject* md = PyObject_GetAttr(node, s_metadata);
      if (!md) goto node_done;
      {
        const int md_true = PyObject_IsTrue(md);
        if (md_true < 0) { Py_DECREF(md); goto node_done; }
        if (md_true) {
          name = PyObject_GetAttr(md, s_name);
          labels = name ? PyObject_GetAttr(md, s_labels) : NULL;
          if (labels) {
            const int lt = PyObject_IsTrue(labels);
            if (lt < 0) { Py_DECREF(md); goto node_done; }
            if (!lt) { Py_DECREF(labels); labels = PyDict_New(); }
          }
        } else {
          name = PyUnicode_FromString("");
          labels = PyDict_New();
        }
        Py_DECREF(md);
        if (!name || !labels) goto node_done;
      }
      /* taints: check-gpu-node.py:208-211 */
      taints = PyList_New(0);
      if (!taints) goto node_done;
      {
        PyObject* spec = PyObject_GetAttr(node, s_spec);
        if (!spec) goto node_done;
        const int st = PyObject_IsTrue(spec);
        PyObject* src = NULL;
        if (st > 0) {
          src = PyObject_GetAttr(spec, s_taints);
          if (!src) PyErr_Clear(); /* getattr(node.spec, "taints", None) */
        }
        Py_DECREF(spec);
        if (st < 0) goto node_done;
        if (src) {
          const int tt = PyObject_IsTrue(src);
          if (tt < 0) { Py_DECREF(src); goto node_done; }
          if (tt) {
            PyObject* it = PyObject_GetIter(src);
            if (!it) { Py_DECREF(src); goto node_done; }
            PyObject* t;
            int bad = 0;
            while (!bad && (t = PyIter_Next(it))) {
              PyObject* k = PyObject_GetAttr(t, s_key);
              PyObject* v = k ? PyObject_GetAttr(t, s_value) : NULL;
              PyObject* e = v ? PyObject_GetAttr(t, s_effect) : NULL;
              PyObject* d = e ? PyDict_New() : NULL;
              bad = !d || PyDict_SetItem(d, s_key, k) || PyDict_SetItem(d, s_value, v) || PyDict_SetItem(d, s_effect, e) ||
                    PyList_Append(taints, d);
              Py_XDECREF(k); Py_XDECREF(v); Py_XDECREF(e); Py_XDECREF(d);
              Py_DECREF(t);
            }
            Py_DECREF(it);
            if (bad || PyErr_Occurred()) { Py_DECREF(src); goto node_done; }
          }
          Py_DECREF(src);
        }
      }
      /* gpus / gpu_breakdown: from the device, or from the bignum table */
      {
        PyObject* wrec = NULL;
        if (have_wide) {
          PyObject* key = PyLong_FromLongLong(gi[i]);
          wrec = key ? PyDict_GetItemWithError(wide, key) : NULL; /* borrowed */
          Py_XDECREF(key);
          if (!wrec && PyErr_Occurred()) goto node_done;
        }
        if (wrec) {
          PyObject* gp = PyTuple_GetItem(wrec, 0);
          PyObject* src = PyTuple_GetItem(wrec, 1);
          if (!gp || !src) goto node_done;
          gpus = gp;
          Py_INCREF(gpus);
          bd = PyDict_Copy(src);
        } else {
          gpus = PyLong_FromLongLong(gt[i]);
          bd = PyDict_New();
          for (int k = 0; k < 4 && bd; k++)
            if (f & (8u << k)) {
              PyObject* v = PyLong_FromLongLong(gv[4 * i + k]);
              if (!v || PyDict_SetItem(bd, s_key_names[k], v)) { Py_XDECREF(v); Py_CLEAR(bd); break; }
              Py_DECREF(v);
            }
        }
        if (!gpus || !bd) goto node_done;
      }
      rec = PyDict_New();
      if (!rec || PyDict_SetItem(rec, s_name, name) || PyDict_SetItem(rec, s_ready, (f & 2) ? Py_True : Py_False) ||
          PyDict_SetItem(rec, s_gpus, gpus) || PyDict_SetItem(rec, s_gpu_breakdown, bd) || PyDict_SetItem(rec, s_labels, labels) ||
          PyDict_SetItem(rec, s_taints, taints))
        goto node_done;
      if ((f & 2) && PyList_Append(ready_list, rec)) goto node_done;
      PyList_SET_ITEM(gpu_list, i, rec); /* steals rec */
      rec = NULL;
      ok = 1;
    node_done:
      Py_XDECREF(name); Py_XDECREF(labels); Py_XDECREF(taints); Py_XDECREF(gpus); Py_XDECREF(bd); Py_XDECREF(rec);
      if (!ok) goto done;
    }
    result = PyTuple_Pack(2, gpu_list, ready_list);
  }
done:
  Py_XDECREF(gpu_list);
  Py_XDECREF(ready_list);
  Py_XDECREF(seq);
  PyBuffer_Release(&idx);
  PyBuffer_Release(&flags);
  PyBuffer_Release(&totals);
  PyBuffer_Release(&vals);
  return result;
}

static PyMethodDef methods[] = {
    {"build_records", build_records, METH_VARARGS, "(gpu_nodes, ready_gpu_nodes) as lists of dicts"},
    {"emit_summary_json", emit_summary_json, METH_VARARGS, "the --json document as text"},{"pack_nodes", pack_nodes, METH_VARARGS, "V1Node objects -> (rows bytes, row_ptr bytes, n)"},
                                {NULL, NULL, 0, NULL}};
static struct PyModuleDef moddef = {PyModuleDef_HEAD_INIT, "_fastpack", "PNL-v2 packer (C)", -1, methods};

PyMODINIT_FUNC PyInit__fastpack(void) {
  s_status = PyUnicode_InternFromString("status");
  s_capacity = PyUnicode_InternFromString("capacity");
  s_conditions = PyUnicode_InternFromString("conditions");
  s_type = PyUnicode_InternFromString("type");
  s_items = PyUnicode_InternFromString("items");
  s_metadata = PyUnicode_InternFromString("metadata");
  s_name = PyUnicode_InternFromString("name");
  s_labels = PyUnicode_InternFromString("labels");
  s_spec = PyUnicode_InternFromString("spec");
  s_taints = PyUnicode_InternFromString("taints");
  s_key = PyUnicode_InternFromString("key");
  s_value = PyUnicode_InternFromString("value");
  s_effect = PyUnicode_InternFromString("effect");
  s_ready = PyUnicode_InternFromString("ready");
  s_gpus = PyUnicode_InternFromString("gpus");
  s_gpu_breakdown = PyUnicode_InternFromString("gpu_breakdown");
  for (int k = 0; k < 4; k++) s_key_names[k] = PyUnicode_InternFromString(KEY_NAMES[k]);
  return PyModule_Create(&moddef);
}
