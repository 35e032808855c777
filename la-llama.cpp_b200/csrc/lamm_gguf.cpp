// lamm_gguf.cpp -- GGUF weights straight into (sharded) device memory.
//
// The on-disk block layout of a GGUF tensor IS the device layout of this library (DESIGN "Data layout
// in HBM"): a 2-D weight tensor is registered row range by row range from the file mapping, with no
// intermediate host copy and no repacking.  This file only parses the container; the format follows
// gguf_init_from_file (llama.cpp-b2430/ggml.c:20655-20850) and the writer gguf_write_to_file /
// gguf_write_to_buf (ggml.c:21280-21420):
//
//   "GGUF" | u32 version (2, 3) | u64 n_tensors | u64 n_kv
//   n_kv x   { str key | u32 type | value }          str = u64 length + bytes (no terminator)
//                                                   array value = u32 element type | u64 n | elements
//   n_tensors x { str name | u32 n_dims | u64 ne[n_dims] | u32 ggml type | u64 offset }
//   padding to `general.alignment` (u32 kv, default 32, ggml.h:250)
//   data section: tensor i at data_start + offset[i]
//
// Host C++ only.  Every read is bounds-checked against the mapping: a truncated or corrupt file is an
// error, never a fault.
#include "../../include/lammcu.h"

#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace lamm { void set_error(const char *fmt, ...); }      // lammcu_api.cu: what lammcu_last_error() returns

namespace {

enum : uint32_t {          // enum gguf_type, ggml.h:2214-2229
  GT_U8 = 0, GT_I8, GT_U16, GT_I16, GT_U32, GT_I32, GT_F32, GT_BOOL, GT_STR, GT_ARR, GT_U64, GT_I64, GT_F64, GT_COUNT
};

size_t scalar_size(uint32_t t) {
  switch (t) {
  case GT_U8: case GT_I8: case GT_BOOL: return 1;
  case GT_U16: case GT_I16: return 2;
  case GT_U32: case GT_I32: case GT_F32: return 4;
  case GT_U64: case GT_I64: case GT_F64: return 8;
  default: return 0;
  }
}

struct Kv {
  std::string key;
  uint32_t type = 0, arr_type = 0;
  uint64_t arr_n = 0;
  const uint8_t *val = nullptr;       // scalar: the value; string: its bytes; array: first element
  uint64_t len = 0;                   // string length
};

struct Tensor {
  std::string name;
  int type = 0, n_dims = 0;
  int64_t ne[4] = {1, 1, 1, 1};
  uint64_t offset = 0, nbytes = 0;
};

// bytes of one row of ne0 elements; 0 for a type whose block size is unknown here
uint64_t row_bytes(int type, int64_t ne0) {
  if (lammcu_type_size(type)) return (uint64_t)lammcu_row_size(type, ne0);
  switch (type) {                     // plain types outside the product's table (ggml.c:458-775)
  case 1: return (uint64_t)ne0 * 2;   // F16
  case 24: return (uint64_t)ne0;      // I8
  case 25: return (uint64_t)ne0 * 2;  // I16
  case 26: return (uint64_t)ne0 * 4;  // I32
  default: return 0;
  }
}

}  // namespace

struct lammcu_gguf {
  int fd = -1;
  const uint8_t *map = nullptr;
  uint64_t size = 0;
  uint32_t version = 0, alignment = 32;
  uint64_t data_start = 0;
  std::vector<Kv> kv;
  std::vector<Tensor> tensors;
  std::string error;
};

namespace {

struct Cursor {
  const uint8_t *p;
  uint64_t size, at = 0;
  bool ok = true;
  bool need(uint64_t n) {
    if (!ok || n > size - at) ok = false;
    return ok;
  }
  template <typename T> T get() {
    T v{};
    if (need(sizeof(T))) { std::memcpy(&v, p + at, sizeof(T));  at += sizeof(T); }
    return v;
  }
  const uint8_t *bytes(uint64_t n) {
    if (!need(n)) return nullptr;
    const uint8_t *r = p + at;
    at += n;
    return r;
  }
  std::string str() {
    const uint64_t n = get<uint64_t>();
    const uint8_t *b = bytes(n);
    return b ? std::string((const char *)b, (size_t)n) : std::string();
  }
};

bool skip_value(Cursor &c, uint32_t type, Kv *kv, int depth) {
  if (type == GT_STR) {
    const uint64_t n = c.get<uint64_t>();
    const uint8_t *b = c.bytes(n);
    if (kv) { kv->val = b;  kv->len = n; }
    return c.ok;
  }
  if (type == GT_ARR) {
    if (depth > 0) return false;                       // nested arrays: ggml.c rejects them too
    const uint32_t et = c.get<uint32_t>();
    const uint64_t n = c.get<uint64_t>();
    if (kv) { kv->arr_type = et;  kv->arr_n = n;  kv->val = c.ok ? c.p + c.at : nullptr; }
    if (et == GT_STR) {
      for (uint64_t i = 0; i < n && c.ok; i++) skip_value(c, GT_STR, nullptr, 1);
      return c.ok;
    }
    const size_t es = scalar_size(et);
    if (!es || (n && es > UINT64_MAX / n)) return false;
    c.bytes(n * es);
    return c.ok;
  }
  const size_t s = scalar_size(type);
  if (!s) return false;
  const uint8_t *b = c.bytes(s);
  if (kv) kv->val = b;
  return c.ok;
}

int fail(lammcu_gguf *g, const char *msg) {
  lamm::set_error("gguf: %s", msg);
  if (g) {
    if (g->map) munmap((void *)g->map, (size_t)g->size);
    if (g->fd >= 0) close(g->fd);
    delete g;
  }
  return LAMMCU_ERR_ARG;
}

const Kv *find_kv(const lammcu_gguf *g, const char *key, uint32_t type) {
  for (const Kv &k : g->kv)
    if (k.key == key) return k.type == type ? &k : nullptr;
  return nullptr;
}

}  // namespace

extern "C" {

int lammcu_gguf_open(const char *path, lammcu_gguf_t *out) {
  if (!path || !out) return LAMMCU_ERR_ARG;
  *out = nullptr;
  lammcu_gguf *g = new lammcu_gguf();
  g->fd = open(path, O_RDONLY);
  if (g->fd < 0) return fail(g, "cannot open the file");
  struct stat st;
  if (fstat(g->fd, &st) != 0 || st.st_size < 24) return fail(g, "file too short for a GGUF header");
  g->size = (uint64_t)st.st_size;
  void *m = mmap(nullptr, (size_t)g->size, PROT_READ, MAP_PRIVATE, g->fd, 0);
  if (m == MAP_FAILED) { g->map = nullptr;  return fail(g, "mmap failed"); }
  g->map = (const uint8_t *)m;
  Cursor c{g->map, g->size};
  const uint8_t *magic = c.bytes(4);
  if (!magic || std::memcmp(magic, "GGUF", 4) != 0) return fail(g, "bad magic");
  g->version = c.get<uint32_t>();
  if (g->version < 2 || g->version > 3) return fail(g, "unsupported GGUF version (2 and 3 are read)");
  const uint64_t n_tensors = c.get<uint64_t>(), n_kv = c.get<uint64_t>();
  // each kv / tensor record takes at least 12 / 24 bytes: bounds the counts before any allocation
  if (!c.ok || n_kv > g->size / 12 || n_tensors > g->size / 24) return fail(g, "implausible header counts");
  g->kv.resize((size_t)n_kv);
  for (Kv &k : g->kv) {
    k.key = c.str();
    k.type = c.get<uint32_t>();
    if (!c.ok || k.type >= GT_COUNT || !skip_value(c, k.type, &k, 0)) return fail(g, "corrupt key/value section");
  }
  if (const Kv *a = find_kv(g, "general.alignment", GT_U32)) {
    std::memcpy(&g->alignment, a->val, 4);
    if (g->alignment == 0 || (g->alignment & (g->alignment - 1))) return fail(g, "general.alignment is not a power of two");
  }
  g->tensors.resize((size_t)n_tensors);
  for (Tensor &t : g->tensors) {
    t.name = c.str();
    const uint32_t nd = c.get<uint32_t>();
    if (!c.ok || nd > 4) return fail(g, "corrupt tensor info");
    t.n_dims = (int)nd;
    for (uint32_t j = 0; j < nd; j++) {
      const uint64_t v = c.get<uint64_t>();
      if (v > (uint64_t)INT64_MAX) return fail(g, "tensor dimension out of range");
      t.ne[j] = (int64_t)v;
    }
    t.type = (int)c.get<uint32_t>();
    t.offset = c.get<uint64_t>();
    if (!c.ok) return fail(g, "corrupt tensor info");
  }
  g->data_start = (c.at + g->alignment - 1) / g->alignment * g->alignment;
  if (g->data_start > g->size && n_tensors) return fail(g, "file ends before the data section");
  for (Tensor &t : g->tensors) {
    const int bs = lammcu_blck_size(t.type);
    if (bs > 1 && t.ne[0] % bs) return fail(g, "tensor row is not a whole number of blocks");
    // every factor is bounded by the file size BEFORE it is multiplied in: no product can wrap (hostile dims are up to
    // 2^63 each, and lammcu_row_size itself wraps in size_t for ne0 near 2^62)
    if ((uint64_t)t.ne[0] > g->size * 256) return fail(g, "tensor larger than the file");   // >= 1 byte per 256 elements
    const uint64_t rb = row_bytes(t.type, t.ne[0]);
    uint64_t nb = rb;
    for (int j = 1; j < 4 && nb; j++) {
      if (nb > g->size || (t.ne[j] && (uint64_t)t.ne[j] > g->size / nb)) return fail(g, "tensor larger than the file");
      nb *= (uint64_t)t.ne[j];
    }
    if (nb > g->size) return fail(g, "tensor larger than the file");
    t.nbytes = nb;                                        // 0: a ggml type this library has no size for
    if (t.offset % g->alignment) return fail(g, "tensor offset is not aligned");
    if (t.offset > g->size - g->data_start || t.nbytes > g->size - g->data_start - t.offset)
      return fail(g, "tensor data outside the file");
  }
  *out = g;
  return LAMMCU_OK;
}

int lammcu_gguf_close(lammcu_gguf_t g) {
  if (!g) return LAMMCU_ERR_ARG;
  if (g->map) munmap((void *)g->map, (size_t)g->size);
  if (g->fd >= 0) close(g->fd);
  delete g;
  return LAMMCU_OK;
}

int64_t lammcu_gguf_tensor_count(lammcu_gguf_t g) { return g ? (int64_t)g->tensors.size() : -1; }
uint32_t lammcu_gguf_alignment(lammcu_gguf_t g) { return g ? g->alignment : 0; }
uint64_t lammcu_gguf_data_offset(lammcu_gguf_t g) { return g ? g->data_start : 0; }

int lammcu_gguf_tensor(lammcu_gguf_t g, int64_t i, const char **name, int *type, int *n_dims, int64_t *ne4,
                       const void **data, uint64_t *nbytes) {
  if (!g || i < 0 || i >= (int64_t)g->tensors.size()) return LAMMCU_ERR_ARG;
  const Tensor &t = g->tensors[(size_t)i];
  if (name) *name = t.name.c_str();
  if (type) *type = t.type;
  if (n_dims) *n_dims = t.n_dims;
  if (ne4) for (int j = 0; j < 4; j++) ne4[j] = t.ne[j];
  if (data) *data = g->map + g->data_start + t.offset;
  if (nbytes) *nbytes = t.nbytes;
  return LAMMCU_OK;
}

int64_t lammcu_gguf_find(lammcu_gguf_t g, const char *name) {
  if (!g || !name) return -1;
  for (size_t i = 0; i < g->tensors.size(); i++)
    if (g->tensors[i].name == name) return (int64_t)i;
  return -1;
}

int lammcu_gguf_get_u32(lammcu_gguf_t g, const char *key, uint32_t *out) {
  if (!g || !key || !out) return LAMMCU_ERR_ARG;
  const Kv *k = find_kv(g, key, GT_U32);
  if (!k) return LAMMCU_ERR_ARG;
  std::memcpy(out, k->val, 4);
  return LAMMCU_OK;
}

int lammcu_gguf_get_f32(lammcu_gguf_t g, const char *key, float *out) {
  if (!g || !key || !out) return LAMMCU_ERR_ARG;
  const Kv *k = find_kv(g, key, GT_F32);
  if (!k) return LAMMCU_ERR_ARG;
  std::memcpy(out, k->val, 4);
  return LAMMCU_OK;
}

int64_t lammcu_gguf_get_str(lammcu_gguf_t g, const char *key, char *buf, int64_t cap) {
  if (!g || !key) return -1;
  const Kv *k = find_kv(g, key, GT_STR);
  if (!k) return -1;
  if (buf && cap > 0) {
    const uint64_t n = k->len < (uint64_t)(cap - 1) ? k->len : (uint64_t)(cap - 1);
    std::memcpy(buf, k->val, (size_t)n);
    buf[n] = 0;
  }
  return (int64_t)k->len;
}

int lammcu_gguf_rows(lammcu_gguf_t g, int64_t i, int64_t row_begin, int64_t row_end, const void **data,
                     uint64_t *nbytes) {
  if (!g || i < 0 || i >= (int64_t)g->tensors.size()) return LAMMCU_ERR_ARG;
  const Tensor &t = g->tensors[(size_t)i];
  const uint64_t rb = row_bytes(t.type, t.ne[0]);
  if (t.n_dims > 2 || t.ne[2] != 1 || t.ne[3] != 1 || !rb) return LAMMCU_ERR_ARG;
  if (row_begin < 0 || row_end > t.ne[1] || row_begin >= row_end) return LAMMCU_ERR_ARG;
  if (data) *data = g->map + g->data_start + t.offset + (uint64_t)row_begin * rb;
  if (nbytes) *nbytes = (uint64_t)(row_end - row_begin) * rb;
  return LAMMCU_OK;
}

int lammcu_gguf_upload_rows(lammcu_gguf_t g, int64_t i, int64_t row_begin, int64_t row_end, lammcu_weight_t *out) {
  if (!g || !out || i < 0 || i >= (int64_t)g->tensors.size()) return LAMMCU_ERR_ARG;
  const Tensor &t = g->tensors[(size_t)i];
  if (!lammcu_type_supported(t.type) || lammcu_gguf_rows(g, i, row_begin, row_end, nullptr, nullptr) != LAMMCU_OK) {
    lamm::set_error("gguf: tensor %lld is not a 2-D weight of a supported type, or rows [%lld, %lld) are outside it",
                    (long long)i, (long long)row_begin, (long long)row_end);
    return LAMMCU_ERR_ARG;
  }
  // the file mapping is the host copy: its rows go to the device as they are (weights are keyed by this address)
  return lammcu_weight_register(g->map + g->data_start + t.offset, t.type, t.ne[1], t.ne[0],
                                (int64_t)lammcu_row_size(t.type, t.ne[0]), row_begin, row_end, out);
}

}  // extern "C"
