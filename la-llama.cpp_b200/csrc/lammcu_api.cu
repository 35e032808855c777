// lammcu_api.cu -- the C-ABI of include/lammcu.h: device/stream ownership,
// the weight registry (upload once per src0->data), workspaces, and dispatch
// between the decode GEMV and the prefill GEMM.  No torch, no C++ types in the
// exported signatures.  There is no CPU fallback: without a device every
// compute entry point returns LAMMCU_ERR_NO_DEVICE.
#include "lamm_common.cuh"
#include "lamm_sibling.h"
#include "lamm_xop.cuh"
#include <algorithm>
#include <chrono>

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

namespace lamm {

unsigned long long g_launch_count = 0;

static thread_local char t_err[512] = "";
void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
}

#define CU_TRY(expr)                                                                 \
  do {                                                                               \
    cudaError_t e_ = (expr);                                                         \
    if (e_ != cudaSuccess) {                                                         \
      set_error("%s failed: %s", #expr, cudaGetErrorString(e_));                     \
      return LAMMCU_ERR_CUDA;                                                        \
    }                                                                                \
  } while (0)

struct DeviceState {
  bool inited = false;
  int device = -1;
  int sms = 0;
  cudaStream_t stream = nullptr;
  unsigned int *sched = nullptr;   // ring of self-resetting scheduler counter pairs
  unsigned int sched_next = 0;
  void *ws = nullptr;  size_t ws_bytes = 0;      // quantized-activation workspace
  void *xin = nullptr; size_t xin_bytes = 0;     // host path: X on device
  void *cout = nullptr; size_t cout_bytes = 0;   // host path: C on device
  void *hx = nullptr;  size_t hx_bytes = 0;      // pinned staging
  void *hc = nullptr;  size_t hc_bytes = 0;
  void *gws = nullptr; size_t gws_bytes = 0;     // prefill GEMM operand scratch
};
static DeviceState g_dev;
static std::mutex g_mu;
static int g_gemv_max_n = 8;
static int g_decode_kernel = 1;     // N == 1: the single-launch kernel of lamm_decode.cu
constexpr unsigned int kSchedSlots = 4096;

int sm_count_cached() { return g_dev.sms > 0 ? g_dev.sms : 148; }

static unsigned long long *g_trace = nullptr;
static long long g_trace_cap = 0, g_trace_used = 0;
unsigned long long *next_trace_slot(int grid) {
  if (!g_trace || g_trace_used + (long long)grid * 8 > g_trace_cap) return nullptr;
  unsigned long long *p = g_trace + g_trace_used;
  g_trace_used += (long long)grid * 8;
  return p;
}

unsigned int *next_sched_slot() {
  if (!g_dev.sched) {
    set_error("lammcu not initialised (call lammcu_init)");
    return nullptr;
  }
  unsigned int i = g_dev.sched_next++ % kSchedSlots;
  return g_dev.sched + 2 * i;
}

// The current CUDA device is per host thread: ggml calls the operator from its worker pool, an application may
// call from any thread.  Every entry point binds the calling thread to the library's device once.
static thread_local bool t_bound = false;
static int ensure_init() {
  if (!g_dev.inited) {
    int rc = lammcu_init(-1);
    if (rc) return rc;
  }
  if (!t_bound) {
    cudaError_t e = cudaSetDevice(g_dev.device);
    if (e != cudaSuccess) { set_error("cudaSetDevice(%d): %s", g_dev.device, cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
    t_bound = true;
  }
  return LAMMCU_OK;
}

static int grow(void **p, size_t *have, size_t need, bool pinned) {
  if (*have >= need) return LAMMCU_OK;
  if (*p) {
    if (pinned) cudaFreeHost(*p); else cudaFree(*p);
    *p = nullptr; *have = 0;
  }
  size_t n = need + need / 4 + 4096;
  cudaError_t e = pinned ? cudaMallocHost(p, n) : cudaMalloc(p, n);
  if (e != cudaSuccess) {
    set_error("allocation of %zu bytes failed: %s", n, cudaGetErrorString(e));
    return LAMMCU_ERR_NOMEM;
  }
  *have = n;
  return LAMMCU_OK;
}

void *gemm_workspace(size_t bytes) {
  if (grow(&g_dev.gws, &g_dev.gws_bytes, bytes, false) != LAMMCU_OK) return nullptr;
  return g_dev.gws;
}

}  // namespace lamm

using namespace lamm;

// resident decode server of the host path (see "the resident-server host path" below)
static unsigned int g_srv_life_us = [] {
  const char *e = getenv("LAMM_SERVER_LIFE_US");
  const int v = e ? atoi(e) : 5000;
  return (unsigned int)(v < 10 ? 10 : v > 4000000 ? 4000000 : v);
}();

// 1 (default): a call whose activation row equals the previous call's does not send it again
static const int g_srv_x_reuse = [] { const char *e = getenv("LAMM_SERVER_X_REUSE"); return e ? atoi(e) : 1; }();

constexpr size_t kSrvMaxK = 1 << 16;         // floats in the pinned activation row
constexpr size_t kSrvMaxM = 1 << 18;         // floats in each of the kSrvSlots result buffers
// 1 (default): after a job, also post the jobs of the weights that followed it with the same row last time
static const int g_srv_spec = [] { const char *e = getenv("LAMM_SERVER_SPEC"); return e ? atoi(e) : 1; }();

struct ServerState {
  bool ready = false;        // buffers allocated
  bool running = false;      // a server launch is (believed to be) resident
  int type = -1;             // weight format of the resident instance
  unsigned int seq = 0;      // tag of the last job posted
  int64_t x_k = 0;           // K of the row in x_host that the RUNNING instance answered last (0: none)
  volatile unsigned long long *mbox = nullptr;    // pinned: the job
  float *x_host = nullptr;                         // pinned: activation row
  volatile unsigned long long *c_ll = nullptr;     // pinned: result pairs
  volatile unsigned int *done = nullptr;           // pinned: [1] gave-up tag
  unsigned long long *mbox_dev = nullptr, *x_ll = nullptr;
  unsigned int *ctrs = nullptr;                    // [2] status
  unsigned long long *trace = nullptr;             // LAMM_SERVER_TRACE=1: [64][8] device time stamps
  // sibling speculation: what was learned / posted (lamm_sibling.h; see mul_mat_host_server)
  lamm::SibTracker<lammcu_weight> sib;
  int64_t row_k = 0;                               // K of the row that x_host holds (whatever instance saw it)
  unsigned int launch_seq0 = 0;                    // first tag of the instance believed to be resident
};

static ServerState g_srv;
// The resident instance has exited (idle / lifetime) iff it left the tag it gave up waiting for: tags only grow and a
// new instance always starts above the previous one's exit tag.
static bool server_exited() { return g_srv.done[1] != 0u && (int)(g_srv.done[1] - g_srv.launch_seq0) >= 0; }
// ... and it never served job s
static bool server_missed(unsigned int s) { return server_exited() && (int)(s - g_srv.done[1]) >= 0; }

extern "C" { static void server_stop(); }

struct lammcu_weight {
  const void *host;     // registry key
  int type;
  int64_t M, K, row_begin, row_end;
  int64_t stride;       // bytes between device rows (packed)
  void *dev;
  size_t bytes;
  int refs;             // explicit holders (lammcu_weight_register); entries made by the operator path have 0
  bool detached;        // evicted from the registry while held: freed by the last lammcu_weight_release
  // resident-server host path: the weight whose call FOLLOWED this one's with a bit-identical activation row last time
  // (ggml hands wq/wk/wv one src1, gate/up another); valid while sib_gen == g_weight_gen
  lammcu_weight *sib = nullptr;
  unsigned long long sib_gen = 0;
};
static unsigned long long g_weight_gen = 1;   // bumped whenever a weight may have gone away: every sibling link dies
static std::map<const void *, lammcu_weight *> g_weights;

// drop an entry from the registry (g_mu held): freed now, or by its last holder's release
static void evict_entry(lammcu_weight *h) {
  g_weight_gen++;
  if (h->refs > 0) { h->detached = true; return; }
  cudaFree(h->dev);
  delete h;
}

extern "C" {

const char *lammcu_version(void) { return "lammcu 0.1 (sm_100a)"; }
const char *lammcu_last_error(void) { return t_err; }

int lammcu_device_count(void) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e));
    return LAMMCU_ERR_NO_DEVICE;
  }
  return n;
}

int lammcu_init(int device) {
  std::lock_guard<std::mutex> lk(g_mu);
  int n = lammcu_device_count();
  if (n <= 0) {
    if (n == 0) set_error("no CUDA device");
    return LAMMCU_ERR_NO_DEVICE;
  }
  if (device < 0) {
    if (g_dev.inited) return LAMMCU_OK;
    CU_TRY(cudaGetDevice(&device));
  }
  if (device >= n) {
    set_error("device %d out of range (%d devices)", device, n);
    return LAMMCU_ERR_ARG;
  }
  if (g_dev.inited && g_dev.device == device) return LAMMCU_OK;
  if (g_dev.inited) {
    set_error("lammcu already bound to device %d; call lammcu_shutdown first", g_dev.device);
    return LAMMCU_ERR_ARG;
  }
  CU_TRY(cudaSetDevice(device));
  cudaDeviceProp prop;
  CU_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    return LAMMCU_ERR_NO_DEVICE;
  }
  g_dev.device = device;
  g_dev.sms = prop.multiProcessorCount;
  CU_TRY(cudaStreamCreateWithFlags(&g_dev.stream, cudaStreamNonBlocking));
  CU_TRY(cudaMalloc(&g_dev.sched, kSchedSlots * 2 * sizeof(unsigned int)));
  CU_TRY(cudaMemset(g_dev.sched, 0, kSchedSlots * 2 * sizeof(unsigned int)));
  CU_TRY(cudaDeviceSynchronize());
  g_dev.inited = true;
  return LAMMCU_OK;
}

int lammcu_shutdown(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_dev.inited) return LAMMCU_OK;
  cudaSetDevice(g_dev.device);
  server_stop();
  cudaDeviceSynchronize();
  if (g_srv.ready) {
    cudaFreeHost((void *)g_srv.mbox);  cudaFreeHost(g_srv.x_host);  cudaFreeHost((void *)g_srv.c_ll);  cudaFreeHost((void *)g_srv.done);
    if (g_srv.trace) cudaFree(g_srv.trace);
    cudaFree(g_srv.mbox_dev);  cudaFree(g_srv.x_ll);  cudaFree(g_srv.ctrs);
    g_srv = ServerState();
  }
  for (auto &kv : g_weights) {
    cudaFree(kv.second->dev);
    delete kv.second;
  }
  g_weights.clear();
  if (g_dev.ws) cudaFree(g_dev.ws);
  if (g_dev.xin) cudaFree(g_dev.xin);
  if (g_dev.cout) cudaFree(g_dev.cout);
  if (g_dev.hx) cudaFreeHost(g_dev.hx);
  if (g_dev.hc) cudaFreeHost(g_dev.hc);
  if (g_dev.gws) cudaFree(g_dev.gws);
  if (g_dev.sched) cudaFree(g_dev.sched);
  if (g_dev.stream) cudaStreamDestroy(g_dev.stream);
  g_dev = DeviceState();
  return LAMMCU_OK;
}

int lammcu_sm_count(void) { return g_dev.inited ? g_dev.sms : 0; }
uint64_t lammcu_launch_count(void) { return g_launch_count; }
void lammcu_launch_count_reset(void) { g_launch_count = 0; }

// ---- type table -------------------------------------------------------------
int lammcu_type_supported(int t) {
  return t == LAMMCU_TYPE_F32 || t == LAMMCU_TYPE_F16 || t == LAMMCU_TYPE_Q2_K || t == LAMMCU_TYPE_Q4_0 ||
         t == LAMMCU_TYPE_Q4_1 || t == LAMMCU_TYPE_Q5_0 || t == LAMMCU_TYPE_Q5_1 ||
         t == LAMMCU_TYPE_Q8_0 || is_kquant_ext(t);
}
int lammcu_blck_size(int t) { return type_info(t).blck; }
size_t lammcu_type_size(int t) { return (size_t)type_info(t).size; }
int lammcu_vec_dot_type(int t) { return type_info(t).vec_dot; }
size_t lammcu_row_size(int t, int64_t k) {
  const TypeInfo ti = type_info(t);
  return ti.blck ? (size_t)(k / ti.blck) * ti.size : 0;
}

int lammcu_set_gemv_max_n(int n) {
  if (n < 0) return LAMMCU_ERR_ARG;
  g_gemv_max_n = n;
  return LAMMCU_OK;
}
int lammcu_get_gemv_max_n(void) { return g_gemv_max_n; }
int lammcu_set_decode_kernel(int enable) {
  g_decode_kernel = enable ? 1 : 0;
  return LAMMCU_OK;
}
int lammcu_get_decode_kernel(void) { return g_decode_kernel; }
int lammcu_set_trace(void *dev_buf, int64_t n_u64) {
  g_trace = (unsigned long long *)dev_buf;
  g_trace_cap = dev_buf ? n_u64 : 0;
  g_trace_used = 0;
  return LAMMCU_OK;
}

// ---- device operands ----------------------------------------------------------
static cudaStream_t pick_stream(void *s) { return s ? (cudaStream_t)s : g_dev.stream; }

int lammcu_quantize_act_dev(int vt, const float *x, int64_t ldx, void *xq, int64_t n_rows,
                            int64_t k, void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!x || !xq) { set_error("quantize_act: NULL operand"); return LAMMCU_ERR_ARG; }
  return launch_quantize_act(vt, x, ldx, xq, n_rows, k, pick_stream(stream));
}

int lammcu_mul_mat_q_dev(int w_type, const void *w, int64_t w_stride, const void *xq,
                         int64_t xq_stride, float *c, int64_t ldc, int64_t M, int64_t N,
                         int64_t K, void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!w || !xq || !c) { set_error("mul_mat_q: NULL operand"); return LAMMCU_ERR_ARG; }
  if (!lammcu_type_supported(w_type)) { set_error("mul_mat_q: type %d not supported", w_type); return LAMMCU_ERR_TYPE; }
  cudaStream_t st = pick_stream(stream);
  if (N == 1 && g_decode_kernel && decode_supported(w_type, w, w_stride, xq, M, K))
    return launch_decode(w_type, w, w_stride, nullptr, xq, nullptr, c, M, K, st);
  if (N <= g_gemv_max_n) return launch_gemv(w_type, w, w_stride, xq, xq_stride, c, ldc, M, N, K, st);
  return launch_gemm(w_type, w, w_stride, xq, xq_stride, c, ldc, M, N, K, st);
}

int lammcu_mul_mat_dev(int w_type, const void *w, int64_t w_stride, const float *x, int64_t ldx,
                       float *c, int64_t ldc, int64_t M, int64_t N, int64_t K, void *xq_ws,
                       void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!w || !x || !c) { set_error("mul_mat: NULL operand"); return LAMMCU_ERR_ARG; }
  if (!lammcu_type_supported(w_type)) { set_error("mul_mat: type %d not supported", w_type); return LAMMCU_ERR_TYPE; }
  if (w_type == LAMMCU_TYPE_F32)   // F32 x F32: X is used as it is (src/loongarch_matmul.cpp:89-94)
    return lammcu_mul_mat_q_dev(w_type, w, w_stride, x, ldx * 4, c, ldc, M, N, K, stream);
  const int vt = type_info(w_type).vec_dot;
  if (K <= 0 || K % type_info(w_type).blck || N <= 0 || (w_type == LAMMCU_TYPE_F16 && K % 8)) {
    set_error("mul_mat: bad shape N=%lld K=%lld", (long long)N, (long long)K);
    return LAMMCU_ERR_SHAPE;
  }
  const size_t rs = lammcu_row_size(vt, K);
  // decode: one launch, the quantized row only leaves the SM when the caller asks for it
  if (N == 1 && g_decode_kernel && decode_supported(w_type, w, w_stride, x, M, K) &&
      (!xq_ws || ((uintptr_t)xq_ws & 3) == 0))
    return launch_decode(w_type, w, w_stride, x, nullptr, xq_ws, c, M, K, pick_stream(stream));
  // prefill on the scaled-operand GEMM: quantize + operand prep in one launch, Xq stays in registers
  // unless the caller passed a workspace for it
  if (N > g_gemv_max_n && !get_gemm_exact() && gemm_tc_supported(w_type, w, w_stride, M, N, K) &&
      (((uintptr_t)x | (uintptr_t)(ldx * 4)) & 15) == 0 && (!xq_ws || ((uintptr_t)xq_ws & 3) == 0)) {
    void *ws = gemm_workspace(gemm_so_ws_bytes(M, N, K));
    if (!ws) return LAMMCU_ERR_NOMEM;
    return launch_gemm_so(w_type, w, w_stride, xq_ws, (int64_t)rs, c, ldc, M, N, K, ws, pick_stream(stream), x, ldx);
  }
  if (!xq_ws) {
    rc = grow(&g_dev.ws, &g_dev.ws_bytes, rs * (size_t)N, false);
    if (rc) return rc;
    xq_ws = g_dev.ws;
  }
  rc = launch_quantize_act(vt, x, ldx, xq_ws, N, K, pick_stream(stream));
  if (rc) return rc;
  return lammcu_mul_mat_q_dev(w_type, w, w_stride, xq_ws, (int64_t)rs, c, ldc, M, N, K, stream);
}

// ---- the no-allocation entry point (no counterpart in the reference) --
// Same computation as lammcu_mul_mat_dev for 1 <= N <= 8 on the two-launch path (activation-quantize kernel ->
// Xq scratch -> gemv_stream_kernel), but the Xq scratch is a file-scope device array instead of a cudaMalloc'ed
// workspace: callers under CUDA graph capture, tight serving loops or harnesses that forbid allocation can use it.
// Every GPU that loads the library holds the array until exit.  In host code `g_noalloc_xq` is only a host-side
// stand-in for the symbol, not a device address: the device address is looked up once per device with
// cudaGetSymbolAddress (that device current) and only that pointer is handed to kernels.
constexpr size_t kNoAllocScratchBytes = LAMMCU_NOALLOC_MAX_SCRATCH;
static __device__ __align__(256) unsigned char g_noalloc_xq[kNoAllocScratchBytes];
static void *g_noalloc_ptr = nullptr;      // device address of g_noalloc_xq on device g_noalloc_dev
static int g_noalloc_dev = -1;

extern "C" int lammcu_mul_mat_noalloc_dev(int w_type, const void *w, int64_t w_stride, const float *x, int64_t ldx,
                                          float *c, int64_t ldc, int64_t M, int64_t N, int64_t K, void *stream) {
  // lammcu_init allocates (stream, scheduler counters): it must have happened before, not in here
  if (!g_dev.inited) { set_error("mul_mat_noalloc: call lammcu_init first (this entry point allocates nothing)"); return LAMMCU_ERR_ARG; }
  if (!w || !x || !c) { set_error("mul_mat_noalloc: NULL operand"); return LAMMCU_ERR_ARG; }
  if (w_type == LAMMCU_TYPE_F32 || !lammcu_type_supported(w_type)) {
    set_error("mul_mat_noalloc: type %d not supported (F32 weights need no scratch: use lammcu_mul_mat_dev)", w_type);
    return LAMMCU_ERR_TYPE;
  }
  const TypeInfo ti = type_info(w_type);
  if (M <= 0 || K <= 0 || K % ti.blck || N < 1 || N > 8) {
    set_error("mul_mat_noalloc: bad shape M=%lld N=%lld K=%lld (1 <= N <= 8, K %% %d == 0)", (long long)M, (long long)N,
              (long long)K, ti.blck);
    return LAMMCU_ERR_SHAPE;
  }
  const size_t rs = lammcu_row_size(ti.vec_dot, K);
  if (rs * (size_t)N > kNoAllocScratchBytes) {
    set_error("mul_mat_noalloc: %lld rows of %zu quantized bytes need %zu bytes of scratch, the static array holds %zu",
              (long long)N, rs, rs * (size_t)N, kNoAllocScratchBytes);
    return LAMMCU_ERR_SHAPE;
  }
  {
    std::lock_guard<std::mutex> lk(g_mu);
    if (g_noalloc_dev != g_dev.device) {
      cudaError_t e = cudaSetDevice(g_dev.device);
      void *ptr = nullptr;
      if (e == cudaSuccess) e = cudaGetSymbolAddress(&ptr, g_noalloc_xq);
      if (e != cudaSuccess || !ptr) {
        set_error("mul_mat_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        return LAMMCU_ERR_CUDA;
      }
      g_noalloc_ptr = ptr;
      g_noalloc_dev = g_dev.device;
    }
  }
  cudaStream_t st = pick_stream(stream);
  // nothing in the array has to start at zero: the quantize kernel writes every byte of the N rows it produces
  // (Q8_K: bsums of all-zero super-blocks included), and the GEMV reads exactly those rows
  int rc = launch_quantize_act(ti.vec_dot, x, ldx, g_noalloc_ptr, N, K, st);
  if (rc) return rc;
  return launch_gemv(w_type, w, w_stride, g_noalloc_ptr, (int64_t)rs, c, ldc, M, N, K, st);
}

// ---- row-sharded multi-GPU decode ---------------------------------------------------
struct lammcu_comm {
  int rank, world;
  unsigned int *ctrs;                          // local: [0] CTA arrival tickets, [2] status
};

int lammcu_ipc_alloc(size_t bytes, void **dev_ptr, void *handle64) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!dev_ptr || !handle64 || !bytes) { set_error("ipc_alloc: bad argument"); return LAMMCU_ERR_ARG; }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  void *p = nullptr;
  CU_TRY(cudaMalloc(&p, bytes));
  CU_TRY(cudaMemset(p, 0, bytes));
  CU_TRY(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    set_error("cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
    return LAMMCU_ERR_CUDA;
  }
  memcpy(handle64, &h, 64);
  *dev_ptr = p;
  return LAMMCU_OK;
}

int lammcu_ipc_open(const void *handle64, void **dev_ptr) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!dev_ptr || !handle64) { set_error("ipc_open: bad argument"); return LAMMCU_ERR_ARG; }
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  CU_TRY(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return LAMMCU_OK;
}

int lammcu_ipc_close(void *dev_ptr) {
  CU_TRY(cudaIpcCloseMemHandle(dev_ptr));
  return LAMMCU_OK;
}

int lammcu_ipc_free(void *dev_ptr) {
  CU_TRY(cudaFree(dev_ptr));
  return LAMMCU_OK;
}

int lammcu_memcpy_async(void *dst, const void *src, size_t bytes, void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  CU_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, pick_stream(stream)));
  return LAMMCU_OK;
}

int lammcu_comm_create(int rank, int world, lammcu_comm_t *out) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!out || world < 1 || world > LAMMCU_MAX_RANKS || rank < 0 || rank >= world) {
    set_error("comm_create: bad argument (rank %d of %d)", rank, world);
    return LAMMCU_ERR_ARG;
  }
  lammcu_comm *c = new lammcu_comm();
  c->rank = rank;
  c->world = world;
  cudaError_t e = cudaMalloc(&c->ctrs, 256);
  if (e == cudaSuccess) e = cudaMemset(c->ctrs, 0, 256);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    set_error("comm_create: %s", cudaGetErrorString(e));
    delete c;
    return LAMMCU_ERR_CUDA;
  }
  *out = c;
  return LAMMCU_OK;
}

int lammcu_comm_destroy(lammcu_comm_t c) {
  if (!c) return LAMMCU_ERR_ARG;
  cudaFree(c->ctrs);
  delete c;
  return LAMMCU_OK;
}

int lammcu_comm_status(lammcu_comm_t c, int *timed_out) {
  if (!c || !timed_out) return LAMMCU_ERR_ARG;
  unsigned int v = 0;
  CU_TRY(cudaMemcpy(&v, c->ctrs + 2, sizeof(v), cudaMemcpyDeviceToHost));
  *timed_out = (int)v;
  return LAMMCU_OK;
}

int lammcu_comm_sync(lammcu_comm_t c, void *stream) {
  if (!c) return LAMMCU_ERR_ARG;
  CU_TRY(cudaStreamSynchronize(pick_stream(stream)));
  int bad = 0;
  int rc = lammcu_comm_status(c, &bad);
  if (rc) return rc;
  if (bad) {
    set_error("comm_sync: rank %d of %d gave up polling a peer (a rank died or never launched its shard)", c->rank, c->world);
    return LAMMCU_ERR_TIMEOUT;
  }
  return LAMMCU_OK;
}

size_t lammcu_ll_bytes(int64_t M) { return (size_t)M * 8; }

int lammcu_mul_mat_sharded_dev(lammcu_comm_t comm, int w_type, const void *w_shard, int64_t w_stride,
                               const void *x, int x_back, void *const *c_ll_all, int64_t row_begin,
                               int64_t rows, int64_t K, void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!comm || !w_shard || !x || !c_ll_all || rows <= 0 || row_begin < 0 || x_back < 0) {
    set_error("mul_mat_sharded: bad argument");
    return LAMMCU_ERR_ARG;
  }
  if (!decode_supported(w_type, w_shard, w_stride, x, rows, K)) {
    set_error("mul_mat_sharded: needs a quantized type, K %% 256 == 0 and 16-byte aligned operands");
    return LAMMCU_ERR_SHAPE;
  }
  ShardInfo sh = {};
  for (int g = 0; g < comm->world; g++) {
    if (!c_ll_all[g] || ((uintptr_t)c_ll_all[g] & 15)) { set_error("mul_mat_sharded: output %d NULL or misaligned", g); return LAMMCU_ERR_ALIGN; }
    sh.ll_peer[g] = (unsigned long long *)c_ll_all[g];
  }
  if (x_back > 0) {
    sh.x_ll = (const unsigned long long *)x;
    sh.x_back = x_back;
  }
  sh.arrive_ctr = comm->ctrs;
  sh.status = comm->ctrs + 2;
  sh.n_peer = comm->world;
  sh.my_rank = comm->rank;
  sh.row_begin = (int)row_begin;
  return launch_decode(w_type, w_shard, w_stride, (const float *)x, nullptr, nullptr, nullptr, rows, K,
                       pick_stream(stream), &sh);
}

// sibling matrices (q/k/v, gate/up) that share the activation row: ONE sharded launch, one sequence number
int lammcu_mul_mat_sharded_batch_dev(lammcu_comm_t comm, int w_type, int count, const void *const *w_shards,
                                     int64_t w_stride, const void *x, int x_back, void *const *c_ll_all,
                                     const int64_t *row_begin, const int64_t *rows, int64_t K, void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!comm || !w_shards || !x || !c_ll_all || !row_begin || !rows || count < 1 || count > 4 || x_back < 0) {
    set_error("mul_mat_sharded_batch: bad argument");
    return LAMMCU_ERR_ARG;
  }
  if (count == 1)
    return lammcu_mul_mat_sharded_dev(comm, w_type, w_shards[0], w_stride, x, x_back, c_ll_all, row_begin[0], rows[0], K, stream);
  ShardInfo sh = {};
  BatchInfo bt = {};
  bt.count = count;
  int64_t total = 0;
  for (int j = 0; j < count; j++) {
    if (!w_shards[j] || rows[j] <= 0 || row_begin[j] < 0) { set_error("mul_mat_sharded_batch: bad operand %d", j); return LAMMCU_ERR_ARG; }
    if (!decode_supported(w_type, w_shards[j], w_stride, x, rows[j], K) ||
        (((uintptr_t)total * (uintptr_t)w_stride) & 15) != 0 || (((uintptr_t)w_shards[j] ^ (uintptr_t)w_shards[0]) & 15) != 0) {
      set_error("mul_mat_sharded_batch: needs a quantized type, K %% 256 == 0, 16-byte aligned operands and row counts "
                "that keep every matrix at the same 16-byte phase");
      return LAMMCU_ERR_SHAPE;
    }
    for (int g = 0; g < comm->world; g++) {
      void *v = c_ll_all[j * comm->world + g];
      if (!v || ((uintptr_t)v & 15)) { set_error("mul_mat_sharded_batch: output %d/%d NULL or misaligned", j, g); return LAMMCU_ERR_ALIGN; }
      sh.ll_batch[j][g] = (unsigned long long *)v;
    }
    sh.row_begin_b[j] = row_begin[j];
    bt.w[j] = w_shards[j];
    bt.M[j] = rows[j];
    total += rows[j];
  }
  if (total > 0x3fffffff) { set_error("mul_mat_sharded_batch: too many rows"); return LAMMCU_ERR_SHAPE; }
  for (int g = 0; g < comm->world; g++) sh.ll_peer[g] = sh.ll_batch[0][g];
  if (x_back > 0) {
    sh.x_ll = (const unsigned long long *)x;
    sh.x_back = x_back;
  }
  sh.arrive_ctr = comm->ctrs;
  sh.status = comm->ctrs + 2;
  sh.n_peer = comm->world;
  sh.my_rank = comm->rank;
  sh.row_begin = (int)row_begin[0];
  return launch_decode(w_type, w_shards[0], w_stride, (const float *)x, nullptr, nullptr, nullptr, total, K,
                       pick_stream(stream), &sh, &bt);
}

// ---- row-sharded multi-GPU prefill ----------------------------------------------------------------
// Every rank runs the GEMM on its row shard; the epilogue stores the C tile straight into the full
// [N][ldc] result of every rank through peer-mapped pointers (SURVEY 8e "preferred").  Completion:
// a publish launch after the GEMM raises flag[my_rank] = sequence number on every rank (the kernel
// boundary has made the GEMM's peer stores visible; one system fence in front of the flag), and a
// wait launch polls this rank's flags until all ranks have published.  ~10 us on a >= 100 us GEMM.
struct FlagPtrs { unsigned int *p[LAMMCU_MAX_RANKS]; };
__global__ void gemm_publish_kernel(unsigned int *seq_ctr, FlagPtrs flags, int rank, int world) {
  __shared__ unsigned int seq;
  if (threadIdx.x == 0) seq = atomicAdd(seq_ctr, 1u) + 1u;
  __syncthreads();
  __threadfence_system();
  if ((int)threadIdx.x < world)
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(flags.p[threadIdx.x] + rank), "r"(seq) : "memory");
}
__global__ void gemm_wait_kernel(const unsigned int *seq_ctr, const unsigned int *flags, int world, unsigned int *status) {
  const unsigned int seq = *reinterpret_cast<const volatile unsigned int *>(seq_ctr);   // stream order: after the publish
  if ((int)threadIdx.x < world) {
    unsigned int v, spins = 0;
    unsigned long long t0 = 0;
    while (true) {
      asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + threadIdx.x) : "memory");
      if ((int)(v - seq) >= 0) break;
      __nanosleep(200);
      if ((++spins & 1023u) == 0) {
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (t0 == 0) t0 = now;
        else if (now - t0 > 5000000000ull) { *status = 1u; break; }
      }
    }
  }
  __threadfence_system();
}

int lammcu_mul_mat_sharded_gemm_dev(lammcu_comm_t comm, int w_type, const void *w_shard, int64_t w_stride,
                                    const float *x, int64_t ldx, void *const *c_all, int64_t ldc,
                                    void *const *flags_all, int64_t row_begin, int64_t rows, int64_t N, int64_t K,
                                    void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!comm || !w_shard || !x || !c_all || !flags_all || rows <= 0 || row_begin < 0 || N <= 0 || row_begin + rows > ldc) {
    set_error("mul_mat_sharded_gemm: bad argument");
    return LAMMCU_ERR_ARG;
  }
  GemmPeers peers = {};
  FlagPtrs fl = {};
  peers.n = comm->world;
  for (int g = 0; g < comm->world; g++) {
    if (!c_all[g] || !flags_all[g]) { set_error("mul_mat_sharded_gemm: buffer %d is NULL", g); return LAMMCU_ERR_ARG; }
    peers.c[g] = (float *)c_all[g] + row_begin;
    fl.p[g] = (unsigned int *)flags_all[g];
  }
  cudaStream_t st = pick_stream(stream);
  float *c_local = peers.c[comm->rank];
  gemm_set_peers(&peers);
  rc = lammcu_mul_mat_dev(w_type, w_shard, w_stride, x, ldx, c_local, ldc, rows, N, K, nullptr, stream);
  const bool fused = gemm_peers_used();
  gemm_set_peers(nullptr);
  if (rc) return rc;
  if (!fused) {
    // the dispatch took a kernel without the peer-store epilogue (GEMV sizes, exact mode): copy the slab out
    for (int g = 0; g < comm->world; g++)
      if (g != comm->rank)
        CU_TRY(cudaMemcpy2DAsync(peers.c[g], (size_t)ldc * 4, c_local, (size_t)ldc * 4, (size_t)rows * 4, (size_t)N,
                                 cudaMemcpyDeviceToDevice, st));
  }
  gemm_publish_kernel<<<1, 32, 0, st>>>(comm->ctrs + 4, fl, comm->rank, comm->world);
  gemm_wait_kernel<<<1, 32, 0, st>>>(comm->ctrs + 4, fl.p[comm->rank], comm->world, comm->ctrs + 2);
  g_launch_count += 2;
  CU_TRY(cudaGetLastError());
  return LAMMCU_OK;
}

// LL vector -> plain f32 (waits for every element's tag): the last consumer of a sharded chain
__global__ void ll_unpack_kernel(const unsigned int *arrive_ctr, int grid_fixed, int back, const unsigned long long *ll,
                                 float *out, int64_t M, unsigned int *status) {
  // stream order: every earlier sharded launch has completed, so tickets / grid = launches so far
  const unsigned int tag = *reinterpret_cast<const volatile unsigned int *>(arrive_ctr) / (unsigned int)grid_fixed -
                           (unsigned int)back + 1u;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += (int64_t)gridDim.x * blockDim.x) {
    unsigned long long v;
    unsigned int spins = 0;
    unsigned long long t0 = 0;
    while (true) {
      asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(ll + i) : "memory");
      if ((unsigned int)(v >> 32) == tag) break;
      if ((++spins & 1023u) == 0) {
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (t0 == 0) t0 = now;
        else if (now - t0 > 2000000000ull) { *status = 1u; break; }
      }
    }
    out[i] = __uint_as_float((unsigned int)v);
  }
}

int lammcu_ll_unpack_dev(lammcu_comm_t comm, const void *c_ll, int back, float *out, int64_t M, void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!comm || !c_ll || !out || M <= 0 || back < 1) { set_error("ll_unpack: bad argument"); return LAMMCU_ERR_ARG; }
  const int blocks = (int)std::min<int64_t>((M + 255) / 256, 148);
  ll_unpack_kernel<<<blocks, 256, 0, pick_stream(stream)>>>(comm->ctrs, sm_count_cached(), back,
                                                           (const unsigned long long *)c_ll, out, M, comm->ctrs + 2);
  g_launch_count++;
  CU_TRY(cudaGetLastError());
  return LAMMCU_OK;
}

int lammcu_set_gemm_exact(int on) { set_gemm_exact(on); return LAMMCU_OK; }
int lammcu_get_gemm_exact(void) { return get_gemm_exact(); }

int lammcu_mul_mat_batch_dev(int w_type, int count, const void *const *w_dev, int64_t w_stride,
                             const int64_t *M, const float *x_dev, float *const *c_dev, int64_t K,
                             void *xq_ws_dev, void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  if (count < 1 || !w_dev || !M || !x_dev || !c_dev) { set_error("mul_mat_batch: bad argument"); return LAMMCU_ERR_ARG; }
  for (int j = 0; j < count; j++)
    if (!w_dev[j] || !c_dev[j] || M[j] <= 0) { set_error("mul_mat_batch: bad operand %d", j); return LAMMCU_ERR_ARG; }
  // one launch when the decode kernel can take the whole set; else one call per matrix
  bool fuse = count >= 2 && count <= 4 && g_decode_kernel;
  int64_t total = 0;
  for (int j = 0; fuse && j < count; j++) {
    fuse = decode_supported(w_type, w_dev[j], w_stride, x_dev, M[j], K) &&
           (((uintptr_t)total * (uintptr_t)w_stride) & 15) == 0 &&          // same (address & 15) for every row
           (((uintptr_t)w_dev[j] & 15) == ((uintptr_t)w_dev[0] & 15));
    total += M[j];
  }
  if (fuse && total <= 0x3fffffff) {
    BatchInfo bt = {};
    bt.count = count;
    for (int j = 0; j < count; j++) { bt.w[j] = w_dev[j]; bt.c[j] = c_dev[j]; bt.M[j] = M[j]; }
    return launch_decode(w_type, w_dev[0], w_stride, x_dev, nullptr, nullptr, c_dev[0], total, K, pick_stream(stream),
                         nullptr, &bt);
  }
  for (int j = 0; j < count; j++) {
    rc = lammcu_mul_mat_dev(w_type, w_dev[j], w_stride, x_dev, K, c_dev[j], M[j], M[j], 1, K, xq_ws_dev, stream);
    if (rc) return rc;
  }
  return LAMMCU_OK;
}

// ---- the op in front of the matmul, fused (lamm_xop.cuh) --------------------------------------
static int xop_from(const lammcu_xop *op, XOp *out) {
  if (!op || (op->kind != LAMMCU_XOP_RMS_NORM_MUL && op->kind != LAMMCU_XOP_SILU_MUL)) {
    set_error("xop: NULL or unknown kind (1 = rms_norm x weight, 2 = silu(gate) x up)");
    return LAMMCU_ERR_ARG;
  }
  out->kind = op->kind;  out->aux = op->aux;  out->ld_aux = op->ld_aux;  out->eps = op->eps;  out->silu = nullptr;
  return LAMMCU_OK;
}

int lammcu_quantize_act_xop_dev(const lammcu_xop *op, int vt, const float *x, int64_t ldx, void *xq, int64_t N,
                                int64_t K, void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  XOp xo;
  if ((rc = xop_from(op, &xo))) return rc;
  if (!x || !xq) { set_error("quantize_act_xop: NULL operand"); return LAMMCU_ERR_ARG; }
  return launch_xop_quantize(xo, vt, x, ldx, xq, N, K, pick_stream(stream));
}

int lammcu_mul_mat_xop_dev(const lammcu_xop *op, int w_type, const void *w, int64_t w_stride, const float *x,
                           int64_t ldx, float *c, int64_t ldc, int64_t M, int64_t N, int64_t K, void *xq_ws,
                           void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  XOp xo;
  if ((rc = xop_from(op, &xo))) return rc;
  if (!w || !x || !c) { set_error("mul_mat_xop: NULL operand"); return LAMMCU_ERR_ARG; }
  if (!lammcu_type_supported(w_type)) { set_error("mul_mat_xop: type %d not supported", w_type); return LAMMCU_ERR_TYPE; }
  const int vt = type_info(w_type).vec_dot;
  const int kmul = (w_type == LAMMCU_TYPE_F32 || w_type == LAMMCU_TYPE_F16) ? 8 : type_info(w_type).blck;
  if (K <= 0 || K % kmul || N <= 0 || M <= 0) { set_error("mul_mat_xop: bad shape M=%lld N=%lld K=%lld", (long long)M, (long long)N, (long long)K); return LAMMCU_ERR_SHAPE; }
  cudaStream_t st = pick_stream(stream);
  // decode: ONE launch -- the op is applied while every CTA stages (and quantizes) the row
  if (N == 1 && g_decode_kernel && decode_supported(w_type, w, w_stride, x, M, K) && (((uintptr_t)xo.aux) & 15) == 0 &&
      (!xq_ws || ((uintptr_t)xq_ws & 3) == 0))
    return launch_decode(w_type, w, w_stride, x, nullptr, xq_ws, c, M, K, st, nullptr, nullptr, &xo);
  // otherwise: op + quantize in one launch, then the matmul on the quantized rows
  const size_t rs = (w_type == LAMMCU_TYPE_F32) ? (size_t)K * 4 : lammcu_row_size(vt, K);
  if (!xq_ws) {
    rc = grow(&g_dev.ws, &g_dev.ws_bytes, rs * (size_t)N, false);
    if (rc) return rc;
    xq_ws = g_dev.ws;
  }
  rc = launch_xop_quantize(xo, vt, x, ldx, xq_ws, N, K, st);
  if (rc) return rc;
  return lammcu_mul_mat_q_dev(w_type, w, w_stride, xq_ws, (int64_t)rs, c, ldc, M, N, K, stream);
}

int lammcu_mul_mat_batch_xop_dev(const lammcu_xop *op, int w_type, int count, const void *const *w_dev, int64_t w_stride,
                                 const int64_t *M, const float *x_dev, float *const *c_dev, int64_t K, void *xq_ws_dev,
                                 void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  XOp xo;
  if ((rc = xop_from(op, &xo))) return rc;
  if (count < 1 || !w_dev || !M || !x_dev || !c_dev) { set_error("mul_mat_batch_xop: bad argument"); return LAMMCU_ERR_ARG; }
  for (int j = 0; j < count; j++)
    if (!w_dev[j] || !c_dev[j] || M[j] <= 0) { set_error("mul_mat_batch_xop: bad operand %d", j); return LAMMCU_ERR_ARG; }
  if (!lammcu_type_supported(w_type)) { set_error("mul_mat_batch_xop: type %d not supported", w_type); return LAMMCU_ERR_TYPE; }
  bool fuse = count >= 2 && count <= 4 && g_decode_kernel && (((uintptr_t)xo.aux) & 15) == 0;
  int64_t total = 0;
  for (int j = 0; fuse && j < count; j++) {
    fuse = decode_supported(w_type, w_dev[j], w_stride, x_dev, M[j], K) &&
           (((uintptr_t)total * (uintptr_t)w_stride) & 15) == 0 && (((uintptr_t)w_dev[j] & 15) == ((uintptr_t)w_dev[0] & 15));
    total += M[j];
  }
  if (fuse && total <= 0x3fffffff) {
    BatchInfo bt = {};
    bt.count = count;
    for (int j = 0; j < count; j++) { bt.w[j] = w_dev[j]; bt.c[j] = c_dev[j]; bt.M[j] = M[j]; }
    return launch_decode(w_type, w_dev[0], w_stride, x_dev, nullptr, nullptr, c_dev[0], total, K, pick_stream(stream),
                         nullptr, &bt, &xo);
  }
  // not fusable: the op + quantize once, then one matmul per matrix on the shared quantized row
  const int vt = type_info(w_type).vec_dot;
  const int kmul = (w_type == LAMMCU_TYPE_F32 || w_type == LAMMCU_TYPE_F16) ? 8 : type_info(w_type).blck;
  if (K <= 0 || K % kmul) { set_error("mul_mat_batch_xop: bad K=%lld", (long long)K); return LAMMCU_ERR_SHAPE; }
  const size_t rs = (w_type == LAMMCU_TYPE_F32) ? (size_t)K * 4 : lammcu_row_size(vt, K);
  if (!xq_ws_dev) {
    rc = grow(&g_dev.ws, &g_dev.ws_bytes, rs, false);
    if (rc) return rc;
    xq_ws_dev = g_dev.ws;
  }
  rc = launch_xop_quantize(xo, vt, x_dev, K, xq_ws_dev, 1, K, pick_stream(stream));
  if (rc) return rc;
  for (int j = 0; j < count; j++) {
    rc = lammcu_mul_mat_q_dev(w_type, w_dev[j], w_stride, xq_ws_dev, (int64_t)rs, c_dev[j], M[j], M[j], 1, K, stream);
    if (rc) return rc;
  }
  return LAMMCU_OK;
}

// ---- persistent decode chain (lamm_decode.cu) ---------------------------------------------------
struct lammcu_chain { ChainPlan *plan; };

int lammcu_chain_create(int w_type, int n_ops, const lammcu_chain_op *ops, lammcu_chain_t *out) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!ops || !out || n_ops < 1) { set_error("chain_create: bad argument"); return LAMMCU_ERR_ARG; }
  if (w_type == LAMMCU_TYPE_F32 || !lammcu_type_supported(w_type)) { set_error("chain_create: type %d not supported", w_type); return LAMMCU_ERR_TYPE; }
  std::vector<ChainOp> v((size_t)n_ops);
  for (int i = 0; i < n_ops; i++) {
    ChainOp &o = v[(size_t)i];
    o.count = ops[i].count;
    for (int j = 0; j < 4; j++) { o.w[j] = ops[i].w_dev[j];  o.M[j] = ops[i].M[j];  o.c[j] = ops[i].c_dev[j]; }
    o.x = ops[i].x_dev;  o.K = ops[i].K;  o.w_stride = ops[i].w_stride_bytes;
  }
  ChainPlan *pl = nullptr;
  if ((rc = chain_create(w_type, n_ops, v.data(), &pl))) return rc;
  lammcu_chain *h = new lammcu_chain();
  h->plan = pl;
  *out = h;
  return LAMMCU_OK;
}

int lammcu_chain_run_dev(lammcu_chain_t ch, void *stream) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!ch || !ch->plan) { set_error("chain_run: bad handle"); return LAMMCU_ERR_ARG; }
  if (!stream || (cudaStream_t)stream == g_dev.stream) {      // the persistent kernel needs every SM: no resident server
    std::lock_guard<std::mutex> lk(g_mu);
    server_stop();
  }
  return chain_run(ch->plan, pick_stream(stream));
}

int lammcu_chain_status(lammcu_chain_t ch, int *timed_out) {
  if (!ch || !ch->plan || !timed_out) { set_error("chain_status: bad argument"); return LAMMCU_ERR_ARG; }
  return chain_status(ch->plan, timed_out);
}

int lammcu_chain_destroy(lammcu_chain_t ch) {
  if (!ch) return LAMMCU_ERR_ARG;
  chain_destroy(ch->plan);
  delete ch;
  return LAMMCU_OK;
}

// ---- host operands --------------------------------------------------------------
// hold: 1 = an explicit registration (the caller gets a reference and releases it), 0 = the operator path (the
// registry owns the entry).  refresh: the host bytes may have changed since the last call (src0 is not a leaf
// weight): re-upload them into the existing device copy instead of trusting it.
static int weight_get(const void *w_host, int w_type, int64_t M, int64_t K, int64_t w_stride, int64_t row_begin,
                      int64_t row_end, int hold, int refresh, lammcu_weight_t *out) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!w_host || !out) { set_error("weight_register: NULL argument"); return LAMMCU_ERR_ARG; }
  if (!lammcu_type_supported(w_type)) { set_error("weight_register: type %d not supported", w_type); return LAMMCU_ERR_TYPE; }
  const TypeInfo ti = type_info(w_type);
  if (M <= 0 || K <= 0 || K % ti.blck || row_begin < 0 || row_end > M || row_begin >= row_end) {
    set_error("weight_register: bad shape M=%lld K=%lld rows [%lld,%lld)", (long long)M, (long long)K,
              (long long)row_begin, (long long)row_end);
    return LAMMCU_ERR_SHAPE;
  }
  const size_t rs = lammcu_row_size(w_type, K);
  if ((size_t)w_stride < rs) { set_error("weight_register: row stride smaller than a row"); return LAMMCU_ERR_SHAPE; }
  const int64_t rows = row_end - row_begin;
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_weights.find(w_host);
  if (it != g_weights.end()) {
    lammcu_weight *h = it->second;
    if (h->type == w_type && h->M == M && h->K == K && h->row_begin == row_begin && h->row_end == row_end) {
      if (refresh) {
        server_stop();                                 // a resident instance may still be reading the old bytes
        cudaError_t e = cudaStreamSynchronize(g_dev.stream);
        if (e == cudaSuccess)
          e = cudaMemcpy2D(h->dev, rs, (const char *)w_host + row_begin * w_stride, (size_t)w_stride, rs, (size_t)rows,
                           cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { set_error("weight refresh: %s", cudaGetErrorString(e)); return LAMMCU_ERR_CUDA; }
      }
      if (hold) h->refs++;
      *out = h;
      return LAMMCU_OK;
    }
    // same address, different tensor: the old one is stale (a holder of its handle keeps it alive until release)
    server_stop();
    cudaStreamSynchronize(g_dev.stream);
    g_weights.erase(it);
    evict_entry(h);
  }
  lammcu_weight *h = new lammcu_weight();
  h->host = w_host; h->type = w_type; h->M = M; h->K = K;
  h->row_begin = row_begin; h->row_end = row_end;
  h->stride = (int64_t)rs; h->refs = hold ? 1 : 0; h->detached = false;
  h->bytes = rs * (size_t)rows;
  // +256: the GEMV's bulk copies round every row span out to 16-byte boundaries
  cudaError_t e = cudaMalloc(&h->dev, h->bytes + 256);
  if (e != cudaSuccess) {
    set_error("weight_register: cudaMalloc(%zu): %s", h->bytes + 256, cudaGetErrorString(e));
    delete h;
    return LAMMCU_ERR_NOMEM;
  }
  e = cudaMemcpy2D(h->dev, rs, (const char *)w_host + row_begin * w_stride, (size_t)w_stride, rs,
                   (size_t)rows, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemset((char *)h->dev + h->bytes, 0, 256);
  // A pageable-memory cudaMemcpy returns once the LAST chunk is staged, not when its DMA has landed, and
  // the memset is asynchronous on the legacy stream; the compute stream is non-blocking, so order by hand.
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    set_error("weight_register: upload: %s", cudaGetErrorString(e));
    cudaFree(h->dev);
    delete h;
    return LAMMCU_ERR_CUDA;
  }
  g_weights[w_host] = h;
  *out = h;
  return LAMMCU_OK;
}

int lammcu_weight_register(const void *w_host, int w_type, int64_t M, int64_t K,
                           int64_t w_stride, int64_t row_begin, int64_t row_end,
                           lammcu_weight_t *out) {
  return weight_get(w_host, w_type, M, K, w_stride, row_begin, row_end, 1, 0, out);
}

int lammcu_weight_for_op(const void *w_host, int w_type, int64_t M, int64_t K, int64_t w_stride, int refresh,
                         lammcu_weight_t *out) {
  return weight_get(w_host, w_type, M, K, w_stride, 0, M, 0, refresh, out);
}

int lammcu_weight_lookup(const void *w_host, lammcu_weight_t *out) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_weights.find(w_host);
  if (it == g_weights.end()) { set_error("weight_lookup: not registered"); return LAMMCU_ERR_ARG; }
  *out = it->second;
  return LAMMCU_OK;
}

int lammcu_weight_release(lammcu_weight_t w) {
  if (!w) return LAMMCU_ERR_ARG;
  std::lock_guard<std::mutex> lk(g_mu);
  if (w->detached) {                                   // evicted while held (same address, new tensor; forget)
    if (w->refs <= 0) { set_error("weight_release: bad handle"); return LAMMCU_ERR_ARG; }
    if (--w->refs > 0) return LAMMCU_OK;
    cudaFree(w->dev);
    delete w;
    return LAMMCU_OK;
  }
  auto it = g_weights.find(w->host);
  if (it == g_weights.end() || it->second != w || w->refs <= 0) { set_error("weight_release: bad handle"); return LAMMCU_ERR_ARG; }
  if (--w->refs > 0) return LAMMCU_OK;
  server_stop();
  cudaStreamSynchronize(g_dev.stream);
  cudaFree(w->dev);
  g_weights.erase(it);
  delete w;
  return LAMMCU_OK;
}

// Drop registry entries whatever their reference count: the host tensor is about to be freed (model
// unload), so a later tensor at the same address must not hit the stale device copy.
int lammcu_weight_forget(const void *w_host) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_weights.find(w_host);
  if (it == g_weights.end()) return LAMMCU_OK;
  server_stop();
  if (g_dev.inited) cudaStreamSynchronize(g_dev.stream);
  cudaFree(it->second->dev);
  delete it->second;
  g_weights.erase(it);
  return LAMMCU_OK;
}
int lammcu_weight_forget_all(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_weights.empty()) return LAMMCU_OK;
  server_stop();
  if (g_dev.inited) cudaStreamSynchronize(g_dev.stream);
  for (auto &kv : g_weights) {
    cudaFree(kv.second->dev);
    delete kv.second;
  }
  g_weights.clear();
  return LAMMCU_OK;
}

const void *lammcu_weight_dev_ptr(lammcu_weight_t w) { return w ? w->dev : nullptr; }
int64_t lammcu_weight_rows(lammcu_weight_t w) { return w ? w->row_end - w->row_begin : 0; }
int64_t lammcu_weight_bytes(lammcu_weight_t w) { return w ? (int64_t)w->bytes : 0; }

// D2H of the shard's column range of every output row
// Small results (decode / GEMV sizes) are written by the kernel straight into pinned host memory
// (device-accessible under UVA): one DMA and its launch latency less per call.  Large (prefill)
// results keep the coalesced D2H copy.
// host-path phase timers (diagnostics: lammcu_host_profile)
static double g_hp[8] = {0, 0, 0, 0, 0, 0, 0, 0};
static inline double now_us() {
  return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static bool zero_copy_out(int64_t N) { return N <= (int64_t)lammcu_get_gemv_max_n(); }

static int out_buffer(lammcu_weight_t w, int64_t N, float **c_dev) {
  const size_t bytes = (size_t)N * (size_t)(w->row_end - w->row_begin) * 4;
  int rc = grow(&g_dev.hc, &g_dev.hc_bytes, bytes, true);
  if (rc) return rc;
  if (zero_copy_out(N)) { *c_dev = (float *)g_dev.hc; return LAMMCU_OK; }
  if ((rc = grow(&g_dev.cout, &g_dev.cout_bytes, bytes, false))) return rc;
  *c_dev = (float *)g_dev.cout;
  return LAMMCU_OK;
}

static int copy_c_back(lammcu_weight_t w, int64_t N, float *c_host, int64_t ldc) {
  const int64_t rows = w->row_end - w->row_begin;
  if (!zero_copy_out(N))
    CU_TRY(cudaMemcpyAsync(g_dev.hc, g_dev.cout, (size_t)N * rows * 4, cudaMemcpyDeviceToHost, g_dev.stream));
  const double t0 = now_us();
  CU_TRY(cudaStreamSynchronize(g_dev.stream));
  const double t1 = now_us();
  for (int64_t j = 0; j < N; j++)
    memcpy(c_host + j * ldc + w->row_begin, (const float *)g_dev.hc + j * rows, (size_t)rows * 4);
  g_hp[3] += t1 - t0;  g_hp[4] += now_us() - t1;  g_hp[5] += 1;
  return LAMMCU_OK;
}

// ---- the resident-server host path (decode-sized calls) ---------------------------------------
// A synchronous call that launches a kernel pays ~10 us of launch + completion latency on top of
// ~5 us of work (scratch/hostlat.cu measures the floor on this box).  In mode 1 an N == 1 call is a
// JOB for the resident decode server (lamm_decode.cu: decode_server_kernel): the calling thread
// copies X into a pinned row, posts 16 {word, tag} pairs, polls one pinned word and copies C out.
// No CUDA API call on the steady-state path.  The server gives up after an idle period and is
// relaunched on demand; anything else that needs the stream stops it first (server_stop).
//   mode 0: copy-engine H2D, one kernel launch per call, cudaStreamSynchronize
//   mode 1: resident server (default)
static int g_host_mode = [] {
  const char *e = getenv("LAMM_HOST_MODE");
  const int m = e ? atoi(e) : 1;
  return (m < 0 || m > 1) ? 1 : m;
}();
static unsigned int g_srv_idle_us = [] {
  const char *e = getenv("LAMM_SERVER_IDLE_US");
  const int v = e ? atoi(e) : 200;
  return (unsigned int)(v < 1 ? 1 : v > 1000000 ? 1000000 : v);
}();

static int server_setup() {
  if (g_srv.ready) return LAMMCU_OK;
  void *p = nullptr;
  constexpr size_t kMboxBytes = (size_t)kSrvSlots * kSrvWords * 8;
  static_assert(kSrvSlots == 4 && kSrvWords == 16, "decode_server_kernel reads the mailbox as 2 x 32 pairs");
  CU_TRY(cudaMallocHost(&p, kMboxBytes));            g_srv.mbox = (volatile unsigned long long *)p;
  CU_TRY(cudaMallocHost(&p, kSrvMaxK * 4));          g_srv.x_host = (float *)p;
  CU_TRY(cudaMallocHost(&p, kSrvSlots * kSrvMaxM * 8));  g_srv.c_ll = (volatile unsigned long long *)p;
  CU_TRY(cudaMallocHost(&p, 256));                   g_srv.done = (volatile unsigned int *)p;
  CU_TRY(cudaMalloc(&p, kMboxBytes));                g_srv.mbox_dev = (unsigned long long *)p;
  CU_TRY(cudaMalloc(&p, kSrvMaxK * 8));              g_srv.x_ll = (unsigned long long *)p;
  CU_TRY(cudaMalloc(&p, 256));                       g_srv.ctrs = (unsigned int *)p;
  if (getenv("LAMM_SERVER_TRACE")) {
    CU_TRY(cudaMalloc(&p, 64 * 8 * 8));
    g_srv.trace = (unsigned long long *)p;
    CU_TRY(cudaMemset(p, 0, 64 * 8 * 8));
  }
  memset((void *)g_srv.mbox, 0, kMboxBytes);         // tag 0 is never issued
  memset((void *)g_srv.done, 0, 256);
  memset((void *)g_srv.c_ll, 0, kSrvSlots * kSrvMaxM * 8);
  CU_TRY(cudaMemset(g_srv.mbox_dev, 0, kMboxBytes));
  CU_TRY(cudaMemset(g_srv.x_ll, 0, kSrvMaxK * 8));
  CU_TRY(cudaMemset(g_srv.ctrs, 0, 256));
  CU_TRY(cudaDeviceSynchronize());
  // job tags are 32-bit and only ever compared for equality / signed distance; LAMM_SERVER_SEQ0 starts them elsewhere
  // (test knob: tests/test_gpu_parity.py crosses the restart below in a few hundred calls)
  const char *seq0 = getenv("LAMM_SERVER_SEQ0");
  g_srv.seq = seq0 ? (unsigned int)strtoul(seq0, nullptr, 10) : 0u;
  g_srv.ready = true;
  return LAMMCU_OK;
}

// Post `n` jobs under the next n tags (slot = tag % kSrvSlots).  The LAST one is written first and the first one last:
// the poll that finds the job the server is waiting for then finds its followers too and accepts them in the same
// PCIe round trip.  Tags validate each slot, so a poll that lands between two writes only misses the look-ahead.
static void server_post_n(const unsigned int (*words)[kSrvWords], int n) {
  const unsigned int s0 = g_srv.seq + 1;
  g_srv.seq += (unsigned int)n;
  __atomic_thread_fence(__ATOMIC_RELEASE);           // X (and everything before) is visible before the job
  for (int j = n - 1; j >= 0; j--) {
    const unsigned int s = s0 + (unsigned int)j;
    volatile unsigned long long *slot = g_srv.mbox + (size_t)(s % kSrvSlots) * kSrvWords;
    for (int i = 0; i < kSrvWords; i++) slot[i] = ((unsigned long long)s << 32) | words[j][i];
    __atomic_thread_fence(__ATOMIC_RELEASE);
  }
}
static void server_post(const unsigned int *words) {
  unsigned int one[1][kSrvWords];
  memcpy(one[0], words, sizeof(one[0]));
  server_post_n(one, 1);
}

static int server_launch(int type, unsigned int seq0) {
  ServerParams sp = {};
  sp.mbox_host = (const unsigned long long *)g_srv.mbox;
  sp.mbox_dev = g_srv.mbox_dev;
  sp.x_host = g_srv.x_host;
  sp.x_ll = g_srv.x_ll;
  sp.c_ll_host = (unsigned long long *)g_srv.c_ll;
  sp.c_slot_pairs = (unsigned int)kSrvMaxM;
  sp.done_host = g_srv.done;
  sp.status = g_srv.ctrs + 2;
  sp.seq0 = seq0;                                    // the job the caller is waiting for
  sp.idle_ns = g_srv_idle_us * 1000u;
  sp.life_ns = g_srv_life_us * 1000u;
  sp.trace = g_srv.trace;
  int rc = launch_decode_server(type, sp, g_dev.stream);
  if (rc) return rc;
  g_srv.running = true;
  g_srv.launch_seq0 = seq0;
  g_srv.type = type;
  g_srv.x_k = 0;                                     // a fresh instance has nothing staged
  g_srv.sib.drop_posted();                           // ... and never looks at older tags
  return LAMMCU_OK;
}

// values of `n` result pairs whose tags all equal `tag` -> out; returns how many leading pairs were ready
static size_t ll_take_scalar(const volatile unsigned long long *ll, size_t n, unsigned int tag, float *out) {
  size_t i = 0;
  for (; i < n; i++) {
    const unsigned long long v = ll[i];
    if ((unsigned int)(v >> 32) != tag) break;
    const uint32_t bits = (uint32_t)v;
    memcpy(out + i, &bits, 4);
  }
  return i;
}
#if defined(__x86_64__)
#include <immintrin.h>
__attribute__((target("avx2"))) static size_t ll_take_avx2(const volatile unsigned long long *ll, size_t n, unsigned int tag,
                                                          float *out) {
  const __m256i want = _mm256_set1_epi32((int)tag);
  const __m256i pick = _mm256_setr_epi32(0, 2, 4, 6, 1, 3, 5, 7);     // values to the low half, tags to the high half
  size_t i = 0;
  for (; i + 8 <= n; i += 8) {
    const __m256i a = _mm256_permutevar8x32_epi32(_mm256_loadu_si256((const __m256i *)(const void *)(ll + i)), pick);
    const __m256i b = _mm256_permutevar8x32_epi32(_mm256_loadu_si256((const __m256i *)(const void *)(ll + i + 4)), pick);
    const __m256i tags = _mm256_permute2x128_si256(a, b, 0x31);
    if (_mm256_movemask_epi8(_mm256_cmpeq_epi32(tags, want)) != -1) break;
    _mm256_storeu_si256((__m256i *)(out + i), _mm256_permute2x128_si256(a, b, 0x20));
  }
  return i + ll_take_scalar(ll + i, n - i < 8 ? n - i : 8, tag, out + i);
}
#endif
static size_t ll_take(const volatile unsigned long long *ll, size_t n, unsigned int tag, float *out) {
#if defined(__x86_64__)
  static const bool avx2 = __builtin_cpu_supports("avx2");
  if (avx2) return ll_take_avx2(ll, n, tag, out);
#endif
  return ll_take_scalar(ll, n, tag, out);
}

// Make the resident instance exit and wait for it (a no-op when none is running).
static void server_stop() {
  // every path that frees, replaces or rewrites a weight stops the server first: the sibling links learned so far die
  // here (they hold raw handles), whether or not an instance is resident; one token later they are learned again
  g_weight_gen++;
  g_srv.sib.reset();
  if (!g_srv.ready || !g_srv.running) return;
  unsigned int words[kSrvWords] = {2u};
  server_post(words);
  cudaStreamSynchronize(g_dev.stream);
  g_srv.running = false;
  g_srv.sib.reset();
}

// N == 1, operands the decode kernel takes
//
// SIBLING SPECULATION (LAMM_SERVER_SPEC=1, default).  ggml calls wq, wk, wv with one src1 and gate, up with another, one
// synchronous lamm_mul_mat after the other; each costs a job hand-over (PCIe poll, republish, result write-back) of
// several microseconds around ~2 us of streaming.  The path learns, per weight, which weight was called NEXT with a
// bit-identical row (w->sib); the next time that weight is called it posts the jobs of up to two predicted followers
// behind its own, flagged "row unchanged" (they reuse the quantized row in the CTAs' shared memory), so the server runs
// them back to back while the host is still copying the first result out.  A following call is a HIT when it is for
// the predicted weight AND its row memcmp-equals the row those jobs used: it only collects the result.  Anything else
// is a miss: the speculative results are ignored (their jobs still run, a few microseconds of device time) and the
// call is posted normally.  Results are the same bits either way: same kernel, same weights, same quantized row.
static int mul_mat_host_server(lammcu_weight_t w, const float *x_host, float *c_host) {
  const int64_t K = w->K, rows = w->row_end - w->row_begin;
  int rc = server_setup();
  if (rc) return rc;
  if ((size_t)rows > kSrvMaxM || (size_t)K > kSrvMaxK) { set_error("mul_mat_host: shape too large for the server buffers"); return LAMMCU_ERR_SHAPE; }
  unsigned int jobs[3][kSrvWords];
  unsigned int *words = jobs[0];
  if ((rc = decode_server_job(w->type, w->dev, w->stride, rows, K, words))) return rc;
  if (g_srv.running && g_srv.type != w->type) server_stop();
  // Tags restart from 1 long before they could wrap (4e9 jobs = half a day of uninterrupted decoding): tag 0 means
  // "nothing" in the mailbox, in pre_tag and in done[1].  Everything stale then carries a huge tag, never a small one.
  if (g_srv.seq > 0xffffff00u) {
    server_stop();
    // no pair anywhere may still carry a small tag from the first hours of the process (a result entry beyond the
    // rows of every later job in its slot would): wipe every tagged buffer once
    memset((void *)g_srv.mbox, 0, (size_t)kSrvSlots * kSrvWords * 8);
    memset((void *)g_srv.c_ll, 0, kSrvSlots * kSrvMaxM * 8);
    CU_TRY(cudaMemset(g_srv.mbox_dev, 0, (size_t)kSrvSlots * kSrvWords * 8));
    CU_TRY(cudaMemset(g_srv.x_ll, 0, kSrvMaxK * 8));
    CU_TRY(cudaDeviceSynchronize());
    g_srv.seq = 0;
    g_srv.done[1] = 0;
    g_srv.launch_seq0 = 0;
  }
  if (g_srv.running && server_exited()) {            // it idled out (or reached its lifetime) since the last call
    g_srv.running = false;
    g_srv.sib.drop_posted();
    g_srv.x_k = 0;
    // it left an EXIT word under the tag it was waiting for in the device mailbox: never issue that tag again, a new
    // instance's CTAs would take the stale word for their first job
    if ((int)(g_srv.done[1] - g_srv.seq) > 0) g_srv.seq = g_srv.done[1];
  }
  const double t0 = now_us();
  // ggml hands q/k/v (and gate/up) the same src1: when this row is bit-identical to the one the running
  // instance answered last, its quantized copy is still in every CTA's shared memory and the job only
  // says so (word 15) -- no copy to pinned memory, no PCIe pull, no quantize.  A job that ends up on a
  // new instance (below) always carries the row.
  const bool row_eq = g_srv_x_reuse && g_srv.row_k == K && memcmp(g_srv.x_host, x_host, (size_t)K * 4) == 0;
  const bool same = row_eq && g_srv.running && g_srv.x_k == K;
  // learn: the previous call's weight was followed by this one with the same row (or not)
  g_srv.sib.learn(w, row_eq, g_weight_gen);
  unsigned int s = 0;
  // hit: the job is already posted (possibly already answered); a misprediction empties the queue -- whatever those
  // jobs produce is ignored
  const bool hit = g_srv.sib.take_hit(w, same, &s);
  const double t1 = now_us();
  if (!hit) {
    if (!row_eq) {
      memcpy(g_srv.x_host, x_host, (size_t)K * 4);
      g_srv.row_k = K;
    }
    words[15] = same ? 1u : 0u;
    g_srv.x_k = 0;
    // the followers this weight had last time (same type / K: they reuse the staged row and the shared-memory plan)
    int n = 1;
    lammcu_weight *fol[lamm::SibTracker<lammcu_weight>::kMaxFollowers];
    if (g_srv_spec && g_srv_x_reuse && g_srv.running) {
      const int nf = g_srv.sib.followers(w, g_weight_gen, (int64_t)kSrvMaxM, fol);
      for (int j = 0; j < nf; j++) {
        if (decode_server_job(fol[j]->type, fol[j]->dev, fol[j]->stride, fol[j]->row_end - fol[j]->row_begin, K, jobs[n])) break;
        jobs[n][15] = 1u;
        n++;
      }
    }
    server_post_n(jobs, n);
    s = g_srv.seq - (unsigned int)(n - 1);
    g_srv.sib.posted(s + 1u, fol, n - 1);
    // An instance that gave up waiting for tag s has left an EXIT word with that tag in the device
    // mailbox: the job is re-posted under a fresh tag before a new instance is launched.
    if (g_srv.running && server_missed(s)) {
      g_srv.running = false;
      g_srv.sib.drop_posted();
      words[15] = 0u;
      server_post(words);
      s = g_srv.seq;
    }
    if (!g_srv.running && (rc = server_launch(w->type, s))) return rc;
  }
  const double t2 = now_us();
  // the result arrives as {value, tag} pairs, chunk by chunk: take what is there, spin on the rest
  float *out = c_host + w->row_begin;
  size_t got = 0;
  unsigned int spins = 0;
  double t_first = 0;
  while (true) {
    const volatile unsigned long long *buf = g_srv.c_ll + (size_t)(s % kSrvSlots) * kSrvMaxM;
    got += ll_take(buf + got, (size_t)rows - got, s, out + got);
    if (got == (size_t)rows) break;
    __builtin_ia32_pause();
    asm volatile("" ::: "memory");                                      // re-read the pairs
    if (server_missed(s)) {                                             // the idle / lifetime exit raced with the post
      g_srv.sib.drop_posted();
      words[15] = 0u;                                                   // (x_host holds this row: a hit implies row_eq)
      server_post(words);
      s = g_srv.seq;
      got = 0;
      if ((rc = server_launch(w->type, s))) return rc;
      continue;
    }
    if ((++spins & 0x3fffu) == 0) {
      const double now = now_us();
      if (t_first == 0) t_first = now;
      else if (now - t_first > 5e6) {
        cudaError_t e = cudaStreamSynchronize(g_dev.stream);
        g_srv.running = false;
        g_srv.sib.drop_posted();
        set_error("mul_mat_host: the decode server did not answer job %u (row %zu of %lld, %s)", s, got,
                  (long long)rows, cudaGetErrorString(e));
        return LAMMCU_ERR_CUDA;
      }
    }
  }
  const double t3 = now_us();
  g_srv.x_k = K;                                     // the instance that answered holds this row
  g_hp[0] += t1 - t0;  g_hp[2] += t2 - t1;  g_hp[3] += t3 - t2;  g_hp[5] += 1;  g_hp[6] += (hit || words[15]) ? 1 : 0;
  g_hp[7] += hit ? 1 : 0;
  return LAMMCU_OK;
}

int lammcu_set_host_mode(int mode) {
  if (mode < 0 || mode > 1) { set_error("set_host_mode: 0 or 1"); return LAMMCU_ERR_ARG; }
  std::lock_guard<std::mutex> lk(g_mu);
  server_stop();
  if (g_dev.inited) cudaStreamSynchronize(g_dev.stream);
  g_host_mode = mode;
  return LAMMCU_OK;
}
int lammcu_get_host_mode(void) { return g_host_mode; }
int lammcu_set_server_timeouts(unsigned int idle_us, unsigned int life_us) {
  if (idle_us < 1 || idle_us > 1000000 || life_us < 10 || life_us > 4000000) {
    set_error("set_server_timeouts: idle 1..1e6 us, life 10..4e6 us");
    return LAMMCU_ERR_ARG;
  }
  std::lock_guard<std::mutex> lk(g_mu);
  server_stop();
  g_srv_idle_us = idle_us;
  g_srv_life_us = life_us;
  return LAMMCU_OK;
}
/* diagnostics (LAMM_SERVER_TRACE=1): device time stamps of the last 64 server jobs, [64][8] ns:
 * job seen, X republished, X quantized, CTA 0's rows done, CTA 0's chunk written */
int lammcu_server_trace(unsigned long long *out512) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_srv.trace || !out512) { set_error("server_trace: tracing is off"); return LAMMCU_ERR_ARG; }
  server_stop();
  CU_TRY(cudaMemcpy(out512, g_srv.trace, 64 * 8 * 8, cudaMemcpyDeviceToHost));
  return LAMMCU_OK;
}
int lammcu_host_quiesce(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  server_stop();
  return LAMMCU_OK;
}

int lammcu_mul_mat_host(lammcu_weight_t w, const float *x_host, int64_t N, int64_t ldx,
                        float *c_host, int64_t ldc, void *xq_out_host) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!w || !x_host || !c_host || N <= 0) { set_error("mul_mat_host: bad argument"); return LAMMCU_ERR_ARG; }
  std::lock_guard<std::mutex> lk(g_mu);
  const int64_t K = w->K, rows = w->row_end - w->row_begin;
  const size_t xbytes = (size_t)N * K * 4;
  if ((rc = grow(&g_dev.hx, &g_dev.hx_bytes, xbytes, true))) return rc;
  if ((rc = grow(&g_dev.xin, &g_dev.xin_bytes, xbytes, false))) return rc;
  if (N == 1 && g_host_mode == 1 && !xq_out_host && g_decode_kernel && (size_t)K <= kSrvMaxK &&
      (size_t)rows <= kSrvMaxM && decode_supported(w->type, w->dev, w->stride, nullptr, rows, K))
    return mul_mat_host_server(w, x_host, c_host);
  server_stop();                                   // the launches below queue behind a resident instance
  float *c_dev = nullptr;
  if ((rc = out_buffer(w, N, &c_dev))) return rc;
  const double t0 = now_us();
  for (int64_t j = 0; j < N; j++) memcpy((float *)g_dev.hx + j * K, x_host + j * ldx, (size_t)K * 4);
  const double t1 = now_us();
  CU_TRY(cudaMemcpyAsync(g_dev.xin, g_dev.hx, xbytes, cudaMemcpyHostToDevice, g_dev.stream));
  const double t2 = now_us();
  g_hp[0] += t1 - t0;  g_hp[1] += t2 - t1;
  void *ws = nullptr;
  if (xq_out_host && w->type != LAMMCU_TYPE_F32) {
    const size_t qb = lammcu_row_size(type_info(w->type).vec_dot, K) * (size_t)N;
    if ((rc = grow(&g_dev.ws, &g_dev.ws_bytes, qb, false))) return rc;
    ws = g_dev.ws;
  }
  rc = lammcu_mul_mat_dev(w->type, w->dev, w->stride, (const float *)g_dev.xin, K, c_dev,
                          rows, rows, N, K, ws, g_dev.stream);
  g_hp[2] += now_us() - t2;
  if (rc) return rc;
  if (xq_out_host && w->type != LAMMCU_TYPE_F32) {
    const size_t qb = lammcu_row_size(type_info(w->type).vec_dot, K) * (size_t)N;
    CU_TRY(cudaMemcpyAsync(xq_out_host, g_dev.ws, qb, cudaMemcpyDeviceToHost, g_dev.stream));
  }
  return copy_c_back(w, N, c_host, ldc);
}

/* diagnostics: accumulated microseconds {x memcpy, H2D enqueue, kernel enqueue, sync wait, C memcpy, calls} */
void lammcu_host_profile(double *out8, int reset) {
  for (int i = 0; i < 8; i++) { if (out8) out8[i] = g_hp[i]; if (reset) g_hp[i] = 0; }
}

int lammcu_mul_mat_q_host(lammcu_weight_t w, const void *xq_host, int64_t xq_stride, int64_t N,
                          float *c_host, int64_t ldc) {
  int rc = ensure_init();
  if (rc) return rc;
  if (!w || !xq_host || !c_host || N <= 0) { set_error("mul_mat_q_host: bad argument"); return LAMMCU_ERR_ARG; }
  std::lock_guard<std::mutex> lk(g_mu);
  server_stop();
  const int64_t K = w->K, rows = w->row_end - w->row_begin;
  const int vt = type_info(w->type).vec_dot;
  const size_t rs = (vt == LAMMCU_TYPE_F32) ? (size_t)K * 4 : lammcu_row_size(vt, K);
  if ((size_t)xq_stride < rs) { set_error("mul_mat_q_host: activation stride smaller than a row"); return LAMMCU_ERR_SHAPE; }
  if ((rc = grow(&g_dev.ws, &g_dev.ws_bytes, rs * (size_t)N, false))) return rc;
  float *c_dev = nullptr;
  if ((rc = out_buffer(w, N, &c_dev))) return rc;
  CU_TRY(cudaMemcpy2DAsync(g_dev.ws, rs, xq_host, (size_t)xq_stride, rs, (size_t)N, cudaMemcpyHostToDevice,
                           g_dev.stream));
  rc = lammcu_mul_mat_q_dev(w->type, w->dev, w->stride, g_dev.ws, (int64_t)rs, c_dev, rows,
                            rows, N, K, g_dev.stream);
  if (rc) return rc;
  return copy_c_back(w, N, c_host, ldc);
}

}  // extern "C"
