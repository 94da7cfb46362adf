// C ABI of libdimreduce4gpu.so (include/dimreduce4gpu.h): the persistent device context, the
// fit / transform orchestration and the reference's legacy host-pointer entry points.
//
// What it replaces in the reference:
//   truncated_svd_float/double   src/tsvd.cu:427-452   (H2D copy + dispatch)
//   truncated_svd_matrix<T>      src/tsvd.cu:454-485   (algorithm string dispatch, try/catch)
//   cusolver_tsvd / power_tsvd   src/tsvd.cu:315-333, 346-414
//   get_tsvd_attr                src/tsvd.cu:186-291
//   pca_float/double             src/pca.cu:22-116
//   DeviceContext, Matrix<T>     src/device_context.cuh:8-30, src/matrix.cuh:16-303
// Output conventions are the CPU backend's (src/cpu_backend.cpp:277-332) because that is the
// implementation parity is judged against.
//
// Pipeline of one fit (all on one stream, no host round trip until the caller synchronises):
//   [0] pilot shift   (PCA only) column means of the first rows  -> shift c        (K1)
//   [1] streamed pass G_c = (X-c)^T (X-c), s_c = sum(x-c), q_c = sum (x-c)^2, n    (K2 [+K1])
//   [2] one allreduce of [G|s|q|n] across ranks; C = G_c - s_c s_c^T / n           (NCCL, K3)
//   [3] eigensolve of C: blocked Jacobi ("cusolver") or subspace iteration (else)  (K4 / K5)
//   [4] sort, sign, truncate, sigma / explained variance / ratio / mean            (K6)
// transform is a second streamed pass (K7), shard-local.
#include <sys/mman.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <map>
#include <memory>
#include <thread>
#include <vector>

#include "common.cuh"
#include "nccl_dyn.h"
#include "../../include/dimreduce4gpu.h"

namespace dr4g {

// ---------------------------------------------------------------- errors
static thread_local char g_err[1024] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }

void fail(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  throw Error(buf);
}

// ---------------------------------------------------------------- arena
void Arena::reserve(size_t bytes) {
  if (bytes <= cap) return;
  if (base) {
    DR4G_CUDA(cudaDeviceSynchronize());
    DR4G_CUDA(cudaFree(base));
    base = nullptr;
    cap = 0;
  }
  size_t want = bytes + (bytes >> 3) + (1u << 20);
  DR4G_CUDA(cudaMalloc(&base, want));
  cap = want;
}

void* Arena::take(size_t bytes, size_t align) {
  size_t o = (off + align - 1) / align * align;
  if (o + bytes > cap) fail("workspace arena overflow: need %zu at offset %zu, capacity %zu", bytes, o, cap);
  off = o + bytes;
  return base + o;
}

void Arena::release() {
  if (base) cudaFree(base);
  base = nullptr;
  cap = off = 0;
}

static size_t up256(size_t b) { return (b + 255) / 256 * 256; }

// ---------------------------------------------------------------- context
struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) DR4G_CUDA(cudaSetDevice(dev));
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

// Streaming (out-of-core) fit, SURVEY 8f-1: [G | s | q | n] of (x - shift) accumulated over row chunks.
// The shift is fixed by the first chunk (it has to be the same for every chunk and every rank).
struct StreamFit {
  double* acc = nullptr;      // packed accumulators, d*d + 2d + 1 doubles
  double* shift_d = nullptr;  // d doubles
  void* shift_t = nullptr;    // d elements of the input type
  int d = 0, center = 0, elem = 0;
  long long rows = 0;         // rows accumulated on this rank
  void release() {
    if (acc) cudaFree(acc);
    if (shift_d) cudaFree(shift_d);
    if (shift_t) cudaFree(shift_t);
    acc = shift_d = nullptr;
    shift_t = nullptr;
    d = center = elem = 0;
    rows = 0;
  }
};

// Pageable host memory <-> device through pinned bounce buffers.  cudaMemcpyAsync on a pageable pointer is staged by
// the driver on ONE thread (~6-12 GB/s measured through the estimator's numpy arrays, against ~55 GB/s of the PCIe
// link from pinned memory), and the reference's wrapper hands the library exactly such arrays
// (dimreduce4gpu/truncated_svd.py:72-86).  Here a few host threads fill a ring of pinned 32 MB buffers while the
// previous buffers are in flight on the copy engine.
struct HostStage {
  // four 8 MB buffers (round 2 began with three of 32 MB: a 51 MB array - the reference's own bench shape 50,000 x 256 -
  // was two sub-chunks, so the host copy and the DMA hardly overlapped; 8 MB is ~0.15 ms on the link)
  static constexpr int NB = 4;
  static constexpr size_t kSub = (size_t)8 << 20;
  void* buf[NB] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t done[NB] = {nullptr, nullptr, nullptr, nullptr};
  bool busy[NB] = {false, false, false, false};
  int next = 0;
  // a D2H sub-chunk that is in flight into buf[b] and still has to be copied out to the caller's array
  struct Out { void* dst; size_t bytes; } out[NB] = {};
  void ensure() {
    for (int b = 0; b < NB; ++b) {
      if (!buf[b]) DR4G_CUDA(cudaHostAlloc(&buf[b], kSub, cudaHostAllocDefault));
      if (!done[b]) DR4G_CUDA(cudaEventCreateWithFlags(&done[b], cudaEventDisableTiming));
    }
  }
  void release() {
    for (int b = 0; b < NB; ++b) {
      if (buf[b]) cudaFreeHost(buf[b]);
      if (done[b]) cudaEventDestroy(done[b]);
      buf[b] = nullptr; done[b] = nullptr; busy[b] = false; out[b] = Out{};
    }
  }
};

static int stage_threads() {
  static const int n = [] {
    const char* e = getenv("DR4G_STAGE_THREADS");
    int v = e ? atoi(e) : 0;
    if (v <= 0) {
      const unsigned hc = std::thread::hardware_concurrency();
      v = hc >= 16 ? 6 : (hc >= 8 ? 4 : 2);
    }
    return v > 16 ? 16 : v;
  }();
  return n;
}

// Persistent helpers for the staging copies (a std::thread per slice and sub-chunk cost more than the copy itself once the
// sub-chunks are 8 MB).  One job at a time (the callers hold the context's mutex, the pool its own): the caller takes
// slice 0, the workers the others; also used to touch fresh output pages from several cores at once.
class CopyPool {
 public:
  static CopyPool& get() {
    static CopyPool* p = new CopyPool(stage_threads() - 1);   // leaked on purpose: workers may outlive static destructors
    return *p;
  }
  // src == nullptr: write-touch every page of dst (contents kept) - takes the page faults of a fresh array early
  static void do_slice(char* d, const char* s, size_t len) {
    if (s) { std::memcpy(d, s, len); return; }
    for (size_t o = 0; o < len; o += 4096) { volatile char* q = d + o; *q = *q; }
  }
  void copy(void* dst, const void* src, size_t bytes) {
    const int nt = bytes < ((size_t)1 << 20) ? 1 : (int)workers_.size() + 1;
    if (nt <= 1) { do_slice(static_cast<char*>(dst), static_cast<const char*>(src), bytes); return; }
    std::lock_guard<std::mutex> job(job_mu_);
    const size_t per = ((bytes + nt - 1) / nt + 4095) & ~(size_t)4095;
    {
      std::lock_guard<std::mutex> lk(mu_);
      dst_ = static_cast<char*>(dst); src_ = static_cast<const char*>(src); bytes_ = bytes; per_ = per;
      pending_ = (int)workers_.size();
      ++gen_;
    }
    cv_.notify_all();
    do_slice(static_cast<char*>(dst), static_cast<const char*>(src), per < bytes ? per : bytes);
    std::unique_lock<std::mutex> lk(mu_);
    done_cv_.wait(lk, [&] { return pending_ == 0; });
  }

 private:
  explicit CopyPool(int n) {
    for (int i = 0; i < (n < 0 ? 0 : n); ++i) workers_.emplace_back([this, i] { run(i + 1); });
    for (auto& t : workers_) t.detach();
  }
  void run(int slice) {
    unsigned long long seen = 0;
    for (;;) {
      std::unique_lock<std::mutex> lk(mu_);
      cv_.wait(lk, [&] { return gen_ != seen; });
      seen = gen_;
      char* d = dst_; const char* s = src_; const size_t bytes = bytes_, per = per_;
      lk.unlock();
      const size_t off = (size_t)slice * per;
      if (off < bytes) do_slice(d + off, s ? s + off : nullptr, bytes - off < per ? bytes - off : per);
      lk.lock();
      if (--pending_ == 0) done_cv_.notify_one();
    }
  }
  std::vector<std::thread> workers_;
  std::mutex mu_, job_mu_;
  std::condition_variable cv_, done_cv_;
  unsigned long long gen_ = 0;
  int pending_ = 0;
  char* dst_ = nullptr;
  const char* src_ = nullptr;
  size_t bytes_ = 0, per_ = 0;
};

static void parallel_memcpy(void* dst, const void* src, size_t bytes) { CopyPool::get().copy(dst, src, bytes); }
static void parallel_touch(void* dst, size_t bytes) { CopyPool::get().copy(dst, nullptr, bytes); }

// Fresh output arrays (the wrapper's np.zeros: untouched calloc pages) take a page fault per 4 KB when they are first
// written - 256 MB of scores cost ~10 ms of a 36 ms call.  While the input is still on its way to the device, helper
// threads ask the kernel to back the arrays with huge pages where a 2 MB aligned piece exists (MADV_HUGEPAGE) and to
// populate them writable in one go (MADV_POPULATE_WRITE, Linux >= 5.14): contents are kept, only the faults move out
// of the copy-out.  Kernels without it (EINVAL) fall back to write-touching the pages from the copy pool.
#ifndef MADV_POPULATE_WRITE
#define MADV_POPULATE_WRITE 23
#endif
struct Prefault {
  std::vector<std::thread> th;
  std::atomic<int> failed{0};
  void start(void* p, size_t bytes, int nthreads) {
    static const bool off = getenv("DR4G_NO_MADVISE") != nullptr;
    if (off || !p || bytes < ((size_t)4 << 20)) { failed = 1; return; }
    const uintptr_t a = (reinterpret_cast<uintptr_t>(p) + 4095) & ~(uintptr_t)4095;
    const uintptr_t e = (reinterpret_cast<uintptr_t>(p) + bytes) & ~(uintptr_t)4095;
    if (e <= a) return;
    const uintptr_t m2 = ((uintptr_t)2 << 20) - 1, a2 = (a + m2) & ~m2, e2 = e & ~m2;
    if (e2 > a2) madvise(reinterpret_cast<void*>(a2), e2 - a2, MADV_HUGEPAGE);   // advice only: failure is harmless
    const size_t per = (((e - a) + nthreads - 1) / nthreads + m2) & ~(size_t)m2;
    for (int t = 0; t < nthreads; ++t) {
      const uintptr_t lo = a + (uintptr_t)t * per;
      if (lo >= e) break;
      const uintptr_t hi = lo + per < e ? lo + per : e;
      th.emplace_back([this, lo, hi] {
        for (uintptr_t q = lo; q < hi; q += (size_t)16 << 20) {
          const size_t len = hi - q < ((size_t)16 << 20) ? hi - q : ((size_t)16 << 20);
          if (madvise(reinterpret_cast<void*>(q), len, MADV_POPULATE_WRITE) != 0) { failed = 1; return; }
        }
      });
    }
  }
  bool finish() {   // true: every page is populated
    for (auto& t : th) if (t.joinable()) t.join();
    th.clear();
    return failed.load() == 0;
  }
  ~Prefault() { finish(); }
};

// DR4G_LEGACY_TRACE=1: host-side timeline of a legacy call (microseconds since entry) on stderr
struct LegacyTrace {
  bool on;
  std::chrono::steady_clock::time_point t0;
  std::vector<std::pair<const char*, double>> ev;
  LegacyTrace() : on(getenv("DR4G_LEGACY_TRACE") != nullptr), t0(std::chrono::steady_clock::now()) {}
  void mark(const char* what) {
    if (on) ev.emplace_back(what, std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count());
  }
  ~LegacyTrace() {
    if (!on) return;
    fprintf(stderr, "[dr4g legacy trace]");
    for (auto& e : ev) fprintf(stderr, " %s=%.0f", e.first, e.second);
    fprintf(stderr, "\n");
  }
};
static thread_local LegacyTrace* g_ltrace = nullptr;
#define LTRACE(x) do { if (g_ltrace) g_ltrace->mark(x); } while (0)

static bool is_pageable_host(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return true;
  }
  return a.type == cudaMemoryTypeUnregistered;
}

struct CtxFull : Ctx {
  std::mutex mu;
  HostStage stage;
  Arena io;  // device copies of inputs/outputs for the legacy host-pointer symbols
  StreamFit sf;
  cudaStream_t copy_stream = nullptr, out_stream = nullptr;  // H2D / D2H of the out-of-core legacy path
  cudaEvent_t switch_ev = nullptr;                            // orders a newly selected stream behind the old one
};

static CtxFull* ctx_new(int device) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess) fail("no usable CUDA device: %s", cudaGetErrorString(e));
  if (device < 0 || device >= count) fail("gpu_id %d out of range (device count %d)", device, count);
  DeviceGuard g(device);
  cudaDeviceProp prop;
  DR4G_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    fail("libdimreduce4gpu.so is built for sm_100a (B200); device %d is sm_%d%d", device, prop.major, prop.minor);
  std::unique_ptr<CtxFull> c(new CtxFull);
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  DR4G_CUDA(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
  c->stream = c->own_stream;
  for (auto& ev : c->stage_ev) DR4G_CUDA(cudaEventCreate(&ev));
  return c.release();
}

static void ctx_delete(CtxFull* c) {
  if (!c) return;
  DeviceGuard g(c->device);
  cudaStreamSynchronize(c->stream);
  if (c->comm) {
    try { nccl_api()->CommDestroy(c->comm); } catch (...) {}
  }
  for (auto& ev : c->stage_ev) if (ev) cudaEventDestroy(ev);
  c->arena.release();
  c->io.release();
  if (c->proj_buf) cudaFree(c->proj_buf);
  if (c->pad_buf) cudaFree(c->pad_buf);
  if (c->fold_buf) cudaFree(c->fold_buf);
  for (int r = 0; r < kP2PMaxWorld; ++r) {
    if (!c->p2p_peer[r]) continue;
    if (r == c->rank) cudaFree(c->p2p_peer[r]);
    else cudaIpcCloseMemHandle(c->p2p_peer[r]);
    c->p2p_peer[r] = nullptr;
  }
  if (c->p2p_status_host) cudaFreeHost(const_cast<int*>(c->p2p_status_host));
  c->stage.release();
  if (c->switch_ev) cudaEventDestroy(c->switch_ev);
  c->sf.release();
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  if (c->out_stream) cudaStreamDestroy(c->out_stream);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  delete c;
}

// per-device contexts behind the legacy symbols (created on first use, live for the process)
static std::mutex g_legacy_mu;
static std::map<int, CtxFull*> g_legacy;

static CtxFull* legacy_ctx(int device) {
  std::lock_guard<std::mutex> lk(g_legacy_mu);
  auto it = g_legacy.find(device);
  if (it != g_legacy.end()) return it->second;
  CtxFull* c = ctx_new(device);
  g_legacy[device] = c;
  return c;
}

// ---------------------------------------------------------------- fit
static bool is_exact(const params& p) {
  // src/tsvd.cu:457-471: "cusolver" -> exact eigensolve, anything else -> power
  return p.algorithm != nullptr && std::strcmp(p.algorithm, "cusolver") == 0;
}

static void allreduce_f64(Ctx& c, double* buf, size_t count) {
  if (c.world <= 1) return;
  NcclApi* api = nccl_api();
  int rc = api->AllReduce(buf, buf, count, kNcclFloat64, kNcclSum, c.comm, c.stream);
  if (rc != 0) fail("ncclAllReduce failed: %s", api->GetErrorString(rc));
}

static void allgather_f64(Ctx& c, const double* send, double* recv, size_t count) {
  NcclApi* api = nccl_api();
  int rc = api->AllGather(send, recv, count, kNcclFloat64, c.comm, c.stream);
  if (rc != 0) fail("ncclAllGather failed: %s", api->GetErrorString(rc));
}

constexpr int kOversample = 10;  // same block oversampling as the CPU backend's randomized path (src/cpu_backend.cpp:170)


// Row folding of narrow fp32 inputs (finalize.cu): f consecutive rows are streamed as one row of f d <= 256 columns, so
// the pass runs on the CTA-pair tcgen05 kernel at its best shape whatever d is (the single-CTA kernel needs the same
// time per ROW for every d <= 128: 2.5 TB/s at d = 128, 0.8 at d = 32, 0.2 at d = 8; folded: 3.6 TB/s for all of them).
// f d must keep the 16-byte row pitch, so d % 4 != 0 folds by multiples of 2 or 4 and needs no padded copy either.
// Small inputs stay as they are (100,000 x 64: 54 us unfolded, 66 us folded).  0: do not fold.
static int gram_fold_factor(const Ctx& c, const float* dX, long long n, int d) {
  if (!c.fold || c.gram_path != 0 || d < 1 || d > 128) return 0;
  if ((reinterpret_cast<uintptr_t>(dX) & 15) != 0) return 0;
  if ((double)n * (double)d < 16777216.0) return 0;
  const int step = d % 4 == 0 ? 1 : (d % 2 == 0 ? 2 : 4);
  const int f = (256 / d) / step * step;
  if (f < 2 || f * d <= 128) return 0;
  if (n / f < 1 || !gram_tc_eligible(c, dX, n / f, f * d)) return 0;
  return f;
}
// fp64: the same fold onto gram_f64.cu's 128 x 128 DMMA tiles.  That kernel spends the fp64 pipe on a full 128-wide tile
// whatever d is (4M x 64: 11.4 ms, 8M x 32: 22 ms, 16M x 16: 44 ms for 2 GiB each); folded to 256 columns they all take the
// three tiles of a 256-column problem on n / f rows (~4.3 ms).  f d must be even (16-byte rows for cp.async).
static int gram_fold_factor(const Ctx& c, const double* dX, long long n, int d) {
  if (!c.fold || c.gram_path != 0 || d < 1 || d > 128) return 0;
  if ((reinterpret_cast<uintptr_t>(dX) & 15) != 0) return 0;
  if ((double)n * (double)d < 8388608.0) return 0;
  const int step = d % 2 == 0 ? 1 : 2;
  const int f = (256 / d) / step * step;
  if (f < 2 || f * d <= 128) return 0;
  if (n / f < 1 || !gram_f64_eligible(c, dX, n / f, f * d)) return 0;
  return f;
}
template <typename T>
static void fit_impl(Ctx& c, const T* dX, long long n_local, int d, int k, int center, const params& p, T* dQ,
                     T* dw, T* dev, T* devr, T* dmean, StreamFit* sf = nullptr, int phase = 0) {
  // phase 0: whole fit.  phase 1 (sf): pilot shift on the first chunk, pass over this chunk, sf->acc += result.
  // phase 2 (sf): no pass, the accumulated [G|s|q|n] goes through the exchange, eigensolve and outputs.
  if (d <= 0 || k <= 0 || n_local < 0) fail("fit: bad shape n_local=%lld d=%d k=%d", n_local, d, k);
  if (k > d) k = d;
  // "cusolver" -> exact eigensolve, anything else -> 'power' (src/tsvd.cu:457-471) - except that small problems take the
  // exact path whatever the string says, as the CPU backend (the parity oracle) does: use_exact = algorithm ==
  // "cusolver" || min(n, d) <= 256 (src/cpu_backend.cpp:347,369).  The exact solve of a d <= 256 covariance costs less
  // than 1 ms here, and BASELINE config 1 (TruncatedSVD default algorithm on 10,000 x 100) then matches the reference at
  // 1e-4 instead of at subspace-iteration accuracy.  Option power_min_dim = 0 switches the rule off (tests of K5).
  const bool exact = is_exact(p) || d <= c.power_min_dim || (c.world <= 1 && phase == 0 && n_local <= c.power_min_dim);
  const size_t dd = (size_t)d * d;
  const size_t packed_count = dd + 2 * (size_t)d + 1;
  const size_t ext_count = packed_count + (size_t)d;   // + this rank's pilot shift: what travels in the exchange
  const int dp = jacobi_padded_dim(d);
  int l = k + kOversample;
  if (l > d) l = d;
  const int lp = power_padded_block(l);

  // workspace
  size_t need = 2 * up256(ext_count * 8) + up256((size_t)d * 8) * 4 + up256((size_t)d * sizeof(T)) + up256(dd * 8) + 4096;
  need += up256(((size_t)d + 1) * 8);
  if (c.world > 1 && center) need += up256((size_t)c.world * ext_count * 8);
  const bool use_jacobi = c.eig_solver == 1;
  if (exact && use_jacobi)
    need += 2 * up256((size_t)dp * dp * 8) + up256((size_t)dp * 8) + up256(jacobi_workspace_bytes(dp));
  else if (exact)
    need += up256(eigh_topk_workspace_bytes(d, k)) + up256((size_t)k * 8) + up256((size_t)k * d * 8);
  else
    need += up256(power_workspace_bytes(d, l)) + up256((size_t)lp * 8) + up256((size_t)lp * d * 8);
  need += up256((size_t)(exact ? dp : lp) * 8) + up256((size_t)k * 4) + 64 * 256;
  // Tiny fp32 inputs (<= 1 MB: the reference's own test shapes, 4 x 3 ... 256 x 256) take the generic kernel - fp64
  // products of the fp32 input, fp64 accumulation - instead of the tensor-core split: the pass is launch-bound either
  // way, and the split's ~1e-5 of G shows where a rank-deficient input is reconstructed from ALL its components
  // (tests/test_gpu_reference_suite.py: 2.1e-6 against scikit-learn's 7.5e-7 and the reference's 1.75 x bound).
  const bool tiny = sizeof(T) == 4 && c.gram_path == 0 && (double)n_local * (double)d <= 262144.0;
  bool want_tc = false;
  if constexpr (sizeof(T) == 4)
    want_tc = !tiny && c.gram_path != 1 && gram_tc_eligible(c, reinterpret_cast<const float*>(dX), n_local, d);
  if (want_tc) need += up256(gram_tc_workspace_bytes(c, d));
  bool want_f64 = false;
  if constexpr (sizeof(T) == 8)
    want_f64 = c.gram_path != 5 && phase != 2 && gram_f64_eligible(c, reinterpret_cast<const double*>(dX), n_local, d);
  if (want_f64) need += up256(gram_f64_workspace_bytes(c, n_local, d));
  // Shapes the tensor-core / DMMA kernels cannot read in place (fp32: d % 4 != 0, fp64: odd d, or a base pointer that is
  // not 16-byte aligned - rows without a 16-byte pitch): the pass runs over zero-padded CHUNK COPIES of X (row pitch dpad)
  // and the padded moments are cut back to d afterwards.  These shapes used to take the generic kernel: 36 GB/s at
  // 500,000 x 1001 fp32, 25x slower than d = 1000.
  int fold = 0;
  if constexpr (sizeof(T) == 4) {
    if (phase != 2) fold = gram_fold_factor(c, reinterpret_cast<const float*>(dX), n_local, d);
    if (fold) {
      const size_t df = (size_t)fold * d;
      need += up256(df * df * 8) + up256(2 * df * 8) + up256(df * 4) + up256(dd * 8) + up256((size_t)2 * d * 8) +
              up256(gram_tc_workspace_bytes(c, (int)df)) + 2048;
    }
  } else {
    if (phase != 2) fold = gram_fold_factor(c, reinterpret_cast<const double*>(dX), n_local, d);
    if (fold) {
      const size_t df = (size_t)fold * d;
      need += up256(df * df * 8) + up256(df * 8) + up256(dd * 8) + up256(gram_f64_workspace_bytes(c, n_local / fold, (int)df)) + 2048;
      want_f64 = false;   // the folded pass replaces the unfolded one
    }
  }
  int dpad = 0;
  long long pad_rows = 0;
  // (fp64 with d <= 128 stays on the generic kernel: the 128-wide tiles of gram_f64 are mostly padding there and the
  // copy costs more than it saves - 1M x 99: 2.8 ms generic, 3.5 ms padded)
  if (phase != 2 && !fold && !tiny && !want_tc && !want_f64 && c.gram_path == 0 && n_local > 0 && d >= 2 && (sizeof(T) == 4 || d > 128)) {
    const int q4 = sizeof(T) == 4 ? 4 : 2;
    dpad = (d + q4 - 1) / q4 * q4;
    pad_rows = (long long)(((size_t)128 << 20) / ((size_t)dpad * sizeof(T)));
    if (pad_rows < 1024) pad_rows = 1024;
    if (pad_rows > n_local) pad_rows = n_local;
    need += up256((size_t)pad_rows * dpad * sizeof(T)) + 2 * up256((size_t)dpad * dpad * 8) + up256((size_t)2 * dpad * 8) +
            up256((size_t)dpad * sizeof(T)) + 2048;
    if constexpr (sizeof(T) == 4) need += up256(gram_tc_workspace_bytes(c, dpad));
    else need += up256(gram_f64_workspace_bytes(c, pad_rows, dpad));
  }
  c.arena.reserve(need);
  c.arena.reset();

  double* packed = c.arena.take_n<double>(ext_count);   // [G | s | q | n | c]
  double* G = packed;
  double* s = packed + dd;
  double* q = s + d;
  double* cnt = q + d;
  double* pilot = c.arena.take_n<double>((size_t)d + 1);
  double* shift_d = cnt + 1;                            // the pilot shift travels with the moments
  T* shift_t = c.arena.take_n<T>(d);
  if (sf) {
    shift_d = sf->shift_d;
    shift_t = static_cast<T*>(sf->shift_t);
  }
  double* mean_d = c.arena.take_n<double>(d);
  double* total_var = c.arena.take_n<double>(1);

  const bool timing = c.timing != 0;
  auto mark = [&](int i) { if (timing) DR4G_CUDA(cudaEventRecord(c.stage_ev[i], c.stream)); };
  c.stage_valid = false;
  mark(0);

  // [0] pilot shift: column means of this rank's first rows.  Every rank keeps ITS OWN shift (no collective before the
  // pass); the exchange re-bases the moments to a common shift (finalize.cu: rebase_sum_*).
  const T* shift_arg = nullptr;
  const double* shift_d_arg = nullptr;
  if (center && sf && !(phase == 1 && sf->rows == 0)) {
    // streaming: the shift of the first chunk stays in force
    if constexpr (sizeof(T) == 8) shift_arg = reinterpret_cast<const T*>(shift_d);
    else shift_arg = shift_t;
    shift_d_arg = shift_d;
  } else if (center) {
    long long rows = n_local < c.pilot_rows ? n_local : c.pilot_rows;
    DR4G_CUDA(cudaMemsetAsync(pilot, 0, ((size_t)d + 1) * 8, c.stream));
    DR4G_CUDA(cudaMemsetAsync(shift_d, 0, (size_t)d * 8, c.stream));
    launch_colstats<T>(c, dX, rows, d, nullptr, pilot, shift_d /* scratch for sumsq */);
    launch_set_scalar(c, pilot + d, (double)rows);
    if constexpr (sizeof(T) == 4)
      launch_pilot_finish(c, pilot, d, reinterpret_cast<float*>(shift_t), shift_d);
    else
      launch_pilot_finish(c, pilot, d, nullptr, shift_d);
    if constexpr (sizeof(T) == 8) shift_arg = reinterpret_cast<const T*>(shift_d);
    else shift_arg = shift_t;
    shift_d_arg = shift_d;
  }
  mark(1);

  // [1] the streamed pass
  if (phase == 2) {
    DR4G_CUDA(cudaMemcpyAsync(packed, sf->acc, packed_count * 8, cudaMemcpyDeviceToDevice, c.stream));
    DR4G_CUDA(cudaMemcpyAsync(cnt + 1, sf->shift_d, (size_t)d * 8, cudaMemcpyDeviceToDevice, c.stream));
    shift_d_arg = center ? cnt + 1 : nullptr;
  } else {
  DR4G_CUDA(cudaMemsetAsync(packed, 0, packed_count * 8, c.stream));
  if (!center) DR4G_CUDA(cudaMemsetAsync(cnt + 1, 0, (size_t)d * 8, c.stream));
  bool used_tc = false;
  c.last_gram_kernel = 0;
  bool padded_done = false;
  c.last_fold = 0;
  if (fold > 0) {
    if constexpr (sizeof(T) == 4) {
      const int df = fold * d;
      const long long nf = n_local / fold, tail = n_local - nf * fold;
      const float* Xf = reinterpret_cast<const float*>(dX);
      double* Gf = c.arena.take_n<double>((size_t)df * df);
      double* sqf = c.arena.take_n<double>((size_t)2 * df);
      float* shf = c.arena.take_n<float>(df);
      double* Gt = c.arena.take_n<double>(dd);
      double* sqt = c.arena.take_n<double>((size_t)2 * d);
      void* ws = c.arena.take(gram_tc_workspace_bytes(c, df));
      DR4G_CUDA(cudaMemsetAsync(sqf, 0, (size_t)2 * df * 8, c.stream));
      const float* shf_arg = nullptr;
      if (shift_arg) {
        launch_tile_shift(c, reinterpret_cast<const float*>(shift_arg), d, fold, shf);
        shf_arg = shf;
      }
      launch_gram_tc(c, Xf, nf, df, shf_arg, Gf, sqf, sqf + df, ws);
      if (tail > 0) {   // the last rows do not fill a folded row: fp64 products on the generic kernel
        DR4G_CUDA(cudaMemsetAsync(Gt, 0, dd * 8, c.stream));
        DR4G_CUDA(cudaMemsetAsync(sqt, 0, (size_t)2 * d * 8, c.stream));
        const float* Xt = Xf + (size_t)nf * df;
        launch_colstats<float>(c, Xt, tail, d, reinterpret_cast<const float*>(shift_arg), sqt, sqt + d);
        launch_gram_generic<float>(c, Xt, tail, d, reinterpret_cast<const float*>(shift_arg), Gt);
        launch_mirror_upper(c, Gt, d);
      }
      launch_unfold_moments(c, Gf, sqf, sqf + df, d, fold, tail > 0 ? Gt : nullptr, tail > 0 ? sqt : nullptr,
                            tail > 0 ? sqt + d : nullptr, G, s, q);
      padded_done = true;
      used_tc = true;
      c.last_fold = fold;
    } else {
      const int df = fold * d;
      const long long nf = n_local / fold, tail = n_local - nf * fold;
      const double* Xd = reinterpret_cast<const double*>(dX);
      const double* shd = reinterpret_cast<const double*>(shift_arg);
      double* Gf = c.arena.take_n<double>((size_t)df * df);
      double* shf = c.arena.take_n<double>(df);
      double* Gt = c.arena.take_n<double>(dd);
      void* ws = c.arena.take(gram_f64_workspace_bytes(c, nf, df));
      // s and q: one column-statistics pass over the rows as they are (K1); G: the folded rows on the DMMA tiles
      launch_colstats<double>(c, Xd, n_local, d, shd, s, q);
      DR4G_CUDA(cudaMemsetAsync(Gf, 0, (size_t)df * df * 8, c.stream));
      const double* shf_arg = nullptr;
      if (shd) {
        launch_tile_shift(c, shd, d, fold, shf);
        shf_arg = shf;
      }
      launch_gram_f64(c, Xd, nf, df, shf_arg, Gf, ws);
      launch_mirror_upper(c, Gf, df);   // a d x d diagonal block may straddle two 128-column tiles
      if (tail > 0) {
        DR4G_CUDA(cudaMemsetAsync(Gt, 0, dd * 8, c.stream));
        launch_gram_generic<double>(c, Xd + (size_t)nf * df, tail, d, shd, Gt);
        launch_mirror_upper(c, Gt, d);
      }
      launch_unfold_moments(c, Gf, nullptr, nullptr, d, fold, tail > 0 ? Gt : nullptr, nullptr, nullptr, G, nullptr, nullptr);
      padded_done = true;
      c.last_gram_kernel = 4;
      c.last_fold = fold;
    }
  }
  if (dpad > 0) {
    T* Xp = c.arena.take_n<T>((size_t)pad_rows * dpad);
    double* Gp = c.arena.take_n<double>((size_t)dpad * dpad);
    double* Gacc = c.arena.take_n<double>((size_t)dpad * dpad);
    double* sq = c.arena.take_n<double>((size_t)2 * dpad);
    T* shp = c.arena.take_n<T>(dpad);
    bool ok = true;
    if constexpr (sizeof(T) == 4) ok = gram_tc_eligible(c, reinterpret_cast<const float*>(Xp), pad_rows, dpad);
    else ok = gram_f64_eligible(c, reinterpret_cast<const double*>(Xp), pad_rows, dpad);
    if (ok) {
      void* ws = nullptr;
      if constexpr (sizeof(T) == 4) ws = c.arena.take(gram_tc_workspace_bytes(c, dpad));
      else ws = c.arena.take(gram_f64_workspace_bytes(c, pad_rows, dpad));
      DR4G_CUDA(cudaMemsetAsync(Gacc, 0, (size_t)dpad * dpad * 8, c.stream));
      DR4G_CUDA(cudaMemsetAsync(sq, 0, (size_t)2 * dpad * 8, c.stream));
      const T* shp_arg = nullptr;
      if (shift_arg) {
        launch_pad_rows<T>(c, shift_arg, 1, d, dpad, shp);
        shp_arg = shp;
      }
      for (long long r0 = 0; r0 < n_local; r0 += pad_rows) {
        const long long rows = n_local - r0 < pad_rows ? n_local - r0 : pad_rows;
        launch_pad_rows<T>(c, dX + (size_t)r0 * d, rows, d, dpad, Xp);
        if constexpr (sizeof(T) == 4) {
          launch_gram_tc(c, reinterpret_cast<const float*>(Xp), rows, dpad, reinterpret_cast<const float*>(shp_arg), Gp, sq,
                         sq + dpad, ws);
        } else {
          launch_colstats<T>(c, Xp, rows, dpad, shp_arg, sq, sq + dpad);
          DR4G_CUDA(cudaMemsetAsync(Gp, 0, (size_t)dpad * dpad * 8, c.stream));
          launch_gram_f64(c, reinterpret_cast<const double*>(Xp), rows, dpad, reinterpret_cast<const double*>(shp_arg), Gp, ws);
          launch_mirror_upper(c, Gp, dpad);
          c.last_gram_kernel = 4;
        }
        launch_add_f64(c, Gacc, Gp, (size_t)dpad * dpad);
      }
      launch_unpad_moments(c, Gacc, sq, sq + dpad, d, dpad, G, s, q);
      padded_done = true;
      used_tc = sizeof(T) == 4;
    }
  }
  if (padded_done) {
  } else if constexpr (sizeof(T) == 4) {
    const float* Xf = reinterpret_cast<const float*>(dX);
    if (want_tc) {
      void* slabs = c.arena.take(gram_tc_workspace_bytes(c, d));
      launch_gram_tc(c, Xf, n_local, d, reinterpret_cast<const float*>(shift_arg), G, s, q, slabs);
      used_tc = true;
    } else if (c.gram_path >= 2) {
      fail("gram_path=2 (tcgen05) requested but the shape is not eligible: n=%lld d=%d", n_local, d);
    }
  }
  if (!used_tc && !padded_done) {
    launch_colstats<T>(c, dX, n_local, d, shift_arg, s, q);
    bool done = false;
    if constexpr (sizeof(T) == 8) {
      if (want_f64) {
        void* ws = c.arena.take(gram_f64_workspace_bytes(c, n_local, d));
        launch_gram_f64(c, reinterpret_cast<const double*>(dX), n_local, d, reinterpret_cast<const double*>(shift_arg), G, ws);
        done = true;
      }
    }
    if (!done) launch_gram_generic<T>(c, dX, n_local, d, shift_arg, G);
    launch_mirror_upper(c, G, d);
    if (done) c.last_gram_kernel = 4;
  }
  c.last_gram_path = (!used_tc && c.last_gram_kernel == 4) ? 5 : used_tc ? (c.last_gram_kernel == 3 ? 4 : c.last_gram_kernel == 2 ? 3 : 2) : 1;
  launch_set_scalar(c, cnt, (double)n_local);
  }
  if (phase == 1) {
    launch_add_f64(c, sf->acc, packed, packed_count);
    sf->rows += n_local;
    return;
  }
  mark(2);

  // [2] combine ranks, centre.  PCA: ONE all-gather of [G | s | q | n | c_r], re-based to the common shift and summed in
  // rank order on every rank (bit-identical everywhere).  TruncatedSVD has no shift: one all-reduce.
  const double* moments = packed;
  c.last_exchange = 0;
  if (c.world > 1 && center) {
    double* merged = c.arena.take_n<double>(ext_count);
    const double* gathered = nullptr;
    if (c.p2p_status_host && *c.p2p_status_host != 0)
      fail("peer-memory exchange: rank %d never arrived in an earlier exchange (waited 4 s)", *c.p2p_status_host - 1);
    if (c.p2p_ready && c.exchange != 1 && (long long)ext_count <= c.p2p_cap) {
      // ONE kernel: push the partials into every peer's buffer over NVLink, release a flag, wait for everybody's (p2p.cu)
      gathered = launch_p2p_allgather(c, packed, (long long)ext_count);
      c.last_exchange = 2;
    } else {
      double* g = c.arena.take_n<double>((size_t)c.world * ext_count);
      allgather_f64(c, packed, g, ext_count);
      gathered = g;
      c.last_exchange = 1;
    }
    launch_rebase_sum(c, gathered, c.world, d, merged);
    moments = merged;
    cnt = merged + dd + 2 * (size_t)d;
    shift_d_arg = cnt + 1;
  } else {
    allreduce_f64(c, packed, packed_count);
    if (c.world > 1) c.last_exchange = 1;
  }
  double* C = c.arena.take_n<double>(dd);
  launch_center(c, moments, d, center, shift_d_arg, C, mean_d, total_var);
  mark(3);

  // [3] eigensolve + [4] outputs
  if (exact && !use_jacobi) {
    void* ws = c.arena.take(eigh_topk_workspace_bytes(d, k));
    double* W = c.arena.take_n<double>(k);
    double* Vt = c.arena.take_n<double>((size_t)k * d);
    launch_eigh_topk(c, C, d, k, W, Vt, ws);
    mark(4);
    launch_postproc<T>(c, W, k, Vt, d, /*vectors_in_rows=*/true, /*check_mass=*/false, d, k, cnt, total_var,
                       center ? mean_d : nullptr, dQ, dw, dev, devr, center ? dmean : nullptr, p.whiten);
  } else if (exact) {
    double* A = c.arena.take_n<double>((size_t)dp * dp);
    double* V = c.arena.take_n<double>((size_t)dp * dp);
    double* W = c.arena.take_n<double>(dp);
    void* ws = c.arena.take(jacobi_workspace_bytes(dp));
    launch_pad_copy(c, C, d, A, dp);
    const double tol = c.jacobi_tol > 0 ? c.jacobi_tol : 1e-13;
    launch_jacobi(c, A, V, dp, W, ws, tol, c.jacobi_sweeps);
    mark(4);
    launch_postproc<T>(c, W, dp, V, dp, /*vectors_in_rows=*/false, /*check_mass=*/dp != d, d, k, cnt, total_var,
                       center ? mean_d : nullptr, dQ, dw, dev, devr, center ? dmean : nullptr, p.whiten);
  } else {
    void* ws = c.arena.take(power_workspace_bytes(d, l));
    double* W = c.arena.take_n<double>(lp);
    double* Vt = c.arena.take_n<double>((size_t)lp * d);
    const int seed = p.random_state;
    const double tol = p.tol > 0 ? (double)p.tol : 0.0;
    launch_power(c, C, d, k, l, p.n_iter, tol, seed, W, Vt, ws);
    mark(4);
    launch_postproc<T>(c, W, lp, Vt, d, /*vectors_in_rows=*/true, /*check_mass=*/true, d, k, cnt, total_var,
                       center ? mean_d : nullptr, dQ, dw, dev, devr, center ? dmean : nullptr, p.whiten);
  }
  mark(5);
  c.stage_valid = timing;
}

// ---------------------------------------------------------------- streaming fit
template <typename T>
static void partial_fit_impl(CtxFull& c, const T* dX, long long n, int d, int center) {
  if (d <= 0 || n < 0) fail("partial_fit: bad shape n=%lld d=%d", n, d);
  StreamFit& sf = c.sf;
  if (!sf.acc) {
    const size_t count = (size_t)d * d + 2 * (size_t)d + 1;
    DR4G_CUDA(cudaMalloc(&sf.acc, count * 8));
    DR4G_CUDA(cudaMalloc(&sf.shift_d, (size_t)d * 8));
    DR4G_CUDA(cudaMalloc(&sf.shift_t, (size_t)d * 8));
    DR4G_CUDA(cudaMemsetAsync(sf.acc, 0, count * 8, c.stream));
    DR4G_CUDA(cudaMemsetAsync(sf.shift_d, 0, (size_t)d * 8, c.stream));
    DR4G_CUDA(cudaMemsetAsync(sf.shift_t, 0, (size_t)d * 8, c.stream));
    sf.d = d;
    sf.center = center;
    sf.elem = (int)sizeof(T);
    sf.rows = 0;
  } else if (sf.d != d || sf.center != center || sf.elem != (int)sizeof(T)) {
    fail("partial_fit: chunk (d=%d center=%d %d-byte elements) does not match the accumulated state (d=%d center=%d %d-byte); "
         "call dr4g_partial_fit_reset first", d, center, (int)sizeof(T), sf.d, sf.center, sf.elem);
  }
  if (n == 0) return;
  params p{};
  p.algorithm = const_cast<char*>("cusolver");
  fit_impl<T>(c, dX, n, d, 1, center, p, nullptr, nullptr, nullptr, nullptr, nullptr, &sf, 1);
}

template <typename T>
static void partial_fit_finalize_impl(CtxFull& c, int k, const params& p, T* dQ, T* dw, T* dev, T* devr, T* dmean) {
  StreamFit& sf = c.sf;
  if (!sf.acc || sf.elem != (int)sizeof(T)) fail("partial_fit_finalize: nothing accumulated for this element type");
  if (sf.rows <= 0 && c.world <= 1) fail("partial_fit_finalize: no rows accumulated");
  fit_impl<T>(c, nullptr, sf.rows, sf.d, k, sf.center, p, dQ, dw, dev, devr, dmean, &sf, 2);
}

// Copy-outs of staged D2H sub-chunks that have completed (all of them when `all`), oldest first.
static void stage_drain(HostStage& st, bool all) {
  for (int i = 0; i < HostStage::NB; ++i) {
    const int b = (st.next + i) % HostStage::NB;
    if (!st.busy[b]) continue;
    if (!all && i > 0) break;               // only the buffer that is about to be reused has to be free
    DR4G_CUDA(cudaEventSynchronize(st.done[b]));
    if (st.out[b].dst) parallel_memcpy(st.out[b].dst, st.buf[b], st.out[b].bytes);
    st.out[b] = HostStage::Out{};
    st.busy[b] = false;
  }
}

// host -> device; pageable sources go through the pinned ring
static void h2d_copy(CtxFull& c, void* dst, const void* src, size_t bytes, bool pageable, cudaStream_t s) {
  if (!pageable) {
    DR4G_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, s));
    return;
  }
  HostStage& st = c.stage;
  st.ensure();
  for (size_t off = 0; off < bytes; off += HostStage::kSub) {
    const size_t len = bytes - off < HostStage::kSub ? bytes - off : HostStage::kSub;
    const int b = st.next;
    stage_drain(st, false);
    parallel_memcpy(st.buf[b], static_cast<const char*>(src) + off, len);
    DR4G_CUDA(cudaMemcpyAsync(static_cast<char*>(dst) + off, st.buf[b], len, cudaMemcpyHostToDevice, s));
    DR4G_CUDA(cudaEventRecord(st.done[b], s));
    st.busy[b] = true;
    st.next = (b + 1) % HostStage::NB;
  }
}

// device -> host; pageable destinations go through the pinned ring (the copy-out of a sub-chunk happens when its buffer
// comes round again, or in stage_drain(all) at the end)
static void d2h_copy(CtxFull& c, void* dst, const void* src, size_t bytes, bool pageable, cudaStream_t s) {
  if (!pageable) {
    DR4G_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, s));
    return;
  }
  HostStage& st = c.stage;
  st.ensure();
  for (size_t off = 0; off < bytes; off += HostStage::kSub) {
    const size_t len = bytes - off < HostStage::kSub ? bytes - off : HostStage::kSub;
    const int b = st.next;
    stage_drain(st, false);
    DR4G_CUDA(cudaMemcpyAsync(st.buf[b], static_cast<const char*>(src) + off, len, cudaMemcpyDeviceToHost, s));
    DR4G_CUDA(cudaEventRecord(st.done[b], s));
    st.busy[b] = true;
    st.out[b] = HostStage::Out{static_cast<char*>(dst) + off, len};
    st.next = (b + 1) % HostStage::NB;
  }
}

// Two sweeps over a host-resident X in row chunks: [1] H2D (copy stream) overlapped with the chunk's Gram pass
// (partial fit), then the eigensolve; [2] H2D again, projection, D2H of the scores (third stream), all three
// overlapped through two chunk buffers.  Device footprint: 2 chunks of X and of the two score matrices.
template <typename T>
static void legacy_pipelined(CtxFull& c, bool is_pca, bool resident, const T* X, long long n, int d, int k, long long chunk,
                             const params& p, T* Q, T* w, T* U, T* Xt, T* ev, T* evr, T* mean) {
  // resident: the chunks land in one device copy of X (one H2D sweep; the pass on chunk i runs under the copy of
  // chunk i+1, the D2H of the scores of chunk i under the projection of chunk i+1); else two chunk buffers.
  if (!c.copy_stream) DR4G_CUDA(cudaStreamCreateWithFlags(&c.copy_stream, cudaStreamNonBlocking));
  if (!c.out_stream) DR4G_CUDA(cudaStreamCreateWithFlags(&c.out_stream, cudaStreamNonBlocking));
  static const bool no_stage = getenv("DR4G_NO_STAGING") != nullptr;
  // small arrays stay on the driver's own pageable path (the first use pins the 32 MB ring, once per process); from
  // a few MB on the ring wins: the driver stages a pageable copy on one thread, and fresh output arrays (np.zeros of
  // the wrapper) take their page faults there too (50,000 x 256, k = 64: 5 of 10.7 ms)
  // (an input of one sub-chunk cannot overlap its host copy with its DMA: 10,000 x 128 got slower, 1.6 -> 2.3 ms)
  const size_t stage_min_in = (size_t)12 << 20, stage_min_out = (size_t)4 << 20;
  const bool x_pageable = !no_stage && (size_t)n * d * sizeof(T) >= stage_min_in && is_pageable_host(X);
  const bool z_pageable = !no_stage && Xt && (size_t)n * k * sizeof(T) >= stage_min_out && is_pageable_host(Xt);
  const bool u_pageable = !no_stage && U && (size_t)n * k * sizeof(T) >= stage_min_out && is_pageable_host(U);
  const size_t cx = (size_t)(resident ? n : 2 * chunk) * d * sizeof(T), cz = (size_t)(resident ? n : 2 * chunk) * k * sizeof(T);
  c.io.reserve(up256(cx) + 2 * up256(cz) + up256((size_t)k * d * sizeof(T)) + 3 * up256((size_t)k * sizeof(T)) +
               up256((size_t)d * sizeof(T)) + 8192);
  c.io.reset();
  T* dQ = c.io.take_n<T>((size_t)k * d);
  T* dw = c.io.take_n<T>(k);
  T* dev = c.io.take_n<T>(k);
  T* devr = c.io.take_n<T>(k);
  T* dmean = c.io.take_n<T>(d);
  T* xall = c.io.take_n<T>((size_t)(resident ? n : 2 * chunk) * d);
  T* zall = c.io.take_n<T>((size_t)(resident ? n : 2 * chunk) * k);
  T* uall = c.io.take_n<T>((size_t)(resident ? n : 2 * chunk) * k);
  auto xbuf = [&](long long r0, int b) { return resident ? xall + (size_t)r0 * d : xall + (size_t)b * chunk * d; };
  auto zbuf = [&](long long r0, int b) { return resident ? zall + (size_t)r0 * k : zall + (size_t)b * chunk * k; };
  auto ubuf = [&](long long r0, int b) { return resident ? uall + (size_t)r0 * k : uall + (size_t)b * chunk * k; };
  cudaEvent_t loaded[2], xfree[2], projected[2], zfree[2];
  for (int b = 0; b < 2; ++b) {
    DR4G_CUDA(cudaEventCreateWithFlags(&loaded[b], cudaEventDisableTiming));
    DR4G_CUDA(cudaEventCreateWithFlags(&xfree[b], cudaEventDisableTiming));
    DR4G_CUDA(cudaEventCreateWithFlags(&projected[b], cudaEventDisableTiming));
    DR4G_CUDA(cudaEventCreateWithFlags(&zfree[b], cudaEventDisableTiming));
  }
  auto cleanup = [&]() {
    for (int b = 0; b < 2; ++b) {
      cudaEventDestroy(loaded[b]);
      cudaEventDestroy(xfree[b]);
      cudaEventDestroy(projected[b]);
      cudaEventDestroy(zfree[b]);
    }
  };
  LTRACE("setup");
  Prefault pf_z, pf_u;
  if (z_pageable) pf_z.start(Xt, (size_t)n * k * sizeof(T), 2);
  if (u_pageable) pf_u.start(U, (size_t)n * k * sizeof(T), 2);
  try {
    // the copy stream must not overtake earlier work of this context that still reads the io arena
    DR4G_CUDA(cudaStreamSynchronize(c.stream));
    // accumulators of the context are reused when the shape allows (cudaMalloc / cudaFree synchronise)
    if (c.sf.acc && c.sf.d == d && c.sf.elem == (int)sizeof(T)) {
      DR4G_CUDA(cudaMemsetAsync(c.sf.acc, 0, ((size_t)d * d + 2 * (size_t)d + 1) * 8, c.stream));
      c.sf.center = is_pca ? 1 : 0;
      c.sf.rows = 0;
    } else {
      c.sf.release();
    }
    const long long nchunks = (n + chunk - 1) / chunk;
    // [1] fit
    for (long long i = 0; i < nchunks; ++i) {
      const int b = (int)(i & 1);
      const long long r0 = i * chunk, rows = n - r0 < chunk ? n - r0 : chunk;
      if (!resident && i >= 2) DR4G_CUDA(cudaStreamWaitEvent(c.copy_stream, xfree[b], 0));
      h2d_copy(c, xbuf(r0, b), X + (size_t)r0 * d, (size_t)rows * d * sizeof(T), x_pageable, c.copy_stream);
      DR4G_CUDA(cudaEventRecord(loaded[b], c.copy_stream));
      DR4G_CUDA(cudaStreamWaitEvent(c.stream, loaded[b], 0));
      partial_fit_impl<T>(c, xbuf(r0, b), rows, d, is_pca ? 1 : 0);
      if (!resident) DR4G_CUDA(cudaEventRecord(xfree[b], c.stream));
    }
    LTRACE("pass_issued");
    partial_fit_finalize_impl<T>(c, k, p, dQ, dw, dev, devr, dmean);
    LTRACE("finalize_issued");
    // while the device is busy with the eigensolve: take the page faults of the (fresh) score arrays on all helper threads
    if (z_pageable && !pf_z.finish()) parallel_touch(Xt, (size_t)n * k * sizeof(T));
    if (u_pageable && !pf_u.finish()) parallel_touch(U, (size_t)n * k * sizeof(T));
    LTRACE("prefaulted");
    // [2] scores
    for (long long i = 0; Xt && i < nchunks; ++i) {
      const int b = (int)(i & 1);
      const long long r0 = i * chunk, rows = n - r0 < chunk ? n - r0 : chunk;
      if (!resident) {
        DR4G_CUDA(cudaStreamWaitEvent(c.copy_stream, xfree[b], 0));   // last reader: the pass or the projection of chunk i-2
        h2d_copy(c, xbuf(r0, b), X + (size_t)r0 * d, (size_t)rows * d * sizeof(T), x_pageable, c.copy_stream);
        DR4G_CUDA(cudaEventRecord(loaded[b], c.copy_stream));
        DR4G_CUDA(cudaStreamWaitEvent(c.stream, loaded[b], 0));
        if (i >= 2) DR4G_CUDA(cudaStreamWaitEvent(c.stream, zfree[b], 0));
      }
      launch_project<T>(c, xbuf(r0, b), rows, d, k, dQ, is_pca ? dmean : nullptr, dw, zbuf(r0, b), U ? ubuf(r0, b) : nullptr);
      DR4G_CUDA(cudaEventRecord(projected[b], c.stream));
      if (!resident) DR4G_CUDA(cudaEventRecord(xfree[b], c.stream));
      DR4G_CUDA(cudaStreamWaitEvent(c.out_stream, projected[b], 0));
      d2h_copy(c, Xt + (size_t)r0 * k, zbuf(r0, b), (size_t)rows * k * sizeof(T), z_pageable, c.out_stream);
      if (U) d2h_copy(c, U + (size_t)r0 * k, ubuf(r0, b), (size_t)rows * k * sizeof(T), u_pageable, c.out_stream);
      if (!resident) DR4G_CUDA(cudaEventRecord(zfree[b], c.out_stream));
    }
    LTRACE("scores_issued");
    auto out = [&](T* dst, const T* src, size_t count) {
      if (dst) DR4G_CUDA(cudaMemcpyAsync(dst, src, count * sizeof(T), cudaMemcpyDefault, c.stream));
    };
    out(Q, dQ, (size_t)k * d);
    out(w, dw, k);
    out(ev, dev, k);
    out(evr, devr, k);
    if (is_pca) out(mean, dmean, d);
    LTRACE("small_outputs");
    stage_drain(c.stage, true);   // copy-outs of the last staged score sub-chunks
    DR4G_CUDA(cudaStreamSynchronize(c.copy_stream));
    DR4G_CUDA(cudaStreamSynchronize(c.stream));
    DR4G_CUDA(cudaStreamSynchronize(c.out_stream));
    LTRACE("synced");
  } catch (...) {
    for (int b = 0; b < HostStage::NB; ++b) { c.stage.busy[b] = false; c.stage.out[b] = HostStage::Out{}; }
    cudaStreamSynchronize(c.copy_stream);
    cudaStreamSynchronize(c.stream);
    cudaStreamSynchronize(c.out_stream);
    c.sf.release();
    cleanup();
    throw;
  }
  cleanup();
}

// Host-resident X (pageable or pinned) -> fit + scores on the device -> host outputs.  Chunk pipeline on three streams
// (H2D | pass, projection | D2H).  One device copy of X filled chunk by chunk when X, its scores and the fit's
// workspace fit into the memory that is FREE now (other tenants of the device - a torch caching allocator in the same
// process - are respected); otherwise out of core: two chunk buffers and a second H2D sweep for the
// scores.  DR4G_STREAM_ROWS forces the out-of-core path with that chunk size.  With a communicator on the context the
// pass is the row-sharded one (pilot + packed allreduce), X being this rank's shard.
template <typename T>
static void host_fit_transform(CtxFull& c, bool is_pca, const T* X, long long n, int d, int k, const params& p, T* Q, T* w,
                               T* U, T* Xt, T* ev, T* evr, T* mean) {
  const size_t xbytes = (size_t)n * d * sizeof(T);
  const size_t zbytes = (size_t)n * k * sizeof(T);
  const char* env = getenv("DR4G_STREAM_ROWS");
  const long long forced = env ? atoll(env) : 0;
  // what the resident path would have to allocate on top of what the context already owns: the io arena (grown by
  // Arena::reserve's 12.5 % + 1 MB policy; the old block is released first) and the fit workspace (packed accumulators,
  // covariance, eigensolver scratch ~ 6 d^2 doubles, Gram slabs ~ 40 MB)
  const size_t io_need = up256(xbytes) + 2 * up256(zbytes) + up256((size_t)k * d * sizeof(T)) + ((size_t)1 << 20);
  const size_t io_grow = io_need > c.io.cap ? io_need + (io_need >> 3) + ((size_t)1 << 20) - c.io.cap : 0;
  const size_t ws_need = (size_t)d * d * 8 * 6 + ((size_t)48 << 20);
  const size_t ws_grow = ws_need > c.arena.cap ? ws_need - c.arena.cap : 0;
  // cudaMemGetInfo costs ~0.1 ms (a fifth of a 4 x 3 call): asked only when the call has to grow the arenas by more than
  // 256 MB; a smaller growth that does not fit is caught below (cudaMalloc failure -> out-of-core pipeline)
  size_t free_b = ~(size_t)0, total_b = 0;
  if (forced <= 0 && io_grow + ws_grow > ((size_t)256 << 20)) DR4G_CUDA(cudaMemGetInfo(&free_b, &total_b));
  LTRACE("meminfo");
  bool resident = forced <= 0 && (double)io_grow + (double)ws_grow <= 0.92 * (double)free_b;
  auto chunk_rows = [&](bool res) {
    long long chunk = forced > 0 ? forced : (long long)(((size_t)(res ? 128 : 256) << 20) / ((size_t)d * sizeof(T)));
    if (chunk < 1024 && forced <= 0) chunk = 1024;
    return chunk > n ? n : chunk;
  };
  if (resident) {
    try {
      legacy_pipelined<T>(c, is_pca, true, X, n, d, k, chunk_rows(true), p, Q, w, U, Xt, ev, evr, mean);
      return;
    } catch (const Error& e) {
      // lost a race for device memory: fall back to the out-of-core pipeline instead of failing
      if (std::strstr(e.what(), "cudaMalloc") == nullptr) throw;
      cudaGetLastError();
      resident = false;
    }
  }
  legacy_pipelined<T>(c, is_pca, false, X, n, d, k, chunk_rows(false), p, Q, w, U, Xt, ev, evr, mean);
}

// ---------------------------------------------------------------- legacy host-pointer entry points
static bool is_device_pointer(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

template <typename T>
static void legacy_entry(bool is_pca, const T* X, T* Q, T* w, T* U, T* Xt, T* ev, T* evr, T* mean, params p) {
  const char* name = is_pca ? (sizeof(T) == 4 ? "pca_float" : "pca_double")
                            : (sizeof(T) == 4 ? "truncated_svd_float" : "truncated_svd_double");
  g_err[0] = 0;
  try {
    // same silent-return guards as the CPU backend (src/cpu_backend.cpp:343,364) plus a message
    if (!X || !Q || !w || !U || !Xt || (is_pca && !mean)) fail("%s: null argument", name);
    const long long n = p.X_n;
    const int d = p.X_m;
    int k = p.k;
    if (k > n) k = (int)n;
    if (k > d) k = d;
    if (n <= 0 || d <= 0 || k <= 0) fail("%s: empty problem (X_n=%lld X_m=%d k=%d)", name, n, d, p.k);

    LegacyTrace ltrace;
    g_ltrace = ltrace.on ? &ltrace : nullptr;
    struct Unset { ~Unset() { g_ltrace = nullptr; } } unset;
    CtxFull* c = legacy_ctx(p.gpu_id);
    std::lock_guard<std::mutex> lk(c->mu);
    DeviceGuard guard(c->device);
    LTRACE("ctx");
    if (p.verbose)
      fprintf(stderr, "[dimreduce4gpu] %s: algorithm=%s n=%lld d=%d k=%d device=%d\n", name,
              is_exact(p) ? "cusolver (blocked Jacobi)" : "power (subspace iteration)", n, d, k, c->device);

    const size_t zbytes = (size_t)n * k * sizeof(T);
    if (!is_device_pointer(X)) {
      host_fit_transform<T>(*c, is_pca, X, n, d, k, p, Q, w, U, Xt, ev, evr, mean);
      return;
    }
    // X already on the device (a caller handing device pointers to the legacy symbols): fit + projection in place
    c->io.reserve(2 * up256(zbytes) + up256((size_t)k * d * sizeof(T)) + 3 * up256((size_t)k * sizeof(T)) +
                  up256((size_t)d * sizeof(T)) + 4096);
    c->io.reset();
    T* dQ = c->io.take_n<T>((size_t)k * d);
    T* dw = c->io.take_n<T>(k);
    T* dev = c->io.take_n<T>(k);
    T* devr = c->io.take_n<T>(k);
    T* dmean = c->io.take_n<T>(d);
    T* dZ = c->io.take_n<T>((size_t)n * k);
    T* dU = c->io.take_n<T>((size_t)n * k);
    const T* dX = X;
    fit_impl<T>(*c, dX, n, d, k, is_pca ? 1 : 0, p, dQ, dw, dev, devr, dmean);
    launch_project<T>(*c, dX, n, d, k, dQ, is_pca ? dmean : nullptr, dw, dZ, dU);

    auto out = [&](T* dst, const T* src, size_t count) {
      if (dst) DR4G_CUDA(cudaMemcpyAsync(dst, src, count * sizeof(T), cudaMemcpyDefault, c->stream));
    };
    out(Q, dQ, (size_t)k * d);
    out(w, dw, k);
    out(ev, dev, k);
    out(evr, devr, k);
    if (is_pca) out(mean, dmean, d);
    out(Xt, dZ, (size_t)n * k);
    out(U, dU, (size_t)n * k);
    DR4G_CUDA(cudaStreamSynchronize(c->stream));
  } catch (const std::exception& e) {
    // reference: print and return with outputs untouched (src/tsvd.cu:473-484); never exit()
    set_last_error("%s", e.what());
    fprintf(stderr, "[dimreduce4gpu] %s failed: %s\n", name, e.what());
    cudaGetLastError();
  }
}

}  // namespace dr4g

// ================================================================== extern "C"
using namespace dr4g;

struct dr4g_ctx : CtxFull {};

#define DR4G_API_BEGIN \
  g_err[0] = 0;        \
  try {
#define DR4G_API_END                      \
  }                                       \
  catch (const std::exception& e) {       \
    set_last_error("%s", e.what());       \
    cudaGetLastError();                   \
    return 1;                             \
  }                                       \
  return 0;

extern "C" {

void truncated_svd_float(const float* X, float* Q, float* w, float* U, float* X_transformed,
                         float* explained_variance, float* explained_variance_ratio, params p) {
  legacy_entry<float>(false, X, Q, w, U, X_transformed, explained_variance, explained_variance_ratio, nullptr, p);
}
void truncated_svd_double(const double* X, double* Q, double* w, double* U, double* X_transformed,
                          double* explained_variance, double* explained_variance_ratio, params p) {
  legacy_entry<double>(false, X, Q, w, U, X_transformed, explained_variance, explained_variance_ratio, nullptr, p);
}
void pca_float(const float* X, float* Q, float* w, float* U, float* X_transformed, float* explained_variance,
               float* explained_variance_ratio, float* mean, params p) {
  legacy_entry<float>(true, X, Q, w, U, X_transformed, explained_variance, explained_variance_ratio, mean, p);
}
void pca_double(const double* X, double* Q, double* w, double* U, double* X_transformed, double* explained_variance,
                double* explained_variance_ratio, double* mean, params p) {
  legacy_entry<double>(true, X, Q, w, U, X_transformed, explained_variance, explained_variance_ratio, mean, p);
}

const char* dr4g_last_error(void) { return last_error(); }
int dr4g_abi_version(void) { return 1; }

dr4g_ctx* dr4g_ctx_create(int device) {
  g_err[0] = 0;
  try {
    return static_cast<dr4g_ctx*>(ctx_new(device));
  } catch (const std::exception& e) {
    set_last_error("%s", e.what());
    cudaGetLastError();
    return nullptr;
  }
}

void dr4g_ctx_destroy(dr4g_ctx* ctx) { ctx_delete(ctx); }

int dr4g_ctx_set_stream(dr4g_ctx* ctx, void* cuda_stream) {
  DR4G_API_BEGIN
  if (!ctx) fail("null context");
  std::lock_guard<std::mutex> lk(ctx->mu);
  const cudaStream_t next = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  if (next != ctx->stream) {
    // the arena and the accumulators are reused by whatever runs next: order the new stream behind the work that is
    // still queued on the old one
    DeviceGuard g(ctx->device);
    if (!ctx->switch_ev) DR4G_CUDA(cudaEventCreateWithFlags(&ctx->switch_ev, cudaEventDisableTiming));
    DR4G_CUDA(cudaEventRecord(ctx->switch_ev, ctx->stream));
    DR4G_CUDA(cudaStreamWaitEvent(next, ctx->switch_ev, 0));
    ctx->stream = next;
  }
  DR4G_API_END
}

int dr4g_sync(dr4g_ctx* ctx) {
  DR4G_API_BEGIN
  if (!ctx) fail("null context");
  DeviceGuard g(ctx->device);
  DR4G_CUDA(cudaStreamSynchronize(ctx->stream));
  if (ctx->p2p_status_host && *ctx->p2p_status_host != 0)
    fail("peer-memory exchange: rank %d never arrived (waited 4 s); results of the last sharded fit are invalid", *ctx->p2p_status_host - 1);
  DR4G_API_END
}

int dr4g_set_option(dr4g_ctx* ctx, const char* key, double value) {
  DR4G_API_BEGIN
  if (!ctx || !key) fail("null argument");
  const std::string k(key);
  if (k == "gram_path") ctx->gram_path = (int)value;
  else if (k == "gram_split") ctx->gram_split = (int)value;
  else if (k == "proj_path") ctx->proj_path = (int)value;
  else if (k == "power_min_dim") ctx->power_min_dim = (int)value;
  else if (k == "flush_rows") ctx->flush_rows = (int)value;
  else if (k == "pilot_rows") ctx->pilot_rows = (int)value;
  else if (k == "jacobi_tol") ctx->jacobi_tol = value;
  else if (k == "jacobi_sweeps") ctx->jacobi_sweeps = (int)value;
  else if (k == "timing") ctx->timing = (int)value;
  else if (k == "eig_solver") ctx->eig_solver = (int)value;
  else if (k == "fold") ctx->fold = (int)value;
  else if (k == "exchange") ctx->exchange = (int)value;
  else fail("unknown option '%s'", key);
  DR4G_API_END
}

int dr4g_nccl_unique_id(void* out128) {
  DR4G_API_BEGIN
  if (!out128) fail("null argument");
  NcclUniqueId id;
  NcclApi* api = nccl_api();
  int rc = api->GetUniqueId(&id);
  if (rc != 0) fail("ncclGetUniqueId failed: %s", api->GetErrorString(rc));
  std::memcpy(out128, &id, sizeof(id));
  DR4G_API_END
}

int dr4g_comm_init(dr4g_ctx* ctx, const void* unique_id128, int rank, int world) {
  DR4G_API_BEGIN
  if (!ctx || !unique_id128) fail("null argument");
  if (world < 1 || rank < 0 || rank >= world) fail("bad rank/world %d/%d", rank, world);
  DeviceGuard g(ctx->device);
  NcclApi* api = nccl_api();
  if (ctx->comm) { api->CommDestroy(ctx->comm); ctx->comm = nullptr; }
  ctx->rank = 0;
  ctx->world = 1;
  if (world > 1) {
    NcclUniqueId id;
    std::memcpy(&id, unique_id128, sizeof(id));
    int rc = api->CommInitRank(&ctx->comm, world, id, rank);
    if (rc != 0) fail("ncclCommInitRank failed: %s", api->GetErrorString(rc));
    ctx->rank = rank;
    ctx->world = world;
  }
  DR4G_API_END
}

int dr4g_comm_world(dr4g_ctx* ctx) { return ctx ? ctx->world : 0; }

int dr4g_p2p_prepare(dr4g_ctx* ctx, long long capacity_doubles, void* out_handle64) {
  DR4G_API_BEGIN
  if (!ctx || !out_handle64) fail("null argument");
  if (ctx->world < 2 || ctx->world > kP2PMaxWorld) fail("dr4g_p2p_prepare: needs a communicator of 2..%d ranks (dr4g_comm_init first)", kP2PMaxWorld);
  if (capacity_doubles < 1) fail("dr4g_p2p_prepare: bad capacity");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  std::lock_guard<std::mutex> lk(ctx->mu);
  DeviceGuard g(ctx->device);
  if (ctx->p2p_peer[ctx->rank]) fail("dr4g_p2p_prepare: already prepared");
  const size_t bytes = p2p_buffer_doubles(ctx->world, capacity_doubles) * sizeof(double);
  void* buf = nullptr;
  DR4G_CUDA(cudaMalloc(&buf, bytes));
  DR4G_CUDA(cudaMemset(buf, 0, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, buf);
  if (e != cudaSuccess) { cudaFree(buf); cudaGetLastError(); fail("cudaIpcGetMemHandle: %s", cudaGetErrorString(e)); }
  if (!ctx->p2p_status_host) {
    int* hs = nullptr;
    DR4G_CUDA(cudaHostAlloc(&hs, 64, cudaHostAllocMapped));
    *hs = 0;
    int* ds = nullptr;
    DR4G_CUDA(cudaHostGetDevicePointer(&ds, hs, 0));
    ctx->p2p_status_host = hs;
    ctx->p2p_status = ds;
  }
  ctx->p2p_peer[ctx->rank] = static_cast<double*>(buf);
  ctx->p2p_cap = capacity_doubles;
  std::memcpy(out_handle64, &h, sizeof(h));
  DR4G_API_END
}

int dr4g_p2p_connect(dr4g_ctx* ctx, const void* handles) {
  DR4G_API_BEGIN
  if (!ctx || !handles) fail("null argument");
  std::lock_guard<std::mutex> lk(ctx->mu);
  DeviceGuard g(ctx->device);
  if (!ctx->p2p_peer[ctx->rank]) fail("dr4g_p2p_connect: dr4g_p2p_prepare first");
  for (int r = 0; r < ctx->world; ++r) {
    if (r == ctx->rank || ctx->p2p_peer[r]) continue;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, static_cast<const char*>(handles) + (size_t)r * sizeof(h), sizeof(h));
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { cudaGetLastError(); fail("cudaIpcOpenMemHandle (rank %d): %s", r, cudaGetErrorString(e)); }
    ctx->p2p_peer[r] = static_cast<double*>(p);
  }
  ctx->p2p_ready = true;
  DR4G_API_END
}

int dr4g_last_exchange_path(dr4g_ctx* ctx) { return ctx ? ctx->last_exchange : 0; }

#define DR4G_FIT(NAME, T)                                                                                   \
  int NAME(dr4g_ctx* ctx, const T* dX, long long n_local, int d, int k, int center, params p, T* dQ, T* dw,  \
           T* dev, T* devr, T* dmean) {                                                                     \
    DR4G_API_BEGIN                                                                                          \
    if (!ctx || !dX || !dQ || !dw) fail(#NAME ": null argument");                                           \
    std::lock_guard<std::mutex> lk(ctx->mu);                                                                \
    DeviceGuard g(ctx->device);                                                                             \
    fit_impl<T>(*ctx, dX, n_local, d, k, center, p, dQ, dw, dev, devr, dmean);                              \
    DR4G_API_END                                                                                            \
  }
DR4G_FIT(dr4g_fit_float, float)
DR4G_FIT(dr4g_fit_double, double)

// ---- host-resident shard in, host results out, on THIS context (communicator honoured) -------------
#define DR4G_HOST_FT(NAME, T)                                                                               \
  int NAME(dr4g_ctx* ctx, const T* X, long long n_local, int d, int k, int center, params p, T* Q, T* w, T* U, \
           T* Xt, T* ev, T* evr, T* mean) {                                                                \
    DR4G_API_BEGIN                                                                                          \
    if (!ctx || !X || !Q || !w || (center && !mean)) fail(#NAME ": null argument");                         \
    if (n_local <= 0 || d <= 0 || k <= 0) fail(#NAME ": empty problem");                                    \
    if (k > d) k = d;                                                                                       \
    if (ctx->world <= 1 && k > n_local) k = (int)n_local;                                                   \
    std::lock_guard<std::mutex> lk(ctx->mu);                                                                \
    DeviceGuard g(ctx->device);                                                                             \
    cudaStream_t keep = ctx->stream;                                                                        \
    ctx->stream = ctx->own_stream; /* blocking call with its own copy streams: never the caller's stream */  \
    try {                                                                                                   \
      host_fit_transform<T>(*ctx, center != 0, X, n_local, d, k, p, Q, w, U, Xt, ev, evr, mean);            \
    } catch (...) {                                                                                         \
      ctx->stream = keep;                                                                                   \
      throw;                                                                                                \
    }                                                                                                       \
    ctx->stream = keep;                                                                                     \
    DR4G_API_END                                                                                            \
  }
DR4G_HOST_FT(dr4g_fit_transform_host_float, float)
DR4G_HOST_FT(dr4g_fit_transform_host_double, double)

// ---- streaming fit ----------------------------------------------------------------
int dr4g_partial_fit_reset(dr4g_ctx* ctx) {
  DR4G_API_BEGIN
  if (!ctx) fail("null context");
  std::lock_guard<std::mutex> lk(ctx->mu);
  DeviceGuard g(ctx->device);
  DR4G_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->sf.release();
  DR4G_API_END
}

long long dr4g_partial_fit_rows(dr4g_ctx* ctx) { return ctx ? ctx->sf.rows : -1; }

#define DR4G_PARTIAL(NAME, FNAME, T)                                                                        \
  int NAME(dr4g_ctx* ctx, const T* dX, long long n, int d, int center) {                                    \
    DR4G_API_BEGIN                                                                                          \
    if (!ctx || (!dX && n > 0)) fail(#NAME ": null argument");                                              \
    std::lock_guard<std::mutex> lk(ctx->mu);                                                                \
    DeviceGuard g(ctx->device);                                                                             \
    partial_fit_impl<T>(*ctx, dX, n, d, center);                                                            \
    DR4G_API_END                                                                                            \
  }                                                                                                         \
  int FNAME(dr4g_ctx* ctx, int k, params p, T* dQ, T* dw, T* dev, T* devr, T* dmean) {                      \
    DR4G_API_BEGIN                                                                                          \
    if (!ctx || !dQ || !dw) fail(#FNAME ": null argument");                                                 \
    std::lock_guard<std::mutex> lk(ctx->mu);                                                                \
    DeviceGuard g(ctx->device);                                                                             \
    partial_fit_finalize_impl<T>(*ctx, k, p, dQ, dw, dev, devr, dmean);                                     \
    DR4G_API_END                                                                                            \
  }
DR4G_PARTIAL(dr4g_partial_fit_float, dr4g_partial_fit_finalize_float, float)
DR4G_PARTIAL(dr4g_partial_fit_double, dr4g_partial_fit_finalize_double, double)

#define DR4G_TRANSFORM(NAME, T)                                                                            \
  int NAME(dr4g_ctx* ctx, const T* dX, long long n, int d, int k, const T* dQ, const T* dmean, const T* dw, \
           T* dZ, T* dU) {                                                                                 \
    DR4G_API_BEGIN                                                                                         \
    if (!ctx || !dX || !dQ || !dZ) fail(#NAME ": null argument");                                          \
    if (d <= 0 || k <= 0 || n < 0) fail(#NAME ": bad shape");                                              \
    std::lock_guard<std::mutex> lk(ctx->mu);                                                               \
    DeviceGuard g(ctx->device);                                                                            \
    launch_project<T>(*ctx, dX, n, d, k, dQ, dmean, dw, dZ, dU);                                           \
    DR4G_API_END                                                                                           \
  }
DR4G_TRANSFORM(dr4g_transform_float, float)
DR4G_TRANSFORM(dr4g_transform_double, double)

#define DR4G_INVERSE(NAME, T)                                                                                \
  int NAME(dr4g_ctx* ctx, const T* dZ, long long n, int d, int k, const T* dQ, const T* dmean, T* dX) {      \
    DR4G_API_BEGIN                                                                                           \
    if (!ctx || !dZ || !dQ || !dX) fail(#NAME ": null argument");                                            \
    if (d <= 0 || k <= 0 || n < 0) fail(#NAME ": bad shape");                                                \
    std::lock_guard<std::mutex> lk(ctx->mu);                                                                 \
    DeviceGuard g(ctx->device);                                                                              \
    launch_inverse<T>(*ctx, dZ, n, d, k, dQ, dmean, dX);                                                     \
    DR4G_API_END                                                                                             \
  }
DR4G_INVERSE(dr4g_inverse_transform_float, float)
DR4G_INVERSE(dr4g_inverse_transform_double, double)

#define DR4G_CSR(NAME, T)                                                                                    \
  int NAME(dr4g_ctx* ctx, const long long* dindptr, const int* dindices, const T* ddata, long long row0,     \
           long long rows, int d, T* dout) {                                                                 \
    DR4G_API_BEGIN                                                                                           \
    if (!ctx || !dindptr || !dout || (rows > 0 && (!dindices || !ddata))) fail(#NAME ": null argument");      \
    if (d <= 0 || rows < 0 || row0 < 0) fail(#NAME ": bad shape");                                           \
    std::lock_guard<std::mutex> lk(ctx->mu);                                                                 \
    DeviceGuard g(ctx->device);                                                                              \
    launch_csr_to_dense<T>(*ctx, dindptr, dindices, ddata, row0, rows, d, dout);                             \
    DR4G_API_END                                                                                             \
  }
DR4G_CSR(dr4g_csr_to_dense_float, float)
DR4G_CSR(dr4g_csr_to_dense_double, double)

#define DR4G_COLSTATS(NAME, T)                                                                         \
  int NAME(dr4g_ctx* ctx, const T* dX, long long n, int d, const T* dshift, double* dsum, double* dsumsq) { \
    DR4G_API_BEGIN                                                                                     \
    if (!ctx || !dX || !dsum || !dsumsq) fail(#NAME ": null argument");                                \
    std::lock_guard<std::mutex> lk(ctx->mu);                                                           \
    DeviceGuard g(ctx->device);                                                                        \
    DR4G_CUDA(cudaMemsetAsync(dsum, 0, (size_t)d * 8, ctx->stream));                                   \
    DR4G_CUDA(cudaMemsetAsync(dsumsq, 0, (size_t)d * 8, ctx->stream));                                 \
    launch_colstats<T>(*ctx, dX, n, d, dshift, dsum, dsumsq);                                          \
    DR4G_API_END                                                                                       \
  }
DR4G_COLSTATS(dr4g_colstats_float, float)
DR4G_COLSTATS(dr4g_colstats_double, double)

int dr4g_gram_float(dr4g_ctx* ctx, const float* dX, long long n, int d, const float* dshift, double* dG) {
  DR4G_API_BEGIN
  if (!ctx || !dX || !dG) fail("dr4g_gram_float: null argument");
  std::lock_guard<std::mutex> lk(ctx->mu);
  DeviceGuard g(ctx->device);
  DR4G_CUDA(cudaMemsetAsync(dG, 0, (size_t)d * d * 8, ctx->stream));
  if (ctx->gram_path != 1 && gram_tc_eligible(*ctx, dX, n, d)) {
    const size_t ws = gram_tc_workspace_bytes(*ctx, d);
    ctx->arena.reserve(ws + 4096);
    ctx->arena.reset();
    launch_gram_tc(*ctx, dX, n, d, dshift, dG, nullptr, nullptr, ctx->arena.take(ws));
    ctx->last_gram_path = ctx->last_gram_kernel == 3 ? 4 : ctx->last_gram_kernel == 2 ? 3 : 2;
  } else {
    if (ctx->gram_path >= 2) fail("gram_path=2 (tcgen05) requested but the shape is not eligible: n=%lld d=%d", n, d);
    launch_gram_generic<float>(*ctx, dX, n, d, dshift, dG);
    launch_mirror_upper(*ctx, dG, d);
    ctx->last_gram_path = 1;
  }
  DR4G_API_END
}

int dr4g_gram_double(dr4g_ctx* ctx, const double* dX, long long n, int d, const double* dshift, double* dG) {
  DR4G_API_BEGIN
  if (!ctx || !dX || !dG) fail("dr4g_gram_double: null argument");
  std::lock_guard<std::mutex> lk(ctx->mu);
  DeviceGuard g(ctx->device);
  DR4G_CUDA(cudaMemsetAsync(dG, 0, (size_t)d * d * 8, ctx->stream));
  if (ctx->gram_path != 5 && gram_f64_eligible(*ctx, dX, n, d)) {
    const size_t ws = gram_f64_workspace_bytes(*ctx, n, d);
    ctx->arena.reserve(ws + 4096);
    ctx->arena.reset();
    launch_gram_f64(*ctx, dX, n, d, dshift, dG, ctx->arena.take(ws));
    ctx->last_gram_path = 5;
  } else {
    launch_gram_generic<double>(*ctx, dX, n, d, dshift, dG);
    ctx->last_gram_path = 1;
  }
  launch_mirror_upper(*ctx, dG, d);
  DR4G_API_END
}

int dr4g_eigh_double(dr4g_ctx* ctx, double* dA, int d, double* dW, double* dV) {
  DR4G_API_BEGIN
  if (!ctx || !dA || !dW || !dV) fail("dr4g_eigh_double: null argument");
  if (d <= 0) fail("dr4g_eigh_double: bad dimension");
  std::lock_guard<std::mutex> lk(ctx->mu);
  DeviceGuard g(ctx->device);
  const int dp = jacobi_padded_dim(d);
  const size_t need = 2 * up256((size_t)dp * dp * 8) + up256((size_t)dp * 8) + up256(jacobi_workspace_bytes(dp)) + 4096;
  ctx->arena.reserve(need);
  ctx->arena.reset();
  double* A = ctx->arena.take_n<double>((size_t)dp * dp);
  double* V = ctx->arena.take_n<double>((size_t)dp * dp);
  double* W = ctx->arena.take_n<double>(dp);
  void* ws = ctx->arena.take(jacobi_workspace_bytes(dp));
  launch_pad_copy(*ctx, dA, d, A, dp);
  launch_jacobi(*ctx, A, V, dp, W, ws, ctx->jacobi_tol > 0 ? ctx->jacobi_tol : 1e-13, ctx->jacobi_sweeps);
  // un-pad: eigenpairs that live in the padding have zero mass on the first d coordinates and a
  // zero eigenvalue; the caller gets the d x d leading block and the first d eigenvalues of the
  // padded problem, which by construction (block-diagonal padding, rotations never mix the two
  // blocks) are exactly the eigenpairs of A.
  DR4G_CUDA(cudaMemcpy2DAsync(dV, (size_t)d * 8, V, (size_t)dp * 8, (size_t)d * 8, d, cudaMemcpyDeviceToDevice, ctx->stream));
  DR4G_CUDA(cudaMemcpyAsync(dW, W, (size_t)d * 8, cudaMemcpyDeviceToDevice, ctx->stream));
  DR4G_API_END
}

int dr4g_eigh_topk_double(dr4g_ctx* ctx, double* dA, int d, int k, double* dW, double* dVt) {
  DR4G_API_BEGIN
  if (!ctx || !dA || !dW || !dVt) fail("dr4g_eigh_topk_double: null argument");
  if (d <= 0 || k <= 0 || k > d) fail("dr4g_eigh_topk_double: bad dimensions d=%d k=%d", d, k);
  std::lock_guard<std::mutex> lk(ctx->mu);
  DeviceGuard g(ctx->device);
  const size_t ws = eigh_topk_workspace_bytes(d, k);
  ctx->arena.reserve(ws + 4096);
  ctx->arena.reset();
  launch_eigh_topk(*ctx, dA, d, k, dW, dVt, ctx->arena.take(ws));
  DR4G_API_END
}

int dr4g_eigh_topk_static_double(dr4g_ctx* ctx, double* dA, int d, int k, double* dW, double* dVt) {
  DR4G_API_BEGIN
  if (!ctx || !dA || !dW || !dVt) fail("dr4g_eigh_topk_static_double: null argument");
  std::lock_guard<std::mutex> lk(ctx->mu);
  DeviceGuard g(ctx->device);
  launch_eigh_topk_static(*ctx, dA, d, k, dW, dVt);   // no arena, no cudaMalloc / cudaFree: workspace is a static device array
  DR4G_API_END
}

int dr4g_eigh_topk_static_max_d(void) { return eigh_topk_static_max_d(); }

int dr4g_stage_times(dr4g_ctx* ctx, double* ms, int n) {
  if (!ctx || !ms || !ctx->stage_valid) return 0;
  DeviceGuard g(ctx->device);
  int m = n < kNumStages ? n : kNumStages;
  for (int i = 0; i < m; ++i) {
    float t = 0.f;
    cudaError_t e = i < kNumStages - 1 ? cudaEventElapsedTime(&t, ctx->stage_ev[i], ctx->stage_ev[i + 1])
                                       : cudaEventElapsedTime(&t, ctx->stage_ev[0], ctx->stage_ev[kNumStages - 1]);
    if (e != cudaSuccess) { cudaGetLastError(); return 0; }
    ms[i] = t;
  }
  return m;
}

int dr4g_last_gram_path(dr4g_ctx* ctx) { return ctx ? ctx->last_gram_path : 0; }
int dr4g_last_proj_path(dr4g_ctx* ctx) { return ctx ? ctx->last_proj_path : 0; }
int dr4g_last_fold(dr4g_ctx* ctx) { return ctx ? ctx->last_fold : 0; }
long long dr4g_launch_count(dr4g_ctx* ctx) { return ctx ? ctx->launches : 0; }

}  // extern "C"
