// Shared host-side plumbing for libdimreduce4gpu.so: error handling, the persistent device
// context (replaces the reference's per-call DeviceContext, src/device_context.cuh:8-30, and
// per-object cudaMalloc Matrix<T>, src/matrix.cuh:16-303) and kernel launcher prototypes.
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace dr4g {

// ---------------------------------------------------------------- errors
void set_last_error(const char* fmt, ...);
const char* last_error();

struct Error : std::runtime_error {
  using std::runtime_error::runtime_error;
};

[[noreturn]] void fail(const char* fmt, ...);

#define DR4G_CUDA(expr)                                                                      \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess)                                                                   \
      ::dr4g::fail("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

// ---------------------------------------------------------------- NCCL (dlopen'ed, see nccl_dyn.cu)
struct NcclApi;
NcclApi* nccl_api();  // throws if libnccl.so.2 cannot be loaded

// ---------------------------------------------------------------- context
constexpr int kNumStages = 6;

struct Arena {
  char* base = nullptr;
  size_t cap = 0;
  size_t off = 0;
  void reset() { off = 0; }
  // grow-only; growing synchronises the device (only happens while shapes are still warming up)
  void reserve(size_t bytes);
  void* take(size_t bytes, size_t align = 256);
  template <typename T>
  T* take_n(size_t n) { return static_cast<T*>(take(n * sizeof(T))); }
  void release();
};

constexpr int kP2PMaxWorld = 16;   // ranks of one NVSwitch box the peer-memory exchange (p2p.cu) can address

struct Ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;
  Arena arena;        // fit/transform workspace
  // pinned staging for the legacy host-pointer symbols
  void* pinned[2] = {nullptr, nullptr};
  size_t pinned_cap = 0;
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t copy_ev[2] = {nullptr, nullptr};
  cudaEvent_t comp_ev[2] = {nullptr, nullptr};
  // stage timing
  cudaEvent_t stage_ev[kNumStages + 1] = {};
  bool stage_valid = false;
  // NCCL
  void* comm = nullptr;
  int rank = 0, world = 1;
  // peer-memory exchange (p2p.cu): every rank's exchange buffer mapped here (own buffer at [rank])
  bool p2p_ready = false;
  long long p2p_cap = 0;                 // doubles per slot
  double* p2p_peer[kP2PMaxWorld] = {};
  unsigned long long p2p_epoch = 0;
  int* p2p_status = nullptr;             // device view of a mapped host word: != 0 after a wait that gave up
  volatile int* p2p_status_host = nullptr;
  int exchange = 0;                      // 0 auto (peer memory when set up and the payload fits), 1 NCCL only
  int last_exchange = 0;                 // 0 none, 1 NCCL, 2 peer memory
  // options
  int gram_path = 0;          // 0 auto, 1 generic, 2 tcgen05 (pair kernel where eligible), 3 tcgen05 single-CTA kernel only,
                              // 5 first-generation 64 x 64 DMMA kernel for fp64 input too (cross-check of gram_f64.cu)
  int gram_split = 0;         // pair kernel operand split: 0 auto (bf16x3), 1 3xTF32 (kind::tf32), 2 bf16x3 (kind::f16)
  int flush_rows = 1024;      // rows between TMEM flushes in the tcgen05 Gram
  int pilot_rows = 4096;      // rows sampled for the pilot shift
  int power_min_dim = 256;    // min(n, d) <= this: exact path whatever `algorithm` says (src/cpu_backend.cpp:347)
  double jacobi_tol = 0.0;    // 0 -> per-dtype default
  int jacobi_sweeps = 30;
  int eig_solver = 0;         // exact path: 0 tridiagonal route (top-k), 1 blocked Jacobi (all pairs), 2 tridiagonal route without the cluster-resident T1
  int timing = 1;             // record per-stage events
  int fold = 1;               // narrow fp32 inputs (d <= 128): stream f rows as one row of f d <= 256 columns (api.cu)
  int last_fold = 0;          // fold factor of the last pass (0: not folded)
  // telemetry
  int last_gram_path = 0;
  int last_gram_kernel = 0;   // tcgen05 path: 1 single-CTA kernel (gram_tc.cu), 2 CTA-pair kernel 3xTF32, 3 CTA-pair kernel bf16x3 (gram_tc2.cu)
  long long launches = 0;
  // K7 on tcgen05: pre-split components image (grow-only, private to the projection)
  void* proj_buf = nullptr;
  size_t proj_cap = 0;
  void* pad_buf = nullptr;    // K7: zero-padded chunk copy of X + padded components / mean (shapes without a 16-byte row pitch)
  size_t pad_cap = 0;
  void* fold_buf = nullptr;   // K7 on row-folded narrow inputs: block-diagonal components, tiled mean / sigma (fixed 160 KB)
  int proj_path = 0;          // 0 auto (tensor cores where eligible, tiny fp32 inputs on CUDA cores), 1 CUDA-core kernels only, 2 tensor cores wherever eligible
  int last_proj_path = 0;     // 1 CUDA cores, 2 tcgen05, 3 fp64 DMMA
  // tensor-map encoder (driver entry point fetched lazily)
  void* encode_tiled = nullptr;
};

// ---------------------------------------------------------------- kernel launchers (one .cu each)
// K1: column sums / sums of squares of (x - shift), double outputs (accumulated into dsum/dsumsq,
// which the caller zeroes).
template <typename T>
void launch_colstats(Ctx& c, const T* dX, long long n, int d, const T* dshift, double* dsum, double* dsumsq);

// Gram, generic path (DMMA, fp64 accumulate).  Accumulates the upper block triangle into dG
// (double d x d, zeroed by caller); mirror_upper() completes the symmetric matrix.
template <typename T>
void launch_gram_generic(Ctx& c, const T* dX, long long n, int d, const T* dshift, double* dG);
void launch_mirror_upper(Ctx& c, double* dG, int d);

// Gram, fp64 input (gram_f64.cu): 128 x 128 DMMA tiles, cp.async ring, persistent schedule, fixed summation order.
// Writes the upper block triangle (by 128) of dG; mirror_upper() completes it.
// K7, fp64 input (project_f64.cu): 128-row tiles on the DMMA pipe, k <= 128
bool project_f64_eligible(const Ctx& c, const double* dX, long long n, int d, int k, const double* dQ);
void launch_project_f64(Ctx& c, const double* dX, long long n, int d, int k, const double* dQ, const double* dmean,
                        const double* dw, double* dZ, double* dU, int ldz = 0);   // ldz: row pitch of Z / U (0 = k)
bool gram_f64_eligible(const Ctx& c, const double* dX, long long n, int d);
size_t gram_f64_workspace_bytes(const Ctx& c, long long n, int d);
void launch_gram_f64(Ctx& c, const double* dX, long long n, int d, const double* dshift, double* dG, void* workspace);

// Gram + column statistics, tcgen05 path (fp32 input, 3xTF32).  Returns false if the shape is not
// eligible (caller falls back to the generic path).  Writes the full symmetric dG (double) and,
// if dsum != nullptr, the column sums / sums of squares of (x - stat_shift).
bool gram_tc_eligible(const Ctx& c, const float* dX, long long n, int d);
size_t gram_tc_workspace_bytes(const Ctx& c, int d);  // per-CTA fp32 partial-tile slabs
void launch_gram_tc(Ctx& c, const float* dX, long long n, int d, const float* dshift, double* dG, double* dsum,
                    double* dsumsq, void* workspace);

// CTA-pair variant (gram_tc2.cu): d > 128, d % 32 == 0.  Dispatched by launch_gram_tc.
bool gram_pair_eligible(const Ctx& c, const float* dX, long long n, int d);
size_t gram_pair_workspace_bytes(const Ctx& c, int d);
void launch_gram_pair(Ctx& c, const float* dX, long long n, int d, const float* dshift, double* dG, double* dsum,
                      double* dsumsq, void* workspace);
void* tensor_map_encoder(Ctx& c);  // cuTensorMapEncodeTiled, fetched lazily through the runtime

// K3: finalize.  packed = [G (d*d) | s (d) | q (d) | n (1)] doubles.
// dsum_and_count = [column sums (d) | row count (1)] of the pilot sample (all-reduced across ranks)
void launch_pilot_finish(Ctx& c, const double* dsum_and_count, int d, float* shift_f, double* shift_d);
void launch_set_scalar(Ctx& c, double* dst, double value);
void launch_add_f64(Ctx& c, double* acc, const double* x, size_t count);  // acc += x (streaming fit)
// gathered: world x (d*d + 3d + 1) doubles [G | s | q | n | c_r] of every rank; out (same layout) <- moments re-based to the
// common shift and summed in rank order (finalize.cu)
void launch_rebase_sum(Ctx& c, const double* gathered, int world, int d, double* out);
// exchange over NVLink peer memory (p2p.cu)
size_t p2p_buffer_doubles(int world, long long cap);
const double* launch_p2p_allgather(Ctx& c, const double* src, long long count);
void launch_center(Ctx& c, const double* packed, int d, int center, const double* shift_d,
                   double* dC /* d x d, may alias packed G */, double* dmean, double* dtotal_var);

// K4: blocked cyclic Jacobi (double).  A (dp x dp, row-major, symmetric; destroyed), V out.
// dp must be a multiple of 2*kJacobiBlock; use jacobi_padded_dim().
int jacobi_padded_dim(int d);
size_t jacobi_workspace_bytes(int dp);
void launch_jacobi(Ctx& c, double* dA, double* dV, int dp, double* dW, void* workspace, double tol, int max_sweeps);
void launch_pad_copy(Ctx& c, const double* src, int d, double* dst, int dp);
// row folding of narrow fp32 inputs (finalize.cu): shift tiled f times; moments of the folded problem -> G, s, q
void launch_tile_shift(Ctx& c, const float* shift, int d, int f, float* out);
void launch_tile_shift(Ctx& c, const double* shift, int d, int f, double* out);
void launch_unfold_moments(Ctx& c, const double* Gf, const double* sf, const double* qf, int d, int f, const double* Gt,
                           const double* st, const double* qt, double* G, double* s, double* q);
template <typename T>
void launch_pad_rows(Ctx& c, const T* src, long long rows, int d, int dp, T* dst);   // rows x d -> rows x dp, zero pad
void launch_unpad_moments(Ctx& c, const double* Gp, const double* sp, const double* qp, int d, int dp, double* G, double* s,
                          double* q);

// K4 (exact path): top-k eigenpairs via Householder tridiagonalisation + Sturm multisection +
// inverse iteration + back-transformation.  dA (d x d, symmetric, row-major) is destroyed; dW gets
// the k largest eigenvalues in descending order, dVt (k x d) the unit eigenvectors in rows.
size_t eigh_topk_workspace_bytes(int d, int k);
void launch_eigh_topk(Ctx& c, double* dA, int d, int k, double* dW, double* dVt, void* workspace);
// the same with the workspace in a file-scope device array (no allocation at call time; d <= eigh_topk_static_max_d())
void launch_eigh_topk_static(Ctx& c, double* dA, int d, int k, double* dW, double* dVt);
int eigh_topk_static_max_d();

// K5: subspace iteration on the d x d Gram/covariance ("power" algorithm).
size_t power_workspace_bytes(int d, int l);
int power_padded_block(int l);  // rows of dW_out / dVt_out
void launch_power(Ctx& c, const double* dC, int d, int k, int l, int n_iter, double tol, int seed,
                  double* dW_out /* l */, double* dVt_out /* l x d rows = Ritz vectors */, void* workspace);

// K6: sort descending, sign-normalise, truncate, derive outputs.
template <typename T>
void launch_postproc(Ctx& c, const double* dW, int nw, const double* dV, int ldv, bool vectors_in_rows,
                     bool check_mass, int d, int k, const double* packed_n /* pointer to n */, const double* dtotal_var,
                     const double* dmean_d, T* dQ, T* dw, T* dev, T* devr, T* dmean, bool whiten = false);

// K7: projection and its inverse.
template <typename T>
void launch_project(Ctx& c, const T* dX, long long n, int d, int k, const T* dQ, const T* dmean,
                    const T* dw, T* dZ, T* dU);
// tcgen05 projection (project_tc.cu): fp32, k <= 128 per launch, d % 4 == 0, 16-byte aligned X.
bool project_tc_eligible(const Ctx& c, const float* dX, long long n, int d, int k, const float* dZ, const float* dU);
void launch_project_tc(Ctx& c, const float* dX, long long n, int d, int k, const float* dQ, const float* dmean,
                       const float* dw, float* dZ, float* dU, int ldz = 0);      // ldz: row pitch of Z / U (0 = k)
template <typename T>
void launch_inverse(Ctx& c, const T* dZ, long long n, int d, int k, const T* dQ, const T* dmean, T* dX);

// CSR rows [row0, row0 + rows) -> dense row-major tile (csr.cu); indptr int64, indices int32, all device pointers
template <typename T>
void launch_csr_to_dense(Ctx& c, const long long* indptr, const int* indices, const T* data, long long row0, long long rows,
                         int d, T* out);

// cudaFuncSetAttribute acts on the current device only: a call site remembers, per device, the largest dynamic
// shared-memory size it has requested (a per-process flag broke the second device of a process).
struct PerDeviceMax {
  size_t v[64] = {};
  bool need(int device, size_t want) {
    if (device < 0 || device >= 64) return true;
    if (want > v[device]) { v[device] = want; return true; }
    return false;
  }
};

inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

}  // namespace dr4g
