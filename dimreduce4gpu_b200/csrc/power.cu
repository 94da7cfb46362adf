// K5 - algorithm="power": on-device block subspace iteration on the d x d Gram / covariance
// matrix with CholeskyQR2 orthonormalisation and a Rayleigh-Ritz step.
//
// Replaces power_tsvd (reference src/tsvd.cu:346-414): per-component power iteration with
// Hotelling deflation, k * n_iter tiny GEMVs and two host synchronisations per iteration
// (cublasSdot src/tsvd.cu:387, cublasSnrm2 src/matrix.cu:207-217).  Here the whole block of
// l = k + oversample vectors is iterated together and nothing returns to the host: every kernel
// is enqueued up front and a device-side `done` flag turns the remaining iterations into no-ops
// once the Ritz values have converged to `tol` (the reference's stopping rule, src/tsvd.cu:388-390,
// applied to the block).  n_iter is the iteration cap, random_state seeds the start block
// (reference: Matrix::random, src/matrix.cuh:207-220; bitwise RNG parity is not required, the
// parity oracle is the CPU backend).
//
//   Q0 = orth(randn(d, l));  repeat: Y = C Q;  theta_j = q_j . y_j;  Q = orth(Y)
//   B = Q^T C Q;  B = W diag(lambda) W^T (K4 tridiagonal solver, all l pairs);  Ritz vectors = Q W
#include <algorithm>
#include <cstdlib>

#include "common.cuh"

namespace dr4g {
namespace {

// ---- small fp64 GEMM with arbitrary strides: C(m,n) = sum_k A(m,k) B(k,n) -----------------
// Everything here is a few MFLOP on matrices that live in L2, so the kernels are latency-bound:
// TM = 32 tiles when 64 x 64 tiles would leave most SMs idle, the next k-step's operands are
// fetched into registers while the current one is multiplied (one barrier per step, two shared
// buffers), and a long contraction with a tiny output (S = Y^T Y) is split over blockIdx.z into
// partial products that the consumer (chol_inv_kernel) adds up in a fixed order.
constexpr int GK = 16;
constexpr int kSplits = 16;  // partial products of S = Y^T Y (split contraction)

template <int TM>
__global__ void __launch_bounds__(256)
dgemm_kernel(int M, int N, int K, const double* __restrict__ A, long long sa_m, long long sa_k,
             const double* __restrict__ B, long long sb_k, long long sb_n, double* __restrict__ C,
             long long sc_m, long long sc_n, const int* __restrict__ done, int kchunk, long long split_stride) {
  if (done && *done) return;
  constexpr int PT = TM / 16;          // thread tile PT x PT
  constexpr int NL = TM * GK / 256;    // operand elements per thread and k-step
  __shared__ double As[2][GK][TM + 1];
  __shared__ double Bs[2][GK][TM + 1];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TM;
  const int kbeg = blockIdx.z * kchunk, kend = min(K, kbeg + kchunk);
  C += (long long)blockIdx.z * split_stride;
  // element e of a GK x TM operand tile: the fast index follows the unit stride of the operand
  int akk[NL], amm[NL], bkk[NL], bnn[NL];
#pragma unroll
  for (int i = 0; i < NL; ++i) {
    const int e = tid + 256 * i;
    if (sa_k == 1) { akk[i] = e % GK; amm[i] = e / GK; } else { amm[i] = e % TM; akk[i] = e / TM; }
    if (sb_k == 1) { bkk[i] = e % GK; bnn[i] = e / GK; } else { bnn[i] = e % TM; bkk[i] = e / TM; }
  }
  double ra[NL], rb[NL];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int i = 0; i < NL; ++i) {
      const int m = m0 + amm[i], ka = k0 + akk[i];
      ra[i] = (m < M && ka < kend) ? A[m * sa_m + ka * sa_k] : 0.0;
      const int n = n0 + bnn[i], kb = k0 + bkk[i];
      rb[i] = (n < N && kb < kend) ? B[kb * sb_k + n * sb_n] : 0.0;
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int i = 0; i < NL; ++i) {
      As[buf][akk[i]][amm[i]] = ra[i];
      Bs[buf][bkk[i]][bnn[i]] = rb[i];
    }
  };
  double acc[PT][PT] = {};
  fetch(kbeg);
  stash(0);
  __syncthreads();
  int buf = 0;
  for (int k0 = kbeg; k0 < kend; k0 += GK) {
    const bool more = k0 + GK < kend;
    if (more) fetch(k0 + GK);
#pragma unroll
    for (int kk = 0; kk < GK; ++kk) {
      double a[PT], b[PT];
#pragma unroll
      for (int i = 0; i < PT; ++i) a[i] = As[buf][kk][ty + 16 * i];
#pragma unroll
      for (int j = 0; j < PT; ++j) b[j] = Bs[buf][kk][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < PT; ++i)
#pragma unroll
        for (int j = 0; j < PT; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
    }
    if (more) stash(buf ^ 1);
    __syncthreads();
    buf ^= 1;
  }
#pragma unroll
  for (int i = 0; i < PT; ++i)
#pragma unroll
    for (int j = 0; j < PT; ++j) {
      const int m = m0 + ty + 16 * i, n = n0 + tx + 16 * j;
      if (m < M && n < N) C[m * sc_m + n * sc_n] = acc[i][j];
    }
}

// splits > 1: C is an array of `splits` partial products (split_stride doubles apart)
void dgemm(Ctx& c, int M, int N, int K, const double* A, long long sa_m, long long sa_k, const double* B,
           long long sb_k, long long sb_n, double* C, long long sc_m, long long sc_n, const int* done,
           int splits = 1, long long split_stride = 0) {
  const int kchunk = ceil_div(ceil_div(K, splits), GK) * GK;
  const bool small = (long long)ceil_div(N, 64) * ceil_div(M, 64) * splits < c.sm_count;
  if (small) {
    dim3 grid(ceil_div(N, 32), ceil_div(M, 32), splits);
    dgemm_kernel<32><<<grid, 256, 0, c.stream>>>(M, N, K, A, sa_m, sa_k, B, sb_k, sb_n, C, sc_m, sc_n, done, kchunk,
                                                 split_stride);
  } else {
    dim3 grid(ceil_div(N, 64), ceil_div(M, 64), splits);
    dgemm_kernel<64><<<grid, 256, 0, c.stream>>>(M, N, K, A, sa_m, sa_k, B, sb_k, sb_n, C, sc_m, sc_n, done, kchunk,
                                                 split_stride);
  }
  c.launches++;
}

// ---- start block: counter-based normal deviates --------------------------------------------
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__global__ void randn_kernel(double* __restrict__ Y, long long count, unsigned long long seed) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const unsigned long long a = mix64(seed * 0x100000001B3ull + 2 * (unsigned long long)i);
  const unsigned long long b = mix64(seed * 0x100000001B3ull + 2 * (unsigned long long)i + 1);
  const double u1 = ((double)(a >> 11) + 1.0) * (1.0 / 9007199254740993.0);
  const double u2 = (double)(b >> 11) * (1.0 / 9007199254740992.0);
  Y[i] = sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
}

// ---- Cholesky of the l x l Gram of the block and inverse of the factor (one CTA) -----------
// S = R^T R (R upper).  Rinv = R^-1 so that Q = Y Rinv.  Pivots that have collapsed (columns of Y
// linearly dependent on earlier ones: rank-deficient data) give a zero column in Rinv, i.e. the
// direction is dropped instead of amplified.  S arrives as `splits` partial products, is summed in
// a fixed order into shared memory (row stride l + 1), factored there, and R^-1 is built row by row
// from the bottom (one warp per column, lanes over the inner sum).
constexpr int kCholThreads = 256;

__global__ void __launch_bounds__(kCholThreads)
chol_inv_kernel(const double* __restrict__ Sparts, int splits, int l, double* __restrict__ Rinv,
                const int* __restrict__ done, double* __restrict__ scratch /* used when 2 l (l+1) doubles exceed smem */) {
  if (done && *done) return;
  extern __shared__ double csm[];
  const int ld = l + 1;
  double* S = scratch ? scratch : csm;   // l x ld
  double* Ri = S + (size_t)l * ld;       // l x ld
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ double piv;
  __shared__ double maxdiag;
  for (int e = tid; e < l * l; e += kCholThreads) {
    double v = 0.0;
    for (int z = 0; z < splits; ++z) v += Sparts[(size_t)z * l * l + e];
    const int r = e / l, cc = e - r * l;
    S[r * ld + cc] = v;
    Ri[r * ld + cc] = 0.0;
  }
  __syncthreads();
  if (warp == 0) {
    double m = 0;
    for (int i = lane; i < l; i += 32) m = fmax(m, S[i * ld + i]);
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) maxdiag = m;
  }
  __syncthreads();
  // right-looking Cholesky, upper factor stored in place (row j holds R[j][j..])
  for (int j = 0; j < l; ++j) {
    if (tid == 0) {
      const double p = S[j * ld + j];
      piv = (p > 1e-26 * maxdiag && p > 0) ? sqrt(p) : 0.0;
    }
    __syncthreads();
    const double rjj = piv;
    const double rinv = rjj > 0 ? 1.0 / rjj : 0.0;
    // scaled row j is needed by the update: each thread scales the entries it reads (no extra barrier),
    // the row itself is rewritten by the threads of the first update row after the barrier below
    if (rjj > 0) {
      for (int r = j + 1 + warp; r < l; r += kCholThreads / 32) {
        const double srj = S[j * ld + r] * rinv;
        for (int cc = r + lane; cc < l; cc += 32) S[r * ld + cc] -= srj * (S[j * ld + cc] * rinv);
      }
    }
    __syncthreads();
    for (int cc = j + tid; cc < l; cc += kCholThreads) S[j * ld + cc] = rjj > 0 ? S[j * ld + cc] * rinv : 0.0;
  }
  __syncthreads();
  // R^-1 (upper), rows from the bottom: Ri[r][c] = (delta_rc - sum_{t=r+1..c} R[r][t] Ri[t][c]) / R[r][r]
  for (int r = l - 1; r >= 0; --r) {
    const double rrr = S[r * ld + r];
    const double inv = rrr > 0 ? 1.0 / rrr : 0.0;
    for (int cc = r + warp; cc < l; cc += kCholThreads / 32) {
      double acc = 0.0;
      for (int t = r + 1 + lane; t <= cc; t += 32) acc += S[r * ld + t] * Ri[t * ld + cc];
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) Ri[r * ld + cc] = ((r == cc ? 1.0 : 0.0) - acc) * inv;
    }
    __syncthreads();
  }
  for (int e = tid; e < l * l; e += kCholThreads) {
    const int r = e / l, cc = e - r * l;
    Rinv[e] = Ri[r * ld + cc];
  }
}

// 1 / sqrt(p) for normal positive p: exponent folded to [1, 4), fp32 estimate, two Newton steps in fp64
// (~9 dependent fp64 operations instead of the ~60 of sqrt() followed by a division; dependent fp64
// latency is what the factorisation's pivot chain is made of).  Relative error ~2e-16.
__device__ __forceinline__ double fast_rsqrt(double p) {
  const long long bits = __double_as_longlong(p);
  const int e2 = (int)(((bits >> 52) & 0x7ff) - 1023) & ~1;          // even, <= exponent
  const double m = __longlong_as_double(bits - ((long long)e2 << 52));  // [1, 4)
  double y = (double)rsqrtf((float)m);
  const double hm = 0.5 * m;
  y = y * fma(-hm * y, y, 1.5);
  y = y * fma(-hm * y, y, 1.5);
  return y * __longlong_as_double((long long)(1023 - e2 / 2) << 52);
}

__device__ long long g_chol_trace[16];  // DR4G_CHOL_TRACE=1: phase time stamps of thread 0

// sum of the split-contraction partial products (many CTAs: all loads of an element in flight at once)
__global__ void __launch_bounds__(256)
sum_partials_kernel(const double* __restrict__ parts, int splits, int count, double* __restrict__ out,
                    const int* __restrict__ done) {
  if (done && *done) return;
  const int e = blockIdx.x * 256 + threadIdx.x;
  if (e >= count) return;
  double part[kSplits];
#pragma unroll
  for (int z = 0; z < kSplits; ++z) part[z] = z < splits ? parts[(size_t)z * count + e] : 0.0;
  double acc = 0.0;
#pragma unroll
  for (int z = 0; z < kSplits; ++z) acc += part[z];
  out[e] = acc;
}

// Register-resident variant for l <= min(8 P, 32 Q): thread (warp w, lane) keeps the elements (w + 8 p, lane + 32 q)
// of the trailing matrix in registers for the whole factorisation; only the pivot row goes through shared memory
// (two buffers, one barrier per step).  Row tests are warp-uniform (real branches, finished rows cost nothing),
// column tests are predicates, indices are compile-time offsets: ~60 instructions per warp and step.  (A first
// version with a flat element map spent 1100 instructions per step on index unpacking and selects, and one with
// a branch per element 3300 cycles of exposed latency; the shared-memory kernel above ~250 us at l = 74.)
// R^-1 is then solved two lanes per column, columns private to their lanes (no block barriers).
template <int P, int Q>
__global__ void __launch_bounds__(kCholThreads)
chol_inv_reg_kernel(const double* __restrict__ S, int l, double* __restrict__ Rinv, const int* __restrict__ done,
                    int trace) {
  if (done && *done) return;
  extern __shared__ double csm[];
  const int ld = l + 1;
  double* R = csm;                        // l x ld, upper factor
  double* Ri = R + (size_t)l * ld;        // l x ld, upper inverse
  double* rowj = Ri + (size_t)l * ld;     // 2 x 32 Q pivot-row buffers (zero beyond l)
  double* rinvs = rowj + 2 * 32 * Q;      // l reciprocal pivots
  __shared__ double maxdiag_s;
  __shared__ double pivs[2];              // reciprocal pivot per row buffer
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool tr = trace && tid == 0;
  if (tr) g_chol_trace[0] = clock64();
  double v[P][Q];
  double dmax = 0.0;
  for (int e = tid; e < 2 * 32 * Q; e += kCholThreads) rowj[e] = 0.0;
#pragma unroll
  for (int p = 0; p < P; ++p)
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      const int r = warp + 8 * p, cc = lane + 32 * q;
      const bool in = r < l && cc < l;
      v[p][q] = (in && cc >= r) ? S[r * l + cc] : 0.0;   // only the upper triangle is used
      if (in && cc == r) dmax = fmax(dmax, v[p][q]);
      if (in && cc < r) Rinv[r * l + cc] = 0.0;          // the inverse is upper triangular
    }
  for (int o = 16; o > 0; o >>= 1) dmax = fmax(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
  if (tid == 0) maxdiag_s = 0.0;
  __syncthreads();
  if (lane == 0 && dmax > 0) atomicMax(reinterpret_cast<unsigned long long*>(&maxdiag_s), (unsigned long long)__double_as_longlong(dmax));
  __syncthreads();
  if (tr) g_chol_trace[1] = clock64();
  const double thr = 1e-26 * maxdiag_s;
  auto publish_pivot = [&](int buf, double pv) { pivs[buf] = (pv > thr && pv > 1e-290) ? fast_rsqrt(pv) : 0.0; };
  if (warp == 0) {
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      const int cc = lane + 32 * q;
      if (cc < l) rowj[cc] = v[0][q];
    }
    if (lane == 0) publish_pivot(0, v[0][0]);
  }
  __syncthreads();
  // step j: R[j][j..] = S[j][j..] / r_jj, trailing S[r][c] -= R[j][r] R[j][c].  The step is a latency chain
  // (barrier -> pivot row -> update of row j+1 -> its rsqrt -> barrier): all shared-memory operands are
  // fetched up front, row j+1 is updated and published first, the other rows follow branch-free.
  // Measured at l = 74 (B200, trace hook): 1460 cycles per step = ~200 dependent instructions of one warp
  // (operands 190, row update through the unrolled owner test 600, rsqrt + publication 520, barrier 100);
  // a panel-blocked factorisation (one barrier per 8 columns) is the next step if this ever matters.
  for (int j = 0; j < l; ++j) {
    const double* rj = rowj + (j & 1) * (32 * Q);
    double* rnext = rowj + ((j + 1) & 1) * (32 * Q);
    const double rinv = pivs[j & 1];
    double b[Q], a[P];
#pragma unroll
    for (int q = 0; q < Q; ++q) b[q] = rj[lane + 32 * q] * rinv;   // entries below j are stale but finite and unused
#pragma unroll
    for (int p = 0; p < P; ++p) a[p] = rj[warp + 8 * p] * rinv;     // zero beyond l
    // row j+1 (one warp, warp-uniform test)
    if (((j + 1) & 7) == warp && j + 1 < l) {
      double nx[Q];
#pragma unroll
      for (int p = 0; p < P; ++p)
        if (warp + 8 * p == j + 1) {
#pragma unroll
          for (int q = 0; q < Q; ++q) {
            if (lane + 32 * q >= j + 1) v[p][q] = fma(-a[p], b[q], v[p][q]);
            nx[q] = v[p][q];
          }
        }
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int cc = lane + 32 * q;
        if (cc >= j + 1 && cc < l) rnext[cc] = nx[q];
        if (cc == j + 1) publish_pivot((j + 1) & 1, nx[q]);
      }
    }
    // rows below j+1
#pragma unroll
    for (int p = 0; p < P; ++p) {
      const int r = warp + 8 * p;
      const bool act = r > j + 1;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const double nv = fma(-a[p], b[q], v[p][q]);
        v[p][q] = (act && lane + 32 * q >= r) ? nv : v[p][q];
      }
    }
    // row j of the factor
    if ((j & 7) == warp) {
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const int cc = lane + 32 * q;
        if (cc >= j && cc < l) R[j * ld + cc] = b[q];
        if (cc == j) rinvs[j] = rinv;
      }
    }
    __syncthreads();
  }
  if (tr) g_chol_trace[2] = clock64();
  // R^-1 (upper): two lanes per column cc, from the diagonal upwards:
  // Ri[r][c] = (delta_rc - sum_{t=r+1..c} R[r][t] Ri[t][c]) / R[r][r], with 1 / R[r][r] kept from the factorisation.
  // All lanes of a warp walk the same rows (those above their column idle) so that the shuffles are warp-uniform.
  if ((tid & ~31) < 2 * l) {
    const int cc = tid < 2 * l ? tid >> 1 : -1, h = tid & 1;
    const int rtop = min(l - 1, ((tid & ~31) + 31) >> 1);
    for (int r = rtop; r >= 0; --r) {
      double acc = 0.0;
      for (int t = r + 1 + h; t <= cc; t += 2) acc = fma(R[r * ld + t], Ri[t * ld + cc], acc);
      acc += __shfl_xor_sync(0xffffffffu, acc, 1);
      if (h == 0 && r <= cc) {
        const double val = ((r == cc ? 1.0 : 0.0) - acc) * rinvs[r];   // rinvs[r] = 0 for a dropped pivot
        Ri[r * ld + cc] = val;
        Rinv[r * l + cc] = val;
      }
      __syncwarp();
    }
  }
  if (tr) g_chol_trace[3] = clock64();
}

// theta_j = q_j . y_j ; convergence of the leading k Ritz estimates against the previous iteration
constexpr int kRitzThreads = 1024;

__global__ void __launch_bounds__(kRitzThreads)
ritz_check_kernel(const double* __restrict__ Q, const double* __restrict__ Y, int d, int l, int k,
                  double* __restrict__ theta_prev, double tol, int* __restrict__ done, int iter) {
  if (*done) return;
  __shared__ int bad;
  if (threadIdx.x == 0) bad = 0;
  __syncthreads();
  // a thread walks one column (consecutive threads: consecutive columns, coalesced), kRitzThreads / kc row slices
  const int kc = k <= 32 ? 32 : (k <= 64 ? 64 : (k <= 128 ? 128 : 256));
  const int slices = kRitzThreads / kc;
  __shared__ double part[kRitzThreads];
  for (int j0 = 0; j0 < k; j0 += kc) {
    const int j = j0 + (int)(threadIdx.x % kc), sl = threadIdx.x / kc;
    double acc = 0;
    if (j < k)
      for (int r = sl; r < d; r += slices) acc += Q[(size_t)r * l + j] * Y[(size_t)r * l + j];
    part[threadIdx.x] = acc;
    __syncthreads();
    if (sl == 0 && j < k) {
      double th = 0;
      for (int q = 0; q < slices; ++q) th += part[q * kc + (threadIdx.x % kc)];
      const double pr = theta_prev[j];
      if (iter == 0 || !(fabs(th - pr) <= tol * fabs(pr))) bad = 1;
      theta_prev[j] = th;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0 && !bad) *done = 1;
}

}  // namespace

constexpr size_t kCholSmemMax = 200 * 1024;

struct PowerWs {
  double *Y, *Z, *Q, *S, *Ssum, *Rinv, *theta, *chol_scratch, *B, *Wl, *Vl;
  void* eig_ws;
  int* done;
};

static size_t r256(size_t b) { return (b + 255) / 256 * 256; }
static size_t chol_bytes(int l) { return (size_t)2 * l * (l + 1) * 8; }

static PowerWs carve(void* workspace, int d, int l) {
  PowerWs w;
  char* p = static_cast<char*>(workspace);
  auto take = [&](size_t bytes) { void* r = p; p += r256(bytes); return r; };
  w.Y = (double*)take((size_t)d * l * 8);
  w.Z = (double*)take((size_t)d * l * 8);
  w.Q = (double*)take((size_t)d * l * 8);
  w.S = (double*)take((size_t)kSplits * l * l * 8);
  w.Ssum = (double*)take((size_t)l * l * 8);
  w.Rinv = (double*)take((size_t)l * l * 8);
  w.theta = (double*)take((size_t)l * 8);
  w.chol_scratch = chol_bytes(l) > kCholSmemMax ? (double*)take(chol_bytes(l)) : nullptr;
  w.B = (double*)take((size_t)l * l * 8);
  w.Wl = (double*)take((size_t)l * 8);
  w.Vl = (double*)take((size_t)l * l * 8);
  w.eig_ws = take(eigh_topk_workspace_bytes(l, l));
  w.done = (int*)take(256);
  return w;
}

size_t power_workspace_bytes(int d, int l) {
  return 3 * r256((size_t)d * l * 8) + r256((size_t)kSplits * l * l * 8) + 4 * r256((size_t)l * l * 8) +
         2 * r256((size_t)l * 8) + (chol_bytes(l) > kCholSmemMax ? r256(chol_bytes(l)) : 0) +
         r256(eigh_topk_workspace_bytes(l, l)) + 256 + 256;
}

int power_padded_block(int l) { return l; }

// Q = orth(Y) by CholeskyQR2: Y -> Z -> Q.  Every launch is guarded by `done` (may be null).
static void cholqr2(Ctx& c, PowerWs& w, int d, int l, const int* done) {
  constexpr int kRegP = 10, kRegQ = 3;           // register-resident factorisation up to l = 80
  const bool reg = l <= 8 * kRegP && l <= 32 * kRegQ;
  const size_t smem = reg ? chol_bytes(l) + (size_t)(2 * 32 * kRegQ + l) * 8 : (w.chol_scratch ? 0 : chol_bytes(l));
  static PerDeviceMax smem_set[2];
  if (smem > 48 * 1024 && smem_set[reg].need(c.device, smem)) {
    if (reg) DR4G_CUDA(cudaFuncSetAttribute(chol_inv_reg_kernel<kRegP, kRegQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else DR4G_CUDA(cudaFuncSetAttribute(chol_inv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  }
  const int splits = std::max(1, std::min(kSplits, d / 64));
  for (int pass = 0; pass < 2; ++pass) {
    const double* src = pass == 0 ? w.Y : w.Z;
    double* dst = pass == 0 ? w.Z : w.Q;
    dgemm(c, l, l, d, src, 1, l, src, l, 1, w.S, l, 1, done, splits, (long long)l * l);  // S = src^T src (partials)
    static const int trace = getenv("DR4G_CHOL_TRACE") ? atoi(getenv("DR4G_CHOL_TRACE")) : 0;
    if (reg) {
      sum_partials_kernel<<<ceil_div(l * l, 256), 256, 0, c.stream>>>(w.S, splits, l * l, w.Ssum, done);
      c.launches++;
      chol_inv_reg_kernel<kRegP, kRegQ><<<1, kCholThreads, smem, c.stream>>>(w.Ssum, l, w.Rinv, done, trace);
      static int printed = 0;
      if (trace && printed < 4) {
        long long h[16];
        DR4G_CUDA(cudaStreamSynchronize(c.stream));
        DR4G_CUDA(cudaMemcpyFromSymbol(h, g_chol_trace, sizeof(h)));
        fprintf(stderr, "chol trace l=%d (cycles, thread 0): load %lld, factor %lld, inverse (column 0 only) %lld\n", l,
                h[1] - h[0], h[2] - h[1], h[3] - h[2]);
        ++printed;
      }
    }
    else chol_inv_kernel<<<1, kCholThreads, smem, c.stream>>>(w.S, splits, l, w.Rinv, done, w.chol_scratch);
    c.launches++;
    dgemm(c, d, l, l, src, l, 1, w.Rinv, l, 1, dst, l, 1, done);   // dst = src Rinv
  }
}

__global__ void symmetrize_kernel(const double* __restrict__ B, int l, double* __restrict__ A) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
  if (c >= l) return;
  A[(size_t)r * l + c] = 0.5 * (B[(size_t)r * l + c] + B[(size_t)c * l + r]);
}

void launch_power(Ctx& c, const double* dC, int d, int k, int l, int n_iter, double tol, int seed,
                  double* dW_out /* l */, double* dVt_out /* l x d */, void* workspace) {
  PowerWs w = carve(workspace, d, l);
  DR4G_CUDA(cudaMemsetAsync(w.done, 0, 256, c.stream));
  DR4G_CUDA(cudaMemsetAsync(w.theta, 0, (size_t)l * 8, c.stream));
  const long long cnt = (long long)d * l;
  randn_kernel<<<ceil_div(cnt, 256), 256, 0, c.stream>>>(w.Y, cnt, (unsigned long long)(unsigned)seed);
  c.launches++;
  cholqr2(c, w, d, l, nullptr);
  if (n_iter < 1) n_iter = 1;
  for (int it = 0; it < n_iter; ++it) {
    dgemm(c, d, l, d, dC, d, 1, w.Q, l, 1, w.Y, l, 1, w.done);   // Y = C Q
    ritz_check_kernel<<<1, kRitzThreads, 0, c.stream>>>(w.Q, w.Y, d, l, k, w.theta, tol, w.done, it);
    c.launches++;
    cholqr2(c, w, d, l, w.done);                                 // Q = orth(Y) unless converged
  }
  // Rayleigh-Ritz on span(Q): B = Q^T C Q (l x l), all l eigenpairs by the tridiagonal solver (K4)
  dgemm(c, d, l, d, dC, d, 1, w.Q, l, 1, w.Y, l, 1, nullptr);      // Y = C Q
  dgemm(c, l, l, d, w.Q, 1, l, w.Y, l, 1, w.S, l, 1, nullptr);     // S = Q^T Y
  {
    dim3 grid(ceil_div(l, 128), l);
    symmetrize_kernel<<<grid, 128, 0, c.stream>>>(w.S, l, w.B);
    c.launches++;
  }
  launch_eigh_topk(c, w.B, l, l, dW_out, w.Vl, w.eig_ws);          // Vl rows = eigenvectors of B, descending
  // Ritz vectors as rows: Vt[j][col] = sum_i Vl[j][i] Q[col][i]
  dgemm(c, l, d, l, w.Vl, l, 1, w.Q, 1, l, dVt_out, d, 1, nullptr);
  DR4G_CUDA(cudaGetLastError());
}

}  // namespace dr4g
