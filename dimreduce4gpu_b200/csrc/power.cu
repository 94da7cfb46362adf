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
//   B = Q^T C Q;  B = W diag(lambda) W^T (K4 Jacobi);  Ritz vectors = Q W
#include "common.cuh"

namespace dr4g {
namespace {

// ---- small fp64 GEMM with arbitrary strides: C(m,n) = sum_k A(m,k) B(k,n) -----------------
constexpr int GT = 64, GK = 16;

__global__ void __launch_bounds__(256)
dgemm_kernel(int M, int N, int K, const double* __restrict__ A, long long sa_m, long long sa_k,
             const double* __restrict__ B, long long sb_k, long long sb_n, double* __restrict__ C,
             long long sc_m, long long sc_n, const int* __restrict__ done) {
  if (done && *done) return;
  __shared__ double As[GK][GT + 1];
  __shared__ double Bs[GK][GT + 1];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * GT, n0 = blockIdx.x * GT;
  double acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += GK) {
    for (int e = tid; e < GK * GT; e += 256) {
      int kk, mm;
      if (sa_k == 1) { kk = e % GK; mm = e / GK; } else { mm = e % GT; kk = e / GT; }
      const int m = m0 + mm, k = k0 + kk;
      As[kk][mm] = (m < M && k < K) ? A[m * sa_m + k * sa_k] : 0.0;
    }
    for (int e = tid; e < GK * GT; e += 256) {
      int kk, nn;
      if (sb_k == 1) { kk = e % GK; nn = e / GK; } else { nn = e % GT; kk = e / GT; }
      const int n = n0 + nn, k = k0 + kk;
      Bs[kk][nn] = (n < N && k < K) ? B[k * sb_k + n * sb_n] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < GK; ++kk) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty + 16 * i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = m0 + ty + 16 * i, n = n0 + tx + 16 * j;
      if (m < M && n < N) C[m * sc_m + n * sc_n] = acc[i][j];
    }
}

void dgemm(Ctx& c, int M, int N, int K, const double* A, long long sa_m, long long sa_k, const double* B,
           long long sb_k, long long sb_n, double* C, long long sc_m, long long sc_n, const int* done) {
  dim3 grid(ceil_div(N, GT), ceil_div(M, GT));
  dgemm_kernel<<<grid, 256, 0, c.stream>>>(M, N, K, A, sa_m, sa_k, B, sb_k, sb_n, C, sc_m, sc_n, done);
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
// direction is dropped instead of amplified.
__global__ void __launch_bounds__(256)
chol_inv_kernel(double* __restrict__ S, int l, double* __restrict__ Rinv, const int* __restrict__ done) {
  if (done && *done) return;
  const int tid = threadIdx.x;
  __shared__ double piv;
  __shared__ double maxdiag;
  if (tid == 0) {
    double m = 0;
    for (int i = 0; i < l; ++i) m = fmax(m, S[(size_t)i * l + i]);
    maxdiag = m;
  }
  __syncthreads();
  // right-looking Cholesky, upper factor stored in place (row j holds R[j][j..])
  for (int j = 0; j < l; ++j) {
    if (tid == 0) {
      const double p = S[(size_t)j * l + j];
      piv = (p > 1e-26 * maxdiag && p > 0) ? sqrt(p) : 0.0;
    }
    __syncthreads();
    const double rjj = piv;
    for (int c = j + tid; c < l; c += 256) S[(size_t)j * l + c] = (rjj > 0) ? S[(size_t)j * l + c] / rjj : 0.0;
    __syncthreads();
    if (rjj > 0) {
      const int rem = l - j - 1;
      for (int e = tid; e < rem * rem; e += 256) {
        const int r = j + 1 + e / rem, c = j + 1 + e % rem;
        if (c >= r) S[(size_t)r * l + c] -= S[(size_t)j * l + r] * S[(size_t)j * l + c];
      }
    }
    __syncthreads();
  }
  // Rinv (upper): column c solved independently by back substitution R x = e_c
  for (int c = tid; c < l; c += 256) {
    for (int r = l - 1; r >= 0; --r) {
      double v = 0.0;
      if (r <= c) {
        const double rrr = S[(size_t)r * l + r];
        if (rrr > 0) {
          double acc = (r == c) ? 1.0 : 0.0;
          for (int t = r + 1; t <= c; ++t) acc -= S[(size_t)r * l + t] * Rinv[(size_t)t * l + c];
          v = acc / rrr;
        }
      }
      Rinv[(size_t)r * l + c] = v;
    }
  }
}

// theta_j = q_j . y_j ; convergence of the leading k Ritz estimates against the previous iteration
__global__ void __launch_bounds__(256)
ritz_check_kernel(const double* __restrict__ Q, const double* __restrict__ Y, int d, int l, int k,
                  double* __restrict__ theta_prev, double tol, int* __restrict__ done, int iter) {
  if (*done) return;
  __shared__ double red[256];
  __shared__ int bad;
  if (threadIdx.x == 0) bad = 0;
  __syncthreads();
  for (int j = 0; j < k; ++j) {
    double acc = 0;
    for (int r = threadIdx.x; r < d; r += 256) acc += Q[(size_t)r * l + j] * Y[(size_t)r * l + j];
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      const double th = red[0], pr = theta_prev[j];
      if (iter == 0 || !(fabs(th - pr) <= tol * fabs(pr))) bad = 1;
      theta_prev[j] = th;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0 && !bad) *done = 1;
}

}  // namespace

struct PowerWs {
  double *Y, *Z, *Q, *S, *Rinv, *theta, *jac_A, *jac_V;
  void* jac_ws;
  int* done;
  int lp;
};

static size_t r256(size_t b) { return (b + 255) / 256 * 256; }

static PowerWs carve(void* workspace, int d, int l) {
  PowerWs w;
  char* p = static_cast<char*>(workspace);
  auto take = [&](size_t bytes) { void* r = p; p += r256(bytes); return r; };
  w.lp = jacobi_padded_dim(l);
  w.Y = (double*)take((size_t)d * l * 8);
  w.Z = (double*)take((size_t)d * l * 8);
  w.Q = (double*)take((size_t)d * l * 8);
  w.S = (double*)take((size_t)l * l * 8);
  w.Rinv = (double*)take((size_t)l * l * 8);
  w.theta = (double*)take((size_t)l * 8);
  w.jac_A = (double*)take((size_t)w.lp * w.lp * 8);
  w.jac_V = (double*)take((size_t)w.lp * w.lp * 8);
  w.jac_ws = take(jacobi_workspace_bytes(w.lp));
  w.done = (int*)take(256);
  return w;
}

size_t power_workspace_bytes(int d, int l) {
  const int lp = jacobi_padded_dim(l);
  return 3 * r256((size_t)d * l * 8) + 2 * r256((size_t)l * l * 8) + r256((size_t)l * 8) +
         2 * r256((size_t)lp * lp * 8) + r256(jacobi_workspace_bytes(lp)) + 256 + 256;
}

int power_padded_block(int l) { return jacobi_padded_dim(l); }

// Q = orth(Y) by CholeskyQR2: Y -> Z -> Q.  Every launch is guarded by `done` (may be null).
static void cholqr2(Ctx& c, PowerWs& w, int d, int l, const int* done) {
  for (int pass = 0; pass < 2; ++pass) {
    const double* src = pass == 0 ? w.Y : w.Z;
    double* dst = pass == 0 ? w.Z : w.Q;
    dgemm(c, l, l, d, src, 1, l, src, l, 1, w.S, l, 1, done);      // S = src^T src
    chol_inv_kernel<<<1, 256, 0, c.stream>>>(w.S, l, w.Rinv, done);
    c.launches++;
    dgemm(c, d, l, l, src, l, 1, w.Rinv, l, 1, dst, l, 1, done);   // dst = src Rinv
  }
}

__global__ void symmetrize_pad_kernel(const double* __restrict__ B, int l, double* __restrict__ A, int lp) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
  if (c >= lp) return;
  A[(size_t)r * lp + c] = (r < l && c < l) ? 0.5 * (B[(size_t)r * l + c] + B[(size_t)c * l + r]) : 0.0;
}

void launch_power(Ctx& c, const double* dC, int d, int k, int l, int n_iter, double tol, int seed,
                  double* dW_out /* lp */, double* dVt_out /* lp x d */, void* workspace) {
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
    ritz_check_kernel<<<1, 256, 0, c.stream>>>(w.Q, w.Y, d, l, k, w.theta, tol, w.done, it);
    c.launches++;
    cholqr2(c, w, d, l, w.done);                                 // Q = orth(Y) unless converged
  }
  // Rayleigh-Ritz on span(Q)
  dgemm(c, d, l, d, dC, d, 1, w.Q, l, 1, w.Y, l, 1, nullptr);      // Y = C Q
  dgemm(c, l, l, d, w.Q, 1, l, w.Y, l, 1, w.S, l, 1, nullptr);     // S = Q^T Y
  {
    dim3 grid(ceil_div(w.lp, 128), w.lp);
    symmetrize_pad_kernel<<<grid, 128, 0, c.stream>>>(w.S, l, w.jac_A, w.lp);
    c.launches++;
  }
  launch_jacobi(c, w.jac_A, w.jac_V, w.lp, dW_out, w.jac_ws, 1e-14, 30);
  // Ritz vectors as rows: Vt[j][col] = sum_i Vl[i][j] Q[col][i]
  dgemm(c, w.lp, d, l, w.jac_V, 1, w.lp, w.Q, 1, l, dVt_out, d, 1, nullptr);
  DR4G_CUDA(cudaGetLastError());
}

}  // namespace dr4g
