// K7 - streamed projection Z = (X - mean) * Q^T  (+ U = Z / sigma), and its inverse
// X = Z * Q + mean.
//
// Replaces calculate_u + multiply_diag (reference src/tsvd.cu:152-177,229-260: SGEMM X*V, a
// divide-by-sigma lambda, dgmm by sigma again and three n x k D2H copies) and the host-side numpy
// transform (dimreduce4gpu/truncated_svd.py:116-118).  The mean is subtracted while the X tile is
// staged in shared memory, so centred data is never materialised (reference src/pca.cu:46-49).
// inverse_transform does not exist in the reference; semantics follow sklearn.
//
// One kernel serves both directions:  C[r][m] = sum_t (A[r][t] - a_shift[t]) * B(m,t) + c_shift[m]
//   project: A = X (n x d), B(m,t) = Q[m*d + t],  m < k, t < d
//   inverse: A = Z (n x k), B(m,t) = Q[t*d + m],  m < d, t < k
// CTA = 256 threads, 128 rows x 64 outputs per pass, 32-wide (fp32) / 16-wide (fp64) contraction chunks; each thread owns
// an 8 x 4 register tile.  HBM-bound for k << d: algorithmic bytes = n*(d+k)*sizeof(T).
#include "common.cuh"

namespace dr4g {
namespace {

constexpr int RT = 128;  // rows per CTA
constexpr int MT = 64;   // outputs per pass

template <typename T>
__global__ void __launch_bounds__(256)
rowgemm_kernel(const T* __restrict__ A, long long n, int lda, int tdim, const T* __restrict__ a_shift,
               const T* __restrict__ B, long long bs_m, long long bs_t, int mdim, const T* __restrict__ c_shift,
               const T* __restrict__ scale_div, T* __restrict__ C, T* __restrict__ C2, int ldc) {
  constexpr int TC = sizeof(T) == 8 ? 16 : 32;  // contraction chunk (keeps static smem < 48 KB)
  constexpr int AR = 256 / TC;                  // A-tile rows staged per pass
  __shared__ T As[RT][TC + 1];
  __shared__ T Bs[MT][TC + 1];
  const int tid = threadIdx.x;
  const int ty = tid >> 4, tx = tid & 15;
  const long long row0 = (long long)blockIdx.x * RT;

  for (int m0 = 0; m0 < mdim; m0 += MT) {
    T acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = (T)0;

    for (int t0 = 0; t0 < tdim; t0 += TC) {
      // stage A tile (coalesced along t), centred
      {
        const int tc = tid % TC, tr = tid / TC;
        const int t = t0 + tc;
        const T sh = (a_shift && t < tdim) ? a_shift[t] : (T)0;
#pragma unroll
        for (int i = 0; i < RT / AR; ++i) {
          const long long r = row0 + tr + AR * i;
          As[tr + AR * i][tc] = (r < n && t < tdim) ? A[r * lda + t] - sh : (T)0;
        }
      }
      // stage B tile
      if (bs_t == 1) {
        const int tc = tid % TC, mr = tid / TC;
#pragma unroll
        for (int i = 0; i < MT / AR; ++i) {
          const int m = m0 + mr + AR * i, t = t0 + tc;
          Bs[mr + AR * i][tc] = (m < mdim && t < tdim) ? B[m * bs_m + t] : (T)0;
        }
      } else {
        const int mc = tid & 63, tr = tid >> 6;
#pragma unroll
        for (int i = 0; i < TC / 4; ++i) {
          const int m = m0 + mc, t = t0 + tr + 4 * i;
          Bs[mc][tr + 4 * i] = (m < mdim && t < tdim) ? B[m * bs_m + t * bs_t] : (T)0;
        }
      }
      __syncthreads();
#pragma unroll 8
      for (int t = 0; t < TC; ++t) {
        T a[8], b[4];
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = As[ty * 8 + i][t];
#pragma unroll
        for (int j = 0; j < 4; ++j) b[j] = Bs[tx + 16 * j][t];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] += a[i] * b[j];
      }
      __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int m = m0 + tx + 16 * j;
      if (m >= mdim) continue;
      const T cs = c_shift ? c_shift[m] : (T)0;
      T inv = (T)0;
      if (C2) { const T s = scale_div[m]; inv = (s != (T)0) ? (T)1 / s : (T)0; }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const long long r = row0 + ty * 8 + i;
        if (r >= n) continue;
        const T v = acc[i][j] + cs;
        C[r * ldc + m] = v;
        if (C2) C2[r * ldc + m] = v * inv;
      }
    }
  }
}

}  // namespace

template <typename T>
void launch_project(Ctx& c, const T* dX, long long n, int d, int k, const T* dQ, const T* dmean, const T* dw,
                    T* dZ, T* dU) {
  if (n <= 0) return;
  const long long blocks = (n + RT - 1) / RT;
  rowgemm_kernel<T><<<(unsigned)blocks, 256, 0, c.stream>>>(dX, n, d, d, dmean, dQ, (long long)d, 1LL, k, nullptr,
                                                            dw, dZ, (dU && dw) ? dU : nullptr, k);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

template <typename T>
void launch_inverse(Ctx& c, const T* dZ, long long n, int d, int k, const T* dQ, const T* dmean, T* dX) {
  if (n <= 0) return;
  const long long blocks = (n + RT - 1) / RT;
  rowgemm_kernel<T><<<(unsigned)blocks, 256, 0, c.stream>>>(dZ, n, k, k, nullptr, dQ, 1LL, (long long)d, d, dmean,
                                                            nullptr, dX, nullptr, d);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

template void launch_project<float>(Ctx&, const float*, long long, int, int, const float*, const float*, const float*, float*, float*);
template void launch_project<double>(Ctx&, const double*, long long, int, int, const double*, const double*, const double*, double*, double*);
template void launch_inverse<float>(Ctx&, const float*, long long, int, int, const float*, const float*, float*);
template void launch_inverse<double>(Ctx&, const double*, long long, int, int, const double*, const double*, double*);

}  // namespace dr4g
