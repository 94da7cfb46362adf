// Gram G = (X - shift)^T (X - shift), generic path: any d, any alignment, fp32 or fp64 input,
// products and accumulation in fp64 on the DMMA pipe (mma.sync.m8n8k4.f64).
//
// Replaces the reference's full cublas{S,D}gemm Gram (src/tsvd.cu:324,358 via matrix::multiply,
// src/matrix.cu:95-127).  It is (a) the fp64 product path (pca_double / truncated_svd_double;
// tcgen05 has no f64 kind) and (b) the path for fp32 shapes the tcgen05 kernel does not take
// (rows not 16-byte aligned, tiny n).  Only the upper block triangle is computed (symmetric
// half); launch_mirror_upper fills the rest.
//
// Tiling: CTA = 4 warps (2x2), 64x64 output tile, each warp 32x32 = 4x4 m8n8k4 tiles.  The row
// (contraction) axis is split across gridDim.y; partial tiles are combined with fp64 atomics.
// X is row-major, so both MMA operands come from the same kind of shared tile Xs[k][col]:
//   A(m,k) = Xs_i[k][m], B(k,n) = Xs_j[k][n].
#include "common.cuh"

namespace dr4g {
namespace {

constexpr int TB = 64;      // output tile edge
constexpr int KT = 16;      // rows (contraction) per shared stage
constexpr int PITCH = 72;   // doubles per shared row (64 + 8 -> 2-wavefront minimum on LDS.64)

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

template <typename T>
__global__ void __launch_bounds__(128)
gram_dmma_kernel(const T* __restrict__ X, long long n, int d, const T* __restrict__ shift,
                 double* __restrict__ G, int nblk, long long rows_per_split) {
  // decode upper-triangular tile pair
  int t = blockIdx.x, bi = 0;
  while (t >= nblk - bi) { t -= nblk - bi; ++bi; }
  const int bj = bi + t;
  const bool diag = (bi == bj);

  __shared__ double As[KT][PITCH];
  __shared__ double Bs[KT][PITCH];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp >> 1, wn = warp & 1;
  const int lcol = tid & 63, lrow = tid >> 6;  // loader: column within tile, row parity

  const int ca = bi * TB + lcol, cb = bj * TB + lcol;
  const bool va = ca < d, vb = cb < d;
  const double sha = (shift && va) ? (double)shift[ca] : 0.0;
  const double shb = (shift && vb) ? (double)shift[cb] : 0.0;

  const long long r0 = (long long)blockIdx.y * rows_per_split;
  long long r1 = r0 + rows_per_split;
  if (r1 > n) r1 = n;

  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

  double pa[KT / 2], pb[KT / 2];
  auto fetch = [&](long long rb) {
#pragma unroll
    for (int i = 0; i < KT / 2; ++i) {
      const long long r = rb + lrow + 2 * i;
      const bool vr = r < r1;
      pa[i] = (vr && va) ? (double)X[r * d + ca] - sha : 0.0;
      if (!diag) pb[i] = (vr && vb) ? (double)X[r * d + cb] - shb : 0.0;
    }
  };

  if (r0 < r1) fetch(r0);
  for (long long rb = r0; rb < r1; rb += KT) {
#pragma unroll
    for (int i = 0; i < KT / 2; ++i) {
      As[lrow + 2 * i][lcol] = pa[i];
      if (!diag) Bs[lrow + 2 * i][lcol] = pb[i];
    }
    __syncthreads();
    if (rb + KT < r1) fetch(rb + KT);
    const double(*Bt)[PITCH] = diag ? As : Bs;
#pragma unroll
    for (int k0 = 0; k0 < KT; k0 += 4) {
      double a[4], b[4];
      const int kr = k0 + (lane & 3), cc = lane >> 2;
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kr][wm * 32 + i * 8 + cc];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bt[kr][wn * 32 + j * 8 + cc];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dmma(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
    __syncthreads();
  }

  // C fragment: row = lane/4, cols = 2*(lane%4) + {0,1}
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gr = bi * TB + wm * 32 + i * 8 + (lane >> 2);
    if (gr >= d) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gc = bj * TB + wn * 32 + j * 8 + 2 * (lane & 3);
      if (gc < d) atomicAdd(&G[(size_t)gr * d + gc], acc[i][j][0]);
      if (gc + 1 < d) atomicAdd(&G[(size_t)gr * d + gc + 1], acc[i][j][1]);
    }
  }
}

// lower block triangle <- transpose of the upper one
__global__ void mirror_upper_kernel(double* __restrict__ G, int d) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= d) return;
  if (r / TB > c / TB) G[(size_t)r * d + c] = G[(size_t)c * d + r];
}

}  // namespace

template <typename T>
void launch_gram_generic(Ctx& c, const T* dX, long long n, int d, const T* dshift, double* dG) {
  if (n <= 0 || d <= 0) return;
  const int nblk = ceil_div(d, TB);
  const int pairs = nblk * (nblk + 1) / 2;
  long long want = ((long long)c.sm_count * 8 + pairs - 1) / pairs;
  long long max_by_rows = (n + 255) / 256;
  long long splits = want < max_by_rows ? want : max_by_rows;
  if (splits < 1) splits = 1;
  if (splits > 65535) splits = 65535;
  long long rows_per_split = (n + splits - 1) / splits;
  rows_per_split = (rows_per_split + KT - 1) / KT * KT;
  splits = (n + rows_per_split - 1) / rows_per_split;
  dim3 grid(pairs, (unsigned)splits);
  gram_dmma_kernel<T><<<grid, 128, 0, c.stream>>>(dX, n, d, dshift, dG, nblk, rows_per_split);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

void launch_mirror_upper(Ctx& c, double* dG, int d) {
  dim3 grid(ceil_div(d, 128), d);
  mirror_upper_kernel<<<grid, 128, 0, c.stream>>>(dG, d);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

template void launch_gram_generic<float>(Ctx&, const float*, long long, int, const float*, double*);
template void launch_gram_generic<double>(Ctx&, const double*, long long, int, const double*, double*);

}  // namespace dr4g
