// K7 for fp64 input on the DMMA pipe - streamed projection  Z = (X - mean) Q^T,  U = Z / sigma   (k <= 128, d even).
//
// Replaces the reference's cublasDgemm X V + divide-by-sigma of the double entry points (src/tsvd.cu:152-177,229-260 via
// src/matrix.cu:95-127, reached from pca_double / truncated_svd_double) for BASELINE config 5 (k = 100, d = 4096).
// AI = d k / (4 (d + k)) ~ k / 4 flop/B: 25 at k = 100, so on B200 (HBM 6.5 TB/s x 25 = 160 TF/s against ~35 TF/s of
// fp64) the projection is bound by the fp64 pipe, not by HBM; the CUDA-core rowgemm kernel it replaces ran at 0.3 TB/s.
//
// Tiling.  A CTA of 8 warps owns 128 rows of X at a time (persistent over row tiles); warp w owns rows 16 w .. 16 w + 15
// and ALL components: 2 x NJ m8n8 accumulator tiles per lane (NJ = ceil(k / 8) <= 16, instantiated exactly so that k = 100
// issues 13 column tiles, not 16).  The contraction runs over the features in chunks of 16: a stage is the X tile
// (128 rows x 16 columns) and the Q chunk (8 NJ rows x 16 columns), copied global -> shared with 16-byte cp.async through
// a 4-stage ring; the pitch of 20 doubles makes the fragment loads conflict-free (rows x 20 + k hits 16 distinct 8-byte
// banks per half-warp).  Q (k x d, 3.3 MB at config 5) stays in L2.  The mean is applied as Z = X Q^T - (Q mean) in the
// epilogue (fp64: the cancellation costs |mean| / |x - mean| ulps, far inside the 1e-10 gate); Q mean is formed by a
// tiny kernel first.  Results leave through shared memory so that Z and U rows are written as whole 128-byte lines.
#include "common.cuh"

namespace dr4g {
namespace {

constexpr int PR = 128;        // rows per CTA tile
constexpr int PKC = 16;        // feature columns per stage
constexpr int PPITCH = 20;     // doubles per shared row: 20 % 16 == 4
constexpr int PSTAGES = 4;
constexpr int PTHREADS = 256;

__device__ __forceinline__ void pdmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}
__device__ __forceinline__ void pcp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void pcp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void pcp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// mq[j] = sum_c Q[j][c] mean[c]   (grid = k, block = 256)
__global__ void __launch_bounds__(256)
project_f64_meandot_kernel(const double* __restrict__ Q, const double* __restrict__ mean, int d, double* __restrict__ mq) {
  __shared__ double part[8];
  const double* q = Q + (size_t)blockIdx.x * d;
  double acc = 0.0;
  for (int c = threadIdx.x; c < d; c += 256) acc += q[c] * mean[c];
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < 8; ++w) s += part[w];
    mq[blockIdx.x] = s;
  }
}

template <int NJ>
__global__ void __launch_bounds__(PTHREADS, 1)
project_f64_kernel(const double* __restrict__ X, long long n, int d, int k, const double* __restrict__ Q,
                   const double* __restrict__ mq, const double* __restrict__ sigma, double* __restrict__ Z,
                   double* __restrict__ U, int ldz) {
  constexpr int KP = NJ * 8;
  constexpr int kXs = PR * PPITCH, kQs = KP * PPITCH;          // doubles per stage
  constexpr int kStage = kXs + kQs;
  extern __shared__ __align__(16) double psm[];
  const uint32_t sm_base = static_cast<uint32_t>(__cvta_generic_to_shared(psm));
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cc = lane >> 2, kl = lane & 3;
  const int nst = (d + PKC - 1) / PKC;
  const long long ntiles = (n + PR - 1) / PR;

  // per-lane epilogue constants: this lane's C columns are j * 8 + 2 kl + {0, 1}
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long r0 = tile * PR;
    double acc[2][NJ][2];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < NJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    auto issue = [&](int st) {
      if (st < nst) {
        const int slot = st % PSTAGES;
        const uint32_t sx = sm_base + (uint32_t)(slot * kStage) * 8u;
        const uint32_t sq = sx + (uint32_t)kXs * 8u;
        const int c0 = st * PKC;
        // X tile: 128 rows x 8 chunks of 16 B
        for (int e = tid; e < PR * (PKC / 2); e += PTHREADS) {
          const int rr = e >> 3, ch = e & 7;
          const long long r = r0 + rr;
          const int col = c0 + 2 * ch;
          const bool ok = r < n && col < d;       // d is even: a 16-byte chunk is inside or outside as a whole
          pcp_async16(sx + (uint32_t)(rr * PPITCH + 2 * ch) * 8u, X + (ok ? r * d + col : 0), ok ? 16 : 0);
        }
        // Q chunk: KP rows x 8 chunks (rows >= k zero-filled)
        for (int e = tid; e < KP * (PKC / 2); e += PTHREADS) {
          const int rr = e >> 3, ch = e & 7;
          const int col = c0 + 2 * ch;
          const bool ok = rr < k && col < d;
          pcp_async16(sq + (uint32_t)(rr * PPITCH + 2 * ch) * 8u, Q + (ok ? (size_t)rr * d + col : 0), ok ? 16 : 0);
        }
      }
      pcp_commit();   // one group per stage, possibly empty: keeps the wait count uniform
    };

#pragma unroll
    for (int st = 0; st < PSTAGES - 1; ++st) issue(st);
    for (int st = 0; st < nst; ++st) {
      pcp_wait<PSTAGES - 2>();
      __syncthreads();                 // stage st has landed for everyone; everyone is done with stage st - 1
      issue(st + PSTAGES - 1);
      const double* Xs = psm + (size_t)(st % PSTAGES) * kStage;
      const double* Qs = Xs + kXs;
#pragma unroll
      for (int k0 = 0; k0 < PKC; k0 += 4) {
        const double* ar = Xs + (warp * 16 + cc) * PPITCH + k0 + kl;
        const double* br = Qs + cc * PPITCH + k0 + kl;
        const double a0 = ar[0], a1 = ar[8 * PPITCH];
        double b[NJ];
#pragma unroll
        for (int j = 0; j < NJ; ++j) b[j] = br[j * 8 * PPITCH];
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          pdmma884(acc[0][j][0], acc[0][j][1], a0, b[j]);
          pdmma884(acc[1][j][0], acc[1][j][1], a1, b[j]);
        }
      }
    }
    pcp_wait<0>();
    __syncthreads();                   // the ring is idle: reuse its first bytes as the output tile [128][KP + 2]

    constexpr int OP = KP + 2;         // pitch of the staging tile (even, so double2 stores stay aligned)
    double* out = psm;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const int lr = warp * 16 + i * 8 + cc, lc = j * 8 + 2 * kl;
        double m0 = 0.0, m1 = 0.0;
        if (mq) { m0 = lc < k ? mq[lc] : 0.0; m1 = lc + 1 < k ? mq[lc + 1] : 0.0; }
        *reinterpret_cast<double2*>(&out[lr * OP + lc]) = make_double2(acc[i][j][0] - m0, acc[i][j][1] - m1);
      }
    __syncthreads();
    for (int e = tid; e < PR * k; e += PTHREADS) {
      const int rr = e / k, c = e - rr * k;
      const long long r = r0 + rr;
      if (r >= n) break;
      const double z = out[rr * OP + c];
      Z[(size_t)r * ldz + c] = z;      // ldz: row pitch of Z / U (>= k: a column slice of wider scores)
      if (U) { const double sg = sigma[c]; U[(size_t)r * ldz + c] = sg != 0.0 ? z / sg : 0.0; }   // sigma == 0 -> U column 0 (reference src/tsvd.cu:171-175)
    }
    __syncthreads();                   // the staging tile is read before the next tile's copies land on it
  }
}

template <int NJ>
void launch_nj(Ctx& c, const double* dX, long long n, int d, int k, const double* dQ, const double* mq, const double* dw,
               double* dZ, double* dU, int ldz) {
  constexpr size_t smem = (size_t)PSTAGES * (PR + NJ * 8) * PPITCH * sizeof(double);
  static_assert(smem >= (size_t)PR * (NJ * 8 + 2) * sizeof(double), "the staging tile fits the ring");
  static PerDeviceMax smem_set;
  if (smem_set.need(c.device, smem))
    DR4G_CUDA(cudaFuncSetAttribute(project_f64_kernel<NJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long ntiles = (n + PR - 1) / PR;
  const int grid = (int)(ntiles < c.sm_count ? ntiles : c.sm_count);
  project_f64_kernel<NJ><<<grid, PTHREADS, smem, c.stream>>>(dX, n, d, k, dQ, mq, dw, dZ, dU, ldz);
}

}  // namespace

bool project_f64_eligible(const Ctx& c, const double* dX, long long n, int d, int k, const double* dQ) {
  if (n < 1 || d < 2 || (d % 2) != 0 || k < 1 || k > 128) return false;
  return (reinterpret_cast<uintptr_t>(dX) & 15) == 0 && (reinterpret_cast<uintptr_t>(dQ) & 15) == 0;
}

void launch_project_f64(Ctx& c, const double* dX, long long n, int d, int k, const double* dQ, const double* dmean,
                        const double* dw, double* dZ, double* dU, int ldz) {
  if (ldz <= 0) ldz = k;
  double* mq = nullptr;
  if (dmean) {
    const size_t need = 128 * sizeof(double);
    if (need > c.proj_cap) {   // grow-only, private to the projection (shared with the fp32 tensor-core path's image)
      if (c.proj_buf) {
        DR4G_CUDA(cudaStreamSynchronize(c.stream));
        DR4G_CUDA(cudaFree(c.proj_buf));
        c.proj_buf = nullptr;
        c.proj_cap = 0;
      }
      DR4G_CUDA(cudaMalloc(&c.proj_buf, 4096));
      c.proj_cap = 4096;
    }
    mq = static_cast<double*>(c.proj_buf);
    project_f64_meandot_kernel<<<k, 256, 0, c.stream>>>(dQ, dmean, d, mq);
    c.launches++;
  }
  const double* sg = (dU && dw) ? dw : nullptr;
  double* du = (dU && dw) ? dU : nullptr;
  const int nj = (k + 7) / 8;
  switch (nj) {
#define DR4G_NJ(N) case N: launch_nj<N>(c, dX, n, d, k, dQ, mq, sg, dZ, du, ldz); break;
    DR4G_NJ(1) DR4G_NJ(2) DR4G_NJ(3) DR4G_NJ(4) DR4G_NJ(5) DR4G_NJ(6) DR4G_NJ(7) DR4G_NJ(8)
    DR4G_NJ(9) DR4G_NJ(10) DR4G_NJ(11) DR4G_NJ(12) DR4G_NJ(13) DR4G_NJ(14) DR4G_NJ(15) DR4G_NJ(16)
#undef DR4G_NJ
    default: fail("project_f64: k=%d out of range", k);
  }
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

}  // namespace dr4g
