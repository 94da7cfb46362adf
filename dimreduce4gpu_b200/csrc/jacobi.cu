// K4 - device-resident symmetric eigensolver: blocked cyclic (two-sided) Jacobi in fp64.
//
// Replaces cusolverDn{S,D}syevd (reference src/matrix.cu:231-273, called from src/tsvd.cu:329).
// One cooperative launch runs the whole solve; there is no host round trip until it has converged.
//
// The dp x dp matrix (dp = d padded to a multiple of 32) is cut into nb = dp/16 column blocks.
// A sweep is nb-1 rounds of a round-robin tournament; in a round every block meets exactly one
// partner, giving nb/2 independent 32x32 sub-problems [A_pp A_pq; A_qp A_qq]:
//   phase 1  one CTA per pair loads its sub-problem into shared memory, diagonalises it with
//            parallel-ordered scalar Jacobi (16 disjoint rotations per step, each thread owns one
//            2x2 block so updates are in place), and stores the accumulated 32x32 rotation R;
//   phase 2  T = (A R_all)^T  and  V <- V R_all      (column-block updates, 32-row tiles);
//   phase 3  A = T R_all = R_all^T A R_all           (A is symmetric, so the row update is the
//            column update of the transposed intermediate - one code path, coalesced both times).
// Phases are separated by a grid barrier (monotonic counter, one arrival per CTA).
// Convergence: the squared off-diagonal mass seen by the sub-problems of a sweep, relative to
// ||A||_F^2, is compared with tol^2 by every CTA on the same value -> uniform exit.
#include "common.cuh"

namespace dr4g {
namespace {

constexpr int JB = 16;        // block width
constexpr int JM = 2 * JB;    // sub-problem size
constexpr int JP = JM + 1;    // padded shared pitch
constexpr int kThreads = 256;
constexpr int kInnerMax = 8;

struct Ctrl {
  unsigned int barrier;
  unsigned int pad;
  double normA2;
  double off2[64];  // per sweep
};

__device__ __forceinline__ void grid_barrier(Ctrl* ctrl, unsigned int& target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    target += gridDim.x;
    __threadfence();
    atomicAdd(&ctrl->barrier, 1u);
    while (*reinterpret_cast<volatile unsigned int*>(&ctrl->barrier) < target) { __nanosleep(32); }
    __threadfence();
  }
  __syncthreads();
}

__device__ __forceinline__ void tournament(int r, int a, int m, int& p, int& q) {
  // circle method on m (even) players: player m-1 fixed
  if (a == 0) { p = m - 1; q = r; }
  else { p = (r + a) % (m - 1); q = (r - a + (m - 1)) % (m - 1); }
}

__device__ __forceinline__ void rotation(double app, double aqq, double apq, double& c, double& s) {
  if (apq == 0.0) { c = 1.0; s = 0.0; return; }
  const double theta = (aqq - app) / (2.0 * apq);
  const double t = copysign(1.0, theta) / (fabs(theta) + sqrt(theta * theta + 1.0));
  c = rsqrt(t * t + 1.0);
  s = t * c;
}

__device__ __forceinline__ double block_sum(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0;
#pragma unroll
  for (int w = 0; w < kThreads / 32; ++w) t += red[w];
  return t;
}

// phase 1: diagonalise one 32x32 sub-problem
__device__ void solve_pair(const double* __restrict__ A, int dp, int P, int Q, double* __restrict__ Rout,
                           double (*S)[JP], double (*Rm)[JP], double* red, double* off2_out) {
  const int tid = threadIdx.x;
  for (int e = tid; e < JM * JM; e += kThreads) {
    const int i = e / JM, j = e % JM;
    const int gi = (i < JB ? P * JB + i : Q * JB + i - JB);
    const int gj = (j < JB ? P * JB + j : Q * JB + j - JB);
    S[i][j] = __ldcg(&A[(size_t)gi * dp + gj]);
    Rm[i][j] = (i == j) ? 1.0 : 0.0;
  }
  __syncthreads();
  double o = 0, dg = 0;
  for (int e = tid; e < JM * JM; e += kThreads) {
    const int i = e / JM, j = e % JM;
    const double v = S[i][j];
    if (i == j) dg += v * v; else o += v * v;
  }
  double off2 = block_sum(o, red);
  const double diag2 = block_sum(dg, red);
  if (tid == 0) *off2_out = off2;

  const int a = tid >> 4, b = tid & 15;  // row pair, column pair
  for (int sw = 0; sw < kInnerMax; ++sw) {
    if (!(off2 > 1e-34 * diag2) || off2 == 0.0) break;  // uniform: off2/diag2 are block-wide values
    for (int r = 0; r < JM - 1; ++r) {
      int pa, qa, pb, qb;
      tournament(r, a, JM, pa, qa);
      tournament(r, b, JM, pb, qb);
      double ca, sa, cb, sb;
      rotation(S[pa][pa], S[qa][qa], S[pa][qa], ca, sa);
      rotation(S[pb][pb], S[qb][qb], S[pb][qb], cb, sb);
      const double s00 = S[pa][pb], s01 = S[pa][qb], s10 = S[qa][pb], s11 = S[qa][qb];
      const double r00 = Rm[2 * a][pb], r01 = Rm[2 * a][qb], r10 = Rm[2 * a + 1][pb], r11 = Rm[2 * a + 1][qb];
      __syncthreads();
      // rows: J^T S
      const double t00 = ca * s00 - sa * s10, t01 = ca * s01 - sa * s11;
      const double t10 = sa * s00 + ca * s10, t11 = sa * s01 + ca * s11;
      // columns: (.) J
      S[pa][pb] = cb * t00 - sb * t01;
      S[pa][qb] = sb * t00 + cb * t01;
      S[qa][pb] = cb * t10 - sb * t11;
      S[qa][qb] = sb * t10 + cb * t11;
      Rm[2 * a][pb] = cb * r00 - sb * r01;
      Rm[2 * a][qb] = sb * r00 + cb * r01;
      Rm[2 * a + 1][pb] = cb * r10 - sb * r11;
      Rm[2 * a + 1][qb] = sb * r10 + cb * r11;
      __syncthreads();
    }
    o = 0;
    for (int e = tid; e < JM * JM; e += kThreads) {
      const int i = e / JM, j = e % JM;
      if (i != j) { const double v = S[i][j]; o += v * v; }
    }
    off2 = block_sum(o, red);
  }
  for (int e = tid; e < JM * JM; e += kThreads) Rout[e] = Rm[e / JM][e % JM];
  __syncthreads();
}

// phases 2/3: Out[:, cols(P,Q)] = In[rows, cols(P,Q)] * R, optionally stored transposed
__device__ void col_update(const double* In, double* Out, int dp, int row0, int P, int Q,
                           const double* __restrict__ R, bool transposed, double (*Xs)[JP], double (*Rs)[JP],
                           double (*Os)[JP]) {
  const int tid = threadIdx.x;
  for (int e = tid; e < JM * JM; e += kThreads) {
    const int i = e / JM, j = e % JM;
    const int gj = (j < JB ? P * JB + j : Q * JB + j - JB);
    Xs[i][j] = __ldcg(&In[(size_t)(row0 + i) * dp + gj]);
    Rs[i][j] = __ldcg(&R[e]);
  }
  __syncthreads();
  const int r = tid >> 3, j0 = (tid & 7) * 4;
  double acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
#pragma unroll 8
  for (int k = 0; k < JM; ++k) {
    const double x = Xs[r][k];
    acc0 = fma(x, Rs[k][j0 + 0], acc0);
    acc1 = fma(x, Rs[k][j0 + 1], acc1);
    acc2 = fma(x, Rs[k][j0 + 2], acc2);
    acc3 = fma(x, Rs[k][j0 + 3], acc3);
  }
  if (!transposed) {
    const int gj = (j0 < JB ? P * JB + j0 : Q * JB + j0 - JB);  // 4 outputs never straddle the P/Q seam
    double* o = &Out[(size_t)(row0 + r) * dp + gj];
    o[0] = acc0; o[1] = acc1; o[2] = acc2; o[3] = acc3;
    __syncthreads();
  } else {
    Os[r][j0 + 0] = acc0; Os[r][j0 + 1] = acc1; Os[r][j0 + 2] = acc2; Os[r][j0 + 3] = acc3;
    __syncthreads();
    const int j = tid >> 3, rr = (tid & 7) * 4;
    const int gj = (j < JB ? P * JB + j : Q * JB + j - JB);
    double* o = &Out[(size_t)gj * dp + row0 + rr];
    o[0] = Os[rr + 0][j]; o[1] = Os[rr + 1][j]; o[2] = Os[rr + 2][j]; o[3] = Os[rr + 3][j];
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kThreads)
jacobi_kernel(double* __restrict__ A, double* __restrict__ T, double* __restrict__ V, int dp,
              double* __restrict__ W, double* __restrict__ R, Ctrl* __restrict__ ctrl, double tol, int max_sweeps) {
  __shared__ double sm0[JM][JP];
  __shared__ double sm1[JM][JP];
  __shared__ double sm2[JM][JP];
  __shared__ double red[kThreads / 32];
  __shared__ double off2_s;

  const int nb = dp / JB, npairs = nb / 2, ntiles = dp / JM;
  unsigned int target = 0;

  // ||A||_F^2 and V = I
  {
    double acc = 0;
    const size_t total = (size_t)dp * dp;
    for (size_t e = (size_t)blockIdx.x * kThreads + threadIdx.x; e < total; e += (size_t)gridDim.x * kThreads) {
      const double v = A[e];
      acc += v * v;
      V[e] = (e / dp == e % dp) ? 1.0 : 0.0;
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) atomicAdd(&ctrl->normA2, acc);
  }
  grid_barrier(ctrl, target);
  const double normA2 = *reinterpret_cast<volatile double*>(&ctrl->normA2);

  if (normA2 > 0.0) {
    for (int sweep = 0; sweep < max_sweeps && sweep < 64; ++sweep) {
      for (int round = 0; round < nb - 1; ++round) {
        // phase 1
        for (int pi = blockIdx.x; pi < npairs; pi += gridDim.x) {
          int P, Q;
          tournament(round, pi, nb, P, Q);
          if (P > Q) { const int t = P; P = Q; Q = t; }
          solve_pair(A, dp, P, Q, R + (size_t)pi * JM * JM, sm0, sm1, red, &off2_s);
          if (threadIdx.x == 0) atomicAdd(&ctrl->off2[sweep], off2_s);
        }
        grid_barrier(ctrl, target);
        // phase 2: T = (A R)^T ; V = V R
        for (int w = blockIdx.x; w < 2 * npairs * ntiles; w += gridDim.x) {
          const int which = w / (npairs * ntiles);
          const int rem = w % (npairs * ntiles);
          const int pi = rem / ntiles, tile = rem % ntiles;
          int P, Q;
          tournament(round, pi, nb, P, Q);
          if (P > Q) { const int t = P; P = Q; Q = t; }
          const double* Rp = R + (size_t)pi * JM * JM;
          if (which == 0) col_update(A, T, dp, tile * JM, P, Q, Rp, true, sm0, sm1, sm2);
          else col_update(V, V, dp, tile * JM, P, Q, Rp, false, sm0, sm1, sm2);
        }
        grid_barrier(ctrl, target);
        // phase 3: A = T R
        for (int w = blockIdx.x; w < npairs * ntiles; w += gridDim.x) {
          const int pi = w / ntiles, tile = w % ntiles;
          int P, Q;
          tournament(round, pi, nb, P, Q);
          if (P > Q) { const int t = P; P = Q; Q = t; }
          col_update(T, A, dp, tile * JM, P, Q, R + (size_t)pi * JM * JM, false, sm0, sm1, sm2);
        }
        grid_barrier(ctrl, target);
      }
      const double off2 = *reinterpret_cast<volatile double*>(&ctrl->off2[sweep]);
      if (off2 <= tol * tol * normA2) break;
    }
  }
  for (int i = blockIdx.x * kThreads + threadIdx.x; i < dp; i += gridDim.x * kThreads)
    W[i] = __ldcg(&A[(size_t)i * dp + i]);
}

}  // namespace

int jacobi_padded_dim(int d) { return (d + JM - 1) / JM * JM; }

size_t jacobi_workspace_bytes(int dp) {
  const size_t npairs = dp / JM;
  return (size_t)dp * dp * sizeof(double)            // T
         + npairs * JM * JM * sizeof(double)          // R
         + 1024;                                      // Ctrl
}

void launch_jacobi(Ctx& c, double* dA, double* dV, int dp, double* dW, void* workspace, double tol, int max_sweeps) {
  char* ws = static_cast<char*>(workspace);
  double* T = reinterpret_cast<double*>(ws);
  double* R = T + (size_t)dp * dp;
  Ctrl* ctrl = reinterpret_cast<Ctrl*>(R + (size_t)(dp / JM) * JM * JM);
  DR4G_CUDA(cudaMemsetAsync(ctrl, 0, sizeof(Ctrl), c.stream));

  int per_sm = 0;
  DR4G_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, jacobi_kernel, kThreads, 0));
  if (per_sm < 1) fail("jacobi kernel cannot be made resident");
  const int npairs = dp / JM, ntiles = dp / JM;
  long long items = 2LL * npairs * ntiles;
  long long grid = (long long)per_sm * c.sm_count;
  if (grid > items) grid = items;
  if (grid < 1) grid = 1;
  int grid_i = (int)grid;
  void* args[] = {&dA, &T, &dV, &dp, &dW, &R, &ctrl, &tol, &max_sweeps};
  DR4G_CUDA(cudaLaunchCooperativeKernel((const void*)jacobi_kernel, dim3(grid_i), dim3(kThreads), args, 0, c.stream));
  c.launches++;
}

}  // namespace dr4g
