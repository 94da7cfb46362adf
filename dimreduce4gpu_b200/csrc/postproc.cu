// K6 - descending order, sign normalisation, truncation to the top k and the derived outputs.
//
// Replaces get_tsvd_attr's d-sized work (reference src/tsvd.cu:186-226: geam transpose,
// row_reverse_trunc_q, col_reverse_q, sqrt transform, host-side std::reverse round trip) and the
// explained-variance passes over n x k / n x d data (src/tsvd.cu:262-290), which here are closed
// forms of the eigenvalues and of K1's column statistics.  Output definitions follow the CPU
// backend, which is the parity oracle (src/cpu_backend.cpp:277-332):
//   singular_values_ = sqrt(max(lambda,0)) descending;  components_ rows = eigenvectors;
//   explained_variance_ = sigma^2 / max(1, n-1);  explained_variance_ratio_ = ev / total_var.
// Sign convention: largest-|.| entry of every component row is made positive (sklearn's v-based
// svd_flip); the oracle applies no flip (src/cpu_backend.cpp:302-332), the parity harness applies
// this same convention to it before comparing.
#include <cfloat>

#include "common.cuh"

namespace dr4g {
namespace {

// key_j = lambda_j, or -DBL_MAX for eigenvectors that live in the zero padding
__global__ void keys_kernel(const double* __restrict__ W, int nw, const double* __restrict__ V, int ldv,
                            bool vectors_in_rows, bool check_mass, int d, double* __restrict__ keys) {
  // one warp per eigenpair
  const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (j >= nw) return;
  double key = W[j];
  if (check_mass) {
    double mass = 0;
    for (int r = lane; r < d; r += 32) {
      const double v = vectors_in_rows ? V[(size_t)j * ldv + r] : V[(size_t)r * ldv + j];
      mass += v * v;
    }
    for (int o = 16; o > 0; o >>= 1) mass += __shfl_xor_sync(0xffffffffu, mass, o);
    if (mass < 0.5) key = -DBL_MAX;
  }
  if (!(key == key)) key = -DBL_MAX;  // NaN guard
  if (lane == 0) keys[j] = key;
}

__global__ void rank_kernel(const double* __restrict__ keys, int nw, int k, int* __restrict__ sel) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nw) return;
  const double kj = keys[j];
  int rank = 0;
  for (int i = 0; i < nw; ++i) {
    const double ki = keys[i];
    rank += (ki > kj) || (ki == kj && i < j);
  }
  if (rank < k) sel[rank] = j;
}

template <typename T>
__global__ void gather_kernel(const double* __restrict__ keys, const int* __restrict__ sel,
                              const double* __restrict__ V, int ldv, bool vectors_in_rows, int d, int k,
                              const double* __restrict__ n_ptr, const double* __restrict__ total_var,
                              const double* __restrict__ mean_d, bool whiten, T* __restrict__ Q, T* __restrict__ w,
                              T* __restrict__ ev, T* __restrict__ evr, T* __restrict__ mean) {
  const int i = blockIdx.x;
  const int j = sel[i];
  __shared__ double bestv[256];
  __shared__ int besti[256];
  auto at = [&](int c) -> double {
    return vectors_in_rows ? V[(size_t)j * ldv + c] : V[(size_t)c * ldv + j];
  };
  double bv = -1.0;
  int bi = 0x7fffffff;
  for (int c = threadIdx.x; c < d; c += blockDim.x) {
    const double a = fabs(at(c));
    if (a > bv) { bv = a; bi = c; }
  }
  bestv[threadIdx.x] = bv;
  besti[threadIdx.x] = bi;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      const double ov = bestv[threadIdx.x + o];
      const int oi = besti[threadIdx.x + o];
      if (ov > bestv[threadIdx.x] || (ov == bestv[threadIdx.x] && oi < besti[threadIdx.x])) {
        bestv[threadIdx.x] = ov;
        besti[threadIdx.x] = oi;
      }
    }
    __syncthreads();
  }
  double sgn = (besti[0] < d && at(besti[0]) < 0.0) ? -1.0 : 1.0;
  if (whiten) {
    // opt-in (PCA(whiten=True)): component rows / sqrt(explained_variance), so that the scores X_c Q^T have unit
    // variance.  What the reference's GPU code attempts at src/tsvd.cu:195-207 (it indexes lambda by feature there);
    // its CPU backend - the parity oracle - ignores the flag, so nothing changes when it is off.
    double lam = keys[j];
    const double nn = *n_ptr;
    const double e = (lam > 0 ? lam : 0) / (nn - 1 > 1 ? nn - 1 : 1);
    sgn = e > 0 ? sgn / sqrt(e) : 0.0;
  }
  for (int c = threadIdx.x; c < d; c += blockDim.x) Q[(size_t)i * d + c] = (T)(sgn * at(c));
  if (threadIdx.x == 0) {
    double lam = keys[j];
    if (!(lam > 0)) lam = 0;
    const T sig = (T)sqrt(lam);
    w[i] = sig;
    const double n = *n_ptr;
    const double denom = n - 1 > 1 ? n - 1 : 1;
    const T e = (T)((double)sig * (double)sig / denom);
    if (ev) ev[i] = e;
    if (evr) evr[i] = (T)((double)e / *total_var);
  }
  if (i == 0 && mean && mean_d)
    for (int c = threadIdx.x; c < d; c += blockDim.x) mean[c] = (T)mean_d[c];
}

}  // namespace

template <typename T>
void launch_postproc(Ctx& c, const double* dW, int nw, const double* dV, int ldv, bool vectors_in_rows,
                     bool check_mass, int d, int k, const double* packed_n, const double* dtotal_var, const double* dmean_d,
                     T* dQ, T* dw, T* dev, T* devr, T* dmean, bool whiten) {
  double* keys = c.arena.take_n<double>(nw);
  int* sel = c.arena.take_n<int>(k);
  keys_kernel<<<ceil_div(nw, 4), 128, 0, c.stream>>>(dW, nw, dV, ldv, vectors_in_rows, check_mass, d, keys);
  rank_kernel<<<ceil_div(nw, 128), 128, 0, c.stream>>>(keys, nw, k, sel);
  gather_kernel<T><<<k, 256, 0, c.stream>>>(keys, sel, dV, ldv, vectors_in_rows, d, k, packed_n, dtotal_var,
                                            dmean_d, whiten, dQ, dw, dev, devr, dmean);
  c.launches += 3;
  DR4G_CUDA(cudaGetLastError());
}

template void launch_postproc<float>(Ctx&, const double*, int, const double*, int, bool, bool, int, int, const double*,
                                     const double*, const double*, float*, float*, float*, float*, float*, bool);
template void launch_postproc<double>(Ctx&, const double*, int, const double*, int, bool, bool, int, int, const double*,
                                      const double*, const double*, double*, double*, double*, double*, double*, bool);

}  // namespace dr4g
