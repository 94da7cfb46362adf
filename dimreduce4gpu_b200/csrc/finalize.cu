// K3 - d-sized bookkeeping between the streamed pass and the eigensolve.
//
//  * pilot shift c = column means of a row sample (keeps the rank-1 centring correction
//    G_c - s_c s_c^T / n free of catastrophic cancellation, SURVEY 7.3 hard part 5);
//  * centring: C = G_c - s_c s_c^T / n  (replaces the reference's materialised 1*mu^T and X - 1*mu^T,
//    src/pca.cu:46-49), mean = c + s_c / n (src/pca.cu:37-43), and the total variance
//    sum_j (q_j - s_j^2/n) / max(1, n-1) that the reference gets from two more sweeps over X
//    (src/tsvd.cu:274-290; definition follows the CPU backend, src/cpu_backend.cpp:283-296).
// Packed layout (doubles): [G d*d | s d | q d | n 1 | c d] - the buffer that is exchanged across ranks.  Every rank runs
// its pass with its OWN pilot shift c_r (no collective before the pass); the exchange is ONE all-gather of the packed
// buffers and rebase_sum_* re-bases every rank's moments to the common shift c = sum_r n_r c_r / sum_r n_r:
//   (x - c) = (x - c_r) + delta_r, delta_r = c_r - c
//   G  = sum_r [ G_r + s_r delta_r^T + delta_r s_r^T + n_r delta_r delta_r^T ]
//   s  = sum_r [ s_r + n_r delta_r ],   q = sum_r [ q_r + 2 delta_r o s_r + n_r delta_r^2 ],   n = sum_r n_r
// exact algebra with a tiny correction (the pilots differ by ~sigma / sqrt(pilot_rows)), summed in rank order on every
// rank, so all ranks hold bit-identical results.
#include "common.cuh"

namespace dr4g {
namespace {

__global__ void pilot_finish_kernel(const double* __restrict__ sum, const double* __restrict__ count_or_null,
                                    double n_rows, int d, float* __restrict__ shift_f,
                                    double* __restrict__ shift_d) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= d) return;
  const double cnt = count_or_null ? *count_or_null : n_rows;
  const double m = cnt > 0 ? sum[j] / cnt : 0.0;
  // the shift is rounded to fp32 so that the fp32 kernels and the fp64 bookkeeping use the
  // exact same number
  const float mf = (float)m;
  if (shift_f) shift_f[j] = mf;
  if (shift_d) shift_d[j] = shift_f ? (double)mf : m;
}

__global__ void set_scalar_kernel(double* dst, double v) { *dst = v; }

__global__ void center_kernel(const double* __restrict__ packed, int d, int center,
                              const double* __restrict__ shift_d, double* __restrict__ C) {
  const double* G = packed;
  const double* s = packed + (size_t)d * d;
  const double n = packed[(size_t)d * d + 2 * d];
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= d) return;
  double v = G[(size_t)r * d + c];
  if (center && n > 0) v -= s[r] * s[c] / n;
  C[(size_t)r * d + c] = v;
}

__global__ void moments_kernel(const double* __restrict__ packed, int d, const double* __restrict__ shift_d,
                               double* __restrict__ mean, double* __restrict__ total_var) {
  const double* s = packed + (size_t)d * d;
  const double* q = s + d;
  const double n = packed[(size_t)d * d + 2 * d];
  __shared__ double red[256];
  double acc = 0;
  for (int j = threadIdx.x; j < d; j += blockDim.x) {
    const double sj = s[j];
    if (mean) mean[j] = (shift_d ? shift_d[j] : 0.0) + (n > 0 ? sj / n : 0.0);
    double ss = q[j] - (n > 0 ? sj * sj / n : 0.0);  // sum of squared deviations about the mean
    if (ss < 0) ss = 0;
    acc += ss;
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double denom = n - 1 > 1 ? n - 1 : 1;  // max(1, n-1), src/cpu_backend.cpp:279
    double tv = red[0] / denom;
    if (!(tv > 0)) tv = 1.0;                      // src/cpu_backend.cpp:296
    *total_var = tv;
  }
}

// gathered: world x ext doubles, ext = d*d + 3d + 1.  out: same layout, holds the common shift and the re-based s, q, n.
__global__ void rebase_sum_vec_kernel(const double* __restrict__ gathered, int world, int d, double* __restrict__ out) {
  const size_t dd = (size_t)d * d, ext = dd + 3 * (size_t)d + 1;
  double ntot = 0.0;
  for (int r = 0; r < world; ++r) ntot += gathered[r * ext + dd + 2 * d];
  for (int j = threadIdx.x + blockIdx.x * blockDim.x; j < d; j += blockDim.x * gridDim.x) {
    double c = 0.0;
    for (int r = 0; r < world; ++r) c += gathered[r * ext + dd + 2 * d] * gathered[r * ext + dd + 2 * d + 1 + j];
    c = ntot > 0 ? c / ntot : 0.0;
    double s = 0.0, q = 0.0;
    for (int r = 0; r < world; ++r) {
      const double* g = gathered + r * ext + dd;
      const double nr = g[2 * d], dl = g[2 * d + 1 + j] - c, sr = g[j];
      s += sr + nr * dl;
      q += g[d + j] + 2.0 * dl * sr + nr * dl * dl;
    }
    out[dd + j] = s;
    out[dd + d + j] = q;
    out[dd + 2 * d + 1 + j] = c;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) out[dd + 2 * d] = ntot;
}

__global__ void rebase_sum_gram_kernel(const double* __restrict__ gathered, int world, int d, double* __restrict__ out) {
  const size_t dd = (size_t)d * d, ext = dd + 3 * (size_t)d + 1;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y;
  if (j >= d) return;
  const double ci = out[dd + 2 * d + 1 + i], cj = out[dd + 2 * d + 1 + j];   // common shift (rebase_sum_vec_kernel ran first)
  double acc = 0.0;
  for (int r = 0; r < world; ++r) {
    const double* g = gathered + r * ext;
    const double* v = g + dd;
    const double nr = v[2 * d], di = v[2 * d + 1 + i] - ci, dj = v[2 * d + 1 + j] - cj;
    acc += g[(size_t)i * d + j] + v[i] * dj + di * v[j] + nr * di * dj;
  }
  out[(size_t)i * d + j] = acc;
}

__global__ void pad_copy_kernel(const double* __restrict__ src, int d, double* __restrict__ dst, int dp) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  const int r = blockIdx.y;
  if (c >= dp) return;
  dst[(size_t)r * dp + c] = (r < d && c < d) ? src[(size_t)r * d + c] : 0.0;
}

}  // namespace

void launch_pilot_finish(Ctx& c, const double* dsum_and_count, int d, float* shift_f, double* shift_d) {
  pilot_finish_kernel<<<ceil_div(d, 128), 128, 0, c.stream>>>(dsum_and_count, dsum_and_count + d, 0.0, d,
                                                               shift_f, shift_d);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

__global__ void add_f64_kernel(double* __restrict__ acc, const double* __restrict__ x, size_t count) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) acc[i] += x[i];
}

void launch_add_f64(Ctx& c, double* acc, const double* x, size_t count) {
  add_f64_kernel<<<(unsigned)((count + 255) / 256), 256, 0, c.stream>>>(acc, x, count);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

void launch_set_scalar(Ctx& c, double* dst, double value) {
  set_scalar_kernel<<<1, 1, 0, c.stream>>>(dst, value);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

void launch_center(Ctx& c, const double* packed, int d, int center, const double* shift_d, double* dC,
                   double* dmean, double* dtotal_var) {
  dim3 grid(ceil_div(d, 128), d);
  center_kernel<<<grid, 128, 0, c.stream>>>(packed, d, center, shift_d, dC);
  moments_kernel<<<1, 256, 0, c.stream>>>(packed, d, shift_d, dmean, dtotal_var);
  c.launches += 2;
  DR4G_CUDA(cudaGetLastError());
}

void launch_rebase_sum(Ctx& c, const double* gathered, int world, int d, double* out) {
  rebase_sum_vec_kernel<<<ceil_div(d, 256), 256, 0, c.stream>>>(gathered, world, d, out);
  dim3 grid(ceil_div(d, 128), d);
  rebase_sum_gram_kernel<<<grid, 128, 0, c.stream>>>(gathered, world, d, out);
  c.launches += 2;
  DR4G_CUDA(cudaGetLastError());
}

// rows x d -> rows x dp (dp >= d, zero in the pad columns): the chunk copy that gives the TMA / cp.async kernels the
// 16-byte row pitch and base they need when d % 4 != 0 (fp32) or d is odd (fp64) or the caller's pointer is unaligned
template <typename T>
__global__ void pad_rows_kernel(const T* __restrict__ src, long long rows, int d, int dp, T* __restrict__ dst) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= rows * dp) return;
  const long long r = e / dp;
  const int cidx = (int)(e - r * dp);
  dst[e] = cidx < d ? src[r * d + cidx] : T(0);
}

template <typename T>
void launch_pad_rows(Ctx& c, const T* src, long long rows, int d, int dp, T* dst) {
  const long long total = rows * dp;
  pad_rows_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, c.stream>>>(src, rows, d, dp, dst);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}
template void launch_pad_rows<float>(Ctx&, const float*, long long, int, int, float*);
template void launch_pad_rows<double>(Ctx&, const double*, long long, int, int, double*);

// [G_p (dp x dp) | s_p | q_p] of the padded problem -> G (d x d), s, q
__global__ void unpad_moments_kernel(const double* __restrict__ Gp, const double* __restrict__ sp, const double* __restrict__ qp,
                                     int d, int dp, double* __restrict__ G, double* __restrict__ s, double* __restrict__ q) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j >= d) return;
  G[(size_t)i * d + j] = Gp[(size_t)i * dp + j];
  if (i == 0) { s[j] = sp[j]; q[j] = qp[j]; }
}

void launch_unpad_moments(Ctx& c, const double* Gp, const double* sp, const double* qp, int d, int dp, double* G, double* s,
                          double* q) {
  dim3 grid(ceil_div(d, 128), d);
  unpad_moments_kernel<<<grid, 128, 0, c.stream>>>(Gp, sp, qp, d, dp, G, s, q);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

void launch_pad_copy(Ctx& c, const double* src, int d, double* dst, int dp) {
  dim3 grid(ceil_div(dp, 128), dp);
  pad_copy_kernel<<<grid, 128, 0, c.stream>>>(src, d, dst, dp);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------- row folding (narrow X, api.cu)
// f consecutive rows of an n x d matrix read as ONE row of f d columns: the streamed kernels then see the 256-column
// problem they are built for, whatever d is (their time per row does not depend on d <= 128, so d = 8 ran at 0.2 TB/s).
// The Gram of the folded matrix holds G in its f diagonal d x d blocks (the other blocks are products of different
// rows: ignored), the folded column statistics hold s and q in f pieces.
template <typename T>
__global__ void tile_shift_kernel(const T* __restrict__ shift, int d, int f, T* __restrict__ out) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < f * d) out[e] = shift[e % d];
}

// G = sum_g Gf[g d + i][g d + j] (+ Gt: the rows that did not fill a folded row), s, q likewise; fixed summation order
__global__ void unfold_moments_kernel(const double* __restrict__ Gf, const double* __restrict__ sf, const double* __restrict__ qf,
                                      int d, int f, const double* __restrict__ Gt, const double* __restrict__ st,
                                      const double* __restrict__ qt, double* __restrict__ G, double* __restrict__ s,
                                      double* __restrict__ q) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j >= d) return;
  const size_t df = (size_t)f * d;
  double a = Gt ? Gt[(size_t)i * d + j] : 0.0;
  for (int g = 0; g < f; ++g) a += Gf[((size_t)g * d + i) * df + (size_t)g * d + j];
  G[(size_t)i * d + j] = a;
  if (i == 0 && sf != nullptr) {   // (the fp64 fold takes s and q from a column-statistics pass over the unfolded rows)
    double b = st ? st[j] : 0.0, cq = qt ? qt[j] : 0.0;
    for (int g = 0; g < f; ++g) { b += sf[(size_t)g * d + j]; cq += qf[(size_t)g * d + j]; }
    s[j] = b;
    q[j] = cq;
  }
}

void launch_tile_shift(Ctx& c, const float* shift, int d, int f, float* out) {
  tile_shift_kernel<float><<<ceil_div(f * d, 256), 256, 0, c.stream>>>(shift, d, f, out);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}
void launch_tile_shift(Ctx& c, const double* shift, int d, int f, double* out) {
  tile_shift_kernel<double><<<ceil_div(f * d, 256), 256, 0, c.stream>>>(shift, d, f, out);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

void launch_unfold_moments(Ctx& c, const double* Gf, const double* sf, const double* qf, int d, int f, const double* Gt,
                           const double* st, const double* qt, double* G, double* s, double* q) {
  dim3 grid(ceil_div(d, 128), d);
  unfold_moments_kernel<<<grid, 128, 0, c.stream>>>(Gf, sf, qf, d, f, Gt, st, qt, G, s, q);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

}  // namespace dr4g
