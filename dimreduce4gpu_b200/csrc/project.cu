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


// ---------------------------------------------------------------- fp32 projection, k <= 64
// Z = (X - mean) Q^T, U = Z / sigma.  The generic kernel above reached ~0.6 TB/s (scalar shared-memory
// reads, synchronous staging, half of its 64 output columns idle at k = 32).  This one is a register-
// tiled SIMT GEMM: 256 threads, RT = 8192 / KP rows per CTA (256 rows at KP = 32, 128 at KP = 64), thread
// tile 8 rows x 4 outputs, contraction in chunks of 32 columns.  Shared tiles are stored contraction-
// major (As[t][row], Bs[t][m]) so that a thread's 8 rows / 4 outputs are 128-bit loads; X is read with
// coalesced float4 (a lane = row mapping ran at half the speed: 32 lines per load instruction) and
// transposed into a row-swizzled tile with conflict-free scalar stores, the next chunk's
// global loads are in flight while the current one is multiplied (register prefetch, two smem buffers,
// one barrier per chunk).  AI = k/2 flop/B: at k = 32 the fp32 FMA pipe (~74 TF/s) and HBM need about
// the same time, so this is the CUDA-core ceiling; a tensor-core version would need 3xTF32 again.
template <int KP>
__global__ void __launch_bounds__(256, 2)
project_f32_kernel(const float* __restrict__ X, long long n, int d, int k, const float* __restrict__ mean,
                   const float* __restrict__ Q, const float* __restrict__ sigma, float* __restrict__ Z,
                   float* __restrict__ U) {
  constexpr int RT = 8192 / KP;        // rows per CTA
  constexpr int TC = 32;               // contraction chunk
  constexpr int XF = RT * TC / 256 / 4;  // float4 of X per thread and chunk (8 at KP = 32, 4 at KP = 64)
  extern __shared__ __align__(16) float psm[];
  float* As = psm;                       // [2][TC][RT]
  float* Bs = psm + 2 * TC * RT;         // [2][TC][KP]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tx = tid % (KP / 4), ty = tid / (KP / 4);
  const long long row0 = (long long)blockIdx.x * RT;

  // staging: a warp instruction reads 4 rows x 128 B (lane = (row r4 = lane / 8, float4 c4 = lane % 8)), the
  // values go to As[t][row] transposed; rows are XOR-swizzled in groups of 4 by (t / 4) % 8 so that these
  // scalar stores hit 32 distinct banks while a thread's 8 rows stay two aligned float4
  constexpr int WR = RT / 8;             // rows per warp
  const int c4 = lane & 7, r4 = lane >> 3;
  const int srow0 = warp * WR + r4;      // + 4 j
  const int swz = c4 << 2;               // ((4 c4 + e) / 4) % 8 = c4
  // Q chunk: thread (m = tid % KP, quarter q = tid / KP) loads TC * KP / 256 floats
  constexpr int BF = TC * KP / 256;      // 4 at KP = 32, 8 at KP = 64
  const int bm = tid % KP, bt0 = (tid / KP) * BF;
  const bool full_rows = row0 + RT <= n;

  float4 xa[XF], mu4;
  float bq[BF];
  auto fetch = [&](int t0) {
    const int t = t0 + 4 * c4;
    const bool tin = t + 3 < d;          // d % 4 == 0: a float4 is inside or outside as a whole
#pragma unroll
    for (int j = 0; j < XF; ++j) {
      const long long r = row0 + srow0 + 4 * j;
      xa[j] = (tin && (full_rows || r < n)) ? __ldg(reinterpret_cast<const float4*>(X + r * (long long)d + t))
                                            : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    mu4 = (mean && tin) ? __ldg(reinterpret_cast<const float4*>(mean + t)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int j = 0; j < BF; ++j) {
      const int tq = t0 + bt0 + j;
      bq[j] = (bm < k && tq < d) ? __ldg(Q + (long long)bm * d + tq) : 0.f;
    }
  };
  auto stash = [&](int buf, int t0) {
    float* Ab = As + buf * TC * RT + (4 * c4) * RT;
    float* Bb = Bs + buf * TC * KP;
    const bool tin = t0 + 4 * c4 + 3 < d;
#pragma unroll
    for (int j = 0; j < XF; ++j) {
      const int rl = srow0 + 4 * j;
      const bool in = tin && (full_rows || row0 + rl < n);  // rows past n and columns past d contribute 0
      const int pr = rl ^ swz;
      Ab[0 * RT + pr] = in ? xa[j].x - mu4.x : 0.f;
      Ab[1 * RT + pr] = in ? xa[j].y - mu4.y : 0.f;
      Ab[2 * RT + pr] = in ? xa[j].z - mu4.z : 0.f;
      Ab[3 * RT + pr] = in ? xa[j].w - mu4.w : 0.f;
    }
#pragma unroll
    for (int j = 0; j < BF; ++j) Bb[(bt0 + j) * KP + bm] = bq[j];
  };

  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  fetch(0);
  stash(0, 0);
  __syncthreads();
  int buf = 0;
  for (int t0 = 0; t0 < d; t0 += TC) {
    const bool more = t0 + TC < d;
    if (more) fetch(t0 + TC);
    const float* Ab = As + buf * TC * RT;
    const float* Bb = Bs + buf * TC * KP + tx * 4;
#pragma unroll 8
    for (int t = 0; t < TC; ++t) {
      const int g = ((t >> 2) & 7) << 2;
      const float4 a0 = *reinterpret_cast<const float4*>(Ab + t * RT + ((ty * 8) ^ g));
      const float4 a1 = *reinterpret_cast<const float4*>(Ab + t * RT + ((ty * 8 + 4) ^ g));
      const float4 b = *reinterpret_cast<const float4*>(Bb + t * KP);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    if (more) stash(buf ^ 1, t0 + TC);
    __syncthreads();
    buf ^= 1;
  }
  // epilogue: 4 consecutive outputs per thread and row
  float inv[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int m = tx * 4 + j;
    const float sg = (U && sigma && m < k) ? __ldg(sigma + m) : 0.f;
    inv[j] = sg != 0.f ? 1.f / sg : 0.f;
  }
  const bool vec = (k % 4) == 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const long long r = row0 + ty * 8 + i;
    if (r >= n) continue;
    const int m0 = tx * 4;
    if (vec && m0 + 3 < k) {
      *reinterpret_cast<float4*>(Z + r * k + m0) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
      if (U) *reinterpret_cast<float4*>(U + r * k + m0) =
          make_float4(acc[i][0] * inv[0], acc[i][1] * inv[1], acc[i][2] * inv[2], acc[i][3] * inv[3]);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (m0 + j < k) {
          Z[r * k + m0 + j] = acc[i][j];
          if (U) U[r * k + m0 + j] = acc[i][j] * inv[j];
        }
    }
  }
}

template <int KP>
void launch_project_f32(Ctx& c, const float* dX, long long n, int d, int k, const float* dQ, const float* dmean,
                        const float* dw, float* dZ, float* dU) {
  constexpr int RT = 8192 / KP;
  const size_t smem = (size_t)2 * 32 * (RT + KP) * sizeof(float);
  static PerDeviceMax smem_set;
  if (smem_set.need(c.device, smem))
    DR4G_CUDA(cudaFuncSetAttribute(project_f32_kernel<KP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long blocks = (n + RT - 1) / RT;
  project_f32_kernel<KP><<<(unsigned)blocks, 256, smem, c.stream>>>(dX, n, d, k, dmean, dQ, (dU && dw) ? dw : nullptr,
                                                                   dZ, (dU && dw) ? dU : nullptr);
}

}  // namespace

static inline size_t up256(size_t b) { return (b + 255) / 256 * 256; }

// ---------------------------------------------------------------- row folding of narrow fp32 inputs
// f consecutive rows of X read as one row of f d columns, projected on the block-diagonal image of the components
// (f k x f d): the scores of the folded row ARE the f score rows of the original ones, contiguous in memory, so Z and U
// need no rearrangement.  The tensor-core kernel's time per row does not depend on d, so d = 8 ran at 0.84 TB/s and
// d = 32 at 3.2; folded to 256 columns all of them stream at 4.8 TB/s (B200, tools/fold_proj_probe.py).
__global__ void fold_components_kernel(const float* __restrict__ Q, const float* __restrict__ mean, const float* __restrict__ w,
                                       int d, int k, int f, float* __restrict__ Qf, float* __restrict__ mf, float* __restrict__ wf) {
  const int fd = f * d, fk = f * k;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < fk * fd; e += gridDim.x * blockDim.x) {
    const int r = e / fd, col = e - r * fd;
    const int g = r / k, j = r - g * k;
    Qf[e] = (col / d == g) ? Q[(size_t)j * d + (col - g * d)] : 0.0f;
  }
  if (blockIdx.x == 0) {
    for (int e = threadIdx.x; e < fd; e += blockDim.x) mf[e] = mean ? mean[e % d] : 0.0f;
    for (int e = threadIdx.x; e < fk; e += blockDim.x) wf[e] = w ? w[e % k] : 0.0f;
  }
}

static int project_fold_factor(const Ctx& c, const void* dX, long long n, int d, int k, const void* dZ, const void* dU) {
  if (!c.fold || c.proj_path != 0 || d < 1 || d > 32 || k < 1 || k > d) return 0;
  if (((reinterpret_cast<uintptr_t>(dX) | reinterpret_cast<uintptr_t>(dZ) | reinterpret_cast<uintptr_t>(dU)) & 15) != 0) return 0;
  if ((double)n * (double)d < 16777216.0) return 0;
  const int step = d % 4 == 0 ? 1 : (d % 2 == 0 ? 2 : 4);
  int f = 256 / d < 128 / k ? 256 / d : 128 / k;
  f = f / step * step;
  if (f < 2 || f * d < 64) return 0;
  return f;
}

template <typename T>
void launch_project(Ctx& c, const T* dX, long long n, int d, int k, const T* dQ, const T* dmean, const T* dw,
                    T* dZ, T* dU) {
  if (n <= 0) return;
  c.last_proj_path = 1;
  if constexpr (sizeof(T) == 4) {
    const int f = project_fold_factor(c, dX, n, d, k, dZ, dU);
    if (d <= 32) c.last_fold = f;
    if (f > 1) {
      const int fd = f * d, fk = f * k;
      if (!c.fold_buf) DR4G_CUDA(cudaMalloc(&c.fold_buf, 128 * 256 * 4 + 256 * 4 + 128 * 4 + 1024));
      float* Qf = static_cast<float*>(c.fold_buf);
      float* mf = Qf + 128 * 256;
      float* wf = mf + 256 + 64;
      fold_components_kernel<<<64, 256, 0, c.stream>>>(dQ, dmean, dw, d, k, f, Qf, mf, wf);
      c.launches++;
      DR4G_CUDA(cudaGetLastError());
      const long long nf = n / f, tail = n - nf * f;
      launch_project<float>(c, dX, nf, fd, fk, Qf, dmean ? mf : nullptr, dw ? wf : nullptr, dZ, dU);
      const int path = c.last_proj_path;
      if (tail > 0)
        launch_project<float>(c, dX + (size_t)nf * fd, tail, d, k, dQ, dmean, dw, dZ + (size_t)nf * fk,
                              dU ? dU + (size_t)nf * fk : nullptr);
      c.last_proj_path = path;
      c.last_fold = f;
      return;
    }
  }
  // Tensor-core / DMMA kernels: at most 128 components per launch - wider k goes through in column slices of the score
  // matrices (row pitch k), X being re-read from HBM once per slice; shapes whose rows have no 16-byte pitch (fp32
  // d % 4 != 0, fp64 odd d, unaligned base) go through zero-padded chunk copies of X (pad_rows_kernel) with the
  // components and the mean padded alike.  Both used to fall to the CUDA-core kernels below (0.16 TB/s at k = 200,
  // 0.55 TB/s at d = 1001).
  // Tiny fp32 inputs (<= 1 MB, like the Gram pass in api.cu) stay on the CUDA-core kernels: plain fp32 FMAs (1e-7 of
  // |Z|) instead of the two-term bf16 split (2e-6), at the same launch-bound time.
  const bool tiny = sizeof(T) == 4 && c.proj_path == 0 && (double)n * (double)d <= 262144.0;
  if (!tiny && c.proj_path != 1 && k >= 1 && d >= 2 && (long long)n * d >= (1ll << 16)) {
    constexpr int q4 = sizeof(T) == 4 ? 4 : 2;
    const bool in_place = (d % q4) == 0 && (reinterpret_cast<uintptr_t>(dX) & 15) == 0 && (reinterpret_cast<uintptr_t>(dQ) & 15) == 0;
    const int dp = (d + q4 - 1) / q4 * q4;
    const T* Xs = dX;
    const T* Qs = dQ;
    const T* ms = dmean;
    long long chunk = n;
    T* Xp = nullptr;
    if (!in_place) {
      chunk = (long long)(((size_t)128 << 20) / ((size_t)dp * sizeof(T)));
      if (chunk < 1024) chunk = 1024;
      if (chunk > n) chunk = n;
      const size_t need = up256((size_t)chunk * dp * sizeof(T)) + up256((size_t)k * dp * sizeof(T)) + up256((size_t)dp * sizeof(T));
      if (need > c.pad_cap) {
        if (c.pad_buf) {
          DR4G_CUDA(cudaStreamSynchronize(c.stream));
          DR4G_CUDA(cudaFree(c.pad_buf));
          c.pad_buf = nullptr;
          c.pad_cap = 0;
        }
        DR4G_CUDA(cudaMalloc(&c.pad_buf, need));
        c.pad_cap = need;
      }
      char* base = static_cast<char*>(c.pad_buf);
      Xp = reinterpret_cast<T*>(base);
      T* Qp = reinterpret_cast<T*>(base + up256((size_t)chunk * dp * sizeof(T)));
      T* mp = reinterpret_cast<T*>(base + up256((size_t)chunk * dp * sizeof(T)) + up256((size_t)k * dp * sizeof(T)));
      launch_pad_rows<T>(c, dQ, k, d, dp, Qp);
      Qs = Qp;
      if (dmean) {
        launch_pad_rows<T>(c, dmean, 1, d, dp, mp);
        ms = mp;
      }
    }
    for (long long r0 = 0; r0 < n; r0 += chunk) {
      const long long rows = n - r0 < chunk ? n - r0 : chunk;
      if (!in_place) {
        launch_pad_rows<T>(c, dX + (size_t)r0 * d, rows, d, dp, Xp);
        Xs = Xp;
      } else {
        Xs = dX + (size_t)r0 * d;
      }
      for (int k0 = 0; k0 < k; k0 += 128) {
        const int kc = k - k0 < 128 ? k - k0 : 128;
        T* zs = dZ + (size_t)r0 * k + k0;
        T* us = (dU && dw) ? dU + (size_t)r0 * k + k0 : nullptr;
        if constexpr (sizeof(T) == 4)
          launch_project_tc(c, Xs, rows, dp, kc, Qs + (size_t)k0 * dp, ms, dw ? dw + k0 : nullptr, zs, us, k);
        else
          launch_project_f64(c, Xs, rows, dp, kc, Qs + (size_t)k0 * dp, ms, dw ? dw + k0 : nullptr, zs, us, k);
      }
    }
    c.last_proj_path = sizeof(T) == 4 ? 2 : 3;
    return;
  }
  if constexpr (sizeof(T) == 4) {
    if (!tiny && c.proj_path != 1 && project_tc_eligible(c, dX, n, d, k, dZ, (dU && dw) ? dU : nullptr)) {
      launch_project_tc(c, dX, n, d, k, dQ, dmean, dw, dZ, dU);
      c.last_proj_path = 2;
      return;
    }
    // 16-byte aligned rows (float4 loads of X, float4 stores of Z when k % 4 == 0)
    const bool aligned = (reinterpret_cast<uintptr_t>(dX) & 15) == 0 && (d % 4) == 0 &&
                         (reinterpret_cast<uintptr_t>(dZ) & 15) == 0 && (!dU || (reinterpret_cast<uintptr_t>(dU) & 15) == 0);
    if (aligned && k <= 64) {
      if (k <= 32) launch_project_f32<32>(c, dX, n, d, k, dQ, dmean, dw, dZ, dU);
      else launch_project_f32<64>(c, dX, n, d, k, dQ, dmean, dw, dZ, dU);
      c.launches++;
      DR4G_CUDA(cudaGetLastError());
      return;
    }
  }
  if constexpr (sizeof(T) == 8) {
    if (c.proj_path != 1 && project_f64_eligible(c, dX, n, d, k, dQ)) {
      launch_project_f64(c, dX, n, d, k, dQ, dmean, dw, dZ, dU);
      c.last_proj_path = 3;
      return;
    }
  }
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
