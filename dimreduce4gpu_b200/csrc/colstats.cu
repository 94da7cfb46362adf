// K1 - fused column sum / sum-of-squares reduction.
//
// Replaces the reference's column-mean GEMV-as-GEMM (src/pca.cu:37-42, matrix::multiply
// src/matrix.cu:95-127) and the two extra variance sweeps of get_tsvd_attr (src/tsvd.cu:274-290):
// one read of X gives sum_r (x - shift) and sum_r (x - shift)^2 per column.
//
// Layout: X is row-major n x d.  A warp covers 4 rows x 32 columns per step: lane = 8*rsub + cg,
// each lane loads one 128-bit vector (4 columns), so every row segment is a full 128-byte line.
// Per-thread fp32 partials are spilled into fp64 every kSpill steps; the 4 row sub-lanes are folded
// with warp shuffles, the warps of a CTA through shared memory, CTAs with fp64 atomics.
// HBM-bound: algorithmic bytes = n*d*sizeof(T).
#include "common.cuh"

namespace dr4g {

namespace {

constexpr int kWarps = 8;
constexpr int kSpill = 64;

template <typename T>
struct Vec4;
template <>
struct Vec4<float> {
  using type = float4;
};
template <>
struct Vec4<double> {
  using type = double4;
};

__device__ __forceinline__ float4 ld_stream(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}

// vectorised path: d % 4 == 0 and 16-byte aligned rows (fp32)
__global__ void __launch_bounds__(kWarps * 32)
colstats_f32_vec(const float* __restrict__ X, long long n, int d, const float* __restrict__ shift,
                 double* __restrict__ osum, double* __restrict__ osq, int rows_per_cta) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cg = lane & 7, rsub = lane >> 3;
  const int col = blockIdx.x * 32 + cg * 4;
  const long long r0 = (long long)blockIdx.y * rows_per_cta;
  long long r1 = r0 + rows_per_cta;
  if (r1 > n) r1 = n;
  const bool active = col < d;  // d % 4 == 0 -> whole vector in range

  float4 sh = make_float4(0.f, 0.f, 0.f, 0.f);
  if (active && shift) sh = *reinterpret_cast<const float4*>(shift + col);

  double ds[4] = {0, 0, 0, 0}, dq[4] = {0, 0, 0, 0};
  float fs[4] = {0, 0, 0, 0}, fq[4] = {0, 0, 0, 0};
  int steps = 0;
  if (active) {
    const long long stride = (long long)kWarps * 4;
    long long r = r0 + warp * 4 + rsub;
    // 4 independent loads in flight per thread
    for (; r + 3 * stride < r1; r += 4 * stride) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        v[u] = ld_stream(reinterpret_cast<const float4*>(X + (r + u * stride) * d + col));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float a = v[u].x - sh.x, b = v[u].y - sh.y, c2 = v[u].z - sh.z, e = v[u].w - sh.w;
        fs[0] += a; fs[1] += b; fs[2] += c2; fs[3] += e;
        fq[0] = fmaf(a, a, fq[0]); fq[1] = fmaf(b, b, fq[1]);
        fq[2] = fmaf(c2, c2, fq[2]); fq[3] = fmaf(e, e, fq[3]);
      }
      if (++steps == kSpill) {
        steps = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) { ds[j] += fs[j]; dq[j] += fq[j]; fs[j] = 0.f; fq[j] = 0.f; }
      }
    }
    for (; r < r1; r += stride) {
      float4 v = ld_stream(reinterpret_cast<const float4*>(X + r * d + col));
      float a = v.x - sh.x, b = v.y - sh.y, c2 = v.z - sh.z, e = v.w - sh.w;
      fs[0] += a; fs[1] += b; fs[2] += c2; fs[3] += e;
      fq[0] = fmaf(a, a, fq[0]); fq[1] = fmaf(b, b, fq[1]);
      fq[2] = fmaf(c2, c2, fq[2]); fq[3] = fmaf(e, e, fq[3]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) { ds[j] += fs[j]; dq[j] += fq[j]; }
  }
  // fold the 4 row sub-lanes (lane bits 3,4)
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    ds[j] += __shfl_xor_sync(0xffffffffu, ds[j], 8);
    ds[j] += __shfl_xor_sync(0xffffffffu, ds[j], 16);
    dq[j] += __shfl_xor_sync(0xffffffffu, dq[j], 8);
    dq[j] += __shfl_xor_sync(0xffffffffu, dq[j], 16);
  }
  __shared__ double sm[kWarps][2][32];
  if (rsub == 0) {
#pragma unroll
    for (int j = 0; j < 4; ++j) { sm[warp][0][cg * 4 + j] = ds[j]; sm[warp][1][cg * 4 + j] = dq[j]; }
  }
  __syncthreads();
  if (threadIdx.x < 64) {
    const int which = threadIdx.x >> 5, c = threadIdx.x & 31;
    double t = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) t += sm[w][which][c];
    const int gc = blockIdx.x * 32 + c;
    if (gc < d) atomicAdd((which ? osq : osum) + gc, t);
  }
}

// generic path: any d / alignment / dtype; lane = column within a 32-column strip, warps stride rows
template <typename T>
__global__ void __launch_bounds__(kWarps * 32)
colstats_generic(const T* __restrict__ X, long long n, int d, const T* __restrict__ shift,
                 double* __restrict__ osum, double* __restrict__ osq, int rows_per_cta) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + lane;
  const long long r0 = (long long)blockIdx.y * rows_per_cta;
  long long r1 = r0 + rows_per_cta;
  if (r1 > n) r1 = n;
  double s = 0, q = 0;
  if (col < d) {
    const double sh = shift ? (double)shift[col] : 0.0;
    for (long long r = r0 + warp; r < r1; r += kWarps) {
      double v = (double)X[r * d + col] - sh;
      s += v;
      q = fma(v, v, q);
    }
  }
  __shared__ double sm[kWarps][2][32];
  sm[warp][0][lane] = s;
  sm[warp][1][lane] = q;
  __syncthreads();
  if (threadIdx.x < 64) {
    const int which = threadIdx.x >> 5, c = threadIdx.x & 31;
    double t = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) t += sm[w][which][c];
    const int gc = blockIdx.x * 32 + c;
    if (gc < d) atomicAdd((which ? osq : osum) + gc, t);
  }
}

}  // namespace

template <typename T>
void launch_colstats(Ctx& c, const T* dX, long long n, int d, const T* dshift, double* dsum, double* dsumsq) {
  if (n <= 0 || d <= 0) return;
  const int strips = ceil_div(d, 32);
  // ~4 CTAs per SM overall, at least 64 rows each
  long long want = ((long long)c.sm_count * 4 + strips - 1) / strips;
  long long max_by_rows = (n + 63) / 64;
  long long ny = want < max_by_rows ? want : max_by_rows;
  if (ny < 1) ny = 1;
  if (ny > 65535) ny = 65535;
  int rows_per_cta = (int)((n + ny - 1) / ny);
  rows_per_cta = (rows_per_cta + 31) / 32 * 32;
  ny = (n + rows_per_cta - 1) / rows_per_cta;
  dim3 grid(strips, (unsigned)ny);
  bool vec = false;
  if constexpr (sizeof(T) == 4) {
    vec = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(dX) & 15) == 0) &&
          (!dshift || (reinterpret_cast<uintptr_t>(dshift) & 15) == 0);
    if (vec) {
      colstats_f32_vec<<<grid, kWarps * 32, 0, c.stream>>>(dX, n, d, dshift, dsum, dsumsq, rows_per_cta);
    }
  }
  if (!vec) colstats_generic<T><<<grid, kWarps * 32, 0, c.stream>>>(dX, n, d, dshift, dsum, dsumsq, rows_per_cta);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

template void launch_colstats<float>(Ctx&, const float*, long long, int, const float*, double*, double*);
template void launch_colstats<double>(Ctx&, const double*, long long, int, const double*, double*, double*);

}  // namespace dr4g
