// K2 for fp64 input - Gram G = (X - 1 c^T)^T (X - 1 c^T) on the DMMA pipe (mma.sync.m8n8k4.f64; tcgen05 has no f64 kind).
//
// Replaces the reference's cublasDgemm Gram of the double entry points (src/tsvd.cu:324,358 via matrix::multiply,
// src/matrix.cu:95-127, reached from pca_double / truncated_svd_double, src/pca.cu:76-116, src/tsvd.cu:440-452).
// This is BASELINE config 5's kernel (PCA k=100, 20M x 4096 fp64 over 8 GPUs): arithmetic intensity d/4 = 1024 flop/B,
// bound by the fp64 pipe (64 FMA/clk/SM, tools/fp64_probe.cu), not by HBM.
//
// Tiling.  Output tiles of 128 x 128 over the upper block triangle (tile pairs bi <= bj); a CTA of 8 warps (2 x 4, warp
// tile 64 x 32 = 8 x 4 m8n8 tiles, 64 accumulator doubles per lane) owns one tile at a time and keeps it in registers
// for ALL the rows it is given.  X is row-major, so both operands are [row][column] strips of 16 rows x 128 columns,
// copied global -> shared with 16-byte cp.async (L2 only) through a 4-stage ring; the pitch of 132 doubles makes the
// fragment loads (4 rows x 8 columns per instruction half) conflict-free.  The pilot shift c (PCA) is subtracted when the
// fragments are read (12 DADD per 32 DMMA), so rows past the end are masked in the last stage only.
//
// Schedule (persistent, one CTA per SM, no atomics, fixed summation order).  With P CTAs and `pairs` tile pairs:
//   * floor(pairs / P) full rounds: CTA c takes pair r P + c over all n rows and stores the finished tile straight into
//     G.  All CTAs sweep the same rows at the same time, so a round reads X from HBM once (the strips are shared
//     through L2): d = 4096 -> 528 pairs = 3.57 rounds -> X is streamed ~4.6 times, 0.06 s of HBM time beside ~1.2 s
//     of fp64 work per 2.5M-row shard.
//   * the remaining pairs are split s ways over the rows so that rem * s fills the CTAs to within a few per cent
//     (d = 4096: 84 pairs x 7 splits = 588 units = 3.97 per CTA); their partial tiles go to slabs that
//     gram_f64_reduce_kernel sums in split order.  Small d (few pairs) is all "remaining".
// The lower block triangle is filled by launch_mirror_upper (gram_generic.cu).
#include "common.cuh"

namespace dr4g {
namespace {

constexpr int FT = 128;          // output tile edge
constexpr int FK = 16;           // rows per stage
constexpr int FPITCH = 132;      // doubles per shared row: 132 % 16 == 4 -> (4 rows x 4 columns) of a half-warp hit 16 distinct 8-byte banks
constexpr int FSTAGES = 4;
constexpr int FTHREADS = 256;
constexpr int kStripDoubles = FK * FPITCH;
constexpr size_t kF64Smem = (size_t)FSTAGES * 2 * kStripDoubles * sizeof(double);   // 135,168 B
constexpr int kMaxSplit = 64;

struct F64Plan {
  int nblk, pairs, P, full_rounds, rem, s;
  long long rows_per_split;
};

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

// 16-byte global -> shared copy through L2 only; src_bytes = 0 zero-fills (rows past n, columns past d)
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(FTHREADS, 1)
gram_f64_kernel(const double* __restrict__ X, long long n, int d, const double* __restrict__ shift, double* __restrict__ G,
                double* __restrict__ slab, F64Plan plan) {
  extern __shared__ __align__(16) double fsm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp >> 2, wn = warp & 3;           // 2 x 4 warps, warp tile 64 x 32
  const int cc = lane >> 2, kl = lane & 3;
  const uint32_t sm_base = static_cast<uint32_t>(__cvta_generic_to_shared(fsm));

  // loader mapping: a strip row is 128 doubles = 64 chunks of 16 B; 256 threads cover 4 rows per pass, 4 passes per strip
  const int lchunk = tid & 63, lrow = tid >> 6;

  const int tail_units = plan.rem * plan.s;   // this CTA takes tail units blockIdx.x, blockIdx.x + P, ...
  const int my_tail = tail_units > (int)blockIdx.x ? (tail_units - (int)blockIdx.x + plan.P - 1) / plan.P : 0;
  const int units_total = plan.full_rounds + my_tail;
  for (int u = 0; u < units_total; ++u) {
    // ---- which tile pair, which rows
    int pair;
    long long r0 = 0, r1 = n;
    int tail_unit = -1;
    if (u < plan.full_rounds) {
      pair = u * plan.P + (int)blockIdx.x;
    } else {
      tail_unit = (u - plan.full_rounds) * plan.P + (int)blockIdx.x;
      pair = plan.full_rounds * plan.P + tail_unit / plan.s;
      const int split = tail_unit % plan.s;
      r0 = (long long)split * plan.rows_per_split;
      r1 = r0 + plan.rows_per_split;
      if (r1 > n) r1 = n;
      if (r0 > n) r0 = n;
    }
    int t = pair, bi = 0;
    while (t >= plan.nblk - bi) { t -= plan.nblk - bi; ++bi; }
    const int bj = bi + t;
    const bool diag = bi == bj;

    // shifts of this lane's fragment columns (0 for columns past d: their operands are zero-filled)
    double sha[8], shb[4];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int col = bi * FT + wm * 64 + i * 8 + cc;
      sha[i] = (shift && col < d) ? shift[col] : 0.0;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int col = bj * FT + wn * 32 + j * 8 + cc;
      shb[j] = (shift && col < d) ? shift[col] : 0.0;
    }

    double acc[8][4][2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    const long long nst = (r1 - r0 + FK - 1) / FK;   // stages of this unit
    const int ca = bi * FT + 2 * lchunk, cb = bj * FT + 2 * lchunk;   // d is even: a 16-byte chunk is inside or outside as a whole
    auto issue = [&](long long st) {
      if (st < nst) {
        const int slot = (int)(st % FSTAGES);
        const uint32_t sa = sm_base + (uint32_t)(slot * 2 * kStripDoubles) * 8u;
        const uint32_t sb = sa + (uint32_t)kStripDoubles * 8u;
        const long long rb = r0 + st * FK;
#pragma unroll
        for (int p = 0; p < FK / 4; ++p) {
          const int rr = lrow + 4 * p;
          const long long r = rb + rr;
          const bool rv = r < r1;
          const long long rc = rv ? r : 0;
          const uint32_t off = (uint32_t)(rr * FPITCH + 2 * lchunk) * 8u;
          cp_async16(sa + off, X + rc * d + (ca < d ? ca : 0), (rv && ca < d) ? 16 : 0);
          if (!diag) cp_async16(sb + off, X + rc * d + (cb < d ? cb : 0), (rv && cb < d) ? 16 : 0);
        }
      }
      cp_async_commit();   // one group per stage, possibly empty: keeps the wait count uniform
    };

#pragma unroll
    for (int st = 0; st < FSTAGES - 1; ++st) issue(st);

    for (long long st = 0; st < nst; ++st) {
      cp_async_wait<FSTAGES - 2>();
      __syncthreads();                 // stage st has landed for everyone; everyone is done with stage st - 1
      issue(st + FSTAGES - 1);         // refills the slot of stage st - 1
      const int slot = (int)(st % FSTAGES);
      const double* As = fsm + (size_t)slot * 2 * kStripDoubles;
      const double* Bs = diag ? As : As + kStripDoubles;
      const int valid = (int)((r1 - r0 - st * FK) < FK ? (r1 - r0 - st * FK) : FK);
      const bool tail = valid < FK;
#pragma unroll
      for (int k0 = 0; k0 < FK; k0 += 4) {
        const int kr = k0 + kl;
        const double* ar = As + kr * FPITCH + wm * 64 + cc;
        const double* br = Bs + kr * FPITCH + wn * 32 + cc;
        double a[8], b[4];
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = ar[i * 8] - sha[i];
#pragma unroll
        for (int j = 0; j < 4; ++j) b[j] = br[j * 8] - shb[j];
        if (tail && kr >= valid) {     // rows past the end of the unit were zero-filled: 0 - shift must not count
#pragma unroll
          for (int i = 0; i < 8; ++i) a[i] = 0.0;
#pragma unroll
          for (int j = 0; j < 4; ++j) b[j] = 0.0;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
      }
    }
    cp_async_wait<0>();
    __syncthreads();                   // the ring is idle before the next unit refills it

    // ---- epilogue: C fragment row = lane / 4, columns 2 (lane % 4) + {0, 1}
    if (tail_unit < 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int gr = bi * FT + wm * 64 + i * 8 + cc;
        if (gr >= d) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int gc = bj * FT + wn * 32 + j * 8 + 2 * kl;
          if (gc + 1 < d) *reinterpret_cast<double2*>(&G[(size_t)gr * d + gc]) = make_double2(acc[i][j][0], acc[i][j][1]);
          else if (gc < d) G[(size_t)gr * d + gc] = acc[i][j][0];
        }
      }
    } else {
      double* S = slab + (size_t)tail_unit * FT * FT;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int lr = wm * 64 + i * 8 + cc;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int lc = wn * 32 + j * 8 + 2 * kl;
          *reinterpret_cast<double2*>(&S[lr * FT + lc]) = make_double2(acc[i][j][0], acc[i][j][1]);
        }
      }
    }
  }
}

// G tile <- sum of the tile's s slabs, in split order.  grid = (rem, 64), block = 256: 2 x 128-element rows per block row.
__global__ void __launch_bounds__(256)
gram_f64_reduce_kernel(const double* __restrict__ slab, int d, double* __restrict__ G, F64Plan plan) {
  const int pair = plan.full_rounds * plan.P + (int)blockIdx.x;
  int t = pair, bi = 0;
  while (t >= plan.nblk - bi) { t -= plan.nblk - bi; ++bi; }
  const int bj = bi + t;
  const int lr = (int)blockIdx.y * 2 + ((int)threadIdx.x >> 7), lc = (int)threadIdx.x & 127;
  const int gr = bi * FT + lr, gc = bj * FT + lc;
  if (gr >= d || gc >= d) return;
  const double* p = slab + ((size_t)blockIdx.x * plan.s) * FT * FT + (size_t)lr * FT + lc;
  double acc = 0.0;
  for (int sidx = 0; sidx < plan.s; ++sidx) acc += p[(size_t)sidx * FT * FT];
  G[(size_t)gr * d + gc] = acc;
}

F64Plan make_plan(const Ctx& c, long long n, int d) {
  F64Plan p{};
  p.nblk = (d + FT - 1) / FT;
  p.pairs = p.nblk * (p.nblk + 1) / 2;
  p.P = c.sm_count;
  p.full_rounds = p.pairs / p.P;
  p.rem = p.pairs % p.P;
  p.s = 1;
  p.rows_per_split = n;
  if (p.rem > 0) {
    // split the remaining pairs over the rows: minimise ceil(rem s / P) / s (the length of the tail in passes);
    // a split must keep at least 4 stages of rows
    long long smax = n / (4 * FK);
    if (smax < 1) smax = 1;
    if (smax > kMaxSplit) smax = kMaxSplit;
    double best = 1e30;
    for (int s = 1; s <= (int)smax; ++s) {
      const double tlen = (double)((p.rem * s + p.P - 1) / p.P) / s + 1e-4 * s;   // slight preference for fewer splits
      if (tlen < best) { best = tlen; p.s = s; }
    }
    long long rps = (n + p.s - 1) / p.s;
    rps = (rps + FK - 1) / FK * FK;
    p.rows_per_split = rps;
  }
  return p;
}

}  // namespace

bool gram_f64_eligible(const Ctx& c, const double* dX, long long n, int d) {
  return n >= 1 && d >= 2 && (d % 2) == 0 && (reinterpret_cast<uintptr_t>(dX) & 15) == 0 && c.sm_count >= 1;
}

size_t gram_f64_workspace_bytes(const Ctx& c, long long n, int d) {
  const F64Plan p = make_plan(c, n, d);
  return (size_t)p.rem * p.s * FT * FT * sizeof(double) + 256;
}

// Upper block triangle (by 128) of G; the caller mirrors it (launch_mirror_upper).  dG needs no zeroing for the tiles
// written here, but the caller's memset keeps the lower triangle defined until the mirror runs.
void launch_gram_f64(Ctx& c, const double* dX, long long n, int d, const double* dshift, double* dG, void* workspace) {
  const F64Plan plan = make_plan(c, n, d);
  static PerDeviceMax smem_set;
  if (smem_set.need(c.device, kF64Smem))
    DR4G_CUDA(cudaFuncSetAttribute(gram_f64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kF64Smem));
  int grid = plan.P;
  if (plan.full_rounds == 0 && plan.rem * plan.s < grid) grid = plan.rem * plan.s;
  gram_f64_kernel<<<grid, FTHREADS, kF64Smem, c.stream>>>(dX, n, d, dshift, dG, static_cast<double*>(workspace), plan);
  c.launches++;
  if (plan.rem > 0) {
    gram_f64_reduce_kernel<<<dim3(plan.rem, FT / 2), 256, 0, c.stream>>>(static_cast<const double*>(workspace), d, dG, plan);
    c.launches++;
  }
  DR4G_CUDA(cudaGetLastError());
}

}  // namespace dr4g
