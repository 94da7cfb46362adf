// K2 - streamed Gram + column statistics on the 5th-generation tensor cores (fp32 input, 3xTF32).
//
//   G = (X - 1 c^T)^T (X - 1 c^T),   s = sum_r (x_r - c),   q = sum_r (x_r - c)^2
//
// Replaces the reference's full cublasSgemm Gram (src/tsvd.cu:324,358 via matrix::multiply,
// src/matrix.cu:95-127), the column-mean GEMV-as-GEMM and the materialised centring of pca_float
// (src/pca.cu:37-49) and the variance sweeps of get_tsvd_attr (src/tsvd.cu:274-290): X is read from
// HBM once per job (see below), by TMA, and never rewritten to global memory.
//
// Shape of the computation.  X is row-major n x d, the contraction runs over rows, so for
// D[MxN] += A[MxK] B[KxN] both operands are "MN-major" views of the same kind of shared tile
// Xs[row][col] (M = N = feature columns, K = rows).  tcgen05 kind::tf32 accepts MN-major for both
// operands (instruction-descriptor bits 15/16).  Columns are cut into blocks of 128; the block upper
// triangle of G is cut into JOBS of up to 4 tiles of 128x128 over up to 4 column blocks (2x2
// "super-block" pairs), so that a job's accumulators fill the SM's 512 TMEM columns and every
// loaded X element feeds all MMAs that need it.  d <= 256 is one job: X is read exactly once.
// A CTA owns one job and every ncta-th chunk of KT rows of it (split over rows, all SMs busy);
// partial tiles go to per-CTA fp32 slabs that a second small kernel sums in fp64 (fixed order).
//
// 3xTF32.  y = x - c is split exactly into hi = y with the 13 low mantissa bits cleared and
// lo = y - hi (also cleared to 11 bits), both written to shared memory by the converter warps, and
// each tile gets hi_a^T hi_b + hi_a^T lo_b + lo_a^T hi_b into one fp32 TMEM accumulator.  Because both
// parts are exact tf32 numbers the result does not depend on whether the tensor core truncates or
// rounds its inputs.  The tensor core's fp32 accumulation TRUNCATES (measured on B200: relative
// bias of G about 1e-8 per accumulated row), so the TMEM accumulators are flushed every
// `flush_rows` (default 1024) rows: stored / RED-added (round-to-nearest, in L2) into the CTA's slab.
// Each flush moves the job's accumulators through L2 (about 8 us at d = 256), so the window trades
// ~1.6e-8 of relative bias per row against ~10 % of kernel time at 1024 rows.
//
// Warp roles (320 threads): warps 0-7 converters + epilogue (TMEM lane quadrant = warp % 4),
// warp 8 TMA producer, warp 9 TMEM allocator + MMA issuer (one elected lane).
// Pipeline per stage: full (TMA -> converters), conv (converters -> MMA), empty (MMA commit ->
// TMA); per flush window: acc_full (MMA commit -> epilogue), acc_empty (epilogue -> MMA).
//
// Shared-memory stage (64 KB): [hi | lo], each [column group of 32][KT rows][128 B], swizzled in
// 32-byte chunks (chunk ^= row % 4) exactly as TMA writes it (SWIZZLE_128B_ATOM_32B) and as the UMMA
// descriptor reads it (SWIZZLE_128B_BASE32B, the only layout kind::tf32 takes for MN-major
// operands; LBO = KT*128, SBO = 512).
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace dr4g {
namespace {

constexpr int kBlk = 128;            // column block = MMA M = MMA N
constexpr int kStages = 3;
constexpr int kStageBytes = 65536;   // hi 32 KB + lo 32 KB
constexpr int kHalfBytes = 32768;
constexpr int kConvWarps = 8;
constexpr int kConvThreads = kConvWarps * 32;
constexpr int kThreads = kConvThreads + 64;
constexpr int kBoxRows = 16;         // TMA box = 32 columns x 16 rows = 2 KB
constexpr int kBoxBytes = 32 * 4 * kBoxRows;
constexpr int kMaxJobs = 16;
constexpr int kLookahead = kStages - 1;  // chunks converted past a window end before its epilogue
constexpr uint32_t kTmemCols = 512;

struct TcJob {
  int nblk;       // column blocks this job loads (1..4)
  int nbp;        // nblk rounded up to a power of two (1, 2, 4); KT = 64 / nbp
  int blk[4];     // global column-block indices
  int ntile;      // output tiles (1..4), tile t accumulates in TMEM columns [128 t, 128 t + 128)
  int ta[4];      // index into blk[] of the tile's row block (A operand)
  int tb[4];      // index into blk[] of the tile's column block (B operand); blk[ta] <= blk[tb]
  int stat_mask;  // bit i set: this job owns the column statistics of blk[i]
  int cta_begin;  // first CTA of this job
  int ncta;       // CTAs (row splits) of this job
};

struct TcJobTable {
  int njobs;
  TcJob job[kMaxJobs];
};

// Debug aid (DR4G_GRAM_TRACE=1): CTA 0 records clock64() at the pipeline hand-offs of its first
// kTraceIts chunks: [it][0] TMA: stage free, [1] TMA issued, [2] conv: data landed, [3] conv: converted,
// [4] conv: arrived, [5] MMA: operands ready, [6] MMA: issued + committed.
constexpr int kTraceIts = 96;
__device__ long long g_trace[kTraceIts * 8];
#define DR4G_TRACE(ev) \
  do { if (trace && blockIdx.x == 0 && it < kTraceIts) g_trace[it * 8 + (ev)] = clock64(); } while (0)

__device__ __forceinline__ float clear13(float v) { return __uint_as_float(__float_as_uint(v) & 0xffffe000u); }

__global__ void __launch_bounds__(kThreads, 1)
gram_tc_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ TcJobTable table, long long n, int d,
               const float* __restrict__ shift, float* __restrict__ slab, double* __restrict__ osum,
               double* __restrict__ osq, int flush_rows, int trace) {
  // ---------------------------------------------------------------- which job / which chunks
  int ji = 0;
  while (ji + 1 < table.njobs && (int)blockIdx.x >= table.job[ji + 1].cta_begin) ++ji;
  const TcJob& job = table.job[ji];
  const int split = (int)blockIdx.x - job.cta_begin;
  const int kt = 64 / job.nbp;
  const long long n_chunks = (n + kt - 1) / kt;
  if (split >= job.ncta || split >= n_chunks) return;
  const int my_chunks = (int)((n_chunks - split + job.ncta - 1) / job.ncta);
  const int flush_chunks = flush_rows > kt ? flush_rows / kt : 1;  // chunks per flush window

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ __align__(8) uint64_t bar_full[kStages], bar_conv[kStages], bar_empty[kStages], bar_acc_full, bar_acc_empty;
  __shared__ uint32_t tmem_base_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(ptx::smem_u32(&bar_full[s]), 1);
      ptx::mbar_init(ptx::smem_u32(&bar_conv[s]), kConvWarps);
      ptx::mbar_init(ptx::smem_u32(&bar_empty[s]), 1);
    }
    ptx::mbar_init(ptx::smem_u32(&bar_acc_full), 1);
    ptx::mbar_init(ptx::smem_u32(&bar_acc_empty), kConvWarps);
    ptx::fence_barrier_init();
  }
  if (warp == kConvWarps + 1) {
    ptx::tmem_alloc(ptx::smem_u32(&tmem_base_slot), kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_slot;

  if (warp == kConvWarps) {
    // ================================================================ TMA producer
    if (ptx::elect_one()) {
      ptx::prefetch_tmap(&tmap);
      for (int it = 0; it < my_chunks; ++it) {
        const int s = it % kStages;
        const uint32_t ph = (uint32_t)(it / kStages) & 1u;
        ptx::mbar_wait(ptx::smem_u32(&bar_empty[s]), ph ^ 1u);
        DR4G_TRACE(0);
        const long long row0 = ((long long)split + (long long)it * job.ncta) * kt;
        // boxes wholly outside the matrix are not loaded (the converters write zeros there)
        int boxes = 0;
        for (int bi = 0; bi < job.nblk; ++bi)
          for (int sub = 0; sub < 4; ++sub)
            if (job.blk[bi] * kBlk + sub * 32 < d)
              for (int rc = 0; rc < kt / kBoxRows; ++rc)
                if (row0 + rc * kBoxRows < n) ++boxes;
        const uint32_t full = ptx::smem_u32(&bar_full[s]);
        ptx::mbar_arrive_expect_tx(full, (uint32_t)boxes * kBoxBytes);
        const uint32_t sx = smem_base + (uint32_t)s * kStageBytes;
        for (int bi = 0; bi < job.nblk; ++bi)
          for (int sub = 0; sub < 4; ++sub) {
            const int col = job.blk[bi] * kBlk + sub * 32;
            if (col >= d) continue;
            for (int rc = 0; rc < kt / kBoxRows; ++rc) {
              const long long row = row0 + rc * kBoxRows;
              if (row >= n) continue;
              ptx::tma_load_2d(sx + (uint32_t)(((bi * 4 + sub) * kt + rc * kBoxRows) * 128), &tmap, full, col, (int)row);
            }
          }
        DR4G_TRACE(1);
      }
    }
  } else if (warp == kConvWarps + 1) {
    // ================================================================ MMA issuer
    if (ptx::elect_one()) {
      constexpr uint32_t idesc = ptx::make_idesc_tf32(kBlk, kBlk, true, true);
      const uint32_t lbo = (uint32_t)kt * 128u;
      int window = 0;
      for (int it = 0; it < my_chunks; ++it) {
        const int s = it % kStages;
        const uint32_t ph = (uint32_t)(it / kStages) & 1u;
        const bool window_start = (it % flush_chunks) == 0;
        if (window_start && it > 0) {
          // accumulators of the previous window must have been drained by the epilogue
          ptx::mbar_wait(ptx::smem_u32(&bar_acc_empty), (uint32_t)(window - 1) & 1u);
          ptx::tc_fence_after_sync();
        }
        ptx::mbar_wait(ptx::smem_u32(&bar_conv[s]), ph);
        ptx::tc_fence_after_sync();
        DR4G_TRACE(5);
        const uint32_t sx = smem_base + (uint32_t)s * kStageBytes;
        for (int ks = 0; ks < kt / 8; ++ks) {
          const uint32_t acc = (window_start && ks == 0) ? 0u : 1u;
          for (int t = 0; t < job.ntile; ++t) {
            const uint32_t a_off = (uint32_t)(job.ta[t] * 4 * kt) * 128u + (uint32_t)ks * 1024u;
            const uint32_t b_off = (uint32_t)(job.tb[t] * 4 * kt) * 128u + (uint32_t)ks * 1024u;
            const uint64_t a_hi = ptx::make_smem_desc(sx + a_off, lbo, 512, ptx::kLayoutSw128Base32);
            const uint64_t b_hi = ptx::make_smem_desc(sx + b_off, lbo, 512, ptx::kLayoutSw128Base32);
            const uint64_t a_lo = ptx::make_smem_desc(sx + kHalfBytes + a_off, lbo, 512, ptx::kLayoutSw128Base32);
            const uint64_t b_lo = ptx::make_smem_desc(sx + kHalfBytes + b_off, lbo, 512, ptx::kLayoutSw128Base32);
            const uint32_t dcol = tmem_base + (uint32_t)t * kBlk;
            ptx::mma_tf32_ss(dcol, a_hi, b_hi, idesc, acc);
            ptx::mma_tf32_ss(dcol, a_hi, b_lo, idesc, 1u);
            ptx::mma_tf32_ss(dcol, a_lo, b_hi, idesc, 1u);
          }
        }
        ptx::mma_commit(ptx::smem_u32(&bar_empty[s]));
        DR4G_TRACE(6);
        const bool window_end = ((it + 1) % flush_chunks) == 0 || it == my_chunks - 1;
        if (window_end) {
          ptx::mma_commit(ptx::smem_u32(&bar_acc_full));
          ++window;
        }
      }
    }
  } else {
    // ================================================================ converters + epilogue
    const int tid = threadIdx.x;
    const int quads = 32 * job.nbp;
    const int q = tid % quads, rg = tid / quads, nrg = kConvThreads / quads;  // nrg = 8 / nbp, kt / nrg = 8
    const int bi = q >> 5, sub = (q >> 3) & 3, lc = q & 7;
    const bool has_block = bi < job.nblk;
    const int col = has_block ? job.blk[bi] * kBlk + sub * 32 + lc * 4 : d;
    const bool col_ok = col < d;
    const bool stat_owner = col_ok && osum != nullptr && ((job.stat_mask >> bi) & 1);
    float4 sh = make_float4(0.f, 0.f, 0.f, 0.f);
    if (col_ok && shift) sh = *reinterpret_cast<const float4*>(shift + col);
    double ds[4] = {0, 0, 0, 0}, dq[4] = {0, 0, 0, 0};
    const uint32_t q_off = (uint32_t)((bi * 4 + sub) * kt) * 128u;
    const int quadrant = warp & 3, half = warp >> 2;

    // Window flush: TMEM accumulators -> this CTA's private fp32 slab [tile][row m][col n] in global
    // memory (L2 resident).  The first window stores, later ones use vector RED (fire-and-forget,
    // round-to-nearest adds in L2), so the tensor core's truncating accumulation never runs longer
    // than one window.  reduce_slabs_kernel sums the slabs in fp64 afterwards.
    float* my_slab = slab + (size_t)blockIdx.x * (4 * kBlk * kBlk);
    auto epilogue = [&](int window) {
      ptx::mbar_wait(ptx::smem_u32(&bar_acc_full), (uint32_t)window & 1u);
      ptx::tc_fence_after_sync();
      for (int t = 0; t < job.ntile; ++t) {
#pragma unroll
        for (int part = 0; part < 2; ++part) {
          const int c0 = half * 64 + part * 32;
          uint32_t v[32];
          ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(quadrant * 32) << 16) + (uint32_t)(t * kBlk + c0), v);
          ptx::tmem_ld_wait();
          float* dst = my_slab + ((size_t)t * kBlk + quadrant * 32 + lane) * kBlk + c0;
          if (window == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              asm volatile("st.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(dst + j), "r"(v[j]), "r"(v[j + 1]), "r"(v[j + 2]), "r"(v[j + 3]) : "memory");
          } else {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst + j), "f"(__uint_as_float(v[j])),
                           "f"(__uint_as_float(v[j + 1])), "f"(__uint_as_float(v[j + 2])), "f"(__uint_as_float(v[j + 3]))
                           : "memory");
          }
        }
      }
      ptx::tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&bar_acc_empty));
    };

    int windows_done = 0;
    for (int it = 0; it < my_chunks; ++it) {
      const int s = it % kStages;
      const uint32_t ph = (uint32_t)(it / kStages) & 1u;
      ptx::mbar_wait(ptx::smem_u32(&bar_full[s]), ph);
      if (tid == 0) DR4G_TRACE(2);
      const long long row0 = ((long long)split + (long long)it * job.ncta) * kt;
      const uint32_t sx = smem_base + (uint32_t)s * kStageBytes + q_off;
      if (has_block) {
        float fs[4] = {0.f, 0.f, 0.f, 0.f}, fq[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int r = rg + nrg * i;
          // 32-byte chunks XOR (row % 4): the BASE32B swizzle, as TMA (SWIZZLE_128B_ATOM_32B) laid it out
          const uint32_t a = sx + (uint32_t)r * 128u + (uint32_t)(((((lc >> 1) ^ (r & 3)) << 1) | (lc & 1)) << 4);
          float4 hi = make_float4(0.f, 0.f, 0.f, 0.f), lo = hi;
          if (col_ok && row0 + r < n) {
            float4 x;
            asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w) : "r"(a));
            const float y0 = x.x - sh.x, y1 = x.y - sh.y, y2 = x.z - sh.z, y3 = x.w - sh.w;
            fs[0] += y0; fs[1] += y1; fs[2] += y2; fs[3] += y3;
            fq[0] = fmaf(y0, y0, fq[0]); fq[1] = fmaf(y1, y1, fq[1]);
            fq[2] = fmaf(y2, y2, fq[2]); fq[3] = fmaf(y3, y3, fq[3]);
            hi.x = clear13(y0); hi.y = clear13(y1); hi.z = clear13(y2); hi.w = clear13(y3);
            lo.x = clear13(y0 - hi.x); lo.y = clear13(y1 - hi.y); lo.z = clear13(y2 - hi.z); lo.w = clear13(y3 - hi.w);
          }
          asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a), "f"(hi.x), "f"(hi.y), "f"(hi.z), "f"(hi.w) : "memory");
          asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(a + kHalfBytes), "f"(lo.x), "f"(lo.y), "f"(lo.z), "f"(lo.w) : "memory");
        }
        if (stat_owner) {
#pragma unroll
          for (int j = 0; j < 4; ++j) { ds[j] += (double)fs[j]; dq[j] += (double)fq[j]; }
        }
      }
      if (tid == 0) DR4G_TRACE(3);
      ptx::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&bar_conv[s]));
      if (tid == 0) DR4G_TRACE(4);

      const int e = it - kLookahead;
      if (e >= 0 && (e + 1) % flush_chunks == 0) epilogue(windows_done++);
    }
    // window ends not serviced inside the loop (the last kLookahead chunks and the final partial window)
    for (int e = (my_chunks - kLookahead > 0 ? my_chunks - kLookahead : 0); e < my_chunks; ++e)
      if ((e + 1) % flush_chunks == 0 || e == my_chunks - 1) epilogue(windows_done++);

    if (stat_owner) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        atomicAdd(&osum[col + j], ds[j]);
        atomicAdd(&osq[col + j], dq[j]);
      }
    }
  }

  ptx::tc_fence_before_sync();
  __syncthreads();
  if (warp == kConvWarps + 1) ptx::tmem_dealloc(tmem_base, kTmemCols);
}

// G (fp64, full symmetric) <- sum over a job's CTAs of their fp32 slabs, in a fixed order
// (deterministic).  grid = (tiles of all jobs, 128 rows m), block = 128 threads (column n).
// Tile (A, B), A <= B: D[m][n] = G[A*128+m][B*128+n] and, by symmetry, G[B*128+n][A*128+m].
__global__ void __launch_bounds__(kBlk)
reduce_slabs_kernel(const __grid_constant__ TcJobTable table, const float* __restrict__ slab, long long n, int d,
                    double* __restrict__ G) {
  int ji = 0, t = (int)blockIdx.x;
  while (t >= table.job[ji].ntile) { t -= table.job[ji].ntile; ++ji; }
  const TcJob& job = table.job[ji];
  const int kt = 64 / job.nbp;
  const long long n_chunks = (n + kt - 1) / kt;
  const int active = (int)(n_chunks < job.ncta ? n_chunks : job.ncta);
  const int m = (int)blockIdx.y, nn = (int)threadIdx.x;
  const int ga = job.blk[job.ta[t]] * kBlk + m, gb = job.blk[job.tb[t]] * kBlk + nn;
  if (ga >= d || gb >= d) return;
  // diagonal tiles: the MMA computes both (m, n) and (n, m) with different accumulation orders;
  // keep the upper one so that G is exactly symmetric
  if (job.ta[t] == job.tb[t] && nn < m) return;
  const float* p = slab + ((size_t)job.cta_begin * 4 + t) * (kBlk * kBlk) + (size_t)m * kBlk + nn;
  double acc = 0.0;
  for (int c = 0; c < active; ++c) acc += (double)p[(size_t)c * (4 * kBlk * kBlk)];
  G[(size_t)ga * d + gb] = acc;
  G[(size_t)gb * d + ga] = acc;
}

}  // namespace

// ---------------------------------------------------------------- host side
namespace {

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn(Ctx& c) { return reinterpret_cast<EncodeTiledFn>(tensor_map_encoder(c)); }

}  // namespace

void* tensor_map_encoder(Ctx& c) {
  if (!c.encode_tiled) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    DR4G_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (qres != cudaDriverEntryPointSuccess || !fn) fail("cuTensorMapEncodeTiled is not available in this driver");
    c.encode_tiled = fn;
  }
  return c.encode_tiled;
}

namespace {

// Cut the block upper triangle into jobs: column blocks are paired into super-blocks of 256
// columns; job (I, J), I <= J, owns the tiles between super-blocks I and J.
bool build_jobs(int d, int sm_count, TcJobTable& tab) {
  const int T = (d + kBlk - 1) / kBlk;
  const int S = (T + 1) / 2;
  if (S * (S + 1) / 2 > kMaxJobs) return false;
  tab.njobs = 0;
  int total_cost = 0;
  for (int I = 0; I < S; ++I)
    for (int J = I; J < S; ++J) {
      TcJob j{};
      const int i0 = 2 * I, i1 = (2 * I + 2 < T) ? 2 * I + 2 : T;
      const int j0 = 2 * J, j1 = (2 * J + 2 < T) ? 2 * J + 2 : T;
      for (int b = i0; b < i1; ++b) j.blk[j.nblk++] = b;
      const int na = j.nblk;
      if (J != I)
        for (int b = j0; b < j1; ++b) j.blk[j.nblk++] = b;
      for (int a = 0; a < na; ++a)
        for (int b = (J == I ? a : na); b < j.nblk; ++b) {
          j.ta[j.ntile] = a;
          j.tb[j.ntile] = b;
          ++j.ntile;
        }
      j.nbp = j.nblk <= 1 ? 1 : (j.nblk == 2 ? 2 : 4);
      j.stat_mask = (J == I) ? ((1 << na) - 1) : 0;
      tab.job[tab.njobs++] = j;
      total_cost += j.ntile;
    }
  // CTAs in proportion to the tile count, at least one each
  int ctas = sm_count > tab.njobs ? sm_count : tab.njobs;
  int given = 0;
  for (int i = 0; i < tab.njobs; ++i) {
    int m = (int)((long long)ctas * tab.job[i].ntile / total_cost);
    if (m < 1) m = 1;
    tab.job[i].ncta = m;
    given += m;
  }
  for (int i = 0; given < ctas; i = (i + 1) % tab.njobs) {
    tab.job[i].ncta++;
    given++;
  }
  int begin = 0;
  for (int i = 0; i < tab.njobs; ++i) {
    tab.job[i].cta_begin = begin;
    begin += tab.job[i].ncta;
  }
  return true;
}

}  // namespace

static bool use_pair(const Ctx& c, const float* dX, long long n, int d) {
  return c.gram_path != 3 && gram_pair_eligible(c, dX, n, d);
}

bool gram_tc_eligible(const Ctx& c, const float* dX, long long n, int d) {
  if (use_pair(c, dX, n, d)) return true;
  if (n < 1 || d < 4 || (d % 4) != 0) return false;                 // TMA: 16-byte row pitch
  if ((reinterpret_cast<uintptr_t>(dX) & 15) != 0) return false;    // TMA: 16-byte base
  if (n >= (1ll << 31)) return false;                               // TMA coordinates are int32
  const int T = (d + kBlk - 1) / kBlk, S = (T + 1) / 2;
  return S * (S + 1) / 2 <= kMaxJobs;
}

size_t gram_tc_workspace_bytes(const Ctx& c, int d) {
  TcJobTable tab;
  size_t single = 0;
  if (build_jobs(d, c.sm_count, tab)) {
    const int grid = tab.job[tab.njobs - 1].cta_begin + tab.job[tab.njobs - 1].ncta;
    single = (size_t)grid * 4 * kBlk * kBlk * sizeof(float);
  }
  const size_t pair = (d > kBlk && d % 4 == 0) ? gram_pair_workspace_bytes(c, d) : 0;
  return single > pair ? single : pair;
}

void launch_gram_tc(Ctx& c, const float* dX, long long n, int d, const float* dshift, double* dG, double* dsum,
                    double* dsumsq, void* workspace) {
  if (use_pair(c, dX, n, d)) {
    launch_gram_pair(c, dX, n, d, dshift, dG, dsum, dsumsq, workspace);   // sets last_gram_kernel = 2 (3xTF32) / 3 (bf16x3)
    return;
  }
  c.last_gram_kernel = 1;
  TcJobTable tab;
  if (!build_jobs(d, c.sm_count, tab)) fail("gram_tc: d=%d needs more than %d jobs", d, kMaxJobs);
  CUtensorMap tmap;
  const cuuint64_t gdim[2] = {(cuuint64_t)d, (cuuint64_t)n};
  const cuuint64_t gstride[1] = {(cuuint64_t)d * sizeof(float)};
  const cuuint32_t box[2] = {32, (cuuint32_t)kBoxRows};
  const cuuint32_t estride[2] = {1, 1};
  CUresult r = encode_tiled_fn(c)(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(dX), gdim, gstride, box,
                                 estride, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) fail("cuTensorMapEncodeTiled failed (%d) for n=%lld d=%d", (int)r, n, d);

  const int smem = kStages * kStageBytes + 1024;
  static PerDeviceMax smem_set;
  if (smem_set.need(c.device, (size_t)smem))
    DR4G_CUDA(cudaFuncSetAttribute(gram_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));

  const int flush_rows = c.flush_rows > 0 ? c.flush_rows : 1024;
  const int grid = tab.job[tab.njobs - 1].cta_begin + tab.job[tab.njobs - 1].ncta;
  float* slab = static_cast<float*>(workspace);
  static const bool trace = getenv("DR4G_GRAM_TRACE") != nullptr;
  gram_tc_kernel<<<grid, kThreads, smem, c.stream>>>(tmap, tab, n, d, dshift, slab, dsum, dsumsq, flush_rows, trace ? 1 : 0);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
  if (trace) {
    static long long h[kTraceIts * 8];
    DR4G_CUDA(cudaStreamSynchronize(c.stream));
    DR4G_CUDA(cudaMemcpyFromSymbol(h, g_trace, sizeof(h)));
    fprintf(stderr, "gram_tc trace (CTA 0, clocks relative to chunk 0 'stage free'; n=%lld d=%d)\n", n, d);
    fprintf(stderr, "  it  tma_free tma_issued  landed converted  arrived  mma_ready mma_issued\n");
    for (int it = 0; it < kTraceIts; ++it) {
      fprintf(stderr, "%4d", it);
      for (int e = 0; e < 7; ++e) fprintf(stderr, " %9lld", h[it * 8 + e] ? h[it * 8 + e] - h[0] : -1ll);
      fprintf(stderr, "\n");
    }
  }
  int tiles = 0;
  for (int i = 0; i < tab.njobs; ++i) tiles += tab.job[i].ntile;
  reduce_slabs_kernel<<<dim3(tiles, kBlk), kBlk, 0, c.stream>>>(tab, slab, n, d, dG);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

}  // namespace dr4g
