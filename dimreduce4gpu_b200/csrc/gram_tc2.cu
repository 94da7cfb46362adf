// K2 (d > 128) - streamed Gram + column statistics on CTA PAIRS: tcgen05.mma cta_group::2, M = 256.
//
//   G = (X - 1 c^T)^T (X - 1 c^T),   s = sum_r (x_r - c),   q = sum_r (x_r - c)^2       (fp32 X, 3xTF32)
//
// Same contract as gram_tc.cu (which keeps d <= 128 and shapes whose d is not a multiple of 32);
// replaces the reference's cublasSgemm Gram (src/tsvd.cu:324,358 via src/matrix.cu:95-127) and the
// centring / variance sweeps around it (src/pca.cu:37-49, src/tsvd.cu:274-290).
//
// Why pairs.  Measured on B200 (tools/mma_probe.cu, profiles/r01_mma_probe_tf32_rates.txt): a
// kind::tf32 MMA runs at its floor (N/2 cycles for M = 128 per CTA) from any operand placement, and
// the tensor core has priority on shared memory; what is left for the converter warps (which split
// every fp32 element into two tf32 numbers: 4 B TMA write + 4 B read + 8 B written per element) is
// ~34 B/clk beside 128x128 MMAs, ~54 B/clk beside 128x256 and ~70 B/clk beside a 256x256 pair MMA,
// because each CTA of a pair only supplies half of B.  In a pair every converted element feeds
// 256 (A) or 256 (B) output rows/columns, so conversion stays under the MMA time for every d.
//
// Work decomposition.  Columns are cut into blocks of 128 (T blocks) and super-blocks of 256.  A
// JOB is (row super-block I) x (one or two N-chunks of 256 columns): CTA rank r of the pair owns A
// block 2I + r (its 128 accumulator lanes) and, for every chunk, B block cb + r (its half of the 256
// columns).  Only the block upper triangle by 256 is computed (S = ceil(d/256) diagonal super-tiles,
// S(S-1)/2 off-diagonal ones).  An off-diagonal super-tile takes the three tf32 products
// hi^T hi + hi^T lo + lo^T hi into one accumulator; a DIAGONAL super-tile (A and B are the same
// columns) takes only two, P = hi^T hi and R = hi^T lo, into separate accumulators, because
// lo^T hi = R^T: G = P + R + R^T is formed by reduce_pair_slabs_kernel.  Issued tensor flops are
// (2 S + 3 S(S-1)/2) * 256^2 * 2n: 2.0x the algorithmic 2nd^2 at d = 256, 1.75x at d = 512, 1.63x at
// d = 1024 (3.0 / 2.25 / 1.88 without the symmetric diagonal).
// A job's clusters split the rows (stage chunks it * ncl + split); partial accumulators go to per-CTA
// fp32 slabs that reduce_pair_slabs_kernel sums in fp64 in a fixed order (deterministic).
//
// Shared-memory stage: 16-row x 128-column SLOTS of 8 KB, each written by ONE 3-D TMA box
// (32 columns, 16 rows, 4 column groups) in the layout the MN-major tf32 descriptor needs
// ([group][row][128 B], 32-byte chunks XOR row % 4 = SWIZZLE_128B_ATOM_32B; LBO = 2 KB between
// groups, SBO = 512 B between 4-row atoms).  The first version issued 2 KB 2-D boxes and its
// producer needed ~4100 cycles per 32-row stage (the whole pipeline ran at the TMA issue rate, see
// the trace hook below); a stage is now 2-4 instructions.
//
// Warp roles (512 threads): 0-7 converters (+ column statistics), 8-10 TMA producers, 11 TMEM
// allocation + MMA issue (leader CTA only), 12-15 epilogue (TMEM lane quadrant = warp % 4).
// The fp32 accumulation of the tensor core truncates (about 1e-8 relative per accumulated row), so
// the accumulators are flushed every `flush_rows` rows: the first window is stored, later ones are
// RED-added (round to nearest, in L2) into the CTA's slab.  Jobs with one chunk use both halves
// of TMEM alternately, so their flush overlaps the next window's MMAs; window boundaries are
// staggered over the clusters so that the chip's flushes do not hit L2 at the same time.
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <type_traits>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace dr4g {
namespace {

constexpr int kBlk = 128;
constexpr int kSlotBytes = 8192;         // 16 rows x 128 columns x 4 B
constexpr int kSlotRows = 16;
constexpr int kRingBytes = 196608;       // 3xTF32: stage ring (hi + lo halves)
constexpr int kRingBytesBf16 = 229376;   // bf16x3: landing ring + operand ring (224 KB)
constexpr int kMaxStages = 12;
constexpr int kMaxStagesTf32 = 6;
constexpr int kConvWarps = 8;
constexpr int kConvThreads = kConvWarps * 32;
constexpr int kEpiWarps = 4;
constexpr int kProducers = 3;            // TMA-issuing warps: one box costs its issuer ~700 cycles whatever its size (a 4th warp = 544 threads caps the kernel at 96 registers: spills)
constexpr int kMmaWarp = kConvWarps + kProducers;
constexpr int kThreads = kConvThreads + (kProducers + 1 + kEpiWarps) * 32;
constexpr int kMaxPJobs = 40;
constexpr uint32_t kTmemCols = 512;
constexpr int kSlabFloats = 2 * kBlk * 256;  // per CTA: [chunk][lane][256 columns]

struct PJob {
  int I;            // row super-block: A block of rank r is 2 I + r
  int nchunk;       // N-chunks of 256 columns (1 or 2)
  int cb[2];        // first column block of each chunk (rank r supplies block cb + r)
  int nlb;          // column blocks each CTA loads per 16 rows (1..3); local block 0 is the A block
  int bb[3];        // first global block of local block lb (rank r loads bb + r)
  int lb_of[2];     // local block holding a chunk's B half
  int stats;        // 1: this job owns the column statistics of its A blocks
  int diag;         // 1: diagonal super-tile, two accumulators P = hi^T hi and R = hi^T lo (G = P + R + R^T)
  int cl_begin;     // first cluster of this job
  int ncl;          // clusters (row splits)
};

struct PJobTable {
  int njobs;
  PJob job[kMaxPJobs];
};

constexpr int kTraceIts = 96;
__device__ long long g_trace2[kTraceIts * 8];
__device__ unsigned long long g_cta_time[2 * 160];   // DR4G_GRAM_TRACE bit 32: %globaltimer at the start and end of every CTA
#define DR4G_TRACE2(ev) \
  do { if ((trace & 1) && (int)blockIdx.x == ((trace >> 8) & 0xff) && it >= (trace >> 16) * 64 && it < (trace >> 16) * 64 + kTraceIts) \
         g_trace2[(it - (trace >> 16) * 64) * 8 + (ev)] = clock64(); } while (0)

__device__ __forceinline__ float clear13(float v) { return __uint_as_float(__float_as_uint(v) & 0xffffe000u); }

// ------------------------------------------------------------------------------------------------ converters
// What the converter warps need to know about the kernel's rings and barriers (all shared-memory addresses).
struct ConvArgs {
  uint32_t smem_base, op_base, land_stride, half_bytes;
  uint32_t bar_full, bar_conv_leader, bar_empty, bar_lfree;   // arrays of 8-byte barriers (bar_conv: in the leader CTA)
  int NL, NO, my_chunks, ncl, split, kt, trace;
  long long n;
};

// One group of 4 converter warps (grp = 0 / 1 takes the even / odd stages).  The whole loop is instantiated per
// <operand split, column blocks per stage, column statistics>, so that an off-diagonal job does not carry the fp64
// statistics registers and every variant keeps all of a stage's loads in flight (the first version was one loop for
// every job type: ptxas serialised its eight 128-bit loads into four round trips to fit 120 registers, and the eight
// converter warps - two per scheduler - ran at ~2.5 cycles per instruction, 350 instructions per thread and stage).
// Thread -> fixed position inside a 16-row x 128-column slot: 4 columns (32-column group g, 16-byte chunk lc) and the
// rows rg + 4 h, h = 0..3.  Landing layout (TMA, SWIZZLE_128B_ATOM_32B): [g][row][128 B], 32-byte chunks XOR row % 4.
// bf16 operand layout of a slot: hi 4 KB then lo 4 KB, each [64-column group g >> 1][8-row atom][8 rows][128 B], 16-byte
// chunks XOR row % 8 (SWIZZLE_128B); a thread's 4 columns are 8 bytes.
template <bool BF16, int NLB, bool STATS>
__device__ __forceinline__ void converter_loop(const ConvArgs& a, const int grp, const float4 (&sh)[3], const bool (&col_ok)[3],
                                               const int col0, double* __restrict__ osum, double* __restrict__ osq) {
  constexpr int CPS = NLB == 1 ? 2 : 1;
  constexpr int NS = CPS * NLB;                    // slots per stage
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int lc = tid & 7, g = (tid >> 3) & 3, rg = warp & 3;
  const int trace = a.trace;
  // r = rg + 4 h: r % 4 = rg for every h, so the landing offset is toff0 + 512 h
  const uint32_t toff0 = (uint32_t)g * 2048u + (uint32_t)rg * 128u + (uint32_t)(((((lc >> 1) ^ rg) << 1) | (lc & 1)) << 4);
  // r % 8 = rg + 4 (h & 1), r / 8 = h >> 1: operand offset boff[h & 1] + 1024 (h >> 1)
  uint32_t boff[2];
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    const int r7 = rg + 4 * e;
    boff[e] = (uint32_t)(g >> 1) * 2048u + (uint32_t)r7 * 128u + (uint32_t)((((((g & 1) << 2) | (lc >> 1)) ^ r7) << 4) | ((lc & 1) << 3));
  }
  bool cols_ok = true;
#pragma unroll
  for (int lb = 0; lb < NLB; ++lb) cols_ok = cols_ok && col_ok[lb];
  const bool stat_owner = STATS && col_ok[0] && osum != nullptr;
  double ds[4] = {0, 0, 0, 0}, dq[4] = {0, 0, 0, 0};
  float fs[4] = {0.f, 0.f, 0.f, 0.f}, fq[4] = {0.f, 0.f, 0.f, 0.f};
  int s = grp, so = grp, pending = 0;              // landing slot / operand slot of this group's next stage (NL, NO even)
  uint32_t ph = 0, pho = 0;

  for (int it = grp; it < a.my_chunks; it += 2) {
    if (lane == 0 && rg == 0) DR4G_TRACE2(7);
    ptx::mbar_wait(a.bar_full + (uint32_t)s * 8u, ph);
    if (lane == 0 && rg == 0) DR4G_TRACE2(2);
    if constexpr (BF16) {
      // operand slot free (its MMAs of NO stages ago have completed; normally long ago)
      ptx::mbar_wait(a.bar_empty + (uint32_t)so * 8u, pho ^ 1u);
      if ((trace & 16) && lane == 0 && rg == 0) DR4G_TRACE2(1);
    }
    const long long row0 = ((long long)a.split + (long long)it * a.ncl) * a.kt;
    const uint32_t sx = a.smem_base + (uint32_t)s * a.land_stride + toff0;
    const uint32_t ox = a.op_base + (uint32_t)so * a.land_stride + boff[0];
    const uint32_t bdelta = boff[1] - boff[0];
    const bool fast = cols_ok && row0 + a.kt <= a.n;     // every element of the stage is a real one

    auto body = [&](auto masked_c) {
      constexpr bool MASKED = decltype(masked_c)::value;
      float4 x[NS * 4];
#pragma unroll
      for (int sl = 0; sl < NS; ++sl)
#pragma unroll
        for (int h = 0; h < 4; ++h)
          asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                       : "=f"(x[sl * 4 + h].x), "=f"(x[sl * 4 + h].y), "=f"(x[sl * 4 + h].z), "=f"(x[sl * 4 + h].w)
                       : "r"(sx + (uint32_t)sl * kSlotBytes + (uint32_t)h * 512u));
#pragma unroll
      for (int sl = 0; sl < NS; ++sl) {
        const int lb = sl % NLB, c = sl / NLB;
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          const float4 v = x[sl * 4 + h];
          float y0 = v.x - sh[lb].x, y1 = v.y - sh[lb].y, y2 = v.z - sh[lb].z, y3 = v.w - sh[lb].w;
          if constexpr (MASKED) {
            if (!(col_ok[lb] && row0 + c * kSlotRows + rg + 4 * h < a.n)) { y0 = 0.f; y1 = 0.f; y2 = 0.f; y3 = 0.f; }
          }
          if constexpr (STATS) {
            if (lb == 0) {
              fs[0] += y0; fs[1] += y1; fs[2] += y2; fs[3] += y3;
              fq[0] = fmaf(y0, y0, fq[0]); fq[1] = fmaf(y1, y1, fq[1]);
              fq[2] = fmaf(y2, y2, fq[2]); fq[3] = fmaf(y3, y3, fq[3]);
            }
          }
          if constexpr (BF16) {
            const uint32_t h01 = ptx::pack_bf16x2(y0, y1), h23 = ptx::pack_bf16x2(y2, y3);
            // y - bf16(y) is exact in fp32; its bf16 rounding is the second term
            const uint32_t l01 = ptx::pack_bf16x2(y0 - __uint_as_float(h01 << 16), y1 - __uint_as_float(h01 & 0xffff0000u));
            const uint32_t l23 = ptx::pack_bf16x2(y2 - __uint_as_float(h23 << 16), y3 - __uint_as_float(h23 & 0xffff0000u));
            const uint32_t o = ox + (uint32_t)sl * kSlotBytes + (uint32_t)(h >> 1) * 1024u + ((h & 1) ? bdelta : 0u);
            asm volatile("st.shared.v2.b32 [%0], {%1,%2};" ::"r"(o), "r"(h01), "r"(h23) : "memory");
            asm volatile("st.shared.v2.b32 [%0], {%1,%2};" ::"r"(o + 4096u), "r"(l01), "r"(l23) : "memory");
          } else {
            float4 hi, lo;
            hi.x = clear13(y0); hi.y = clear13(y1); hi.z = clear13(y2); hi.w = clear13(y3);
            lo.x = clear13(y0 - hi.x); lo.y = clear13(y1 - hi.y); lo.z = clear13(y2 - hi.z); lo.w = clear13(y3 - hi.w);
            const uint32_t o = sx + (uint32_t)sl * kSlotBytes + (uint32_t)h * 512u;   // hi in place, lo in the second half
            asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(o), "f"(hi.x), "f"(hi.y), "f"(hi.z), "f"(hi.w) : "memory");
            asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(o + a.half_bytes), "f"(lo.x), "f"(lo.y), "f"(lo.z), "f"(lo.w) : "memory");
          }
        }
      }
    };
    if (trace & 4) {          // probe: no conversion, the barriers still cycle
    } else if (fast) body(std::false_type{});
    else body(std::true_type{});
    if constexpr (BF16) {
      // every load of the landing slot has been consumed: hand it back to its TMA producer
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(a.bar_lfree + (uint32_t)s * 8u);
    }
    if constexpr (STATS) {
      // fp32 partial sums of at most 16 stages x 8 rows, then fp64 (fp64 conversions and adds are slow enough on B200
      // to double the conversion time when done per stage: measured)
      if (stat_owner && (++pending == 16 || it + 2 >= a.my_chunks)) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          ds[j] += (double)fs[j]; dq[j] += (double)fq[j];
          fs[j] = 0.f; fq[j] = 0.f;
        }
        pending = 0;
      }
    }
    if (lane == 0 && rg == 0) DR4G_TRACE2(3);
    // the MMA that reads this stage is issued by the leader CTA but reads THIS CTA's shared memory: make the writes
    // visible to the async proxy (the fence also waits for the stores queued behind the MMA's operand reads), then
    // arrive on the leader's barrier.  After the fence the stores are complete and the data's consumer is this CTA's
    // own tensor core, so the arrive is relaxed (a release at cluster scope cost 1000-3000 cycles here, measured).
    ptx::fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      if (trace & 8) ptx::mbar_arrive_cluster(a.bar_conv_leader + (uint32_t)so * 8u);
      else ptx::mbar_arrive_cluster_relaxed(a.bar_conv_leader + (uint32_t)so * 8u);
    }
    if (lane == 0 && rg == 0) DR4G_TRACE2(4);
    s += 2;
    if (s >= a.NL) { s -= a.NL; ph ^= 1u; }
    so += 2;
    if (so >= a.NO) { so -= a.NO; pho ^= 1u; }
  }
  if constexpr (STATS) {
    if (stat_owner) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        atomicAdd(&osum[col0 + j], ds[j]);
        atomicAdd(&osq[col0 + j], dq[j]);
      }
    }
  }
}


// BF16 = false: 3xTF32 (kind::tf32, hi/lo = two 11-bit halves, fp32-exact products except lo*lo).
// BF16 = true:  bf16x3 (kind::f16, twice the multiply rate, K = 16 per instruction): y = b1 + b2 + (|rest| <= 2^-18 |y|)
//   with b1 = bf16(y), b2 = bf16(y - b1), both rounded to nearest so the dropped parts have zero mean; products
//   b1^T b1 + b1^T b2 + b2^T b1, dropped b2^T b2 <= 2^-18 relative (measured Gram error ~2e-6 of sqrt(G_ii G_jj), below
//   the truncation of the tensor core's own fp32 accumulation).  The converters write 4 B per element (hi + lo) in
//   the MN-major SWIZZLE_128B layout of 16-bit operands ([64-column group][8-row atom][8 rows][128 B]) into the second
//   half of the stage instead of 8 B, and the MMA reads half as many operand bytes: the pass is bound by shared-
//   memory bandwidth (TMA write + converter read + converter write + operand reads), not by the tensor pipe.
template <bool BF16>
__global__ void __launch_bounds__(kThreads, 1)
gram_pair_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ PJobTable table, long long n, int d,
                 const float* __restrict__ shift, float* __restrict__ slab, double* __restrict__ osum,
                 double* __restrict__ osq, int flush_rows, int trace, int* __restrict__ prog, int pace_rows, int op_stages) {
  // ---------------------------------------------------------------- which job / which chunks
  const uint32_t rank = ptx::cluster_ctarank();
  const int cluster = (int)blockIdx.x >> 1;
  int ji = 0;
  while (ji + 1 < table.njobs && cluster >= table.job[ji + 1].cl_begin) ++ji;
  const PJob& job = table.job[ji];
  const int split = cluster - job.cl_begin;
  const int nlb = job.nlb, nchunk = job.nchunk;
  // 16-row slots per local block per stage: stages are kept small (32-48 KB) and the ring deep (4-6),
  // because a stage's way round the ring (TMA issue ~1000 clk + landing ~1100 + conversion + store
  // drain ~1000-1800 + MMA) is several times its MMA time
  const int cps = nlb == 1 ? 2 : 1;
  const int kt = kSlotRows * cps;                        // rows per stage
  const int slots = cps * nlb;
  const uint32_t half_bytes = (uint32_t)slots * kSlotBytes;   // hi half, then lo half
  const uint32_t stage_bytes = 2u * half_bytes;
  // 3xTF32: ONE ring of stages (landing half converted in place = hi, lo half behind it): NL = NO.
  // bf16x3: TWO rings.  The operands (hi + lo bf16 = 4 B per element) go to their own ring of NO `half_bytes` slots and
  // the fp32 landing slots form a ring of NL.  A landing slot is held from the TMA issue to the converters' read (free ->
  // issue ~1000 cycles -> landed 1500-4000 -> read ~600: measured with the trace hook), an operand slot only from the
  // converters' store to the end of its MMAs (~1500), and the pass ran at the pace of the landing ring (6 x 16 KB in
  // flight per SM against that latency: ~1000 cycles per stage, tensor pipe 35 % active, smem and HBM under 50 %).  So
  // nearly all of the shared memory is landing ring (12 x 16 KB or 6 x 24 KB in flight) and the operand ring is short.
  const int NO = BF16 ? ((int)(kRingBytesBf16 / half_bytes) - op_stages >= 4 ? op_stages : 4) : ((int)(kRingBytes / stage_bytes) < kMaxStagesTf32 ? (int)(kRingBytes / stage_bytes) : kMaxStagesTf32);
  int nl_ = BF16 ? (((int)(kRingBytesBf16 / half_bytes) - NO) & ~1) : NO;   // even: a converter group always meets the same slots
  const int NL = nl_ > kMaxStages ? kMaxStages : nl_;
  const uint32_t land_stride = BF16 ? half_bytes : stage_bytes;
  const uint32_t op_stride = land_stride;
  const long long n_chunks = (n + kt - 1) / kt;
  if (split >= job.ncl || split >= n_chunks) {          // both CTAs of a pair take the same branch
    if (prog != nullptr && threadIdx.x == 0 && rank == 0) prog[cluster] = 0x7fffffff;   // pacing: nothing to wait for here
    return;
  }
  const int my_chunks = (int)((n_chunks - split + job.ncl - 1) / job.ncl);
  // Stages per flush window.  `flush_rows` is calibrated for three accumulating MMAs per 8 rows; P of a
  // diagonal job takes one (R, 2^-11 of P, is never flushed before the end), so its window is 4x longer.
  const int diag = job.diag;
  const int fc0 = flush_rows > kt ? flush_rows / kt : 1;
  const int fc = diag ? 4 * fc0 : fc0;
  const int stagger = (split & 3) * (fc >> 2);                    // window phase of this cluster
  const int nacc = (nchunk == 1 && !diag) ? 2 : 1;                // accumulator sets in TMEM
  const int T = (d + kBlk - 1) / kBlk;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t op_base = BF16 ? smem_base + (uint32_t)NL * half_bytes : smem_base;   // 3xTF32: hi in place, lo at + half_bytes
  __shared__ __align__(8) uint64_t bar_full[kMaxStages], bar_conv[kMaxStages], bar_empty[kMaxStages], bar_lfree[kMaxStages];
  __shared__ __align__(8) uint64_t bar_acc_full[2], bar_acc_empty[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ volatile int pace_hold, pace_it;   // cross-job pacing: written by the pair's monitor / by producer 0

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    pace_hold = 0;
    pace_it = 0;
    for (int s = 0; s < kMaxStages; ++s) {
      ptx::mbar_init(ptx::smem_u32(&bar_full[s]), 1);                 // the stage's owner (one producer) + its TMA bytes
      ptx::mbar_init(ptx::smem_u32(&bar_conv[s]), kConvWarps);       // a group of 4 warps per CTA; leader CTA only
      ptx::mbar_init(ptx::smem_u32(&bar_empty[s]), 1);
      ptx::mbar_init(ptx::smem_u32(&bar_lfree[s]), kConvWarps / 2);   // bf16x3: the group of 4 warps that read the landing half
    }
    for (int a = 0; a < 2; ++a) {
      ptx::mbar_init(ptx::smem_u32(&bar_acc_full[a]), 1);
      ptx::mbar_init(ptx::smem_u32(&bar_acc_empty[a]), 2 * kEpiWarps);  // leader CTA only
    }
    ptx::fence_barrier_init();
  }
  if (warp == kMmaWarp) {
    ptx::tmem_alloc_pair(ptx::smem_u32(&tmem_base_slot), kTmemCols);
    ptx::tmem_relinquish_pair();
  }
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::cluster_sync();   // the peer's barriers exist before anything arrives on them
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_slot;
  if ((trace & 32) && threadIdx.x == 0 && blockIdx.x < 160) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g_cta_time[2 * blockIdx.x] = t;
  }

  if (warp >= kConvWarps && warp < kMmaWarp) {
    // ================================================================ TMA producers (each CTA its own blocks)
    // Measured (tools/tma_probe.cu): a thread issues one box per ~700 cycles whatever the box size, so one issuer with
    // 8 KB boxes streams 3.1 TB/s chip-wide and four issuers 6.5 TB/s.  A stage belongs to ONE producer,
    // which does everything for it (wait, expect_tx, all boxes): a first version dealt the BOXES round-robin and had
    // every producer arrive on every stage's barrier - its pipeline trace showed ~750 cycles per stage and producer
    // whether or not the producer owned a box, so three producers advanced no faster than one (1030 cycles per 16-row
    // stage at d = 512, the kernel's pace).
    //
    // Cross-job pacing.  At d > 256 every column block is read by several jobs (d = 512: two, d = 1024: four).  The jobs
    // are balanced to finish together but drift apart by a few per cent of n, far more than the L2 holds, so each job
    // pulled its own copy from HBM: ncu showed dram reads of 2.1x (d = 512) and 3.4x (d = 1024) the matrix, 3.5 TB/s.
    // A first version had the producers themselves publish their progress (a global RED per stage) and poll the other
    // jobs' counters every 8 stages: the polling round trips stalled the TMA issue and the pass got slower whatever
    // the window (5.6 against 4.6 ms at d = 512).  Now the otherwise idle MMA warp of the pair's second CTA is the
    // MONITOR: every ~1 us it publishes the pair's row frontier (stages issued x rows per stage of its job) in prog[cluster],
    // reads the frontiers of all clusters and raises `pace_hold` in both CTAs while this pair is more than `pace_rows`
    // ahead of the slowest one; a producer only reads that shared-memory flag.  (Per JOB totals were not enough: over a
    // long pass the clusters of one job drift apart as well, and at 2^24 rows the dram reads stayed at 1.93x.)  The
    // monitor gives up when the slowest frontier has not moved for ~100 polls (a partly resident grid cannot hang).
    if (ptx::elect_one()) {
      const int p = warp - kConvWarps;
      ptx::prefetch_tmap(&tmap);
      // a producer owns fixed STAGES of the ring (stage s belongs to producer s % np, np dividing the ring length), so it
      // waits for every round of its stages in order: with ownership by iteration (it % 3 on a ring of 4) the parity wait
      // of round r could pass while round r - 2 of that stage was still open (the wait only distinguishes two phases),
      // and two producers armed the same phase of bar_full (cuda-gdb: Warp Illegal Instruction in arrive.expect_tx)
      const int np = NL % 3 == 0 ? 3 : (NL % 2 == 0 ? 2 : 1);
      int s = 0, owner = 0;
      uint32_t ph = 0;
      const bool pacing = prog != nullptr && table.njobs > 1;
      int blk_of[3];
#pragma unroll
      for (int lb = 0; lb < 3; ++lb) blk_of[lb] = lb < nlb ? job.bb[lb] + (int)rank : T;
      for (int it = 0; it < my_chunks; ++it) {
        if (owner == p) {
          if (pacing) {
            while (pace_hold) __nanosleep(200);
            if (p == 0) pace_it = it;
          }
          // 3xTF32: the stage (landing half converted in place + lo half) is free when its MMAs have completed.
          // bf16x3: the operands live in the other half, so the landing half is free as soon as the converters have
          // read it (bar_lfree) and the operand half when its MMAs have completed (bar_empty, waited for by the converters).
          ptx::mbar_wait(ptx::smem_u32(BF16 ? &bar_lfree[s] : &bar_empty[s]), ph ^ 1u);
          if (p == 0 && !(trace & 16)) DR4G_TRACE2(0);
          const long long row0 = ((long long)split + (long long)it * job.ncl) * kt;
          const uint32_t full = ptx::smem_u32(&bar_full[s]);
          const uint32_t sx = smem_base + (uint32_t)s * land_stride;
          int boxes = 0;
          for (int c = 0; c < cps; ++c)
            for (int lb = 0; lb < nlb; ++lb)
              if (row0 + c * kSlotRows < n && blk_of[lb] < T) ++boxes;
          ptx::mbar_arrive_expect_tx(full, (uint32_t)boxes * kSlotBytes);
          for (int c = 0, sl = 0; c < cps; ++c)
            for (int lb = 0; lb < nlb; ++lb, ++sl) {
              const long long row = row0 + c * kSlotRows;
              if (row >= n || blk_of[lb] >= T) continue;
              ptx::tma_load_3d(sx + (uint32_t)sl * kSlotBytes, &tmap, full, 0, (int)row, blk_of[lb] * 4);
            }
          if (p == 0 && !(trace & 16)) DR4G_TRACE2(1);
        }
        if (++owner == np) owner = 0;
        if (++s == NL) { s = 0; ph ^= 1u; owner = 0; }
      }
      if (pacing && p == 0) pace_it = my_chunks;   // tells the monitor that this pair has issued everything
    }
  } else if (warp == kMmaWarp) {
    if (rank == 1 && prog != nullptr && table.njobs > 1 && ptx::elect_one()) {
      // ============================================================== pacing monitor (second CTA of the pair)
      const uint32_t hold_here = ptx::smem_u32(const_cast<int*>(&pace_hold));
      const uint32_t hold_leader = ptx::mapa(hold_here, 0);
      int held = 0, slow_prev = -1, same = 0;
      bool on = true;
      const int nclus = (int)gridDim.x >> 1;
      const int step_rows = job.ncl * kt;       // rows the job's frontier moves per stage of this cluster
      for (;;) {
        const int cur = pace_it;
        const int mine = cur >= my_chunks ? 0x7fffffff : cur * step_rows;
        asm volatile("st.relaxed.gpu.global.s32 [%0], %1;" ::"l"(prog + cluster), "r"(mine) : "memory");
        if (cur >= my_chunks) break;
        int slowest = 0x7fffffff;
        for (int j = 0; j < nclus; ++j) {
          int v;
          asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(prog + j));
          if (v < slowest) slowest = v;
        }
        same = slowest == slow_prev ? same + 1 : 0;
        slow_prev = slowest;
        if (same > 100) on = false;               // the slowest cluster is not running: stop waiting for it
        const int want = (on && mine - slowest > pace_rows) ? 1 : 0;
        if (want != held) {
          held = want;
          asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(hold_leader), "r"(want) : "memory");
          pace_hold = want;
        }
        __nanosleep(1000);
      }
      if (held) {
        asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(hold_leader), "r"(0) : "memory");
        pace_hold = 0;
      }
    }
    // ================================================================ MMA issuer (leader CTA, one lane)
    if (rank == 0 && ptx::elect_one()) {
      // One thread feeds the tensor pipe: no divisions, no constant-bank indexing inside the loop
      // (a first version spent ~1900 cycles per 12-MMA stage on exactly that, measured).
      constexpr uint32_t idesc = BF16 ? ptx::make_idesc_bf16(256, 256, true, true) : ptx::make_idesc_tf32(256, 256, true, true);
      const int nks = (trace & 2) ? 0 : kt / 8;
      const uint32_t boff0 = (uint32_t)job.lb_of[0] * kSlotBytes, boff1 = (uint32_t)job.lb_of[1] * kSlotBytes;
      const uint32_t slot_row_bytes = (uint32_t)nlb * kSlotBytes;   // next 16 rows of the same local block
      // address added below.  tf32: [32-column group][4-row atom][4 rows][128 B]; bf16: [64-column group][8-row atom][8 rows][128 B]
      const uint64_t desc0 = BF16 ? ptx::make_smem_desc(0, 2048, 1024, ptx::kLayoutSw128)
                                  : ptx::make_smem_desc(0, 2048, 512, ptx::kLayoutSw128Base32);
      int s = 0, window = 0, wpos = stagger, set = 0;
      uint32_t ph = 0, racc = 0u;
      uint32_t sx = op_base;
      bool window_start = true;
      for (int it = 0; it < my_chunks; ++it) {
        if (window_start && window >= nacc) {
          // the epilogue warps of both CTAs have drained this accumulator set
          ptx::mbar_wait_cluster(ptx::smem_u32(&bar_acc_empty[set]), (uint32_t)(nacc == 2 ? (window - 2) >> 1 : window - 1) & 1u);
          ptx::tc_fence_after_sync();
        }
        ptx::mbar_wait_cluster(ptx::smem_u32(&bar_conv[s]), ph);
        ptx::tc_fence_after_sync();
        DR4G_TRACE2(5);
        const uint32_t d0 = tmem_base + (uint32_t)set * 256u;
        uint32_t acc = window_start ? 0u : 1u;
        if constexpr (BF16) {
          // one K = 16 instruction per product and 16-row slot; hi at the slot's offset in the second half of the
          // stage, lo 4 KB behind it
          uint32_t row = sx;
#pragma unroll 1
          for (int ks = 0; ks < nks; ks += 2) {
            const uint64_t a_hi = desc0 | (uint64_t)((row >> 4) & 0x3FFF);
            const uint64_t a_lo = desc0 | (uint64_t)(((row + 4096u) >> 4) & 0x3FFF);
            const uint64_t b_hi = desc0 | (uint64_t)(((row + boff0) >> 4) & 0x3FFF);
            const uint64_t b_lo = desc0 | (uint64_t)(((row + boff0 + 4096u) >> 4) & 0x3FFF);
            if (diag) {
              ptx::mma_f16_ss_pair(d0, a_hi, b_hi, idesc, acc);
              ptx::mma_f16_ss_pair(d0 + 256u, a_hi, b_lo, idesc, racc);
              racc = 1u;
            } else {
              ptx::mma_f16_ss_pair(d0, a_hi, b_hi, idesc, acc);
              ptx::mma_f16_ss_pair(d0, a_hi, b_lo, idesc, 1u);
              ptx::mma_f16_ss_pair(d0, a_lo, b_hi, idesc, 1u);
              if (nchunk == 2) {
                const uint64_t c_hi = desc0 | (uint64_t)(((row + boff1) >> 4) & 0x3FFF);
                const uint64_t c_lo = desc0 | (uint64_t)(((row + boff1 + 4096u) >> 4) & 0x3FFF);
                ptx::mma_f16_ss_pair(d0 + 256u, a_hi, c_hi, idesc, acc);
                ptx::mma_f16_ss_pair(d0 + 256u, a_hi, c_lo, idesc, 1u);
                ptx::mma_f16_ss_pair(d0 + 256u, a_lo, c_hi, idesc, 1u);
              }
            }
            acc = 1u;
            row += slot_row_bytes;
          }
        } else {
        uint32_t row = sx;
#pragma unroll 1
        for (int ks = 0; ks < nks; ks += 2) {
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const uint32_t r = row + (uint32_t)h * 1024u;
            const uint64_t a_hi = desc0 | (uint64_t)((r >> 4) & 0x3FFF);
            const uint64_t a_lo = desc0 | (uint64_t)(((r + half_bytes) >> 4) & 0x3FFF);
            const uint64_t b_hi = desc0 | (uint64_t)(((r + boff0) >> 4) & 0x3FFF);
            const uint64_t b_lo = desc0 | (uint64_t)(((r + boff0 + half_bytes) >> 4) & 0x3FFF);
            if (diag) {
              // P (columns 0-255) += hi^T hi, R (columns 256-511) += hi^T lo; lo^T hi = R^T is added by
              // reduce_pair_slabs_kernel: 2 MMAs per 8 rows instead of 3
              ptx::mma_tf32_ss_pair(d0, a_hi, b_hi, idesc, acc);
              ptx::mma_tf32_ss_pair(d0 + 256u, a_hi, b_lo, idesc, racc);
              racc = 1u;
              acc = 1u;
              continue;
            }
            ptx::mma_tf32_ss_pair(d0, a_hi, b_hi, idesc, acc);
            ptx::mma_tf32_ss_pair(d0, a_hi, b_lo, idesc, 1u);
            ptx::mma_tf32_ss_pair(d0, a_lo, b_hi, idesc, 1u);
            if (nchunk == 2) {
              const uint64_t c_hi = desc0 | (uint64_t)(((r + boff1) >> 4) & 0x3FFF);
              const uint64_t c_lo = desc0 | (uint64_t)(((r + boff1 + half_bytes) >> 4) & 0x3FFF);
              ptx::mma_tf32_ss_pair(d0 + 256u, a_hi, c_hi, idesc, acc);
              ptx::mma_tf32_ss_pair(d0 + 256u, a_hi, c_lo, idesc, 1u);
              ptx::mma_tf32_ss_pair(d0 + 256u, a_lo, c_hi, idesc, 1u);
            }
            acc = 1u;
          }
          row += slot_row_bytes;
        }
        }
        ptx::mma_commit_pair(ptx::smem_u32(&bar_empty[s]), 3);   // stage free in both CTAs
        DR4G_TRACE2(6);
        window_start = false;
        if (++wpos == fc || it == my_chunks - 1) {              // window end
          ptx::mma_commit_pair(ptx::smem_u32(&bar_acc_full[set]), 3);
          ++window;
          wpos = 0;
          window_start = true;
          if (nacc == 2) set ^= 1;
        }
        sx += op_stride;
        if (++s == NO) { s = 0; ph ^= 1u; sx = op_base; }
      }
    }
  } else if (warp < kConvWarps) {
    // ================================================================ converters (+ column statistics): converter_loop above
    // Two groups of 4 warps convert alternate stages, so that one group's store drain (the proxy fence waits for the
    // stores queued behind the MMA's operand reads) overlaps the other group's loads.
    const int tid = threadIdx.x;
    const int lc = tid & 7, g = (tid >> 3) & 3;
    float4 sh[3];
    bool col_ok[3];
    int col0[3];
#pragma unroll
    for (int lb = 0; lb < 3; ++lb) {
      sh[lb] = make_float4(0.f, 0.f, 0.f, 0.f);
      col0[lb] = lb < nlb ? (job.bb[lb] + (int)rank) * kBlk + g * 32 + lc * 4 : d;
      col_ok[lb] = col0[lb] < d;
      if (col_ok[lb] && shift) sh[lb] = *reinterpret_cast<const float4*>(shift + col0[lb]);
    }
    ConvArgs ca;
    ca.smem_base = smem_base; ca.op_base = op_base; ca.land_stride = land_stride; ca.half_bytes = half_bytes;
    ca.bar_full = ptx::smem_u32(&bar_full[0]); ca.bar_conv_leader = ptx::mapa(ptx::smem_u32(&bar_conv[0]), 0);
    ca.bar_empty = ptx::smem_u32(&bar_empty[0]); ca.bar_lfree = ptx::smem_u32(&bar_lfree[0]);
    ca.NL = NL; ca.NO = NO; ca.my_chunks = my_chunks; ca.ncl = job.ncl; ca.split = split; ca.kt = kt; ca.trace = trace;
    ca.n = n;
    const int grp = warp >> 2;
    if (nlb == 1) {
      if (job.stats) converter_loop<BF16, 1, true>(ca, grp, sh, col_ok, col0[0], osum, osq);
      else converter_loop<BF16, 1, false>(ca, grp, sh, col_ok, col0[0], osum, osq);
    } else if (nlb == 2) {
      if (job.stats) converter_loop<BF16, 2, true>(ca, grp, sh, col_ok, col0[0], osum, osq);    // merged diagonal + first chunk
      else converter_loop<BF16, 2, false>(ca, grp, sh, col_ok, col0[0], osum, osq);
    }
    else converter_loop<BF16, 3, false>(ca, grp, sh, col_ok, col0[0], osum, osq);
  } else {
    // ================================================================ epilogue: TMEM -> this CTA's slab
    // slab [chunk][accumulator lane m (128)][256 columns] fp32, L2 resident.  First window: store;
    // later windows: vector RED (fire-and-forget, round-to-nearest adds in L2).
    const int quadrant = warp & 3;
    float* my_slab = slab + (size_t)blockIdx.x * kSlabFloats;
    const uint32_t acc_empty_leader = ptx::mapa(ptx::smem_u32(&bar_acc_empty[0]), 0);
    int window = 0;
    // a window of fc stages lasts about fc * kt / 8 * nchunk * 384 cycles; spend ~1/3 of it draining
    const unsigned pace_ns = (unsigned)(fc * kt * nchunk * 48 / 3 / 4 / 2) / (BF16 ? 2u : 1u);  // cycles -> ~ns at 2 GHz, per part
    int wpos = stagger;
    for (int it = 0; it < my_chunks; ++it) {
      if (++wpos != fc && it != my_chunks - 1) continue;
      wpos = 0;
      const int set = nacc == 2 ? (window & 1) : 0;
      ptx::mbar_wait(ptx::smem_u32(&bar_acc_full[set]), (uint32_t)(nacc == 2 ? window >> 1 : window) & 1u);
      ptx::tc_fence_after_sync();
      // accumulators to drain: the chunks of an ordinary job; P of a diagonal job and, at its very
      // last window, R as well (stored once into the second slab plane)
      const bool last = it == my_chunks - 1;
      const int ndrain = diag ? (last ? 2 : 1) : nchunk;
      for (int j = 0; j < ndrain; ++j) {
        float* dst_row = my_slab + ((size_t)j * kBlk + quadrant * 32 + lane) * 256;
        const uint32_t tbase = tmem_base + ((uint32_t)(quadrant * 32) << 16) + (uint32_t)(set + j) * 256u;
        const bool store = window == 0 || (diag && j == 1);
#pragma unroll 1
        for (int part = 0; part < 4; ++part) {
          // With two accumulator sets the drain has a whole window to finish: pace it, so that the
          // burst of global stores / REDs does not back up the LSU queues the converters share.
          if (nacc == 2 && part) __nanosleep(pace_ns);
          uint32_t v[32], w[32];
          ptx::tmem_ld_32x32(tbase + (uint32_t)(part * 64), v);
          ptx::tmem_ld_32x32(tbase + (uint32_t)(part * 64 + 32), w);
          ptx::tmem_ld_wait();
          float* dst = dst_row + part * 64;
          if (store) {
#pragma unroll
            for (int q = 0; q < 32; q += 4) {
              asm volatile("st.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(dst + q), "r"(v[q]), "r"(v[q + 1]), "r"(v[q + 2]), "r"(v[q + 3]) : "memory");
              asm volatile("st.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(dst + 32 + q), "r"(w[q]), "r"(w[q + 1]), "r"(w[q + 2]), "r"(w[q + 3]) : "memory");
            }
          } else {
#pragma unroll
            for (int q = 0; q < 32; q += 4) {
              asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst + q), "f"(__uint_as_float(v[q])),
                           "f"(__uint_as_float(v[q + 1])), "f"(__uint_as_float(v[q + 2])), "f"(__uint_as_float(v[q + 3])) : "memory");
              asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst + 32 + q), "f"(__uint_as_float(w[q])),
                           "f"(__uint_as_float(w[q + 1])), "f"(__uint_as_float(w[q + 2])), "f"(__uint_as_float(w[q + 3])) : "memory");
            }
          }
        }
      }
      ptx::tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(acc_empty_leader + (uint32_t)set * 8u);
      ++window;
    }
  }

  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::cluster_sync();   // no CTA of the pair leaves while the other may still signal it or read its smem
  if (warp == kMmaWarp) ptx::tmem_dealloc_pair(tmem_base, kTmemCols);
  if ((trace & 32) && threadIdx.x == 0 && blockIdx.x < 160) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g_cta_time[2 * blockIdx.x + 1] = t;
  }
}

// G (fp64, full symmetric) <- sum over a job's clusters of the slabs of CTA rank r, in a fixed
// order (deterministic).  grid = (njobs * 4 [rank, chunk], 128 lanes m), block = 256 (column nn).
__global__ void __launch_bounds__(256)
reduce_pair_slabs_kernel(const __grid_constant__ PJobTable table, const float* __restrict__ slab, long long n, int d,
                         double* __restrict__ G) {
  const int ji = (int)blockIdx.x >> 2, r = ((int)blockIdx.x >> 1) & 1, j = (int)blockIdx.x & 1;
  const PJob& job = table.job[ji];
  if (j >= job.nchunk) return;
  const int cps = job.nlb == 1 ? 2 : 1;
  const long long n_chunks = (n + kSlotRows * cps - 1) / (kSlotRows * cps);
  const int active = (int)(n_chunks < job.ncl ? n_chunks : job.ncl);
  const int m = (int)blockIdx.y, nn = (int)threadIdx.x;
  const int ga = (2 * job.I + r) * kBlk + m, gb = job.cb[j] * kBlk + nn;
  if (ga >= d || gb >= d) return;
  const size_t cta_stride = 2 * (size_t)kSlabFloats;   // next cluster, same rank
  const float* p = slab + ((size_t)(job.cl_begin * 2 + r)) * kSlabFloats + ((size_t)j * kBlk + m) * 256 + nn;
  // fixed summation order (deterministic), eight loads in flight
  auto sum_plane = [&](const float* q) {
    double acc = 0.0;
    int c = 0;
    for (; c + 8 <= active; c += 8) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = q[(size_t)(c + u) * cta_stride];
#pragma unroll
      for (int u = 0; u < 8; ++u) acc += (double)v[u];
    }
    for (; c < active; ++c) acc += (double)q[(size_t)c * cta_stride];
    return acc;
  };
  if (job.diag) {
    // plane 0 = P = hi^T hi, plane 1 = R = hi^T lo, G = P + R + R^T.  Written here: M = P / 2 + R for EVERY (a, b) of
    // the super-tile (coalesced); symmetrize_diag_kernel then forms G = M + M^T (exactly symmetric).  Reading
    // R[b][a] here instead cost a 32-byte sector per 4 useful bytes, 74 times per element.
    G[(size_t)ga * d + gb] = 0.5 * sum_plane(p) + sum_plane(p + (size_t)kBlk * 256);
    return;
  }
  // the diagonal chunk holds both (a, b) and (b, a), accumulated in different orders: keep the
  // upper one so that G is exactly symmetric
  if (gb < ga) return;
  const double acc = sum_plane(p);
  G[(size_t)ga * d + gb] = acc;
  G[(size_t)gb * d + ga] = acc;
}

// G = M + M^T on the 256 x 256 diagonal super-tiles, in place: block (pair of 32 x 32 tiles (ti, tj), ti <= tj; I).
__global__ void __launch_bounds__(256)
symmetrize_diag_kernel(double* __restrict__ G, int d, unsigned pr_mask) {
  if (!((pr_mask >> blockIdx.y) & 1u)) return;   // this diagonal super-tile was computed in full (merged job): already symmetric
  __shared__ double A[32][33], B[32][33];
  int t = (int)blockIdx.x, ti = 0;
  while (t >= 8 - ti) { t -= 8 - ti; ++ti; }   // 36 pairs: ti = 0..7, tj = ti..7
  const int tj = ti + t;
  const int base = (int)blockIdx.y * 256;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int rr = ty; rr < 32; rr += 8) {
    const int a = base + ti * 32 + rr, b = base + tj * 32 + tx;        // M[a][b]
    A[rr][tx] = (a < d && b < d) ? G[(size_t)a * d + b] : 0.0;
    const int a2 = base + tj * 32 + rr, b2 = base + ti * 32 + tx;      // M[b'][a'] tile
    B[rr][tx] = (a2 < d && b2 < d) ? G[(size_t)a2 * d + b2] : 0.0;
  }
  __syncthreads();
  for (int rr = ty; rr < 32; rr += 8) {
    const int a = base + ti * 32 + rr, b = base + tj * 32 + tx;
    if (a < d && b < d) G[(size_t)a * d + b] = A[rr][tx] + B[tx][rr];
    if (ti != tj) {
      const int a2 = base + tj * 32 + rr, b2 = base + ti * 32 + tx;
      if (a2 < d && b2 < d) G[(size_t)a2 * d + b2] = A[tx][rr] + B[rr][tx];
    }
  }
}

// G += y y^T, s += y, q += y^2 for a few trailing rows (y = x - c in fp32 like the converters, products in fp64).
// grid = (ceil(d / 256), d), block = 256: thread (j, blockIdx.y = i).
__global__ void __launch_bounds__(256)
gram_tail_rows_kernel(const float* __restrict__ X, long long rows, int d, const float* __restrict__ shift, double* __restrict__ G,
                      double* __restrict__ osum, double* __restrict__ osq) {
  const int j = (int)(blockIdx.x * 256 + threadIdx.x), i = (int)blockIdx.y;
  if (j >= d) return;
  const float ci = shift ? shift[i] : 0.f, cj = shift ? shift[j] : 0.f;
  double acc = 0.0, s = 0.0, q = 0.0;
  for (long long r = 0; r < rows; ++r) {
    const double yi = (double)(X[(size_t)r * d + i] - ci), yj = (double)(X[(size_t)r * d + j] - cj);
    acc += yi * yj;
    s += yj;
    q += yj * yj;
  }
  G[(size_t)i * d + j] += acc;
  if (i == 0 && osum != nullptr) {
    atomicAdd(&osum[j], s);
    atomicAdd(&osq[j], q);
  }
}

}  // namespace

// ---------------------------------------------------------------- host side
namespace {

// Relative cost of a job per row (x10): MMA time is 2 : 3 per chunk, but an off-diagonal job converts
// two blocks per 16 rows against a diagonal job's one per 16 rows, and conversion is not free.
// DR4G_OCOST overrides the off-diagonal figure (tuning hook).
// bf16x3: the pass is bound by shared-memory bytes per 16 rows and CTA (TMA write + converter read + converter write +
// operand reads): diagonal 8 + 8 + 8 + 2 x 8 = 40 KB, one off-diagonal chunk 16 + 16 + 16 + 3 x 8 = 72 KB, two chunks
// 24 + 24 + 24 + 6 x 8 = 120 KB (tensor time 256 / 384 / 768 cycles): 20 : 36 : 60.  DR4G_OCOST_BF16 / DR4G_OCOST2_BF16 override.
int job_cost(const PJob& j, bool bf16) {
  static const int ocost = getenv("DR4G_OCOST") ? atoi(getenv("DR4G_OCOST")) : 40;  // 26..42 swept on B200: 40 is best at d = 512 and 1024
  static const int ocost_b = getenv("DR4G_OCOST_BF16") ? atoi(getenv("DR4G_OCOST_BF16")) : 40;
  static const int ocost2_b = getenv("DR4G_OCOST2_BF16") ? atoi(getenv("DR4G_OCOST2_BF16")) : 56;
  static const int mcost_b = getenv("DR4G_MCOST_BF16") ? atoi(getenv("DR4G_MCOST_BF16")) : 52;
  if (j.diag) return 20;
  if (bf16 && j.stats) return mcost_b;     // merged job: full diagonal super-tile + one off-diagonal chunk, two blocks per 16 rows
  if (bf16) return j.nchunk == 2 ? ocost2_b : ocost_b;
  return ocost * j.nchunk;
}

bool build_pair_jobs(int d, int sm_count, bool bf16, PJobTable& tab) {
  const int T = (d + kBlk - 1) / kBlk;
  const int S = (T + 1) / 2;
  const int clusters = sm_count / 2;
  tab.njobs = 0;
  int total_cost = 0;
  auto push = [&](const PJob& j) {
    if (tab.njobs == kMaxPJobs) return false;
    tab.job[tab.njobs++] = j;
    total_cost += job_cost(j, bf16);
    return true;
  };
  // bf16x3, DR4G_MERGE_DIAG: the diagonal super-tile of row super-block I is computed IN FULL (three products, one
  // accumulator) together with the first off-diagonal chunk in one two-chunk job, so the A blocks are loaded and
  // converted once for both (d = 512: 6 instead of 8 block loads per 16 rows, 8 instead of 7 MMAs)
  static const int merge_env = getenv("DR4G_MERGE_DIAG") ? atoi(getenv("DR4G_MERGE_DIAG")) : 0;
  const bool merge = bf16 && merge_env != 0;
  for (int I = 0; I < S; ++I) {
    int c_first = 2 * I + 2;
    if (merge && c_first < T) {
      PJob mj{};
      mj.I = I;
      mj.nchunk = 2;
      mj.cb[0] = 2 * I;
      mj.cb[1] = c_first;
      mj.nlb = 2;
      mj.bb[0] = 2 * I;
      mj.bb[1] = c_first;
      mj.lb_of[0] = 0;
      mj.lb_of[1] = 1;
      mj.stats = 1;
      mj.diag = 0;
      if (!push(mj)) return false;
      c_first += 2;
    } else {
    // diagonal super-tile: A and B are the same blocks
    PJob dj{};
    dj.I = I;
    dj.nchunk = 1;
    dj.cb[0] = 2 * I;
    dj.nlb = 1;
    dj.bb[0] = 2 * I;
    dj.lb_of[0] = 0;
    dj.stats = 1;
    dj.diag = 1;
    if (!push(dj)) return false;
    }
    // super-tiles right of the diagonal, two per job
    for (int c0 = c_first; c0 < T; c0 += 4) {
      PJob j{};
      j.I = I;
      j.nlb = 1;
      j.bb[0] = 2 * I;
      for (int cb = c0; cb < T && cb < c0 + 4; cb += 2) {
        j.cb[j.nchunk] = cb;
        j.bb[j.nlb] = cb;
        j.lb_of[j.nchunk] = j.nlb++;
        ++j.nchunk;
      }
      if (!push(j)) return false;
    }
  }
  if (tab.njobs > clusters) return false;
  int given = 0;
  for (int i = 0; i < tab.njobs; ++i) {
    int m = (int)((long long)clusters * job_cost(tab.job[i], bf16) / total_cost);
    if (m < 1) m = 1;
    tab.job[i].ncl = m;
    given += m;
  }
  // hand the remaining clusters to the jobs with the most work per cluster
  while (given < clusters) {
    int best = 0;
    double worst = -1.0;
    for (int i = 0; i < tab.njobs; ++i) {
      const double load = (double)job_cost(tab.job[i], bf16) / tab.job[i].ncl;
      if (load > worst) { worst = load; best = i; }
    }
    tab.job[best].ncl++;
    given++;
  }
  while (given > clusters) {  // only when njobs is close to the cluster count
    int best = -1;
    double least = 1e30;
    for (int i = 0; i < tab.njobs; ++i) {
      if (tab.job[i].ncl <= 1) continue;
      const double load = (double)job_cost(tab.job[i], bf16) / (tab.job[i].ncl - 1);
      if (load < least) { least = load; best = i; }
    }
    if (best < 0) return false;
    tab.job[best].ncl--;
    given--;
  }
  int begin = 0;
  for (int i = 0; i < tab.njobs; ++i) {
    tab.job[i].cl_begin = begin;
    begin += tab.job[i].ncl;
  }
  return true;
}

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

}  // namespace

bool gram_pair_eligible(const Ctx& c, const float* dX, long long n, int d) {
  if (n < 1 || d <= kBlk || (d % 4) != 0) return false;            // TMA: 16-byte row pitch
  if ((d % 32) != 0 && n < 2) return false;                         // ragged d: the last row goes through gram_tail_rows_kernel
  if ((reinterpret_cast<uintptr_t>(dX) & 15) != 0) return false;    // TMA: 16-byte base
  if (n >= (1ll << 31)) return false;                               // TMA coordinates are int32
  if (c.sm_count < 2) return false;
  PJobTable tab;
  return build_pair_jobs(d, c.sm_count, c.gram_split != 1, tab);
}

size_t gram_pair_workspace_bytes(const Ctx& c, int d) {
  return (size_t)(c.sm_count / 2) * 2 * kSlabFloats * sizeof(float) + 1024;   // + the clusters' row frontiers (pacing)
}

void launch_gram_pair(Ctx& c, const float* dX, long long n, int d, const float* dshift, double* dG, double* dsum,
                      double* dsumsq, void* workspace) {
  const bool bf16 = c.gram_split != 1;   // 0 auto / 2: bf16x3, 1: 3xTF32
  PJobTable tab;
  if (!build_pair_jobs(d, c.sm_count, bf16, tab)) fail("gram_pair: d=%d does not fit the job table", d);
  // X as a 3-D tensor: (column within a 32-column group, row, column group); one box = 32 x 16 x 4
  CUtensorMap tmap;
  // d % 32 != 0 (784, 1000, ...): the last 32-column group is only partly inside a row; its other columns are the first
  // columns of the NEXT row (in bounds, masked by the converters) - except in the very last row, which therefore stays
  // out of the tensor map and is added by gram_tail_rows_kernel.  (These shapes used to fall to the single-CTA kernel:
  // 0.15 TB/s at 500,000 x 516.)
  const long long n_all = n;
  if ((d % 32) != 0) n = n - 1;
  const cuuint64_t gdim[3] = {32, (cuuint64_t)n, (cuuint64_t)((d + 31) / 32)};
  const cuuint64_t gstride[2] = {(cuuint64_t)d * sizeof(float), 32 * sizeof(float)};
  const cuuint32_t box[3] = {32, (cuuint32_t)kSlotRows, 4};
  const cuuint32_t estride[3] = {1, 1, 1};
  CUresult r = reinterpret_cast<EncodeTiledFn>(tensor_map_encoder(c))(
      &tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(dX), gdim, gstride, box, estride,
      CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) fail("cuTensorMapEncodeTiled (3-D) failed (%d) for n=%lld d=%d", (int)r, n, d);

  const int smem = (bf16 ? kRingBytesBf16 : kRingBytes) + 1024;
  static const int op_env = getenv("DR4G_OP_STAGES") ? atoi(getenv("DR4G_OP_STAGES")) : 4;   // operand ring of the bf16x3 kernel (even)
  const int op_stages = op_env >= 8 ? 8 : (op_env >= 6 ? 6 : (op_env >= 4 ? 4 : 2));
  static PerDeviceMax smem_set[2];
  if (smem_set[bf16 ? 1 : 0].need(c.device, (size_t)smem)) {
    if (bf16) DR4G_CUDA(cudaFuncSetAttribute(gram_pair_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    else DR4G_CUDA(cudaFuncSetAttribute(gram_pair_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  }

  static const int trace = getenv("DR4G_GRAM_TRACE") ? atoi(getenv("DR4G_GRAM_TRACE")) : 0;  // bit 0 trace (CTA = bits 8-15, first iteration = bits 16.. x 64), 2 no MMA, 4 no conversion, 16 converter stamps, 32 per-job end times
  const int flush_rows = c.flush_rows > 0 ? c.flush_rows : 1024;
  const int clusters = tab.job[tab.njobs - 1].cl_begin + tab.job[tab.njobs - 1].ncl;
  float* slab = static_cast<float*>(workspace);
  // cross-job pacing: the clusters' row frontiers live behind the slabs; a pair may run ahead of the slowest one by 48 MB
  // of rows (swept 8..96 MB on B200: 8 and 16 MB stall the pipelines - 2x the time at 2^23 x 512 -, 48..96 MB cost
  // nothing; dram reads 2.0x -> 1.16x of the matrix at d = 512, 3.6x -> 1.3x at d = 1024; DR4G_PACE_MB=0 switches it off)
  static const int pace_mb = getenv("DR4G_PACE_MB") ? atoi(getenv("DR4G_PACE_MB")) : 48;
  int* prog = nullptr;
  int pace_rows = 0;
  if (pace_mb > 0 && tab.njobs > 1) {
    prog = reinterpret_cast<int*>(reinterpret_cast<char*>(workspace) + (size_t)(c.sm_count / 2) * 2 * kSlabFloats * sizeof(float));
    DR4G_CUDA(cudaMemsetAsync(prog, 0, 1024, c.stream));
    pace_rows = (int)(((long long)pace_mb << 20) / ((long long)d * 4));
    if (pace_rows < 1024) pace_rows = 1024;
  }
  cudaLaunchConfig_t lc = {};
  lc.gridDim = dim3(2 * clusters, 1, 1);
  lc.blockDim = dim3(kThreads, 1, 1);
  lc.dynamicSmemBytes = smem;
  lc.stream = c.stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  lc.attrs = at;
  lc.numAttrs = 1;
  if (bf16)
    DR4G_CUDA(cudaLaunchKernelEx(&lc, gram_pair_kernel<true>, tmap, tab, n, d, dshift, slab, dsum, dsumsq, flush_rows, trace, prog, pace_rows, op_stages));
  else
    DR4G_CUDA(cudaLaunchKernelEx(&lc, gram_pair_kernel<false>, tmap, tab, n, d, dshift, slab, dsum, dsumsq, flush_rows, trace, prog, pace_rows, op_stages));
  c.last_gram_kernel = bf16 ? 3 : 2;
  c.launches++;
  if (trace & 32) {
    static unsigned long long ht[2 * 160];
    DR4G_CUDA(cudaStreamSynchronize(c.stream));
    DR4G_CUDA(cudaMemcpyFromSymbol(ht, g_cta_time, sizeof(ht)));
    unsigned long long t0 = ~0ull;
    for (int b = 0; b < 2 * clusters; ++b) if (ht[2 * b] < t0) t0 = ht[2 * b];
    fprintf(stderr, "gram_pair job times (us after the first CTA started; n=%lld d=%d)\n", n, d);
    for (int j = 0; j < tab.njobs; ++j) {
      double lo = 1e30, hi = 0, sum = 0;
      for (int cl = tab.job[j].cl_begin; cl < tab.job[j].cl_begin + tab.job[j].ncl; ++cl) {
        const double e = (double)(ht[2 * (2 * cl) + 1] - t0) * 1e-3;
        lo = e < lo ? e : lo; hi = e > hi ? e : hi; sum += e;
      }
      fprintf(stderr, "  job %2d I=%d diag=%d chunks=%d clusters=%2d: end min %8.1f mean %8.1f max %8.1f\n", j, tab.job[j].I,
              tab.job[j].diag, tab.job[j].nchunk, tab.job[j].ncl, lo, sum / tab.job[j].ncl, hi);
    }
  }
  if (trace & 1) {
    static long long h[kTraceIts * 8];
    DR4G_CUDA(cudaStreamSynchronize(c.stream));
    DR4G_CUDA(cudaMemcpyFromSymbol(h, g_trace2, sizeof(h)));
    fprintf(stderr, "gram_pair trace (CTA 0, clocks relative to chunk 0 'stage free'; n=%lld d=%d)\n", n, d);
    fprintf(stderr, (trace & 16) ? "  it  alu_done  op_free    landed converted  arrived  mma_ready mma_issued  loop_top\n"
                                 : "  it  tma_free tma_issued  landed converted  arrived  mma_ready mma_issued  loop_top\n");
    long long base = 0;
    for (int e = 0; e < kTraceIts * 8; ++e) if (h[e] && (base == 0 || h[e] < base)) base = h[e];
    for (int it = 0; it < kTraceIts; ++it) {
      fprintf(stderr, "%4d", it);
      for (int e = 0; e < 8; ++e) fprintf(stderr, " %9lld", h[it * 8 + e] ? h[it * 8 + e] - base : -1ll);
      fprintf(stderr, "\n");
    }
  }
  reduce_pair_slabs_kernel<<<dim3(tab.njobs * 4, kBlk), 256, 0, c.stream>>>(tab, slab, n, d, dG);
  unsigned pr_mask = 0;
  for (int i = 0; i < tab.njobs; ++i) if (tab.job[i].diag) pr_mask |= 1u << tab.job[i].I;
  symmetrize_diag_kernel<<<dim3(36, (d + 255) / 256), 256, 0, c.stream>>>(dG, d, pr_mask);
  c.launches++;
  c.launches++;
  if (n < n_all) {
    gram_tail_rows_kernel<<<dim3((d + 255) / 256, d), 256, 0, c.stream>>>(dX + (size_t)n * d, n_all - n, d, dshift, dG, dsum, dsumsq);
    c.launches++;
  }
  DR4G_CUDA(cudaGetLastError());
}

}  // namespace dr4g
