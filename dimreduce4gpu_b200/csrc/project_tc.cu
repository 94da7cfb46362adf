// K7 on the tensor cores - streamed projection  Z = (X - mean) Q^T,  U = Z / sigma   (fp32 X, k <= 128, d % 4 == 0).
//
// Replaces calculate_u + multiply_diag (reference src/tsvd.cu:152-177,229-260: cublasSgemm X V, a divide-by-sigma
// lambda, dgmm by sigma again) and the host-side numpy transform (dimreduce4gpu/truncated_svd.py:116-118).  The CUDA-core
// kernel in project.cu was bound by FFMA issue (0.31 of HBM at k = 32); with AI = k/2 flop/B the projection is
// HBM-bound once the multiply-adds run on tcgen05.
//
// Shape.  D[M x N] += A[M x K] B[K x N] with M = 128 rows of X (one tile), K = features, N = components: X is row-major,
// so A is K-major, and components_ (k x d row-major) is a K-major B.  fp32 accuracy comes from the same two-term bf16
// split as the Gram (gram_tc2.cu): x - mean = a1 + a2, q = b1 + b2, products a1 b1 + a1 b2 + a2 b1 (dropped a2 b2 and
// third terms ~2^-17 relative per product, zero mean; measured ~5e-6 of |z|).  B is stacked [b1 ; b2] along N, so one
// instruction A1 x [B1 | B2] (N = 2 KP) covers two of the three products and A2 x B1 (N = KP) the third: each operand
// tile of A is read from shared memory once.  The epilogue adds the two accumulator halves.
//
// Data path (persistent CTA per SM, 13 warps):
//   warps 0-7   loaders/converters: every thread copies ITS 8 x 16 B of a stage global -> shared with cp.async (L2 only)
//               into a thread-private landing ring three stages deep (96 KB in flight per SM; a register prefetch of one
//               stage, 32 KB, left the kernel at 0.64 of HBM), reads them back, subtracts the mean, splits into bf16
//               hi / lo and stores into the K-major SWIZZLE_128B operand tiles (128 rows x 64 columns per stage).  The
//               landing data is private to the thread that copied it: no barrier, only cp.async.wait_group
//   warp 8      one lane: bulk-copies the stage's slice of the pre-split components image (cp.async.bulk, 2 KP rows x
//               128 B) and issues the tcgen05.mma (kind::f16) of the stage; owns the TMEM allocation
//   warps 9-12  epilogue: tcgen05.ld of the finished 128 x 2 KP accumulator (double-buffered in TMEM), z = D1 + D2,
//               128-bit stores of Z and U rows
// Shared memory: operand ring (A hi 16 KB + A lo 16 KB + B 2 KP x 128 B per stage, 2-3 stages) + landing ring
// (3 x 32 KB).  Shared-memory traffic is ~18 B per element of X (landing write + read, operand write + read, B slices).
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace dr4g {
namespace {

constexpr int PT_ROWS = 128;          // rows per tile = MMA M
constexpr int PT_KC = 64;             // feature columns per stage (one 128-byte bf16 row)
constexpr int PT_CONV_WARPS = 8;
constexpr int PT_MMA_WARP = 8;
constexpr int PT_EPI_WARPS = 4;
constexpr int PT_THREADS = (PT_CONV_WARPS + 1 + PT_EPI_WARPS) * 32;   // 416
constexpr int PT_MAX_STAGES = 4;
constexpr int PT_PF = 3;                                              // landing ring depth (stages in flight)
constexpr uint32_t PT_LAND_BYTES = PT_ROWS * PT_KC * 4;              // 32 KB per landing stage
constexpr uint32_t PT_A_BYTES = PT_ROWS * PT_KC * 2;                  // 16 KB per half

__device__ __forceinline__ void bulk_copy_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

__device__ __forceinline__ void pt_cp_async16(uint32_t dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void pt_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void pt_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Q (k x d fp32) -> per 64-column chunk the shared-memory image of B = [hi rows 0..KP-1 ; lo rows KP..2KP-1], each row
// 64 bf16 = 128 B, K-major SWIZZLE_128B (row r at (r / 8) * 1024 + (r % 8) * 128, 16-byte chunk c at c ^ (r % 8)).
// Rows >= k of each half and columns >= d are zero.  grid = chunks, block = 256.
__global__ void __launch_bounds__(256)
project_prep_kernel(const float* __restrict__ Q, int d, int k, int KP, uint8_t* __restrict__ img) {
  const int c = blockIdx.x;
  const size_t chunk_bytes = (size_t)2 * KP * 128;
  uint8_t* out = img + (size_t)c * chunk_bytes;
  for (int e = threadIdx.x; e < KP * 16; e += blockDim.x) {   // (row, float4 index)
    const int row = e >> 4, c4 = e & 15;
    const int col = c * PT_KC + c4 * 4;
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row < k && col < d) q = *reinterpret_cast<const float4*>(Q + (size_t)row * d + col);   // d % 4 == 0
    const uint32_t h01 = ptx::pack_bf16x2(q.x, q.y), h23 = ptx::pack_bf16x2(q.z, q.w);
    const uint32_t l01 = ptx::pack_bf16x2(q.x - __uint_as_float(h01 << 16), q.y - __uint_as_float(h01 & 0xffff0000u));
    const uint32_t l23 = ptx::pack_bf16x2(q.z - __uint_as_float(h23 << 16), q.w - __uint_as_float(h23 & 0xffff0000u));
    auto off = [&](int r) {
      return (size_t)(r >> 3) * 1024 + (size_t)(r & 7) * 128 + (size_t)((((c4 >> 1) ^ (r & 7)) << 4) | ((c4 & 1) << 3));
    };
    *reinterpret_cast<uint2*>(out + off(row)) = make_uint2(h01, h23);
    *reinterpret_cast<uint2*>(out + off(KP + row)) = make_uint2(l01, l23);
  }
}

template <int KP>
__global__ void __launch_bounds__(PT_THREADS, 1)
project_tc_kernel(const float* __restrict__ X, long long n, int d, int k, const float* __restrict__ mean,
                  const uint8_t* __restrict__ bimg, const float* __restrict__ sigma, float* __restrict__ Z,
                  float* __restrict__ U, int nstages, int vec4, int ldz) {
  constexpr uint32_t B_BYTES = 2 * KP * 128;                  // [hi ; lo] slice of the components image
  constexpr uint32_t STAGE_BYTES = 2 * PT_A_BYTES + B_BYTES;
  constexpr uint32_t ACC_COLS = 2 * KP;                       // one accumulator buffer: [a1 b1 + a2 b1 | a1 b2]
  constexpr uint32_t TMEM_COLS = 2 * ACC_COLS < 32 ? 32 : 2 * ACC_COLS;   // power of two >= 32: 128 / 256 / 512

  extern __shared__ uint8_t psmem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(psmem_raw) + 1023u) & ~1023u;
  __shared__ __align__(8) uint64_t bar_full[PT_MAX_STAGES], bar_empty[PT_MAX_STAGES], bar_acc_full[2], bar_acc_empty[2];
  __shared__ uint32_t tmem_base_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long ntiles = (n + PT_ROWS - 1) / PT_ROWS;
  const int nch = (d + PT_KC - 1) / PT_KC;
  const long long my_tiles = ntiles > (long long)blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < PT_MAX_STAGES; ++s) {
      ptx::mbar_init(ptx::smem_u32(&bar_full[s]), PT_CONV_WARPS + 1);   // 8 converter warps + the B slice (expect_tx)
      ptx::mbar_init(ptx::smem_u32(&bar_empty[s]), 1);                  // tcgen05.commit
    }
    for (int a = 0; a < 2; ++a) {
      ptx::mbar_init(ptx::smem_u32(&bar_acc_full[a]), 1);
      ptx::mbar_init(ptx::smem_u32(&bar_acc_empty[a]), PT_EPI_WARPS);
    }
    ptx::fence_barrier_init();
  }
  if (warp == PT_MMA_WARP) {
    ptx::tmem_alloc(ptx::smem_u32(&tmem_base_slot), TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = tmem_base_slot;

  if (warp < PT_CONV_WARPS) {
    // ================================================================ loaders / converters
    // thread -> float4 c4 (0..15) of the 64-column chunk and rows r + 16 j (j = 0..7): a half-warp reads one row's
    // 256 B, a warp two rows (full 128-byte lines), and a half-warp's 8-byte stores cover one 128-byte operand row
    const int tid = threadIdx.x;
    const int c4 = tid & 15, r = tid >> 4;
    uint32_t soff[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int row = r + 16 * j;
      soff[j] = (uint32_t)(row >> 3) * 1024u + (uint32_t)(row & 7) * 128u +
                (uint32_t)((((c4 >> 1) ^ (row & 7)) << 4) | ((c4 & 1) << 3));
    }
    // landing ring behind the operand ring: slot p holds [j][thread] float4 (a warp's 32 x 16 B are contiguous)
    const uint32_t land_base = smem_base + (uint32_t)nstages * STAGE_BYTES;
    const long long total = my_tiles * nch;
    // (tile, chunk) of the next stage to copy
    long long ftile = blockIdx.x;
    int fc = 0;
    long long fit = 0;
    auto issue = [&]() {
      if (fit < total) {
        const int col = fc * PT_KC + c4 * 4;
        const bool cin = col < d;                          // d % 4 == 0: a float4 is inside or outside as a whole
        const long long row0 = ftile * PT_ROWS + r;
        const float* base = X + row0 * (long long)d + (cin ? col : 0);
        const uint32_t dst = land_base + (uint32_t)(fit % PT_PF) * PT_LAND_BYTES + (uint32_t)tid * 16u;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          // rows past n / columns past d are zero-filled; a row past n only feeds an output row that is never stored
          const bool ok = cin && row0 + 16 * j < n;
          pt_cp_async16(dst + (uint32_t)j * 4096u, ok ? base + (long long)(16 * j) * d : X, ok ? 16 : 0);
        }
        if (++fc == nch) { fc = 0; ftile += gridDim.x; }
      }
      pt_cp_async_commit();   // one group per stage, possibly empty: keeps the wait count uniform
      ++fit;
    };
#pragma unroll
    for (int p = 0; p < PT_PF; ++p) issue();
    int c = 0, s = 0;
    uint32_t ph = 0;
    for (long long it = 0; it < total; ++it) {
      pt_cp_async_wait<PT_PF - 1>();                       // this thread's copies of stage `it` have landed
      const uint32_t src = land_base + (uint32_t)(it % PT_PF) * PT_LAND_BYTES + (uint32_t)tid * 16u;
      float4 x[8];
#pragma unroll
      for (int j = 0; j < 8; ++j)
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x[j].x), "=f"(x[j].y), "=f"(x[j].z), "=f"(x[j].w) : "r"(src + (uint32_t)j * 4096u));
      const int col = c * PT_KC + c4 * 4;
      const float4 mu = (mean && col < d) ? __ldg(reinterpret_cast<const float4*>(mean + col)) : make_float4(0.f, 0.f, 0.f, 0.f);
      uint32_t hp[8][2], lp[8][2];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float y0 = x[j].x - mu.x, y1 = x[j].y - mu.y, y2 = x[j].z - mu.z, y3 = x[j].w - mu.w;
        hp[j][0] = ptx::pack_bf16x2(y0, y1);
        hp[j][1] = ptx::pack_bf16x2(y2, y3);
        lp[j][0] = ptx::pack_bf16x2(y0 - __uint_as_float(hp[j][0] << 16), y1 - __uint_as_float(hp[j][0] & 0xffff0000u));
        lp[j][1] = ptx::pack_bf16x2(y2 - __uint_as_float(hp[j][1] << 16), y3 - __uint_as_float(hp[j][1] & 0xffff0000u));
      }
      issue();                                             // refill this landing slot (its data is in registers now)
      ptx::mbar_wait(ptx::smem_u32(&bar_empty[s]), ph ^ 1u);
      const uint32_t sa = smem_base + (uint32_t)s * STAGE_BYTES;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        asm volatile("st.shared.v2.b32 [%0], {%1,%2};" ::"r"(sa + soff[j]), "r"(hp[j][0]), "r"(hp[j][1]) : "memory");
        asm volatile("st.shared.v2.b32 [%0], {%1,%2};" ::"r"(sa + PT_A_BYTES + soff[j]), "r"(lp[j][0]), "r"(lp[j][1]) : "memory");
      }
      ptx::fence_proxy_async_smem();   // generic-proxy stores -> visible to the tensor core's async-proxy reads
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&bar_full[s]));
      if (++s == nstages) { s = 0; ph ^= 1u; }
      if (++c == nch) c = 0;
    }
    pt_cp_async_wait<0>();
  } else if (warp == PT_MMA_WARP) {
    // ================================================================ B slices + MMA issue (one lane)
    if (ptx::elect_one()) {
      constexpr uint32_t idesc_wide = ptx::make_idesc_bf16(PT_ROWS, 2 * KP, false, false);   // A1 x [B1 | B2]
      constexpr uint32_t idesc_narrow = ptx::make_idesc_bf16(PT_ROWS, KP, false, false);     // A2 x B1
      const uint64_t desc0 = ptx::make_smem_desc(0, 16, 1024, ptx::kLayoutSw128);            // K-major: SBO = 8 rows x 128 B
      const long long total = my_tiles * nch;
      // B slices run up to nstages ahead of the MMAs: slice it goes into stage it % nstages once that stage is free
      long long issued = 0;
      int bs = 0, bc = 0;
      uint32_t bph = 0;
      auto issue_b = [&]() {
        ptx::mbar_wait(ptx::smem_u32(&bar_empty[bs]), bph ^ 1u);
        const uint32_t full = ptx::smem_u32(&bar_full[bs]);
        ptx::mbar_arrive_expect_tx(full, B_BYTES);
        bulk_copy_g2s(smem_base + (uint32_t)bs * STAGE_BYTES + 2 * PT_A_BYTES, bimg + (size_t)bc * B_BYTES, B_BYTES, full);
        if (++bs == nstages) { bs = 0; bph ^= 1u; }
        if (++bc == nch) bc = 0;
        ++issued;
      };
      int s = 0, c = 0, acc = 0;
      uint32_t ph = 0, aph[2] = {0u, 0u};
      long long tiles_done = 0;
      for (long long it = 0; it < total; ++it) {
        while (issued < total && issued < it + nstages) issue_b();
        if (c == 0 && tiles_done >= 2) {
          // the epilogue warps have drained this accumulator buffer
          ptx::mbar_wait(ptx::smem_u32(&bar_acc_empty[acc]), aph[acc]);
          aph[acc] ^= 1u;
          ptx::tc_fence_after_sync();
        }
        ptx::mbar_wait(ptx::smem_u32(&bar_full[s]), ph);
        ptx::tc_fence_after_sync();
        const uint32_t sa = smem_base + (uint32_t)s * STAGE_BYTES;
        const uint32_t d0 = tmem_base + (uint32_t)acc * ACC_COLS;
#pragma unroll
        for (int ks = 0; ks < PT_KC / 16; ++ks) {
          const uint32_t koff = (uint32_t)ks * 32u;   // 16 bf16 along K inside the 128-byte swizzled row
          const uint64_t a_hi = desc0 | (uint64_t)(((sa + koff) >> 4) & 0x3FFF);
          const uint64_t a_lo = desc0 | (uint64_t)(((sa + PT_A_BYTES + koff) >> 4) & 0x3FFF);
          const uint64_t b = desc0 | (uint64_t)(((sa + 2 * PT_A_BYTES + koff) >> 4) & 0x3FFF);
          ptx::mma_f16_ss(d0, a_hi, b, idesc_wide, (c == 0 && ks == 0) ? 0u : 1u);
          ptx::mma_f16_ss(d0, a_lo, b, idesc_narrow, 1u);
        }
        ptx::mma_commit(ptx::smem_u32(&bar_empty[s]));           // stage free once these MMAs have read it
        if (++c == nch) {
          c = 0;
          ptx::mma_commit(ptx::smem_u32(&bar_acc_full[acc]));    // tile finished
          acc ^= 1;
          ++tiles_done;
        }
        if (++s == nstages) { s = 0; ph ^= 1u; }
      }
    }
  } else {
    // ================================================================ epilogue: TMEM -> Z, U
    const int q = warp & 3;                               // TMEM lane quadrant this warp may read
    int acc = 0;
    uint32_t aph[2] = {0u, 0u};
    long long tile = blockIdx.x;
    for (long long t = 0; t < my_tiles; ++t, tile += gridDim.x) {
      ptx::mbar_wait(ptx::smem_u32(&bar_acc_full[acc]), aph[acc]);
      aph[acc] ^= 1u;
      ptx::tc_fence_after_sync();
      const long long row = tile * PT_ROWS + q * 32 + lane;
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)acc * ACC_COLS;
#pragma unroll
      for (int part = 0; part < KP / 32; ++part) {
        uint32_t v[32], w[32];
        ptx::tmem_ld_32x32(tbase + (uint32_t)(part * 32), v);
        ptx::tmem_ld_32x32(tbase + (uint32_t)(KP + part * 32), w);
        ptx::tmem_ld_wait();
        if (row < n) {
          float* zr = Z + row * (long long)ldz + part * 32;      // ldz: row pitch of Z / U (>= k: a column slice of wider scores)
          float* ur = U ? U + row * (long long)ldz + part * 32 : nullptr;
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const int m = part * 32 + j;
            if (m >= k) break;
            float z0 = __uint_as_float(v[j]) + __uint_as_float(w[j]);
            float z1 = __uint_as_float(v[j + 1]) + __uint_as_float(w[j + 1]);
            float z2 = __uint_as_float(v[j + 2]) + __uint_as_float(w[j + 2]);
            float z3 = __uint_as_float(v[j + 3]) + __uint_as_float(w[j + 3]);
            if (vec4) {        // k % 4 == 0 and 16-byte aligned outputs: m + 3 < k
              *reinterpret_cast<float4*>(zr + j) = make_float4(z0, z1, z2, z3);
              if (ur) {
                const float4 sg = __ldg(reinterpret_cast<const float4*>(sigma + m));
                *reinterpret_cast<float4*>(ur + j) = make_float4(sg.x != 0.f ? z0 / sg.x : 0.f, sg.y != 0.f ? z1 / sg.y : 0.f,
                                                                 sg.z != 0.f ? z2 / sg.z : 0.f, sg.w != 0.f ? z3 / sg.w : 0.f);
              }
            } else {           // any k (2, 3, 10, 50, ...): the row pitch is not a multiple of 16 bytes
              const float zz[4] = {z0, z1, z2, z3};
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (m + e < k) {
                  zr[j + e] = zz[e];
                  if (ur) { const float sg = __ldg(sigma + m + e); ur[j + e] = sg != 0.f ? zz[e] / sg : 0.f; }
                }
            }
          }
        }
      }
      ptx::tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&bar_acc_empty[acc]));
      acc ^= 1;
    }
  }

  ptx::tc_fence_before_sync();
  __syncthreads();
  if (warp == PT_MMA_WARP) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}

template <int KP>
void launch_kp(Ctx& c, const float* dX, long long n, int d, int k, const float* dmean, const uint8_t* img, const float* dw,
               float* dZ, float* dU, int ldz) {
  constexpr uint32_t stage = 2 * PT_A_BYTES + 2 * KP * 128;
  constexpr uint32_t land = PT_PF * PT_LAND_BYTES;
  int nstages = (int)((225u * 1024u - land) / stage);   // 3 / 2 / 2 operand stages at KP = 32 / 64 / 128
  if (nstages > PT_MAX_STAGES) nstages = PT_MAX_STAGES;
  const size_t smem = (size_t)nstages * stage + land + 1024;
  static PerDeviceMax smem_set;
  if (smem_set.need(c.device, smem))
    DR4G_CUDA(cudaFuncSetAttribute(project_tc_kernel<KP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long ntiles = (n + PT_ROWS - 1) / PT_ROWS;
  const int grid = (int)(ntiles < c.sm_count ? ntiles : c.sm_count);
  const int vec4 = (k % 4) == 0 && (ldz % 4) == 0 && (reinterpret_cast<uintptr_t>(dZ) & 15) == 0 && (!dU || (reinterpret_cast<uintptr_t>(dU) & 15) == 0) &&
                   (!dw || (reinterpret_cast<uintptr_t>(dw) & 15) == 0);
  project_tc_kernel<KP><<<grid, PT_THREADS, smem, c.stream>>>(dX, n, d, k, dmean, img, dw, dZ, dU, nstages, vec4, ldz);
}

}  // namespace

bool project_tc_eligible(const Ctx& c, const float* dX, long long n, int d, int k, const float* dZ, const float* dU) {
  (void)dZ;
  (void)dU;
  if (n < 1 || d < 4 || (d % 4) != 0 || k < 1 || k > 128) return false;     // any k: unaligned output rows take scalar stores
  return (reinterpret_cast<uintptr_t>(dX) & 15) == 0;
}

void launch_project_tc(Ctx& c, const float* dX, long long n, int d, int k, const float* dQ, const float* dmean,
                       const float* dw, float* dZ, float* dU, int ldz) {
  if (ldz <= 0) ldz = k;
  const int KP = k <= 32 ? 32 : (k <= 64 ? 64 : 128);
  const int nch = (d + PT_KC - 1) / PT_KC;
  const size_t img_bytes = (size_t)nch * 2 * KP * 128;
  if (img_bytes > c.proj_cap) {   // grow-only, private to the projection (the fit arena may be in use by the same stream)
    if (c.proj_buf) {
      DR4G_CUDA(cudaStreamSynchronize(c.stream));
      DR4G_CUDA(cudaFree(c.proj_buf));
      c.proj_buf = nullptr;
      c.proj_cap = 0;
    }
    DR4G_CUDA(cudaMalloc(&c.proj_buf, img_bytes + (img_bytes >> 2)));
    c.proj_cap = img_bytes + (img_bytes >> 2);
  }
  uint8_t* img = static_cast<uint8_t*>(c.proj_buf);
  project_prep_kernel<<<nch, 256, 0, c.stream>>>(dQ, d, k, KP, img);
  const float* sg = (dU && dw) ? dw : nullptr;
  float* du = (dU && dw) ? dU : nullptr;
  if (KP == 32) launch_kp<32>(c, dX, n, d, k, dmean, img, sg, dZ, du, ldz);
  else if (KP == 64) launch_kp<64>(c, dX, n, d, k, dmean, img, sg, dZ, du, ldz);
  else launch_kp<128>(c, dX, n, d, k, dmean, img, sg, dZ, du, ldz);
  c.launches += 2;
  DR4G_CUDA(cudaGetLastError());
}

}  // namespace dr4g
