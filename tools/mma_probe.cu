// GPU probe (timing only, results are not checked): how fast does tcgen05.mma kind::tf32 run on
// B200 for the operand placements the Gram kernel can choose from, alone and with converter-like
// shared-memory traffic beside it?   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o
// tools/build/mma_probe tools/mma_probe.cu ; run under gpurun.
//
// variants: A operand from smem MN-major (BASE32B), smem K-major (SW128) or TMEM; B MN-major or
// K-major; N = 128 / 256; cta_group 1 / 2; `noise` warps looping LDS.128 + 2 x STS.128.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../dimreduce4gpu_b200/csrc/ptx_sm100.cuh"

using namespace dr4g;

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e_ = (x);                                                          \
    if (e_ != cudaSuccess) {                                                       \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(1);                                                                     \
    }                                                                              \
  } while (0)

struct Cfg {
  int a_mode;   // 0 smem MN-major, 1 smem K-major, 2 TMEM
  int b_mode;   // 0 MN-major, 1 K-major
  int n;        // MMA N
  int cg;       // cta_group
  int per_step; // MMAs per k-step
  int steps;    // k-steps per commit group (one "stage")
  int groups;   // commit groups
  int noise;    // noise warps (0..8)
  int noise_sts; // STS.128 per LDS.128 in the noise loop
  int kind;     // 0 tf32, 1 bf16 (K = 16, K-major timing only)
  int drot;     // accumulator rotation: 0 every 3 MMAs, 1 every MMA, 2 never
};

template <int cg>
__device__ __forceinline__ void mma_ss_f16(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc) {
  if constexpr (cg == 1)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
}
template <int cg>
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc) {
  if constexpr (cg == 1)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
}
template <int cg>
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc) {
  if constexpr (cg == 1)
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(b), "r"(idesc) : "memory");
  else
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, 1, 0;\n\ttcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(b), "r"(idesc) : "memory");
}
template <int cg>
__device__ __forceinline__ void commit(uint32_t bar) {
  if constexpr (cg == 1)
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
  else
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

constexpr int kStage = 65536;
constexpr int kStages = 3;

template <int CG, int AM, int BM, int NN, int KIND, int DROT, int PS>
__global__ void __launch_bounds__(320, 1) probe(Cfg c0, long long* out_cycles, long long* out_noise, volatile int* gflag) {
  Cfg c = c0;
  c.cg = CG;
  c.steps = 4;
  extern __shared__ uint8_t raw[];
  const uint32_t base = (ptx::smem_u32(raw) + 1023u) & ~1023u;
  __shared__ __align__(8) uint64_t bar[4];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t rank = 0;
  if constexpr (CG == 2) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));

  // fill shared memory with small tf32-exact numbers
  for (uint32_t i = threadIdx.x; i < kStages * kStage / 4; i += blockDim.x)
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(base + i * 4), "f"((float)((i * 2654435761u >> 24) & 7) * 0.125f));
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) ptx::mbar_init(ptx::smem_u32(&bar[i]), 1);
    ptx::fence_barrier_init();
  }
  if (warp == 9) {
    if constexpr (CG == 1) {
      ptx::tmem_alloc(ptx::smem_u32(&tmem_slot), 512);
      ptx::tmem_relinquish();
    } else {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(ptx::smem_u32(&tmem_slot)), "r"(512) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
  }
  ptx::fence_proxy_async_smem();
  ptx::tc_fence_before_sync();
  __syncthreads();
  if constexpr (CG == 2) {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
  ptx::tc_fence_after_sync();
  const uint32_t tmem = tmem_slot;

  if (warp == 9) {
    if (rank == 0 && ptx::elect_one()) {
      constexpr bool a_mn = AM == 0, b_mn = BM == 0;
      constexpr int m = CG == 2 ? 256 : 128;
      constexpr uint32_t idesc = KIND == 1 ? ((1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NN >> 3) << 17) | ((uint32_t)(m >> 4) << 24))
                                           : ptx::make_idesc_tf32(m, NN, a_mn, b_mn);
      constexpr uint32_t lbo = 32u * 128u;
      constexpr int ntile = AM == 2 ? (512 - 64) / NN : 512 / NN;
      const uint32_t a_tmem = tmem + 448;
      long long t0 = clock64();
      for (int g = 0; g < c.groups; ++g) {
        const uint32_t sx = base + (uint32_t)(g % kStages) * kStage;
        if (g >= 2) ptx::mbar_wait(ptx::smem_u32(&bar[g & 1]), (uint32_t)((g - 2) >> 1) & 1u);
        const uint64_t da = a_mn ? ptx::make_smem_desc(sx, lbo, 512, ptx::kLayoutSw128Base32) : ptx::make_smem_desc(sx, 16, 1024, ptx::kLayoutSw128);
        const uint64_t db = b_mn ? ptx::make_smem_desc(sx, lbo, 512, ptx::kLayoutSw128Base32) : ptx::make_smem_desc(sx, 16, 1024, ptx::kLayoutSw128);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
          for (int i = 0; i < PS; ++i) {
            constexpr uint32_t dummy = 0; (void)dummy;
            const uint32_t half = (i % 3 == 1) ? 32768u : 0u;
            const uint32_t halfb = (i % 3 == 2) ? 32768u : 0u;
            const uint32_t ablk = (uint32_t)((i / 3) & 1), bblk = (uint32_t)(((i / 3) >> 1) & 1);
            const uint32_t aoff = half + ablk * 16384u + (uint32_t)ks * (a_mn ? 1024u : 32u);
            const uint32_t boff = halfb + (NN == 256 ? 0u : bblk * 16384u) + (uint32_t)ks * (b_mn ? 1024u : 32u);
            const int di = DROT == 0 ? i / 3 : (DROT == 1 ? i : 0);
            const uint32_t dcol = tmem + (uint32_t)(di % ntile) * (uint32_t)NN;
            if (AM == 2) mma_ts<CG>(dcol, a_tmem + (uint32_t)(ks & 7) * 8u, db + (boff >> 4), idesc);
            else if (KIND == 1) mma_ss_f16<CG>(dcol, da + (aoff >> 4), db + (boff >> 4), idesc);
            else mma_ss<CG>(dcol, da + (aoff >> 4), db + (boff >> 4), idesc);
          }
        }
        commit<CG>(ptx::smem_u32(&bar[g & 1]));
      }
      for (int g = (c.groups >= 2 ? c.groups - 2 : 0); g < c.groups; ++g)
        ptx::mbar_wait(ptx::smem_u32(&bar[g & 1]), (uint32_t)(g >> 1) & 1u);
      long long t1 = clock64();
      out_cycles[blockIdx.x] = t1 - t0;
      gflag[blockIdx.x / c.cg] = 1;
      __threadfence();
    } else if (rank != 0 && lane == 0) {
      // peer CTA of a pair: wait until the leader is done (cluster barrier below)
    }
    __syncwarp();
    if (rank != 0 || c.cg == 1) { /* nothing */ }
  } else if (warp < c.noise) {
    // converter-like traffic: LDS.128 + noise_sts x STS.128 per iteration on the stage buffers
    long long it = 0;
    uint32_t a = base + (uint32_t)threadIdx.x * 16u;
    float4 acc = make_float4(0, 0, 0, 0);
    volatile int* fl = gflag + blockIdx.x / c.cg;
    while (it < (1ll << 22)) {
      if ((it & 3) == 0 && *fl) break;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float4 x;
        const uint32_t addr = a + (uint32_t)((it * 8 + u) & 15) * 4096u + (uint32_t)(u & 1) * 65536u;
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w) : "r"(addr));
        acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
        x.x = __uint_as_float(__float_as_uint(x.x) & 0xffffe000u);
        for (int s = 0; s < c.noise_sts; ++s)
          asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr + (uint32_t)s * 32768u), "f"(x.x), "f"(x.y), "f"(x.z), "f"(x.w) : "memory");
      }
      ++it;
    }
    if (lane == 0) atomicAdd((unsigned long long*)&out_noise[blockIdx.x], (unsigned long long)it);
    if (acc.x == 123.456f) out_noise[0] = 0;
  }
  ptx::tc_fence_before_sync();
  __syncthreads();
  if constexpr (CG == 2) {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
  if (warp == 9) {
    if constexpr (CG == 1) ptx::tmem_dealloc(tmem, 512);
    else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
  }
}

struct Named { const char* name; Cfg c; const void* fn; };
template <int CG, int AM, int BM, int NN, int KIND, int DROT, int PS>
Named mk(const char* name, int noise, int ns) {
  const int smem = kStages * kStage + 1024;
  auto* f = probe<CG, AM, BM, NN, KIND, DROT, PS>;
  CK(cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  return {name, {AM, BM, NN, CG, PS, 4, 64, noise, ns, KIND, DROT}, (const void*)f};
}

int main(int argc, char** argv) {
  int sm = 0;
  CK(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0));
  const int smem = kStages * kStage + 1024;
  long long *d_cyc, *d_noise;
  int* d_flag;
  CK(cudaMalloc(&d_flag, sm * sizeof(int)));
  CK(cudaMalloc(&d_cyc, sm * sizeof(long long)));
  CK(cudaMalloc(&d_noise, sm * sizeof(long long)));
  std::vector<Named> v;
  // --- issue / dependency structure, no noise
  v.push_back(mk<1, 1, 1, 32, 0, 0, 12>("tf32 SS K/K   N= 32 cg1 drot0", 0, 2));
  v.push_back(mk<1, 1, 1, 64, 0, 0, 12>("tf32 SS K/K   N= 64 cg1 drot0", 0, 2));
  v.push_back(mk<1, 1, 1, 128, 0, 0, 12>("tf32 SS K/K   N=128 cg1 drot0", 0, 2));
  v.push_back(mk<1, 1, 1, 256, 0, 0, 12>("tf32 SS K/K   N=256 cg1 drot0", 0, 2));
  v.push_back(mk<1, 1, 1, 128, 0, 1, 12>("tf32 SS K/K   N=128 cg1 drot1", 0, 2));
  v.push_back(mk<1, 1, 1, 256, 0, 1, 12>("tf32 SS K/K   N=256 cg1 drot1", 0, 2));
  v.push_back(mk<1, 1, 1, 128, 0, 2, 12>("tf32 SS K/K   N=128 cg1 drot2", 0, 2));
  v.push_back(mk<1, 1, 1, 256, 0, 2, 12>("tf32 SS K/K   N=256 cg1 drot2", 0, 2));
  v.push_back(mk<1, 1, 1, 128, 1, 0, 12>("bf16 SS K/K   N=128 cg1 drot0", 0, 2));
  v.push_back(mk<1, 1, 1, 256, 1, 0, 12>("bf16 SS K/K   N=256 cg1 drot0", 0, 2));
  v.push_back(mk<1, 0, 0, 128, 0, 0, 9>("tf32 SS MN/MN N=128 cg1 drot0 (current)", 0, 2));
  v.push_back(mk<1, 0, 0, 256, 0, 0, 6>("tf32 SS MN/MN N=256 cg1 drot0", 0, 2));
  v.push_back(mk<1, 1, 0, 256, 0, 0, 6>("tf32 SS K/MN  N=256 cg1 drot0", 0, 2));
  v.push_back(mk<1, 2, 0, 128, 0, 0, 9>("tf32 TS TM/MN N=128 cg1 drot0", 0, 2));
  v.push_back(mk<1, 2, 0, 256, 0, 0, 6>("tf32 TS TM/MN N=256 cg1 drot0", 0, 2));
  v.push_back(mk<1, 2, 1, 256, 0, 0, 6>("tf32 TS TM/K  N=256 cg1 drot0", 0, 2));
  // --- with converter-like noise
  v.push_back(mk<1, 0, 0, 128, 0, 0, 9>("tf32 SS MN/MN N=128 cg1 (current)", 8, 2));
  v.push_back(mk<1, 0, 0, 256, 0, 0, 6>("tf32 SS MN/MN N=256 cg1", 8, 2));
  v.push_back(mk<1, 1, 1, 256, 0, 0, 12>("tf32 SS K/K   N=256 cg1", 8, 2));
  v.push_back(mk<1, 2, 0, 256, 0, 0, 6>("tf32 TS TM/MN N=256 cg1", 8, 2));
  v.push_back(mk<1, 0, 0, 256, 0, 0, 6>("tf32 SS MN/MN N=256 cg1", 4, 2));
  v.push_back(mk<1, 0, 0, 256, 0, 0, 6>("tf32 SS MN/MN N=256 cg1 (LDS+1STS)", 8, 1));
  // --- CTA pairs
  v.push_back(mk<2, 0, 0, 256, 0, 0, 6>("tf32 SS MN/MN N=256 cg2 M=256", 0, 2));
  v.push_back(mk<2, 1, 1, 256, 0, 0, 12>("tf32 SS K/K   N=256 cg2 M=256", 0, 2));
  v.push_back(mk<2, 2, 0, 256, 0, 0, 6>("tf32 TS TM/MN N=256 cg2 M=256", 0, 2));
  v.push_back(mk<2, 1, 1, 256, 1, 0, 12>("bf16 SS K/K   N=256 cg2 M=256", 0, 2));
  v.push_back(mk<2, 0, 0, 256, 0, 0, 6>("tf32 SS MN/MN N=256 cg2 M=256", 8, 2));
  v.push_back(mk<2, 1, 1, 256, 0, 0, 12>("tf32 SS K/K   N=256 cg2 M=256", 8, 2));
  v.push_back(mk<2, 2, 0, 256, 0, 0, 6>("tf32 TS TM/MN N=256 cg2 M=256", 8, 2));
  v.push_back(mk<2, 0, 0, 256, 0, 0, 6>("tf32 SS MN/MN N=256 cg2 M=256", 4, 2));
  v.push_back(mk<2, 0, 0, 256, 0, 0, 6>("tf32 SS MN/MN N=256 cg2 M=256 (LDS+1STS)", 8, 1));
  std::vector<long long> cyc(sm), noi(sm);
  for (auto& nv : v) {
    const Cfg& c = nv.c;
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(sm, 1, 1);
    lc.blockDim = dim3(320, 1, 1);
    lc.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = c.cg;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    lc.attrs = at;
    lc.numAttrs = c.cg == 2 ? 1 : 0;
    Cfg cc = c;
    volatile int* fl = d_flag;
    void* args[] = {&cc, &d_cyc, &d_noise, &fl};
    for (int rep = 0; rep < 2; ++rep) {
      CK(cudaMemset(d_noise, 0, sm * sizeof(long long)));
      CK(cudaMemset(d_flag, 0, sm * sizeof(int)));
      CK(cudaLaunchKernelExC(&lc, nv.fn, args));
      CK(cudaDeviceSynchronize());
    }
    CK(cudaMemcpy(cyc.data(), d_cyc, sm * sizeof(long long), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(noi.data(), d_noise, sm * sizeof(long long), cudaMemcpyDeviceToHost));
    double sum = 0, nsum = 0;
    int cnt = 0;
    for (int i = 0; i < sm; i += c.cg) { sum += (double)cyc[i]; nsum += (double)noi[i]; ++cnt; }
    const double avg = sum / cnt;
    const double mmas = (double)c.groups * c.steps * c.per_step;
    const double ideal = (double)c.n / 2.0;
    const double nbytes = nsum / cnt * 8.0 * (1 + c.noise_sts) * 512.0;
    printf("%-44s noise=%d cyc/MMA=%6.1f (floor %5.1f, %5.1f%% of peak) noise B/clk=%5.1f\n", nv.name, c.noise,
           avg / mmas, ideal, 100.0 * ideal / (avg / mmas), nbytes / avg);
    fflush(stdout);
  }
  return 0;
}
