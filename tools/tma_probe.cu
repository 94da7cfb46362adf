// GPU probe (timing only): HBM -> shared-memory streaming rate of TMA on B200 for the box shapes the
// Gram kernels can use to read a row-major fp32 matrix once.  Each CTA keeps `depth` boxes in flight
// (no consumer: a buffer is reused as soon as its box has landed).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/build/tma_probe tools/tma_probe.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../dimreduce4gpu_b200/csrc/ptx_sm100.cuh"

using namespace dr4g;

#define CK(x)                                                                         \
  do {                                                                                \
    cudaError_t e_ = (x);                                                             \
    if (e_ != cudaSuccess) {                                                          \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(1);                                                                        \
    }                                                                                 \
  } while (0)

struct Cfg {
  int mode;        // 0: 2-D map (col, row); 1: 3-D map (0, row, col/32); 2: 1-D bulk copy (contiguous bytes)
  int box_rows;    // rows per box
  int box_cols;    // columns per box
  int col_blocks;  // boxes per row stripe (d / box_cols)
  int depth;       // boxes in flight per CTA
  int box_bytes;
  long long n_boxes;  // total boxes
  int d;
  int issuers;     // warps issuing (each its own boxes and buffers)
};

__global__ void __launch_bounds__(128, 1) tma_probe(const __grid_constant__ CUtensorMap tmap, Cfg c, const float* X,
                                                     long long* cycles) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (ptx::smem_u32(raw) + 1023u) & ~1023u;
  __shared__ __align__(8) uint64_t bar_all[4 * 32];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4 * 32; ++i) ptx::mbar_init(ptx::smem_u32(&bar_all[i]), 1);
    ptx::fence_barrier_init();
  }
  __syncthreads();
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0 && w < c.issuers) {
    uint64_t* bar = bar_all + w * 32;
    const long long t0 = clock64();
    long long i = 0;
    for (long long b = (long long)blockIdx.x * c.issuers + w; b < c.n_boxes; b += (long long)gridDim.x * c.issuers, ++i) {
      const int s = (int)(i % c.depth);
      if (i >= c.depth) ptx::mbar_wait(ptx::smem_u32(&bar[s]), (uint32_t)((i / c.depth - 1) & 1));
      const uint32_t full = ptx::smem_u32(&bar[s]);
      ptx::mbar_arrive_expect_tx(full, (uint32_t)c.box_bytes);
      const uint32_t dst = base + (uint32_t)(w * c.depth + s) * (uint32_t)c.box_bytes;
      const long long stripe = b / c.col_blocks;
      const int cb = (int)(b % c.col_blocks);
      const int row0 = (int)(stripe * c.box_rows), col0 = cb * c.box_cols;
      if (c.mode == 0) ptx::tma_load_2d(dst, &tmap, full, col0, row0);
      else if (c.mode == 1) ptx::tma_load_3d(dst, &tmap, full, 0, row0, col0 / 32);
      else
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                     "l"(X + ((long long)row0 * c.d + col0)), "r"(c.box_bytes), "r"(full)
                     : "memory");
    }
    // drain
    for (long long j = (i > c.depth ? i - c.depth : 0); j < i; ++j)
      ptx::mbar_wait(ptx::smem_u32(&bar[j % c.depth]), (uint32_t)((j / c.depth) & 1));
    if (w == 0) cycles[blockIdx.x] = clock64() - t0;
  }
}

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  const long long n = 1000000;
  const int d = 256;
  float* X;
  CK(cudaMalloc(&X, (size_t)n * d * 4 + (1 << 20)));
  CK(cudaMemset(X, 0, (size_t)n * d * 4));
  long long* cyc;
  CK(cudaMalloc(&cyc, 148 * 8));
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  auto enc = reinterpret_cast<EncodeTiledFn>(fn);
  const int smem = 200 * 1024;
  CK(cudaFuncSetAttribute(tma_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));

  struct V { const char* name; int mode, rows, cols, depth; CUtensorMapSwizzle sw; CUtensorMapL2promotion l2; int issuers = 1; };
  const CUtensorMapSwizzle A32 = CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, NONE = CU_TENSOR_MAP_SWIZZLE_NONE;
  const CUtensorMapL2promotion P256 = CU_TENSOR_MAP_L2_PROMOTION_L2_256B, P128 = CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                               P0 = CU_TENSOR_MAP_L2_PROMOTION_NONE;
  std::vector<V> vs = {
      {"3-D box 32x16x4 (8 KB, 128 cols) swz ATOM32B  depth 12", 1, 16, 128, 12, A32, P256},
      {"3-D box 32x16x4 (8 KB, 128 cols) swz ATOM32B  depth 24", 1, 16, 128, 24, A32, P256},
      {"3-D box 32x16x4 (8 KB) swz ATOM32B no L2 promo depth 24", 1, 16, 128, 24, A32, P0},
      {"3-D box 32x64x4 (32 KB, 128 cols) swz ATOM32B depth 6", 1, 64, 128, 6, A32, P256},
      {"2-D box 32x64   (8 KB,  32 cols) swz ATOM32B  depth 24", 0, 64, 32, 24, A32, P256},
      {"2-D box 128x16  (8 KB, 128 cols) no swizzle   depth 24", 0, 16, 128, 24, NONE, P256},
      {"2-D box 128x64  (32 KB,128 cols) no swizzle   depth 6", 0, 64, 128, 6, NONE, P256},
      {"2-D box 256x8   (8 KB, full rows) no swizzle  depth 24", 0, 8, 256, 24, NONE, P256},
      {"2-D box 256x32  (32 KB, full rows) no swizzle depth 6", 0, 32, 256, 6, NONE, P256},
      {"2-D box 256x32  (32 KB, full rows) no swz 128B promo d6", 0, 32, 256, 6, NONE, P128},
      {"3-D box 32x16x4 (8 KB) swz ATOM32B  2 issuers x depth 12", 1, 16, 128, 12, A32, P256, 2},
      {"3-D box 32x16x4 (8 KB) swz ATOM32B  4 issuers x depth 6", 1, 16, 128, 6, A32, P256, 4},
      {"3-D box 32x32x4 (16 KB) swz ATOM32B 1 issuer  x depth 12", 1, 32, 128, 12, A32, P256, 1},
      {"3-D box 32x32x4 (16 KB) swz ATOM32B 2 issuers x depth 6", 1, 32, 128, 6, A32, P256, 2},
      {"3-D box 32x8x4  (4 KB) swz ATOM32B  4 issuers x depth 8", 1, 8, 128, 8, A32, P256, 4},
      {"3-D box 32x16x4 (8 KB) swz ATOM32B  1 issuer  x depth 3", 1, 16, 128, 3, A32, P256, 1},
      {"3-D box 32x64x4 (32 KB) swz ATOM32B 1 issuer  x depth 2", 1, 64, 128, 2, A32, P256, 1},
      {"3-D box 32x64x4 (32 KB) swz ATOM32B 1 issuer  x depth 3", 1, 64, 128, 3, A32, P256, 1},
      {"1-D bulk 8 KB contiguous (8 full rows)        depth 24", 2, 8, 256, 24, NONE, P0},
      {"1-D bulk 32 KB contiguous (32 full rows)      depth 6", 2, 32, 256, 6, NONE, P0},
      {"1-D bulk 32 KB contiguous (32 full rows)      depth 3", 2, 32, 256, 3, NONE, P0},
      {"1-D bulk 64 KB contiguous (64 full rows)      depth 3", 2, 64, 256, 3, NONE, P0},
  };
  for (auto& v : vs) {
    CUtensorMap tmap;
    memset(&tmap, 0, sizeof(tmap));
    CUresult r = CUDA_SUCCESS;
    if (v.mode == 1) {
      const cuuint64_t gdim[3] = {32, (cuuint64_t)n, (cuuint64_t)(d / 32)};
      const cuuint64_t gstride[2] = {(cuuint64_t)d * 4, 128};
      const cuuint32_t box[3] = {32, (cuuint32_t)v.rows, (cuuint32_t)(v.cols / 32)};
      const cuuint32_t es[3] = {1, 1, 1};
      r = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, X, gdim, gstride, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, v.sw, v.l2,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else if (v.mode == 0) {
      const cuuint64_t gdim[2] = {(cuuint64_t)d, (cuuint64_t)n};
      const cuuint64_t gstride[1] = {(cuuint64_t)d * 4};
      const cuuint32_t box[2] = {(cuuint32_t)v.cols, (cuuint32_t)v.rows};
      const cuuint32_t es[2] = {1, 1};
      r = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, X, gdim, gstride, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, v.sw, v.l2,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    if (r != CUDA_SUCCESS) {
      printf("%-58s encode failed (%d)\n", v.name, (int)r);
      continue;
    }
    Cfg c;
    c.mode = v.mode;
    c.box_rows = v.rows;
    c.box_cols = v.cols;
    c.col_blocks = d / v.cols;
    c.depth = v.depth;
    c.box_bytes = v.rows * v.cols * 4;
    c.n_boxes = (n / v.rows) * c.col_blocks;
    c.d = d;
    c.issuers = v.issuers;
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
      CK(cudaEventRecord(e0));
      tma_probe<<<148, 128, smem>>>(tmap, c, X, cyc);
      CK(cudaEventRecord(e1));
      CK(cudaDeviceSynchronize());
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (ms < best) best = ms;
    }
    const double bytes = (double)c.n_boxes * c.box_bytes;
    printf("%-58s %7.3f ms  %7.1f GB/s  (%.1f B/clk/SM at 1965 MHz)\n", v.name, best, bytes / best / 1e6,
           bytes / best / 1e6 * 1e9 / 148 / 1.965e9 / 1e3 * 1e3 / 1e3);
    fflush(stdout);
  }
  return 0;
}
