// K4 - device-resident symmetric eigensolver for the top-k eigenpairs (fp64), tridiagonal route:
//   T1  Householder tridiagonalisation  A = Q T Q^T          (one cooperative launch, 1 grid barrier / column)
//   T2  eigenvalues of T by Sturm-count multisection          (one warp per eigenvalue, 32 shifts per step)
//   T3  eigenvectors of T by inverse iteration                (pivoted tridiagonal LU; clusters re-orthogonalised)
//   T4  back-transformation  v = Q z                          (one CTA per vector)
// Replaces cusolverDn{S,D}syevd (reference src/matrix.cu:231-273, called from src/tsvd.cu:329).
// Only the k leading pairs are formed: everything the outputs need beyond them (the total variance)
// comes from the column statistics, not from the spectrum.  The blocked Jacobi solver (jacobi.cu)
// stays for dr4g_eigh_double (all pairs) and the small Rayleigh-Ritz problem of the 'power' path;
// at d = 256 it needed 18 ms per fit on B200 against 0.65 ms for the Gram pass, which is why the
// exact path moved here.
//
// T1 keeps the full symmetric trailing matrix in global memory (L2 resident: 8 d^2 bytes) and owns
// rows cyclically, one warp per row.  Iteration s eliminates column s:
//   A  every CTA (redundantly): corrected row s  r = A[s,:] - v_s w - w_s v  (previous reflector's
//      rank-2 update applied on the fly), d_s = r_s, new reflector (v', tau') from r[s+1:], e_s = beta
//   B  own rows i > s: A[i,c] -= v_i w_c + w_i v_c (c > s), and in the same pass p'_i = tau' A[i,:] v'
//   -- grid barrier --
//   C  every CTA (redundantly): w' = p' - (tau'/2)(p'.v') v'
// so the rank-2 update of one column and the matrix-vector product of the next share one sweep
// over the rows and one barrier.
#include <cfloat>
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace dr4g {
namespace {

constexpr int kTriThreads = 256;
constexpr int kTriWarps = kTriThreads / 32;

struct TriCtrl {
  unsigned int barrier;
  unsigned int pad;
};

// Branch-free fp64 reciprocal: hardware seed (rcp.approx.ftz.f64, ~2^-23) + two Newton steps.  The
// IEEE division is a ~160-cycle call-like sequence on sm_100 that also defeats ILP across
// independent recurrences; this one is ~5 dependent instructions.  |x| must be a normal number.
__device__ __forceinline__ double fast_rcp(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  r = fma(r, fma(-x, r, 1.0), r);
  r = fma(r, fma(-x, r, 1.0), r);
  return r;
}

// sqrt(x) for normal positive x the same way: rsqrt.approx.ftz.f64 seed + two Newton steps, sqrt = x * rsqrt
// (the IEEE sqrt is another call-like sequence on the per-step critical path of the tridiagonalisation).
__device__ __forceinline__ double fast_sqrt(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double hx = 0.5 * x;
  y = y * fma(-hx * y, y, 1.5);
  y = y * fma(-hx * y, y, 1.5);
  const double r = x * y;
  return fma(fma(-r, r, x), 0.5 * y, r);   // one correction step on the root itself
}

__device__ __forceinline__ void tri_grid_barrier(TriCtrl* ctrl, unsigned int& target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    target += gridDim.x;
    __threadfence();
    atomicAdd(&ctrl->barrier, 1u);
    while (*reinterpret_cast<volatile unsigned int*>(&ctrl->barrier) < target) {}
    __threadfence();
  }
  __syncthreads();
}

__device__ __forceinline__ double tri_block_sum(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0;
#pragma unroll
  for (int w = 0; w < kTriWarps; ++w) t += red[w];
  return t;
}

// one-barrier variant: partials go to alternating buffers, so a warp that races ahead into the next
// reduction writes the other buffer (a third reduction cannot start before everyone passed the second)
__device__ __forceinline__ double tri_block_sum1(double v, double (*red)[kTriWarps], int& par) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) red[par][threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0;
#pragma unroll
  for (int w = 0; w < kTriWarps; ++w) t += red[par][w];
  par ^= 1;
  return t;
}

// A: d x d symmetric, row-major, destroyed.  Out: diag[d], off[d] (off[s] couples s and s+1; off[d-1] = 0),
// tau[d], Vh[d x d] (row s = reflector of column s, entries s+1.. with Vh[s][s+1] = 1).
// xbuf: 4 x d doubles of exchange space: p' (two parities) and the published next row (two parities).
// kRowsInSmem: every CTA keeps its rows of the trailing matrix in shared memory (no L2 round trip
// in the row sweep); otherwise they stay in A (large d).
template <bool kRowsInSmem>
__global__ void __launch_bounds__(kTriThreads)
tridiag_kernel(double* __restrict__ A, int d, double* __restrict__ diag, double* __restrict__ off,
               double* __restrict__ tau, double* __restrict__ Vh, double* __restrict__ xbuf, TriCtrl* ctrl) {
  extern __shared__ double tri_sm[];
  double* v = tri_sm;          // previous reflector (full length d, zero outside its support)
  double* w = v + d;           // previous w
  double* vn = w + d;          // new reflector
  double* row = vn + d;        // row s: raw, then corrected
  double* pl = row + d;        // p of the previous iteration
  double* R = pl + d;          // own rows (kRowsInSmem): slot (i / nw) * kTriWarps + warp
  __shared__ double red[kTriWarps];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gw = blockIdx.x * kTriWarps + warp, nw = gridDim.x * kTriWarps;
  unsigned int target = 0;
  double* pbuf = xbuf;
  double* rbuf = xbuf + 2 * (size_t)d;

  for (int c = tid; c < d; c += kTriThreads) { v[c] = 0.0; w[c] = 0.0; vn[c] = 0.0; pl[c] = 0.0; }
  if (kRowsInSmem) {
    for (int i = gw; i < d; i += nw) {
      double* Ri = R + (size_t)((i / nw) * kTriWarps + warp) * d;
      for (int c = lane; c < d; c += 32) Ri[c] = A[(size_t)i * d + c];
    }
  }
  __syncthreads();

  double tprev = 0.0;  // tau of reflector s-1 (every CTA computes the same value)
  for (int s = 0; s < d - 1; ++s) {
    // ---- (after the barrier) fetch p of reflector s-1 and the published row s in one round trip
    if (s == 0) {
      for (int c = tid; c < d; c += kTriThreads) row[c] = A[c];
    } else {
      const double* pp = pbuf + (size_t)((s - 1) & 1) * d;
      const double* rp = rbuf + (size_t)(s & 1) * d;
      for (int c = s + tid; c < d; c += kTriThreads) { pl[c] = __ldcg(&pp[c]); row[c] = __ldcg(&rp[c]); }
    }
    // w of reflector s-1:  w = p - (tau/2)(p.v) v   (vn still holds reflector s-1, tprev its tau)
    double pv = 0.0;
    for (int c = s + tid; c < d; c += kTriThreads) pv += pl[c] * vn[c];
    pv = tri_block_sum(pv, red);
    const double kk = 0.5 * tprev * pv;
    for (int c = tid; c < d; c += kTriThreads) {
      const double vv = vn[c];
      v[c] = vv;
      w[c] = c >= s ? pl[c] - kk * vv : 0.0;
    }
    __syncthreads();
    // corrected row s and the reflector that eliminates r[s+2:]
    const double vs = v[s], ws = w[s];
    double nrm2 = 0.0;
    for (int c = s + tid; c < d; c += kTriThreads) {
      const double r = row[c] - vs * w[c] - ws * v[c];
      row[c] = r;
      if (c > s + 1) nrm2 += r * r;
    }
    nrm2 = tri_block_sum(nrm2, red);
    const double x0 = row[s + 1];
    double beta = x0, t = 0.0, scale = 0.0;
    if (nrm2 > 0.0 && s < d - 2) {
      const double nrm = sqrt(x0 * x0 + nrm2);
      beta = x0 >= 0.0 ? -nrm : nrm;
      t = (beta - x0) / beta;
      scale = 1.0 / (x0 - beta);
    }
    for (int c = tid; c < d; c += kTriThreads) vn[c] = c > s + 1 ? row[c] * scale : (c == s + 1 ? 1.0 : 0.0);
    if (blockIdx.x == 0) {
      if (tid == 0) { diag[s] = row[s]; off[s] = beta; tau[s] = t; }
      for (int c = tid; c < d; c += kTriThreads) Vh[(size_t)s * d + c] = c > s + 1 ? row[c] * scale : (c == s + 1 ? 1.0 : 0.0);
    }
    __syncthreads();
    // ---- rank-2 update of own rows fused with p' = tau' A' v'; the owner of row s+1 publishes it
    for (int i = s + 1 + ((gw - (s + 1)) % nw + nw) % nw; i < d; i += nw) {
      const double vi = v[i], wi = w[i];
      double* Ai = kRowsInSmem ? R + (size_t)((i / nw) * kTriWarps + warp) * d : A + (size_t)i * d;
      double* pub = (i == s + 1) ? rbuf + (size_t)((s + 1) & 1) * d : nullptr;
      double dot = 0.0;
      int c = s + 1 + lane;
      if (!kRowsInSmem) {
        // rows in L2 / HBM (d > ~1000): four independent loads in flight per lane - the one-load-per-iteration loop
        // paid the full memory latency 128 times per row at d = 4096 (100 us per column, measured)
        // (round 2: sixteen per lane - with four, 8 warps x 1 KB in flight per SM is ~1.2 TB/s chip-wide at ~1 us of L2
        // latency, and the sweep of a 45 MB trailing matrix took 32 us)
        for (; c + 32 * 15 < d; c += 32 * 16) {
          double a[16];
#pragma unroll
          for (int u = 0; u < 16; ++u) a[u] = __ldcg(&Ai[c + 32 * u]);
#pragma unroll
          for (int u = 0; u < 16; ++u) {
            a[u] = a[u] - vi * w[c + 32 * u] - wi * v[c + 32 * u];
            __stcg(&Ai[c + 32 * u], a[u]);
            if (pub) __stcg(&pub[c + 32 * u], a[u]);
            dot = fma(a[u], vn[c + 32 * u], dot);
          }
        }
        for (; c + 96 < d; c += 128) {
          double a[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) a[u] = __ldcg(&Ai[c + 32 * u]);
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            a[u] = a[u] - vi * w[c + 32 * u] - wi * v[c + 32 * u];
            __stcg(&Ai[c + 32 * u], a[u]);
            if (pub) __stcg(&pub[c + 32 * u], a[u]);
            dot = fma(a[u], vn[c + 32 * u], dot);
          }
        }
      }
      for (; c < d; c += 32) {
        const double a = (kRowsInSmem ? Ai[c] : __ldcg(&Ai[c])) - vi * w[c] - wi * v[c];
        if (kRowsInSmem) Ai[c] = a; else __stcg(&Ai[c], a);
        if (pub) __stcg(&pub[c], a);
        dot = fma(a, vn[c], dot);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
      if (lane == 0) __stcg(&pbuf[(size_t)(s & 1) * d + i], t * dot);
    }
    tprev = t;
    tri_grid_barrier(ctrl, target);
  }
  // last diagonal entry: row d-1 as published in iteration d-2 (that reflector is the identity: no correction)
  if (blockIdx.x == 0 && tid == 0) {
    const int s = d - 1;
    diag[s] = d > 1 ? __ldcg(&rbuf[(size_t)(s & 1) * d + s]) : A[0];
    off[s] = 0.0;
    tau[s] = 0.0;
  }
}

// ---------------------------------------------------------------- T1, cluster-resident variant (d <= 512)
// Same algorithm inside ONE thread-block cluster (1, 8 or 16 CTAs): the trailing matrix lives in the
// cluster's shared memory (rows cyclic over CTAs), the per-column exchange of p' goes through
// distributed shared memory (each owner pushes its entry into every CTA), the hardware cluster
// barrier replaces the atomic-counter grid barrier, and only the published row travels through L2.
__device__ __forceinline__ unsigned cluster_rank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned cluster_size() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void st_cluster_f64(double* local_ptr, unsigned cta, double val) {
  const unsigned la = static_cast<unsigned>(__cvta_generic_to_shared(local_ptr));
  unsigned ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(cta));
  asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(ra), "d"(val) : "memory");
}

// 8-byte store into the shared memory of CTA `cta` that also counts its bytes on that CTA's mbarrier
// (st.async): the data's arrival is the signal, no release fence / cluster barrier on the producer side.
__device__ __forceinline__ void st_async_f64(double* local_ptr, uint64_t* local_bar, unsigned cta, double val) {
  const unsigned la = static_cast<unsigned>(__cvta_generic_to_shared(local_ptr));
  const unsigned lb = static_cast<unsigned>(__cvta_generic_to_shared(local_bar));
  unsigned ra, rb;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(la), "r"(cta));
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rb) : "r"(lb), "r"(cta));
  asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.f64 [%0], %1, [%2];" ::"r"(ra), "d"(val), "r"(rb)
               : "memory");
}

// Bulk copy of `bytes` (multiple of 16, 16-byte aligned) from this CTA's shared memory into the shared memory of CTA
// `cta`, counted on that CTA's mbarrier: one instruction per destination instead of one st.async per element.
__device__ __forceinline__ void bulk_push(double* dst_local, const double* src_local, unsigned bytes, uint64_t* bar_local,
                                          unsigned cta) {
  const unsigned ls = static_cast<unsigned>(__cvta_generic_to_shared(src_local));
  const unsigned ld = static_cast<unsigned>(__cvta_generic_to_shared(dst_local));
  const unsigned lb = static_cast<unsigned>(__cvta_generic_to_shared(bar_local));
  unsigned rd, rb;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rd) : "r"(ld), "r"(cta));
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rb) : "r"(lb), "r"(cta));
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(rd),
               "r"(ls), "r"(bytes), "r"(rb)
               : "memory");
}

// Debug aid (DR4G_TRI_TRACE=1): clock64() of CTA 0, thread 0 at the phase boundaries of steps 100..103.
__device__ long long g_tri_trace[4 * 16];
#define TRI_TRACE(ev) \
  do { if ((trace & 1) && rank == 0 && tid == 0 && s >= 100 && s < 104) g_tri_trace[(s - 100) * 16 + (ev)] = clock64(); } while (0)

// RW > 0: at most RW own rows per warp (d <= 128 RW in a 16-CTA cluster: RW = 2 up to d = 256, RW = 4 up to d = 512);
// they stay in registers (lane: columns 2 lane + 64 j, j < 2 RW) with a fixed row -> warp assignment, because the rows
// phase was bound by shared-memory bandwidth (per element of an own row: the row read and written back plus w, v, v'
// = 40 B; with the rows in registers only w, v, v' remain and they are shared by the warp's RW rows).  d = 512, step
// 100: rows phase 5900 of 11000 cycles with the rows in shared memory.  RW = 0: rows in shared memory.
template <int RW>
__global__ void __launch_bounds__(kTriThreads)
tridiag_cluster_kernel(const double* __restrict__ A, int d, double* __restrict__ diag, double* __restrict__ off,
                       double* __restrict__ tau, double* __restrict__ Vh, double* __restrict__ xbuf, int trace) {
  extern __shared__ __align__(16) double tri_sm[];
  double* v = tri_sm;
  double* w = v + d;
  double* vn = w + d;
  double* rowp = vn + d;       // row s, two parities (pushed into every CTA by the row's owner through DSMEM)
  double* pl = rowp + 2 * (size_t)d;  // p, two parities (pushed into every CTA by the row owners)
  double* R = pl + 2 * (size_t)d;  // own rows: slot i / cs
  __shared__ double red2[2][kTriWarps];
  // data of step s+1 (p' and row s+1, 16 (d-s-1) bytes per CTA) lands in the parity-(s+1) buffers and is counted
  // on xbar[(s+1) & 1]: the consumer waits for the bytes, there is no cluster barrier inside the loop
  __shared__ __align__(8) uint64_t xbar[2];
  unsigned xph = 0u;   // bit b: phase parity of xbar[b]
  int par = 0;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rank = (int)cluster_rank(), cs = (int)cluster_size();
  (void)xbuf;
  if (tid == 0) {
    ptx::mbar_init(ptx::smem_u32(&xbar[0]), 1);
    ptx::mbar_init(ptx::smem_u32(&xbar[1]), 1);
    ptx::fence_barrier_init();
  }

  for (int c = tid; c < d; c += kTriThreads) {
    v[c] = 0.0; w[c] = 0.0; vn[c] = 0.0; pl[c] = 0.0; pl[d + c] = 0.0;
    rowp[c] = A[c];            // row 0
    rowp[d + c] = 0.0;
  }
  constexpr bool REG = RW > 0;
  constexpr int NR = REG ? RW : 1, NJ = REG ? 2 * RW : 1;
  double2 rr[NR][NJ];   // REG: slot q = warp + 8 k, columns 2 lane + 64 j
  if constexpr (REG) {
#pragma unroll
    for (int k = 0; k < NR; ++k) {
      const int i = rank + cs * (warp + kTriWarps * k);
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        const int c = 2 * lane + 64 * j;
        rr[k][j] = (i < d && c < d) ? *reinterpret_cast<const double2*>(A + (size_t)i * d + c) : make_double2(0.0, 0.0);
      }
    }
  } else {
    for (int i = rank + cs * warp; i < d; i += cs * kTriWarps) {
      double* Ri = R + (size_t)(i / cs) * d;
      for (int c = lane; c < d; c += 32) Ri[c] = A[(size_t)i * d + c];
    }
  }
  cluster_barrier();  // also orders the zero-initialisation of pl before any remote push

  double tprev = 0.0, sprev = 0.0;   // tau and 1 / (x0 - beta) of reflector s-1
  int qfirst = 0, owner = 0, oslot = 0;   // row s belongs to CTA `owner`, slot `oslot` (s = owner + cs oslot)
  const int nslots = (d - rank + cs - 1) / cs;   // own rows: rank, rank + cs, ...
  for (int s = 0; s < d - 1; ++s) {
    const double* pcur = pl + (size_t)((s + 1) & 1) * d;  // parity (s-1)&1
    TRI_TRACE(0);
    if (s > 0) {
      // wait for this step's operands (pushed by every CTA during its step s-1), then one CTA barrier: the own
      // rows in R are handed from warp to warp between steps
      ptx::mbar_wait_cluster(ptx::smem_u32(&xbar[s & 1]), (xph >> (s & 1)) & 1u);
      xph ^= 1u << (s & 1);
      __syncthreads();
    }
    const bool last = s == d - 2;      // nothing is exchanged after the last step
    // bytes of step s+1: p' entries s+1.. (8 each) and row s+1 from the even column (s+1) & ~1 (16-byte aligned bulk copy)
    if (tid == 0 && !last)
      ptx::mbar_arrive_expect_tx(ptx::smem_u32(&xbar[(s + 1) & 1]), (uint32_t)(8 * (d - s - 1) + 8 * (d - ((s + 1) & ~1))));
    // row s arrived in this CTA's shared memory before the last cluster barrier (an L2 round trip through a
    // published global copy cost ~1100 cycles per step here)
    double* row = rowp + (size_t)(s & 1) * d;
    TRI_TRACE(1);
    double t = 0.0, beta = 0.0, scale = 0.0;
    if constexpr (REG) {
      // The reflector is kept UN-normalised in `vn`: u = corrected row s from column s+1 on, v' = scale (u - beta e_{s+1}).
      // Everything the rows phase needs from it is linear in u, so the rows phase starts right behind the barrier of the
      // norm's reduction and the chain sqrt -> beta -> tau, 1 / (x0 - beta) (~300 dependent cycles) runs under its loads;
      // the normalised reflector is formed one step later, on the fly, where it is read as `v` (saves the sweep that
      // wrote it and one CTA barrier per step).  vn holds u_{s-1} here; sprev = its scale; its unit entry is column s.
      auto vprev_at = [&](int c) { return c == s ? (s > 0 ? 1.0 : 0.0) : (c > s ? vn[c] * sprev : 0.0); };
      double pv = 0.0;
      for (int c = s + tid; c < d; c += kTriThreads) pv += pcur[c] * vprev_at(c);
      pv = tri_block_sum1(pv, red2, par);
      TRI_TRACE(2);
      const double kk = 0.5 * tprev * pv;
      const double vs = s > 0 ? 1.0 : 0.0, ws = pcur[s] - kk * vs;
      double nrm2 = 0.0;
      for (int c = tid; c < d; c += kTriThreads) {
        const double vv = vprev_at(c);
        const double wc = c >= s ? pcur[c] - kk * vv : 0.0;
        v[c] = vv;
        w[c] = wc;
        double u = 0.0;
        if (c >= s) {
          const double r = row[c] - vs * wc - ws * vv;
          row[c] = r;
          if (c > s) u = r;
          if (c > s + 1) nrm2 += r * r;
        }
        vn[c] = u;
        if (rank == 0 && s > 0) Vh[(size_t)(s - 1) * d + c] = vv;   // reflector s-1, normalised, exact zeros below its 1
      }
      nrm2 = tri_block_sum1(nrm2, red2, par);   // its barrier also publishes v, w, vn and the corrected row
      TRI_TRACE(3);
      const double x0 = row[s + 1];
      beta = x0;
      if (nrm2 > 0.0 && s < d - 2) {
        const double y2 = fma(x0, x0, nrm2);
        const double nrm = y2 > 1e-290 && y2 < 1e290 ? fast_sqrt(y2) : sqrt(y2);
        beta = x0 >= 0.0 ? -nrm : nrm;
        t = (beta - x0) * fast_rcp(beta);
        scale = fast_rcp(x0 - beta);
      }
      if (rank == 0 && tid == 0) { diag[s] = row[s]; off[s] = beta; tau[s] = t; }
      sprev = scale;
    } else {
      double pv = 0.0;
      for (int c = s + tid; c < d; c += kTriThreads) pv += pcur[c] * vn[c];
      pv = tri_block_sum1(pv, red2, par);
      TRI_TRACE(2);
      const double kk = 0.5 * tprev * pv;
      // v <- previous reflector, w = p - kk v, corrected row r = row - v_s w - w_s v: one sweep, every
      // thread needs only its own column besides the scalars v_s and w_s
      const double vs = vn[s], ws = pcur[s] - kk * vs;
      double nrm2 = 0.0;
      for (int c = tid; c < d; c += kTriThreads) {
        const double vv = vn[c];
        const double wc = c >= s ? pcur[c] - kk * vv : 0.0;
        v[c] = vv;
        w[c] = wc;
        if (c >= s) {
          const double r = row[c] - vs * wc - ws * vv;
          row[c] = r;
          if (c > s + 1) nrm2 += r * r;
        }
      }
      nrm2 = tri_block_sum1(nrm2, red2, par);
      TRI_TRACE(3);
      const double x0 = row[s + 1];
      beta = x0;
      if (nrm2 > 0.0 && s < d - 2) {
        const double y2 = fma(x0, x0, nrm2);
        const double nrm = y2 > 1e-290 && y2 < 1e290 ? fast_sqrt(y2) : sqrt(y2);
        beta = x0 >= 0.0 ? -nrm : nrm;
        t = (beta - x0) * fast_rcp(beta);
        scale = fast_rcp(x0 - beta);
      }
      for (int c = tid; c < d; c += kTriThreads) vn[c] = c > s + 1 ? row[c] * scale : (c == s + 1 ? 1.0 : 0.0);
      if (rank == 0) {
        if (tid == 0) { diag[s] = row[s]; off[s] = beta; tau[s] = t; }
        for (int c = tid; c < d; c += kTriThreads) Vh[(size_t)s * d + c] = c > s + 1 ? row[c] * scale : (c == s + 1 ? 1.0 : 0.0);
      }
    __syncthreads();
    }
    TRI_TRACE(4);
    // own rows i > s: rank-2 update fused with p'_i; push p'_i into every CTA; publish row s+1
    double* pnext = pl + (size_t)(s & 1) * d;
    // own slots q (row i = rank + cs q) with i > s; qfirst advances with s (no division in the loop)
    while (rank + cs * qfirst <= s) ++qfirst;
    const int cbeg = (s + 1) & ~63;          // lanes take aligned pairs of columns: c = cbeg + 2 lane + 64 j
    if constexpr (REG) {
      int irow[NR];
      bool act[NR];
      bool any = false;
#pragma unroll
      for (int k = 0; k < NR; ++k) {
        irow[k] = rank + cs * (warp + kTriWarps * k);
        act[k] = irow[k] > s && irow[k] < d;
        any = any || act[k];
      }
      if (any) {                                // warp-uniform
        double vi[NR], wi[NR], dx[NR], dy[NR], asel[NR];
#pragma unroll
        for (int k = 0; k < NR; ++k) {
          vi[k] = act[k] ? v[irow[k]] : 0.0;
          wi[k] = act[k] ? w[irow[k]] : 0.0;
          dx[k] = 0.0;
          dy[k] = 0.0;
          asel[k] = 0.0;
        }
        const int jsel = (s + 1) >> 6, lsel = ((s + 1) & 63) >> 1;   // where column s+1 of an own row lives
        // Branch-free over the lanes: columns <= s of an own row are dead (never read again) and v' is exactly zero
        // there, so every lane updates its pair and adds it to the dot; an inactive row has vi = wi = 0 and stays
        // as it is.  (Per-lane tests compiled to divergent regions: ~290 cycles per 64 columns.)
        const int jbeg = (s + 1) >> 6;
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          const int c = 2 * lane + 64 * j;
          if (j >= jbeg && 64 * j < d) {          // warp-uniform
            const bool inb = c < d;
            const int cc = inb ? c : 0;
            const double2 wv = *reinterpret_cast<const double2*>(w + cc);
            const double2 vv = *reinterpret_cast<const double2*>(v + cc);
            double2 nv = *reinterpret_cast<const double2*>(vn + cc);
            nv.x = inb ? nv.x : 0.0;
            nv.y = inb ? nv.y : 0.0;
#pragma unroll
            for (int k = 0; k < NR; ++k) {
              double2 a = rr[k][j];
              a.x = fma(-wi[k], vv.x, fma(-vi[k], wv.x, a.x));
              a.y = fma(-wi[k], vv.y, fma(-vi[k], wv.y, a.y));
              dx[k] = fma(a.x, nv.x, dx[k]);
              dy[k] = fma(a.y, nv.y, dy[k]);
              if (j == jsel) asel[k] = ((s + 1) & 1) ? a.y : a.x;   // warp-uniform test; used by lane lsel only
              if (inb) rr[k][j] = a;
            }
          }
        }
        // A' v' = scale (A' u - beta A'[:, s+1]): the lane that holds column s+1 takes the second term
        double dot[NR];
#pragma unroll
        for (int k = 0; k < NR; ++k) dot[k] = (dx[k] + dy[k]) - (lane == lsel ? beta * asel[k] : 0.0);
        const double tsc = t * scale;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
          for (int k = 0; k < NR; ++k) dot[k] += __shfl_xor_sync(0xffffffffu, dot[k], o);
        }
        if (lane < cs && !last) {
#pragma unroll
          for (int k = 0; k < NR; ++k)
            if (act[k]) st_async_f64(&pnext[irow[k]], &xbar[(s + 1) & 1], (unsigned)lane, tsc * dot[k]);
        }
      }
      // the warp that holds row s+1 stages it in shared memory for the push below (after the last step: for diag[d-1])
      if (rank == (owner + 1 == cs ? 0 : owner + 1)) {
        const int os = owner + 1 == cs ? oslot + 1 : oslot;     // slot of row s+1
        if ((os & (kTriWarps - 1)) == warp) {
          const int kk = os / kTriWarps;
#pragma unroll
          for (int j = 0; j < NJ; ++j) {
            const int c = 2 * lane + 64 * j;
            double2 e = rr[0][j];
#pragma unroll
            for (int k = 1; k < NR; ++k) if (k == kk) e = rr[k][j];
            if (c < d) *reinterpret_cast<double2*>(R + c) = e;
          }
          ptx::fence_proxy_async_smem();   // the bulk copy below reads R through the async proxy
        }
      }
    } else {
    // two own rows per pass (independent dot chains and shuffle trees; w, v, v' are loaded once)
    for (int q = qfirst + warp; q < nslots; q += 2 * kTriWarps) {
      const bool two = q + kTriWarps < nslots;
      const int i0 = rank + cs * q, i1 = i0 + cs * kTriWarps;
      const double vi0 = v[i0], wi0 = w[i0];
      const double vi1 = two ? v[i1] : 0.0, wi1 = two ? w[i1] : 0.0;
      double* R0 = R + (size_t)q * d;
      double* R1 = two ? R0 + (size_t)kTriWarps * d : R0;
      double dot0 = 0.0, dot1 = 0.0;
      for (int c = cbeg + 2 * lane; c < d; c += 64) {
        const double2 wv = *reinterpret_cast<const double2*>(w + c);
        const double2 vv = *reinterpret_cast<const double2*>(v + c);
        const double2 nv = *reinterpret_cast<const double2*>(vn + c);
        double2 a = *reinterpret_cast<const double2*>(R0 + c);
        double2 b = *reinterpret_cast<const double2*>(R1 + c);
        if (c > s) {
          a.x = a.x - vi0 * wv.x - wi0 * vv.x;
          b.x = b.x - vi1 * wv.x - wi1 * vv.x;
          dot0 = fma(a.x, nv.x, dot0);
          dot1 = fma(b.x, nv.x, dot1);
        }
        if (c + 1 > s) {
          a.y = a.y - vi0 * wv.y - wi0 * vv.y;
          b.y = b.y - vi1 * wv.y - wi1 * vv.y;
          dot0 = fma(a.y, nv.y, dot0);
          dot1 = fma(b.y, nv.y, dot1);
        }
        *reinterpret_cast<double2*>(R0 + c) = a;
        if (two) *reinterpret_cast<double2*>(R1 + c) = b;
      }
      if (!(trace & 8)) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        dot0 += __shfl_xor_sync(0xffffffffu, dot0, o);
        dot1 += __shfl_xor_sync(0xffffffffu, dot1, o);
      }
      }
      if (lane < cs && !last) {
        st_async_f64(&pnext[i0], &xbar[(s + 1) & 1], (unsigned)lane, t * dot0);
        if (two) st_async_f64(&pnext[i1], &xbar[(s + 1) & 1], (unsigned)lane, t * dot1);
      }
    }
    }
    tprev = t;
    if (++owner == cs) { owner = 0; ++oslot; }
    // the owner of row s+1 pushes it (columns > s) into every CTA of the cluster, all its threads sharing the stores
    if (owner == rank && !last) {
      if constexpr (!REG) ptx::fence_proxy_async_smem();   // the row was updated in R by generic stores of several warps
      __syncthreads();
      const double* Rs = REG ? R : R + (size_t)oslot * d;
      double* dst = rowp + (size_t)((s + 1) & 1) * d;
      // one bulk copy per destination CTA (round 1 pushed element by element: 16 x (d - s) st.async from one CTA,
      // ~2700 cycles of the step at d = 512); the extra element at an odd s + 1 lands in a dead slot of the buffer
      const int c0 = (s + 1) & ~1;
      if (tid < cs) bulk_push(dst + c0, Rs + c0, (unsigned)(8 * (d - c0)), &xbar[(s + 1) & 1], (unsigned)tid);
    }
    TRI_TRACE(5);
    TRI_TRACE(6);   // pushes issued; the wait for the next step's operands is event 1 of the next step
  }
  if constexpr (REG) {
    // reflector d-2 (the identity: tau = 0, unit entry in column d-1) was never read back as `v`: its row of Vh
    if (rank == 0 && d >= 2)
      for (int c = tid; c < d; c += kTriThreads) Vh[(size_t)(d - 2) * d + c] = c == d - 1 ? 1.0 : 0.0;
  }
  cluster_barrier();   // no CTA may exit while its shared memory can still be a target
  if (d == 1) {
    if (tid == 0) { diag[0] = A[0]; off[0] = 0.0; tau[0] = 0.0; }
  } else if (REG) {
    // row d-1 was staged in R by its owner during the last step (the cluster barrier above orders the store)
    if ((d - 1) % cs == rank && tid == 0) {
      diag[d - 1] = R[d - 1];
      off[d - 1] = 0.0;
      tau[d - 1] = 0.0;
    }
  } else if ((d - 1) % cs == rank && tid == 0) {
    const int s = d - 1;
    diag[s] = R[(size_t)(s / cs) * d + s];
    off[s] = 0.0;
    tau[s] = 0.0;
  }
}

}  // namespace

namespace {

// ---------------------------------------------------------------- T2: eigenvalues by multisection
// Number of eigenvalues of T smaller than x (Sturm count on the LDL^T pivots, LAPACK dlaebz style).
__device__ __forceinline__ int sturm_count(const double* __restrict__ dg, const double* __restrict__ e2, int d,
                                           double x, double pivmin) {
  int cnt = 0;
  double q = dg[0] - x;
  if (fabs(q) < pivmin) q = -pivmin;
  cnt += q < 0.0;
  for (int i = 1; i < d; ++i) {
    q = dg[i] - x - e2[i - 1] * fast_rcp(q);
    if (fabs(q) < pivmin) q = -pivmin;
    cnt += q < 0.0;
  }
  return cnt;
}

// One CTA (4 warps) per wanted eigenvalue (j-th largest, j = blockIdx.x).  lam[j] descending.
// info[0] = ||T||_1-ish scale (max Gershgorin radius bound), written by warp 0 for T3.
constexpr int kBsThreads = 256, kBsU = 1;   // sections per pass = kBsThreads * kBsU

__global__ void __launch_bounds__(kBsThreads)
bisect_kernel(const double* __restrict__ diag, const double* __restrict__ off, int d, int k,
              double* __restrict__ lam, double* __restrict__ info) {
  extern __shared__ double bs_sm[];
  double* dg = bs_sm;
  double* e2 = dg + d;
  __shared__ double s_lo[kBsThreads / 32], s_hi[kBsThreads / 32], s_emax[kBsThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double lo = DBL_MAX, hi = -DBL_MAX, emax = 0.0;
  for (int i = tid; i < d; i += blockDim.x) {
    const double a = diag[i];
    const double el = i > 0 ? fabs(off[i - 1]) : 0.0, er = i < d - 1 ? fabs(off[i]) : 0.0;
    dg[i] = a;
    e2[i] = i < d - 1 ? off[i] * off[i] : 0.0;
    lo = fmin(lo, a - el - er);
    hi = fmax(hi, a + el + er);
    emax = fmax(emax, er * er);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    emax = fmax(emax, __shfl_xor_sync(0xffffffffu, emax, o));
  }
  if (lane == 0) { s_lo[warp] = lo; s_hi[warp] = hi; s_emax[warp] = emax; }
  __syncthreads();
  const int nwarp = blockDim.x / 32;
  for (int i = 0; i < nwarp; ++i) { lo = fmin(lo, s_lo[i]); hi = fmax(hi, s_hi[i]); emax = fmax(emax, s_emax[i]); }
  const double tnorm = fmax(fabs(lo), fabs(hi));
  const double pivmin = fmax(DBL_MIN * fmax(1.0, emax), DBL_MIN);
  // widen the Gershgorin interval a little (dstebz)
  lo -= 2.0 * DBL_EPSILON * tnorm * d + 2.0 * pivmin;
  hi += 2.0 * DBL_EPSILON * tnorm * d + 2.0 * pivmin;
  if (blockIdx.x == 0 && tid == 0) { info[0] = tnorm; info[1] = pivmin; }

  const int j = blockIdx.x;   // one CTA per wanted eigenvalue
  if (j >= k) return;
  const int idx = d - 1 - j;  // 0-based index in ascending order: count(x) <= idx  <=>  x <= lambda_idx
  if (!(tnorm > 0.0)) {       // zero matrix
    if (tid == 0) lam[j] = 0.0;
    return;
  }
  // Sturm count in PRODUCT form on T / ||T||: p_0 = 1, p_1 = d_0 - x, p_i = (d_{i-1} - x) p_{i-1} - e_{i-2}^2 p_{i-2}; the
  // pivots of the LDL^T form are q_i = p_i / p_{i-1}, so the count of negative pivots is the count of sign changes of
  // p.  The dependent chain is ONE fma per row (the product e^2 p_{i-2} is a step old) instead of fma + reciprocal
  // (~5 dependent fp64 operations): 255 steps of ~130 cycles were 82 us of the 0.73 ms eigensolve at d = 256.  The pair
  // (p_{i-1}, p_i) is rescaled by a power of two every 8 rows (|d - x| <= 2, e^2 <= 1 after the scaling by ||T||: at most
  // 3^8 of growth in between); an exact zero counts as a negative pivot, like dlaebz's -pivmin.
  __syncthreads();            // everybody has read the Gershgorin bounds off dg / e2
  const double inv = 1.0 / tnorm;
  for (int i = tid; i < d; i += blockDim.x) { dg[i] *= inv; e2[i] *= inv * inv; }
  __syncthreads();
  lo *= inv;
  hi *= inv;
  // kBsThreads sections per pass, one recurrence per thread; a pass narrows the interval 513-fold
  __shared__ int s_first[kBsThreads / 32];
  const int S = kBsU * kBsThreads;
  const double atol = 1e-13;   // of ||T||: five 513-section passes from the Gershgorin interval; T3 refines by Rayleigh quotient
  for (int it = 0; it < 12; ++it) {
    const double width = hi - lo;
    if (!(width > atol)) break;   // CTA-uniform: lo, hi are identical in every thread
    const double x = lo + width * (double)(tid + 1) / (double)(S + 1);
    double pp = 1.0, pc = dg[0] - x;
    if (pc == 0.0) pc = -1e-300;
    int cnt = pc < 0.0;
    int i = 1;
    auto step = [&](int r) {
      double pn = fma(dg[r] - x, pc, -(e2[r - 1] * pp));
      if (pn == 0.0) pn = pc < 0.0 ? 1e-300 : -1e-300;          // virtual sign opposite to p_{i-1}: a negative pivot
      cnt += (__double2hiint(pn) ^ __double2hiint(pc)) < 0;     // sign change
      pp = pc;
      pc = pn;
    };
    for (; i + 8 <= d; i += 8) {
#pragma unroll
      for (int u = 0; u < 8; ++u) step(i + u);
      // rescale by 2^-(exponent of the larger one)
      const int ec = (__double2hiint(pc) >> 20) & 0x7ff, ep = (__double2hiint(pp) >> 20) & 0x7ff;
      const int em = ec > ep ? ec : ep;
      const double sc = __hiloint2double((2046 - (em < 1 ? 1 : (em > 2045 ? 2045 : em))) << 20, 0);
      pc *= sc;
      pp *= sc;
      if (pc == 0.0) pc = (cnt & 1) ? -1e-300 : 1e-300;          // underflow: keep the sign the count implies
    }
    for (; i < d; ++i) step(i);
    const int fl = cnt > idx ? 0 : 1;   // counts are monotone in x up to rounding: first shift right of lambda_idx
    const unsigned m = __ballot_sync(0xffffffffu, fl == 0);
    const int lane1 = m ? __ffs(m) - 1 : 0;
    if (lane == 0) s_first[warp] = m ? warp * 32 + lane1 : S;
    __syncthreads();
    int first = S;
    for (int w = 0; w < nwarp; ++w) first = min(first, s_first[w]);
    __syncthreads();
    const double nlo = first == 0 ? lo : lo + width * (double)first / (double)(S + 1);
    const double nhi = first == S ? hi : lo + width * (double)(first + 1) / (double)(S + 1);
    lo = nlo;
    hi = nhi;
  }
  if (tid == 0) lam[j] = 0.5 * (lo + hi) * tnorm;
}

// ---------------------------------------------------------------- T3: eigenvectors by inverse iteration
// Two tiers of closeness (LAPACK dstein has one: gap < 1e-3 ||T|| -> sequential modified Gram-Schmidt):
//   * TIGHT cluster, gap <= 1e-7 ||T|| (degenerate or nearly so): the CTA of the first member computes the cluster's
//     vectors one after the other, re-orthogonalising against the earlier ones inside every iteration, as dstein does.
//   * LOOSE cluster, 1e-7 ||T|| < gap <= 1e-3 ||T||: every member (every tight sub-cluster) gets its own CTA.  Their
//     residuals are ~eps ||T||, so two of these vectors are orthogonal to ~2 eps ||T|| / gap <= 2e-9 already; one
//     classical Gram-Schmidt step against the ORIGINAL earlier members (invit_orth_kernel, all members in parallel)
//     leaves errors of second order (the p = i term of z_i' . z_j' cancels E_ij exactly).  A PCA spectrum with thousands
//     of samples per feature is one big loose cluster (config 5: 100 eigenvalues 3e-4 ||T|| apart): processed
//     sequentially the stage took 180 ms at d = 4096.
// Z: k x d (rows), unit vectors of T.
constexpr int kInvThreads = 512;   // thread 0 carries the tridiagonal solves; the others matter in the Gram-Schmidt passes of tight clusters
constexpr double kTightGap = 1e-7, kLooseGap = 1e-3;

__device__ __forceinline__ double inv_block_sum(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0;
#pragma unroll
  for (int w = 0; w < kInvThreads / 32; ++w) t += red[w];
  return t;
}

__device__ __forceinline__ double hash_uniform(unsigned a, unsigned b) {
  unsigned x = a * 0x9E3779B1u ^ (b + 0x7F4A7C15u) * 0x85EBCA77u;
  x ^= x >> 15; x *= 0x2C1B3C6Du; x ^= x >> 12; x *= 0x297A2D39u; x ^= x >> 15;
  return ((double)x + 0.5) * (1.0 / 4294967296.0) - 0.5;
}

__global__ void __launch_bounds__(kInvThreads)
invit_kernel(const double* __restrict__ diag, const double* __restrict__ off, int d, int k,
             const double* __restrict__ lam, const double* __restrict__ info, double* __restrict__ lam_out,
             double* __restrict__ Z, double* __restrict__ coef_g) {
  const int j0 = blockIdx.x;
  const double tnorm = info[0];
  const double ortol = kTightGap * tnorm;
  if (j0 > 0 && fabs(lam[j0 - 1] - lam[j0]) <= ortol) return;  // not the head of a cluster
  int j1 = j0 + 1;
  while (j1 < k && fabs(lam[j1 - 1] - lam[j1]) <= ortol) ++j1;

  extern __shared__ double iv_sm[];
  double* a = iv_sm;        // U diagonal
  double* b = a + d;        // U first super-diagonal
  double* c = b + d;        // multipliers
  // (the second super-diagonal of U is not stored: it is off[i+1] where rows were swapped and 0 elsewhere, so that
  //  the kernel needs 52 d bytes of shared memory and d = 4096 - BASELINE config 5 - fits the 227 KB limit)
  double* z = c + d;        // iterate
  double* dg0 = z + d;      // diag - x (read-only during the factorisation)
  double* of0 = dg0 + d;    // off
  int* piv = reinterpret_cast<int*>(of0 + d);
  // Gram-Schmidt coefficients of a tight cluster (fewer than k): behind piv (d ints) in shared memory, or - when
  // 6.5 d + k doubles pass the 227 KB (d = 4096 with k > ~2300) - in a global array of k doubles, clusters being disjoint
  double* coef = coef_g ? coef_g + j0 : of0 + d + (d + 1) / 2;
  __shared__ double red[kInvThreads / 32];
  const int tid = threadIdx.x;
  const double eps = DBL_EPSILON;
  const double tiny = fmax(eps * tnorm, DBL_MIN * 1e16);

  double prev = 0.0;
  for (int j = j0; j < j1; ++j) {
    // perturb repeated values so that cluster members get distinct shifts (dstein)
    double x = lam[j];
    if (j > j0 && prev - x < 10.0 * eps * tnorm) x = prev - 10.0 * eps * tnorm;
    prev = x;
    // ---- factor T - x I = P L U (partial pivoting), sequential; a[] ends up holding 1 / pivot.  The running
    // diagonal / super-diagonal entries are carried in registers (a shared-memory store -> load round trip per
    // step was most of the chain); the inputs come from read-only copies so their loads can run ahead.
    for (int i = tid; i < d; i += kInvThreads) {
      dg0[i] = diag[i] - x;
      of0[i] = i < d - 1 ? off[i] : 0.0;
      z[i] = hash_uniform((unsigned)i, (unsigned)j * 7919u + 13u);
    }
    __syncthreads();
    if (tid == 0) {
      const double* __restrict__ dgr = dg0;
      const double* __restrict__ ofr = of0;
      double ai = dgr[0], bi = ofr[0];
      for (int i = 0; i < d - 1; ++i) {
        const double ci = ofr[i];
        const double an = dgr[i + 1], bn = ofr[i + 1];   // of0[d-1] = 0
        if (fabs(ai) >= fabs(ci) || fabs(ci) < tiny) {
          double pv = ai;
          if (fabs(pv) < tiny) pv = pv < 0 ? -tiny : tiny;
          const double ip = fast_rcp(pv);
          const double m = ci * ip;
          a[i] = ip;
          b[i] = bi;
          c[i] = m;
          piv[i] = 0;
          ai = fma(-m, bi, an);
          bi = bn;
        } else {
          const double ip = fast_rcp(ci);
          const double m = ai * ip;
          a[i] = ip;
          b[i] = an;
          c[i] = m;
          piv[i] = 1;
          ai = fma(-m, an, bi);
          bi = i < d - 2 ? -m * bn : 0.0;
        }
      }
      double pv = ai;
      if (fabs(pv) < tiny) pv = pv < 0 ? -tiny : tiny;
      a[d - 1] = fast_rcp(pv);
      b[d - 1] = 0.0;
      c[d - 1] = 0.0;
      piv[d - 1] = 0;
    }
    __syncthreads();
    for (int iter = 0; iter < 3; ++iter) {
      if (tid == 0) {
        if (iter > 0) {  // apply P, L to the right-hand side (the first solve uses U only, as dstein does)
          double zi = z[0];
          for (int i = 0; i < d - 1; ++i) {
            const double zn = z[i + 1];
            if (piv[i] == 0) { z[i] = zi; zi = fma(-c[i], zi, zn); }
            else { z[i] = zn; zi = fma(-c[i], zn, zi); }
          }
          z[d - 1] = zi;
        }
        double z1 = 0.0, z2 = 0.0;   // z[i+1], z[i+2]; b[d-1] = 0, second super-diagonal 0 in the last two rows
        for (int i = d - 1; i >= 0; --i) {  // U z = y
          const double d2i = (piv[i] != 0 && i < d - 2) ? of0[i + 1] : 0.0;   // off the dependent chain
          const double t = fma(-d2i, z2, fma(-b[i], z1, z[i]));
          const double zi = t * a[i];
          z[i] = zi;
          z2 = z1;
          z1 = zi;
        }
      }
      __syncthreads();
      // re-orthogonalise against the cluster's earlier vectors, normalise.  Classical Gram-Schmidt, twice ("twice is
      // enough"): all j - j0 coefficients of a pass are formed at once, one warp per earlier vector with four vectors in
      // flight, then one update sweep - the modified Gram-Schmidt loop this replaces paid an L2 round trip and two block
      // barriers per earlier vector and iteration (k = d = 256 on a rank-deficient matrix, a zero cluster of ~170: 43 ms)
      if (j > j0) {
        const int lane = tid & 31, warp = tid >> 5;
        constexpr int NW = kInvThreads / 32;
        for (int pass = 0; pass < 2; ++pass) {
          for (int p0 = j0 + 4 * warp; p0 < j; p0 += 4 * NW) {
            double acc[4] = {0.0, 0.0, 0.0, 0.0};
            for (int i = lane; i < d; i += 32) {
              const double zi = z[i];
#pragma unroll
              for (int u = 0; u < 4; ++u)
                if (p0 + u < j) acc[u] = fma(zi, Z[(size_t)(p0 + u) * d + i], acc[u]);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
              for (int u = 0; u < 4; ++u) acc[u] += __shfl_xor_sync(0xffffffffu, acc[u], o);
            }
            if (lane == 0) {
#pragma unroll
              for (int u = 0; u < 4; ++u) if (p0 + u < j) coef[p0 + u - j0] = acc[u];
            }
          }
          __syncthreads();
          for (int i = tid; i < d; i += kInvThreads) {
            double a0 = z[i], a1 = 0.0, a2 = 0.0, a3 = 0.0;
            int p = j0;
            for (; p + 4 <= j; p += 4) {
              a0 = fma(-coef[p - j0], Z[(size_t)p * d + i], a0);
              a1 = fma(-coef[p + 1 - j0], Z[(size_t)(p + 1) * d + i], a1);
              a2 = fma(-coef[p + 2 - j0], Z[(size_t)(p + 2) * d + i], a2);
              a3 = fma(-coef[p + 3 - j0], Z[(size_t)(p + 3) * d + i], a3);
            }
            for (; p < j; ++p) a0 = fma(-coef[p - j0], Z[(size_t)p * d + i], a0);
            z[i] = (a0 + a1) + (a2 + a3);
          }
          __syncthreads();
        }
      }
      double mx = 0.0;
      for (int i = tid; i < d; i += kInvThreads) mx = fmax(mx, fabs(z[i]));
      // max via sum of squares after a power-of-two-free rescale: two passes keep it overflow safe
      __syncthreads();
      red[tid >> 5] = 0.0;
      __syncthreads();
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      if ((tid & 31) == 0) red[tid >> 5] = mx;
      __syncthreads();
      double gmx = 0.0;
      for (int wv = 0; wv < kInvThreads / 32; ++wv) gmx = fmax(gmx, red[wv]);
      __syncthreads();
      const double inv = gmx > 0.0 && isfinite(gmx) ? 1.0 / gmx : 1.0;
      double ss = 0.0;
      for (int i = tid; i < d; i += kInvThreads) { const double t = z[i] * inv; z[i] = t; ss += t * t; }
      ss = inv_block_sum(ss, red);
      const double rn = ss > 0.0 ? rsqrt(ss) : 0.0;
      for (int i = tid; i < d; i += kInvThreads) z[i] *= rn;
      __syncthreads();
    }
    // Rayleigh quotient z^T T z refines the multisection value (lam itself stays read-only: the
    // other CTAs derive the cluster boundaries from it)
    double rq = 0.0;
    for (int i = tid; i < d; i += kInvThreads) {
      double tz = diag[i] * z[i];
      if (i > 0) tz += off[i - 1] * z[i - 1];
      if (i < d - 1) tz += off[i] * z[i + 1];
      rq += z[i] * tz;
      Z[(size_t)j * d + i] = z[i];
    }
    rq = inv_block_sum(rq, red);
    if (tid == 0) lam_out[j] = rq;
    __threadfence_block();
    __syncthreads();
  }
}

// Second half of T3: one classical Gram-Schmidt step inside every LOOSE cluster.  CTA j: z_j' = z_j - sum_p (z_p . z_j) z_p
// over the earlier members p of its loose cluster (original vectors, read from Zin), normalised, written to Zout; the
// other vectors are copied.  grid = k, block = 256.
__global__ void __launch_bounds__(256)
invit_orth_kernel(const double* __restrict__ lam, const double* __restrict__ info, int d, int k,
                  const double* __restrict__ Zin, double* __restrict__ Zout) {
  extern __shared__ double io_sm[];
  double* coef = io_sm;                 // k coefficients
  __shared__ double red[8];
  const int j = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double ltol = kLooseGap * info[0];
  int L = j;
  while (L > 0 && fabs(lam[L - 1] - lam[L]) <= ltol) --L;
  const double* zj = Zin + (size_t)j * d;
  if (L == j) {
    for (int i = tid; i < d; i += 256) Zout[(size_t)j * d + i] = zj[i];
    return;
  }
  // dots: one warp per earlier member
  for (int p = L + warp; p < j; p += 8) {
    const double* zp = Zin + (size_t)p * d;
    double dot = 0.0;
    for (int i = lane; i < d; i += 32) dot = fma(zp[i], zj[i], dot);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    if (lane == 0) coef[p - L] = dot;
  }
  __syncthreads();
  double ss = 0.0;
  for (int i = tid; i < d; i += 256) {
    double z = zj[i];
    for (int p = L; p < j; ++p) z = fma(-coef[p - L], Zin[(size_t)p * d + i], z);
    Zout[(size_t)j * d + i] = z;
    ss = fma(z, z, ss);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  if (lane == 0) red[warp] = ss;
  __syncthreads();
  double t = 0.0;
#pragma unroll
  for (int w = 0; w < 8; ++w) t += red[w];
  const double rn = t > 0.0 ? rsqrt(t) : 0.0;
  for (int i = tid; i < d; i += 256) Zout[(size_t)j * d + i] *= rn;
}

// ---------------------------------------------------------------- T4: v = H_0 H_1 ... H_{d-3} z
__global__ void __launch_bounds__(256)
backtransform_kernel(const double* __restrict__ Vh, const double* __restrict__ tau, int d, int k,
                     double* __restrict__ Z) {
  // One CTA per vector, TWO reflectors per block reduction (as backtransform_pair_kernel does per warp): with
  // a = v_s.z, b = v_{s-1}.z, g = v_{s-1}.v_s reduced together, H_{s-1} H_s z = z - ca v_s - cb v_{s-1},
  // ca = tau_s a, cb = tau_{s-1} (b - ca g).  One reflector per reduction took two CTA barriers each (441 us at d = 784).
  extern __shared__ double bt_sm[];
  double* z = bt_sm;
  __shared__ double red[2][3][8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, j = blockIdx.x;
  for (int i = tid; i < d; i += 256) z[i] = Z[(size_t)j * d + i];
  __syncthreads();
  int s = d - 3, par = 0;
  if (s >= 0 && ((d - 2) & 1)) {   // odd count: one single reflector first
    const double* vr = Vh + (size_t)s * d;
    double dot = 0.0;
    for (int i = s + 1 + tid; i < d; i += 256) dot += vr[i] * z[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    if (lane == 0) red[par][0][warp] = dot;
    __syncthreads();
    double tot = 0.0;
#pragma unroll
    for (int wv = 0; wv < 8; ++wv) tot += red[par][0][wv];
    tot *= tau[s];
    for (int i = s + 1 + tid; i < d; i += 256) z[i] -= tot * vr[i];
    __syncthreads();
    par ^= 1;
    --s;
  }
  // rows are prefetched one pair ahead (registers) for d <= 256 * kBtPer, so the L2 latency is off the dependent chain
  constexpr int kBtPer = 4;
  const bool pre = d <= 256 * kBtPer;
  double na[kBtPer], nb[kBtPer];
  auto fetch = [&](int sp) {   // reflectors sp and sp - 1 (T1 stores exact zeros below a reflector's leading 1)
#pragma unroll
    for (int u = 0; u < kBtPer; ++u) {
      const int i = tid + 256 * u;
      const bool in = i < d && sp >= 1;
      na[u] = in ? __ldg(&Vh[(size_t)sp * d + i]) : 0.0;
      nb[u] = in ? __ldg(&Vh[(size_t)(sp - 1) * d + i]) : 0.0;
    }
  };
  auto i0 = [&](int from) { return tid >= from ? tid : tid + (from - tid + 255) / 256 * 256; };   // first i >= from with i = tid mod 256
  if (pre) fetch(s);
  for (; s >= 1; s -= 2) {
    const double ta = tau[s], tb = tau[s - 1];
    double ca[kBtPer], cb[kBtPer];
    double a = 0.0, b = 0.0, g = 0.0;
    const double* va = Vh + (size_t)s * d;
    const double* vb = va - d;
    if (pre) {
#pragma unroll
      for (int u = 0; u < kBtPer; ++u) { ca[u] = na[u]; cb[u] = nb[u]; }
      fetch(s - 2);
#pragma unroll
      for (int u = 0; u < kBtPer; ++u) {
        const int i = tid + 256 * u;
        const double zi = i < d ? z[i] : 0.0;
        a = fma(ca[u], zi, a);
        b = fma(cb[u], zi, b);
        g = fma(ca[u], cb[u], g);
      }
    } else {
      // v_{s-1} starts at column s, v_s at s + 1 (zeros below); a thread keeps the elements i = tid mod 256 for good
      for (int i = i0(s); i < d; i += 256) {
        const double x = va[i], y = vb[i], zi = z[i];
        a = fma(x, zi, a);
        b = fma(y, zi, b);
        g = fma(x, y, g);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      b += __shfl_xor_sync(0xffffffffu, b, o);
      g += __shfl_xor_sync(0xffffffffu, g, o);
    }
    if (lane == 0) { red[par][0][warp] = a; red[par][1][warp] = b; red[par][2][warp] = g; }
    __syncthreads();
    double A = 0.0, B = 0.0, G = 0.0;
#pragma unroll
    for (int wv = 0; wv < 8; ++wv) { A += red[par][0][wv]; B += red[par][1][wv]; G += red[par][2][wv]; }
    par ^= 1;
    const double fa = ta * A;
    const double fb = tb * fma(-fa, G, B);
    if (pre) {
#pragma unroll
      for (int u = 0; u < kBtPer; ++u) {
        const int i = tid + 256 * u;
        if (i < d) z[i] = fma(-fb, cb[u], fma(-fa, ca[u], z[i]));
      }
    } else {
      for (int i = i0(s); i < d; i += 256) z[i] = fma(-fb, vb[i], fma(-fa, va[i], z[i]));
    }
    // every thread reads and writes only ITS elements of z; the partials alternate between two buffers, so one barrier
    // per pair is enough (a warp that races ahead writes the other buffer; a third pair cannot start before the second's barrier)
  }
  __syncthreads();
  for (int i = tid; i < d; i += 256) Z[(size_t)j * d + i] = z[i];
}

// T4, d <= 512: one WARP per vector, z and the reflector row in registers (element i = lane + 32 u),
// no block barrier on the dependent chain (the CTA-per-vector version needs two per reflector).
// Four warps per CTA so that the fp64 work spreads over many SMs (fp64 issue is scarce on B200: a
// 32-warp CTA with the reflectors staged in shared memory was slower, measured).  The next reflector
// row and tau are fetched one step ahead.
template <int PER>
__global__ void __launch_bounds__(128)
backtransform_warp_kernel(const double* __restrict__ Vh, const double* __restrict__ tau, int d, int k,
                          double* __restrict__ Z) {
  const int lane = threadIdx.x & 31, j = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (j >= k) return;
  double z[PER], nxt[PER];
#pragma unroll
  for (int u = 0; u < PER; ++u) {
    const int i = lane + 32 * u;
    z[u] = i < d ? Z[(size_t)j * d + i] : 0.0;
    nxt[u] = 0.0;
  }
  auto fetch = [&](int s) {
    const double* vr = Vh + (size_t)s * d;
#pragma unroll
    for (int u = 0; u < PER; ++u) {
      const int i = lane + 32 * u;
      nxt[u] = (i > s && i < d) ? __ldg(&vr[i]) : 0.0;
    }
  };
  double tnext = 0.0;
  if (d >= 3) { fetch(d - 3); tnext = __ldg(&tau[d - 3]); }
  for (int s = d - 3; s >= 0; --s) {
    double cur[PER];
#pragma unroll
    for (int u = 0; u < PER; ++u) cur[u] = nxt[u];
    const double t = tnext;
    if (s > 0) { fetch(s - 1); tnext = __ldg(&tau[s - 1]); }
    double dot = 0.0;
#pragma unroll
    for (int u = 0; u < PER; ++u) dot = fma(cur[u], z[u], dot);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    dot *= t;
#pragma unroll
    for (int u = 0; u < PER; ++u) z[u] = fma(-dot, cur[u], z[u]);
  }
#pragma unroll
  for (int u = 0; u < PER; ++u) {
    const int i = lane + 32 * u;
    if (i < d) Z[(size_t)j * d + i] = z[u];
  }
}

// T4 for d <= 256: one warp per vector, TWO reflectors per shuffle tree.  With a = v_s.z, b = v_{s-1}.z and
// g = v_{s-1}.v_s reduced together, H_{s-1} H_s z = z - ca v_s - cb v_{s-1}, ca = tau_s a, cb = tau_{s-1} (b - ca g).
// The single-reflector kernel above spent ~560 cycles per reflector (an 8-long dependent dot, a 5-level shuffle
// tree, and an L2 fetch only one step ahead); this one prefetches two pairs ahead (three register sets in rotation).
template <int PER>
struct ReflPair {
  double a[PER], b[PER];
  double ta, tb;
};

template <int PER>
__global__ void __launch_bounds__(128)
backtransform_pair_kernel(const double* __restrict__ Vh, const double* __restrict__ tau, int d, int k,
                          double* __restrict__ Z) {
  const int lane = threadIdx.x & 31, j = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (j >= k) return;
  double z[PER];
#pragma unroll
  for (int u = 0; u < PER; ++u) {
    const int i = lane + 32 * u;
    z[u] = i < d ? Z[(size_t)j * d + i] : 0.0;
  }
  int s = d - 3;   // next reflector to apply (H_s, s = d-3 ... 0)
  if (s >= 0 && ((d - 2) & 1)) {   // odd count: one single reflector first
    const double* vr = Vh + (size_t)s * d;
    double cur[PER], dot = 0.0;
#pragma unroll
    for (int u = 0; u < PER; ++u) {
      const int i = lane + 32 * u;
      cur[u] = (i > s && i < d) ? __ldg(&vr[i]) : 0.0;
      dot = fma(cur[u], z[u], dot);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    dot *= __ldg(&tau[s]);
#pragma unroll
    for (int u = 0; u < PER; ++u) z[u] = fma(-dot, cur[u], z[u]);
    --s;
  }
  bool inb[PER];   // T1 stores the reflector rows with exact zeros up to their leading 1: no per-element mask
#pragma unroll
  for (int u = 0; u < PER; ++u) inb[u] = lane + 32 * u < d;
  auto fetch = [&](ReflPair<PER>& P, int sp) {   // reflectors sp and sp-1
    if (sp < 1) return;
    const double* va = Vh + (size_t)sp * d + lane;
    const double* vb = va - d;
#pragma unroll
    for (int u = 0; u < PER; ++u) {
      P.a[u] = inb[u] ? __ldg(&va[32 * u]) : 0.0;
      P.b[u] = inb[u] ? __ldg(&vb[32 * u]) : 0.0;
    }
    P.ta = __ldg(&tau[sp]);
    P.tb = __ldg(&tau[sp - 1]);
  };
  auto apply = [&](const ReflPair<PER>& P) {
    double a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0, g0 = 0.0, g1 = 0.0;
#pragma unroll
    for (int u = 0; u < PER; u += 2) {
      a0 = fma(P.a[u], z[u], a0);
      b0 = fma(P.b[u], z[u], b0);
      g0 = fma(P.a[u], P.b[u], g0);
      if (u + 1 < PER) {
        a1 = fma(P.a[u + 1], z[u + 1], a1);
        b1 = fma(P.b[u + 1], z[u + 1], b1);
        g1 = fma(P.a[u + 1], P.b[u + 1], g1);
      }
    }
    double a = a0 + a1, b = b0 + b1, g = g0 + g1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      b += __shfl_xor_sync(0xffffffffu, b, o);
      g += __shfl_xor_sync(0xffffffffu, g, o);
    }
    const double ca = P.ta * a;
    const double cb = P.tb * fma(-ca, g, b);
#pragma unroll
    for (int u = 0; u < PER; ++u) z[u] = fma(-cb, P.b[u], fma(-ca, P.a[u], z[u]));
  };
  ReflPair<PER> P0, P1, P2;
  fetch(P0, s);
  fetch(P1, s - 2);
  while (s >= 1) {
    fetch(P2, s - 4);
    apply(P0);
    s -= 2;
    if (s < 1) break;
    fetch(P0, s - 4);
    apply(P1);
    s -= 2;
    if (s < 1) break;
    fetch(P1, s - 4);
    apply(P2);
    s -= 2;
  }
#pragma unroll
  for (int u = 0; u < PER; ++u) {
    const int i = lane + 32 * u;
    if (i < d) Z[(size_t)j * d + i] = z[u];
  }
}

}  // namespace

// ---------------------------------------------------------------- host side
static size_t tri_r256(size_t b) { return (b + 255) / 256 * 256; }

size_t eigh_topk_workspace_bytes(int d, int k) {
  return tri_r256((size_t)d * d * 8)      // Vh
         + 3 * tri_r256((size_t)d * 8)    // diag, off, tau
         + tri_r256((size_t)4 * d * 8)    // p' / published-row exchange
         + tri_r256((size_t)k * 8)        // lam
         + tri_r256((size_t)k * d * 8)    // inverse-iteration vectors before the loose-cluster Gram-Schmidt step
         + tri_r256((size_t)k * 8)        // Gram-Schmidt coefficients of tight clusters when they do not fit shared memory
         + 256 + 256;                     // info, ctrl
}

void launch_eigh_topk(Ctx& c, double* dA, int d, int k, double* dW, double* dVt, void* workspace) {
  if (k > d) k = d;
  char* p = static_cast<char*>(workspace);
  auto take = [&](size_t bytes) { char* q = p; p += tri_r256(bytes); return q; };
  double* Vh = reinterpret_cast<double*>(take((size_t)d * d * 8));
  double* diag = reinterpret_cast<double*>(take((size_t)d * 8));
  double* off = reinterpret_cast<double*>(take((size_t)d * 8));
  double* tau = reinterpret_cast<double*>(take((size_t)d * 8));
  double* xbuf = reinterpret_cast<double*>(take((size_t)4 * d * 8));
  double* info = reinterpret_cast<double*>(take(256));
  TriCtrl* ctrl = reinterpret_cast<TriCtrl*>(take(256));
  double* lam = reinterpret_cast<double*>(take((size_t)k * 8));
  double* Zraw = reinterpret_cast<double*>(take((size_t)k * d * 8));
  double* coef_ws = reinterpret_cast<double*>(take((size_t)k * 8));
  DR4G_CUDA(cudaMemsetAsync(ctrl, 0, sizeof(TriCtrl), c.stream));

  // T1: d <= 512 runs inside one thread-block cluster (hardware barrier, DSMEM exchange); larger d uses
  // the cooperative grid with an atomic barrier, rows in shared memory whenever they fit
  if (d <= 512 && d % 2 == 0 && c.eig_solver != 2) {   // the cluster kernel sweeps rows in aligned column pairs
    static const int cs_env = getenv("DR4G_TRI_CS") ? atoi(getenv("DR4G_TRI_CS")) : 0;  // tuning hook
    // 16 beats 8 at d = 256 (1.16 vs 1.28 ms); 8 beats 1 for 32 < d <= 128 (whole eigensolve, round-2 sweep on B200:
    // d = 64: 0.273 -> 0.210 ms, d = 100: 0.482 -> 0.303, d = 128 k = 32: 0.698 -> 0.415; 16 is no better than 8 there)
    const int cs = cs_env > 0 ? cs_env : (d <= 32 ? 1 : (d <= 128 ? 8 : 16));
    const size_t smem = ((size_t)7 * d + (size_t)((d + cs - 1) / cs) * d) * sizeof(double);
    static PerDeviceMax smem_set, nonportable_set;
    if (smem > 48 * 1024 && smem_set.need(c.device, smem)) {
      DR4G_CUDA(cudaFuncSetAttribute(tridiag_cluster_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      DR4G_CUDA(cudaFuncSetAttribute(tridiag_cluster_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      DR4G_CUDA(cudaFuncSetAttribute(tridiag_cluster_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    if (cs > 8 && nonportable_set.need(c.device, 1)) {
      DR4G_CUDA(cudaFuncSetAttribute(tridiag_cluster_kernel<0>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
      DR4G_CUDA(cudaFuncSetAttribute(tridiag_cluster_kernel<2>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
      DR4G_CUDA(cudaFuncSetAttribute(tridiag_cluster_kernel<4>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(cs);
    cfg.blockDim = dim3(kTriThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = c.stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    static const int tri_trace = getenv("DR4G_TRI_TRACE") ? atoi(getenv("DR4G_TRI_TRACE")) : 0;
    static const int no_reg = getenv("DR4G_TRI_NOREG") ? atoi(getenv("DR4G_TRI_NOREG")) : 0;
    const int own = (d + cs - 1) / cs;   // own rows per CTA
    const int rw = no_reg ? 0 : (d <= 256 && own <= 2 * kTriWarps ? 2 : (d <= 512 && own <= 4 * kTriWarps ? 4 : 0));
    if (rw == 2)
      DR4G_CUDA(cudaLaunchKernelEx(&cfg, tridiag_cluster_kernel<2>, (const double*)dA, d, diag, off, tau, Vh, xbuf, tri_trace));
    else if (rw == 4)
      DR4G_CUDA(cudaLaunchKernelEx(&cfg, tridiag_cluster_kernel<4>, (const double*)dA, d, diag, off, tau, Vh, xbuf, tri_trace));
    else
      DR4G_CUDA(cudaLaunchKernelEx(&cfg, tridiag_cluster_kernel<0>, (const double*)dA, d, diag, off, tau, Vh, xbuf, tri_trace));
    if (tri_trace & 1) {
      long long h[64];
      DR4G_CUDA(cudaStreamSynchronize(c.stream));
      DR4G_CUDA(cudaMemcpyFromSymbol(h, g_tri_trace, sizeof(h)));
      fprintf(stderr, "tridiag trace d=%d (cycles from step start): operands_arrived pv_done nrm_done vn_done rows_done pushes_issued\n", d);
      for (int i = 0; i < 4; ++i) {
        fprintf(stderr, "  step %d:", 100 + i);
        for (int e = 1; e < 7; ++e) fprintf(stderr, " %6lld", h[i * 16 + e] - h[i * 16]);
        if (i < 3) fprintf(stderr, "  | next step starts at %6lld", h[(i + 1) * 16] - h[i * 16]);
        fprintf(stderr, "\n");
      }
    }
  } else {
    int grid = (d + 2 * kTriWarps - 1) / (2 * kTriWarps);
    if (grid > c.sm_count) grid = c.sm_count;
    if (grid < 1) grid = 1;
    const int nw = grid * kTriWarps;
    const size_t vec_smem = (size_t)5 * d * sizeof(double);
    const size_t rows_smem = (size_t)((d + nw - 1) / nw) * kTriWarps * d * sizeof(double);
    const bool in_smem = vec_smem + rows_smem <= 200 * 1024;
    const size_t tri_smem = vec_smem + (in_smem ? rows_smem : 0);
    const void* tri_fn = in_smem ? (const void*)tridiag_kernel<true> : (const void*)tridiag_kernel<false>;
    static PerDeviceMax tri_smem_set[2];
    if (tri_smem > 48 * 1024 && tri_smem_set[in_smem].need(c.device, tri_smem))
      DR4G_CUDA(cudaFuncSetAttribute(tri_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tri_smem));
    void* args[] = {&dA, &d, &diag, &off, &tau, &Vh, &xbuf, &ctrl};
    DR4G_CUDA(cudaLaunchCooperativeKernel(tri_fn, dim3(grid), dim3(kTriThreads), args, tri_smem, c.stream));
  }
  // T2
  const size_t bs_smem = (size_t)2 * d * sizeof(double);
  static PerDeviceMax bs_smem_set;
  if (bs_smem > 48 * 1024 && bs_smem_set.need(c.device, bs_smem))
    DR4G_CUDA(cudaFuncSetAttribute(bisect_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bs_smem));
  bisect_kernel<<<k, kBsThreads, bs_smem, c.stream>>>(diag, off, d, k, lam, info);
  // T3
  size_t iv_smem = (size_t)6 * d * sizeof(double) + (size_t)((d + 1) / 2) * sizeof(double);
  if (iv_smem > 227 * 1024)
    fail("eigh_topk: d=%d needs %zu bytes of shared memory for the inverse iteration (limit 227 KB, d <= 4469)", d, iv_smem);
  double* coef_g = nullptr;
  // (k <= 512: 4 KB of shared memory at most; beyond that the coefficients would cost resident CTAs per SM - at
  // k = d = 2100 one instead of two - for the sake of tight clusters, which are rare)
  if (k <= 512 && iv_smem + (size_t)k * sizeof(double) <= 227 * 1024) iv_smem += (size_t)k * sizeof(double);
  else coef_g = coef_ws;
  static PerDeviceMax iv_smem_set;
  if (iv_smem > 48 * 1024 && iv_smem_set.need(c.device, iv_smem))
    DR4G_CUDA(cudaFuncSetAttribute(invit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)iv_smem));
  invit_kernel<<<k, kInvThreads, iv_smem, c.stream>>>(diag, off, d, k, lam, info, dW, Zraw, coef_g);
  invit_orth_kernel<<<k, 256, (size_t)k * sizeof(double), c.stream>>>(lam, info, d, k, Zraw, dVt);
  // T4
  const size_t bt_smem = (size_t)d * sizeof(double);
  if (d <= 256) backtransform_pair_kernel<8><<<(k + 3) / 4, 128, 0, c.stream>>>(Vh, tau, d, k, dVt);
  else if (d <= 512) backtransform_warp_kernel<16><<<(k + 3) / 4, 128, 0, c.stream>>>(Vh, tau, d, k, dVt);
  else backtransform_kernel<<<k, 256, bt_smem, c.stream>>>(Vh, tau, d, k, dVt);
  c.launches += 5;
  DR4G_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------- allocation-free variant (dr4g_eigh_topk_static_double)
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the
// same T1-T4 pipeline with its workspace (reflectors, tridiagonal, multisection values, inverse-iteration vectors) in a
// file-scope device array instead of the context's grow-only arena.  d <= kStaticEighMaxD; every GPU that loads the
// library holds the array (1,088 KiB) until exit.  All calls share it: calls that overlap on the GPU race.  Nothing in
// it has to start at zero except the grid-barrier counter, which launch_eigh_topk clears itself on the stream.
constexpr int kStaticEighMaxD = 256;
constexpr size_t kStaticEighBytes = (size_t)2 * kStaticEighMaxD * kStaticEighMaxD * 8 + 65536;
static __device__ __align__(256) unsigned char g_eigh_static_ws[kStaticEighBytes];

int eigh_topk_static_max_d() { return kStaticEighMaxD; }

void launch_eigh_topk_static(Ctx& c, double* dA, int d, int k, double* dW, double* dVt) {
  if (d < 1 || d > kStaticEighMaxD)
    fail("dr4g_eigh_topk_static_double: d=%d is beyond the static workspace (d <= %d); use dr4g_eigh_topk_double", d, kStaticEighMaxD);
  if (k < 1 || k > d) fail("dr4g_eigh_topk_static_double: bad k=%d for d=%d", k, d);
  if (eigh_topk_workspace_bytes(d, k) > kStaticEighBytes)
    fail("dr4g_eigh_topk_static_double: d=%d k=%d needs %zu bytes, the static workspace has %zu", d, k,
         eigh_topk_workspace_bytes(d, k), kStaticEighBytes);
  // In host code the array's name is a host-side stand-in, not a device address: the address comes from
  // cudaGetSymbolAddress, once per device, with that device current (the caller's DeviceGuard).
  constexpr int kMaxDev = 64;
  static void* addr[kMaxDev] = {};
  static std::mutex mu;
  if (c.device < 0 || c.device >= kMaxDev) fail("dr4g_eigh_topk_static_double: device index %d out of range", c.device);
  void* ws = nullptr;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (!addr[c.device]) {
      void* p = nullptr;
      DR4G_CUDA(cudaGetSymbolAddress(&p, g_eigh_static_ws));
      if (!p) fail("dr4g_eigh_topk_static_double: cudaGetSymbolAddress returned a null address");
      addr[c.device] = p;
    }
    ws = addr[c.device];
  }
  launch_eigh_topk(c, dA, d, k, dW, dVt, ws);
}

}  // namespace dr4g
