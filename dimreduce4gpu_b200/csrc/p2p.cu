// The exchange step of the row-sharded fit as ONE kernel over NVLink peer memory (replaces the ncclAllGather
// of the packed partials [G | s | q | n | c_r] when every rank of the communicator sits on one NVSwitch box).
//
// Every rank owns an exchange buffer (cudaMalloc + cudaIpcGetMemHandle; the peers map it with cudaIpcOpenMemHandle):
//   [parity 0: world slots | parity 1: world slots | flags 2 x world (128 B apart) | CTA counters 2]
// An exchange of `count` doubles with epoch e (parity e & 1), one launch per rank:
//   A  the CTAs PUSH the rank's partials into slot `rank` of every peer's buffer (and of its own) with plain 16-byte
//      stores over NVLink - the sender knows where the data goes, nobody has to ask for it;
//   B  each thread fences at system scope, the last CTA to finish (device-scope counter) releases flag[parity][rank] = e in
//      every peer's buffer;
//   C  the CTAs wait until the world flags of the LOCAL buffer carry e: all partials have landed here.
// The re-basing kernels (finalize.cu) then read the local buffer exactly as they read the all-gather's output, in rank
// order, so the ranks' models stay bit-identical.  Two parities make the reuse safe without a second handshake: a peer
// reaches epoch e + 2 only after it has seen this rank's flag of epoch e + 1, which this rank posts after (stream
// order) it has consumed the slots of epoch e.  The grid is small enough to be resident as a whole (the wait of step C
// must not keep CTAs of step A off the SMs), and the wait gives up after ~4 s and raises a status word instead of
// hanging the device.  NCCL stays the fallback (other topologies, larger payloads than the buffer) and keeps the
// all-reduce of the uncentred (TruncatedSVD) fit.
#include <cstring>

#include "common.cuh"

namespace dr4g {
namespace {

constexpr int kP2PGrid = 64, kP2PThreads = 256, kFlagStride = 16;   // flags 128 B apart (unsigned long long units)

struct P2PPeers {
  double* base[kP2PMaxWorld];
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(kP2PThreads)
p2p_allgather_kernel(const double* __restrict__ src, long long count, int rank, int world, P2PPeers peers, long long cap,
                     unsigned long long epoch, int* __restrict__ status) {
  const int par = (int)(epoch & 1ull);
  const size_t slot_off = (size_t)par * world * cap + (size_t)rank * count;   // slots packed at stride `count`
  const size_t flags_off = (size_t)2 * world * cap;
  // A: push (count is odd in general - d^2 + 3 d + 1 - so 8-byte stores; 2 MB per peer at d = 512)
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (int r = 0; r < world; ++r) {
    double* dst = peers.base[(rank + r) % world] + slot_off;   // start with the own buffer, then round the ring
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) dst[i] = src[i];
  }
  // B: everything this CTA wrote is visible system-wide before its arrival is
  __threadfence_system();
  __syncthreads();
  __shared__ int s_last;
  unsigned long long* local_flags = reinterpret_cast<unsigned long long*>(peers.base[rank] + flags_off);
  unsigned long long* ctr = local_flags + (size_t)2 * world * kFlagStride + par * kFlagStride;
  if (threadIdx.x == 0) {
    const unsigned long long old = atomicAdd(ctr, 1ull);
    s_last = old == (unsigned long long)gridDim.x - 1ull;
    if (s_last) *ctr = 0ull;   // next use of this parity: two exchanges away
  }
  __syncthreads();
  if (s_last && threadIdx.x < world) {
    __threadfence_system();
    unsigned long long* f = reinterpret_cast<unsigned long long*>(peers.base[threadIdx.x] + flags_off) +
                            ((size_t)par * world + rank) * kFlagStride;
    st_release_sys(f, epoch);
  }
  // C: wait for every rank's flag in the local buffer
  if (threadIdx.x < world) {
    const unsigned long long* f = local_flags + ((size_t)par * world + threadIdx.x) * kFlagStride;
    unsigned long long t0 = 0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    unsigned spins = 0;
    while (ld_acquire_sys(f) < epoch) {
      __nanosleep(64);
      if ((++spins & 1023u) == 0) {
        unsigned long long t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > 4000000000ull) {   // a peer never arrived: flag the failure instead of hanging the device
          atomicExch(status, 1 + threadIdx.x);
          break;
        }
      }
    }
  }
  __syncthreads();
}

}  // namespace

size_t p2p_buffer_doubles(int world, long long cap) {
  return (size_t)2 * world * cap + (size_t)2 * world * kFlagStride + 2 * kFlagStride;
}

// Pushes `count` doubles of `src` into every rank's buffer and waits for everybody's; returns the local parity region
// ([world][count], rank order), valid until the exchange after the next one.
const double* launch_p2p_allgather(Ctx& c, const double* src, long long count) {
  if (!c.p2p_ready || count > c.p2p_cap || c.world > kP2PMaxWorld) fail("p2p exchange not available for %lld doubles", count);
  P2PPeers peers;
  for (int r = 0; r < kP2PMaxWorld; ++r) peers.base[r] = r < c.world ? c.p2p_peer[r] : nullptr;
  const unsigned long long epoch = ++c.p2p_epoch;
  p2p_allgather_kernel<<<kP2PGrid, kP2PThreads, 0, c.stream>>>(src, count, c.rank, c.world, peers, c.p2p_cap, epoch, c.p2p_status);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
  return c.p2p_peer[c.rank] + (size_t)(epoch & 1ull) * c.world * c.p2p_cap;
}

}  // namespace dr4g
