// CSR input without a host-side densify.
//
// The reference turns scipy CSR input into a dense array on the HOST (dimreduce4gpu/truncated_svd.py:67-70, pca.py:44-47)
// and then uploads n*d values (its cuSPARSE handle, src/device_context.cuh:14, is never used).  Here only the CSR arrays
// cross PCIe (nnz values + nnz column indices + n+1 row pointers); a row chunk is expanded on the device straight into
// the row-major tile the streamed kernels read (K2 Gram via dr4g_partial_fit_*, K7 projection), so a matrix whose dense
// form does not fit in HBM is still fitted chunk by chunk.  With d <= a few thousand the dense tensor-core Gram of a
// chunk costs less than scattered atomics on a sparse outer product would (2 nnz_r^2 atomics per row).
#include "common.cuh"

namespace dr4g {
namespace {

// one warp per row: zero the row with 128-bit stores, then scatter its non-zeros (duplicates are summed, as
// scipy's toarray() does)
template <typename T>
__global__ void __launch_bounds__(256)
csr_to_dense_kernel(const long long* __restrict__ indptr, const int* __restrict__ indices, const T* __restrict__ data,
                    long long row0, long long rows, int d, T* __restrict__ out) {
  const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (r >= rows) return;
  T* o = out + r * (long long)d;
  for (int c = lane; c < d; c += 32) o[c] = (T)0;
  __syncwarp();
  const long long b = indptr[row0 + r], e = indptr[row0 + r + 1];
  for (long long i = b + lane; i < e; i += 32) {
    const int c = indices[i];
    if (c >= 0 && c < d) atomicAdd(&o[c], data[i]);
  }
}

}  // namespace

template <typename T>
void launch_csr_to_dense(Ctx& c, const long long* indptr, const int* indices, const T* data, long long row0, long long rows,
                         int d, T* out) {
  if (rows <= 0) return;
  csr_to_dense_kernel<T><<<(unsigned)((rows + 7) / 8), 256, 0, c.stream>>>(indptr, indices, data, row0, rows, d, out);
  c.launches++;
  DR4G_CUDA(cudaGetLastError());
}

template void launch_csr_to_dense<float>(Ctx&, const long long*, const int*, const float*, long long, long long, int, float*);
template void launch_csr_to_dense<double>(Ctx&, const long long*, const int*, const double*, long long, long long, int, double*);

}  // namespace dr4g
