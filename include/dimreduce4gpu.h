/* dimreduce4gpu.h - C ABI of libdimreduce4gpu.so (B200 / sm_100a build).
 *
 * This is the drop-in boundary for the PCA / TruncatedSVD fit+transform hot path.  Everything is
 * `extern "C"`, plain pointers and sizes; no C++/torch types cross it.
 *
 * Part 1 ("legacy symbols") is what the reference's ctypes layer binds today and is preserved
 * bit-for-bit:
 *     params                  reference include/tsvd.h:24-35, include/pca.h:24-35,
 *                             dimreduce4gpu/lib_dimreduce4gpu.py:6-18 (ctypes mirror)
 *     truncated_svd_float     reference include/tsvd.h:48   (src/tsvd.cu:427-438)
 *     truncated_svd_double    reference include/tsvd.h:49   (src/tsvd.cu:440-452)
 *     pca_float               reference include/pca.h:48    (src/pca.cu:22-62)
 *     pca_double              reference include/pca.h:49    (src/pca.cu:76-116)
 *   The two *_float symbols are the reference's CI contract (ci/expected_symbols.txt:4-5).
 *   Layout convention: the one the reference's Python wrapper and CPU backend use
 *   (dimreduce4gpu/truncated_svd.py:72-86, src/cpu_backend.cpp:340-345): X is row-major
 *   X_n (samples) x X_m (features) in HOST memory; Q is k x X_m, U and X_transformed are X_n x k.
 *   Output conventions are the CPU backend's (src/cpu_backend.cpp:277-332), see DESIGN.md.
 *
 * Part 2 ("dr4g_* symbols") is additive: device-resident, row-sharded operation that the
 * reference does not have (it is single-device, src/tsvd.cu:429).  The legacy symbols are thin
 * wrappers over it (H2D -> fit -> transform -> D2H).
 *
 * Error convention: the legacy symbols return void like the reference; on failure they print one
 * line to stderr and leave the caller's (pre-zeroed) outputs untouched, never exit() (the reference
 * calls exit(-1), src/utils.cuh:36-50).  dr4g_last_error() returns the thread's last message.
 * dr4g_* functions return 0 on success, non-zero on failure.
 */
#ifndef DIMREDUCE4GPU_H
#define DIMREDUCE4GPU_H

#include <stdbool.h>
#include <stdint.h>

#if defined(__GNUC__)
#define DR4G_EXPORT __attribute__((visibility("default")))
#else
#define DR4G_EXPORT
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* 48 bytes, align 8: X_n@0 X_m@4 k@8 algorithm@16 n_iter@24 random_state@28 tol@32 verbose@36
 * gpu_id@40 whiten@44.  Passed BY VALUE.  `algorithm` is borrowed for the call only. */
typedef struct params {
  int X_n;          /* samples (rows)     */
  int X_m;          /* features (columns) */
  int k;            /* components         */
  char* algorithm;  /* "cusolver" -> exact eigensolve; anything else -> "power" (src/tsvd.cu:457-471) */
  int n_iter;
  int random_state;
  float tol;
  int verbose;
  int gpu_id;
  bool whiten;      /* accepted, ignored: the CPU backend (the parity oracle) ignores it */
} params;

/* ---------------------------------------------------------------- Part 1: legacy symbols */
DR4G_EXPORT void truncated_svd_float(const float* X, float* Q, float* w, float* U, float* X_transformed,
                         float* explained_variance, float* explained_variance_ratio, params p);
DR4G_EXPORT void truncated_svd_double(const double* X, double* Q, double* w, double* U, double* X_transformed,
                          double* explained_variance, double* explained_variance_ratio, params p);
DR4G_EXPORT void pca_float(const float* X, float* Q, float* w, float* U, float* X_transformed,
               float* explained_variance, float* explained_variance_ratio, float* mean, params p);
DR4G_EXPORT void pca_double(const double* X, double* Q, double* w, double* U, double* X_transformed,
                double* explained_variance, double* explained_variance_ratio, double* mean, params p);

/* ---------------------------------------------------------------- Part 2: dr4g_* symbols */
typedef struct dr4g_ctx dr4g_ctx;

/* Thread-local message of the last failure on this thread ("" if none). */
DR4G_EXPORT const char* dr4g_last_error(void);
/* ABI version of this header (bumped when a dr4g_* signature changes). */
DR4G_EXPORT int dr4g_abi_version(void);

/* Persistent per-device context: private stream, workspace arena, optional NCCL communicator.
 * Replaces the reference's per-call DeviceContext (src/device_context.cuh:8-30). */
DR4G_EXPORT dr4g_ctx* dr4g_ctx_create(int device);
DR4G_EXPORT void dr4g_ctx_destroy(dr4g_ctx* ctx);
/* Run this context's work on a caller-owned CUDA stream (a cudaStream_t; NULL -> private stream). */
DR4G_EXPORT int dr4g_ctx_set_stream(dr4g_ctx* ctx, void* cuda_stream);
/* Block until everything enqueued on the context's stream is done; returns deferred errors. */
DR4G_EXPORT int dr4g_sync(dr4g_ctx* ctx);
/* Tunables / test hooks.  Keys: "gram_path" (0 auto - fp32 inputs of <= 262144 elements take the generic kernel, narrow ones
 * are row-folded, see "fold" -, 1 generic DMMA kernel, 2 tcgen05 kernels, 3 tcgen05 single-CTA kernel only),
 * "power_min_dim" (min(n, d) <= this takes the exact path whatever params.algorithm says, default 256 as
 * src/cpu_backend.cpp:347; 0 = off),
 * "proj_path" (projection: 0 auto = tensor cores where eligible except fp32 inputs of <= 262144 elements, 1 CUDA-core kernels,
 * 2 tensor cores wherever eligible),
 * "fold" (1 = default: inputs with d <= 128 and n d >= 2^24 (fp32) / 2^23 (fp64) are streamed f rows at a time as rows of
 * f d <= 256 columns; 0 = off),
 * "gram_split" (CTA-pair kernel operands: 0 auto = 2 bf16x3 on kind::f16, 1 = 3xTF32 on kind::tf32),
 * "flush_rows" (rows between TMEM accumulator flushes), "pilot_rows", "jacobi_tol", "jacobi_sweeps",
 * "eig_solver" (exact path: 0 tridiagonal route for the top k, 1 blocked Jacobi), "timing". */
DR4G_EXPORT int dr4g_set_option(dr4g_ctx* ctx, const char* key, double value);

/* Row-sharded operation: rank 0 obtains a 128-byte NCCL unique id, the host layer ships it to
 * every rank, each rank calls dr4g_comm_init.  After that dr4g_fit_* exchanges the packed
 * [Gram | column sums | column sums of squares | row count | pilot shift] partials ONCE per fit: PCA all-gathers them and
 * every rank re-bases them to the common shift and sums them in rank order (bit-identical models, no broadcast);
 * TruncatedSVD (no shift) all-reduces them. */
DR4G_EXPORT int dr4g_nccl_unique_id(void* out128);
DR4G_EXPORT int dr4g_comm_init(dr4g_ctx* ctx, const void* unique_id128, int rank, int world);
DR4G_EXPORT int dr4g_comm_world(dr4g_ctx* ctx);
/* Exchange over NVLink peer memory instead of NCCL (one NVSwitch box, one process per GPU; csrc/p2p.cu): after dr4g_comm_init
 * every rank calls dr4g_p2p_prepare (allocates its exchange buffer for payloads of up to capacity_doubles = d^2 + 3 d + 1 doubles
 * and returns its 64-byte cudaIpcMemHandle), the host layer all-gathers the handles, every rank calls dr4g_p2p_connect with the
 * world x 64 bytes in rank order.  From then on the centred (PCA) fit exchanges its partials with ONE kernel that pushes them
 * into every peer's buffer, releases a flag and waits for the peers' flags; larger payloads, the uncentred fit and
 * option "exchange" = 1 keep NCCL.  dr4g_last_exchange_path: 0 none, 1 NCCL, 2 peer memory. */
DR4G_EXPORT int dr4g_p2p_prepare(dr4g_ctx* ctx, long long capacity_doubles, void* out_handle64);
DR4G_EXPORT int dr4g_p2p_connect(dr4g_ctx* ctx, const void* handles);
DR4G_EXPORT int dr4g_last_exchange_path(dr4g_ctx* ctx);

/* fit on a device-resident row shard.  dX: n_local x d row-major on ctx's device.  center=1 ->
 * PCA (covariance of centred data), 0 -> TruncatedSVD.  p supplies algorithm / n_iter /
 * random_state / tol / verbose (X_n, X_m, k, gpu_id in p are ignored here).  Outputs are device
 * pointers: dQ k x d (components_), dw k (singular_values_), dev k, devr k, dmean d (may be NULL
 * when center=0).  Asynchronous on ctx's stream. */
DR4G_EXPORT int dr4g_fit_float(dr4g_ctx* ctx, const float* dX, long long n_local, int d, int k, int center,
                   params p, float* dQ, float* dw, float* dev, float* devr, float* dmean);
DR4G_EXPORT int dr4g_fit_double(dr4g_ctx* ctx, const double* dX, long long n_local, int d, int k, int center,
                    params p, double* dQ, double* dw, double* dev, double* devr, double* dmean);

/* Host-resident row shard in, host results out, on THIS context: what the legacy symbols do behind
 * truncated_svd.py:98-107 / pca.py:73-86 (H2D of X, fit, projection, D2H of the scores), but on a caller-owned context
 * so that a communicator set with dr4g_comm_init is honoured: X is then this rank's shard (n_local rows) of a
 * row-sharded matrix and Q, w, ev, evr, mean are the global model; U / Xt (n_local x k, each may be NULL) are the
 * shard's scores.  All pointers are host memory (pinned or pageable).  Blocking.  Out of core when the shard does not
 * fit in free device memory (second H2D sweep for the scores). */
DR4G_EXPORT int dr4g_fit_transform_host_float(dr4g_ctx* ctx, const float* X, long long n_local, int d, int k, int center,
                                  params p, float* Q, float* w, float* U, float* Xt, float* ev, float* evr, float* mean);
DR4G_EXPORT int dr4g_fit_transform_host_double(dr4g_ctx* ctx, const double* X, long long n_local, int d, int k, int center,
                                   params p, double* Q, double* w, double* U, double* Xt, double* ev, double* evr,
                                   double* mean);

/* Streaming (out-of-core) fit, SURVEY 8f-1: the packed accumulators [G | s | q | n] of the fit are summed
 * over row chunks, so X never has to be device-resident as a whole (the reference's host-pointer symbols,
 * truncated_svd.py:98-107 / src/tsvd.cu:427-452, copy all of X to the device first).
 *   dr4g_partial_fit_*          one pass over a device-resident chunk (n x d); the first chunk fixes d, center,
 *                               the element type and, for center=1, this rank's pilot shift (no collective).
 *   dr4g_partial_fit_finalize_* exchange + eigensolve + outputs, exactly as dr4g_fit_* after its pass
 *                               (one collective on the packed accumulators when a communicator is set);
 *                               the accumulated state stays valid (more chunks may follow).
 *   dr4g_partial_fit_reset      drop the accumulated state;  dr4g_partial_fit_rows: rows accumulated so far.
 * Asynchronous on ctx's stream: a chunk buffer may be reused once work queued after the call has completed. */
DR4G_EXPORT int dr4g_partial_fit_float(dr4g_ctx* ctx, const float* dX, long long n, int d, int center);
DR4G_EXPORT int dr4g_partial_fit_double(dr4g_ctx* ctx, const double* dX, long long n, int d, int center);
DR4G_EXPORT int dr4g_partial_fit_finalize_float(dr4g_ctx* ctx, int k, params p, float* dQ, float* dw, float* dev,
                                    float* devr, float* dmean);
DR4G_EXPORT int dr4g_partial_fit_finalize_double(dr4g_ctx* ctx, int k, params p, double* dQ, double* dw, double* dev,
                                     double* devr, double* dmean);
DR4G_EXPORT int dr4g_partial_fit_reset(dr4g_ctx* ctx);
DR4G_EXPORT long long dr4g_partial_fit_rows(dr4g_ctx* ctx);

/* transform on a device-resident row shard (no communication):
 *   dZ[n x k] = (dX - dmean) * dQ^T   (dmean NULL -> no centring, the reference's transform,
 *                                      dimreduce4gpu/truncated_svd.py:116-118)
 *   dU[n x k] = dZ / dw               (only if dU and dw are non-NULL; 0 where dw == 0,
 *                                      src/tsvd.cu:171-175) */
DR4G_EXPORT int dr4g_transform_float(dr4g_ctx* ctx, const float* dX, long long n, int d, int k, const float* dQ,
                         const float* dmean, const float* dw, float* dZ, float* dU);
DR4G_EXPORT int dr4g_transform_double(dr4g_ctx* ctx, const double* dX, long long n, int d, int k, const double* dQ,
                          const double* dmean, const double* dw, double* dZ, double* dU);
/* inverse_transform: dX[n x d] = dZ * dQ (+ dmean).  Not in the reference. */
DR4G_EXPORT int dr4g_inverse_transform_float(dr4g_ctx* ctx, const float* dZ, long long n, int d, int k,
                                 const float* dQ, const float* dmean, float* dX);
DR4G_EXPORT int dr4g_inverse_transform_double(dr4g_ctx* ctx, const double* dZ, long long n, int d, int k,
                                  const double* dQ, const double* dmean, double* dX);

/* CSR input: expand rows [row0, row0 + rows) of a device-resident CSR matrix (dindptr: n+1 int64,
 * dindices: nnz int32, ddata: nnz values) into the dense row-major tile dout (rows x d) that dr4g_partial_fit_* and
 * dr4g_transform_* read.  Replaces the reference's host-side densify (dimreduce4gpu/truncated_svd.py:67-70,
 * pca.py:44-47): only the CSR arrays cross PCIe.  Asynchronous on ctx's stream. */
DR4G_EXPORT int dr4g_csr_to_dense_float(dr4g_ctx* ctx, const long long* dindptr, const int* dindices, const float* ddata,
                            long long row0, long long rows, int d, float* dout);
DR4G_EXPORT int dr4g_csr_to_dense_double(dr4g_ctx* ctx, const long long* dindptr, const int* dindices, const double* ddata,
                             long long row0, long long rows, int d, double* dout);

/* Building blocks, exported for parity tests and micro-benchmarks (device pointers).
 *   colstats: per-column sum and sum of squares of (x - shift) in double; shift may be NULL.
 *   gram:     dG[d x d] (double, row-major, full symmetric) = (X - shift)^T (X - shift).
 *   eigh:     symmetric eigen-decomposition by blocked cyclic Jacobi; dA (d x d) is destroyed,
 *             dW gets the d eigenvalues (unsorted), dV (d x d row-major) eigenvectors in columns. */
DR4G_EXPORT int dr4g_colstats_float(dr4g_ctx* ctx, const float* dX, long long n, int d, const float* dshift,
                        double* dsum, double* dsumsq);
DR4G_EXPORT int dr4g_colstats_double(dr4g_ctx* ctx, const double* dX, long long n, int d, const double* dshift,
                         double* dsum, double* dsumsq);
DR4G_EXPORT int dr4g_gram_float(dr4g_ctx* ctx, const float* dX, long long n, int d, const float* dshift, double* dG);
DR4G_EXPORT int dr4g_gram_double(dr4g_ctx* ctx, const double* dX, long long n, int d, const double* dshift, double* dG);
DR4G_EXPORT int dr4g_eigh_double(dr4g_ctx* ctx, double* dA, int d, double* dW, double* dV);
/*   eigh_topk: the k largest eigenpairs of the symmetric dA (d x d, destroyed) by the tridiagonal
 *             route the exact fit uses (replaces cusolverDn?syevd, src/matrix.cu:231-273): dW gets k
 *             eigenvalues in descending order, dVt (k x d, row-major) unit eigenvectors in rows. */
DR4G_EXPORT int dr4g_eigh_topk_double(dr4g_ctx* ctx, double* dA, int d, int k, double* dW, double* dVt);
/*   eigh_topk_static: dr4g_eigh_topk_double WITHOUT any allocation at call time.  No counterpart in the reference; added
 *             for callers that cannot call cudaMalloc (CUDA graph capture, tight
 *             serving loops, harnesses that forbid it).  The workspace (reflectors, tridiagonal, inverse-iteration
 *             vectors) lives in a file-scope device array of 1,088 KiB that every GPU loading the library holds until
 *             exit, so d <= dr4g_eigh_topk_static_max_d() = 256 (any 1 <= k <= d); larger d is rejected with an error
 *             (dr4g_last_error).  Same inputs, outputs and results as dr4g_eigh_topk_double; asynchronous on the
 *             context's stream.  All calls on a device share the array: calls that overlap on the GPU (two contexts /
 *             streams) race - serialise them.  Calls no cudaMalloc / cudaFree / allocating routine. */
DR4G_EXPORT int dr4g_eigh_topk_static_double(dr4g_ctx* ctx, double* dA, int d, int k, double* dW, double* dVt);
DR4G_EXPORT int dr4g_eigh_topk_static_max_d(void);

/* Per-stage device times (ms) of the last dr4g_fit_* on this context, valid after dr4g_sync:
 * [0] pilot shift, [1] Gram + column stats, [2] exchange between ranks + centring,
 * [3] eigensolve (Jacobi or subspace iteration), [4] sort/flip/outputs, [5] total.
 * Returns the number of entries written (<= n). */
DR4G_EXPORT int dr4g_stage_times(dr4g_ctx* ctx, double* ms, int n);
/* Which Gram kernel the last fit used (1 generic DMMA, 2 tcgen05 single-CTA, 3 tcgen05 CTA pairs 3xTF32, 4 tcgen05 CTA pairs bf16x3), and launches it issued. */
DR4G_EXPORT int dr4g_last_gram_path(dr4g_ctx* ctx);
/* Which projection kernel the last transform used: 1 CUDA cores (project.cu), 2 tcgen05 (project_tc.cu). */
DR4G_EXPORT int dr4g_last_proj_path(dr4g_ctx* ctx);
/* Row-fold factor of the last streamed pass (fit, or transform of an input with d <= 32): narrow fp32 inputs with n d >= 2^24 are
 * read as n/f rows of f d <= 256 columns - the fit (d <= 128) on the CTA-pair tcgen05 Gram kernel, the projection (d <= 32) against
 * the block-diagonal image of the components; option "fold" = 0 switches it off.  0: not folded. */
DR4G_EXPORT int dr4g_last_fold(dr4g_ctx* ctx);
DR4G_EXPORT long long dr4g_launch_count(dr4g_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* DIMREDUCE4GPU_H */
