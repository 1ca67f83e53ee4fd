/*
 * deepfm_b200.h -- C ABI of libdfm_cuda.so, the sm_100a kernel library of the
 * DeepFM framework.  Every entry point
 *   - takes raw device pointers (the Python layer passes tensor.data_ptr()),
 *   - enqueues work on the given stream and returns without synchronising,
 *   - returns DFM_OK or a negative DFM_ERR_* code; dfm_last_error() gives the
 *     text for the calling host thread.
 *
 * Citations of the form HVD.py:NNN point at
 * /root/reference/2-hvd-gpu/DeepFM-hvd-tfrecord-vectorized-map.py, the TF1
 * script whose graph these kernels replace.
 */
#ifndef DEEPFM_B200_H_
#define DEEPFM_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DFM_OK 0
#define DFM_ERR_ARG (-1)
#define DFM_ERR_CUDA (-2)
#define DFM_ERR_UNSUPPORTED (-3)

typedef void* dfm_stream_t; /* cudaStream_t */

const char* dfm_last_error(void);
int dfm_version(void);

/* ------------------------------------------------------------------------
 * Generic bf16 GEMM on tcgen05/TMEM fed by TMA (building block + unit-test
 * surface for the fused MLP kernels).
 *   C[M,N] (fp32, row-major, leading dim ldc) = A[M,K] * B[N,K]^T (+ bias[N]) (relu)
 * A and B are bf16, K contiguous ("K-major"); lda/ldb in elements, multiples of 8.
 * b_mn_major != 0: B is given as [K,N] with N contiguous (ldb = its leading dim).
 * a_mn_major != 0: A is given as [K,M] with M contiguous.
 * Replaces the cuBLAS SGEMM behind tf.contrib.layers.fully_connected, HVD.py:224.
 * ---------------------------------------------------------------------- */
int dfm_gemm_bf16(const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major,
                  float* C, int ldc, int M, int N, int K, const float* bias, int relu,
                  dfm_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DEEPFM_B200_H_ */
