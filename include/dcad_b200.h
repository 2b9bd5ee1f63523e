/* C ABI of libdcad_b200.so — the hand-written sm_100a kernels behind the B200-native federated DDoS-detection
 * framework.  Every entry point takes raw device pointers plus a CUDA stream handle (void* == cudaStream_t),
 * launches asynchronously on that stream and returns 0 on success, a positive cudaError_t, or a negative
 * DCAD_ERR_* code.  Unless an entry point says otherwise it allocates nothing: callers own all buffers.
 *
 * "reference" in the comments below is /root/reference (client1.py, client2.py, server.py); HF is the
 * transformers DistilBERT implementation the reference delegates its model body to.
 */
#ifndef DCAD_B200_H_
#define DCAD_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- error codes (negative; positive values are cudaError_t) ---- */
#define DCAD_ERR_BAD_SHAPE (-1)
#define DCAD_ERR_BAD_ALIGNMENT (-2)
#define DCAD_ERR_BAD_ARGUMENT (-3)
#define DCAD_ERR_UNSUPPORTED_TILE (-4)
#define DCAD_ERR_NO_DRIVER_ENTRY (-5)
#define DCAD_ERR_TENSOR_MAP (-6)
#define DCAD_ERR_TOO_LARGE (-7)
#define DCAD_ERR_TIMEOUT (-8)

/* ---- GEMM epilogues ---- */
#define DCAD_EPI_NONE 0      /* D = acc                                   */
#define DCAD_EPI_BIAS 1      /* D = acc + bias[n]                         */
#define DCAD_EPI_BIAS_GELU 2 /* pre = acc + bias; aux_out = pre; D = gelu(pre)   (HF FFN lin1 + GELU(erf)) */
#define DCAD_EPI_DGELU 3     /* D = acc * gelu'(aux[m,n])                 (backward of the above)          */

/* operand majors */
#define DCAD_MAJOR_K 0  /* operand stored [MN][K]: reduction dimension contiguous */
#define DCAD_MAJOR_MN 1 /* operand stored [K][MN]: M (or N) dimension contiguous  */

int dcad_version(void);

/* bf16 x bf16 -> fp32-accumulated GEMM on tcgen05 tensor cores (TMA-fed, TMEM accumulators, persistent).
 *   D[M,N] = epilogue( sum_k A(m,k) * B(n,k) )
 * A is [M][lda] when a_major == DCAD_MAJOR_K, [K][lda] when DCAD_MAJOR_MN; likewise B with N.
 * D is bf16 (out_fp32 == 0) or fp32 (out_fp32 == 1; accumulate != 0 adds into D).
 * block_n: 0 = auto, else one of 32/64/96/128/192/256 (MN-major B: 64/128/192/256).  max_ctas: 0 = #SMs.
 * Replaces the cuBLAS calls behind nn.Linear forward/backward in the reference (client1.py:61,109).
 */
int dcad_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K, int a_major, int b_major, int lda,
                   int ldb, int ldd, const float* bias, const void* aux, void* aux_out, int ldaux, int epi,
                   int out_fp32, int accumulate, int block_n, int max_ctas, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DCAD_B200_H_ */
