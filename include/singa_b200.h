/* singa_b200 native library — C ABI.
 *
 * Every entry point launches sm_100a kernels on the CUDA stream passed as `stream` (a cudaStream_t
 * cast to void*; NULL = legacy default stream) and returns 0 on success or a non-zero code; the text
 * of the last failure on the calling thread is returned by sg_last_error().  Pointers are device
 * pointers unless stated otherwise.  "bf16" buffers hold __nv_bfloat16.  Leading dimensions (ld*) are in
 * elements.  Unless an entry point says otherwise it allocates no memory.
 *
 * Reference citations (path:line under the reference tree) name the code whose behaviour the entry
 * point provides; none of that code is reused.
 */
#ifndef SINGA_B200_H_
#define SINGA_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

const char* sg_last_error(void);
/* Library/ABI version and device capability probe (returns compute capability major*10+minor, or <0). */
int sg_version(void);
int sg_device_cc(int device);

/* ------------------------------------------------------------------------------------------------
 * tcgen05 GEMM  (reference: mshadow dot() -> cblas/cublas sgemm, include/mshadow/tensor_expr_engine-inl.hpp:271-357;
 *                used by InnerProduct src/worker/layer.cc:186-213 and Convolution layer.cc:63-123)
 *
 *   acc[m][n] = sum_k A(m,k) * B(n,k)          bf16 operands, fp32 accumulation in TMEM
 *   A: a_mn_major == 0 -> stored [M][K] (row stride lda), == 1 -> stored [K][M] (row stride lda)
 *   B: b_mn_major == 0 -> stored [N][K] (row stride ldb), == 1 -> stored [K][N] (row stride ldb)
 *   lda/ldb must be multiples of 8 elements and the base pointers 16-byte aligned (TMA).
 *
 * Epilogue, applied in registers after tcgen05.ld, in this order:
 *   v = acc (+ bias[n]); v = act(v); v *= act'(aux[m][n]); v += d32_old (accumulate); store fp32 and/or bf16.
 * ---------------------------------------------------------------------------------------------- */
enum { SG_ACT_NONE = 0, SG_ACT_RELU = 1, SG_ACT_STANH = 2 };

typedef struct SgGemmEpilogue {
  float* d32;            /* optional fp32 output [M][ldd32] */
  long long ldd32;
  void* d16;             /* optional bf16 output [M][ldd16] */
  long long ldd16;
  const float* bias;     /* optional [N] */
  int act;               /* SG_ACT_* applied to acc+bias */
  int grad_act;          /* SG_ACT_*: multiply by the activation derivative evaluated from aux (= activation OUTPUT) */
  const void* aux;       /* [M][ldaux], bf16 if aux_bf16 else fp32 */
  long long ldaux;
  int aux_bf16;
  int accumulate;        /* d32 += result instead of = */
  int perm_c;            /* >0: column j=(kk_idx*perm_c + c) is stored at column c*perm_kk + kk_idx (conv wgrad -> {F, C*k*k}) */
  int perm_kk;
} SgGemmEpilogue;

int sg_gemm_bf16(const void* A, int a_mn_major, long long lda, const void* B, int b_mn_major,
                 long long ldb, int M, int N, int K, const SgGemmEpilogue* ep, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SINGA_B200_H_ */
