/* libvisitron_b200 -- C ABI of the B200-native VISITRON encoder hot path.
 *
 * The reference (alexa/visitron) is 100 % Python and has NO FFI for this path: every op below
 * replaces a stock PyTorch/ATen/cuBLAS call made from the reference modules.  Each entry point
 * cites the reference lines it replaces (paths relative to the reference root;
 * hf/ = transformers/pytorch_transformers/, vs/ = tasks/viewpoint_select/, tb/ = tasks/turn_based/).
 * INTEGRATION.md shows the ctypes binding a maintainer would add.
 *
 * Conventions: plain pointers (DEVICE memory unless stated) and sizes, no torch types; every call is
 * asynchronous on `stream`; returns 0 or a negative vt_status; vt_last_error() gives the
 * thread-local message.  bf16 = 16-bit brain float, row-major everywhere, `ld*` in elements.
 * There is no CPU fallback: every call fails with VT_ERR_NO_DEVICE off sm_100.
 */
#ifndef VISITRON_B200_H
#define VISITRON_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* vt_stream; /* cudaStream_t */

typedef enum {
  VT_OK = 0,
  VT_ERR_BAD_ARGUMENT = -1, /* shape / alignment / envelope violation */
  VT_ERR_CUDA = -2,         /* a CUDA runtime / driver call failed */
  VT_ERR_NO_DEVICE = -3,    /* not an sm_100 device */
  VT_ERR_UNSUPPORTED = -4
} vt_status;

const char* vt_last_error(void);
/* Checks compute capability 10.0, resolves cuTensorMapEncodeTiled, raises smem limits. */
int vt_init(int device);
int vt_num_sms(void);
/* Kernel launches issued by this library since the last reset (bench.py's gpu_launches). */
int64_t vt_launch_count(void);
void vt_reset_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * Dense projections: one tcgen05/TMEM GEMM fed by TMA.
 *   D[M,N] = epilogue( alpha * sum_k A(m,k) * B(n,k) )
 * a_major / b_major: 0 = the contraction dim k is contiguous in memory (A is [M,K], B is [N,K],
 * i.e. nn.Linear's weight [out,in]); 1 = the M / N dim is contiguous (A stored as [K,M], B stored
 * as [K,N]).  Forward uses (0,0); data-gradient uses (0,1) with B = the same weight; weight-
 * gradient uses (1,1) with A = dY [rows,out], B = X [rows,in] and k = rows.
 * Replaces F.linear / torch.matmul at: oscar/modeling_bert.py:43-45 (Q,K,V), hf/modeling_bert.py:342
 * (attention output), :388 (FFN up + gelu :136-142), :401 (FFN down), :465 (pooler), :481,:499
 * (MLM transform / tied decoder), vs/encoder.py:277-279 (img_embedding + location_embeds),
 * vs/encoder.py:324 (token head) and their autograd backward.
 * Epilogue order: acc*alpha -> +bias[n] -> (store aux_out) -> activation -> dropout -> +residual -> store.
 * ------------------------------------------------------------------------------------------ */
enum { VT_ACT_NONE = 0, VT_ACT_GELU = 1, VT_ACT_TANH = 2, VT_ACT_GELU_BWD = 3, VT_ACT_TANH_BWD = 4 };

typedef struct {
  int out_fp32;           /* 0: D is bf16, 1: D is fp32 */
  int accumulate;         /* 1: D += result with fp32 atomics (requires out_fp32; enables split_k) */
  int split_k;            /* 0 = choose (only splits when accumulate), else >= 1; > 1 needs accumulate */
  int act;                /* VT_ACT_* ; *_BWD multiply by the activation derivative evaluated at aux */
  float alpha;
  const float* alpha_dev; /* optional device scalar multiplied into alpha (e.g. upstream loss grad) */
  const float* bias;      /* fp32 [N] or NULL */
  const void* residual;   /* bf16 [*, ldr] or NULL, indexed with the remapped output row */
  int64_t ldr;
  const void* aux;        /* bf16 [M, ldaux]: GELU_BWD: pre-activation; TANH_BWD: the tanh output */
  int64_t ldaux;
  void* aux_out;          /* bf16 [M, ldaux_out] or NULL: pre-activation (after bias) */
  int64_t ldaux_out;
  float dropout_p;        /* 0 disables */
  uint64_t seed;
  uint32_t dropout_stream;
  /* output row remap: row m is written to (m / rows_in) * rows_out + row_off + m % rows_in
   * (rows_in == 0 disables).  Used to write image rows behind the text rows of each sample
   * (replaces torch.cat, vs/encoder.py:287) and to scatter the pooler gradient to row 0. */
  int64_t remap_rows_in, remap_rows_out, remap_row_off;
} vt_gemm_epilogue;

int vt_gemm_bf16(const void* A, int64_t lda, int a_major, const void* B, int64_t ldb, int b_major,
                 void* D, int64_t ldd, int64_t M, int64_t N, int64_t K, const vt_gemm_epilogue* ep,
                 vt_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* VISITRON_B200_H */
