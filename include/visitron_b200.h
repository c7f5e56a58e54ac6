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
 * Device-resident step state: what lets a whole training step run without a host <-> device sync and be captured in a
 * CUDA graph (the reference's loop syncs only at its .item() reads, vs/pretrain.py:195-201).
 *
 * Seeds.  Every op with dropout takes a 64-bit `seed` BY VALUE; a graph bakes it in at capture.  With a seed source set, the
 * kernels XOR the 64-bit word at `seed_dev` (device memory) into that seed when they run, so each replay draws fresh masks.
 * vt_seed_advance is the first node of a captured step: state[0] += 0x9E3779B97F4A7C15, state[1] = splitmix64(state[0]);
 * the seed source is &state[1].  The source is process-wide (forward runs on the caller's thread, backward on autograd's);
 * NULL (the default) restores by-value seeds.
 *
 * Row limit.  The pre-training heads run on the labelled rows only, and how many there are is known on the device only
 * (vt_compact_labels).  While a row limit is set ON THE CALLING THREAD, every launch below that walks activation rows --
 * vt_gemm_bf16 (M when A is [M,K]; the contraction K when A is [K,M], i.e. the weight-gradient form), vt_layernorm_fwd/bwd,
 * vt_colsum_bf16, vt_gather_rows, vt_scatter_add_rows, vt_gelu_bwd, vt_softmax_ce_fwd/bwd -- is sized on the host for the
 * caller's bound (the capacity) and clips itself on the device to rows_dev[0] (the count rounded up to 128; padding rows are
 * zeros so that they add nothing to a contraction); the cross-entropy kernels use rows_dev[1] (the true count) for the rows
 * that exist and for the mean.  NULL clears the limit.
 * ------------------------------------------------------------------------------------------ */
void vt_set_seed_source(const uint64_t* seed_dev);
int vt_seed_advance(uint64_t* state /* device uint64[2] */, vt_stream stream);
void vt_set_row_limit(const int32_t* rows_dev /* device int32[2] = {rows padded to 128, rows} */);
/* Keep-bit store of the attention-probability dropout (oscar/modeling_bert.py:64 `self.dropout(attention_probs)`; autograd keeps
 * the mask of that nn.Dropout between forward and backward).  While a store is set ON THE CALLING THREAD, vt_attention_fwd*
 * (dropout_p > 0) also writes the keep decisions it draws -- uint32 [B * heads][4 * ceil(S / 128)][S]: bit c of word (bh, g, i)
 * = key 32 g + c of query row i kept -- and vt_attention_bwd* reads them instead of running the counter generator again.  Same
 * mask either way (the bits ARE the generator's output); NULL (the default) regenerates.  vt_attention_keep_bits_words = words. */
void vt_set_attention_keep_bits(uint32_t* bits);
int64_t vt_attention_keep_bits_words(int64_t B, int64_t S, int64_t heads);
/* Item schedule of the persistent attention backward on packed rows: with samples of different lengths a (key tile, head, sample)
 * item costs its sample's number of query tiles, and round-robin assignment leaves the busiest CTA ~30 % over the mean.
 * vt_attention_bwd_schedule writes int32 sched[ctas + 1], ctas = vt_attention_bwd_schedule_ctas(B, S, heads): CTA i works on the
 * items [sched[i], sched[i + 1]) -- contiguous ranges of equal total cost -- from the device-resident sample starts (no host sync).
 * While set ON THE CALLING THREAD, vt_attention_bwd_packed follows it (ignored if its grid differs from `ctas`); NULL: round-robin. */
int64_t vt_attention_bwd_schedule_ctas(int64_t B, int64_t S, int64_t heads);
int vt_attention_bwd_schedule(const int32_t* start, int64_t B, int64_t S, int64_t heads, int32_t* sched, vt_stream stream);
void vt_set_attention_bwd_schedule(const int32_t* sched, int64_t ctas);
/* Compaction of the labelled rows (replaces torch.nonzero(labels != -1) + labels[idx] in front of the MLM / token heads,
 * vs/encoder.py:387-389,381-385, whose data-dependent sizes cost a sync per step).  For each of the two label arrays
 * (int64 [n]; either may be NULL): idx[r] / lab[r] = position / label of the r-th label != ignore_index in ascending order
 * (torch.nonzero's order), padding rows up to the next multiple of 128 get idx = -1, lab = 0;
 * counts[4 a + {0,1,2}] = {padded count, count, labels outside [0, classes) or beyond capacity}.  capacity: rows of idx / lab. */
int vt_compact_labels(const int64_t* labels_a, int64_t classes_a, const int64_t* labels_b, int64_t classes_b, int64_t n,
                      int64_t ignore_index, int64_t capacity, int64_t* idx_a, int64_t* lab_a, int64_t* idx_b, int64_t* lab_b,
                      int32_t* counts /* device int32[8] */, vt_stream stream);

/* Packed rows (padding-free execution of the pre-training step).  The loader pads every sample to 511 + 256 positions
 * (tasks/viewpoint_select/data_loader_pretrain.py:636-712); padded positions are masked as keys and carry no label, so neither the
 * loss nor any gradient depends on them.  vt_pack_plan keeps position (b, s) iff its mask is non-zero (bias != -10000), it carries a
 * label in either array, or s == 0, and lays the kept rows out sample after sample:
 *   row_src [capacity] padded index of packed row r (-1 on the layout's own padding rows),  row_dst [B*S] the inverse (-1: dropped),
 *   row_b [capacity] the sample of a packed row,  start [B+1] first packed row of every sample,  cls_rows [B] = start as int64,
 *   counts {rows kept rounded up to 128 plus 128, rows kept}: the vt_set_row_limit block of every launch of the step,
 *   bias_packed / labels_*_packed [capacity]: the key bias and the labels in packed order (-1 on padding rows).
 * capacity: rows of every activation buffer, a multiple of 128, >= roundup(B*S, 128) + 128.  The *_packed entry points below take
 * these maps; everything else runs under the row limit.  vt_zero_pad_rows clears rows [counts[1], counts[0]) of a buffer (the rows a
 * per-sample kernel -- attention -- does not write, and which must be finite zeros where a GEMM contracts over rows). */
int vt_pack_plan(const float* bias, const int64_t* labels_a, const int64_t* labels_b, int64_t B, int64_t S, int64_t capacity,
                 int32_t* row_src, int32_t* row_dst, int32_t* row_b, int32_t* start, int32_t* counts, float* bias_packed,
                 int64_t* labels_a_packed, int64_t* labels_b_packed, int64_t* cls_rows, vt_stream stream);
int vt_zero_pad_rows(void* buf, int64_t pitch_bytes, int64_t row_bytes, const int32_t* counts, vt_stream stream);

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
enum { VT_ACT_NONE = 0, VT_ACT_GELU = 1, VT_ACT_TANH = 2, VT_ACT_GELU_BWD = 3, VT_ACT_TANH_BWD = 4,
       /* the encoder's FFN pair: the forward GEMM writes gelu(x) to D and gelu'(x) -- evaluated on the fp32 accumulator -- to aux_out
        * (instead of x), so that the backward GEMM's epilogue is one multiply by aux (VT_ACT_MUL_AUX) instead of re-evaluating gelu' */
       VT_ACT_GELU_SAVE_GRAD = 5, VT_ACT_MUL_AUX = 6 };

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
  /* optional second map (packed rows, vt_pack_plan): the row computed above is looked up in remap_index (int32); a negative entry
   * drops the row.  With remap_rows_in == 0 the logical row m itself is looked up. */
  const int32_t* remap_index;
  /* optional, a_major == 1 only (A is [K, M]: the weight-gradient form dW = dY^T X): colsum[m] += alpha * sum_k A[k, m], fp32,
   * i.e. the bias gradient of the same layer (torch autograd's sum over rows of dY for F.linear, oscar/modeling_bert.py:43-45)
   * from the dY tiles the GEMM has in shared memory anyway.  Always accumulates (split-K safe); needs block N <= 192. */
  float* colsum;
} vt_gemm_epilogue;

/* Scheduling inside (no effect on the interface; results are bit-identical for the [M, K] forms): launches with at least 4096 rows of a
 * K-major A, and weight-gradient launches (a_major = b_major = 1, accumulate, no colsum) over at least 4096 rows, run on CTA pairs
 * (tcgen05.mma.cta_group::2, one 256 x 256 tile per two SMs); everything else on one CTA per 128 x BN tile. */
int vt_gemm_bf16(const void* A, int64_t lda, int a_major, const void* B, int64_t ldb, int b_major,
                 void* D, int64_t ldd, int64_t M, int64_t N, int64_t K, const vt_gemm_epilogue* ep,
                 vt_stream stream);

/* Debug hook for tests: force the GEMM tile width (0 = heuristic; 64/128/192/256). */
void vt_debug_force_block_n(int bn);
/* CTA pairs (tcgen05.mma.cta_group::2, two SMs per 256 x 256 tile) in vt_gemm_bf16: -1 = where they pay (default), 0 = never,
 * 1 = wherever supported (tests / A-B timing). */
void vt_debug_pair_mode(int mode);
/* The persistent GEMM launches one CTA per SM.  While a gradient all-reduce runs beside the backward GEMMs its CTAs hold
 * a few SMs; a full-width GEMM grid would then run its last CTAs as a second wave.  n > 0 makes the GEMM leave n SMs free
 * (visitron_b200/parallel.py sets it for the span of the overlapped exchange); 0 restores the full grid. */
void vt_reserve_sms(int n);

/* ------------------------------------------------------------------------------------------
 * Row kernels (HBM-bound, one warp per 768-wide row, 16-byte vectors).  hidden must be 768.
 * Dropout masks are never stored: they are regenerated from (seed, dropout_stream, element index);
 * p is quantised to 1/65536.  `remap_*` as in vt_gemm_epilogue.
 * ------------------------------------------------------------------------------------------ */

/* y[remap(m)] = dropout(BertLayerNorm(x[m]))   -- hf/modeling_bert.py:227-240 (TF style, biased
 * variance, eps inside the sqrt); x already holds dense-output + residual (hf/modeling_bert.py:344,403).
 * With remap: the optional image-embedding LayerNorm + dropout + torch.cat (vs/encoder.py:280-287). */
int vt_layernorm_fwd(const void* x, int64_t ldx, const float* gamma, const float* beta, float eps, int64_t rows,
                     int64_t hidden, void* y, int64_t ldy, float dropout_p, uint64_t seed, uint32_t dropout_stream,
                     int64_t remap_rows_in, int64_t remap_rows_out, int64_t remap_row_off, vt_stream stream);

/* Rows of fp32 workspace [vt_row_partials_rows(rows)][4][768] the LayerNorm / embedding backward need. */
int64_t vt_row_partials_rows(int64_t rows);

/* Backward of the site  x = dropout_out(h) + res ; y = LN(x)  (autograd of hf/modeling_bert.py:341-345,
 * 400-404):  dx = LN'(dy) [bf16, the residual branch], dx_dropped = dx * mask_out [bf16, the operand
 * of the dense layer's data/weight gradient GEMMs; may be NULL when out_dropout_p == 0],
 * dgamma/dbeta/dbias (+)= column sums (dbias = colsum(dx_dropped) is the dense bias gradient).
 * in_dropout_*: mask applied to dy first (sites where dropout FOLLOWS the LayerNorm, vs/encoder.py:280-284);
 * dy is then read at the remapped row. */
int vt_layernorm_bwd(const void* dy, int64_t lddy, const void* x, int64_t ldx, const float* gamma, float eps, int64_t rows,
                     int64_t hidden, void* dx, int64_t lddx, void* dx_dropped, int64_t lddxd, float* partials,
                     float* dgamma, float* dbeta, float* dbias, float in_dropout_p, uint32_t in_dropout_stream,
                     float out_dropout_p, uint32_t out_dropout_stream, uint64_t seed, int64_t remap_rows_in,
                     int64_t remap_rows_out, int64_t remap_row_off, vt_stream stream);

/* BertEmbeddings.forward (hf/modeling_bert.py:256-271): y[s*out_rows_per_sample + t] =
 * dropout(LN(word[ids] + pos[pos_ids or t] + type[type_ids or 0])); ids int64 [batch, T]. */
int vt_embed_ln_fwd(const int64_t* ids, const int64_t* type_ids, const int64_t* pos_ids, const float* word,
                    const float* pos, const float* type, const float* gamma, const float* beta, float eps,
                    int64_t batch, int64_t T, int64_t out_rows_per_sample, int64_t hidden, void* y, float dropout_p,
                    uint64_t seed, uint32_t dropout_stream, vt_stream stream);
/* the same into packed rows: position (s, t) is written to row row_dst[s * out_rows_per_sample + t], or skipped when negative */
int vt_embed_ln_fwd_packed(const int64_t* ids, const int64_t* type_ids, const int64_t* pos_ids, const float* word,
                           const float* pos, const float* type, const float* gamma, const float* beta, float eps,
                           int64_t batch, int64_t T, int64_t out_rows_per_sample, int64_t hidden, void* y, float dropout_p,
                           uint64_t seed, uint32_t dropout_stream, const int32_t* row_dst, vt_stream stream);
/* Its backward: scatter-adds (fp32 atomics) into dword/dpos/dtype, accumulates dgamma/dbeta. */
int vt_embed_ln_bwd(const void* dy, const int64_t* ids, const int64_t* type_ids, const int64_t* pos_ids, const float* word,
                    const float* pos, const float* type, const float* gamma, float eps, int64_t batch, int64_t T,
                    int64_t out_rows_per_sample, int64_t hidden, int64_t type_vocab, float* dword, float* dpos,
                    float* dtype, float* dgamma, float* dbeta, float* partials, float dropout_p, uint64_t seed,
                    uint32_t dropout_stream, vt_stream stream);

int vt_embed_ln_bwd_packed(const void* dy, const int64_t* ids, const int64_t* type_ids, const int64_t* pos_ids, const float* word,
                           const float* pos, const float* type, const float* gamma, float eps, int64_t batch, int64_t T,
                           int64_t out_rows_per_sample, int64_t hidden, int64_t type_vocab, float* dword, float* dpos,
                           float* dtype, float* dgamma, float* dbeta, float* partials, float dropout_p, uint64_t seed,
                           uint32_t dropout_stream, const int32_t* row_dst, vt_stream stream);

/* out[c] (+)= sum_r x[r, c]  for bf16 x [rows, ncols] -- bias gradients of the QKV and FFN-up layers.
 * partials: fp32 [vt_colsum_partials_rows(rows)][roundup8(ncols)]. */
int64_t vt_colsum_partials_rows(int64_t rows);
int vt_colsum_bf16(const void* x, int64_t ldx, int64_t rows, int64_t ncols, float* partials, float* out, int accumulate,
                   vt_stream stream);
/* The same without a caller workspace and without any allocation (no cudaMalloc / cudaFree / allocating library call), for callers
 * that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid it.  No counterpart in the
 * reference.  The partial sums go to a static device array inside the library
 * (4.6 MiB on every GPU that loads it, until exit): ncols (rounded up to 8) must not exceed vt_colsum_bf16_static_max_cols() = 4096,
 * larger calls are rejected with VT_ERR_BAD_ARGUMENT (vt_last_error says why).  Every call shares that array: calls that may overlap on the GPU (two streams with no
 * ordering between them) race; calls on one stream, or ordered by events, are safe.  Same result as vt_colsum_bf16. */
int64_t vt_colsum_bf16_static_max_cols(void);
int vt_colsum_bf16_static(const void* x, int64_t ldx, int64_t rows, int64_t ncols, float* out, int accumulate, vt_stream stream);

/* fp32 [rows, cols] -> bf16 copy (dst, may be NULL) and/or bf16 transpose (dst_t, may be NULL):
 * the per-step cast of the fp32 master weights into tensor-core operands. */
int vt_cast_f32_bf16(const float* src, int64_t lds, int64_t rows, int64_t cols, void* dst, int64_t ldd, void* dst_t,
                     int64_t ldt, vt_stream stream);
/* out[r] = bf16([a[r, 0:d0] | b[r, 0:d1] | 0...]) : img_feats and img_location_embeddings become ONE
 * GEMM operand so that img_embedding(x) + location_embeds(l) (vs/encoder.py:277-279) is one contraction. */
int vt_pad_cast_concat(const float* a, int64_t d0, const float* b, int64_t d1, int64_t rows, void* out, int64_t ldo,
                       vt_stream stream);
/* the same with either source in bf16 (a_bf16 / b_bf16 = 1): region features kept in a bf16 cache skip the fp32 detour */
int vt_pad_cast_concat_mixed(const void* a, int a_bf16, int64_t d0, const void* b, int b_bf16, int64_t d1, int64_t rows, void* out,
                             int64_t ldo, vt_stream stream);
/* y[m] = x[remap(m)] * dropout_mask(m) */
int vt_dropout_rows(const void* x, int64_t ldx, int64_t rows, int64_t hidden, void* y, int64_t ldy, float dropout_p,
                    uint64_t seed, uint32_t dropout_stream, int64_t remap_rows_in, int64_t remap_rows_out,
                    int64_t remap_row_off, vt_stream stream);
/* y[m] = x[row_dst[remap(m)]] * dropout_mask(m), zeros where that packed row does not exist */
int vt_dropout_rows_packed(const void* x, int64_t ldx, int64_t rows, int64_t hidden, void* y, int64_t ldy, float dropout_p,
                           uint64_t seed, uint32_t dropout_stream, int64_t remap_rows_in, int64_t remap_rows_out,
                           int64_t remap_row_off, const int32_t* row_dst, vt_stream stream);
/* dst[r] = src[idx[r]] ; dst[idx[r]] += src[r]   (compaction of the labelled rows for the MLM / token heads);
 * idx[r] < 0 (a padding row of vt_compact_labels): gathered as zeros, skipped by the scatter */
int vt_gather_rows(const void* src, int64_t lds, const int64_t* idx, int64_t n, int64_t hidden, void* dst, int64_t ldd,
                   vt_stream stream);
int vt_scatter_add_rows(const void* src, int64_t lds, const int64_t* idx, int64_t n, int64_t hidden, void* dst, int64_t ldd,
                        vt_stream stream);

/* ------------------------------------------------------------------------------------------
 * Fused masked-softmax self-attention (oscar/modeling_bert.py:52-72), head dim 64.
 * qkv: bf16 [B*S, ld] = [Q | K | V] (each heads*64 wide) as written by the fused QKV GEMM.
 * bias: fp32 [B, S] additive key bias from vt_mask_bias.  ctx: bf16 [B*S, heads*64].
 * lse: fp32 [B, heads, S, 2] = (row max, 1 / sum exp) saved for backward.
 * dsum_ws: fp32 scratch of vt_attention_bwd_workspace(B, S, heads) floats, 16-byte aligned: rowsum(dO.O) [B, heads, S], then the
 * per-key-tile bf16 dQ partials [ceil(S/128)][B*S][heads*64] that the backward adds in a fixed order (deterministic, no atomics).
 * ------------------------------------------------------------------------------------------ */
/* bias = (1 - mask) * -10000 in fp32 for any mask value (vs/encoder.py:226-241).
 * mask_dtype: 0 int64, 1 uint8/bool, 2 int32, 3 fp32. */
int vt_mask_bias(const void* mask, int mask_dtype, int64_t n, float* bias, vt_stream stream);
/* Optional plan of the key tiles worth visiting: key_tiles[b] bit t = the 64-key tile t of sample b holds a key whose bias is within
 * 5000 of the sample's largest (an unmasked key).  The probabilities of the other tiles are exactly 0 in the reference's fp32 softmax
 * (exp(-10000 + ...) underflows), so the kernels skip them without changing a bit of ctx or of any gradient.  S <= 2048.
 * Pass NULL as key_tiles to visit every tile.  vt_attention_fwd (64-key tiles) uses the plan; vt_attention_bwd accepts it and visits
 * every one of its 128-key tiles (at the loader's padding they are almost never dead as a whole; measured, no gain). */
int vt_attention_key_tiles(const float* bias, int64_t B, int64_t S, uint32_t* key_tiles, vt_stream stream);
int vt_attention_fwd(const void* qkv, int64_t ld, const float* bias, const uint32_t* key_tiles, int64_t B, int64_t S, int64_t heads,
                     int64_t head_dim, void* ctx, float* lse, float dropout_p, uint64_t seed, uint32_t dropout_stream, vt_stream stream);
/* Packed rows (vt_pack_plan): sample b owns rows [start[b], start[b+1]) of qkv / ctx / dctx / dqkv (rows_cap rows each) and of the
 * packed bias; S stays the bound of a sample's length and the stride of lse / dsum ([B, heads, S, .]).  Query and key tiles past a
 * sample's end are not visited; row_b [rows_cap] gives the sample of a row (-1: padding row, its dQ is written as 0). */
int vt_attention_key_tiles_packed(const float* bias_packed, const int32_t* start, int64_t B, int64_t S, uint32_t* key_tiles, vt_stream stream);
int vt_attention_fwd_packed(const void* qkv, int64_t ld, const float* bias_packed, const uint32_t* key_tiles, const int32_t* start,
                            int64_t rows_cap, int64_t B, int64_t S, int64_t heads, int64_t head_dim, void* ctx, float* lse, float dropout_p,
                            uint64_t seed, uint32_t dropout_stream, vt_stream stream);
int64_t vt_attention_bwd_workspace_packed(int64_t rows_cap, int64_t B, int64_t S, int64_t heads);
int vt_attention_bwd_packed(const void* qkv, int64_t ld, const float* bias_packed, const uint32_t* key_tiles, const int32_t* start,
                            const int32_t* row_b, int64_t rows_cap, const void* ctx, const void* dctx, const float* lse, float* dsum_ws,
                            int64_t B, int64_t S, int64_t heads, int64_t head_dim, void* dqkv, float dropout_p, uint64_t seed,
                            uint32_t dropout_stream, vt_stream stream);
int64_t vt_attention_bwd_workspace(int64_t B, int64_t S, int64_t heads);
int vt_attention_bwd(const void* qkv, int64_t ld, const float* bias, const uint32_t* key_tiles, const void* ctx, const void* dctx,
                     const float* lse, float* dsum_ws, int64_t B, int64_t S, int64_t heads, int64_t head_dim, void* dqkv,
                     float dropout_p, uint64_t seed, uint32_t dropout_stream, vt_stream stream);

/* ------------------------------------------------------------------------------------------
 * Pre-training heads (vs/encoder.py:377-441).
 * ------------------------------------------------------------------------------------------ */
/* Row-wise CrossEntropy + argmax over fp32 logits [rows, ld] (every row labelled, labels in [0, ncols)).
 * double_softmax = 1 reproduces token_head's Softmax -> CrossEntropyLoss (vs/encoder.py:323-326,381-385).
 * stats: fp32 [rows, 2] saved for backward. */
int vt_softmax_ce_fwd(const float* logits, int64_t ld, int64_t rows, int64_t ncols, const int64_t* labels, int double_softmax,
                      float* row_loss, int64_t* argmax, float* stats, vt_stream stream);
/* dlogits (bf16 [rows, ld_out], pad columns zeroed) = scale * (*scale_dev) * dLoss_row/dlogits */
int vt_softmax_ce_bwd(const float* logits, int64_t ld, int64_t rows, int64_t ncols, const int64_t* labels, int double_softmax,
                      const float* stats, float scale, const float* scale_dev, void* dlogits, int64_t ld_out, vt_stream stream);
/* NextActionPrediction (vs/encoder.py:142-158) + CrossEntropy(ignore_index=-1) + argmax (:391-398):
 * scores = log_softmax(pooled W^T + b) fp32 [B, actions]; labels may be NULL. */
int vt_action_head_fwd(const float* pooled, const float* W, const float* bias, int64_t B, int64_t hidden, int64_t actions,
                       const int64_t* labels, float* scores, float* row_loss, int64_t* argmax, vt_stream stream);
/* dpooled = ..., dW += ..., dbias += ...; gradient scale = (*g_dev) / (*n_valid_dev) */
int vt_action_head_bwd(const float* pooled, const float* W, int64_t B, int64_t hidden, int64_t actions, const int64_t* labels,
                       const float* scores, const float* n_valid_dev, const float* g_dev, float* dpooled, float* dW, float* dbias,
                       vt_stream stream);
/* out8 = {loss, mask_loss, next_loss, token_loss, words_accuracy, action_accuracy, token_accuracy, n_valid_actions}
 * (vs/encoder.py:396-441; means over labelled rows, 0/0 -> nan like the reference; n_act == 0 means
 * next_action was None: next_loss = 0 and action_accuracy = 0, vs/encoder.py:393-395,412-418). */
int vt_pretrain_finalize(const float* mlm_loss, const int64_t* mlm_arg, const int64_t* mlm_lab, int64_t n_mlm, const float* tok_loss,
                         const int64_t* tok_arg, const int64_t* tok_lab, int64_t n_tok, const float* act_loss, const int64_t* act_arg,
                         const int64_t* act_lab, int64_t n_act, float* out8, vt_stream stream);

/* the same with the two label counts read from vt_compact_labels' counts (no host knowledge of them); a label outside its class
 * range turns the loss NaN (torch's CrossEntropyLoss raises a device assert there) */
int vt_pretrain_finalize_counted(const float* mlm_loss, const int64_t* mlm_arg, const int64_t* mlm_lab, const float* tok_loss,
                                 const int64_t* tok_arg, const int64_t* tok_lab, const float* act_loss, const int64_t* act_arg,
                                 const int64_t* act_lab, int64_t n_act, const int32_t* counts, float* out8, vt_stream stream);

/* dx (bf16) = dy * (1 - y^2)  -- backward of the pooler's tanh (hf/modeling_bert.py:461-467) */
int vt_tanh_bwd(const float* dy, const float* y, int64_t n, void* dx, vt_stream stream);
/* dx (bf16) = dy * gelu'(pre), flat over n elements -- backward of the MLM transform activation (hf/modeling_bert.py:480-485) */
int vt_gelu_bwd(const void* dy, const void* pre, int64_t n, void* dx, vt_stream stream);
/* module-boundary casts: bf16 -> fp32 ; bf16(x) (+ addend) */
int vt_cast_bf16_f32(const void* x, int64_t n, float* y, vt_stream stream);
int vt_cast_f32_bf16_add(const float* x, const void* addend, int64_t n, void* y, vt_stream stream);

/* ------------------------------------------------------------------------------------------
 * Fine-tuning heads around the encoder, fp32 end to end so that the
 * action argmax agrees with the reference exactly.
 * ------------------------------------------------------------------------------------------ */
/* y[B,N] = act([x1 ; x2] W^T + bias): nn.Linear on a decoder step's batch, the input optionally the
 * concatenation of two tensors (replaces torch.cat + linear_out, vs/agent_models.py:351-353).
 * act: 0 none, 1 tanh.  Replaces SoftDotAttention.linear_in / linear_out (vs/agent_models.py:323-326),
 * decoder2action (tb/agent_models.py:297,318), question_linear (vs/agent_models.py:455-459,503). */
int vt_small_linear_fwd(const float* x1, int64_t ldx1, int64_t K1, const float* x2, int64_t ldx2, int64_t K2, const float* W,
                        int64_t ldw, const float* bias, int act, float* y, int64_t ldy, int64_t B, int64_t N, vt_stream stream);
/* dx1 / dx2 / dW / dbias may be NULL (not needed); dW, dbias are overwritten, not accumulated.  workspace: fp32 scratch of
 * vt_small_linear_bwd_workspace(B, N, K1, K2) floats (0 for N <= 256: may be NULL) for the chunked, fixed-order dx reduction. */
int64_t vt_small_linear_bwd_workspace(int64_t B, int64_t N, int64_t K1, int64_t K2);
int vt_small_linear_bwd(const float* dy, int64_t lddy, const float* y, int64_t ldy, int act, const float* x1, int64_t ldx1, int64_t K1,
                        const float* x2, int64_t ldx2, int64_t K2, const float* W, int64_t ldw, float* dx1, int64_t lddx1, float* dx2,
                        int64_t lddx2, float* dW, int64_t lddw, float* dbias, int64_t B, int64_t N, float* workspace, vt_stream stream);
/* Context splits P the two calls below use; workspace floats: fwd B*P*(2+D), bwd B*L + B*P*D. */
int64_t vt_softdot_splits(int64_t B, int64_t L);
/* SoftDotAttention core (vs/agent_models.py:335-349): logit[b,l] = ctx[b,l,:].target[b,:], positions with mask != 0
 * set to -inf IN the returned logit (the reference's masked_fill_ aliases it), attn = softmax_l(logit),
 * weighted[b,:] = sum_l attn[b,l] ctx[b,l,:].  ctx: [B,L,D] contiguous, D % 4 == 0, D <= 2176. */
int vt_softdot_fwd(const float* ctx, const float* target, const uint8_t* mask, int64_t B, int64_t L, int64_t D, float* logit, float* attn,
                   float* weighted, float* workspace, vt_stream stream);
/* Backward: upstream gradients of weighted, of attn (output_prob=True) and of the returned logit (output_prob=False),
 * each optional; dctx optional (candidate / view features carry no gradient in the reference). */
int vt_softdot_bwd(const float* ctx, const float* target, const uint8_t* mask, const float* attn, const float* dweighted, const float* dattn,
                   const float* dlogit_ext, int64_t B, int64_t L, int64_t D, float* dctx, float* dtarget, float* workspace, vt_stream stream);
/* Action-logit masking + CrossEntropyLoss(ignore_index, mean) + argmax (vs/agent.py:396-408, tb/agent.py:150-152,316-327):
 * columns >= valid_len[b] and, where forbid[b], column forbid_col ("forward" = 4 in tb/agent.py:71-80) become -inf in `masked` (may alias logits);
 * out3 = {loss, rows not ignored, sum of their losses}; lse fp32 [B] saved for backward.  A step whose rows are ALL ignored gives
 * nan like torch, or 0 with empty_is_zero (rollouts: the reference leaves the episode loop before such a step, vs/agent.py:466-468). */
int vt_head_ce_fwd(const float* logits, int64_t ld, int64_t B, int64_t C, const int64_t* valid_len, const uint8_t* forbid,
                   int64_t forbid_col, int empty_is_zero, const int64_t* target, int64_t ignore_index, float* masked, int64_t ldm, float* lse, int64_t* argmax, float* out3,
                   vt_stream stream);
int vt_head_ce_bwd(const float* masked, int64_t ldm, int64_t B, int64_t C, const int64_t* target, int64_t ignore_index, const float* lse,
                   const float* out3, const float* g, float* dlogits, int64_t ldd, vt_stream stream);
/* BCEWithLogitsLoss(pos_weight, reduction none) masked mean (vs/classifier/agent.py:158-163,493-503): out2 = {loss, n_valid}. */
int vt_bce_logits_fwd(const float* q, const float* target, const uint8_t* valid, int64_t B, float pos_weight, float* per_row, float* out2,
                      vt_stream stream);
int vt_bce_logits_bwd(const float* q, const float* target, const uint8_t* valid, int64_t B, float pos_weight, const float* out2,
                      const float* g, float* dq, vt_stream stream);

/* ------------------------------------------------------------------------------------------
 * The agents' recurrent middle, hidden size 512, fp32 state.
 * ------------------------------------------------------------------------------------------ */
/* OscarEncoder's nn.LSTM(768 -> 512, 1 layer, batch_first) over all T encoder positions (vs/agent_models.py:223-230,283-287 ==
 * tb/agent_models.py:148-155,208-212) behind pack_padded_sequence / pad_packed_sequence (:284,305).
 * gates [B,T,2048]: IN the input projection x W_ih^T + b_ih + b_hh of every step (one vt_gemm_bf16, gate order i,f,g,o as
 * nn.LSTM), OUT the activated gates (saved for backward).  w_hh: fp32 [2048,512].  lengths: device int64 [B] or NULL (= T):
 * sample b advances for t < lengths[b], ctx [B,T,512] is 0 behind its end, (h_last, c_last) [B,512] are its last valid state.
 * c_all [B,T,512] fp32 and hprev_bf16 [B,T,512] (= bf16(h_{t-1}), the operand of dW_hh) are saved for backward.
 * One 16-CTA thread-block cluster per 8 samples; W_hh stays in shared memory for the whole sequence. */
int vt_lstm_seq_fwd(float* gates, const float* w_hh, const int64_t* lengths, int64_t B, int64_t T, int64_t hidden, float* ctx,
                    float* c_all, void* hprev_bf16, float* h_last, float* c_last, vt_stream stream);
/* Backward through the recurrence: dA_bf16 [B,T,2048] = gradient of the gate pre-activations (0 behind a sample's end).  The
 * caller finishes with GEMMs over all steps: dW_ih += dA^T X, dW_hh += dA^T hprev, dX = dA W_ih, db_ih = db_hh = colsum(dA).
 * dh_last / dc_last: gradients of the returned final state (NULL = none). */
int vt_lstm_seq_bwd(const float* gates, const float* c_all, const float* w_hh, const int64_t* lengths, const float* dctx,
                    const float* dh_last, const float* dc_last, int64_t B, int64_t T, int64_t hidden, void* dA_bf16, vt_stream stream);
/* nn.LSTMCell of the decoders (vs/agent_models.py:380,414 ; tb/agent_models.py:295,313): gates = ga + gb, the two halves
 * x W_ih^T + b_ih and h W_hh^T + b_hh from vt_small_linear_fwd; gates_act [B,4H] keeps the activated gates for backward. */
int vt_lstm_cell_fwd(const float* ga, const float* gb, const float* c0, int64_t B, int64_t hidden, float* gates_act, float* h1, float* c1,
                     vt_stream stream);
/* dgates [B,4H] (the gradient of both halves), dc0 [B,H]; dh1 / dc1 may be NULL. */
int vt_lstm_cell_bwd(const float* gates_act, const float* c0, const float* c1, const float* dh1, const float* dc1, int64_t B, int64_t hidden,
                     float* dgates, float* dc0, vt_stream stream);
/* nn.Dropout on a flat fp32 array (the agents' self.drop, p = 0.5, vs/agent_models.py:216,379): y = x * keep / (1 - p); the
 * backward is the same call on the gradient -- the mask is a function of (seed, dropout_stream, index), never stored. */
int vt_dropout_f32(const float* x, int64_t n, float dropout_p, uint64_t seed, uint32_t dropout_stream, float* y, vt_stream stream);

/* nn.Embedding of the turn-based decoder's previous action (tb/agent_models.py:292,308): y[b] = W[idx[b]] (fp32 [V,D]); the backward
 * OVERWRITES dW with the per-row sums, samples in ascending order (deterministic). */
int vt_embedding_f32_fwd(const float* W, const int64_t* idx, int64_t B, int64_t V, int64_t D, float* y, vt_stream stream);
int vt_embedding_f32_bwd(const float* dy, const int64_t* idx, int64_t B, int64_t V, int64_t D, float* dW, vt_stream stream);
/* Action feedback of one rollout step ON THE DEVICE (vs/agent.py:410-441 and tb/agent.py:329-346 do it on the host
 * behind a .cpu() / .item() sync per decoder step).  mode 0: a = target (teacher forcing); 1: argmax of the masked logits; 2: a sample
 * of softmax(masked logits) drawn with the counter RNG keyed on (seed, rng_stream, step, sample).  ended[b] |= a is the stop action
 * (stop_len[b] - 1, i.e. the <end> candidate, when stop_len is given; else stop_const) or the ignore id.  a_out [B]: the action.
 * next_target_eff (optional): ended[b] ? ignore_index : next_target[b] -- the teacher's answer for a finished episode. */
int vt_rollout_feedback(const float* masked_logits, int64_t ld, int64_t B, int64_t C, const int64_t* target, const int64_t* stop_len,
                        int64_t stop_const, int64_t ignore_index, int mode, uint64_t seed, uint32_t rng_stream, uint32_t step, uint8_t* ended,
                        int64_t* a_out, const int64_t* next_target, int64_t* next_target_eff, vt_stream stream);

/* ---- the step after backward: multi-tensor AdamW and gradient-norm clipping ----
 * Replace transformers/pytorch_transformers/optimization.py:141-189 (AdamW.step, ~10 ATen launches per parameter) and
 * torch.nn.utils.clip_grad_norm_ at tasks/viewpoint_select/agent.py:514-518.
 * A segment table lives in DEVICE memory (the caller builds it once and reuses it while the pointers stay put): */
typedef struct {
  float* p;            /* parameter (fp32 master), n elements */
  float* g;            /* its gradient */
  float* m;            /* exp_avg */
  float* v;            /* exp_avg_sq */
  void* p_bf16;        /* optional bf16 copy of p (the tensor-core operand of the next forward), written in the same pass; NULL: none */
  int64_t n;
  int32_t group;       /* parameter group 0..7: selects step_size[] / lr_decay[] */
  int32_t first_chunk; /* prefix sum of ceil(n / vt_opt_chunk()) over the preceding segments */
} vt_opt_segment;
int64_t vt_opt_chunk(void);
/* One launch for every segment: m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; p -= step_size[group] m / (sqrt(v) + eps);
 * p -= lr_decay[group] p  (step_size = lr sqrt(1-b2^t)/(1-b1^t) with correct_bias, else lr; lr_decay = lr * weight_decay; host arrays).
 * beta1 / beta2 / eps are doubles so that (1 - beta) is formed exactly as the reference's Python scalars form it.
 * grad_scale_dev: optional device scalar multiplied into every gradient on the fly (the clip coefficient below). */
int vt_adamw_step(const void* segments_dev, int64_t n_segments, int64_t n_chunks, int64_t n_groups, const float* step_size,
                  const float* lr_decay, double beta1, double beta2, double eps, const float* grad_scale_dev, vt_stream stream);
/* norm_and_coef[0] = L2 norm over all segments' gradients (deterministic), [1] = min(1, max_norm / (norm + 1e-6));
 * partials: fp32 [n_chunks] workspace; scale_in_place != 0 multiplies the gradients by the coefficient (clip_grad_norm_). */
int vt_clip_grad_norm(const void* segments_dev, int64_t n_segments, int64_t n_chunks, float max_norm, float* partials,
                      float* norm_and_coef, int scale_in_place, vt_stream stream);

/* ---- the step before forward: on-device masked-LM masking ----
 * Replaces PretrainDataset._mask_tokens, tasks/viewpoint_select/data_loader_pretrain.py:549-612, for a whole batch in one
 * launch: masked = (u1 < p_mask and id not special) or token_class != -1; label = id where masked and token_class == -1,
 * else -1; id' = mask_id if (u2 < 0.8 and masked) or token_class != -1, else a random word if u3 < 0.5 and masked, else id;
 * attention = id != pad_id.  special_lut: uint8 [vocab], 1 for tokenizer.all_special_ids; token_classes may be null
 * (masked_token_prediction off).  One Philox4x32-7 call per token keyed on (seed, token index). */
int vt_mlm_mask(const int64_t* ids, const int64_t* token_classes, const uint8_t* special_lut, int64_t n, int64_t vocab,
                int64_t mask_id, int64_t pad_id, float p_mask, uint64_t seed, int64_t* ids_out, int64_t* labels,
                uint8_t* attention_mask, vt_stream stream);
/* The draws vt_mlm_mask uses for (seed, n, vocab): u fp32 [3][n] (u1, u2, u3), r int64 [n]; for the parity tests. */
int vt_mlm_draws(int64_t n, int64_t vocab, uint64_t seed, float* u, int64_t* r, vt_stream stream);
/* Ragged per-sample region rows -> the padded [samples, regions, *] feature / location blocks of the model batch, replacing the per-sample
 * slicing, torch.zeros and torch.cat of PretrainDataset._preprocess_item (data_loader_pretrain.py:656-690): sample s owns rows
 * [offsets[s], offsets[s+1]) of the packed fp32 arrays; more than `regions` rows keeps the LAST `regions`, fewer pads with zeros.
 * attention_tail (optional, int64, row pitch ld_attention) receives 1 for a kept row and 0 for padding. */
int vt_pack_regions(const float* feats, int64_t d_feat, const float* loc, int64_t d_loc, const int64_t* offsets, int64_t samples,
                    int64_t regions, float* out_feats, float* out_loc, int64_t* attention_tail, int64_t ld_attention, vt_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* VISITRON_B200_H */
