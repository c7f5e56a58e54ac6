// Loss / scoring kernels of the pre-training heads (vs/encoder.py:377-441):
//   * row-wise softmax cross-entropy + argmax over fp32 logits (masked-LM head), with the
//     reference's DOUBLE softmax for the detector-class head (Softmax module followed by
//     CrossEntropyLoss, vs/encoder.py:323-326,381-385) and its backward (writes bf16 dlogits);
//   * the 36-way next-action head: Linear + LogSoftmax + CrossEntropy + argmax in one warp per
//     sample (vs/encoder.py:142-158,391-398) and its backward;
//   * the 7-scalar loss / accuracy assembly (vs/encoder.py:396-441);
//   * small elementwise helpers (tanh backward, bf16 <-> fp32 casts).
#include "common.cuh"

namespace vt {

void count_launch(int n);

constexpr int CE_THREADS = 256;

struct MaxArg {
  float m;
  int idx;
};
__device__ __forceinline__ MaxArg better(MaxArg a, MaxArg b) {  // larger value; ties -> lower index (torch.argmax)
  return (b.m > a.m || (b.m == a.m && b.idx < a.idx)) ? b : a;
}

template <typename T>
__device__ __forceinline__ T block_reduce(T v, T (*op)(T, T), T* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    T other;
    unsigned char* src = reinterpret_cast<unsigned char*>(&v);
    unsigned char* dst = reinterpret_cast<unsigned char*>(&other);
#pragma unroll
    for (int w = 0; w < (int)(sizeof(T) / 4); ++w) reinterpret_cast<uint32_t*>(dst)[w] = __shfl_xor_sync(0xffffffffu, reinterpret_cast<uint32_t*>(src)[w], o);
    v = op(v, other);
  }
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  T r = scratch[0];
  for (int w = 1; w < (int)(blockDim.x >> 5); ++w) r = op(r, scratch[w]);
  return r;
}
__device__ __forceinline__ float fadd(float a, float b) { return a + b; }
__device__ __forceinline__ float fmaxop(float a, float b) { return fmaxf(a, b); }

// One CTA per row.  stats[row] = {lse, lse2}: lse = logsumexp(x); double softmax: lse2 = logsumexp(softmax(x)).
__global__ void __launch_bounds__(CE_THREADS)
softmax_ce_fwd_kernel(const float* __restrict__ logits, long long ld, int ncols, const long long* __restrict__ labels, int double_softmax,
                      float* __restrict__ row_loss, long long* __restrict__ argmax, float* __restrict__ stats,
                      const int32_t* __restrict__ limit) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  __shared__ MaxArg s_ma[CE_THREADS / 32];
  __shared__ float s_f[CE_THREADS / 32];
  const long long row = blockIdx.x;
  if (limit && row >= (long long)__ldg(limit + 1)) return;   // device-counted rows: only the labelled ones exist
  const float* x = logits + row * ld;
  MaxArg ma = {-INFINITY, 0x7fffffff};
  for (int j = threadIdx.x; j < ncols; j += CE_THREADS) ma = better(ma, MaxArg{x[j], j});
  ma = block_reduce<MaxArg>(ma, better, s_ma);
  float s = 0.f;
  for (int j = threadIdx.x; j < ncols; j += CE_THREADS) s += __expf(x[j] - ma.m);
  s = block_reduce<float>(s, fadd, s_f);
  const float lse = ma.m + logf(s);
  long long y = labels[row];
  y = y < 0 ? 0 : (y >= ncols ? ncols - 1 : y);   // memory safety only: out-of-range labels are reported by vt_compact_labels
  float loss, lse2 = 0.f;
  if (!double_softmax) {
    loss = lse - x[y];
  } else {
    // probabilities are in [0,1]: no max subtraction needed for the second log-sum-exp
    float s2 = 0.f;
    for (int j = threadIdx.x; j < ncols; j += CE_THREADS) s2 += __expf(__expf(x[j] - lse));
    s2 = block_reduce<float>(s2, fadd, s_f);
    lse2 = logf(s2);
    loss = lse2 - __expf(x[y] - lse);
  }
  if (threadIdx.x == 0) {
    row_loss[row] = loss;
    argmax[row] = ma.idx;
    stats[2 * row] = lse;
    stats[2 * row + 1] = lse2;
  }
}

// dlogits[row, j] = g * dL/dx_j, g = scale * (*scale_dev); pad columns [ncols, ld_out) are zeroed.
__global__ void __launch_bounds__(CE_THREADS)
softmax_ce_bwd_kernel(const float* __restrict__ logits, long long ld, int ncols, const long long* __restrict__ labels, int double_softmax,
                      const float* __restrict__ stats, float scale, const float* __restrict__ scale_dev, bf16* __restrict__ dlogits,
                      long long ld_out, const int32_t* __restrict__ limit) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  __shared__ float s_f[CE_THREADS / 32];
  const long long row = blockIdx.x;
  const float* x = logits + row * ld;
  bf16* d = dlogits + row * ld_out;
  if (limit) {
    // device-counted rows {padded, true}: rows past the padded count do not exist; the padding rows [true, padded) are
    // written as zeros (they are contraction rows of the weight-gradient GEMM); the mean's 1 / n comes from the device
    const long long n_pad = __ldg(limit), n = __ldg(limit + 1);
    if (row >= n_pad) return;
    if (row >= n) {
      for (int j = threadIdx.x; j < ld_out; j += CE_THREADS) d[j] = __float2bfloat16(0.f);
      return;
    }
    scale /= (float)n;
  }
  const float g = scale * (scale_dev ? __ldg(scale_dev) : 1.0f);
  const float lse = stats[2 * row], lse2 = stats[2 * row + 1];
  int y = (int)labels[row];
  y = y < 0 ? 0 : (y >= ncols ? ncols - 1 : y);
  if (!double_softmax) {
    for (int j = threadIdx.x; j < ld_out; j += CE_THREADS) {
      float v = 0.f;
      if (j < ncols) v = (__expf(x[j] - lse) - (j == y ? 1.f : 0.f)) * g;
      d[j] = __float2bfloat16(v);
    }
  } else {
    // L = lse2 - p_y,  p = softmax(x),  q = softmax(p):  dL/dp_k = q_k - [k == y];  dL/dx_j = p_j (dL/dp_j - sum_k dL/dp_k p_k)
    float dot = 0.f;
    for (int j = threadIdx.x; j < ncols; j += CE_THREADS) {
      float pj = __expf(x[j] - lse);
      dot += (__expf(pj - lse2) - (j == y ? 1.f : 0.f)) * pj;
    }
    dot = block_reduce<float>(dot, fadd, s_f);
    for (int j = threadIdx.x; j < ld_out; j += CE_THREADS) {
      float v = 0.f;
      if (j < ncols) {
        float pj = __expf(x[j] - lse);
        v = pj * ((__expf(pj - lse2) - (j == y ? 1.f : 0.f)) - dot) * g;
      }
      d[j] = __float2bfloat16(v);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// next-action head: one warp per sample.  scores = log_softmax(pooled . W^T + b)  (returned),
// loss = CrossEntropy(scores, y) = -log_softmax(scores)[y]  (the reference applies log-softmax twice)
// ---------------------------------------------------------------------------------------------
constexpr int MAX_ACTIONS = 64;

__global__ void __launch_bounds__(128)
action_head_fwd_kernel(const float* __restrict__ pooled, const float* __restrict__ W, const float* __restrict__ bias, int B, int Hd, int A,
                       const long long* __restrict__ labels, float* __restrict__ scores, float* __restrict__ row_loss,
                       long long* __restrict__ argmax) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  const int lane = threadIdx.x & 31;
  const int s = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (s >= B) return;
  const float* x = pooled + (long long)s * Hd;
  float mine0 = -INFINITY, mine1 = -INFINITY;  // lane owns actions lane and lane + 32
  // four actions per pass: their loads and shuffle reductions interleave (one action at a time was a chain of dependent
  // load -> fma -> 5 shuffles, 36 times: 137 us for a 32 x 768 x 36 product)
  for (int a0 = 0; a0 < A; a0 += 4) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int k = lane; k < Hd; k += 32) {
      const float xv = x[k];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (a0 + u < A) acc[u] = fmaf(xv, __ldg(W + (long long)(a0 + u) * Hd + k), acc[u]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int u = 0; u < 4; ++u) acc[u] += __shfl_xor_sync(0xffffffffu, acc[u], o);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int a = a0 + u;
      if (a < A && (a & 31) == lane) { if (a < 32) mine0 = acc[u] + bias[a]; else mine1 = acc[u] + bias[a]; }
    }
  }
  float m = warp_max(fmaxf(mine0, mine1));
  float e = (lane < A ? __expf(mine0 - m) : 0.f) + (lane + 32 < A ? __expf(mine1 - m) : 0.f);
  const float lse = m + logf(warp_sum(e));
  const float ls0 = mine0 - lse, ls1 = mine1 - lse;  // log-probabilities
  float e2 = (lane < A ? __expf(ls0) : 0.f) + (lane + 32 < A ? __expf(ls1) : 0.f);
  const float lse2 = logf(warp_sum(e2));
  if (lane < A) scores[(long long)s * A + lane] = ls0;
  if (lane + 32 < A) scores[(long long)s * A + lane + 32] = ls1;
  // argmax with lowest-index tie break
  float bv = lane < A ? ls0 : -INFINITY;
  int bi = lane;
  if (lane + 32 < A && ls1 > bv) { bv = ls1; bi = lane + 32; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
  }
  const long long y = labels ? labels[s] : -1;
  float ly = 0.f;
  if (y >= 0) {
    float cand = (y < 32) ? ls0 : ls1;
    ly = __shfl_sync(0xffffffffu, cand, (int)(y & 31));
  }
  if (lane == 0) {
    row_loss[s] = (y >= 0) ? (lse2 - ly) : 0.f;
    argmax[s] = bi;
  }
}

// dscore[s, a] = (exp(score[s,a]) - [a == y_s]) * g / n_valid ; then dpooled = dscore . W, dW = dscore^T . pooled, db = colsum
__global__ void __launch_bounds__(256)
action_head_bwd_kernel(const float* __restrict__ pooled, const float* __restrict__ W, int B, int Hd, int A, const long long* __restrict__ labels,
                       const float* __restrict__ scores, const float* __restrict__ n_valid_dev, const float* __restrict__ g_dev,
                       float* __restrict__ dpooled, float* __restrict__ dW, float* __restrict__ dbias) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  extern __shared__ float sd[];  // [B][A] dscore
  const float nv = __ldg(n_valid_dev);
  const float g = (g_dev ? __ldg(g_dev) : 1.0f) / nv;  // nv == 0 -> inf/nan like the reference's empty mean
  for (int i = threadIdx.x; i < B * A; i += blockDim.x) {
    const int s = i / A, a = i % A;
    const long long y = labels[s];
    sd[i] = (y >= 0) ? (__expf(scores[i]) - (a == y ? 1.f : 0.f)) * g : 0.f;
  }
  __syncthreads();
  // this CTA owns the hidden columns [k0, k1)
  const int per = (Hd + gridDim.x - 1) / gridDim.x;
  const int k0 = blockIdx.x * per, k1 = min(Hd, k0 + per);
  for (int i = threadIdx.x; i < B * (k1 - k0); i += blockDim.x) {
    const int s = i / (k1 - k0), k = k0 + i % (k1 - k0);
    float acc = 0.f;
    for (int a = 0; a < A; ++a) acc = fmaf(sd[s * A + a], __ldg(W + (long long)a * Hd + k), acc);
    dpooled[(long long)s * Hd + k] = acc;
  }
  for (int i = threadIdx.x; i < A * (k1 - k0); i += blockDim.x) {
    const int a = i / (k1 - k0), k = k0 + i % (k1 - k0);
    float acc = 0.f;
    for (int s = 0; s < B; ++s) acc = fmaf(sd[s * A + a], pooled[(long long)s * Hd + k], acc);
    dW[(long long)a * Hd + k] += acc;
  }
  if (blockIdx.x == 0)
    for (int a = threadIdx.x; a < A; a += blockDim.x) {
      float acc = 0.f;
      for (int s = 0; s < B; ++s) acc += sd[s * A + a];
      dbias[a] += acc;
    }
}

// ---------------------------------------------------------------------------------------------
// 7-tuple assembly (vs/encoder.py:396-441).  out = {loss, mask_loss, next_loss, token_loss,
// words_accuracy, action_accuracy, token_accuracy, n_valid_actions}
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
pretrain_finalize_kernel(const float* __restrict__ mlm_loss, const long long* __restrict__ mlm_arg, const long long* __restrict__ mlm_lab, long long n_mlm,
                         const float* __restrict__ tok_loss, const long long* __restrict__ tok_arg, const long long* __restrict__ tok_lab, long long n_tok,
                         const float* __restrict__ act_loss, const long long* __restrict__ act_arg, const long long* __restrict__ act_lab, long long n_act,
                         const int32_t* __restrict__ counts, float* __restrict__ out) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  __shared__ float sf[8];
  bool bad = false;
  if (counts) {   // vt_compact_labels' {mlm padded, mlm, mlm out of range, -, token padded, token, token out of range, -}
    n_mlm = counts[1];
    n_tok = counts[5];
    bad = counts[2] != 0 || counts[6] != 0;
  }
  float a = 0.f, c = 0.f;
  for (long long i = threadIdx.x; i < n_mlm; i += blockDim.x) { a += mlm_loss[i]; c += (mlm_arg[i] == mlm_lab[i]) ? 1.f : 0.f; }
  const float mlm_sum = block_reduce<float>(a, fadd, sf), mlm_ok = block_reduce<float>(c, fadd, sf);
  a = 0.f; c = 0.f;
  for (long long i = threadIdx.x; i < n_tok; i += blockDim.x) { a += tok_loss[i]; c += (tok_arg[i] == tok_lab[i]) ? 1.f : 0.f; }
  const float tok_sum = block_reduce<float>(a, fadd, sf), tok_ok = block_reduce<float>(c, fadd, sf);
  a = 0.f; c = 0.f;
  float v = 0.f;
  for (long long i = threadIdx.x; i < n_act; i += blockDim.x) {
    const bool valid = act_lab[i] >= 0;
    a += valid ? act_loss[i] : 0.f;
    v += valid ? 1.f : 0.f;
    c += (act_arg[i] == act_lab[i]) ? 1.f : 0.f;
  }
  const float act_sum = block_reduce<float>(a, fadd, sf), act_ok = block_reduce<float>(c, fadd, sf), act_valid = block_reduce<float>(v, fadd, sf);
  if (threadIdx.x == 0) {
    const float mask_loss = mlm_sum / (float)n_mlm;     // 0/0 -> nan, as CrossEntropyLoss over an all-ignored batch
    const float token_loss = tok_sum / (float)n_tok;
    // n_act == 0: next_action was None -> next_loss = 0, action_accuracy = 0 (vs/encoder.py:393-395,412-418);
    // labels given but all ignored -> 0/0 = nan like CrossEntropyLoss
    const float next_loss = n_act > 0 ? act_sum / act_valid : 0.f;
    // a label outside [0, classes) makes torch's CrossEntropyLoss raise a device assert; here the loss turns NaN (loud, no sync)
    out[0] = bad ? __int_as_float(0x7fc00000) : mask_loss + next_loss + token_loss;
    out[1] = mask_loss;
    out[2] = next_loss;
    out[3] = token_loss;
    out[4] = mlm_ok / (float)n_mlm;
    out[5] = n_act > 0 ? act_ok / (float)n_act : 0.f;
    out[6] = tok_ok / (float)n_tok;
    out[7] = act_valid;
  }
}

// ---------------------------------------------------------------------------------------------
// Labelled-row compaction on the device (replaces torch.nonzero + fancy indexing, whose size-dependent output forced a
// host <-> device sync at the top of every step).  One CTA per label array: each thread counts a contiguous chunk, a block
// scan gives its offset, the labelled rows are written in ascending row order (the order torch.nonzero gives, so every
// downstream sum keeps its order).  counts[4 a ..] = {n rounded up to 128 (capped), n, labels outside [0, classes), 0};
// the padding rows [n, padded) get idx = -1 / label = 0.
// ---------------------------------------------------------------------------------------------
struct CompactArgs {
  const long long* labels[2];
  long long* idx[2];
  long long* lab[2];
  long long classes[2];
};
__global__ void __launch_bounds__(1024)
compact_labels_kernel(CompactArgs a, long long n, long long ignore, long long cap, int32_t* __restrict__ counts) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  __shared__ int s_warp[32];
  __shared__ int s_bad;
  const int w = blockIdx.x;
  const long long* lab = a.labels[w];
  if (!lab) {
    if (threadIdx.x < 4) counts[4 * w + threadIdx.x] = 0;
    return;
  }
  if (threadIdx.x == 0) s_bad = 0;
  const long long per = (n + blockDim.x - 1) / blockDim.x;
  const long long i0 = min(n, per * threadIdx.x), i1 = min(n, i0 + per);
  int cnt = 0, bad = 0;
  for (long long i = i0; i < i1; ++i) {
    const long long y = lab[i];
    if (y != ignore) {
      ++cnt;
      if (y < 0 || y >= a.classes[w]) ++bad;
    }
  }
  // block-wide exclusive scan of cnt
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int v = lane < (int)(blockDim.x >> 5) ? s_warp[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += t;
    }
    s_warp[lane] = v;   // inclusive over warps
  }
  if (bad) atomicAdd(&s_bad, bad);
  __syncthreads();
  const int total = s_warp[(blockDim.x >> 5) - 1];
  long long pos = (long long)(inc - cnt) + (warp ? s_warp[warp - 1] : 0);
  for (long long i = i0; i < i1; ++i) {
    const long long y = lab[i];
    if (y != ignore) {
      if (pos < cap) { a.idx[w][pos] = i; a.lab[w][pos] = y; }
      ++pos;
    }
  }
  const long long n_lab = min((long long)total, cap);
  const long long n_pad = min(cap, (n_lab + 127) / 128 * 128);
  for (long long r = n_lab + threadIdx.x; r < n_pad; r += blockDim.x) { a.idx[w][r] = -1; a.lab[w][r] = 0; }
  if (threadIdx.x == 0) {
    counts[4 * w] = (int32_t)n_pad;
    counts[4 * w + 1] = (int32_t)n_lab;
    counts[4 * w + 2] = s_bad + (total > cap ? 1 : 0);
    counts[4 * w + 3] = 0;
  }
}

__global__ void tanh_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y, long long n, bf16* __restrict__ dx) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dx[i] = __float2bfloat16(dy[i] * (1.0f - y[i] * y[i]));
}

// dx = dy * gelu'(pre)   (backward of the MLM transform's activation, hf/modeling_bert.py:480-485)
__global__ void gelu_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ pre, long long n8, bf16* __restrict__ dx,
                                long long row8, const int32_t* __restrict__ limit) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  if (limit) n8 = min(n8, (long long)__ldg(limit) * row8);   // row8 = 16-byte chunks per row
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    float d[8], x[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(dy) + i), d);
    unpack8(__ldg(reinterpret_cast<const uint4*>(pre) + i), x);
#pragma unroll
    for (int j = 0; j < 8; ++j) d[j] *= gelu_grad(x[j]);
    reinterpret_cast<uint4*>(dx)[i] = pack8(d);
  }
}

__global__ void cast_bf16_f32_kernel(const bf16* __restrict__ x, long long n8, float* __restrict__ y) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    float v[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(x) + i), v);
    reinterpret_cast<float4*>(y)[2 * i] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(y)[2 * i + 1] = make_float4(v[4], v[5], v[6], v[7]);
  }
}
// y = bf16(x) (+ bf16 add of `addend` when given)
__global__ void cast_f32_bf16_flat_kernel(const float* __restrict__ x, const bf16* __restrict__ addend, long long n8, bf16* __restrict__ y) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    float4 a = __ldg(reinterpret_cast<const float4*>(x) + 2 * i), b = __ldg(reinterpret_cast<const float4*>(x) + 2 * i + 1);
    float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    if (addend) {
      float w[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(addend) + i), w);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += w[j];
    }
    reinterpret_cast<uint4*>(y)[i] = pack8(v);
  }
}

}  // namespace vt

using namespace vt;

extern "C" int vt_softmax_ce_fwd(const float* logits, int64_t ld, int64_t rows, int64_t ncols, const int64_t* labels, int double_softmax,
                                 float* row_loss, int64_t* argmax, float* stats, vt_stream stream) {
  VT_REQUIRE(rows >= 0 && ncols > 0 && ncols < (1ll << 31) && ld >= ncols, "vt_softmax_ce_fwd: bad sizes");
  if (rows == 0) return VT_OK;
  launch_pdl(softmax_ce_fwd_kernel, (unsigned)rows, CE_THREADS, 0, (cudaStream_t)stream, logits, ld, (int)ncols, (const long long*)labels, double_softmax,
                                                                                 row_loss, (long long*)argmax, stats, row_limit());
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_softmax_ce_bwd(const float* logits, int64_t ld, int64_t rows, int64_t ncols, const int64_t* labels, int double_softmax,
                                 const float* stats, float scale, const float* scale_dev, void* dlogits, int64_t ld_out, vt_stream stream) {
  VT_REQUIRE(rows >= 0 && ncols > 0 && ld >= ncols && ld_out >= ncols, "vt_softmax_ce_bwd: bad sizes");
  if (rows == 0) return VT_OK;
  launch_pdl(softmax_ce_bwd_kernel, (unsigned)rows, CE_THREADS, 0, (cudaStream_t)stream, logits, ld, (int)ncols, (const long long*)labels, double_softmax,
                                                                                 stats, scale, scale_dev, (bf16*)dlogits, ld_out, row_limit());
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_action_head_fwd(const float* pooled, const float* W, const float* bias, int64_t B, int64_t hidden, int64_t actions,
                                  const int64_t* labels, float* scores, float* row_loss, int64_t* argmax, vt_stream stream) {
  VT_REQUIRE(actions > 0 && actions <= MAX_ACTIONS, "vt_action_head_fwd: action space %lld outside the supported envelope (<= 64)", (long long)actions);
  if (B == 0) return VT_OK;
  launch_pdl(action_head_fwd_kernel, (unsigned)((B + 3) / 4), 128, 0, (cudaStream_t)stream, pooled, W, bias, (int)B, (int)hidden, (int)actions,
                                                                                    (const long long*)labels, scores, row_loss, (long long*)argmax);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_action_head_bwd(const float* pooled, const float* W, int64_t B, int64_t hidden, int64_t actions, const int64_t* labels,
                                  const float* scores, const float* n_valid_dev, const float* g_dev, float* dpooled, float* dW, float* dbias,
                                  vt_stream stream) {
  VT_REQUIRE(actions > 0 && actions <= MAX_ACTIONS && B * actions * 4 <= 160 * 1024, "vt_action_head_bwd: batch x actions too large for one CTA's shared memory");
  if (B == 0) return VT_OK;
  const size_t smem = (size_t)B * actions * sizeof(float);
  static bool attr = false;
  if (!attr) { VT_CUDA(cudaFuncSetAttribute(action_head_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024)); attr = true; }
  launch_pdl(action_head_bwd_kernel, 24, 256, smem, (cudaStream_t)stream, pooled, W, (int)B, (int)hidden, (int)actions, (const long long*)labels, scores,
                                                                 n_valid_dev, g_dev, dpooled, dW, dbias);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

static int finalize(const float* mlm_loss, const int64_t* mlm_arg, const int64_t* mlm_lab, int64_t n_mlm, const float* tok_loss,
                    const int64_t* tok_arg, const int64_t* tok_lab, int64_t n_tok, const float* act_loss, const int64_t* act_arg,
                    const int64_t* act_lab, int64_t n_act, const int32_t* counts, float* out8, vt_stream stream) {
  launch_pdl(pretrain_finalize_kernel, 1, 256, 0, (cudaStream_t)stream, mlm_loss, (const long long*)mlm_arg, (const long long*)mlm_lab, (long long)n_mlm, tok_loss,
                                                               (const long long*)tok_arg, (const long long*)tok_lab, (long long)n_tok, act_loss,
                                                               (const long long*)act_arg, (const long long*)act_lab, (long long)n_act, counts, out8);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_pretrain_finalize(const float* mlm_loss, const int64_t* mlm_arg, const int64_t* mlm_lab, int64_t n_mlm, const float* tok_loss,
                                    const int64_t* tok_arg, const int64_t* tok_lab, int64_t n_tok, const float* act_loss, const int64_t* act_arg,
                                    const int64_t* act_lab, int64_t n_act, float* out8, vt_stream stream) {
  return finalize(mlm_loss, mlm_arg, mlm_lab, n_mlm, tok_loss, tok_arg, tok_lab, n_tok, act_loss, act_arg, act_lab, n_act, nullptr, out8, stream);
}

extern "C" int vt_pretrain_finalize_counted(const float* mlm_loss, const int64_t* mlm_arg, const int64_t* mlm_lab, const float* tok_loss,
                                            const int64_t* tok_arg, const int64_t* tok_lab, const float* act_loss, const int64_t* act_arg,
                                            const int64_t* act_lab, int64_t n_act, const int32_t* counts, float* out8, vt_stream stream) {
  VT_REQUIRE(counts, "vt_pretrain_finalize_counted: counts (vt_compact_labels) required");
  return finalize(mlm_loss, mlm_arg, mlm_lab, 0, tok_loss, tok_arg, tok_lab, 0, act_loss, act_arg, act_lab, n_act, counts, out8, stream);
}

extern "C" int vt_compact_labels(const int64_t* labels_a, int64_t classes_a, const int64_t* labels_b, int64_t classes_b, int64_t n, int64_t ignore_index,
                                 int64_t capacity, int64_t* idx_a, int64_t* lab_a, int64_t* idx_b, int64_t* lab_b, int32_t* counts, vt_stream stream) {
  VT_REQUIRE(counts && n >= 0 && n < (1ll << 31) && capacity > 0 && capacity % 128 == 0, "vt_compact_labels: counts required, capacity a positive multiple of 128");
  VT_REQUIRE((!labels_a || (idx_a && lab_a)) && (!labels_b || (idx_b && lab_b)), "vt_compact_labels: output arrays required for every label array");
  CompactArgs a;
  a.labels[0] = (const long long*)labels_a; a.labels[1] = (const long long*)labels_b;
  a.idx[0] = (long long*)idx_a; a.idx[1] = (long long*)idx_b;
  a.lab[0] = (long long*)lab_a; a.lab[1] = (long long*)lab_b;
  a.classes[0] = classes_a; a.classes[1] = classes_b;
  launch_pdl(compact_labels_kernel, 2, 1024, 0, (cudaStream_t)stream, a, (long long)n, (long long)ignore_index, (long long)capacity, counts);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_tanh_bwd(const float* dy, const float* y, int64_t n, void* dx, vt_stream stream) {
  if (n == 0) return VT_OK;
  launch_pdl(tanh_bwd_kernel, (unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream, dy, y, n, (bf16*)dx);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_gelu_bwd(const void* dy, const void* pre, int64_t n, void* dx, vt_stream stream) {
  VT_REQUIRE(n % 8 == 0 && ((uintptr_t)dy & 15) == 0 && ((uintptr_t)pre & 15) == 0 && ((uintptr_t)dx & 15) == 0, "vt_gelu_bwd: n %% 8 == 0 and 16-byte alignment required");
  if (n == 0) return VT_OK;
  long long n8 = n / 8, g = (n8 + 255) / 256, cap = (long long)num_sms() * 16;
  launch_pdl(gelu_bwd_kernel, (unsigned)(g < cap ? g : cap), 256, 0, (cudaStream_t)stream, (const bf16*)dy, (const bf16*)pre, n8, (bf16*)dx, (long long)(768 / 8), row_limit());
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_cast_bf16_f32(const void* x, int64_t n, float* y, vt_stream stream) {
  VT_REQUIRE(n % 8 == 0 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)y & 15) == 0, "vt_cast_bf16_f32: n %% 8 == 0 and 16-byte alignment required");
  if (n == 0) return VT_OK;
  long long n8 = n / 8, g = (n8 + 255) / 256, cap = (long long)num_sms() * 16;
  launch_pdl(cast_bf16_f32_kernel, (unsigned)(g < cap ? g : cap), 256, 0, (cudaStream_t)stream, (const bf16*)x, n8, y);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_cast_f32_bf16_add(const float* x, const void* addend, int64_t n, void* y, vt_stream stream) {
  VT_REQUIRE(n % 8 == 0 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)y & 15) == 0 && ((uintptr_t)addend & 15) == 0, "vt_cast_f32_bf16_add: n %% 8 == 0 and 16-byte alignment required");
  if (n == 0) return VT_OK;
  long long n8 = n / 8, g = (n8 + 255) / 256, cap = (long long)num_sms() * 16;
  launch_pdl(cast_f32_bf16_flat_kernel, (unsigned)(g < cap ? g : cap), 256, 0, (cudaStream_t)stream, x, (const bf16*)addend, n8, (bf16*)y);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}
