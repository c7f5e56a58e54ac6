// Fine-tuning heads around the encoder, fp32, warp-level:
//   SoftDotAttention        vs/agent_models.py:313-357 == tb/agent_models.py:238-274
//     candidate scorer      vs/agent_models.py:382,426  (logit[b,c] = cand[b,c,:] . (W_in h~[b]))
//     context attention     tb/agent_models.py:297-318  (h~ = tanh(W_out [softmax(ctx W_in h) ctx ; h]))
//   decoder2action / question_linear   tb/agent_models.py:297,318 ; vs/agent_models.py:455-459,503
//   masking + CrossEntropy(ignore_index) + argmax   vs/agent.py:396-408 ; tb/agent.py:150-152,316-327
//   BCEWithLogits(pos_weight) masked mean           vs/classifier/agent.py:158-163,493-503
// Batches are a decoder step's (4-64 rows), so nothing here is a tensor-core problem: the kernels are bound by
// streaming the context / candidate / weight tensors once from HBM (a [b,511,512] fp32 context is 1 MB per sample)
// and by launch latency; every reduction runs in a fixed order, so results are deterministic and the argmax is taken
// on fp32 logits (ties to the lowest index, like torch.argmax).
#include "common.cuh"

namespace vt {

void count_launch(int n);

// ---------------------------------------------------------------------------------------------
// small-batch linear:  y[b, n] = act( sum_k [x1 ; x2](b, k) * W[n, k] + bias[n] )
// one warp per output column n and tile of LIN_BT batch rows; W row streamed once per tile with 16-byte loads
// ---------------------------------------------------------------------------------------------
constexpr int LIN_BT = 8;
constexpr int LIN_WARPS = 8;

__device__ __forceinline__ float act_grad_from_y(int act, float y) { return act == 1 ? 1.0f - y * y : 1.0f; }

__global__ void __launch_bounds__(LIN_WARPS * 32)
small_linear_fwd_kernel(const float* __restrict__ x1, long long ldx1, int K1, const float* __restrict__ x2, long long ldx2, int K2,
                        const float* __restrict__ W, long long ldw, const float* __restrict__ bias, int act, float* __restrict__ y,
                        long long ldy, int B, int N, int vec) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * LIN_WARPS + warp;
  const int b0 = blockIdx.y * LIN_BT;
  if (n >= N) return;
  float acc[LIN_BT];
#pragma unroll
  for (int j = 0; j < LIN_BT; ++j) acc[j] = 0.f;
  const float* w = W + (long long)n * ldw;
  for (int seg = 0; seg < 2; ++seg) {
    const float* x = seg ? x2 : x1;
    const long long ldx = seg ? ldx2 : ldx1;
    const int K = seg ? K2 : K1;
    const float* ws = w + (seg ? K1 : 0);
    if (K == 0) continue;
    if (vec) {
      for (int k = lane * 4; k < K; k += 128) {
        const float4 wv = __ldg(reinterpret_cast<const float4*>(ws + k));
#pragma unroll
        for (int j = 0; j < LIN_BT; ++j) {
          if (b0 + j < B) {
            const float4 xv = __ldg(reinterpret_cast<const float4*>(x + (long long)(b0 + j) * ldx + k));
            acc[j] = fmaf(wv.x, xv.x, fmaf(wv.y, xv.y, fmaf(wv.z, xv.z, fmaf(wv.w, xv.w, acc[j]))));
          }
        }
      }
    } else {
      for (int k = lane; k < K; k += 32) {
        const float wv = __ldg(ws + k);
#pragma unroll
        for (int j = 0; j < LIN_BT; ++j)
          if (b0 + j < B) acc[j] = fmaf(wv, __ldg(x + (long long)(b0 + j) * ldx + k), acc[j]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < LIN_BT; ++j) acc[j] = warp_sum(acc[j]);
  if (lane == 0) {
    const float bv = bias ? __ldg(bias + n) : 0.f;
#pragma unroll
    for (int j = 0; j < LIN_BT; ++j) {
      if (b0 + j < B) {
        float v = acc[j] + bv;
        if (act == 1) v = tanhf(v);
        y[(long long)(b0 + j) * ldy + n] = v;
      }
    }
  }
}

// dx[b, k] = sum_n dy'(b, n) W[n, k]  with  dy' = dy * act'(y).
// One CTA per (32 columns k, tile of LIN_BT rows): 8 threads x float4 cover the columns, the 32 thread groups stride the
// output features n (W is read once per row tile, 128 contiguous bytes per group), and the groups' partial sums are added
// through shared memory in group order, so the result is deterministic.  (One thread per output looping over all n left a
// decoder step's dx on 4 CTAs: 2.1 ms for [32, 2052] x [2052, 512].)
// The output features are further cut into chunks of DX_NCH rows of W, one CTA each (blockIdx.z): the LSTMCell's [2048, 2116]
// weight left a decoder step's dx on 65 CTAs that each walked 2048 rows (131 us of 21 MB at 6.5 TB/s = 3 us); with the chunks a
// step's dx runs on ~500 CTAs.  Chunk partials go to a workspace [chunk][row][column] and small_linear_dx_reduce_kernel adds
// them in chunk order: still deterministic.
constexpr int DX_KT = 32, DX_NG = 32, DX_NCH = 256;

__global__ void __launch_bounds__(256)
small_linear_dx_kernel(const float* __restrict__ dy, long long lddy, const float* __restrict__ y, long long ldy, int act,
                       const float* __restrict__ W, long long ldw, int koff, float* __restrict__ dx, long long lddx, int B, int N, int K, int vec) {
  __shared__ float red[DX_NG][LIN_BT][DX_KT + 1];
  const int kq = threadIdx.x & 7, g = threadIdx.x >> 3;
  const int k = blockIdx.x * DX_KT + kq * 4;
  const int b0 = blockIdx.y * LIN_BT;
  float acc[LIN_BT][4];
#pragma unroll
  for (int j = 0; j < LIN_BT; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
  const int kw = max(0, min(4, K - k));
  const int n_lo = blockIdx.z * DX_NCH, n_hi = min(N, n_lo + DX_NCH);
  if (gridDim.z > 1) dx += (long long)blockIdx.z * B * lddx;      // this chunk's slice of the workspace
  for (int n = n_lo + g; n < n_hi; n += DX_NG) {
    float wv[4];
    const float* wp = W + (long long)n * ldw + koff + k;
    if (vec && kw == 4) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(wp));
      wv[0] = t.x; wv[1] = t.y; wv[2] = t.z; wv[3] = t.w;
    } else {
#pragma unroll
      for (int c = 0; c < 4; ++c) wv[c] = c < kw ? __ldg(wp + c) : 0.f;
    }
#pragma unroll
    for (int j = 0; j < LIN_BT; ++j) {
      if (b0 + j < B) {
        float gr = __ldg(dy + (long long)(b0 + j) * lddy + n);
        if (act) gr *= act_grad_from_y(act, __ldg(y + (long long)(b0 + j) * ldy + n));
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[j][c] = fmaf(gr, wv[c], acc[j][c]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < LIN_BT; ++j)
#pragma unroll
    for (int c = 0; c < 4; ++c) red[g][j][kq * 4 + c] = acc[j][c];
  __syncthreads();
  const int j = threadIdx.x >> 5, c = threadIdx.x & 31;   // 8 rows x 32 columns
  const int ko = blockIdx.x * DX_KT + c;
  if (b0 + j < B && ko < K) {
    float sum = 0.f;
#pragma unroll 8
    for (int gg = 0; gg < DX_NG; ++gg) sum += red[gg][j][c];
    dx[(long long)(b0 + j) * lddx + ko] = sum;
  }
}

__global__ void __launch_bounds__(256)
small_linear_dx_reduce_kernel(const float* __restrict__ ws, int chunks, long long slice, int B, int K, float* __restrict__ dx, long long lddx) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= (long long)B * K) return;
  float s = 0.f;
  for (int c = 0; c < chunks; ++c) s += ws[c * slice + i];
  dx[(i / K) * lddx + i % K] = s;
}

// dW[n, koff + k] = sum_b dy'(b, n) x[b, k];  dbias[n] = sum_b dy'(b, n)   (one thread per (n, k); b in order: deterministic)
__global__ void __launch_bounds__(256)
small_linear_dw_kernel(const float* __restrict__ dy, long long lddy, const float* __restrict__ y, long long ldy, int act,
                       const float* __restrict__ x, long long ldx, int koff, float* __restrict__ dW, long long lddw, float* __restrict__ dbias,
                       int B, int N, int K) {
  const int k = blockIdx.x * 256 + threadIdx.x;
  const int n = blockIdx.y;
  if (k >= K) return;
  float acc = 0.f, accb = 0.f;
  for (int b = 0; b < B; ++b) {
    float g = __ldg(dy + (long long)b * lddy + n);
    if (act) g *= act_grad_from_y(act, __ldg(y + (long long)b * ldy + n));
    acc = fmaf(g, __ldg(x + (long long)b * ldx + k), acc);
    accb += g;
  }
  dW[(long long)n * lddw + koff + k] = acc;
  if (dbias && koff == 0 && k == 0) dbias[n] = accb;
}

// ---------------------------------------------------------------------------------------------
// SoftDotAttention core:  logit[b,l] = ctx[b,l,:] . target[b,:]  (masked -> -inf),  attn = softmax_l(logit),
//                         weighted[b,:] = sum_l attn[b,l] ctx[b,l,:]
// Split over the context rows (grid = splits x batch) so that a decoder step's handful of samples still fills the
// SMs: every CTA keeps an online-softmax partial (max, sum, unnormalised weighted row), a second kernel merges the
// partials in a fixed order.  ctx is read from HBM once (the second touch of a row hits L1).
// ---------------------------------------------------------------------------------------------
constexpr int SD_WARPS = 8;
constexpr int SD_THREADS = SD_WARPS * 32;

template <int NV4>  // float4 chunks per lane: D <= 128 * NV4
__global__ void __launch_bounds__(SD_THREADS)
softdot_fwd_partial_kernel(const float* __restrict__ ctx, const float* __restrict__ target, const uint8_t* __restrict__ mask, int L, int D,
                           int rows_per_cta, float* __restrict__ logit, float* __restrict__ part_ms, float* __restrict__ part_w) {
  extern __shared__ __align__(16) float sd_smem[];
  float* s_t = sd_smem;        // [D] target, later the CTA's weighted partial
  __shared__ float s_m[SD_WARPS], s_s[SD_WARPS];
  const int b = blockIdx.y, p = blockIdx.x, P = gridDim.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int k = threadIdx.x; k < D; k += SD_THREADS) s_t[k] = target[(long long)b * D + k];
  __syncthreads();
  const int l0 = p * rows_per_cta, l1 = min(L, l0 + rows_per_cta);
  float m = -INFINITY, s = 0.f;
  float4 w[NV4];
#pragma unroll
  for (int i = 0; i < NV4; ++i) w[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int l = l0 + warp; l < l1; l += SD_WARPS) {
    const float* row = ctx + ((long long)b * L + l) * D;
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < NV4; ++i) {
      const int k = lane * 4 + 128 * i;
      if (k < D) {
        const float4 c = __ldg(reinterpret_cast<const float4*>(row + k));
        const float4 t = *reinterpret_cast<const float4*>(s_t + k);
        dot = fmaf(c.x, t.x, fmaf(c.y, t.y, fmaf(c.z, t.z, fmaf(c.w, t.w, dot))));
      }
    }
    dot = warp_sum(dot);
    if (mask && mask[(long long)b * L + l]) dot = -INFINITY;
    if (lane == 0) logit[(long long)b * L + l] = dot;
    if (dot == -INFINITY) continue;
    float pr = 1.f;
    if (dot > m) {
      const float sc = expf(m - dot);  // 0 on the first row (m = -inf)
      s *= sc;
#pragma unroll
      for (int i = 0; i < NV4; ++i) { w[i].x *= sc; w[i].y *= sc; w[i].z *= sc; w[i].w *= sc; }
      m = dot;
    } else {
      pr = expf(dot - m);
    }
    s += pr;
#pragma unroll
    for (int i = 0; i < NV4; ++i) {
      const int k = lane * 4 + 128 * i;
      if (k < D) {
        const float4 c = __ldg(reinterpret_cast<const float4*>(row + k));
        w[i].x = fmaf(pr, c.x, w[i].x); w[i].y = fmaf(pr, c.y, w[i].y); w[i].z = fmaf(pr, c.z, w[i].z); w[i].w = fmaf(pr, c.w, w[i].w);
      }
    }
  }
  if (lane == 0) { s_m[warp] = m; s_s[warp] = s; }
  __syncthreads();  // also: every warp is done reading s_t
  float M = -INFINITY;
#pragma unroll
  for (int i = 0; i < SD_WARPS; ++i) M = fmaxf(M, s_m[i]);
  float S = 0.f;
#pragma unroll
  for (int i = 0; i < SD_WARPS; ++i) S += s_m[i] == -INFINITY ? 0.f : s_s[i] * expf(s_m[i] - M);
  const float sc = m == -INFINITY ? 0.f : expf(m - M);
  for (int wi = 0; wi < SD_WARPS; ++wi) {  // fixed order: deterministic
    if (warp == wi) {
#pragma unroll
      for (int i = 0; i < NV4; ++i) {
        const int k = lane * 4 + 128 * i;
        if (k < D) {
          float4 a = wi == 0 ? make_float4(0.f, 0.f, 0.f, 0.f) : *reinterpret_cast<float4*>(s_t + k);
          a.x = fmaf(w[i].x, sc, a.x); a.y = fmaf(w[i].y, sc, a.y); a.z = fmaf(w[i].z, sc, a.z); a.w = fmaf(w[i].w, sc, a.w);
          *reinterpret_cast<float4*>(s_t + k) = a;
        }
      }
    }
    __syncthreads();
  }
  float* pw = part_w + ((long long)b * P + p) * D;
  for (int k = threadIdx.x; k < D; k += SD_THREADS) pw[k] = s_t[k];
  if (threadIdx.x == 0) { part_ms[2 * ((long long)b * P + p)] = M; part_ms[2 * ((long long)b * P + p) + 1] = S; }
}

__global__ void __launch_bounds__(SD_THREADS)
softdot_fwd_combine_kernel(const float* __restrict__ logit, const float* __restrict__ part_ms, const float* __restrict__ part_w, int L, int D,
                           int P, float* __restrict__ attn, float* __restrict__ weighted) {
  const int b = blockIdx.x;
  float M = -INFINITY;
  for (int p = 0; p < P; ++p) M = fmaxf(M, part_ms[2 * ((long long)b * P + p)]);
  float S = 0.f;
  for (int p = 0; p < P; ++p) {
    const float mp = part_ms[2 * ((long long)b * P + p)];
    S += mp == -INFINITY ? 0.f : part_ms[2 * ((long long)b * P + p) + 1] * expf(mp - M);
  }
  const float inv = 1.0f / S;  // every row masked: 1/0 -> inf, attn = nan like softmax of all -inf (vs/agent_models.py:345)
  for (int k = threadIdx.x; k < D; k += SD_THREADS) {
    float a = 0.f;
    for (int p = 0; p < P; ++p) {
      const float mp = part_ms[2 * ((long long)b * P + p)];
      if (mp != -INFINITY) a = fmaf(part_w[((long long)b * P + p) * D + k], expf(mp - M), a);
    }
    weighted[(long long)b * D + k] = a * inv;
  }
  for (int l = threadIdx.x; l < L; l += SD_THREADS) attn[(long long)b * L + l] = expf(logit[(long long)b * L + l] - M) * inv;
}

// backward, step 1:  da[b,l] = ctx[b,l,:] . dweighted[b,:]  (+ dattn[b,l])
__global__ void __launch_bounds__(SD_THREADS)
softdot_bwd_da_kernel(const float* __restrict__ ctx, const float* __restrict__ dweighted, const float* __restrict__ dattn, int L, int D,
                      int rows_per_cta, float* __restrict__ da) {
  extern __shared__ __align__(16) float sd_smem[];
  const int b = blockIdx.y, p = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (dweighted) {
    for (int k = threadIdx.x; k < D; k += SD_THREADS) sd_smem[k] = dweighted[(long long)b * D + k];
    __syncthreads();
  }
  const int l0 = p * rows_per_cta, l1 = min(L, l0 + rows_per_cta);
  for (int l = l0 + warp; l < l1; l += SD_WARPS) {
    float dot = 0.f;
    if (dweighted) {
      const float* row = ctx + ((long long)b * L + l) * D;
      for (int k = lane * 4; k < D; k += 128) {
        const float4 c = __ldg(reinterpret_cast<const float4*>(row + k));
        const float4 t = *reinterpret_cast<const float4*>(sd_smem + k);
        dot = fmaf(c.x, t.x, fmaf(c.y, t.y, fmaf(c.z, t.z, fmaf(c.w, t.w, dot))));
      }
      dot = warp_sum(dot);
    }
    if (lane == 0) da[(long long)b * L + l] = dot + (dattn ? dattn[(long long)b * L + l] : 0.f);
  }
}

// backward, step 2:  dlogit = attn (da - sum_l attn da) (+ dlogit_ext off the masked positions);
//   dctx[b,l,:] = dlogit target + attn dweighted;   dtarget partial = sum_l dlogit ctx[b,l,:]
template <int NV4>
__global__ void __launch_bounds__(SD_THREADS)
softdot_bwd_main_kernel(const float* __restrict__ ctx, const float* __restrict__ target, const uint8_t* __restrict__ mask,
                        const float* __restrict__ attn, const float* __restrict__ da, const float* __restrict__ dweighted,
                        const float* __restrict__ dlogit_ext, int L, int D, int rows_per_cta, float* __restrict__ dctx,
                        float* __restrict__ part_dt) {
  extern __shared__ __align__(16) float sd_smem[];
  float* s_t = sd_smem;       // [D] target, later the dtarget partial
  float* s_dw = sd_smem + D;  // [D] dweighted
  __shared__ float s_red[SD_WARPS];
  const int b = blockIdx.y, p = blockIdx.x, P = gridDim.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int k = threadIdx.x; k < D; k += SD_THREADS) {
    s_t[k] = target[(long long)b * D + k];
    s_dw[k] = dweighted ? dweighted[(long long)b * D + k] : 0.f;
  }
  float c = 0.f;
  if (da) {
    for (int l = threadIdx.x; l < L; l += SD_THREADS) {
      const float a = attn[(long long)b * L + l];
      if (a != 0.f) c = fmaf(a, da[(long long)b * L + l], c);
    }
    c = warp_sum(c);
    if (lane == 0) s_red[warp] = c;
  }
  __syncthreads();
  if (da) {
    c = 0.f;
#pragma unroll
    for (int i = 0; i < SD_WARPS; ++i) c += s_red[i];
  }
  const int l0 = p * rows_per_cta, l1 = min(L, l0 + rows_per_cta);
  float4 w[NV4];
#pragma unroll
  for (int i = 0; i < NV4; ++i) w[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int l = l0 + warp; l < l1; l += SD_WARPS) {
    const long long bl = (long long)b * L + l;
    const float a = attn[bl];
    float dl = (da && a != 0.f) ? a * (da[bl] - c) : 0.f;
    if (dlogit_ext && !(mask && mask[bl])) dl += dlogit_ext[bl];
    const float* row = ctx + bl * D;
#pragma unroll
    for (int i = 0; i < NV4; ++i) {
      const int k = lane * 4 + 128 * i;
      if (k < D) {
        const float4 cv = __ldg(reinterpret_cast<const float4*>(row + k));
        w[i].x = fmaf(dl, cv.x, w[i].x); w[i].y = fmaf(dl, cv.y, w[i].y); w[i].z = fmaf(dl, cv.z, w[i].z); w[i].w = fmaf(dl, cv.w, w[i].w);
        if (dctx) {
          const float4 t = *reinterpret_cast<const float4*>(s_t + k);
          const float4 g = *reinterpret_cast<const float4*>(s_dw + k);
          *reinterpret_cast<float4*>(dctx + bl * D + k) =
              make_float4(fmaf(dl, t.x, a * g.x), fmaf(dl, t.y, a * g.y), fmaf(dl, t.z, a * g.z), fmaf(dl, t.w, a * g.w));
        }
      }
    }
  }
  __syncthreads();
  for (int wi = 0; wi < SD_WARPS; ++wi) {
    if (warp == wi) {
#pragma unroll
      for (int i = 0; i < NV4; ++i) {
        const int k = lane * 4 + 128 * i;
        if (k < D) {
          float4 acc = wi == 0 ? make_float4(0.f, 0.f, 0.f, 0.f) : *reinterpret_cast<float4*>(s_t + k);
          acc.x += w[i].x; acc.y += w[i].y; acc.z += w[i].z; acc.w += w[i].w;
          *reinterpret_cast<float4*>(s_t + k) = acc;
        }
      }
    }
    __syncthreads();
  }
  float* pd = part_dt + ((long long)b * P + p) * D;
  for (int k = threadIdx.x; k < D; k += SD_THREADS) pd[k] = s_t[k];
}

__global__ void __launch_bounds__(256)
sum_partials_kernel(const float* __restrict__ part, int P, int D, float* __restrict__ out) {
  const int b = blockIdx.y, k = blockIdx.x * 256 + threadIdx.x;
  if (k >= D) return;
  float a = 0.f;
  for (int p = 0; p < P; ++p) a += part[((long long)b * P + p) * D + k];
  out[(long long)b * D + k] = a;
}

// ---------------------------------------------------------------------------------------------
// Action-logit masking + CrossEntropyLoss(ignore_index, reduction = mean) + argmax over [B, C] fp32 logits:
//   columns >= valid_len[b] -> -inf (utils.length2mask + masked_fill_, vs/agent.py:396-403);
//   column forbid_col (model_actions.index("forward") = 4, tb/agent.py:71-80) -> -inf where forbid[b] (no navigable location:
//   "forward" is impossible, tb/agent.py:316-318).
// One CTA (a decoder step has at most a few hundred rows); one warp per row; out3 = {loss, n_valid, sum of row losses}.
// ---------------------------------------------------------------------------------------------
constexpr int HCE_WARPS = 16;

__global__ void __launch_bounds__(HCE_WARPS * 32)
head_ce_fwd_kernel(const float* __restrict__ logits, long long ld, int B, int C, const long long* __restrict__ valid_len,
                   const uint8_t* __restrict__ forbid0, int fcol, int empty_zero, const long long* __restrict__ target, long long ignore_index,
                   float* __restrict__ masked, long long ldm, float* __restrict__ lse_out, long long* __restrict__ argmax,
                   float* __restrict__ out3) {
  __shared__ float s_loss[HCE_WARPS];
  __shared__ int s_cnt[HCE_WARPS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float loss_acc = 0.f;
  int cnt = 0;
  for (int b = warp; b < B; b += HCE_WARPS) {
    const int vl = valid_len ? (int)min((long long)C, valid_len[b]) : C;
    const bool f0 = forbid0 && forbid0[b];
    float mx = -INFINITY;
    int arg = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      float v = logits[(long long)b * ld + c];
      if (c >= vl || (f0 && c == fcol)) v = -INFINITY;
      masked[(long long)b * ldm + c] = v;
      if (arg == 0x7fffffff || v > mx) { mx = v; arg = c; }  // c ascends within a lane: ties keep the lowest index
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
    }
    float se = 0.f;
    for (int c = lane; c < C; c += 32) {
      float v = logits[(long long)b * ld + c];
      if (c >= vl || (f0 && c == fcol)) v = -INFINITY;
      se += expf(v - mx);
    }
    se = warp_sum(se);
    const float lse = mx + logf(se);
    const long long t = target[b];
    if (lane == 0) {
      lse_out[b] = lse;
      argmax[b] = arg;
      if (t != ignore_index) {
        float vt_ = (t >= 0 && t < C) ? logits[(long long)b * ld + t] : NAN;
        if (t >= vl || (f0 && t == fcol)) vt_ = -INFINITY;
        loss_acc += lse - vt_;
        cnt += 1;
      }
    }
  }
  if (lane == 0) { s_loss[warp] = loss_acc; s_cnt[warp] = cnt; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f;
    int n = 0;
    for (int i = 0; i < HCE_WARPS; ++i) { tot += s_loss[i]; n += s_cnt[i]; }
    // 0/0 = nan when every row is ignored, like torch -- unless the caller is a rollout, which the reference leaves before
    // such a step (`if ended.all(): break`, vs/agent.py:466-468): the step then adds nothing
    out3[0] = (n == 0 && empty_zero) ? 0.f : tot / (float)n;
    out3[1] = (float)n;
    out3[2] = tot;
  }
}

// dlogits[b,c] = g / n_valid * (softmax(masked)[b,c] - [c == target[b]]) on the rows that are not ignored, else 0
__global__ void __launch_bounds__(256)
head_ce_bwd_kernel(const float* __restrict__ masked, long long ldm, int B, int C, const long long* __restrict__ target, long long ignore_index,
                   const float* __restrict__ lse, const float* __restrict__ out3, const float* __restrict__ g, float* __restrict__ dlogits,
                   long long ldd) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= (long long)B * C) return;
  const int b = (int)(i / C), c = (int)(i % C);
  const long long t = target[b];
  float v = 0.f;
  if (t != ignore_index) {
    const float scale = __ldg(g) / out3[1];
    v = (expf(masked[(long long)b * ldm + c] - lse[b]) - (c == t ? 1.f : 0.f)) * scale;
  }
  dlogits[(long long)b * ldd + c] = v;
}

// BCEWithLogitsLoss(pos_weight, reduction = none) * valid, summed / max(n_valid, 1)  (vs/classifier/agent.py:158-163,493-503:
// sum / count, or 0 when nothing is valid).  One CTA; out2 = {loss, n_valid}.
__global__ void __launch_bounds__(256)
bce_logits_fwd_kernel(const float* __restrict__ q, const float* __restrict__ target, const uint8_t* __restrict__ valid, int B, float pos_weight,
                      float* __restrict__ per_row, float* __restrict__ out2) {
  __shared__ float s_l[8];
  __shared__ int s_n[8];
  float acc = 0.f;
  int n = 0;
  for (int b = threadIdx.x; b < B; b += 256) {
    const float x = q[b], y = target[b];
    const float w = 1.0f + (pos_weight - 1.0f) * y;
    const float l = (1.0f - y) * x + w * (log1pf(expf(-fabsf(x))) + fmaxf(-x, 0.f));
    per_row[b] = l;
    if (!valid || valid[b]) { acc += l; n += 1; }
  }
  // fixed-order reduction: lanes, then warps
  for (int o = 16; o > 0; o >>= 1) { acc += __shfl_down_sync(0xffffffffu, acc, o); n += __shfl_down_sync(0xffffffffu, n, o); }
  if ((threadIdx.x & 31) == 0) { s_l[threadIdx.x >> 5] = acc; s_n[threadIdx.x >> 5] = n; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = 0.f;
    int cnt = 0;
    for (int i = 0; i < 8; ++i) { tot += s_l[i]; cnt += s_n[i]; }
    out2[0] = cnt > 0 ? tot / (float)cnt : 0.f;
    out2[1] = (float)cnt;
  }
}

__global__ void __launch_bounds__(256)
bce_logits_bwd_kernel(const float* __restrict__ q, const float* __restrict__ target, const uint8_t* __restrict__ valid, int B, float pos_weight,
                      const float* __restrict__ out2, const float* __restrict__ g, float* __restrict__ dq) {
  const int b = blockIdx.x * 256 + threadIdx.x;
  if (b >= B) return;
  float v = 0.f;
  if ((!valid || valid[b]) && out2[1] > 0.f) {
    const float x = q[b], y = target[b];
    const float w = 1.0f + (pos_weight - 1.0f) * y;
    const float sig_neg = 1.0f / (1.0f + expf(x));  // sigmoid(-x)
    v = ((1.0f - y) - w * sig_neg) * __ldg(g) / out2[1];
  }
  dq[b] = v;
}

}  // namespace vt

using namespace vt;

static int pick_splits(int64_t B, int64_t L) {
  int64_t P = (2 * (int64_t)num_sms() + B - 1) / B;
  const int64_t maxP = (L + SD_WARPS - 1) / SD_WARPS;
  if (P > maxP) P = maxP;
  if (P > 64) P = 64;
  if (P < 1) P = 1;
  return (int)P;
}

extern "C" int64_t vt_softdot_splits(int64_t B, int64_t L) { return pick_splits(B, L); }

extern "C" int vt_small_linear_fwd(const float* x1, int64_t ldx1, int64_t K1, const float* x2, int64_t ldx2, int64_t K2, const float* W,
                                   int64_t ldw, const float* bias, int act, float* y, int64_t ldy, int64_t B, int64_t N, vt_stream stream) {
  VT_REQUIRE(x1 && W && y && B >= 0 && N > 0 && K1 > 0 && K2 >= 0 && (K2 == 0 || x2), "vt_small_linear_fwd: bad arguments");
  VT_REQUIRE(act == 0 || act == 1, "vt_small_linear_fwd: act must be 0 (none) or 1 (tanh)");
  VT_REQUIRE(ldw >= K1 + K2 && ldx1 >= K1 && ldx2 >= K2 && ldy >= N, "vt_small_linear_fwd: leading dimension smaller than the row");
  if (B == 0) return VT_OK;
  const int vec = (K1 % 4 == 0 && K2 % 4 == 0 && ldw % 4 == 0 && ldx1 % 4 == 0 && (K2 == 0 || ldx2 % 4 == 0) && ((uintptr_t)x1 & 15) == 0 &&
                   ((uintptr_t)W & 15) == 0 && (K2 == 0 || ((uintptr_t)x2 & 15) == 0)) ? 1 : 0;
  dim3 grid((unsigned)((N + LIN_WARPS - 1) / LIN_WARPS), (unsigned)((B + LIN_BT - 1) / LIN_BT));
  small_linear_fwd_kernel<<<grid, LIN_WARPS * 32, 0, (cudaStream_t)stream>>>(x1, ldx1, (int)K1, x2, ldx2, (int)K2, W, ldw, bias, act, y, ldy, (int)B,
                                                                              (int)N, vec);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int64_t vt_small_linear_bwd_workspace(int64_t B, int64_t N, int64_t K1, int64_t K2) {
  const int64_t chunks = (N + DX_NCH - 1) / DX_NCH;
  return chunks <= 1 ? 0 : chunks * B * (K1 > K2 ? K1 : K2);
}

extern "C" int vt_small_linear_bwd(const float* dy, int64_t lddy, const float* y, int64_t ldy, int act, const float* x1, int64_t ldx1, int64_t K1,
                                   const float* x2, int64_t ldx2, int64_t K2, const float* W, int64_t ldw, float* dx1, int64_t lddx1, float* dx2,
                                   int64_t lddx2, float* dW, int64_t lddw, float* dbias, int64_t B, int64_t N, float* workspace, vt_stream stream) {
  VT_REQUIRE(dy && x1 && W && B >= 0 && N > 0 && K1 > 0 && K2 >= 0, "vt_small_linear_bwd: bad arguments");
  VT_REQUIRE(N <= DX_NCH || workspace || (!dx1 && !dx2), "vt_small_linear_bwd: more than %d output features need the workspace (vt_small_linear_bwd_workspace)", DX_NCH);
  VT_REQUIRE(act == 0 || (act == 1 && y), "vt_small_linear_bwd: tanh backward needs the forward output");
  if (B == 0) return VT_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int launches = 0;
  for (int seg = 0; seg < 2; ++seg) {
    const int64_t K = seg ? K2 : K1;
    if (K == 0) continue;
    float* dx = seg ? dx2 : dx1;
    const int koff = seg ? (int)K1 : 0;
    if (dx) {
      const int chunks = (int)((N + DX_NCH - 1) / DX_NCH);
      dim3 grid((unsigned)((K + DX_KT - 1) / DX_KT), (unsigned)((B + LIN_BT - 1) / LIN_BT), (unsigned)chunks);
      const int vec = (((uintptr_t)W & 15) == 0 && ldw % 4 == 0 && koff % 4 == 0) ? 1 : 0;
      const long long lddx = seg ? lddx2 : lddx1;
      if (chunks == 1) {
        small_linear_dx_kernel<<<grid, 256, 0, st>>>(dy, lddy, y, ldy, act, W, ldw, koff, dx, lddx, (int)B, (int)N, (int)K, vec);
        ++launches;
      } else {   // partials [chunk][B][K] (row pitch K), then the fixed-order sum
        small_linear_dx_kernel<<<grid, 256, 0, st>>>(dy, lddy, y, ldy, act, W, ldw, koff, workspace, K, (int)B, (int)N, (int)K, vec);
        small_linear_dx_reduce_kernel<<<(unsigned)((B * K + 255) / 256), 256, 0, st>>>(workspace, chunks, B * K, (int)B, (int)K, dx, lddx);
        launches += 2;
      }
    }
    if (dW) {
      dim3 grid((unsigned)((K + 255) / 256), (unsigned)N);
      small_linear_dw_kernel<<<grid, 256, 0, st>>>(dy, lddy, y, ldy, act, seg ? x2 : x1, seg ? ldx2 : ldx1, koff, dW, lddw, dbias, (int)B, (int)N, (int)K);
      ++launches;
    }
  }
  count_launch(launches);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

static int softdot_check(const char* who, int64_t B, int64_t L, int64_t D, const void* ctx) {
  VT_REQUIRE(B > 0 && L > 0 && D > 0 && B <= 65535, "%s: bad sizes B=%lld L=%lld D=%lld", who, (long long)B, (long long)L, (long long)D);
  VT_REQUIRE(D % 4 == 0 && D <= 2176, "%s: feature width %lld outside the supported envelope (multiple of 4, <= 2176: 512 / 2048 / 2052)", who, (long long)D);
  VT_REQUIRE(((uintptr_t)ctx & 15) == 0, "%s: context must be 16-byte aligned", who);
  return VT_OK;
}

extern "C" int vt_softdot_fwd(const float* ctx, const float* target, const uint8_t* mask, int64_t B, int64_t L, int64_t D, float* logit, float* attn,
                              float* weighted, float* workspace, vt_stream stream) {
  int rc = softdot_check("vt_softdot_fwd", B, L, D, ctx);
  if (rc) return rc;
  VT_REQUIRE(target && logit && attn && weighted && workspace, "vt_softdot_fwd: null pointer");
  const int P = pick_splits(B, L);
  const int rpc = (int)((L + P - 1) / P);
  float* part_ms = workspace;                      // [B, P, 2]
  float* part_w = workspace + (size_t)B * P * 2;   // [B, P, D]
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = (size_t)D * sizeof(float);
  dim3 grid((unsigned)P, (unsigned)B);
  if (D <= 512) softdot_fwd_partial_kernel<4><<<grid, SD_THREADS, smem, st>>>(ctx, target, mask, (int)L, (int)D, rpc, logit, part_ms, part_w);
  else softdot_fwd_partial_kernel<17><<<grid, SD_THREADS, smem, st>>>(ctx, target, mask, (int)L, (int)D, rpc, logit, part_ms, part_w);
  softdot_fwd_combine_kernel<<<(unsigned)B, SD_THREADS, 0, st>>>(logit, part_ms, part_w, (int)L, (int)D, P, attn, weighted);
  count_launch(2);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_softdot_bwd(const float* ctx, const float* target, const uint8_t* mask, const float* attn, const float* dweighted,
                              const float* dattn, const float* dlogit_ext, int64_t B, int64_t L, int64_t D, float* dctx, float* dtarget,
                              float* workspace, vt_stream stream) {
  int rc = softdot_check("vt_softdot_bwd", B, L, D, ctx);
  if (rc) return rc;
  VT_REQUIRE(target && attn && dtarget && workspace, "vt_softdot_bwd: null pointer");
  const int P = pick_splits(B, L);
  const int rpc = (int)((L + P - 1) / P);
  float* da = workspace;                          // [B, L]
  float* part_dt = workspace + (size_t)B * L;     // [B, P, D]
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid((unsigned)P, (unsigned)B);
  const bool need_da = dweighted || dattn;
  int launches = 2;
  if (need_da) {
    softdot_bwd_da_kernel<<<grid, SD_THREADS, (size_t)D * sizeof(float), st>>>(ctx, dweighted, dattn, (int)L, (int)D, rpc, da);
    ++launches;
  }
  const size_t smem = 2 * (size_t)D * sizeof(float);
  if (D <= 512)
    softdot_bwd_main_kernel<4><<<grid, SD_THREADS, smem, st>>>(ctx, target, mask, attn, need_da ? da : nullptr, dweighted, dlogit_ext, (int)L, (int)D, rpc,
                                                               dctx, part_dt);
  else
    softdot_bwd_main_kernel<17><<<grid, SD_THREADS, smem, st>>>(ctx, target, mask, attn, need_da ? da : nullptr, dweighted, dlogit_ext, (int)L, (int)D, rpc,
                                                                dctx, part_dt);
  sum_partials_kernel<<<dim3((unsigned)((D + 255) / 256), (unsigned)B), 256, 0, st>>>(part_dt, P, (int)D, dtarget);
  count_launch(launches);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_head_ce_fwd(const float* logits, int64_t ld, int64_t B, int64_t C, const int64_t* valid_len, const uint8_t* forbid0,
                              int64_t forbid_col, int empty_is_zero, const int64_t* target, int64_t ignore_index, float* masked, int64_t ldm, float* lse, int64_t* argmax, float* out3,
                              vt_stream stream) {
  VT_REQUIRE(logits && target && masked && lse && argmax && out3 && B > 0 && C > 0 && ld >= C && ldm >= C, "vt_head_ce_fwd: bad arguments");
  head_ce_fwd_kernel<<<1, HCE_WARPS * 32, 0, (cudaStream_t)stream>>>(logits, ld, (int)B, (int)C, (const long long*)valid_len, forbid0, (int)forbid_col, empty_is_zero,
                                                                      (const long long*)target, (long long)ignore_index, masked, ldm, lse,
                                                                      (long long*)argmax, out3);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_head_ce_bwd(const float* masked, int64_t ldm, int64_t B, int64_t C, const int64_t* target, int64_t ignore_index, const float* lse,
                              const float* out3, const float* g, float* dlogits, int64_t ldd, vt_stream stream) {
  VT_REQUIRE(masked && target && lse && out3 && g && dlogits && B > 0 && C > 0, "vt_head_ce_bwd: bad arguments");
  head_ce_bwd_kernel<<<(unsigned)((B * C + 255) / 256), 256, 0, (cudaStream_t)stream>>>(masked, ldm, (int)B, (int)C, (const long long*)target,
                                                                                        (long long)ignore_index, lse, out3, g, dlogits, ldd);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_bce_logits_fwd(const float* q, const float* target, const uint8_t* valid, int64_t B, float pos_weight, float* per_row, float* out2,
                                 vt_stream stream) {
  VT_REQUIRE(q && target && per_row && out2 && B > 0, "vt_bce_logits_fwd: bad arguments");
  bce_logits_fwd_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(q, target, valid, (int)B, pos_weight, per_row, out2);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_bce_logits_bwd(const float* q, const float* target, const uint8_t* valid, int64_t B, float pos_weight, const float* out2,
                                 const float* g, float* dq, vt_stream stream) {
  VT_REQUIRE(q && target && out2 && g && dq && B > 0, "vt_bce_logits_bwd: bad arguments");
  bce_logits_bwd_kernel<<<(unsigned)((B + 255) / 256), 256, 0, (cudaStream_t)stream>>>(q, target, valid, (int)B, pos_weight, out2, g, dq);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}
