// HBM-bound row kernels: one warp per 768-wide row, 16-byte vector loads/stores, fp32 statistics.
//   * weight cast fp32 -> bf16 (optionally also transposed)
//   * image-feature pad/cast (img_feats | location embeddings -> one bf16 operand row)
//   * embedding gather + sum + LayerNorm + dropout (forward) and its backward (scatter-add)
//   * LayerNorm forward (+dropout, + row remap) and backward (+ dropout-masked copy, + dgamma /
//     dbeta / dbias partial sums)
//   * column sums (bias gradients), partial-sum reduction, row gather / scatter-add
// BertLayerNorm is the TF flavour: biased variance, eps inside the sqrt (hf/modeling_bert.py:227-240).
#include <mutex>

#include "common.cuh"

namespace vt {

void count_launch(int n);

constexpr int H = 768;           // hidden size of the supported envelope (bert-base)
constexpr int CH = H / (32 * 8); // 8-element chunks per lane = 3
constexpr int ROW_WARPS = 8;     // warps (rows in flight) per CTA

static int ln_bwd_grid(long long rows) {
  long long g = (rows + ROW_WARPS - 1) / ROW_WARPS;
  long long cap = (long long)num_sms();   // exactly one resident CTA per SM: the fewest partial rows to reduce afterwards
  return (int)(g < cap ? (g < 1 ? 1 : g) : cap);
}

static int row_grid(long long rows) {
  long long g = (rows + ROW_WARPS - 1) / ROW_WARPS;
  long long cap = (long long)num_sms() * 8;
  return (int)(g < cap ? (g < 1 ? 1 : g) : cap);
}

// lane l owns elements [8*(l+32c), 8*(l+32c)+8) for c = 0..2: a warp instruction touches 512
// contiguous bytes of bf16 (1 KiB of fp32).
__device__ __forceinline__ int chunk_col(int lane, int c) { return 8 * (lane + 32 * c); }

__device__ __forceinline__ void load_row_bf16(const bf16* row, int lane, float (*x)[8]) {
#pragma unroll
  for (int c = 0; c < CH; ++c) unpack8(__ldg(reinterpret_cast<const uint4*>(row + chunk_col(lane, c))), x[c]);
}
__device__ __forceinline__ void store_row_bf16(bf16* row, int lane, const float (*x)[8]) {
#pragma unroll
  for (int c = 0; c < CH; ++c) *reinterpret_cast<uint4*>(row + chunk_col(lane, c)) = pack8(x[c]);
}
__device__ __forceinline__ void load_vec_f32(const float* v, int lane, float (*x)[8]) {
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    float4 a = __ldg(reinterpret_cast<const float4*>(v + chunk_col(lane, c)));
    float4 b = __ldg(reinterpret_cast<const float4*>(v + chunk_col(lane, c) + 4));
    x[c][0] = a.x; x[c][1] = a.y; x[c][2] = a.z; x[c][3] = a.w;
    x[c][4] = b.x; x[c][5] = b.y; x[c][6] = b.z; x[c][7] = b.w;
  }
}
__device__ __forceinline__ void add_vec_f32(const float* v, int lane, float (*x)[8]) {
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    float4 a = __ldg(reinterpret_cast<const float4*>(v + chunk_col(lane, c)));
    float4 b = __ldg(reinterpret_cast<const float4*>(v + chunk_col(lane, c) + 4));
    x[c][0] += a.x; x[c][1] += a.y; x[c][2] += a.z; x[c][3] += a.w;
    x[c][4] += b.x; x[c][5] += b.y; x[c][6] += b.z; x[c][7] += b.w;
  }
}

// mean and 1/sqrt(var + eps) of one row held across the warp (two-pass, as the reference computes it)
__device__ __forceinline__ void row_stats(const float (*x)[8], float eps, float& mean, float& rstd) {
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < CH; ++c)
#pragma unroll
    for (int j = 0; j < 8; ++j) s += x[c][j];
  mean = warp_sum(s) * (1.0f / H);
  float v = 0.f;
#pragma unroll
  for (int c = 0; c < CH; ++c)
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float d = x[c][j] - mean;
      v += d * d;
    }
  rstd = rsqrtf(warp_sum(v) * (1.0f / H) + eps);
}

__device__ __forceinline__ void apply_dropout_row(const DropoutCfg& d, long long m, int lane, float (*x)[8]) {
  if (!d.thresh16) return;
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    float mult[8];
    dropout_mult8(d, (uint64_t)m * H + chunk_col(lane, c), mult);
#pragma unroll
    for (int j = 0; j < 8; ++j) x[c][j] *= mult[j];
  }
}

__device__ __forceinline__ long long remap_row(long long m, long long rin, long long rout, long long off) {
  return rin > 0 ? (m / rin) * rout + off + (m % rin) : m;
}

// sum the per-warp accumulators of a CTA and write one partial row: part[blockIdx.x][slot][H]
template <int NACC>
__device__ __forceinline__ void flush_partials(float (*acc)[CH][8], float* part, int lane, int warp) {
  __shared__ float red[ROW_WARPS][H];
#pragma unroll
  for (int a = 0; a < NACC; ++a) {
    __syncthreads();
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) red[warp][chunk_col(lane, c) + j] = acc[a][c][j];
    __syncthreads();
    for (int col = threadIdx.x; col < H; col += blockDim.x) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < ROW_WARPS; ++w) s += red[w][col];
      part[((long long)blockIdx.x * NACC + a) * H + col] = s;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// LayerNorm forward:  y[remap(m)] = dropout(LN(x[m]) * gamma + beta)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ROW_WARPS * 32)
ln_fwd_kernel(const bf16* __restrict__ x, long long ldx, const float* __restrict__ gamma, const float* __restrict__ beta,
              float eps, long long rows, bf16* __restrict__ y, long long ldy, DropoutCfg drop, long long rin, long long rout,
              long long roff, const int32_t* __restrict__ limit) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  drop = resolve_dropout(drop);
  rows = limited_rows(rows, limit);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float g[CH][8], b[CH][8];
  load_vec_f32(gamma, lane, g);
  load_vec_f32(beta, lane, b);
  // the next row is requested (packed, 12 registers) before the current one is reduced: twice the bytes in flight per warp
  const long long stride = (long long)gridDim.x * ROW_WARPS;
  long long m = (long long)blockIdx.x * ROW_WARPS + warp;
  uint4 xq[CH];
  if (m < rows) {
#pragma unroll
    for (int c = 0; c < CH; ++c) xq[c] = __ldg(reinterpret_cast<const uint4*>(x + m * ldx + chunk_col(lane, c)));
  }
  for (; m < rows; m += stride) {
    float v[CH][8];
#pragma unroll
    for (int c = 0; c < CH; ++c) unpack8(xq[c], v[c]);
    if (m + stride < rows) {
#pragma unroll
      for (int c = 0; c < CH; ++c) xq[c] = __ldg(reinterpret_cast<const uint4*>(x + (m + stride) * ldx + chunk_col(lane, c)));
    }
    float mean, rstd;
    row_stats(v, eps, mean, rstd);
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) v[c][j] = (v[c][j] - mean) * rstd * g[c][j] + b[c][j];
    apply_dropout_row(drop, m, lane, v);
    store_row_bf16(y + remap_row(m, rin, rout, roff) * ldy, lane, v);
  }
}

// ---------------------------------------------------------------------------------------------
// LayerNorm backward.
//   forward site A (encoder blocks):  x = dropout_out(h) + res ; y = LN(x)
//       dx  = LN'(dy)                (-> residual branch)      dxd = dx * mask_out   (-> GEMM branch)
//   forward site B (image rows):      y = dropout_in(LN(x))    dy is masked with mask_in first.
//   partial sums per CTA: [dgamma, dbeta, colsum(dxd)]  (the last one is the bias gradient of the
//   GEMM that produced h).
// ---------------------------------------------------------------------------------------------
// One CTA (8 warps) per SM at ~230 registers: the next row's x and dy are prefetched (packed, 24 registers)
// while the current row is processed, which doubles the bytes in flight per warp.
// Tried and measured equal, not kept (same-box A/B, tools/r2_call24.sh): 16 warps per CTA with the column sums in per-warp shared-memory
// slices (42.1 vs 42.6 us incl. the partial reduction) -- the kernel is not short of rows in flight; under ncu it moves 4.1 TB/s
// over its 26 us, the rest of the case's time is the launch ramp and the reduction of the partials.
// NACC = 3 also accumulates the column sums of dx (the bias gradient of the Linear in front of the LayerNorm); NACC = 2 leaves
// that to the weight-gradient GEMM's colsum side product and spends the 24 registers on a second row of prefetch (DEPTH = 2).
template <int NACC, int DEPTH>
__global__ void __launch_bounds__(ROW_WARPS * 32, 1)
ln_bwd_kernel(const bf16* __restrict__ dy, long long lddy, const bf16* __restrict__ x, long long ldx,
              const float* __restrict__ gamma, float eps, long long rows, bf16* __restrict__ dx, long long lddx,
              bf16* __restrict__ dxd, long long lddxd, float* __restrict__ partials, DropoutCfg drop_in, DropoutCfg drop_out,
              long long rin, long long rout, long long roff, const int32_t* __restrict__ limit) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  drop_in = resolve_dropout(drop_in);
  drop_out = resolve_dropout(drop_out);
  rows = limited_rows(rows, limit);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float g[CH][8];
  load_vec_f32(gamma, lane, g);
  float acc[NACC][CH][8];
#pragma unroll
  for (int a = 0; a < NACC; ++a)
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[a][c][j] = 0.f;
  const long long stride = (long long)gridDim.x * ROW_WARPS;
  long long m = (long long)blockIdx.x * ROW_WARPS + warp;
  uint4 xq[DEPTH][CH], dq[DEPTH][CH];
#pragma unroll
  for (int k = 0; k < DEPTH; ++k) {
    const long long mk = m + k * stride;
    if (mk < rows) {
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        xq[k][c] = __ldg(reinterpret_cast<const uint4*>(x + mk * ldx + chunk_col(lane, c)));
        dq[k][c] = __ldg(reinterpret_cast<const uint4*>(dy + remap_row(mk, rin, rout, roff) * lddy + chunk_col(lane, c)));
      }
    }
  }
  for (; m < rows; m += stride) {
    float v[CH][8], d[CH][8];
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      unpack8(xq[0][c], v[c]);
      unpack8(dq[0][c], d[c]);
    }
#pragma unroll
    for (int k = 0; k + 1 < DEPTH; ++k)
#pragma unroll
      for (int c = 0; c < CH; ++c) { xq[k][c] = xq[k + 1][c]; dq[k][c] = dq[k + 1][c]; }
    const long long mn = m + DEPTH * stride;
    if (mn < rows) {
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        xq[DEPTH - 1][c] = __ldg(reinterpret_cast<const uint4*>(x + mn * ldx + chunk_col(lane, c)));
        dq[DEPTH - 1][c] = __ldg(reinterpret_cast<const uint4*>(dy + remap_row(mn, rin, rout, roff) * lddy + chunk_col(lane, c)));
      }
    }
    apply_dropout_row(drop_in, m, lane, d);
    float mean, rstd;
    row_stats(v, eps, mean, rstd);
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float xh = (v[c][j] - mean) * rstd;
        acc[0][c][j] += d[c][j] * xh;
        acc[1][c][j] += d[c][j];
        float dg = d[c][j] * g[c][j];
        s1 += dg;
        s2 += dg * xh;
        v[c][j] = xh;
        d[c][j] = dg;
      }
    s1 = warp_sum(s1) * (1.0f / H);
    s2 = warp_sum(s2) * (1.0f / H);
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) d[c][j] = rstd * (d[c][j] - s1 - v[c][j] * s2);
    if (dx) store_row_bf16(dx + m * lddx, lane, d);
    if (dxd) {
      apply_dropout_row(drop_out, m, lane, d);
      store_row_bf16(dxd + m * lddxd, lane, d);
    }
    if (NACC > 2) {
#pragma unroll
      for (int c = 0; c < CH; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[NACC - 1][c][j] += d[c][j];
    }
  }
  flush_partials<NACC>(acc, partials, lane, warp);
}

// ---------------------------------------------------------------------------------------------
// Embeddings: y[s*S_out + t] = dropout(LN(word[ids] + pos[p] + type[tt]))   (hf/modeling_bert.py:256-271)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ROW_WARPS * 32)
embed_ln_fwd_kernel(const long long* __restrict__ ids, const long long* __restrict__ type_ids, const long long* __restrict__ pos_ids,
                    const float* __restrict__ word, const float* __restrict__ pos, const float* __restrict__ type,
                    const float* __restrict__ gamma, const float* __restrict__ beta, float eps, long long batch, long long T,
                    long long S_out, bf16* __restrict__ y, DropoutCfg drop, const int32_t* __restrict__ row_dst) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  drop = resolve_dropout(drop);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float g[CH][8], b[CH][8];
  load_vec_f32(gamma, lane, g);
  load_vec_f32(beta, lane, b);
  const long long rows = batch * T;
  for (long long m = (long long)blockIdx.x * ROW_WARPS + warp; m < rows; m += (long long)gridDim.x * ROW_WARPS) {
    const long long t = m % T, s = m / T;
    long long orow = s * S_out + t;
    if (row_dst) {                       // packed rows (vt_pack_plan): where the row lives, or nowhere
      orow = row_dst[orow];
      if (orow < 0) continue;
    }
    const long long id = ids[m], tt = type_ids ? type_ids[m] : 0, p = pos_ids ? pos_ids[m] : t;
    float v[CH][8];
    load_vec_f32(word + id * H, lane, v);
    add_vec_f32(pos + p * H, lane, v);
    add_vec_f32(type + tt * H, lane, v);
    float mean, rstd;
    row_stats(v, eps, mean, rstd);
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) v[c][j] = (v[c][j] - mean) * rstd * g[c][j] + b[c][j];
    apply_dropout_row(drop, m, lane, v);
    store_row_bf16(y + orow * H, lane, v);
  }
}

__device__ __forceinline__ void red_add_row(float* dst, int lane, const float (*d)[8]) {
#pragma unroll
  for (int c = 0; c < CH; ++c) {
    float* p = dst + chunk_col(lane, c);
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(d[c][0]), "f"(d[c][1]), "f"(d[c][2]), "f"(d[c][3]) : "memory");
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p + 4), "f"(d[c][4]), "f"(d[c][5]), "f"(d[c][6]), "f"(d[c][7]) : "memory");
  }
}

// backward: scatter-adds into the fp32 word / position / token-type gradient tables.  Token types
// 0 and 1 (the only values the reference emits) are accumulated in registers and flushed once per
// CTA -- 2 x 768 addresses would otherwise take every row's atomics.
__global__ void __launch_bounds__(ROW_WARPS * 32)
embed_ln_bwd_kernel(const bf16* __restrict__ dy, const long long* __restrict__ ids, const long long* __restrict__ type_ids,
                    const long long* __restrict__ pos_ids, const float* __restrict__ word, const float* __restrict__ pos,
                    const float* __restrict__ type, const float* __restrict__ gamma, float eps, long long batch, long long T,
                    long long S_out, float* __restrict__ dword, float* __restrict__ dpos, float* __restrict__ dtype,
                    float* __restrict__ partials, DropoutCfg drop, const int32_t* __restrict__ row_dst) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  drop = resolve_dropout(drop);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float g[CH][8];
  load_vec_f32(gamma, lane, g);
  float acc[4][CH][8];  // dgamma, dbeta, dtype[0], dtype[1]
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[a][c][j] = 0.f;
  const long long rows = batch * T;
  for (long long m = (long long)blockIdx.x * ROW_WARPS + warp; m < rows; m += (long long)gridDim.x * ROW_WARPS) {
    const long long t = m % T, s = m / T;
    long long orow = s * S_out + t;
    if (row_dst) {                       // a dropped row reached no loss: it has no gradient
      orow = row_dst[orow];
      if (orow < 0) continue;
    }
    const long long id = ids[m], tt = type_ids ? type_ids[m] : 0, p = pos_ids ? pos_ids[m] : t;
    float v[CH][8], d[CH][8];
    load_vec_f32(word + id * H, lane, v);
    add_vec_f32(pos + p * H, lane, v);
    add_vec_f32(type + tt * H, lane, v);
    load_row_bf16(dy + orow * H, lane, d);
    apply_dropout_row(drop, m, lane, d);
    float mean, rstd;
    row_stats(v, eps, mean, rstd);
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float xh = (v[c][j] - mean) * rstd;
        acc[0][c][j] += d[c][j] * xh;
        acc[1][c][j] += d[c][j];
        float dg = d[c][j] * g[c][j];
        s1 += dg;
        s2 += dg * xh;
        v[c][j] = xh;
        d[c][j] = dg;
      }
    s1 = warp_sum(s1) * (1.0f / H);
    s2 = warp_sum(s2) * (1.0f / H);
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) d[c][j] = rstd * (d[c][j] - s1 - v[c][j] * s2);
    red_add_row(dword + id * H, lane, d);
    red_add_row(dpos + p * H, lane, d);
    if (tt == 0 || tt == 1) {
#pragma unroll
      for (int c = 0; c < CH; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          acc[2][c][j] += (tt == 0) ? d[c][j] : 0.f;
          acc[3][c][j] += (tt == 1) ? d[c][j] : 0.f;
        }
    } else {
      red_add_row(dtype + tt * H, lane, d);
    }
  }
  flush_partials<4>(acc, partials, lane, warp);
}

// ---------------------------------------------------------------------------------------------
// out[col] (+)= sum_r part[r][col]     (second stage of every column reduction; deterministic)
// ---------------------------------------------------------------------------------------------
__global__ void reduce_partials_kernel(const float* __restrict__ part, long long nrows, long long row_stride, long long ncols,
                                       float* __restrict__ out, int accumulate) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  // grid (column blocks of 32, row splits): with more than one row split the splits meet in `out` through fp32 atomics
  // (accumulate mode only; `out` then already holds a value to add to)
  __shared__ float red[8][33];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const long long col = (long long)blockIdx.x * 32 + tx;
  float s = 0.f;
  if (col < ncols)
    for (long long r = (long long)blockIdx.y * 8 + ty; r < nrows; r += 8ll * gridDim.y) s += part[r * row_stride + col];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && col < ncols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][tx];
    if (gridDim.y > 1) atomicAdd(out + col, t);
    else out[col] = accumulate ? out[col] + t : t;
  }
}

// the NACC column sums a LayerNorm / embedding backward leaves in part[cta][a][H]: one launch, blockIdx.z = a
struct ReduceOuts { float* p[4]; };
__global__ void reduce_partials_multi_kernel(const float* __restrict__ part, long long nrows, int nacc, ReduceOuts outs) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  __shared__ float red[8][33];
  float* out = outs.p[blockIdx.z];
  if (!out) return;
  const int tx = threadIdx.x, ty = threadIdx.y;
  const long long col = (long long)blockIdx.x * 32 + tx;
  float s = 0.f;
  for (long long r = (long long)blockIdx.y * 8 + ty; r < nrows; r += 8ll * gridDim.y) s += part[(r * nacc + blockIdx.z) * H + col];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][tx];
    atomicAdd(out + col, t);
  }
}

static inline dim3 reduce_grid(long long ncols, long long nrows, int accumulate) {
  int splits = 1;
  if (accumulate && nrows > 64) splits = (int)((nrows + 63) / 64 < 16 ? (nrows + 63) / 64 : 16);
  return dim3((unsigned)((ncols + 31) / 32), (unsigned)splits);
}

// column sums of a bf16 matrix [rows, ncols] (ncols % 8 == 0): part[blockIdx.y][col]
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const bf16* __restrict__ x, long long ldx, long long rows, long long ncols, float* __restrict__ part,
                      const int32_t* __restrict__ limit) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  rows = limited_rows(rows, limit);
  const long long col = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (col >= ncols) return;
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
    float v[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(x + r * ldx + col)), v);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] += v[j];
  }
  float* p = part + (long long)blockIdx.y * ncols + col;
  *reinterpret_cast<float4*>(p) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
}

// ---------------------------------------------------------------------------------------------
// weight cast fp32 -> bf16, optionally also the transpose (32x32 tiles through shared memory)
// ---------------------------------------------------------------------------------------------
__global__ void cast_transpose_kernel(const float* __restrict__ src, long long lds, long long rows, long long cols,
                                      bf16* __restrict__ dst, long long ldd, bf16* __restrict__ dst_t, long long ldt) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  __shared__ float tile[32][33];
  const long long c0 = (long long)blockIdx.x * 32, r0 = (long long)blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    long long r = r0 + i, c = c0 + threadIdx.x;
    float v = (r < rows && c < cols) ? src[r * lds + c] : 0.f;
    tile[i][threadIdx.x] = v;
    if (dst && r < rows && c < cols) dst[r * ldd + c] = __float2bfloat16(v);
  }
  if (!dst_t) return;
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    long long c = c0 + i, r = r0 + threadIdx.x;  // transposed: row index c, column index r
    if (c < cols && r < rows) dst_t[c * ldt + r] = __float2bfloat16(tile[threadIdx.x][i]);
  }
}

// out[r, 0:d0] = a[r], out[r, d0:d0+d1] = b[r], out[r, d0+d1:ldo] = 0   (fp32 or bf16 -> bf16; a bf16 feature cache skips the fp32 detour)
__device__ __forceinline__ float ld_f32(const float* p) { return *p; }
__device__ __forceinline__ float ld_f32(const bf16* p) { return __bfloat162float(*p); }
template <typename TA, typename TB>
__global__ void pad_cast_concat_kernel(const TA* __restrict__ a, long long d0, const TB* __restrict__ b, long long d1,
                                       long long rows, bf16* __restrict__ out, long long ldo) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    const TA* ar = a + r * d0;
    const TB* br = b ? b + r * d1 : nullptr;
    bf16* o = out + r * ldo;
    for (long long c = threadIdx.x * 2; c < ldo; c += blockDim.x * 2) {
      float v0 = 0.f, v1 = 0.f;
      long long c1 = c + 1;
      if (c < d0) v0 = ld_f32(ar + c); else if (br && c < d0 + d1) v0 = ld_f32(br + (c - d0));
      if (c1 < d0) v1 = ld_f32(ar + c1); else if (br && c1 < d0 + d1) v1 = ld_f32(br + (c1 - d0));
      *reinterpret_cast<uint32_t*>(o + c) = pack_bf16(v0, v1);
    }
  }
}

// y[m] = x[remap(m)] * dropout_mask(m)    (768-wide rows; gathers the image rows of dX0 into a compact operand)
__global__ void __launch_bounds__(ROW_WARPS * 32)
dropout_rows_kernel(const bf16* __restrict__ x, long long ldx, long long rows, bf16* __restrict__ y, long long ldy, DropoutCfg drop,
                    long long rin, long long rout, long long roff, const int32_t* __restrict__ row_dst) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  drop = resolve_dropout(drop);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (long long m = (long long)blockIdx.x * ROW_WARPS + warp; m < rows; m += (long long)gridDim.x * ROW_WARPS) {
    float v[CH][8];
    long long src = remap_row(m, rin, rout, roff);
    if (row_dst) src = row_dst[src];     // packed rows: the padded position's packed row, or none (the row was dropped: zeros)
    if (src < 0) {
#pragma unroll
      for (int c = 0; c < CH; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) v[c][j] = 0.f;
      store_row_bf16(y + m * ldy, lane, v);
      continue;
    }
    load_row_bf16(x + src * ldx, lane, v);
    apply_dropout_row(drop, m, lane, v);
    store_row_bf16(y + m * ldy, lane, v);
  }
}

// dst[r] = src[idx[r]]   /  dst[idx[r]] += src[r]      (768-wide bf16 rows; idx rows are unique; idx < 0 = a padding row of a
// device-counted compaction (vt_compact_labels): gathered as zeros, skipped by the scatter)
__global__ void __launch_bounds__(ROW_WARPS * 32)
gather_rows_kernel(const bf16* __restrict__ src, long long lds, const long long* __restrict__ idx, long long n, bf16* __restrict__ dst, long long ldd,
                   const int32_t* __restrict__ limit) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  n = limited_rows(n, limit);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (long long r = (long long)blockIdx.x * ROW_WARPS + warp; r < n; r += (long long)gridDim.x * ROW_WARPS) {
    const long long i = idx[r];
    const bf16* s = src + i * lds;
    bf16* d = dst + r * ldd;
#pragma unroll
    for (int c = 0; c < CH; ++c)
      *reinterpret_cast<uint4*>(d + chunk_col(lane, c)) = i >= 0 ? __ldg(reinterpret_cast<const uint4*>(s + chunk_col(lane, c))) : make_uint4(0u, 0u, 0u, 0u);
  }
}
__global__ void __launch_bounds__(ROW_WARPS * 32)
scatter_add_rows_kernel(const bf16* __restrict__ src, long long lds, const long long* __restrict__ idx, long long n, bf16* __restrict__ dst, long long ldd,
                        const int32_t* __restrict__ limit) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  n = limited_rows(n, limit);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (long long r = (long long)blockIdx.x * ROW_WARPS + warp; r < n; r += (long long)gridDim.x * ROW_WARPS) {
    if (idx[r] < 0) continue;
    float a[CH][8], b[CH][8];
    load_row_bf16(src + r * lds, lane, a);
    bf16* d = dst + idx[r] * ldd;
    load_row_bf16(d, lane, b);
#pragma unroll
    for (int c = 0; c < CH; ++c)
#pragma unroll
      for (int j = 0; j < 8; ++j) a[c][j] += b[c][j];
    store_row_bf16(d, lane, a);
  }
}

}  // namespace vt

using namespace vt;

static inline bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

extern "C" int vt_layernorm_fwd(const void* x, int64_t ldx, const float* gamma, const float* beta, float eps, int64_t rows,
                                int64_t hidden, void* y, int64_t ldy, float dropout_p, uint64_t seed, uint32_t dropout_stream,
                                int64_t remap_rows_in, int64_t remap_rows_out, int64_t remap_row_off, vt_stream stream) {
  VT_REQUIRE(hidden == H, "vt_layernorm_fwd: hidden size %lld outside the supported envelope (768)", (long long)hidden);
  VT_REQUIRE(aligned16(x) && aligned16(y) && aligned16(gamma) && aligned16(beta) && ldx % 8 == 0 && ldy % 8 == 0, "vt_layernorm_fwd: 16-byte alignment required");
  if (rows == 0) return VT_OK;
  launch_pdl(ln_fwd_kernel, row_grid(rows), ROW_WARPS * 32, 0, (cudaStream_t)stream, (const bf16*)x, ldx, gamma, beta, eps, rows, (bf16*)y, ldy, make_dropout(dropout_p, seed, dropout_stream), remap_rows_in,
      remap_rows_out, remap_row_off, row_limit());
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int64_t vt_row_partials_rows(int64_t rows) { return row_grid(rows); }

extern "C" int vt_layernorm_bwd(const void* dy, int64_t lddy, const void* x, int64_t ldx, const float* gamma, float eps, int64_t rows,
                                int64_t hidden, void* dx, int64_t lddx, void* dx_dropped, int64_t lddxd, float* partials,
                                float* dgamma, float* dbeta, float* dbias, float in_dropout_p, uint32_t in_dropout_stream,
                                float out_dropout_p, uint32_t out_dropout_stream, uint64_t seed, int64_t remap_rows_in,
                                int64_t remap_rows_out, int64_t remap_row_off, vt_stream stream) {
  VT_REQUIRE(hidden == H, "vt_layernorm_bwd: hidden size %lld outside the supported envelope (768)", (long long)hidden);
  VT_REQUIRE(aligned16(x) && aligned16(dy) && aligned16(dx) && aligned16(dx_dropped) && aligned16(gamma) && aligned16(partials), "vt_layernorm_bwd: 16-byte alignment required");
  VT_REQUIRE(partials, "vt_layernorm_bwd: partials workspace (vt_row_partials_rows(rows) x 3 x 768 fp32) required");
  if (rows == 0) return VT_OK;
  const int grid = ln_bwd_grid(rows);
  cudaStream_t st = (cudaStream_t)stream;
  const int nacc = dbias ? 3 : 2;
  if (dbias)
    launch_pdl(ln_bwd_kernel<3, 1>, grid, ROW_WARPS * 32, 0, st, (const bf16*)dy, lddy, (const bf16*)x, ldx, gamma, eps, rows, (bf16*)dx, lddx,
                                                 (bf16*)dx_dropped, lddxd, partials, make_dropout(in_dropout_p, seed, in_dropout_stream),
                                                 make_dropout(out_dropout_p, seed, out_dropout_stream), remap_rows_in, remap_rows_out,
                                                 remap_row_off, row_limit());
  else
    launch_pdl(ln_bwd_kernel<2, 2>, grid, ROW_WARPS * 32, 0, st, (const bf16*)dy, lddy, (const bf16*)x, ldx, gamma, eps, rows, (bf16*)dx, lddx,
                                                 (bf16*)dx_dropped, lddxd, partials, make_dropout(in_dropout_p, seed, in_dropout_stream),
                                                 make_dropout(out_dropout_p, seed, out_dropout_stream), remap_rows_in, remap_rows_out,
                                                 remap_row_off, row_limit());
  count_launch(1);
  VT_CHECK_LAUNCH();
  // second stage: accumulate into the fp32 gradient buffers
  ReduceOuts outs = {{dgamma, dbeta, dbias, nullptr}};
  const int splits = grid > 64 ? (grid + 63) / 64 < 16 ? (grid + 63) / 64 : 16 : 1;
  launch_pdl(reduce_partials_multi_kernel, dim3(H / 32, splits, nacc), dim3(32, 8), 0, st, partials, grid, nacc, outs);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

static int embed_fwd_impl(const int64_t* ids, const int64_t* type_ids, const int64_t* pos_ids, const float* word,
                          const float* pos, const float* type, const float* gamma, const float* beta, float eps,
                          int64_t batch, int64_t T, int64_t out_rows_per_sample, int64_t hidden, void* y, float dropout_p,
                          uint64_t seed, uint32_t dropout_stream, const int32_t* row_dst, vt_stream stream) {
  VT_REQUIRE(hidden == H, "vt_embed_ln_fwd: hidden size %lld outside the supported envelope (768)", (long long)hidden);
  VT_REQUIRE(aligned16(word) && aligned16(pos) && aligned16(type) && aligned16(y) && aligned16(gamma) && aligned16(beta), "vt_embed_ln_fwd: 16-byte alignment required");
  if (batch * T == 0) return VT_OK;
  launch_pdl(embed_ln_fwd_kernel, row_grid(batch * T), ROW_WARPS * 32, 0, (cudaStream_t)stream, (const long long*)ids, (const long long*)type_ids, (const long long*)pos_ids, word, pos, type, gamma, beta, eps, batch, T,
      out_rows_per_sample, (bf16*)y, make_dropout(dropout_p, seed, dropout_stream), row_dst);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_embed_ln_fwd(const int64_t* ids, const int64_t* type_ids, const int64_t* pos_ids, const float* word,
                               const float* pos, const float* type, const float* gamma, const float* beta, float eps,
                               int64_t batch, int64_t T, int64_t out_rows_per_sample, int64_t hidden, void* y, float dropout_p,
                               uint64_t seed, uint32_t dropout_stream, vt_stream stream) {
  return embed_fwd_impl(ids, type_ids, pos_ids, word, pos, type, gamma, beta, eps, batch, T, out_rows_per_sample, hidden, y, dropout_p, seed,
                        dropout_stream, nullptr, stream);
}
extern "C" int vt_embed_ln_fwd_packed(const int64_t* ids, const int64_t* type_ids, const int64_t* pos_ids, const float* word,
                                      const float* pos, const float* type, const float* gamma, const float* beta, float eps,
                                      int64_t batch, int64_t T, int64_t out_rows_per_sample, int64_t hidden, void* y, float dropout_p,
                                      uint64_t seed, uint32_t dropout_stream, const int32_t* row_dst, vt_stream stream) {
  VT_REQUIRE(row_dst, "vt_embed_ln_fwd_packed: row_dst required");
  return embed_fwd_impl(ids, type_ids, pos_ids, word, pos, type, gamma, beta, eps, batch, T, out_rows_per_sample, hidden, y, dropout_p, seed,
                        dropout_stream, row_dst, stream);
}

static int embed_bwd_impl(const void* dy, const int64_t* ids, const int64_t* type_ids, const int64_t* pos_ids, const float* word,
                          const float* pos, const float* type, const float* gamma, float eps, int64_t batch, int64_t T,
                          int64_t out_rows_per_sample, int64_t hidden, int64_t type_vocab, float* dword, float* dpos,
                          float* dtype, float* dgamma, float* dbeta, float* partials, float dropout_p, uint64_t seed,
                          uint32_t dropout_stream, const int32_t* row_dst, vt_stream stream) {
  VT_REQUIRE(hidden == H, "vt_embed_ln_bwd: hidden size %lld outside the supported envelope (768)", (long long)hidden);
  VT_REQUIRE(aligned16(dy) && aligned16(dword) && aligned16(dpos) && aligned16(dtype) && aligned16(partials), "vt_embed_ln_bwd: 16-byte alignment required");
  if (batch * T == 0) return VT_OK;
  const int grid = row_grid(batch * T);
  cudaStream_t st = (cudaStream_t)stream;
  launch_pdl(embed_ln_bwd_kernel, grid, ROW_WARPS * 32, 0, st, (const bf16*)dy, (const long long*)ids, (const long long*)type_ids,
                                                       (const long long*)pos_ids, word, pos, type, gamma, eps, batch, T,
                                                       out_rows_per_sample, dword, dpos, dtype, partials,
                                                       make_dropout(dropout_p, seed, dropout_stream), row_dst);
  count_launch(1);
  VT_CHECK_LAUNCH();
  ReduceOuts outs = {{dgamma, dbeta, dtype, type_vocab > 1 ? dtype + H : nullptr}};
  const int splits = grid > 64 ? (grid + 63) / 64 < 16 ? (grid + 63) / 64 : 16 : 1;
  launch_pdl(reduce_partials_multi_kernel, dim3(H / 32, splits, 4), dim3(32, 8), 0, st, partials, grid, 4, outs);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_embed_ln_bwd(const void* dy, const int64_t* ids, const int64_t* type_ids, const int64_t* pos_ids, const float* word,
                               const float* pos, const float* type, const float* gamma, float eps, int64_t batch, int64_t T,
                               int64_t out_rows_per_sample, int64_t hidden, int64_t type_vocab, float* dword, float* dpos,
                               float* dtype, float* dgamma, float* dbeta, float* partials, float dropout_p, uint64_t seed,
                               uint32_t dropout_stream, vt_stream stream) {
  return embed_bwd_impl(dy, ids, type_ids, pos_ids, word, pos, type, gamma, eps, batch, T, out_rows_per_sample, hidden, type_vocab, dword, dpos, dtype,
                        dgamma, dbeta, partials, dropout_p, seed, dropout_stream, nullptr, stream);
}
extern "C" int vt_embed_ln_bwd_packed(const void* dy, const int64_t* ids, const int64_t* type_ids, const int64_t* pos_ids, const float* word,
                                      const float* pos, const float* type, const float* gamma, float eps, int64_t batch, int64_t T,
                                      int64_t out_rows_per_sample, int64_t hidden, int64_t type_vocab, float* dword, float* dpos,
                                      float* dtype, float* dgamma, float* dbeta, float* partials, float dropout_p, uint64_t seed,
                                      uint32_t dropout_stream, const int32_t* row_dst, vt_stream stream) {
  VT_REQUIRE(row_dst, "vt_embed_ln_bwd_packed: row_dst required");
  return embed_bwd_impl(dy, ids, type_ids, pos_ids, word, pos, type, gamma, eps, batch, T, out_rows_per_sample, hidden, type_vocab, dword, dpos, dtype,
                        dgamma, dbeta, partials, dropout_p, seed, dropout_stream, row_dst, stream);
}

extern "C" int64_t vt_colsum_partials_rows(int64_t rows) {
  long long g = (rows + 63) / 64;
  long long cap = (long long)num_sms() * 2;
  return g < cap ? (g < 1 ? 1 : g) : cap;
}

extern "C" int vt_colsum_bf16(const void* x, int64_t ldx, int64_t rows, int64_t ncols, float* partials, float* out, int accumulate,
                              vt_stream stream) {
  const long long npad = (ncols + 7) / 8 * 8;  // pad columns are read (must exist, ldx >= npad) but never written to `out`
  VT_REQUIRE(ldx % 8 == 0 && ldx >= npad && aligned16(x) && aligned16(partials), "vt_colsum_bf16: ldx must be a multiple of 8 covering the padded row and pointers 16-byte aligned");
  if (ncols == 0) return VT_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int gy = (int)vt_colsum_partials_rows(rows);
  const int threads = 128;
  const int gx = (int)((npad / 8 + threads - 1) / threads);
  launch_pdl(colsum_partial_kernel, dim3(gx, gy), threads, 0, st, (const bf16*)x, ldx, rows, npad, partials, row_limit());
  launch_pdl(reduce_partials_kernel, reduce_grid(ncols, gy, accumulate), dim3(32, 8), 0, st, partials, gy, npad, ncols, out, accumulate);
  count_launch(2);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

// ---------------------------------------------------------------------------------------------
// vt_colsum_bf16_static: vt_colsum_bf16 for callers that cannot allocate (no counterpart in the reference).
// The first-stage partial sums live in a file-scope device array instead of a caller workspace.  The
// array is one per GPU that loads the library (4.6 MiB each, held until exit) and is shared by every call: the first stage writes
// every partial the second stage reads (nothing has to start at zero, stale contents of an earlier call are never read), but two
// calls that overlap on the GPU (different streams without an ordering between them) race on it.
// ---------------------------------------------------------------------------------------------
constexpr long long COLSUM_STATIC_MAX_COLS = 4096;                                   // padded columns per call
constexpr long long COLSUM_STATIC_FLOATS = 296 * COLSUM_STATIC_MAX_COLS;             // = 2 CTAs x 148 SMs rows of partials
static __device__ __align__(256) float g_colsum_scratch[COLSUM_STATIC_FLOATS];

// In host code `g_colsum_scratch` is only a handle: the device address is looked up once per device, with that device current.
static int colsum_scratch_ptr(float** out) {
  constexpr int MAX_DEV = 64;
  static float* ptrs[MAX_DEV] = {};
  static std::mutex mu;
  int dev = -1;
  VT_CUDA(cudaGetDevice(&dev));
  VT_REQUIRE(dev >= 0 && dev < MAX_DEV, "vt_colsum_bf16_static: device index %d outside the table", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    float* p = nullptr;
    VT_CUDA(cudaGetSymbolAddress((void**)&p, g_colsum_scratch));
    VT_REQUIRE(p, "vt_colsum_bf16_static: cudaGetSymbolAddress returned a null address");
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return VT_OK;
}

extern "C" int64_t vt_colsum_bf16_static_max_cols(void) { return COLSUM_STATIC_MAX_COLS; }

extern "C" int vt_colsum_bf16_static(const void* x, int64_t ldx, int64_t rows, int64_t ncols, float* out, int accumulate, vt_stream stream) {
  const long long npad = (ncols + 7) / 8 * 8;
  VT_REQUIRE(x && out && rows >= 0 && ncols >= 0, "vt_colsum_bf16_static: null pointer or negative size");
  VT_REQUIRE(npad <= COLSUM_STATIC_MAX_COLS, "vt_colsum_bf16_static: %lld columns exceed the static scratch (at most %lld; use vt_colsum_bf16 with a workspace)",
             (long long)ncols, COLSUM_STATIC_MAX_COLS);
  VT_REQUIRE(ldx % 8 == 0 && ldx >= npad && aligned16(x), "vt_colsum_bf16_static: ldx must be a multiple of 8 covering the padded row and x 16-byte aligned");
  if (ncols == 0) return VT_OK;
  float* partials = nullptr;
  int rc = colsum_scratch_ptr(&partials);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  long long gy = vt_colsum_partials_rows(rows);
  if (gy * npad > COLSUM_STATIC_FLOATS) gy = COLSUM_STATIC_FLOATS / npad;   // a GPU with more SMs than the array was sized for
  const int threads = 128;
  const int gx = (int)((npad / 8 + threads - 1) / threads);
  launch_pdl(colsum_partial_kernel, dim3(gx, (unsigned)gy), threads, 0, st, (const bf16*)x, ldx, rows, npad, partials, row_limit());
  launch_pdl(reduce_partials_kernel, reduce_grid(ncols, gy, accumulate), dim3(32, 8), 0, st, partials, gy, npad, ncols, out, accumulate);
  count_launch(2);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_cast_f32_bf16(const float* src, int64_t lds, int64_t rows, int64_t cols, void* dst, int64_t ldd, void* dst_t,
                                int64_t ldt, vt_stream stream) {
  VT_REQUIRE(src && (dst || dst_t), "vt_cast_f32_bf16: null pointer");
  if (rows == 0 || cols == 0) return VT_OK;
  dim3 grid((unsigned)((cols + 31) / 32), (unsigned)((rows + 31) / 32));
  launch_pdl(cast_transpose_kernel, grid, dim3(32, 8), 0, (cudaStream_t)stream, src, lds, rows, cols, (bf16*)dst, ldd, (bf16*)dst_t, ldt);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_pad_cast_concat_mixed(const void* a, int a_bf16, int64_t d0, const void* b, int b_bf16, int64_t d1, int64_t rows, void* out,
                                        int64_t ldo, vt_stream stream) {
  VT_REQUIRE(a && out && ldo % 2 == 0 && ldo >= d0 + d1 && ((uintptr_t)out & 3) == 0, "vt_pad_cast_concat: bad arguments");
  if (rows == 0) return VT_OK;
  long long cap = (long long)num_sms() * 8;
  const int grid = (int)(rows < cap ? rows : cap);
  cudaStream_t st = (cudaStream_t)stream;
  bf16* o = (bf16*)out;
  if (a_bf16 && b_bf16) launch_pdl(pad_cast_concat_kernel<bf16, bf16>, grid, 256, 0, st, (const bf16*)a, d0, (const bf16*)b, d1, rows, o, ldo);
  else if (a_bf16) launch_pdl(pad_cast_concat_kernel<bf16, float>, grid, 256, 0, st, (const bf16*)a, d0, (const float*)b, d1, rows, o, ldo);
  else if (b_bf16) launch_pdl(pad_cast_concat_kernel<float, bf16>, grid, 256, 0, st, (const float*)a, d0, (const bf16*)b, d1, rows, o, ldo);
  else launch_pdl(pad_cast_concat_kernel<float, float>, grid, 256, 0, st, (const float*)a, d0, (const float*)b, d1, rows, o, ldo);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_pad_cast_concat(const float* a, int64_t d0, const float* b, int64_t d1, int64_t rows, void* out, int64_t ldo,
                                  vt_stream stream) {
  return vt_pad_cast_concat_mixed(a, 0, d0, b, 0, d1, rows, out, ldo, stream);
}

static int dropout_rows_impl(const void* x, int64_t ldx, int64_t rows, int64_t hidden, void* y, int64_t ldy, float dropout_p,
                             uint64_t seed, uint32_t dropout_stream, int64_t remap_rows_in, int64_t remap_rows_out,
                             int64_t remap_row_off, const int32_t* row_dst, vt_stream stream) {
  VT_REQUIRE(hidden == H && ldx % 8 == 0 && ldy % 8 == 0 && aligned16(x) && aligned16(y), "vt_dropout_rows: bad arguments");
  if (rows == 0) return VT_OK;
  launch_pdl(dropout_rows_kernel, row_grid(rows), ROW_WARPS * 32, 0, (cudaStream_t)stream, (const bf16*)x, ldx, rows, (bf16*)y, ldy, make_dropout(dropout_p, seed, dropout_stream), remap_rows_in, remap_rows_out, remap_row_off, row_dst);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}
extern "C" int vt_dropout_rows(const void* x, int64_t ldx, int64_t rows, int64_t hidden, void* y, int64_t ldy, float dropout_p,
                               uint64_t seed, uint32_t dropout_stream, int64_t remap_rows_in, int64_t remap_rows_out,
                               int64_t remap_row_off, vt_stream stream) {
  return dropout_rows_impl(x, ldx, rows, hidden, y, ldy, dropout_p, seed, dropout_stream, remap_rows_in, remap_rows_out, remap_row_off, nullptr, stream);
}
extern "C" int vt_dropout_rows_packed(const void* x, int64_t ldx, int64_t rows, int64_t hidden, void* y, int64_t ldy, float dropout_p,
                                      uint64_t seed, uint32_t dropout_stream, int64_t remap_rows_in, int64_t remap_rows_out,
                                      int64_t remap_row_off, const int32_t* row_dst, vt_stream stream) {
  VT_REQUIRE(row_dst, "vt_dropout_rows_packed: row_dst required");
  return dropout_rows_impl(x, ldx, rows, hidden, y, ldy, dropout_p, seed, dropout_stream, remap_rows_in, remap_rows_out, remap_row_off, row_dst, stream);
}

extern "C" int vt_gather_rows(const void* src, int64_t lds, const int64_t* idx, int64_t n, int64_t hidden, void* dst, int64_t ldd,
                              vt_stream stream) {
  VT_REQUIRE(hidden == H && lds % 8 == 0 && ldd % 8 == 0 && aligned16(src) && aligned16(dst), "vt_gather_rows: bad arguments");
  if (n == 0) return VT_OK;
  launch_pdl(gather_rows_kernel, row_grid(n), ROW_WARPS * 32, 0, (cudaStream_t)stream, (const bf16*)src, lds, (const long long*)idx, n, (bf16*)dst, ldd, row_limit());
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_scatter_add_rows(const void* src, int64_t lds, const int64_t* idx, int64_t n, int64_t hidden, void* dst, int64_t ldd,
                                   vt_stream stream) {
  VT_REQUIRE(hidden == H && lds % 8 == 0 && ldd % 8 == 0 && aligned16(src) && aligned16(dst), "vt_scatter_add_rows: bad arguments");
  if (n == 0) return VT_OK;
  launch_pdl(scatter_add_rows_kernel, row_grid(n), ROW_WARPS * 32, 0, (cudaStream_t)stream, (const bf16*)src, lds, (const long long*)idx, n, (bf16*)dst, ldd, row_limit());
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}
