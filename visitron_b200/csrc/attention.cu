// Fused masked-softmax self-attention for the short mixed text+region sequences (S <= 1024, head
// dim 64), forward and backward; replaces oscar/modeling_bert.py:52-72 (scores, /sqrt(d), +mask,
// softmax, dropout, .V, head merge) and its autograd backward.  No S x S tensor ever reaches HBM:
// the forward keeps only the per-row log-sum-exp, the backward recomputes the probabilities.
//
//   * Q, K, V are read in place from the fused QKV projection output [rows, 3*768] (row stride ld).
//   * The additive mask is the reference's fp32 bias (1 - mask) * -10000 for ARBITRARY mask values
//     (fine-tuning passes 255) added to the fp32 scores before the softmax.
//   * Probability dropout uses a counter RNG keyed on (batch*head, query row, key column) so the
//     backward regenerates the identical mask.
//
// Round-1 implementation: 64x64 tiles, 4 warps, mma.sync.m16n8k16 bf16 with fp32 accumulation and
// cp.async double buffering.  Backward = one pass per 64-query tile for dQ and one pass per 64-key
// tile for dK/dV (no atomics, deterministic).
#include "common.cuh"

namespace vt {

void count_launch(int n);

constexpr int HD = 64;            // head dim
constexpr int TQ = 64, TK = 64;   // tile sizes
constexpr int TILE_BYTES = 64 * 128;
constexpr int ATT_THREADS = 128;
constexpr float LOG2E = 1.4426950408889634f;

struct AttnParams {
  const bf16* qkv;      // [B*S, ld]: Q at col h*64, K at hidden + h*64, V at 2*hidden + h*64
  long long ld;
  int B, S, heads, hidden;
  const float* bias;    // [B, S] additive key bias
  float scale;          // 1/sqrt(d)
  bf16* out;            // fwd: ctx [B*S, hidden]
  float* lse;           // [B, heads, S, 2]: row max m and 1 / sum exp(s - m).  Kept separate: with the
                        // fine-tuning mask bias (+2.54e6) a fused m + log(l) would lose log(l) to fp32 rounding.
  // backward
  const bf16* dout;     // [B*S, hidden]
  const float* dsum;    // [B, heads, S]  rowsum(dO * O)
  bf16* dqkv;           // [B*S, ld]
  DropoutCfg drop;
};

__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// byte address of the 16-byte chunk holding element (r, c) (c % 8 == 0) of a swizzled 64x64 bf16 tile
__device__ __forceinline__ uint32_t tile_addr(uint32_t base, int r, int c) { return base + r * 128 + ((((c >> 3) ^ r) & 7) << 4); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  int sz = valid ? 16 : 0;  // src-size 0 => zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// rows [row0, row0+64) x 64 columns starting at col0 of a [.., ld] matrix -> swizzled tile; rows >= row_end are zero filled
__device__ __forceinline__ void load_tile(uint32_t sbase, const bf16* g, long long ld, long long row0, long long row_end, int col0) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int id = threadIdx.x + ATT_THREADS * i;
    int r = id >> 3, ch = id & 7;
    bool ok = row0 + r < row_end;
    const bf16* src = g + (ok ? (row0 + r) : row0) * ld + col0 + ch * 8;
    cp_async16(sbase + r * 128 + (((ch ^ r) & 7) << 4), src, ok);
  }
}

__device__ __forceinline__ void ldsm4(uint32_t addr, uint32_t* r) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm4t(uint32_t addr, uint32_t* r) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// A fragments (16 rows starting at r0, all 64 k columns) of a row-major tile
__device__ __forceinline__ void load_a_frags(uint32_t tile, int r0, int lane, uint32_t (*a)[4]) {
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) ldsm4(tile_addr(tile, r0 + (lane & 15), ks * 16 + (lane >> 4) * 8), a[ks]);
}

// acc[16 x 64] += A(16 x 64, fragments) * X^T where X is a row-major [64 n-rows][64 k-cols] tile  (Q.K^T, dO.V^T)
__device__ __forceinline__ void mma_a_xt(float (*acc)[4], const uint32_t (*a)[4], uint32_t xtile, int lane) {
  const int mi = lane >> 3;
#pragma unroll
  for (int ks = 0; ks < 4; ++ks)
#pragma unroll
    for (int np = 0; np < 4; ++np) {
      uint32_t b[4];
      ldsm4(tile_addr(xtile, np * 16 + (lane & 7) + ((mi >> 1) << 3), ks * 16 + (mi & 1) * 8), b);
      mma16816(acc[2 * np], a[ks], b[0], b[1]);
      mma16816(acc[2 * np + 1], a[ks], b[2], b[3]);
    }
}

// acc[16 x 64] += A(16 x 64(k), fragments) * X where X is a row-major [64 k-rows][64 n-cols] tile  (P.V, dS.K, P^T.dO, dS^T.Q)
__device__ __forceinline__ void mma_a_x(float (*acc)[4], const uint32_t (*a)[4], uint32_t xtile, int lane) {
  const int mi = lane >> 3;
#pragma unroll
  for (int ks = 0; ks < 4; ++ks)
#pragma unroll
    for (int np = 0; np < 4; ++np) {
      uint32_t b[4];
      ldsm4t(tile_addr(xtile, ks * 16 + (lane & 7) + ((mi & 1) << 3), np * 16 + ((mi >> 1) << 3)), b);
      mma16816(acc[2 * np], a[ks], b[0], b[1]);
      mma16816(acc[2 * np + 1], a[ks], b[2], b[3]);
    }
}

// C-fragment (16 x 64 fp32) -> A-fragment (bf16) re-use: no data movement between lanes needed
__device__ __forceinline__ void c_to_a(const float (*c)[4], uint32_t (*a)[4]) {
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    a[ks][0] = pack_bf16(c[2 * ks][0], c[2 * ks][1]);
    a[ks][1] = pack_bf16(c[2 * ks][2], c[2 * ks][3]);
    a[ks][2] = pack_bf16(c[2 * ks + 1][0], c[2 * ks + 1][1]);
    a[ks][3] = pack_bf16(c[2 * ks + 1][2], c[2 * ks + 1][3]);
  }
}

__device__ __forceinline__ void zero_acc(float (*c)[4]) {
#pragma unroll
  for (int n = 0; n < 8; ++n) c[n][0] = c[n][1] = c[n][2] = c[n][3] = 0.f;
}

// dropout keep-multipliers for one thread's 16x64 C fragment: rows (i0, i0+8), key columns j0 + 8n + 2t (+1)
__device__ __forceinline__ void apply_attn_dropout(const DropoutCfg& d, uint32_t bh, int i0, int j0, int lane, float (*c)[4]) {
  if (!d.thresh16) return;
  const uint32_t t = lane & 3;
#pragma unroll
  for (int half = 0; half < 2; ++half) {     // 32-column groups of the tile
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {         // row g / row g + 8
      uint4 r = attn_dropout_rand(d, bh, (uint32_t)(i0 + 8 * rr), (uint32_t)((j0 >> 5) + half), t);
      uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int n = half * 4 + q;
        c[n][2 * rr] *= ((w[q] & 0xffffu) >= d.thresh16) ? d.scale : 0.f;
        c[n][2 * rr + 1] *= ((w[q] >> 16) >= d.thresh16) ? d.scale : 0.f;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ATT_THREADS)
attn_fwd_kernel(const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sQ = s_u32(smem), sK = sQ + TILE_BYTES, sV = sK + 2 * TILE_BYTES;
  float* sBias = reinterpret_cast<float*>(smem + 5 * TILE_BYTES);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int S = p.S;
  const long long row_base = (long long)b * S;
  const int nkt = (S + TK - 1) / TK;
  const int g = lane >> 2, t = lane & 3;

  load_tile(sQ, p.qkv, p.ld, row_base + qt * TQ, row_base + S, h * HD);
  load_tile(sK, p.qkv, p.ld, row_base, row_base + S, p.hidden + h * HD);
  load_tile(sV, p.qkv, p.ld, row_base, row_base + S, 2 * p.hidden + h * HD);
  cp_async_commit();
  for (int j = threadIdx.x; j < nkt * TK; j += ATT_THREADS) sBias[j] = j < S ? p.bias[(long long)b * S + j] : -INFINITY;

  uint32_t qf[4][4];
  float o[8][4];
  zero_acc(o);
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  const int i0 = qt * TQ + warp * 16 + g;  // this thread's first query row (second is i0 + 8)

  for (int kt = 0; kt < nkt; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nkt) {
      load_tile(sK + (buf ^ 1) * TILE_BYTES, p.qkv, p.ld, row_base + (kt + 1) * TK, row_base + S, p.hidden + h * HD);
      load_tile(sV + (buf ^ 1) * TILE_BYTES, p.qkv, p.ld, row_base + (kt + 1) * TK, row_base + S, 2 * p.hidden + h * HD);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (kt == 0) load_a_frags(sQ, warp * 16, lane, qf);

    float s[8][4];
    zero_acc(s);
    mma_a_xt(s, qf, sK + buf * TILE_BYTES, lane);

    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int n = 0; n < 8; ++n) {
      const float b0 = sBias[kt * TK + 8 * n + 2 * t], b1 = sBias[kt * TK + 8 * n + 2 * t + 1];
      s[n][0] = fmaf(s[n][0], p.scale, b0); s[n][1] = fmaf(s[n][1], p.scale, b1);
      s[n][2] = fmaf(s[n][2], p.scale, b0); s[n][3] = fmaf(s[n][3], p.scale, b1);
      mx[0] = fmaxf(mx[0], fmaxf(s[n][0], s[n][1]));
      mx[1] = fmaxf(mx[1], fmaxf(s[n][2], s[n][3]));
    }
    float alpha[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      const float m_new = fmaxf(m_run[r], mx[r]);
      alpha[r] = exp2f((m_run[r] - m_new) * LOG2E);
      m_run[r] = m_new;
    }
    float rs[2] = {0.f, 0.f};
#pragma unroll
    for (int n = 0; n < 8; ++n) {
      s[n][0] = exp2f((s[n][0] - m_run[0]) * LOG2E); s[n][1] = exp2f((s[n][1] - m_run[0]) * LOG2E);
      s[n][2] = exp2f((s[n][2] - m_run[1]) * LOG2E); s[n][3] = exp2f((s[n][3] - m_run[1]) * LOG2E);
      rs[0] += s[n][0] + s[n][1];
      rs[1] += s[n][2] + s[n][3];
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) l_run[r] = l_run[r] * alpha[r] + rs[r];
#pragma unroll
    for (int n = 0; n < 8; ++n) {
      o[n][0] *= alpha[0]; o[n][1] *= alpha[0];
      o[n][2] *= alpha[1]; o[n][3] *= alpha[1];
    }
    apply_attn_dropout(p.drop, (uint32_t)(b * p.heads + h), i0, kt * TK, lane, s);
    uint32_t pf[4][4];
    c_to_a(s, pf);
    mma_a_x(o, pf, sV + buf * TILE_BYTES, lane);
    __syncthreads();
  }

#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int i = i0 + 8 * r;
    if (i < S) {
      const float inv = 1.0f / l_run[r];
      bf16* dst = p.out + (row_base + i) * p.hidden + h * HD + 2 * t;
#pragma unroll
      for (int n = 0; n < 8; ++n) *reinterpret_cast<uint32_t*>(dst + 8 * n) = pack_bf16(o[n][2 * r] * inv, o[n][2 * r + 1] * inv);
      if (t == 0) *reinterpret_cast<float2*>(p.lse + 2 * (((long long)b * p.heads + h) * S + i)) = make_float2(m_run[r], inv);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// backward prep: dsum[b, h, i] = sum_d dO[i, h, d] * O[i, h, d]
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
attn_bwd_prep_kernel(const bf16* __restrict__ dout, const bf16* __restrict__ out, float* __restrict__ dsum, int B, int S, int heads, int hidden) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long rows = (long long)B * S;
  for (long long m = (long long)blockIdx.x * 8 + warp; m < rows; m += (long long)gridDim.x * 8) {
    const long long b = m / S, i = m % S;
    for (int c = lane; c < hidden / 8; c += 32) {  // 8 lanes cover one 64-wide head
      float a[8], o[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(dout + m * hidden + c * 8)), a);
      unpack8(__ldg(reinterpret_cast<const uint4*>(out + m * hidden + c * 8)), o);
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) s += a[j] * o[j];
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
      if ((lane & 7) == 0) dsum[(b * heads + c / 8) * S + i] = s;
    }
  }
}

// Shared by both backward kernels: one warp's 16 query rows against one 64-key tile.
//   in : qf, dof (A fragments of Q and dO), sKt / sVt (key / value tiles), per-row lse and dsum
//   out: pr  = P (dropout applied, for dV), ds = dS * scale (for dQ / dK)       both fp32 C fragments
__device__ __forceinline__ void bwd_tile(const AttnParams& p, const uint32_t (*qf)[4], const uint32_t (*dof)[4], uint32_t sKt, uint32_t sVt,
                                         const float* sBias, int j0, uint32_t bh, int i0, const float2* lse2, const float* dsum2, int lane,
                                         float (*pr)[4], float (*ds)[4]) {
  const int t = lane & 3;
  zero_acc(pr);
  mma_a_xt(pr, qf, sKt, lane);
  zero_acc(ds);
  mma_a_xt(ds, dof, sVt, lane);
  apply_attn_dropout(p.drop, bh, i0, j0, lane, ds);  // dP = mask * (dO . V^T)
#pragma unroll
  for (int n = 0; n < 8; ++n) {
    const float b0 = sBias[j0 + 8 * n + 2 * t], b1 = sBias[j0 + 8 * n + 2 * t + 1];
    const float p0 = exp2f((fmaf(pr[n][0], p.scale, b0) - lse2[0].x) * LOG2E) * lse2[0].y, p1 = exp2f((fmaf(pr[n][1], p.scale, b1) - lse2[0].x) * LOG2E) * lse2[0].y;
    const float p2 = exp2f((fmaf(pr[n][2], p.scale, b0) - lse2[1].x) * LOG2E) * lse2[1].y, p3 = exp2f((fmaf(pr[n][3], p.scale, b1) - lse2[1].x) * LOG2E) * lse2[1].y;
    ds[n][0] = p0 * (ds[n][0] - dsum2[0]) * p.scale; ds[n][1] = p1 * (ds[n][1] - dsum2[0]) * p.scale;
    ds[n][2] = p2 * (ds[n][2] - dsum2[1]) * p.scale; ds[n][3] = p3 * (ds[n][3] - dsum2[1]) * p.scale;
    pr[n][0] = p0; pr[n][1] = p1; pr[n][2] = p2; pr[n][3] = p3;
  }
  apply_attn_dropout(p.drop, bh, i0, j0, lane, pr);
}

// ---------------------------------------------------------------------------------------------
// backward, dQ: one CTA per (64-query tile, head, sample), loop over key tiles
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ATT_THREADS)
attn_bwd_dq_kernel(const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sQ = s_u32(smem), sdO = sQ + TILE_BYTES, sK = sdO + TILE_BYTES, sV = sK + 2 * TILE_BYTES;
  float* sBias = reinterpret_cast<float*>(smem + 6 * TILE_BYTES);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int S = p.S;
  const long long row_base = (long long)b * S;
  const int nkt = (S + TK - 1) / TK;
  const int g = lane >> 2, t = lane & 3;

  load_tile(sQ, p.qkv, p.ld, row_base + qt * TQ, row_base + S, h * HD);
  load_tile(sdO, p.dout, p.hidden, row_base + qt * TQ, row_base + S, h * HD);
  load_tile(sK, p.qkv, p.ld, row_base, row_base + S, p.hidden + h * HD);
  load_tile(sV, p.qkv, p.ld, row_base, row_base + S, 2 * p.hidden + h * HD);
  cp_async_commit();
  for (int j = threadIdx.x; j < nkt * TK; j += ATT_THREADS) sBias[j] = j < S ? p.bias[(long long)b * S + j] : -INFINITY;

  const int i0 = qt * TQ + warp * 16 + g;
  float2 lse2[2];
  float dsum2[2];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int i = min(i0 + 8 * r, S - 1);
    lse2[r] = *reinterpret_cast<const float2*>(p.lse + 2 * (((long long)b * p.heads + h) * S + i));
    dsum2[r] = p.dsum[((long long)b * p.heads + h) * S + i];
  }
  uint32_t qf[4][4], dof[4][4];
  float dq[8][4];
  zero_acc(dq);
  for (int kt = 0; kt < nkt; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nkt) {
      load_tile(sK + (buf ^ 1) * TILE_BYTES, p.qkv, p.ld, row_base + (kt + 1) * TK, row_base + S, p.hidden + h * HD);
      load_tile(sV + (buf ^ 1) * TILE_BYTES, p.qkv, p.ld, row_base + (kt + 1) * TK, row_base + S, 2 * p.hidden + h * HD);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (kt == 0) {
      load_a_frags(sQ, warp * 16, lane, qf);
      load_a_frags(sdO, warp * 16, lane, dof);
    }
    float pr[8][4], ds[8][4];
    bwd_tile(p, qf, dof, sK + buf * TILE_BYTES, sV + buf * TILE_BYTES, sBias, kt * TK, (uint32_t)(b * p.heads + h), i0, lse2, dsum2, lane, pr, ds);
    uint32_t dsf[4][4];
    c_to_a(ds, dsf);
    mma_a_x(dq, dsf, sK + buf * TILE_BYTES, lane);
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int i = i0 + 8 * r;
    if (i < S) {
      bf16* dst = p.dqkv + (row_base + i) * p.ld + h * HD + 2 * t;
#pragma unroll
      for (int n = 0; n < 8; ++n) *reinterpret_cast<uint32_t*>(dst + 8 * n) = pack_bf16(dq[n][2 * r], dq[n][2 * r + 1]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// backward, dK / dV: one CTA per (64-key tile, head, sample), loop over query tiles
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void store_c_tile(uint32_t tile, int r0, int lane, const float (*c)[4]) {
  const int g = lane >> 2, t = lane & 3;
#pragma unroll
  for (int n = 0; n < 8; ++n) {
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(tile_addr(tile, r0 + g, 8 * n) + t * 4), "r"(pack_bf16(c[n][0], c[n][1])) : "memory");
    asm volatile("st.shared.b32 [%0], %1;" ::"r"(tile_addr(tile, r0 + g + 8, 8 * n) + t * 4), "r"(pack_bf16(c[n][2], c[n][3])) : "memory");
  }
}

// A fragments of X^T for output rows [m0, m0+16) where X is a row-major [64 k-rows][64 m-cols] tile
__device__ __forceinline__ void load_at_frags(uint32_t tile, int m0, int lane, uint32_t (*a)[4]) {
  const int mi = lane >> 3;
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) ldsm4t(tile_addr(tile, ks * 16 + (lane & 7) + ((mi >> 1) << 3), m0 + ((mi & 1) << 3)), a[ks]);
}

__global__ void __launch_bounds__(ATT_THREADS)
attn_bwd_dkv_kernel(const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sK = s_u32(smem), sV = sK + TILE_BYTES, sQ = sV + TILE_BYTES, sdO = sQ + 2 * TILE_BYTES, sP = sdO + 2 * TILE_BYTES, sdS = sP + TILE_BYTES;
  float* sBias = reinterpret_cast<float*>(smem + 8 * TILE_BYTES);  // 64 entries (this key tile)
  float2* sLse = reinterpret_cast<float2*>(sBias + TK);             // [2][64] (m, 1/l)
  float* sDsum = reinterpret_cast<float*>(sLse + 2 * TQ);           // [2][64]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int kt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int S = p.S;
  const long long row_base = (long long)b * S;
  const int nqt = (S + TQ - 1) / TQ;
  const int g = lane >> 2, t = lane & 3;
  const long long stat_base = ((long long)b * p.heads + h) * S;

  load_tile(sK, p.qkv, p.ld, row_base + kt * TK, row_base + S, p.hidden + h * HD);
  load_tile(sV, p.qkv, p.ld, row_base + kt * TK, row_base + S, 2 * p.hidden + h * HD);
  load_tile(sQ, p.qkv, p.ld, row_base, row_base + S, h * HD);
  load_tile(sdO, p.dout, p.hidden, row_base, row_base + S, h * HD);
  cp_async_commit();
  if (threadIdx.x < TK) {
    const int j = kt * TK + threadIdx.x;
    sBias[threadIdx.x] = j < S ? p.bias[(long long)b * S + j] : -INFINITY;
  }
  if (threadIdx.x < TQ) {
    const int i = min((int)threadIdx.x, S - 1);
    sLse[threadIdx.x] = *reinterpret_cast<const float2*>(p.lse + 2 * (stat_base + i));
    sDsum[threadIdx.x] = p.dsum[stat_base + i];
  }

  float dk[8][4], dv[8][4];
  zero_acc(dk);
  zero_acc(dv);
  for (int qt = 0; qt < nqt; ++qt) {
    const int buf = qt & 1;
    if (qt + 1 < nqt) {
      load_tile(sQ + (buf ^ 1) * TILE_BYTES, p.qkv, p.ld, row_base + (qt + 1) * TQ, row_base + S, h * HD);
      load_tile(sdO + (buf ^ 1) * TILE_BYTES, p.dout, p.hidden, row_base + (qt + 1) * TQ, row_base + S, h * HD);
      cp_async_commit();
      if (threadIdx.x < TQ) {
        const int i = min((qt + 1) * TQ + (int)threadIdx.x, S - 1);
        sLse[(buf ^ 1) * TQ + threadIdx.x] = *reinterpret_cast<const float2*>(p.lse + 2 * (stat_base + i));
        sDsum[(buf ^ 1) * TQ + threadIdx.x] = p.dsum[stat_base + i];
      }
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    uint32_t qf[4][4], dof[4][4];
    load_a_frags(sQ + buf * TILE_BYTES, warp * 16, lane, qf);
    load_a_frags(sdO + buf * TILE_BYTES, warp * 16, lane, dof);
    const int i0 = qt * TQ + warp * 16 + g;
    float2 lse2[2] = {sLse[buf * TQ + warp * 16 + g], sLse[buf * TQ + warp * 16 + g + 8]};
    float dsum2[2] = {sDsum[buf * TQ + warp * 16 + g], sDsum[buf * TQ + warp * 16 + g + 8]};
    float pr[8][4], ds[8][4];
    bwd_tile(p, qf, dof, sK, sV, sBias - kt * TK, kt * TK, (uint32_t)(b * p.heads + h), i0, lse2, dsum2, lane, pr, ds);
    // query rows beyond the sequence carry garbage statistics: zero their contribution
    if (i0 >= S) {
#pragma unroll
      for (int n = 0; n < 8; ++n) pr[n][0] = pr[n][1] = ds[n][0] = ds[n][1] = 0.f;
    }
    if (i0 + 8 >= S) {
#pragma unroll
      for (int n = 0; n < 8; ++n) pr[n][2] = pr[n][3] = ds[n][2] = ds[n][3] = 0.f;
    }
    store_c_tile(sP, warp * 16, lane, pr);
    store_c_tile(sdS, warp * 16, lane, ds);
    __syncthreads();
    uint32_t af[4][4];
    load_at_frags(sP, warp * 16, lane, af);           // (P^T)[key rows of this warp][query cols]
    mma_a_x(dv, af, sdO + buf * TILE_BYTES, lane);     // dV += P^T . dO
    load_at_frags(sdS, warp * 16, lane, af);
    mma_a_x(dk, af, sQ + buf * TILE_BYTES, lane);      // dK += dS^T . Q
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int j = kt * TK + warp * 16 + g + 8 * r;
    if (j < S) {
      bf16* dkp = p.dqkv + (row_base + j) * p.ld + p.hidden + h * HD + 2 * t;
      bf16* dvp = p.dqkv + (row_base + j) * p.ld + 2 * p.hidden + h * HD + 2 * t;
#pragma unroll
      for (int n = 0; n < 8; ++n) {
        *reinterpret_cast<uint32_t*>(dkp + 8 * n) = pack_bf16(dk[n][2 * r], dk[n][2 * r + 1]);
        *reinterpret_cast<uint32_t*>(dvp + 8 * n) = pack_bf16(dv[n][2 * r], dv[n][2 * r + 1]);
      }
    }
  }
}

// bias[i] = (1 - (float)mask[i]) * -10000     (vs/encoder.py:238-241; any integer / float mask value)
template <typename T>
__global__ void mask_bias_kernel(const T* __restrict__ mask, long long n, float* __restrict__ bias) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) bias[i] = (1.0f - (float)mask[i]) * -10000.0f;
}

}  // namespace vt

using namespace vt;

static int check_attn(const char* who, const void* qkv, int64_t ld, int64_t B, int64_t S, int64_t heads, int64_t head_dim) {
  VT_REQUIRE(head_dim == HD, "%s: head dim %lld outside the supported envelope (64)", who, (long long)head_dim);
  VT_REQUIRE(B > 0 && S > 0 && heads > 0 && S <= 4096, "%s: bad sizes B=%lld S=%lld heads=%lld", who, (long long)B, (long long)S, (long long)heads);
  VT_REQUIRE(ld % 8 == 0 && ld >= 3 * heads * HD && ((uintptr_t)qkv & 15) == 0, "%s: qkv must be 16-byte aligned with ld %% 8 == 0", who);
  VT_REQUIRE(B <= 65535 && heads <= 65535, "%s: grid limits", who);
  return VT_OK;
}

extern "C" int vt_mask_bias(const void* mask, int mask_dtype, int64_t n, float* bias, vt_stream stream) {
  VT_REQUIRE(mask && bias, "vt_mask_bias: null pointer");
  if (n == 0) return VT_OK;
  const int th = 256;
  const int gr = (int)((n + th - 1) / th);
  cudaStream_t st = (cudaStream_t)stream;
  switch (mask_dtype) {
    case 0: mask_bias_kernel<long long><<<gr, th, 0, st>>>((const long long*)mask, n, bias); break;
    case 1: mask_bias_kernel<unsigned char><<<gr, th, 0, st>>>((const unsigned char*)mask, n, bias); break;
    case 2: mask_bias_kernel<int><<<gr, th, 0, st>>>((const int*)mask, n, bias); break;
    case 3: mask_bias_kernel<float><<<gr, th, 0, st>>>((const float*)mask, n, bias); break;
    default: VT_REQUIRE(false, "vt_mask_bias: unknown mask dtype code %d", mask_dtype);
  }
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

namespace vt {

int attention_bwd_prep(const void* dctx, const void* ctx, float* dsum, int64_t B, int64_t S, int64_t heads, cudaStream_t st) {
  const long long rows = B * S;
  long long pg = (rows + 7) / 8, cap = (long long)num_sms() * 8;
  attn_bwd_prep_kernel<<<(int)(pg < cap ? pg : cap), 256, 0, st>>>((const bf16*)dctx, (const bf16*)ctx, dsum, (int)B, (int)S, (int)heads, (int)(heads * HD));
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

int attention_fwd_legacy(const void* qkv, int64_t ld, const float* bias, int64_t B, int64_t S, int64_t heads, void* ctx, float* lse,
                         float dropout_p, uint64_t seed, uint32_t dropout_stream, cudaStream_t stream) {
  int rc = check_attn("vt_attention_fwd", qkv, ld, B, S, heads, HD);
  if (rc) return rc;
  AttnParams p = {};
  p.qkv = (const bf16*)qkv; p.ld = ld; p.B = (int)B; p.S = (int)S; p.heads = (int)heads; p.hidden = (int)(heads * HD);
  p.bias = bias; p.scale = 0.125f; p.out = (bf16*)ctx; p.lse = lse;
  p.drop = make_dropout(dropout_p, seed, dropout_stream);
  const int nt = (int)((S + TK - 1) / TK);
  const size_t smem = 5 * TILE_BYTES + (size_t)nt * TK * sizeof(float);
  static bool attr = false;
  if (!attr) { VT_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 5 * TILE_BYTES + 4096 * 4)); attr = true; }
  attn_fwd_kernel<<<dim3(nt, (unsigned)heads, (unsigned)B), ATT_THREADS, smem, (cudaStream_t)stream>>>(p);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

int attention_bwd_legacy(const void* qkv, int64_t ld, const float* bias, const void* ctx, const void* dctx, const float* lse, float* dsum_ws,
                         int64_t B, int64_t S, int64_t heads, void* dqkv, float dropout_p, uint64_t seed, uint32_t dropout_stream,
                         cudaStream_t stream) {
  int rc = check_attn("vt_attention_bwd", qkv, ld, B, S, heads, HD);
  if (rc) return rc;
  VT_REQUIRE(((uintptr_t)dqkv & 15) == 0 && ((uintptr_t)ctx & 15) == 0 && ((uintptr_t)dctx & 15) == 0, "vt_attention_bwd: 16-byte alignment required");
  AttnParams p = {};
  p.qkv = (const bf16*)qkv; p.ld = ld; p.B = (int)B; p.S = (int)S; p.heads = (int)heads; p.hidden = (int)(heads * HD);
  p.bias = bias; p.scale = 0.125f; p.lse = const_cast<float*>(lse);
  p.dout = (const bf16*)dctx; p.dsum = dsum_ws; p.dqkv = (bf16*)dqkv;
  p.drop = make_dropout(dropout_p, seed, dropout_stream);
  cudaStream_t st = (cudaStream_t)stream;
  const long long rows = B * S;
  long long pg = (rows + 7) / 8, cap = (long long)num_sms() * 8;
  attn_bwd_prep_kernel<<<(int)(pg < cap ? pg : cap), 256, 0, st>>>((const bf16*)dctx, (const bf16*)ctx, dsum_ws, (int)B, (int)S, (int)heads, p.hidden);
  const int nt = (int)((S + TK - 1) / TK);
  static bool attr = false;
  if (!attr) {
    VT_CUDA(cudaFuncSetAttribute(attn_bwd_dq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 6 * TILE_BYTES + 4096 * 4));
    VT_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * TILE_BYTES + 1024 * 4));
    attr = true;
  }
  attn_bwd_dkv_kernel<<<dim3(nt, (unsigned)heads, (unsigned)B), ATT_THREADS, 8 * TILE_BYTES + (TK + 6 * TQ) * sizeof(float), st>>>(p);
  attn_bwd_dq_kernel<<<dim3(nt, (unsigned)heads, (unsigned)B), ATT_THREADS, 6 * TILE_BYTES + (size_t)nt * TK * sizeof(float), st>>>(p);
  count_launch(3);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

}  // namespace vt
