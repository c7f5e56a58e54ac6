// Fused masked-softmax self-attention on the 5th-generation tensor cores (tcgen05 + TMEM + TMA),
// forward and backward, head dim 64; replaces oscar/modeling_bert.py:52-72 and its autograd backward.
//
// Forward  (attn_fwd_tc):  one CTA per (128-query tile, head, sample); 64-key tiles stream through a
//   2-stage TMA ring.  S = Q.K^T accumulates in TMEM; four softmax warps (one thread per query row)
//   pull the row out with tcgen05.ld, add the fp32 mask bias, run the online softmax, apply the
//   counter-RNG dropout and write P (bf16) into 128B-swizzled shared memory, from where a second
//   tcgen05.mma accumulates O += P.V in TMEM.  O is rescaled in TMEM only when a row's running max
//   grows by more than 2^8 (lazy rescale: the stale max is a valid softmax reference).
// Backward (attn_bwd_tc<MODE>): recomputes S and dP = dO.V^T per 128x128 tile in TMEM, eight
//   element-wise warps turn them into P (dropout applied) and dS = P (dP - rowsum(dO.O)) / sqrt(d)
//   in shared memory, and further tcgen05.mma's accumulate in TMEM:
//     MODE 0 (one CTA per key tile):    dV += P^T.dO,  dK += dS^T.Q   (operands read MN-major in place)
//     MODE 1 (one CTA per query tile):  dQ += dS.K
//   Two kernels, no atomics: results are deterministic.
// Both share the dropout counter layout of the mma.sync kernels in attention.cu (kept for S > 2048
// and as the A/B reference), so forward and backward of either implementation are interchangeable.
#include "common.cuh"
#include "ptx.cuh"

namespace vt {

using namespace ptx;

void count_launch(int n);
int resolve_driver_entry_points();
int make_map(CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner, uint32_t box_outer);
int attention_fwd_legacy(const void* qkv, int64_t ld, const float* bias, int64_t B, int64_t S, int64_t heads, void* ctx, float* lse,
                         float dropout_p, uint64_t seed, uint32_t dropout_stream, cudaStream_t st);
int attention_bwd_legacy(const void* qkv, int64_t ld, const float* bias, const void* ctx, const void* dctx, const float* lse, float* dsum_ws,
                         int64_t B, int64_t S, int64_t heads, void* dqkv, float dropout_p, uint64_t seed, uint32_t dropout_stream,
                         cudaStream_t st);
int attention_bwd_prep(const void* dctx, const void* ctx, float* dsum, int64_t B, int64_t S, int64_t heads, cudaStream_t st);

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kRescaleTau = 5.545177444f;  // ln(2^8): rescale O only when the row max grew by more than this

struct AttnTcParams {
  int B, S, heads, hidden;
  const float* bias;   // [B, S]
  float scale;
  bf16* out;           // fwd: ctx [B*S, hidden]
  float* lse;          // [B, heads, S, 2] = (reference max, 1 / sum)
  const float* dsum;   // [B, heads, S]
  bf16* dqkv;          // [B*S, ld]
  long long ld;
  DropoutCfg drop;
};

// keep-bits of 32 consecutive key columns [32*jgroup, +32) of query row i (bit c = column 32*jgroup + c),
// in the counter layout of attn_dropout_rand (common.cuh)
__device__ __forceinline__ uint32_t attn_keep_mask32(const DropoutCfg& d, uint32_t bh, uint32_t i, uint32_t jgroup) {
  uint32_t keep = 0;
#pragma unroll
  for (uint32_t t = 0; t < 4; ++t) {
    const uint4 r = attn_dropout_rand(d, bh, i, jgroup, t);
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      keep |= ((w[q] & 0xffffu) >= d.thresh16 ? 1u : 0u) << (8 * q + 2 * t);
      keep |= ((w[q] >> 16) >= d.thresh16 ? 1u : 0u) << (8 * q + 2 * t + 1);
    }
  }
  return keep;
}

// ---------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------
constexpr int FQ = 128, FK = 64;
constexpr int FWD_THREADS = 192;  // warps 0-3 softmax, warp 4 TMA, warp 5 MMA + TMEM
constexpr int FWD_SMEM_TILES = 16384 + 2 * 8192 + 2 * 8192 + 16384;  // Q, K[2], V[2], P
constexpr int FWD_TMEM_COLS = 128;                                    // S: [0,64)  O: [64,128)

__global__ void __launch_bounds__(FWD_THREADS, 2)
attn_fwd_tc(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV, const AttnTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base, sK = base + 16384, sV = sK + 2 * 8192, sP = sV + 2 * 8192;
  const uint32_t bars = sP + 16384;
  const uint32_t q_full = bars, kv_full = bars + 8, kv_empty = bars + 24, s_full = bars + 40, s_free = bars + 48, p_ready = bars + 56,
                 o_full = bars + 64, tmem_slot = bars + 72;
  float* sBias = reinterpret_cast<float*>(smem_raw + (bars + 128 - smem_u32(smem_raw)));
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int S = p.S;
  const int nkt = (S + FK - 1) / FK;
  const int row_base = b * S;

  if (warp == 4 && lane == 0) {
    prefetch_tensormap(&tmQ);
    prefetch_tensormap(&tmKV);
    mbar_init(q_full, 1);
    mbar_init(kv_full, 1); mbar_init(kv_full + 8, 1);
    mbar_init(kv_empty, 1); mbar_init(kv_empty + 8, 1);
    mbar_init(s_full, 1);
    mbar_init(s_free, 128);
    mbar_init(p_ready, 128);
    mbar_init(o_full, 1);
    fence_barrier_init();
  }
  if (warp == 5) {
    tmem_alloc(tmem_slot, FWD_TMEM_COLS);
    tmem_relinquish();
  }
  for (int j = threadIdx.x; j < nkt * FK; j += FWD_THREADS) sBias[j] = j < S ? p.bias[(long long)b * S + j] : -INFINITY;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 4) {
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, 16384);
      tma_load_2d(sQ, &tmQ, q_full, h * 64, row_base + qt * FQ);
      for (int j = 0; j < nkt; ++j) {
        const int st = j & 1;
        if (j >= 2) mbar_wait(kv_empty + 8 * st, (uint32_t)(((j >> 1) - 1) & 1));
        mbar_arrive_expect_tx(kv_full + 8 * st, 16384);
        tma_load_2d(sK + st * 8192, &tmKV, kv_full + 8 * st, p.hidden + h * 64, row_base + j * FK);
        tma_load_2d(sV + st * 8192, &tmKV, kv_full + 8 * st, 2 * p.hidden + h * 64, row_base + j * FK);
      }
    }
  } else if (warp == 5) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, FK, 0, 0);   // S[128 q, 64 keys] = Q (K-major) x K (K-major)
      constexpr uint32_t idesc_o = make_idesc_bf16(128, 64, 0, 1);   // O[128 q, 64 d]   = P (K-major) x V (MN-major)
      auto issue_s = [&](int j) {
        const uint64_t a = make_smem_desc(sQ, 16, 1024), bd = make_smem_desc(sK + (j & 1) * 8192, 16, 1024);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16(tmem_base, a + (uint64_t)(2 * k), bd + (uint64_t)(2 * k), idesc_s, k > 0 ? 1u : 0u);
        umma_commit(s_full);
      };
      mbar_wait(q_full, 0);
      mbar_wait(kv_full, 0);
      tc_fence_after();
      issue_s(0);
      for (int j = 0; j < nkt; ++j) {
        if (j + 1 < nkt) {
          mbar_wait(kv_full + 8 * ((j + 1) & 1), (uint32_t)(((j + 1) >> 1) & 1));
          mbar_wait(s_free, (uint32_t)(j & 1));
          tc_fence_after();
          issue_s(j + 1);
        }
        mbar_wait(p_ready, (uint32_t)(j & 1));
        tc_fence_after();
        const uint64_t a = make_smem_desc(sP, 16, 1024), bd = make_smem_desc(sV + (j & 1) * 8192, 8192, 1024);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16(tmem_base + 64, a + (uint64_t)(2 * k), bd + (uint64_t)(128 * k), idesc_o, (j > 0 || k > 0) ? 1u : 0u);
        umma_commit(o_full);
        umma_commit(kv_empty + 8 * (j & 1));
      }
    }
  } else {
    const int r = warp * 32 + lane;
    const int qi = qt * FQ + r;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
    const uint32_t bh = (uint32_t)(b * p.heads + h);
    float m_used = -INFINITY, l = 0.f;
    for (int j = 0; j < nkt; ++j) {
      mbar_wait(s_full, (uint32_t)(j & 1));
      tc_fence_after();
      uint32_t sv[64];
      tmem_ld_32x32(lane_addr, sv);
      tmem_ld_32x32(lane_addr + 32, sv + 32);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(s_free);
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < 64; ++c) {
        const float v = fmaf(__uint_as_float(sv[c]), p.scale, sBias[j * FK + c]);
        sv[c] = __float_as_uint(v);
        mx = fmaxf(mx, v);
      }
      bool resc = false;
      float alpha = 1.f;
      if (mx > m_used + kRescaleTau) {
        alpha = ex2_approx((m_used - mx) * kLog2e);
        l *= alpha;
        m_used = mx;
        resc = j > 0;
      }
      const float mb = m_used * kLog2e;
      float rs = 0.f;
#pragma unroll
      for (int c = 0; c < 64; ++c) {
        const float e = ex2_approx(fmaf(__uint_as_float(sv[c]), kLog2e, -mb));
        rs += e;
        sv[c] = __float_as_uint(e);
      }
      l += rs;
      if (p.drop.thresh16) {
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const uint32_t keep = attn_keep_mask32(p.drop, bh, (uint32_t)qi, (uint32_t)(j * 2 + g));
#pragma unroll
          for (int c = 0; c < 32; ++c) sv[32 * g + c] = __float_as_uint(((keep >> c) & 1u) ? __uint_as_float(sv[32 * g + c]) * p.drop.scale : 0.f);
        }
      }
      if (j > 0) {
        mbar_wait(o_full, (uint32_t)((j - 1) & 1));
        tc_fence_after();
        if (__any_sync(0xffffffffu, resc)) {
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            uint32_t ov[32];
            tmem_ld_32x32(lane_addr + 64 + 32 * hh, ov);
            tmem_ld_wait();
#pragma unroll
            for (int c = 0; c < 32; ++c) ov[c] = __float_as_uint(__uint_as_float(ov[c]) * alpha);
            tmem_st_32x32(lane_addr + 64 + 32 * hh, ov);
          }
          tmem_st_wait();
        }
      }
      // P row (64 keys = 128 B) into the K-major 128B-swizzled tile
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        const uint32_t addr = sP + r * 128 + (((uint32_t)ch ^ (uint32_t)(r & 7)) << 4);
        st_shared_v4(addr, pack_bf16(__uint_as_float(sv[8 * ch]), __uint_as_float(sv[8 * ch + 1])),
                     pack_bf16(__uint_as_float(sv[8 * ch + 2]), __uint_as_float(sv[8 * ch + 3])),
                     pack_bf16(__uint_as_float(sv[8 * ch + 4]), __uint_as_float(sv[8 * ch + 5])),
                     pack_bf16(__uint_as_float(sv[8 * ch + 6]), __uint_as_float(sv[8 * ch + 7])));
      }
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(p_ready);
    }
    mbar_wait(o_full, (uint32_t)((nkt - 1) & 1));
    tc_fence_after();
    const float inv = 1.0f / l;
    bf16* dst = p.out + ((long long)row_base + qi) * p.hidden + h * 64;
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      uint32_t ov[32];
      tmem_ld_32x32(lane_addr + 64 + 32 * hh, ov);
      tmem_ld_wait();
      if (qi < S) {
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(ov[8 * ch + e]) * inv;
          *reinterpret_cast<uint4*>(dst + 32 * hh + 8 * ch) = pack8(f);
        }
      }
    }
    if (qi < S) *reinterpret_cast<float2*>(p.lse + 2 * (((long long)b * p.heads + h) * S + qi)) = make_float2(m_used, inv);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_dealloc(tmem_base, FWD_TMEM_COLS);
}

// ---------------------------------------------------------------------------------------------
// backward
// ---------------------------------------------------------------------------------------------
constexpr int BWD_THREADS = 320;  // warps 0-7 element-wise, warp 8 TMA, warp 9 MMA + TMEM
constexpr int BWD_TMEM_COLS = 512;  // S [0,128)  dP [128,256)  acc0 [256,320)  acc1 [320,384)
constexpr int TILE16K = 16384;

template <int MODE>  // 0: dK/dV (outer = key tile, inner = query tiles)   1: dQ (outer = query tile, inner = key tiles)
__global__ void __launch_bounds__(BWD_THREADS, 1)
attn_bwd_tc(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO, const AttnTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  // outer operands: MODE 0: K_j, V_j      MODE 1: Q_i, dO_i
  // inner ring:     MODE 0: Q_i, dO_i     MODE 1: K_j, V_j
  const uint32_t sOutA = base, sOutB = base + TILE16K;
  const uint32_t sInA = base + 2 * TILE16K, sInB = base + 4 * TILE16K;   // 2 stages each
  const uint32_t sDS = base + 6 * TILE16K;                               // [128 q, 128 k] bf16 as two 64-key halves
  const uint32_t sPD = base + 8 * TILE16K;                               // MODE 0 only
  const uint32_t bars = base + (MODE == 0 ? 10 : 8) * TILE16K;
  const uint32_t out_full = bars, in_full = bars + 8, in_empty = bars + 24, sdp_full = bars + 40, sdp_free = bars + 48, pds_ready = bars + 56,
                 pds_free = bars + 64, done = bars + 72, tmem_slot = bars + 80;
  float* sBias = reinterpret_cast<float*>(smem_raw + (bars + 128 - smem_u32(smem_raw)));
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int ot = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int S = p.S;
  const int nin = (S + 127) / 128;
  const int row_base = b * S;
  const int hid = p.hidden;

  if (warp == 8 && lane == 0) {
    prefetch_tensormap(&tmQKV);
    prefetch_tensormap(&tmDO);
    mbar_init(out_full, 1);
    mbar_init(in_full, 1); mbar_init(in_full + 8, 1);
    mbar_init(in_empty, 1); mbar_init(in_empty + 8, 1);
    mbar_init(sdp_full, 1);
    mbar_init(sdp_free, 256);
    mbar_init(pds_ready, 256);
    mbar_init(pds_free, 1);
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 9) {
    tmem_alloc(tmem_slot, BWD_TMEM_COLS);
    tmem_relinquish();
  }
  if (MODE == 0) {
    for (int j = threadIdx.x; j < 128; j += BWD_THREADS) {
      const int k = ot * 128 + j;
      sBias[j] = k < S ? p.bias[(long long)b * S + k] : -INFINITY;
    }
  } else {
    for (int j = threadIdx.x; j < nin * 128; j += BWD_THREADS) sBias[j] = j < S ? p.bias[(long long)b * S + j] : -INFINITY;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 8) {
    if (lane == 0) {
      mbar_arrive_expect_tx(out_full, 2 * TILE16K);
      if (MODE == 0) {
        tma_load_2d(sOutA, &tmQKV, out_full, hid + h * 64, row_base + ot * 128);
        tma_load_2d(sOutB, &tmQKV, out_full, 2 * hid + h * 64, row_base + ot * 128);
      } else {
        tma_load_2d(sOutA, &tmQKV, out_full, h * 64, row_base + ot * 128);
        tma_load_2d(sOutB, &tmDO, out_full, h * 64, row_base + ot * 128);
      }
      for (int t = 0; t < nin; ++t) {
        const int st = t & 1;
        if (t >= 2) mbar_wait(in_empty + 8 * st, (uint32_t)(((t >> 1) - 1) & 1));
        mbar_arrive_expect_tx(in_full + 8 * st, 2 * TILE16K);
        if (MODE == 0) {
          tma_load_2d(sInA + st * TILE16K, &tmQKV, in_full + 8 * st, h * 64, row_base + t * 128);
          tma_load_2d(sInB + st * TILE16K, &tmDO, in_full + 8 * st, h * 64, row_base + t * 128);
        } else {
          tma_load_2d(sInA + st * TILE16K, &tmQKV, in_full + 8 * st, hid + h * 64, row_base + t * 128);
          tma_load_2d(sInB + st * TILE16K, &tmQKV, in_full + 8 * st, 2 * hid + h * 64, row_base + t * 128);
        }
      }
    }
  } else if (warp == 9) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_t = make_idesc_bf16(128, 64, 1, 1);  // acc[128 keys, 64] += X^T (MN-major A) x Y (MN-major B)
      constexpr uint32_t idesc_q = make_idesc_bf16(128, 64, 0, 1);  // dQ[128 q, 64]    += dS (K-major A) x K (MN-major B)
      auto issue_sdp = [&](int t) {
        const int st = t & 1;
        // S = Q.K^T ; dP = dO.V^T   (A: query-side tile, B: key-side tile, both K-major over d = 64)
        const uint32_t q = MODE == 0 ? sInA + st * TILE16K : sOutA, k = MODE == 0 ? sOutA : sInA + st * TILE16K;
        const uint32_t dO = MODE == 0 ? sInB + st * TILE16K : sOutB, v = MODE == 0 ? sOutB : sInB + st * TILE16K;
        const uint64_t qa = make_smem_desc(q, 16, 1024), kb = make_smem_desc(k, 16, 1024);
        const uint64_t da = make_smem_desc(dO, 16, 1024), vb = make_smem_desc(v, 16, 1024);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) umma_bf16(tmem_base, qa + (uint64_t)(2 * kk), kb + (uint64_t)(2 * kk), idesc_s, kk > 0 ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) umma_bf16(tmem_base + 128, da + (uint64_t)(2 * kk), vb + (uint64_t)(2 * kk), idesc_s, kk > 0 ? 1u : 0u);
        umma_commit(sdp_full);
      };
      mbar_wait(out_full, 0);
      mbar_wait(in_full, 0);
      tc_fence_after();
      issue_sdp(0);
      for (int t = 0; t < nin; ++t) {
        if (t + 1 < nin) {
          mbar_wait(in_full + 8 * ((t + 1) & 1), (uint32_t)(((t + 1) >> 1) & 1));
          mbar_wait(sdp_free, (uint32_t)(t & 1));
          tc_fence_after();
          issue_sdp(t + 1);
        }
        mbar_wait(pds_ready, (uint32_t)(t & 1));
        tc_fence_after();
        const int st = t & 1;
        if (MODE == 0) {
          // dV += Pd^T.dO_t ; dK += dS^T.Q_t : A = the [128 q, 128 k] tile read MN-major (M = keys: two 64-key chunks 16 KiB
          // apart, K = query rows 128 B apart), B = dO_t / Q_t read MN-major (K = query rows)
          const uint64_t pa = make_smem_desc(sPD, TILE16K, 1024), sa = make_smem_desc(sDS, TILE16K, 1024);
          const uint64_t dob = make_smem_desc(sInB + st * TILE16K, TILE16K, 1024), qb = make_smem_desc(sInA + st * TILE16K, TILE16K, 1024);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) umma_bf16(tmem_base + 320, pa + (uint64_t)(128 * kk), dob + (uint64_t)(128 * kk), idesc_t, (t > 0 || kk > 0) ? 1u : 0u);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) umma_bf16(tmem_base + 256, sa + (uint64_t)(128 * kk), qb + (uint64_t)(128 * kk), idesc_t, (t > 0 || kk > 0) ? 1u : 0u);
        } else {
          // dQ += dS.K_t : A = dS K-major over 128 keys (two 64-key atoms), B = K_t MN-major (K = key rows)
          const uint64_t kb = make_smem_desc(sInA + st * TILE16K, TILE16K, 1024);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint64_t sa = make_smem_desc(sDS + (kk >> 2) * TILE16K + (kk & 3) * 32, 16, 1024);
            umma_bf16(tmem_base + 256, sa, kb + (uint64_t)(128 * kk), idesc_q, (t > 0 || kk > 0) ? 1u : 0u);
          }
        }
        umma_commit(pds_free);
        umma_commit(in_empty + 8 * st);
      }
      umma_commit(done);
    }
  } else {
    const int quad = warp & 3, half = warp >> 2;
    const int r = quad * 32 + lane;               // query row within the tile
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16);
    const uint32_t bh = (uint32_t)(b * p.heads + h);
    const long long stat_base = ((long long)b * p.heads + h) * S;
    float2 lse = make_float2(0.f, 0.f);
    float dsum = 0.f;
    if (MODE == 1) {
      const int gi = min(ot * 128 + r, S - 1);
      lse = *reinterpret_cast<const float2*>(p.lse + 2 * (stat_base + gi));
      dsum = p.dsum[stat_base + gi];
    }
    for (int t = 0; t < nin; ++t) {
      const int qtile = MODE == 0 ? t : ot, ktile = MODE == 0 ? ot : t;
      const int gi = qtile * 128 + r;
      if (MODE == 0) {
        const int gc = min(gi, S - 1);
        lse = *reinterpret_cast<const float2*>(p.lse + 2 * (stat_base + gc));
        dsum = p.dsum[stat_base + gc];
      }
      const float* bias = MODE == 0 ? sBias + half * 64 : sBias + t * 128 + half * 64;
      mbar_wait(sdp_full, (uint32_t)(t & 1));
      tc_fence_after();
      uint32_t sv[64], dv[64];
      tmem_ld_32x32(lane_addr + half * 64, sv);
      tmem_ld_32x32(lane_addr + half * 64 + 32, sv + 32);
      tmem_ld_32x32(lane_addr + 128 + half * 64, dv);
      tmem_ld_32x32(lane_addr + 128 + half * 64 + 32, dv + 32);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(sdp_free);
      const bool row_ok = gi < S;
      const float mb = lse.x * kLog2e;
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        uint32_t keep = 0xffffffffu;
        float dscale = 1.f;
        if (p.drop.thresh16) {
          keep = attn_keep_mask32(p.drop, bh, (uint32_t)gi, (uint32_t)(ktile * 4 + half * 2 + g));
          dscale = p.drop.scale;
        }
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          const int cc = 32 * g + c;
          const float v = fmaf(__uint_as_float(sv[cc]), p.scale, bias[cc]);
          float pr = ex2_approx(fmaf(v, kLog2e, -mb)) * lse.y;
          const float m = ((keep >> c) & 1u) ? dscale : 0.f;
          const float dp = __uint_as_float(dv[cc]) * m;
          float ds = pr * (dp - dsum) * p.scale;
          pr *= m;
          if (!row_ok) { pr = 0.f; ds = 0.f; }
          sv[cc] = __float_as_uint(pr);
          dv[cc] = __float_as_uint(ds);
        }
      }
      if (t > 0) mbar_wait(pds_free, (uint32_t)((t - 1) & 1));
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        const uint32_t off = (uint32_t)half * TILE16K + r * 128 + (((uint32_t)ch ^ (uint32_t)(r & 7)) << 4);
        st_shared_v4(sDS + off, pack_bf16(__uint_as_float(dv[8 * ch]), __uint_as_float(dv[8 * ch + 1])),
                     pack_bf16(__uint_as_float(dv[8 * ch + 2]), __uint_as_float(dv[8 * ch + 3])),
                     pack_bf16(__uint_as_float(dv[8 * ch + 4]), __uint_as_float(dv[8 * ch + 5])),
                     pack_bf16(__uint_as_float(dv[8 * ch + 6]), __uint_as_float(dv[8 * ch + 7])));
        if (MODE == 0)
          st_shared_v4(sPD + off, pack_bf16(__uint_as_float(sv[8 * ch]), __uint_as_float(sv[8 * ch + 1])),
                       pack_bf16(__uint_as_float(sv[8 * ch + 2]), __uint_as_float(sv[8 * ch + 3])),
                       pack_bf16(__uint_as_float(sv[8 * ch + 4]), __uint_as_float(sv[8 * ch + 5])),
                       pack_bf16(__uint_as_float(sv[8 * ch + 6]), __uint_as_float(sv[8 * ch + 7])));
      }
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(pds_ready);
    }
    mbar_wait(done, 0);
    tc_fence_after();
    // accumulators -> global: row = outer-tile row (key for MODE 0, query for MODE 1); this warp takes columns [32*half, +32)
    const int go = ot * 128 + r;
    uint32_t a0[32];
    tmem_ld_32x32(lane_addr + 256 + 32 * half, a0);
    tmem_ld_wait();
    bf16* row = p.dqkv + ((long long)row_base + go) * p.ld + h * 64 + 32 * half;
    if (go < S) {
      bf16* d0 = row + (MODE == 0 ? hid : 0);   // dK or dQ
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        float f[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(a0[8 * ch + e]);
        *reinterpret_cast<uint4*>(d0 + 8 * ch) = pack8(f);
      }
    }
    if (MODE == 0) {
      tmem_ld_32x32(lane_addr + 320 + 32 * half, a0);
      tmem_ld_wait();
      if (go < S) {
        bf16* d1 = row + 2 * hid;               // dV
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(a0[8 * ch + e]);
          *reinterpret_cast<uint4*>(d1 + 8 * ch) = pack8(f);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem_base, BWD_TMEM_COLS);
}

static int g_attn_impl = 0;  // 0 auto, 1 mma.sync kernels, 2 tcgen05 kernels

}  // namespace vt

using namespace vt;

extern "C" void vt_debug_attention_impl(int impl) { vt::g_attn_impl = impl; }

static int check_attn_tc(const char* who, const void* qkv, int64_t ld, int64_t B, int64_t S, int64_t heads, int64_t head_dim) {
  VT_REQUIRE(head_dim == 64, "%s: head dim %lld outside the supported envelope (64)", who, (long long)head_dim);
  VT_REQUIRE(B > 0 && S > 0 && heads > 0 && S <= 4096, "%s: bad sizes B=%lld S=%lld heads=%lld", who, (long long)B, (long long)S, (long long)heads);
  VT_REQUIRE(ld % 8 == 0 && ld >= 3 * heads * 64 && ((uintptr_t)qkv & 15) == 0, "%s: qkv must be 16-byte aligned with ld %% 8 == 0", who);
  VT_REQUIRE(B <= 65535 && heads <= 65535 && B * S < (1ll << 31), "%s: grid limits", who);
  return VT_OK;
}

static bool use_tc(int64_t S) { return g_attn_impl == 2 || (g_attn_impl == 0 && S <= 2048); }

extern "C" int vt_attention_fwd(const void* qkv, int64_t ld, const float* bias, int64_t B, int64_t S, int64_t heads, int64_t head_dim,
                                void* ctx, float* lse, float dropout_p, uint64_t seed, uint32_t dropout_stream, vt_stream stream) {
  int rc = check_attn_tc("vt_attention_fwd", qkv, ld, B, S, heads, head_dim);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (!use_tc(S)) return attention_fwd_legacy(qkv, ld, bias, B, S, heads, ctx, lse, dropout_p, seed, dropout_stream, st);
  rc = resolve_driver_entry_points();
  if (rc) return rc;
  VT_REQUIRE(((uintptr_t)ctx & 15) == 0, "vt_attention_fwd: ctx must be 16-byte aligned");
  const int64_t hidden = heads * 64;
  CUtensorMap tq, tkv;
  rc = make_map(&tq, qkv, (uint64_t)(3 * hidden), (uint64_t)(B * S), (uint64_t)ld, 64, FQ);
  if (rc) return rc;
  rc = make_map(&tkv, qkv, (uint64_t)(3 * hidden), (uint64_t)(B * S), (uint64_t)ld, 64, FK);
  if (rc) return rc;
  AttnTcParams p = {};
  p.B = (int)B; p.S = (int)S; p.heads = (int)heads; p.hidden = (int)hidden;
  p.bias = bias; p.scale = 0.125f; p.out = (bf16*)ctx; p.lse = lse; p.ld = ld;
  p.drop = make_dropout(dropout_p, seed, dropout_stream);
  const int nkt = (int)((S + FK - 1) / FK);
  const size_t smem = 1024 + FWD_SMEM_TILES + 128 + (size_t)nkt * FK * sizeof(float);
  static bool attr = false;
  if (!attr) {
    VT_CUDA(cudaFuncSetAttribute(attn_fwd_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, 1024 + FWD_SMEM_TILES + 128 + 2048 * 4 + 256));
    attr = true;
  }
  attn_fwd_tc<<<dim3((unsigned)((S + FQ - 1) / FQ), (unsigned)heads, (unsigned)B), FWD_THREADS, smem, st>>>(tq, tkv, p);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_attention_bwd(const void* qkv, int64_t ld, const float* bias, const void* ctx, const void* dctx, const float* lse,
                                float* dsum_ws, int64_t B, int64_t S, int64_t heads, int64_t head_dim, void* dqkv, float dropout_p,
                                uint64_t seed, uint32_t dropout_stream, vt_stream stream) {
  int rc = check_attn_tc("vt_attention_bwd", qkv, ld, B, S, heads, head_dim);
  if (rc) return rc;
  VT_REQUIRE(((uintptr_t)dqkv & 15) == 0 && ((uintptr_t)ctx & 15) == 0 && ((uintptr_t)dctx & 15) == 0, "vt_attention_bwd: 16-byte alignment required");
  cudaStream_t st = (cudaStream_t)stream;
  if (!use_tc(S)) return attention_bwd_legacy(qkv, ld, bias, ctx, dctx, lse, dsum_ws, B, S, heads, dqkv, dropout_p, seed, dropout_stream, st);
  rc = resolve_driver_entry_points();
  if (rc) return rc;
  const int64_t hidden = heads * 64;
  CUtensorMap tq, tdo;
  rc = make_map(&tq, qkv, (uint64_t)(3 * hidden), (uint64_t)(B * S), (uint64_t)ld, 64, 128);
  if (rc) return rc;
  rc = make_map(&tdo, dctx, (uint64_t)hidden, (uint64_t)(B * S), (uint64_t)hidden, 64, 128);
  if (rc) return rc;
  rc = attention_bwd_prep(dctx, ctx, dsum_ws, B, S, heads, st);
  if (rc) return rc;
  AttnTcParams p = {};
  p.B = (int)B; p.S = (int)S; p.heads = (int)heads; p.hidden = (int)hidden;
  p.bias = bias; p.scale = 0.125f; p.lse = const_cast<float*>(lse); p.dsum = dsum_ws; p.dqkv = (bf16*)dqkv; p.ld = ld;
  p.drop = make_dropout(dropout_p, seed, dropout_stream);
  const int nt = (int)((S + 127) / 128);
  const size_t smem0 = 1024 + 10 * TILE16K + 128 + 128 * sizeof(float);
  const size_t smem1 = 1024 + 8 * TILE16K + 128 + (size_t)nt * 128 * sizeof(float);
  static bool attr = false;
  if (!attr) {
    VT_CUDA(cudaFuncSetAttribute(attn_bwd_tc<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(1024 + 10 * TILE16K + 128 + 512)));
    VT_CUDA(cudaFuncSetAttribute(attn_bwd_tc<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(1024 + 8 * TILE16K + 128 + 2048 * 4)));
    attr = true;
  }
  const dim3 grid((unsigned)nt, (unsigned)heads, (unsigned)B);
  attn_bwd_tc<0><<<grid, BWD_THREADS, smem0, st>>>(tq, tdo, p);
  attn_bwd_tc<1><<<grid, BWD_THREADS, smem1, st>>>(tq, tdo, p);
  count_launch(2);
  VT_CHECK_LAUNCH();
  return VT_OK;
}
