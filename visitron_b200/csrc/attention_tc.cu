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
// (The round-1 mma.sync kernels that served as the A/B reference are gone: the path never exceeds S = 767.)
#include <stdlib.h>

#include "common.cuh"
#include "ptx.cuh"

// Timing-ablation switches (tools/attn_dbg.sh): compiled in only with -DVT_ATTN_ABLATE (VT_BUILD_ABLATE=1 python -m
// visitron_b200.build); in the product build every VT_DBG(bit) is a compile-time 0 and the branches vanish.
#ifdef VT_ATTN_ABLATE
#define VT_DBG(bit) ((p.dbg & (bit)) != 0)
#else
#define VT_DBG(bit) false
#endif

namespace vt {

using namespace ptx;

void count_launch(int n);
int resolve_driver_entry_points();
int make_map(CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner, uint32_t box_outer);

// bias[i] = (1 - (float)mask[i]) * -10000     (vs/encoder.py:238-241; any integer / float mask value)
template <typename T>
__global__ void mask_bias_kernel(const T* __restrict__ mask, long long n, float* __restrict__ bias) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) bias[i] = (1.0f - (float)mask[i]) * -10000.0f;
}

// backward prep: dsum[b, h, i] = sum_d dO[i, h, d] * O[i, h, d]
__global__ void __launch_bounds__(256)
attn_bwd_prep_kernel(const bf16* __restrict__ dout, const bf16* __restrict__ out, float* __restrict__ dsum, int B, int S, int heads, int hidden,
                     const int* __restrict__ start, const int* __restrict__ row_b, long long rows_cap) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long rows = start ? rows_cap : (long long)B * S;
  for (long long m = (long long)blockIdx.x * 8 + warp; m < rows; m += (long long)gridDim.x * 8) {
    long long b = m / S, i = m % S;
    if (start) {           // packed rows: (sample, position) of row m; padding rows belong to nobody
      b = row_b[m];
      if (b < 0) continue;
      i = m - start[b];
    }
    for (int c0 = lane; c0 < hidden / 8; c0 += 96) {  // 8 lanes cover one 64-wide head; three 32-chunk groups (a 768-wide row) loaded at once
      uint4 qa[3], qo[3];
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const int c = c0 + 32 * u;
        if (c < hidden / 8) {
          qa[u] = __ldg(reinterpret_cast<const uint4*>(dout + m * hidden + c * 8));
          qo[u] = __ldg(reinterpret_cast<const uint4*>(out + m * hidden + c * 8));
        }
      }
#pragma unroll
      for (int u = 0; u < 3; ++u) {
        const int c = c0 + 32 * u;
        if (c - lane < hidden / 8) {      // warp-uniform (hidden is a multiple of 64): the shuffles below need the whole warp
          float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
          if (c < hidden / 8) { unpack8(qa[u], a); unpack8(qo[u], o); }
          float s = 0.f;
#pragma unroll
          for (int j = 0; j < 8; ++j) s += a[j] * o[j];
          s += __shfl_xor_sync(0xffffffffu, s, 1);
          s += __shfl_xor_sync(0xffffffffu, s, 2);
          s += __shfl_xor_sync(0xffffffffu, s, 4);
          if ((lane & 7) == 0 && c < hidden / 8) dsum[(b * heads + c / 8) * S + i] = s;
        }
      }
    }
  }
}

constexpr uint32_t MMA_POLL_NS = 0;  // the MMA-issuing lane spins: a nanosleep here sits on the critical path of every tile
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kRescaleTau = 5.545177444f;  // ln(2^8): rescale O only when the row max grew by more than this

struct AttnTcParams {
  int B, S, heads, hidden;
  const float* bias;   // [B, S]
  const uint32_t* key_tiles;   // [B] bit t = the 64-key tile t of the sample has an unmasked key (vt_attention_key_tiles); null = all
  float scale;
  bf16* out;           // fwd: ctx [B*S, hidden]
  float* lse;          // [B, heads, S, 2] = (reference max, 1 / sum)
  const float* dsum;   // [B, heads, S]
  bf16* dq_ws;         // [key tiles, B*S, hidden] bf16 dQ partials (summed in fp32 by attn_dq_reduce_kernel)
  bf16* dqkv;          // [B*S, ld]
  long long ld;
  DropoutCfg drop;
  // Packed rows (padding-free execution): sample b owns rows [start[b], start[b+1]) of the activation buffers instead of
  // [b S, (b+1) S); S stays the bound (and the stride of lse / dsum).  null = the dense layout.
  const int* start;    // [B + 1] or null
  const int* row_b;    // [rows_cap] sample of a packed row, -1 for a padding row (backward prep / dQ reduction) or null
  long long rows_cap;  // rows of one dQ workspace slice (B S when dense)
  // optional keep-bit store of the probability dropout (vt_set_attention_keep_bits): [B * heads][4 * ceil(S / 128) groups][S rows]
  // words, bit c of word (bh, g, i) = key 32 g + c of query row i was kept.  The forward writes it next to drawing the mask, the
  // backward reads it instead of running the generator again (16 % of the backward's time at p = 0.1).  null: regenerate.
  uint32_t* keep_bits;
  const int* sched;    // backward: [ctas + 1] item ranges of the CTAs (vt_attention_bwd_schedule) or null = round-robin
  int dbg;             // timing ablations only (VT_ATTN_DBG, results are garbage): 1 skip the element-wise arithmetic, 2 skip the exponentials
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
// HV threads per query row (HV = 1: one thread owns the row's 64 keys of a tile; HV = 2: two threads own 32 each and agree on
// the row maximum through shared memory): softmax warps = 4 HV, TMEM lane quadrant = warp & 3, column half = warp >> 2
constexpr int FWD_QB = 1;         // Q buffers (2 lets the TMA lane run ahead across the items of a persistent CTA)
constexpr int FWD_KV = 3;         // K and V ring depths (separate rings: S runs two key tiles ahead of P.V)
constexpr bool FWD_PERSISTENT = false;
constexpr int FWD_SMEM_TILES = FWD_QB * 16384 + 2 * FWD_KV * 8192 + 2 * 16384;  // Q[QB], K[KV], V[KV], P[2]
constexpr int FWD_ITEMS_PER_CTA = 2;   // query tiles per CTA: 2 measured 197 -> 191 us (dense) / 171 -> 168 us (key-tile plan), 3 and 6 lose to
                                       // wave quantisation (209 / 230 us); VT_FWD_ITEMS_PER_CTA overrides for A/B runs
constexpr int FWD_TMEM_COLS = 256;                                            // S: [0,64) [64,128)   O: [128,192)
constexpr int FWD_MAX_S = 2048;

// Pipeline.  Timing ablations (tools/attn_dbg.sh) showed what bound the first version (S and P single-buffered): 131 of its
// 204 us remained with the arithmetic AND the MMAs removed, i.e. the round trip softmax warps -> mbarrier -> MMA-issuing lane
// -> tcgen05.commit -> mbarrier -> softmax warps once per key tile.  Hence:
//  * S is double-buffered in TMEM and P in shared memory: the MMA lane issues S(g+2) as soon as the softmax warps have pulled
//    S(g) into registers and P.V(g) as soon as P(g) is staged; the softmax warps only wait for events two tiles old;
//  * K and V stream through separate FWD_KV-deep rings; a CTA walks a contiguous range of (query tile, head, sample) items.
//    Measured at b=32, S=767 (p=0.1 / p=0): one item per CTA with 3-deep rings 200 / 144 us; persistent (two CTAs per SM, Q
//    double-buffered, which leaves shared memory for 2-deep rings only) 218 / 171 us: the TMA latency of a 2-deep ring costs
//    more than the per-CTA launch / TMEM allocation / cold loads, so FWD_PERSISTENT is off;
//  * HV threads per query row: HV = 2 gives every scheduler 4 softmax warps and was meant to hide latency, but with S / P
//    double-buffered there is little left to hide and its duplicated per-tile bookkeeping costs more than it gains, so the
//    host launches HV = 1.
// ---------------------------------------------------------------------------------------------
// Key tiles that are masked out as a whole are skipped.  The reference adds -10000 to the scores of masked keys
// (vs/encoder.py:226-241), so their probabilities underflow to exactly 0 in fp32 as soon as the row has one unmasked key:
// a 64-key tile without any such key contributes nothing to the row sums, to O, or to any gradient, and leaving it out
// changes no bit of the result.  The batches of this workload are padded to fixed lengths ([text <= 511 | regions <= 256]),
// so typically a quarter of the key tiles go.  key_tile_plan_kernel marks a tile as live when its largest bias is within
// 5000 of the sample's largest bias (robust for any mask value; a sample with every key masked keeps all its tiles and
// gets the same uniform softmax as the reference).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
key_tile_plan_kernel(const float* __restrict__ bias, int S, const int* __restrict__ start, uint32_t* __restrict__ plan) {
  __shared__ float tmax[32];
  const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long base = start ? start[b] : (long long)b * S;   // packed rows: the sample's own range and length
  if (start) S = start[b + 1] - start[b];
  const int nkt = (S + 63) / 64;
  for (int t = warp; t < 32; t += 4) {
    float m = -INFINITY;
    if (t < nkt) {
      const int k0 = t * 64 + lane, k1 = k0 + 32;
      if (k0 < S) m = bias[base + k0];
      if (k1 < S) m = fmaxf(m, bias[base + k1]);
    }
    m = warp_max(m);
    if (lane == 0) tmax[t] = m;
  }
  __syncthreads();
  if (warp == 0) {
    const float m = warp_max(tmax[lane]);
    const uint32_t live = __ballot_sync(0xffffffffu, tmax[lane] > m - 5000.0f && lane < nkt);
    if (lane == 0) plan[b] = live ? live : (nkt >= 32 ? 0xffffffffu : (1u << nkt) - 1u);
  }
}

// walks the live key tiles of a CTA's items in order (same sequence in the TMA, MMA and softmax roles).  An item is a (128-query
// tile, head, sample); with packed rows a sample may be shorter than S, and the query tiles past its end are skipped by every
// role alike -- `lit` counts the items that were NOT skipped and drives the parities of the per-item barriers.
struct TileCursor {
  const uint32_t* plan;
  const int* start;
  int item0, my_items, nqt, heads, S;
  int it, lit;
  int qt, h, b, sb, len;     // the current item: query tile, head, sample, the sample's first row and length
  uint32_t full, rem;
  __device__ __forceinline__ void load() {
    while (it < my_items) {
      const int id = item0 + it;
      qt = id % nqt; h = (id / nqt) % heads; b = id / (nqt * heads);
      sb = start ? __ldg(start + b) : b * S;
      len = start ? __ldg(start + b + 1) - sb : S;
      if (qt * FQ < len) {
        const int nkt = (len + FK - 1) / FK;
        const uint32_t dense = nkt >= 32 ? 0xffffffffu : (1u << nkt) - 1u;
        full = plan ? (__ldg(plan + b) & dense) : dense;
        if (!full) full = dense;
        rem = full;
        return;
      }
      ++it;
    }
    full = rem = 0;
  }
  __device__ __forceinline__ void init(const AttnTcParams& p, int i0, int n, int nqt_) {
    plan = p.key_tiles; start = p.start; item0 = i0; my_items = n; nqt = nqt_; heads = p.heads; S = p.S;
    it = 0; lit = 0;
    load();
  }
  __device__ __forceinline__ bool done() const { return it >= my_items; }
  __device__ __forceinline__ int tile() const { return __ffs((int)rem) - 1; }
  __device__ __forceinline__ bool first() const { return rem == full; }
  __device__ __forceinline__ bool last() const { return (rem & (rem - 1)) == 0; }
  __device__ __forceinline__ void next_item() { ++it; ++lit; load(); }
  __device__ __forceinline__ void next() {
    rem &= rem - 1;
    if (!rem) next_item();
  }
};

template <int HV>
__global__ void __launch_bounds__((4 * HV + 2) * 32, 2)
attn_fwd_tc(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV, const AttnTcParams p) {
  constexpr int FWD_SM_WARPS = 4 * HV, FWD_THREADS = (FWD_SM_WARPS + 2) * 32, COLS = FK / HV;   // COLS: keys per thread and tile
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base, sK = base + FWD_QB * 16384, sV = sK + FWD_KV * 8192, sP = sV + FWD_KV * 8192;
  const uint32_t bars = sP + 2 * 16384;
  static_assert(FWD_QB <= 2 && FWD_KV <= 4, "barrier slots");
  const uint32_t q_full = bars, q_empty = bars + 16, k_full = bars + 32, k_empty = bars + 64, v_full = bars + 96, v_empty = bars + 128,
                 s_full = bars + 160, s_free = bars + 176, p_ready = bars + 192, p_free = bars + 208, o_free = bars + 224, tmem_slot = bars + 232;
  float* sXch = reinterpret_cast<float*>(smem_raw + (bars + 256 - smem_u32(smem_raw)));   // [3][2 halves][128 rows]
  float* sBias = sXch + 768;                                                               // [key tiles * 64]
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int S = p.S;
  const int nkt = (S + FK - 1) / FK, nqt = (S + FQ - 1) / FQ;
  const int n_items = nqt * p.heads * p.B;
  const int item0 = (int)((long long)n_items * blockIdx.x / gridDim.x), item1 = (int)((long long)n_items * (blockIdx.x + 1) / gridDim.x);
  const int my_items = item1 - item0;

  if (warp == FWD_SM_WARPS && lane == 0) {
    prefetch_tensormap(&tmQ);
    prefetch_tensormap(&tmKV);
    for (int i = 0; i < FWD_QB; ++i) { mbar_init(q_full + 8 * i, 1); mbar_init(q_empty + 8 * i, 1); }
    for (int i = 0; i < FWD_KV; ++i) {
      mbar_init(k_full + 8 * i, 1); mbar_init(k_empty + 8 * i, 1);
      mbar_init(v_full + 8 * i, 1); mbar_init(v_empty + 8 * i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(s_full + 8 * i, 1); mbar_init(s_free + 8 * i, FWD_SM_WARPS);
      mbar_init(p_ready + 8 * i, FWD_SM_WARPS); mbar_init(p_free + 8 * i, 1);
    }
    mbar_init(o_free, FWD_SM_WARPS);
    fence_barrier_init();
  }
  if (warp == FWD_SM_WARPS + 1) {
    tmem_alloc(tmem_slot, FWD_TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_sync();   // the prologue above overlapped the tail of the previous kernel; global memory is touched only below
  const uint32_t tmem_base = *tmem_slot_ptr;
  const uint32_t tmem_o = tmem_base + 128;

  if (warp == FWD_SM_WARPS) {
    // ---------------- TMA producer: runs ahead of the consumers across items, bounded by the rings ----------------
    if (lane == 0) {
      TileCursor c;
      c.init(p, item0, my_items, nqt);
      for (int g = 0; !c.done(); ++g, c.next()) {
        const int h = c.h;
        if (c.first()) {
          const int it = c.lit, qb = it % FWD_QB;
          if (it >= FWD_QB) mbar_wait_backoff(q_empty + 8 * qb, (uint32_t)((it / FWD_QB - 1) & 1), 32);
          mbar_arrive_expect_tx(q_full + 8 * qb, 16384);
          tma_load_2d(sQ + qb * 16384, &tmQ, q_full + 8 * qb, h * 64, c.sb + c.qt * FQ);
        }
        const int j = c.tile(), row_base = c.sb;
        const int st = g % FWD_KV;
        const uint32_t ph = (uint32_t)((g / FWD_KV - 1) & 1);
        if (g >= FWD_KV) mbar_wait_backoff(k_empty + 8 * st, ph, 32);
        mbar_arrive_expect_tx(k_full + 8 * st, 8192);
        tma_load_2d(sK + st * 8192, &tmKV, k_full + 8 * st, p.hidden + h * 64, row_base + j * FK);
        if (g >= FWD_KV) mbar_wait_backoff(v_empty + 8 * st, ph, 32);
        mbar_arrive_expect_tx(v_full + 8 * st, 8192);
        tma_load_2d(sV + st * 8192, &tmKV, v_full + 8 * st, 2 * p.hidden + h * 64, row_base + j * FK);
      }
    }
  } else if (warp == FWD_SM_WARPS + 1) {
    // ---------------- MMA issuer ----------------
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, FK, 0, 0);   // S[128 q, 64 keys] = Q (K-major) x K (K-major)
      constexpr uint32_t idesc_o = make_idesc_bf16(128, 64, 0, 1);   // O[128 q, 64 d]   = P (K-major) x V (MN-major)
      TileCursor ci, cc;   // ci: the next S to issue (runs two tiles ahead), cc: the next P.V
      ci.init(p, item0, my_items, nqt);
      cc.init(p, item0, my_items, nqt);
      int gi = 0;
      auto issue_s = [&]() {   // S(gi) -> TMEM buffer gi & 1; frees its K stage (and the item's Q buffer after its last tile)
        const int it = ci.lit, st = gi % FWD_KV, qb = it % FWD_QB;
        if (ci.first()) mbar_wait_backoff(q_full + 8 * qb, (uint32_t)((it / FWD_QB) & 1), MMA_POLL_NS);
        mbar_wait_backoff(k_full + 8 * st, (uint32_t)((gi / FWD_KV) & 1), MMA_POLL_NS);
        tc_fence_after();
        const uint64_t a = make_smem_desc(sQ + qb * 16384, 16, 1024), bd = make_smem_desc(sK + st * 8192, 16, 1024);
        if (!VT_DBG(128)) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tmem_base + 64 * (gi & 1), a + (uint64_t)(2 * k), bd + (uint64_t)(2 * k), idesc_s, k > 0 ? 1u : 0u);
        }
        umma_commit(s_full + 8 * (gi & 1));
        umma_commit(k_empty + 8 * st);
        if (ci.last()) umma_commit(q_empty + 8 * qb);
        ++gi;
        ci.next();
      };
      if (!ci.done()) issue_s();
      if (!ci.done()) issue_s();
      for (int g = 0; !cc.done(); ++g, cc.next()) {
        const int bf = g & 1;
        if (!ci.done()) {   // S(g + 2) reuses S(g)'s TMEM buffer: the softmax warps have pulled S(g) into registers
          mbar_wait_backoff(s_free + 8 * bf, (uint32_t)((g >> 1) & 1), MMA_POLL_NS);
          issue_s();
        }
        mbar_wait_backoff(p_ready + 8 * bf, (uint32_t)((g >> 1) & 1), MMA_POLL_NS);
        const int vst = g % FWD_KV;
        mbar_wait_backoff(v_full + 8 * vst, (uint32_t)((g / FWD_KV) & 1), MMA_POLL_NS);
        const bool first = cc.first();
        if (first && cc.lit > 0) mbar_wait_backoff(o_free, (uint32_t)((cc.lit - 1) & 1), MMA_POLL_NS);   // previous item's O drained
        tc_fence_after();
        const uint64_t a = make_smem_desc(sP + bf * 16384, 16, 1024), bd = make_smem_desc(sV + vst * 8192, 8192, 1024);
        if (!VT_DBG(16)) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tmem_o, a + (uint64_t)(2 * k), bd + (uint64_t)(128 * k), idesc_o, (!first || k > 0) ? 1u : 0u);
        }
        umma_commit(p_free + 8 * bf);      // P buffer bf consumed, O holds the item's live tiles so far
        umma_commit(v_empty + 8 * vst);
      }
    }
  } else {
    // ---------------- softmax warps ----------------
    const int quad = warp & 3, half = warp >> 2;
    const int r = quad * 32 + lane;
    const int sm_tid = (int)threadIdx.x;   // 0 .. 255
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16);
    const uint32_t lane_o = tmem_o + ((uint32_t)(quad * 32) << 16) + COLS * half;   // this thread's columns of O
    const DropoutCfg dcfg = resolve_dropout(p.drop);
    int g = 0, b_loaded = -1;
    TileCursor cs;
    for (cs.init(p, item0, my_items, nqt); !cs.done(); cs.next_item()) {
      const int qt = cs.qt, h = cs.h, b = cs.b;
      const int row_base = cs.sb, len = cs.len;
      const int qi = qt * FQ + r;
      const uint32_t bh = (uint32_t)(b * p.heads + h);
      if (b != b_loaded) {   // a contiguous item range changes sample at most a few times per CTA
        asm volatile("bar.sync 9, %0;" ::"n"(FWD_SM_WARPS * 32) : "memory");   // every softmax warp has left the previous sample's last tile
        for (int k = sm_tid; k < nkt * FK; k += FWD_SM_WARPS * 32) sBias[k] = k < len ? p.bias[(long long)row_base + k] : -INFINITY;
        asm volatile("bar.sync 9, %0;" ::"n"(FWD_SM_WARPS * 32) : "memory");
        b_loaded = b;
      }
      float m_used = -INFINITY, l = 0.f;   // l: sum over this thread's columns only
      const uint32_t live = cs.full;
      for (uint32_t rem = live; rem; rem &= rem - 1, ++g) {
        const int j = __ffs((int)rem) - 1;   // the tile's real index: bias columns and dropout counters do not move
        const int bf = g & 1;
        mbar_wait(s_full + 8 * bf, (uint32_t)((g >> 1) & 1));
        tc_fence_after();
        uint32_t sv[COLS];
#pragma unroll
        for (int c = 0; c < COLS; c += 32) tmem_ld_32x32(lane_addr + 64 * bf + COLS * half + c, sv + c);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_free + 8 * bf);
        const float* bias = sBias + j * FK + COLS * half;
        float mx = -INFINITY;
        if (!VT_DBG(1)) {
#pragma unroll
          for (int c = 0; c < COLS; ++c) {
            const float v = fmaf(__uint_as_float(sv[c]), p.scale, bias[c]);
            sv[c] = __float_as_uint(v);
            mx = fmaxf(mx, v);
          }
        }
        // the row maximum of the whole tile: exchange with the thread that holds the other 32 keys of this row
        if (HV == 2) {
          float* xch = sXch + bf * 256;
          xch[half * 128 + r] = mx;
          asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory");
          mx = fmaxf(mx, xch[(half ^ 1) * 128 + r]);
        }
        bool resc = false;
        float alpha = 1.f;
        if (mx > m_used + kRescaleTau) {
          alpha = ex2_approx((m_used - mx) * kLog2e);
          l *= alpha;
          m_used = mx;
          resc = rem != live;
        }
        const float mb = m_used * kLog2e;
        float rs = 0.f;
        if (!VT_DBG(1)) {
#pragma unroll
          for (int c = 0; c < COLS; ++c) {
            const float x = fmaf(__uint_as_float(sv[c]), kLog2e, -mb);
            const float e = VT_DBG(2) ? x : ex2_approx(x);
            rs += e;
            sv[c] = __float_as_uint(e);
          }
        }
        l += rs;
        if (dcfg.thresh16 && !VT_DBG(1)) {
          // counter layout of attn_dropout_rand: call t of 32-column group g covers columns 8 q + 2 t (low 16-bit lane of word q)
          // and 8 q + 2 t + 1 (high lane); keep iff lane >= thresh16.  The 1 / (1 - p) scale is applied once, to O.
          const uint32_t thi = dcfg.thresh16 << 16;
#pragma unroll
          for (int gq = 0; gq < COLS / 32; ++gq) {
            const uint32_t grp = (uint32_t)(j * 2 + (HV == 2 ? half : gq));
            uint32_t kept = 0xffffffffu;
#pragma unroll
            for (uint32_t t = 0; t < 4; ++t) {
              const uint4 rr = attn_dropout_rand(dcfg, bh, (uint32_t)qi, grp, t);
              const uint32_t w[4] = {rr.x, rr.y, rr.z, rr.w};
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const int c0 = 32 * gq + 8 * q + 2 * (int)t;
                if ((w[q] << 16) < thi) { sv[c0] = 0u; kept &= ~(1u << (8 * q + 2 * (int)t)); }
                if (w[q] < thi) { sv[c0 + 1] = 0u; kept &= ~(2u << (8 * q + 2 * (int)t)); }
              }
            }
            // lanes = consecutive rows: one 128-byte store per warp and group
            if (p.keep_bits && qi < len) p.keep_bits[((size_t)bh * (size_t)(4 * ((S + 127) / 128)) + grp) * (size_t)S + (size_t)qi] = kept;
          }
        }
        if (__any_sync(0xffffffffu, resc)) {
          // O must hold exactly the item's tiles 0 .. j-1 while it is rescaled: P.V(g-1) complete (p_free of the other buffer),
          // and P.V(g) cannot start before this warp's p_ready below.  Same rows, same decision in the partner warp: each
          // rescales its 32 columns.
          mbar_wait(p_free + 8 * (bf ^ 1), (uint32_t)(((g - 1) >> 1) & 1));
          tc_fence_after();
#pragma unroll
          for (int hh = 0; hh < COLS / 16; ++hh) {
            uint32_t ov[16];
            tmem_ld_32x16(lane_o + 16 * hh, ov);
            tmem_ld_wait();
#pragma unroll
            for (int c = 0; c < 16; ++c) ov[c] = __float_as_uint(__uint_as_float(ov[c]) * alpha);
            tmem_st_32x16(lane_o + 16 * hh, ov);
          }
          tmem_st_wait();
        }
        if (g >= 2) mbar_wait(p_free + 8 * bf, (uint32_t)(((g - 2) >> 1) & 1));   // P.V(g-2) has read this P buffer
        // this thread's 32 keys (64 B) of row r of the K-major 128B-swizzled P tile
#pragma unroll
        for (int ch = 0; ch < COLS / 8; ++ch) {
          const uint32_t addr = sP + bf * 16384 + r * 128 + (((uint32_t)(COLS / 8 * half + ch) ^ (uint32_t)(r & 7)) << 4);
          st_shared_v4(addr, pack_bf16(__uint_as_float(sv[8 * ch]), __uint_as_float(sv[8 * ch + 1])),
                       pack_bf16(__uint_as_float(sv[8 * ch + 2]), __uint_as_float(sv[8 * ch + 3])),
                       pack_bf16(__uint_as_float(sv[8 * ch + 4]), __uint_as_float(sv[8 * ch + 5])),
                       pack_bf16(__uint_as_float(sv[8 * ch + 6]), __uint_as_float(sv[8 * ch + 7])));
        }
        fence_proxy_async();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_ready + 8 * bf);
      }
      // item epilogue.  Row sum = the two halves' sums (same reference maximum in both); third exchange buffer, so that it
      // cannot collide with the tile buffers of the neighbouring tiles
      if (HV == 2) {
        float* xch = sXch + 512;
        xch[half * 128 + r] = l;
        asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory");
        l += xch[(half ^ 1) * 128 + r];
      }
      mbar_wait(p_free + 8 * ((g - 1) & 1), (uint32_t)(((g - 1) >> 1) & 1));   // the item's last P.V
      tc_fence_after();
      uint32_t ov[COLS];
#pragma unroll
      for (int c = 0; c < COLS; c += 32) tmem_ld_32x32(lane_o + c, ov + c);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(o_free);   // the next item's first P.V may overwrite O
      const float inv = 1.0f / l;
      const float onorm = dcfg.thresh16 ? inv * dcfg.scale : inv;
      if (qi < len) {
        bf16* dst = p.out + ((long long)row_base + qi) * p.hidden + h * 64 + COLS * half;
#pragma unroll
        for (int ch = 0; ch < COLS / 8; ++ch) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(ov[8 * ch + e]) * onorm;
          *reinterpret_cast<uint4*>(dst + 8 * ch) = pack8(f);
        }
        if (half == 0) *reinterpret_cast<float2*>(p.lse + 2 * (((long long)b * p.heads + h) * S + qi)) = make_float2(m_used, inv);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == FWD_SM_WARPS + 1) tmem_dealloc(tmem_base, FWD_TMEM_COLS);
}

// ---------------------------------------------------------------------------------------------
// backward: ONE pass over the (query tile, key tile) pairs.  A CTA owns a 128-key tile (K_j, V_j resident, dK_j / dV_j
// accumulating in TMEM) and streams the query tiles: S and dP are recomputed per pair, the eight element-wise warps turn
// them into P (dropout applied) and dS once, and three accumulations consume them:
//     dV_j += P^T.dO_i      dK_j += dS^T.Q_i      dQ_i partial = dS.K_j
// The dQ partial of the pair goes to a bf16 workspace slice [key tile][row][hidden] (plain stores, no atomics; the partials
// are the largest HBM stream of the backward, fp32 slices cost 90 us of stores per layer at b=32);
// attn_dq_reduce_kernel adds the key-tile slices in fp32 in a fixed order, so the result is deterministic.
// Constant factors are kept out of the per-element work: dS is staged without the 1/sqrt(d) and 1/(1-p) factors, P
// without 1/(1-p); they are applied to the 64-wide outputs.
// ---------------------------------------------------------------------------------------------
constexpr int BWD_EW_WARPS = 16;     // element-wise warps: TMEM quadrant = warp & 3, 32-key column group = warp >> 2
constexpr int BWD_THREADS = (BWD_EW_WARPS + 2) * 32;  // + TMA warp + MMA / TMEM warp
constexpr int BWD_TMEM_COLS = 512;  // S [0,128)  dP [128,256)  dK [256,320)  dV [320,384)  dQ partial x2 [384,448) [448,512)
constexpr int TILE16K = 16384;
constexpr int BWD_NST = 3;          // Q / dO ring depth
constexpr int BWD_SMEM_TILES = (8 + 2 * BWD_NST) * TILE16K;  // K[2], V[2], Q[NST], dO[NST], dS (2 halves), P (2 halves)

// The kernel is PERSISTENT: one CTA per SM walks the (key tile, head, sample) items round-robin.  A CTA needs all 512
// TMEM columns and 224 KB of shared memory, so two cannot share an SM; as one-item CTAs the launch, TMEM allocation and the
// cold K / V / Q / dO loads of every item (~8k cycles for ~12k of work) were exposed.  Here the TMA warp runs ahead across
// item boundaries (K / V double-buffered), so the next item's operands are resident when the current one drains.
// Registers: a CTA is allocated in whole warpgroups, so 18 warps cost as much as 20 and cap every thread at 96 registers
// (a setmaxnreg split does not help: the MMA-issuing warp itself needs ~90).  The element-wise loop is arranged to fit: the
// next tile's row statistics are requested only after the pair's arithmetic, when the S / dP values are dead.
// This CTA's items (128-key tile, head, sample), round-robin over the grid, in order; with packed rows the key tiles past a
// sample's end are skipped by every role alike.  `lit` numbers the items that were not skipped (K / V buffer and accumulator
// parities), `nin` is the item's number of query tiles.
struct BwdItemCursor {      // loop-invariant inputs (p, nin_max, n_items) are passed in, not stored: the element-wise warps run at a 96-register cap
  int id, lit;
  int kt, h, b, sb, len, nin;
  __device__ __forceinline__ bool valid(int n_items) const { return id < n_items; }
  __device__ __forceinline__ void settle(const AttnTcParams& p, int nin_max, int n_items, int step) {     // moves to the first non-empty item at or after id
    while (id < n_items) {
      kt = id % nin_max; h = (id / nin_max) % p.heads; b = id / (nin_max * p.heads);
      sb = p.start ? __ldg(p.start + b) : b * p.S;
      len = p.start ? __ldg(p.start + b + 1) - sb : p.S;
      if (kt * 128 < len) { nin = (len + 127) / 128; return; }
      id += step;
    }
  }
  __device__ __forceinline__ void init(const AttnTcParams& p, int nin_max, int n_items, int begin, int step) {
    id = begin; lit = 0;
    settle(p, nin_max, n_items, step);
  }
  __device__ __forceinline__ void next(const AttnTcParams& p, int nin_max, int n_items, int step) { id += step; ++lit; settle(p, nin_max, n_items, step); }
};

// BITS = 1: the probability-dropout decisions are read from AttnTcParams::keep_bits (written by the forward) instead of being drawn
// again; also the variant for dropout_p == 0 (keep_bits null: every bit set).  BITS = 0 draws them with the counter generator.
// Tried on this variant and measured slower or equal, not kept (same-box A/B, tools/r2_call19.sh / r2_call20.sh): draining the dQ
// partial at the top of the pair two pairs behind (412 vs 397 us), staging chunk 0 behind chunk 1 (no change), skipping the
// arithmetic of key groups past a packed sample's end (305 -> 302 us), and in the forward letting the warps whose rows lie past the
// sample's end only keep the barrier protocol (165 -> 188 us: the polling warps take more issue slots than the arithmetic did);
// four interleaved chains for the forward's running maximum / sum instead of one (+1 %: the single chain already uses FMNMX3).
template <int BITS>
__global__ void __launch_bounds__(BWD_THREADS, 1)
attn_bwd_fused(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO, const AttnTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sK = base, sV = base + 2 * TILE16K;                      // 2 buffers each (item parity)
  const uint32_t sQ = base + 4 * TILE16K, sDO = sQ + BWD_NST * TILE16K;  // BWD_NST stages each
  const uint32_t sDS = sDO + BWD_NST * TILE16K;                          // [128 q, 128 k] bf16 as two 64-key halves
  const uint32_t sPD = sDS + 2 * TILE16K;
  const uint32_t bars = sPD + 2 * TILE16K;
  const uint32_t sdp_full = bars, sdp_free = bars + 8, pds_ready = bars + 16, pds_free = bars + 24, acc_full = bars + 32, acc_free = bars + 40,
                 dq_full = bars + 48, dq_free = bars + 64, tmem_slot = bars + 80, kv_full = bars + 96, kv_empty = bars + 112,
                 in_full = bars + 128, in_empty = in_full + 8 * BWD_NST;
  float* sBias = reinterpret_cast<float*>(smem_raw + (bars + 256 - smem_u32(smem_raw)));  // [2][128]
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int nin = (p.S + 127) / 128;     // key / query tiles of a full-length sample (the item numbering's bound)
  const int hid = p.hidden;
  // This CTA's items: round-robin over the grid, or -- with a schedule (vt_attention_bwd_schedule: packed rows, where a sample's
  // items cost its own number of query tiles) -- the contiguous range [sched[cta], sched[cta + 1]) of equal total cost.  Below,
  // n_items is the END of the CTA's range.
  int n_items = nin * p.heads * p.B, it_begin = (int)blockIdx.x, it_step = (int)gridDim.x;   // (the schedule is read behind pdl_sync)

  if (warp == BWD_EW_WARPS && lane == 0) {
    prefetch_tensormap(&tmQKV);
    prefetch_tensormap(&tmDO);
    for (int i = 0; i < 2; ++i) { mbar_init(kv_full + 8 * i, 1); mbar_init(kv_empty + 8 * i, 1); }
    for (int i = 0; i < BWD_NST; ++i) { mbar_init(in_full + 8 * i, 1); mbar_init(in_empty + 8 * i, 1); }
    mbar_init(sdp_full, 1);
    mbar_init(sdp_free, BWD_EW_WARPS);
    mbar_init(pds_ready, BWD_EW_WARPS);
    mbar_init(pds_free, 1);
    mbar_init(acc_full, 1);
    mbar_init(acc_free, BWD_EW_WARPS);
    mbar_init(dq_full, 1); mbar_init(dq_full + 8, 1);
    mbar_init(dq_free, BWD_EW_WARPS); mbar_init(dq_free + 8, BWD_EW_WARPS);
    fence_barrier_init();
  }
  if (warp == BWD_EW_WARPS + 1) {
    tmem_alloc(tmem_slot, BWD_TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_sync();   // the prologue above overlapped the tail of the previous kernel; global memory is touched only below
  if (p.sched) {
    it_begin = __ldg(p.sched + blockIdx.x);
    n_items = __ldg(p.sched + blockIdx.x + 1);
    it_step = 1;
  }
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == BWD_EW_WARPS) {
    // ---------------- TMA producer ----------------
    if (lane == 0) {
      int n = 0;
      BwdItemCursor c;
      for (c.init(p, nin, n_items, it_begin, it_step); c.valid(n_items); c.next(p, nin, n_items, it_step)) {
        const int kt = c.kt, h = c.h, it = c.lit;
        const int row_base = c.sb, kvb = it & 1;
        if (it >= 2) mbar_wait_backoff(kv_empty + 8 * kvb, (uint32_t)(((it >> 1) - 1) & 1), 64);
        mbar_arrive_expect_tx(kv_full + 8 * kvb, 2 * TILE16K);
        tma_load_2d(sK + kvb * TILE16K, &tmQKV, kv_full + 8 * kvb, hid + h * 64, row_base + kt * 128);
        tma_load_2d(sV + kvb * TILE16K, &tmQKV, kv_full + 8 * kvb, 2 * hid + h * 64, row_base + kt * 128);
        for (int t = 0; t < c.nin; ++t, ++n) {
          const int st = n % BWD_NST;
          if (n >= BWD_NST) mbar_wait_backoff(in_empty + 8 * st, (uint32_t)((n / BWD_NST - 1) & 1), 64);
          if (VT_DBG(512) && n >= BWD_NST) { mbar_arrive(in_full + 8 * st); continue; }
          mbar_arrive_expect_tx(in_full + 8 * st, 2 * TILE16K);
          tma_load_2d(sQ + st * TILE16K, &tmQKV, in_full + 8 * st, h * 64, row_base + t * 128);
          tma_load_2d(sDO + st * TILE16K, &tmDO, in_full + 8 * st, h * 64, row_base + t * 128);
        }
      }
    }
  } else if (warp == BWD_EW_WARPS + 1) {
    // ---------------- MMA issuer ----------------
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, 128, 0, 0);
      constexpr uint32_t idesc_t = make_idesc_bf16(128, 64, 1, 1);  // acc[128 keys, 64] += X^T (MN-major A) x Y (MN-major B)
      constexpr uint32_t idesc_q = make_idesc_bf16(128, 64, 0, 1);  // dQ[128 q, 64]    = dS (K-major A) x K (MN-major B)
      // pair n of this CTA = (live item cur.lit, query tile t): K / V buffer cur.lit & 1, ring stage n % BWD_NST.  `nxt` runs one
      // pair ahead: S / dP of pair n + 1 are issued before the accumulations of pair n.
      auto issue_sdp = [&](int n, int kvb) {
        const int st = n % BWD_NST;
        // S = Q.K^T ; dP = dO.V^T   (A: query-side tile, B: key-side tile, both K-major over d = 64)
        const uint64_t qa = make_smem_desc(sQ + st * TILE16K, 16, 1024), kb = make_smem_desc(sK + kvb * TILE16K, 16, 1024);
        const uint64_t da = make_smem_desc(sDO + st * TILE16K, 16, 1024), vb = make_smem_desc(sV + kvb * TILE16K, 16, 1024);
        if (!VT_DBG(128)) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) umma_bf16(tmem_base, qa + (uint64_t)(2 * kk), kb + (uint64_t)(2 * kk), idesc_s, kk > 0 ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) umma_bf16(tmem_base + 128, da + (uint64_t)(2 * kk), vb + (uint64_t)(2 * kk), idesc_s, kk > 0 ? 1u : 0u);
        }
        umma_commit(sdp_full);
      };
      BwdItemCursor cur, nxt;
      cur.init(p, nin, n_items, it_begin, it_step);
      nxt = cur;
      int t = 0, t_nxt = 0;     // query tile of pair n / of pair n + 1
      if (cur.valid(n_items)) {
        mbar_wait_backoff(kv_full, 0, MMA_POLL_NS);
        mbar_wait_backoff(in_full, 0, MMA_POLL_NS);
        tc_fence_after();
        issue_sdp(0, 0);
      }
      for (int n = 0; cur.valid(n_items); ++n) {
        const int it = cur.lit, kvb = it & 1;
        // advance the look-ahead cursor to pair n + 1
        if (++t_nxt == nxt.nin) { nxt.next(p, nin, n_items, it_step); t_nxt = 0; }
        if (nxt.valid(n_items)) {
          const int it1 = nxt.lit;
          if (t_nxt == 0) mbar_wait_backoff(kv_full + 8 * (it1 & 1), (uint32_t)((it1 >> 1) & 1), MMA_POLL_NS);
          mbar_wait_backoff(in_full + 8 * ((n + 1) % BWD_NST), (uint32_t)(((n + 1) / BWD_NST) & 1), MMA_POLL_NS);
          mbar_wait_backoff(sdp_free, (uint32_t)(n & 1), MMA_POLL_NS);
          tc_fence_after();
          issue_sdp(n + 1, it1 & 1);
        }
        mbar_wait_backoff(pds_ready, (uint32_t)(n & 1), MMA_POLL_NS);
        if (n >= 2) mbar_wait_backoff(dq_free + 8 * (n & 1), (uint32_t)(((n >> 1) - 1) & 1), MMA_POLL_NS);
        if (t == 0 && it > 0) mbar_wait_backoff(acc_free, (uint32_t)((it - 1) & 1), MMA_POLL_NS);  // dK / dV of the previous item read out
        tc_fence_after();
        const int st = n % BWD_NST, dqb = n & 1;
        // dV += P^T.dO_t ; dK += dS^T.Q_t : A = the [128 q, 128 k] tile read MN-major (M = keys: two 64-key chunks 16 KiB
        // apart, K = query rows 128 B apart), B = dO_t / Q_t read MN-major (K = query rows)
        const uint64_t pa = make_smem_desc(sPD, TILE16K, 1024), sa = make_smem_desc(sDS, TILE16K, 1024);
        const uint64_t dob = make_smem_desc(sDO + st * TILE16K, TILE16K, 1024), qb = make_smem_desc(sQ + st * TILE16K, TILE16K, 1024);
        if (!VT_DBG(16)) {
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) umma_bf16(tmem_base + 320, pa + (uint64_t)(128 * kk), dob + (uint64_t)(128 * kk), idesc_t, (t > 0 || kk > 0) ? 1u : 0u);
        }
        if (!VT_DBG(32)) {
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) umma_bf16(tmem_base + 256, sa + (uint64_t)(128 * kk), qb + (uint64_t)(128 * kk), idesc_t, (t > 0 || kk > 0) ? 1u : 0u);
        }
        // dQ_t partial = dS.K_j : A = dS K-major over 128 keys (two 64-key atoms), B = K_j MN-major (K = key rows)
        const uint64_t kmn = make_smem_desc(sK + kvb * TILE16K, TILE16K, 1024);
        if (!VT_DBG(64))
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint64_t dsa = make_smem_desc(sDS + (kk >> 2) * TILE16K + (kk & 3) * 32, 16, 1024);
          umma_bf16(tmem_base + 384 + 64 * dqb, dsa, kmn + (uint64_t)(128 * kk), idesc_q, kk > 0 ? 1u : 0u);
        }
        umma_commit(pds_free);
        umma_commit(in_empty + 8 * st);
        umma_commit(dq_full + 8 * dqb);
        if (t == cur.nin - 1) {
          umma_commit(acc_full);
          umma_commit(kv_empty + 8 * kvb);
        }
        // pair n + 1 becomes the current pair
        if (++t == cur.nin) { cur = nxt; t = 0; }
      }
    }
  } else {
    // ---------------- element-wise warps ----------------
    const int quad = warp & 3, cq = warp >> 2;    // rows [32 quad, +32) of the tile, key columns [32 cq, +32)
    const int r = quad * 32 + lane;               // query row within the tile
    const int ew_tid = (int)threadIdx.x;          // 0 .. 511
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16);
    const DropoutCfg dcfg = resolve_dropout(p.drop);
    const bool drop = dcfg.thresh16 != 0;
    const float dscale = drop ? dcfg.scale : 1.0f;
    const float inv_dscale = 1.0f / dscale;
    const uint32_t thi = dcfg.thresh16 << 16;   // 0 without dropout: every unsigned lane compares >= 0, i.e. is kept
    const uint32_t st_off = (uint32_t)(cq >> 1) * TILE16K + (uint32_t)r * 128u;  // row r of the 64-key half this group lies in
    const uint32_t ch0 = (uint32_t)(cq & 1) * 4u;                                 // first 16-byte chunk of the group in that row

    // dQ partial of pair n: TMEM -> workspace (this warp: 32 rows x 16 of the 64 columns; 32 bytes per lane = one sector)
    auto drain_dq = [&](int n, bf16* dst) {
      const int dqb = n & 1;
      mbar_wait(dq_full + 8 * dqb, (uint32_t)((n >> 1) & 1));
      tc_fence_after();
      uint32_t a[16];
      tmem_ld_32x16(lane_addr + 384 + 64 * dqb + 16 * cq, a);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(dq_free + 8 * dqb);
      if (dst && !VT_DBG(256)) {
        uint4* d = reinterpret_cast<uint4*>(dst);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(a[8 * j + e]);
          d[j] = pack8(f);
        }
      }
    };

    // dK / dV accumulators of item `item` -> global (row = key; this warp takes columns [16 cq, +16))
    bf16* acc_row = nullptr;
    bool acc_pending = false;
    auto drain_acc = [&](int item) {
      mbar_wait(acc_full, (uint32_t)(item & 1));
      tc_fence_after();
      uint32_t a0[16], a1[16];
      tmem_ld_32x16(lane_addr + 256 + 16 * cq, a0);
      tmem_ld_32x16(lane_addr + 320 + 16 * cq, a1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_free);
      if (acc_row) {
        const float fk = p.scale * dscale, fv = dscale;
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(a0[8 * ch + e]) * fk;
          *reinterpret_cast<uint4*>(acc_row + hid + 8 * ch) = pack8(f);
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(a1[8 * ch + e]) * fv;
          *reinterpret_cast<uint4*>(acc_row + 2 * hid + 8 * ch) = pack8(f);
        }
      }
      acc_pending = false;
    };
    int n = 0;
    float2 lse_n = make_float2(0.f, 0.f);   // row statistics of the NEXT tile (fetched a tile, and across items an item, ahead)
    float dsum_n = 0.f;
    uint32_t kw_n = 0xffffffffu;    // keep bits of this thread's 32 keys in the NEXT tile's row (BITS; all ones without dropout)
    const int kb_groups = 4 * nin;
    bf16* pend = nullptr;  // where the dQ partial of pair n - 1 goes (null: row past the end of the sample)
    // mask bias of the item's 128 keys: copied into the other half of sBias one item ahead with cp.async (no register is
    // held across the item and the global-load latency is not exposed at the item boundary)
    auto fetch_bias = [&](const BwdItemCursor& c2, int slot) {   // c2: the item the slot is for (nothing to do past the last one)
      if (ew_tid < 128 && c2.valid(n_items)) {
        const int k = c2.kt * 128 + ew_tid;
        float* dst = sBias + slot * 128 + ew_tid;
        if (k < c2.len) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(p.bias + (long long)c2.sb + k) : "memory");
        else *dst = -INFINITY;
      }
    };
    BwdItemCursor cur;
    cur.init(p, nin, n_items, it_begin, it_step);
    fetch_bias(cur, 0);
    const int nin_max = nin;
    for (; cur.valid(n_items); cur.next(p, nin_max, n_items, it_step)) {
      const int kt = cur.kt, h = cur.h, b = cur.b, it = cur.lit;
      const int S = cur.len;         // everything below is bounded by the SAMPLE's length (== p.S in the dense layout)
      const int nin = cur.nin;
      const int row_base = cur.sb;
      const uint32_t bh = (uint32_t)(b * p.heads + h);
      const long long stat_base = ((long long)b * p.heads + h) * p.S;
      float* sB = sBias + (it & 1) * 128;
      asm volatile("cp.async.wait_all;" ::: "memory");
      asm volatile("bar.sync 1, %0;" ::"n"(BWD_EW_WARPS * 32) : "memory");  // element-wise warps only
      long long next_stat = -1;        // where the NEXT live item's first-tile row statistics are (bias / statistics run an item ahead)
      int next_kb = 0;
      const int kb_off = BITS ? ((int)bh * kb_groups + kt * 4 + cq) * p.S : 0;   // this thread's word column of the keep-bit store
      {
        BwdItemCursor nxt = cur;
        nxt.next(p, nin_max, n_items, it_step);
        fetch_bias(nxt, (it + 1) & 1);  // that half was last read in item it - 1, which every warp has left
        if (nxt.valid(n_items)) {
          next_stat = ((long long)nxt.b * p.heads + nxt.h) * p.S + min(r, nxt.len - 1);
          if (BITS) next_kb = ((nxt.b * p.heads + nxt.h) * kb_groups + nxt.kt * 4 + cq) * p.S + min(r, nxt.len - 1);
        }
      }
      const float* bias = sB + cq * 32;
      // this warp's 16 of the 64 dQ columns in this key tile's slice of the workspace
      bf16* ws = p.dq_ws + ((long long)kt * p.rows_cap + row_base) * hid + h * 64 + 16 * cq;
      auto load_stats = [&](int t, float2& lse, float& ds) {
        const int gc = min(t * 128 + r, S - 1);
        lse = *reinterpret_cast<const float2*>(p.lse + 2 * (stat_base + gc));
        ds = p.dsum[stat_base + gc];
        if (BITS && p.keep_bits) kw_n = __ldg(p.keep_bits + kb_off + gc);
      };
      if (it == 0) load_stats(0, lse_n, dsum_n);   // later items: requested during the previous item's last pair
      for (int t = 0; t < nin; ++t, ++n) {
        const int gi = t * 128 + r;
        const float2 lse = lse_n;
        const float dsp = dsum_n * inv_dscale;
        const uint32_t kw = kw_n;
        // P = exp2(v log2e - mb) / sum with mb rounded exactly as in the forward (the F6 mask bias of 2.54e6 leaves 0.25 ulp:
        // any other rounding of the reference point would rescale the whole row); rows past the end of the sample get P = dS = 0
        const float mb = lse.x * kLog2e;
        const float inv = gi < S ? lse.y : 0.f;
        mbar_wait(sdp_full, (uint32_t)(n & 1));
        tc_fence_after();
        // Two halves of 16 columns, chosen to match the dropout counter layout (half hf = Philox calls 2 hf and 2 hf + 1 =
        // columns 8 q + 4 hf .. + 3 of each 8-column chunk q): only 16 + 16 accumulator values are live at a time, and the
        // first half is already packed to bf16 when the second is loaded (the whole tile at once needed > 96 registers).
        if constexpr (BITS) {
          // Without the generator there is no counter layout to follow: 8-column chunks (one 16-byte row piece of the staged P / dS
          // each), all four chunks' loads in flight before the second is needed, every chunk staged as soon as it is computed
          // (16 packed registers never accumulate), so the only exposed TMEM round trip of the pair is the first.
          uint32_t sv[4][8], dv[4][8];
          auto ld_chunk = [&](int q) {
            tmem_ld_32x8(lane_addr + cq * 32 + 8 * q, sv[q]);
            tmem_ld_32x8(lane_addr + 128 + cq * 32 + 8 * q, dv[q]);
          };
          ld_chunk(0);
          ld_chunk(1);
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (q < 2) ld_chunk(q + 2);
            if (q == 2) {
              tmem_ld_wait();
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(sdp_free);
            }
            uint32_t pp[4], dd[4];
#pragma unroll
            for (int j2 = 0; j2 < 4; ++j2) {
              float pe[2], de[2];
#pragma unroll
              for (int e = 0; e < 2; ++e) {
                const int cc = 8 * q + 2 * j2 + e;     // column within the warp's 32
                const int ix = 2 * j2 + e;
                const float v = fmaf(__uint_as_float(sv[q][ix]), p.scale, bias[cc]);
                const float x = fmaf(v, kLog2e, -mb);
                const float pr = (VT_DBG(2) ? x : ex2_approx(x)) * inv;
                const bool keep = ((kw >> cc) & 1u) != 0u;
                const float dpm = keep ? __uint_as_float(dv[q][ix]) : 0.f;
                pe[e] = keep ? pr : 0.f;
                de[e] = pr * (dpm - dsp);
              }
              pp[j2] = pack_bf16(pe[0], pe[1]);
              dd[j2] = pack_bf16(de[0], de[1]);
            }
            if (q == 0) {
              // next tile's statistics (see below), requested as early as their registers are free
              if (t + 1 < nin) {
                load_stats(t + 1, lse_n, dsum_n);
              } else if (next_stat >= 0) {
                const long long sb2 = next_stat;
                lse_n = *reinterpret_cast<const float2*>(p.lse + 2 * sb2);
                dsum_n = p.dsum[sb2];
                if (p.keep_bits) kw_n = __ldg(p.keep_bits + next_kb);
              }
              if (n > 0) mbar_wait(pds_free, (uint32_t)((n - 1) & 1));
            }
            const uint32_t off = st_off + (((ch0 + (uint32_t)q) ^ (uint32_t)(r & 7)) << 4);
            st_shared_v4(sDS + off, dd[0], dd[1], dd[2], dd[3]);
            st_shared_v4(sPD + off, pp[0], pp[1], pp[2], pp[3]);
          }
        } else {
        uint32_t pk_p[2][8], pk_d[2][8];   // [half][2 q + (0: columns +0,+1 | 1: columns +2,+3)]
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t sv[16], dv[16];         // index 4 q + j  <->  column 8 q + 4 hf + j
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            tmem_ld_32x4(lane_addr + cq * 32 + 8 * q + 4 * hf, sv + 4 * q);
            tmem_ld_32x4(lane_addr + 128 + cq * 32 + 8 * q + 4 * hf, dv + 4 * q);
          }
          tmem_ld_wait();
          if (hf == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(sdp_free);
          }
          if (VT_DBG(1)) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { pk_p[hf][i] = sv[i]; pk_d[hf][i] = dv[i]; }
            continue;
          }
#pragma unroll
          for (int t2 = 0; t2 < 2; ++t2) {
            const uint32_t tt = (uint32_t)(2 * hf + t2);
            uint4 rr = make_uint4(0u, 0u, 0u, 0u);
            if (!BITS && drop) rr = attn_dropout_rand(dcfg, bh, (uint32_t)gi, (uint32_t)(kt * 4 + cq), tt);
            const uint32_t w[4] = {rr.x, rr.y, rr.z, rr.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              float pe[2], de[2];
#pragma unroll
              for (int e = 0; e < 2; ++e) {
                const int cc = 8 * q + 2 * (int)tt + e;     // column within the warp's 32
                const int ix = 4 * q + 2 * t2 + e;
                const float v = fmaf(__uint_as_float(sv[ix]), p.scale, bias[cc]);
                const float x = fmaf(v, kLog2e, -mb);
                const float pr = (VT_DBG(2) ? x : ex2_approx(x)) * inv;
                const bool keep = BITS ? ((kw >> cc) & 1u) != 0u : (e == 0 ? ((w[q] << 16) >= thi) : (w[q] >= thi));
                const float dpm = keep ? __uint_as_float(dv[ix]) : 0.f;
                pe[e] = keep ? pr : 0.f;
                de[e] = pr * (dpm - dsp);
              }
              pk_p[hf][2 * q + t2] = pack_bf16(pe[0], pe[1]);
              pk_d[hf][2 * q + t2] = pack_bf16(de[0], de[1]);
            }
          }
        }
        // next tile's statistics: requested here, where the S / dP registers are dead, and covered by the stores, the
        // barrier round trip and the dQ drain below
        if (t + 1 < nin) {
          load_stats(t + 1, lse_n, dsum_n);
        } else if (next_stat >= 0) {   // the next item's first tile (its load latency was exposed at every item start)
          const long long sb2 = next_stat;
          lse_n = *reinterpret_cast<const float2*>(p.lse + 2 * sb2);
          dsum_n = p.dsum[sb2];
          if (BITS) kw_n = __ldg(p.keep_bits + next_kb);
        }
        if (n > 0) mbar_wait(pds_free, (uint32_t)((n - 1) & 1));
#pragma unroll
        for (uint32_t ch = 0; ch < 4; ++ch) {
          const uint32_t off = st_off + (((ch0 + ch) ^ (uint32_t)(r & 7)) << 4);
          st_shared_v4(sDS + off, pk_d[0][2 * ch], pk_d[0][2 * ch + 1], pk_d[1][2 * ch], pk_d[1][2 * ch + 1]);
          st_shared_v4(sPD + off, pk_p[0][2 * ch], pk_p[0][2 * ch + 1], pk_p[1][2 * ch], pk_p[1][2 * ch + 1]);
        }
        }   // !BITS
        fence_proxy_async();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(pds_ready);
        if (t == 0 && acc_pending) drain_acc(it - 1);   // the previous item's dK / dV (see the end of the item loop)
        if (n > 0) drain_dq(n - 1, pend);  // while the MMA warp works on this pair
        pend = gi < S ? ws + (long long)gi * hid : nullptr;
      }
      // item epilogue, DEFERRED: the dK / dV accumulators of this item are read out during the next item's first pair (after
      // its P / dS are staged), when the MMAs that complete them have long retired -- waiting for them here idled the
      // element-wise warps at every item boundary (3.6 % of the stall samples).  The MMA lane starts the next item's
      // dK / dV accumulation only after acc_free, i.e. after that read-out.
      const int go = kt * 128 + r;
      acc_row = go < S ? p.dqkv + ((long long)row_base + go) * p.ld + h * 64 + 16 * cq : nullptr;
      acc_pending = true;
    }
    if (acc_pending) drain_acc(cur.lit - 1);
    if (n > 0) drain_dq(n - 1, pend);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == BWD_EW_WARPS + 1) tmem_dealloc(tmem_base, BWD_TMEM_COLS);
}

// Cost-balanced item ranges for the persistent backward on packed rows.  Item (key tile kt, head h, sample b) costs the sample's
// number of query tiles (+ 1 for the item switch) when kt lies inside the sample and nothing otherwise; round-robin assignment
// left the busiest CTA 31 % over the mean on the benchmark's batch (lengths 389 .. 686).  CTA i gets the items whose cost
// prefix lies in [i, i + 1) * total / ctas: sched[i] = first item id (ids as in BwdItemCursor), sched[ctas] = all items.
__global__ void __launch_bounds__(256)
attn_bwd_schedule_kernel(const int* __restrict__ start, int B, int S, int heads, int ctas, int* __restrict__ sched) {
  extern __shared__ long long pre[];   // [B + 1] cost prefix over samples
  const int nin = (S + 127) / 128;
  if (threadIdx.x == 0) {
    long long acc = 0;
    pre[0] = 0;
    for (int b = 0; b < B; ++b) {
      const int nt = (start[b + 1] - start[b] + 127) / 128;
      acc += (long long)heads * nt * (nt + 1);
      pre[b + 1] = acc;
    }
  }
  __syncthreads();
  const long long total = pre[B];
  for (int i = threadIdx.x; i <= ctas; i += blockDim.x) {
    if (i == ctas || total == 0) { sched[i] = i == ctas ? nin * heads * B : 0; continue; }
    const long long T = total * i / ctas;
    int lo = 0, hi = B;            // the sample with pre[b] <= T < pre[b + 1]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (pre[mid] <= T) lo = mid; else hi = mid;
    }
    while (lo + 1 < B && pre[lo + 1] <= T) ++lo;   // samples without items (never with [CLS] kept, but cheap)
    const int nt = (start[lo + 1] - start[lo] + 127) / 128;
    int k = nt > 0 ? (int)((T - pre[lo]) / (nt + 1)) : 0;   // live item of the sample, numbered h * nt + kt
    if (nt > 0 && k >= heads * nt) k = heads * nt - 1;
    sched[i] = nt > 0 ? (lo * heads + k / nt) * nin + k % nt : lo * heads * nin;
  }
}

// dQ[row, c] = factor * sum over key tiles of the workspace slices, in key-tile order (deterministic) -> bf16 into dqkv[:, 0:hidden]
__global__ void __launch_bounds__(256)
attn_dq_reduce_kernel(const bf16* __restrict__ ws, long long slice, int nkt, long long rows, int hidden, float factor, bf16* __restrict__ dqkv,
                      long long ld, const int* __restrict__ start, const int* __restrict__ row_b) {
  pdl_sync();   // everything below is ordered after the previous kernel on the stream (common.cuh)
  const long long per_row = hidden / 8;
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= rows * per_row) return;
  const long long row = i / per_row;
  const int c = (int)(i % per_row) * 8;
  if (start) {   // packed rows: a row has partials from its own sample's key tiles only; padding rows get dQ = 0
    const int b = row_b[row];
    nkt = b < 0 ? 0 : (start[b + 1] - start[b] + 127) / 128;
  }
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  // groups of 8 slices with every load of the group in flight before the first add (S <= 1024 is one group); same order of additions
  for (int k0 = 0; k0 < nkt; k0 += 8) {
    uint4 q[8];
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (k0 + u < nkt) q[u] = __ldcs(reinterpret_cast<const uint4*>(ws + (long long)(k0 + u) * slice + row * hidden + c));
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (k0 + u < nkt) {
        float f[8];
        unpack8(q[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] *= factor;
  *reinterpret_cast<uint4*>(dqkv + row * ld + c) = pack8(acc);
}

}  // namespace vt

using namespace vt;

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

extern "C" int64_t vt_attention_bwd_workspace(int64_t B, int64_t S, int64_t heads) {
  const int64_t dsum_floats = (B * heads * S + 3) & ~(int64_t)3;
  return dsum_floats + ((S + 127) / 128) * B * S * heads * 64 / 2;  // the dQ partials are bf16
}

static int check_attn_tc(const char* who, const void* qkv, int64_t ld, int64_t B, int64_t S, int64_t heads, int64_t head_dim) {
  VT_REQUIRE(head_dim == 64, "%s: head dim %lld outside the supported envelope (64)", who, (long long)head_dim);
  VT_REQUIRE(B > 0 && S > 0 && heads > 0, "%s: bad sizes B=%lld S=%lld heads=%lld", who, (long long)B, (long long)S, (long long)heads);
  VT_REQUIRE(S <= FWD_MAX_S, "%s: S = %lld outside the supported envelope (<= %d; the path's position tables end at 767)", who, (long long)S, FWD_MAX_S);
  VT_REQUIRE(ld % 8 == 0 && ld >= 3 * heads * 64 && ((uintptr_t)qkv & 15) == 0, "%s: qkv must be 16-byte aligned with ld %% 8 == 0", who);
  VT_REQUIRE(B <= 65535 && heads <= 65535 && B * S < (1ll << 31), "%s: grid limits", who);
  return VT_OK;
}

static int attn_dbg() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("VT_ATTN_DBG"); v = e ? atoi(e) : 0; }
  return v;
}

static int key_tiles_impl(const float* bias, const int32_t* start, int64_t B, int64_t S, uint32_t* key_tiles, vt_stream stream) {
  VT_REQUIRE(bias && key_tiles && B > 0 && S > 0, "vt_attention_key_tiles: bad arguments");
  VT_REQUIRE(S <= 2048, "vt_attention_key_tiles: a plan is 32 tiles of 64 keys (S <= 2048); pass no plan for longer sequences");
  launch_pdl(key_tile_plan_kernel, (unsigned)B, 128, 0, (cudaStream_t)stream, bias, (int)S, (const int*)start, key_tiles);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_attention_key_tiles(const float* bias, int64_t B, int64_t S, uint32_t* key_tiles, vt_stream stream) {
  return key_tiles_impl(bias, nullptr, B, S, key_tiles, stream);
}
extern "C" int vt_attention_key_tiles_packed(const float* bias, const int32_t* start, int64_t B, int64_t S, uint32_t* key_tiles, vt_stream stream) {
  VT_REQUIRE(start, "vt_attention_key_tiles_packed: start required");
  return key_tiles_impl(bias, start, B, S, key_tiles, stream);
}

static int attention_fwd_impl(const void* qkv, int64_t ld, const float* bias, const uint32_t* key_tiles, const int32_t* start, int64_t rows_cap,
                              int64_t B, int64_t S, int64_t heads, int64_t head_dim, void* ctx, float* lse, float dropout_p, uint64_t seed,
                              uint32_t dropout_stream, vt_stream stream) {
  int rc = check_attn_tc("vt_attention_fwd", qkv, ld, B, S, heads, head_dim);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  rc = resolve_driver_entry_points();
  if (rc) return rc;
  VT_REQUIRE(((uintptr_t)ctx & 15) == 0, "vt_attention_fwd: ctx must be 16-byte aligned");
  VT_REQUIRE(S <= FWD_MAX_S, "vt_attention_fwd: the tcgen05 kernel keeps the mask bias of a sample in shared memory (S <= %d)", FWD_MAX_S);
  const int64_t hidden = heads * 64;
  CUtensorMap tq, tkv;
  const int64_t rows = start ? rows_cap : B * S;     // rows of the activation buffers (packed: the caller's capacity)
  VT_REQUIRE(rows > 0 && rows < (1ll << 31), "vt_attention_fwd: bad row capacity");
  rc = make_map(&tq, qkv, (uint64_t)(3 * hidden), (uint64_t)rows, (uint64_t)ld, 64, FQ);
  if (rc) return rc;
  rc = make_map(&tkv, qkv, (uint64_t)(3 * hidden), (uint64_t)rows, (uint64_t)ld, 64, FK);
  if (rc) return rc;
  AttnTcParams p = {};
  p.B = (int)B; p.S = (int)S; p.heads = (int)heads; p.hidden = (int)hidden;
  p.start = (const int*)start; p.rows_cap = rows;
  p.bias = bias; p.key_tiles = key_tiles; p.scale = 0.125f; p.out = (bf16*)ctx; p.lse = lse; p.ld = ld;
  p.drop = make_dropout(dropout_p, seed, dropout_stream);
  p.keep_bits = p.drop.thresh16 ? attention_keep_bits() : nullptr;
  p.dbg = attn_dbg();
  const int nkt = (int)((S + FK - 1) / FK);
  const size_t smem = 1024 + FWD_SMEM_TILES + 256 + 3072 + (size_t)nkt * FK * sizeof(float);
  static bool attr = false;
  if (!attr) {
    VT_CUDA(cudaFuncSetAttribute(attn_fwd_tc<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 1024 + FWD_SMEM_TILES + 256 + 3072 + FWD_MAX_S * 4 + 256));
    VT_CUDA(cudaFuncSetAttribute(attn_fwd_tc<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 1024 + FWD_SMEM_TILES + 256 + 3072 + FWD_MAX_S * 4 + 256));
    attr = true;
  }
  const long long n_items = ((S + FQ - 1) / FQ) * heads * B;
  static const int per_cta = [] { const char* e = getenv("VT_FWD_ITEMS_PER_CTA"); const int v = e ? atoi(e) : FWD_ITEMS_PER_CTA; return v > 0 ? v : 1; }();
  // each CTA a contiguous range of items; small launches (fewer CTAs than the two-per-SM slots) keep one item per CTA
  // ... and take two only when that does not cost a wave: ceil(items / k / slots) * k item-times, slots = two CTAs per SM
  const long long slots = 2ll * usable_sms();
  auto waves = [&](long long k) { return ((n_items + k - 1) / k + slots - 1) / slots * k; };
  const long long per = (n_items >= per_cta * slots && waves(per_cta) <= waves(1)) ? per_cta : 1;
  const long long ctas = (FWD_PERSISTENT || VT_DBG(32768)) ? 2ll * usable_sms() : (n_items + per - 1) / per;
  // one thread per row executes 28 % fewer instructions than two (ncu: 96.6 M vs 123.4 M warp instructions; the per-tile
  // bookkeeping and the maximum exchange are not duplicated) and measured faster with and without dropout
  // (197 vs 213 us at p = 0.1, 143 vs 147 us at p = 0); the two-thread variant stays selectable for ablations
  const bool two = VT_DBG(16384);
  if (two) launch_pdl(attn_fwd_tc<2>, (unsigned)(n_items < ctas ? n_items : ctas), 10 * 32, smem, st, tq, tkv, p);
  else launch_pdl(attn_fwd_tc<1>, (unsigned)(n_items < ctas ? n_items : ctas), 6 * 32, smem, st, tq, tkv, p);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_attention_fwd(const void* qkv, int64_t ld, const float* bias, const uint32_t* key_tiles, int64_t B, int64_t S, int64_t heads,
                                int64_t head_dim, void* ctx, float* lse, float dropout_p, uint64_t seed, uint32_t dropout_stream,
                                vt_stream stream) {
  return attention_fwd_impl(qkv, ld, bias, key_tiles, nullptr, 0, B, S, heads, head_dim, ctx, lse, dropout_p, seed, dropout_stream, stream);
}
extern "C" int vt_attention_fwd_packed(const void* qkv, int64_t ld, const float* bias, const uint32_t* key_tiles, const int32_t* start,
                                       int64_t rows_cap, int64_t B, int64_t S, int64_t heads, int64_t head_dim, void* ctx, float* lse,
                                       float dropout_p, uint64_t seed, uint32_t dropout_stream, vt_stream stream) {
  VT_REQUIRE(start, "vt_attention_fwd_packed: start required");
  return attention_fwd_impl(qkv, ld, bias, key_tiles, start, rows_cap, B, S, heads, head_dim, ctx, lse, dropout_p, seed, dropout_stream, stream);
}

static int attention_bwd_impl(const void* qkv, int64_t ld, const float* bias, const uint32_t* key_tiles, const int32_t* start, const int32_t* row_b,
                              int64_t rows_cap, const void* ctx, const void* dctx, const float* lse, float* dsum_ws, int64_t B, int64_t S,
                              int64_t heads, int64_t head_dim, void* dqkv, float dropout_p, uint64_t seed, uint32_t dropout_stream,
                              vt_stream stream) {
  int rc = check_attn_tc("vt_attention_bwd", qkv, ld, B, S, heads, head_dim);
  if (rc) return rc;
  VT_REQUIRE(((uintptr_t)dqkv & 15) == 0 && ((uintptr_t)ctx & 15) == 0 && ((uintptr_t)dctx & 15) == 0, "vt_attention_bwd: 16-byte alignment required");
  cudaStream_t st = (cudaStream_t)stream;
  rc = resolve_driver_entry_points();
  if (rc) return rc;
  const int64_t hidden = heads * 64;
  CUtensorMap tq, tdo;
  const int64_t rows_all = start ? rows_cap : B * S;
  VT_REQUIRE(rows_all > 0 && rows_all < (1ll << 31) && (!start || row_b), "vt_attention_bwd: bad row capacity / row map");
  rc = make_map(&tq, qkv, (uint64_t)(3 * hidden), (uint64_t)rows_all, (uint64_t)ld, 64, 128);
  if (rc) return rc;
  rc = make_map(&tdo, dctx, (uint64_t)hidden, (uint64_t)rows_all, (uint64_t)hidden, 64, 128);
  if (rc) return rc;
  {
    const long long prows = rows_all;
    long long pg = (prows + 7) / 8, cap = (long long)num_sms() * 8;
    attn_bwd_prep_kernel<<<(int)(pg < cap ? pg : cap), 256, 0, st>>>((const bf16*)dctx, (const bf16*)ctx, dsum_ws, (int)B, (int)S, (int)heads, (int)hidden,
                                                                     (const int*)start, (const int*)row_b, (long long)rows_all);
    count_launch(1);
    VT_CHECK_LAUNCH();
  }
  AttnTcParams p = {};
  p.B = (int)B; p.S = (int)S; p.heads = (int)heads; p.hidden = (int)hidden;
  const int nt = (int)((S + 127) / 128);
  // workspace: [B, heads, S] rowsum(dO.O), then the dQ partial slices [key tiles][B*S][hidden] (16-byte aligned)
  const size_t dsum_floats = ((size_t)(B * heads * S) + 3) & ~(size_t)3;
  // key_tiles is accepted and not used: the backward owns 128-key tiles, and at the loader's padding ([text <= 511 | regions <= 256])
  // a 128-key tile is dead as a whole for ~3 % of the items only.  Numbering the live items per sample (prefix table in shared
  // memory, binary search per item) was built and measured: 441 -> 445 us per layer, and the extra registers cost the dense case 6 %.
  (void)key_tiles;
  p.start = (const int*)start; p.row_b = (const int*)row_b; p.rows_cap = rows_all;
  p.bias = bias; p.scale = 0.125f; p.lse = const_cast<float*>(lse); p.dsum = dsum_ws; p.dq_ws = reinterpret_cast<bf16*>(dsum_ws + dsum_floats);
  p.dqkv = (bf16*)dqkv; p.ld = ld;
  p.drop = make_dropout(dropout_p, seed, dropout_stream);
  p.dbg = attn_dbg();
  VT_REQUIRE(((uintptr_t)dsum_ws & 15) == 0, "vt_attention_bwd: workspace must be 16-byte aligned");
  const size_t smem = 1024 + BWD_SMEM_TILES + 256 + 2 * 128 * sizeof(float);
  static bool attr = false;
  if (!attr) {
    VT_CUDA(cudaFuncSetAttribute(attn_bwd_fused<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(1024 + BWD_SMEM_TILES + 256 + 1024)));
    VT_CUDA(cudaFuncSetAttribute(attn_bwd_fused<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(1024 + BWD_SMEM_TILES + 256 + 1024)));
    attr = true;
  }
  const long long n_items = (long long)nt * heads * B;
  const int sms = usable_sms();  // leaves the SMs of an overlapped gradient exchange alone (vt_reserve_sms)
  p.keep_bits = p.drop.thresh16 ? attention_keep_bits() : nullptr;
  VT_REQUIRE(!p.keep_bits || B * heads * 4 * nt * S < (1ll << 31), "vt_attention_bwd: keep-bit store too large for 32-bit word offsets");
  {   // a schedule built for another grid (the SM reserve changed in between) is ignored: round-robin is always valid
    int sched_ctas = 0;
    const int32_t* sc = attention_bwd_schedule(&sched_ctas);
    p.sched = (start && sc && sched_ctas == (int)(n_items < sms ? n_items : sms)) ? sc : nullptr;
  }
  if (p.keep_bits || !p.drop.thresh16) launch_pdl(attn_bwd_fused<1>, (unsigned)(n_items < sms ? n_items : sms), BWD_THREADS, smem, st, tq, tdo, p);
  else launch_pdl(attn_bwd_fused<0>, (unsigned)(n_items < sms ? n_items : sms), BWD_THREADS, smem, st, tq, tdo, p);
  const long long rows = rows_all;
  const long long items = rows * (hidden / 8);
  launch_pdl(attn_dq_reduce_kernel, (unsigned)((items + 255) / 256), 256, 0, st, p.dq_ws, rows * hidden, nt, rows, (int)hidden,
                                                                         p.scale * (p.drop.thresh16 ? p.drop.scale : 1.0f), (bf16*)dqkv, ld,
                                                                         (const int*)start, (const int*)row_b);
  count_launch(2);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_attention_bwd(const void* qkv, int64_t ld, const float* bias, const uint32_t* key_tiles, const void* ctx, const void* dctx,
                                const float* lse, float* dsum_ws, int64_t B, int64_t S, int64_t heads, int64_t head_dim, void* dqkv,
                                float dropout_p, uint64_t seed, uint32_t dropout_stream, vt_stream stream) {
  return attention_bwd_impl(qkv, ld, bias, key_tiles, nullptr, nullptr, 0, ctx, dctx, lse, dsum_ws, B, S, heads, head_dim, dqkv, dropout_p, seed,
                            dropout_stream, stream);
}
extern "C" int64_t vt_attention_bwd_schedule_ctas(int64_t B, int64_t S, int64_t heads) {
  const int64_t n_items = ((S + 127) / 128) * heads * B;
  const int64_t sms = usable_sms();
  return n_items < sms ? n_items : sms;
}
extern "C" int vt_attention_bwd_schedule(const int32_t* start, int64_t B, int64_t S, int64_t heads, int32_t* sched, vt_stream stream) {
  VT_REQUIRE(start && sched && B > 0 && B <= 4000 && S > 0 && heads > 0, "vt_attention_bwd_schedule: bad arguments (B <= 4000)");
  const int ctas = (int)vt_attention_bwd_schedule_ctas(B, S, heads);
  attn_bwd_schedule_kernel<<<1, 256, (size_t)(B + 1) * sizeof(long long), (cudaStream_t)stream>>>((const int*)start, (int)B, (int)S, (int)heads, ctas, sched);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}
extern "C" int64_t vt_attention_bwd_workspace_packed(int64_t rows_cap, int64_t B, int64_t S, int64_t heads) {
  const int64_t dsum_floats = (B * heads * S + 3) & ~(int64_t)3;
  return dsum_floats + ((S + 127) / 128) * rows_cap * heads * 64 / 2;
}
extern "C" int vt_attention_bwd_packed(const void* qkv, int64_t ld, const float* bias, const uint32_t* key_tiles, const int32_t* start,
                                       const int32_t* row_b, int64_t rows_cap, const void* ctx, const void* dctx, const float* lse, float* dsum_ws,
                                       int64_t B, int64_t S, int64_t heads, int64_t head_dim, void* dqkv, float dropout_p, uint64_t seed,
                                       uint32_t dropout_stream, vt_stream stream) {
  VT_REQUIRE(start && row_b, "vt_attention_bwd_packed: start and row_b required");
  return attention_bwd_impl(qkv, ld, bias, key_tiles, start, row_b, rows_cap, ctx, dctx, lse, dsum_ws, B, S, heads, head_dim, dqkv, dropout_p, seed,
                            dropout_stream, stream);
}
