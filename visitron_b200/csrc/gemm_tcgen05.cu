// bf16 GEMM for sm_100a: TMA -> 128B-swizzled shared memory -> tcgen05.mma (fp32 accumulators in
// TMEM) -> tcgen05.ld epilogue.  One persistent CTA per SM, warp-specialised:
//   warp 0      TMA producer (one elected lane)
//   warp 1      MMA issuer   (one lane; tcgen05.mma is a single-thread instruction)
//   warp 2      TMEM allocator / deallocator
//   warps 4-11  epilogue: two warps per 32-lane TMEM quadrant, each taking half of the tile's columns
// Three mbarrier pipelines: smem full/empty (TMA <-> MMA), TMEM full/empty (MMA <-> epilogue; two
// accumulator stages so the epilogue of tile i overlaps the main loop of tile i+1).
//
// Operands may be K-major (contraction contiguous) or MN-major (row-major [K, MN]); the latter is
// what lets data-gradient GEMMs reuse the forward weight and weight-gradient GEMMs read dY and X
// in place, with no transposed copies in HBM (see include/visitron_b200.h).
#include <atomic>
#include <cstdlib>

#include "common.cuh"
#include "ptx.cuh"

// Timing-ablation switches (results are garbage): compiled in only with -DVT_ATTN_ABLATE (VT_BUILD_ABLATE=1), read from VT_GEMM_DBG:
// 1 epilogue only drains TMEM (no staging, arithmetic or stores), 2 no global stores, 4 no MMAs, 8 no TMA loads
#ifdef VT_ATTN_ABLATE
#define VT_GDBG(bit) ((p.dbg & (bit)) != 0)
#else
#define VT_GDBG(bit) false
#endif

namespace vt {

using namespace ptx;

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 bytes = one swizzle span
constexpr int UMMA_K = 16;
constexpr int NUM_EPI_WARPS = 8;
constexpr int GEMM_THREADS = 128 + NUM_EPI_WARPS * 32;

struct GemmParams {
  int M, N, K;
  int m_blocks, n_blocks, k_blocks, splits, k_blocks_per_split;
  void* D;
  long long ldd;
  int out_fp32, accumulate, act, vec_ok;
  float alpha;
  const float* alpha_dev;
  const float* bias;
  const bf16* residual;
  long long ldr;
  const bf16* aux;
  long long ldaux;
  bf16* aux_out;
  long long ldaux_out;
  DropoutCfg drop;
  long long remap_in, remap_out, remap_off;
  // device-resident row count of the A operand (vt_set_row_limit): limits M when A is [M, K], K when A is [K, M]
  const int32_t* rows_dev;
  // optional output row map (packed rows): logical row m is written to row remap_index[m], or nowhere when that is negative
  const int32_t* remap_index;
  // weight-gradient form only (A is [K, M]): colsum[m] += alpha * sum over k of A[k, m] -- the bias gradient of the layer whose
  // weight gradient this GEMM is, taken from the A tiles that are in shared memory anyway (16 more columns of ones on the B side)
  float* colsum;
  int dbg;
};

// CS = 1: the column-sum variant.  Every stage's B tile is followed by a block of bf16 ones (written once; TMA never touches it), so
// that for the first column block of a row block the MMA simply runs with N = BN + 16: accumulator columns BN .. BN + 15 all
// receive sum_k A[k, m].  (A separate N = 16 MMA per k step re-read the whole A tile from shared memory: +33 % on those tiles.)
// PAIR = 1: two CTAs of a cluster (the two SMs of a TPC) work on one 256 x BN tile with tcgen05.mma.cta_group::2: each stages its own
// 128 rows of A and HALF of the B tile (BN / 2 columns), the leader issues one M = 256 MMA that reads both shared memories and
// writes rows 0-127 of the accumulator to its own TMEM and rows 128-255 to the peer's.  Per SM the B traffic (L2 -> shared memory
// and shared memory -> tensor core) halves and the ring gets 6 stages instead of 4.
// What bounds the K = 768 GEMMs (timing ablations, VT_GEMM_DBG, FFN-up with the GELU + gelu' epilogue at M = 24.5 k: 133 us whole, 77 us
// with the epilogue reduced to draining TMEM, 95 us for the epilogue alone without MMAs and loads): the epilogue's ~17 arithmetic
// instructions per element, issued by two warps per scheduler at 77 % issue utilisation.  Tried without gain: 16 epilogue warps
// (same time: issue-bound, not latency-bound), packed f32x2 arithmetic (FFMA2 issues at half rate), CTA pairs for these shapes.
// The data-gradient GEMM with the saved-gelu' multiply ([M,768]x[768,3072], 119 us) is different: its epilogue alone takes 103 us with or
// without the stores at ~12 % issue utilisation, i.e. it waits for the per-element operand; pulling the next tile's slice of it into L2 a
// tile ahead (cp.async.bulk.prefetch.L2, 256 B per row) changed nothing (121.9 vs 119.3 us), so the loads' queueing is not DRAM page
// locality either -- open.
template <int BN, int CS = 0, int PAIR = 0>
struct GemmCfg {
  static_assert(!CS || BN <= 192, "the column-sum variant needs 16 accumulator columns next to the tile");
  static_assert(!PAIR || (!CS && BN == 256), "CTA pairs: 256 x 256 tiles, no column sums");
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = (PAIR ? BN / 2 : BN) * BK * 2;
  static constexpr int ONES_BYTES = CS ? 64 * BK * 2 : 0;   // one MN-major 64-column chunk (a K-major B needs only 16 rows of it)
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES + ONES_BYTES;
  static constexpr int ACC_STRIDE = !CS ? BN : (BN + 16 <= 128 ? 128 : 256);   // TMEM columns per accumulator stage
  static constexpr int EPI_BYTES = NUM_EPI_WARPS * 32 * 32 * 4;  // per-warp fp32 [32][32] transpose buffer (XOR-swizzled)
  static constexpr int BAR_BYTES = 256;
  static constexpr int BUDGET = 227 * 1024 - 1024 - BAR_BYTES - EPI_BYTES;  // 1024: slack for 1024-B alignment
  static constexpr int STAGES = BUDGET / STAGE_BYTES > 8 ? 8 : BUDGET / STAGE_BYTES;
  static constexpr int TMEM_NEED = 2 * ACC_STRIDE;
  static constexpr int TMEM_COLS = (TMEM_NEED <= 32) ? 32 : (TMEM_NEED <= 64) ? 64 : (TMEM_NEED <= 128) ? 128 : (TMEM_NEED <= 256) ? 256 : 512;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024;
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Epilogue kinds.  EPI_GENERIC reads every switch from GemmParams at run time (any output type, accumulate, row remap,
// unaligned operands); the others are the encoder's hot combinations with the switches as template constants -- less
// code (the generic epilogue alone is ~3500 SASS instructions, enough to miss in the instruction cache with eight
// epilogue warps at different places), fewer registers, and the per-element operand (residual or saved pre-activation)
// of the NEXT 32-column chunk is prefetched into registers while the current chunk is processed.
// The fast kinds require: bf16 output, no accumulate / split-K, no row remap, 16-byte aligned operands, N % 8 == 0.
enum { EPI_BIAS = 1, EPI_GELU = 2, EPI_GELU_BWD = 4, EPI_DROP = 8, EPI_RESID = 16, EPI_GENERIC = 32, EPI_GELU_SG = 64, EPI_MUL_AUX = 128 };

template <int EPI>
struct EpiK {
  static constexpr bool generic = (EPI & EPI_GENERIC) != 0;
  static constexpr bool ext = !generic && (EPI & (EPI_GELU_BWD | EPI_MUL_AUX | EPI_RESID)) != 0;  // prefetched per-element operand
};

// Epilogue of 8 consecutive columns of one output row, in the COALESCED layout: after the in-warp shared-memory
// transpose lane L owns columns [8 (L & 3), +8) of row (L >> 2) + 8 i, i = 0..3, of the 32 x 32 chunk, so one warp-wide
// 16-byte access covers 8 rows x 64 contiguous bytes (whole 32-byte sectors) of every bf16 operand -- saved
// pre-activation, residual, output -- instead of 32 rows x 16 bytes.  x arrives as the raw accumulator.
template <int EPI>
__device__ __forceinline__ void epilogue_rows8(const GemmParams& p, const DropoutCfg& drop, int M, int m, int n0, float* x, float alpha, const float* bias8,
                                               const uint4& ext) {
  constexpr bool G = EpiK<EPI>::generic;
  if (m >= M) return;
  const int ncols = G ? min(8, p.N - n0) : 8;
  if (n0 >= p.N) return;
  long long orow = m;
  if (G && p.remap_in > 0) orow = (long long)(m / p.remap_in) * p.remap_out + p.remap_off + (m % p.remap_in);
  if (G && p.remap_index) {
    orow = __ldg(p.remap_index + orow);
    if (orow < 0) return;
  }
  const bool vec = G ? (p.vec_ok && ncols == 8) : true;
  // x arrives as the raw accumulator: alpha and the bias are one fma (alpha == 1: the same bits as acc + bias)
  if (G ? (p.bias != nullptr) : (EPI & EPI_BIAS) != 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = fmaf(x[j], alpha, bias8[j]);
  } else if (alpha != 1.0f) {
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] *= alpha;
  }
  const bool save_grad = G ? (p.act == VT_ACT_GELU_SAVE_GRAD) : (EPI & EPI_GELU_SG) != 0;
  if (save_grad) {   // D = gelu(x), aux_out = gelu'(x)
    float dy[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) gelu_fwd_grad(x[j], x[j], dy[j]);
    bf16* ao = p.aux_out + (long long)m * p.ldaux_out + n0;
    if (VT_GDBG(2) && dy[0] != 1234.5f) {
    } else if (vec) {
      *reinterpret_cast<uint4*>(ao) = pack8(dy);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j < ncols) ao[j] = __float2bfloat16(dy[j]);
    }
  } else if (G ? (p.aux_out != nullptr) : (EPI & EPI_GELU) != 0) {
    bf16* ao = p.aux_out + (long long)m * p.ldaux_out + n0;
    if (vec) {
      *reinterpret_cast<uint4*>(ao) = pack8(x);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j < ncols) ao[j] = __float2bfloat16(x[j]);
    }
  }
  if (save_grad) {
  } else if (G ? (p.act == VT_ACT_GELU) : (EPI & EPI_GELU) != 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = gelu_fwd(x[j]);
  } else if (G && p.act == VT_ACT_TANH) {
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = tanhf(x[j]);
  } else if (G ? (p.act == VT_ACT_GELU_BWD || p.act == VT_ACT_TANH_BWD || p.act == VT_ACT_MUL_AUX) : (EPI & (EPI_GELU_BWD | EPI_MUL_AUX)) != 0) {
    float a[8];
    if (!G) {
      unpack8(ext, a);
    } else {
      const bf16* ax = p.aux + (long long)m * p.ldaux + n0;
      if (vec) {
        unpack8(__ldg(reinterpret_cast<const uint4*>(ax)), a);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = (j < ncols) ? __bfloat162float(ax[j]) : 0.f;
      }
    }
    if (G ? (p.act == VT_ACT_MUL_AUX) : (EPI & EPI_MUL_AUX) != 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] *= a[j];
    } else if (!G || p.act == VT_ACT_GELU_BWD) {
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] *= gelu_grad(a[j]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] *= (1.0f - a[j] * a[j]);
    }
  }
  if ((G || (EPI & EPI_DROP) != 0) && drop.thresh16) {
    // linear index of the LOGICAL element (m, n); N % 8 == 0 is enforced on the host
    float mult[8];
    dropout_mult8(drop, (uint64_t)m * (uint64_t)p.N + (uint64_t)n0, mult);
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] *= mult[j];
  }
  if (G ? (p.residual != nullptr) : (EPI & EPI_RESID) != 0) {
    float t[8];
    if (!G) {
      unpack8(ext, t);
    } else {
      const bf16* r = p.residual + orow * p.ldr + n0;
      if (vec) {
        unpack8(__ldg(reinterpret_cast<const uint4*>(r)), t);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) t[j] = (j < ncols) ? __bfloat162float(r[j]) : 0.f;
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] += t[j];
  }
  if (G && p.out_fp32) {
    float* d = reinterpret_cast<float*>(p.D) + orow * p.ldd + n0;
    if (p.accumulate) {
      if (vec) {
        red_add_v4(d, x[0], x[1], x[2], x[3]);
        red_add_v4(d + 4, x[4], x[5], x[6], x[7]);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j < ncols) atomicAdd(d + j, x[j]);
      }
    } else {
      if (vec) {
        reinterpret_cast<float4*>(d)[0] = make_float4(x[0], x[1], x[2], x[3]);
        reinterpret_cast<float4*>(d)[1] = make_float4(x[4], x[5], x[6], x[7]);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j < ncols) d[j] = x[j];
      }
    }
  } else {
    bf16* d = reinterpret_cast<bf16*>(p.D) + orow * p.ldd + n0;
    if (VT_GDBG(2) && x[0] != 1234.5f) {
    } else if (vec) {
      *reinterpret_cast<uint4*>(d) = pack8(x);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j < ncols) d[j] = __float2bfloat16(x[j]);
    }
  }
}

template <int BN, int A_MN, int B_MN, int EPI, int CS = 0, int PAIR = 0>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
  static_assert(!CS || A_MN, "column sums are taken over the rows of a [K, M] A operand");
  using Cfg = GemmCfg<BN, CS, PAIR>;
  const uint32_t rank = PAIR ? pair_rank() : 0u;   // CTA of the pair: 0 = leader (issues the MMAs, owns the full / accumulator-empty barriers)
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t epi_base = smem_base + STAGES * Cfg::STAGE_BYTES;
  const uint32_t bar_base = epi_base + Cfg::EPI_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmA);
    prefetch_tensormap(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), NUM_EPI_WARPS * (PAIR ? 2 : 1));   // pair: the leader's barrier counts both CTAs' epilogue warps
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    if constexpr (PAIR) { tmem_alloc_pair(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish_pair(); }
    else { tmem_alloc(tmem_slot, Cfg::TMEM_COLS); tmem_relinquish(); }
  }
  if (CS && warp == 3) {   // bf16 1.0 everywhere behind every stage's B tile: no swizzle or major can misplace an element
    constexpr uint32_t one2 = 0x3f803f80u;
    for (int s = 0; s < STAGES; ++s) {
      const uint32_t ob = smem_base + s * Cfg::STAGE_BYTES + Cfg::A_BYTES + Cfg::B_BYTES;
      for (int i = lane; i < Cfg::ONES_BYTES / 16; i += 32) st_shared_v4(ob + 16u * (uint32_t)i, one2, one2, one2, one2);
    }
    fence_proxy_async();
  }
  tc_fence_before();
  if constexpr (PAIR) pair_sync();   // the peer's barriers exist before anything (its TMA, the leader's commits) can arrive on them
  else __syncthreads();
  tc_fence_after();
  pdl_sync();   // the prologue above overlapped the tail of the previous kernel; global memory is touched only below
  const uint32_t tmem_base = *tmem_slot_ptr;

  // tile counts: the host's, or recomputed from the device-resident row count of A (sync-free steps / CUDA graphs: the number
  // of labelled rows the pre-training heads run on is only known on the device); the grid was sized for the host's bound
  int M = p.M, m_blocks = p.m_blocks, k_blocks = p.k_blocks, kbps = p.k_blocks_per_split, splits = p.splits;
  if (p.rows_dev) {
    const int lim = __ldg(p.rows_dev);
    if (A_MN) {
      const int K = min(p.K, lim);
      k_blocks = (K + BK - 1) / BK;
      kbps = (k_blocks + p.splits - 1) / p.splits;
      splits = kbps > 0 ? (k_blocks + kbps - 1) / kbps : 0;
    } else {
      M = min(p.M, lim);
      m_blocks = (M + BM - 1) / BM;
    }
  }
  if constexpr (PAIR) m_blocks = (M + 2 * BM - 1) / (2 * BM);   // a pair's tile is 256 rows: this CTA's 128-row block is 2 * (tile row) + rank
  const int mn_tiles = m_blocks * p.n_blocks;
  const int total_tiles = mn_tiles * splits;
  const int t_first = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, t_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  auto row_block = [&](int mn) { return PAIR ? 2 * (mn / p.n_blocks) + (int)rank : mn / p.n_blocks; };

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = t_first; t < total_tiles; t += t_step) {
        const int split = t / mn_tiles, mn = t % mn_tiles;
        const int m_blk = row_block(mn), n_blk = mn % p.n_blocks;
        const int kb0 = split * kbps;
        const int kb1 = min(kb0 + kbps, k_blocks);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
          const uint32_t sb = sa + Cfg::A_BYTES;
          if constexpr (PAIR) {
            // both CTAs load into their own shared memory; all bytes of the stage (2 x (A + half B)) are counted on the LEADER's barrier
            const uint32_t lbar = pair_map(full_bar(stage), 0);
            if (rank == 0) mbar_arrive_expect_tx(full_bar(stage), 2 * (Cfg::A_BYTES + Cfg::B_BYTES));
            if (A_MN) {
#pragma unroll
              for (int c = 0; c < BM / 64; ++c) tma_load_2d_pair(sa + c * (64 * BK * 2), &tmA, lbar, m_blk * BM + c * 64, kb * BK);
            } else {
              tma_load_2d_pair(sa, &tmA, lbar, kb * BK, m_blk * BM);
            }
            const int n0 = n_blk * BN + (int)rank * (BN / 2);
            if (B_MN) {
#pragma unroll
              for (int c = 0; c < BN / 128; ++c) tma_load_2d_pair(sb + c * (64 * BK * 2), &tmB, lbar, n0 + c * 64, kb * BK);
            } else {
              tma_load_2d_pair(sb, &tmB, lbar, kb * BK, n0);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1u; }
            continue;
          }
          if (VT_GDBG(8)) { mbar_arrive(full_bar(stage)); if (++stage == STAGES) { stage = 0; phase ^= 1u; } continue; }
          mbar_arrive_expect_tx(full_bar(stage), Cfg::A_BYTES + Cfg::B_BYTES);   // the ones block behind B is not TMA's
          if (A_MN) {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c) tma_load_2d(sa + c * (64 * BK * 2), &tmA, full_bar(stage), m_blk * BM + c * 64, kb * BK);
          } else {
            tma_load_2d(sa, &tmA, full_bar(stage), kb * BK, m_blk * BM);
          }
          if (B_MN) {
#pragma unroll
            for (int c = 0; c < BN / 64; ++c) tma_load_2d(sb + c * (64 * BK * 2), &tmB, full_bar(stage), n_blk * BN + c * 64, kb * BK);
          } else {
            tma_load_2d(sb, &tmB, full_bar(stage), kb * BK, n_blk * BN);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(PAIR ? 2 * BM : BM, BN, A_MN, B_MN);
      constexpr uint32_t idesc_cs = make_idesc_bf16(BM, BN + 16, A_MN, B_MN);   // the tile plus 16 columns of ones
      // K-major, 128B swizzle: 8-row groups 1024 B apart (SBO); LBO unused (encoded 1).
      // MN-major, 128B swizzle: 64-element MN chunks one TMA box (64*BK*2 B) apart (LBO),
      //                         8-row K groups 1024 B apart (SBO).
      constexpr uint32_t A_LBO = A_MN ? 64 * BK * 2 : 16, B_LBO = B_MN ? 64 * BK * 2 : 16;
      constexpr uint32_t A_KSTEP = A_MN ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;  // descriptor units of 16 B
      constexpr uint32_t B_KSTEP = B_MN ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int t = t_first; t < total_tiles; t += t_step) {
        const int split = t / mn_tiles;
        const int kb0 = split * kbps;
        const int kb1 = min(kb0 + kbps, k_blocks);
        const uint32_t idesc_t = (CS && ((t % mn_tiles) % p.n_blocks) == 0) ? idesc_cs : idesc;   // once per (row block of A, split)
        mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * Cfg::ACC_STRIDE);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = smem_base + stage * Cfg::STAGE_BYTES;
          const uint64_t a_desc = make_smem_desc(sa, A_LBO, 1024);
          const uint64_t b_desc = make_smem_desc(sa + Cfg::A_BYTES, B_LBO, 1024);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            if (VT_GDBG(4)) break;
            if constexpr (PAIR) umma_bf16_pair(d_tmem, a_desc + (uint64_t)(k * A_KSTEP), b_desc + (uint64_t)(k * B_KSTEP), idesc_t, (kb > kb0 || k > 0) ? 1u : 0u);
            else umma_bf16(d_tmem, a_desc + (uint64_t)(k * A_KSTEP), b_desc + (uint64_t)(k * B_KSTEP), idesc_t, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          if constexpr (PAIR) umma_commit_pair(empty_bar(stage));   // frees the stage in BOTH CTAs
          else umma_commit(empty_bar(stage));
          if (++stage == STAGES) { stage = 0; phase ^= 1u; }
        }
        if constexpr (PAIR) umma_commit_pair(tfull_bar(acc));       // both CTAs' epilogue warps read their half of the rows
        else umma_commit(tfull_bar(acc));
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1u;
      }
    }
  } else if (warp >= 4) {
    constexpr bool EXT = EpiK<EPI>::ext;
    constexpr bool G = EpiK<EPI>::generic;
    const int q = warp & 3;            // TMEM lane quadrant this warp may read
    const int half = (warp - 4) >> 2;  // which half of the tile's columns
    constexpr int CHUNKS = BN / 32;
    constexpr int CH_PER_HALF = (CHUNKS + 1) / 2;
    const int c_begin = half * CH_PER_HALF, c_end = min(CHUNKS, c_begin + CH_PER_HALF);
    const uint32_t stg = epi_base + (uint32_t)(warp - 4) * 4096u;
    // transposed ownership (see epilogue_rows8): row r8 + 8 i, 16-byte fp32 chunks 2 c4 and 2 c4 + 1 of the staged row
    const int r8 = lane >> 2, c4 = lane & 3;
    float alpha = p.alpha;
    if (p.alpha_dev) alpha *= __ldg(p.alpha_dev);
    const DropoutCfg drop = resolve_dropout(p.drop);
    const bf16* ext_base = nullptr;
    long long ext_ld = 0;
    if (EXT) {
      ext_base = (EPI & EPI_RESID) ? p.residual : p.aux;
      ext_ld = (EPI & EPI_RESID) ? p.ldr : p.ldaux;
    }
    // the per-element operand of chunk c of tile (m_blk, n_blk) in the coalesced layout: 4 rows x 16 bytes per lane
    auto load_ext = [&](uint4* e, int m_blk, int n_blk, int c) {
      const int n0 = n_blk * BN + c * 32 + 8 * c4;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int m = m_blk * BM + q * 32 + r8 + 8 * i;
        e[i] = (m < M && n0 < p.N) ? __ldg(reinterpret_cast<const uint4*>(ext_base + (long long)m * ext_ld + n0)) : make_uint4(0u, 0u, 0u, 0u);
      }
    };
    int acc = 0;
    uint32_t acc_phase = 0;
    uint4 ext[4] = {}, ext_next[4] = {};
    const uint32_t tempty_leader = PAIR ? pair_map(tempty_bar(0), 0) : 0u;
    if (EXT && t_first < total_tiles) {
      const int mn = t_first % mn_tiles;
      load_ext(ext, row_block(mn), mn % p.n_blocks, c_begin);
    }
    for (int t = t_first; t < total_tiles; t += t_step) {
      const int mn = t % mn_tiles;
      const int m_blk = row_block(mn), n_blk = mn % p.n_blocks;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * Cfg::ACC_STRIDE);
      uint32_t v[32];
      tmem_ld_32x32(t_addr + (uint32_t)(c_begin * 32), v);
      if (CS && n_blk == 0 && half == 0) {
        // the column sums of this row block of A over the tile's k range: one value per TMEM lane, i.e. per m
        uint32_t cv[4];
        tmem_ld_32x4(t_addr + (uint32_t)BN, cv);
        tmem_ld_wait();
        const int m = m_blk * BM + q * 32 + lane;
        if (m < M) atomicAdd(p.colsum + m, __uint_as_float(cv[0]) * alpha);
      }
      tmem_ld_wait();
      // With a per-element operand the chunk loop is unrolled over the half tile: rolled up, the hand-over `ext = ext_next` at the end
      // of a chunk is a register move that waits for the prefetch issued in the SAME chunk -- one full memory latency per chunk (the
      // MUL_AUX epilogue alone took 103 us at 12 % issue utilisation); unrolled, the compiler renames the registers and the loads
      // really run a chunk ahead.
      constexpr int CHUNK_UNROLL = EXT ? CH_PER_HALF : 1;
#pragma unroll CHUNK_UNROLL
      for (int c = c_begin; c < c_end; ++c) {
        if (VT_GDBG(1)) {
          if (c + 1 < c_end) { tmem_ld_32x32(t_addr + (uint32_t)((c + 1) * 32), v); tmem_ld_wait(); }
          else {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { if constexpr (PAIR) mbar_arrive_remote(pair_map(tempty_bar(acc), 0)); else mbar_arrive(tempty_bar(acc)); }
          }
          continue;
        }
        // row-per-lane (TMEM) layout -> staging: lane = row, 16-byte chunk j at physical chunk j ^ (row & 7)
#pragma unroll
        for (int j = 0; j < 8; ++j)
          st_shared_v4(stg + (uint32_t)lane * 128u + ((uint32_t)(j ^ (lane & 7)) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        __syncwarp();
        if (c + 1 < c_end) {
          tmem_ld_32x32(t_addr + (uint32_t)((c + 1) * 32), v);  // in flight while this chunk is post-processed
          if (EXT) load_ext(ext_next, m_blk, n_blk, c + 1);
        } else {
          // the accumulator stage is drained: hand it back to the MMA warp before the post-processing
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if constexpr (PAIR) mbar_arrive_remote(tempty_leader + 8u * (uint32_t)acc);
            else mbar_arrive(tempty_bar(acc));
          }
          if (EXT && t + t_step < total_tiles) {
            const int mn2 = (t + t_step) % mn_tiles;
            load_ext(ext_next, row_block(mn2), mn2 % p.n_blocks, c_begin);
          }
        }
        const int n0 = n_blk * BN + c * 32 + 8 * c4;
        float bias8[8];
        if (G ? (p.bias != nullptr) : (EPI & EPI_BIAS) != 0) {
          if ((!G || p.vec_ok) && n0 + 8 <= p.N) {
            const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + n0)), b1 = __ldg(reinterpret_cast<const float4*>(p.bias + n0) + 1);
            bias8[0] = b0.x; bias8[1] = b0.y; bias8[2] = b0.z; bias8[3] = b0.w;
            bias8[4] = b1.x; bias8[5] = b1.y; bias8[6] = b1.z; bias8[7] = b1.w;
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) bias8[j] = (n0 + j < p.N) ? __ldg(p.bias + n0 + j) : 0.f;
          }
        }
        // all four rows out of the staging buffer first, then the arithmetic of the four rows (independent chains)
        uint4 lo[4], hi[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int row = r8 + 8 * i;
          const uint32_t ra = stg + (uint32_t)row * 128u;
          lo[i] = ld_shared_v4(ra + ((uint32_t)((2 * c4) ^ (row & 7)) << 4));
          hi[i] = ld_shared_v4(ra + ((uint32_t)((2 * c4 + 1) ^ (row & 7)) << 4));
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float x[8] = {__uint_as_float(lo[i].x), __uint_as_float(lo[i].y), __uint_as_float(lo[i].z), __uint_as_float(lo[i].w),
                        __uint_as_float(hi[i].x), __uint_as_float(hi[i].y), __uint_as_float(hi[i].z), __uint_as_float(hi[i].w)};
          epilogue_rows8<EPI>(p, drop, M, m_blk * BM + q * 32 + r8 + 8 * i, n0, x, alpha, bias8, ext[i]);
        }
        if (EXT) {
#pragma unroll
          for (int i = 0; i < 4; ++i) ext[i] = ext_next[i];
        }
        if (c + 1 < c_end) tmem_ld_wait();
        __syncwarp();  // staging buffer is rewritten by the next chunk
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1u;
    }
  }

  tc_fence_before();
  if constexpr (PAIR) pair_sync();   // the leader's MMAs read the peer's shared memory and write its TMEM: neither CTA leaves before both are done
  else __syncthreads();
  if (warp == 2) {
    if constexpr (PAIR) tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS);
    else tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static std::atomic<long long> g_launches{0};

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int resolve_driver_entry_points() {
  if (g_encode) return VT_OK;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  VT_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  if (qres != cudaDriverEntryPointSuccess || !fn) {
    set_error("cuTensorMapEncodeTiled not available from the driver");
    return VT_ERR_CUDA;
  }
  g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  return VT_OK;
}

// 2-D bf16 tensor map: `inner` contiguous elements per row, `outer` rows, row pitch ld elements.
int make_map(CUtensorMap* m, const void* base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner, uint32_t box_outer) {
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) base=%p inner=%llu outer=%llu ld=%llu box=%ux%u", (int)r, base,
              (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)ld, box_inner, box_outer);
    return VT_ERR_CUDA;
  }
  return VT_OK;
}

template <int BN, int A_MN, int B_MN, int EPI = EPI_GENERIC, int CS = 0>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid, cudaStream_t st) {
  using Cfg = GemmCfg<BN, CS>;
  static bool attr_set = false;
  auto kern = gemm_bf16_kernel<BN, A_MN, B_MN, EPI, CS>;
  if (!attr_set) {
    VT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set = true;
  }
  launch_pdl(kern, grid, GEMM_THREADS, Cfg::SMEM_BYTES, st, ta, tb, p);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

// CTA-pair launch (see GemmCfg): a cluster of two CTAs per 256 x 256 tile, grid = 2 x min(tiles, SMs / 2)
template <int B_MN, int EPI, int A_MN = 0>
static int launch_gemm_pair(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int sms, cudaStream_t st) {
  using Cfg = GemmCfg<256, 0, 1>;
  static bool attr_set = false;
  auto kern = gemm_bf16_kernel<256, A_MN, B_MN, EPI, 0, 1>;
  if (!attr_set) {
    VT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set = true;
  }
  const long long tiles = (long long)((p.M + 2 * BM - 1) / (2 * BM)) * p.n_blocks * p.splits;
  const long long pairs = tiles < sms / 2 ? tiles : sms / 2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(2 * pairs));
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  VT_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, p));
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

template <int BN>
static int dispatch_major(int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid, cudaStream_t st) {
  if constexpr (BN <= 192) {
    if (p.colsum) return b_mn ? launch_gemm<BN, 1, 1, EPI_GENERIC, 1>(ta, tb, p, grid, st) : launch_gemm<BN, 1, 0, EPI_GENERIC, 1>(ta, tb, p, grid, st);
  }
  if (!a_mn && !b_mn) return launch_gemm<BN, 0, 0>(ta, tb, p, grid, st);
  if (!a_mn && b_mn) return launch_gemm<BN, 0, 1>(ta, tb, p, grid, st);
  if (a_mn && b_mn) return launch_gemm<BN, 1, 1>(ta, tb, p, grid, st);
  return launch_gemm<BN, 1, 0>(ta, tb, p, grid, st);
}

static int g_force_bn = 0;
// -1: CTA pairs where they pay (large M), 0: never, 1: wherever the kernel supports them (vt_debug_pair_mode; VT_GEMM_PAIR at load time)
static int g_pair_mode = [] { const char* e = getenv("VT_GEMM_PAIR"); return e ? atoi(e) : -1; }();

}  // namespace vt

using namespace vt;

extern "C" void vt_debug_force_block_n(int bn) { vt::g_force_bn = bn; }
extern "C" void vt_debug_pair_mode(int mode) { vt::g_pair_mode = mode; }
extern "C" int64_t vt_launch_count(void) { return (int64_t)vt::g_launches.load(); }
extern "C" void vt_reset_launch_count(void) { vt::g_launches.store(0); }

extern "C" int vt_gemm_bf16(const void* A, int64_t lda, int a_major, const void* B, int64_t ldb, int b_major, void* D,
                            int64_t ldd, int64_t M, int64_t N, int64_t K, const vt_gemm_epilogue* ep, vt_stream stream) {
  int rc = resolve_driver_entry_points();
  if (rc) return rc;
  VT_REQUIRE(ep, "vt_gemm_bf16: null epilogue");
  if (M == 0 || N == 0) return VT_OK;   // e.g. an objective without a single labelled row: nothing to compute (pointers may be null)
  if (K == 0) {                          // empty contraction (weight gradient over zero rows): the product is zero
    if (ep->accumulate) return VT_OK;
    VT_REQUIRE(D, "vt_gemm_bf16: null output");
    VT_CUDA(cudaMemset2DAsync(D, (size_t)ldd * (ep->out_fp32 ? 4 : 2), 0, (size_t)N * (ep->out_fp32 ? 4 : 2), (size_t)M, (cudaStream_t)stream));
    return VT_OK;
  }
  VT_REQUIRE(A && B && D, "vt_gemm_bf16: null pointer");
  VT_REQUIRE(M > 0 && N > 0 && K > 0 && M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "vt_gemm_bf16: bad sizes M=%lld N=%lld K=%lld", (long long)M, (long long)N, (long long)K);
  VT_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "vt_gemm_bf16: lda/ldb must be multiples of 8 elements (TMA 16-byte pitch), got %lld %lld", (long long)lda, (long long)ldb);
  VT_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "vt_gemm_bf16: A/B must be 16-byte aligned");
  VT_REQUIRE(lda >= (a_major ? M : K) && ldb >= (b_major ? N : K), "vt_gemm_bf16: leading dimension smaller than the row");
  VT_REQUIRE(!ep->accumulate || ep->out_fp32, "vt_gemm_bf16: accumulate needs fp32 output");
  VT_REQUIRE(ep->split_k <= 1 || ep->accumulate, "vt_gemm_bf16: split_k > 1 needs accumulate");
  VT_REQUIRE(ep->dropout_p <= 0.f || N % 8 == 0, "vt_gemm_bf16: dropout epilogue needs N %% 8 == 0");
  VT_REQUIRE(ep->dropout_p < 1.f, "vt_gemm_bf16: dropout_p must be < 1");
  VT_REQUIRE(!((ep->act == VT_ACT_GELU_BWD || ep->act == VT_ACT_TANH_BWD || ep->act == VT_ACT_MUL_AUX) && !ep->aux), "vt_gemm_bf16: *_BWD / MUL_AUX activation needs aux");
  VT_REQUIRE(!(ep->act == VT_ACT_GELU_SAVE_GRAD && !ep->aux_out), "vt_gemm_bf16: VT_ACT_GELU_SAVE_GRAD needs aux_out");
  VT_REQUIRE(ep->act >= VT_ACT_NONE && ep->act <= VT_ACT_MUL_AUX, "vt_gemm_bf16: unknown activation %d", ep->act);
  const bool splitting = ep->split_k > 1 || (ep->split_k == 0 && ep->accumulate);
  VT_REQUIRE(!splitting || (!ep->bias && ep->act == VT_ACT_NONE && !ep->residual && ep->dropout_p <= 0.f && !ep->aux_out),
             "vt_gemm_bf16: split-K supports only the alpha epilogue");

  const int sms = usable_sms();  // see vt_reserve_sms
  // Weight gradients dW[M, N] = A^T B over K = rows, both operands MN-major: the timing ablations (VT_GEMM_DBG, tools/r2_call35.sh)
  // show them bound by operand traffic -- the TMA loads alone take 84 of 108 us for dW[3072, 768] at 24.5 k rows, 18 TB/s out of
  // L2 with 128 x 192 tiles.  A CTA pair's 256 x 256 tile moves 40 % fewer bytes per flop.  Not with the column-sum side product
  // (its 16 extra B columns do not fit a pair's N <= 256): those launches stay single-CTA.
  const bool wgrad_pair = a_major && b_major && ep->accumulate && ep->out_fp32 && ep->split_k != 1 && !ep->colsum && !g_force_bn && sms >= 2 &&
                          (g_pair_mode == 1 || (g_pair_mode < 0 && K >= 4096 && M >= 512 && N >= 512));
  int BN = 256;
  const int m_blocks = (int)((M + BM - 1) / BM);
  if (g_force_bn) {
    BN = g_force_bn;
  } else {
    if (N <= 64) BN = 64;
    else if (N <= 128) BN = 128;
    else if (wgrad_pair) BN = 256;
    else if (ep->accumulate && ep->split_k != 1) BN = (N % 192 == 0) ? 192 : 256;  // weight gradients: split-K fills the SMs
    else {
      // widest tile that keeps every SM busy with the least padded columns (192 divides 768/2304/3072)
      const int cand[4] = {256, 192, 128, 64};
      double best = -1.0;
      for (int i = 0; i < 4; ++i) {
        const long long nb = (N + cand[i] - 1) / cand[i];
        const double fill = (double)N / (double)(nb * cand[i]);
        const long long tiles = (long long)m_blocks * nb;
        const double occ = tiles >= sms ? 1.0 : (double)tiles / sms;
        const double mma_eff = cand[i] >= 128 ? 1.0 : 0.7;  // 128x64 tiles are shared-memory bound
        const double score = fill * occ * mma_eff;
        if (score > best + 1e-9) { best = score; BN = cand[i]; }
      }
    }
  }
  if (ep->colsum) {
    VT_REQUIRE(a_major == 1, "vt_gemm_bf16: colsum is the column sum of a [K, M] A operand (weight-gradient form)");
    if (BN == 256 && !g_force_bn) BN = (N % 192 == 0) ? 192 : 128;
    VT_REQUIRE(BN != 256, "vt_gemm_bf16: colsum needs block N <= 192 (TMEM columns)");
  }
  VT_REQUIRE(BN == 64 || BN == 128 || BN == 192 || BN == 256, "vt_gemm_bf16: bad block N %d", BN);

  GemmParams p;
  p.M = (int)M; p.N = (int)N; p.K = (int)K;
  p.m_blocks = m_blocks;
  p.n_blocks = (int)((N + BN - 1) / BN);
  p.k_blocks = (int)((K + BK - 1) / BK);
  int splits = ep->split_k;
  if (splits <= 0) {
    splits = 1;
    if (ep->accumulate) {
      // the split count whose work items fill whole waves of the SMs best (96 tiles: 3 splits = 288 items = 1.95 waves,
      // where floor(148 / 96) = 1 would leave a third of the SMs idle); ties go to fewer splits (less red.add traffic)
      const long long tiles = wgrad_pair ? (long long)((p.m_blocks + 1) / 2) * p.n_blocks : (long long)p.m_blocks * p.n_blocks;
      const int units = wgrad_pair ? sms / 2 : sms;   // CTA pairs: a work item occupies two SMs
      const int max_splits = p.k_blocks / 8 < 1 ? 1 : (p.k_blocks / 8 > 16 ? 16 : p.k_blocks / 8);
      double best = -1.0;
      for (int s = 1; s <= max_splits; ++s) {
        const long long items = tiles * s;
        const long long waves = (items + units - 1) / units;
        const double eff = (double)items / (double)(waves * units) - 0.004 * (s - 1);
        if (eff > best + 1e-9) { best = eff; splits = s; }
      }
    }
  }
  if (splits > p.k_blocks) splits = p.k_blocks;
  p.k_blocks_per_split = (p.k_blocks + splits - 1) / splits;
  p.splits = (p.k_blocks + p.k_blocks_per_split - 1) / p.k_blocks_per_split;
  p.D = D; p.ldd = ldd;
  p.out_fp32 = ep->out_fp32; p.accumulate = ep->accumulate; p.act = ep->act;
  p.alpha = ep->alpha; p.alpha_dev = ep->alpha_dev; p.bias = ep->bias;
  p.residual = (const bf16*)ep->residual; p.ldr = ep->ldr;
  p.aux = (const bf16*)ep->aux; p.ldaux = ep->ldaux;
  p.aux_out = (bf16*)ep->aux_out; p.ldaux_out = ep->ldaux_out;
  p.drop = make_dropout(ep->dropout_p, ep->seed, ep->dropout_stream);
  p.remap_in = ep->remap_rows_in; p.remap_out = ep->remap_rows_out; p.remap_off = ep->remap_row_off;
  p.rows_dev = row_limit();
  p.remap_index = ep->remap_index;
  p.colsum = ep->colsum;
  static const int dbg = [] { const char* e = getenv("VT_GEMM_DBG"); return e ? atoi(e) : 0; }();
  p.dbg = dbg;
  // 16-byte vector path: every row slice the epilogue touches must be 16-byte aligned
  const int ob = p.out_fp32 ? 4 : 2;
  bool vec = ((uintptr_t)D % 16 == 0) && ((ldd * ob) % 16 == 0);
  if (p.bias) vec = vec && ((uintptr_t)p.bias % 16 == 0);
  if (p.residual) vec = vec && ((uintptr_t)p.residual % 16 == 0) && (p.ldr % 8 == 0);
  if (p.aux) vec = vec && ((uintptr_t)p.aux % 16 == 0) && (p.ldaux % 8 == 0);
  if (p.aux_out) vec = vec && ((uintptr_t)p.aux_out % 16 == 0) && (p.ldaux_out % 8 == 0);
  p.vec_ok = vec ? 1 : 0;

  // CTA pairs: [M, K] A operand, 256-wide tiles, one pass over K, no output row maps; by default only where a launch is many waves of
  // tiles (below ~4 k rows the 256-row granularity and the cluster launch cost more than the halved B traffic gains)
  const bool pair_ok = !a_major && BN == 256 && p.splits == 1 && !p.accumulate && p.remap_in <= 0 && !p.remap_index && sms >= 2;
  // measured at M = 24.5 k (tools/r2_call25.sh and later): +3 .. +6 % where the main loop bounds the tile (K >= 2304, or N >= 2304 with a light
  // epilogue); the small square shapes [M,768]x[768,768] are neutral (dropout + residual epilogue) to +6 % (plain) since the accumulator
  // hand-back is a relaxed arrive (it cost them 4 % before)
  const bool pair = wgrad_pair || (pair_ok && (g_pair_mode == 1 || (g_pair_mode < 0 && M >= 4096 && !g_force_bn)));
  CUtensorMap ta, tb;
  if (a_major) rc = make_map(&ta, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, BK);
  else rc = make_map(&ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM);
  if (rc) return rc;
  if (b_major) rc = make_map(&tb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, BK);
  else rc = make_map(&tb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, (uint32_t)(pair ? BN / 2 : BN));
  if (rc) return rc;

  long long total = (long long)p.m_blocks * p.n_blocks * p.splits;
  int grid = (int)(total < sms ? total : sms);
  cudaStream_t st = (cudaStream_t)stream;
  // the encoder's hot epilogues as compile-time kinds (see EPI_*); everything else takes the generic kernel
  if (wgrad_pair) return launch_gemm_pair<1, EPI_GENERIC, 1>(ta, tb, p, sms, st);
  if (pair) {
    const bool fast = !p.out_fp32 && p.vec_ok && N % 8 == 0;
    const bool bias = p.bias != nullptr, drop = p.drop.thresh16 != 0, res = p.residual != nullptr, auxo = p.aux_out != nullptr;
    if (fast && !b_major) {
      if (bias && p.act == VT_ACT_NONE && !res && !drop && !auxo) return launch_gemm_pair<0, EPI_BIAS>(ta, tb, p, sms, st);
      if (bias && p.act == VT_ACT_GELU_SAVE_GRAD && !res && !drop && auxo) return launch_gemm_pair<0, EPI_BIAS | EPI_GELU_SG>(ta, tb, p, sms, st);
      if (bias && p.act == VT_ACT_NONE && res && !auxo) return launch_gemm_pair<0, EPI_BIAS | EPI_DROP | EPI_RESID>(ta, tb, p, sms, st);
    } else if (fast && b_major && !bias && !drop && !auxo) {
      if (p.act == VT_ACT_MUL_AUX && !res) return launch_gemm_pair<1, EPI_MUL_AUX>(ta, tb, p, sms, st);
      if (p.act == VT_ACT_NONE && res) return launch_gemm_pair<1, EPI_RESID>(ta, tb, p, sms, st);
      if (p.act == VT_ACT_NONE && !res) return launch_gemm_pair<1, 0>(ta, tb, p, sms, st);
    }
    return b_major ? launch_gemm_pair<1, EPI_GENERIC>(ta, tb, p, sms, st) : launch_gemm_pair<0, EPI_GENERIC>(ta, tb, p, sms, st);
  }
  const bool fast_ok = BN == 256 && !p.out_fp32 && !p.accumulate && p.remap_in <= 0 && !p.remap_index && p.vec_ok && N % 8 == 0 && p.splits == 1;
  if (fast_ok) {
    const bool bias = p.bias != nullptr, drop = p.drop.thresh16 != 0, res = p.residual != nullptr, auxo = p.aux_out != nullptr;
    if (!a_major && !b_major) {
      if (bias && p.act == VT_ACT_NONE && !res && !drop && !auxo) return launch_gemm<256, 0, 0, EPI_BIAS>(ta, tb, p, grid, st);
      if (bias && p.act == VT_ACT_GELU && !res && !drop && auxo) return launch_gemm<256, 0, 0, EPI_BIAS | EPI_GELU>(ta, tb, p, grid, st);
      if (bias && p.act == VT_ACT_GELU_SAVE_GRAD && !res && !drop && auxo) return launch_gemm<256, 0, 0, EPI_BIAS | EPI_GELU_SG>(ta, tb, p, grid, st);
      if (bias && p.act == VT_ACT_NONE && res && !auxo) return launch_gemm<256, 0, 0, EPI_BIAS | EPI_DROP | EPI_RESID>(ta, tb, p, grid, st);
    } else if (!a_major && b_major && !bias && !drop && !auxo) {
      if (p.act == VT_ACT_GELU_BWD && !res) return launch_gemm<256, 0, 1, EPI_GELU_BWD>(ta, tb, p, grid, st);
      if (p.act == VT_ACT_MUL_AUX && !res) return launch_gemm<256, 0, 1, EPI_MUL_AUX>(ta, tb, p, grid, st);
      if (p.act == VT_ACT_NONE && res) return launch_gemm<256, 0, 1, EPI_RESID>(ta, tb, p, grid, st);
      if (p.act == VT_ACT_NONE && !res) return launch_gemm<256, 0, 1, 0>(ta, tb, p, grid, st);
    }
  }
  switch (BN) {
    case 64: return dispatch_major<64>(a_major, b_major, ta, tb, p, grid, st);
    case 128: return dispatch_major<128>(a_major, b_major, ta, tb, p, grid, st);
    case 192: return dispatch_major<192>(a_major, b_major, ta, tb, p, grid, st);
    default: return dispatch_major<256>(a_major, b_major, ta, tb, p, grid, st);
  }
}
