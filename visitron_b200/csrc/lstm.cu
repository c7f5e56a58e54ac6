// The recurrent middle of the fine-tuning path: OscarEncoder's nn.LSTM(768 -> 512) over the 511 encoder
// positions (vs/agent_models.py:223-230,283-287 == tb/agent_models.py:148-155,208-212) and the decoders' LSTMCell
// (vs/agent_models.py:380,414 ; tb/agent_models.py:295,313), forward and backward.
//
// Sequence LSTM.  The input projection x W_ih^T + b_ih + b_hh of ALL steps is one tcgen05 GEMM (gemm_tcgen05.cu) done by
// the caller; what is sequential is h_{t-1} W_hh^T, 511 dependent [B,512] x [512,2048] products.  cuDNN's answer is a
// persistent kernel with the weights on chip; here that is a thread-block CLUSTER of 16 CTAs per group of 8 samples:
//   * CTA j owns hidden units [32 j, 32 j + 32): the 128 rows of W_hh that make their four gates (128 KB as bf16) stay in
//     its shared memory for the whole sequence; the cell state c of (unit, sample) lives in one thread's register;
//   * per step every warp multiplies its 16 gate rows by h_{t-1} of the 8 samples with mma.sync.m16n8k16 (h as a bf16 hi + lo
//     pair, so the recurrence sees h to ~16 bits) -- a [128 x 512] x [512 x 8] product is far too small and too
//     latency-critical for a tcgen05 / TMEM round trip; the 16-row tile is ordered gate-major x 4 units, so the four gates
//     of a unit meet in two lanes of one warp and are combined with one shuffle, without shared memory;
//   * the new h slice (32 units x 8 samples) is pushed into the next-step h buffer of all 16 CTAs through distributed
//     shared memory, and ONE cluster barrier per step (arrive.release / wait.acquire) orders it; h is double-buffered.
// The backward kernel mirrors it: CTA j owns dh_{t-1} of its 32 units, keeps the matching 32 columns of W_hh (as W_hh^T rows)
// on chip, every CTA computes the gate pre-activation gradients dA_t of its units, broadcasts them (bf16) to the cluster, and
// dh_{t-1} = dA_t W_hh is 32 mma.sync per warp over a 256-gate slice of the contraction, reduced through shared memory in a
// fixed order (deterministic).  dA of all steps goes to HBM as bf16 [B*T, 2048]: the weight gradients dW_ih = dA^T X,
// dW_hh = dA^T H_prev and the input gradient dX = dA W_ih are tcgen05 GEMMs over all steps, the bias gradient a column sum.
// Samples are independent: ceil(B / 8) clusters.  pack_padded_sequence semantics (vs/agent_models.py:284-305): sample b is
// advanced for t < lengths[b] only, ctx is 0 behind its end, (h, c) returned are those of its last valid step.
#include "common.cuh"

namespace vt {

void count_launch(int n);

namespace lstm {

constexpr int H = 512, G = 4 * H;
constexpr int CL = 16;            // CTAs per cluster
constexpr int U = H / CL;         // hidden units per CTA = 32
constexpr int NS = 8;             // samples per cluster (the N of m16n8k16)
constexpr int THREADS = 256;
constexpr int WROW = H * 2 + 16;  // bytes per shared-memory row of 512 bf16, padded: rows 260 words apart -> ldmatrix / LDS.32 conflict-free
constexpr int DROW = G * 2 + 16;  // bytes per row of 2048 bf16, same padding rule
constexpr int FWD_SMEM = 4 * U * WROW + 2 * 2 * NS * WROW + 2 * (2 * NS * U * 2);          // W slice, h[2 buffers][hi,lo][8], staging x 2
constexpr int BWD_SMEM = U * DROW + 2 * NS * DROW + 8 * U * NS * 4 + 2 * (NS * 4 * U * 2);  // W^T slice, dA[2 buffers][8], partials, staging x 2

__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t map_remote(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_remote_v4(uint32_t raddr, uint4 v) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(raddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// remote 16-byte store whose completion is counted (in bytes) on an mbarrier of the DESTINATION CTA: the receiver waits for
// "all 16 slices of this step have landed" on its own barrier -- no cluster-wide barrier and, unlike barrier.cluster's
// release, no wait for this thread's global stores (the saved activations) to drain every step
__device__ __forceinline__ void st_async_v4(uint32_t raddr, uint4 v, uint32_t rbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];"
               ::"r"(raddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(rbar) : "memory");
}
__device__ __forceinline__ void mbar_init_(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_expect_(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  do {
    // default semantics (acquire at CTA scope), as for TMA / multicast arrivals: the data is in THIS CTA's shared memory once the
    // phase completes.  (.acquire.cluster made ptxas emit CCTL.IVALL -- an L1 invalidate -- in every poll: 21 % of the forward
    // kernel's stall samples, profiles/r2_ncu_lstm.txt.)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t* a) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]) : "r"(addr));
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void mma_bf16(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + __expf(-x)); }


struct SeqParams {
  float* gates;            // [B, T, 4H] in: x W_ih^T + b_ih + b_hh (gate order i, f, g, o as nn.LSTM); out: the ACTIVATED gates (saved for backward)
  const float* w_hh;       // [4H, H] fp32 master
  const long long* lengths;  // [B] valid steps per sample, or null (all T)
  int B, T;
  float* ctx;              // [B, T, H]  h_t, 0 behind the sample's end (pad_packed_sequence)
  float* c_all;            // [B, T, H]  c_t (saved for backward)
  bf16* hprev;             // [B, T, H]  bf16(h_{t-1}): the operand of dW_hh = dA^T H_prev
  float* h_last;           // [B, H]
  float* c_last;           // [B, H]
  // backward
  const float* dctx;       // [B, T, H]
  const float* dh_last;    // [B, H] or null
  const float* dc_last;    // [B, H] or null
  bf16* dA;                // [B, T, 4H] gate pre-activation gradients (0 behind the sample's end)
};

// ---------------------------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(THREADS, 1)
lstm_seq_fwd_kernel(const SeqParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* sW = smem;                                   // [128 rows][WROW]: row 16 w + 4 gate + ul  <->  W_hh[gate * 512 + 32 j + 4 w + ul, :]
  uint8_t* sH = sW + 4 * U * WROW;                      // [buffer 2][hi / lo 2][sample 8][WROW]
  uint8_t* sStage = sH + 2 * 2 * NS * WROW;             // [hi / lo 2][sample 8][32 units] bf16
  __shared__ __align__(8) unsigned long long s_bar[2];  // one per h buffer: counts the 16 x 1 KB slices of a step
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t j = cluster_rank();
  const int sample0 = (blockIdx.x / CL) * NS;
  const uint32_t bar_u = smem_addr(s_bar);
  constexpr uint32_t STEP_BYTES = CL * 2 * NS * U * 2;  // what one step delivers to a CTA: 16 CTAs x (hi, lo) x 8 samples x 32 units bf16
  if (tid == 0) {
    mbar_init_(bar_u, 1);
    mbar_init_(bar_u + 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_expect_(bar_u + 8, STEP_BYTES);                 // step 1's h arrives in buffer 1
  }

  // W slice -> shared memory (fp32 -> bf16), once
  for (int i = tid; i < 4 * U * (H / 4); i += THREADS) {
    const int row = i / (H / 4), c4 = i % (H / 4);
    const int w = row >> 4, r = row & 15, gate = r >> 2, ul = r & 3;
    const float4 v = __ldg(reinterpret_cast<const float4*>(p.w_hh + (long long)(gate * H + U * (int)j + 4 * w + ul) * H) + c4);
    uint2 o = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
    *reinterpret_cast<uint2*>(sW + row * WROW + c4 * 8) = o;
  }
  for (int i = tid; i < 2 * 2 * NS * WROW / 4; i += THREADS) reinterpret_cast<uint32_t*>(sH)[i] = 0u;   // h_{-1} = 0
  __syncthreads();
  cluster_sync_all();   // every CTA's buffers are initialised before anyone pushes h into them

  // this thread's (unit, sample): the four gates of unit 4 warp + ul meet in lanes (lane, lane ^ 16)
  const int r0 = lane >> 2, g0 = r0 >> 2, ul = r0 & 3;
  const int n = 2 * (lane & 3) + g0;                    // sample within the cluster's 8
  const int ucta = 4 * warp + ul;                       // unit within the CTA's 32
  const int unit = U * (int)j + ucta;
  const int b = sample0 + n;
  const bool exists = b < p.B;
  const int len = exists ? (p.lengths ? (int)min((long long)p.T, max(0ll, p.lengths[b])) : p.T) : 0;
  float c = 0.f, h = 0.f;
  const long long gbase = exists ? ((long long)b * p.T) * G + unit : 0;
  const long long hbase = exists ? ((long long)b * p.T) * H + unit : 0;
  float gx[4] = {0.f, 0.f, 0.f, 0.f};
  if (len > 0) {
#pragma unroll
    for (int q = 0; q < 4; ++q) gx[q] = p.gates[gbase + q * H];
  }
  const uint32_t sW_u = smem_addr(sW), sH_u = smem_addr(sH), sStage_u = smem_addr(sStage);
  // ldmatrix row address of this lane: row (lane % 8) + 8 ((lane / 8) % 2) of the warp's 16, k offset 8 (lane / 16)
  const uint32_t a_lane = sW_u + (uint32_t)(16 * warp + (lane & 7) + 8 * ((lane >> 3) & 1)) * WROW + (uint32_t)(lane >> 4) * 16u;
  // B fragments of one k-step with ONE ldmatrix.x4: matrices (h hi, k 0-7), (hi, k 8-15), (lo, k 0-7), (lo, k 8-15), rows = samples
  const uint32_t b_lane = (uint32_t)(lane & 7) * WROW + (uint32_t)((lane >> 3) & 1) * 16u + (uint32_t)(lane >> 4) * (NS * WROW);

  for (int t = 0; t < p.T; ++t) {
    const uint32_t hb = sH_u + (uint32_t)(t & 1) * (2 * NS * WROW);
    if (t > 0) {
      // buffer t & 1 holds step t's h once all 16 slices have landed; its k-th use has parity k & 1
      mbar_wait_cluster(bar_u + 8 * (t & 1), (uint32_t)(((t - 1) >> 1) & 1));
      // arm the buffer's NEXT use (step t + 2).  Safe this early: a peer sends for step t + 2 only after it has received this
      // CTA's step-(t + 1) slice, which is sent below, after this point
      if (tid == 0 && t + 2 < p.T) mbar_expect_(bar_u + 8 * (t & 1), STEP_BYTES);
    } else if (tid == 0 && 2 < p.T) {
      mbar_expect_(bar_u, STEP_BYTES);                   // step 2's h arrives in buffer 0
    }
    // eight independent accumulation chains (4 k-quarters x hi / lo): one chain of 64 dependent MMAs was the step's critical path
    float accq[8][4];
#pragma unroll
    for (int c = 0; c < 8; ++c) accq[c][0] = accq[c][1] = accq[c][2] = accq[c][3] = 0.f;
#pragma unroll
    for (int kk = 0; kk < H / 16; ++kk) {
      uint32_t a[4];
      ldmatrix_x4(a_lane + kk * 32, a);
      uint32_t bq[4];
      ldmatrix_x4(hb + b_lane + kk * 32, bq);
      mma_bf16(accq[kk & 3], a, bq[0], bq[1]);          // h hi
      mma_bf16(accq[4 + (kk & 3)], a, bq[2], bq[3]);    // h lo
    }
    float acc[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[q] = ((accq[0][q] + accq[1][q]) + (accq[2][q] + accq[3][q])) + ((accq[4][q] + accq[5][q]) + (accq[6][q] + accq[7][q]));
    // gate exchange: lanes with g0 == 0 finish column n0 = 2 (lane % 4), their partners (lane ^ 16) column n0 + 1
    const float s0 = g0 == 0 ? acc[1] : acc[0], s1 = g0 == 0 ? acc[3] : acc[2];
    const float r_a = __shfl_xor_sync(0xffffffffu, s0, 16), r_b = __shfl_xor_sync(0xffffffffu, s1, 16);
    float pi, pf, pg, po;   // recurrent parts of the pre-activations
    if (g0 == 0) { pi = acc[0]; pg = acc[2]; pf = r_a; po = r_b; }
    else         { pf = acc[1]; po = acc[3]; pi = r_a; pg = r_b; }
    const bool active = t < len;
    // next step's input projections: requested now, consumed after the next step's MMAs
    float gn[4] = {0.f, 0.f, 0.f, 0.f};
    if (t + 1 < len) {
#pragma unroll
      for (int q = 0; q < 4; ++q) gn[q] = p.gates[gbase + (long long)(t + 1) * G + q * H];
    }
    if (active) {
      p.hprev[hbase + (long long)t * H] = __float2bfloat16(h);
      const float ig = sigmoidf_(gx[0] + pi), fg = sigmoidf_(gx[1] + pf), gg = tanhf(gx[2] + pg), og = sigmoidf_(gx[3] + po);
      c = fg * c + ig * gg;
      h = og * tanhf(c);
      float* gs = p.gates + gbase + (long long)t * G;
      gs[0] = ig; gs[H] = fg; gs[2 * H] = gg; gs[3 * H] = og;
      p.c_all[hbase + (long long)t * H] = c;
      p.ctx[hbase + (long long)t * H] = h;
    } else if (exists) {
      p.ctx[hbase + (long long)t * H] = 0.f;
      p.hprev[hbase + (long long)t * H] = __float2bfloat16(0.f);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) gx[q] = gn[q];
    if (t + 1 == p.T) break;
    // h (frozen for finished samples) as a bf16 hi + lo pair -> staging -> every CTA's next-step buffer
    const bf16 hi = __float2bfloat16(h);
    const bf16 lo = __float2bfloat16(h - __bfloat162float(hi));
    // staging is double-buffered by step parity: a thread still copying step t's slots cannot meet step t + 1's writes
    uint8_t* stg = sStage + (t & 1) * (2 * NS * U * 2);
    reinterpret_cast<bf16*>(stg)[(0 * NS + n) * U + ucta] = hi;
    reinterpret_cast<bf16*>(stg)[(1 * NS + n) * U + ucta] = lo;
    __syncthreads();
    {
      // 2 x 8 x 64 B = 64 chunks of 16 B, to 16 CTAs: 4 remote stores per thread, each counted on the receiver's barrier
      const uint32_t dst_cta = (uint32_t)(tid >> 4);
      const uint32_t nb = sH_u + (uint32_t)((t + 1) & 1) * (2 * NS * WROW);
      const uint32_t rbar = map_remote(bar_u + 8 * ((t + 1) & 1), dst_cta);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int ch = (tid & 15) + 16 * k;           // chunk: hl = ch / 32, sample = (ch / 4) % 8, part = ch % 4
        const int hl = ch >> 5, sn = (ch >> 2) & 7, part = ch & 3;
        const uint4 v = *reinterpret_cast<const uint4*>(stg + ((hl * NS + sn) * U + part * 8) * 2);
        const uint32_t local = nb + (uint32_t)(hl * NS + sn) * WROW + (uint32_t)(U * (int)j + part * 8) * 2u;
        st_async_v4(map_remote(local, dst_cta), v, rbar);
      }
    }
    // (no barrier here: the next step's wait on the h buffer completes only when every thread of every CTA has sent,
    //  i.e. has read its staging slots; the staging buffer is rewritten after that wait)
  }
  if (exists) {
    p.h_last[(long long)b * H + unit] = h;
    p.c_last[(long long)b * H + unit] = c;
  }
  cluster_sync_all();   // no CTA exits while a peer may still address its shared memory
}


// ---------------------------------------------------------------------------------------------
// backward: t = T-1 .. 0.  Thread (unit u = tid / 8, sample n = tid % 8) carries dc and receives dh_{t} of its pair.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(THREADS, 1)
lstm_seq_bwd_kernel(const SeqParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* sWT = smem;                                  // [32 rows k][DROW]: row k  <->  W_hh[:, 32 j + k]  (W_hh^T slice)
  uint8_t* sD = sWT + U * DROW;                         // [buffer 2][sample 8][DROW]: dA_t of ALL 2048 gates, bf16
  float* sPart = reinterpret_cast<float*>(sD + 2 * NS * DROW);    // [warp 8][32 units][8 samples]
  uint8_t* sStage = reinterpret_cast<uint8_t*>(sPart + 8 * U * NS);   // [parity 2][sample 8][gate 4][32 units] bf16
  __shared__ __align__(8) unsigned long long s_bar[2];  // one per dA buffer: counts the 16 x 2 KB slices of a step
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t j = cluster_rank();
  const int sample0 = (blockIdx.x / CL) * NS;
  const uint32_t bar_u = smem_addr(s_bar);
  constexpr uint32_t STEP_BYTES = CL * NS * 4 * U * 2;   // 16 CTAs x 8 samples x 4 gates x 32 units bf16
  if (tid == 0) {
    mbar_init_(bar_u, 1);
    mbar_init_(bar_u + 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // dA_t (t = T-1 down to 1) lands in buffer t & 1: arm the first use of both buffers
    if (p.T >= 2) mbar_expect_(bar_u + 8 * ((p.T - 1) & 1), STEP_BYTES);
    if (p.T >= 3) mbar_expect_(bar_u + 8 * ((p.T - 2) & 1), STEP_BYTES);
  }

  // W_hh^T slice: sWT[k][g] = W_hh[g][32 j + k]; a warp reads 32 consecutive floats of one row of W_hh (128 B) per step
  for (int g = warp; g < G; g += THREADS / 32) {
    const float v = __ldg(p.w_hh + (long long)g * H + U * (int)j + lane);
    *reinterpret_cast<bf16*>(sWT + lane * DROW + g * 2) = __float2bfloat16(v);
  }
  __syncthreads();
  cluster_sync_all();

  const int ucta = tid >> 3, n = tid & 7;
  const int unit = U * (int)j + ucta;
  const int b = sample0 + n;
  const bool exists = b < p.B;
  const int len = exists ? (p.lengths ? (int)min((long long)p.T, max(0ll, p.lengths[b])) : p.T) : 0;
  const long long gbase = exists ? ((long long)b * p.T) * G + unit : 0;
  const long long hbase = exists ? ((long long)b * p.T) * H + unit : 0;
  float dc = 0.f, dh_rec = 0.f;
  const uint32_t sWT_u = smem_addr(sWT), sD_u = smem_addr(sD);
  // A fragment rows: m-tile mt covers units 16 mt .. + 15; this warp contracts over gates [256 warp, + 256)
  const uint32_t a_lane = sWT_u + (uint32_t)((lane & 7) + 8 * ((lane >> 3) & 1)) * DROW + (uint32_t)(lane >> 4) * 16u + (uint32_t)warp * 512u;
  // B fragments of TWO k-steps with one ldmatrix.x4: matrices (k 0-7), (k 8-15), (k 16-23), (k 24-31), rows = samples
  const uint32_t b_lane = (uint32_t)(lane & 7) * DROW + (uint32_t)(lane >> 3) * 16u + (uint32_t)warp * 512u;

  for (int t = p.T - 1; t >= 0; --t) {
    const bool active = t < len;
    float da[4] = {0.f, 0.f, 0.f, 0.f};
    if (active) {
      const float* gs = p.gates + gbase + (long long)t * G;
      const float ig = gs[0], fg = gs[H], gg = gs[2 * H], og = gs[3 * H];
      const float ct = p.c_all[hbase + (long long)t * H];
      const float cp = t > 0 ? p.c_all[hbase + (long long)(t - 1) * H] : 0.f;
      float dh = p.dctx[hbase + (long long)t * H] + dh_rec;
      if (t == len - 1) {
        if (p.dh_last) dh += p.dh_last[(long long)b * H + unit];
        if (p.dc_last) dc += p.dc_last[(long long)b * H + unit];
      }
      const float tc = tanhf(ct);
      const float dct = dc + dh * og * (1.f - tc * tc);
      da[0] = dct * gg * ig * (1.f - ig);       // i
      da[1] = dct * cp * fg * (1.f - fg);       // f
      da[2] = dct * ig * (1.f - gg * gg);       // g
      da[3] = dh * tc * og * (1.f - og);        // o
      dc = dct * fg;
    }
    if (exists) {
      bf16* dst = p.dA + gbase + (long long)t * G;
#pragma unroll
      for (int q = 0; q < 4; ++q) dst[q * H] = __float2bfloat16(da[q]);
    }
    if (t == 0) break;      // dh_{-1} is nobody's gradient (h_{-1} = 0 is a constant)
    uint8_t* stg = sStage + (t & 1) * (NS * 4 * U * 2);    // double-buffered by step parity (see the forward kernel)
#pragma unroll
    for (int q = 0; q < 4; ++q) reinterpret_cast<bf16*>(stg)[(n * 4 + q) * U + ucta] = __float2bfloat16(da[q]);
    __syncthreads();
    {
      // 8 samples x 4 gates x 64 B = 128 chunks of 16 B, to 16 CTAs: 8 remote stores per thread, counted on the receiver's barrier
      const uint32_t dst_cta = (uint32_t)(tid >> 4);
      const uint32_t nb = sD_u + (uint32_t)(t & 1) * (NS * DROW);
      const uint32_t rbar = map_remote(bar_u + 8 * (t & 1), dst_cta);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int ch = (tid & 15) + 16 * k;           // chunk: sample = ch / 16, gate = (ch / 4) % 4, part = ch % 4
        const int sn = ch >> 4, q = (ch >> 2) & 3, part = ch & 3;
        const uint4 v = *reinterpret_cast<const uint4*>(stg + ((sn * 4 + q) * U + part * 8) * 2);
        const uint32_t local = nb + (uint32_t)sn * DROW + (uint32_t)(q * H + U * (int)j + part * 8) * 2u;
        st_async_v4(map_remote(local, dst_cta), v, rbar);
      }
    }
    // all 16 slices of dA_t have landed; the buffer's k-th use (t = T-1-2k or T-2-2k) has parity k & 1
    mbar_wait_cluster(bar_u + 8 * (t & 1), (uint32_t)(((p.T - 1 - t) >> 1) & 1));
    // its next use is dA_{t-2}: a peer sends that only after it has received this CTA's dA_{t-1}, sent in the next iteration
    if (tid == 0 && t - 2 >= 1) mbar_expect_(bar_u + 8 * (t & 1), STEP_BYTES);
    // dh_{t-1}[unit, sample] = sum_g W_hh[g, unit] dA_t[sample, g]: this warp's 256-gate slice, both 16-unit tiles
    const uint32_t db = sD_u + (uint32_t)(t & 1) * (NS * DROW);
    float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
    float acc2[4] = {0.f, 0.f, 0.f, 0.f}, acc3[4] = {0.f, 0.f, 0.f, 0.f};   // four independent chains
#pragma unroll
    for (int k2 = 0; k2 < 8; ++k2) {
      uint32_t a0[4], a1[4], a2[4], a3[4], bq[4];
      ldmatrix_x4(a_lane + k2 * 64, a0);
      ldmatrix_x4(a_lane + 16 * DROW + k2 * 64, a1);
      ldmatrix_x4(a_lane + k2 * 64 + 32, a2);
      ldmatrix_x4(a_lane + 16 * DROW + k2 * 64 + 32, a3);
      ldmatrix_x4(db + b_lane + k2 * 64, bq);
      mma_bf16(acc0, a0, bq[0], bq[1]);
      mma_bf16(acc1, a1, bq[0], bq[1]);
      mma_bf16(acc2, a2, bq[2], bq[3]);
      mma_bf16(acc3, a3, bq[2], bq[3]);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) { acc0[q] += acc2[q]; acc1[q] += acc3[q]; }
    {
      // fragment (row = lane / 4 (+ 8), col = 2 (lane % 4) (+ 1)) -> sPart[warp][unit][sample]
      float* pw = sPart + warp * (U * NS);
      const int r = lane >> 2, cc = 2 * (lane & 3);
      pw[r * NS + cc] = acc0[0]; pw[r * NS + cc + 1] = acc0[1];
      pw[(r + 8) * NS + cc] = acc0[2]; pw[(r + 8) * NS + cc + 1] = acc0[3];
      pw[(r + 16) * NS + cc] = acc1[0]; pw[(r + 16) * NS + cc + 1] = acc1[1];
      pw[(r + 24) * NS + cc] = acc1[2]; pw[(r + 24) * NS + cc + 1] = acc1[3];
    }
    __syncthreads();
    float sum = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) sum += sPart[w * (U * NS) + tid];   // [unit][sample] flat index == tid; fixed order
    dh_rec = sum;
    // (the next iteration's __syncthreads orders these reads before sPart is written again)
  }
  cluster_sync_all();
}


// ---------------------------------------------------------------------------------------------
// LSTMCell of the decoders (one step): the two gate halves x W_ih^T + b_ih and h W_hh^T + b_hh come from the step-sized
// linears of agent_heads.cu; this is the element-wise part and its backward.  gates_act keeps (i, f, g, o) activated.
// ---------------------------------------------------------------------------------------------
__global__ void lstm_cell_fwd_kernel(const float* __restrict__ ga, const float* __restrict__ gb, const float* __restrict__ c0, int B, int Hh,
                                     float* __restrict__ act, float* __restrict__ h1, float* __restrict__ c1) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * Hh) return;
  const int s = i / Hh, u = i % Hh;
  const long long g = (long long)s * 4 * Hh + u;
  const float ig = sigmoidf_(ga[g] + gb[g]), fg = sigmoidf_(ga[g + Hh] + gb[g + Hh]);
  const float gg = tanhf(ga[g + 2 * Hh] + gb[g + 2 * Hh]), og = sigmoidf_(ga[g + 3 * Hh] + gb[g + 3 * Hh]);
  const float c = fg * c0[i] + ig * gg;
  act[g] = ig; act[g + Hh] = fg; act[g + 2 * Hh] = gg; act[g + 3 * Hh] = og;
  c1[i] = c;
  h1[i] = og * tanhf(c);
}

__global__ void lstm_cell_bwd_kernel(const float* __restrict__ act, const float* __restrict__ c0, const float* __restrict__ c1,
                                     const float* __restrict__ dh1, const float* __restrict__ dc1, int B, int Hh,
                                     float* __restrict__ dgates, float* __restrict__ dc0) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * Hh) return;
  const int s = i / Hh, u = i % Hh;
  const long long g = (long long)s * 4 * Hh + u;
  const float ig = act[g], fg = act[g + Hh], gg = act[g + 2 * Hh], og = act[g + 3 * Hh];
  const float tc = tanhf(c1[i]);
  const float dh = dh1 ? dh1[i] : 0.f;
  const float dct = (dc1 ? dc1[i] : 0.f) + dh * og * (1.f - tc * tc);
  dgates[g] = dct * gg * ig * (1.f - ig);
  dgates[g + Hh] = dct * c0[i] * fg * (1.f - fg);
  dgates[g + 2 * Hh] = dct * ig * (1.f - gg * gg);
  dgates[g + 3 * Hh] = dh * tc * og * (1.f - og);
  dc0[i] = dct * fg;
}

// y = x * keep / (1 - p) over a flat fp32 array: nn.Dropout of the agents' modules (vs/agent_models.py:216,379 ; p = 0.5); the
// backward is the same call on the gradient (the mask is regenerated from (seed, stream, index), never stored)
__global__ void dropout_f32_kernel(const float* __restrict__ x, long long n, DropoutCfg drop, float* __restrict__ y) {
  drop = resolve_dropout(drop);
  for (long long g8 = (long long)blockIdx.x * blockDim.x + threadIdx.x; g8 * 8 < n; g8 += (long long)gridDim.x * blockDim.x) {
    float m[8];
    dropout_mult8(drop, (uint64_t)g8 * 8, m);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const long long i = g8 * 8 + e;
      if (i < n) y[i] = x[i] * m[e];
    }
  }
}


// nn.Embedding lookup of the turn-based decoder's previous action (tb/agent_models.py:292,308): y[b] = W[idx[b]]; backward
// dW[v] = sum over b with idx[b] == v of dy[b], samples in ascending order (deterministic, no atomics; the vocabulary is 8)
__global__ void embedding_f32_fwd_kernel(const float* __restrict__ W, const long long* __restrict__ idx, int B, int V, int D, float* __restrict__ y) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * D) return;
  long long v = idx[i / D];
  v = v < 0 ? 0 : (v >= V ? V - 1 : v);
  y[i] = W[v * D + i % D];
}
__global__ void embedding_f32_bwd_kernel(const float* __restrict__ dy, const long long* __restrict__ idx, int B, int V, int D, float* __restrict__ dW) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V * D) return;
  const int v = i / D, d = i % D;
  float acc = 0.f;
  for (int b = 0; b < B; ++b)
    if (idx[b] == v) acc += dy[(long long)b * D + d];
  dW[i] = acc;
}

// Action feedback of one rollout step, on the device (vs/agent.py:410-441 ; tb/agent.py:329-346 do this on the host after a
// .cpu() / .item() sync per step): a_t = the teacher's action | argmax of the masked logits | a sample of softmax(masked logits);
// ended[b] |= a_t is the stop action (candidate_leng[b] - 1 when stop_len is given, else the constant <end>) or the ignore id;
// next_target_eff[b] = ended[b] ? ignore : next_target[b]  (what _teacher_action(obs, ended) returns for a finished episode).
__global__ void rollout_feedback_kernel(const float* __restrict__ logits, long long ld, int B, int C, const long long* __restrict__ target,
                                        const long long* __restrict__ stop_len, long long stop_const, long long ignore, int mode,
                                        DropoutCfg rng, unsigned step, unsigned char* __restrict__ ended, long long* __restrict__ a_out,
                                        const long long* __restrict__ next_target, long long* __restrict__ next_target_eff) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  rng = resolve_dropout(rng);
  const float* x = logits + (long long)b * ld;
  long long a = target[b];
  if (mode != 0) {
    float mx = -INFINITY;
    int arg = 0;
    for (int c = 0; c < C; ++c)
      if (x[c] > mx) { mx = x[c]; arg = c; }          // ties keep the lowest index, like torch.max
    a = arg;
    if (mode == 2) {
      float se = 0.f;
      for (int c = 0; c < C; ++c) se += __expf(x[c] - mx);
      const uint4 r = philox4x32<7>(make_uint4((unsigned)b, step, 0x5a3c1eu, rng.stream), make_uint2(rng.seed_lo, rng.seed_hi));
      const float u = (float)(r.x >> 8) * (1.0f / 16777216.0f) * se;   // uniform in [0, sum)
      float cum = 0.f;
      for (int c = 0; c < C; ++c) {
        const float e = __expf(x[c] - mx);
        if (e > 0.f) { a = c; cum += e; if (u < cum) break; }
      }
    }
  }
  const long long stop = stop_len ? stop_len[b] - 1 : stop_const;
  const bool fin = ended[b] || a == stop || a == ignore;
  ended[b] = fin ? 1 : 0;
  a_out[b] = a;
  if (next_target_eff) next_target_eff[b] = fin ? ignore : next_target[b];
}

}  // namespace lstm
}  // namespace vt

using namespace vt;
using namespace vt::lstm;

static int launch_seq(void (*kern)(const SeqParams), int smem_bytes, const SeqParams& p, cudaStream_t st, const char* who) {
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  if (e != cudaSuccess) return cuda_fail(e, who, __FILE__, __LINE__);
  e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);   // 16 CTAs per cluster (8 is the portable limit)
  if (e != cudaSuccess) return cuda_fail(e, who, __FILE__, __LINE__);
  cudaLaunchConfig_t cfg = {};
  const int clusters = (p.B + NS - 1) / NS;
  cfg.gridDim = dim3((unsigned)(clusters * CL));
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = (size_t)smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  e = cudaLaunchKernelEx(&cfg, kern, p);
  if (e != cudaSuccess) return cuda_fail(e, who, __FILE__, __LINE__);
  count_launch(1);
  return VT_OK;
}

extern "C" int vt_lstm_seq_fwd(float* gates, const float* w_hh, const int64_t* lengths, int64_t B, int64_t T, int64_t hidden, float* ctx,
                               float* c_all, void* hprev_bf16, float* h_last, float* c_last, vt_stream stream) {
  VT_REQUIRE(hidden == H, "vt_lstm_seq_fwd: hidden size %lld outside the supported envelope (512, vs/params.py rnn_dim)", (long long)hidden);
  VT_REQUIRE(gates && w_hh && ctx && c_all && hprev_bf16 && h_last && c_last && B > 0 && T > 0 && B * T < (1ll << 31), "vt_lstm_seq_fwd: bad arguments");
  VT_REQUIRE(((uintptr_t)w_hh & 15) == 0, "vt_lstm_seq_fwd: w_hh must be 16-byte aligned");
  SeqParams p = {};
  p.gates = gates; p.w_hh = w_hh; p.lengths = (const long long*)lengths; p.B = (int)B; p.T = (int)T;
  p.ctx = ctx; p.c_all = c_all; p.hprev = (bf16*)hprev_bf16; p.h_last = h_last; p.c_last = c_last;
  return launch_seq(lstm_seq_fwd_kernel, FWD_SMEM, p, (cudaStream_t)stream, "vt_lstm_seq_fwd");
}

extern "C" int vt_lstm_seq_bwd(const float* gates, const float* c_all, const float* w_hh, const int64_t* lengths, const float* dctx,
                               const float* dh_last, const float* dc_last, int64_t B, int64_t T, int64_t hidden, void* dA_bf16, vt_stream stream) {
  VT_REQUIRE(hidden == H, "vt_lstm_seq_bwd: hidden size %lld outside the supported envelope (512)", (long long)hidden);
  VT_REQUIRE(gates && c_all && w_hh && dctx && dA_bf16 && B > 0 && T > 0 && B * T < (1ll << 31), "vt_lstm_seq_bwd: bad arguments");
  SeqParams p = {};
  p.gates = const_cast<float*>(gates); p.c_all = const_cast<float*>(c_all); p.w_hh = w_hh; p.lengths = (const long long*)lengths;
  p.B = (int)B; p.T = (int)T; p.dctx = dctx; p.dh_last = dh_last; p.dc_last = dc_last; p.dA = (bf16*)dA_bf16;
  return launch_seq(lstm_seq_bwd_kernel, BWD_SMEM, p, (cudaStream_t)stream, "vt_lstm_seq_bwd");
}

extern "C" int vt_lstm_cell_fwd(const float* ga, const float* gb, const float* c0, int64_t B, int64_t hidden, float* gates_act, float* h1,
                                float* c1, vt_stream stream) {
  VT_REQUIRE(ga && gb && c0 && gates_act && h1 && c1 && B >= 0 && hidden > 0 && B * hidden < (1ll << 31), "vt_lstm_cell_fwd: bad arguments");
  if (B == 0) return VT_OK;
  const int n = (int)(B * hidden);
  lstm_cell_fwd_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(ga, gb, c0, (int)B, (int)hidden, gates_act, h1, c1);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_lstm_cell_bwd(const float* gates_act, const float* c0, const float* c1, const float* dh1, const float* dc1, int64_t B,
                                int64_t hidden, float* dgates, float* dc0, vt_stream stream) {
  VT_REQUIRE(gates_act && c0 && c1 && dgates && dc0 && B >= 0 && hidden > 0 && B * hidden < (1ll << 31), "vt_lstm_cell_bwd: bad arguments");
  if (B == 0) return VT_OK;
  const int n = (int)(B * hidden);
  lstm_cell_bwd_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(gates_act, c0, c1, dh1, dc1, (int)B, (int)hidden, dgates, dc0);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_dropout_f32(const float* x, int64_t n, float dropout_p, uint64_t seed, uint32_t dropout_stream, float* y, vt_stream stream) {
  VT_REQUIRE(x && y && n >= 0 && dropout_p >= 0.f && dropout_p < 1.f, "vt_dropout_f32: bad arguments");
  if (n == 0) return VT_OK;
  const long long groups = (n + 7) / 8;
  long long g = (groups + 255) / 256, cap = (long long)num_sms() * 8;
  dropout_f32_kernel<<<(unsigned)(g < cap ? g : cap), 256, 0, (cudaStream_t)stream>>>(x, (long long)n, make_dropout(dropout_p, seed, dropout_stream), y);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_embedding_f32_fwd(const float* W, const int64_t* idx, int64_t B, int64_t V, int64_t D, float* y, vt_stream stream) {
  VT_REQUIRE(W && idx && y && B >= 0 && V > 0 && D > 0 && B * D < (1ll << 31), "vt_embedding_f32_fwd: bad arguments");
  if (B == 0) return VT_OK;
  const int n = (int)(B * D);
  embedding_f32_fwd_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(W, (const long long*)idx, (int)B, (int)V, (int)D, y);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_embedding_f32_bwd(const float* dy, const int64_t* idx, int64_t B, int64_t V, int64_t D, float* dW, vt_stream stream) {
  VT_REQUIRE(dy && idx && dW && B >= 0 && V > 0 && D > 0 && V * D < (1ll << 31), "vt_embedding_f32_bwd: bad arguments");
  const int n = (int)(V * D);
  embedding_f32_bwd_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(dy, (const long long*)idx, (int)B, (int)V, (int)D, dW);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_rollout_feedback(const float* masked_logits, int64_t ld, int64_t B, int64_t C, const int64_t* target, const int64_t* stop_len,
                                   int64_t stop_const, int64_t ignore_index, int mode, uint64_t seed, uint32_t rng_stream, uint32_t step,
                                   uint8_t* ended, int64_t* a_out, const int64_t* next_target, int64_t* next_target_eff, vt_stream stream) {
  VT_REQUIRE(masked_logits && target && ended && a_out && B > 0 && C > 0 && ld >= C && mode >= 0 && mode <= 2, "vt_rollout_feedback: bad arguments");
  VT_REQUIRE(!next_target_eff || next_target, "vt_rollout_feedback: next_target_eff needs next_target");
  rollout_feedback_kernel<<<(unsigned)((B + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
      masked_logits, ld, (int)B, (int)C, (const long long*)target, (const long long*)stop_len, (long long)stop_const, (long long)ignore_index, mode,
      make_dropout(0.5f, seed, rng_stream), step, ended, (long long*)a_out, (const long long*)next_target, (long long*)next_target_eff);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}
