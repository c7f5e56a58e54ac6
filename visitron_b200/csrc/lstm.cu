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
constexpr int FWD_SMEM = 4 * U * WROW + 2 * 2 * NS * WROW + 2 * NS * U * 2;          // W slice, h[2 buffers][hi,lo][8], staging
constexpr int BWD_SMEM = U * DROW + 2 * NS * DROW + 8 * U * NS * 4 + NS * 4 * U * 2;  // W^T slice, dA[2 buffers][8], partials, staging

__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t map_remote(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_remote_v4(uint32_t raddr, uint4 v) {
  asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(raddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
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

}  // namespace lstm
}  // namespace vt
