// Shared device/host helpers for libvisitron_b200 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/visitron_b200.h"

namespace vt {

typedef __nv_bfloat16 bf16;

// ---------------------------------------------------------------------------------------------
// error plumbing (thread-local message, negative status codes, see include/visitron_b200.h)
// ---------------------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define VT_CUDA(call)                                                        \
  do {                                                                       \
    cudaError_t e__ = (call);                                                \
    if (e__ != cudaSuccess) return vt::cuda_fail(e__, #call, __FILE__, __LINE__); \
  } while (0)

#define VT_CHECK_LAUNCH() VT_CUDA(cudaGetLastError())

#define VT_REQUIRE(cond, ...)            \
  do {                                   \
    if (!(cond)) {                       \
      vt::set_error(__VA_ARGS__);        \
      return VT_ERR_BAD_ARGUMENT;        \
    }                                    \
  } while (0)

int num_sms();
int usable_sms();

// ---------------------------------------------------------------------------------------------
// Programmatic dependent launch (opt-in: VT_PDL=1).  A step is ~380 launches on one stream; launch_pdl() lets the next grid
// become resident while the previous one is still running, so that its launch latency and prologue (barrier init, TMEM
// allocation, descriptor prefetch) overlap the previous grid's tail.  pdl_sync(), which EVERY thread of a kernel launched
// this way executes before its first global-memory access (and before any early return -- a grid that finished without
// waiting would let its own dependent start ahead of the grid before it), blocks until the previous grid has completed and
// its writes are visible, then releases the dependent of this grid: stream order is unchanged and at most two grids are
// resident at a time.  MEASURED (profiles/r1_pdl_ab.txt, same box, alternating runs): 24.4-24.6 ms per step with it against
// 24.2-24.3 without -- the step runs at the board's power limit (sw_power_cap, SM clocks 1.71-1.78 of 1.965 GHz), so closing
// the bubbles between kernels only lowers the clocks.  Hence off by default; without the launch attribute
// griddepcontrol.wait / launch_dependents are no-ops and launch_pdl() is a plain launch.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_sync() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
bool pdl_enabled();
template <typename... KA, typename... A>
inline void launch_pdl(void (*kern)(KA...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, A&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  (void)cudaLaunchKernelEx(&cfg, kern, static_cast<KA>(args)...);   // a failure is picked up by VT_CHECK_LAUNCH() behind every call
}

// ---------------------------------------------------------------------------------------------
// small device utilities
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float2 unpack_bf16(uint32_t v) {
  __nv_bfloat162 t = *reinterpret_cast<__nv_bfloat162*>(&v);
  return __bfloat1622float2(t);
}
__device__ __forceinline__ void unpack8(const uint4& q, float* f) {
  float2 a = unpack_bf16(q.x), b = unpack_bf16(q.y), c = unpack_bf16(q.z), d = unpack_bf16(q.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 q;
  q.x = pack_bf16(f[0], f[1]); q.y = pack_bf16(f[2], f[3]);
  q.z = pack_bf16(f[4], f[5]); q.w = pack_bf16(f[6], f[7]);
  return q;
}

// Standard normal CDF, Phi(x) = 0.5 (1 + erf(x / sqrt 2)), to 1e-6 absolute: the lower tail is exp2 of a degree-5
// polynomial in |x| (weighted least-squares fit of log2 Phi(-a) on a in [0, 6], clamped beyond, where Phi(-6) = 1e-9),
// the upper side its complement.  One MUFU and seven FMA-pipe instructions -- the FFN-up epilogue applies it to every
// element of [M, 3072] and has to stay below the tensor-core time of its tile; erff() costs about three times as much.
__device__ __forceinline__ float normal_cdf(float x) {
  const float a = fminf(fabsf(x), 6.0f);
  float q = fmaf(-5.021095858e-04f, a, 7.289528381e-03f);
  q = fmaf(q, a, -5.234604329e-02f);
  q = fmaf(q, a, -4.594255686e-01f);
  q = fmaf(q, a, -1.151043534e+00f);
  q = fmaf(q, a, -1.000002623e+00f);
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(q));
  return x >= 0.f ? 1.0f - e : e;
}
// gelu(x) = x * 0.5 * (1 + erf(x / sqrt(2)))          (hf/modeling_bert.py:136-142)
__device__ __forceinline__ float gelu_fwd(float x) { return x * normal_cdf(x); }
// d/dx gelu(x) = Phi(x) + x * exp(-x^2/2) / sqrt(2 pi)
// (the 1 / sqrt(2 pi) rides in the exponent: exp2(-x^2 log2(e) / 2 + log2(1 / sqrt(2 pi))), one multiply less per element)
__device__ __forceinline__ float gelu_grad(float x) {
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fmaf(x * x, -0.72134752044448170f, -1.3257480647361592f)));
  return fmaf(x, e, normal_cdf(x));
}

// gelu(x) and d/dx gelu(x) together (one CDF evaluation)
__device__ __forceinline__ void gelu_fwd_grad(float x, float& y, float& dy) {
  const float c = normal_cdf(x);
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fmaf(x * x, -0.72134752044448170f, -1.3257480647361592f)));
  y = x * c;
  dy = fmaf(x, e, c);
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-7 counter RNG.  Dropout masks are never stored: forward and backward regenerate them
// from (seed, stream id, element index).  One call yields 8 x 16-bit lanes; an element is kept iff
// its lane >= thresh16, thresh16 = round(p * 65536)  (p is thereby quantised to 1/65536).
// ---------------------------------------------------------------------------------------------
template <int ROUNDS>
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < ROUNDS; ++i) {
    uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

// 7 rounds everywhere: the fewest for which Philox4x32 is Crush-resistant (Salmon et al. 2011).  (Host-computed round keys
// passed as kernel parameters were tried: -3 us on the attn-out GEMM's dropout epilogue, but the larger parameter block cost
// the attention kernels 13 us per layer in a same-box A/B, so the keys are bumped in registers.)
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) { return philox4x32<7>(ctr, key); }

struct DropoutCfg {
  uint32_t thresh16;   // 0 => dropout disabled
  float scale;         // 1 / (1 - thresh16/65536)
  uint32_t seed_lo, seed_hi;
  uint32_t stream;     // distinguishes the dropout sites of one step (layer, op)
  // Optional DEVICE-resident 64-bit seed (vt_set_seed_source) XORed into (seed_lo, seed_hi) by the kernel: a CUDA graph bakes
  // every by-value argument in at capture, so a replayed step can only get a fresh dropout mask from device memory.
  const uint32_t* seed_dev;
};

const uint32_t* seed_source();   // api.cu: the pointer set by vt_set_seed_source (null = seeds are passed by value only)
const int32_t* attention_bwd_schedule(int* ctas);   // api.cu: the calling thread's item schedule for vt_attention_bwd_packed or null
uint32_t* attention_keep_bits();  // api.cu: the calling thread's keep-bit store for the next attention launches (vt_set_attention_keep_bits) or null
const int32_t* row_limit();      // api.cu: the calling thread's device-resident row count {rows_padded, rows} (vt_set_row_limit) or null

// every kernel that draws dropout masks calls this once on its copy of the configuration
__device__ __forceinline__ DropoutCfg resolve_dropout(DropoutCfg d) {
  if (d.thresh16 && d.seed_dev) {
    d.seed_lo ^= __ldg(d.seed_dev);
    d.seed_hi ^= __ldg(d.seed_dev + 1);
  }
  return d;
}
// rows of a launch under a device-resident row limit: min(host rows, limit[which]); which = 0 padded count, 1 true count
__device__ __forceinline__ long long limited_rows(long long rows, const int32_t* limit, int which = 0) {
  if (!limit) return rows;
  const long long l = (long long)__ldg(limit + which);
  return l < rows ? l : rows;
}

inline DropoutCfg make_dropout(float p, uint64_t seed, uint32_t stream) {
  DropoutCfg d;
  d.seed_dev = seed_source();
  double q = p <= 0.f ? 0.0 : (double)p;
  uint32_t t = (uint32_t)(q * 65536.0 + 0.5);
  if (t > 65535u) t = 65535u;
  d.thresh16 = t;
  d.scale = (float)(1.0 / (1.0 - (double)t / 65536.0));
  d.seed_lo = (uint32_t)seed;
  d.seed_hi = (uint32_t)(seed >> 32);
  d.stream = stream;
  return d;
}

// 8 consecutive elements starting at linear index idx (idx % 8 == 0): keep-multipliers (0 or scale).
__device__ __forceinline__ void dropout_mult8(const DropoutCfg& d, uint64_t idx, float* m) {
  uint64_t g = idx >> 3;
  uint4 r = philox4x32_10(make_uint4((uint32_t)g, (uint32_t)(g >> 32), 0x5eedu, d.stream), make_uint2(d.seed_lo, d.seed_hi));
  uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m[2 * i] = ((w[i] & 0xffffu) >= d.thresh16) ? d.scale : 0.f;
    m[2 * i + 1] = ((w[i] >> 16) >= d.thresh16) ? d.scale : 0.f;
  }
}

// Attention-probability dropout: the 16-bit lane of element (row i, col j) of head bh.  The counter
// is laid out so that a thread holding the mma.sync accumulator fragment of a 16x32 score block
// (cols 8n+2t, 8n+2t+1 for n = 4g..4g+3, t = lane%4) needs exactly ONE Philox call per row per
// 32-column group; it is nevertheless a pure function of (bh, i, j).
__device__ __forceinline__ uint4 attn_dropout_rand(const DropoutCfg& d, uint32_t bh, uint32_t i, uint32_t jgroup, uint32_t t) {
  // 7 rounds: the fewest for which Philox4x32 is Crush-resistant (Salmon et al. 2011); the generator is a large share
  // of the attention kernels' issue slots (one 16-bit lane per probability)
  return philox4x32<7>(make_uint4(jgroup * 4u + t, i, bh, d.stream), make_uint2(d.seed_lo, d.seed_hi));
}

}  // namespace vt
