// Multi-tensor AdamW (decoupled weight decay, optional bias correction) and gradient-norm clipping: the step after
// backward.  Replaces transformers/pytorch_transformers/optimization.py:141-189 (AdamW.step: ~10 ATen launches per
// parameter, 200 parameters) and torch.nn.utils.clip_grad_norm_ at tasks/viewpoint_select/agent.py:514-518.
//
// All parameters of all groups are updated by ONE launch: a device-resident segment table (parameter / gradient / moment
// pointers, length, group) is cut into chunks of OPT_CHUNK elements, one CTA per chunk; per-group scalars (lr, lr * decay,
// bias-corrected step size) travel in the kernel parameters.  HBM-bound: 16 B read + 12 B written per element.
#include "common.cuh"

namespace vt {

void count_launch(int n);

constexpr int OPT_THREADS = 256;
constexpr int OPT_CHUNK = 8192;   // elements per CTA: 8 float4 per thread
constexpr int OPT_MAX_GROUPS = 8;

struct OptSeg {        // mirrors vt_opt_segment
  float* p;
  float* g;
  float* m;
  float* v;
  bf16* p_bf16;        // optional bf16 shadow of p (the tensor-core operand): written in the same pass, so no cast launch follows
  long long n;
  int group;
  int first_chunk;     // prefix sum of ceil(n / OPT_CHUNK)
};

struct AdamGroups {
  float step_size[OPT_MAX_GROUPS];   // lr * sqrt(1 - beta2^t) / (1 - beta1^t)  (or lr without bias correction)
  float lr_decay[OPT_MAX_GROUPS];    // lr * weight_decay
};

__device__ __forceinline__ int find_segment(const OptSeg* __restrict__ segs, int n_segs, int chunk) {
  int lo = 0, hi = n_segs - 1;
  while (lo < hi) {   // last segment whose first_chunk <= chunk
    const int mid = (lo + hi + 1) >> 1;
    if (segs[mid].first_chunk <= chunk) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// the reference's order of operations, element by element (optimization.py:165-187):
//   m = m*b1 + (1-b1)*g ; v = v*b2 + (1-b2)*g*g ; p -= step_size * m / (sqrt(v) + eps) ; p -= lr*decay * p
// (1 - beta) is formed in double on the host like the reference's Python scalars: 1 - float(0.999) is off by 1.3e-5 relative
struct AdamHyper { float b1, b2, om1, om2, eps; };

__device__ __forceinline__ void adamw_one(float& p, float g, float& m, float& v, const AdamHyper& h, float step, float lrd) {
  const float eps = h.eps;
  m = fmaf(h.om1, g, m * h.b1);
  v = fmaf(h.om2 * g, g, v * h.b2);
  p = p - step * (m / (sqrtf(v) + eps));
  if (lrd > 0.f) p = p - lrd * p;
}

__global__ void __launch_bounds__(OPT_THREADS)
adamw_kernel(const OptSeg* __restrict__ segs, int n_segs, AdamGroups gr, AdamHyper hp, const float* __restrict__ gscale) {
  const int si = find_segment(segs, n_segs, (int)blockIdx.x);
  const OptSeg s = segs[si];
  const long long off = (long long)((int)blockIdx.x - s.first_chunk) * OPT_CHUNK;
  const long long n = min((long long)OPT_CHUNK, s.n - off);
  const float step = gr.step_size[s.group], lrd = gr.lr_decay[s.group];
  const float gs = gscale ? *gscale : 1.0f;   // clip coefficient of vt_clip_coef (1 when the norm is below the bound)
  float* p = s.p + off; const float* g = s.g + off; float* m = s.m + off; float* v = s.v + off;
  bf16* pb = s.p_bf16 ? s.p_bf16 + off : nullptr;
  const bool vec = (((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0 && ((uintptr_t)pb & 7) == 0;
  if (vec && n == OPT_CHUNK) {
    // full chunk: 8 float4 per thread in batches of 4, so that 16 independent 16-byte loads are in flight per thread
#pragma unroll
    for (int it = 0; it < OPT_CHUNK / 4 / OPT_THREADS; it += 4) {
      float4 pv[4], mv[4], vv[4], gv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = (it + u) * OPT_THREADS + threadIdx.x;
        pv[u] = reinterpret_cast<float4*>(p)[i]; mv[u] = reinterpret_cast<float4*>(m)[i]; vv[u] = reinterpret_cast<float4*>(v)[i];
        gv[u] = __ldcs(reinterpret_cast<const float4*>(g) + i);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = (it + u) * OPT_THREADS + threadIdx.x;
        adamw_one(pv[u].x, gv[u].x * gs, mv[u].x, vv[u].x, hp, step, lrd);
        adamw_one(pv[u].y, gv[u].y * gs, mv[u].y, vv[u].y, hp, step, lrd);
        adamw_one(pv[u].z, gv[u].z * gs, mv[u].z, vv[u].z, hp, step, lrd);
        adamw_one(pv[u].w, gv[u].w * gs, mv[u].w, vv[u].w, hp, step, lrd);
        reinterpret_cast<float4*>(p)[i] = pv[u]; reinterpret_cast<float4*>(m)[i] = mv[u]; reinterpret_cast<float4*>(v)[i] = vv[u];
        if (pb) reinterpret_cast<uint2*>(pb)[i] = make_uint2(pack_bf16(pv[u].x, pv[u].y), pack_bf16(pv[u].z, pv[u].w));
      }
    }
  } else if (vec) {
    const long long n4 = n >> 2;
    for (long long i = threadIdx.x; i < n4; i += OPT_THREADS) {
      float4 pv = reinterpret_cast<float4*>(p)[i], mv = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
      const float4 gv = __ldcs(reinterpret_cast<const float4*>(g) + i);
      adamw_one(pv.x, gv.x * gs, mv.x, vv.x, hp, step, lrd);
      adamw_one(pv.y, gv.y * gs, mv.y, vv.y, hp, step, lrd);
      adamw_one(pv.z, gv.z * gs, mv.z, vv.z, hp, step, lrd);
      adamw_one(pv.w, gv.w * gs, mv.w, vv.w, hp, step, lrd);
      reinterpret_cast<float4*>(p)[i] = pv; reinterpret_cast<float4*>(m)[i] = mv; reinterpret_cast<float4*>(v)[i] = vv;
      if (pb) reinterpret_cast<uint2*>(pb)[i] = make_uint2(pack_bf16(pv.x, pv.y), pack_bf16(pv.z, pv.w));
    }
    for (long long i = (n4 << 2) + threadIdx.x; i < n; i += OPT_THREADS) {
      adamw_one(p[i], g[i] * gs, m[i], v[i], hp, step, lrd);
      if (pb) pb[i] = __float2bfloat16(p[i]);
    }
  } else {
    for (long long i = threadIdx.x; i < n; i += OPT_THREADS) {
      adamw_one(p[i], g[i] * gs, m[i], v[i], hp, step, lrd);
      if (pb) pb[i] = __float2bfloat16(p[i]);
    }
  }
}

// sum of squares of every gradient: one partial per chunk (fixed order inside the CTA), summed in chunk order by
// clip_coef_kernel, so the norm is deterministic
__global__ void __launch_bounds__(OPT_THREADS)
grad_sumsq_kernel(const OptSeg* __restrict__ segs, int n_segs, float* __restrict__ partials) {
  __shared__ float red[OPT_THREADS / 32];
  const int si = find_segment(segs, n_segs, (int)blockIdx.x);
  const OptSeg s = segs[si];
  const long long off = (long long)((int)blockIdx.x - s.first_chunk) * OPT_CHUNK;
  const long long n = min((long long)OPT_CHUNK, s.n - off);
  const float* g = s.g + off;
  float acc = 0.f;
  if (((uintptr_t)g & 15) == 0) {
    const long long n4 = n >> 2;
    for (long long i = threadIdx.x; i < n4; i += OPT_THREADS) {
      const float4 t = reinterpret_cast<const float4*>(g)[i];
      acc = fmaf(t.x, t.x, fmaf(t.y, t.y, fmaf(t.z, t.z, fmaf(t.w, t.w, acc))));
    }
    for (long long i = (n4 << 2) + threadIdx.x; i < n; i += OPT_THREADS) acc = fmaf(g[i], g[i], acc);
  } else {
    for (long long i = threadIdx.x; i < n; i += OPT_THREADS) acc = fmaf(g[i], g[i], acc);
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < OPT_THREADS / 32; ++w) t += red[w];
    partials[blockIdx.x] = t;
  }
}

// out[0] = total L2 norm, out[1] = clip coefficient min(1, max_norm / (norm + 1e-6))  (torch.nn.utils.clip_grad_norm_)
__global__ void __launch_bounds__(256)
clip_coef_kernel(const float* __restrict__ partials, int n, float max_norm, float* __restrict__ out) {
  __shared__ double red[256];
  double acc = 0.0;
  const int per = (n + 255) / 256;   // contiguous slices in thread order: a fixed summation order
  for (int i = threadIdx.x * per; i < min(n, (int)(threadIdx.x + 1) * per); ++i) acc += (double)partials[i];
  red[threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < 256; ++i) t += red[i];
    const float norm = (float)sqrt(t);
    const float c = max_norm / (norm + 1e-6f);
    out[0] = norm;
    out[1] = c < 1.0f ? c : 1.0f;
  }
}

__global__ void __launch_bounds__(OPT_THREADS)
scale_grads_kernel(const OptSeg* __restrict__ segs, int n_segs, const float* __restrict__ coef) {
  const float c = *coef;
  if (c >= 1.0f) return;
  const int si = find_segment(segs, n_segs, (int)blockIdx.x);
  const OptSeg s = segs[si];
  const long long off = (long long)((int)blockIdx.x - s.first_chunk) * OPT_CHUNK;
  const long long n = min((long long)OPT_CHUNK, s.n - off);
  float* g = s.g + off;
  for (long long i = threadIdx.x; i < n; i += OPT_THREADS) g[i] *= c;
}

}  // namespace vt

using namespace vt;

extern "C" int64_t vt_opt_chunk(void) { return OPT_CHUNK; }

extern "C" int vt_adamw_step(const void* segments_dev, int64_t n_segments, int64_t n_chunks, int64_t n_groups, const float* step_size,
                             const float* lr_decay, double beta1, double beta2, double eps, const float* grad_scale_dev, vt_stream stream) {
  VT_REQUIRE(segments_dev && n_segments > 0 && n_chunks > 0, "vt_adamw_step: empty segment table");
  VT_REQUIRE(n_groups > 0 && n_groups <= OPT_MAX_GROUPS && step_size && lr_decay, "vt_adamw_step: 1..%d parameter groups", OPT_MAX_GROUPS);
  VT_REQUIRE(n_chunks < (1ll << 31), "vt_adamw_step: too many chunks");
  AdamGroups gr = {};
  for (int i = 0; i < n_groups; ++i) { gr.step_size[i] = step_size[i]; gr.lr_decay[i] = lr_decay[i]; }
  const AdamHyper hp = {(float)beta1, (float)beta2, (float)(1.0 - beta1), (float)(1.0 - beta2), (float)eps};
  adamw_kernel<<<(unsigned)n_chunks, OPT_THREADS, 0, (cudaStream_t)stream>>>((const OptSeg*)segments_dev, (int)n_segments, gr, hp, grad_scale_dev);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_clip_grad_norm(const void* segments_dev, int64_t n_segments, int64_t n_chunks, float max_norm, float* partials,
                                 float* norm_and_coef, int scale_in_place, vt_stream stream) {
  VT_REQUIRE(segments_dev && n_segments > 0 && n_chunks > 0 && partials && norm_and_coef, "vt_clip_grad_norm: bad arguments");
  VT_REQUIRE(n_chunks < (1ll << 31), "vt_clip_grad_norm: too many chunks");
  cudaStream_t st = (cudaStream_t)stream;
  grad_sumsq_kernel<<<(unsigned)n_chunks, OPT_THREADS, 0, st>>>((const OptSeg*)segments_dev, (int)n_segments, partials);
  clip_coef_kernel<<<1, 256, 0, st>>>(partials, (int)n_chunks, max_norm, norm_and_coef);
  int launches = 2;
  if (scale_in_place) {
    scale_grads_kernel<<<(unsigned)n_chunks, OPT_THREADS, 0, st>>>((const OptSeg*)segments_dev, (int)n_segments, norm_and_coef + 1);
    ++launches;
  }
  count_launch(launches);
  VT_CHECK_LAUNCH();
  return VT_OK;
}
