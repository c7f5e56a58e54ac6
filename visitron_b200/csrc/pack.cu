// Padding-free execution of the pre-training step: the loader pads every sample to 511 text + 256 region
// slots (tasks/viewpoint_select/data_loader_pretrain.py:636-712) and ~30 % of the rows of a batch are padding.  Padded rows are
// masked as keys (their probability underflows to exactly 0 in the reference's fp32 softmax) and carry no labels, so they
// influence neither the loss nor any gradient: the encoder can run on the rows that exist.  Which rows those are is decided on
// the device (no host sync): pack_plan_kernel compacts the [B, S] layout sample by sample into packed rows
// [sample 0's rows | sample 1's rows | ...], and every kernel of the step takes its row range from the plan.
#include "common.cuh"

namespace vt {

void count_launch(int n);

struct PackPlanArgs {
  const float* bias;          // [B*S] additive key bias of vt_mask_bias: a row is attended iff its mask != 0, i.e. bias != -10000
  const long long* lab_a;     // [B*S] or null: a labelled row is kept even if it is masked as a key
  const long long* lab_b;
  int B, S, cap;
  int* row_src;               // [cap]   padded index of packed row r (-1: padding row of the packed layout)
  int* row_dst;               // [B*S]   packed row of padded index i (-1: dropped)
  int* row_b;                 // [cap]   sample of packed row r (-1: padding row)
  int* start;                 // [B+1]   first packed row of every sample; start[B] = rows kept
  int* counts;                // [4]     {rows kept rounded up to 128, + 128 (the row limit of the step's launches), rows kept, 0, 0}
  float* bias_p;              // [cap]   bias in packed order
  long long* lab_a_p;         // [cap]   labels in packed order (-1 on padding rows), or null
  long long* lab_b_p;
  long long* cls_rows;        // [B]     packed row of every sample's first position ([CLS]: the pooler's input)
};

// One CTA: M = B*S <= ~100 k flags, a contiguous chunk per thread, block scan, ordered write (packed order == padded order with the
// dropped rows removed, so every sample's rows stay contiguous and in position order).
__global__ void __launch_bounds__(1024)
pack_plan_kernel(const PackPlanArgs a) {
  __shared__ int s_warp[32];
  const long long M = (long long)a.B * a.S;
  const long long per = (M + blockDim.x - 1) / blockDim.x;
  const long long i0 = min(M, per * threadIdx.x), i1 = min(M, i0 + per);
  auto keep = [&](long long i) {
    return a.bias[i] != -10000.0f || (a.lab_a && a.lab_a[i] != -1) || (a.lab_b && a.lab_b[i] != -1) || (i % a.S) == 0;
  };
  int cnt = 0;
  for (long long i = i0; i < i1; ++i) cnt += keep(i) ? 1 : 0;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int v = lane < (int)(blockDim.x >> 5) ? s_warp[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += t;
    }
    s_warp[lane] = v;
  }
  __syncthreads();
  const int total = s_warp[(blockDim.x >> 5) - 1];
  int pos = (inc - cnt) + (warp ? s_warp[warp - 1] : 0);
  for (long long i = i0; i < i1; ++i) {
    const int b = (int)(i / a.S);
    if ((i % a.S) == 0) { a.start[b] = pos; a.cls_rows[b] = pos; }     // position 0 is always kept
    if (keep(i)) {
      a.row_src[pos] = (int)i;
      a.row_b[pos] = b;
      a.row_dst[i] = pos;
      a.bias_p[pos] = a.bias[i];
      if (a.lab_a_p) a.lab_a_p[pos] = a.lab_a[i];
      if (a.lab_b_p) a.lab_b_p[pos] = a.lab_b[i];
      ++pos;
    } else {
      a.row_dst[i] = -1;
    }
  }
  const int limit = min(a.cap, (total + 127) / 128 * 128 + 128);
  for (int r = total + threadIdx.x; r < a.cap; r += blockDim.x) {
    a.row_src[r] = -1;
    a.row_b[r] = -1;
    a.bias_p[r] = 0.f;
    if (a.lab_a_p) a.lab_a_p[r] = -1;
    if (a.lab_b_p) a.lab_b_p[r] = -1;
  }
  if (threadIdx.x == 0) {
    a.start[a.B] = total;
    a.counts[0] = limit;
    a.counts[1] = total;
    a.counts[2] = 0;
    a.counts[3] = 0;
  }
}

// rows [counts[1], counts[0]) of a row-major buffer := 0 (at most 255 rows: the packed layout's padding rows, which a kernel that
// works sample by sample -- attention -- never writes, and which must be finite zeros for the GEMMs that contract over rows)
__global__ void __launch_bounds__(256)
zero_pad_rows_kernel(unsigned char* __restrict__ buf, long long pitch_bytes, long long row_bytes, const int* __restrict__ counts) {
  const long long row = (long long)counts[1] + blockIdx.x;
  if (row >= counts[0]) return;
  uint4* dst = reinterpret_cast<uint4*>(buf + row * pitch_bytes);
  for (long long i = threadIdx.x; i < row_bytes / 16; i += blockDim.x) dst[i] = make_uint4(0u, 0u, 0u, 0u);
}

}  // namespace vt

using namespace vt;

extern "C" int vt_pack_plan(const float* bias, const int64_t* labels_a, const int64_t* labels_b, int64_t B, int64_t S, int64_t capacity,
                            int32_t* row_src, int32_t* row_dst, int32_t* row_b, int32_t* start, int32_t* counts, float* bias_packed,
                            int64_t* labels_a_packed, int64_t* labels_b_packed, int64_t* cls_rows, vt_stream stream) {
  VT_REQUIRE(bias && row_src && row_dst && row_b && start && counts && bias_packed && cls_rows, "vt_pack_plan: null pointer");
  VT_REQUIRE(B > 0 && S > 0 && B * S < (1ll << 30) && capacity >= (B * S + 127) / 128 * 128 + 128 && capacity % 128 == 0,
             "vt_pack_plan: capacity must be a multiple of 128 and at least roundup(B*S, 128) + 128");
  VT_REQUIRE((!labels_a || labels_a_packed) && (!labels_b || labels_b_packed), "vt_pack_plan: packed label outputs required");
  PackPlanArgs a;
  a.bias = bias; a.lab_a = (const long long*)labels_a; a.lab_b = (const long long*)labels_b;
  a.B = (int)B; a.S = (int)S; a.cap = (int)capacity;
  a.row_src = row_src; a.row_dst = row_dst; a.row_b = row_b; a.start = start; a.counts = counts; a.bias_p = bias_packed;
  a.lab_a_p = (long long*)labels_a_packed; a.lab_b_p = (long long*)labels_b_packed; a.cls_rows = (long long*)cls_rows;
  pack_plan_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(a);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_zero_pad_rows(void* buf, int64_t pitch_bytes, int64_t row_bytes, const int32_t* counts, vt_stream stream) {
  VT_REQUIRE(buf && counts && row_bytes % 16 == 0 && pitch_bytes % 16 == 0 && ((uintptr_t)buf & 15) == 0, "vt_zero_pad_rows: 16-byte alignment required");
  zero_pad_rows_kernel<<<256, 256, 0, (cudaStream_t)stream>>>((unsigned char*)buf, pitch_bytes, row_bytes, counts);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}
