// On-device masked-LM masking of a pre-training batch: the step BEFORE forward.
// Replaces PretrainDataset._mask_tokens, tasks/viewpoint_select/data_loader_pretrain.py:549-612 (three torch.bernoulli
// passes, a randint and five boolean index assignments per SAMPLE on the host CPU) with one launch over the whole batch:
//   masked   = (u1 < p_mask and id not special) or token_class != -1          (masked_token_prediction forces those)
//   label    = id if masked and token_class == -1 else -1
//   id'      = [MASK] if (u2 < 0.8 and masked) or token_class != -1 ; else r if u3 < 0.5 and masked ; else id
//   attention= id != [PAD]
// u1, u2, u3 (24-bit uniforms) and the random word r come from ONE Philox4x32-7 call per token keyed on (seed, token index);
// vt_mlm_draws writes the very same draws out, so the parity tests run the oracle on them and compare bit for bit.
#include "common.cuh"

namespace vt {

void count_launch(int n);

struct MlmDraw { float u1, u2, u3; long long r; };

__device__ __forceinline__ MlmDraw mlm_draw(uint64_t seed, long long i, long long vocab) {
  const uint4 w = philox4x32<7>(make_uint4((uint32_t)i, (uint32_t)((unsigned long long)i >> 32), 0x6d6c6du, 0u),
                                make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
  MlmDraw d;
  d.u1 = (float)(w.x >> 8) * (1.0f / 16777216.0f);
  d.u2 = (float)(w.y >> 8) * (1.0f / 16777216.0f);
  d.u3 = (float)(w.z >> 8) * (1.0f / 16777216.0f);
  d.r = (long long)(w.w % (uint32_t)vocab);
  return d;
}

__global__ void __launch_bounds__(256)
mlm_mask_kernel(const long long* __restrict__ ids, const long long* __restrict__ token_classes, const uint8_t* __restrict__ special,
                long long n, long long vocab, long long mask_id, long long pad_id, float p_mask, uint64_t seed,
                long long* __restrict__ ids_out, long long* __restrict__ labels, uint8_t* __restrict__ attention) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    const long long id = ids[i];
    const bool tc = token_classes && token_classes[i] != -1;
    const bool is_special = id >= 0 && id < vocab ? special[id] != 0 : false;
    const MlmDraw d = mlm_draw(seed, i, vocab);
    const bool masked = (d.u1 < p_mask && !is_special) || tc;
    const bool replaced = (d.u2 < 0.8f && masked) || tc;
    const bool random = d.u3 < 0.5f && masked && !replaced;
    labels[i] = (masked && !tc) ? id : -1;
    ids_out[i] = replaced ? mask_id : (random ? d.r : id);
    attention[i] = id != pad_id ? 1 : 0;
  }
}

__global__ void __launch_bounds__(256)
mlm_draws_kernel(long long n, long long vocab, uint64_t seed, float* __restrict__ u, long long* __restrict__ r) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    const MlmDraw d = mlm_draw(seed, i, vocab);
    u[i] = d.u1; u[n + i] = d.u2; u[2 * n + i] = d.u3;
    r[i] = d.r;
  }
}

// Ragged region features -> the fixed [B, R, *] block of the model batch (data_loader_pretrain.py:656-690): sample b owns rows
// [offsets[b], offsets[b+1]) of the packed arrays; if it has more than R rows the LAST R are kept, otherwise the tail is zero
// padding with attention 0.  One CTA per output row, both arrays (features, location embeddings) in the same launch; 8-byte
// vector copies (row pitches of 2054 and 2052 floats are 8- but not 16-byte aligned).  HBM-bound: read + write of every row.
template <typename V>
__global__ void __launch_bounds__(128)
pack_regions_kernel(const float* __restrict__ a, int da, const float* __restrict__ b, int db, const long long* __restrict__ offsets,
                    int R, float* __restrict__ out_a, float* __restrict__ out_b, long long* __restrict__ att, long long ld_att) {
  constexpr int W = sizeof(V) / 4;
  const int s = blockIdx.x / R, r = blockIdx.x % R;
  const long long lo = offsets[s], n = offsets[s + 1] - lo;
  const long long kept = n < R ? n : R, src = lo + (n - kept) + r;
  const bool valid = r < kept;
  V* oa = reinterpret_cast<V*>(out_a + (long long)blockIdx.x * da);
  V* ob = reinterpret_cast<V*>(out_b + (long long)blockIdx.x * db);
  V zero;
  memset(&zero, 0, sizeof(V));
  if (valid) {
    const V* ia = reinterpret_cast<const V*>(a + src * da);
    const V* ib = reinterpret_cast<const V*>(b + src * db);
    for (int i = threadIdx.x; i < da / W; i += 128) oa[i] = __ldg(ia + i);
    for (int i = threadIdx.x; i < db / W; i += 128) ob[i] = __ldg(ib + i);
  } else {
    for (int i = threadIdx.x; i < da / W; i += 128) oa[i] = zero;
    for (int i = threadIdx.x; i < db / W; i += 128) ob[i] = zero;
  }
  if (att && threadIdx.x == 0) att[(long long)s * ld_att + r] = valid ? 1 : 0;
}

}  // namespace vt

using namespace vt;

extern "C" int vt_mlm_mask(const int64_t* ids, const int64_t* token_classes, const uint8_t* special_lut, int64_t n, int64_t vocab,
                           int64_t mask_id, int64_t pad_id, float p_mask, uint64_t seed, int64_t* ids_out, int64_t* labels,
                           uint8_t* attention_mask, vt_stream stream) {
  VT_REQUIRE(n >= 0 && vocab > 0 && vocab < (1ll << 32) && p_mask >= 0.f && p_mask <= 1.f, "vt_mlm_mask: bad size, vocabulary size or probability");
  if (n == 0) return VT_OK;   // an empty block has no storage behind its pointers
  VT_REQUIRE(ids && special_lut && ids_out && labels && attention_mask, "vt_mlm_mask: null pointer");
  long long g = (n + 255) / 256, cap = (long long)num_sms() * 8;
  mlm_mask_kernel<<<(unsigned)(g < cap ? g : cap), 256, 0, (cudaStream_t)stream>>>(
      (const long long*)ids, (const long long*)token_classes, special_lut, n, vocab, mask_id, pad_id, p_mask, seed, (long long*)ids_out,
      (long long*)labels, attention_mask);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_mlm_draws(int64_t n, int64_t vocab, uint64_t seed, float* u, int64_t* r, vt_stream stream) {
  VT_REQUIRE(n >= 0 && vocab > 0 && vocab < (1ll << 32), "vt_mlm_draws: bad arguments");
  if (n == 0) return VT_OK;
  VT_REQUIRE(u && r, "vt_mlm_draws: null pointer");
  long long g = (n + 255) / 256, cap = (long long)num_sms() * 8;
  mlm_draws_kernel<<<(unsigned)(g < cap ? g : cap), 256, 0, (cudaStream_t)stream>>>(n, vocab, seed, u, (long long*)r);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}

extern "C" int vt_pack_regions(const float* feats, int64_t d_feat, const float* loc, int64_t d_loc, const int64_t* offsets, int64_t samples,
                               int64_t regions, float* out_feats, float* out_loc, int64_t* attention_tail, int64_t ld_attention,
                               vt_stream stream) {
  VT_REQUIRE(samples >= 0 && regions >= 0 && d_feat > 0 && d_loc > 0 && d_feat < (1 << 30) && d_loc < (1 << 30), "vt_pack_regions: bad sizes");
  if (samples == 0 || regions == 0) return VT_OK;
  VT_REQUIRE(offsets && out_feats && out_loc, "vt_pack_regions: null pointer");   // feats / loc may be null when every sample is empty
  VT_REQUIRE(samples * regions < (1ll << 31), "vt_pack_regions: too many rows for one launch");
  const unsigned grid = (unsigned)(samples * regions);
  const bool v2 = d_feat % 2 == 0 && d_loc % 2 == 0 && ((uintptr_t)feats | (uintptr_t)loc | (uintptr_t)out_feats | (uintptr_t)out_loc) % 8 == 0;
  if (v2)
    pack_regions_kernel<float2><<<grid, 128, 0, (cudaStream_t)stream>>>(feats, (int)d_feat, loc, (int)d_loc, (const long long*)offsets, (int)regions,
                                                                        out_feats, out_loc, (long long*)attention_tail, ld_attention);
  else
    pack_regions_kernel<float><<<grid, 128, 0, (cudaStream_t)stream>>>(feats, (int)d_feat, loc, (int)d_loc, (const long long*)offsets, (int)regions,
                                                                       out_feats, out_loc, (long long*)attention_tail, ld_attention);
  count_launch(1);
  VT_CHECK_LAUNCH();
  return VT_OK;
}
