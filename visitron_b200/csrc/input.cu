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
