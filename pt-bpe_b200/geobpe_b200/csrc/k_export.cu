// f-3 token-stream export: Tokenizer.tokenize (tokenizer.py:379-392) + BPE.quantize (bpe.py:928-956) for ALL
// backbones in one pass over the token arrays.  Token j of backbone b emits its motif id (position of its token in
// BPE._tokens = tid[j]) and, unless it is the backbone's last token, the ids of the three glue values that follow a
// residue-aligned token: DIHEDRAL omega, DIHEDRAL phi, BOND_ANGLE C:1N:1CA, each as V + cum_bin_count(type) + bin
// (res_init counts only omega / phi / C:1N:1CA, bpe.py:905-915).  A backbone with m tokens emits 4m - 3 ids, so token j
// writes at 4 j - 3 b.
#include "common.cuh"

__global__ void __launch_bounds__(256) k_token_stream(const u32* __restrict__ tid, const u32* __restrict__ glue, const u32* __restrict__ pos,
                                                      const u32* __restrict__ bb, i64 T, u32 V, u32 off_omega, u32 off_phi, u32 off_cn,
                                                      i64* __restrict__ out) {
  for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < T; j += (i64)gridDim.x * blockDim.x) {
    const u32 g = glue[j];
    const i64 o = 4 * j - 3 * (i64)bb[pos[j]];
    out[o] = (i64)tid[j];
    if (g & GB_LAST_FLAG) continue;
    out[o + 1] = (i64)(V + off_omega + (g & 0x3FFu));
    out[o + 2] = (i64)(V + off_phi + ((g >> 20) & 0x3FFu));
    out[o + 3] = (i64)(V + off_cn + ((g >> 10) & 0x3FFu));
  }
}
extern "C" int gb_token_stream(const uint32_t* tid, const uint32_t* glue, const uint32_t* pos, const uint32_t* bb, int64_t T,
                               uint32_t V, uint32_t off_omega, uint32_t off_phi, uint32_t off_cn, int64_t* out, void* stream) {
  if (T <= 0) return 0;
  int sms = gb_sm_count();
  k_token_stream<<<gb_blocks_for(T, 256 * 4, (sms > 0 ? sms : 148) * 8), 256, 0, (cudaStream_t)stream>>>(tid, glue, pos, bb, T, V, off_omega,
                                                                                                            off_phi, off_cn, out);
  GB_LAUNCH_CK();
  return 0;
}
