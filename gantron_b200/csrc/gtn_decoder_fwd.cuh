// Shared declarations of the persistent decoder forward kernels:
//   gtn_decoder_fwd2.cu  autoregressive inference (and the sequential fall-back of the teacher-forced pass)
//   gtn_decoder_fwd3.cu  teacher-forced training pass, software-pipelined
#pragma once
#include <cstdlib>
#include "gtn_decoder2.cuh"
#include "gtn_internal.h"
#include "../../include/gantron_b200.h"

#define BFP(x) (reinterpret_cast<bf16*>(const_cast<void*>(static_cast<const void*>(x))))

namespace gtn {

typedef gtn_decoder_fwd_params DecFwd;

struct Fwd2Maps {
  CUtensorMap w_att, w_dec, w_pre1, w_pre2, w_proj, uloc;
  CUtensorMap pre, ctx, ha, hd, xmel, xp1, xp2;
  CUtensorMap qpart;      // fp32 [Ha/32][B][A] partial attention queries (pipelined training kernel: one box per sequence)
  CUtensorMap xmel_c, xp1_c, hd_c, ctx_c;   // 64-row boxes of the same buffers (batch-chunked prenet / projection phases, B > 64)
};

static constexpr int JH_MAX = 128;     // encoder positions per CTA of a pair (T_in <= 256)
static constexpr int TIN_MAX = 256;
static constexpr int KS_MAX = 63;

GTN_DEVICE float block_reduce2(float v, float* scr, bool is_max) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = is_max ? warp_max(v) : warp_sum(v);
  __syncthreads();
  if (lane == 0) scr[warp] = v;
  __syncthreads();
  float r = scr[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) r = is_max ? fmaxf(r, scr[w]) : r + scr[w];
  return r;
}

template <typename MT>
GTN_DEVICE void ld8(const MT* p, float (&o)[8]);
template <>
GTN_DEVICE void ld8<float>(const float* p, float (&o)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}
template <>
GTN_DEVICE void ld8<bf16>(const bf16* p, float (&o)[8]) {
  const uint4 q = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    o[2 * i] = __uint_as_float(w[i] << 16);
    o[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}

int make_tma_map_3d(CUtensorMap* out, const void* base, long long d0, long long d1, long long d2, long long stride1_elems,
                    long long stride2_elems, int box1);
int make_tma_map_f32_3d(CUtensorMap* out, const void* base, long long d0, long long d1, long long d2, long long stride1_elems,
                        long long stride2_elems, int box0, int box1, int box2);

// teacher-forced pass, pipelined kernel (gtn_decoder_fwd3.cu); FWD3_NOT_COVERED when the configuration is not covered by it
static constexpr int FWD3_NOT_COVERED = 0x7fff0001;
int launch_train_fwd3(const Fwd2Maps& maps, const DecFwd& p, cudaStream_t st);

}  // namespace gtn
