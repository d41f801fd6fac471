// Small device helpers shared by the persistent decoder kernels (gtn_decoder_fwd2.cu / gtn_decoder_bwd2.cu):
// release / acquire words of the grid barrier and the activation-plane writer.
//
// Precision planes.  prec==1: operands are plain bf16.  prec==3 ("fp32 mode"): every fp32 value v is
// split into three bf16 planes hi+lo+lo2 (24 mantissa bits) and K is expanded six-fold so that the same
// bf16 engine accumulates  hi*hi + lo*hi + lo2*hi + hi*lo + lo*lo + hi*lo2  (all terms >= 2^-24 rel.).
//   weight sub-segments  s=0..5 hold  (hi, lo, lo2, hi, lo, hi)
//   activation sub-segs  s=0..5 hold  (hi, hi, hi,  lo, lo, lo2)
#pragma once
#include "gtn_common.cuh"

namespace gtn {

typedef __nv_bfloat16 bf16;

GTN_DEVICE uint32_t ld_acquire_u32(const unsigned int* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
GTN_DEVICE void red_release_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Write activation value v for logical index k of a segment whose padded logical length is kpad.
// prec==1: x[k]; prec==3: six sub-segments (hi,hi,hi,lo,lo,lo2) of length kpad each.
GTN_DEVICE void store_act(bf16* x, int k, int kpad, int prec, float v) {
  const bf16 hi = __float2bfloat16_rn(v);
  if (prec == 1) { x[k] = hi; return; }
  const float r1 = v - __bfloat162float(hi);
  const bf16 lo = __float2bfloat16_rn(r1);
  const bf16 lo2 = __float2bfloat16_rn(r1 - __bfloat162float(lo));
  x[k] = hi; x[kpad + k] = hi; x[2 * kpad + k] = hi;
  x[3 * kpad + k] = lo; x[4 * kpad + k] = lo;
  x[5 * kpad + k] = lo2;
}

GTN_DEVICE float tanh_acc(float x) { return tanhf(x); }
GTN_DEVICE float sigmoid_acc(float x) { return 1.0f / (1.0f + expf(-x)); }

}  // namespace gtn
