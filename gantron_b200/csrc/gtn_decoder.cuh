// Shared device machinery of the persistent decoder kernels (forward: gtn_decoder_fwd.cu, backward:
// gtn_decoder_bwd.cu): grid barrier, the skinny-batch GEMM engine, activation-plane writers.
//
// GEMM engine.  Every per-frame matrix product of the decoder (LSTM gates, their data gradients, the
// mel/gate projection, the inference-time prenet) has the shape
//        OUT[rows, B] = W[rows, K] . X[B, K]^T          B <= 64 per pass, K in the thousands
// i.e. weight-streaming bound.  A CTA owns 16*MT weight rows; its 8 warps split K; each warp keeps the
// full (16*MT x 8*NT) fp32 accumulator tile in registers and feeds mma.sync.m16n8k16 (bf16 in, fp32
// accumulate) straight from 16-byte global loads (weights: read-only path; activations: L2-only loads
// because other CTAs wrote them in the previous phase).  K is consumed in chunks of 32: a thread's 16 B
// load supplies the same 8 consecutive k of both operands, so no shared-memory staging or ldmatrix is
// needed (the k-permutation inside a chunk is identical for A and B).  Partial tiles are then reduced
// across the 8 warps through shared memory.
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

// Grid-wide barrier for a cooperative (co-resident) launch.  `epoch` is a per-thread running target.
GTN_DEVICE void grid_sync(unsigned int* counter, unsigned int& epoch, unsigned int nblocks) {
  __syncthreads();
  epoch += nblocks;
  if (threadIdx.x == 0) {
    __threadfence();
    red_release_add(counter, 1u);
    long long t0 = clock64();
    while (ld_acquire_u32(counter) < epoch) {
      if (clock64() - t0 > 20000000000LL) __trap();   // ~10 s: never hang the GPU on a logic error
    }
    __threadfence();
  }
  __syncthreads();
}

GTN_DEVICE void mma_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

GTN_DEVICE uint4 ldg_nc_v4(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
GTN_DEVICE uint4 ldcg_v4(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }

// One activation segment of the K axis: X[b][k], k < 32*nch, row stride ld (bf16 elements).
struct Seg {
  const bf16* x;
  int ld;
  int nch;
};

// Accumulate acc += W[rows, chunks c_begin..c_end) . X^T for this warp.
//   wrow[mt][h]: this thread's weight row (tile mt, half h: row lane/4 + 8h) already offset by 8*(lane%4);
//                chunk c lives at wrow + 32*c (weights are packed contiguously across segments).
//   segs: activation segments in the same K order; B rows valid, batch offset b0.
template <int MT, int NT>
GTN_DEVICE void warp_gemm(float (&acc)[MT][NT][4], const bf16* const (&wrow)[MT][2], const Seg* segs, int nseg,
                          int c_begin, int c_end, int B, int b0, int lane) {
  if (c_begin >= c_end) return;
  const int q8 = (lane & 3) * 8;
  const int rr = lane >> 2;
  // locate c_begin
  int si = 0, cs = c_begin;
  while (si < nseg - 1 && cs >= segs[si].nch) { cs -= segs[si].nch; ++si; }
  uint4 a_cur[MT][2], x_cur[NT], a_nxt[MT][2], x_nxt[NT];
  auto load = [&](int c, int s_i, int c_s, uint4 (&a)[MT][2], uint4 (&x)[NT]) {
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      a[mt][0] = ldg_nc_v4(wrow[mt][0] + (size_t)c * 32);
      a[mt][1] = ldg_nc_v4(wrow[mt][1] + (size_t)c * 32);
    }
    const bf16* xb = segs[s_i].x + (size_t)c_s * 32 + q8;
    const int ld = segs[s_i].ld;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int row = b0 + nt * 8 + rr;
      x[nt] = row < B ? ldcg_v4(xb + (size_t)row * ld) : make_uint4(0u, 0u, 0u, 0u);
    }
  };
  load(c_begin, si, cs, a_cur, x_cur);
  for (int c = c_begin; c < c_end; ++c) {
    int sn = si, cn = cs + 1;
    if (cn >= segs[sn].nch && sn < nseg - 1) { cn = 0; ++sn; }
    if (c + 1 < c_end) load(c + 1, sn, cn, a_nxt, x_nxt);
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        mma_16816(acc[mt][nt], a_cur[mt][0].x, a_cur[mt][1].x, a_cur[mt][0].y, a_cur[mt][1].y, x_cur[nt].x, x_cur[nt].y);
        mma_16816(acc[mt][nt], a_cur[mt][0].z, a_cur[mt][1].z, a_cur[mt][0].w, a_cur[mt][1].w, x_cur[nt].z, x_cur[nt].w);
      }
    }
    if (c + 1 < c_end) {
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) { a_cur[mt][0] = a_nxt[mt][0]; a_cur[mt][1] = a_nxt[mt][1]; }
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) x_cur[nt] = x_nxt[nt];
    }
    si = sn; cs = cn;
  }
}

// Spill a warp's accumulator tile to the reduction buffer red[warp][mt][nt][c][lane].
template <int MT, int NT>
GTN_DEVICE void store_partials(float* red, const float (&acc)[MT][NT][4], int warp, int lane) {
  float* base = red + (size_t)warp * MT * NT * 128;
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int c = 0; c < 4; ++c) base[((mt * NT + nt) * 4 + c) * 32 + lane] = acc[mt][nt][c];
}
// Sum over the 8 warps of tile element (mt, row r in [0,16), col in [0, 8*NT)).
template <int MT, int NT>
GTN_DEVICE float reduce_partial(const float* red, int mt, int r, int col) {
  const int nt = col >> 3, l = (r & 7) * 4 + ((col & 7) >> 1), c = (col & 1) + ((r >> 3) << 1);
  const float* p = red + ((mt * NT + nt) * 4 + c) * 32 + l;
  float s = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) s += p[(size_t)w * MT * NT * 128];
  return s;
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
