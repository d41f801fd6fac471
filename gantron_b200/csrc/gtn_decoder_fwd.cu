// Persistent cooperative decoder kernel (forward): ONE launch runs every frame of
//   * Decoder.inference  (reference model.py:456-497): prenet -> decode() -> gate test, autoregressive
//   * Decoder.forward    (reference model.py:419-454): the teacher-forced loop of decode() calls
// with a grid-wide barrier between the dependent phases of a frame instead of ~40 kernel launches
// and (at inference) a host sync per frame.  decode() is model.py:378-417.
//
// Phases of one frame (B = batch, H = LSTM width, 8 hidden units -> 32 gate rows per CTA):
//   [infer only] prenet layer 1, prenet layer 2          (model.py:102-105, dropout always on)
//   A  attention LSTM: gates = W_att . [prenet | ctx | h_att]  -> cell -> dropout -> h_att'
//      + this CTA's partial of the attention query  Wq[:, units] . h_att'[units]     (model.py:391-394, 62)
//   B  location-sensitive attention, one CTA per sequence: sum query partials, conv1d(2->F,k) over
//      [prev | cumulative] weights, dense F->A, e = v.tanh(q + loc + processed_memory), masked softmax
//      (warp-shuffle reductions), context = w . memory, cum += w                      (model.py:27-31,62-89,403)
//   C  decoder LSTM: gates = W_dec . [h_att' | ctx' | h_dec] -> cell -> dropout -> h_dec' (model.py:404-409)
//   [infer only] D  mel/gate projection on [h_dec' | ctx'] + stop test                (model.py:411-416,486)
// In training the projection is deferred to one tcgen05 GEMM over all frames (nothing in the loop
// depends on it) and the per-frame activations needed by the backward kernel are saved.
//
// The time-constant columns of `memory` (noise / speaker / emotion: model.py:275-278,655-663) are not
// re-read per frame: their context is  const * sum_j w_j.
#include "gtn_decoder.cuh"
#include "gtn_internal.h"
#include "../../include/gantron_b200.h"

namespace gtn {

typedef gtn_decoder_fwd_params DecFwd;
#define BFP(x) (reinterpret_cast<bf16*>(const_cast<void*>(static_cast<const void*>(x))))

struct FwdSmem {
  float* red;    // [8][2][NT][4][32]
  float* hs;     // [8*NT][8]
  float* wq_s;   // [A][8]   (slice of the query weight for this CTA's hidden units)
  float* wcT;    // [2][KS][F]
  float* wdT;    // [F][A]
  float* v_s;    // [A]
  float* q_s;    // [A]
  float* wcat;   // [2][Tin+KS-1]
  float* f_s;    // [Tin][F]
  float* e_s;    // [Tin]
  float* scr;    // [64]
};

template <int NT>
__host__ __device__ inline size_t fwd_smem_floats(int Tin, int A, int F, int KS) {
  return (size_t)8 * 2 * NT * 128 + 8 * NT * 8 + (size_t)A * 8 + 2 * KS * F + (size_t)F * A + A + A +
         2 * (Tin + KS - 1) + (size_t)Tin * F + Tin + 64;
}

template <int NT>
GTN_DEVICE FwdSmem carve(float* s, const DecFwd& p) {
  FwdSmem m;
  m.red = s; s += 8 * 2 * NT * 128;
  m.hs = s; s += 8 * NT * 8;
  m.wq_s = s; s += p.A * 8;
  m.wcT = s; s += 2 * p.KS * p.F;
  m.wdT = s; s += p.F * p.A;
  m.v_s = s; s += p.A;
  m.q_s = s; s += p.A;
  m.wcat = s; s += 2 * (p.Tin + p.KS - 1);
  m.f_s = s; s += p.Tin * p.F;
  m.e_s = s; s += p.Tin;
  m.scr = s;
  return m;
}

GTN_DEVICE float block_reduce(float v, float* scr, bool is_max) {
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

// ----------------------------------------------------------------------------------------------
// LSTM phase (attention_rnn when is_att, else decoder_rnn)
// ----------------------------------------------------------------------------------------------
template <int NT>
GTN_DEVICE void lstm_phase(const DecFwd& p, const FwdSmem& sm, int t, bool is_att) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int H = is_att ? p.Ha : p.Hd;
  const int Hp = is_att ? p.Hap : p.Hdp;
  const int mult = p.mult;
  const int cur = t & 1, nxt = cur ^ 1;
  Seg segs[3];
  int Ktot;
  const bf16* W;
  const float* bias;
  if (is_att) {
    const int ldp = mult * p.Pp;
    segs[0].x = p.train ? BFP(p.pre_all) + (size_t)t * p.B * ldp : BFP(p.xp2); segs[0].ld = ldp; segs[0].nch = ldp / 32;
    segs[1].x = BFP(p.xctx); segs[1].ld = mult * p.Dp; segs[1].nch = mult * p.Dp / 32;
    segs[2].x = BFP(p.xh_a) + (size_t)cur * p.B * mult * p.Hap; segs[2].ld = mult * p.Hap; segs[2].nch = mult * p.Hap / 32;
    Ktot = mult * (p.Pp + p.Dp + p.Hap);
    W = BFP(p.w_att); bias = p.b_att;
  } else {
    segs[0].x = BFP(p.xh_a) + (size_t)nxt * p.B * mult * p.Hap; segs[0].ld = mult * p.Hap; segs[0].nch = mult * p.Hap / 32;
    segs[1].x = BFP(p.xctx); segs[1].ld = mult * p.Dp; segs[1].nch = mult * p.Dp / 32;
    segs[2].x = BFP(p.xh_d) + (size_t)cur * p.B * mult * p.Hdp; segs[2].ld = mult * p.Hdp; segs[2].nch = mult * p.Hdp / 32;
    Ktot = mult * (p.Hap + p.Dp + p.Hdp);
    W = BFP(p.w_dec); bias = p.b_dec;
  }
  const int nch = Ktot / 32;
  const int c_begin = (int)((long long)warp * nch / 8), c_end = (int)((long long)(warp + 1) * nch / 8);
  bf16* xh_next = (is_att ? BFP(p.xh_a) : BFP(p.xh_d)) + (size_t)nxt * p.B * mult * Hp;
  float* cst = is_att ? p.c_a : p.c_d;
  const float pdrop = is_att ? p.p_att : p.p_dec;
  const unsigned thr = is_att ? p.thr_att : p.thr_dec;
  const float dscale = 1.0f / (1.0f - pdrop);
  const unsigned site = is_att ? 3u : 4u;
  float* gates_all = is_att ? p.gates_a : p.gates_d;
  float* c_all = is_att ? p.c_a_all : p.c_d_all;
  float* h_all = is_att ? p.h_a_all : p.h_d_all;

  for (int ug = blockIdx.x; ug < H / 8; ug += gridDim.x) {
    const int u0 = ug * 8;
    if (is_att) {  // query-weight slice of these 8 units
      for (int i = threadIdx.x; i < p.A * 8; i += blockDim.x) sm.wq_s[i] = p.wq[(size_t)(i >> 3) * p.Ha + u0 + (i & 7)];
    }
    const int rr = lane >> 2, q8 = (lane & 3) * 8;
    const bf16* wrow[2][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int h = 0; h < 2; ++h) wrow[mt][h] = W + (size_t)((2 * mt + h) * H + u0 + rr) * Ktot + q8;

    for (int b0 = 0; b0 < p.B; b0 += 8 * NT) {
      float acc[2][NT][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[mt][nt][c] = 0.f;
      warp_gemm<2, NT>(acc, wrow, segs, 3, c_begin, c_end, p.B, b0, lane);
      __syncthreads();   // previous users of red / hs are done
      store_partials<2, NT>(sm.red, acc, warp, lane);
      __syncthreads();
      const int nb = min(8 * NT, p.B - b0);
      for (int idx = threadIdx.x; idx < nb * 8; idx += blockDim.x) {
        const int u = idx & 7, bl = idx >> 3, b = b0 + bl, unit = u0 + u;
        const float gi = reduce_partial<2, NT>(sm.red, 0, u, bl) + bias[unit];
        const float gf = reduce_partial<2, NT>(sm.red, 0, u + 8, bl) + bias[H + unit];
        const float gg = reduce_partial<2, NT>(sm.red, 1, u, bl) + bias[2 * H + unit];
        const float go = reduce_partial<2, NT>(sm.red, 1, u + 8, bl) + bias[3 * H + unit];
        const float i_ = sigmoid_acc(gi), f_ = sigmoid_acc(gf), g_ = tanh_acc(gg), o_ = sigmoid_acc(go);
        const size_t ci = (size_t)b * H + unit;
        const float c_new = f_ * cst[ci] + i_ * g_;
        cst[ci] = c_new;
        float h = o_ * tanh_acc(c_new);
        if (p.use_dropout && thr != 0u) {
          const uint32_t bits = philox_bits(p.seed, site, (uint32_t)t, (uint32_t)ci);
          h = bits >= thr ? h * dscale : 0.f;
        }
        store_act(xh_next + (size_t)b * mult * Hp, unit, Hp, p.prec, h);
        if (is_att) sm.hs[bl * 8 + u] = h;
        if (p.train) {
          const size_t g0 = ((size_t)t * p.B + b) * 4 * H + unit;
          gates_all[g0] = i_; gates_all[g0 + H] = f_; gates_all[g0 + 2 * H] = g_; gates_all[g0 + 3 * H] = o_;
          c_all[((size_t)t * p.B + b) * H + unit] = c_new;
          h_all[((size_t)t * p.B + b) * H + unit] = h;
        }
      }
      if (is_att) {
        __syncthreads();
        for (int idx = threadIdx.x; idx < nb * p.A; idx += blockDim.x) {
          const int bl = idx / p.A, a = idx - bl * p.A;
          float s = 0.f;
#pragma unroll
          for (int u = 0; u < 8; ++u) s += sm.wq_s[a * 8 + u] * sm.hs[bl * 8 + u];
          p.qpart[((size_t)ug * p.B + b0 + bl) * p.A + a] = s;
        }
      }
    }
    __syncthreads();
  }
}

// ----------------------------------------------------------------------------------------------
// generic 16-row tile linear used by the inference-only phases (prenet layers, projection)
//   kind 0: prenet1 (x = xmel)  kind 1: prenet2 (x = xp1)  kind 2: projection (x = [h_dec' | ctx'])
// ----------------------------------------------------------------------------------------------
template <int NT>
GTN_DEVICE void small_linear_phase(const DecFwd& p, const FwdSmem& sm, int t, int kind) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mult = p.mult;
  Seg segs[2];
  int nseg, Ktot, rows;
  const bf16* W;
  if (kind == 0) {
    segs[0].x = BFP(p.xmel); segs[0].ld = mult * p.NMp; segs[0].nch = mult * p.NMp / 32; nseg = 1;
    Ktot = mult * p.NMp; rows = p.P; W = BFP(p.w_pre1);
  } else if (kind == 1) {
    segs[0].x = BFP(p.xp1); segs[0].ld = mult * p.Pp; segs[0].nch = mult * p.Pp / 32; nseg = 1;
    Ktot = mult * p.Pp; rows = p.P; W = BFP(p.w_pre2);
  } else {
    const int nxt = (t & 1) ^ 1;
    segs[0].x = BFP(p.xh_d) + (size_t)nxt * p.B * mult * p.Hdp; segs[0].ld = mult * p.Hdp; segs[0].nch = mult * p.Hdp / 32;
    segs[1].x = BFP(p.xctx); segs[1].ld = mult * p.Dp; segs[1].nch = mult * p.Dp / 32; nseg = 2;
    Ktot = mult * (p.Hdp + p.Dp); rows = p.NM + 1; W = BFP(p.w_proj);
  }
  const int nch = Ktot / 32;
  const int c_begin = (int)((long long)warp * nch / 8), c_end = (int)((long long)(warp + 1) * nch / 8);
  const int ntiles = (rows + 15) / 16;   // weight matrices are zero-padded to 16*ntiles rows
  const int rr = lane >> 2, q8 = (lane & 3) * 8;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const bf16* wrow[1][2];
    wrow[0][0] = W + (size_t)(tile * 16 + rr) * Ktot + q8;
    wrow[0][1] = W + (size_t)(tile * 16 + rr + 8) * Ktot + q8;
    for (int b0 = 0; b0 < p.B; b0 += 8 * NT) {
      float acc[1][NT][4];
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[0][nt][c] = 0.f;
      warp_gemm<1, NT>(acc, wrow, segs, nseg, c_begin, c_end, p.B, b0, lane);
      __syncthreads();
      store_partials<1, NT>(sm.red, acc, warp, lane);
      __syncthreads();
      const int nb = min(8 * NT, p.B - b0);
      for (int idx = threadIdx.x; idx < nb * 16; idx += blockDim.x) {
        const int r = idx & 15, bl = idx >> 4, b = b0 + bl, row = tile * 16 + r;
        if (row >= rows) continue;
        float val = reduce_partial<1, NT>(sm.red, 0, r, bl);
        if (kind <= 1) {
          val = fmaxf(val, 0.f);
          const uint32_t bits = philox_bits(p.seed, kind == 0 ? 1u : 2u, (uint32_t)t, (uint32_t)(b * p.P + row));
          val = bits >= 0x80000000u ? val * 2.0f : 0.f;     // p = 0.5, always on (model.py:104)
          store_act((kind == 0 ? BFP(p.xp1) : BFP(p.xp2)) + (size_t)b * mult * p.Pp, row, p.Pp, p.prec, val);
        } else {
          val += p.b_proj[row];
          if (row < p.NM) {
            p.mel_out[((size_t)b * p.NM + row) * p.n_steps + t] = val;
            store_act(BFP(p.xmel) + (size_t)b * mult * p.NMp, row, p.NMp, p.prec, val);
          } else {
            p.gate_out[(size_t)b * p.n_steps + t] = val;
            if (sigmoid_acc(val) > p.gate_threshold && p.done[b] == 0) { p.done[b] = 1; p.lengths_out[b] = t + 1; }
          }
        }
      }
    }
    if (kind == 2 && tile == p.NM / 16) {   // the CTA that owns the gate row decides whether everyone is done
      __syncthreads();
      int all = 1;
      for (int b = threadIdx.x; b < p.B; b += blockDim.x) all &= (p.done[b] != 0);
      all = __syncthreads_and(all);
      if (threadIdx.x == 0 && all && !p.ignore_stop) *p.stop = 1;
    }
    __syncthreads();
  }
}

// ----------------------------------------------------------------------------------------------
// attention phase: one CTA per sequence
// ----------------------------------------------------------------------------------------------
template <typename MT>
GTN_DEVICE float ld_mem(const MT* p);
template <>
GTN_DEVICE float ld_mem<float>(const float* p) { return __ldg(p); }
template <>
GTN_DEVICE float ld_mem<bf16>(const bf16* p) { return __bfloat162float(*p); }

template <typename MT>
GTN_DEVICE void attention_phase(const DecFwd& p, const FwdSmem& sm, int t) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Tin = p.Tin, A = p.A, F = p.F, KS = p.KS, half = (p.KS - 1) / 2, Wd = p.Tin + p.KS - 1;
  const int nq = p.Ha / 8;
  const MT* mem = reinterpret_cast<const MT*>(p.mem_enc);
  const MT* pm = reinterpret_cast<const MT*>(p.pm);
  for (int b = blockIdx.x; b < p.B; b += gridDim.x) {
    const int L = p.lengths ? min(p.lengths[b], Tin) : Tin;
    float* wprev = p.w_prev + (size_t)b * p.ldT;
    float* wcum = p.w_cum + (size_t)b * p.ldT;
    for (int a = threadIdx.x; a < A; a += blockDim.x) {
      float s = 0.f;
      for (int c = 0; c < nq; ++c) s += __ldcg(&p.qpart[((size_t)c * p.B + b) * A + a]);
      sm.q_s[a] = s;
      if (p.train) p.q_all[((size_t)t * p.B + b) * A + a] = s;
    }
    for (int i = threadIdx.x; i < 2 * Wd; i += blockDim.x) {
      const int ch = i / Wd, j = i - ch * Wd - half;
      const float val = (j >= 0 && j < Tin) ? (ch == 0 ? wprev[j] : wcum[j]) : 0.f;
      sm.wcat[i] = val;
      if (p.train && ch == 1 && j >= 0 && j < Tin) p.cum_all[((size_t)t * p.B + b) * Tin + j] = val;
    }
    __syncthreads();
    // location conv (2 -> F, kernel KS, zero padded): f[j][c]
    for (int idx = threadIdx.x; idx < Tin * F; idx += blockDim.x) {
      const int j = idx / F, c = idx - j * F;
      float s = 0.f;
      for (int ch = 0; ch < 2; ++ch) {
        const float* wc = sm.wcT + ch * KS * F + c;
        const float* xw = sm.wcat + ch * Wd + j;
        for (int k = 0; k < KS; ++k) s += wc[k * F] * xw[k];
      }
      sm.f_s[idx] = s;
      if (p.train) p.f_all[((size_t)t * p.B + b) * Tin * F + idx] = s;
    }
    __syncthreads();
    // energies: one warp per encoder position
    for (int j = warp; j < Tin; j += 8) {
      if (j >= L) { if (lane == 0) sm.e_s[j] = -INFINITY; continue; }
      float part = 0.f;
      const float* fj = sm.f_s + j * F;
      const MT* pmj = pm + ((size_t)b * Tin + j) * A;
      for (int a = lane; a < A; a += 32) {
        float s = sm.q_s[a] + ld_mem<MT>(pmj + a);
        for (int c = 0; c < F; ++c) s += sm.wdT[c * A + a] * fj[c];
        const float th = tanh_acc(s);
        if (p.train) p.th_all[(((size_t)t * p.B + b) * Tin + j) * A + a] = th;
        part += sm.v_s[a] * th;
      }
      part = warp_sum(part);
      if (lane == 0) sm.e_s[j] = part;
    }
    __syncthreads();
    // masked softmax over encoder positions
    float mx = -INFINITY;
    for (int j = threadIdx.x; j < Tin; j += blockDim.x) mx = fmaxf(mx, sm.e_s[j]);
    mx = block_reduce(mx, sm.scr, true);
    float sum = 0.f;
    for (int j = threadIdx.x; j < Tin; j += blockDim.x) {
      const float e = sm.e_s[j] == -INFINITY ? 0.f : expf(sm.e_s[j] - mx);
      sm.e_s[j] = e;
      sum += e;
    }
    sum = block_reduce(sum, sm.scr, false);
    const float inv = 1.0f / sum;
    float wsum = 0.f;
    float* aout = p.align_out + ((size_t)b * p.n_steps + t) * Tin;
    for (int j = threadIdx.x; j < Tin; j += blockDim.x) {
      const float w = sm.e_s[j] * inv;
      sm.e_s[j] = w;
      wsum += w;
      wprev[j] = w;
      wcum[j] += w;
      aout[j] = w;
    }
    wsum = block_reduce(wsum, sm.scr, false);   // also orders the e_s writes before the reads below
    // context over the time-varying (encoder) columns
    const int mult = p.mult;
    bf16* xc = BFP(p.xctx) + (size_t)b * mult * p.Dp;
    float* call = p.train ? p.ctx_all + ((size_t)t * p.B + b) * p.D : nullptr;
    for (int d = threadIdx.x; d < p.E; d += blockDim.x) {
      const MT* mp = mem + (size_t)b * Tin * p.E + d;
      float acc = 0.f;
      int j = 0;
      for (; j + 4 <= L; j += 4) {
        const float m0 = ld_mem<MT>(mp + (size_t)j * p.E), m1 = ld_mem<MT>(mp + (size_t)(j + 1) * p.E);
        const float m2 = ld_mem<MT>(mp + (size_t)(j + 2) * p.E), m3 = ld_mem<MT>(mp + (size_t)(j + 3) * p.E);
        acc += sm.e_s[j] * m0; acc += sm.e_s[j + 1] * m1; acc += sm.e_s[j + 2] * m2; acc += sm.e_s[j + 3] * m3;
      }
      for (; j < L; ++j) acc += sm.e_s[j] * ld_mem<MT>(mp + (size_t)j * p.E);
      store_act(xc, d, p.Dp, p.prec, acc);
      if (call) call[d] = acc;
    }
    // time-constant columns: context = const * sum_j w_j
    for (int i = threadIdx.x; i < p.Cc; i += blockDim.x) {
      const float val = p.mem_const[(size_t)b * p.Cc + i] * wsum;
      store_act(xc, p.E + i, p.Dp, p.prec, val);
      if (call) call[p.E + i] = val;
    }
    __syncthreads();
  }
}

template <int NT>
__global__ void __launch_bounds__(256, 1) decoder_fwd_kernel(const DecFwd p) {
  extern __shared__ float smem_f[];
  const FwdSmem sm = carve<NT>(smem_f, p);
  // persistent caches of the small attention weights (transposed for conflict-free access)
  for (int i = threadIdx.x; i < 2 * p.KS * p.F; i += blockDim.x) {
    const int ch = i / (p.KS * p.F), rem = i - ch * p.KS * p.F, k = rem / p.F, c = rem - k * p.F;
    sm.wcT[i] = p.wconv[((size_t)c * 2 + ch) * p.KS + k];
  }
  for (int i = threadIdx.x; i < p.F * p.A; i += blockDim.x) {
    const int c = i / p.A, a = i - c * p.A;
    sm.wdT[i] = p.wdense[(size_t)a * p.F + c];
  }
  for (int i = threadIdx.x; i < p.A; i += blockDim.x) sm.v_s[i] = p.v[i];
  __syncthreads();

  unsigned int epoch = 0;
  int t = 0;
  for (; t < p.n_steps; ++t) {
    if (!p.train) {
      small_linear_phase<NT>(p, sm, t, 0);
      grid_sync(p.barrier, epoch, gridDim.x);
      small_linear_phase<NT>(p, sm, t, 1);
      grid_sync(p.barrier, epoch, gridDim.x);
    }
    lstm_phase<NT>(p, sm, t, true);
    grid_sync(p.barrier, epoch, gridDim.x);
    if (p.mem_bf16) attention_phase<bf16>(p, sm, t); else attention_phase<float>(p, sm, t);
    grid_sync(p.barrier, epoch, gridDim.x);
    lstm_phase<NT>(p, sm, t, false);
    grid_sync(p.barrier, epoch, gridDim.x);
    if (!p.train) {
      small_linear_phase<NT>(p, sm, t, 2);
      grid_sync(p.barrier, epoch, gridDim.x);
      if (__ldcg(p.stop) != 0) { ++t; break; }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && p.steps_done) *p.steps_done = t;
}

template <int NT>
static int launch_fwd(const DecFwd& p, int grid, cudaStream_t st) {
  const size_t smem = fwd_smem_floats<NT>(p.Tin, p.A, p.F, p.KS) * sizeof(float);
  GTN_REQUIRE(smem <= 227 * 1024, "decoder forward needs %zu B of shared memory (T_in too large)", smem);
  GTN_CUDA_OK(cudaFuncSetAttribute(decoder_fwd_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  GTN_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, decoder_fwd_kernel<NT>, 256, smem));
  GTN_REQUIRE(per_sm >= 1, "decoder forward kernel does not fit on an SM");
  void* args[] = {const_cast<DecFwd*>(&p)};
  GTN_CUDA_OK(cudaLaunchCooperativeKernel((const void*)decoder_fwd_kernel<NT>, dim3(grid), dim3(256), args, smem, st));
  count_launch();
  return 0;
}

}  // namespace gtn

using namespace gtn;

extern "C" int gtn_decoder_forward(const gtn_decoder_fwd_params* hp, void* stream) {
  GTN_REQUIRE(hp != nullptr, "null params");
  DecFwd p = *hp;
  GTN_REQUIRE(p.prec == 1 || p.prec == 3, "prec must be 1 (bf16) or 3 (fp32 via 3 bf16 planes)");
  p.mult = p.prec == 1 ? 1 : 6;
  GTN_REQUIRE(p.Ha % 8 == 0 && p.Hd % 8 == 0, "LSTM widths must be multiples of 8");
  GTN_REQUIRE(p.Hap % 32 == 0 && p.Hdp % 32 == 0 && p.Dp % 32 == 0 && p.Pp % 32 == 0 && p.NMp % 32 == 0,
              "padded K segments must be multiples of 32");
  GTN_REQUIRE(p.Hap >= p.Ha && p.Hdp >= p.Hd && p.Dp >= p.D && p.Pp >= p.P && p.NMp >= p.NM, "padded dims too small");
  GTN_REQUIRE(p.D == p.E + p.Cc, "D must equal E + Cc");
  GTN_REQUIRE(p.KS % 2 == 1, "location kernel size must be odd");
  GTN_REQUIRE(p.B >= 1 && p.Tin >= 1 && p.n_steps >= 1, "empty problem");
  auto thr = [](float pr) -> unsigned { double t = (double)pr * 4294967296.0; return t >= 4294967295.0 ? 0xFFFFFFFFu : (unsigned)t; };
  p.thr_att = p.use_dropout ? thr(p.p_att) : 0u;
  p.thr_dec = p.use_dropout ? thr(p.p_dec) : 0u;
  int dev = 0, sms = 0, coop = 0;
  GTN_CUDA_OK(cudaGetDevice(&dev));
  GTN_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  GTN_CUDA_OK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
  GTN_REQUIRE(coop, "device does not support cooperative launch");
  const int grid = sms;
  cudaStream_t st = (cudaStream_t)stream;
  if (p.B <= 8) return launch_fwd<1>(p, grid, st);
  if (p.B <= 16) return launch_fwd<2>(p, grid, st);
  if (p.B <= 32) return launch_fwd<4>(p, grid, st);
  return launch_fwd<8>(p, grid, st);
}
