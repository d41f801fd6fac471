// Persistent decode kernel for autoregressive inference at BATCH 1 (Decoder.inference, reference model.py:456-497; one
// decode() = model.py:378-417 per frame).  At batch 1 a frame is ~19 M MAC of mat-vecs against 36 MB of bf16 LSTM weights:
// the only way under the HBM / L2 streaming time is to keep every weight ON CHIP for the whole utterance, and the only way
// to make that fit is to stop materialising the attention context:
//
//   * context(t) = sum_j w_j(t) memory[j] enters the frame only through three LINEAR maps (attention LSTM input, decoder LSTM
//     input, mel / gate projection).  Their products with the encoder memory are pre-multiplied once per utterance
//     (host: M_a = W_ih_att[:, ctx] memory^T, M_d, P_c; [rows][T_in]), so that per frame the context costs K = T_in instead of
//     K = 512 and the 154 KB memory tile is never read again.  The time-constant columns (noise / speaker / emotion) are
//     per-sequence biases as in the other kernels.
//   * Each of the 148 CTAs owns 7 hidden units of BOTH LSTM cells = 28 gate rows x (K_att 1472 + K_dec 2240) bf16 = 203 KB.
//     They are parked in TENSOR MEMORY (256 KB per SM), used as plain storage: lane = gate row, the four lane quarters hold
//     the four quarters of K, read back with tcgen05.ld 64 weights per thread and instruction.  Shared memory stays free for
//     the prenet (W2, 128 KB), the location filter (32 KB) and the fp32 activation vectors; W1 lives in registers.
//   * Three grid barriers per frame (h_att + query; energies; h_dec + mel / gate).  Reductions over CTAs (query, mel, gate)
//     are 64-bit FIXED-POINT atomics: exact, hence independent of the arrival order (bit-reproducible runs).
//   * One attention position per CTA (T_in <= 2 x CTAs): location filter, tanh, energy; softmax redundantly in every CTA.
//   * The prenet (with its always-on Philox dropout, model.py:104) runs redundantly in every CTA; so does the stop test.
//
// Arithmetic: bf16 weights (as the bf16 mode of the batched kernels), fp32 activations and accumulation, fp32 cell state.
#include <cstdlib>
#include "gtn_decoder2.cuh"
#include "gtn_internal.h"
#include "../../include/gantron_b200.h"

namespace gtn {

typedef gtn_decoder_infer1_params Inf1;

#define PROFI(slot)                                                                         \
  do {                                                                                      \
    if (p.prof != nullptr && blockIdx.x == 0 && threadIdx.x == 0 && t < 16) p.prof[t * 16 + (slot)] = clock64(); \
  } while (0)
static constexpr int I1_THREADS = 512;
static constexpr int I1_WARPS = I1_THREADS / 32;
static constexpr double I1_FIX = 4294967296.0;          // 2^32: fixed-point scale of the cross-CTA sums

struct SmemI1 {
  uint8_t* w2;        // bf16, chunked [P/8][P][8]: row-contiguous 16-byte chunks (conflict-free with lanes over rows)
  float* u;           // [64][128] location filter, transposed: u[ch * KS + k][a] (lanes over a: conflict-free), fp32
  float* xa;          // [Ka] = [prenet P | w(t-1) Tp64 | h_att(t-1) Ha | 0]
  float* xd;          // [Kd] = [h_att(t) Ha | w(t) Tp64 | h_dec(t-1) Hd | 0]
  float* part;        // [16][32] per-warp partial gate sums
  float* gates;       // [32]
  float* mel;         // [NM + 1 padded to 96]
  float* h1;          // [P]
  float* wprev;       // [Tp64 + 2 * 32] previous weights, zero halo of 32 on both sides
  float* wcum;        // [Tp64 + 2 * 32]
  float* e;           // [Tp64]
  float* q;           // [128]
  float* pm;          // [2][128] processed_memory rows of this CTA's positions
  float* v;           // [128]
  float* wq;          // [128][8]  Wq[a][own units]
  float* wp;          // [96][8]   W_proj[row][own h_dec units]
  float* pc;          // [Tp64]    P_c[row = cta][:] (cta <= NM)
  float* red;         // [32]
  uint32_t* tmem_slot;
};

GTN_DEVICE float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
GTN_DEVICE float bf_hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }

GTN_DEVICE void red_add_s64(long long* addr, long long v) {
  asm volatile("red.global.add.u64 [%0], %1;" ::"l"(addr), "l"(v) : "memory");
}

// grid barrier (co-resident cooperative launch), thread 0 carries release / acquire
GTN_DEVICE void grid_sync_i1(unsigned int* counter, unsigned int& epoch, unsigned int nblocks) {
  __syncthreads();
  epoch += nblocks;
  if (threadIdx.x == 0) {
    red_release_add(counter, 1u);
    long long t0 = clock64();
    while (ld_acquire_u32(counter) < epoch) {
      if (clock64() - t0 > 20000000000LL) __trap();
    }
  }
  __syncthreads();
}

// y[rho] partials of  W[rho][:] . x  for the 32 gate rows of this CTA: the weights sit in tensor memory (lane = row, lane
// quarter q = K quarter, `nch` chunks of 32 columns = 64 bf16 per quarter starting at column col0); warp w reads quarter w & 3,
// chunks w >> 2, w >> 2 + 4, ...; x is the matching fp32 vector in shared memory (broadcast reads).
GTN_DEVICE void matvec_tmem(uint32_t tmem_base, int col0, int nch, const float* x, float* part) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3, sub = warp >> 2;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  for (int c = sub; c < nch; c += I1_WARPS / 4) {
    uint32_t r[32];
    tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(col0 + 32 * c), r);
    tmem_ld_wait();
    const float4* xv = reinterpret_cast<const float4*>(x + ((size_t)q * nch + c) * 64);
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      const float4 x0 = xv[i], x1 = xv[i + 1];
      a0 += bf_lo(r[2 * i]) * x0.x + bf_lo(r[2 * i + 1]) * x0.z;
      a1 += bf_hi(r[2 * i]) * x0.y + bf_hi(r[2 * i + 1]) * x0.w;
      a2 += bf_lo(r[2 * i + 2]) * x1.x + bf_lo(r[2 * i + 3]) * x1.z;
      a3 += bf_hi(r[2 * i + 2]) * x1.y + bf_hi(r[2 * i + 3]) * x1.w;
    }
  }
  part[warp * 32 + lane] = (a0 + a1) + (a2 + a3);
}

GTN_DEVICE float block_sum_i1(float v, float* red) {
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float s = 0.f;
#pragma unroll
  for (int w = 0; w < I1_WARPS; ++w) s += red[w];
  return s;
}
GTN_DEVICE float block_max_i1(float v, float* red) {
  v = warp_max(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float s = red[0];
#pragma unroll
  for (int w = 1; w < I1_WARPS; ++w) s = fmaxf(s, red[w]);
  return s;
}

__global__ void __launch_bounds__(I1_THREADS, 1) decoder_infer1_kernel(const Inf1 p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int cta = blockIdx.x, ncta = gridDim.x;
  const int P = p.P, Ha = p.Ha, Hd = p.Hd, Tin = p.Tin, Tp = p.Tp, A = p.A, KS = p.KS, hk = (p.KS - 1) / 2, NM = p.NM;
  const int Ka = p.cols_a * 8, Kd = p.cols_d * 8;               // 4 quarters x cols x 2 bf16
  const int nch_a = p.cols_a / 32, nch_d = p.cols_d / 32;
  const int nu = p.nu, u0 = cta * nu;
  const int nua = max(0, min(nu, Ha - u0)), nud = max(0, min(nu, Hd - u0));

  SmemI1 sm;
  {
    uint8_t* q = base;
    sm.w2 = q; q += (size_t)P * P * 2;
    float* f = reinterpret_cast<float*>(q);
    sm.u = f; f += 128 * 64;
    sm.xa = f; f += Ka;
    sm.xd = f; f += Kd;
    sm.part = f; f += I1_WARPS * 32;
    sm.gates = f; f += 32;
    sm.mel = f; f += 96;
    sm.h1 = f; f += P;
    sm.wprev = f; f += Tp + 64;
    sm.wcum = f; f += Tp + 64;
    sm.e = f; f += Tp;
    sm.q = f; f += 128;
    sm.pm = f; f += 2 * 128;
    sm.v = f; f += 128;
    sm.wq = f; f += 128 * 8;
    sm.wp = f; f += 96 * 8;
    sm.pc = f; f += Tp;
    sm.red = f; f += 32;
    sm.tmem_slot = reinterpret_cast<uint32_t*>(f);
  }
  if (warp == 0) tmem_alloc(sm.tmem_slot, 512);
  // ---- one-time loads --------------------------------------------------------------------------------------------------
  {
    // W2 (prenet layer 2), bf16 [P][P] -> 16-byte chunks [k / 8][row]
    const uint4* src = reinterpret_cast<const uint4*>(p.w2);
    uint4* dst = reinterpret_cast<uint4*>(sm.w2);
    const int nck = P / 8;
    for (int i = tid; i < P * nck; i += I1_THREADS) {
      const int r = i / nck, kc = i - r * nck;
      dst[(size_t)kc * P + r] = __ldg(src + i);
    }
    for (int i = tid; i < 128 * 64; i += I1_THREADS) sm.u[(i & 63) * 128 + (i >> 6)] = p.u[i];      // [a][kk] -> [kk][a]
    for (int i = tid; i < Ka; i += I1_THREADS) sm.xa[i] = 0.f;
    for (int i = tid; i < Kd; i += I1_THREADS) sm.xd[i] = 0.f;
    for (int i = tid; i < Tp + 64; i += I1_THREADS) { sm.wprev[i] = 0.f; sm.wcum[i] = 0.f; }
    for (int i = tid; i < 96; i += I1_THREADS) sm.mel[i] = 0.f;               // go frame (model.py:281-294)
    for (int i = tid; i < 128; i += I1_THREADS) sm.v[i] = i < A ? p.v[i] : 0.f;
    for (int i = tid; i < 128 * 8; i += I1_THREADS) {
      const int a = i >> 3, u = i & 7;
      sm.wq[i] = (a < A && u < nua) ? p.wq[(size_t)a * Ha + u0 + u] : 0.f;
    }
    for (int i = tid; i < 96 * 8; i += I1_THREADS) {
      const int r = i >> 3, u = i & 7;
      sm.wp[i] = (r <= NM && u < nud) ? p.wp_h[(size_t)r * Hd + u0 + u] : 0.f;
    }
    for (int i = tid; i < Tp; i += I1_THREADS) sm.pc[i] = (cta <= NM && i < Tin) ? p.pc[(size_t)cta * Tin + i] : 0.f;
    for (int i = tid; i < 2 * 128; i += I1_THREADS) {
      const int s = i >> 7, a = i & 127, j = cta + s * ncta;
      sm.pm[i] = (j < Tin && a < A) ? p.pm[(size_t)j * A + a] : 0.f;
    }
  }
  // W1 (prenet layer 1) in registers: thread (row = tid % P, half = tid / P) holds half of the row's NMp bf16 (zero padded)
  uint32_t w1r[24];
  {
    const int r = tid % P, half = tid / P;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(p.w1) + ((size_t)r * p.NMp + half * (p.NMp / 2)) / 2;
#pragma unroll
    for (int i = 0; i < 24; ++i) w1r[i] = (tid < 2 * P && i < p.NMp / 4) ? __ldg(src + i) : 0u;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *sm.tmem_slot;
  if (warp < 4) {
    // this CTA's gate rows -> tensor memory: image [4 quarters][32 lanes][cols_a + cols_d] 32-bit words
    const int ncols = p.cols_a + p.cols_d;
    const uint32_t* img = reinterpret_cast<const uint32_t*>(p.wimg) + (((size_t)cta * 4 + warp) * 32 + lane) * ncols;
    for (int c = 0; c < ncols; c += 32) {
      uint32_t r[32];
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const uint4 v4 = __ldg(reinterpret_cast<const uint4*>(img + c + i));
        r[i] = v4.x; r[i + 1] = v4.y; r[i + 2] = v4.z; r[i + 3] = v4.w;
      }
      tmem_st_32x32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c, r);
    }
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  unsigned int epoch = 0;
  const unsigned int nblocks = gridDim.x;
  float c_att = 0.f, c_dec = 0.f;                        // cell states of unit tid (tid < nu)
  const int Tw = Tp;                                     // width of the w segment in xa / xd
  float* xa_w = sm.xa + P;                               // w(t-1)
  float* xa_h = sm.xa + P + Tw;                          // h_att(t-1)
  float* xd_ha = sm.xd;                                  // h_att(t)
  float* xd_w = sm.xd + Ha;                              // w(t)
  float* xd_hd = sm.xd + Ha + Tw;                        // h_dec(t-1)
  int t = 0;
  bool stopped = false;
  for (; t < p.n_steps && !stopped; ++t) {
    const int cur = t & 1;
    PROFI(0);
    // ---- (1) prenet of the previous mel frame (model.py:102-105): both layers in every CTA --------------------------------
    {
      const int r = tid % P, half = tid / P;
      float s = 0.f;
      if (tid < 2 * P) {
        const float* m = sm.mel + half * (p.NMp / 2);
#pragma unroll
        for (int i = 0; i < 24; ++i) s += bf_lo(w1r[i]) * m[2 * i] + bf_hi(w1r[i]) * m[2 * i + 1];
      }
      sm.part[tid] = s;                                   // (part has 512 floats)
      __syncthreads();
      if (tid < P) {
        float hv = fmaxf(sm.part[tid] + sm.part[tid + P], 0.f);
        const uint32_t bits = philox_bits(p.seed, 1u, (uint32_t)t, (uint32_t)tid);
        sm.h1[tid] = bits >= 0x80000000u ? hv * 2.0f : 0.f;            // p = 0.5, always on (model.py:104)
      }
      __syncthreads();
      // layer 2: thread (row r, K half): 16-byte chunks [k / 8][row], lanes over rows
      float s2 = 0.f;
      if (tid < 2 * P) {
        const int nck = P / 8, kc0 = half * (nck / 2);
        const uint4* wch = reinterpret_cast<const uint4*>(sm.w2);
        float b0 = 0.f, b1 = 0.f;
        for (int kc = kc0; kc < kc0 + nck / 2; ++kc) {
          const uint4 w = wch[(size_t)kc * P + r];
          const float4 x0 = *reinterpret_cast<const float4*>(sm.h1 + kc * 8), x1 = *reinterpret_cast<const float4*>(sm.h1 + kc * 8 + 4);
          b0 += bf_lo(w.x) * x0.x + bf_hi(w.x) * x0.y + bf_lo(w.y) * x0.z + bf_hi(w.y) * x0.w;
          b1 += bf_lo(w.z) * x1.x + bf_hi(w.z) * x1.y + bf_lo(w.w) * x1.z + bf_hi(w.w) * x1.w;
        }
        s2 = b0 + b1;
      }
      __syncthreads();
      sm.part[tid] = s2;
      __syncthreads();
      if (tid < P) {
        float pv = fmaxf(sm.part[tid] + sm.part[tid + P], 0.f);
        const uint32_t bits = philox_bits(p.seed, 2u, (uint32_t)t, (uint32_t)tid);
        sm.xa[tid] = bits >= 0x80000000u ? pv * 2.0f : 0.f;
      }
      __syncthreads();
    }
    PROFI(1);
    // ---- (2) attention LSTM (model.py:391-394) ---------------------------------------------------------------------------
    matvec_tmem(tmem, 0, nch_a, sm.xa, sm.part);
    __syncthreads();
    PROFI(2);
    if (tid < 32) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < I1_WARPS; ++w) s += sm.part[w * 32 + tid];
      sm.gates[tid] = s;
    }
    __syncthreads();
    if (tid < nua) {
      const int unit = u0 + tid;
      // frame 0 sees the zero initial context (model.py:314-316): bias only, no constant-column term
      const float4 gb = t == 0 ? make_float4(p.b_att[unit], p.b_att[Ha + unit], p.b_att[2 * Ha + unit], p.b_att[3 * Ha + unit])
                               : reinterpret_cast<const float4*>(p.gbias_a)[unit];
      const float i_ = sigmoid_acc(sm.gates[tid] + gb.x), f_ = sigmoid_acc(sm.gates[nu + tid] + gb.y);
      const float g_ = tanh_acc(sm.gates[2 * nu + tid] + gb.z), o_ = sigmoid_acc(sm.gates[3 * nu + tid] + gb.w);
      c_att = f_ * c_att + i_ * g_;
      const float h = o_ * tanh_acc(c_att);
      p.h_att_g[(size_t)cur * Ha + unit] = h;
      sm.red[tid] = h;
    }
    __syncthreads();
    if (tid < A) {
      // partial attention query of the own units, exact fixed-point sum over the CTAs
      float s = 0.f;
      for (int u = 0; u < nua; ++u) s += sm.wq[tid * 8 + u] * sm.red[u];
      red_add_s64(p.qacc + (size_t)cur * 128 + tid, (long long)llrint((double)s * I1_FIX));
    }
    PROFI(3);
    grid_sync_i1(p.barrier, epoch, nblocks);
    PROFI(4);
    if (cta == 0 && tid < 128 + 96) {
      // zero the accumulators of the NEXT frame: every CTA is past this frame's barrier 1, hence done reading the previous
      // frame's sums; the next adds come after two more barriers
      if (tid < 128) p.qacc[(size_t)(cur ^ 1) * 128 + tid] = 0;
      else p.macc[(size_t)(cur ^ 1) * 96 + tid - 128] = 0;
    }
    // ---- (4) h_att(t) and the query ----------------------------------------------------------------------------------------
    for (int i = tid; i < Ha; i += I1_THREADS) {
      const float h = __ldcg(&p.h_att_g[(size_t)cur * Ha + i]);
      xd_ha[i] = h;
      xa_h[i] = h;                                          // recurrent input of the next frame
    }
    if (tid < 128) sm.q[tid] = (float)((double)__ldcg(&p.qacc[(size_t)cur * 128 + tid]) * (1.0 / I1_FIX));
    __syncthreads();
    PROFI(5);
    // ---- (5) energies of this CTA's encoder positions (model.py:62-68): e_j = v . tanh(q + loc_j + pm_j) ---------------------
    for (int s = 0; s < 2; ++s) {
      const int j = cta + s * ncta;
      if (j >= Tin) break;                                   // (uniform over the CTA)
      float val = 0.f;
      if (tid < 2 * 128) {
        const int a = tid & 127, ch = tid >> 7;
        const float* ucol = sm.u + (size_t)ch * KS * 128 + a;
        const float* win = (ch == 0 ? sm.wprev : sm.wcum) + 32 + j - hk;       // zero halo of 32 on both sides
        float l0 = 0.f, l1 = 0.f;
        int k = 0;
        for (; k + 1 < KS; k += 2) { l0 += ucol[k * 128] * win[k]; l1 += ucol[(k + 1) * 128] * win[k + 1]; }
        if (k < KS) l0 += ucol[k * 128] * win[k];
        val = l0 + l1;
      }
      sm.part[tid] = val;
      __syncthreads();
      float term = 0.f;
      if (tid < A) term = sm.v[tid] * tanh_acc(sm.q[tid] + sm.part[tid] + sm.part[tid + 128] + sm.pm[s * 128 + tid]);
      const float ej = block_sum_i1(term, sm.red);
      if (tid == 0) p.e_g[(size_t)cur * Tp + j] = ej;
      __syncthreads();
    }
    PROFI(6);
    grid_sync_i1(p.barrier, epoch, nblocks);
    PROFI(7);
    // ---- (7) softmax over the encoder positions (model.py:84-89), redundantly in every CTA ------------------------------------
    {
      float ev = -INFINITY;
      if (tid < Tin) ev = __ldcg(&p.e_g[(size_t)cur * Tp + tid]);
      if (Tin > I1_THREADS) __trap();
      const float mx = block_max_i1(ev, sm.red);
      const float ex = tid < Tin ? expf(ev - mx) : 0.f;
      const float sum = block_sum_i1(ex, sm.red);
      const float w = ex / sum;
      if (tid < Tp) {
        const float wv = tid < Tin ? w : 0.f;
        sm.wprev[32 + tid] = wv;
        sm.wcum[32 + tid] += wv;                            // model.py:403
        xd_w[tid] = wv;
        xa_w[tid] = wv;                                      // context term of the next frame's attention LSTM
        if (cta == 0 && tid < Tin) p.align_out[(size_t)t * Tin + tid] = wv;
      }
      __syncthreads();
    }
    PROFI(8);
    // ---- (8) decoder LSTM (model.py:404-409) and the mel / gate projection partials (model.py:411-416) -------------------------
    matvec_tmem(tmem, p.cols_a, nch_d, sm.xd, sm.part);
    __syncthreads();
    PROFI(9);
    if (tid < 32) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < I1_WARPS; ++w) s += sm.part[w * 32 + tid];
      sm.gates[tid] = s;
    }
    __syncthreads();
    if (tid < nud) {
      const int unit = u0 + tid;
      const float4 gb = reinterpret_cast<const float4*>(p.gbias_d)[unit];
      const float i_ = sigmoid_acc(sm.gates[tid] + gb.x), f_ = sigmoid_acc(sm.gates[nu + tid] + gb.y);
      const float g_ = tanh_acc(sm.gates[2 * nu + tid] + gb.z), o_ = sigmoid_acc(sm.gates[3 * nu + tid] + gb.w);
      c_dec = f_ * c_dec + i_ * g_;
      const float h = o_ * tanh_acc(c_dec);
      p.h_dec_g[(size_t)cur * Hd + unit] = h;
      sm.red[tid] = h;
    }
    __syncthreads();
    {
      // projection rows 0..NM (NM = gate): own h_dec units; CTA r additionally adds row r's context term P_c[r] . w(t)
      float s = 0.f;
      if (tid <= NM) {
        for (int u = 0; u < nud; ++u) s += sm.wp[tid * 8 + u] * sm.red[u];
        red_add_s64(p.macc + (size_t)cur * 96 + tid, (long long)llrint((double)s * I1_FIX));
      }
      float c = 0.f;
      if (cta <= NM && tid < Tp) c = sm.pc[tid] * xd_w[tid];
      const float cs = block_sum_i1(c, sm.red);
      if (cta <= NM && tid == 0) red_add_s64(p.macc + (size_t)cur * 96 + cta, (long long)llrint((double)cs * I1_FIX));
    }
    PROFI(10);
    grid_sync_i1(p.barrier, epoch, nblocks);
    PROFI(11);
    // ---- (0') h_dec(t), mel(t), gate(t), stop test (model.py:486) ---------------------------------------------------------------
    for (int i = tid; i < Hd; i += I1_THREADS) xd_hd[i] = __ldcg(&p.h_dec_g[(size_t)cur * Hd + i]);
    if (tid < 96) {
      float mv = 0.f;
      if (tid <= NM) mv = (float)((double)__ldcg(&p.macc[(size_t)cur * 96 + tid]) * (1.0 / I1_FIX)) + p.pbias[tid];
      if (tid < NM) {
        sm.mel[tid] = mv;
        if (cta == 0) p.mel_out[(size_t)tid * p.n_steps + t] = mv;
      } else if (tid == NM) {
        sm.mel[NM] = 0.f;
        sm.gates[0] = mv;
        if (cta == 0) p.gate_out[t] = mv;
      } else {
        sm.mel[tid] = 0.f;
      }
    }
    __syncthreads();
    if (!p.ignore_stop && sigmoid_acc(sm.gates[0]) > p.gate_threshold) stopped = true;      // identical in every CTA
    __syncthreads();
  }
  if (cta == 0 && tid == 0) {
    *p.steps_done = t;
    p.done[0] = stopped ? 1 : 0;
    p.lengths_out[0] = t;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

static size_t infer1_smem_bytes(const Inf1& p) {
  const int Ka = p.cols_a * 8, Kd = p.cols_d * 8;
  return (size_t)p.P * p.P * 2 +
         (size_t)(128 * 64 + Ka + Kd + I1_WARPS * 32 + 32 + 96 + p.P + 2 * (p.Tp + 64) + p.Tp + 128 + 256 + 128 + 128 * 8 + 96 * 8 + p.Tp + 32) * 4 +
         64 + 128;
}

}  // namespace gtn

using namespace gtn;

extern "C" int gtn_decoder_infer1_ctas(void) {
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  return sms;
}

extern "C" int gtn_decoder_infer1(const gtn_decoder_infer1_params* hp, void* stream) {
  GTN_REQUIRE(hp != nullptr, "null params");
  Inf1 p = *hp;
  const int ncta = gtn_decoder_infer1_ctas();
  GTN_REQUIRE(ncta > 0 && p.ncta == ncta, "params.ncta (%d) must equal gtn_decoder_infer1_ctas() (%d)", p.ncta, ncta);
  GTN_REQUIRE(p.nu >= 1 && p.nu <= 8 && p.nu * ncta >= p.Ha && p.nu * ncta >= p.Hd, "units per CTA (%d) must cover the LSTM widths", p.nu);
  GTN_REQUIRE(p.cols_a % 32 == 0 && p.cols_d % 32 == 0 && p.cols_a + p.cols_d <= 512, "weight image does not fit tensor memory");
  GTN_REQUIRE(p.cols_a * 8 >= p.P + p.Tp + p.Ha && p.cols_d * 8 >= p.Ha + p.Tp + p.Hd, "K layout too small");
  GTN_REQUIRE(p.Tp % 64 == 0 && p.Tp >= p.Tin && p.Tin >= 1 && p.Tin <= 2 * ncta && p.Tin <= I1_THREADS, "T_in out of range");
  GTN_REQUIRE(p.P % 64 == 0 && 2 * p.P <= I1_THREADS && p.NMp % 8 == 0 && p.NMp <= 96 && p.NM + 1 <= 96 && p.NM <= p.NMp,
              "prenet / mel dims out of range");
  GTN_REQUIRE(p.A <= 128 && 2 * p.KS <= 64 && p.KS % 2 == 1 && (p.KS - 1) / 2 <= 32, "attention dims out of range");
  GTN_REQUIRE(p.Ha % 64 == 0 && p.Hd % 64 == 0 && p.NM + 1 <= ncta, "LSTM widths must be multiples of 64");
  const size_t smem = infer1_smem_bytes(p);
  GTN_REQUIRE(smem <= 227 * 1024, "decode kernel: %zu bytes of shared memory needed", smem);
  GTN_CUDA_OK(cudaFuncSetAttribute(decoder_infer1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ncta);
  cfg.blockDim = dim3(I1_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeCooperative;
  attrs[0].val.cooperative = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = profiler_attached() ? 0 : 1;
  GTN_CUDA_OK(cudaLaunchKernelEx(&cfg, decoder_infer1_kernel, p));
  count_launch();
  return 0;
}

// ------------------------------------------------------------------------------------------------------------------------
// gtn_decoder_infer1_static: the same decode for callers that cannot allocate device memory (CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The kernel's cross-CTA workspace (hidden-state exchange, energies, fixed-point sums, barrier counter) lives in
// ONE file-scope device array of a stated maximum size; every GPU that loads the library holds it (40,704 bytes) until exit.
// Every call on a device shares the array: it is cleared at the start of each call (the kernel's barrier epochs and sums
// start at zero), and two calls that overlap on the GPU (different streams) RACE -- serialise them.
// ------------------------------------------------------------------------------------------------------------------------
namespace gtn {

static constexpr int I1S_HMAX = GTN_INFER1_STATIC_MAX_H;        // max attention_rnn_dim / decoder_rnn_dim
static constexpr int I1S_TPMAX = GTN_INFER1_STATIC_MAX_TP;      // max T_in rounded up to 64
struct Inf1Scratch {                // (the kernel indexes [2][Ha] / [2][Hd] / [2][Tp] densely from each member's start)
  float h_att[2][I1S_HMAX];
  float h_dec[2][I1S_HMAX];
  float e[2][I1S_TPMAX];
  long long qacc[2][128];
  long long macc[2][96];
  unsigned int barrier[64];
};
static __device__ __align__(256) unsigned char g_infer1_scratch[sizeof(Inf1Scratch)];

// The array's NAME is a host-side stand-in in host code, not a device address: its address on the CURRENT device comes from
// cudaGetSymbolAddress, once per device.
static int infer1_scratch_ptr(Inf1Scratch** out) {
  static void* ptrs[64] = {};
  int dev = 0;
  GTN_CUDA_OK(cudaGetDevice(&dev));
  GTN_REQUIRE(dev >= 0 && dev < 64, "device ordinal %d out of range", dev);
  if (ptrs[dev] == nullptr) {
    void* ptr = nullptr;
    GTN_CUDA_OK(cudaGetSymbolAddress(&ptr, g_infer1_scratch));
    GTN_REQUIRE(ptr != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    ptrs[dev] = ptr;
  }
  *out = reinterpret_cast<Inf1Scratch*>(ptrs[dev]);
  return 0;
}

}  // namespace gtn

extern "C" int gtn_decoder_infer1_static(const gtn_decoder_infer1_params* hp, void* stream) {
  using namespace gtn;
  GTN_REQUIRE(hp != nullptr, "null params");
  GTN_REQUIRE(hp->Ha >= 1 && hp->Ha <= I1S_HMAX && hp->Hd >= 1 && hp->Hd <= I1S_HMAX,
              "gtn_decoder_infer1_static: LSTM widths (%d, %d) exceed the static scratch maximum of %d units; use gtn_decoder_infer1",
              hp->Ha, hp->Hd, I1S_HMAX);
  GTN_REQUIRE(hp->Tp >= 1 && hp->Tp <= I1S_TPMAX,
              "gtn_decoder_infer1_static: padded encoder length %d exceeds the static scratch maximum of %d positions; use gtn_decoder_infer1",
              hp->Tp, I1S_TPMAX);
  Inf1Scratch* s = nullptr;
  int rc = infer1_scratch_ptr(&s);
  if (rc) return rc;
  // the array keeps the previous call's contents: exchange buffers, sums and the barrier counter must start at zero
  GTN_CUDA_OK(cudaMemsetAsync(s, 0, sizeof(Inf1Scratch), (cudaStream_t)stream));
  gtn_decoder_infer1_params q = *hp;
  q.h_att_g = &s->h_att[0][0];
  q.h_dec_g = &s->h_dec[0][0];
  q.e_g = &s->e[0][0];
  q.qacc = &s->qacc[0][0];
  q.macc = &s->macc[0][0];
  q.barrier = &s->barrier[0];
  return gtn_decoder_infer1(&q, stream);
}
