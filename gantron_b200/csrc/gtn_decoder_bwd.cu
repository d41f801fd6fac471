// Persistent cooperative decoder kernel (backward through time).  See include/gantron_b200.h for the
// contract and gtn_decoder.cuh for the GEMM engine.  Reverse of decode() (reference model.py:378-417)
// for t = T-1 .. 0, replacing autograd's replay of the per-frame tape built by model.py:444-449.
//
// Gradient flow of one frame (names as in the forward kernel):
//   d[h_dec'|ctx']  <- projection (hoisted GEMM, dproj_all)                       model.py:411-416
//   decoder LSTM cell backward (dropout mask regenerated from Philox)             model.py:406-409
//   dX_dec = dGates_dec . W_dec   -> d h_att', d ctx', d h_dec(t-1)               (phase 1)
//   attention backward: ctx = w.memory, softmax, e = v.tanh(q+loc+pm), location dense+conv,
//                       cum += w                                                   model.py:62-89,396-403 (phase 2)
//   query backward + attention LSTM cell backward                                  model.py:391-394 (phase 3)
//   dX_att = dGates_att . W_att   -> d prenet(t), d ctx(t-1), d h_att(t-1)        (phase 1 of frame t-1)
#include "gtn_decoder.cuh"
#include "gtn_internal.h"
#include "../../include/gantron_b200.h"

namespace gtn {

typedef gtn_decoder_bwd_params DecBwd;
#define BFP(x) (reinterpret_cast<bf16*>(const_cast<void*>(static_cast<const void*>(x))))

struct BwdSmem {
  float *red, *w_s, *dw_s, *dctx_s, *du_s, *df_s, *f_s, *wcat, *wc_s, *wd_s, *v_s, *part_s, *dv_acc, *dwd_acc, *dwc_acc, *scr;
};

template <int NT>
__host__ __device__ inline size_t bwd_smem_floats(int Tin, int D, int A, int F, int KS) {
  return (size_t)8 * 2 * NT * 128 + Tin + Tin + D + 32 * A + (size_t)Tin * (F + 1) + 32 * F + 2 * (Tin + KS - 1) +
         2 * F * KS + (size_t)A * F + A + 8 * A + A + (size_t)A * F + 2 * F * KS + 64;
}

template <int NT>
GTN_DEVICE BwdSmem carve_bwd(float* s, const DecBwd& p) {
  BwdSmem m;
  m.red = s; s += 8 * 2 * NT * 128;
  m.w_s = s; s += p.Tin;
  m.dw_s = s; s += p.Tin;
  m.dctx_s = s; s += p.D;
  m.du_s = s; s += 32 * p.A;
  m.df_s = s; s += p.Tin * (p.F + 1);
  m.f_s = s; s += 32 * p.F;
  m.wcat = s; s += 2 * (p.Tin + p.KS - 1);
  m.wc_s = s; s += 2 * p.F * p.KS;
  m.wd_s = s; s += p.A * p.F;
  m.v_s = s; s += p.A;
  m.part_s = s; s += 8 * p.A;
  m.dv_acc = s; s += p.A;
  m.dwd_acc = s; s += p.A * p.F;
  m.dwc_acc = s; s += 2 * p.F * p.KS;
  m.scr = s;
  return m;
}

GTN_DEVICE float block_sum(float v, float* scr) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) scr[warp] = v;
  __syncthreads();
  float r = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) r += scr[w];
  return r;
}

// LSTM cell backward for one (b, unit); returns nothing, writes the four pre-activation gate gradients.
GTN_DEVICE void cell_backward(const DecBwd& p, int t, int b, int unit, int H, float dh, const float* gates_all,
                              const float* c_all, float* dc_carry, float* dg_all, bf16* xdg, unsigned site,
                              unsigned thr, float pdrop) {
  const size_t ci = (size_t)b * H + unit;
  if (p.use_dropout && thr != 0u) {
    const uint32_t bits = philox_bits(p.seed, site, (uint32_t)t, (uint32_t)ci);
    dh = bits >= thr ? dh * (1.0f / (1.0f - pdrop)) : 0.f;
  }
  const size_t g0 = ((size_t)t * p.B + b) * 4 * H + unit;
  const float i_ = gates_all[g0], f_ = gates_all[g0 + H], g_ = gates_all[g0 + 2 * H], o_ = gates_all[g0 + 3 * H];
  const float c = c_all[((size_t)t * p.B + b) * H + unit];
  const float cprev = t > 0 ? c_all[((size_t)(t - 1) * p.B + b) * H + unit] : 0.f;
  const float tc = tanh_acc(c);
  const float d_o = dh * tc;
  const float dc = dh * o_ * (1.f - tc * tc) + dc_carry[ci];
  dc_carry[ci] = dc * f_;
  const float dgi = dc * g_ * i_ * (1.f - i_);
  const float dgf = dc * cprev * f_ * (1.f - f_);
  const float dgg = dc * i_ * (1.f - g_ * g_);
  const float dgo = d_o * o_ * (1.f - o_);
  dg_all[g0] = dgi; dg_all[g0 + H] = dgf; dg_all[g0 + 2 * H] = dgg; dg_all[g0 + 3 * H] = dgo;
  bf16* xr = xdg + (size_t)b * p.mult * 4 * H;
  store_act(xr, unit, 4 * H, p.prec, dgi);
  store_act(xr, H + unit, 4 * H, p.prec, dgf);
  store_act(xr, 2 * H + unit, 4 * H, p.prec, dgg);
  store_act(xr, 3 * H + unit, 4 * H, p.prec, dgo);
}

// decoder LSTM cell backward of frame t, spread over every thread of the grid
GTN_DEVICE void dec_cell_phase(const DecBwd& p, int t) {
  const int n = p.B * p.Hd;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += gridDim.x * blockDim.x) {
    const int b = idx / p.Hd, u = idx - b * p.Hd;
    const float dh = p.dproj_all[((size_t)t * p.B + b) * (p.Hd + p.D) + u] + __ldcg(&p.carry_hd[idx]);
    cell_backward(p, t, b, u, p.Hd, dh, p.gates_d, p.c_d_all, p.dc_d, p.dgd_all, BFP(p.xdg_d), 4u, p.thr_dec, p.p_dec);
  }
}

// data-gradient GEMMs: decoder LSTM of frame t_dec (if >= 0) and attention LSTM of frame t_att (if >= 0)
template <int NT>
GTN_DEVICE void dgrad_phase(const DecBwd& p, const BwdSmem& sm, int t_dec, int t_att) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int R_att = p.P + p.D + p.Ha, R_dec = p.Ha + p.D + p.Hd;
  const int n_att = t_att >= 0 ? (R_att + 31) / 32 : 0;
  const int n_dec = t_dec >= 0 ? (R_dec + 31) / 32 : 0;
  const int rr = lane >> 2, q8 = (lane & 3) * 8;
  for (int pass = blockIdx.x; pass < n_att + n_dec; pass += gridDim.x) {
    const bool att = pass < n_att;
    const int row0 = (att ? pass : pass - n_att) * 32;
    const int H = att ? p.Ha : p.Hd;
    const int Ktot = p.mult * 4 * H;
    const bf16* WT = att ? BFP(p.wT_att) : BFP(p.wT_dec);
    Seg seg;
    seg.x = att ? BFP(p.xdg_a) : BFP(p.xdg_d); seg.ld = Ktot; seg.nch = Ktot / 32;
    const int nch = Ktot / 32;
    const int c_begin = (int)((long long)warp * nch / 8), c_end = (int)((long long)(warp + 1) * nch / 8);
    const bf16* wrow[2][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int h = 0; h < 2; ++h) wrow[mt][h] = WT + (size_t)(row0 + 16 * mt + 8 * h + rr) * Ktot + q8;
    for (int b0 = 0; b0 < p.B; b0 += 8 * NT) {
      float acc[2][NT][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[mt][nt][c] = 0.f;
      warp_gemm<2, NT>(acc, wrow, &seg, 1, c_begin, c_end, p.B, b0, lane);
      __syncthreads();
      store_partials<2, NT>(sm.red, acc, warp, lane);
      __syncthreads();
      const int nb = min(8 * NT, p.B - b0);
      for (int idx = threadIdx.x; idx < nb * 32; idx += blockDim.x) {
        const int r = idx & 31, bl = idx >> 5, b = b0 + bl, row = row0 + r;
        const float val = reduce_partial<2, NT>(sm.red, r >> 4, r & 15, bl);
        if (att) {
          if (row < p.P) p.dpre_all[((size_t)t_att * p.B + b) * p.P + row] = val;
          else if (row < p.P + p.D) p.dctx_att[(size_t)b * p.D + row - p.P] = val;
          else if (row < R_att) p.carry_ha[(size_t)b * p.Ha + row - p.P - p.D] = val;
        } else {
          if (row < p.Ha) p.dha_dec[(size_t)b * p.Ha + row] = val;
          else if (row < p.Ha + p.D) p.dctx_dec[(size_t)b * p.D + row - p.Ha] = val;
          else if (row < R_dec) p.carry_hd[(size_t)b * p.Hd + row - p.Ha - p.D] = val;
        }
      }
    }
    __syncthreads();
  }
}

// query backward GEMM (d h_att' += Wq^T dq) fused with the attention LSTM cell backward of frame t
template <int NT>
GTN_DEVICE void query_cell_phase(const DecBwd& p, const BwdSmem& sm, int t) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Ktot = p.mult * p.Ap, nch = Ktot / 32;
  const int c_begin = (int)((long long)warp * nch / 8), c_end = (int)((long long)(warp + 1) * nch / 8);
  const int rr = lane >> 2, q8 = (lane & 3) * 8;
  Seg seg;
  seg.x = BFP(p.xdq); seg.ld = Ktot; seg.nch = nch;
  for (int pass = blockIdx.x; pass < (p.Ha + 31) / 32; pass += gridDim.x) {
    const int u0 = pass * 32;
    const bf16* wrow[2][2];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        int row = u0 + 16 * mt + 8 * h + rr;
        if (row >= p.Ha) row = p.Ha - 1;     // clamp (results of clamped rows are discarded)
        wrow[mt][h] = BFP(p.wqT) + (size_t)row * Ktot + q8;
      }
    for (int b0 = 0; b0 < p.B; b0 += 8 * NT) {
      float acc[2][NT][4];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
          for (int c = 0; c < 4; ++c) acc[mt][nt][c] = 0.f;
      warp_gemm<2, NT>(acc, wrow, &seg, 1, c_begin, c_end, p.B, b0, lane);
      __syncthreads();
      store_partials<2, NT>(sm.red, acc, warp, lane);
      __syncthreads();
      const int nb = min(8 * NT, p.B - b0);
      for (int idx = threadIdx.x; idx < nb * 32; idx += blockDim.x) {
        const int r = idx & 31, bl = idx >> 5, b = b0 + bl, unit = u0 + r;
        if (unit >= p.Ha) continue;
        const size_t ci = (size_t)b * p.Ha + unit;
        const float dh = reduce_partial<2, NT>(sm.red, r >> 4, r & 15, bl) + __ldcg(&p.dha_dec[ci]) + __ldcg(&p.carry_ha[ci]);
        cell_backward(p, t, b, unit, p.Ha, dh, p.gates_a, p.c_a_all, p.dc_a, p.dga_all, BFP(p.xdg_a), 3u, p.thr_att, p.p_att);
      }
    }
    __syncthreads();
  }
}

template <typename MT>
GTN_DEVICE float ld_mem_b(const MT* p);
template <>
GTN_DEVICE float ld_mem_b<float>(const float* p) { return __ldg(p); }
template <>
GTN_DEVICE float ld_mem_b<bf16>(const bf16* p) { return __bfloat162float(*p); }

// attention backward of frame t, one CTA per sequence
template <typename MT>
GTN_DEVICE void attn_bwd_phase(const DecBwd& p, const BwdSmem& sm, int t) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Tin = p.Tin, A = p.A, F = p.F, KS = p.KS, half = (p.KS - 1) / 2, Wd = p.Tin + p.KS - 1, D = p.D;
  const MT* mem = reinterpret_cast<const MT*>(p.mem_enc);
  const int NA = (A + 31) / 32;
  for (int b = blockIdx.x; b < p.B; b += gridDim.x) {
    const int L = p.lengths ? min(p.lengths[b], Tin) : Tin;
    const size_t tb = (size_t)t * p.B + b, bt = (size_t)b * p.T + t;
    for (int j = threadIdx.x; j < Tin; j += blockDim.x) sm.w_s[j] = p.align[bt * Tin + j];
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
      const float val = p.dproj_all[tb * (p.Hd + D) + p.Hd + d] + __ldcg(&p.dctx_dec[(size_t)b * D + d]) +
                        __ldcg(&p.dctx_att[(size_t)b * D + d]);
      sm.dctx_s[d] = val;
      p.dctx_all[bt * D + d] = val;
    }
    for (int i = threadIdx.x; i < 2 * Wd; i += blockDim.x) {
      const int ch = i / Wd, j = i - ch * Wd - half;
      float val = 0.f;
      if (j >= 0 && j < Tin) val = ch == 0 ? (t > 0 ? p.align[(bt - 1) * Tin + j] : 0.f) : p.cum_all[tb * Tin + j];
      sm.wcat[i] = val;
    }
    __syncthreads();
    // gradient reaching every w_j through the time-constant memory columns
    float part = 0.f;
    for (int i = threadIdx.x; i < p.Cc; i += blockDim.x) part += sm.dctx_s[p.E + i] * p.mem_const[(size_t)b * p.Cc + i];
    const float dcs = block_sum(part, sm.scr);
    // d w_j
    for (int j = warp; j < Tin; j += 8) {
      float s = 0.f;
      if (j < L) {
        const MT* mj = mem + ((size_t)b * Tin + j) * p.E;
        for (int d = lane; d < p.E; d += 32) s += sm.dctx_s[d] * ld_mem_b<MT>(mj + d);
        s = warp_sum(s);
        s += dcs + p.dwprev[(size_t)b * p.ldT + j] + p.dcum[(size_t)b * p.ldT + j];
        if (p.dalign) s += p.dalign[bt * Tin + j];
      }
      if (lane == 0) sm.dw_s[j] = s;
    }
    __syncthreads();
    // softmax backward: de_j = w_j (dw_j - sum_k w_k dw_k)
    float dot = 0.f;
    for (int j = threadIdx.x; j < Tin; j += blockDim.x) dot += sm.w_s[j] * sm.dw_s[j];
    dot = block_sum(dot, sm.scr);
    for (int j = threadIdx.x; j < Tin; j += blockDim.x) sm.dw_s[j] = sm.w_s[j] * (sm.dw_s[j] - dot);
    __syncthreads();
    float dq_r[4] = {0.f, 0.f, 0.f, 0.f}, dv_r[4] = {0.f, 0.f, 0.f, 0.f};   // A <= 128
    for (int jb = 0; jb < Tin; jb += 32) {
      const int nj = min(32, Tin - jb);
      for (int jj = warp; jj < nj; jj += 8) {
        const int j = jb + jj;
        const float de = sm.dw_s[j];
        const bool valid = j < L;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int a = lane + 32 * i;
          if (i < NA && a < A) {
            float du = 0.f;
            if (valid) {
              const float th = p.th_all[(tb * Tin + j) * A + a];
              du = de * sm.v_s[a] * (1.f - th * th);
              dv_r[i] += de * th;
              dq_r[i] += du;
              p.dpm[((size_t)b * Tin + j) * A + a] += du;
            }
            sm.du_s[jj * A + a] = du;
          }
        }
      }
      for (int idx = threadIdx.x; idx < nj * F; idx += blockDim.x) sm.f_s[idx] = p.f_all[(tb * Tin + jb) * F + idx];
      __syncthreads();
      // d f[j][c] = sum_a Wd[a][c] du[j][a]
      for (int idx = threadIdx.x; idx < nj * F; idx += blockDim.x) {
        const int jj = idx / F, c = idx - jj * F;
        float s = 0.f;
        const float* du = sm.du_s + jj * A;
        for (int a = 0; a < A; ++a) s += sm.wd_s[a * F + c] * du[a];
        sm.df_s[(jb + jj) * (F + 1) + c] = s;
      }
      // d Wd[a][c] += sum_j du[j][a] f[j][c]
      for (int idx = threadIdx.x; idx < A * F; idx += blockDim.x) {
        const int a = idx / F, c = idx - a * F;
        float s = 0.f;
        for (int jj = 0; jj < nj; ++jj) s += sm.du_s[jj * A + a] * sm.f_s[jj * F + c];
        sm.dwd_acc[idx] += s;
      }
      __syncthreads();
    }
    // combine the per-warp partials of dq and dv
#pragma unroll
    for (int i = 0; i < 4; ++i) { const int a = lane + 32 * i; if (i < NA && a < A) sm.part_s[warp * A + a] = dq_r[i]; }
    __syncthreads();
    for (int a = threadIdx.x; a < A; a += blockDim.x) {
      float s = 0.f;
      for (int w = 0; w < 8; ++w) s += sm.part_s[w * A + a];
      p.dq_all[tb * A + a] = s;
      store_act(BFP(p.xdq) + (size_t)b * p.mult * p.Ap, a, p.Ap, p.prec, s);
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) { const int a = lane + 32 * i; if (i < NA && a < A) sm.part_s[warp * A + a] = dv_r[i]; }
    __syncthreads();
    for (int a = threadIdx.x; a < A; a += blockDim.x) {
      float s = 0.f;
      for (int w = 0; w < 8; ++w) s += sm.part_s[w * A + a];
      sm.dv_acc[a] += s;
    }
    // location conv backward: gradients w.r.t. the previous weights (ch 0) and the cumulative weights (ch 1)
    for (int idx = threadIdx.x; idx < 2 * Tin; idx += blockDim.x) {
      const int ch = idx / Tin, pos = idx - ch * Tin;
      float s = 0.f;
      for (int k = 0; k < KS; ++k) {
        const int j = pos - k + half;
        if (j >= 0 && j < Tin) {
          const float* df = sm.df_s + j * (F + 1);
          const float* wc = sm.wc_s + ch * KS + k;
          for (int c = 0; c < F; ++c) s += wc[c * 2 * KS] * df[c];
        }
      }
      if (ch == 0) p.dwprev[(size_t)b * p.ldT + pos] = s;
      else p.dcum[(size_t)b * p.ldT + pos] += s;
    }
    // d Wc[c][ch][k] += sum_j df[j][c] wcat[ch][j + k]
    for (int idx = threadIdx.x; idx < F * 2 * KS; idx += blockDim.x) {
      const int c = idx % F, ck = idx / F, ch = ck / KS, k = ck - ch * KS;
      float s = 0.f;
      const float* xw = sm.wcat + ch * Wd + k;
      for (int j = 0; j < Tin; ++j) s += sm.df_s[j * (F + 1) + c] * xw[j];
      sm.dwc_acc[(c * 2 + ch) * KS + k] += s;
    }
    __syncthreads();
  }
}

template <int NT>
__global__ void __launch_bounds__(256, 1) decoder_bwd_kernel(const DecBwd p) {
  extern __shared__ float smem_f[];
  const BwdSmem sm = carve_bwd<NT>(smem_f, p);
  for (int i = threadIdx.x; i < 2 * p.F * p.KS; i += blockDim.x) { sm.wc_s[i] = p.wconv[i]; sm.dwc_acc[i] = 0.f; }
  for (int i = threadIdx.x; i < p.A * p.F; i += blockDim.x) { sm.wd_s[i] = p.wdense[i]; sm.dwd_acc[i] = 0.f; }
  for (int i = threadIdx.x; i < p.A; i += blockDim.x) { sm.v_s[i] = p.v[i]; sm.dv_acc[i] = 0.f; }
  __syncthreads();
  unsigned int epoch = 0;
  dec_cell_phase(p, p.T - 1);
  grid_sync(p.barrier, epoch, gridDim.x);
  for (int t = p.T - 1; t >= 0; --t) {
    dgrad_phase<NT>(p, sm, t, t + 1 < p.T ? t + 1 : -1);
    grid_sync(p.barrier, epoch, gridDim.x);
    if (p.mem_bf16) attn_bwd_phase<bf16>(p, sm, t); else attn_bwd_phase<float>(p, sm, t);
    grid_sync(p.barrier, epoch, gridDim.x);
    query_cell_phase<NT>(p, sm, t);
    if (t > 0) dec_cell_phase(p, t - 1);
    grid_sync(p.barrier, epoch, gridDim.x);
  }
  dgrad_phase<NT>(p, sm, -1, 0);
  // per-CTA partial sums of the small attention weight gradients
  for (int i = threadIdx.x; i < p.A; i += blockDim.x) p.dv_part[(size_t)blockIdx.x * p.A + i] = sm.dv_acc[i];
  for (int i = threadIdx.x; i < p.A * p.F; i += blockDim.x) p.dwd_part[(size_t)blockIdx.x * p.A * p.F + i] = sm.dwd_acc[i];
  for (int i = threadIdx.x; i < 2 * p.F * p.KS; i += blockDim.x) p.dwc_part[(size_t)blockIdx.x * 2 * p.F * p.KS + i] = sm.dwc_acc[i];
}

template <int NT>
static int launch_bwd(const DecBwd& p, int grid, cudaStream_t st) {
  const size_t smem = bwd_smem_floats<NT>(p.Tin, p.D, p.A, p.F, p.KS) * sizeof(float);
  GTN_REQUIRE(smem <= 227 * 1024, "decoder backward needs %zu B of shared memory (T_in too large)", smem);
  GTN_CUDA_OK(cudaFuncSetAttribute(decoder_bwd_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  void* args[] = {const_cast<DecBwd*>(&p)};
  GTN_CUDA_OK(cudaLaunchCooperativeKernel((const void*)decoder_bwd_kernel<NT>, dim3(grid), dim3(256), args, smem, st));
  count_launch();
  return 0;
}

}  // namespace gtn

using namespace gtn;

extern "C" int gtn_decoder_bwd_grid(void) {
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  return sms;
}

extern "C" int gtn_decoder_backward(const gtn_decoder_bwd_params* hp, void* stream) {
  GTN_REQUIRE(hp != nullptr, "null params");
  DecBwd p = *hp;
  GTN_REQUIRE(p.prec == 1 || p.prec == 3, "prec must be 1 or 3");
  p.mult = p.prec == 1 ? 1 : 6;
  GTN_REQUIRE(p.Ha % 8 == 0 && p.Hd % 8 == 0 && p.Ap % 32 == 0 && p.Ap >= p.A && p.A <= 128,
              "unsupported dims (Ha/Hd multiples of 8, attention_dim <= 128)");
  GTN_REQUIRE(p.D == p.E + p.Cc && p.T >= 1 && p.B >= 1, "bad dims");
  auto thr = [](float pr) -> unsigned { double t = (double)pr * 4294967296.0; return t >= 4294967295.0 ? 0xFFFFFFFFu : (unsigned)t; };
  p.thr_att = p.use_dropout ? thr(p.p_att) : 0u;
  p.thr_dec = p.use_dropout ? thr(p.p_dec) : 0u;
  const int grid = gtn_decoder_bwd_grid();
  GTN_REQUIRE(grid > 0 && p.grid == grid, "params.grid (%d) must equal gtn_decoder_bwd_grid() (%d)", p.grid, grid);
  cudaStream_t st = (cudaStream_t)stream;
  if (p.B <= 8) return launch_bwd<1>(p, grid, st);
  if (p.B <= 16) return launch_bwd<2>(p, grid, st);
  if (p.B <= 32) return launch_bwd<4>(p, grid, st);
  return launch_bwd<8>(p, grid, st);
}
