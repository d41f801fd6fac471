// HBM-bound elementwise / reduction kernels of the Postnet, Discriminator and loss paths (sm_100a).
//
// Every activation tensor here is CHANNELS-LAST: rows = (batch, time) pairs, C contiguous channels per row,
// row stride ld (elements).  A thread owns one group of 8 consecutive channels (one 16-byte bf16 vector or
// two float4) and walks rows with a block-wide stride, so per-channel parameters stay in registers and
// per-channel reductions (BatchNorm statistics, bias gradients) finish with one shared-memory pass and one
// atomicAdd per channel per CTA.  Grids are sized as a multiple of the SM count.
//
// Dropout keep decisions are Philox bits of element (row * C + c) of (seed, site) -- the same numbering the
// implicit-GEMM conv epilogue uses (gtn_gemm.cu) and oracle/philox.py reproduces on the host.
//
// Reference ops replaced (all /root/reference):
//   bn_act_fwd            BatchNorm1d + tanh + F.dropout of Postnet.forward          model.py:146-151
//   bn_act_bwd_*          autograd through the same                                   model.py:146-151
//   drop_tanh_bwd         autograd through Dropout -> Tanh of DiscConv1d              layers.py:53-57
//   disc_head_*           final Conv1d(n_mel -> 1, k=1) of the Discriminator          model.py:512
//   window_mean_*         per-sample masked mean of Discriminator.adversarial_loss    model.py:529-540
//   taco_loss_*           MSELoss x2 + BCEWithLogitsLoss of Tacotron2Loss.forward     loss_function.py:10-22
#include "gtn_common.cuh"
#include "gtn_internal.h"
#include "../../include/gantron_b200.h"

namespace gtn {

typedef __nv_bfloat16 bf16;

struct Vec8 { float v[8]; };

GTN_DEVICE Vec8 load8(const void* base, int is_bf16, size_t off) {
  Vec8 r;
  if (is_bf16) {
    const uint4 q = *reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(base) + off);
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      r.v[2 * i] = __uint_as_float(w[i] << 16);
      r.v[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
    }
  } else {
    const float4 a = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(base) + off);
    const float4 b = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(base) + off + 4);
    r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w; r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  }
  return r;
}
GTN_DEVICE void store8(void* base, int is_bf16, size_t off, const Vec8& x) {
  if (is_bf16) {
    uint4 q = make_uint4(pack_bf16x2(x.v[0], x.v[1]), pack_bf16x2(x.v[2], x.v[3]), pack_bf16x2(x.v[4], x.v[5]),
                         pack_bf16x2(x.v[6], x.v[7]));
    *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(base) + off) = q;
  } else {
    float* p = reinterpret_cast<float*>(base) + off;
    *reinterpret_cast<float4*>(p) = make_float4(x.v[0], x.v[1], x.v[2], x.v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(x.v[4], x.v[5], x.v[6], x.v[7]);
  }
}
// keep flags (bit i = keep element e0 + i) for 8 consecutive elements, e0 % 8 == 0
GTN_DEVICE uint32_t keep8(unsigned long long seed, uint32_t site, unsigned long long e0, uint32_t thr) {
  const uint4 a = philox_bits4(seed, site, 0u, (uint32_t)(e0 >> 2));
  const uint4 b = philox_bits4(seed, site, 0u, (uint32_t)(e0 >> 2) + 1u);
  return (a.x >= thr ? 1u : 0u) | (a.y >= thr ? 2u : 0u) | (a.z >= thr ? 4u : 0u) | (a.w >= thr ? 8u : 0u) |
         (b.x >= thr ? 16u : 0u) | (b.y >= thr ? 32u : 0u) | (b.z >= thr ? 64u : 0u) | (b.w >= thr ? 128u : 0u);
}
static unsigned int drop_threshold(float p) {
  if (p <= 0.f) return 0u;
  double t = (double)p * 4294967296.0;
  return t >= 4294967295.0 ? 0xFFFFFFFFu : (unsigned int)t;
}

// blockDim.x == G * rows_per_iter exactly (make_geom), so every thread owns (channel group, row lane)
struct EltGeom {
  long long rows;
  int C, G;               // channels, channel groups of 8
  int rows_per_iter;      // blockDim.x / G
};

// ------------------------------------------------------------------------------------------------
// BatchNorm finalize: stats[2][C] (sum, sum of squares over `count` samples) -> mean, rstd; optional
// running-stat update (momentum, unbiased variance) exactly as nn.BatchNorm1d in training mode.
// ------------------------------------------------------------------------------------------------
__global__ void bn_finalize_kernel(const float* __restrict__ stats, int C, float count, float eps, float momentum,
                                   float* mean, float* rstd, float* running_mean, float* running_var) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float m = stats[c] / count;
  float var = stats[C + c] / count - m * m;
  var = fmaxf(var, 0.f);
  mean[c] = m;
  rstd[c] = rsqrtf(var + eps);
  if (running_mean != nullptr) {
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m;
    const float unb = count > 1.f ? var * count / (count - 1.f) : var;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unb;
  }
}

// ------------------------------------------------------------------------------------------------
// forward: out = dropout(act(y * scale[c] + shift[c]))   (scale/shift fold BatchNorm; act 0 none / 2 tanh)
// ------------------------------------------------------------------------------------------------
// tanh of the BatchNorm + tanh layers.  bf16 activations (bf16 mode): the hardware approximation (relative error ~2^-11,
// below the bf16 rounding of the stored activation; the decoder kernels use it too) -- libm's tanhf made these kernels
// instruction-bound (50 % issue utilisation at a quarter of the HBM rate).  fp32 activations (fp32 mode): libm.
GTN_DEVICE float tanh_mode(float x, int approx) {
  if (approx) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
  }
  return tanhf(x);
}

__global__ void __launch_bounds__(256)
bn_act_fwd_kernel(const void* __restrict__ y, int y_bf16, long long ldy, EltGeom g, const float* __restrict__ mean,
                  const float* __restrict__ rstd, const float* __restrict__ gamma, const float* __restrict__ beta,
                  int act, uint32_t thr, float dscale, unsigned long long seed, uint32_t site, void* out, int out_bf16,
                  long long ldo) {
  const int cg = threadIdx.x % g.G, r0 = threadIdx.x / g.G;
  const int c0 = cg * 8;
  float sc[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float s = (mean != nullptr) ? gamma[c0 + i] * rstd[c0 + i] : 1.f;
    sc[i] = s;
    sh[i] = (mean != nullptr) ? beta[c0 + i] - mean[c0 + i] * s : 0.f;
  }
  for (long long row = (long long)blockIdx.x * g.rows_per_iter + r0; row < g.rows; row += (long long)gridDim.x * g.rows_per_iter) {
    Vec8 x = load8(y, y_bf16, (size_t)row * ldy + c0);
    const uint32_t keep = thr ? keep8(seed, site, (unsigned long long)row * g.C + c0, thr) : 0xFFu;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float z = x.v[i] * sc[i] + sh[i];
      if (act == 2) z = tanh_mode(z, out_bf16);
      else if (act == 1) z = fmaxf(z, 0.f);
      x.v[i] = (keep >> i) & 1u ? z * dscale : 0.f;
    }
    store8(out, out_bf16, (size_t)row * ldo + c0, x);
  }
}

// ------------------------------------------------------------------------------------------------
// backward, pass 1: dz = dx * keep * dscale * act'(z);  dgamma[c] += sum dz * yhat,  dbeta[c] += sum dz
// ------------------------------------------------------------------------------------------------
GTN_DEVICE void block_channel_reduce(float (&a)[8], float (&b)[8], float* smem, int G, int rpi, int cg, int r0,
                                     float* out_a, float* out_b) {
  // smem: [rpi][G][16]
  float* mine = smem + ((size_t)r0 * G + cg) * 16;
#pragma unroll
  for (int i = 0; i < 8; ++i) { mine[i] = a[i]; mine[8 + i] = b[i]; }
  __syncthreads();
  for (int idx = threadIdx.x; idx < G * 16; idx += blockDim.x) {
    float s = 0.f;
    for (int r = 0; r < rpi; ++r) s += smem[(size_t)r * G * 16 + idx];
    const int cgi = idx >> 4, k = idx & 15;
    if (k < 8) atomicAdd(&out_a[cgi * 8 + k], s);
    else if (out_b != nullptr) atomicAdd(&out_b[cgi * 8 + k - 8], s);
  }
}

__global__ void __launch_bounds__(256)
bn_act_bwd_reduce_kernel(const void* __restrict__ dx, int dx_bf16, long long lddx, const void* __restrict__ y, int y_bf16,
                         long long ldy, EltGeom g, const float* __restrict__ mean, const float* __restrict__ rstd,
                         const float* __restrict__ gamma, const float* __restrict__ beta, int act, uint32_t thr,
                         float dscale, unsigned long long seed, uint32_t site, float* dgamma, float* dbeta) {
  extern __shared__ float red_smem[];
  const int cg = threadIdx.x % g.G, r0 = threadIdx.x / g.G;
  const int c0 = cg * 8;
  float m[8], rs[8], ga[8], be[8], sg[8], sb[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    m[i] = mean[c0 + i]; rs[i] = rstd[c0 + i]; ga[i] = gamma[c0 + i]; be[i] = beta[c0 + i];
    sg[i] = 0.f; sb[i] = 0.f;
  }
  {
    for (long long row = (long long)blockIdx.x * g.rows_per_iter + r0; row < g.rows; row += (long long)gridDim.x * g.rows_per_iter) {
      const Vec8 d = load8(dx, dx_bf16, (size_t)row * lddx + c0);
      const Vec8 yy = load8(y, y_bf16, (size_t)row * ldy + c0);
      const uint32_t keep = thr ? keep8(seed, site, (unsigned long long)row * g.C + c0, thr) : 0xFFu;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float yh = (yy.v[i] - m[i]) * rs[i];
        float dz = (keep >> i) & 1u ? d.v[i] * dscale : 0.f;
        if (act == 2) { const float a = tanh_mode(yh * ga[i] + be[i], y_bf16); dz *= 1.f - a * a; }
        else if (act == 1) { if (yh * ga[i] + be[i] <= 0.f) dz = 0.f; }
        sg[i] += dz * yh;
        sb[i] += dz;
      }
    }
  }
  block_channel_reduce(sg, sb, red_smem, g.G, g.rows_per_iter, cg, r0, dgamma, dbeta);
}

// pass 2: dy = gamma * rstd * (dz - dbeta / N - yhat * dgamma / N);  dbias[c] += sum dy
__global__ void __launch_bounds__(256)
bn_act_bwd_apply_kernel(const void* __restrict__ dx, int dx_bf16, long long lddx, const void* __restrict__ y, int y_bf16,
                        long long ldy, EltGeom g, const float* __restrict__ mean, const float* __restrict__ rstd,
                        const float* __restrict__ gamma, const float* __restrict__ beta, int act, uint32_t thr,
                        float dscale, unsigned long long seed, uint32_t site, const float* __restrict__ dgamma,
                        const float* __restrict__ dbeta, float inv_n, void* dy, int dy_bf16, long long lddy, float* dbias) {
  extern __shared__ float red_smem[];
  const int cg = threadIdx.x % g.G, r0 = threadIdx.x / g.G;
  const int c0 = cg * 8;
  float m[8], rs[8], ga[8], be[8], k1[8], k2[8], sb[8], dummy[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    m[i] = mean[c0 + i]; rs[i] = rstd[c0 + i]; ga[i] = gamma[c0 + i]; be[i] = beta[c0 + i];
    k1[i] = dbeta[c0 + i] * inv_n; k2[i] = dgamma[c0 + i] * inv_n; sb[i] = 0.f; dummy[i] = 0.f;
  }
  {
    for (long long row = (long long)blockIdx.x * g.rows_per_iter + r0; row < g.rows; row += (long long)gridDim.x * g.rows_per_iter) {
      const Vec8 d = load8(dx, dx_bf16, (size_t)row * lddx + c0);
      const Vec8 yy = load8(y, y_bf16, (size_t)row * ldy + c0);
      const uint32_t keep = thr ? keep8(seed, site, (unsigned long long)row * g.C + c0, thr) : 0xFFu;
      Vec8 o;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float yh = (yy.v[i] - m[i]) * rs[i];
        float dz = (keep >> i) & 1u ? d.v[i] * dscale : 0.f;
        if (act == 2) { const float a = tanh_mode(yh * ga[i] + be[i], y_bf16); dz *= 1.f - a * a; }
        else if (act == 1) { if (yh * ga[i] + be[i] <= 0.f) dz = 0.f; }
        o.v[i] = ga[i] * rs[i] * (dz - k1[i] - yh * k2[i]);
        sb[i] += o.v[i];
      }
      store8(dy, dy_bf16, (size_t)row * lddy + c0, o);
    }
  }
  if (dbias != nullptr) {
    block_channel_reduce(sb, dummy, red_smem, g.G, g.rows_per_iter, cg, r0, dbias, nullptr);
  }
}

// ------------------------------------------------------------------------------------------------
// Discriminator layer backward through  a = tanh(dropout(pre)) :  dpre = dout * (1 - a^2) * keep * dscale
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
drop_tanh_bwd_kernel(const void* __restrict__ dout, int dout_bf16, long long lddo, const void* __restrict__ a, int a_bf16,
                     long long lda, EltGeom g, uint32_t thr, float dscale, unsigned long long seed, uint32_t site,
                     void* dpre, int dpre_bf16, long long lddp, float* dbias) {
  extern __shared__ float red_smem[];
  const int cg = threadIdx.x % g.G, r0 = threadIdx.x / g.G;
  const int c0 = cg * 8;
  float sb[8], dummy[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { sb[i] = 0.f; dummy[i] = 0.f; }
  {
    for (long long row = (long long)blockIdx.x * g.rows_per_iter + r0; row < g.rows; row += (long long)gridDim.x * g.rows_per_iter) {
      const Vec8 d = load8(dout, dout_bf16, (size_t)row * lddo + c0);
      const Vec8 aa = load8(a, a_bf16, (size_t)row * lda + c0);
      const uint32_t keep = thr ? keep8(seed, site, (unsigned long long)row * g.C + c0, thr) : 0xFFu;
      Vec8 o;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        o.v[i] = (keep >> i) & 1u ? d.v[i] * (1.f - aa.v[i] * aa.v[i]) * dscale : 0.f;
        sb[i] += o.v[i];
      }
      store8(dpre, dpre_bf16, (size_t)row * lddp + c0, o);
    }
  }
  if (dbias != nullptr) {
    block_channel_reduce(sb, dummy, red_smem, g.G, g.rows_per_iter, cg, r0, dbias, nullptr);
  }
}

// ------------------------------------------------------------------------------------------------
// Discriminator head: out[row] = bias + sum_c w[c] x[row][c]  (one warp per row), and its backward
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
disc_head_fwd_kernel(const void* __restrict__ x, int x_bf16, long long ldx, long long rows, int C,
                     const float* __restrict__ w, const float* __restrict__ bias, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = warp; row < rows; row += nw) {
    float s = 0.f;
    for (int c = lane; c < C; c += 32) {
      const float xv = x_bf16 ? __bfloat162float(reinterpret_cast<const bf16*>(x)[row * ldx + c])
                              : reinterpret_cast<const float*>(x)[row * ldx + c];
      s += xv * w[c];
    }
    s = warp_sum(s);
    if (lane == 0) out[row] = s + bias[0];
  }
}
// dx[row][c] = dout[row] * w[c];  dw[c] += sum_row dout[row] x[row][c];  db += sum_row dout[row]
__global__ void __launch_bounds__(256)
disc_head_bwd_kernel(const float* __restrict__ dout, const void* __restrict__ x, int x_bf16, long long ldx, long long rows,
                     int C, const float* __restrict__ w, void* dx, int dx_bf16, long long lddx, float* dw, float* db) {
  // thread -> channel c = threadIdx.x % Cp, row lane = threadIdx.x / Cp  (Cp = C rounded to... simple 1 elem/thread)
  const int tpr = C < 256 ? C : 256;        // threads per row; wider heads (LinearDiscriminator: 512) loop over the channels
  const int rpi = blockDim.x / tpr;
  const int c0 = threadIdx.x % tpr, r0 = threadIdx.x / tpr;
  if (r0 < rpi) {
    for (int c = c0; c < C; c += tpr) {
      float acc_w = 0.f, acc_b = 0.f;
      const float wc = w[c];
      for (long long row = (long long)blockIdx.x * rpi + r0; row < rows; row += (long long)gridDim.x * rpi) {
        const float g = dout[row];
        const float xv = x_bf16 ? __bfloat162float(reinterpret_cast<const bf16*>(x)[row * ldx + c])
                                : reinterpret_cast<const float*>(x)[row * ldx + c];
        acc_w += g * xv;
        if (c == 0) acc_b += g;
        const float d = g * wc;
        if (dx_bf16) reinterpret_cast<bf16*>(dx)[row * lddx + c] = __float2bfloat16_rn(d);
        else reinterpret_cast<float*>(dx)[row * lddx + c] = d;
      }
      if (dw != nullptr) atomicAdd(&dw[c], acc_w);
      if (db != nullptr && c == 0) atomicAdd(db, acc_b);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// adversarial loss reduction: loss = (1/B) sum_b mean(out[b][0 : ceil(len_b / window)])
// ------------------------------------------------------------------------------------------------
__global__ void window_mean_fwd_kernel(const float* __restrict__ out, int B, int P, const long long* __restrict__ lengths,
                                       int window, float* loss) {
  float s = 0.f;
  for (int b = threadIdx.x >> 5; b < B; b += blockDim.x >> 5) {
    long long n = (lengths[b] + window - 1) / window;
    n = n < 1 ? 1 : (n > P ? P : n);
    float t = 0.f;
    for (int p = threadIdx.x & 31; p < n; p += 32) t += out[(size_t)b * P + p];
    t = warp_sum(t);
    s += t / (float)n;
  }
  __shared__ float scr[32];
  if ((threadIdx.x & 31) == 0) scr[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float r = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) r += scr[w];
    loss[0] = r / (float)B;
  }
}
__global__ void window_mean_bwd_kernel(const float* __restrict__ gloss, int B, int P, const long long* __restrict__ lengths,
                                       int window, float* __restrict__ dout) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * P) return;
  const int b = idx / P, p = idx - b * P;
  long long n = (lengths[b] + window - 1) / window;
  n = n < 1 ? 1 : (n > P ? P : n);
  dout[idx] = p < n ? gloss[0] / ((float)n * (float)B) : 0.f;
}

// ------------------------------------------------------------------------------------------------
// Tacotron2Loss: sums[0] = sum (mel - tgt)^2, sums[1] = sum (mel_post - tgt)^2, sums[2] = sum bce(gate, gtgt)
// mel / mel_post are addressed as [b][c][t] through explicit strides (they are channels-last views).
// ------------------------------------------------------------------------------------------------
struct LossView { const float* p; long long sb, sc, st; };

__global__ void __launch_bounds__(256)
taco_loss_fwd_kernel(LossView mel, LossView post, LossView tgt, int B, int C, int T, const float* __restrict__ gate,
                     long long gsb, const float* __restrict__ gate_tgt, long long tsb, float* sums) {
  const long long n = (long long)B * C * T;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    // iterate with c fastest: the predictions are channels-last in memory
    const int c = (int)(i % C);
    const long long bt = i / C;
    const int t = (int)(bt % T), b = (int)(bt / T);
    const float y = tgt.p[b * tgt.sb + c * tgt.sc + t * tgt.st];
    const float d0 = mel.p[b * mel.sb + c * mel.sc + t * mel.st] - y;
    const float d1 = post.p[b * post.sb + c * post.sc + t * post.st] - y;
    s0 += d0 * d0; s1 += d1 * d1;
    if (c == 0) {
      const float x = gate[b * gsb + t], z = gate_tgt[b * tsb + t];
      s2 += fmaxf(x, 0.f) - x * z + log1pf(expf(-fabsf(x)));
    }
  }
  s0 = warp_sum(s0); s1 = warp_sum(s1); s2 = warp_sum(s2);
  __shared__ float scr[3][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { scr[0][warp] = s0; scr[1][warp] = s1; scr[2][warp] = s2; }
  __syncthreads();
  if (threadIdx.x < 3) {
    float r = 0.f;
    for (int w = 0; w < 8; ++w) r += scr[threadIdx.x][w];
    atomicAdd(&sums[threadIdx.x], r);
  }
}
// gradients: dmel = g[0] * 2 (mel - tgt) / n, dpost likewise, dgate = g[1] * (sigmoid(x) - z) / (B T)
// outputs are dense channels-last [b][t][c] (dmel, dpost) and [b][t] (dgate)
__global__ void __launch_bounds__(256)
taco_loss_bwd_kernel(LossView mel, LossView post, LossView tgt, int B, int C, int T, const float* __restrict__ gate,
                     long long gsb, const float* __restrict__ gate_tgt, long long tsb, const float* __restrict__ g_mel,
                     const float* __restrict__ g_gate, float* dmel, float* dpost, float* dgate) {
  const long long n = (long long)B * C * T;
  const float km = 2.f * g_mel[0] / (float)n, kg = g_gate[0] / ((float)B * (float)T);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const long long bt = i / C;
    const int t = (int)(bt % T), b = (int)(bt / T);
    const float y = tgt.p[b * tgt.sb + c * tgt.sc + t * tgt.st];
    dmel[i] = km * (mel.p[b * mel.sb + c * mel.sc + t * mel.st] - y);
    dpost[i] = km * (post.p[b * post.sb + c * post.sc + t * post.st] - y);
    if (c == 0) {
      const float x = gate[b * gsb + t], z = gate_tgt[b * tsb + t];
      dgate[bt] = kg * (1.f / (1.f + expf(-x)) - z);
    }
  }
}


// ------------------------------------------------------------------------------------------------
// Guided-attention loss (loss_function.py:25-52): mean_b BCE(align[b].T[:ne_b, :nd_b], guide_b),
// guide_b[n][f] = exp(-(f - c_n)^2 / (width * 3)), c_n = n * (nd_b - 1) // (ne_b - 1)   (integer floor).
// align is [B][T][Tin].  fwd accumulates loss[0]; bwd writes d(align) (zero outside the valid rectangle).
// ------------------------------------------------------------------------------------------------
GTN_DEVICE float guide_value(int n, int f, int ne, int nd, float inv_w) {
  const int c = (int)(((long long)n * (nd - 1)) / (ne > 1 ? ne - 1 : 1));
  const float d = (float)f - (float)c;
  return expf(-d * d * inv_w);
}
__global__ void __launch_bounds__(256)
attn_guide_fwd_kernel(const float* __restrict__ align, int B, int T, int Tin, const long long* __restrict__ enc_len,
                      const long long* __restrict__ dec_len, float inv_w, float* loss) {
  const long long n_all = (long long)B * T * Tin;
  float s = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_all; i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i % Tin);
    const long long bf = i / Tin;
    const int f = (int)(bf % T), b = (int)(bf / T);
    const int ne = (int)min((long long)Tin, enc_len[b]), nd = (int)min((long long)T, dec_len[b]);
    if (n < ne && f < nd) {
      const float w = align[i], g = guide_value(n, f, ne, nd, inv_w);
      const float lw = fmaxf(logf(w), -100.f), l1w = fmaxf(log1pf(-w), -100.f);     // torch's BCE clamps the logs
      s += -(g * lw + (1.f - g) * l1w) / ((float)ne * (float)nd);
    }
  }
  s = warp_sum(s);
  __shared__ float scr[8];
  if ((threadIdx.x & 31) == 0) scr[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float r = 0.f;
    for (int w = 0; w < 8; ++w) r += scr[w];
    atomicAdd(loss, r / (float)B);
  }
}
__global__ void __launch_bounds__(256)
attn_guide_bwd_kernel(const float* __restrict__ align, int B, int T, int Tin, const long long* __restrict__ enc_len,
                      const long long* __restrict__ dec_len, float inv_w, const float* __restrict__ gloss, float* dalign) {
  const long long n_all = (long long)B * T * Tin;
  const float gl = gloss[0] / (float)B;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_all; i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i % Tin);
    const long long bf = i / Tin;
    const int f = (int)(bf % T), b = (int)(bf / T);
    const int ne = (int)min((long long)Tin, enc_len[b]), nd = (int)min((long long)T, dec_len[b]);
    float d = 0.f;
    if (n < ne && f < nd) {
      const float w = align[i], g = guide_value(n, f, ne, nd, inv_w);
      // d/dw of -(g log w + (1-g) log(1-w)); zero where torch's clamp at -100 is active
      const float a = logf(w) > -100.f ? -g / w : 0.f;
      const float c = log1pf(-w) > -100.f ? (1.f - g) / (1.f - w) : 0.f;
      d = gl * (a + c) / ((float)ne * (float)nd);
    }
    dalign[i] = d;
  }
}

static int sm_count() {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) sms = 148;
  }
  return sms;
}

static int make_geom(EltGeom* g, long long rows, int C, int* threads, int* blocks) {
  GTN_REQUIRE(C % 8 == 0 && C >= 8 && C <= 2048, "channel count %d must be a multiple of 8 in [8, 2048]", C);
  g->rows = rows; g->C = C; g->G = C / 8;
  g->rows_per_iter = 256 / g->G;
  GTN_REQUIRE(g->rows_per_iter >= 1, "too many channels for one CTA row");
  *threads = g->G * g->rows_per_iter;
  long long want = (rows + g->rows_per_iter - 1) / g->rows_per_iter;
  const long long cap = (long long)sm_count() * 8;
  *blocks = (int)(want < 1 ? 1 : (want > cap ? cap : want));
  return 0;
}

}  // namespace gtn

using namespace gtn;

extern "C" {

int gtn_bn_finalize(const float* stats, int C, long long count, float eps, float momentum, float* mean, float* rstd,
                    float* running_mean, float* running_var, void* stream) {
  GTN_REQUIRE(C > 0 && count > 0, "empty BatchNorm");
  bn_finalize_kernel<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(stats, C, (float)count, eps, momentum, mean, rstd,
                                                                       running_mean, running_var);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

int gtn_bn_act_fwd(const void* y, int y_dtype, long long ldy, long long rows, int C, const float* mean, const float* rstd,
                   const float* gamma, const float* beta, int act, float drop_p, unsigned long long seed, int site,
                   void* out, int out_dtype, long long ldo, void* stream) {
  EltGeom g; int th, bl;
  if (int rc = make_geom(&g, rows, C, &th, &bl)) return rc;
  GTN_REQUIRE(ldy % 8 == 0 && ldo % 8 == 0, "row strides must be multiples of 8 elements");
  const uint32_t thr = drop_threshold(drop_p);
  bn_act_fwd_kernel<<<bl, th, 0, (cudaStream_t)stream>>>(y, y_dtype, ldy, g, mean, rstd, gamma, beta, act, thr,
                                                         drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f, seed, (uint32_t)site,
                                                         out, out_dtype, ldo);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

static int bn_act_bwd_impl(const void* dx, int dx_dtype, long long lddx, const void* y, int y_dtype, long long ldy, long long rows,
                           int C, const float* mean, const float* rstd, const float* gamma, const float* beta, int act,
                           float drop_p, unsigned long long seed, int site, float* dgamma, float* dbeta, void* dy, int dy_dtype,
                           long long lddy, float* dbias, void* stream, bool batch_stats) {
  EltGeom g; int th, bl;
  if (int rc = make_geom(&g, rows, C, &th, &bl)) return rc;
  GTN_REQUIRE(lddx % 8 == 0 && ldy % 8 == 0 && lddy % 8 == 0, "row strides must be multiples of 8 elements");
  const uint32_t thr = drop_threshold(drop_p);
  const float ds = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  const size_t smem = (size_t)g.rows_per_iter * g.G * 16 * sizeof(float);
  cudaStream_t st = (cudaStream_t)stream;
  GTN_CUDA_OK(cudaMemsetAsync(dgamma, 0, C * sizeof(float), st));
  GTN_CUDA_OK(cudaMemsetAsync(dbeta, 0, C * sizeof(float), st));
  if (dbias) GTN_CUDA_OK(cudaMemsetAsync(dbias, 0, C * sizeof(float), st));
  bn_act_bwd_reduce_kernel<<<bl, th, smem, st>>>(dx, dx_dtype, lddx, y, y_dtype, ldy, g, mean, rstd, gamma, beta, act, thr,
                                                 ds, seed, (uint32_t)site, dgamma, dbeta);
  GTN_CUDA_OK(cudaGetLastError());
  bn_act_bwd_apply_kernel<<<bl, th, smem, st>>>(dx, dx_dtype, lddx, y, y_dtype, ldy, g, mean, rstd, gamma, beta, act, thr,
                                                ds, seed, (uint32_t)site, dgamma, dbeta, batch_stats ? 1.f / (float)rows : 0.f,
                                                dy, dy_dtype, lddy, dbias);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch(2);
  return 0;
}

int gtn_bn_act_bwd(const void* dx, int dx_dtype, long long lddx, const void* y, int y_dtype, long long ldy, long long rows,
                   int C, const float* mean, const float* rstd, const float* gamma, const float* beta, int act,
                   float drop_p, unsigned long long seed, int site, float* dgamma, float* dbeta, void* dy, int dy_dtype,
                   long long lddy, float* dbias, void* stream) {
  return bn_act_bwd_impl(dx, dx_dtype, lddx, y, y_dtype, ldy, rows, C, mean, rstd, gamma, beta, act, drop_p, seed, site, dgamma,
                         dbeta, dy, dy_dtype, lddy, dbias, stream, true);
}

// module.eval(): mean / rstd are the running statistics, i.e. constants: the batch-coupling terms vanish (inv_n = 0)
int gtn_bn_act_bwd_eval(const void* dx, int dx_dtype, long long lddx, const void* y, int y_dtype, long long ldy, long long rows,
                        int C, const float* mean, const float* rstd, const float* gamma, const float* beta, int act,
                        float drop_p, unsigned long long seed, int site, float* dgamma, float* dbeta, void* dy, int dy_dtype,
                        long long lddy, float* dbias, void* stream) {
  return bn_act_bwd_impl(dx, dx_dtype, lddx, y, y_dtype, ldy, rows, C, mean, rstd, gamma, beta, act, drop_p, seed, site, dgamma,
                         dbeta, dy, dy_dtype, lddy, dbias, stream, false);
}

int gtn_drop_tanh_bwd(const void* dout, int dout_dtype, long long lddo, const void* a, int a_dtype, long long lda,
                      long long rows, int C, float drop_p, unsigned long long seed, int site, void* dpre, int dpre_dtype,
                      long long lddp, float* dbias, void* stream) {
  EltGeom g; int th, bl;
  if (int rc = make_geom(&g, rows, C, &th, &bl)) return rc;
  GTN_REQUIRE(lddo % 8 == 0 && lda % 8 == 0 && lddp % 8 == 0, "row strides must be multiples of 8 elements");
  const size_t smem = (size_t)g.rows_per_iter * g.G * 16 * sizeof(float);
  cudaStream_t st = (cudaStream_t)stream;
  if (dbias) GTN_CUDA_OK(cudaMemsetAsync(dbias, 0, C * sizeof(float), st));
  drop_tanh_bwd_kernel<<<bl, th, smem, st>>>(dout, dout_dtype, lddo, a, a_dtype, lda, g, drop_threshold(drop_p),
                                             drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f, seed, (uint32_t)site, dpre,
                                             dpre_dtype, lddp, dbias);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

int gtn_disc_head_fwd(const void* x, int x_dtype, long long ldx, long long rows, int C, const float* w, const float* bias,
                      float* out, void* stream) {
  GTN_REQUIRE(rows > 0 && C > 0, "empty head");
  long long blocks = (rows + 7) / 8;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  disc_head_fwd_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(x, x_dtype, ldx, rows, C, w, bias, out);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

int gtn_disc_head_bwd(const float* dout, const void* x, int x_dtype, long long ldx, long long rows, int C, const float* w,
                      void* dx, int dx_dtype, long long lddx, float* dw, float* db, void* stream) {
  GTN_REQUIRE(rows > 0 && C > 0 && C <= 8192, "head width must be <= 8192");
  cudaStream_t st = (cudaStream_t)stream;
  if (dw) GTN_CUDA_OK(cudaMemsetAsync(dw, 0, C * sizeof(float), st));
  if (db) GTN_CUDA_OK(cudaMemsetAsync(db, 0, sizeof(float), st));
  const int rpi = 256 / (C < 256 ? C : 256);
  long long blocks = (rows + rpi - 1) / rpi;
  const long long cap = (long long)sm_count() * 4;
  if (blocks > cap) blocks = cap;
  disc_head_bwd_kernel<<<(int)blocks, 256, 0, st>>>(dout, x, x_dtype, ldx, rows, C, w, dx, dx_dtype, lddx, dw, db);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

int gtn_window_mean_fwd(const float* out, int B, int P, const long long* lengths, int window, float* loss, void* stream) {
  GTN_REQUIRE(B > 0 && P > 0 && window > 0, "empty adversarial loss");
  window_mean_fwd_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(out, B, P, lengths, window, loss);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

int gtn_window_mean_bwd(const float* gloss, int B, int P, const long long* lengths, int window, float* dout, void* stream) {
  GTN_REQUIRE(B > 0 && P > 0 && window > 0, "empty adversarial loss");
  window_mean_bwd_kernel<<<(B * P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(gloss, B, P, lengths, window, dout);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

int gtn_taco_loss_fwd(const float* mel, const float* post, const float* tgt, const long long* strides9, int B, int C, int T,
                      const float* gate, long long gate_sb, const float* gate_tgt, long long tgt_sb, float* sums,
                      void* stream) {
  GTN_REQUIRE(B > 0 && C > 0 && T > 0, "empty loss");
  LossView m{mel, strides9[0], strides9[1], strides9[2]}, p{post, strides9[3], strides9[4], strides9[5]},
      t{tgt, strides9[6], strides9[7], strides9[8]};
  cudaStream_t st = (cudaStream_t)stream;
  GTN_CUDA_OK(cudaMemsetAsync(sums, 0, 3 * sizeof(float), st));
  const long long n = (long long)B * C * T;
  long long blocks = (n + 1023) / 1024;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  taco_loss_fwd_kernel<<<(int)blocks, 256, 0, st>>>(m, p, t, B, C, T, gate, gate_sb, gate_tgt, tgt_sb, sums);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

int gtn_taco_loss_bwd(const float* mel, const float* post, const float* tgt, const long long* strides9, int B, int C, int T,
                      const float* gate, long long gate_sb, const float* gate_tgt, long long tgt_sb, const float* g_mel,
                      const float* g_gate, float* dmel, float* dpost, float* dgate, void* stream) {
  GTN_REQUIRE(B > 0 && C > 0 && T > 0, "empty loss");
  LossView m{mel, strides9[0], strides9[1], strides9[2]}, p{post, strides9[3], strides9[4], strides9[5]},
      t{tgt, strides9[6], strides9[7], strides9[8]};
  const long long n = (long long)B * C * T;
  long long blocks = (n + 1023) / 1024;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  taco_loss_bwd_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(m, p, t, B, C, T, gate, gate_sb, gate_tgt, tgt_sb,
                                                                      g_mel, g_gate, dmel, dpost, dgate);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

int gtn_attn_guide_loss_fwd(const float* align, int B, int T, int Tin, const long long* enc_len, const long long* dec_len,
                            float width, float* loss, void* stream) {
  GTN_REQUIRE(B > 0 && T > 0 && Tin > 0 && width > 0.f, "empty attention loss");
  cudaStream_t st = (cudaStream_t)stream;
  GTN_CUDA_OK(cudaMemsetAsync(loss, 0, sizeof(float), st));
  const long long n = (long long)B * T * Tin;
  long long blocks = (n + 1023) / 1024;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  attn_guide_fwd_kernel<<<(int)blocks, 256, 0, st>>>(align, B, T, Tin, enc_len, dec_len, 1.f / (width * 3.f), loss);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

int gtn_attn_guide_loss_bwd(const float* align, int B, int T, int Tin, const long long* enc_len, const long long* dec_len,
                            float width, const float* gloss, float* dalign, void* stream) {
  GTN_REQUIRE(B > 0 && T > 0 && Tin > 0 && width > 0.f, "empty attention loss");
  const long long n = (long long)B * T * Tin;
  long long blocks = (n + 1023) / 1024;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  attn_guide_bwd_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(align, B, T, Tin, enc_len, dec_len,
                                                                       1.f / (width * 3.f), gloss, dalign);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

}  // extern "C"
