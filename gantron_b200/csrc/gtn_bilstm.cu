// Persistent bidirectional LSTM of the text encoder (reference model.py:183-185, 199-213: nn.LSTM(512, 256, 1, batch_first,
// bidirectional) on packed sequences), forward and backward through time, bf16 weights / fp32 state.
//
// The input projection x W_ih^T + b (both directions, all time steps) is ONE batched tcgen05 GEMM on the host side; what is
// left is 150 dependent steps of a [B, 256] x [256, 1024] product per direction.  One cluster of 4 CTAs runs one direction for
// one group of 16 sequences: CTA r owns hidden units [64 r, 64 r + 64).  Its slice of W_hh (forward: 256 gate rows x 256;
// backward: 64 unit rows x 1024) = 128 KB lives in REGISTERS as mma.sync A fragments for the whole launch (8 warps x 32 lanes
// x 128 registers); the only per-step traffic is the 2 KB hidden-state block (8 KB gate-gradient block in the backward pass)
// that every CTA bulk-copies into its three peers' shared memory, completing on the receiver's mbarrier.
// (mma.sync m16n8k16, not tcgen05: the per-step tile is 32 rows x 16 sequences per warp and the accumulators are wanted in
// registers -- row order [i | f] / [g | o] puts all four gates of a cell into ONE thread -- so tensor memory would only add a
// round trip to a 2k-cycle step.)
//
// Packed-sequence semantics (model.py:202-213): sequence b takes part in steps t < L_b only; its outputs beyond L_b are zero,
// the reverse direction starts from a zero state at t = L_b - 1.
#include <cstdlib>
#include "gtn_decoder2.cuh"
#include "gtn_internal.h"
#include "../../include/gantron_b200.h"

namespace gtn {

static constexpr int LH = 256;             // hidden units per direction
static constexpr int LCL = 4;              // CTAs per cluster
static constexpr int LU = LH / LCL;        // units per CTA
static constexpr int LNB = 16;             // sequences per cluster
static constexpr int HROW = LU + 8;        // bf16 elements per (sequence) row of a hidden-state block: 144 B pitch, conflict-free
static constexpr int HBLK = LNB * HROW;    // elements of one CTA's block
static constexpr int GROW = 4 * LU + 8;    // gate-gradient block row (backward)
static constexpr int GBLK = LNB * GROW;

GTN_DEVICE void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
GTN_DEVICE uint32_t ldg_bf16x2(const bf16* p) { return __ldg(reinterpret_cast<const uint32_t*>(p)); }

// one block -> the three peers' copy of it, signalling the receivers' mbarrier
GTN_DEVICE void send_block(const void* src_local, uint32_t bytes, uint64_t* bar_local, int rank) {
#pragma unroll
  for (int d = 1; d < LCL; ++d) {
    const uint32_t dst = (uint32_t)((rank + d) % LCL);
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(map_to_rank(smem_u32(src_local), dst)), "r"(smem_u32(src_local)), "r"(bytes),
                   "r"(map_to_rank(smem_u32(bar_local), dst)) : "memory");
  }
}

// ------------------------------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 1) bilstm_fwd_kernel(const gtn_bilstm_params p) {
  __shared__ __align__(16) bf16 hs[2][LCL * HBLK];          // h(t) of the cluster's 16 sequences: [buffer][owner CTA][seq][unit]
  __shared__ uint64_t full[2];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = (int)cluster_ctarank(), cl = (int)cluster_idx();
  const int ngroups = (p.B + LNB - 1) / LNB;
  const int dir = cl / ngroups, b0 = (cl % ngroups) * LNB;
  const int T = p.T, B = p.B;
  const int r4 = lane >> 2, c2 = (lane & 3) * 2;
  // A fragments of this warp's 8 units: tile 0 rows [i (8) | f (8)], tile 1 rows [g | o]; k = previous hidden unit
  uint32_t wa[2][16][4];
  {
    const int unit = rank * LU + warp * 8 + r4;
    const bf16* W = reinterpret_cast<const bf16*>(p.whh) + (size_t)dir * 4 * LH * LH;
#pragma unroll
    for (int tl = 0; tl < 2; ++tl) {
      const bf16* rlo = W + ((size_t)(2 * tl) * LH + unit) * LH;            // gate 2 tl     (i or g)
      const bf16* rhi = W + ((size_t)(2 * tl + 1) * LH + unit) * LH;        // gate 2 tl + 1 (f or o)
#pragma unroll
      for (int ks = 0; ks < 16; ++ks) {
        wa[tl][ks][0] = ldg_bf16x2(rlo + ks * 16 + c2);
        wa[tl][ks][1] = ldg_bf16x2(rhi + ks * 16 + c2);
        wa[tl][ks][2] = ldg_bf16x2(rlo + ks * 16 + 8 + c2);
        wa[tl][ks][3] = ldg_bf16x2(rhi + ks * 16 + 8 + c2);
      }
    }
  }
  for (int i = tid; i < 2 * LCL * HBLK; i += 256) (&hs[0][0])[i] = __float2bfloat16_rn(0.f);
  if (tid == 0) { mbar_init(&full[0], 1); mbar_init(&full[1], 1); mbar_fence_init(); }
  __syncthreads();
  cluster_sync_all();
  // this thread's four cells: unit u, sequences b0 + nt * 8 + c2 + {0, 1}
  const int u = rank * LU + warp * 8 + r4;
  int len[2][2];
  float c[2][2];
#pragma unroll
  for (int nt = 0; nt < 2; ++nt)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int b = b0 + nt * 8 + c2 + j;
      len[nt][j] = b < B ? min(p.lengths[b], T) : 0;
      c[nt][j] = 0.f;
    }
  const float* xp = p.xproj + (size_t)dir * 4 * LH + u;
  auto load_x = [&](int t, float (&x)[2][2][4]) {
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int b = b0 + nt * 8 + c2 + j;
#pragma unroll
        for (int g = 0; g < 4; ++g)
          x[nt][j][g] = (b < B && t < len[nt][j]) ? __ldg(xp + ((size_t)b * T + t) * (8 * LH) + g * LH) : 0.f;
      }
  };
  float xg[2][2][4];
  load_x(dir == 0 ? 0 : T - 1, xg);
  for (int s = 0; s < T; ++s) {
    const int t = dir == 0 ? s : T - 1 - s;
    const int cur = s & 1, nxt = cur ^ 1;
    // the next step's input projection is requested at the top of the step: its round trip runs underneath the wait for the
    // peers' hidden state, the MMAs and the cells (requested behind the MMAs, a third of the kernel's stall samples sat on it)
    float xn[2][2][4];
    if (s + 1 < T) load_x(dir == 0 ? t + 1 : t - 1, xn);
    // ---- gates = xproj + W_hh h(previous step)
    float acc[2][2][4];
#pragma unroll
    for (int tl = 0; tl < 2; ++tl)
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[tl][nt][k] = 0.f;
    if (s > 0) {
      wait_bar(&full[cur], (uint32_t)((s >> 1) & 1) ^ (cur == 0 ? 1u : 0u));
      const bf16* hb = &hs[cur][0];
#pragma unroll
      for (int ks = 0; ks < 16; ++ks) {
        const int k = ks * 16 + c2, own = k >> 6, kk = k & 63;
        // (k and k + 8 lie in the same owner's block: 64 units per owner)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const bf16* row = hb + own * HBLK + (nt * 8 + r4) * HROW + kk;
          const uint32_t bb0 = *reinterpret_cast<const uint32_t*>(row), bb1 = *reinterpret_cast<const uint32_t*>(row + 8);
          mma16816(acc[0][nt], wa[0][ks], bb0, bb1);
          mma16816(acc[1][nt], wa[1][ks], bb0, bb1);
        }
      }
    }
    // ---- cells (all four gates of a cell sit in this thread: rows r4 / r4 + 8 of the two tiles)
    bf16* hown = &hs[nxt][rank * HBLK];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int bl = nt * 8 + c2 + j, b = b0 + bl;
        float h = 0.f;
        if (t < len[nt][j]) {
          const float i_ = sigmoidf_(acc[0][nt][j] + xg[nt][j][0]), f_ = sigmoidf_(acc[0][nt][2 + j] + xg[nt][j][1]);
          const float g_ = tanhf_(acc[1][nt][j] + xg[nt][j][2]), o_ = sigmoidf_(acc[1][nt][2 + j] + xg[nt][j][3]);
          const float cn = f_ * c[nt][j] + i_ * g_;
          c[nt][j] = cn;
          h = o_ * tanhf_(cn);
          if (p.gates != nullptr) {
            const size_t cell = (((size_t)dir * T + t) * B + b) * LH + u;
            reinterpret_cast<float4*>(p.gates)[cell] = make_float4(i_, f_, g_, o_);
            p.c_all[cell] = cn;
          }
        } else {
          c[nt][j] = 0.f;
        }
        if (b < B) p.out[((size_t)b * T + t) * (2 * LH) + dir * LH + u] = h;
        hown[bl * HROW + warp * 8 + r4] = __float2bfloat16_rn(h);
      }
    if (s + 1 < T) {
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
          for (int g = 0; g < 4; ++g) xg[nt][j][g] = xn[nt][j][g];
      // ---- h(t) -> the peers (their buffer `nxt` was last read one step ago, before they sent the block I have received)
      fence_proxy_async_smem();
      __syncthreads();
      if (tid == 0) {
        mbar_expect_tx(&full[nxt], (uint32_t)((LCL - 1) * HBLK * 2));
        send_block(hown, (uint32_t)(HBLK * 2), &full[nxt], rank);
      }
    }
  }
  cluster_sync_all();        // no CTA exits while a peer may still write into its shared memory
}

// ------------------------------------------------------------------------------------------------------------------------
// backward through time.  d h(t) = d out(t) + W_hh^T d gates(next processed step); CTA r owns units [64 r, 64 r + 64): its
// A fragments are rows of W_hh^T (units) over K = the 1024 gate rows, ordered kk = owner * 256 + gate * 64 + local unit so
// that every owner's d gates block is contiguous.  Warp w: unit tile w & 3 (16 units), K half w >> 2.
// ------------------------------------------------------------------------------------------------------------------------
struct BwdSmem {
  bf16 gs[2][LCL * GBLK];               // d gates of the step just processed: [buffer][owner][seq][gate * 64 + unit]
  float part[2][LU][LNB + 1];           // partial W_hh^T dG of the two K halves
  uint64_t full[2];
};

__global__ void __launch_bounds__(256, 1) bilstm_bwd_kernel(const gtn_bilstm_params p) {
  extern __shared__ __align__(16) uint8_t bwd_smem_raw[];
  BwdSmem& S = *reinterpret_cast<BwdSmem*>(bwd_smem_raw);
  auto& gs = S.gs;
  auto& part = S.part;
  uint64_t* full = S.full;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = (int)cluster_ctarank(), cl = (int)cluster_idx();
  const int ngroups = (p.B + LNB - 1) / LNB;
  const int dir = cl / ngroups, b0 = (cl % ngroups) * LNB;
  const int T = p.T, B = p.B;
  const int r4 = lane >> 2, c2 = (lane & 3) * 2;
  const int ut = warp & 3, kh = warp >> 2;
  uint32_t wa[32][4];
  {
    const bf16* W = reinterpret_cast<const bf16*>(p.whh) + (size_t)dir * 4 * LH * LH;      // [gate * H + unit'][unit]
    const int ulo = rank * LU + ut * 16 + r4, uhi = ulo + 8;
#pragma unroll
    for (int ks = 0; ks < 32; ++ks) {
      // kk = kh * 512 + ks * 16 + {c2, c2 + 1, c2 + 8, c2 + 9}  ->  torch gate row = gate * H + owner * 64 + local
      uint32_t v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int kk0 = kh * 512 + ks * 16 + c2 + (q >> 1) * 8, urow = (q & 1) ? uhi : ulo;
        float w2[2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int kk = kk0 + e, owner = kk >> 8, gate = (kk >> 6) & 3, ul = kk & 63;
          w2[e] = __bfloat162float(W[((size_t)gate * LH + owner * LU + ul) * LH + urow]);
        }
        v[q] = pack_bf16x2(w2[0], w2[1]);
      }
      wa[ks][0] = v[0]; wa[ks][1] = v[1]; wa[ks][2] = v[2]; wa[ks][3] = v[3];
    }
  }
  for (int i = tid; i < 2 * LCL * GBLK; i += 256) (&gs[0][0])[i] = __float2bfloat16_rn(0.f);
  if (tid == 0) { mbar_init(&full[0], 1); mbar_init(&full[1], 1); mbar_fence_init(); }
  __syncthreads();
  cluster_sync_all();
  // cell ownership in the elementwise part: thread = (local unit tid >> 2, sequences 4 (tid & 3) .. + 3)
  const int ul = tid >> 2, bq = (tid & 3) * 4, u = rank * LU + ul;
  int len[4];
  float dc[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int b = b0 + bq + j;
    len[j] = b < B ? min(p.lengths[b], T) : 0;
    dc[j] = 0.f;
  }
  for (int s = 0; s < T; ++s) {
    // the forward pass of this direction ran t = 0 .. T-1 (dir 0) or T-1 .. 0 (dir 1): go back the other way
    const int t = dir == 0 ? T - 1 - s : s;
    const int tprev = dir == 0 ? t - 1 : t + 1;            // the step whose cell state fed this one
    const int cur = s & 1, nxt = cur ^ 1;
    // operands of this step's cell backward: saves of the forward pass and the incoming gradient.  Their addresses depend on
    // (t, sequence, unit) only, so they are requested BEFORE the wait for the previous step's gate gradients: the HBM / L2
    // round trip runs underneath the exchange and the MMAs instead of behind them (ncu: 38 % of the kernel's stall samples
    // sat on the first use of c)
    float4 g4[4];
    float ccv[4], cpv[4], dov[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int b = b0 + bq + j;
      g4[j] = make_float4(0.f, 0.f, 0.f, 0.f); ccv[j] = 0.f; cpv[j] = 0.f; dov[j] = 0.f;
      if (t < len[j]) {
        const size_t cell = (((size_t)dir * T + t) * B + b) * LH + u;
        g4[j] = __ldg(reinterpret_cast<const float4*>(p.gates) + cell);
        ccv[j] = __ldg(p.c_all + cell);
        const bool first = dir == 0 ? t == 0 : t == len[j] - 1;       // the step that started from the zero state
        cpv[j] = first ? 0.f : __ldg(p.c_all + (((size_t)dir * T + tprev) * B + b) * LH + u);
        dov[j] = __ldg(p.dout + ((size_t)b * T + t) * (2 * LH) + dir * LH + u);
      }
    }
    // ---- W_hh^T dG(step processed before)
    float acc[2][4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[nt][k] = 0.f;
    if (s > 0) {
      wait_bar(&full[cur], (uint32_t)((s >> 1) & 1) ^ (cur == 0 ? 1u : 0u));
      const bf16* gb = &gs[cur][0];
#pragma unroll
      for (int ks = 0; ks < 32; ++ks) {
        const int kk = kh * 512 + ks * 16 + c2, owner = kk >> 8, kl = kk & 255;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const bf16* row = gb + owner * GBLK + (nt * 8 + r4) * GROW + kl;
          mma16816(acc[nt], wa[ks], *reinterpret_cast<const uint32_t*>(row), *reinterpret_cast<const uint32_t*>(row + 8));
        }
      }
    }
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      part[kh][ut * 16 + r4][nt * 8 + c2] = acc[nt][0];
      part[kh][ut * 16 + r4][nt * 8 + c2 + 1] = acc[nt][1];
      part[kh][ut * 16 + r4 + 8][nt * 8 + c2] = acc[nt][2];
      part[kh][ut * 16 + r4 + 8][nt * 8 + c2 + 1] = acc[nt][3];
    }
    __syncthreads();
    // ---- cell backward of (unit, 4 sequences)
    bf16* gown = &gs[nxt][rank * GBLK];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int bl = bq + j, b = b0 + bl;
      float dgi = 0.f, dgf = 0.f, dgg = 0.f, dgo = 0.f;
      if (t < len[j]) {
        const float4 g = g4[j];
        const float cc = ccv[j];
        const bool first = dir == 0 ? t == 0 : t == len[j] - 1;       // the step that started from the zero state
        const float cp = cpv[j];
        const float dh = part[0][ul][bl] + part[1][ul][bl] + dov[j];
        const float tc = tanhf_(cc);
        const float d_o = dh * tc;
        const float dcc = dh * g.w * (1.f - tc * tc) + dc[j];
        dc[j] = first ? 0.f : dcc * g.y;
        dgi = dcc * g.z * g.x * (1.f - g.x);
        dgf = dcc * cp * g.y * (1.f - g.y);
        dgg = dcc * g.x * (1.f - g.z * g.z);
        dgo = d_o * g.w * (1.f - g.w);
      } else {
        dc[j] = 0.f;
      }
      bf16* gr = gown + bl * GROW + ul;
      gr[0] = __float2bfloat16_rn(dgi); gr[LU] = __float2bfloat16_rn(dgf);
      gr[2 * LU] = __float2bfloat16_rn(dgg); gr[3 * LU] = __float2bfloat16_rn(dgo);
      if (b < B) {
        // d gates of all steps, torch gate-row order, for the batched weight / input gradient GEMMs
        bf16* o = reinterpret_cast<bf16*>(p.dgates) + ((size_t)b * T + t) * (8 * LH) + dir * 4 * LH + u;
        o[0] = __float2bfloat16_rn(dgi); o[LH] = __float2bfloat16_rn(dgf);
        o[2 * LH] = __float2bfloat16_rn(dgg); o[3 * LH] = __float2bfloat16_rn(dgo);
      }
    }
    if (s + 1 < T) {
      fence_proxy_async_smem();
      __syncthreads();
      if (tid == 0) {
        mbar_expect_tx(&full[nxt], (uint32_t)((LCL - 1) * GBLK * 2));
        send_block(gown, (uint32_t)(GBLK * 2), &full[nxt], rank);
      }
    }
  }
  cluster_sync_all();
}

static int bilstm_launch(const gtn_bilstm_params& p, bool backward, cudaStream_t st) {
  GTN_REQUIRE(p.H == LH, "encoder LSTM kernel: hidden size per direction must be %d (got %d)", LH, p.H);
  GTN_REQUIRE(p.B >= 1 && p.T >= 1 && p.whh && p.lengths && (!backward || (p.gates && p.c_all)), "bad BiLSTM arguments");
  const int ngroups = (p.B + LNB - 1) / LNB;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * ngroups * LCL);
  cfg.blockDim = dim3(256);
  cfg.stream = st;
  cudaLaunchAttribute attrs[1];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = LCL; attrs[0].val.clusterDim.y = 1; attrs[0].val.clusterDim.z = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  // (clusters are independent of one another: no co-residency requirement, any number of waves is fine)
  if (backward) {
    GTN_CUDA_OK(cudaFuncSetAttribute(bilstm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(BwdSmem)));
    cfg.dynamicSmemBytes = sizeof(BwdSmem);
    GTN_CUDA_OK(cudaLaunchKernelEx(&cfg, bilstm_bwd_kernel, p));
  } else {
    GTN_CUDA_OK(cudaLaunchKernelEx(&cfg, bilstm_fwd_kernel, p));
  }
  count_launch();
  return 0;
}

}  // namespace gtn

using namespace gtn;

extern "C" int gtn_bilstm_forward(const gtn_bilstm_params* p, void* stream) {
  GTN_REQUIRE(p != nullptr && p->xproj && p->out, "null BiLSTM forward arguments");
  return bilstm_launch(*p, false, (cudaStream_t)stream);
}
extern "C" int gtn_bilstm_backward(const gtn_bilstm_params* p, void* stream) {
  GTN_REQUIRE(p != nullptr && p->dout && p->dgates, "null BiLSTM backward arguments");
  return bilstm_launch(*p, true, (cudaStream_t)stream);
}
