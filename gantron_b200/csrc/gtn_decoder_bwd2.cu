// Persistent clustered decoder kernel, backward through time (second generation).  Reverse of decode()
// (reference model.py:378-417) for t = T-1 .. 0, replacing autograd's replay of the per-frame tape built by
// model.py:444-449.  Same machinery as gtn_decoder_fwd2.cu (gtn_decoder2.cuh): clusters of 4 CTAs split K of
// every per-frame GEMM, operands arrive by TMA, tcgen05.mma accumulates in TMEM, partial tiles are exchanged
// through distributed shared memory.
//
// Per frame, three grid-synchronised phases:
//   P1  data-gradient GEMMs, one 128-row tile per cluster:
//         dX_dec = W_dec^T . dGates_dec(t)   rows [h_att | ctx_enc | h_dec]     (model.py:406-407)
//         dX_att = W_att^T . dGates_att(t+1) rows [ctx_enc | h_att]             (model.py:391-392)
//       the h_dec rows feed the decoder LSTM cell backward of frame t-1 directly in the epilogue; the prenet rows and
//       the time-constant context columns are not computed per frame (deferred to K = T*B GEMMs on the host).
//   P2  attention backward, one CTA pair per sequence: context bmm, softmax, e = v.tanh(q + loc + pm), and the
//       pre-multiplied location filter U: d[prev | cum] = dth . U and dU += dth^T . im2col both on tcgen05, reusing
//       the SAME shared-memory tiles (a K-major tile with rows = positions is an MN-major tile with K = positions);
//       dU stays in TMEM for the whole pass                                      (model.py:27-31,62-89,396-403)
//   P3  query backward (Wq^T dq) + attention LSTM cell backward, spread over all CTAs by hidden unit (model.py:391-394)
// All WEIGHT gradients are deferred: the kernel stores per-frame gate / query gradients and the host issues a
// handful of K = T*B tcgen05 GEMMs afterwards (gtn_gemm_tn_bf16).
#include <cstdlib>
#include <cstring>
#include <type_traits>
#include "gtn_decoder2.cuh"
#include "gtn_internal.h"
#include "../../include/gantron_b200.h"

namespace gtn {

typedef gtn_decoder_bwd2_params DecBwd2;
#define BFP(x) (reinterpret_cast<bf16*>(const_cast<void*>(static_cast<const void*>(x))))

// dev instrumentation: CTA 0 -> prof[0..511], CTA 48 (first h_dec tile at the default dims) -> prof[512..1023]
#define PROFB(slot)                                                                         \
  do {                                                                                      \
    if (p.prof != nullptr && (blockIdx.x == 0 || blockIdx.x == 48) && threadIdx.x == 0 && it < 16)                 \
      p.prof[(blockIdx.x ? 512 : 0) + it * 32 + (slot)] = clock64();                        \
  } while (0)
#define PROFB_T(thr, slot)                                                                  \
  do {                                                                                      \
    if (p.prof != nullptr && blockIdx.x == 0 && threadIdx.x == (thr) && it < 16) p.prof[it * 32 + (slot)] = clock64(); \
  } while (0)

struct Bwd2Maps {
  CUtensorMap wT_att, wT_dec, ut;      // weights (A operands) and U^T (B operand of d[prev|cum] = dth . U)
  CUtensorMap xdg_a, xdg_d;            // gate gradients as X operands
  CUtensorMap wq;                      // bf16 Wq [A][Ha] seen as fp32 pairs: box = one rank's quarter of the units x all A rows
  CUtensorMap mem;                     // bf16 encoder memory [B][Tin][E]: box = 64 columns x (half a sequence rounded up to 8 rows)
};

static constexpr int B2_TIN_MAX = 256;
static constexpr int B2_JH = 128;
static constexpr int P2_RING_BYTES = 6 * W_TILE_BYTES;
static constexpr int WQ_UNITS = 16;        // hidden units per CTA in P3 (attention_rnn_dim / grid, rounded up) must not exceed this     // dth hi (2 blocks) + dth lo (2) + im2col hi + lo

struct SmemB {
  Ring ring;
  float* recv;         // [2][4][128][NC4]   (P2: G_s [128][64] swizzled)
  uint8_t* ut_s;       // [4][8 KB]: U^T hi a<64, hi a>=64, lo a<64, lo a>=64  (N = 64 rows kk, K = 64 a)
  float* dctx_s;       // [E <= 1024]
  float* w_s;          // [Tin]
  float* dw_s;         // [Tin]   dw, then de
  float* wcat;         // [2][JH + KS - 1]
  float* dwc_s;        // [2][JH + KS - 1]   d wcat (own window)
  float* halo_s;       // [2][2][KS]         halo contributions received from the peer
  float* v_s;          // [128]
  float* part_s;       // [8][64]
  float* dqp_s;        // [4][128]           dq partials of the cluster's two sequences, slot = source rank
  float* dv_acc;       // [128]
  float* dot_s;        // [4]                softmax dot partial: own, peer (double buffered)
  float* scr;          // [64]
  float* dqf_s;        // [2][128]  full d query of the cluster's two sequences (cell backward)
  uint64_t* acc_full;
  uint64_t* g_full;
  uint64_t* ut_full;
  uint64_t* recv_full;  // [2] bulk-copy exchange: one per receive buffer
  uint64_t* mem_full;   // the encoder-memory tile of this CTA's positions has landed in the ring (bulk copy)
  uint64_t* th_full;    // the frame's tanh tile (th_all rows of this CTA's positions) has landed in the ring
  uint64_t* wq_full;    // this rank's quarter of Wq has landed in the ring
  uint64_t* du_done;    // the dU MMAs of the frame have retired (their operand tiles in the ring may be overwritten)
  uint64_t* dw_done;    // dw = mem . dctx on the tensor core has retired
  uint32_t* tmem_slot;
  uint32_t ring_bytes;
};

template <int N>
__host__ __device__ constexpr size_t bwd2_fixed_bytes() {
  constexpr int NC4 = (N < 64 ? N : 64) / 4;
  // N <= 64: one exchange per frame and a grid barrier before the next one, so a single receive buffer is safe and the
  // 32 KB saved buys two more TMA stages; N > 64: back-to-back chunk exchanges need the double buffer
  constexpr size_t recv_b = (size_t)(N > 64 ? 2 : 1) * 4 * TILE_M * NC4 * 4;
  return (recv_b < 32768 ? 32768 : recv_b) + 4 * 8192 +
         (size_t)(1024 + 2 * B2_TIN_MAX + 4 * (B2_JH + 62) + 4 * 32 + 128 + 8 * 64 + 512 + 128 + 4 + 64 + 256) * 4 + 64 * 8 + 64;
}

template <int N>
GTN_DEVICE SmemB carveB(uint8_t* base, int S) {
  constexpr int NC4 = (N < 64 ? N : 64) / 4;
  constexpr size_t recv_b = (size_t)(N > 64 ? 2 : 1) * 4 * TILE_M * NC4 * 4;
  SmemB m;
  m.ring.base = base;
  m.ring.S = S;
  m.ring.stage_bytes = W_TILE_BYTES + N * KB * 2;
  m.ring.prod_it = m.ring.mma_it = 0;
  size_t ring_b = (size_t)S * m.ring.stage_bytes;
  if (ring_b < (size_t)P2_RING_BYTES) ring_b = P2_RING_BYTES;
  m.ring_bytes = (uint32_t)ring_b;
  uint8_t* p = base + ring_b;
  m.recv = reinterpret_cast<float*>(p); p += (recv_b < 32768 ? 32768 : recv_b);
  m.ut_s = p; p += 4 * 8192;
  float* f = reinterpret_cast<float*>(p);
  m.dctx_s = f; f += 1024;
  m.w_s = f; f += B2_TIN_MAX;
  m.dw_s = f; f += B2_TIN_MAX;
  m.wcat = f; f += 2 * (B2_JH + 62);
  m.dwc_s = f; f += 2 * (B2_JH + 62);
  m.halo_s = f; f += 4 * 32;
  m.v_s = f; f += 128;
  m.part_s = f; f += 8 * 64;
  m.dqp_s = f; f += 512;
  m.dv_acc = f; f += 128;
  m.dot_s = f; f += 4;
  m.scr = f; f += 64;
  m.dqf_s = f; f += 256;
  uint64_t* b = reinterpret_cast<uint64_t*>(f);
  m.ring.full = b; b += 16;
  m.ring.empty = b; b += 16;
  m.acc_full = b++;
  m.g_full = b++;
  m.ut_full = b++;
  m.recv_full = b; b += 2;
  m.mem_full = b++;
  m.th_full = b++;
  m.wq_full = b++;
  m.du_done = b++;
  m.dw_done = b++;
  m.tmem_slot = reinterpret_cast<uint32_t*>(b + 1);
  return m;
}

// split grid barrier: ARRIVE, then work that does not depend on the other CTAs, then WAIT (thread 0 carries the fences)
// proxy_fence: this CTA wrote global data with ordinary stores that ANOTHER CTA reads through TMA after the barrier (the gate
// gradients, barrier 2).  P1's results (barrier 1) are only read by ordinary loads; skipping the fence there matters because a
// fence.proxy.async issued while the memory tile's bulk copy is in flight waits for it (3.3k cycles per frame).
GTN_DEVICE void grid_arrive2(unsigned int* counter, bool proxy_fence = true, long long* stamp = nullptr) {
  __syncthreads();
  if (threadIdx.x == 0) {
    if (stamp) stamp[0] = clock64();
    if (proxy_fence) fence_proxy_async();
    red_release_add(counter, 1u);
    if (stamp) stamp[1] = clock64();
  }
}
GTN_DEVICE void grid_wait2(unsigned int* counter, unsigned int target) {
  if (threadIdx.x == 0) {
    long long t0 = clock64();
    while (ld_acquire_u32(counter) < target) {
      if (clock64() - t0 > 20000000000LL) __trap();
    }
    fence_proxy_async();
  }
  __syncthreads();
}

// Where the attention backward keeps its tiles inside the (then idle) TMA ring.  `fast` layout (one sequence per CTA pair):
// im2col tiles at the END of the ring and the bf16 encoder-memory tile of this CTA's positions (bulk-copied once per frame,
// while P1's epilogue and barrier 1 are in flight) from byte 32 K on -- behind the staging area of P1's split-K exchange;
// the dth tiles overwrite it after the d(weights) reduction has consumed it.
static constexpr uint32_t MEM_TILE_OFF = 2 * W_TILE_BYTES;
// The frame's tanh tile (fp32 [positions][A], one contiguous run of th_all) is bulk-copied behind the dth tiles once the
// d(weights) reduction has consumed the memory tile: the dth stage then reads shared memory instead of waiting three
// dependent L2 round trips (7.5k -> ~3k cycles).  This rank's quarter of Wq (bf16, A rows x units/4) lands at the start of
// the ring by ONE TMA box once the attention MMAs have retired: the cells' Wq^T dq product reads it from shared memory
// instead of streaming 64 KB per frame through the LSU (6k -> ~1.5k cycles).
static constexpr uint32_t TH_TILE_OFF = 4 * W_TILE_BYTES;
// mma: the encoder-memory tile arrives as SWIZZLE_128B k-block tiles (TMA) and dw = mem . dctx runs on tcgen05 (M = positions,
// N = 8: the context gradient as hi / lo bf16 rows, K = E) instead of 16 FMA chains per warp (3.7k -> ~1k cycles)
struct P2Tiles { bool th; bool wq; bool mma; };

GTN_DEVICE void red_add_f2(float* addr, float a, float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

GTN_DEVICE float block_sum2(float v, float* scr) {
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

// LSTM cell backward for one (b, unit) of frame t, split into an operand fetch (issued early so that several cells'
// global loads overlap) and the arithmetic.  dh = gradient w.r.t. the (post-dropout) hidden output.
struct CellIn {
  float4 g;            // i, f, g, o (post-activation) of frame t
  float c, cprev, dc;  // c(t), c(t-1), carried d c
};
GTN_DEVICE CellIn cell_load(const DecBwd2& p, int t, int b, int unit, int H, const float* gates_all, const float* c_all,
                            const float* dc_carry) {
  CellIn in;
  const size_t cell = ((size_t)t * p.B + b) * H + unit;
  in.g = __ldg(reinterpret_cast<const float4*>(gates_all) + cell);
  in.c = __ldg(&c_all[cell]);
  in.cprev = t > 0 ? __ldg(&c_all[cell - (size_t)p.B * H]) : 0.f;
  in.dc = dc_carry[(size_t)b * H + unit];
  return in;
}
// Writes the four pre-activation gate gradients (gate-major K order g*H + unit) as GEMM operand and as fp32 save.
template <bool F32>
GTN_DEVICE void cell_apply(const DecBwd2& p, const CellIn& in, int t, int b, int unit, int H, float dh, float* dc_carry,
                           float* dg_all, bf16* xdg, unsigned site, unsigned thr, float pdrop) {
  constexpr int PREC = F32 ? 3 : 1, MULT = F32 ? 6 : 1;
  const size_t ci = (size_t)b * H + unit;
  if (p.use_dropout && thr != 0u) {
    const uint32_t bits = philox_bits(p.seed, site, (uint32_t)t, (uint32_t)ci);
    dh = bits >= thr ? dh * (1.0f / (1.0f - pdrop)) : 0.f;
  }
  const float4 g = in.g;
  const float tc = F32 ? tanh_acc(in.c) : tanh_fast(in.c);
  const float d_o = dh * tc;
  const float dc = dh * g.w * (1.f - tc * tc) + in.dc;
  dc_carry[ci] = dc * g.y;
  const float dgi = dc * g.z * g.x * (1.f - g.x);
  const float dgf = dc * in.cprev * g.y * (1.f - g.y);
  const float dgg = dc * g.x * (1.f - g.z * g.z);
  const float dgo = d_o * g.w * (1.f - g.w);
  const size_t g0 = ((size_t)t * p.B + b) * 4 * H + unit;
  if (!F32) {
    bf16* dgb = reinterpret_cast<bf16*>(dg_all);          // bf16 mode: the saves are GEMM operands only
    dgb[g0] = __float2bfloat16_rn(dgi); dgb[g0 + H] = __float2bfloat16_rn(dgf);
    dgb[g0 + 2 * H] = __float2bfloat16_rn(dgg); dgb[g0 + 3 * H] = __float2bfloat16_rn(dgo);
  } else {
    dg_all[g0] = dgi; dg_all[g0 + H] = dgf; dg_all[g0 + 2 * H] = dgg; dg_all[g0 + 3 * H] = dgo;
  }
  bf16* xr = xdg + (size_t)b * MULT * 4 * H;
  store_act(xr, unit, 4 * H, PREC, dgi);
  store_act(xr, H + unit, 4 * H, PREC, dgf);
  store_act(xr, 2 * H + unit, 4 * H, PREC, dgg);
  store_act(xr, 3 * H + unit, 4 * H, PREC, dgo);
}
// Two ADJACENT units (unit even) at once: in bf16 mode the eight bf16 saves / GEMM operands of the pair go out as four + four
// 32-bit stores instead of sixteen 16-bit ones, and the carried dc as one float2.
struct CellGrad { float dgi, dgf, dgg, dgo, dcn; };
// bits: the element's 32 random bits (the caller draws one Philox block for the pair of adjacent units)
template <bool F32>
GTN_DEVICE CellGrad cell_math(const DecBwd2& p, const CellIn& in, uint32_t bits, float dh, unsigned thr, float pdrop) {
  if (p.use_dropout && thr != 0u) dh = bits >= thr ? dh * (1.0f / (1.0f - pdrop)) : 0.f;
  const float4 g = in.g;
  const float tc = F32 ? tanh_acc(in.c) : tanh_fast(in.c);
  const float d_o = dh * tc;
  const float dc = dh * g.w * (1.f - tc * tc) + in.dc;
  CellGrad r;
  r.dcn = dc * g.y;
  r.dgi = dc * g.z * g.x * (1.f - g.x);
  r.dgf = dc * in.cprev * g.y * (1.f - g.y);
  r.dgg = dc * g.x * (1.f - g.z * g.z);
  r.dgo = d_o * g.w * (1.f - g.w);
  return r;
}
GTN_DEVICE uint2 cell_pair_bits(const DecBwd2& p, int t, int b, int unit, int H, unsigned site, unsigned thr) {
  // ci is even, so ci and ci + 1 share one Philox block (element e = component e & 3 of block e >> 2): one draw for the pair
  const size_t ci = (size_t)b * H + unit;
  uint2 bits = make_uint2(0u, 0u);
  if (p.use_dropout && thr != 0u) {
    const uint4 r = philox_bits4(p.seed, site, (uint32_t)t, (uint32_t)(ci >> 2));
    const bool lo = (ci & 2) == 0;
    bits = make_uint2(lo ? r.x : r.z, lo ? r.y : r.w);
  }
  return bits;
}
// pre: the pair's dropout bits drawn earlier by the caller (cell_pair_bits; bf16 mode only), or nullptr
template <bool F32>
GTN_DEVICE void cell_apply2(const DecBwd2& p, const CellIn (&in)[2], int t, int b, int unit, int H, float dh0, float dh1,
                            float* dc_carry, float* dg_all, bf16* xdg, unsigned site, unsigned thr, float pdrop,
                            const uint2* pre = nullptr) {
  if (F32) {                                      // fp32 mode: per-unit path (three planes per operand)
    cell_apply<F32>(p, in[0], t, b, unit, H, dh0, dc_carry, dg_all, xdg, site, thr, pdrop);
    cell_apply<F32>(p, in[1], t, b, unit + 1, H, dh1, dc_carry, dg_all, xdg, site, thr, pdrop);
    return;
  }
  const size_t ci = (size_t)b * H + unit;
  const uint2 bits = pre != nullptr ? *pre : cell_pair_bits(p, t, b, unit, H, site, thr);
  const CellGrad a = cell_math<F32>(p, in[0], bits.x, dh0, thr, pdrop), c = cell_math<F32>(p, in[1], bits.y, dh1, thr, pdrop);
  *reinterpret_cast<float2*>(dc_carry + ci) = make_float2(a.dcn, c.dcn);
  const uint32_t pi = pack_bf16x2(a.dgi, c.dgi), pf = pack_bf16x2(a.dgf, c.dgf), pg = pack_bf16x2(a.dgg, c.dgg),
                 po = pack_bf16x2(a.dgo, c.dgo);
  uint32_t* dgb = reinterpret_cast<uint32_t*>(reinterpret_cast<bf16*>(dg_all) + ((size_t)t * p.B + b) * 4 * H + unit);
  dgb[0] = pi; dgb[H >> 1] = pf; dgb[H] = pg; dgb[3 * (H >> 1)] = po;
  uint32_t* xr = reinterpret_cast<uint32_t*>(xdg + (size_t)b * 4 * H + unit);
  xr[0] = pi; xr[H >> 1] = pf; xr[H] = pg; xr[3 * (H >> 1)] = po;
}
template <bool F32>
GTN_DEVICE void cell_backward2(const DecBwd2& p, int t, int b, int unit, int H, float dh, const float* gates_all,
                               const float* c_all, float* dc_carry, float* dg_all, bf16* xdg, unsigned site, unsigned thr,
                               float pdrop) {
  const CellIn in = cell_load(p, t, b, unit, H, gates_all, c_all, dc_carry);
  cell_apply<F32>(p, in, t, b, unit, H, dh, dc_carry, dg_all, xdg, site, thr, pdrop);
}

// decoder LSTM cell backward of frame t (prologue only: frame T-1 has no recurrent carry), spread over the grid
template <bool F32>
GTN_DEVICE void dec_cell_prologue(const DecBwd2& p, int t) {
  const int n = p.B * p.Hd;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += gridDim.x * blockDim.x) {
    const int b = idx / p.Hd, u = idx - b * p.Hd;
    const float dh = p.dproj_all[((size_t)t * p.B + b) * (p.Hd + p.D) + u];
    cell_backward2<F32>(p, t, b, u, p.Hd, dh, p.gates_d, p.c_d_all, p.dc_d, p.dgd_all, BFP(p.xdg_d), 4u, p.thr_dec, p.p_dec);
  }
}

GTN_DEVICE void attn_bwd_window(const DecBwd2& p, SmemB& sm, int t, int b, int j0, int jn, int tid, int nthr);

template <int N, bool F32>
GTN_DEVICE void dgrad_phase2(const DecBwd2& p, const Bwd2Maps& maps, SmemB& sm, int t, int it, int rank, int cid, int ncl,
                             uint32_t tmem_base, uint32_t& acc_par, uint32_t& round, int res_n, int win_b, int win_j0, int win_jn) {
  constexpr int CH = N < 64 ? N : 64;
  constexpr int NC4 = CH / 4;
  const int R_dec = p.Ha + p.E + p.Hd, R_att = p.E + p.Ha;
  const int nD = (R_dec + TILE_M - 1) / TILE_M;
  const int nA = (t + 1 < p.T) ? (R_att + TILE_M - 1) / TILE_M : 0;
  const int cur = it & 1, nxt = cur ^ 1;
  for (int tile = cid; tile < nD + nA; tile += ncl) {
    const bool dec = tile < nD;
    const int row0 = (dec ? tile : tile - nD) * TILE_M;
    GemmJob job;
    job.wmap = dec ? &maps.wT_dec : &maps.wT_att;
    job.row0 = row0;
    job.nseg = 1;
    job.seg[0] = dec ? XSeg{&maps.xdg_d, (F32 ? 6 : 1) * 4 * p.Hd / KB, cur} : XSeg{&maps.xdg_a, (F32 ? 6 : 1) * 4 * p.Ha / KB, 0};
    job.nkb = job.seg[0].nkb;
    job.res_n = tile == cid ? res_n : 0;
    job.res_slot0 = 0;
    const bool has_work = gemm_issue<N>(job, sm.ring, tmem_base, sm.acc_full, rank, tmem_base);
    PROFB(1);
    __syncwarp();
    // warps 4..7 have nothing to do until the accumulator is complete: they fetch the attention backward's window meanwhile
    // (fast layout; ordered before its readers by the barriers of the exchange below)
    if ((threadIdx.x >> 5) >= 4 && win_b >= 0 && tile == cid) attn_bwd_window(p, sm, t, win_b, win_j0, win_jn, threadIdx.x - 128, 128);
    if ((threadIdx.x >> 5) >= 4 && has_work) {
      wait_bar(sm.acc_full, acc_par);
      tc_fence_after();
    }
    PROFB_T(128, 2);
    if (has_work) acc_par ^= 1u;
    for (int c0 = 0; c0 < N; c0 += CH) {
      float* recv = sm.recv + (size_t)(N > 64 ? (round & 1u) : 0u) * 4 * TILE_M * NC4;
      const int bbase = c0 + rank * NC4;
      if constexpr (N > 64) {
        // two receive buffers: the other one is the staging area (see exchange_bulk)
        exchange_bulk<CH, false, true>(has_work, tmem_base, c0, recv, sm.recv + (size_t)((round & 1u) ^ 1u) * 4 * TILE_M * NC4,
                                       &sm.recv_full[round & 1u], (round >> 1) & 1u, rank);
      } else if (nD + nA <= ncl) {
        // one receive buffer, one exchange per frame with grid barriers in between: stage in the (idle) TMA ring
        exchange_bulk<CH, false, true>(has_work, tmem_base, c0, recv, reinterpret_cast<float*>(sm.ring.base), &sm.recv_full[0], round & 1u, rank);
      } else {
        exchange_chunk<CH, true>(has_work, tmem_base, c0, recv, rank);
      }
      ++round;
      PROFB(3);
      for (int idx = threadIdx.x; idx < TILE_M * NC4; idx += blockDim.x) {
        const int r = idx % TILE_M, bl = idx / TILE_M, b = bbase + bl, row = row0 + r;
        if (b >= p.B) continue;
        // (bulk exchanges deliver column-major blocks: lanes over rows read consecutive words; the DSMEM-store fall-back
        // keeps the swizzled row-major layout)
        const float val = (N > 64 || nD + nA <= ncl) ? recv_sum_cm<NC4>(recv, r, bl) : recv_sum<NC4, true>(recv, r, bl);
        if (dec) {
          if (row < p.Ha) p.dha_dec[(size_t)b * p.Ha + row] = val;
          else if (row < p.Ha + p.E) p.dctx_dec[(size_t)b * p.Ep + row - p.Ha] = val;
          else if (row < R_dec) p.dhd_rec[(size_t)b * p.Hd + row - p.Ha - p.E] = val;   // consumed by P3 (cell backward of t-1)
        } else {
          if (row < p.E) p.dctx_att[(size_t)b * p.Ep + row] = val;
          else if (row < R_att) p.carry_ha[(size_t)b * p.Ha + row - p.E] = val;
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Cell backward of the cluster's TWO sequences, run right after their attention backward (this used to be a third,
// unit-sliced phase behind a grid barrier; per sequence it needs no other cluster's results):
//   * attention LSTM cell of frame t:     dh = Wq^T dq + dha_dec + carry_ha                   (model.py:391-394, 62)
//   * decoder LSTM cell of frame t - 1:   dh = dhd_rec (P1 of this frame) + dproj(t-1)         (model.py:404-409)
// A thread owns two ADJACENT units, so every save / gradient access of a warp is one contiguous 128..512-byte run.
// Wq ([A][Ha], bf16 in bf16 mode) streams from L2 against the d query of both sequences held in shared memory.
// ------------------------------------------------------------------------------------------------
// Work split over the whole CLUSTER: rank r takes quarter r of the hidden units for BOTH sequences of the cluster, so
// that a Wq value fetched from L2 is used for two sequences (the Wq stream is the largest single item of the phase:
// 128 KB per CTA per frame through a ~30 B/clk port; 64 KB this way).  Thread = (unit pair pr, half h): during the Wq
// product h selects the half of the attention dimension (both sequences), partial sums meet in shared memory, and for the
// cell arithmetic h selects the sequence.  dq2 = [2][128] full d query of the two sequences, part = [2][2][128][2] scratch.
// operands of one round of quad_cells: fetched ahead of the pair's second cluster barrier (none of them depends on it)
struct QuadOps {
  CellIn ca[2], cd[2];
  float2 adda, addd;
};
GTN_DEVICE void quad_cells_load(const DecBwd2& p, int t, int b0, int b1, int rank, int rd, QuadOps& o) {
  const int Hq_a = ((p.Ha / CL + 1) / 2) * 2, ua_beg = rank * Hq_a, ua_end = min(p.Ha, (rank + 1) * Hq_a);
  const int Hq_d = ((p.Hd / CL + 1) / 2) * 2, ud_beg = rank * Hq_d, ud_end = min(p.Hd, (rank + 1) * Hq_d);
  const int pr = threadIdx.x & 127, h = threadIdx.x >> 7;
  const int b = h == 0 ? b0 : b1;
  const bool seq_ok = b < p.B;
  const int ua = ua_beg + rd * 256 + 2 * pr, ud = ud_beg + rd * 256 + 2 * pr;
  o.adda = make_float2(0.f, 0.f);
  o.addd = make_float2(0.f, 0.f);
  if (t > 0 && ud < ud_end && seq_ok) {
    o.cd[0] = cell_load(p, t - 1, b, ud, p.Hd, p.gates_d, p.c_d_all, p.dc_d);
    o.cd[1] = cell_load(p, t - 1, b, ud + 1, p.Hd, p.gates_d, p.c_d_all, p.dc_d);
    const float2 r = __ldcg(reinterpret_cast<const float2*>(p.dhd_rec + (size_t)b * p.Hd + ud));
    const float* q = p.dproj_all + ((size_t)(t - 1) * p.B + b) * (p.Hd + p.D) + ud;     // row stride Hd + D may be odd
    o.addd = make_float2(r.x + __ldg(q), r.y + __ldg(q + 1));
  }
  if (ua < ua_end && seq_ok) {
    o.ca[0] = cell_load(p, t, b, ua, p.Ha, p.gates_a, p.c_a_all, p.dc_a);
    o.ca[1] = cell_load(p, t, b, ua + 1, p.Ha, p.gates_a, p.c_a_all, p.dc_a);
    const size_t ci = (size_t)b * p.Ha + ua;
    const float2 x = __ldcg(reinterpret_cast<const float2*>(p.dha_dec + ci)), y = __ldcg(reinterpret_cast<const float2*>(p.carry_ha + ci));
    o.adda = make_float2(x.x + y.x, x.y + y.y);
  }
}
// rounds of 256 units per rank the cells need (1 at the default dims: then the operands can be fetched early)
GTN_DEVICE int quad_rounds(const DecBwd2& p, int t) {
  const int Hq_a = ((p.Ha / CL + 1) / 2) * 2, Hq_d = ((p.Hd / CL + 1) / 2) * 2;
  return max((Hq_a + 255) / 256, t > 0 ? (Hq_d + 255) / 256 : 0);
}

template <bool F32>
GTN_DEVICE void quad_cells(const DecBwd2& p, const float* dq2, float* part, int t, int it, int b0, int b1, int rank,
                           const uint32_t* wq_s, uint64_t* wq_full, uint32_t wq_par, const QuadOps* pre) {
  const int nxt = (it & 1) ^ 1;
  bf16* xnext_d = BFP(p.xdg_d) + (size_t)nxt * p.B * (F32 ? 6 : 1) * 4 * p.Hd;
  const int Hq_a = ((p.Ha / CL + 1) / 2) * 2, ua_beg = rank * Hq_a, ua_end = min(p.Ha, (rank + 1) * Hq_a);
  const int Hq_d = ((p.Hd / CL + 1) / 2) * 2, ud_beg = rank * Hq_d, ud_end = min(p.Hd, (rank + 1) * Hq_d);
  const int na = (max(ua_end - ua_beg, 0) + 255) / 256, nd = t > 0 ? (max(ud_end - ud_beg, 0) + 255) / 256 : 0;
  const int pr = threadIdx.x & 127, h = threadIdx.x >> 7;
  const int b = h == 0 ? b0 : b1;
  const bool seq_ok = b < p.B;
  const int A2 = p.A / 2, a_lo = h * A2;
  for (int rd = 0; rd < max(na, nd); ++rd) {
    const int ua = ua_beg + rd * 256 + 2 * pr;
    const int ud = ud_beg + rd * 256 + 2 * pr;
    const bool oka_u = rd < na && ua < ua_end;                       // unit pair exists (Wq product: both sequences)
    const bool oka = oka_u && seq_ok, okd = rd < nd && ud < ud_end && seq_ok;
    QuadOps o;
    if (rd == 0 && pre != nullptr) o = *pre;
    else quad_cells_load(p, t, b0, b1, rank, rd, o);
    // the cells' dropout bits depend on (seed, site, frame, element) only: drawn here, in the shadow of the Wq tile's arrival
    uint2 bits_a = make_uint2(0u, 0u), bits_d = make_uint2(0u, 0u);
    if (!F32) {
      if (oka) bits_a = cell_pair_bits(p, t, b, ua, p.Ha, 3u, p.thr_att);
      if (okd) bits_d = cell_pair_bits(p, t - 1, b, ud, p.Hd, 4u, p.thr_dec);
    }
    // Wq^T dq over this thread's half of the attention dimension, for both sequences
    float s00 = 0.f, s01 = 0.f, s10 = 0.f, s11 = 0.f;                // [sequence][unit of the pair]
    if (!F32 && wq_s != nullptr) {
      // Wq quarter in shared memory: [A][Hq / 2] words, word pr = this thread's unit pair (a warp reads 128 contiguous bytes)
      wait_bar(wq_full, wq_par);
      PROFB(27);
      if (oka_u) {
        const int ldw = (((p.Ha / CL + 1) / 2) * 2) >> 1;
        const uint32_t* w = wq_s + (size_t)a_lo * ldw + pr;
        const float* d0 = dq2 + a_lo;
        const float* d1 = dq2 + 128 + a_lo;
#pragma unroll 16
        for (int a = 0; a < A2; ++a) {
          const uint32_t wv = w[a * ldw];
          const float w0 = __uint_as_float(wv << 16), w1 = __uint_as_float(wv & 0xFFFF0000u);
          s00 += w0 * d0[a]; s01 += w1 * d0[a];
          s10 += w0 * d1[a]; s11 += w1 * d1[a];
        }
      }
    } else if (oka_u) {
      const float* d0 = dq2 + a_lo;
      const float* d1 = dq2 + 128 + a_lo;
      const int ld = p.Ha >> 1;
      if constexpr (!F32) {
        const uint32_t* w = reinterpret_cast<const uint32_t*>(p.wq_au) + (size_t)a_lo * ld + (ua >> 1);
#pragma unroll 32
        for (int a = 0; a < A2; ++a) {
          // 32 rows in flight, WITHOUT L1 allocation: with allocating loads (__ldg) 32 in flight ran slower than 16 (the ~28 KB
          // of L1 left beside the shared-memory carve-out cannot hold 8 warps x 32 lines)
          uint32_t wv;
          asm volatile("ld.global.nc.L1::no_allocate.b32 %0, [%1];" : "=r"(wv) : "l"(w + (size_t)a * ld));
          const float w0 = __uint_as_float(wv << 16), w1 = __uint_as_float(wv & 0xFFFF0000u);
          s00 += w0 * d0[a]; s01 += w1 * d0[a];
          s10 += w0 * d1[a]; s11 += w1 * d1[a];
        }
      } else {
        const float2* w = reinterpret_cast<const float2*>(p.wq_au) + (size_t)a_lo * ld + (ua >> 1);
#pragma unroll 8
        for (int a = 0; a < A2; ++a) {
          const float2 wv = __ldg(w + (size_t)a * ld);
          s00 += wv.x * d0[a]; s01 += wv.y * d0[a];
          s10 += wv.x * d1[a]; s11 += wv.y * d1[a];
        }
      }
    }
    PROFB(25);
    // part[h][seq][pr] (float2): this thread's half-sums for both sequences
    __syncthreads();                                                 // previous round's readers are done
    reinterpret_cast<float2*>(part)[(h * 2 + 0) * 128 + pr] = make_float2(s00, s01);
    reinterpret_cast<float2*>(part)[(h * 2 + 1) * 128 + pr] = make_float2(s10, s11);
    __syncthreads();
    const float2 h0 = reinterpret_cast<const float2*>(part)[(0 * 2 + h) * 128 + pr];
    const float2 h1 = reinterpret_cast<const float2*>(part)[(1 * 2 + h) * 128 + pr];
    const float s0 = h0.x + h1.x, s1 = h0.y + h1.y;                  // sequence h, units ua, ua + 1
    PROFB(26);
    if (okd) {
      cell_apply2<F32>(p, o.cd, t - 1, b, ud, p.Hd, o.addd.x, o.addd.y, p.dc_d, p.dgd_all, xnext_d, 4u, p.thr_dec, p.p_dec,
                       F32 ? nullptr : &bits_d);
    }
    if (oka) {
      cell_apply2<F32>(p, o.ca, t, b, ua, p.Ha, s0 + o.adda.x, s1 + o.adda.y, p.dc_a, p.dga_all, BFP(p.xdg_a), 3u, p.thr_att, p.p_att,
                       F32 ? nullptr : &bits_a);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// P2: attention backward, one CTA pair per sequence
// ------------------------------------------------------------------------------------------------
template <typename MT>
GTN_DEVICE void ld8b(const MT* p, float (&o)[8]);
template <>
GTN_DEVICE void ld8b<float>(const float* p, float (&o)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w; o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}
template <>
GTN_DEVICE void ld8b<bf16>(const bf16* p, float (&o)[8]) {
  const uint4 q = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    o[2 * i] = __uint_as_float(w[i] << 16);
    o[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}

// Window, im2col tiles and the frame's alignment row of ONE sequence: everything of the attention backward that depends on
// the forward pass's saves only.  In the `fast` layout it runs between ARRIVE and WAIT of barrier 1.
// the frame's alignment row and the location-conv window [prev | cum] of one sequence -> shared memory (threads tid of nthr)
GTN_DEVICE void attn_bwd_window(const DecBwd2& p, SmemB& sm, int t, int b, int j0, int jn, int tid, int nthr) {
  const int Tin = p.Tin, KS = p.KS, hk = (p.KS - 1) / 2;
  const int Wd = jn + KS - 1;
  const size_t tb = (size_t)t * p.B + b, bt = (size_t)b * p.T + t;
  for (int j = tid; j < Tin; j += nthr) sm.w_s[j] = p.align[bt * Tin + j];
  for (int i = tid; i < 2 * Wd; i += nthr) {
    const int ch = i / Wd, jj = i - ch * Wd, j = j0 + jj - hk;
    float val = 0.f;
    if (j >= 0 && j < Tin) val = ch == 0 ? (t > 0 ? p.align[(bt - 1) * Tin + j] : 0.f) : p.cum_all[tb * Tin + j];
    sm.wcat[i] = val;
  }
}
// window_loaded: attn_bwd_window already ran for this frame (warps 4..7, while P1's operands stream)
GTN_DEVICE void attn_bwd_prepare(const DecBwd2& p, SmemB& sm, int t, int b, bool active, int j0, int jn, uint8_t* im_hi, uint8_t* im_lo,
                                 bool window_loaded = false) {
  const int KS = p.KS;
  const int Wd = jn + KS - 1;
  if (active && !window_loaded) attn_bwd_window(p, sm, t, b, j0, jn, threadIdx.x, blockDim.x);
  if (!window_loaded) __syncthreads();
  if (p.prof != nullptr && blockIdx.x == 0 && threadIdx.x == 0 && (p.T - 1 - t) < 16) p.prof[(p.T - 1 - t) * 32 + 14] = clock64();
  if (active) {
    build_im2col(sm.wcat, Wd, KS, jn, im_hi, im_lo);
    // zero the K-rows of the im2col tiles beyond jn (they are the K extent of dU += dth^T . im2col)
    for (int i = threadIdx.x; i < (TILE_M - jn) * 8; i += blockDim.x) {
      const int r = jn + (i >> 3), chunk = i & 7;
      const uint32_t off = sw128_off(r, chunk);
      *reinterpret_cast<uint4*>(im_hi + off) = make_uint4(0u, 0u, 0u, 0u);
      *reinterpret_cast<uint4*>(im_lo + off) = make_uint4(0u, 0u, 0u, 0u);
    }
  }
}

// prepared: attn_bwd_prepare already ran for this frame (fast layout); mem_tile: the encoder-memory tile is in the ring
template <bool F32>
GTN_DEVICE void attn_bwd_phase2(const DecBwd2& p, const Bwd2Maps& maps, SmemB& sm, int t, int it, int rank, int cid, int ncl,
                                uint32_t tmem_g, uint32_t tmem_du, uint32_t& g_par, uint32_t& xround, bool prepared, bool mem_tile,
                                uint32_t& mem_par, const P2Tiles tiles, uint32_t& tile_par) {
  typedef typename std::conditional<F32, float, bf16>::type MT;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Tin = p.Tin, A = p.A, KS = p.KS, hk = (p.KS - 1) / 2, E = p.E;
  const int half = rank & 1, peer = rank ^ 1;
  const int pair = cid * 2 + (rank >> 1), npairs = ncl * 2;
  const int JH = (Tin + 1) / 2;
  const int j0 = half * JH, jn = min(Tin, j0 + JH) - j0;
  const int Wd = jn + KS - 1;
  const MT* mem = reinterpret_cast<const MT*>(p.mem_enc);
  uint8_t* dth_hi = sm.ring.base;                        // 2 blocks of [128 rows j][64 a] (a < 64, a >= 64)
  uint8_t* dth_lo = sm.ring.base + 2 * W_TILE_BYTES;
  uint8_t* im_hi = sm.ring.base + (prepared ? sm.ring_bytes - 2 * W_TILE_BYTES : 4 * W_TILE_BYTES);
  uint8_t* im_lo = im_hi + W_TILE_BYTES;
  float* G_s = sm.recv;                                  // [128][63]: odd row stride = conflict-free for row-per-lane stores and the diagonal reads (2 KS <= 62 columns are used)
  const int niter = (p.B + npairs - 1) / npairs;
  for (int itb = 0; itb < niter; ++itb) {
    const int b = pair + itb * npairs;
    const bool active = b < p.B;
    const int L = active ? reinterpret_cast<const int*>(sm.scr + 48)[itb] : 0;   // staged once at kernel start
    const size_t tb = (size_t)t * p.B + b, bt = (size_t)b * p.T + t;
    const uint32_t xb = xround & 1u;
    ++xround;
    if (active) {
      // all loads of the (up to four) elements a thread owns are issued before the first store: one L2 round trip, not one per trip
      float cv[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int d = threadIdx.x + q * 256;
        cv[q] = 0.f;
        if (d < E)
          cv[q] = p.dproj_all[tb * (p.Hd + p.D) + p.Hd + d] + __ldcg(&p.dctx_dec[(size_t)b * p.Ep + d]) +
                  __ldcg(&p.dctx_att[(size_t)b * p.Ep + d]);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int d = threadIdx.x + q * 256;
        if (d < E) {
          sm.dctx_s[d] = cv[q];
          if (half == 0) p.dctx_all[bt * E + d] = cv[q];
          if (tiles.mma) {
            // B operand of dw: per 64-column k-block an [8 rows][64] K-major SWIZZLE_128B tile in the (idle) receive buffer,
            // row 0 = bf16(dctx), row 1 = the residual; rows 2..7 are never read back (their D columns are ignored)
            uint8_t* bt_ = reinterpret_cast<uint8_t*>(sm.recv) + (d >> 6) * 1024 + (d & 7) * 2;
            const int c = (d & 63) >> 3;
            const float hi = bf16_round(cv[q]);
            *reinterpret_cast<__nv_bfloat16*>(bt_ + (c << 4)) = __float2bfloat16_rn(hi);
            *reinterpret_cast<__nv_bfloat16*>(bt_ + 128 + ((c ^ 1) << 4)) = __float2bfloat16_rn(cv[q] - hi);
          }
        }
      }
      if (tiles.mma) fence_proxy_async_smem();
    }
    // what reaches w_j through the next frame's location features and the upstream alignment gradient: fetched together with
    // the context gradient (one L2 round trip instead of two); this CTA wrote dwprev / dcum of its positions itself
    float dw_extra = 0.f;
    if (active && (int)threadIdx.x < jn && j0 + (int)threadIdx.x < L) {
      const int j = j0 + threadIdx.x;
      dw_extra = __ldcg(&p.dwprev[(size_t)b * p.ldT + j]) + __ldcg(&p.dcum[(size_t)b * p.ldT + j]);
      if (p.dalign) dw_extra += p.dalign[bt * Tin + j];
    }
    if (!prepared) attn_bwd_prepare(p, sm, t, b, active, j0, jn, im_hi, im_lo);
    __syncthreads();
    PROFB(10);
    float dot_part = 0.f;
    if (tiles.mma && mem_tile) {
      // dw[j] = sum_d mem[j][d] dctx[d] on tcgen05: A = the memory tiles (128 rows from each tile base: the rows beyond the
      // box belong to the next tile, their results are not read), B = the context gradient tiles, D = first columns of tmem_g
      const int box1 = ((((Tin + 1) / 2) + 7) / 8) * 8;
      if (active && __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0) == 1) {
        wait_bar(sm.mem_full, mem_par);
        tc_fence_after();
        const uint32_t idesc = umma_idesc_bf16(TILE_M, 8);
        const uint32_t a0 = smem_u32(sm.ring.base + MEM_TILE_OFF), b0 = smem_u32(sm.recv);
        if (elect_one()) {
          for (int kb = 0; kb < E / KB; ++kb) {
            const uint64_t da = umma_desc_sw128(a0 + (uint32_t)kb * box1 * 128u), db = umma_desc_sw128(b0 + (uint32_t)kb * 1024u);
#pragma unroll
            for (int k = 0; k < KB / 16; ++k)
              umma_bf16(tmem_g, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((kb | k) != 0));
          }
          umma_commit(sm.dw_done);
        }
        __syncwarp();
      }
      PROFB(22);
      if (active) {
        wait_bar(sm.dw_done, (tile_par >> 3) & 1u);
        tc_fence_after();
        if (warp >= 4) {
          const int wq = warp & 3, jl = wq * 32 + lane;
          uint32_t r[32];
          tmem_ld_32x32(tmem_g + ((uint32_t)(wq * 32) << 16), r);
          tmem_ld_wait();
          if (jl < jn) sm.dw_s[j0 + jl] = (j0 + jl < L) ? __uint_as_float(r[0]) + __uint_as_float(r[1]) : 0.f;
        }
        tc_fence_before();
      }
    } else if (active && mem_tile) {
      // d w_j = mem[j][:] . dctx over this CTA's positions, from the tile in shared memory: one warp per position, lanes over
      // the columns (16 bytes per lane and trip: conflict-free)
      wait_bar(sm.mem_full, mem_par);
      PROFB(22);
      const bf16* mt = reinterpret_cast<const bf16*>(sm.ring.base + MEM_TILE_OFF);
      // warp w owns positions w, w + 8, ... (at most 16): ALL of its rows accumulate at once (independent FMA chains; the
      // one-row-at-a-time version spent 6.7k cycles in dependent FMAs and shuffles), then one 16-value butterfly
      float acc[16];
#pragma unroll
      for (int r = 0; r < 16; ++r) acc[r] = 0.f;
      for (int d = lane * 8; d < E; d += 256) {
        const float4 c0 = *reinterpret_cast<const float4*>(sm.dctx_s + d), c1 = *reinterpret_cast<const float4*>(sm.dctx_s + d + 4);
#pragma unroll
        for (int r = 0; r < 16; ++r) {
          const int jl = warp + 8 * r;
          if (jl < jn && j0 + jl < L) {
            const uint4 q = *reinterpret_cast<const uint4*>(mt + (size_t)jl * E + d);
            const float s0 = __uint_as_float(q.x << 16) * c0.x + __uint_as_float(q.y << 16) * c0.z;
            const float s1 = __uint_as_float(q.x & 0xFFFF0000u) * c0.y + __uint_as_float(q.y & 0xFFFF0000u) * c0.w;
            const float s2 = __uint_as_float(q.z << 16) * c1.x + __uint_as_float(q.w << 16) * c1.z;
            const float s3 = __uint_as_float(q.z & 0xFFFF0000u) * c1.y + __uint_as_float(q.w & 0xFFFF0000u) * c1.w;
            acc[r] += (s0 + s1) + (s2 + s3);
          }
        }
      }
      // butterfly reduce-scatter of the 16 row sums over the 32 lanes: lane l ends up with row ((l >> 1) & 15)
#define GTN_BFLY(OFF, NN)                                                                   \
      {                                                                                     \
        const bool up = (lane & (OFF)) != 0;                                                \
        _Pragma("unroll") for (int i = 0; i < (NN); ++i) {                                  \
          const float keep = up ? acc[i + (NN)] : acc[i];                                   \
          const float send = up ? acc[i] : acc[i + (NN)];                                   \
          acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, (OFF));                        \
        }                                                                                   \
      }
      GTN_BFLY(16, 8) GTN_BFLY(8, 4) GTN_BFLY(4, 2) GTN_BFLY(2, 1)
#undef GTN_BFLY
      acc[0] += __shfl_xor_sync(0xffffffffu, acc[0], 1);
      {
        // bits of the row index, most significant first, were chosen by lane bits 4, 3, 2, 1
        const int r = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
        const int jl = warp + 8 * r;
        if ((lane & 1) == 0 && jl < jn) sm.dw_s[j0 + jl] = acc[0];
      }
    } else if (active) {
      // d w_j over this CTA's positions: one warp per position, lanes over the time-varying columns
      constexpr int RD = 4;
      for (int jl0 = warp; jl0 < jn; jl0 += 8 * RD) {
        // RD positions per warp in flight (rows jl0, +8, ...): all loads first, then the reductions
        float s[RD];
#pragma unroll
        for (int r = 0; r < RD; ++r) s[r] = 0.f;
        for (int d = lane * 8; d < E; d += 256) {
          float mv[RD][8];
#pragma unroll
          for (int r = 0; r < RD; ++r) {
            const int j = j0 + jl0 + 8 * r;
            if (jl0 + 8 * r < jn && j < L) ld8b<MT>(mem + ((size_t)b * Tin + j) * E + d, mv[r]);
            else {
#pragma unroll
              for (int e = 0; e < 8; ++e) mv[r][e] = 0.f;
            }
          }
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float dc = sm.dctx_s[d + e];
#pragma unroll
            for (int r = 0; r < RD; ++r) s[r] += dc * mv[r][e];
          }
        }
#pragma unroll
        for (int r = 0; r < RD; ++r) {
          const float tot = warp_sum(s[r]);
          if (lane == 0 && jl0 + 8 * r < jn) sm.dw_s[j0 + jl0 + 8 * r] = tot;
        }
      }
    }
    if (mem_tile) mem_par ^= 1u;
    PROFB(23);
    __syncthreads();
    if (tiles.th && active && threadIdx.x == 0 && jn > 0) {
      // the memory tile has been consumed (generic-proxy reads, ordered by the barrier above): the tanh tile takes its place
      fence_proxy_async_smem();
      const uint32_t bytes = (uint32_t)jn * A * 4u;
      mbar_expect_tx(sm.th_full, bytes);
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem_u32(sm.ring.base + TH_TILE_OFF)), "l"(p.th_all + (tb * Tin + j0) * A), "r"(bytes),
                     "r"(smem_u32(sm.th_full)) : "memory");
    }
    if (active && (int)threadIdx.x < jn && j0 + (int)threadIdx.x < L) sm.dw_s[j0 + threadIdx.x] += dw_extra;   // (jn <= 128 < blockDim)
    __syncthreads();
    if (active) {
      for (int jl = threadIdx.x; jl < jn; jl += blockDim.x) dot_part += sm.w_s[j0 + jl] * sm.dw_s[j0 + jl];
    }
    PROFB(24);
    dot_part = block_sum2(dot_part, sm.scr);
    if (active && threadIdx.x == 0) {
      sm.dot_s[xb * 2] = dot_part;
      st_cluster_f1(map_to_rank(smem_u32(&sm.dot_s[xb * 2 + 1]), (uint32_t)peer), dot_part);
    }
    PROFB(11);
    cluster_sync_all();                                              // #1: softmax dot product of the pair
    PROFB(12);
    if (active) {
      const float dot = sm.dot_s[xb * 2] + sm.dot_s[xb * 2 + 1];
      // softmax backward: de_j = w_j (dw_j - sum_k w_k dw_k)
      for (int jl = threadIdx.x; jl < jn; jl += blockDim.x) {
        const int j = j0 + jl;
        sm.dw_s[j] = sm.w_s[j] * (sm.dw_s[j] - dot);
      }
    }
    __syncthreads();
    if (active) {
      // thread = (two adjacent attention columns, one of four row groups): th / dpm traffic is coalesced along the
      // attention dimension, dq[a] = sum_j dth[j][a] and dv[a] += sum_j de_j th[j][a] accumulate in registers over rows.
      // dth goes to shared memory as bf16 hi / lo planes: K-major tile rows = positions (block = a / 64), which is also
      // the MN-major operand of dU.  Rows beyond jn / masked positions are written as zeros.
      const int ap = (threadIdx.x & 63) * 2, rg = threadIdx.x >> 6;
      const bool a_ok = ap < A;
      const float v0 = sm.v_s[ap], v1 = sm.v_s[ap + 1];
      const uint32_t toff = (uint32_t)(ap >> 6) * W_TILE_BYTES + (uint32_t)((ap & 7) * 2);
      const int chunk = (ap & 63) >> 3;
      float dq0 = 0.f, dq1 = 0.f, dv0 = 0.f, dv1 = 0.f;
      PROFB(7);
      if (tiles.th && jn > 0) wait_bar(sm.th_full, tile_par & 1u);
      PROFB(8);
      // A thread's rows are jr + 8 k with jr = rg or rg + 4: (row & 7) is then fixed, so the swizzled tile offset, the tanh
      // tile, the d(energies) and the d(processed memory) addresses all advance by constants (this loop was a fifth of the
      // kernel's instructions).  Rows at or beyond nvalid (padding / the other half / beyond Tin) hold zeros.
      const int nvalid = max(0, min(jn, L - j0));
      const int row8 = 8 * A;
      // rows(TH, FULL): RW rows k0 .. k0 + RW - 1 of residue jr; TH = tanh tile in shared memory, FULL = all RW rows valid
      constexpr int RW = 5;
      auto rows = [&](auto TH, auto FULL, int k0, int nr, uint8_t* hi, uint8_t* lo, const float* th_g, uint32_t th_sa,
                      float* dpmp, const float* dep) {
        float2 th[RW];
        float de[RW];
#pragma unroll
        for (int r = 0; r < RW; ++r) {
          const bool ok = FULL.value || k0 + r < nr;
          th[r] = make_float2(0.f, 0.f);
          de[r] = 0.f;
          if (ok) {
            if (TH.value) {
              asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(th[r].x), "=f"(th[r].y) : "r"(th_sa));
            } else {
              th[r] = __ldg(reinterpret_cast<const float2*>(th_g));
            }
            de[r] = dep[r * 8];
          }
          th_sa += 8u * 16u;                 // (chunk-major tanh tile: 8 rows further = 8 x 16 bytes)
          th_g += 8 * 4;
        }
#pragma unroll
        for (int r = 0; r < RW; ++r) {
          const float e0 = de[r] * v0, e1 = de[r] * v1;
          const float d0 = e0 * (1.f - th[r].x * th[r].x), d1 = e1 * (1.f - th[r].y * th[r].y);
          // d processed_memory accumulates over the frames: a fire-and-forget L2 reduction instead of load + add + store
          // (this thread is the only writer of the address, frames are ordered by the grid barriers: same sums, no read)
          if (FULL.value || k0 + r < nr) red_add_f2(dpmp, d0, d1);
          dpmp += row8;
          dq0 += d0; dq1 += d1;
          dv0 += de[r] * th[r].x; dv1 += de[r] * th[r].y;
          const uint32_t hp = pack_bf16x2(d0, d1);
          const float h0 = __uint_as_float(hp << 16), h1 = __uint_as_float(hp & 0xFFFF0000u);
          if (FULL.value || k0 + r < TILE_M / 8) {
            *reinterpret_cast<uint32_t*>(hi + r * 1024) = hp;
            if (F32) *reinterpret_cast<uint32_t*>(lo + r * 1024) = pack_bf16x2(d0 - h0, d1 - h1);
          }
        }
      };
#pragma unroll 1
      for (int s = 0; s < 2; ++s) {
        const int jr = rg + 4 * s;
        const uint32_t sw = toff + (uint32_t)(jr * 128 + ((chunk ^ jr) << 4));
        const int nr = a_ok ? (nvalid - jr + 7) >> 3 : 0;              // this thread's valid rows: k < nr (warp-uniform but for a_ok)
#pragma unroll 1
        for (int k0 = 0; k0 < TILE_M / 8; k0 += RW) {
          uint8_t* hi = dth_hi + sw + k0 * 1024;
          uint8_t* lo = dth_lo + sw + k0 * 1024;
          const int row = jr + 8 * k0;
          // tanh tile of the half, chunk-major as the forward kernels save it: element (row, a) at ((a / 4) jn + row) 4 + a % 4
          const size_t th_el = ((size_t)(ap >> 2) * jn + row) * 4 + (ap & 3);
          const float* th_g = p.th_all + (tb * Tin + j0) * A + th_el;
          const uint32_t th_sa = smem_u32(sm.ring.base + TH_TILE_OFF) + (uint32_t)th_el * 4u;
          float* dpmp = p.dpm + (((size_t)b * Tin + j0 + row) * A + ap);
          const float* dep = sm.dw_s + j0 + row;
          if (k0 + RW <= nr) {
            if (tiles.th) rows(std::true_type{}, std::true_type{}, k0, nr, hi, lo, th_g, th_sa, dpmp, dep);
            else rows(std::false_type{}, std::true_type{}, k0, nr, hi, lo, th_g, th_sa, dpmp, dep);
          } else if (k0 < nr) {
            if (tiles.th) rows(std::true_type{}, std::false_type{}, k0, nr, hi, lo, th_g, th_sa, dpmp, dep);
            else rows(std::false_type{}, std::false_type{}, k0, nr, hi, lo, th_g, th_sa, dpmp, dep);
          } else {
#pragma unroll
            for (int r = 0; r < RW; ++r) {
              if (k0 + r < TILE_M / 8) {
                *reinterpret_cast<uint32_t*>(hi + r * 1024) = 0u;
                if (F32) *reinterpret_cast<uint32_t*>(lo + r * 1024) = 0u;
              }
            }
          }
        }
      }
      PROFB(9);
      fence_proxy_async_smem();
      // combine the four row groups: part_s[rg][0..127] (dq), then the same for dv
      sm.part_s[rg * 128 + ap] = dq0;
      sm.part_s[rg * 128 + ap + 1] = dq1;
      __syncthreads();
      float dqv = 0.f;
      if (threadIdx.x < 128) {
#pragma unroll
        for (int w = 0; w < 4; ++w) dqv += sm.part_s[w * 128 + threadIdx.x];
      }
      __syncthreads();
      sm.part_s[rg * 128 + ap] = dv0;
      sm.part_s[rg * 128 + ap + 1] = dv1;
      __syncthreads();
      if (threadIdx.x < 128) {
        const int a = threadIdx.x;
        float dvv = 0.f;
#pragma unroll
        for (int w = 0; w < 4; ++w) dvv += sm.part_s[w * 128 + a];
        sm.dv_acc[a] += dvv;
        // every CTA of the cluster receives every partial (slot = source rank): the cell backward is split over the cluster
#pragma unroll
        for (int r = 0; r < CL; ++r) st_cluster_f1(map_to_rank(smem_u32(&sm.dqp_s[rank * 128 + a]), (uint32_t)r), dqv);
      }
    }
    __syncthreads();
    PROFB(13);
    // warp 1 issues in warp-uniform control flow (one elected lane; see gemm_issue)
    if (active && __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0) == 1) {
      tc_fence_after();
      // G[j][kk] = sum_a dth[j][a] U[a][kk]   (M = positions, N = 64 filter taps, K = attention dim, 2 k-blocks)
      const uint32_t idesc = umma_idesc_bf16(TILE_M, 64);
      const uint32_t uts = smem_u32(sm.ut_s);
      const uint32_t a_hi = smem_u32(dth_hi), a_lo = smem_u32(dth_lo);
      const uint32_t acc_du = (uint32_t)(!(it == 0 && itb == 0));
      // dU[a][kk] += sum_j dth[j][a] im2col[j][kk]   (M = attention dim: 2 MN slices 16 KB apart, N = taps, K = positions)
      const uint32_t idesc_mn = umma_idesc_bf16(TILE_M, 64) | (1u << 15) | (1u << 16);
      const uint32_t b_hi = smem_u32(im_hi), b_lo = smem_u32(im_lo);
      if (elect_one()) {
#pragma unroll
        // bf16 mode: dth enters as ONE bf16 plane (like every other gradient operand of the mode: the gate gradients of P1, the
        // deferred weight-gradient GEMMs); the filter / im2col operands keep their residual planes.  fp32 mode: three products.
        for (int pl = 0; pl < 3; ++pl) {             // (dth_hi, U_hi), (dth_lo, U_hi), (dth_hi, U_lo)
          if (!F32 && pl == 1) continue;
          const uint32_t abase = pl == 1 ? a_lo : a_hi;
          const uint32_t bbase = uts + (pl == 2 ? 2u * 8192u : 0u);
#pragma unroll
          for (int kb = 0; kb < 2; ++kb) {
            const uint64_t da = umma_desc_sw128(abase + kb * W_TILE_BYTES), db = umma_desc_sw128(bbase + kb * 8192u);
#pragma unroll
            for (int k = 0; k < KB / 16; ++k)
              umma_bf16(tmem_g, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)((pl | kb | k) != 0));
          }
        }
        umma_commit(sm.g_full);
#pragma unroll
        for (int pl = 0; pl < 3; ++pl) {             // (dth_hi, im_hi), (dth_lo, im_hi), (dth_hi, im_lo)
          if (!F32 && pl == 1) continue;
          const uint32_t abase = pl == 1 ? a_lo : a_hi;
          const uint32_t bbase = pl == 2 ? b_lo : b_hi;
#pragma unroll
          for (int k = 0; k < TILE_M / 16; ++k) {
            const uint64_t da = umma_desc_sw128_mn_lbo(abase + k * 2048u, W_TILE_BYTES);
            const uint64_t db = umma_desc_sw128_mn_lbo(bbase + k * 2048u, W_TILE_BYTES);
            umma_bf16(tmem_du, da, db, idesc_mn, (pl | k) != 0 ? 1u : acc_du);
          }
        }
        // (dU itself is only read at the very end; the commit tells when its operand tiles in the ring may be overwritten)
        if (tiles.wq) umma_commit(sm.du_done);
      }
    }
    __syncwarp();
    if (active) {
      wait_bar(sm.g_full, g_par);
      g_par ^= 1u;
      tc_fence_after();
      PROFB(30);
      if (warp >= 4) {
        const int wq = warp & 3;
        const int jl = wq * 32 + lane;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t r[32];
          tmem_ld_32x32(tmem_g + ((uint32_t)(wq * 32) << 16) + (uint32_t)(h * 32), r);
          tmem_ld_wait();
#pragma unroll
          for (int c = 0; c < 32; ++c)
            if (h * 32 + c < 62) G_s[jl * 63 + h * 32 + c] = __uint_as_float(r[c]);
        }
      }
    }
    tc_fence_before();
    __syncthreads();
    if (tiles.wq && threadIdx.x == 0) {
      // this rank's quarter of Wq -> start of the ring (the dth / im2col tiles are dead once the dU MMAs have retired)
      if (active) wait_bar(sm.du_done, (tile_par >> 1) & 1u);
      fence_proxy_async_smem();
      const int Hq = ((p.Ha / CL + 1) / 2) * 2;
      mbar_expect_tx(sm.wq_full, (uint32_t)A * Hq * 2u);
      tma_load_3d(sm.ring.base, &maps.wq, sm.wq_full, rank * (Hq / 2), 0, 0);
    }
    PROFB(15);
    if (active) {
      // d wcat[ch][jj] = sum_k G[jj - k][ch*KS + k]  over this CTA's positions
      for (int i = threadIdx.x; i < 2 * Wd; i += blockDim.x) {
        const int ch = i / Wd, jj = i - ch * Wd;
        // G[jj - k][ch KS + k] for the k that keep the row inside [0, jn): a run of addresses 62 words apart
        const int klo = max(0, jj - jn + 1), khi = min(KS - 1, jj);
        const float* q = G_s + (jj - klo) * 63 + ch * KS + klo;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        int k = klo;
        for (; k + 3 <= khi; k += 4, q -= 4 * 62) { s0 += q[0]; s1 += q[-62]; s2 += q[-124]; s3 += q[-186]; }
        for (; k <= khi; ++k, q -= 62) s0 += q[0];
        sm.dwc_s[i] = (s0 + s1) + (s2 + s3);
      }
    }
    __syncthreads();
    if (active) {
      // positions outside this CTA's range belong to the peer: send them (window index jj <-> position j0 + jj - hk)
      float* hal = sm.halo_s + xb * 2 * 32;
      const uint32_t peer_h = map_to_rank(smem_u32(hal), (uint32_t)peer);
      for (int i = threadIdx.x; i < 2 * hk; i += blockDim.x) {
        const int ch = i / hk, q = i - ch * hk;
        // half 0 sends its upper halo (jj = hk + jn + q), half 1 its lower halo (jj = q)
        const int jj = half == 0 ? hk + jn + q : q;
        st_cluster_f1(peer_h + (uint32_t)((ch * 32 + q) * 4), sm.dwc_s[ch * Wd + jj]);
      }
    }
    PROFB(16);
    QuadOps qops;
    const int qb0 = cid * 2 + itb * npairs;
    const bool qpre = qb0 < p.B && quad_rounds(p, t) == 1;
    if (qpre) quad_cells_load(p, t, qb0, qb0 + 1, rank, 0, qops);
    cluster_sync_all();                                              // #2: halos and dq partials of the pair
    PROFB(17);
    if (active) {
      const float* hal = sm.halo_s + xb * 2 * 32;
      for (int i = threadIdx.x; i < 2 * jn; i += blockDim.x) {
        const int ch = i / jn, jl = i - ch * jn, j = j0 + jl;
        float s = sm.dwc_s[ch * Wd + hk + jl];
        // the peer's halo covers my first hk positions (I am half 1) or my last hk positions (I am half 0)
        if (half == 1 && jl < hk && Tin > JH) s += hal[ch * 32 + jl];
        if (half == 0) {
          const int q = jl - (jn - hk);            // peer's lower halo q <-> its position j0' + q - hk = my position jn - hk + q
          if (q >= 0 && q < hk && Tin > JH) s += hal[ch * 32 + q];
        }
        if (ch == 0) p.dwprev[(size_t)b * p.ldT + j] = s;
        else p.dcum[(size_t)b * p.ldT + j] += s;
      }
    }
    {
      // full d query of BOTH sequences of the cluster (pair 0: ranks 0, 1; pair 1: ranks 2, 3); the pair's first CTA saves its own
      const int sq = threadIdx.x >> 7, a = threadIdx.x & 127;
      const float s = a < A ? sm.dqp_s[(2 * sq) * 128 + a] + ((Tin > JH) ? sm.dqp_s[(2 * sq + 1) * 128 + a] : 0.f) : 0.f;
      sm.dqf_s[sq * 128 + a] = s;
      if (active && half == 0 && sq == (rank >> 1) && a < A) p.dq_all[tb * A + a] = s;
    }
    __syncthreads();
    PROFB(19);
    {
      const int b0 = cid * 2 + itb * npairs, b1 = b0 + 1;
      if (b0 < p.B) {
        quad_cells<F32>(p, sm.dqf_s, sm.dctx_s, t, it, b0, b1, rank,
                        tiles.wq ? reinterpret_cast<const uint32_t*>(sm.ring.base) : nullptr, sm.wq_full, (tile_par >> 2) & 1u,
                        qpre ? &qops : nullptr);
      } else if (tiles.wq) {
        wait_bar(sm.wq_full, (tile_par >> 2) & 1u);                      // the copy must not outlive the frame
      }
    }
    // parities of the three tile barriers: bit 0 tanh tile, bit 1 dU commit, bit 2 Wq tile (each completes once per frame when used)
    if (tiles.th && active && jn > 0) tile_par ^= 1u;
    if (tiles.wq && active) tile_par ^= 2u;
    if (tiles.wq) tile_par ^= 4u;
    if (tiles.mma && mem_tile && active) tile_par ^= 8u;
    __syncthreads();
  }
}

// The saves of the forward pass stream from HBM exactly once and their addresses are known a whole frame ahead: at the
// top of frame t every CTA asks L2 for what ITS P2 (attention backward of its sequence) and P3 (cell backward of its
// hidden units) will read, so that those phases see L2 latency instead of a dependent HBM round trip.
GTN_DEVICE void prefetch_l2(const void* ptr) { asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr)); }

GTN_DEVICE void prefetch_frame(const DecBwd2& p, int t, int rank, int cid, int ncl) {
  const int Tin = p.Tin, A = p.A;
  // ---- P2 operands of this CTA's sequence(s)
  const int half = rank & 1, pair = cid * 2 + (rank >> 1), npairs = ncl * 2;
  const int JH = (Tin + 1) / 2, j0 = half * JH, jn = min(Tin, j0 + JH) - j0;
  for (int b = pair; b < p.B; b += npairs) {
    const size_t tb = (size_t)t * p.B + b, bt = (size_t)b * p.T + t;
    const char* th = reinterpret_cast<const char*>(p.th_all + (tb * Tin + j0) * A);
    const int th_lines = (jn * A * 4 + 127) / 128;
    for (int i = threadIdx.x; i < th_lines; i += blockDim.x) prefetch_l2(th + (size_t)i * 128);
    const int row_lines = (Tin * 4 + 127) / 128 + 1;
    if ((int)threadIdx.x < row_lines) {
      prefetch_l2(reinterpret_cast<const char*>(p.align + bt * Tin) + threadIdx.x * 128);
      prefetch_l2(reinterpret_cast<const char*>(p.cum_all + tb * Tin) + threadIdx.x * 128);
      if (p.dalign) prefetch_l2(reinterpret_cast<const char*>(p.dalign + bt * Tin) + threadIdx.x * 128);
    }
    const int dp_lines = (p.E * 4 + 127) / 128 + 1;
    if ((int)threadIdx.x >= 32 && (int)threadIdx.x < 32 + dp_lines)
      prefetch_l2(reinterpret_cast<const char*>(p.dproj_all + tb * (p.Hd + p.D) + p.Hd) + (threadIdx.x - 32) * 128);
  }
  // ---- cell-backward operands of this CTA's half of the hidden units of its sequence(s): contiguous runs
  const int Hh_a = ((p.Ha / 2 + 1) / 2) * 2, Hh_d = ((p.Hd / 2 + 1) / 2) * 2;
  for (int b = pair; b < p.B; b += npairs) {
    const size_t ca = ((size_t)t * p.B + b) * p.Ha + half * Hh_a;
    const int tid = threadIdx.x;
    if (tid < 128) {
      if (tid * 8 < Hh_a) prefetch_l2(reinterpret_cast<const float4*>(p.gates_a) + ca + tid * 8);            // 8 cells x 16 B per line
      if (tid < 32 && tid * 32 < Hh_a && t > 0) prefetch_l2(p.c_a_all + ca - (size_t)p.B * p.Ha + tid * 32);
      if (tid >= 32 && tid < 64 && (tid - 32) * 32 < Hh_a) prefetch_l2(p.c_a_all + ca + (tid - 32) * 32);
    } else if (t > 0) {
      const int i = tid - 128;
      const size_t cd = ((size_t)(t - 1) * p.B + b) * p.Hd + half * Hh_d;
      if (i * 8 < Hh_d) prefetch_l2(reinterpret_cast<const float4*>(p.gates_d) + cd + i * 8);
      if (i < 32 && i * 32 < Hh_d && t > 1) prefetch_l2(p.c_d_all + cd - (size_t)p.B * p.Hd + i * 32);
      if (i >= 32 && i < 64 && (i - 32) * 32 < Hh_d) prefetch_l2(p.c_d_all + cd + (i - 32) * 32);
      if (i >= 64 && i < 96 && (i - 64) * 32 < Hh_d) prefetch_l2(p.dproj_all + ((size_t)(t - 1) * p.B + b) * (p.Hd + p.D) + half * Hh_d + (i - 64) * 32);
    }
  }
}

template <int N, bool F32>
__global__ void __launch_bounds__(256, 1) decoder_bwd2_kernel(const __grid_constant__ Bwd2Maps maps, const DecBwd2 p, const int S) {
  extern __shared__ uint8_t smem_raw[];
  // align by pointer arithmetic (no integer round trip) so that the compiler keeps the shared address space: LDS / STS, not generic LD / ST
  uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  SmemB sm = carveB<N>(base, S);
  const int warp = threadIdx.x >> 5;
  const int rank = (int)cluster_ctarank(), cid = (int)cluster_idx(), ncl = (int)cluster_count();
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(&sm.ring.full[s], 1); mbar_init(&sm.ring.empty[s], 1); }
    mbar_init(sm.acc_full, 1);
    mbar_init(sm.g_full, 1);
    mbar_init(sm.ut_full, 1);
    mbar_init(&sm.recv_full[0], 1);
    mbar_init(&sm.recv_full[1], 1);
    mbar_init(sm.mem_full, 1);
    mbar_init(sm.th_full, 1);
    mbar_init(sm.wq_full, 1);
    mbar_init(sm.du_done, 1);
    mbar_init(sm.dw_done, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc(sm.tmem_slot, 512);
  for (int i = threadIdx.x; i < 128; i += blockDim.x) { sm.v_s[i] = i < p.A ? p.v[i] : 0.f; sm.dv_acc[i] = 0.f; }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *sm.tmem_slot;
  const uint32_t tmem_g = tmem_base + 256u, tmem_du = tmem_base + 320u;
  if (threadIdx.x == 0) {
    // U^T hi / lo planes, two 64-wide k-blocks each: resident for the whole kernel
    mbar_expect_tx(sm.ut_full, 4 * 8192);
#pragma unroll
    for (int i = 0; i < 4; ++i) tma_load_3d(sm.ut_s + i * 8192, &maps.ut, sm.ut_full, i * KB, 0, 0);
    wait_bar(sm.ut_full, 0);
  }
  int res_n = 0;
  {
    // every cluster owns one 128-row tile of W_dec^T / W_att^T for the whole launch: park the leading weight k-blocks of
    // this CTA's K quarter in the free TMEM columns
    const int R_dec = p.Ha + p.E + p.Hd, R_att = p.E + p.Ha;
    const int nD = (R_dec + TILE_M - 1) / TILE_M, nA = (R_att + TILE_M - 1) / TILE_M;
    if (nD + nA <= ncl && cid < nD + nA) {
      const bool dec = cid < nD;
      const int H = dec ? p.Hd : p.Ha;
      const int nkb = (F32 ? 6 : 1) * 4 * H / KB, kb0 = rank * nkb / CL;
      res_n = min((rank + 1) * nkb / CL - kb0, res_slots<N>());
      load_resident<N>(dec ? p.wT_dec2 : p.wT_att2, (long long)nkb * KB, dec ? R_dec : R_att, (dec ? cid : cid - nD) * TILE_M, kb0,
                       res_n, 0, tmem_base);
    }
  }
  if (threadIdx.x < 8) {
    // encoder lengths of the sequences this CTA pair serves (constant for the launch; a global load per frame otherwise)
    const int b = cid * 2 + (rank >> 1) + (int)threadIdx.x * ncl * 2;
    reinterpret_cast<int*>(sm.scr + 48)[threadIdx.x] = b < p.B ? (p.lengths ? min(p.lengths[b], p.Tin) : p.Tin) : 0;
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();

  unsigned int epoch = 0;
  uint32_t acc_par = 0, g_par = 0, round = 0, xround = 0, mem_par = 0;
  const unsigned int nblocks = gridDim.x;
  // `fast` layout: one data-gradient tile per cluster and one sequence per CTA pair.  Then the part of the attention backward
  // that only reads the forward pass's saves runs between ARRIVE and WAIT of barrier 1, and (bf16 memory) the encoder-memory
  // tile of this CTA's positions -- 75 x 512 x 2 B, constant over the frames but too big to stay on chip next to the TMA ring --
  // is bulk-copied into the idle ring once per frame instead of being read row by row through L1 (dw: 7.2k -> ~1k cycles).
  const int R_dec_ = p.Ha + p.E + p.Hd, R_att_ = p.E + p.Ha;
  const bool fast = ((R_dec_ + TILE_M - 1) / TILE_M + (R_att_ + TILE_M - 1) / TILE_M) <= ncl && p.B <= 2 * ncl;
  const int fJH = (p.Tin + 1) / 2, fj0 = (rank & 1) * fJH, fjn = min(p.Tin, fj0 + fJH) - fj0;
  const int fb = cid * 2 + (rank >> 1);
  const bool factive = fb < p.B;
  const bool mem_tile_bulk = !F32 && fast && (p.E % 8) == 0 &&
                             MEM_TILE_OFF + (uint32_t)fJH * p.E * 2u + 2u * W_TILE_BYTES <= sm.ring_bytes;
  uint8_t* const f_im = sm.ring.base + sm.ring_bytes - 2 * W_TILE_BYTES;
  P2Tiles tiles;
  tiles.th = fast && (p.A % 4) == 0 && TH_TILE_OFF + (uint32_t)fJH * p.A * 4u + 2u * W_TILE_BYTES <= sm.ring_bytes;
  tiles.wq = !F32 && fast && (p.flags & 4) != 0 && (uint32_t)p.A * (((p.Ha / CL + 1) / 2) * 2) * 2u <= sm.ring_bytes;
  const int mem_box1 = ((fJH + 7) / 8) * 8;
  tiles.mma = !F32 && fast && (p.flags & 8) != 0 && (p.E % KB) == 0 &&
              MEM_TILE_OFF + (uint32_t)(p.E / KB) * mem_box1 * 128u + 2u * W_TILE_BYTES <= sm.ring_bytes;
  const bool mem_tile = mem_tile_bulk || tiles.mma;
  uint32_t tile_par = 0;
  dec_cell_prologue<F32>(p, p.T - 1);
  grid_sync2(p.barrier, epoch, nblocks);
  int it = 0;
  for (int t = p.T - 1; t >= 0; --t, ++it) {
    PROFB(0);
    if (p.flags & 1) prefetch_frame(p, t, rank, cid, ncl);
    // (this cluster runs a P1 tile this frame <=> its idle warps can fetch the attention window there)
    const bool win_p1 = fast && factive &&
                        cid < (R_dec_ + TILE_M - 1) / TILE_M + (t + 1 < p.T ? (R_att_ + TILE_M - 1) / TILE_M : 0);
    dgrad_phase2<N, F32>(p, maps, sm, t, it, rank, cid, ncl, tmem_base, acc_par, round, res_n, win_p1 ? fb : -1, fj0, fjn);
    PROFB(4);
    if (fast) {
      grid_arrive2(p.barrier, /*proxy_fence=*/false, (p.prof != nullptr && blockIdx.x == 0 && it < 16) ? p.prof + it * 32 + 28 : nullptr);
      epoch += nblocks;
      if (tiles.mma && factive && fjn > 0) {
        // the memory tile as E / 64 SWIZZLE_128B k-block tiles of mem_box1 rows (warp 0, one elected lane, uniform control flow)
        if (__shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0) == 0) {
          if (elect_one()) {
            mbar_expect_tx(sm.mem_full, (uint32_t)(p.E / KB) * mem_box1 * 128u);
            for (int kb = 0; kb < p.E / KB; ++kb)
              tma_load_3d(sm.ring.base + MEM_TILE_OFF + (size_t)kb * mem_box1 * 128, &maps.mem, sm.mem_full, kb * KB, fj0, fb);
          }
          __syncwarp();
        }
      } else if (mem_tile && factive && threadIdx.x == 0 && fjn > 0) {
        // (P1's MMAs have drained the ring: every thread passed the exchange's syncs behind the accumulator wait)
        const uint32_t bytes = (uint32_t)fjn * p.E * 2u;
        mbar_expect_tx(sm.mem_full, bytes);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(sm.ring.base + MEM_TILE_OFF)),
                       "l"(reinterpret_cast<const bf16*>(p.mem_enc) + ((size_t)fb * p.Tin + fj0) * p.E), "r"(bytes),
                       "r"(smem_u32(sm.mem_full)) : "memory");
      }
      attn_bwd_prepare(p, sm, t, fb, factive, fj0, fjn, f_im, f_im + W_TILE_BYTES, /*window_loaded=*/win_p1);
      PROFB(6);
      grid_wait2(p.barrier, epoch);
    } else {
      grid_sync2(p.barrier, epoch, nblocks);
    }
    PROFB(5);
    attn_bwd_phase2<F32>(p, maps, sm, t, it, rank, cid, ncl, tmem_g, tmem_du, g_par, xround, fast, !F32 && mem_tile && fjn > 0, mem_par,
                         tiles, tile_par);
    PROFB(20);
    grid_sync2(p.barrier, epoch, nblocks);
    PROFB(21);
  }
  // per-CTA partial sums of the small attention weight gradients
  for (int i = threadIdx.x; i < p.A; i += blockDim.x) p.dv_part[(size_t)blockIdx.x * p.A + i] = sm.dv_acc[i];
  {
    // dU accumulator [128 lanes = attention dim][64 taps] -> dU_part[cta][a][kk]; make sure the last MMAs retired
    if (__shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0) == 1 && elect_one()) umma_commit(sm.g_full);
    __syncwarp();
    const int pair = cid * 2 + (rank >> 1);
    const bool used = pair < p.B;                     // CTAs that never ran an attention step hold no accumulator
    wait_bar(sm.g_full, g_par);
    tc_fence_after();
    if (warp >= 4) {
      const int wq = warp & 3, lane = threadIdx.x & 31;
      const int a = wq * 32 + lane;
      float* out = p.dU_part + ((size_t)blockIdx.x * TILE_M + a) * 64;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_du + ((uint32_t)(wq * 32) << 16) + (uint32_t)(h * 32), r);
        tmem_ld_wait();
#pragma unroll
        for (int c = 0; c < 32; ++c) out[h * 32 + c] = used ? __uint_as_float(r[c]) : 0.f;
      }
    }
  }
  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
int make_tma_map_3d(CUtensorMap* out, const void* base, long long d0, long long d1, long long d2, long long stride1_elems,
                    long long stride2_elems, int box1);
int make_tma_map_f32_3d(CUtensorMap* out, const void* base, long long d0, long long d1, long long d2, long long stride1_elems,
                        long long stride2_elems, int box0, int box1, int box2);

template <int N, bool F32>
static int bwd2_config(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attrs, int& S, size_t& smem) {
  const size_t stage = (size_t)W_TILE_BYTES + (size_t)N * KB * 2;
  const size_t fixed = bwd2_fixed_bytes<N>() + 1024;
  S = (int)((227 * 1024 - fixed) / stage);
  if (S > 12) S = 12;
  GTN_REQUIRE(S >= 2, "decoder backward: not enough shared memory for the TMA ring (N=%d)", N);
  size_t ring_b = (size_t)S * stage;
  if (ring_b < (size_t)P2_RING_BYTES) ring_b = P2_RING_BYTES;
  smem = fixed + ring_b;
  GTN_REQUIRE(smem <= 227 * 1024, "decoder backward: shared memory budget exceeded (%zu B, N=%d)", smem, N);
  GTN_CUDA_OK(cudaFuncSetAttribute(decoder_bwd2_kernel<N, F32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cfg = cudaLaunchConfig_t{};
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = smem;
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = CL; attrs[0].val.clusterDim.y = 1; attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeCooperative;
  attrs[1].val.cooperative = 1;
  cfg.attrs = attrs;
  return 0;
}

template <int N>
static int bwd2_grid(int want, int* ncl_out) {
  cudaLaunchConfig_t cfg; cudaLaunchAttribute attrs[2]; int S; size_t smem;
  int rc = bwd2_config<N, false>(cfg, attrs, S, smem);
  if (rc) return rc;
  cfg.numAttrs = 1;
  cfg.gridDim = dim3(CL * want);
  int max_cl = 0;
  GTN_CUDA_OK(cudaOccupancyMaxActiveClusters(&max_cl, decoder_bwd2_kernel<N, false>, &cfg));
  GTN_REQUIRE(max_cl >= 1, "decoder backward: no cluster of %d CTAs fits on this device", CL);
  *ncl_out = max_cl < want ? max_cl : want;
  return 0;
}

template <int N, bool F32>
static int launch_bwd2(const Bwd2Maps& maps, const DecBwd2& p, int ncl, cudaStream_t st) {
  cudaLaunchConfig_t cfg; cudaLaunchAttribute attrs[2]; int S; size_t smem;
  int rc = bwd2_config<N, F32>(cfg, attrs, S, smem);
  if (rc) return rc;
  cfg.stream = st;
  cfg.numAttrs = profiler_attached() ? 1 : 2;   // profilers that cannot replay cooperative launches
  // each CTA pair stages the encoder lengths of at most 8 sequences (scr + 48): B <= 16 * clusters
  GTN_REQUIRE((p.B + 2 * ncl - 1) / (2 * ncl) <= 8, "decoder backward: batch %d needs more than 8 sequences per CTA pair on %d clusters",
              p.B, ncl);
  cfg.gridDim = dim3(CL * ncl);
  GTN_CUDA_OK(cudaLaunchKernelEx(&cfg, decoder_bwd2_kernel<N, F32>, maps, p, S));
  count_launch();
  return 0;
}

static int pick_n(int B) { return B <= 32 ? 32 : (B <= 64 ? 64 : (B <= 128 ? 128 : 256)); }

}  // namespace gtn

using namespace gtn;

extern "C" int gtn_decoder_bwd2_grid(int B, int Ha, int Hd, int E) {
  const int want0 = ((Ha + E + Hd + TILE_M - 1) / TILE_M) + ((E + Ha + TILE_M - 1) / TILE_M);
  const int want = want0 > 32 ? 32 : want0;
  int ncl = 0, rc;
  const int N = pick_n(B);
  if (N == 32) rc = bwd2_grid<32>(want, &ncl);
  else if (N == 64) rc = bwd2_grid<64>(want, &ncl);
  else if (N == 128) rc = bwd2_grid<128>(want, &ncl);
  else rc = bwd2_grid<256>(want, &ncl);
  return rc ? -1 : CL * ncl;
}

extern "C" int gtn_decoder_backward2(const gtn_decoder_bwd2_params* hp, void* stream) {
  GTN_REQUIRE(hp != nullptr, "null params");
  DecBwd2 p = *hp;
  GTN_REQUIRE(p.prec == 1 || p.prec == 3, "prec must be 1 or 3");
  p.mult = p.prec == 1 ? 1 : 6;
  GTN_REQUIRE(p.Ha % 32 == 0 && p.Hd % 32 == 0 && p.A <= 128 && p.A % 8 == 0, "unsupported dims");
  GTN_REQUIRE(p.E % 8 == 0 && p.E <= 1024 && p.Ep >= p.E && p.D >= p.E, "E must be a multiple of 8, <= 1024");
  GTN_REQUIRE(p.KS % 2 == 1 && 2 * p.KS <= 64, "location kernel size must be odd and <= 31");
  GTN_REQUIRE(p.Tin >= 1 && p.Tin <= B2_TIN_MAX && p.T >= 1 && p.B >= 1 && p.B <= 256, "bad dims");
  auto thr = [](float pr) -> unsigned { double t = (double)pr * 4294967296.0; return t >= 4294967295.0 ? 0xFFFFFFFFu : (unsigned)t; };
  p.thr_att = p.use_dropout ? thr(p.p_att) : 0u;
  p.thr_dec = p.use_dropout ? thr(p.p_dec) : 0u;
  const int grid = gtn_decoder_bwd2_grid(p.B, p.Ha, p.Hd, p.E);
  GTN_REQUIRE(grid > 0 && p.grid == grid, "params.grid (%d) must equal gtn_decoder_bwd2_grid() (%d)", p.grid, grid);
  GTN_REQUIRE(p.wq_au != nullptr, "wq_au pack missing");
  GTN_REQUIRE(p.dhd_rec != nullptr, "dhd_rec workspace missing");
  const int N = pick_n(p.B);
  const int m = p.mult;
  Bwd2Maps maps;
  memset(&maps, 0, sizeof(maps));
  int rc = 0;
  rc |= make_tma_map_3d(&maps.wT_att, p.wT_att2, (long long)m * 4 * p.Ha, p.E + p.Ha, 1, (long long)m * 4 * p.Ha, 0, TILE_M);
  rc |= make_tma_map_3d(&maps.wT_dec, p.wT_dec2, (long long)m * 4 * p.Hd, p.Ha + p.E + p.Hd, 1, (long long)m * 4 * p.Hd, 0, TILE_M);
  rc |= make_tma_map_3d(&maps.ut, p.ut_loc, 4 * KB, 64, 1, 4 * KB, 0, 64);
  rc |= make_tma_map_3d(&maps.xdg_a, p.xdg_a, (long long)m * 4 * p.Ha, p.B, 1, (long long)m * 4 * p.Ha, 0, N);
  rc |= make_tma_map_3d(&maps.xdg_d, p.xdg_d, (long long)m * 4 * p.Hd, p.B, 2, (long long)m * 4 * p.Hd, (long long)p.B * m * 4 * p.Hd, N);
  if (rc) return rc;
  {
    // Wq quarter box (bf16 pairs as fp32 words): one rank's units x all A rows, see TH_TILE_OFF
    const int Hq = ((p.Ha / CL + 1) / 2) * 2;
    p.flags &= ~4;
    if (p.prec == 1 && Hq % 8 == 0 && Hq <= 256 && p.Ha % 8 == 0) {
      rc = make_tma_map_f32_3d(&maps.wq, p.wq_au, p.Ha / 2, p.A, 1, p.Ha / 2, (long long)(p.Ha / 2) * p.A, Hq / 2, p.A, 1);
      if (rc) return rc;
      p.flags |= 4;
    }
    // encoder-memory k-block tiles for dw on the tensor core (bf16 memory only)
    p.flags &= ~8;
    const int box1 = ((((p.Tin + 1) / 2) + 7) / 8) * 8;
    if (p.prec == 1 && p.mem_bf16 && p.E % KB == 0 && box1 <= 256) {
      rc = make_tma_map_3d(&maps.mem, p.mem_enc, p.E, p.Tin, p.B, p.E, (long long)p.Tin * p.E, box1);
      if (rc) return rc;
      p.flags |= 8;
    }
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int ncl = grid / CL;
  // precision is a compile-time property of the kernels (half the code: the per-frame path is instruction-cache cold every frame)
  GTN_REQUIRE((p.prec == 3) == (p.mem_bf16 == 0) && ((p.flags & 2) != 0) == (p.prec == 1),
              "decoder backward: fp32 mode <-> fp32 encoder memory and fp32 gate-gradient saves");
  if (p.prec == 3) {
    if (N == 32) return launch_bwd2<32, true>(maps, p, ncl, st);
    if (N == 64) return launch_bwd2<64, true>(maps, p, ncl, st);
    if (N == 128) return launch_bwd2<128, true>(maps, p, ncl, st);
    return launch_bwd2<256, true>(maps, p, ncl, st);
  }
  if (N == 32) return launch_bwd2<32, false>(maps, p, ncl, st);
  if (N == 64) return launch_bwd2<64, false>(maps, p, ncl, st);
  if (N == 128) return launch_bwd2<128, false>(maps, p, ncl, st);
  return launch_bwd2<256, false>(maps, p, ncl, st);
}
