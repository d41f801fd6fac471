// Persistent clustered decoder kernel, forward (second generation).  ONE launch runs every frame of
//   * Decoder.inference  (reference model.py:456-497): prenet -> decode() -> gate test, autoregressive
//   * Decoder.forward    (reference model.py:419-454): the teacher-forced loop of decode() calls
// decode() is model.py:378-417.  See gtn_decoder2.cuh for the cluster-split tcgen05 GEMM engine.
//
// Grid = NC clusters of 4 CTAs (one CTA per SM, all co-resident).  Phases of a frame:
//   [infer] P1, P2  prenet layers (relu + always-on dropout, model.py:102-105) as 128-row GEMM tiles
//   A   attention LSTM: cluster c owns hidden units [32c, 32c+32) (all four gates = one 128-row tile of
//       W_att, rows permuted gate-major per tile); fused cell + dropout; partial attention query
//       Wq[:, units] . h_att' for its 32 units                                   (model.py:391-394, 62)
//   B   location-sensitive attention, one CTA PAIR per sequence (encoder positions split in halves):
//       the location conv (2->F, k) and dense (F->A) are pre-multiplied into one 128 x 64 filter U and run
//       as a tcgen05 GEMM over an im2col tile of [prev | cumulative] weights (hi/lo bf16 planes, fp32
//       accumulate in TMEM), e = v.tanh(q + loc + processed_memory), energies exchanged through DSMEM,
//       masked softmax, context over the time-varying memory columns (split by column halves), cum += w
//                                                                                (model.py:27-31,62-89,403)
//   C   decoder LSTM (model.py:404-409)
//   [infer] D  mel/gate projection + stop test (model.py:411-416,486)
// Grid barriers: after A, after B (training: C(t) and A(t+1) share an interval since A does not read h_dec).
// Training additionally saves what the backward kernel needs.  The time-constant columns of `memory`
// (noise / speaker / emotion: model.py:275-278,655-663) are not re-read per frame: context = const * sum_j w_j.
#include "gtn_decoder_fwd.cuh"

namespace gtn {

// FAST = the kernel instantiation for batched bf16 inference on the cluster-local tail + all-at-once attention paths (the host
// checks the conditions): everything else -- training saves, the one-at-a-time attention, the grid-wide prenet / projection
// phases, the small-batch prenet -- is compiled out.  Half the code, 5 % faster: this kernel's per-frame path is several
// times the instruction cache, and unexecuted code still spreads it out.
#define P_TRAIN (!FAST && p.train)
#define GTN_HD __host__ __device__ __forceinline__

#define PROF(slot)                                                                          \
  do {                                                                                      \
    if (p.prof != nullptr && blockIdx.x == 0 && threadIdx.x == 0 && t < 16) p.prof[t * 32 + (slot)] = clock64(); \
  } while (0)
#define PROF_T(thr, slot)                                                                   \
  do {                                                                                      \
    if (p.prof != nullptr && blockIdx.x == 0 && threadIdx.x == (thr) && t < 16) p.prof[t * 32 + (slot)] = clock64(); \
  } while (0)

enum { K_PRE1 = 0, K_PRE2 = 1, K_ATT = 2, K_DEC = 3, K_PROJ = 4 };

struct Smem2 {
  Ring ring;
  uint8_t* u_s;        // [2][16 KB]  U_hi, U_lo (UMMA B operand of the location GEMM)
  float* recv;         // [2][4][128][NC4]
  float* hs;           // [32][NC4]
  float* q_s;          // [A]
  float* v_s;          // [A]
  float* wcat;         // [2][JH + KS - 1]
  float* e_buf;        // [2][Tin_pad]
  float* e_part;       // [2][128]
  float* w_s;          // [Tin_pad]
  float* red_s;        // [8][256]
  float* scr;          // [64]
  uint64_t* acc_full;
  uint64_t* att_full;
  uint64_t* u_full;
  uint64_t* recv_full;  // [2] bulk-copy exchange: one per receive buffer
  uint32_t* tmem_slot;
  int res_a, res_d;    // weight k-blocks of this CTA's attention / decoder LSTM K quarter that are resident in TMEM
  // batched attention of autoregressive inference (attention_infer_multi): up to AM sequences per CTA pair at once
  float* am_wcat;      // [AM][2][JH + KS - 1]
  float* am_ebuf;      // [AM][Tin_pad]   (the peer CTA writes its half)
  uint64_t* am_att_full;   // [AM]
  uint64_t* am_mem_full;   // [3]
};

// sequences per CTA pair the batched attention holds at once (accumulators: tensor-memory columns 256, 0, 128, 384)
template <int N>
__host__ __device__ constexpr int am_max() { return N >= 256 ? 4 : (N >= 128 ? 2 : 1); }
static constexpr int AM_WCAT = 2 * (JH_MAX + KS_MAX - 1);

template <int N>
__host__ __device__ constexpr size_t fwd2_fixed_bytes() {
  constexpr int NC4 = (N < 64 ? N : 64) / 4;
  return (size_t)2 * W_TILE_BYTES                        // u_s
         + (size_t)2 * 4 * TILE_M * NC4 * 4               // recv
         + (size_t)(NC4 * 32 + 128 + 128 + 2 * (JH_MAX + KS_MAX - 1) + 2 * TIN_MAX + 2 * 128 + TIN_MAX + 8 * 256 + 64) * 4
         + (size_t)am_max<N>() * (AM_WCAT + TIN_MAX) * 4  // am_wcat, am_ebuf
         + 64 * 8 + 64;                                   // barriers + tmem slot
}

template <int N>
GTN_DEVICE Smem2 carve2(uint8_t* base, int S) {
  constexpr int NC4 = (N < 64 ? N : 64) / 4;
  Smem2 m;
  m.ring.base = base;
  m.ring.S = S;
  m.ring.stage_bytes = W_TILE_BYTES + N * KB * 2;
  m.ring.prod_it = m.ring.mma_it = 0;
  m.res_a = m.res_d = 0;
  uint8_t* p = base + (size_t)S * m.ring.stage_bytes;
  m.u_s = p; p += 2 * W_TILE_BYTES;
  float* f = reinterpret_cast<float*>(p);
  m.recv = f; f += 2 * 4 * TILE_M * NC4;
  m.hs = f; f += NC4 * 32;
  m.q_s = f; f += 128;
  m.v_s = f; f += 128;
  m.wcat = f; f += 2 * (JH_MAX + KS_MAX - 1);
  m.e_buf = f; f += 2 * TIN_MAX;
  m.e_part = f; f += 2 * 128;
  m.w_s = f; f += TIN_MAX;
  m.red_s = f; f += 8 * 256;
  m.scr = f; f += 64;
  m.am_wcat = f; f += am_max<N>() * AM_WCAT;
  m.am_ebuf = f; f += am_max<N>() * TIN_MAX;
  uint64_t* b = reinterpret_cast<uint64_t*>(f);
  m.ring.full = b; b += 16;
  m.ring.empty = b; b += 16;
  m.acc_full = b++;
  m.att_full = b++;
  m.u_full = b++;
  m.recv_full = b; b += 2;
  m.am_att_full = b; b += 4;
  m.am_mem_full = b; b += 3;
  m.tmem_slot = reinterpret_cast<uint32_t*>(b + 1);
  return m;
}

// ------------------------------------------------------------------------------------------------
// GEMM phases: prenet layers, the two LSTM cells, the mel/gate projection
// ------------------------------------------------------------------------------------------------
// Operands of the GEMM phase `kind` of frame t; returns the number of 128-row tiles.
template <bool FAST>
GTN_DEVICE int make_job(const DecFwd& p, const Fwd2Maps& maps, int kind, int t, GemmJob& job) {
  const int mult = p.mult, cur = t & 1, nxt = cur ^ 1;
  int ntiles;
  switch (kind) {
    case K_PRE1:
      job.wmap = &maps.w_pre1; ntiles = (p.P + TILE_M - 1) / TILE_M; job.nseg = 1;
      job.seg[0] = {&maps.xmel, mult * p.NMp / KB, 0};
      break;
    case K_PRE2:
      job.wmap = &maps.w_pre2; ntiles = (p.P + TILE_M - 1) / TILE_M; job.nseg = 1;
      job.seg[0] = {&maps.xp1, mult * p.Pp / KB, 0};
      break;
    case K_ATT:
      job.wmap = &maps.w_att; ntiles = p.Ha / 32; job.nseg = 3;
      job.seg[0] = P_TRAIN ? XSeg{&maps.pre, mult * p.Pp / KB, t} : XSeg{&maps.xp2, mult * p.Pp / KB, 0};
      job.seg[1] = {&maps.ctx, mult * p.Ep / KB, 0};
      job.seg[2] = {&maps.ha, mult * p.Hap / KB, cur};
      break;
    case K_DEC:
      job.wmap = &maps.w_dec; ntiles = p.Hd / 32; job.nseg = 3;
      job.seg[0] = {&maps.ha, mult * p.Hap / KB, nxt};
      job.seg[1] = {&maps.ctx, mult * p.Ep / KB, 0};
      job.seg[2] = {&maps.hd, mult * p.Hdp / KB, cur};
      break;
    default:
      job.wmap = &maps.w_proj; ntiles = 1; job.nseg = 2;
      job.seg[0] = {&maps.hd, mult * p.Hdp / KB, nxt};
      job.seg[1] = {&maps.ctx, mult * p.Ep / KB, 0};
      break;
  }
  job.nkb = job.seg[0].nkb + (job.nseg > 1 ? job.seg[1].nkb : 0) + (job.nseg > 2 ? job.seg[2].nkb : 0);
  return ntiles;
}

// Prenet layers and mel / gate projection of autoregressive inference at B > 64, BATCH-CHUNKED: these phases have only 2
// (prenet: 256 rows) and 1 (projection: 81 rows) row tiles, so with one tile per cluster 30 / 31 of the 32 clusters idle while
// one cluster streams a 128 x 256 tile (28k / 54k cycles per phase at B = 256).  Here a work item is (row tile, 64 sequences):
// MMA N = 64 on 64-row boxes of the same activation buffers, one split-K exchange and one 128 x 16 epilogue per CTA and item.
// The stop decision moves behind the phase's grid barrier (every CTA evaluates done[] itself, see the main loop).
GTN_DEVICE void gemm_phase_chunked(const DecFwd& p, const Fwd2Maps& maps, Smem2& sm, int kind, int t, int rank, int cid, int ncl,
                                   uint32_t tmem_base, uint32_t& acc_par, uint32_t& round, float* recv_base) {
  constexpr int CH = 64, NC4 = CH / 4;
  const int mult = p.mult;
  GemmJob job;
  const int ntiles = make_job<false>(p, maps, kind, t, job);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    if (i < job.nseg) { const CUtensorMap* mp = job.seg[i].map; job.seg[i].map = mp == &maps.xmel ? &maps.xmel_c : (mp == &maps.xp1 ? &maps.xp1_c : (mp == &maps.hd ? &maps.hd_c : &maps.ctx_c)); }
  }
  job.res_n = 0;
  const int nch = (p.B + CH - 1) / CH;
  const int rows = kind == K_PROJ ? p.NM + 1 : p.P;
  for (int item = cid; item < ntiles * nch; item += ncl) {
    const int tile = item / nch, chunk = item - tile * nch;
    job.row0 = tile * TILE_M;
    job.x_row0 = chunk * CH;
    const bool has_work = gemm_issue<CH>(job, sm.ring, tmem_base, sm.acc_full, rank, tmem_base);
    __syncwarp();
    if ((threadIdx.x >> 5) >= 4 && has_work) {
      wait_bar(sm.acc_full, acc_par);
      tc_fence_after();
    }
    if (has_work) acc_par ^= 1u;
    float* recv = recv_base + (size_t)(round & 1u) * 4 * TILE_M * NC4;
    const uint32_t xr = round;
    ++round;
    float* stage = recv_base + (size_t)((xr & 1u) ^ 1u) * 4 * TILE_M * NC4;
    exchange_bulk<CH>(has_work, tmem_base, 0, recv, stage, &sm.recv_full[xr & 1u], (xr >> 1) & 1u, rank);
    const int bbase = chunk * CH + rank * NC4;
    for (int idx = threadIdx.x; idx < TILE_M * NC4; idx += blockDim.x) {
      const int bl = idx % NC4, r = idx / NC4, b = bbase + bl, row = tile * TILE_M + r;
      if (b >= p.B || row >= rows) continue;
      float val = recv_sum<NC4, true>(recv, r, bl);
      if (kind != K_PROJ) {
        val = fmaxf(val, 0.f);
        const uint32_t bits = philox_bits(p.seed, kind == K_PRE1 ? 1u : 2u, (uint32_t)t, (uint32_t)(b * p.P + row));
        val = bits >= 0x80000000u ? val * 2.0f : 0.f;     // p = 0.5, always on (model.py:104)
        store_act((kind == K_PRE1 ? BFP(p.xp1) : BFP(p.xp2)) + (size_t)b * mult * p.Pp, row, p.Pp, p.prec, val);
      } else {
        val += p.pbias[(size_t)b * (p.NM + 1) + row];
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
}

template <int N, bool FAST>
GTN_DEVICE void gemm_phase(const DecFwd& p, const Fwd2Maps& maps, Smem2& sm, int kind, int t, int rank, int cid, int ncl,
                           uint32_t tmem_base, uint32_t& acc_par, uint32_t& round) {
  constexpr int CH = N < 64 ? N : 64;
  constexpr int NC4 = CH / 4;
  if constexpr (N > 64 && !FAST) {
    if (kind != K_ATT && kind != K_DEC) {
      gemm_phase_chunked(p, maps, sm, kind, t, rank, cid, ncl, tmem_base, acc_par, round, sm.recv);
      return;
    }
  }
  const int mult = p.mult, cur = t & 1, nxt = cur ^ 1;
  GemmJob job;
  const int ntiles = make_job<FAST>(p, maps, kind, t, job);
  const bool is_lstm = FAST || kind == K_ATT || kind == K_DEC;
  const bool is_att = kind == K_ATT;
  const int H = is_att ? p.Ha : p.Hd, Hp = is_att ? p.Hap : p.Hdp;

  for (int tile = cid; tile < ntiles; tile += ncl) {
    job.row0 = tile * TILE_M;
    const int ps = kind == K_ATT ? 1 : (kind == K_DEC ? 9 : (kind == K_PROJ ? 14 : 20));
    if (tile == cid && kind == K_ATT) { job.res_n = sm.res_a; job.res_slot0 = 0; }
    else if (tile == cid && kind == K_DEC) { job.res_n = sm.res_d; job.res_slot0 = sm.res_a; }
    else job.res_n = 0;
    // batched decode at B > 64: weights + encoder memory + processed memory (~90 MB per frame) do not all stay in L2 (ncu: 50 %
    // hit rate, 68 MB per frame from DRAM); the weights are kept (evict_last), the memory chunks stream (evict_first)
    const bool has_work = gemm_issue<N>(job, sm.ring, tmem_base, sm.acc_full, rank, tmem_base,
                                        (FAST && N > 64 && (p.reserved0 & 8) == 0) ? l2_policy_evict_last() : 0ull);
    PROF(ps);                                    // producer: all TMA loads issued
    __syncwarp();
    float wqr[32];
    if (is_att) {
      // query-weight column of this thread (a = tid % 128) for the tile's 32 units: Wq[a][32 tile + u].  Fetched every frame
      // underneath the accumulator wait (L1 / L2 hits): held across the frame it costs 32 registers in every other phase,
      // and this kernel's spills go to L2 (the shared-memory carve-out leaves the local-memory window almost no L1)
      const int a = threadIdx.x & 127;
#pragma unroll
      for (int u = 0; u < 32; ++u) wqr[u] = a < p.A ? __ldg(&p.wq_t[((size_t)tile * 32 + u) * p.A + a]) : 0.f;
    }
    if ((threadIdx.x >> 5) >= 4 && has_work) {
      wait_bar(sm.acc_full, acc_par);
      tc_fence_after();
    }
    PROF_T(128, ps + 1);                         // epilogue warp: accumulator complete
    if (has_work) acc_par ^= 1u;
    // operands of the fused epilogue that do not depend on the GEMM are fetched while the exchange is in flight
    constexpr int CPT = 32 * NC4 / 256;          // LSTM cells per thread (NC4 = 8 or 16)
    float4 gb[CPT];
    float cprev[CPT];
    float* cst = is_att ? p.c_a : p.c_d;
    auto prefetch = [&](int c0) {
      const int bbase = c0 + rank * NC4;
      if (is_lstm) {
        const float4* gbias = reinterpret_cast<const float4*>(is_att ? p.gbias_a : p.gbias_d);
#pragma unroll
        for (int q = 0; q < CPT; ++q) {
          const int idx = threadIdx.x + q * 256;
          const int ul = idx & 31, bl = idx >> 5, b = bbase + bl, unit = tile * 32 + ul;   // lanes run over hidden units
          gb[q] = make_float4(0.f, 0.f, 0.f, 0.f);
          cprev[q] = 0.f;
          if (b < p.B) {
            // frame 0 of the attention LSTM sees the zero initial context (model.py:314-316): no constant-column term
            if (is_att && t == 0) gb[q] = make_float4(p.b_att[unit], p.b_att[H + unit], p.b_att[2 * H + unit], p.b_att[3 * H + unit]);
            else gb[q] = __ldg(&gbias[(size_t)b * H + unit]);
            cprev[q] = cst[(size_t)b * H + unit];
          }
        }
      }
    };
    // PIPE (batched inference, more than one 64-column chunk per tile): chunk c + 1 is on its way while the cells of chunk c
    // run.  The sends are staged in the idle TMA ring (the accumulator is complete, so every stage has been consumed)
    // instead of the other receive buffer; a cluster barrier behind the cells of chunk c tells every CTA that receive buffer
    // c & 1 and staging area c & 1 are free for chunk c + 2.
    constexpr bool PIPE = FAST && N > 64;
    const uint32_t round0 = round;
    auto send = [&](int ci) {
      const uint32_t xr = round0 + (uint32_t)ci;
      exchange_send<CH, false, true>(has_work, tmem_base, ci * CH, sm.recv + (size_t)(xr & 1u) * 4 * TILE_M * NC4,
                        reinterpret_cast<float*>(sm.ring.base) + (size_t)(ci & 1) * 4 * TILE_M * NC4, &sm.recv_full[xr & 1u], rank);
    };
    if (PIPE) {
      prefetch(0);
      send(0);
      if (N > CH) send(1);
    }
    for (int c0 = 0; c0 < N; c0 += CH) {
      float* recv = sm.recv + (size_t)(round & 1u) * 4 * TILE_M * NC4;
      ++round;
      const int bbase = c0 + rank * NC4;
      if (!PIPE) {
        prefetch(c0);
        // (`round` was advanced when recv was picked: this exchange is number round - 1)
        const uint32_t xr = round - 1u;
        float* stage = sm.recv + (size_t)((xr & 1u) ^ 1u) * 4 * TILE_M * NC4;
        if (FAST) exchange_bulk<CH, false, true>(has_work, tmem_base, c0, recv, stage, &sm.recv_full[xr & 1u], (xr >> 1) & 1u, rank);
        else exchange_bulk<CH>(has_work, tmem_base, c0, recv, stage, &sm.recv_full[xr & 1u], (xr >> 1) & 1u, rank);
      } else {
        const uint32_t xr = round - 1u;
        wait_bar(&sm.recv_full[xr & 1u], (xr >> 1) & 1u);
        tc_fence_after();
      }
      PROF(ps + 2);                              // DSMEM exchange + cluster barrier done
      if (is_lstm) {
        const float pdrop = is_att ? p.p_att : p.p_dec;
        const unsigned thr = is_att ? p.thr_att : p.thr_dec;
        const float dscale = 1.0f / (1.0f - pdrop);
        const unsigned site = is_att ? 3u : 4u;
        bf16* xh_next = (is_att ? BFP(p.xh_a) : BFP(p.xh_d)) + (size_t)nxt * p.B * mult * Hp;
        float* gates_all = is_att ? p.gates_a : p.gates_d;
        float* c_all = is_att ? p.c_a_all : p.c_d_all;
        float* h_all = is_att ? p.h_a_all : p.h_d_all;
#pragma unroll
        for (int q = 0; q < CPT; ++q) {
          const int idx = threadIdx.x + q * 256;
          const int ul = idx & 31, bl = idx >> 5, b = bbase + bl, unit = tile * 32 + ul;
          float h = 0.f;
          if (b < p.B) {
            // bf16 mode: hardware tanh (tanh.approx, rel. error ~2^-11, below the bf16 rounding of every operand; the energies
            // and the backward's tanh(c) use it too); fp32 mode: libm
            const bool fast = FAST || p.prec == 1;
            // (FAST: column-major receive blocks, conflict-free for lanes over units)
            const float zi = (FAST ? recv_sum_cm<NC4>(recv, ul, bl) : recv_sum<NC4, true>(recv, ul, bl)) + gb[q].x;
            const float zf = (FAST ? recv_sum_cm<NC4>(recv, 32 + ul, bl) : recv_sum<NC4, true>(recv, 32 + ul, bl)) + gb[q].y;
            const float zg = (FAST ? recv_sum_cm<NC4>(recv, 64 + ul, bl) : recv_sum<NC4, true>(recv, 64 + ul, bl)) + gb[q].z;
            const float zo = (FAST ? recv_sum_cm<NC4>(recv, 96 + ul, bl) : recv_sum<NC4, true>(recv, 96 + ul, bl)) + gb[q].w;
            const float i_ = fast ? 0.5f * tanh_fast(0.5f * zi) + 0.5f : sigmoid_acc(zi);
            const float f_ = fast ? 0.5f * tanh_fast(0.5f * zf) + 0.5f : sigmoid_acc(zf);
            const float g_ = fast ? tanh_fast(zg) : tanh_acc(zg);
            const float o_ = fast ? 0.5f * tanh_fast(0.5f * zo) + 0.5f : sigmoid_acc(zo);
            const size_t ci = (size_t)b * H + unit;
            const float c_new = f_ * cprev[q] + i_ * g_;
            cst[ci] = c_new;
            h = o_ * (fast ? tanh_fast(c_new) : tanh_acc(c_new));
            if (p.use_dropout && thr != 0u) {
              const uint32_t bits = philox_bits(p.seed, site, (uint32_t)t, (uint32_t)ci);
              h = bits >= thr ? h * dscale : 0.f;
            }
            store_act(xh_next + (size_t)b * mult * Hp, unit, Hp, p.prec, h);
            if (P_TRAIN) {
              const size_t cell = ((size_t)t * p.B + b) * H + unit;
              reinterpret_cast<float4*>(gates_all)[cell] = make_float4(i_, f_, g_, o_);
              c_all[cell] = c_new;
              h_all[cell] = h;
            }
          }
          // (float4 column groups XOR-permuted by the unit: the lanes of a warp run over units, 16-way bank conflict otherwise)
          if (is_att) sm.hs[ul * NC4 + (bl ^ (((ul >> 1) & (NC4 / 4 - 1)) << 2))] = h;
        }
        if (is_att) {
          if (P_TRAIN) PROF(14);
          __syncthreads();
          if (P_TRAIN) PROF(15);
          // partial query of this tile: qpart[tile][b][a] = sum_u Wq[a][32 tile + u] h[b][u]; thread = (a, half of the columns)
          constexpr int HB = NC4 / 2;
          const int a = threadIdx.x & 127, g = threadIdx.x >> 7;
          float acc[HB];
#pragma unroll
          for (int j = 0; j < HB; ++j) acc[j] = 0.f;
#pragma unroll
          for (int u = 0; u < 32; ++u) {
            const float4* hr = reinterpret_cast<const float4*>(sm.hs + u * NC4);
#pragma unroll
            for (int j4 = 0; j4 < HB / 4; ++j4) {
              const float4 hv = hr[(g * (HB / 4) + j4) ^ ((u >> 1) & (NC4 / 4 - 1))];
              acc[4 * j4] += wqr[u] * hv.x; acc[4 * j4 + 1] += wqr[u] * hv.y;
              acc[4 * j4 + 2] += wqr[u] * hv.z; acc[4 * j4 + 3] += wqr[u] * hv.w;
            }
          }
          if (P_TRAIN) PROF(16);
          if (a < p.A) {
#pragma unroll
            for (int j = 0; j < HB; ++j) {
              const int b = bbase + g * HB + j;
              if (b < p.B) p.qpart[((size_t)tile * p.B + b) * p.A + a] = acc[j];
            }
          }
        }
      } else {
        const int rows = kind == K_PROJ ? p.NM + 1 : p.P;
        for (int idx = threadIdx.x; idx < TILE_M * NC4; idx += blockDim.x) {
          const int bl = idx % NC4, r = idx / NC4, b = bbase + bl, row = tile * TILE_M + r;
          if (b >= p.B || row >= rows) continue;
          float val = recv_sum<NC4, true>(recv, r, bl);
          if (kind != K_PROJ) {
            val = fmaxf(val, 0.f);
            const uint32_t bits = philox_bits(p.seed, kind == K_PRE1 ? 1u : 2u, (uint32_t)t, (uint32_t)(b * p.P + row));
            val = bits >= 0x80000000u ? val * 2.0f : 0.f;     // p = 0.5, always on (model.py:104)
            store_act((kind == K_PRE1 ? BFP(p.xp1) : BFP(p.xp2)) + (size_t)b * mult * p.Pp, row, p.Pp, p.prec, val);
          } else {
            val += p.pbias[(size_t)b * (p.NM + 1) + row];
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
      if (PIPE && c0 + CH < N) {
        prefetch(c0 + CH);
        if (c0 + 2 * CH < N) { cluster_sync_relaxed(); send(c0 / CH + 2); }
        else __syncthreads();                      // sm.hs of this chunk has been read
      }
    }
    if (PIPE && tile + ncl < ntiles) cluster_sync_all();   // the peers have received out of the ring before it is refilled
    if (!FAST && kind == K_PROJ) {
      // the cluster that owns the projection decides whether every sequence has fired its gate (model.py:486)
      cluster_sync_all();                                  // the four ranks' done[] writes
      if (rank == 0) {
        __threadfence();
        int all = 1;
        for (int b = threadIdx.x; b < p.B; b += blockDim.x) all &= (__ldcg(&p.done[b]) != 0);
        all = __syncthreads_and(all);
        if (threadIdx.x == 0 && all && !p.ignore_stop) *p.stop = 1;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Prenet of autoregressive inference at small batch (B <= PRE_LOCAL_BMAX), computed inside EVERY cluster with plain FMAs:
// at B = 1 a prenet layer is a 256 x 80..256 mat-vec, and running it as two cluster-split GEMM phases costs two TMA ->
// tcgen05 -> DSMEM-exchange chains plus two grid barriers (~20k of the frame's ~85k cycles).  Here each CTA computes all
// of layer 1 (32 KB of bf16 weights from L2) and its quarter of the layer-2 rows; the four quarters meet in global memory
// behind one cluster barrier.  All clusters write the same bits to xp2, so no grid barrier is needed before the attention
// LSTM GEMM reads it.  Same operands as the GEMM path (bf16 planes of weights and activations, fp32 accumulate); only the
// summation order differs.  (model.py:97-106)
// ------------------------------------------------------------------------------------------------
static constexpr int PRE_LOCAL_BMAX = 4;

GTN_DEVICE void store_act_f(float* x, int k, int kpad, int prec, float v) {     // store_act into an fp32 staging buffer
  const bf16 hi = __float2bfloat16_rn(v);
  const float fh = __bfloat162float(hi);
  if (prec == 1) { x[k] = fh; return; }
  const float r1 = v - fh;
  const bf16 lo = __float2bfloat16_rn(r1);
  const float fl = __bfloat162float(lo);
  const float fl2 = __bfloat162float(__float2bfloat16_rn(r1 - fl));
  x[k] = fh; x[kpad + k] = fh; x[2 * kpad + k] = fh;
  x[3 * kpad + k] = fl; x[4 * kpad + k] = fl;
  x[5 * kpad + k] = fl2;
}

// y[row][b] = sum_k W[row][k] xs[b][k] for rows [row_begin, row_end), K a multiple of 64; a group of G lanes owns a row,
// U rows per group are in flight at once (the loads are L2 latency bound).  Raw sums go to y[b * ldy + row].  Deliberately
// NOT inlined: both prenet layers run the same (instruction-cache warm) copy.
__device__ __noinline__ void matvec_rows(const bf16* W, int K, const float* xs, int B, int row_begin, int row_end, float* y,
                                         int ldy) {
  constexpr int U = 8;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int C = K >> 3;                                   // 16-byte chunks per row
  const int G = C >= 32 ? 32 : (C >= 16 ? 16 : 8);
  const int RP = 32 / G, sub = lane / G, gl = lane % G;
  for (int r0 = row_begin + warp * RP * U; r0 < row_end; r0 += 8 * RP * U) {
    float acc[U][PRE_LOCAL_BMAX];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int b = 0; b < PRE_LOCAL_BMAX; ++b) acc[u][b] = 0.f;
    for (int c = gl; c < C; c += G) {
      uint4 q[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int row = r0 + u * RP + sub;
        q[u] = row < row_end ? __ldg(reinterpret_cast<const uint4*>(W + (size_t)row * K) + c) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int b = 0; b < PRE_LOCAL_BMAX; ++b) {
        if (b >= B) break;
        const float4 x0 = *reinterpret_cast<const float4*>(xs + (size_t)b * K + c * 8);
        const float4 x1 = *reinterpret_cast<const float4*>(xs + (size_t)b * K + c * 8 + 4);
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const uint32_t w[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
          float a = acc[u][b];
          a += __uint_as_float(w[0] << 16) * x0.x; a += __uint_as_float(w[0] & 0xFFFF0000u) * x0.y;
          a += __uint_as_float(w[1] << 16) * x0.z; a += __uint_as_float(w[1] & 0xFFFF0000u) * x0.w;
          a += __uint_as_float(w[2] << 16) * x1.x; a += __uint_as_float(w[2] & 0xFFFF0000u) * x1.y;
          a += __uint_as_float(w[3] << 16) * x1.z; a += __uint_as_float(w[3] & 0xFFFF0000u) * x1.w;
          acc[u][b] = a;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int row = r0 + u * RP + sub;
#pragma unroll
      for (int b = 0; b < PRE_LOCAL_BMAX; ++b) {
        if (b >= B) break;
        float a = acc[u][b];
        for (int off = G >> 1; off > 0; off >>= 1) a += __shfl_xor_sync(0xffffffffu, a, off);
        if (gl == 0 && row < row_end) y[b * ldy + row] = a;
      }
    }
  }
}

GTN_DEVICE void prenet_local(const DecFwd& p, Smem2& sm, int t, int rank) {
  const int mult = p.mult, K1 = mult * p.NMp, K2 = mult * p.Pp, P = p.P, B = p.B;
  float* x1 = reinterpret_cast<float*>(sm.ring.base);      // [B][K1]  (the TMA ring is idle between the GEMM phases)
  float* x2 = x1 + (size_t)B * K1;                         // [B][K2]
  float* y = x2 + (size_t)B * K2;                          // [B][P] raw sums of the layer in flight
  for (int i = threadIdx.x; i < B * K1; i += blockDim.x) x1[i] = __bfloat162float(__ldcg(BFP(p.xmel) + i));
  for (int i = threadIdx.x; i < B * K2; i += blockDim.x) x2[i] = 0.f;      // padded columns P..Pp of every plane
  __syncthreads();
  // the mat-vecs only drop raw sums into shared memory; relu + dropout + plane split run afterwards, one (row, b) per
  // thread, as ONE compact loop (inlined into the unrolled mat-vec the Philox code is replicated U x B times, and a code
  // path that runs once per frame pays an instruction-cache miss for every line of it)
  auto act = [&](float v, uint32_t site, int b, int row) {
    v = fmaxf(v, 0.f);
    const uint32_t bits = philox_bits(p.seed, site, (uint32_t)t, (uint32_t)(b * P + row));
    return bits >= 0x80000000u ? v * 2.0f : 0.f;           // p = 0.5, always on (model.py:104)
  };
  matvec_rows(BFP(p.w_pre1), K1, x1, B, 0, P, y, P);
  __syncthreads();
  for (int i = threadIdx.x; i < B * P; i += blockDim.x) {
    const int b = i / P, row = i - b * P;
    store_act_f(x2 + (size_t)b * K2, row, p.Pp, p.prec, act(y[i], 1u, b, row));
  }
  __syncthreads();
  PROF(29);
  const int q = (P + CL - 1) / CL, r_lo = rank * q, r_hi = min(P, (rank + 1) * q);
  matvec_rows(BFP(p.w_pre2), K2, x2, B, r_lo, r_hi, y, P);
  __syncthreads();
  for (int i = threadIdx.x; i < B * (r_hi - r_lo); i += blockDim.x) {
    const int b = i / (r_hi - r_lo), row = r_lo + i - b * (r_hi - r_lo);
    store_act(BFP(p.xp2) + (size_t)b * K2, row, p.Pp, p.prec, act(y[b * P + row], 2u, b, row));
  }
  PROF(30);
  fence_proxy_async();          // xp2 (global) is read by TMA next; the staging area is overwritten by TMA
  cluster_sync_all();
}

// ------------------------------------------------------------------------------------------------
// "Tail" of a frame of batched autoregressive inference (bf16 mode, B > PRE_LOCAL_BMAX): mel / gate projection of frame t
// and both prenet layers of frame t + 1 INSIDE each cluster for the cluster's own ceil(B / clusters) <= 8 sequences
// (model.py:411-416, 486, 97-106).  As three grid-wide GEMM phases these cost three TMA -> tcgen05 -> DSMEM chains and three
// grid barriers per frame (47k of 221k cycles at B = 256) although they are three small products whose inputs, once
// h_dec(t) is complete, never leave the sequence.  Here they are warp-level tensor-core products (mma.sync m16n8k16: 16
// weight rows x the cluster's 8 sequences, fp32 accumulate) with BOTH operands loaded straight from L2 in fragment order
// -- a lane's 16 bytes of a row are 8 consecutive K elements, used as the k-pairs of two consecutive MMAs; weights and
// activations use the same permutation of K, so the sum is unchanged -- every load of a warp in flight at once, no shared
// memory staging.  The four CTAs of a cluster split the 16-row tiles, the warps of a CTA split K (partials summed through
// shared memory), the layers meet in global memory behind two cluster barriers, and ONE grid barrier follows (xp2 is a
// TMA operand of every cluster's next GEMM).  A first version with FMA mat-vecs unrolled over rows x sequences was 3x
// SLOWER than the GEMM phases: 90 KB of straight-line code that runs once per frame is bound by instruction fetch.
// Same operands as the GEMM path (bf16 weights and activations, fp32 accumulate); only the summation order differs.
// ------------------------------------------------------------------------------------------------
static constexpr int TAIL_BMAX = 8;     // sequences per cluster = N of the MMA
static constexpr int TAIL_UNITS = 16;   // (tile, K part) partial tiles of one layer per CTA (red_s holds 16 x 128 floats)

GTN_DEVICE void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// c[16 rows][8 sequences] += sum over K slabs [s0, s1) (32 elements each) of W[row0 + r][k] x[n][k]; rows >= nrows and
// sequences >= nb contribute zeros.  x[n][k] = k < ksplit ? xa[n * lda + k] : xb[n * ldb + k - ksplit] (bf16, written by other
// CTAs during this launch: L2 loads).  Fragment owner: g = lane / 4 (row g and g + 8 of A, column g of B), q = lane % 4.
GTN_DEVICE void tail_mma_rows(float (&c)[4], const bf16* W, int ldw, int row0, int nrows, const bf16* xa, int lda, const bf16* xb,
                              int ldb, int ksplit, int nb, int s0, int s1) {
  constexpr int UB = 6;                                    // slabs in flight per lane (18 x 16-byte loads)
  const int lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
  const bool r0ok = row0 + g < nrows, r1ok = row0 + g + 8 < nrows, nok = g < nb;
  const bf16* w0 = W + (size_t)(row0 + g) * ldw + 8 * q;
  const bf16* w1 = w0 + (size_t)8 * ldw;
  const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll 1
  for (int sb = s0; sb < s1; sb += UB) {
    uint4 wa[UB], wb[UB], xv[UB];
#pragma unroll
    for (int j = 0; j < UB; ++j) {
      const int k = (sb + j) * 32;
      const bool on = sb + j < s1;
      wa[j] = (on && r0ok) ? __ldg(reinterpret_cast<const uint4*>(w0 + k)) : z4;
      wb[j] = (on && r1ok) ? __ldg(reinterpret_cast<const uint4*>(w1 + k)) : z4;
      const bf16* xs = k < ksplit ? xa + (size_t)g * lda + k : xb + (size_t)g * ldb + (k - ksplit);
      xv[j] = (on && nok) ? __ldcg(reinterpret_cast<const uint4*>(xs + 8 * q)) : z4;
    }
#pragma unroll
    for (int j = 0; j < UB; ++j) {
      mma_bf16_16816(c, wa[j].x, wb[j].x, wa[j].y, wb[j].y, xv[j].x, xv[j].y);
      mma_bf16_16816(c, wa[j].z, wb[j].z, wa[j].w, wb[j].w, xv[j].z, xv[j].w);
    }
  }
}

// One layer for this CTA: tiles [tile0, tile0 + ntl) of 16 rows (tile index step tstep between them), K split into KP parts
// over the warps; partial tiles -> part[unit][16][8] (unit = kp * ntl + tl).  Ends with a CTA barrier.
GTN_DEVICE void tail_layer(float* part, const bf16* W, int ldw, int nrows, int tile0, int tstep, int ntl, int KP, const bf16* xa,
                           int lda, const bf16* xb, int ldb, int ksplit, int K, int nb) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, q = lane & 3;
  const int nsl = K / 32;
  for (int unit = warp; unit < ntl * KP; unit += 8) {
    const int tl = unit % ntl, kp = unit / ntl;
    float c[4] = {0.f, 0.f, 0.f, 0.f};
    tail_mma_rows(c, W, ldw, (tile0 + tl * tstep) * 16, nrows, xa, lda, xb, ldb, ksplit, nb, kp * nsl / KP, (kp + 1) * nsl / KP);
    float* d = part + (size_t)unit * 128;
    *reinterpret_cast<float2*>(d + g * 8 + 2 * q) = make_float2(c[0], c[1]);
    *reinterpret_cast<float2*>(d + (g + 8) * 8 + 2 * q) = make_float2(c[2], c[3]);
  }
  __syncthreads();
}

// Can the tail run for this configuration?  (bf16 planes only; K segments in 32-element slabs; tile counts within part[])
GTN_HD bool tail_covers(const DecFwd& p, int nbc) {
  const int ntp = ((p.P + 15) / 16 + CL - 1) / CL;         // prenet tiles per CTA
  return p.prec == 1 && nbc <= TAIL_BMAX && (p.NM + 1 + 15) / 16 <= 2 * CL && 2 * ntp <= TAIL_UNITS && p.Hdp % 32 == 0 &&
         p.Ep % 32 == 0 && p.NMp % 32 == 0 && p.Pp % 32 == 0;
}

// do_proj: projection + gate test of frame t, then the prenet of frame t + 1 (skipped behind the last frame);
// !do_proj: only the prenet of frame 0 (go frame, xmel is zero).
GTN_DEVICE void tail_phase(const DecFwd& p, Smem2& sm, int t, bool do_proj, int rank, int cid, int ncl) {
  const int nbc = (p.B + ncl - 1) / ncl;
  const int b0 = cid * nbc;
  const int nb = min(p.B, b0 + nbc) - b0;
  if (nb <= 0) return;                                     // (cluster-uniform)
  const int P = p.P, R = p.NM + 1, Kp = p.Hdp + p.Ep;
  float* part = sm.red_s;                                  // [TAIL_UNITS][16][8]
  if (do_proj) {
    // 16-row tiles rank, rank + 4 of the R = n_mel + 1 rows; K in four parts over the warps
    const int nt = (R + 15) / 16, ntl = rank + CL < nt ? 2 : (rank < nt ? 1 : 0);
    tail_layer(part, BFP(p.w_proj), Kp, R, rank, CL, ntl, 4, BFP(p.xh_d) + ((size_t)((t & 1) ^ 1) * p.B + b0) * p.Hdp, p.Hdp,
               BFP(p.xctx) + (size_t)b0 * p.Ep, p.Ep, p.Hdp, Kp, nb);
    for (int i = threadIdx.x; i < ntl * 128; i += blockDim.x) {
      const int tl = i >> 7, r = (i >> 3) & 15, n = i & 7, row = (rank + tl * CL) * 16 + r, b = b0 + n;
      if (row >= R || n >= nb) continue;
      float val = p.pbias[(size_t)b * R + row];
#pragma unroll
      for (int kp = 0; kp < 4; ++kp) val += part[(size_t)(kp * ntl + tl) * 128 + r * 8 + n];
      if (row < p.NM) {
        p.mel_out[((size_t)b * p.NM + row) * p.n_steps + t] = val;
        BFP(p.xmel)[(size_t)b * p.NMp + row] = __float2bfloat16_rn(val);
      } else {
        p.gate_out[(size_t)b * p.n_steps + t] = val;
        if (sigmoid_acc(val) > p.gate_threshold && p.done[b] == 0) { p.done[b] = 1; p.lengths_out[b] = t + 1; }
      }
    }
    cluster_sync_all();                                    // the row tiles of mel(t) meet in global memory
    if (t + 1 >= p.n_steps) return;
  }
  const int tp = do_proj ? t + 1 : 0;                      // frame the prenet output belongs to (Philox step)
  const int nt = (P + 15) / 16, tpr = (nt + CL - 1) / CL, tile0 = rank * tpr, ntl = max(0, min(nt, tile0 + tpr) - tile0);
  for (int layer = 0; layer < 2; ++layer) {
    const int K = layer == 0 ? p.NMp : p.Pp;
    const bf16* xin = layer == 0 ? BFP(p.xmel) + (size_t)b0 * p.NMp : BFP(p.xp1) + (size_t)b0 * p.Pp;
    tail_layer(part, layer == 0 ? BFP(p.w_pre1) : BFP(p.w_pre2), K, P, tile0, 1, ntl, 2, xin, K, xin, K, K, K, nb);
    bf16* xout = layer == 0 ? BFP(p.xp1) : BFP(p.xp2);
    for (int i = threadIdx.x; i < ntl * 128; i += blockDim.x) {
      const int tl = i >> 7, r = (i >> 3) & 15, n = i & 7, row = (tile0 + tl) * 16 + r, b = b0 + n;
      if (row >= P || n >= nb) continue;
      float v = fmaxf(part[(size_t)tl * 128 + r * 8 + n] + part[(size_t)(ntl + tl) * 128 + r * 8 + n], 0.f);
      const uint32_t bits = philox_bits(p.seed, layer == 0 ? 1u : 2u, (uint32_t)tp, (uint32_t)(b * P + row));
      v = bits >= 0x80000000u ? v * 2.0f : 0.f;            // p = 0.5, always on (model.py:104)
      xout[(size_t)b * p.Pp + row] = __float2bfloat16_rn(v);
    }
    if (layer == 0) cluster_sync_all();                    // layer 2 reads every row of layer 1 (also fences part[])
  }
}

// ------------------------------------------------------------------------------------------------
// attention phase: one CTA pair per sequence
// ------------------------------------------------------------------------------------------------
template <typename MT>
GTN_DEVICE void attention_phase2(const DecFwd& p, Smem2& sm, int t, int rank, int cid, int ncl, uint32_t tmem_att,
                                 uint32_t& att_par, uint32_t& eround) {
  constexpr bool FAST = false;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Tin = p.Tin, A = p.A, KS = p.KS, hk = (p.KS - 1) / 2;
  const int half = rank & 1, peer = rank ^ 1;
  const int pair = cid * 2 + (rank >> 1), npairs = ncl * 2;
  const int JH = (Tin + 1) / 2;
  const int j0 = half * JH, jn = min(Tin, j0 + JH) - j0;
  const int Wd = jn + KS - 1;
  const MT* mem = reinterpret_cast<const MT*>(p.mem_enc);
  const MT* pm = reinterpret_cast<const MT*>(p.pm);
  uint8_t* a_hi = sm.ring.base;                          // im2col tiles alias ring stage 0 (idle during this phase)
  uint8_t* a_lo = sm.ring.base + W_TILE_BYTES;
  const int niter = (p.B + npairs - 1) / npairs;
  for (int it = 0; it < niter; ++it) {
    const int b = pair + it * npairs;
    const bool active = b < p.B;
    const int L = active ? reinterpret_cast<const int*>(sm.scr + 48)[it] : 0;   // sequence lengths: staged once at kernel start
    float* e_buf = sm.e_buf + (size_t)(eround & 1u) * TIN_MAX;
    ++eround;
    // this thread's slice of processed_memory (row = encoder position, 64 attention columns) does not depend on the frame's
    // recurrent state: requested first so that its latency hides behind the query gather / im2col / MMA  (bf16 only)
    constexpr bool PM_PRE = sizeof(MT) == 2;
    uint4 pmq[PM_PRE ? 8 : 1];
    {
      const int wq_ = warp & 3, a0_ = (warp >> 2) * 64, jl_ = wq_ * 32 + lane;
      if (PM_PRE && active && jl_ < jn && (j0 + jl_) < L) {
        if (p.reserved0 & 32) {             // chunk-major [B][A / 8][Tin][8]: a warp's positions are contiguous
          const uint4* src = reinterpret_cast<const uint4*>(pm) + ((size_t)b * (A >> 3) + (a0_ >> 3)) * Tin + (j0 + jl_);
#pragma unroll
          for (int g = 0; g < 8; ++g)
            if (a0_ + g * 8 < A) pmq[g] = __ldg(src + (size_t)g * Tin);
        } else {
          const uint4* src = reinterpret_cast<const uint4*>(pm + ((size_t)b * Tin + (j0 + jl_)) * A + a0_);
#pragma unroll
          for (int g = 0; g < 8; ++g)
            if (a0_ + g * 8 < A) pmq[g] = __ldg(src + g);
        }
      }
    }
    if (active) {
      const int ntq = p.Ha / 32;
      for (int a = threadIdx.x; a < A; a += blockDim.x) {
        // every partial of the query in flight at once: one L2 round trip instead of ntq / 8
        const float* qp = p.qpart + (size_t)b * A + a;
        const size_t qs = (size_t)p.B * A;
        float s = 0.f;
        for (int c0 = 0; c0 < ntq; c0 += 32) {
          float qv[32];
#pragma unroll
          for (int c = 0; c < 32; ++c) qv[c] = (c0 + c < ntq) ? __ldcg(qp + (size_t)(c0 + c) * qs) : 0.f;
#pragma unroll
          for (int c = 0; c < 16; ++c) qv[c] += qv[c + 16];
#pragma unroll
          for (int c = 0; c < 8; ++c) qv[c] += qv[c + 8];
#pragma unroll
          for (int c = 0; c < 4; ++c) qv[c] += qv[c + 4];
          s += (qv[0] + qv[1]) + (qv[2] + qv[3]);
        }
        sm.q_s[a] = s;
        if (P_TRAIN && half == 0) p.q_all[((size_t)t * p.B + b) * A + a] = s;
      }
      if (P_TRAIN) PROF(7);
      const float* wprev = p.w_prev + (size_t)b * p.ldT;
      const float* wcum = p.w_cum + (size_t)b * p.ldT;
      for (int i = threadIdx.x; i < 2 * Wd; i += blockDim.x) {
        const int ch = i / Wd, jj = i - ch * Wd, j = j0 + jj - hk;
        const float val = (j >= 0 && j < Tin) ? __ldcg(ch == 0 ? &wprev[j] : &wcum[j]) : 0.f;
        sm.wcat[i] = val;
        if (P_TRAIN && ch == 1 && jj >= hk && jj < hk + jn) p.cum_all[((size_t)t * p.B + b) * Tin + j] = val;
      }
    }
    if (P_TRAIN) PROF(12);
    __syncthreads();
    if (P_TRAIN) PROF(20);
    if (active) {
      // im2col: row r = local encoder position, K index kk = ch*KS + k  ->  wcat[ch][r + k]; bf16 hi / lo planes
      build_im2col(sm.wcat, Wd, KS, jn, a_hi, a_lo);
      fence_proxy_async_smem();
    }
    if (P_TRAIN) PROF(21);
    __syncthreads();
    if (P_TRAIN) PROF(22);
    // warp 1 issues in warp-uniform control flow (one elected lane; see gemm_issue)
    if (active && __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0) == 1) {
      tc_fence_after();
      const uint32_t idesc = umma_idesc_bf16(TILE_M, 128);
      const uint64_t dah = umma_desc_sw128(smem_u32(a_hi)), dal = umma_desc_sw128(smem_u32(a_lo));
      const uint64_t duh = umma_desc_sw128(smem_u32(sm.u_s)), dul = umma_desc_sw128(smem_u32(sm.u_s + W_TILE_BYTES));
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < KB / 16; ++k) umma_bf16(tmem_att, dah + (uint64_t)(2 * k), duh + (uint64_t)(2 * k), idesc, (uint32_t)(k > 0));
#pragma unroll
        for (int k = 0; k < KB / 16; ++k) umma_bf16(tmem_att, dal + (uint64_t)(2 * k), duh + (uint64_t)(2 * k), idesc, 1u);
#pragma unroll
        for (int k = 0; k < KB / 16; ++k) umma_bf16(tmem_att, dah + (uint64_t)(2 * k), dul + (uint64_t)(2 * k), idesc, 1u);
        umma_commit(sm.att_full);
      }
    }
    __syncwarp();
    PROF(24);                                    // q / window / im2col built, MMAs issued
    if (active) {
      wait_bar(sm.att_full, att_par);
      att_par ^= 1u;
      tc_fence_after();
      // thread = (row j = TMEM lane, 64-column half of the attention dimension)
      const int wq = warp & 3, a0 = (warp >> 2) * 64;
      const int jl = wq * 32 + lane;
      const bool valid = jl < jn && (j0 + jl) < L;
      float part = 0.f;
      const MT* pmj = pm + ((size_t)b * Tin + (j0 + jl)) * A + a0;
      // (tanh save: chunk-major inside the half's block, see gtn_decoder_fwd3.cu)
      float* thb = P_TRAIN ? p.th_all + (((size_t)t * p.B + b) * Tin + j0) * A : nullptr;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_att + ((uint32_t)(wq * 32) << 16) + (uint32_t)(a0 + h * 32), r);
        tmem_ld_wait();
        if (valid) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const int ab = h * 32 + g * 8;
            if (a0 + ab < A) {
              float pmv[8];
              if (PM_PRE) {
                const uint4 qq = pmq[h * 4 + g];
                const uint32_t ww[4] = {qq.x, qq.y, qq.z, qq.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  pmv[2 * i] = __uint_as_float(ww[i] << 16);
                  pmv[2 * i + 1] = __uint_as_float(ww[i] & 0xFFFF0000u);
                }
              } else {
                ld8<MT>(pmj + ab, pmv);
              }
              float th[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const float s = __uint_as_float(r[g * 8 + e]) + sm.q_s[a0 + ab + e] + pmv[e];
                th[e] = p.prec == 1 ? tanh_fast(s) : tanh_acc(s);
                part += sm.v_s[a0 + ab + e] * th[e];
              }
              if (thb) {
                float4* thc = reinterpret_cast<float4*>(thb) + (size_t)((a0 + ab) >> 2) * jn + jl;
                thc[0] = make_float4(th[0], th[1], th[2], th[3]);
                thc[jn] = make_float4(th[4], th[5], th[6], th[7]);
              }
            }
          }
        }
      }
      sm.e_part[(warp >> 2) * 128 + jl] = valid ? part : -INFINITY;
    }
    tc_fence_before();
    __syncthreads();
    if (active) {
      const uint32_t peer_e = map_to_rank(smem_u32(e_buf), (uint32_t)peer);
      for (int jl = threadIdx.x; jl < jn; jl += blockDim.x) {
        const float e = sm.e_part[jl] + sm.e_part[128 + jl];      // -inf + -inf = -inf for masked positions
        e_buf[j0 + jl] = e;
        st_cluster_f1(peer_e + (uint32_t)((j0 + jl) * 4), e);
      }
    }
    PROF(25);                                    // energies done
    cluster_sync_all();
    PROF(26);
    if (active) {
      // masked softmax over all encoder positions (both halves compute it redundantly)
      float mx = -INFINITY;
      for (int j = threadIdx.x; j < Tin; j += blockDim.x) mx = fmaxf(mx, e_buf[j]);
      mx = block_reduce2(mx, sm.scr, true);
      float sum = 0.f;
      for (int j = threadIdx.x; j < Tin; j += blockDim.x) {
        const float e = e_buf[j] == -INFINITY ? 0.f : expf(e_buf[j] - mx);
        sm.w_s[j] = e;
        sum += e;
      }
      sum = block_reduce2(sum, sm.scr, false);
      const float inv = 1.0f / sum;
      float wsum = 0.f;
      for (int j = threadIdx.x; j < Tin; j += blockDim.x) {
        const float w = sm.w_s[j] * inv;
        sm.w_s[j] = w;
        wsum += w;
      }
      wsum = block_reduce2(wsum, sm.scr, false);     // also orders the w_s writes before the reads below
      float* wprev = p.w_prev + (size_t)b * p.ldT;
      float* wcum = p.w_cum + (size_t)b * p.ldT;
      float* aout = p.align_out + ((size_t)b * p.n_steps + t) * Tin;
      for (int jl = threadIdx.x; jl < jn; jl += blockDim.x) {
        const int j = j0 + jl;
        const float w = sm.w_s[j];
        wprev[j] = w;
        wcum[j] = sm.wcat[Wd + hk + jl] + w;            // cumulative weights before the step live in wcat ch. 1
        aout[j] = w;
      }
      PROF(27);                                  // softmax + weight outputs done
      // context over this CTA's half of the time-varying columns: 8 columns per thread, warps split j
      const int E8 = p.E / 8, Eh8 = (E8 + 1) / 2;
      const int c8_0 = half * Eh8, c8_n = min(E8, c8_0 + Eh8) - c8_0;      // in units of 8 columns
      const int mult = p.mult;
      bf16* xc = BFP(p.xctx) + (size_t)b * mult * p.Ep;
      float* call = P_TRAIN ? p.ctx_all + ((size_t)t * p.B + b) * p.D : nullptr;
      for (int cb = 0; cb < c8_n; cb += 32) {
        const int c8 = cb + lane;
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
        if (c8 < c8_n) {
          const MT* mp = mem + (size_t)b * Tin * p.E + (size_t)(c8_0 + c8) * 8;
          int j = warp;
          for (; j + 24 < L; j += 32) {              // four rows of this warp in flight
            float mv[4][8];
#pragma unroll
            for (int r = 0; r < 4; ++r) ld8<MT>(mp + (size_t)(j + 8 * r) * p.E, mv[r]);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              const float w = sm.w_s[j + 8 * r];
#pragma unroll
              for (int e = 0; e < 8; ++e) acc[e] += w * mv[r][e];
            }
          }
          for (; j < L; j += 8) {
            float mv[8];
            ld8<MT>(mp + (size_t)j * p.E, mv);
            const float w = sm.w_s[j];
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] += w * mv[e];
          }
        }
        __syncthreads();
#pragma unroll
        for (int e = 0; e < 8; ++e) sm.red_s[warp * 256 + lane * 8 + e] = acc[e];
        __syncthreads();
        for (int i = threadIdx.x; i < 256; i += blockDim.x) {
          const int c8i = cb + (i >> 3);
          if (c8i < c8_n) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) s += sm.red_s[w * 256 + i];
            const int d = (c8_0 + c8i) * 8 + (i & 7);
            store_act(xc, d, p.Ep, p.prec, s);
            if (call) call[d] = s;
          }
        }
      }
      // time-constant columns: context = const * sum_j w_j  (second CTA of the pair)
      if (call && (half == 1 || Tin == 1)) {
        // (saved for the weight-gradient GEMMs only: the per-frame GEMMs see these columns as a per-sequence bias)
        for (int i = threadIdx.x; i < p.Cc; i += blockDim.x) call[p.E + i] = p.mem_const[(size_t)b * p.Cc + i] * wsum;
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------
// Attention phase of autoregressive inference, ALL sequences of the CTA pair at once (bf16 memory).  attention_phase2 serves
// its sequences one after the other, each through a dozen latency-bound stages with CTA / cluster barriers between them, and
// reads the encoder-memory tile row by row through L1 with four rows per warp in flight (9 B/clk per SM): 22k cycles per
// sequence, 87k of a 221k-cycle frame at B = 256.  Here every stage runs once for all sequences:
//   Q  queries (32 partials per element in flight) and [prev | cum] windows of all sequences        1 barrier
//   I  im2col tiles of all sequences (idle ring / receive buffers)                                   1 barrier
//   M  location GEMMs of all sequences (one accumulator each); the first encoder-memory chunks are requested
//   E  energies per sequence as its accumulator completes (next sequence's processed_memory rows prefetched)
//      pair exchange of all energies through DSMEM                                                   1 cluster barrier
//   S  masked softmax: one WARP per sequence (shuffles only); weight state, alignments               1 barrier
//   C  context: a CTA takes its HALF OF THE POSITIONS (a contiguous run of memory rows) for all columns, streamed through
//      three shared-memory slots by bulk copies (chunks of a third of the rows, issued as slots free up), so the phase is
//      bound by the SM's L2 port instead of by load latency; partial sums to the column's owner       1 cluster barrier
//   X  pair sum, context columns written as the next GEMM's operand
// Same operands as attention_phase2 (hi / lo planes of the window, bf16 memory, fp32 accumulate); the softmax and context
// sums run in a different order.  (model.py:62-89, 403)
// ------------------------------------------------------------------------------------------------
struct AttnMulti {
  bool on;
  int nmax;                 // sequences per pair this launch needs
  int cr;                   // memory rows per chunk
  uint32_t slot_bytes;
  uint8_t* slot0;           // [3][slot_bytes]
  uint8_t *im0, *im1, *im2, *im3;   // im2col tiles, hi plane; the lo plane follows at + plane_bytes (no array: a run-time index
  uint32_t plane_bytes;             // would put the struct in local memory)
  GTN_DEVICE uint8_t* im(int it) const { return it == 0 ? im0 : (it == 1 ? im1 : (it == 2 ? im2 : im3)); }
};

// Placement of the batched attention's buffers as OFFSETS (ring base / receive-buffer base), computable on the host too:
// the launcher picks the FAST kernel instantiation only if the phase fits.
struct AttnPlan {
  bool on;
  int nmax, cr;
  uint32_t slot_bytes, plane_bytes;
  int o0, o1, o2, o3;       // im2col tile of sequence i: >= 0 offset in the ring, < 0: -(offset in the receive buffers) - 1
};
template <int N>
GTN_HD AttnPlan attn_multi_plan(const DecFwd& p, int S, int ncl) {
  constexpr int NC4 = (N < 64 ? N : 64) / 4;
  AttnPlan a;
  a.nmax = (p.B + 2 * ncl - 1) / (2 * ncl);
  const int JH = (p.Tin + 1) / 2, rows8 = (JH + 7) & ~7;
  a.cr = (JH + 2) / 3;
  a.slot_bytes = (uint32_t)(a.cr * p.E * 2 + 127) & ~127u;
  a.plane_bytes = (uint32_t)rows8 * 128u;
  a.on = !p.train && p.mem_bf16 && p.prec == 1 && (p.reserved0 & 2) == 0 && a.nmax <= am_max<N>() && p.E <= 512 && p.E % 8 == 0 &&
         p.Ha / 32 <= 32;
  // im2col tiles: behind the memory slots in the ring, then in the receive buffers (both idle during the phase); a tile is
  // read as 128 rows by the MMA whatever its row count, so 16 KB behind its base must lie inside the CTA's shared memory
  const uint32_t ring_bytes = (uint32_t)S * (uint32_t)(W_TILE_BYTES + N * KB * 2), recv_bytes = 2u * 4u * TILE_M * NC4 * 4u;
  uint32_t r1 = (3u * a.slot_bytes + 1023u) & ~1023u, r2 = 0u;
  int o[4];
#pragma unroll
  for (int it = 0; it < 4; ++it) {
    o[it] = 0;
    if (it >= a.nmax) continue;
    if (r1 + 2 * a.plane_bytes <= ring_bytes) { o[it] = (int)r1; r1 += 2 * a.plane_bytes; }
    else if (r2 + 2 * a.plane_bytes <= recv_bytes) { o[it] = -(int)r2 - 1; r2 += 2 * a.plane_bytes; }
    else a.on = false;
  }
  a.o0 = o[0]; a.o1 = o[1]; a.o2 = o[2]; a.o3 = o[3];
  if (3u * a.slot_bytes > ring_bytes) a.on = false;
  return a;
}

template <int N>
GTN_DEVICE AttnMulti attn_multi_layout(const DecFwd& p, const Smem2& sm, int S, int ncl) {
  const AttnPlan a = attn_multi_plan<N>(p, S, ncl);
  AttnMulti am;
  am.on = a.on; am.nmax = a.nmax; am.cr = a.cr; am.slot_bytes = a.slot_bytes; am.plane_bytes = a.plane_bytes;
  am.slot0 = sm.ring.base;
  uint8_t* rb = reinterpret_cast<uint8_t*>(sm.recv);
  am.im0 = a.o0 >= 0 ? sm.ring.base + a.o0 : rb + (-a.o0 - 1);
  am.im1 = a.o1 >= 0 ? sm.ring.base + a.o1 : rb + (-a.o1 - 1);
  am.im2 = a.o2 >= 0 ? sm.ring.base + a.o2 : rb + (-a.o2 - 1);
  am.im3 = a.o3 >= 0 ? sm.ring.base + a.o3 : rb + (-a.o3 - 1);
  return am;
}

GTN_DEVICE uint32_t am_tmem_col(int it) { return it == 0 ? 256u : (it == 1 ? 0u : (it == 2 ? 128u : 384u)); }

template <int N>
GTN_DEVICE void attention_infer_multi(const DecFwd& p, Smem2& sm, int S, int t, int rank, int cid, int ncl,
                                      uint32_t tmem_base, uint32_t& att_par, uint32_t& mc) {
  // (the layout is a few integer operations: recomputed per frame instead of ten values living in registers / local memory
  // across the whole frame)
  const AttnMulti am = attn_multi_layout<N>(p, sm, S, ncl);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Tin = p.Tin, A = p.A, KS = p.KS, hk = (p.KS - 1) / 2, E = p.E;
  const int half = rank & 1, peer = rank ^ 1;
  const int pair = cid * 2 + (rank >> 1), npairs = ncl * 2;
  const int JH = (Tin + 1) / 2;
  const int j0 = half * JH, jn = min(Tin, j0 + JH) - j0;
  const int Wd = jn + KS - 1;
  const bf16* mem = reinterpret_cast<const bf16*>(p.mem_enc);
  const bf16* pm = reinterpret_cast<const bf16*>(p.pm);
  int nact = 0;                                            // sequences of this pair (a prefix of the slots)
  for (int it = 0; it < am.nmax; ++it) nact += (pair + it * npairs) < p.B ? 1 : 0;
  // (a pair without sequences still walks through the phase: the cluster barriers below are shared with the other pair)
  float* q_m = sm.hs;                                      // [AM][128]        (idle outside the attention-LSTM epilogue)
  float* w_m = sm.wcat;                                    // [AM][Tin_pad]    (wcat | e_buf | e_part | w_s of the one-at-a-time path)
  float* epart = sm.red_s;                                 // [AM][2][128]     (dead before the context partials land)
  float* ctx_own = sm.red_s;                               // [AM][256]
  float* ctx_peer = sm.red_s + 4 * 256;                    // [AM][256]        (written by the peer CTA)
  const int* len_s = reinterpret_cast<const int*>(sm.scr + 48);
  const int nchunk = jn > 0 ? (jn + am.cr - 1) / am.cr : 0, total = nact * nchunk;
  const uint32_t mc0 = mc;

  // ---- Q: queries and windows of all sequences
  const int wq = warp & 3, a0 = (warp >> 2) * 64, jl = wq * 32 + lane;         // energies: thread = (position, 64-column half)
  uint4 pmn[8];
  auto pm_fetch = [&](int it) {
    const int b = pair + it * npairs;
    if (it < nact && jl < jn && (j0 + jl) < len_s[it]) {
      if (p.reserved0 & 32) {               // chunk-major [B][A / 8][Tin][8]
        const uint4* src = reinterpret_cast<const uint4*>(pm) + ((size_t)b * (A >> 3) + (a0 >> 3)) * Tin + (j0 + jl);
#pragma unroll
        for (int g = 0; g < 8; ++g)
          if (a0 + g * 8 < A) pmn[g] = __ldg(src + (size_t)g * Tin);
      } else {
        const uint4* src = reinterpret_cast<const uint4*>(pm + ((size_t)b * Tin + (j0 + jl)) * A + a0);
#pragma unroll
        for (int g = 0; g < 8; ++g)
          if (a0 + g * 8 < A) pmn[g] = __ldg(src + g);
      }
    }
  };
  pm_fetch(0);
  {
    const int ntq = p.Ha / 32;
    const size_t qs = (size_t)p.B * A;
    for (int i = threadIdx.x; i < nact * 128; i += blockDim.x) {
      const int it = i >> 7, a = i & 127, b = pair + it * npairs;
      float s = 0.f;
      if (a < A) {
        const float* qp = p.qpart + (size_t)b * A + a;
        float qv[32];
#pragma unroll
        for (int c = 0; c < 32; ++c) qv[c] = c < ntq ? __ldcg(qp + (size_t)c * qs) : 0.f;
#pragma unroll
        for (int c = 0; c < 16; ++c) qv[c] += qv[c + 16];
#pragma unroll
        for (int c = 0; c < 8; ++c) qv[c] += qv[c + 8];
#pragma unroll
        for (int c = 0; c < 4; ++c) qv[c] += qv[c + 4];
        s = (qv[0] + qv[1]) + (qv[2] + qv[3]);
      }
      q_m[i] = s;
    }
    for (int i = threadIdx.x; i < nact * 2 * Wd; i += blockDim.x) {
      const int it = i / (2 * Wd), r = i - it * 2 * Wd, ch = r / Wd, jj = r - ch * Wd, j = j0 + jj - hk;
      const int b = pair + it * npairs;
      const float* src = (ch == 0 ? p.w_prev : p.w_cum) + (size_t)b * p.ldT;
      const float wv = (j >= 0 && j < Tin) ? __ldcg(src + j) : 0.f;
      sm.am_wcat[it * AM_WCAT + r] = wv;
      // hi / lo bf16 planes of the window element, split ONCE (an element appears in ~KS rows of the im2col tile):
      // low half = hi plane bits, high half = lo plane bits.  The memory slots are free until the chunks are requested.
      const bf16 whi = __float2bfloat16_rn(wv);
      const bf16 wlo = __float2bfloat16_rn(wv - __bfloat162float(whi));
      reinterpret_cast<uint32_t*>(am.slot0)[it * AM_WCAT + r] = (uint32_t)__bfloat16_as_ushort(whi) | ((uint32_t)__bfloat16_as_ushort(wlo) << 16);
    }
  }
  __syncthreads();
  // ---- I: im2col tiles
  // row r = local encoder position, K index kk = ch KS + k  ->  window element ch Wd + r + k; 16-byte chunk = 8 K elements
  for (int i = threadIdx.x; i < nact * jn * 8; i += blockDim.x) {
    const int it = i / (jn * 8), rc = i - it * jn * 8, r = rc >> 3, chunk = rc & 7;
    const uint32_t* sp = reinterpret_cast<const uint32_t*>(am.slot0) + it * AM_WCAT + r;
    uint32_t sv[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int kk = chunk * 8 + e;
      sv[e] = kk < 2 * KS ? sp[kk + (kk >= KS ? Wd - KS : 0)] : 0u;
    }
    uint4 hi, lo;
    hi.x = __byte_perm(sv[0], sv[1], 0x5410); lo.x = __byte_perm(sv[0], sv[1], 0x7632);
    hi.y = __byte_perm(sv[2], sv[3], 0x5410); lo.y = __byte_perm(sv[2], sv[3], 0x7632);
    hi.z = __byte_perm(sv[4], sv[5], 0x5410); lo.z = __byte_perm(sv[4], sv[5], 0x7632);
    hi.w = __byte_perm(sv[6], sv[7], 0x5410); lo.w = __byte_perm(sv[6], sv[7], 0x7632);
    uint8_t* tile = am.im(it);
    const uint32_t off = sw128_off(r, chunk);
    *reinterpret_cast<uint4*>(tile + off) = hi;
    *reinterpret_cast<uint4*>(tile + am.plane_bytes + off) = lo;
  }
  fence_proxy_async_smem();
  __syncthreads();
  // ---- M: location GEMMs (warp 1), first memory chunks (warp 0); warp-uniform control flow, one elected lane issues
  auto issue_chunk = [&](int gi) {                         // chunk gi of this frame: sequence gi / nchunk, rows (gi % nchunk) cr ..
    const int it = gi / nchunk, ch = gi - it * nchunk, b = pair + it * npairs;
    const int r0 = ch * am.cr, nr = min(am.cr, jn - r0);
    const uint32_t slot = (mc0 + (uint32_t)gi) % 3u, bytes = (uint32_t)nr * (uint32_t)E * 2u;
    mbar_expect_tx(&sm.am_mem_full[slot], bytes);
    if (N > 64 && (p.reserved0 & 8) == 0) {
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                   ::"r"(smem_u32(am.slot0 + (size_t)slot * am.slot_bytes)), "l"(mem + ((size_t)b * Tin + j0 + r0) * E), "r"(bytes),
                     "r"(smem_u32(&sm.am_mem_full[slot])), "l"(l2_policy_evict_first()) : "memory");
    } else {
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem_u32(am.slot0 + (size_t)slot * am.slot_bytes)), "l"(mem + ((size_t)b * Tin + j0 + r0) * E), "r"(bytes),
                     "r"(smem_u32(&sm.am_mem_full[slot])) : "memory");
    }
  };
  {
    const int warp_u = __shfl_sync(0xffffffffu, warp, 0);
    if (warp_u == 1) {
      tc_fence_after();
      const uint32_t idesc = umma_idesc_bf16(TILE_M, 128);
      const uint64_t duh = umma_desc_sw128(smem_u32(sm.u_s)), dul = umma_desc_sw128(smem_u32(sm.u_s + W_TILE_BYTES));
      for (int it = 0; it < nact; ++it) {
        const uint32_t tacc = tmem_base + am_tmem_col(it);
        const uint64_t dah = umma_desc_sw128(smem_u32(am.im(it))), dal = umma_desc_sw128(smem_u32(am.im(it) + am.plane_bytes));
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < KB / 16; ++k) umma_bf16(tacc, dah + (uint64_t)(2 * k), duh + (uint64_t)(2 * k), idesc, (uint32_t)(k > 0));
#pragma unroll
          for (int k = 0; k < KB / 16; ++k) umma_bf16(tacc, dal + (uint64_t)(2 * k), duh + (uint64_t)(2 * k), idesc, 1u);
#pragma unroll
          for (int k = 0; k < KB / 16; ++k) umma_bf16(tacc, dah + (uint64_t)(2 * k), dul + (uint64_t)(2 * k), idesc, 1u);
          umma_commit(&sm.am_att_full[it]);
        }
        __syncwarp();
      }
    } else if (warp_u == 0) {
      if (elect_one())
        for (int gi = 0; gi < total && gi < 3; ++gi) issue_chunk(gi);
      __syncwarp();
    }
  }
  PROF(24);
  // ---- E: energies, sequence by sequence as the accumulators complete
  for (int it = 0; it < nact; ++it) {
    uint4 pmq[8];
#pragma unroll
    for (int g = 0; g < 8; ++g) pmq[g] = pmn[g];
    pm_fetch(it + 1);
    wait_bar(&sm.am_att_full[it], att_par);
    tc_fence_after();
    const bool valid = jl < jn && (j0 + jl) < len_s[it];
    const float* qv = q_m + it * 128;
    float part = 0.f;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_base + am_tmem_col(it) + ((uint32_t)(wq * 32) << 16) + (uint32_t)(a0 + h * 32), r);
      tmem_ld_wait();
      if (valid) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int ab = h * 32 + g * 8;
          if (a0 + ab < A) {
            const uint4 qq = pmq[h * 4 + g];
            const uint32_t ww[4] = {qq.x, qq.y, qq.z, qq.w};
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const float pmv = (e & 1) ? __uint_as_float(ww[e >> 1] & 0xFFFF0000u) : __uint_as_float(ww[e >> 1] << 16);
              const float sv = __uint_as_float(r[g * 8 + e]) + qv[a0 + ab + e] + pmv;
              part += sm.v_s[a0 + ab + e] * tanh_fast(sv);
            }
          }
        }
      }
    }
    epart[(it * 2 + (warp >> 2)) * 128 + jl] = valid ? part : -INFINITY;
  }
  att_par ^= 1u;
  tc_fence_before();
  __syncthreads();
  {
    const uint32_t peer_e = map_to_rank(smem_u32(sm.am_ebuf), (uint32_t)peer);
    for (int i = threadIdx.x; i < nact * jn; i += blockDim.x) {
      const int it = i / jn, l = i - it * jn;
      const float e = epart[(it * 2) * 128 + l] + epart[(it * 2 + 1) * 128 + l];      // -inf + -inf = -inf for masked positions
      sm.am_ebuf[it * TIN_MAX + j0 + l] = e;
      st_cluster_f1(peer_e + (uint32_t)((it * TIN_MAX + j0 + l) * 4), e);
    }
  }
  PROF(25);
  cluster_sync_all();
  PROF(26);
  // ---- S: masked softmax over all encoder positions, one warp per sequence (both CTAs of the pair compute it redundantly)
  if (warp < nact) {
    const int it = warp, b = pair + it * npairs;
    const float* eb = sm.am_ebuf + it * TIN_MAX;
    float ev[TIN_MAX / 32];
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < TIN_MAX / 32; ++i) {
      const int j = lane + 32 * i;
      ev[i] = j < Tin ? eb[j] : -INFINITY;
      mx = fmaxf(mx, ev[i]);
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < TIN_MAX / 32; ++i) {
      ev[i] = ev[i] == -INFINITY ? 0.f : expf(ev[i] - mx);
      sum += ev[i];
    }
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
    float* wprev = p.w_prev + (size_t)b * p.ldT;
    float* wcum = p.w_cum + (size_t)b * p.ldT;
    float* aout = p.align_out + ((size_t)b * p.n_steps + t) * Tin;
    const float* cum_before = sm.am_wcat + it * AM_WCAT + Wd + hk;        // channel 1 of the window, position j0 + l
#pragma unroll
    for (int i = 0; i < TIN_MAX / 32; ++i) {
      const int j = lane + 32 * i;
      if (j < Tin) {
        const float w = ev[i] * inv;
        w_m[it * TIN_MAX + j] = w;
        if (j >= j0 && j < j0 + jn) {
          wprev[j] = w;
          wcum[j] = cum_before[j - j0] + w;                               // model.py:403
          aout[j] = w;
        }
      }
    }
  }
  __syncthreads();
  PROF(27);
  // ---- C: partial context over this CTA's positions, all E columns: thread = one bf16 column pair
  {
    const int E2 = E >> 1;                                 // 32-bit words per memory row
    const int Eh = ((E / 8 + 1) / 2) * 8;                  // columns [0, Eh) belong to half 0, [Eh, E) to half 1
    const bool has_col = (int)threadIdx.x < E2;
    const int c = 2 * threadIdx.x;
    const uint32_t peer_ctx = map_to_rank(smem_u32(ctx_peer), (uint32_t)peer);
    int gi = 0;
    for (int it = 0; it < nact; ++it) {
      float acc0 = 0.f, acc1 = 0.f;
      const float* wv = w_m + it * TIN_MAX + j0;
      for (int ch = 0; ch < nchunk; ++ch, ++gi) {
        const uint32_t k = mc0 + (uint32_t)gi, slot = k % 3u;
        wait_bar(&sm.am_mem_full[slot], (k / 3u) & 1u);
        const uint32_t* rows = reinterpret_cast<const uint32_t*>(am.slot0 + (size_t)slot * am.slot_bytes) + threadIdx.x;
        const int r0 = ch * am.cr, nr = min(am.cr, jn - r0);
        if (has_col) {
#pragma unroll 5
          for (int r = 0; r < nr; ++r) {
            const uint32_t m2 = rows[(size_t)r * E2];
            const float w = wv[r0 + r];
            acc0 += w * __uint_as_float(m2 << 16);
            acc1 += w * __uint_as_float(m2 & 0xFFFF0000u);
          }
        }
        __syncthreads();                                   // the slot has been read by every thread
        if (gi + 3 < total && __shfl_sync(0xffffffffu, warp, 0) == 0) {
          if (elect_one()) issue_chunk(gi + 3);
          __syncwarp();
        }
      }
      if (has_col) {
        const bool mine = (c < Eh) == (half == 0);
        const int cl = c < Eh ? c : c - Eh;                // index inside the owner's column range
        if (mine) { ctx_own[it * 256 + cl] = acc0; ctx_own[it * 256 + cl + 1] = acc1; }
        else {
          st_cluster_f1(peer_ctx + (uint32_t)((it * 256 + cl) * 4), acc0);
          st_cluster_f1(peer_ctx + (uint32_t)((it * 256 + cl + 1) * 4), acc1);
        }
      }
    }
    mc = mc0 + (uint32_t)total;
    cluster_sync_all();
    // ---- X: pair sum -> context operand of the next GEMMs
    const int c_lo = half == 0 ? 0 : Eh, ncol = half == 0 ? Eh : E - Eh;
    for (int i = threadIdx.x; i < nact * ncol; i += blockDim.x) {
      const int it = i / ncol, cl = i - it * ncol, b = pair + it * npairs;
      const float sv = half == 0 ? ctx_own[it * 256 + cl] + ctx_peer[it * 256 + cl] : ctx_peer[it * 256 + cl] + ctx_own[it * 256 + cl];
      BFP(p.xctx)[(size_t)b * p.Ep + c_lo + cl] = __float2bfloat16_rn(sv);
    }
  }
  __syncthreads();
}

// Host and device agree on what the FAST instantiation covers
template <int N>
GTN_HD bool fwd2_fast_covers(const DecFwd& p, int S, int ncl) {
  return !p.train && p.B > 4 /* PRE_LOCAL_BMAX */ && (p.reserved0 & 3) == 0 && tail_covers(p, (p.B + ncl - 1) / ncl) &&
         attn_multi_plan<N>(p, S, ncl).on;
}

template <int N, bool FAST>
__global__ void __launch_bounds__(256, 1)
decoder_fwd2_kernel(const __grid_constant__ Fwd2Maps maps, const DecFwd p, const int S) {
  extern __shared__ uint8_t smem_raw[];
  // align by pointer arithmetic (no integer round trip) so that the compiler keeps the shared address space: LDS / STS, not generic LD / ST
  uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  Smem2 sm = carve2<N>(base, S);
  const int warp = threadIdx.x >> 5;
  const int rank = (int)cluster_ctarank(), cid = (int)cluster_idx(), ncl = (int)cluster_count();

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(&sm.ring.full[s], 1); mbar_init(&sm.ring.empty[s], 1); }
    mbar_init(sm.acc_full, 1);
    mbar_init(sm.att_full, 1);
    mbar_init(sm.u_full, 1);
    mbar_init(&sm.recv_full[0], 1);
    mbar_init(&sm.recv_full[1], 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc(sm.tmem_slot, 512);
  for (int i = threadIdx.x; i < p.A; i += blockDim.x) sm.v_s[i] = p.v[i];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *sm.tmem_slot;
  const uint32_t tmem_att = tmem_base + 256u;
  if (threadIdx.x == 0) {
    // the pre-multiplied location filter, hi and lo planes: resident for the whole kernel
    mbar_expect_tx(sm.u_full, 2 * W_TILE_BYTES);
    tma_load_3d(sm.u_s, &maps.uloc, sm.u_full, 0, 0, 0);
    tma_load_3d(sm.u_s + W_TILE_BYTES, &maps.uloc, sm.u_full, KB, 0, 0);
    wait_bar(sm.u_full, 0);
  }
  bool am_on;
  int am_nmax;
  {
    const AttnMulti am0 = attn_multi_layout<N>(p, sm, S, ncl);
    am_on = am0.on; am_nmax = am0.nmax;
  }
  if (threadIdx.x == 0 && am_on) {
    for (int i = 0; i < 4; ++i) mbar_init(&sm.am_att_full[i], 1);
    for (int i = 0; i < 3; ++i) mbar_init(&sm.am_mem_full[i], 1);
    mbar_fence_init();
  }
  // (the batched attention's fourth accumulator takes the tensor-memory columns of the last four resident slots)
  if (p.Ha / 32 == ncl && p.Hd / 32 == ncl && !(am_on && am_nmax > 3)) {
    // one attention-LSTM tile and one decoder-LSTM tile per cluster for the whole launch: park weight k-blocks of this
    // CTA's K quarters in the free TMEM columns, split so that each phase's remaining stages fit the ring in one round
    const int nka = p.mult * (p.Pp + p.Ep + p.Hap) / KB, nkd = p.mult * (p.Hap + p.Ep + p.Hdp) / KB;
    const int ka0 = rank * nka / CL, kd0 = rank * nkd / CL;
    sm.res_a = min((rank + 1) * nka / CL - ka0, (res_slots<N>() * 2 + 2) / 5);
    sm.res_d = min((rank + 1) * nkd / CL - kd0, res_slots<N>() - sm.res_a);
    load_resident<N>(p.w_att2, (long long)nka * KB, 4 * p.Ha, cid * TILE_M, ka0, sm.res_a, 0, tmem_base);
    load_resident<N>(p.w_dec2, (long long)nkd * KB, 4 * p.Hd, cid * TILE_M, kd0, sm.res_d, sm.res_a, tmem_base);
  }
  if (threadIdx.x < 8) {
    // encoder lengths of the sequences this CTA pair serves (constant for the launch; a global load per frame otherwise)
    const int b = cid * 2 + (rank >> 1) + (int)threadIdx.x * ncl * 2;
    reinterpret_cast<int*>(sm.scr + 48)[threadIdx.x] = b < p.B ? (p.lengths ? min(p.lengths[b], p.Tin) : p.Tin) : 0;
  }
  tc_fence_before();
  cluster_sync_all();     // peers' barriers / smem are initialised before any DSMEM traffic
  tc_fence_after();

  unsigned int epoch = 0;
  uint32_t acc_par = 0, att_par = 0, round = 0, eround = 0, am_par = 0, am_mc = 0;
  const unsigned int nblocks = gridDim.x;
  int t = 0;
  // The GEMM phases of a frame go through ONE call site of gemm_phase (runtime `kind`): five specialised inlined copies made
  // the per-frame code path larger than the instruction cache holds, and a phase that runs once per frame then runs cold.
  const bool pre_local = !FAST && !P_TRAIN && p.B <= PRE_LOCAL_BMAX;
  // batched inference, bf16 mode: projection + prenet inside each cluster for its own sequences (tail_phase)
  const int nbc = (p.B + ncl - 1) / ncl;
  const bool tail = FAST || (!P_TRAIN && !pre_local && (p.reserved0 & 1) == 0 && tail_covers(p, nbc));
  if (FAST && !fwd2_fast_covers<N>(p, S, ncl)) __trap();   // (the launcher checked the same predicate)
  const int k_first = (P_TRAIN || pre_local || tail) ? K_ATT : K_PRE1, k_last = (P_TRAIN || tail) ? K_DEC : K_PROJ;
  bool stopped = false;
  if (tail) {
    tail_phase(p, sm, 0, false, rank, cid, ncl);           // prenet of the go frame
    grid_sync2(p.barrier, epoch, nblocks);
  }
  for (; t < p.n_steps && !stopped; ++t) {
    if (tail) PROF(28);
    if constexpr (!FAST) {
      if (pre_local) {
        PROF(28);
        prenet_local(p, sm, t, rank);
        PROF(31);
      }
    }
#pragma unroll 1
    for (int kind = k_first; kind <= k_last; ++kind) {
      if (kind == K_PRE1) PROF(28);
      if (kind == K_ATT) PROF(0);
      gemm_phase<N, FAST>(p, maps, sm, kind, t, rank, cid, ncl, tmem_base, acc_par, round);
      PROF(kind == K_PRE1 ? 29 : (kind == K_PRE2 ? 31 : (kind == K_ATT ? 4 : (kind == K_DEC ? 13 : 19))));
      if (kind == K_DEC && P_TRAIN) break;        // C(t) and A(t+1) share an interval: no barrier in between
      grid_sync2(p.barrier, epoch, nblocks);
      PROF(kind == K_PRE1 ? 30 : (kind == K_ATT ? 5 : (kind == K_DEC ? 18 : 23)));
      if (kind == K_ATT) {
        if (FAST || am_on) attention_infer_multi<N>(p, sm, S, t, rank, cid, ncl, tmem_base, am_par, am_mc);
        if constexpr (!FAST) {
          if (am_on) {}
          else if (p.mem_bf16) attention_phase2<bf16>(p, sm, t, rank, cid, ncl, tmem_att, att_par, eround);
          else attention_phase2<float>(p, sm, t, rank, cid, ncl, tmem_att, att_par, eround);
        }
        PROF(6);
        grid_sync2(p.barrier, epoch, nblocks);
        PROF(8);
      }
      if (!FAST && kind == K_PROJ) {
        if constexpr (N > 64) {
          // batch-chunked projection: done[] was written by several clusters before the barrier above; every CTA decides
          int all = 1;
          for (int b = threadIdx.x; b < p.B; b += blockDim.x) all &= (__ldcg(&p.done[b]) != 0);
          all = __syncthreads_and(all);
          if (all && !p.ignore_stop) {
            stopped = true;
            if (blockIdx.x == 0 && threadIdx.x == 0) *p.stop = 1;
          }
        } else if (__ldcg(p.stop) != 0) {
          stopped = true;
        }
      }
    }
    if (tail) {
      tail_phase(p, sm, t, true, rank, cid, ncl);
      PROF(19);
      grid_sync2(p.barrier, epoch, nblocks);
      PROF(23);
      // done[] was written by every cluster before the barrier above; every CTA decides (model.py:486)
      int all = 1;
      for (int b = threadIdx.x; b < p.B; b += blockDim.x) all &= (__ldcg(&p.done[b]) != 0);
      all = __syncthreads_and(all);
      if (all && !p.ignore_stop) {
        stopped = true;
        if (blockIdx.x == 0 && threadIdx.x == 0) *p.stop = 1;
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && p.steps_done) *p.steps_done = t;
  tc_fence_before();
  cluster_sync_all();     // no CTA exits while a peer may still write into its shared memory
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
template <int N>
static int launch_fwd2(const Fwd2Maps& maps, const DecFwd& p, int max_clusters_wanted, cudaStream_t st) {
  const size_t stage = (size_t)W_TILE_BYTES + (size_t)N * KB * 2;
  const size_t fixed = fwd2_fixed_bytes<N>() + 1024;
  int S = (int)((227 * 1024 - fixed) / stage);
  if (S > 12) S = 12;
  GTN_REQUIRE(S >= 2, "decoder forward: not enough shared memory for the TMA ring (N=%d)", N);
  const size_t smem = fixed + (size_t)S * stage;
  auto kern = decoder_fwd2_kernel<N, false>;
  auto kern_fast = decoder_fwd2_kernel<N, true>;
  GTN_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  GTN_CUDA_OK(cudaFuncSetAttribute(kern_fast, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = CL; attrs[0].val.clusterDim.y = 1; attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeCooperative;
  attrs[1].val.cooperative = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  cfg.gridDim = dim3(CL * max_clusters_wanted);
  int max_cl = 0, max_cl_fast = 0;
  GTN_CUDA_OK(cudaOccupancyMaxActiveClusters(&max_cl, kern, &cfg));
  GTN_CUDA_OK(cudaOccupancyMaxActiveClusters(&max_cl_fast, kern_fast, &cfg));
  GTN_REQUIRE(max_cl >= 1, "decoder forward: no cluster of %d CTAs fits on this device", CL);
  const int ncl = max_cl < max_clusters_wanted ? max_cl : max_clusters_wanted;
  // each CTA pair stages the encoder lengths of at most 8 sequences (scr + 48): B <= 16 * clusters
  GTN_REQUIRE((p.B + 2 * ncl - 1) / (2 * ncl) <= 8, "decoder forward: batch %d needs more than 8 sequences per CTA pair on %d clusters",
              p.B, ncl);
  cfg.gridDim = dim3(CL * ncl);
  cfg.numAttrs = profiler_attached() ? 1 : 2;   // profilers that cannot replay cooperative launches
  // batched bf16 inference on the cluster-local tail and the all-at-once attention: the instantiation without the rest
  const bool fast = max_cl_fast >= ncl && !getenv("GTN_FWD2_GENERIC") && fwd2_fast_covers<N>(p, S, ncl);
  GTN_CUDA_OK(cudaLaunchKernelEx(&cfg, fast ? kern_fast : kern, maps, p, S));
  count_launch();
  return 0;
}

}  // namespace gtn

using namespace gtn;

extern "C" int gtn_decoder_forward2(const gtn_decoder_fwd_params* hp, void* stream) {
  GTN_REQUIRE(hp != nullptr, "null params");
  DecFwd p = *hp;
  GTN_REQUIRE(p.prec == 1 || p.prec == 3, "prec must be 1 (bf16) or 3 (fp32 via 3 bf16 planes)");
  p.mult = p.prec == 1 ? 1 : 6;
  GTN_REQUIRE(p.Ha % 32 == 0 && p.Hd % 32 == 0, "LSTM widths must be multiples of 32");
  GTN_REQUIRE(p.Hap % 64 == 0 && p.Hdp % 64 == 0 && p.Dp % 64 == 0 && p.Pp % 64 == 0 && p.NMp % 64 == 0,
              "padded K segments must be multiples of 64");
  GTN_REQUIRE(p.Hap >= p.Ha && p.Hdp >= p.Hd && p.Dp >= p.D && p.Pp >= p.P && p.NMp >= p.NM, "padded dims too small");
  GTN_REQUIRE(p.D == p.E + p.Cc && p.E % 8 == 0, "D must equal E + Cc and E must be a multiple of 8");
  GTN_REQUIRE(p.KS % 2 == 1 && 2 * p.KS <= 64, "location kernel size must be odd and <= 31");
  GTN_REQUIRE(p.A <= 128 && p.A % 8 == 0, "attention_dim must be a multiple of 8 and <= 128");
  GTN_REQUIRE(p.Tin >= 1 && p.Tin <= TIN_MAX, "T_in must be in [1, %d]", TIN_MAX);
  GTN_REQUIRE(p.B >= 1 && p.B <= 256 && p.n_steps >= 1, "batch must be in [1, 256]");
  GTN_REQUIRE(p.NM + 1 <= TILE_M, "n_mel + 1 must fit one 128-row tile");
  GTN_REQUIRE(p.w_att2 && p.w_dec2 && p.u_loc && p.wq_t && p.gbias_a && p.gbias_d, "second-generation weight packs missing");
  GTN_REQUIRE(p.Ep % 64 == 0 && p.Ep >= p.E && (p.train || p.pbias), "Ep must be E padded to 64 (and pbias set for inference)");
  auto thr = [](float pr) -> unsigned { double t = (double)pr * 4294967296.0; return t >= 4294967295.0 ? 0xFFFFFFFFu : (unsigned)t; };
  p.thr_att = p.use_dropout ? thr(p.p_att) : 0u;
  p.thr_dec = p.use_dropout ? thr(p.p_dec) : 0u;
  p.reserved0 = hp->reserved0 & 32;                 // caller flag: chunk-major pm; the other bits are kernel feature switches
  GTN_REQUIRE(!(p.reserved0 & 32) || (p.mem_bf16 && p.A % 8 == 0), "chunk-major pm needs bf16 pm and attention_dim % 8 == 0");
  if (getenv("GTN_INFER_NO_TAIL")) p.reserved0 |= 1;
  if (getenv("GTN_INFER_NO_MULTI_ATTN")) p.reserved0 |= 2;
  if (getenv("GTN_INFER_NO_L2_HINTS")) p.reserved0 |= 8;
  const int N = p.B <= 32 ? 32 : (p.B <= 64 ? 64 : (p.B <= 128 ? 128 : 256));
  const int m = p.mult;
  const long long Ka = (long long)m * (p.Pp + p.Ep + p.Hap), Kd = (long long)m * (p.Hap + p.Ep + p.Hdp);
  Fwd2Maps maps;
  int rc = 0;
  rc |= make_tma_map_3d(&maps.w_att, p.w_att2, Ka, 4 * p.Ha, 1, Ka, 0, TILE_M);
  rc |= make_tma_map_3d(&maps.w_dec, p.w_dec2, Kd, 4 * p.Hd, 1, Kd, 0, TILE_M);
  rc |= make_tma_map_3d(&maps.uloc, p.u_loc, 2 * KB, p.A, 1, 2 * KB, 0, TILE_M);
  // xctx is double buffered [2][B][m Ep] (the pipelined training kernel alternates; this file's kernel uses buffer 0)
  rc |= make_tma_map_3d(&maps.ctx, p.xctx, (long long)m * p.Ep, p.B, 2, (long long)m * p.Ep, (long long)p.B * m * p.Ep, N);
  rc |= make_tma_map_3d(&maps.ha, p.xh_a, (long long)m * p.Hap, p.B, 2, (long long)m * p.Hap, (long long)p.B * m * p.Hap, N);
  rc |= make_tma_map_3d(&maps.hd, p.xh_d, (long long)m * p.Hdp, p.B, 2, (long long)m * p.Hdp, (long long)p.B * m * p.Hdp, N);
  if (p.train) {
    rc |= make_tma_map_3d(&maps.pre, p.pre_all, (long long)m * p.Pp, p.B, p.n_steps + 1, (long long)m * p.Pp,
                          (long long)p.B * m * p.Pp, N);
    if (p.Ha / 32 <= 256)
      rc |= make_tma_map_f32_3d(&maps.qpart, p.qpart, p.A, p.B, p.Ha / 32, p.A, (long long)p.B * p.A, p.A, 1, p.Ha / 32);
    maps.w_pre1 = maps.w_pre2 = maps.w_proj = maps.w_att;
    maps.xmel = maps.xp1 = maps.xp2 = maps.ctx;
  } else {
    const long long Kp = (long long)m * (p.Hdp + p.Ep);
    rc |= make_tma_map_3d(&maps.w_pre1, p.w_pre1, (long long)m * p.NMp, p.P, 1, (long long)m * p.NMp, 0, TILE_M);
    rc |= make_tma_map_3d(&maps.w_pre2, p.w_pre2, (long long)m * p.Pp, p.P, 1, (long long)m * p.Pp, 0, TILE_M);
    rc |= make_tma_map_3d(&maps.w_proj, p.w_proj, Kp, p.NM + 1, 1, Kp, 0, TILE_M);
    rc |= make_tma_map_3d(&maps.xmel, p.xmel, (long long)m * p.NMp, p.B, 1, (long long)m * p.NMp, 0, N);
    rc |= make_tma_map_3d(&maps.xp1, p.xp1, (long long)m * p.Pp, p.B, 1, (long long)m * p.Pp, 0, N);
    rc |= make_tma_map_3d(&maps.xp2, p.xp2, (long long)m * p.Pp, p.B, 1, (long long)m * p.Pp, 0, N);
    maps.pre = maps.xp2;
    maps.qpart = maps.ctx;
  }
  maps.xmel_c = maps.xmel; maps.xp1_c = maps.xp1; maps.hd_c = maps.hd; maps.ctx_c = maps.ctx;
  if (!p.train && N > 64) {
    rc |= make_tma_map_3d(&maps.xmel_c, p.xmel, (long long)m * p.NMp, p.B, 1, (long long)m * p.NMp, 0, 64);
    rc |= make_tma_map_3d(&maps.xp1_c, p.xp1, (long long)m * p.Pp, p.B, 1, (long long)m * p.Pp, 0, 64);
    rc |= make_tma_map_3d(&maps.hd_c, p.xh_d, (long long)m * p.Hdp, p.B, 2, (long long)m * p.Hdp, (long long)p.B * m * p.Hdp, 64);
    rc |= make_tma_map_3d(&maps.ctx_c, p.xctx, (long long)m * p.Ep, p.B, 2, (long long)m * p.Ep, (long long)p.B * m * p.Ep, 64);
  }
  if (rc) return rc;
  const int want = (p.Ha > p.Hd ? p.Ha : p.Hd) / 32;
  cudaStream_t st = (cudaStream_t)stream;
  if (p.train) {
    // teacher-forced pass: the software-pipelined kernel when the shape is covered (B <= 64, one LSTM tile pair per cluster)
    const int rc3 = launch_train_fwd3(maps, p, st);
    if (rc3 != FWD3_NOT_COVERED) return rc3;
  }
  if (N == 32) return launch_fwd2<32>(maps, p, want, st);
  if (N == 64) return launch_fwd2<64>(maps, p, want, st);
  if (N == 128) return launch_fwd2<128>(maps, p, want, st);
  return launch_fwd2<256>(maps, p, want, st);
}
