// Persistent clustered decoder kernel, teacher-forced TRAINING forward, software-pipelined (third generation).
// Same arithmetic, operands, saves and Philox sites as the train = 1 path of gtn_decoder_fwd2.cu (Decoder.forward,
// reference model.py:419-454; decode() model.py:378-417), but the frame is scheduled around its ONLY serial chain:
//
//     attention LSTM A(t)  ->  all CTAs  ->  attention B(t)  ->  all CTAs  ->  A(t+1) ...
//
// The decoder LSTM C(t) (model.py:404-409) feeds nothing but C(t+1) and the mel/gate projection (which runs after the
// loop), so it is taken off that chain and executed one frame late in the shadow of the grid barriers:
//
//   W1  wait barrier 2 of frame t-1           (ctx(t-1), h_dec(t-2) visible)
//       issue the LATE k-blocks of A(t)       (context columns: the only operand that barrier protects) -> acc_A
//       issue the LATE k-blocks of C(t-1)     (context + h_dec(t-2))                                      -> acc_C
//   W2  A(t): split-K exchange, LSTM cells, dropout, saves, partial attention query
//   W3  ARRIVE barrier 1 of frame t
//   W4  C(t-1): split-K exchange, LSTM cells, saves          } work that does not depend on barrier 1, done while the
//   W5  B(t) preparation: im2col of [prev | cum] weights     } other CTAs arrive
//       (kept in shared memory), location GEMM on tcgen05
//   W6  WAIT barrier 1
//   W7  B(t): query gather (bulk copies), energies, masked softmax, context
//   W8  ARRIVE barrier 2 of frame t
//       issue the EARLY k-blocks of A(t+1) and C(t): prenet(t+1), h_att(t) -- operands that were complete before
//       barrier 2 -- while the other CTAs arrive
//
// The TMA loads and tcgen05.mma of the per-frame GEMMs are issued by two DEDICATED warps (8: producer, 9: MMA issuer) that run
// their own instruction stream, gated by "go" mbarriers the workers arrive on; the eight worker warps synchronise among
// themselves on named barrier 1 and only ever wait for a finished accumulator.  So the operand streaming of C(t-1) (10
// k-blocks per CTA) runs underneath the attention-LSTM epilogue W2 instead of in front of it.
//
// Every per-frame GEMM keeps its cluster split (4 CTAs x K quarter, gtn_decoder2.cuh) but a CTA's quarter is now made of
// a quarter of EVERY activation segment, ordered early -> late, so that only 2 (A) + 6 (C) k-blocks per CTA remain behind
// barrier 2; their weight k-blocks are the ones parked in tensor memory, so the post-barrier chain of A(t) is one 16 KB
// activation load + 8 MMAs.  acc_C aliases the tensor-memory columns of the location GEMM (their lifetimes are disjoint).
//
// Covers B <= 64 (one sequence per CTA pair), attention_rnn_dim / 32 == decoder_rnn_dim / 32 == number of clusters;
// anything else runs the sequential kernel of gtn_decoder_fwd2.cu.
#include <type_traits>
#include "gtn_decoder_fwd.cuh"

namespace gtn {

#define PROF3(slot)                                                                         \
  do {                                                                                      \
    if (p.prof != nullptr && blockIdx.x == 0 && threadIdx.x == 0 && t < 16) p.prof[t * 32 + (slot)] = clock64(); \
  } while (0)

#define WSYNC() asm volatile("bar.sync 1, 256;" ::: "memory")      /* the eight worker warps */
static constexpr int NWORK = 256;      // worker threads; warps 8 and 9 are the TMA producer and the MMA issuer
static constexpr int NTHREADS = 384;   // three warpgroups: two of workers (232 registers), one of issuers (40 registers)
// per-CTA skew probe: %globaltimer of every CTA at four points of frame 8 -> prof[512 + 4 cta + k]
#define PROFG(k)                                                                            \
  do {                                                                                      \
    if (p.prof != nullptr && threadIdx.x == 0 && t == 8) {                                  \
      unsigned long long gt_;                                                               \
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt_));                                \
      p.prof[512 + 4 * blockIdx.x + (k)] = (long long)gt_;                                  \
    }                                                                                       \
  } while (0)
static constexpr int KMAX = 64;        // k-blocks of one GEMM per CTA (fp32 mode: 6 planes x 10)

struct KSched {            // this CTA's k-blocks of one per-frame GEMM, in issue order
  int n, n_early;          // [0, n_early): operands complete before the barrier that precedes the GEMM's late part
  short wk[KMAX];          // k-block column in the packed weight matrix
  short xk[KMAX];          // k-block inside its activation segment
  signed char seg[KMAX];   // activation segment 0..2
  signed char slot[KMAX];  // tensor-memory slot of a resident weight k-block, -1 = streamed by TMA
};

struct Smem3 {
  Ring ring;
  uint8_t* u_s;        // [2][16 KB]  U_hi, U_lo (UMMA B operand of the location GEMM)
  float* recv;         // [2][4][128][NC4]
  float* hs;           // [32][NC4]
  float* q_s;          // [128]
  float* v_s;          // [128]
  float* wcat;         // [2][JH + KS - 1]
  float* e_buf;        // [2][Tin_pad]
  float* e_part;       // [2][128]
  float* w_s;          // [Tin_pad]
  float* red_s;        // [8][256]
  float* scr;          // [64]
  float* wprev_s;      // [Tin_pad]  attention weights of the previous frame (all positions, both CTAs of the pair)
  float* wcum_s;       // [Tin_pad]  cumulative attention weights
  float* qg_s;         // [32][128]  partial queries of this sequence (bulk-copied)
  float* wq_s;         // [32][128]  Wq^T rows of the cluster's 32 attention-LSTM units
  uint8_t* im_s;       // [2][16 KB] im2col tiles when they cannot alias a receive buffer (N < 64), else nullptr
  KSched* ks_a;
  KSched* ks_c;
  uint64_t* accA_full;
  uint64_t* accC_full;
  uint64_t* att_full;
  uint64_t* u_full;
  uint64_t* qg_full;
  uint64_t* e_full;     // energies of the peer half have arrived (remote arrive by the peer CTA)
  uint64_t* go_late;    // workers -> issuers: barrier 2 passed (context / h_dec operands visible)
  uint64_t* go_early;   // workers -> issuers: acc_A is free and h_att(t) is visible
  uint64_t* recv_full;  // [2]
  uint32_t* tmem_slot;
};

template <int N>
__host__ __device__ constexpr size_t fwd3_fixed_bytes() {
  constexpr int NC4 = (N < 64 ? N : 64) / 4;
  return (size_t)2 * W_TILE_BYTES + (size_t)2 * 4 * TILE_M * NC4 * 4 +
         (size_t)(NC4 * 32 + 128 + 128 + 2 * (JH_MAX + KS_MAX - 1) + 2 * TIN_MAX + 2 * 128 + TIN_MAX + 8 * 256 + 64 + 2 * TIN_MAX +
                  2 * 32 * 128) * 4 +
         (N < 64 ? (size_t)2 * W_TILE_BYTES : 0) + 2 * ((sizeof(KSched) + 15) / 16 * 16) + 64 * 8 + 64;
}

template <int N>
GTN_DEVICE Smem3 carve3(uint8_t* base, int S) {
  constexpr int NC4 = (N < 64 ? N : 64) / 4;
  Smem3 m;
  m.ring.base = base;
  m.ring.S = S;
  m.ring.stage_bytes = W_TILE_BYTES + N * KB * 2;
  m.ring.prod_it = m.ring.mma_it = 0;
  uint8_t* p = base + (size_t)S * m.ring.stage_bytes;
  m.u_s = p; p += 2 * W_TILE_BYTES;
  m.im_s = nullptr;
  if (N < 64) { m.im_s = p; p += 2 * W_TILE_BYTES; }
  float* f = reinterpret_cast<float*>(p);
  m.recv = f; f += 2 * 4 * TILE_M * NC4;
  m.qg_s = f; f += 32 * 128;
  m.wq_s = f; f += 32 * 128;
  m.hs = f; f += NC4 * 32;
  m.q_s = f; f += 128;
  m.v_s = f; f += 128;
  m.wcat = f; f += 2 * (JH_MAX + KS_MAX - 1);
  m.e_buf = f; f += 2 * TIN_MAX;
  m.e_part = f; f += 2 * 128;
  m.w_s = f; f += TIN_MAX;
  m.red_s = f; f += 8 * 256;
  m.scr = f; f += 64;
  m.wprev_s = f; f += TIN_MAX;
  m.wcum_s = f; f += TIN_MAX;
  uint8_t* q = reinterpret_cast<uint8_t*>(f);
  m.ks_a = reinterpret_cast<KSched*>(q); q += (sizeof(KSched) + 15) / 16 * 16;
  m.ks_c = reinterpret_cast<KSched*>(q); q += (sizeof(KSched) + 15) / 16 * 16;
  uint64_t* b = reinterpret_cast<uint64_t*>(q);
  m.ring.full = b; b += 16;
  m.ring.empty = b; b += 16;
  m.accA_full = b++;
  m.accC_full = b++;
  m.att_full = b++;
  m.u_full = b++;
  m.qg_full = b++;
  m.e_full = b++;
  m.go_late = b++;
  m.go_early = b++;
  m.recv_full = b; b += 2;
  m.tmem_slot = reinterpret_cast<uint32_t*>(b + 1);
  return m;
}

// ---- split grid barrier (co-resident launch): thread 64 (warp 2) carries the release / acquire and the proxy fences, so
// that the TMA / MMA warps 0 and 1 are free to issue between ARRIVE and WAIT ------------------------------------------
GTN_DEVICE void grid_arrive(unsigned int* counter) {
  WSYNC();
  if (threadIdx.x == 64) {
    fence_proxy_async();
    red_release_add(counter, 1u);
  }
}
// `go`: mbarrier of the issuer warps to arrive on as soon as the barrier is seen complete (before the workers' own sync)
GTN_DEVICE void grid_wait(unsigned int* counter, unsigned int target, uint64_t* go = nullptr) {
  if (threadIdx.x == 64) {
    long long t0 = clock64();
    while (ld_acquire_u32(counter) < target) {
      if (clock64() - t0 > 20000000000LL) __trap();
    }
    fence_proxy_async();
    if (go != nullptr) mbar_arrive(go);
  }
  WSYNC();
}
template <int N>
GTN_DEVICE void flush_cell_saves(const DecFwd& p, int tc, int tile, int rank, const float (&sv)[32 * (N < 64 ? N : 64) / 4 / 256][6]) {
  constexpr int NC4 = (N < 64 ? N : 64) / 4;
  constexpr int CPT = 32 * NC4 / 256;
#pragma unroll
  for (int q = 0; q < CPT; ++q) {
    const int idx = threadIdx.x + q * 256;
    const int ul = idx & 31, bl = idx >> 5, b = rank * NC4 + bl, unit = tile * 32 + ul;
    if (b < p.B) {
      const size_t cell = ((size_t)tc * p.B + b) * p.Ha + unit;
      reinterpret_cast<float4*>(p.gates_a)[cell] = make_float4(sv[q][0], sv[q][1], sv[q][2], sv[q][3]);
      p.c_a_all[cell] = sv[q][4];
      p.h_a_all[cell] = sv[q][5];
    }
  }
}

GTN_DEVICE void mbar_arrive_remote(uint32_t remote_bar) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote_bar) : "memory");
}
GTN_DEVICE void wait_bar_cluster(uint64_t* bar, uint32_t parity) {      // acquire at cluster scope: pairs with mbar_arrive_remote
  long long t0 = clock64();
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    if (ok) return;
    if (clock64() - t0 > 8000000000LL) __trap();
  }
}
GTN_DEVICE float block_reduce_w(float v, float* scr, bool is_max) {       // block_reduce2 over the worker warps
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = is_max ? warp_max(v) : warp_sum(v);
  WSYNC();
  if (lane == 0) scr[warp] = v;
  WSYNC();
  float r = scr[0];
#pragma unroll
  for (int w = 1; w < 8; ++w) r = is_max ? fmaxf(r, scr[w]) : r + scr[w];
  return r;
}

// ---- k-block schedule of one GEMM for this CTA -------------------------------------------------------------------------
// Activation segments s = 0..nseg-1 with nkb[s] k-blocks each; rank r takes k-blocks [r nkb / 4, (r + 1) nkb / 4) of EVERY
// segment; early segments first.  wbase[s] = first k-block column of the segment in the packed weight matrix.
GTN_DEVICE void build_sched(KSched& ks, int nseg, const int* nkb, const int* wbase, const bool* late, int rank) {
  int n = 0;
  for (int pass = 0; pass < 2; ++pass) {
    for (int s = 0; s < nseg; ++s) {
      if ((int)late[s] != pass) continue;
      const int lo = rank * nkb[s] / CL, hi = (rank + 1) * nkb[s] / CL;
      for (int k = lo; k < hi && n < KMAX; ++k) {
        ks.wk[n] = (short)(wbase[s] + k); ks.xk[n] = (short)k; ks.seg[n] = (signed char)s; ks.slot[n] = -1;
        ++n;
      }
    }
    if (pass == 0) ks.n_early = n;
  }
  ks.n = n;
}

struct XOps {                       // activation operands of one GEMM at one frame (no arrays: the segment index is a run-time
  const CUtensorMap *m0, *m1, *m2;  // value, and an indexed array would live in local memory)
  int z0, z1, z2;
  GTN_DEVICE const CUtensorMap* map(int s) const { return s == 0 ? m0 : (s == 1 ? m1 : m2); }
  GTN_DEVICE int z(int s) const { return s == 0 ? z0 : (s == 1 ? z1 : z2); }
};

// Issue k-blocks [i0, i1) of the schedule: producer = warp 8, MMA issuer = warp 9, both in warp-uniform control flow (see
// gemm_issue in gtn_decoder2.cuh).  Consecutive resident k-blocks share a ring stage (activation tiles only).
template <int N>
GTN_DEVICE void issue_kblocks(const KSched& ks, int i0, int i1, const CUtensorMap* wmap, int row0, const XOps& x, Ring& rg,
                              uint32_t tmem_acc, uint32_t tmem_base, uint32_t& started, uint64_t* commit_bar) {
  constexpr uint32_t X_BYTES = N * KB * 2;
  constexpr int G = (W_TILE_BYTES + X_BYTES) / X_BYTES;
  const int warp_u = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0) - 8;
  if (warp_u < 0 || warp_u > 1) return;
  const int S = rg.S;
  const uint32_t stage_bytes = rg.stage_bytes;
  uint8_t* const base = rg.base;
  uint64_t* const full = rg.full;
  uint64_t* const empty = rg.empty;
  if (warp_u == 0) {
    const uint32_t it0 = rg.prod_it;
    int s = (int)(it0 % (uint32_t)S);
    uint32_t par = (it0 / (uint32_t)S) & 1u;
    uint32_t nst = 0;
    for (int i = i0; i < i1;) {
      const bool res = ks.slot[i] >= 0;
      int n = 1;
      if (res) while (n < G && i + n < i1 && ks.slot[i + n] >= 0) ++n;
      wait_bar(&empty[s], par ^ 1u);
      uint8_t* st = base + (size_t)s * stage_bytes;
      if (elect_one()) {
        if (res) {
          mbar_expect_tx(&full[s], (uint32_t)n * X_BYTES);
          for (int j = 0; j < n; ++j) {
            const int sg = ks.seg[i + j];
            tma_load_3d(st + (size_t)j * X_BYTES, x.map(sg), &full[s], ks.xk[i + j] * KB, 0, x.z(sg));
          }
        } else {
          const int sg = ks.seg[i];
          mbar_expect_tx(&full[s], W_TILE_BYTES + X_BYTES);
          tma_load_3d(st, wmap, &full[s], ks.wk[i] * KB, row0, 0);
          tma_load_3d(st + W_TILE_BYTES, x.map(sg), &full[s], ks.xk[i] * KB, 0, x.z(sg));
        }
      }
      __syncwarp();
      if (++s == S) { s = 0; par ^= 1u; }
      i += n;
      ++nst;
    }
    rg.prod_it = it0 + nst;
  } else {
    const uint32_t idesc = umma_idesc_bf16(TILE_M, N);
    const uint32_t it0 = rg.mma_it;
    int s = (int)(it0 % (uint32_t)S);
    uint32_t par = (it0 / (uint32_t)S) & 1u;
    uint32_t nst = 0;
    for (int i = i0; i < i1;) {
      const bool res = ks.slot[i] >= 0;
      int n = 1;
      if (res) while (n < G && i + n < i1 && ks.slot[i + n] >= 0) ++n;
      wait_bar(&full[s], par);
      tc_fence_after();
      const uint32_t a = smem_u32(base + (size_t)s * stage_bytes);
      if (res) {
        for (int j = 0; j < n; ++j) {
          const uint32_t ta = tmem_base + res_col<2 * N>(ks.slot[i + j]);     // (acc_A and acc_C take 2 N columns)
          const uint64_t db = umma_desc_sw128(a + (uint32_t)j * X_BYTES);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < KB / 16; ++k)
              umma_bf16_ts(tmem_acc, ta + (uint32_t)(8 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)(started | (k > 0)));
          }
          started = 1u;
        }
      } else {
        const uint64_t da = umma_desc_sw128(a), db = umma_desc_sw128(a + W_TILE_BYTES);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < KB / 16; ++k)
            umma_bf16(tmem_acc, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (uint32_t)(started | (k > 0)));
        }
        started = 1u;
      }
      if (elect_one()) umma_commit(&empty[s]);
      __syncwarp();
      if (++s == S) { s = 0; par ^= 1u; }
      i += n;
      ++nst;
    }
    if (commit_bar != nullptr && elect_one()) umma_commit(commit_bar);
    rg.mma_it = it0 + nst;
    __syncwarp();
  }
}

// Dropout bits of this thread's LSTM cells of frame tc (Philox is a function of (seed, site, frame, element) only, so they
// can be drawn whenever the thread is idle: ~70 dependent instructions per cell).
template <int N, bool IS_ATT>
GTN_DEVICE void draw_cell_bits(const DecFwd& p, int tc, int tile, int rank, uint32_t (&bits)[32 * (N < 64 ? N : 64) / 4 / 256]) {
  constexpr int NC4 = (N < 64 ? N : 64) / 4;
  constexpr int CPT = 32 * NC4 / 256;
  const int H = IS_ATT ? p.Ha : p.Hd;
  const unsigned thr = IS_ATT ? p.thr_att : p.thr_dec;
#pragma unroll
  for (int q = 0; q < CPT; ++q) {
    const int idx = threadIdx.x + q * 256;
    const int ul = idx & 31, bl = idx >> 5, b = rank * NC4 + bl, unit = tile * 32 + ul;
    bits[q] = 0xFFFFFFFFu;
    if (b < p.B && p.use_dropout && thr != 0u)
      bits[q] = philox_bits(p.seed, IS_ATT ? 3u : 4u, (uint32_t)tc, (uint32_t)((size_t)b * H + unit));
  }
}

// ---- LSTM epilogue of one GEMM tile: split-K exchange, cell, dropout, saves, next-frame operand ---------------------------
// pre_bits: the cells' dropout bits if the caller drew them earlier (draw_cell_bits), else nullptr (drawn here, before the
// wait for the accumulator).
// tc = frame index of the cell.  c_reg: this thread's cell states, carried in registers over the whole launch (a thread owns
// the same (unit, sequence) pairs in every frame).
template <int N, bool F32, bool IS_ATT>
GTN_DEVICE void lstm_epilogue(const DecFwd& p, Smem3& sm, int tc, int tile, int rank, bool has_work, uint32_t tmem_acc,
                              uint64_t* acc_full, uint32_t& acc_par, uint32_t& round, float (&c_reg)[32 * (N < 64 ? N : 64) / 4 / 256],
                              float (&sv)[32 * (N < 64 ? N : 64) / 4 / 256][6],
                              const uint32_t (*pre_bits)[32 * (N < 64 ? N : 64) / 4 / 256] = nullptr) {
  constexpr int CH = N < 64 ? N : 64;
  constexpr int NC4 = CH / 4;
  constexpr int CPT = 32 * NC4 / 256;
  static_assert(N <= 64, "one exchange chunk per tile");
  constexpr int mult = F32 ? 6 : 1, PREC = F32 ? 3 : 1;
  const int nxt = (tc & 1) ^ 1;
  const int H = IS_ATT ? p.Ha : p.Hd, Hp = IS_ATT ? p.Hap : p.Hdp;
  const int t = tc;                                  // (PROF3)
  const float pdrop = IS_ATT ? p.p_att : p.p_dec;
  const unsigned thr = IS_ATT ? p.thr_att : p.thr_dec;
  const float dscale = 1.0f / (1.0f - pdrop);
  const unsigned site = IS_ATT ? 3u : 4u;
  // dropout bits: drawn BEFORE the wait for the accumulator (the workers idle ~2k cycles there while the late k-blocks
  // land) instead of ~70 dependent instructions per cell behind the exchange -- unless the caller already has them
  uint32_t dbits[CPT];
  if (pre_bits != nullptr) {
#pragma unroll
    for (int q = 0; q < CPT; ++q) dbits[q] = (*pre_bits)[q];
  } else {
    draw_cell_bits<N, IS_ATT>(p, tc, tile, rank, dbits);
  }
  if ((threadIdx.x >> 5) >= 4 && has_work) {
    wait_bar(acc_full, acc_par);
    tc_fence_after();
  }
  if (has_work) acc_par ^= 1u;
  PROF3(IS_ATT ? 2 : 10);
  float* recv = sm.recv + (size_t)(round & 1u) * 4 * TILE_M * NC4;
  const uint32_t xr = round;
  ++round;
  const int bbase = rank * NC4;
  float4 gb[CPT];
  {
    const float4* gbias = reinterpret_cast<const float4*>(IS_ATT ? p.gbias_a : p.gbias_d);
    const float* b0 = IS_ATT ? p.b_att : p.b_dec;
#pragma unroll
    for (int q = 0; q < CPT; ++q) {
      const int idx = threadIdx.x + q * 256;
      const int ul = idx & 31, bl = idx >> 5, b = bbase + bl, unit = tile * 32 + ul;
      gb[q] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (b < p.B) {
        // frame 0 of the attention LSTM sees the zero initial context (model.py:314-316): no constant-column term
        if (IS_ATT && tc == 0) gb[q] = make_float4(b0[unit], b0[H + unit], b0[2 * H + unit], b0[3 * H + unit]);
        else gb[q] = __ldg(&gbias[(size_t)b * H + unit]);
      }
    }
  }
  {
    float* stage = sm.recv + (size_t)((xr & 1u) ^ 1u) * 4 * TILE_M * NC4;
    exchange_bulk<CH, true, true>(has_work, tmem_acc, 0, recv, stage, &sm.recv_full[xr & 1u], (xr >> 1) & 1u, rank);
  }
  PROF3(IS_ATT ? 3 : 11);
  bf16* xh_next = (IS_ATT ? BFP(p.xh_a) : BFP(p.xh_d)) + (size_t)nxt * p.B * mult * Hp;
  float* gates_all = IS_ATT ? p.gates_a : p.gates_d;
  float* c_all = IS_ATT ? p.c_a_all : p.c_d_all;
  float* h_all = IS_ATT ? p.h_a_all : p.h_d_all;
#pragma unroll
  for (int q = 0; q < CPT; ++q) {
    const int idx = threadIdx.x + q * 256;
    const int ul = idx & 31, bl = idx >> 5, b = bbase + bl, unit = tile * 32 + ul;
    float h = 0.f;
    if (b < p.B) {
      constexpr bool fast = !F32;
      const float zi = recv_sum_cm<NC4>(recv, ul, bl) + gb[q].x, zf = recv_sum_cm<NC4>(recv, 32 + ul, bl) + gb[q].y;
      const float zg = recv_sum_cm<NC4>(recv, 64 + ul, bl) + gb[q].z, zo = recv_sum_cm<NC4>(recv, 96 + ul, bl) + gb[q].w;
      const float i_ = fast ? 0.5f * tanh_fast(0.5f * zi) + 0.5f : sigmoid_acc(zi);
      const float f_ = fast ? 0.5f * tanh_fast(0.5f * zf) + 0.5f : sigmoid_acc(zf);
      const float g_ = fast ? tanh_fast(zg) : tanh_acc(zg);
      const float o_ = fast ? 0.5f * tanh_fast(0.5f * zo) + 0.5f : sigmoid_acc(zo);
      const size_t ci = (size_t)b * H + unit;
      const float c_new = f_ * c_reg[q] + i_ * g_;
      c_reg[q] = c_new;
      h = o_ * (fast ? tanh_fast(c_new) : tanh_acc(c_new));
      if (p.use_dropout && thr != 0u) h = dbits[q] >= thr ? h * dscale : 0.f;
      store_act(xh_next + (size_t)b * mult * Hp, unit, Hp, PREC, h);
      if (IS_ATT) {
        // the saves for the backward pass are not needed by any other CTA: they are written AFTER this CTA has arrived at
        // barrier 1 (flush_cell_saves), so that the barrier's release does not have to wait for them
        sv[q][0] = i_; sv[q][1] = f_; sv[q][2] = g_; sv[q][3] = o_; sv[q][4] = c_new; sv[q][5] = h;
      } else {
        const size_t cell = ((size_t)tc * p.B + b) * H + unit;
        reinterpret_cast<float4*>(gates_all)[cell] = make_float4(i_, f_, g_, o_);
        c_all[cell] = c_new;
        h_all[cell] = h;
      }
    }
    // (float4 column groups XOR-permuted by the unit: the lanes of a warp run over units, 16-way bank conflict otherwise)
    if (IS_ATT) sm.hs[ul * NC4 + (bl ^ (((ul >> 1) & (NC4 / 4 - 1)) << 2))] = h;
  }
  if (IS_ATT) {
    WSYNC();
    // partial query of this tile: qpart[tile][b][a] = sum_u Wq[a][32 tile + u] h[b][u]; thread = (a, half of the columns)
    constexpr int HB = NC4 / 2;
    const int a = threadIdx.x & 127, g = threadIdx.x >> 7;
    float acc[HB];
#pragma unroll
    for (int j = 0; j < HB; ++j) acc[j] = 0.f;
#pragma unroll
    for (int u = 0; u < 32; ++u) {
      const float4* hr = reinterpret_cast<const float4*>(sm.hs + u * NC4);
      const float wq = sm.wq_s[u * 128 + a];
#pragma unroll
      for (int j4 = 0; j4 < HB / 4; ++j4) {
        const float4 hv = hr[(g * (HB / 4) + j4) ^ ((u >> 1) & (NC4 / 4 - 1))];
        acc[4 * j4] += wq * hv.x; acc[4 * j4 + 1] += wq * hv.y;
        acc[4 * j4 + 2] += wq * hv.z; acc[4 * j4 + 3] += wq * hv.w;
      }
    }
    if (a < p.A) {
#pragma unroll
      for (int j = 0; j < HB; ++j) {
        const int b = bbase + g * HB + j;
        if (b < p.B) p.qpart[((size_t)tile * p.B + b) * p.A + a] = acc[j];
      }
    }
  }
}

// ---- attention of frame t for the sequence of this CTA pair --------------------------------------------------------------
struct AttnCtx {
  int b, L, half, peer, j0, jn, Wd;
  bool active;
};

// Part 1 (depends on nothing but the previous frame's attention weights, so it runs in the shadow of barrier 2 of the
// previous frame): window of the previous / cumulative weights out of shared memory, im2col tile, location GEMM on tcgen05.
template <bool F32>
GTN_DEVICE void attn_prepare(const DecFwd& p, Smem3& sm, const AttnCtx& c, int t, uint32_t tmem_att, uint8_t* im) {
  const int Tin = p.Tin, KS = p.KS, hk = (p.KS - 1) / 2;
  uint8_t* a_hi = im;                                    // im2col tiles, bf16 hi / lo planes (32 KB)
  uint8_t* a_lo = im + W_TILE_BYTES;
  if (c.active) {
    for (int i = threadIdx.x; i < 2 * c.Wd; i += 256) {
      const int ch = i / c.Wd, jj = i - ch * c.Wd, j = c.j0 + jj - hk;
      const float val = (j >= 0 && j < Tin) ? (ch == 0 ? sm.wprev_s[j] : sm.wcum_s[j]) : 0.f;
      sm.wcat[i] = val;
      if (ch == 1 && jj >= hk && jj < hk + c.jn) p.cum_all[((size_t)t * p.B + c.b) * Tin + j] = val;
    }
  }
  WSYNC();
  if (c.active) {
    build_im2col(sm.wcat, c.Wd, KS, c.jn, a_hi, a_lo, NWORK);
    fence_proxy_async_smem();
  }
  WSYNC();
  if (c.active && __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0) == 1) {
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
}

// Part 2 (after barrier 1): query, energies, masked softmax, context, weight state update.
template <bool F32>
GTN_DEVICE void attn_main(const DecFwd& p, const CUtensorMap* qmap, Smem3& sm, const AttnCtx& c, int t, uint32_t tmem_att,
                          uint32_t& qg_par, uint32_t& eround, float& ctx_save, float& wsum_out) {
  // (eround doubles as the phase counter of e_full: one remote arrival per frame)
  typedef typename std::conditional<F32, float, bf16>::type MT;
  constexpr int PREC = F32 ? 3 : 1;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Tin = p.Tin, A = p.A;
  const MT* mem = reinterpret_cast<const MT*>(p.mem_enc);
  const MT* pm = reinterpret_cast<const MT*>(p.pm);
  float* e_buf = sm.e_buf + (size_t)(eround & 1u) * TIN_MAX;
  ++eround;
  float* qg = sm.qg_s;                                                        // [ntq][128]
  const int ntq = p.Ha / 32;
  constexpr bool PM_PRE = sizeof(MT) == 2;
  uint4 pmq[8];
  if (c.active) {
    // the partial queries of this sequence (one per attention-LSTM cluster): ONE TMA box [ntq][1][A] (per-lane bulk copies
    // are serialised by the uniform datapath: 32 of them cost 3k cycles)
    if (__shfl_sync(0xffffffffu, warp, 0) == 0 && elect_one()) {
      mbar_expect_tx(sm.qg_full, (uint32_t)(ntq * A * 4));
      asm volatile(
          "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
          ::"r"(smem_u32(qg)), "l"(reinterpret_cast<uint64_t>(qmap)), "r"(smem_u32(sm.qg_full)), "r"(0), "r"(c.b), "r"(0)
          : "memory");
    }
    // this thread's slice of processed_memory (row = encoder position, 64 attention columns): latency hidden by the gather
    {
      const int wq_ = warp & 3, a0_ = (warp >> 2) * 64, jl_ = wq_ * 32 + lane;
      if (sizeof(MT) == 2 && jl_ < c.jn && (c.j0 + jl_) < c.L) {
        if (p.reserved0 & 32) {             // chunk-major [B][A / 8][Tin][8]: a warp's positions are contiguous
          const uint4* src = reinterpret_cast<const uint4*>(pm) + ((size_t)c.b * (A >> 3) + (a0_ >> 3)) * Tin + (c.j0 + jl_);
#pragma unroll
          for (int g = 0; g < 8; ++g)
            if (a0_ + g * 8 < A) pmq[g] = __ldg(src + (size_t)g * Tin);
        } else {
          const uint4* src = reinterpret_cast<const uint4*>(pm + ((size_t)c.b * Tin + (c.j0 + jl_)) * A + a0_);
#pragma unroll
          for (int g = 0; g < 8; ++g)
            if (a0_ + g * 8 < A) pmq[g] = __ldg(src + g);
        }
      }
    }
    PROF3(15);
    wait_bar(sm.qg_full, qg_par);
    qg_par ^= 1u;
    PROF3(24);
    if (threadIdx.x < 128) {
      const int a = threadIdx.x;
      float s = 0.f;
      if (a < A) {
        // fixed summation tree over the partials (same order as the sequential kernel: 32 -> 16 -> 8 -> 4 -> pairs)
        for (int c0 = 0; c0 < ntq; c0 += 32) {
          float qv[32];
#pragma unroll
          for (int k = 0; k < 32; ++k) qv[k] = (c0 + k < ntq) ? qg[(size_t)(c0 + k) * A + a] : 0.f;
#pragma unroll
          for (int k = 0; k < 16; ++k) qv[k] += qv[k + 16];
#pragma unroll
          for (int k = 0; k < 8; ++k) qv[k] += qv[k + 8];
#pragma unroll
          for (int k = 0; k < 4; ++k) qv[k] += qv[k + 4];
          s += (qv[0] + qv[1]) + (qv[2] + qv[3]);
        }
        if (c.half == 0) p.q_all[((size_t)t * p.B + c.b) * A + a] = s;
      }
      sm.q_s[a] = s;
    }
  }
  WSYNC();
  PROF3(16);
  if (c.active) {
    // (the location GEMM of this frame was issued and waited for in the shadow of the previous barrier 2)
    // thread = (row j = TMEM lane, 64-column half of the attention dimension)
    const int wq = warp & 3, a0 = (warp >> 2) * 64;
    const int jl = wq * 32 + lane;
    const bool valid = jl < c.jn && (c.j0 + jl) < c.L;
    float part = 0.f;
    const MT* pmj = pm + ((size_t)c.b * Tin + (c.j0 + jl)) * A + a0;
    // tanh save of this CTA's positions: CHUNK-major inside the half's block, element (jl, a) at ((a / 4) jn + jl) 4 + a % 4.  A
    // thread owns a position, so in a row-major block every 16-byte store of a warp went to 32 different rows (32 tags, half
    // sectors); chunk-major, the 32 positions of a warp are 512 contiguous bytes.  (The backward kernel copies the half's
    // block as one run and indexes it the same way.)
    float* thb = p.th_all + (((size_t)t * p.B + c.b) * Tin + c.j0) * A;
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
              th[e] = F32 ? tanh_acc(s) : tanh_fast(s);
              part += sm.v_s[a0 + ab + e] * th[e];
            }
            float4* thc = reinterpret_cast<float4*>(thb) + (size_t)((a0 + ab) >> 2) * c.jn + jl;
            thc[0] = make_float4(th[0], th[1], th[2], th[3]);
            thc[c.jn] = make_float4(th[4], th[5], th[6], th[7]);
          }
        }
      }
    }
    sm.e_part[(warp >> 2) * 128 + jl] = valid ? part : -INFINITY;
  }
  tc_fence_before();
  WSYNC();
  if (c.active) {
    const uint32_t peer_e = map_to_rank(smem_u32(e_buf), (uint32_t)c.peer);
    for (int jl = threadIdx.x; jl < c.jn; jl += 256) {
      const float e = sm.e_part[jl] + sm.e_part[128 + jl];      // -inf + -inf = -inf for masked positions
      e_buf[c.j0 + jl] = e;
      st_cluster_f1(peer_e + (uint32_t)((c.j0 + jl) * 4), e);
    }
  }
  PROF3(17);
  // the peer half's energies: remote arrive on the peer's mbarrier instead of a cluster barrier (the issuer warps do not
  // take part in the frame loop's synchronisation)
  WSYNC();
  if (c.active) {
    if (threadIdx.x == 0) mbar_arrive_remote(map_to_rank(smem_u32(sm.e_full), (uint32_t)c.peer));
    wait_bar_cluster(sm.e_full, (eround - 1u) & 1u);
  }
  PROF3(18);
  if (c.active) {
    // masked softmax over all encoder positions (both halves compute it redundantly, so that each CTA of the pair holds
    // the complete previous / cumulative weights for the next frame's location window)
    float mx = -INFINITY;
    for (int j = threadIdx.x; j < Tin; j += 256) mx = fmaxf(mx, e_buf[j]);
    mx = block_reduce_w(mx, sm.scr, true);
    float sum = 0.f;
    for (int j = threadIdx.x; j < Tin; j += 256) {
      const float e = e_buf[j] == -INFINITY ? 0.f : expf(e_buf[j] - mx);
      sm.w_s[j] = e;
      sum += e;
    }
    sum = block_reduce_w(sum, sm.scr, false);
    const float inv = 1.0f / sum;
    float wsum = 0.f;
    for (int j = threadIdx.x; j < Tin; j += 256) {
      const float w = sm.w_s[j] * inv;
      sm.w_s[j] = w;
      sm.wprev_s[j] = w;
      sm.wcum_s[j] += w;                                  // model.py:403
      wsum += w;
    }
    wsum = block_reduce_w(wsum, sm.scr, false);     // also orders the w_s writes before the reads below
    float* aout = p.align_out + ((size_t)c.b * p.n_steps + t) * Tin;
    for (int jl = threadIdx.x; jl < c.jn; jl += 256) aout[c.j0 + jl] = sm.w_s[c.j0 + jl];
    PROF3(19);
    // context over this CTA's half of the time-varying columns: 8 columns per thread, warps split j
    const int E8 = p.E / 8, Eh8 = (E8 + 1) / 2;
    const int c8_0 = c.half * Eh8, c8_n = min(E8, c8_0 + Eh8) - c8_0;      // in units of 8 columns
    constexpr int mult = F32 ? 6 : 1;
    bf16* xc = BFP(p.xctx) + ((size_t)(t & 1) * p.B + c.b) * mult * p.Ep;    // ctx(t) -> buffer t & 1
    wsum_out = wsum;
    for (int cb = 0; cb < c8_n; cb += 32) {
      const int c8 = cb + lane;
      float acc[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = 0.f;
      if (c8 < c8_n) {
        const MT* mp = mem + (size_t)c.b * Tin * p.E + (size_t)(c8_0 + c8) * 8;
        int j = warp;
        for (; j + 56 < c.L; j += 64) {              // eight rows of this warp in flight
          float mv[8][8];
#pragma unroll
          for (int r = 0; r < 8; ++r) ld8<MT>(mp + (size_t)(j + 8 * r) * p.E, mv[r]);
#pragma unroll
          for (int r = 0; r < 8; ++r) {
            const float w = sm.w_s[j + 8 * r];
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] += w * mv[r][e];
          }
        }
        for (; j + 24 < c.L; j += 32) {              // four rows in flight
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
        if (j < c.L) {
          // ragged tail (up to three rows) as ONE trip: rows past the end re-read the last row with weight zero (no branches
          // between the loads; one-row trips were dependent L2 round trips)
          float mv[3][8];
#pragma unroll
          for (int r = 0; r < 3; ++r) ld8<MT>(mp + (size_t)min(j + 8 * r, c.L - 1) * p.E, mv[r]);
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            const float w = j + 8 * r < c.L ? sm.w_s[j + 8 * r] : 0.f;
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] += w * mv[r][e];
          }
        }
      }
      WSYNC();
#pragma unroll
      for (int e = 0; e < 8; ++e) sm.red_s[warp * 256 + lane * 8 + e] = acc[e];
      WSYNC();
      for (int i = threadIdx.x; i < 256; i += 256) {
        const int c8i = cb + (i >> 3);
        if (c8i < c8_n) {
          float s = 0.f;
#pragma unroll
          for (int w = 0; w < 8; ++w) s += sm.red_s[w * 256 + i];
          const int d = (c8_0 + c8i) * 8 + (i & 7);
          store_act(xc, d, p.Ep, PREC, s);
          if (c8_n <= 32) ctx_save = s;                    // saved after the barrier arrive (attn_flush_saves)
          else p.ctx_all[((size_t)t * p.B + c.b) * p.D + d] = s;
        }
      }
    }
  }
  WSYNC();
}

// The context save of frame t (weight-gradient GEMM operand of the backward pass): written after the CTA has arrived at
// barrier 2.  Time-constant columns: context = const * sum_j w_j (second CTA of the pair; the per-frame GEMMs see these
// columns as a per-sequence bias).
GTN_DEVICE void attn_flush_saves(const DecFwd& p, const AttnCtx& c, int t, float ctx_save, float wsum) {
  if (!c.active) return;
  const int E8 = p.E / 8, Eh8 = (E8 + 1) / 2;
  const int c8_0 = c.half * Eh8, c8_n = min(E8, c8_0 + Eh8) - c8_0;
  float* call = p.ctx_all + ((size_t)t * p.B + c.b) * p.D;
  if (c8_n <= 32 && (int)(threadIdx.x >> 3) < c8_n) call[(c8_0 + (threadIdx.x >> 3)) * 8 + (threadIdx.x & 7)] = ctx_save;
  if (c.half == 1 || p.Tin == 1) {
    for (int i = threadIdx.x; i < p.Cc; i += 256) call[p.E + i] = p.mem_const[(size_t)c.b * p.Cc + i] * wsum;
  }
}

template <int N, bool F32>
__global__ void __launch_bounds__(NTHREADS, 1)
decoder_fwd3_kernel(const __grid_constant__ Fwd2Maps maps, const DecFwd p, const int S) {
  constexpr int NC4 = (N < 64 ? N : 64) / 4;
  constexpr int CPT = 32 * NC4 / 256;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  Smem3 sm = carve3<N>(base, S);
  const int warp = threadIdx.x >> 5;
  const int rank = (int)cluster_ctarank(), cid = (int)cluster_idx();
  constexpr int mult = F32 ? 6 : 1;
  const int T = p.n_steps;

  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(&sm.ring.full[s], 1); mbar_init(&sm.ring.empty[s], 1); }
    mbar_init(sm.accA_full, 1);
    mbar_init(sm.accC_full, 1);
    mbar_init(sm.att_full, 1);
    mbar_init(sm.u_full, 1);
    mbar_init(sm.qg_full, 1);
    mbar_init(sm.e_full, 1);
    mbar_init(sm.go_late, 1);
    mbar_init(sm.go_early, 1);
    mbar_init(&sm.recv_full[0], 1);
    mbar_init(&sm.recv_full[1], 1);
    mbar_fence_init();
    // k-block schedules.  A(t): [prenet(t) | ctx(t-1) | h_att(t-1)], ctx behind barrier 2; C(t): [h_att(t) | ctx(t) | h_dec(t-1)],
    // ctx and h_dec behind barrier 2 (h_dec(t-1) is produced in the shadow of barrier 1 of frame t).
    {
      const int nkb[3] = {mult * p.Pp / KB, mult * p.Ep / KB, mult * p.Hap / KB};
      const int wb[3] = {0, nkb[0], nkb[0] + nkb[1]};
      const bool late[3] = {false, true, false};
      build_sched(*sm.ks_a, 3, nkb, wb, late, rank);
    }
    {
      const int nkb[3] = {mult * p.Hap / KB, mult * p.Ep / KB, mult * p.Hdp / KB};
      const int wb[3] = {0, nkb[0], nkb[0] + nkb[1]};
      const bool late[3] = {false, true, true};
      build_sched(*sm.ks_c, 3, nkb, wb, late, rank);
    }
    // resident weight k-blocks (tensor memory): all of A first -- its late k-blocks sit on the serial chain and its early
    // ones are streamed in the shadow of barrier 2 only -- then C (streamed underneath the attention-LSTM epilogue)
    int slot = 0;
    const int nslots = res_slots<2 * N>();
    for (int i = sm.ks_a->n - 1; i >= 0 && slot < nslots; --i) sm.ks_a->slot[i] = (signed char)slot++;
    for (int i = sm.ks_c->n - 1; i >= 0 && slot < nslots; --i) sm.ks_c->slot[i] = (signed char)slot++;
  }
  if (warp == 2) tmem_alloc(sm.tmem_slot, 512);
  if (threadIdx.x < NWORK) {
    for (int i = threadIdx.x; i < 128; i += NWORK) sm.v_s[i] = i < p.A ? p.v[i] : 0.f;
    for (int i = threadIdx.x; i < TIN_MAX; i += NWORK) { sm.wprev_s[i] = 0.f; sm.wcum_s[i] = 0.f; }
    // Wq^T rows of the cluster's 32 attention-LSTM units: wq_s[u][a] = Wq[a][32 cid + u]
    for (int i = threadIdx.x; i < 32 * 128; i += NWORK) {
      const int u = i >> 7, a = i & 127;
      sm.wq_s[i] = a < p.A ? p.wq_t[((size_t)cid * 32 + u) * p.A + a] : 0.f;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *sm.tmem_slot;
  const uint32_t tmem_att = tmem_base + 256u;
  const uint32_t tmem_accA = tmem_base, tmem_accC = tmem_base + (uint32_t)N;
  if (threadIdx.x == 0) {
    mbar_expect_tx(sm.u_full, 2 * W_TILE_BYTES);
    tma_load_3d(sm.u_s, &maps.uloc, sm.u_full, 0, 0, 0);
    tma_load_3d(sm.u_s + W_TILE_BYTES, &maps.uloc, sm.u_full, KB, 0, 0);
    wait_bar(sm.u_full, 0);
  }
  {
    const long long lda = (long long)mult * (p.Pp + p.Ep + p.Hap), ldd = (long long)mult * (p.Hap + p.Ep + p.Hdp);
    for (int i = 0; i < sm.ks_a->n; ++i)
      if (sm.ks_a->slot[i] >= 0) load_resident<2 * N>(p.w_att2, lda, 4 * p.Ha, cid * TILE_M, sm.ks_a->wk[i], 1, sm.ks_a->slot[i], tmem_base);
    for (int i = 0; i < sm.ks_c->n; ++i)
      if (sm.ks_c->slot[i] >= 0) load_resident<2 * N>(p.w_dec2, ldd, 4 * p.Hd, cid * TILE_M, sm.ks_c->wk[i], 1, sm.ks_c->slot[i], tmem_base);
  }
  tc_fence_before();
  cluster_sync_all();     // peers' barriers / smem are initialised before any DSMEM traffic
  tc_fence_after();

  const KSched& ka = *sm.ks_a;
  const KSched& kc = *sm.ks_c;

  if (warp >= 8) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    // =====================================================================================================================
    // issuer warps: 8 = TMA producer, 9 = tcgen05.mma issuer.  Per frame t:
    //   go_late(t)  -> late k-blocks of A(t) (commit acc_A), then ALL k-blocks of C(t-1) (commit acc_C)
    //   go_early(t) -> early k-blocks of A(t+1)
    // =====================================================================================================================
    // One call site of issue_kblocks in a loop over "commands" (k = 0: early A(0); then per frame t: late A(t), all of
    // C(t-1), early A(t+1)), each behind the go barrier that makes its operands visible.
    uint32_t startedA = 0, startedC = 0;
    // (warps 10 and 11 only exist to complete the third warpgroup: they go straight to the final cluster barrier)
#pragma unroll 1
    for (int k = 0; k <= 3 * (T + 1) && warp < 10; ++k) {
      const int t = k == 0 ? 0 : (k - 1) / 3, ph = k == 0 ? 2 : (k - 1) % 3;       // ph 0: late A, 1: C, 2: early A
      const int ta = k == 0 ? 0 : (ph == 2 ? t + 1 : t);                           // frame of the A job
      bool run;
      if (ph == 0) {
        wait_bar(sm.go_late, (uint32_t)(t & 1));
        run = t < T;
      } else if (ph == 1) {
        run = t > 0;
      } else {
        run = k == 0 || t + 1 < T;
        if (run && k > 0) wait_bar(sm.go_early, (uint32_t)(t & 1));
      }
      if (!run) continue;
      fence_proxy_async();
      tc_fence_after();
      XOps x;
      if (ph == 1) {        // C(t-1): h_att(t-1) -> buffer ((t-1)&1)^1, ctx(t-1) -> buffer (t-1)&1, h_dec(t-2) -> buffer (t-1)&1
        const int tc = t - 1;
        x.m0 = &maps.ha;  x.z0 = (tc & 1) ^ 1;
        x.m1 = &maps.ctx; x.z1 = tc & 1;
        x.m2 = &maps.hd;  x.z2 = tc & 1;
        startedC = 0;
      } else {              // A(ta): pre(ta), ctx(ta-1) -> buffer (ta&1)^1, h_att(ta-1) -> buffer ta&1
        x.m0 = &maps.pre; x.z0 = ta;
        x.m1 = &maps.ctx; x.z1 = (ta & 1) ^ 1;
        x.m2 = &maps.ha;  x.z2 = ta & 1;
        if (ph == 2) startedA = 0;
      }
      const KSched& ks = ph == 1 ? kc : ka;
      const int i0 = ph == 0 ? ks.n_early : 0, i1 = ph == 2 ? ks.n_early : ks.n;
      issue_kblocks<N>(ks, i0, i1, ph == 1 ? &maps.w_dec : &maps.w_att, cid * TILE_M, x, sm.ring, ph == 1 ? tmem_accC : tmem_accA,
                       tmem_base, ph == 1 ? startedC : startedA, ph == 0 ? sm.accA_full : (ph == 1 ? sm.accC_full : nullptr));
    }
  } else {
    // =====================================================================================================================
    // worker warps 0..7
    // =====================================================================================================================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    AttnCtx ac;                       // the sequence of this CTA pair
    ac.half = rank & 1; ac.peer = rank ^ 1;
    ac.b = cid * 2 + (rank >> 1);
    ac.active = ac.b < p.B;
    ac.L = ac.active ? (p.lengths ? min(p.lengths[ac.b], p.Tin) : p.Tin) : 0;
    {
      const int JH = (p.Tin + 1) / 2;
      ac.j0 = ac.half * JH;
      ac.jn = min(p.Tin, ac.j0 + JH) - ac.j0;
      ac.Wd = ac.jn + p.KS - 1;
    }
    const bool workA = ka.n > 0, workC = kc.n > 0;
    unsigned int epoch = 0;
    const unsigned int nblocks = gridDim.x;
    uint32_t accA_par = 0, accC_par = 0, att_par = 0, qg_par = 0, round = 0, eround = 0;
    float cA[CPT], cC[CPT], svA[CPT][6];
    uint32_t bitsC[CPT];              // dropout bits of the decoder LSTM cells of the frame whose epilogue runs next
    float ctx_save = 0.f, wsum = 0.f;
#pragma unroll
    for (int q = 0; q < CPT; ++q) { cA[q] = 0.f; cC[q] = 0.f; }
    if (threadIdx.x == 64) mbar_arrive(sm.go_late);                 // frame 0 has no barrier in front of it
    // The im2col tiles live in the receive buffer of the latest split-K exchange (its sums have been consumed, and no peer
    // sends into it before the exchange after next; before the first exchange: buffer 1) -- or in their own region (N < 64).
    auto im2col_buf = [&]() -> uint8_t* {
      return N < 64 ? sm.im_s : reinterpret_cast<uint8_t*>(sm.recv + (size_t)((round - 1u) & 1u) * 4 * TILE_M * NC4);
    };
    {
      const int t = 0;
      attn_prepare<F32>(p, sm, ac, 0, tmem_att, im2col_buf());       // zero initial weights (model.py:317-320)
      if (ac.active) { wait_bar(sm.att_full, att_par); att_par ^= 1u; }
      PROF3(13);
    }
    // one extra trip (t == T) runs the decoder LSTM cells of the last frame: every phase has ONE call site, which keeps the
    // per-frame code path small (it is instruction-cache cold every frame)
#pragma unroll 1
    for (int t = 0; t <= T; ++t) {
      PROF3(0);
      PROFG(0);
      if (t < T) {
        // ---- W2: attention LSTM cells of frame t
        lstm_epilogue<N, F32, true>(p, sm, t, cid, rank, workA, tmem_accA, sm.accA_full, accA_par, round, cA, svA);
        PROF3(4);
        PROFG(1);
        // ---- W3
        grid_arrive(p.barrier);
        epoch += nblocks;
        flush_cell_saves<N>(p, t, cid, rank, svA);
        PROF3(5);
      }
      // ---- W4: decoder LSTM cells of frame t - 1, while the other CTAs arrive at barrier 1
      if (t > 0) lstm_epilogue<N, F32, false>(p, sm, t - 1, cid, rank, workC, tmem_accC, sm.accC_full, accC_par, round, cC, svA, &bitsC);
      if (t == T) break;
      PROF3(12);
      // ---- W6
      grid_wait(p.barrier, epoch);
      tc_fence_after();
      PROF3(14);
      PROFG(2);
      // ---- W7
      attn_main<F32>(p, &maps.qpart, sm, ac, t, tmem_att, qg_par, eround, ctx_save, wsum);
      PROF3(20);
      PROFG(3);
      // ---- W8: barrier 2; in its shadow the issuers stream the early k-blocks of A(t+1) and the workers prepare the
      // attention of frame t + 1 (the location GEMM only needs this frame's weights)
      tc_fence_before();
      grid_arrive(p.barrier);
      epoch += nblocks;
      if (threadIdx.x == 64 && t + 1 < T) mbar_arrive(sm.go_early);
      attn_flush_saves(p, ac, t, ctx_save, wsum);
      PROF3(21);
      if (t + 1 < T) {
        tc_fence_after();
        attn_prepare<F32>(p, sm, ac, t + 1, tmem_att, im2col_buf());
      }
      draw_cell_bits<N, false>(p, t, cid, rank, bitsC);     // for C(t), whose epilogue runs in the shadow of the next barrier 1
      PROF3(13);
      grid_wait(p.barrier, epoch, sm.go_late);
      // the location GEMM is complete (its im2col tiles are in a receive buffer the next exchange stages into)
      if (t + 1 < T && ac.active) { wait_bar(sm.att_full, att_par); att_par ^= 1u; }
      tc_fence_after();
      PROF3(23);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && p.steps_done) *p.steps_done = T;
  }
  tc_fence_before();
  cluster_sync_all();     // no CTA exits while a peer may still write into its shared memory
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
template <int N, bool F32>
static int launch_fwd3_n(const Fwd2Maps& maps, const DecFwd& p, cudaStream_t st) {
  const size_t stage = (size_t)W_TILE_BYTES + (size_t)N * KB * 2;
  const size_t fixed = fwd3_fixed_bytes<N>() + 1024;
  int S = (int)((227 * 1024 - fixed) / stage);
  if (S > 12) S = 12;
  if (S < 2) return FWD3_NOT_COVERED;
  const size_t smem = fixed + (size_t)S * stage;
  auto kern = decoder_fwd3_kernel<N, F32>;
  GTN_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(NTHREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attrs[2];
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = CL; attrs[0].val.clusterDim.y = 1; attrs[0].val.clusterDim.z = 1;
  attrs[1].id = cudaLaunchAttributeCooperative;
  attrs[1].val.cooperative = 1;
  cfg.attrs = attrs;
  cfg.numAttrs = 1;
  const int want = p.Ha / 32;
  cfg.gridDim = dim3(CL * want);
  int max_cl = 0;
  GTN_CUDA_OK(cudaOccupancyMaxActiveClusters(&max_cl, kern, &cfg));
  if (max_cl < want) return FWD3_NOT_COVERED;                    // one attention tile + one decoder tile per cluster, all co-resident
  cfg.numAttrs = profiler_attached() ? 1 : 2;       // profilers cannot replay cooperative launches
  GTN_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, maps, p, S));
  count_launch();
  return 0;
}

int launch_train_fwd3(const Fwd2Maps& maps, const DecFwd& p, cudaStream_t st) {
  if (!p.train || getenv("GTN_FWD_SEQUENTIAL")) return FWD3_NOT_COVERED;
  const int ncl = p.Ha / 32;
  if (p.Hd / 32 != ncl || p.B > 2 * ncl || p.B > 64 || p.Ha > 1024) return FWD3_NOT_COVERED;
  const int m = p.mult;
  // every CTA's k-block list must fit the schedule table
  if ((m * (p.Pp + p.Ep + p.Hap) / KB + CL - 1) / CL + 3 > KMAX || (m * (p.Hap + p.Ep + p.Hdp) / KB + CL - 1) / CL + 3 > KMAX) return FWD3_NOT_COVERED;
  if ((p.prec == 3) != (p.mem_bf16 == 0)) return FWD3_NOT_COVERED;      // fp32 mode <-> fp32 encoder memory
  if (p.prec == 3) return p.B <= 32 ? launch_fwd3_n<32, true>(maps, p, st) : launch_fwd3_n<64, true>(maps, p, st);
  return p.B <= 32 ? launch_fwd3_n<32, false>(maps, p, st) : launch_fwd3_n<64, false>(maps, p, st);
}

}  // namespace gtn
