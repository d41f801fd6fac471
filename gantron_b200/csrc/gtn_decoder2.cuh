// Device machinery of the second-generation persistent decoder kernels (gtn_decoder_fwd2.cu / gtn_decoder_bwd2.cu):
// thread-block clusters + distributed shared memory, the cluster-split skinny GEMM engine on tcgen05, and the
// grid-wide barrier.
//
// Skinny GEMM engine.  Every per-frame matrix product of the decoder has the shape
//        OUT[rows, B] = W[rows, K] . X[B, K]^T            B = batch (<= 256), K in the thousands
// and is bound by how fast W and X reach the SMs, not by the tensor pipe.  A cluster of 4 CTAs owns one
// 128-row tile of W and splits K four ways: each CTA streams ITS quarter of the tile's K (W k-blocks 128 x 64 and
// the matching X k-blocks N x 64, bf16, SWIZZLE_128B, through a TMA ring) into tcgen05.mma (M = 128, N = batch,
// fp32 accumulator in TMEM).  Compared with one CTA per row tile this divides the activation traffic per SM by 4
// and lets all 128 clustered SMs pull weights concurrently.  The four partial accumulators are then exchanged
// through distributed shared memory: CTA r receives, from every peer, the columns (batch rows) [r*N/4, (r+1)*N/4)
// of the 128 x N partial tile, sums the four copies and runs the fused epilogue (LSTM cell, dropout, ...) on its
// 128 x N/4 slice.  One cluster barrier per 64 batch columns; the receive buffer is double buffered so that no
// second barrier is needed before the next exchange.
#pragma once
#include "gtn_common.cuh"
#include "gtn_decoder.cuh"

namespace gtn {

static constexpr int CL = 4;           // CTAs per cluster (K split)
static constexpr int TILE_M = 128;     // weight rows per cluster tile (UMMA M)
static constexpr int KB = 64;          // K elements per k-block (one SWIZZLE_128B row of bf16)
static constexpr int W_TILE_BYTES = TILE_M * KB * 2;

GTN_DEVICE uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
GTN_DEVICE uint32_t cluster_idx() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(r));
  return r;
}
GTN_DEVICE uint32_t cluster_count() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%nclusterid.x;" : "=r"(r));
  return r;
}
// full cluster barrier with release/acquire semantics (orders st.shared::cluster before the peers' reads)
GTN_DEVICE void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// cluster barrier WITHOUT memory ordering: enough where it only has to keep a peer from overwriting shared memory this CTA
// has finished READING (the loads' values have been consumed, so they have completed); the release form waits for every
// earlier global store of the CTA to be performed (1.7k cycles behind the LSTM cells' stores)
GTN_DEVICE void cluster_sync_relaxed() {
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
}
GTN_DEVICE uint32_t map_to_rank(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
GTN_DEVICE void st_cluster_f4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
GTN_DEVICE void st_cluster_f1(uint32_t addr, float a) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(a) : "memory");
}
// generic-proxy writes (st.global / st.shared) -> async-proxy reads (TMA, tcgen05.mma operand fetch)
GTN_DEVICE void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }
GTN_DEVICE void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

GTN_DEVICE void wait_bar(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;       // common case on the issue chains: no clock read
  long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 8000000000LL) __trap();   // ~4 s: a pipeline bug must not hang the GPU
  }
}

// Grid-wide barrier for a co-resident launch; thread 0 (the TMA producer thread) carries the proxy fences.
GTN_DEVICE void grid_sync2(unsigned int* counter, unsigned int& epoch, unsigned int nblocks) {
  __syncthreads();
  epoch += nblocks;
  if (threadIdx.x == 0) {
    // red.release.gpu is cumulative over everything this CTA stored before the bar.sync above; ld.acquire.gpu pairs with it
    fence_proxy_async();
    red_release_add(counter, 1u);
    long long t0 = clock64();
    while (ld_acquire_u32(counter) < epoch) {
      if (clock64() - t0 > 20000000000LL) __trap();
    }
    fence_proxy_async();
  }
  __syncthreads();
}

GTN_DEVICE float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// swizzled byte offset of the 16-byte chunk `chunk` (8 bf16) of row `r` inside a K-major SWIZZLE_128B tile
GTN_DEVICE uint32_t sw128_off(int r, int chunk) {
  return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((chunk ^ (r & 7)) << 4));
}

// im2col tile of the location conv input for rows r < jn: K index kk = ch*KS + k  ->  wcat[ch*Wd + r + k]; written as
// two K-major SWIZZLE_128B bf16 tiles (hi plane, lo = residual plane).  Doubles as an MN-major operand with K = rows.
GTN_DEVICE void build_im2col(const float* wcat, int Wd, int KS, int jn, uint8_t* a_hi, uint8_t* a_lo, int nthreads = 0) {
  if (nthreads == 0) nthreads = (int)blockDim.x;
  for (int i = threadIdx.x; i < jn * 8; i += nthreads) {
    const int r = i >> 3, chunk = i & 7;
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float v2[2];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int kk = chunk * 8 + e * 2 + q;
        float v = 0.f;
        if (kk < 2 * KS) {
          const int ch = kk >= KS ? 1 : 0, k = kk - ch * KS;
          v = wcat[ch * Wd + r + k];
        }
        v2[q] = v;
      }
      const float h0 = bf16_round(v2[0]), h1 = bf16_round(v2[1]);
      hi[e] = pack_bf16x2(h0, h1);
      lo[e] = pack_bf16x2(v2[0] - h0, v2[1] - h1);
    }
    const uint32_t off = sw128_off(r, chunk);
    *reinterpret_cast<uint4*>(a_hi + off) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    *reinterpret_cast<uint4*>(a_lo + off) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
  }
}

// MN-major SWIZZLE_128B descriptor with an explicit distance between consecutive 64-element MN slices
GTN_DEVICE uint64_t umma_desc_sw128_mn_lbo(uint32_t smem_addr_bytes, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr_bytes & 0x3FFFFu) >> 4);
  d |= (uint64_t)(lbo_bytes >> 4) << 16;
  d |= (uint64_t)(1024u >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// Sum v[c] over the 32 lanes of a warp for 64 columns; on return lane l holds the totals of columns 2l and 2l+1
// in v[0], v[1] (butterfly reduce-scatter: 62 shuffles instead of 320).
GTN_DEVICE void warp_reduce_scatter64(float (&v)[64]) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int off = 16, n = 32; off >= 1; off >>= 1, n >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < n; ++i) {
      const float keep = up ? v[i + n] : v[i];
      const float send = up ? v[i] : v[i + n];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
}

// ---- ring pipeline state (meaningful in the producer / MMA threads only) ----
struct Ring {
  uint8_t* base;          // S stages of [W tile 16 KB | X tile N*128 B]
  uint64_t* full;
  uint64_t* empty;
  int S;
  uint32_t stage_bytes;
  uint32_t prod_it, mma_it;
};

struct XSeg {
  const CUtensorMap* map;
  int nkb;                // k-blocks in this segment
  int z;                  // third TMA coordinate (time step / double-buffer slot)
};

struct GemmJob {
  const CUtensorMap* wmap;
  int row0;               // first weight row of the tile
  int nkb;                // total k-blocks over K (all 4 ranks)
  int nseg;
  XSeg seg[3];
  int res_n = 0;          // leading k-blocks of THIS CTA's K range whose weights are resident in TMEM (no W load)
  int res_slot0 = 0;      // first resident slot
  int x_row0 = 0;         // first activation row (batch index) of the tile: batch-chunked phases load 64-row boxes
  // segment fields by RUN-TIME index through selects: an indexed array access would put the whole struct in local memory
  GTN_DEVICE const CUtensorMap* seg_map(int i) const { return i == 0 ? seg[0].map : (i == 1 ? seg[1].map : seg[2].map); }
  GTN_DEVICE int seg_nkb(int i) const { return i == 0 ? seg[0].nkb : (i == 1 ? seg[1].nkb : seg[2].nkb); }
  GTN_DEVICE int seg_z(int i) const { return i == 0 ? seg[0].z : (i == 1 ? seg[1].z : seg[2].z); }
};

// ---- TMEM-resident weight k-blocks -------------------------------------------------------------------------------
// A cluster keeps the same 128-row weight tile and a CTA the same K quarter of it for the whole launch, and the SM's L2
// port (~30 B/clk) is what bounds the per-frame GEMM phases.  Tensor memory has 512 columns; beyond the accumulator(s)
// the free columns hold weight k-blocks (128 rows x 64 bf16 = 128 lanes x 32 columns) loaded ONCE, used as the A operand
// of tcgen05.mma straight from TMEM.  Slot i lives at columns N + 32 i (below column 256) or 384 + 32 (i - n1): columns
// [256, 384) belong to the attention GEMMs of both kernels.
template <int N>
GTN_DEVICE constexpr int res_slots() { return (N < 256 ? (256 - N) / 32 : 0) + 4; }
template <int N>
GTN_DEVICE uint32_t res_col(int slot) {
  constexpr int n1 = N < 256 ? (256 - N) / 32 : 0;
  return slot < n1 ? (uint32_t)(N + 32 * slot) : (uint32_t)(384 + 32 * (slot - n1));
}
// warps 4..7 copy `count` k-blocks (first global k-block kb_first) of rows [row0, row0 + 128) of the K-major bf16 matrix W
// (row stride ldw elements) into slots slot0.. ; rows >= nrows are zero.  Ends with tcgen05.wait::st.
template <int N>
GTN_DEVICE void load_resident(const void* W, long long ldw, int nrows, int row0, int kb_first, int count, int slot0,
                              uint32_t tmem_base) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp < 4 || warp >= 8) return;
  const int q = warp & 3, row = row0 + q * 32 + lane;
  for (int i = 0; i < count; ++i) {
    uint32_t r[32];
    if (row < nrows) {
      const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(W) + (size_t)row * ldw + (size_t)(kb_first + i) * KB);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const uint4 v = __ldg(src + c);
        r[4 * c] = v.x; r[4 * c + 1] = v.y; r[4 * c + 2] = v.z; r[4 * c + 3] = v.w;
      }
    } else {
#pragma unroll
      for (int c = 0; c < 32; ++c) r[c] = 0u;
    }
    tmem_st_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + res_col<N>(slot0 + i), r);
  }
  tmem_st_wait();
}

// TMA load of the activation k-block kb (global index over the concatenated K segments) into `dst`
template <int N>
GTN_DEVICE void load_x_kblock(const GemmJob& job, uint8_t* dst, uint64_t* bar, int kb) {
  int kk = kb, si = 0;
  while (si < job.nseg - 1 && kk >= job.seg_nkb(si)) { kk -= job.seg_nkb(si); ++si; }
  tma_load_3d(dst, job.seg_map(si), bar, kk * KB, 0, job.seg_z(si));
}

// Issue one cluster-split GEMM: this CTA's quarter of K.  Producer = warp 0, MMA issuer = warp 1 (one elected lane each).
// A ring stage is [W tile 16 KB | X tile N*128 B].  The first job.res_n k-blocks of the CTA's range have their weights
// resident in TMEM and need only the X tile, so G = stage_bytes / X_BYTES of them share ONE stage (fewer stages per
// GEMM = fewer load -> MMA -> free-slot round trips; the phases are bound by those chains, not by bytes).
// Returns whether this CTA had any k-block (else its partial tile is all zeros).
#ifndef GTN_NPROD
#define GTN_NPROD 1
#endif
template <int N>
GTN_DEVICE bool gemm_issue(const GemmJob& job, Ring& rg, uint32_t tmem_acc, uint64_t* acc_full, int rank,
                           uint32_t tmem_res = 0u, uint64_t w_policy = 0ull) {
  constexpr uint32_t X_BYTES = N * KB * 2;
  constexpr int G = (W_TILE_BYTES + X_BYTES) / X_BYTES;       // resident k-blocks per stage
  const int kb0 = rank * job.nkb / CL, kb1 = (rank + 1) * job.nkb / CL;
  const int res_n = min(job.res_n, kb1 - kb0);
  const int res_stages = (res_n + G - 1) / G;
  const int nstages = res_stages + (kb1 - kb0 - res_n);
  // Both roles run in WARP-UNIFORM control flow (the warp index comes out of a shuffle, every lane walks the loop, one
  // elected lane issues): the TMA / tcgen05 instructions take their operands from uniform registers, and inside a
  // divergent `threadIdx.x == k` branch the compiler wraps every single one in an ELECT / R2UR / branch waterfall
  // (~100 cycles per tcgen05.mma, ~700 per ring stage on the producer side).
  const int warp_u = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  // The issue loops are serial chains on the phase's critical path (measured: ~700 cycles per stage, of which only
  // ~130 per TMA instruction): ring fields are hoisted into locals, slot / parity advance incrementally (no integer
  // division per stage), the activation-segment cursor is walked once, and wait_bar reads no clock when the first
  // try succeeds.
  const int S = rg.S;
  const uint32_t stage_bytes = rg.stage_bytes;
  uint8_t* const base = rg.base;
  uint64_t* const full = rg.full;
  uint64_t* const empty = rg.empty;
  // NPROD producer warps (0, 2, 3) take the ring stages round-robin: a stage costs its issuer several hundred cycles of
  // serial mbarrier / elect / TMA-issue latency, so splitting the stages over warps shortens the chain the MMA waits on.
  constexpr int NPROD = GTN_NPROD;
  const int pw = warp_u == 0 ? 0 : warp_u - 1;             // producer index of warps 0, 2, 3
  if (warp_u != 1 && pw < NPROD) {
    const uint32_t it0 = rg.prod_it;
    int s = (int)(it0 % (uint32_t)S);
    uint32_t par = (it0 / (uint32_t)S) & 1u;
    // activation cursor: segment si, k-block kk inside it, for global k-block kb0
    int si = 0, kk = kb0;
    while (si < job.nseg - 1 && kk >= job.seg_nkb(si)) { kk -= job.seg_nkb(si); ++si; }
    const CUtensorMap* xmap = job.seg_map(si);
    int xz = job.seg_z(si), xn = job.seg_nkb(si);
    // every lane of every producer warp advances the cursor; one elected lane of the stage's warp issues
    auto load_x = [&](bool mine, uint8_t* dst, uint64_t* bar) {
      if (mine && elect_one()) tma_load_3d(dst, xmap, bar, kk * KB, job.x_row0, xz);
      ++kk;
      while (si < job.nseg - 1 && kk >= xn) { kk -= xn; ++si; xmap = job.seg_map(si); xz = job.seg_z(si); xn = job.seg_nkb(si); }
    };
    for (int sg = 0; sg < nstages; ++sg) {
      const bool mine = (sg % NPROD) == pw;
      if (mine) wait_bar(&empty[s], par ^ 1u);
      uint8_t* st = base + (size_t)s * stage_bytes;
      if (sg < res_stages) {
        const int n = min(G, res_n - sg * G);
        if (mine && elect_one()) mbar_expect_tx(&full[s], (uint32_t)n * X_BYTES);
        for (int j = 0; j < n; ++j) load_x(mine, st + (size_t)j * X_BYTES, &full[s]);
      } else {
        const int kb = kb0 + res_n + (sg - res_stages);
        if (mine && elect_one()) {
          mbar_expect_tx(&full[s], W_TILE_BYTES + X_BYTES);
          // w_policy: the caller knows the weights' reuse distance (one frame) against what else streams through L2
          if (w_policy != 0ull) tma_load_3d_hint(st, job.wmap, &full[s], kb * KB, job.row0, 0, w_policy);
          else tma_load_3d(st, job.wmap, &full[s], kb * KB, job.row0, 0);
        }
        load_x(mine, st + W_TILE_BYTES, &full[s]);
      }
      __syncwarp();
      if (++s == S) { s = 0; par ^= 1u; }
    }
    rg.prod_it = it0 + (uint32_t)nstages;          // Ring is per-thread state: every lane keeps its copy in step
  } else if (warp_u == 1) {
    if (kb1 > kb0) {
      const uint32_t idesc = umma_idesc_bf16(TILE_M, N);
      const uint32_t it0 = rg.mma_it;
      int s = (int)(it0 % (uint32_t)S);
      uint32_t par = (it0 / (uint32_t)S) & 1u;
      uint32_t started = 0u;
      for (int sg = 0; sg < nstages; ++sg) {
        wait_bar(&full[s], par);
        tc_fence_after();
        const uint32_t a = smem_u32(base + (size_t)s * stage_bytes);
        if (sg < res_stages) {
          // weights from TMEM: 16 K elements = 8 columns per UMMA
          const int n = min(G, res_n - sg * G);
          for (int j = 0; j < n; ++j) {
            const uint32_t ta = tmem_res + res_col<N>(job.res_slot0 + sg * G + j);
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
      }
      if (elect_one()) umma_commit(acc_full);
      rg.mma_it = it0 + (uint32_t)nstages;
      __syncwarp();
    }
  }
  return kb1 > kb0;
}

// Exchange one chunk (CH = min(N, 64) batch columns starting at column c0) of the partial accumulator:
// warps 4..7 read their TMEM lane quarter and scatter column group r to rank r's receive buffer
// recv[src rank][row 0..127][NC4 = CH/4 columns].  Every thread of the CTA then joins the cluster barrier.
// SWZ: the float4 slots of a row are XOR-permuted by (row >> 1) so that the slice can afterwards be read with threads
// running over consecutive ROWS (coalesced global stores along the row dimension) at a 4-way instead of 16-way conflict.
template <int CH, bool SWZ = false>
GTN_DEVICE void exchange_chunk(bool has_work, uint32_t tmem_acc, int c0, float* recv, int rank) {
  constexpr int NC4 = CH / 4;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp >= 4) {
    const int wq = warp & 3;
    const int row = wq * 32 + lane;
    const uint32_t my = smem_u32(recv + ((size_t)rank * TILE_M + row) * NC4);
#pragma unroll
    for (int h = 0; h < CH / 32; ++h) {
      uint32_t r[32];
      if (has_work) {
        tmem_ld_32x32(tmem_acc + ((uint32_t)(wq * 32) << 16) + (uint32_t)(c0 + h * 32), r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = 0u;
      }
      // columns h*32 .. h*32+31 of the chunk: destination rank = column / NC4
#pragma unroll
      for (int g = 0; g < 32 / NC4; ++g) {
        const int dst = (h * 32 + g * NC4) / NC4;
        const uint32_t a = map_to_rank(my, (uint32_t)dst);
#pragma unroll
        for (int v = 0; v < NC4; v += 4)
          st_cluster_f4(a + (uint32_t)((SWZ ? (v ^ (((row >> 1) & (NC4 / 4 - 1)) << 2)) : v) * 4), __uint_as_float(r[g * NC4 + v]), __uint_as_float(r[g * NC4 + v + 1]),
                        __uint_as_float(r[g * NC4 + v + 2]), __uint_as_float(r[g * NC4 + v + 3]));
      }
    }
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
}

// CTA-wide thread sync of the exchange: bar.sync 0 over the whole CTA, or (WB) named barrier 1 over the first 256 threads
// for kernels whose warps 8+ are dedicated TMA / MMA issuers that never join the workers' syncs.
template <bool WB>
GTN_DEVICE void cta_sync() {
  if (WB) asm volatile("bar.sync 1, 256;" ::: "memory");
  else __syncthreads();
}

// Same exchange through the bulk-copy engine.  Individual st.shared::cluster stores move the 24 KB a CTA owes its peers at
// ~10 B/clk, and the cluster barrier behind them adds ~1k cycles.  Here warps 4..7 drop the three peer-bound blocks into a
// local staging area (plain st.shared), ONE thread sends each as a cp.async.bulk shared::cta -> shared::cluster copy that
// signals the RECEIVER's mbarrier (complete_tx), and every thread waits on its own CTA's mbarrier: no cluster barrier.
// recv / stage / full_bar alternate between two buffers from one exchange to the next, and stage is the buffer this CTA
// received into one exchange earlier.  That is race-free without further synchronisation: peer q overwrites my
// stage[q] block only in its NEXT exchange, which it enters after it has received this exchange's blocks (mine included, so
// my copy out of stage[q] has completed); and I overwrite a peer's buffer only two exchanges after it was filled, which I
// reach only after that peer has sent the exchange in between, i.e. after it finished reading.
template <int CH, bool WB = false, bool CM = false>
GTN_DEVICE void exchange_bulk(bool has_work, uint32_t tmem_acc, int c0, float* recv, float* stage, uint64_t* full_bar,
                              uint32_t parity, int rank) {
  constexpr int NC4 = CH / 4;
  constexpr uint32_t BLK = (uint32_t)TILE_M * NC4 * 4;           // bytes of one [128 rows][NC4] block
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  cta_sync<WB>();         // stage is the buffer the previous exchange's epilogue read: every warp is done with it
  if (threadIdx.x == 0) mbar_expect_tx(full_bar, 3u * BLK);       // arm this exchange's receive (may trail the first bytes)
  if (warp >= 4) {
    const int wq = warp & 3;
    const int row = wq * 32 + lane;
#pragma unroll
    for (int h = 0; h < CH / 32; ++h) {
      uint32_t r[32];
      if (has_work) {
        tmem_ld_32x32(tmem_acc + ((uint32_t)(wq * 32) << 16) + (uint32_t)(c0 + h * 32), r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = 0u;
      }
#pragma unroll
      for (int g = 0; g < 32 / NC4; ++g) {
        const int dst = (h * 32 + g * NC4) / NC4;
        float* blk0 = dst == rank ? recv + (size_t)rank * TILE_M * NC4 : stage + (size_t)dst * TILE_M * NC4;
        if (CM) {
          // column-major block [NC4][128 rows]: a warp's lanes (consecutive rows) write consecutive words, and the LSTM epilogue
          // (lanes over hidden units = rows) reads consecutive words: no bank conflicts on either side
#pragma unroll
          for (int v = 0; v < NC4; ++v) blk0[(size_t)v * TILE_M + row] = __uint_as_float(r[g * NC4 + v]);
        } else {
          float* blk = blk0 + (size_t)row * NC4;
#pragma unroll
          for (int v = 0; v < NC4; v += 4)
            *reinterpret_cast<float4*>(blk + (v ^ (((row >> 1) & (NC4 / 4 - 1)) << 2))) =
                make_float4(__uint_as_float(r[g * NC4 + v]), __uint_as_float(r[g * NC4 + v + 1]),
                            __uint_as_float(r[g * NC4 + v + 2]), __uint_as_float(r[g * NC4 + v + 3]));
        }
      }
    }
    fence_proxy_async_smem();                                     // staging writes -> bulk-copy (async proxy) reads
  }
  tc_fence_before();
  cta_sync<WB>();
  if (__shfl_sync(0xffffffffu, warp, 0) == 0) {
    if (elect_one()) {
#pragma unroll
      for (int d = 1; d < CL; ++d) {
        const uint32_t dst = (uint32_t)((rank + d) % CL);
        asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(map_to_rank(smem_u32(recv + (size_t)rank * TILE_M * NC4), dst)),
                       "r"(smem_u32(stage + (size_t)dst * TILE_M * NC4)), "r"(BLK), "r"(map_to_rank(smem_u32(full_bar), dst))
                     : "memory");
      }
    }
    __syncwarp();
  }
  wait_bar(full_bar, parity);
  tc_fence_after();
}

// First half of exchange_bulk (everything but the wait for this CTA's own receive), with an explicit staging area: lets a
// caller send chunk c + 1 before it consumes chunk c.  `stage` must not be a receive buffer a peer may write into.
template <int CH, bool WB = false, bool CM = false>
GTN_DEVICE void exchange_send(bool has_work, uint32_t tmem_acc, int c0, float* recv, float* stage, uint64_t* full_bar, int rank) {
  constexpr int NC4 = CH / 4;
  constexpr uint32_t BLK = (uint32_t)TILE_M * NC4 * 4;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  cta_sync<WB>();
  if (threadIdx.x == 0) mbar_expect_tx(full_bar, 3u * BLK);
  // all eight warps stage when the chunk has two 32-column halves (a warp reaches TMEM lane quarter warp % 4): warps 0..3
  // take the first half, warps 4..7 the second
  if (warp >= 4 || CH == 64) {
    const int wq = warp & 3;
    const int row = wq * 32 + lane;
    const int h_lo = CH == 64 ? (warp >> 2) : 0, h_hi = CH == 64 ? (warp >> 2) + 1 : CH / 32;
    for (int h = h_lo; h < h_hi; ++h) {
      uint32_t r[32];
      if (has_work) {
        tmem_ld_32x32(tmem_acc + ((uint32_t)(wq * 32) << 16) + (uint32_t)(c0 + h * 32), r);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = 0u;
      }
#pragma unroll
      for (int g = 0; g < 32 / NC4; ++g) {
        const int dst = (h * 32 + g * NC4) / NC4;
        float* blk0 = dst == rank ? recv + (size_t)rank * TILE_M * NC4 : stage + (size_t)dst * TILE_M * NC4;
        if (CM) {
          // column-major block [NC4][128 rows]: a warp's lanes (consecutive rows) write consecutive words, and the LSTM epilogue
          // (lanes over hidden units = rows) reads consecutive words: no bank conflicts on either side
#pragma unroll
          for (int v = 0; v < NC4; ++v) blk0[(size_t)v * TILE_M + row] = __uint_as_float(r[g * NC4 + v]);
        } else {
          float* blk = blk0 + (size_t)row * NC4;
#pragma unroll
          for (int v = 0; v < NC4; v += 4)
            *reinterpret_cast<float4*>(blk + (v ^ (((row >> 1) & (NC4 / 4 - 1)) << 2))) =
                make_float4(__uint_as_float(r[g * NC4 + v]), __uint_as_float(r[g * NC4 + v + 1]),
                            __uint_as_float(r[g * NC4 + v + 2]), __uint_as_float(r[g * NC4 + v + 3]));
        }
      }
    }
    fence_proxy_async_smem();
  }
  tc_fence_before();
  cta_sync<WB>();
  if (__shfl_sync(0xffffffffu, warp, 0) == 0) {
    if (elect_one()) {
#pragma unroll
      for (int d = 1; d < CL; ++d) {
        const uint32_t dst = (uint32_t)((rank + d) % CL);
        asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(map_to_rank(smem_u32(recv + (size_t)rank * TILE_M * NC4), dst)),
                       "r"(smem_u32(stage + (size_t)dst * TILE_M * NC4)), "r"(BLK), "r"(map_to_rank(smem_u32(full_bar), dst))
                     : "memory");
      }
    }
    __syncwarp();
  }
}

// sum of the four partials of element (row, col) of this CTA's slice
// column-major receive blocks (exchange_bulk / exchange_send with CM)
template <int NC4>
GTN_DEVICE float recv_sum_cm(const float* recv, int row, int col) {
  const float* p = recv + (size_t)col * TILE_M + row;
  return (p[0] + p[(size_t)TILE_M * NC4]) + (p[(size_t)2 * TILE_M * NC4] + p[(size_t)3 * TILE_M * NC4]);
}
template <int NC4, bool SWZ = false>
GTN_DEVICE float recv_sum(const float* recv, int row, int col) {
  const float* p = recv + (size_t)row * NC4 + (SWZ ? (col ^ (((row >> 1) & (NC4 / 4 - 1)) << 2)) : col);
  return (p[0] + p[(size_t)TILE_M * NC4]) + (p[(size_t)2 * TILE_M * NC4] + p[(size_t)3 * TILE_M * NC4]);
}

}  // namespace gtn
