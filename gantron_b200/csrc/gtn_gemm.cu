// tcgen05 / TMEM / TMA GEMM for sm_100a: the dense-contraction workhorse of the training path.
//
//   D[m, n] = sum_k A[m, k] * B[n, k]          A, B bf16, K-major (K contiguous); fp32 accumulate in TMEM
//
// One kernel serves four addressing modes (all operands arrive through 3-D TMA tensor maps with
// SWIZZLE_128B and zero OOB fill, so ragged M/N/K/T tails need no special casing):
//   MODE_PLAIN  : (batched) GEMM.  Used for the Prenet linears (reference model.py:102-105), the
//                 attention memory layer (model.py:326), and every *deferred weight-gradient* GEMM of
//                 the LSTM cells / projections (K = T_out * B), and d(memory) = align^T . dctx.
//   MODE_CONV   : implicit-GEMM conv1d, forward and data-gradient: M = C_out, N = time, K = taps x C_in;
//                 the tap shift is a TMA coordinate offset on the time axis (zero padding = OOB fill).
//                 Postnet (model.py:146-151) and Discriminator (model.py:514-527, layers.py:43-60).
//   MODE_WGRAD  : conv1d weight gradient: M = C_out, N = C_in, K = (batch, time), one tap per blockIdx.z.
// and two operand majors: K-major (rows of A/B are K-contiguous) or MN-major ("TN": A is [K][M], B is [K][N],
// both row-major with K as the row index -- the natural layout of every weight gradient, where K is
// (time, batch); the TMA box is then 64 K-rows x 64 MN-elements and the UMMA descriptors use the MN-major
// canonical SWIZZLE_128B layout, so no transposed copy of any activation is ever materialised).
//
// Structure (one 128 x BN output tile per CTA): warp 0 = TMA producer, warp 1 = MMA issuer (a single
// elected thread issues tcgen05.mma, accumulator in TMEM), warp 2 = TMEM allocator, warps 4-7 = epilogue
// (tcgen05.ld -> registers -> fused bias / activation / dropout / BatchNorm-statistics -> global).
// smem ring of `stages` x (A 128x64 + B BNx64) bf16 tiles guarded by full/empty mbarriers.
#include "gtn_common.cuh"
#include "gtn_internal.h"

namespace gtn {

static constexpr int BM = 128;
static constexpr int BK = 64;  // 64 bf16 = 128 B = one SWIZZLE_128B row

struct GemmKernelParams {
  int mode;
  int mn_major;           // 0: A[m][k], B[n][k]   1: A[k][m], B[k][n]
  int b_bcast;            // B operand is shared by all z (batch stride 0)
  int M, N;               // logical output extent per z (rows = M, cols = N)
  int num_k_blocks;
  // MODE_CONV
  int ci_blocks, cin_pad, dil, pad;
  // MODE_WGRAD
  int t_blocks, taps_dil, taps_pad;  // shift = z * taps_dil - taps_pad
  // outputs
  void* C;  long long ldc, strideC;  int c_bf16, c_trans;
  void* C2; long long ldc2, strideC2; int c2_bf16, c2_trans;
  const float* bias_m; const float* bias_n;
  int act;            // 0 none, 1 relu, 2 tanh
  int drop_first;     // 1: dropout is applied BEFORE the activation (DiscConv1d: conv -> dropout -> tanh)
  int accumulate;     // C += D (fp32 C only)
  unsigned int drop_thr; float drop_scale; unsigned long long seed; unsigned int site, step;
  float* stats;       // [2][M] per-row sum / sum of squares over valid n (all z), or nullptr
  int stages;
  // split-K: blockIdx.z = split * zdiv + z; split s covers k-blocks [s * kps, min(nk, (s + 1) * kps)) and ADDS its
  // partial tile to the fp32 output with red.global.add (the host zeroes C first unless it accumulates)
  int zdiv, kps, splits;
};

GTN_DEVICE void wait_or_trap(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 6000000000LL) __trap();   // ~3 s: a pipeline bug must not hang the GPU
  }
}

// Epilogue store of one warp's 32 x 32 sub-tile through shared memory.  After tcgen05.ld a lane owns one output ROW, so a
// direct store instruction touches 32 different rows (32 partial sectors per instruction, or 32 scalar stores per chunk for a
// transposed output).  Staged, every store instruction writes 16 bytes per lane along the output's contiguous direction:
// four full 128-byte lines (fp32) or eight 64-byte runs (bf16) per instruction.  st = this warp's [32][36] float scratch in
// the (idle) TMA ring; org = element index of the sub-tile's first element along the contiguous direction, ld = stride of
// the other direction.  trans: the contiguous direction is the tile's ROW (m) index.
GTN_DEVICE void store_tile32(float* st, const float (&v)[32], int lane, void* C, bool bf16o, bool trans, long long org, long long ld) {
  if (!trans) {
    float4* row = reinterpret_cast<float4*>(st + lane * 36);
#pragma unroll
    for (int j = 0; j < 8; ++j) row[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j) st[j * 36 + lane] = v[j];
  }
  __syncwarp();
  if (!bf16o) {
    float* Cf = reinterpret_cast<float*>(C);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = i * 4 + (lane >> 3), c4 = (lane & 7) * 4;
      *reinterpret_cast<float4*>(Cf + org + (long long)r * ld + c4) = *reinterpret_cast<const float4*>(st + r * 36 + c4);
    }
  } else {
    __nv_bfloat16* Cb = reinterpret_cast<__nv_bfloat16*>(C);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = i * 8 + (lane >> 2), c8 = (lane & 3) * 8;
      const float4 a = *reinterpret_cast<const float4*>(st + r * 36 + c8), b = *reinterpret_cast<const float4*>(st + r * 36 + c8 + 4);
      *reinterpret_cast<uint4*>(Cb + org + (long long)r * ld + c8) =
          make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
    }
  }
  __syncwarp();
}
// the staged path needs 16-byte aligned runs: element alignment of every run start
GTN_DEVICE bool tile32_aligned(const void* C, bool bf16o, long long org, long long ld) {
  const long long a = bf16o ? 7 : 3;
  return (reinterpret_cast<uintptr_t>(C) & 15) == 0 && (org & a) == 0 && (ld & a) == 0;
}

template <int BN>
__global__ void __launch_bounds__(256, 2)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const GemmKernelParams p) {
  extern __shared__ uint8_t smem_raw[];
  // align by pointer arithmetic (no integer round trip) so that the compiler keeps the shared address space: LDS / STS, not generic LD / ST
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  constexpr int A_BYTES = BM * BK * 2;
  constexpr int B_BYTES = BN * BK * 2;
  const int stages = p.stages;
  uint8_t* sA = smem;
  uint8_t* sB = smem + (size_t)stages * A_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(sB + (size_t)stages * B_BYTES);
  uint64_t* empty = full + stages;
  uint64_t* accum_full = empty + stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN;
  const int m0 = blockIdx.y * BM;
  const int z = (int)blockIdx.z % p.zdiv;
  const int split = (int)blockIdx.z / p.zdiv;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(accum_full, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, BN);   // BN fp32 columns x 128 lanes
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;
  const int kb0 = split * p.kps;
  const int kb1 = min(p.num_k_blocks, kb0 + p.kps);

  // Producer and MMA issuer run in WARP-UNIFORM control flow (warp index out of a shuffle, all lanes walk the loop, one
  // elected lane issues): TMA / tcgen05 instructions take uniform-register operands, and under a divergent `lane == 0`
  // branch the compiler wraps each of them in an ELECT / R2UR / branch loop (~100 cycles per tcgen05.mma).
  const int warp_u = __shfl_sync(0xffffffffu, warp, 0);
  if (warp_u == 0) {
    // ------------------------------ TMA producer ------------------------------
    int s = 0;
    uint32_t ph = 0u;
    for (int kb = kb0; kb < kb1; ++kb) {
      wait_or_trap(&empty[s], ph ^ 1u);
      uint8_t* dA = sA + (size_t)s * A_BYTES;
      uint8_t* dB = sB + (size_t)s * B_BYTES;
      if (!p.mn_major) {
        int a0, a1, a2, b0, b1, b2;
        if (p.mode == MODE_PLAIN) {
          a0 = kb * BK; a1 = m0; a2 = z;
          b0 = kb * BK; b1 = n0; b2 = p.b_bcast ? 0 : z;
        } else {  // MODE_CONV
          const int tap = kb / p.ci_blocks, cib = kb - tap * p.ci_blocks;
          a0 = tap * p.cin_pad + cib * BK; a1 = m0; a2 = 0;
          b0 = cib * BK; b1 = n0 + tap * p.dil - p.pad; b2 = z;
        }
        if (elect_one()) {
          mbar_expect_tx(&full[s], A_BYTES + B_BYTES);
          tma_load_3d(dA, &tmA, &full[s], a0, a1, a2);
          tma_load_3d(dB, &tmB, &full[s], b0, b1, b2);
        }
      } else {
        // MN-major: boxes of 64 MN-elements (inner, 128 B) x 64 K-rows, one per 64-wide slice of the tile
        int ka, kbr, za, zb;
        if (p.mode == MODE_PLAIN) {
          ka = kbr = kb * BK; za = z; zb = p.b_bcast ? 0 : z;
        } else {  // MODE_WGRAD: k-block -> (batch, time block); z = tap; the tap shift is a K-row offset of B
          const int bb = kb / p.t_blocks, tb = kb - bb * p.t_blocks;
          ka = tb * BK; kbr = tb * BK + z * p.taps_dil - p.taps_pad; za = zb = bb;
        }
        if (elect_one()) {
          mbar_expect_tx(&full[s], A_BYTES + B_BYTES);
#pragma unroll
          for (int h = 0; h < BM / 64; ++h) tma_load_3d(dA + h * 8192, &tmA, &full[s], m0 + 64 * h, ka, za);
#pragma unroll
          for (int h = 0; h < BN / 64; ++h) tma_load_3d(dB + h * 8192, &tmB, &full[s], n0 + 64 * h, kbr, zb);
        }
      }
      __syncwarp();
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
  } else if (warp_u == 1) {
    // ------------------------------ MMA issuer ------------------------------
    const uint32_t idesc = umma_idesc_bf16(BM, BN) | (p.mn_major ? ((1u << 15) | (1u << 16)) : 0u);
    int s = 0;
    uint32_t ph = 0u;
    for (int kb = kb0; kb < kb1; ++kb) {
      wait_or_trap(&full[s], ph);
      tc_fence_after();
      const uint32_t aA = smem_u32(sA + (size_t)s * A_BYTES), aB = smem_u32(sB + (size_t)s * B_BYTES);
      const uint32_t acc0 = (uint32_t)(kb > kb0);
      if (!p.mn_major) {
        const uint64_t da = umma_desc_sw128(aA), db = umma_desc_sw128(aB);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // advance 16 bf16 = 32 B along K inside the 128 B swizzle row: +2 in the (addr >> 4) field
            umma_bf16(tmem_acc, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, acc0 | (uint32_t)(k > 0));
          }
          umma_commit(&empty[s]);      // frees the smem slot once these MMAs retire
        }
      } else {
        const uint64_t da = umma_desc_sw128_mn(aA), db = umma_desc_sw128_mn(aB);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // 16 K-rows = two 8-row swizzle atoms = 2048 B per UMMA K step (+128 in the (addr >> 4) field)
            umma_bf16(tmem_acc, da + (uint64_t)(128 * k), db + (uint64_t)(128 * k), idesc, acc0 | (uint32_t)(k > 0));
          }
          umma_commit(&empty[s]);
        }
      }
      __syncwarp();
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
    if (elect_one()) umma_commit(accum_full);       // accumulator complete
    __syncwarp();
  }
  {
    // ------------------------------ epilogue ------------------------------
    // ALL eight warps: a warp may only read its own TMEM lane quarter (warp % 4), so warps 4..7 take the even 32-column
    // chunks and warps 0..3 -- whose producer / issuer / allocator roles are over by now -- the odd ones.  With four
    // epilogue warps there was one warp per scheduler and the stores of output-bound shapes ran at ~1.1 TB/s.
    wait_or_trap(accum_full, 0);
    tc_fence_after();
    const int wq = warp & 3;                 // TMEM lane quarter this warp may touch
    const int m = m0 + wq * 32 + lane;       // output row of this thread
    const bool m_ok = m < p.M;
    const float bm = (p.bias_m != nullptr && m_ok) ? p.bias_m[m] : 0.0f;
    float s1 = 0.f, s2 = 0.f;
    float* Cf = reinterpret_cast<float*>(p.C);
    __nv_bfloat16* Cb = reinterpret_cast<__nv_bfloat16*>(p.C);
    float* C2f = reinterpret_cast<float*>(p.C2);
    __nv_bfloat16* C2b = reinterpret_cast<__nv_bfloat16*>(p.C2);
    const long long zc = (long long)z * p.strideC, zc2 = (long long)z * p.strideC2;
    // (the ring is idle once the accumulator is complete: its first bytes are the epilogue warps' staging tiles)
    float* stage = reinterpret_cast<float*>(smem) + warp * (32 * 36);
    const int mt0 = m0 + wq * 32;
    const bool rows_full = mt0 + 32 <= p.M;                 // warp-uniform
#pragma unroll 1
    for (int c = warp >= 4 ? 0 : 1; c < BN / 32; c += 2) {
      uint32_t r[32];
      tmem_ld_32x32(tmem_acc + ((uint32_t)(wq * 32) << 16) + (uint32_t)(c * 32), r);
      tmem_ld_wait();
      const int nb = n0 + c * 32;
      if (nb >= p.N) continue;      // warp-uniform
      if (p.splits > 1) {
        // partial tile of one K split: fp32 reduction in L2 (no bias / activation / second output in this mode)
        if (m_ok) {
          float* dst = Cf + zc + (long long)m * p.ldc + nb;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (nb + j < p.N) atomicAdd(dst + j, __uint_as_float(r[j]));
        }
        continue;
      }
      float v[32];
      // (warp-uniform) nothing fused but the row bias: skip the per-element tests of the general path
      const bool simple = p.bias_n == nullptr && !p.accumulate && p.act == 0 && p.drop_thr == 0u;
      if (simple) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) + bm;
      } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float x = __uint_as_float(r[j]) + bm;
        const int n = nb + j;
        if (p.bias_n != nullptr && n < p.N) x += p.bias_n[n];
        if (p.accumulate && m_ok && n < p.N)
          x += p.c_trans ? Cf[zc + (long long)n * p.ldc + m] : Cf[zc + (long long)m * p.ldc + n];
        if (!p.drop_first) {
          if (p.act == 1) x = fmaxf(x, 0.0f);
          else if (p.act == 2) x = tanhf(x);
        }
        v[j] = x;
      }
      if (p.drop_thr != 0u && m_ok) {
        // element numbering = row-major index in the stored (primary) output of slice z
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int n = nb + j;
          if (n < p.N) {
            const unsigned long long e = p.c_trans ? ((unsigned long long)n * p.M + m) : ((unsigned long long)m * p.N + n);
            const uint32_t bits = philox_bits(p.seed, p.site, p.step, (uint32_t)(e + (unsigned long long)z * p.M * p.N));
            v[j] = bits >= p.drop_thr ? v[j] * p.drop_scale : 0.0f;
          }
        }
      }
      if (p.drop_first) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          if (p.act == 1) v[j] = fmaxf(v[j], 0.0f);
          else if (p.act == 2) v[j] = tanhf(v[j]);
        }
      }
      }
      if (p.stats != nullptr) {
        if (nb + 32 <= p.N) {
#pragma unroll
          for (int j = 0; j < 32; ++j) { s1 += v[j]; s2 += v[j] * v[j]; }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (nb + j < p.N) { s1 += v[j]; s2 += v[j] * v[j]; }
        }
      }
      const bool tile_full = rows_full && nb + 32 <= p.N;
      bool c_done = false, c2_done = false;
      if (tile_full && p.C != nullptr) {
        const long long org = p.c_trans ? zc + (long long)nb * p.ldc + mt0 : zc + (long long)mt0 * p.ldc + nb;
        if (tile32_aligned(p.C, p.c_bf16 != 0, org, p.ldc)) {
          store_tile32(stage, v, lane, p.C, p.c_bf16 != 0, p.c_trans != 0, org, p.ldc);
          c_done = true;
        }
      }
      if (tile_full && p.C2 != nullptr) {
        const long long org = p.c2_trans ? zc2 + (long long)nb * p.ldc2 + mt0 : zc2 + (long long)mt0 * p.ldc2 + nb;
        if (tile32_aligned(p.C2, p.c2_bf16 != 0, org, p.ldc2)) {
          store_tile32(stage, v, lane, p.C2, p.c2_bf16 != 0, p.c2_trans != 0, org, p.ldc2);
          c2_done = true;
        }
      }
      if (!m_ok) continue;
      // primary output
      if (p.C != nullptr && !c_done) {
        if (!p.c_trans) {
          const long long base = zc + (long long)m * p.ldc + nb;
          const bool full32 = (nb + 32 <= p.N);
          if (p.c_bf16) {
            if (full32 && ((base & 7) == 0)) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                uint4 q = make_uint4(pack_bf16x2(v[j], v[j + 1]), pack_bf16x2(v[j + 2], v[j + 3]),
                                     pack_bf16x2(v[j + 4], v[j + 5]), pack_bf16x2(v[j + 6], v[j + 7]));
                *reinterpret_cast<uint4*>(Cb + base + j) = q;
              }
            } else {
              for (int j = 0; j < 32; ++j) if (nb + j < p.N) Cb[base + j] = __float2bfloat16_rn(v[j]);
            }
          } else {
            if (full32 && ((base & 3) == 0)) {
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<float4*>(Cf + base + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            } else {
              for (int j = 0; j < 32; ++j) if (nb + j < p.N) Cf[base + j] = v[j];
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int n = nb + j;
            if (n < p.N) {
              const long long a = zc + (long long)n * p.ldc + m;
              if (p.c_bf16) Cb[a] = __float2bfloat16_rn(v[j]); else Cf[a] = v[j];
            }
          }
        }
      }
      // secondary output (other layout / dtype of the same values)
      if (p.C2 != nullptr && !c2_done) {
        if (!p.c2_trans) {
          const long long base = zc2 + (long long)m * p.ldc2 + nb;
          for (int j = 0; j < 32; ++j)
            if (nb + j < p.N) { if (p.c2_bf16) C2b[base + j] = __float2bfloat16_rn(v[j]); else C2f[base + j] = v[j]; }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int n = nb + j;
            if (n < p.N) {
              const long long a = zc2 + (long long)n * p.ldc2 + m;
              if (p.c2_bf16) C2b[a] = __float2bfloat16_rn(v[j]); else C2f[a] = v[j];
            }
          }
        }
      }
    }
    if (p.stats != nullptr && m_ok) {
      atomicAdd(&p.stats[m], s1);
      atomicAdd(&p.stats[p.M + m], s2);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_acc, BN);
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || sym == nullptr) {
    set_error("cuTensorMapEncodeTiled entry point unavailable (%s)", cudaGetErrorString(e));
    return nullptr;
  }
  fn = reinterpret_cast<EncodeTiledFn>(sym);
  return fn;
}

// 3-D bf16 tensor map: dims (d0 inner, d1, d2), byte strides for d1 / d2, box (64, box1, 1), SWIZZLE_128B.
static int make_map(CUtensorMap* out, const void* base, long long d0, long long d1, long long d2,
                    long long stride1_elems, long long stride2_elems, int box1) {
  // box is always 64 inner elements (128 B = one swizzle row) x box1 rows
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return -3;
  GTN_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base %p not 16-byte aligned", base);
  GTN_REQUIRE((stride1_elems % 8) == 0 && (d2 <= 1 || (stride2_elems % 8) == 0),
              "TMA strides must be multiples of 8 bf16 (got %lld, %lld)", stride1_elems, stride2_elems);
  cuuint64_t dims[3] = {(cuuint64_t)d0, (cuuint64_t)d1, (cuuint64_t)(d2 < 1 ? 1 : d2)};
  cuuint64_t strides[2] = {(cuuint64_t)stride1_elems * 2, (cuuint64_t)(d2 <= 1 ? stride1_elems * d1 : stride2_elems) * 2};
  cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)box1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  GTN_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (dims %lld,%lld,%lld strides %lld,%lld)",
              (int)r, d0, d1, d2, stride1_elems, stride2_elems);
  return 0;
}

int make_tma_map_3d(CUtensorMap* out, const void* base, long long d0, long long d1, long long d2, long long stride1_elems,
                    long long stride2_elems, int box1) {
  return make_map(out, base, d0, d1, d2, stride1_elems, stride2_elems, box1);
}

// 3-D fp32 tensor map without swizzle: dims (d0 inner, d1, d2), element strides of d1 / d2, box (box0, box1, box2)
int make_tma_map_f32_3d(CUtensorMap* out, const void* base, long long d0, long long d1, long long d2, long long stride1_elems,
                        long long stride2_elems, int box0, int box1, int box2) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return -3;
  GTN_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base %p not 16-byte aligned", base);
  GTN_REQUIRE((stride1_elems % 4) == 0 && (stride2_elems % 4) == 0 && (box0 % 4) == 0 && box0 <= 256 && box1 <= 256 && box2 <= 256,
              "fp32 TMA map: strides / inner box must be multiples of 4 floats, boxes <= 256");
  cuuint64_t dims[3] = {(cuuint64_t)d0, (cuuint64_t)d1, (cuuint64_t)d2};
  cuuint64_t strides[2] = {(cuuint64_t)stride1_elems * 4, (cuuint64_t)stride2_elems * 4};
  cuuint32_t box[3] = {(cuuint32_t)box0, (cuuint32_t)box1, (cuuint32_t)box2};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  GTN_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (fp32) failed with CUresult %d", (int)r);
  return 0;
}

template <int BN>
static int launch_bn(const CUtensorMap& ta, const CUtensorMap& tb, GemmKernelParams kp, dim3 grid, cudaStream_t st) {
  // BN <= 128: shallow ring (<= 112 KB) so that TWO CTAs share an SM and one tile's epilogue overlaps the other's main
  // loop; BN = 256: one CTA per SM with a 4-deep ring
  const size_t per_stage = (size_t)BM * BK * 2 + (size_t)BN * BK * 2;
  const int max_stages = BN == 256 ? 4 : (BN == 128 ? 3 : 4);
  const int nkb = kp.kps > 0 ? kp.kps : kp.num_k_blocks;
  int stages = nkb < 2 ? 2 : (nkb > max_stages ? max_stages : nkb);
  kp.stages = stages;
  const size_t smem = stages * per_stage + (2 * stages + 1) * sizeof(uint64_t) + 16 + 1024;
  static size_t configured = 0;
  if (smem > configured) {
    GTN_CUDA_OK(cudaFuncSetAttribute(gemm_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024)));
    configured = 227 * 1024;
  }
  gemm_tc_kernel<BN><<<grid, 256, smem, st>>>(ta, tb, kp);
  GTN_CUDA_OK(cudaGetLastError());
  count_launch();
  return 0;
}

int gemm_launch(const GemmDesc& g, cudaStream_t st) {
  GTN_REQUIRE(g.M > 0 && g.N > 0 && g.num_k_blocks > 0, "empty GEMM (M=%d N=%d kblocks=%d)", g.M, g.N, g.num_k_blocks);
  const int BN = g.block_n;
  GTN_REQUIRE(BN == 64 || BN == 128 || BN == 256, "block_n must be 64/128/256");
  CUtensorMap ta, tb;
  int rc = make_map(&ta, g.A, g.a_d0, g.a_d1, g.a_d2, g.a_s1, g.a_s2, g.mn_major ? BK : BM);
  if (rc) return rc;
  const bool b_bcast = g.mode == MODE_PLAIN && g.batch > 1 && g.b_s2 == 0;
  rc = make_map(&tb, g.B, g.b_d0, g.b_d1, b_bcast ? 1 : g.b_d2, g.b_s1, g.b_s2, g.mn_major ? BK : BN);
  if (rc) return rc;
  GemmKernelParams kp;
  kp.mode = g.mode; kp.mn_major = g.mn_major; kp.b_bcast = b_bcast ? 1 : 0; kp.M = g.M; kp.N = g.N; kp.num_k_blocks = g.num_k_blocks;
  kp.ci_blocks = g.ci_blocks; kp.cin_pad = g.cin_pad; kp.dil = g.dil; kp.pad = g.pad;
  kp.t_blocks = g.t_blocks; kp.taps_dil = g.dil; kp.taps_pad = g.pad;
  kp.C = g.C; kp.ldc = g.ldc; kp.strideC = g.strideC; kp.c_bf16 = g.c_bf16; kp.c_trans = g.c_trans;
  kp.C2 = g.C2; kp.ldc2 = g.ldc2; kp.strideC2 = g.strideC2; kp.c2_bf16 = g.c2_bf16; kp.c2_trans = g.c2_trans;
  kp.bias_m = g.bias_m; kp.bias_n = g.bias_n; kp.act = g.act; kp.drop_first = g.drop_first; kp.accumulate = g.accumulate;
  if (g.drop_p > 0.f) {
    double t = (double)g.drop_p * 4294967296.0;
    kp.drop_thr = t >= 4294967295.0 ? 0xFFFFFFFFu : (unsigned int)t;
    kp.drop_scale = 1.0f / (1.0f - g.drop_p);
  } else { kp.drop_thr = 0; kp.drop_scale = 1.f; }
  kp.seed = g.seed; kp.site = g.site; kp.step = g.step;
  kp.stats = g.stats; kp.stages = 0;
  GTN_REQUIRE(!(g.accumulate && g.c_bf16), "accumulate needs an fp32 primary output");
  const int batch = g.batch < 1 ? 1 : g.batch;
  const int gx = (g.N + BN - 1) / BN, gy = (g.M + BM - 1) / BM;
  // split-K for GEMMs whose tile count cannot fill the 148 SMs (weight gradients with K = T*B and a small M x N):
  // only for a plain fp32, non-transposed primary output without fused epilogue terms
  int splits = 1;
  const long long tiles = (long long)gx * gy * batch;
  const bool splittable = g.mn_major && !g.c_bf16 && !g.c_trans && g.C2 == nullptr && g.bias_m == nullptr &&
                          g.bias_n == nullptr && g.act == 0 && g.drop_p <= 0.f && g.stats == nullptr && g.C != nullptr;
  if (splittable && tiles < 100 && g.num_k_blocks >= 16) {
    long long want = 296 / tiles;
    const long long maxs = g.num_k_blocks / 8;
    if (want > maxs) want = maxs;
    if (want > 1) splits = (int)want;
  }
  kp.kps = (g.num_k_blocks + splits - 1) / splits;
  splits = (g.num_k_blocks + kp.kps - 1) / kp.kps;
  kp.splits = splits; kp.zdiv = batch;
  if (splits > 1 && !g.accumulate) {
    for (int zz = 0; zz < batch; ++zz)
      GTN_CUDA_OK(cudaMemset2DAsync(reinterpret_cast<float*>(g.C) + (long long)zz * g.strideC, (size_t)g.ldc * 4, 0,
                                    (size_t)g.N * 4, (size_t)g.M, st));
  }
  dim3 grid(gx, gy, batch * splits);
  if (BN == 64) return launch_bn<64>(ta, tb, kp, grid, st);
  if (BN == 128) return launch_bn<128>(ta, tb, kp, grid, st);
  return launch_bn<256>(ta, tb, kp, grid, st);
}

}  // namespace gtn
