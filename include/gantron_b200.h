/* gantron_b200 -- C ABI of libgantron_b200.so (hand-written sm_100a kernels for the GANtron
 * decoder / Postnet / Discriminator hot path).
 *
 * The reference (Brechard/GANtron) has no FFI: its boundary is the Python surface of model.py
 * (SURVEY.md section 8b).  These entry points are what a maintainer's `model.py` binds with ctypes to
 * replace the torch op sequences cited next to each function; `gantron_b200/_lib.py` is that
 * binding and INTEGRATION.md shows the reference-side stub.
 *
 * Conventions: plain pointers and sizes only (no torch types); all pointers are DEVICE pointers
 * unless stated; `stream` is a cudaStream_t passed as void*; nothing is allocated inside (callers
 * supply outputs and workspaces); every function returns 0 on success, non-zero on failure with a
 * message retrievable through gtn_last_error().  bf16 buffers are raw uint16 storage.
 */
#ifndef GANTRON_B200_H
#define GANTRON_B200_H
#ifdef __cplusplus
extern "C" {
#endif

/* ---- library / diagnostics ---------------------------------------------------------------- */
const char* gtn_last_error(void);
int gtn_version(void);
/* kernels launched by this library since the last reset (bench.py's "gpu_launches") */
long long gtn_launch_count(void);
void gtn_reset_launch_count(void);
/* fills sm_count, cc_major, cc_minor, cooperative-launch support of the current device */
int gtn_device_info(int* sm_count, int* cc_major, int* cc_minor, int* coop);

/* ---- tcgen05 GEMM family (gtn_gemm.cu) ------------------------------------------------------
 * D[z][m][n] = sum_k A[z][m][k] * B[z][n][k], bf16 operands with K contiguous, fp32 accumulate in
 * TMEM, fused epilogue: + bias_m[m] + bias_n[n] (+ C if accumulate) -> act (0 none,1 relu,2 tanh) ->
 * dropout(p; Philox seed/site/step, element = row-major index in the stored primary output) ->
 * stats[m] += sum_n, stats[M+m] += sum_n of squares -> store to C (and optionally C2) as fp32
 * (dtype 0) or bf16 (dtype 1), row-major [m][n] or transposed [n][m].
 * Replaces: Prenet linears (model.py:102-105), memory_layer (model.py:326), and the per-step
 * weight-gradient accumulations of autograd through model.py:391-416 (as K = T*B GEMMs). */
int gtn_gemm_bf16(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb,
                  int batch, long long strideA, long long strideB,
                  void* C, int c_dtype, long long ldc, long long strideC, int c_trans,
                  void* C2, int c2_dtype, long long ldc2, long long strideC2, int c2_trans,
                  const float* bias_m, const float* bias_n, int act, int accumulate,
                  float drop_p, unsigned long long seed, int site, int step,
                  float* stats, int block_n, void* stream);

/* Implicit-GEMM conv1d (forward and data-gradient form):
 *   out[b][t][co] = bias[co] + sum_{tap,ci} Wp[co][tap*cin_pad + ci] * X[b][t + tap*dil - pad][ci]
 * X channels-last bf16 [b][t][ldx]; Wp bf16 [Cout][taps*cin_pad] (cin_pad = Cin rounded up to 64, zero
 * filled); outputs: C channels-last [b][t][ldc], C2 channels-first [b][co][ldc2] (either may be NULL).
 * Replaces ConvNorm/DiscConv1d convs: model.py:146-151 (Postnet), model.py:506-512 (Discriminator). */
int gtn_conv1d_bf16(const void* Wp, int Cout, int Cin, int cin_pad, int taps, int dil, int pad,
                    const void* X, int nbatch, int T, long long ldx, long long strideXb,
                    void* C, int c_dtype, long long ldc, long long strideCb,
                    void* C2, int c2_dtype, long long ldc2, long long strideC2b,
                    const float* bias_m, int act, float* stats, int block_n, void* stream);

/* conv1d weight gradient: dW[tap][co][ci] (+)= sum_{b,t} dY[b][t][co] * X[b][t + tap*dil - pad][ci]
 * dY, X channels-LAST bf16 ([b][t][ld], the same buffers forward/dgrad use: MN-major UMMA operands, the
 * tap shift is a TMA row offset with zero fill); dW fp32 [taps][Cout][ldw].  Replaces autograd's
 * conv weight gradients of model.py:146-151 / 506-512. */
int gtn_conv1d_wgrad_bf16(const void* dY, const void* X, int nbatch, int Cout, int Cin, int T,
                          long long ldy, long long strideYb, long long ldx, long long strideXb,
                          int taps, int dil, int pad, float* dW, long long ldw, int accumulate,
                          int block_n, void* stream);

/* "TN" GEMM on MN-major operands: D[z][m][n] (+)= sum_k A[z][k][m] * B[z][k][n]; A is [K][lda], B is [K][ldb]
 * (row index = K).  This is the shape of every deferred weight gradient of the decoder (K = time x batch:
 * dW = dGates^T . X replaces the ~800 per-step `mm` + `add_` accumulations autograd performs for
 * model.py:391-392,406-407,413-416) and of d(memory)[b] = align[b]^T . dctx[b] (model.py:88). */
int gtn_gemm_tn_bf16(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb,
                     int batch, long long strideA, long long strideB,
                     void* C, int c_dtype, long long ldc, long long strideC, int c_trans,
                     const float* bias_m, const float* bias_n, int accumulate, int block_n, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GANTRON_B200_H */
