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
/* sizeof() of an ABI struct, for binding self-checks: which = 0 gtn_decoder_fwd_params, 1 gtn_decoder_bwd_params,
 * 2 gtn_decoder_bwd2_params */
int gtn_sizeof(int which);

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
 * Epilogue: + bias -> act -> dropout(drop_p; Philox (seed, site), element = (b*T + t)*Cout + co), or with
 * drop_first = 1: + bias -> dropout -> act (DiscConv1d's Conv -> Dropout -> Tanh, layers.py:53-57);
 * stats (if non-NULL) accumulates per-channel sum / sum of squares of the stored values (BatchNorm).
 * Replaces ConvNorm/DiscConv1d convs: model.py:146-151 (Postnet), model.py:506-512 (Discriminator). */
int gtn_conv1d_bf16(const void* Wp, int Cout, int Cin, int cin_pad, int taps, int dil, int pad,
                    const void* X, int nbatch, int T, long long ldx, long long strideXb,
                    void* C, int c_dtype, long long ldc, long long strideCb,
                    void* C2, int c2_dtype, long long ldc2, long long strideC2b,
                    const float* bias_m, int act, float drop_p, unsigned long long seed, int site, int drop_first,
                    float* stats, int block_n, void* stream);

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

/* ---- Postnet / Discriminator / loss elementwise + reduction kernels (gtn_eltwise.cu) -------------------
 * All activations are channels-last: `rows` = batch*time rows of C channels (C % 8 == 0), row stride ld
 * (elements, % 8 == 0); dtype 0 = fp32, 1 = bf16.  Dropout: Philox (seed, site), element = row*C + c. */

/* BatchNorm1d batch statistics (model.py:146-151, training mode): stats = [sum | sum of squares] over `count`
 * samples per channel (accumulated by the conv epilogue) -> mean, rstd = 1/sqrt(biased var + eps); if
 * running_mean != NULL also the momentum update of the running statistics (unbiased variance). */
int gtn_bn_finalize(const float* stats, int C, long long count, float eps, float momentum, float* mean, float* rstd,
                    float* running_mean, float* running_var, void* stream);
/* out = dropout(act((y - mean) * rstd * gamma + beta)); act 0 none / 1 relu / 2 tanh; mean == NULL skips the affine. */
int gtn_bn_act_fwd(const void* y, int y_dtype, long long ldy, long long rows, int C, const float* mean, const float* rstd,
                   const float* gamma, const float* beta, int act, float drop_p, unsigned long long seed, int site,
                   void* out, int out_dtype, long long ldo, void* stream);
/* Backward of the same through batch statistics (gtn_bn_act_bwd) or through fixed running statistics
 * (gtn_bn_act_bwd_eval: module.eval(), mean / rstd are constants: dy = gamma * rstd * dz): given dx (gradient w.r.t. `out`) and the saved conv output y,
 * writes dgamma, dbeta [C], dy (gradient w.r.t. y) and optionally dbias[c] = sum_rows dy (conv bias gradient). */
int gtn_bn_act_bwd(const void* dx, int dx_dtype, long long lddx, const void* y, int y_dtype, long long ldy, long long rows,
                   int C, const float* mean, const float* rstd, const float* gamma, const float* beta, int act,
                   float drop_p, unsigned long long seed, int site, float* dgamma, float* dbeta, void* dy, int dy_dtype,
                   long long lddy, float* dbias, void* stream);
int gtn_bn_act_bwd_eval(const void* dx, int dx_dtype, long long lddx, const void* y, int y_dtype, long long ldy,
                        long long rows, int C, const float* mean, const float* rstd, const float* gamma,
                        const float* beta, int act, float drop_p, unsigned long long seed, int site, float* dgamma,
                        float* dbeta, void* dy, int dy_dtype, long long lddy, float* dbias, void* stream);
/* DiscConv1d backward through Tanh(Dropout(pre)) (layers.py:53-57): dpre = dout * (1 - a^2) * keep / (1 - p), a = the
 * saved layer output; dbias[c] = sum_rows dpre (optional). */
int gtn_drop_tanh_bwd(const void* dout, int dout_dtype, long long lddo, const void* a, int a_dtype, long long lda,
                      long long rows, int C, float drop_p, unsigned long long seed, int site, void* dpre, int dpre_dtype,
                      long long lddp, float* dbias, void* stream);
/* Final Conv1d(n_mel -> 1, k = 1) of the Discriminator (model.py:512): out[row] = bias + w . x[row][:]; and its
 * backward (dx = dout * w, dw += dout * x, db += dout; dw / db may be NULL for a frozen discriminator). */
int gtn_disc_head_fwd(const void* x, int x_dtype, long long ldx, long long rows, int C, const float* w, const float* bias,
                      float* out, void* stream);
int gtn_disc_head_bwd(const float* dout, const void* x, int x_dtype, long long ldx, long long rows, int C, const float* w,
                      void* dx, int dx_dtype, long long lddx, float* dw, float* db, void* stream);
/* Discriminator.adversarial_loss reduction (model.py:529-540): loss = mean_b mean(out[b][0:ceil(len_b/window)]);
 * lengths are int64 on the device (no host sync, unlike the reference's B `.to(FloatTensor)` round trips). */
int gtn_window_mean_fwd(const float* out, int B, int P, const long long* lengths, int window, float* loss, void* stream);
int gtn_window_mean_bwd(const float* gloss, int B, int P, const long long* lengths, int window, float* dout, void* stream);
/* Tacotron2Loss (loss_function.py:10-22): sums[0..2] = sum (mel-tgt)^2, sum (mel_post-tgt)^2, sum BCE-with-logits(gate);
 * strides9 (HOST array) = element strides (batch, channel, time) of mel, mel_post, target.  Backward writes dense
 * channels-last dmel / dpost [b][t][c] and dgate [b][t], scaled by the device scalars g_mel / g_gate. */
int gtn_taco_loss_fwd(const float* mel, const float* post, const float* tgt, const long long* strides9, int B, int C, int T,
                      const float* gate, long long gate_sb, const float* gate_tgt, long long tgt_sb, float* sums,
                      void* stream);
int gtn_taco_loss_bwd(const float* mel, const float* post, const float* tgt, const long long* strides9, int B, int C, int T,
                      const float* gate, long long gate_sb, const float* gate_tgt, long long tgt_sb, const float* g_mel,
                      const float* g_gate, float* dmel, float* dpost, float* dgate, void* stream);
/* Guided-attention loss (loss_function.py:25-52): loss = mean_b BCE(align[b].T[:enc_len_b, :dec_len_b], guide), align
 * fp32 [B][T][Tin] contiguous, lengths int64 on the device, width = 0.5 in the reference.  bwd writes d(align). */
int gtn_attn_guide_loss_fwd(const float* align, int B, int T, int Tin, const long long* enc_len, const long long* dec_len,
                            float width, float* loss, void* stream);
int gtn_attn_guide_loss_bwd(const float* align, int B, int T, int Tin, const long long* enc_len, const long long* dec_len,
                            float width, const float* gloss, float* dalign, void* stream);

/* ---- persistent clustered decoder, forward (gtn_decoder_fwd2.cu) --------------------------------
 * One launch = all frames of Decoder.inference (model.py:456-497; train = 0) or of the teacher-forced
 * loop of Decoder.forward (model.py:419-454; train = 1).  The struct lives in HOST memory.
 *
 * Precision: prec = 1 -> bf16 operands / fp32 accumulate; prec = 3 -> every fp32 operand is carried as
 * three bf16 planes (K expanded 6x), fp32-accurate.  `mult` below = 1 or 6 (filled in by the library).
 * Packed weights: bf16 [rows][mult * sum(segment pads)], segments in the K order listed, each padded
 * to a multiple of 32 and (prec 3) expanded in place; see gantron_b200/packing.py.
 * Activation work buffers use the same segment layout ([B][mult * pad]) and must be zero-initialised. */
typedef struct gtn_decoder_fwd_params {
  /* dims */
  int B, Tin, ldT, n_steps;
  int E, Cc, D, Dp;          /* memory = [E time-varying | Cc time-constant] columns, D = E + Cc, Dp = pad32(D) */
  int Ha, Hap, Hd, Hdp;      /* attention / decoder LSTM width and pad32 */
  int P, Pp, A, F, KS, NM, NMp;
  int prec, mult;
  int train;                 /* 1 teacher-forced, 0 autoregressive */
  int use_dropout;           /* recurrent-state dropout (model.py:393-394,408-409), i.e. module.training */
  int mem_bf16;              /* element type of mem_enc / pm: 1 bf16, 0 fp32 */
  int ignore_stop;           /* inference: run exactly n_steps frames (benchmarks) */
  float p_att, p_dec, gate_threshold;
  unsigned int thr_att, thr_dec;      /* filled in by the library */
  unsigned long long seed;            /* Philox key; sites 1..4, see oracle/philox.py */
  /* weights */
  const void *w_att;         /* [4Ha][mult*(Pp+Dp+Hap)]  K = [prenet | ctx | h_att]     (attention_rnn.weight_ih|hh) */
  const void *w_dec;         /* [4Hd][mult*(Hap+Dp+Hdp)] K = [h_att | ctx | h_dec]      (decoder_rnn.weight_ih|hh)  */
  const void *w_proj;        /* [pad16(NM+1)][mult*(Hdp+Dp)] rows 0..NM-1 mel, NM gate  (linear_projection, gate_layer) */
  const void *w_pre1;        /* [P][mult*NMp]   (prenet.layers.0) -- inference only */
  const void *w_pre2;        /* [P][mult*Pp]    (prenet.layers.1) -- inference only */
  const float *b_att, *b_dec;            /* [4H] = bias_ih + bias_hh */
  const float *b_proj;                   /* [NM+1] */
  const float *wq, *v, *wconv, *wdense;  /* query_layer [A][Ha], v [A], location_conv [F][2][KS], location_dense [A][F] */
  /* inputs */
  const void *mem_enc;       /* [B][Tin][E] */
  const void *pm;            /* processed_memory [B][Tin][A] (model.py:326) */
  const float *mem_const;    /* [B][Cc] */
  const int *lengths;        /* [B] valid encoder positions, or NULL (inference: no mask) */
  const void *pre_all;       /* train: prenet outputs, bf16 [n_steps+1][B][mult*Pp] */
  /* state / workspace (zero-initialised by the caller) */
  void *xh_a, *xh_d;         /* bf16 [2][B][mult*Hap], [2][B][mult*Hdp] */
  void *xctx, *xmel, *xp1, *xp2; /* bf16 [2][B][mult*Ep] (double buffered), [B][mult*NMp], [B][mult*Pp] x2 */
  float *c_a, *c_d;          /* [B][Ha], [B][Hd] */
  float *w_prev, *w_cum;     /* [B][ldT] */
  float *qpart;              /* [Ha/8][B][A] */
  unsigned int *barrier;     /* 1 word */
  int *done, *lengths_out;   /* [B] */
  int *stop, *steps_done;    /* 1 word each */
  /* outputs */
  float *align_out;          /* [B][n_steps][Tin] */
  float *mel_out;            /* inference: [B][NM][n_steps] */
  float *gate_out;           /* inference: [B][n_steps] */
  /* saved for backward (train) */
  float *gates_a, *c_a_all, *h_a_all;   /* [T][B][4Ha] (i,f,g,o post-activation), [T][B][Ha] x2 (h post-dropout) */
  float *gates_d, *c_d_all, *h_d_all;
  float *ctx_all;                       /* [T][B][D] */
  float *q_all;                         /* [T][B][A]       attention query */
  float *th_all;                        /* [T][B][Tin][A]  tanh(q + loc + processed_memory); inside a (t, b) block the rows of each
                                         * CTA half (j0 = 0 / ceil(Tin/2), jn rows) are stored chunk-major: element (jl, a) at
                                         * j0*A + ((a/4)*jn + jl)*4 + a%4 (written and read by the kernels only) */
  float *f_all;                         /* [T][B][Tin][F]  location-conv outputs */
  float *cum_all;                       /* [T][B][Tin]     cumulative weights BEFORE the step (conv input ch. 1) */
  /* second-generation (clustered tcgen05) kernel: gtn_decoder_forward2 */
  const void *w_att2, *w_dec2;          /* w_att / w_dec with rows permuted per 32-unit tile: row tile*128 + gate*32 + u */
  const void *u_loc;                    /* bf16 [A][128]: location conv x dense pre-multiplied, U[a][ch*KS+k], hi | lo planes */
  const float *wq_t;                    /* [Ha][A] = query_layer weight transposed */
  long long *prof;                      /* optional [16][32] clock64() stamps of CTA 0 for the first 16 frames (NULL = off) */
  /* forward2 drops the time-constant memory columns from every per-frame GEMM: their contribution to the LSTM
   * gates / projection is a per-sequence bias computed once (context_const = mem_const * sum_j w_j, sum_j w_j = 1):
   *   K(att) = [prenet Pp | ctx_enc Ep | h_att Hap], K(dec) = [h_att | ctx_enc Ep | h_dec], K(proj) = [h_dec | ctx_enc Ep],
   *   xctx is [2][B][mult*Ep], Ep = pad64(E): the training kernel alternates the two buffers by frame parity. */
  const float *gbias_a, *gbias_d;       /* [B][H][4] (i,f,g,o): bias_ih + bias_hh + W[:, const cols] . mem_const[b] */
  const float *pbias;                   /* [B][NM+1]: projection / gate bias + W[:, const cols] . mem_const[b] (inference) */
  int Ep, reserved0;                    /* reserved0: flags; bit 5 (32): `pm` is bf16 in chunk-major order [B][A/8][Tin][8]; other bits
                                         * are set by the library */
} gtn_decoder_fwd_params;

/* Clusters of 4 CTAs split K of every per-frame GEMM and run it on tcgen05 with TMA-fed operands
 * (gtn_decoder_fwd2.cu); requires all K pads to be multiples of 64 and qpart sized [Ha/32][B][A].
 * Does not write f_all (the location filter is pre-multiplied; the backward pass works on U); gates_a / gates_d are
 * saved as [T][B][H][4] (i,f,g,o per unit).  (The mma.sync first generation of these kernels, gtn_decoder_forward /
 * gtn_decoder_backward, was removed in round 2; the entry points keep their "2" suffix.) */
int gtn_decoder_forward2(const gtn_decoder_fwd_params* params, void* stream);

/* ---- bidirectional LSTM of the text encoder (gtn_bilstm.cu; reference model.py:183-185, 199-213) -------------
 * nn.LSTM(E, H = 256, bidirectional, batch_first) on packed sequences, one persistent launch per pass.  The caller provides
 * the input projection xproj = x W_ih^T + b_ih + b_hh of both directions, fp32 [B][T][2][4][H] (columns dir * 4H + gate * H +
 * unit, gates i, f, g, o as in torch) and whh = bf16 [2][4H][H] (weight_hh_l0, weight_hh_l0_reverse).
 * forward:  out fp32 [B][T][2H] (zero beyond lengths[b]); training saves gates fp32 [2][T][B][H][4] (post-activation) and
 *           c_all fp32 [2][T][B][H] (both NULL at inference).
 * backward: dout fp32 [B][T][2H] -> dgates bf16 [B][T][2][4][H] (pre-activation gate gradients; zero beyond the lengths); the
 *           caller derives dx, dW_ih, dW_hh and the bias gradients from it with batched GEMMs. */
typedef struct gtn_bilstm_params {
  int B, T, H, reserved0;
  const float *xproj;
  const void *whh;
  const int *lengths;
  float *out, *gates, *c_all;
  const float *dout;
  void *dgates;
} gtn_bilstm_params;

int gtn_bilstm_forward(const gtn_bilstm_params* params, void* stream);
int gtn_bilstm_backward(const gtn_bilstm_params* params, void* stream);

/* ---- batch-1 autoregressive decode with every weight on chip (gtn_decoder_infer1.cu) ---------------
 * Decoder.inference (model.py:456-497) for ONE sequence, bf16 weights / fp32 activations.  All CTAs of the device take
 * part (gtn_decoder_infer1_ctas() = SM count); CTA c owns hidden units [c nu, (c + 1) nu) of both LSTM cells.  The caller
 * pre-multiplies everything linear in the attention context with the encoder memory once per utterance:
 *   wimg   bf16 image of the gate rows for tensor memory, [ncta][4 K-quarters][32 rows][cols_a + cols_d] 32-bit words
 *          (row rho = gate * nu + local unit; K layout of the attention LSTM [prenet P | M_a Tp | W_hh Ha | 0],
 *          M_a = W_ih[:, ctx_enc] . memory_enc^T; decoder LSTM [W_ih[:, :Ha] | M_d Tp | W_hh Hd | 0]; quarter q holds
 *          K in [q K/4, (q + 1) K/4))
 *   pc     fp32 [NM + 1][Tin] = W_proj|gate[:, ctx_enc] . memory_enc^T;  wp_h fp32 [NM + 1][Hd] = W_proj|gate[:, :Hd]
 *   gbias_a / gbias_d [H][4], pbias [NM + 1]: biases + time-constant memory columns (as gtn_decoder_fwd_params)
 * Workspace (zero-initialised): h_att_g / h_dec_g [2][H], e_g [2][Tp], qacc [2][128] / macc [2][96] (64-bit fixed-point
 * cross-CTA sums), barrier. */
typedef struct gtn_decoder_infer1_params {
  int Tin, Tp, n_steps, Ha, Hd, P, A, KS, NM, NMp;
  int nu, ncta, cols_a, cols_d;
  int ignore_stop, reserved0;
  float gate_threshold, reserved1;
  unsigned long long seed;
  const void *wimg, *w1, *w2;            /* w1 bf16 [P][NMp], w2 bf16 [P][P] (prenet.layers.0 / .1) */
  const float *u;                        /* [128][64] location filter U[a][ch * KS + k] */
  const float *wq, *v;                   /* query_layer [A][Ha], v [A] */
  const float *wp_h, *pc, *pbias;
  const float *b_att, *gbias_a, *gbias_d;
  const float *pm;                       /* processed_memory [Tin][A] */
  float *h_att_g, *h_dec_g, *e_g;
  long long *qacc, *macc;
  unsigned int *barrier;
  float *mel_out, *gate_out, *align_out; /* [NM][n_steps], [n_steps], [n_steps][Tin] */
  int *steps_done, *done, *lengths_out;
  long long *prof;                       /* optional [16][16] clock64() stamps of CTA 0 (dev tool), NULL = off */
} gtn_decoder_infer1_params;

int gtn_decoder_infer1_ctas(void);        /* CTAs (= SMs) gtn_decoder_infer1 launches, or -1 */
int gtn_decoder_infer1(const gtn_decoder_infer1_params* params, void* stream);

/* gtn_decoder_infer1 for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  It has NO counterpart in the reference.
 * The workspace members of `params` (h_att_g, h_dec_g, e_g, qacc, macc, barrier) are IGNORED (may be NULL): the
 * kernel's cross-CTA workspace lives in one file-scope `static __device__` array inside the library (40,704 bytes on every GPU that
 * loads the library, held until exit), whose address is taken once per device with cudaGetSymbolAddress.  The caller still
 * passes every input and output.  Limits: Ha, Hd <= GTN_INFER1_STATIC_MAX_H and Tp <= GTN_INFER1_STATIC_MAX_TP, larger sizes
 * are rejected with an error (use gtn_decoder_infer1).  The call allocates nothing (no cudaMalloc / cudaFree): one
 * cudaMemsetAsync of the array on `stream` (sums, exchange buffers and the barrier counter must start at zero; the array
 * keeps the previous call's contents), then the launch.  Every call on a device shares the array: calls that overlap on the
 * GPU (different streams, or a second process thread) RACE -- serialise them on one stream. */
#define GTN_INFER1_STATIC_MAX_H 2048
#define GTN_INFER1_STATIC_MAX_TP 512
int gtn_decoder_infer1_static(const gtn_decoder_infer1_params* params, void* stream);

/* ---- persistent clustered decoder, backward through time (gtn_decoder_bwd2.cu) --------------------
 * Counterpart of gtn_decoder_forward2 (same cluster-split tcgen05 GEMM engine).  Per frame t = T-1 .. 0:
 *   P1  dX_dec = W_dec^T dGates_dec(t) (rows [h_att | ctx_enc | h_dec]) and dX_att = W_att^T dGates_att(t+1)
 *       (rows [ctx_enc | h_att]); the decoder LSTM cell backward of frame t-1 is fused into the h_dec rows' epilogue
 *   P2  attention backward per sequence; d[prev | cum] = dth . U and dU += dth^T . im2col on tcgen05
 *   P3  Wq^T dq + attention LSTM cell backward
 * Not computed per frame (the host derives them with K = T*B GEMMs from dga_all / dgd_all): prenet-output gradients,
 * gradients w.r.t. the time-constant memory columns, every weight gradient.  Replaces autograd's replay of the
 * ~22k-node tape of model.py:444-449. */
typedef struct gtn_decoder_bwd2_params {
  int B, Tin, ldT, T;
  int E, Ep, D, Ha, Hd, A, KS;
  int prec, mult;
  int use_dropout, mem_bf16;
  int grid;                  /* = gtn_decoder_bwd2_grid(B, Ha, Hd, E): sizes dv_part / dU_part */
  float p_att, p_dec;
  unsigned int thr_att, thr_dec;
  unsigned long long seed;
  /* weights */
  const void *wT_att2;       /* bf16 [E + Ha][mult*4Ha]: rows = attention_rnn inputs [ctx_enc | h_att], K = gate rows g*Ha+u */
  const void *wT_dec2;       /* bf16 [Ha + E + Hd][mult*4Hd]: rows = decoder_rnn inputs [h_att | ctx_enc | h_dec] */
  const void *ut_loc;        /* bf16 [64][256]: U^T[kk][a], hi plane (a 0..127) | lo plane; U = location dense x conv */
  const void *wq_au;         /* [A][Ha] query_layer weight (model.py:62): bf16 when prec == 1, fp32 when prec == 3 */
  const float *v;            /* [A] */
  /* forward inputs / saved by gtn_decoder_forward2 */
  const void *mem_enc; const int *lengths;
  const float *gates_a, *c_a_all, *gates_d, *c_d_all;    /* [T][B][H][4], [T][B][H] */
  const float *align;        /* [B][T][Tin] */
  const float *th_all;       /* [T][B][Tin][A], chunk-major inside each half as gtn_decoder_forward2 saves it */
  const float *cum_all;      /* [T][B][Tin] */
  /* upstream gradients */
  const float *dproj_all;    /* [T][B][Hd + D]: d(loss)/d[h_dec' | ctx'] through the mel/gate projection */
  const float *dalign;       /* [B][T][Tin] or NULL */
  /* carried state / workspace (zero-initialised) */
  void *xdg_a;               /* bf16 [B][mult*4Ha] */
  void *xdg_d;               /* bf16 [2][B][mult*4Hd] */
  float *carry_ha, *dha_dec, *dc_a;     /* [B][Ha] */
  float *dc_d;                          /* [B][Hd] */
  float *dhd_rec;                       /* [B][Hd] d h_dec(t-1) through the recurrent weights (P1 -> P3 of a frame) */
  float *dctx_att, *dctx_dec;           /* [B][Ep] */
  float *dwprev, *dcum;                 /* [B][ldT] */
  unsigned int *barrier;
  /* outputs */
  float *dga_all, *dgd_all;  /* [T][B][4Ha], [T][B][4Hd] pre-activation gate gradients (gate-major); bf16 when flags & 2 */
  float *dq_all;             /* [T][B][A] */
  float *dctx_all;           /* [B][T][E] total gradient of the time-varying context columns per frame */
  float *dpm;                /* [B][Tin][A] gradient w.r.t. processed_memory (accumulated over frames) */
  float *dv_part;            /* [grid][A] */
  float *dU_part;            /* [grid][128][64]: per-CTA partial of dU[a][ch*KS+k] */
  long long *prof;           /* optional [16][32] clock64() stamps of CTA 0 for the first 16 processed frames (NULL = off) */
  int flags;                 /* bit 0: prefetch each frame's saves into L2 one frame ahead; bit 1: dga_all / dgd_all are bf16; bit 2 is the library's (cleared on entry) */
} gtn_decoder_bwd2_params;

int gtn_decoder_bwd2_grid(int B, int Ha, int Hd, int E);   /* CTAs gtn_decoder_backward2 launches, or -1 */
int gtn_decoder_backward2(const gtn_decoder_bwd2_params* params, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GANTRON_B200_H */
