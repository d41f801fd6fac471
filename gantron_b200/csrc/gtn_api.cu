// C-ABI entry points (library plumbing + the GEMM family).  See include/gantron_b200.h.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include "gtn_common.cuh"
#include "gtn_internal.h"
#include "../../include/gantron_b200.h"

namespace gtn {

static thread_local char g_err[1024] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

bool profiler_attached() {
  static const bool on = [] {
    for (const char* v : {"GTN_NO_COOP", "CUDA_INJECTION64_PATH", "NV_NSIGHT_INJECTION_PORT_BASE", "NV_NSIGHT_INJECTION_TRANSPORT_TYPE",
                          "NV_COMPUTE_PROFILER_PERFWORKS_DIR", "NVTX_INJECTION64_PATH"})
      if (getenv(v) != nullptr) return true;
    const char* pre = getenv("LD_PRELOAD");
    return pre != nullptr && (strstr(pre, "nsight") != nullptr || strstr(pre, "ncu") != nullptr || strstr(pre, "nvperf") != nullptr);
  }();
  return on;
}

}  // namespace gtn

using namespace gtn;

extern "C" {

const char* gtn_last_error(void) { return g_err; }
int gtn_version(void) { return 100; }
long long gtn_launch_count(void) { return g_launches.load(); }
void gtn_reset_launch_count(void) { g_launches.store(0); }

int gtn_sizeof(int which) {
  if (which == 0) return (int)sizeof(gtn_decoder_fwd_params);
  if (which == 2) return (int)sizeof(gtn_decoder_bwd2_params);
  if (which == 3) return (int)sizeof(gtn_decoder_infer1_params);
  if (which == 4) return (int)sizeof(gtn_bilstm_params);
  return -1;
}

int gtn_device_info(int* sm_count, int* cc_major, int* cc_minor, int* coop) {
  int dev = 0;
  GTN_CUDA_OK(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  GTN_CUDA_OK(cudaGetDeviceProperties(&prop, dev));
  if (sm_count) *sm_count = prop.multiProcessorCount;
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  if (coop) *coop = prop.cooperativeLaunch;
  return 0;
}

int gtn_gemm_bf16(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb,
                  int batch, long long strideA, long long strideB,
                  void* C, int c_dtype, long long ldc, long long strideC, int c_trans,
                  void* C2, int c2_dtype, long long ldc2, long long strideC2, int c2_trans,
                  const float* bias_m, const float* bias_n, int act, int accumulate,
                  float drop_p, unsigned long long seed, int site, int step,
                  float* stats, int block_n, void* stream) {
  GemmDesc g;
  g.mode = MODE_PLAIN; g.M = M; g.N = N; g.num_k_blocks = (K + 63) / 64; g.batch = batch < 1 ? 1 : batch;
  g.block_n = block_n > 0 ? block_n : 128;
  g.A = A; g.a_d0 = K; g.a_d1 = M; g.a_d2 = g.batch; g.a_s1 = lda; g.a_s2 = strideA;
  g.B = B; g.b_d0 = K; g.b_d1 = N; g.b_d2 = g.batch; g.b_s1 = ldb; g.b_s2 = strideB;
  g.C = C; g.ldc = ldc; g.strideC = strideC; g.c_bf16 = c_dtype; g.c_trans = c_trans;
  g.C2 = C2; g.ldc2 = ldc2; g.strideC2 = strideC2; g.c2_bf16 = c2_dtype; g.c2_trans = c2_trans;
  g.bias_m = bias_m; g.bias_n = bias_n; g.act = act; g.accumulate = accumulate;
  g.drop_p = drop_p; g.seed = seed; g.site = (unsigned)site; g.step = (unsigned)step; g.stats = stats;
  return gemm_launch(g, (cudaStream_t)stream);
}

int gtn_conv1d_bf16(const void* Wp, int Cout, int Cin, int cin_pad, int taps, int dil, int pad,
                    const void* X, int nbatch, int T, long long ldx, long long strideXb,
                    void* C, int c_dtype, long long ldc, long long strideCb,
                    void* C2, int c2_dtype, long long ldc2, long long strideC2b,
                    const float* bias_m, int act, float drop_p, unsigned long long seed, int site, int drop_first,
                    float* stats, int block_n, void* stream) {
  GTN_REQUIRE(cin_pad % 64 == 0 && cin_pad >= Cin, "cin_pad must be Cin rounded up to a multiple of 64");
  GemmDesc g;
  g.mode = MODE_CONV; g.M = Cout; g.N = T; g.batch = nbatch; g.block_n = block_n > 0 ? block_n : 128;
  g.ci_blocks = cin_pad / 64; g.cin_pad = cin_pad; g.dil = dil; g.pad = pad;
  g.num_k_blocks = taps * g.ci_blocks;
  g.A = Wp; g.a_d0 = (long long)taps * cin_pad; g.a_d1 = Cout; g.a_d2 = 1; g.a_s1 = (long long)taps * cin_pad; g.a_s2 = 0;
  g.B = X; g.b_d0 = Cin; g.b_d1 = T; g.b_d2 = nbatch; g.b_s1 = ldx; g.b_s2 = strideXb;
  g.C = C; g.ldc = ldc; g.strideC = strideCb; g.c_bf16 = c_dtype; g.c_trans = 1;
  g.C2 = C2; g.ldc2 = ldc2; g.strideC2 = strideC2b; g.c2_bf16 = c2_dtype; g.c2_trans = 0;
  g.bias_m = bias_m; g.act = act; g.stats = stats;
  g.drop_p = drop_p; g.seed = seed; g.site = (unsigned)site; g.step = 0; g.drop_first = drop_first;
  return gemm_launch(g, (cudaStream_t)stream);
}

int gtn_conv1d_wgrad_bf16(const void* dY, const void* X, int nbatch, int Cout, int Cin, int T,
                          long long ldy, long long strideYb, long long ldx, long long strideXb,
                          int taps, int dil, int pad, float* dW, long long ldw, int accumulate,
                          int block_n, void* stream) {
  GemmDesc g;
  g.mode = MODE_WGRAD; g.mn_major = 1; g.M = Cout; g.N = Cin; g.batch = taps; g.block_n = block_n > 0 ? block_n : 128;
  g.t_blocks = (T + 63) / 64; g.num_k_blocks = nbatch * g.t_blocks; g.dil = dil; g.pad = pad;
  g.A = dY; g.a_d0 = Cout; g.a_d1 = T; g.a_d2 = nbatch; g.a_s1 = ldy; g.a_s2 = strideYb;
  g.B = X;  g.b_d0 = Cin;  g.b_d1 = T; g.b_d2 = nbatch; g.b_s1 = ldx; g.b_s2 = strideXb;
  g.C = dW; g.ldc = ldw; g.strideC = (long long)Cout * ldw; g.c_bf16 = 0; g.c_trans = 0;
  g.accumulate = accumulate;
  return gemm_launch(g, (cudaStream_t)stream);
}

int gtn_gemm_tn_bf16(const void* A, const void* B, int M, int N, int K, long long lda, long long ldb,
                     int batch, long long strideA, long long strideB,
                     void* C, int c_dtype, long long ldc, long long strideC, int c_trans,
                     const float* bias_m, const float* bias_n, int accumulate, int block_n, void* stream) {
  GemmDesc g;
  g.mode = MODE_PLAIN; g.mn_major = 1; g.M = M; g.N = N; g.num_k_blocks = (K + 63) / 64; g.batch = batch < 1 ? 1 : batch;
  g.block_n = block_n > 0 ? block_n : 128;
  g.A = A; g.a_d0 = M; g.a_d1 = K; g.a_d2 = g.batch; g.a_s1 = lda; g.a_s2 = strideA;
  g.B = B; g.b_d0 = N; g.b_d1 = K; g.b_d2 = g.batch; g.b_s1 = ldb; g.b_s2 = strideB;
  g.C = C; g.ldc = ldc; g.strideC = strideC; g.c_bf16 = c_dtype; g.c_trans = c_trans;
  g.bias_m = bias_m; g.bias_n = bias_n; g.accumulate = accumulate;
  return gemm_launch(g, (cudaStream_t)stream);
}

}  // extern "C"
