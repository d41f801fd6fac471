// Internal (non-ABI) declarations shared between the translation units of libgantron_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gtn {

enum { MODE_PLAIN = 0, MODE_CONV = 1, MODE_WGRAD = 2 };

// Fully general description of one tcgen05 GEMM launch (see gtn_gemm.cu).
struct GemmDesc {
  int mode = MODE_PLAIN;
  int mn_major = 0;   // operands are [K][M] / [K][N] (MN contiguous) instead of [M][K] / [N][K]
  int M = 0, N = 0, num_k_blocks = 0, batch = 1, block_n = 128;
  // operand A / B as 3-D bf16 tensors: dims (d0 = K-contiguous inner, d1, d2), element strides of d1, d2
  const void* A = nullptr; long long a_d0 = 0, a_d1 = 0, a_d2 = 1, a_s1 = 0, a_s2 = 0;
  const void* B = nullptr; long long b_d0 = 0, b_d1 = 0, b_d2 = 1, b_s1 = 0, b_s2 = 0;
  int ci_blocks = 0, cin_pad = 0, dil = 1, pad = 0, t_blocks = 0;
  void* C = nullptr;  long long ldc = 0, strideC = 0;  int c_bf16 = 0, c_trans = 0;
  void* C2 = nullptr; long long ldc2 = 0, strideC2 = 0; int c2_bf16 = 0, c2_trans = 0;
  const float* bias_m = nullptr; const float* bias_n = nullptr;
  int act = 0, accumulate = 0, drop_first = 0;
  float drop_p = 0.f; unsigned long long seed = 0; unsigned int site = 0, step = 0;
  float* stats = nullptr;
};

int gemm_launch(const GemmDesc& g, cudaStream_t st);

void set_error(const char* fmt, ...);
void count_launch(int n = 1);
// true when a profiler / injection library is attached to the process (Nsight Compute cannot replay cooperative
// launches: the persistent kernels then drop the cooperative attribute; co-residency still holds on an idle GPU)
bool profiler_attached();

}  // namespace gtn
