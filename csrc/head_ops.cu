// Classifier head of the CIFAR network in two + two launches: Reshape(2048) -> Linear(2048, 10) -> LogSoftMax ->
// ClassNLLCriterion (examples/Model.lua:44-52), forward and backward, on the NHWC feature map the last
// conv block produces.  In torch this tail is ~14 tiny kernels (permute copy, GEMM + split-K reduce, softmax,
// NLL, their backward twins, three gradient accumulations); the arithmetic is 2.6 MFLOP.
//
// Features arrive as feat[B, HW, C] (NHWC); the Linear weight keeps the reference's (c, h, w) column order, so
// feature (hw, c) multiplies weight column c*HW + hw.
#include "dl_common.cuh"
#include "dl_host.h"

#define HEAD_THREADS 128
#define HEAD_MAX_CLASSES 16

// ---- forward: one block per sample --------------------------------------------------------------------
template <int J>
__global__ void __launch_bounds__(HEAD_THREADS) head_fwd_kernel(const float* __restrict__ feat, const float* __restrict__ w,
                                                                const float* __restrict__ bias, const long long* __restrict__ target,
                                                                float* __restrict__ logp, float* __restrict__ rowloss, int HW, int C) {
  __shared__ float red[HEAD_THREADS / 32][J];
  const int b = blockIdx.x, F = HW * C;
  float acc[J];
#pragma unroll
  for (int j = 0; j < J; ++j) acc[j] = 0.f;
  const float* f = feat + static_cast<long long>(b) * F;
  for (int idx = threadIdx.x; idx < F; idx += HEAD_THREADS) {
    const int hw = idx / C, c = idx - hw * C;
    const float v = f[idx];
    const float* wc = w + c * HW + hw;
#pragma unroll
    for (int j = 0; j < J; ++j) acc[j] = fmaf(v, __ldg(wc + static_cast<long long>(j) * F), acc[j]);
  }
#pragma unroll
  for (int j = 0; j < J; ++j) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int j = 0; j < J; ++j) red[threadIdx.x >> 5][j] = acc[j];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float z[J], m = -INFINITY;
#pragma unroll
    for (int j = 0; j < J; ++j) {
      float t = bias[j];
      for (int k = 0; k < HEAD_THREADS / 32; ++k) t += red[k][j];
      z[j] = t;
      m = fmaxf(m, t);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < J; ++j) s += expf(z[j] - m);
    const float lse = m + logf(s);
    const long long y = target[b];
#pragma unroll
    for (int j = 0; j < J; ++j) {
      const float lp = z[j] - lse;
      logp[static_cast<long long>(b) * J + j] = lp;
      if (j == y) rowloss[b] = -lp;
    }
    if (y < 0 || y >= J) rowloss[b] = 0.f;
  }
}

// mean of the per-sample losses, fixed order (deterministic)
__global__ void head_loss_kernel(const float* __restrict__ rowloss, int B, float* __restrict__ loss) {
  __shared__ double sh[32];
  double a = 0.0;
  for (int i = threadIdx.x; i < B; i += blockDim.x) a += static_cast<double>(rowloss[i]);
  for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int k = 0; k < static_cast<int>(blockDim.x) / 32; ++k) t += sh[k];
    *loss = static_cast<float>(t / static_cast<double>(B));
  }
}

// dlogits[b][j] = (softmax - onehot) * gscale / B, recomputed from the saved log-probabilities
template <int J>
__device__ __forceinline__ void head_dlogits(const float* logp, const long long* target, int b, float scale, float (&dl)[J]) {
  const long long y = target[b];
#pragma unroll
  for (int j = 0; j < J; ++j) dl[j] = (expf(logp[static_cast<long long>(b) * J + j]) - (j == y ? 1.f : 0.f)) * scale;
}

// ---- backward A: gradient of the features, one block per sample ------------------------------------------
template <int J>
__global__ void __launch_bounds__(HEAD_THREADS) head_bwd_feat_kernel(const float* __restrict__ w, const float* __restrict__ logp,
                                                                     const long long* __restrict__ target,
                                                                     const float* __restrict__ gloss, float* __restrict__ dfeat, int B,
                                                                     int HW, int C) {
  const int b = blockIdx.x, F = HW * C;
  float dl[J];
  head_dlogits<J>(logp, target, b, gloss[0] / static_cast<float>(B), dl);
  for (int idx = threadIdx.x; idx < F; idx += HEAD_THREADS) {
    const int hw = idx / C, c = idx - hw * C;
    const float* wc = w + c * HW + hw;
    float a = 0.f;
#pragma unroll
    for (int j = 0; j < J; ++j) a = fmaf(dl[j], __ldg(wc + static_cast<long long>(j) * F), a);
    dfeat[static_cast<long long>(b) * F + idx] = a;
  }
}

// ---- backward B: weight and bias gradients; thread = one feature, blockIdx.y = one slice of the batch -----------
// The batch is cut into gridDim.y slices so that 128 blocks (not 16) share the work; slices meet in dw / db through
// float atomics (dw / db are the parameters' zeroed-or-accumulating .grad, or were cleared by the host wrapper).
#define HEAD_SLICES 8
template <int J>
__global__ void __launch_bounds__(HEAD_THREADS) head_bwd_w_kernel(const float* __restrict__ feat, const float* __restrict__ logp,
                                                                  const long long* __restrict__ target, const float* __restrict__ gloss,
                                                                  float* __restrict__ dw, float* __restrict__ db, int B, int HW, int C) {
  extern __shared__ float sdl[];  // [slice of B][J]
  const int F = HW * C;
  const int per = (B + gridDim.y - 1) / gridDim.y;
  const int b0 = blockIdx.y * per, b1 = min(B, b0 + per);
  const float scale = gloss[0] / static_cast<float>(B);
  for (int i = b0 + threadIdx.x; i < b1; i += HEAD_THREADS) {
    float dl[J];
    head_dlogits<J>(logp, target, i, scale, dl);
#pragma unroll
    for (int j = 0; j < J; ++j) sdl[(i - b0) * J + j] = dl[j];
  }
  __syncthreads();
  const int idx = blockIdx.x * HEAD_THREADS + threadIdx.x;
  if (idx < F && b1 > b0) {
    float acc[J];
#pragma unroll
    for (int j = 0; j < J; ++j) acc[j] = 0.f;
    for (int b = b0; b < b1; ++b) {
      const float v = feat[static_cast<long long>(b) * F + idx];
#pragma unroll
      for (int j = 0; j < J; ++j) acc[j] = fmaf(sdl[(b - b0) * J + j], v, acc[j]);
    }
    const int hw = idx / C, c = idx - hw * C;
    float* out = dw + c * HW + hw;
#pragma unroll
    for (int j = 0; j < J; ++j) atomicAdd(out + static_cast<long long>(j) * F, acc[j]);
  }
  if (blockIdx.x == 0 && threadIdx.x < J && b1 > b0) {
    float a = 0.f;
    for (int b = b0; b < b1; ++b) a += sdl[(b - b0) * J + threadIdx.x];
    atomicAdd(db + threadIdx.x, a);
  }
}

#define HEAD_DISPATCH(J_, CALL)  \
  switch (J_) {                  \
    case 10: { constexpr int J = 10; CALL; } break; \
    case 16: { constexpr int J = 16; CALL; } break; \
    default: dl_set_error("head: %d classes not instantiated (10 or 16)", J_); return DL_EUNSUPPORTED; \
  }

// feat [B, HW, C] NHWC, w [J, HW*C] with (c, hw) columns, bias [J], target [B] int64 ->
// logp [B, J], loss[1] (mean NLL); rowloss: B floats of scratch.
extern "C" int dl_head_fwd(const float* feat, const float* w, const float* bias, const int64_t* target, float* logp, float* loss,
                           float* rowloss, int B, int HW, int C, int J, void* stream) {
  DL_REQUIRE(feat && w && bias && target && logp && loss && rowloss && B > 0 && HW > 0 && C > 0, "dl_head_fwd: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  HEAD_DISPATCH(J, (head_fwd_kernel<J><<<B, HEAD_THREADS, 0, st>>>(feat, w, bias, reinterpret_cast<const long long*>(target), logp,
                                                                  rowloss, HW, C)));
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  head_loss_kernel<<<1, 256, 0, st>>>(rowloss, B, loss);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}

// gloss: device pointer to the upstream gradient of the (scalar) loss.  dfeat [B, HW, C]; dw [J, HW*C]; db [J];
// accumulate != 0 adds into dw / db (the parameters' .grad).
extern "C" int dl_head_bwd(const float* feat, const float* w, const float* logp, const int64_t* target, const float* gloss,
                           float* dfeat, float* dw, float* db, int B, int HW, int C, int J, int accumulate, void* stream) {
  DL_REQUIRE(feat && w && logp && target && gloss && dfeat && dw && db && B > 0, "dl_head_bwd: bad arguments");
  DL_REQUIRE(static_cast<size_t>(B) * J * sizeof(float) <= 48 * 1024, "dl_head_bwd: batch %d too large for the shared dlogits table", B);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long* tg = reinterpret_cast<const long long*>(target);
  const int F = HW * C;
  HEAD_DISPATCH(J, (head_bwd_feat_kernel<J><<<B, HEAD_THREADS, 0, st>>>(w, logp, tg, gloss, dfeat, B, HW, C)));
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  if (!accumulate) {   // the slices add with atomics: start from zero unless the caller's buffers already accumulate
    DL_CUDA(cudaMemsetAsync(dw, 0, static_cast<size_t>(J) * F * sizeof(float), st));
    DL_CUDA(cudaMemsetAsync(db, 0, static_cast<size_t>(J) * sizeof(float), st));
  }
  const int slices = B >= HEAD_SLICES * 4 ? HEAD_SLICES : 1;
  const int per = (B + slices - 1) / slices;
  HEAD_DISPATCH(J, (head_bwd_w_kernel<J><<<dim3((F + HEAD_THREADS - 1) / HEAD_THREADS, slices), HEAD_THREADS,
                                            static_cast<size_t>(per) * J * sizeof(float), st>>>(feat, logp, tg, gloss, dw, db, B, HW, C)));
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}
