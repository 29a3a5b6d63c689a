// im2col / col2im for 5x5, stride 1, pad 2 convolutions on NHWC activations — the data movement around
// the tensor-core GEMM (csrc/tc_gemm.cu).  The reference's convolution is SpatialConvolutionMM
// (examples/Model.lua:20,26,32,38): im2col + SGEMM; this is the same decomposition with the GEMM on
// tcgen05.  Column order of `col` is (kh, kw, cin), so a row of `col` is 25 contiguous channel vectors and
// every access below is a contiguous run of channels.
//
//   im2col : x[B,H,W,C]            -> col[B*H*W, ld]   (ld >= 25*C, padding columns are written as 0)
//   col2im : dcol[B*H*W, ld]       -> dx[B,H,W,C]      dx[b,h,w,:] = sum over the 25 taps that cover (h,w)
#include "dl_common.cuh"
#include "dl_host.h"

#define CONV_K 5
#define CONV_PAD 2
#define CONV_TAPS 25

template <int VEC>  // VEC = 4: C % 4 == 0, float4 per thread; VEC = 1: scalar
__global__ void __launch_bounds__(256) im2col5x5_kernel(const float* __restrict__ x, float* __restrict__ col, int B, int H, int W,
                                                        int C, long long ld) {
  const int CV = C / VEC;
  const long long row_items = ld / VEC;  // items per col row (including padding columns)
  const long long total = static_cast<long long>(B) * H * W * row_items;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    const long long m = idx / row_items;
    const int r = static_cast<int>(idx - m * row_items);
    const int tap = r / CV, cv = r - tap * CV;
    const int w = static_cast<int>(m % W);
    const long long t = m / W;
    const int h = static_cast<int>(t % H);
    const long long b = t / H;
    const int hi = h + tap / CONV_K - CONV_PAD, wi = w + tap % CONV_K - CONV_PAD;
    const bool in = tap < CONV_TAPS && hi >= 0 && hi < H && wi >= 0 && wi < W;
    if (VEC == 4) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (in) v = reinterpret_cast<const float4*>(x)[((b * H + hi) * W + wi) * CV + cv];
      reinterpret_cast<float4*>(col)[idx] = v;
    } else {
      col[idx] = in ? x[((b * H + hi) * W + wi) * C + cv] : 0.f;
    }
  }
}

// Any C (the first layer has C = 3): one thread produces four consecutive columns of a col row (four scalar
// gathers, one 16-byte store), so the 40 MB col matrix of layer 1 is written with full-width stores.
__global__ void __launch_bounds__(256) im2col5x5_gather4_kernel(const float* __restrict__ x, float* __restrict__ col, int B, int H,
                                                                int W, int C, long long ld) {
  const long long row_items = ld / 4;
  const long long total = static_cast<long long>(B) * H * W * row_items;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const int kc = CONV_TAPS * C;
  for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    const long long m = idx / row_items;
    const int k0 = static_cast<int>(idx - m * row_items) * 4;
    const int w = static_cast<int>(m % W);
    const long long t = m / W;
    const int h = static_cast<int>(t % H);
    const long long b = t / H;
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int k = k0 + e;
      const int tap = k / C, c = k - tap * C;
      const int hi = h + tap / CONV_K - CONV_PAD, wi = w + tap % CONV_K - CONV_PAD;
      const bool in = k < kc && hi >= 0 && hi < H && wi >= 0 && wi < W;
      v[e] = in ? __ldg(x + ((b * H + hi) * W + wi) * C + c) : 0.f;
    }
    reinterpret_cast<float4*>(col)[idx] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

template <int VEC>
__global__ void __launch_bounds__(256) col2im5x5_kernel(const float* __restrict__ dcol, float* __restrict__ dx, int B, int H, int W,
                                                        int C, long long ld) {
  const int CV = C / VEC;
  const long long total = static_cast<long long>(B) * H * W * CV;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    const long long m = idx / CV;
    const int cv = static_cast<int>(idx - m * CV);
    const int w = static_cast<int>(m % W);
    const long long t = m / W;
    const int h = static_cast<int>(t % H);
    const long long b = t / H;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int kh = 0; kh < CONV_K; ++kh) {
      const int ho = h - kh + CONV_PAD;  // the output pixel whose tap (kh, kw) reads input pixel (h, w)
      if (ho < 0 || ho >= H) continue;
#pragma unroll
      for (int kw = 0; kw < CONV_K; ++kw) {
        const int wo = w - kw + CONV_PAD;
        if (wo < 0 || wo >= W) continue;
        const long long row = (b * H + ho) * W + wo;
        const int tap = kh * CONV_K + kw;
        if (VEC == 4) {
          const float4 v = reinterpret_cast<const float4*>(dcol + row * ld)[tap * CV + cv];
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        } else {
          acc.x += dcol[row * ld + tap * C + cv];
        }
      }
    }
    if (VEC == 4) reinterpret_cast<float4*>(dx)[idx] = acc;
    else dx[idx] = acc.x;
  }
}

static int conv_grid(long long items) {
  long long nb = (items + 255) / 256;
  const long long cap = static_cast<long long>(dl_sm_count()) * 16;
  if (nb > cap) nb = cap;
  return nb < 1 ? 1 : static_cast<int>(nb);
}

extern "C" int dl_im2col5x5(const float* x, float* col, int B, int H, int W, int C, long long ld, void* stream) {
  DL_REQUIRE(x && col && B > 0 && H > 0 && W > 0 && C > 0, "dl_im2col5x5: bad arguments");
  DL_REQUIRE(ld >= static_cast<long long>(CONV_TAPS) * C, "dl_im2col5x5: ld %lld is smaller than 25*C", ld);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool vec = C % 4 == 0 && ld % 4 == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(col) % 16 == 0;
  const bool gather4 = !vec && ld % 4 == 0 && reinterpret_cast<uintptr_t>(col) % 16 == 0;
  const long long items = static_cast<long long>(B) * H * W * ((vec || gather4) ? ld / 4 : ld);
  if (vec) im2col5x5_kernel<4><<<conv_grid(items), 256, 0, st>>>(x, col, B, H, W, C, ld);
  else if (gather4) im2col5x5_gather4_kernel<<<conv_grid(items), 256, 0, st>>>(x, col, B, H, W, C, ld);
  else im2col5x5_kernel<1><<<conv_grid(items), 256, 0, st>>>(x, col, B, H, W, C, ld);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}

extern "C" int dl_col2im5x5(const float* dcol, float* dx, int B, int H, int W, int C, long long ld, void* stream) {
  DL_REQUIRE(dcol && dx && B > 0 && H > 0 && W > 0 && C > 0, "dl_col2im5x5: bad arguments");
  DL_REQUIRE(ld >= static_cast<long long>(CONV_TAPS) * C, "dl_col2im5x5: ld %lld is smaller than 25*C", ld);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool vec = C % 4 == 0 && ld % 4 == 0 && reinterpret_cast<uintptr_t>(dx) % 16 == 0 && reinterpret_cast<uintptr_t>(dcol) % 16 == 0;
  const long long items = static_cast<long long>(B) * H * W * (vec ? C / 4 : C);
  if (vec) col2im5x5_kernel<4><<<conv_grid(items), 256, 0, st>>>(dcol, dx, B, H, W, C, ld);
  else col2im5x5_kernel<1><<<conv_grid(items), 256, 0, st>>>(dcol, dx, B, H, W, C, ld);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}
