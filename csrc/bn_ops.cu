// Fused SpatialBatchNormalization (training mode) + ReLU + 2x2/2 max-pooling on NHWC activations,
// forward and backward — the elementwise tail of every block of the reference's CIFAR network
// (examples/Model.lua:21-24: batch-norm eps 1e-3, ReLU, SpatialMaxPooling(2,2,2,2)).
//
// Done op by op this tail streams each activation ~17 times per step; here it is read twice in the
// forward pass (statistics, then normalise + ReLU + pool, writing only the pooled quarter) and twice
// in the backward pass (per-channel reductions, then the input gradient), with the ReLU mask and the
// arg-max of each window recomputed from the saved conv output instead of being stored.
//
//   forward   stats : sum, sum of squares per channel                      -> mean, invstd (+ running statistics)
//             apply : p[b,h/2,w/2,c] = max_{2x2} relu(gamma * (y - mean) * invstd + beta)
//   backward  reduce: dz = dp routed to the window arg-max where z > 0;  sum dz, sum dz * xhat per channel
//             apply : dy = gamma * invstd * (dz - mean(dz) - xhat * mean(dz * xhat));  dgamma, dbeta
#include "dl_common.cuh"
#include "dl_host.h"

#define BN_THREADS 256

// ---- forward statistics --------------------------------------------------------------------------
// Thread t of a block owns channel group cg = t % C4 (4 channels) and walks rows r0 + t / C4, stepping by
// rows_per_pass; block partials are combined through shared memory and written out per block (no atomics).
__global__ void __launch_bounds__(BN_THREADS) bn_stats_kernel(const float4* __restrict__ y, long long M, int C4,
                                                              float4* __restrict__ partial) {
  extern __shared__ float4 sh[];  // [2][BN_THREADS]
  const int cg = threadIdx.x % C4;
  const int lane_row = threadIdx.x / C4;
  const int rows_per_pass = BN_THREADS / C4;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f), q = make_float4(0.f, 0.f, 0.f, 0.f);
  if (lane_row < rows_per_pass) {
    for (long long r = static_cast<long long>(blockIdx.x) * rows_per_pass + lane_row; r < M;
         r += static_cast<long long>(gridDim.x) * rows_per_pass) {
      const float4 v = y[r * C4 + cg];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      q.x += v.x * v.x; q.y += v.y * v.y; q.z += v.z * v.z; q.w += v.w * v.w;
    }
  }
  sh[threadIdx.x] = s;
  sh[BN_THREADS + threadIdx.x] = q;
  __syncthreads();
  if (threadIdx.x < C4) {
    float4 ds = make_float4(0.f, 0.f, 0.f, 0.f), dq = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < rows_per_pass; ++k) {
      const float4 a = sh[k * C4 + threadIdx.x], b = sh[BN_THREADS + k * C4 + threadIdx.x];
      ds.x += a.x; ds.y += a.y; ds.z += a.z; ds.w += a.w;
      dq.x += b.x; dq.y += b.y; dq.z += b.z; dq.w += b.w;
    }
    // per-block partial sums; the finalize kernel adds them in a fixed order (deterministic, no atomics)
    partial[(static_cast<long long>(blockIdx.x) * 2 + 0) * C4 + threadIdx.x] = ds;
    partial[(static_cast<long long>(blockIdx.x) * 2 + 1) * C4 + threadIdx.x] = dq;
  }
}

// Finalize kernels: block = 32 channels x 16 slices.  Slice y adds the partial rows y, y+16, ... of its channel in
// double, the 16 slice sums meet in shared memory and thread y == 0 adds them in a fixed order (deterministic).  One
// thread per channel walking all ~300 partial rows alone took 20-40 us per call; this takes a few.
#define FIN_X 32
#define FIN_Y 16
// reductions use two blocks per SM at most: each block leaves one partial row, summed by a finalize kernel
#define BN_MAX_REDUCE_BLOCKS 296
// All of a thread's partial rows are loaded before the first add: the loads are independent, so the kernel costs ONE
// memory round trip instead of nblocks / FIN_Y dependent ones (ncu: 17 us for a 128-channel finalize, nearly all of it
// waiting on 19 loads in a row; profiles/ncu_kernels_r2.md).  Same summation order as before.
#define FIN_MAX_ROWS ((BN_MAX_REDUCE_BLOCKS + FIN_Y - 1) / FIN_Y)
__device__ __forceinline__ double fin_sum(const float* partial, int nblocks, long long row_stride, int c, double (*sh)[FIN_X]) {
  float v[FIN_MAX_ROWS];
#pragma unroll
  for (int k = 0; k < FIN_MAX_ROWS; ++k) {
    const int b = threadIdx.y + k * FIN_Y;
    v[k] = b < nblocks ? __ldcg(partial + static_cast<long long>(b) * row_stride + c) : 0.f;
  }
  double a = 0.0;
#pragma unroll
  for (int k = 0; k < FIN_MAX_ROWS; ++k) a += static_cast<double>(v[k]);
  sh[threadIdx.y][threadIdx.x] = a;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.y == 0)
    for (int k = 0; k < FIN_Y; ++k) t += sh[k][threadIdx.x];
  __syncthreads();
  return t;
}

__global__ void colsum_finalize_kernel(const float* partial, int nblocks, int C, float* out, int accumulate);

// mean / invstd from the sums; running statistics like nn.BatchNorm2d (momentum, unbiased variance)
__global__ void bn_finalize_kernel(const float* partial, int nblocks, long long M, int C, float eps, float momentum,
                                   float* mean, float* invstd, float* running_mean, float* running_var) {
  __shared__ double sh[FIN_Y][FIN_X];
  const int c = blockIdx.x * FIN_X + threadIdx.x;           // C is a multiple of 4; channels >= C read channel C-1 and are dropped
  const int cc = c < C ? c : C - 1;
  const double sum = fin_sum(partial, nblocks, 2LL * C, cc, sh);
  const double sumsq = fin_sum(partial + C, nblocks, 2LL * C, cc, sh);
  if (c >= C || threadIdx.y != 0) return;
  const double m = sum / static_cast<double>(M);
  double var = sumsq / static_cast<double>(M) - m * m;
  if (var < 0) var = 0;
  mean[c] = static_cast<float>(m);
  invstd[c] = static_cast<float>(1.0 / sqrt(var + static_cast<double>(eps)));
  if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * static_cast<float>(m);
  if (running_var) {
    const double unbiased = M > 1 ? var * static_cast<double>(M) / static_cast<double>(M - 1) : var;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * static_cast<float>(unbiased);
  }
}

__device__ __forceinline__ float4 bn_affine(const float4& v, const float4& sc, const float4& sh) {
  return make_float4(fmaf(v.x, sc.x, sh.x), fmaf(v.y, sc.y, sh.y), fmaf(v.z, sc.z, sh.z), fmaf(v.w, sc.w, sh.w));
}

// ---- forward apply: one thread = one pooled output pixel x 4 channels ------------------------------
__global__ void __launch_bounds__(BN_THREADS) bn_relu_pool_fwd_kernel(const float4* __restrict__ y, float4* __restrict__ p,
                                                                      const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                      const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                      int B, int H, int W, int C4) {
  const int Ho = H / 2, Wo = W / 2;
  const long long total = static_cast<long long>(B) * Ho * Wo * C4;
  for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cg = static_cast<int>(idx % C4);
    long long t = idx / C4;
    const int wo = static_cast<int>(t % Wo);
    t /= Wo;
    const int ho = static_cast<int>(t % Ho);
    const long long b = t / Ho;
    const float4 g = reinterpret_cast<const float4*>(gamma)[cg], be = reinterpret_cast<const float4*>(beta)[cg];
    const float4 mu = reinterpret_cast<const float4*>(mean)[cg], is = reinterpret_cast<const float4*>(invstd)[cg];
    const float4 sc = make_float4(g.x * is.x, g.y * is.y, g.z * is.z, g.w * is.w);
    const float4 sh = make_float4(be.x - mu.x * sc.x, be.y - mu.y * sc.y, be.z - mu.z * sc.z, be.w - mu.w * sc.w);
    const long long base = ((b * H + 2 * ho) * W + 2 * wo) * C4 + cg;
    float4 m = make_float4(0.f, 0.f, 0.f, 0.f);  // relu floor
#pragma unroll
    for (int dh = 0; dh < 2; ++dh)
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        const float4 z = bn_affine(y[base + (static_cast<long long>(dh) * W + dw) * C4], sc, sh);
        m.x = fmaxf(m.x, z.x); m.y = fmaxf(m.y, z.y); m.z = fmaxf(m.z, z.z); m.w = fmaxf(m.w, z.w);
      }
    p[idx] = m;
  }
}

// dz of the four window positions for one channel: the pooled gradient goes to the FIRST position (row-major)
// holding the window maximum, and only if that maximum is positive (ReLU) — the convention of max_pool2d.
__device__ __forceinline__ void route(const float z[4], float dp, float dz[4]) {
  int best = 0;
  float bv = z[0];
#pragma unroll
  for (int k = 1; k < 4; ++k)
    if (z[k] > bv) { bv = z[k]; best = k; }
#pragma unroll
  for (int k = 0; k < 4; ++k) dz[k] = (k == best && bv > 0.f) ? dp : 0.f;
}

// ---- backward reduce: per-channel sum of dz and of dz * xhat ----------------------------------------
__global__ void __launch_bounds__(BN_THREADS) bn_relu_pool_bwd_reduce_kernel(const float4* __restrict__ y, const float4* __restrict__ dp,
                                                                             const float* __restrict__ gamma,
                                                                             const float* __restrict__ beta,
                                                                             const float* __restrict__ mean,
                                                                             const float* __restrict__ invstd, int B, int H, int W,
                                                                             int C4, float4* __restrict__ partial) {
  extern __shared__ float4 sh[];
  const int Ho = H / 2, Wo = W / 2;
  const long long windows = static_cast<long long>(B) * Ho * Wo;
  const int cg = threadIdx.x % C4;
  const int lane_row = threadIdx.x / C4;
  const int rows_per_pass = BN_THREADS / C4;
  const float4 g = reinterpret_cast<const float4*>(gamma)[cg], be = reinterpret_cast<const float4*>(beta)[cg];
  const float4 mu = reinterpret_cast<const float4*>(mean)[cg], is = reinterpret_cast<const float4*>(invstd)[cg];
  const float gs[4] = {g.x * is.x, g.y * is.y, g.z * is.z, g.w * is.w};
  const float bs[4] = {be.x - mu.x * gs[0], be.y - mu.y * gs[1], be.z - mu.z * gs[2], be.w - mu.w * gs[3]};
  const float mus[4] = {mu.x, mu.y, mu.z, mu.w}, iss[4] = {is.x, is.y, is.z, is.w};
  float s[4] = {0.f, 0.f, 0.f, 0.f}, q[4] = {0.f, 0.f, 0.f, 0.f};
  if (lane_row < rows_per_pass) {
    for (long long wi = static_cast<long long>(blockIdx.x) * rows_per_pass + lane_row; wi < windows;
         wi += static_cast<long long>(gridDim.x) * rows_per_pass) {
      const int wo = static_cast<int>(wi % Wo);
      const long long t = wi / Wo;
      const int ho = static_cast<int>(t % Ho);
      const long long b = t / Ho;
      const long long base = ((b * H + 2 * ho) * W + 2 * wo) * C4 + cg;
      float4 v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] = y[base + (static_cast<long long>(k >> 1) * W + (k & 1)) * C4];
      const float4 d = dp[wi * C4 + cg];
      const float dpe[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float ye[4], z[4], dz[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          ye[k] = e == 0 ? v[k].x : e == 1 ? v[k].y : e == 2 ? v[k].z : v[k].w;
          z[k] = fmaf(ye[k], gs[e], bs[e]);
        }
        route(z, dpe[e], dz);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          s[e] += dz[k];
          q[e] += dz[k] * (ye[k] - mus[e]) * iss[e];
        }
      }
    }
  }
  sh[threadIdx.x] = make_float4(s[0], s[1], s[2], s[3]);
  sh[BN_THREADS + threadIdx.x] = make_float4(q[0], q[1], q[2], q[3]);
  __syncthreads();
  if (threadIdx.x < C4) {
    float4 ds = make_float4(0.f, 0.f, 0.f, 0.f), dq = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < rows_per_pass; ++k) {
      const float4 a = sh[k * C4 + threadIdx.x], b = sh[BN_THREADS + k * C4 + threadIdx.x];
      ds.x += a.x; ds.y += a.y; ds.z += a.z; ds.w += a.w;
      dq.x += b.x; dq.y += b.y; dq.z += b.z; dq.w += b.w;
    }
    partial[(static_cast<long long>(blockIdx.x) * 2 + 0) * C4 + threadIdx.x] = ds;
    partial[(static_cast<long long>(blockIdx.x) * 2 + 1) * C4 + threadIdx.x] = dq;
  }
}

// sums of the per-block partials of the backward reduction -> sum_dz, sum_dzx (double), dgamma, dbeta
__global__ void bn_bwd_finalize_kernel(const float* partial, int nblocks, int C, double* sums, float* dgamma, float* dbeta,
                                       int accumulate) {
  __shared__ double sh[FIN_Y][FIN_X];
  const int c = blockIdx.x * FIN_X + threadIdx.x;
  const int cc = c < C ? c : C - 1;
  const double a = fin_sum(partial, nblocks, 2LL * C, cc, sh);
  const double b = fin_sum(partial + C, nblocks, 2LL * C, cc, sh);
  if (c >= C || threadIdx.y != 0) return;
  sums[c] = a;
  sums[C + c] = b;
  // accumulate: add into the caller's gradient buffers (the parameters' .grad) instead of overwriting them
  dbeta[c] = (accumulate ? dbeta[c] : 0.f) + static_cast<float>(a);
  dgamma[c] = (accumulate ? dgamma[c] : 0.f) + static_cast<float>(b);
}

// ---- backward apply: dy for the four positions of each window ----------------------------------------
__global__ void __launch_bounds__(BN_THREADS) bn_relu_pool_bwd_apply_kernel(const float4* __restrict__ y, const float4* __restrict__ dp,
                                                                            float4* __restrict__ dy, const float* __restrict__ gamma,
                                                                            const float* __restrict__ beta,
                                                                            const float* __restrict__ mean,
                                                                            const float* __restrict__ invstd, int B, int H, int W,
                                                                            int C4, const double* sum_dz, const double* sum_dzx,
                                                                            float4* __restrict__ colpart) {
  // colpart != NULL: also leave the per-block column sums of dy (the bias gradient of the convolution in front of this
  // block: dy IS the gradient of its output) so that nobody has to read dy again for them.  The grid stride is a multiple
  // of C4 (256 % C4 == 0), so a thread keeps its channel group for the whole loop and sums in registers.
  extern __shared__ float4 sh_col[];
  float4 csum = make_float4(0.f, 0.f, 0.f, 0.f);
  const int Ho = H / 2, Wo = W / 2;
  const long long windows = static_cast<long long>(B) * Ho * Wo;
  const double invM = 1.0 / (static_cast<double>(windows) * 4.0);
  const long long total = windows * C4;
  for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cg = static_cast<int>(idx % C4);
    const long long wi = idx / C4;
    const int wo = static_cast<int>(wi % Wo);
    const long long t = wi / Wo;
    const int ho = static_cast<int>(t % Ho);
    const long long b = t / Ho;
    const float4 g = reinterpret_cast<const float4*>(gamma)[cg], be = reinterpret_cast<const float4*>(beta)[cg];
    const float4 mu = reinterpret_cast<const float4*>(mean)[cg], is = reinterpret_cast<const float4*>(invstd)[cg];
    const float gs[4] = {g.x * is.x, g.y * is.y, g.z * is.z, g.w * is.w};
    const float bs[4] = {be.x - mu.x * gs[0], be.y - mu.y * gs[1], be.z - mu.z * gs[2], be.w - mu.w * gs[3]};
    const float mus[4] = {mu.x, mu.y, mu.z, mu.w}, iss[4] = {is.x, is.y, is.z, is.w};
    const long long base = ((b * H + 2 * ho) * W + 2 * wo) * C4 + cg;
    float4 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = y[base + (static_cast<long long>(k >> 1) * W + (k & 1)) * C4];
    const float4 d = dp[idx];
    const float dpe[4] = {d.x, d.y, d.z, d.w};
    float out[4][4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float m_dz = static_cast<float>(sum_dz[cg * 4 + e] * invM), m_dzx = static_cast<float>(sum_dzx[cg * 4 + e] * invM);
      float ye[4], z[4], dz[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        ye[k] = e == 0 ? v[k].x : e == 1 ? v[k].y : e == 2 ? v[k].z : v[k].w;
        z[k] = fmaf(ye[k], gs[e], bs[e]);
      }
      route(z, dpe[e], dz);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float xhat = (ye[k] - mus[e]) * iss[e];
        out[k][e] = gs[e] * (dz[k] - m_dz - xhat * m_dzx);
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      dy[base + (static_cast<long long>(k >> 1) * W + (k & 1)) * C4] = make_float4(out[k][0], out[k][1], out[k][2], out[k][3]);
      csum.x += out[k][0]; csum.y += out[k][1]; csum.z += out[k][2]; csum.w += out[k][3];
    }
  }
  if (colpart != nullptr) {
    sh_col[threadIdx.x] = csum;
    __syncthreads();
    if (threadIdx.x < C4) {
      const int rows_per_pass = BN_THREADS / C4;
      float4 ds = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k = 0; k < rows_per_pass; ++k) {
        const float4 a = sh_col[k * C4 + threadIdx.x];
        ds.x += a.x; ds.y += a.y; ds.z += a.z; ds.w += a.w;
      }
      colpart[static_cast<long long>(blockIdx.x) * C4 + threadIdx.x] = ds;
    }
  }
}

static int bn_grid(long long items, int per_block) {
  long long nb = (items + per_block - 1) / per_block;
  const long long cap = static_cast<long long>(dl_sm_count()) * 8;
  if (nb > cap) nb = cap;
  return nb < 1 ? 1 : static_cast<int>(nb);
}

// (BN_MAX_REDUCE_BLOCKS is defined next to the finalize kernels above)
static int bn_reduce_grid(long long rows, int rows_per_pass) {
  long long nb = (rows + static_cast<long long>(rows_per_pass) * 8 - 1) / (static_cast<long long>(rows_per_pass) * 8);
  if (nb > BN_MAX_REDUCE_BLOCKS) nb = BN_MAX_REDUCE_BLOCKS;
  return nb < 1 ? 1 : static_cast<int>(nb);
}

// scratch: per-block partial sums (float) followed by 2*C doubles
extern "C" size_t dl_bn_work_bytes(int C) {
  // reduction partials (2 rows per block) + 2 C doubles + the column-sum partials of the fused conv-bias gradient
  return static_cast<size_t>(BN_MAX_REDUCE_BLOCKS) * 3 * C * sizeof(float) + static_cast<size_t>(C) * 2 * sizeof(double);
}

// y [B,H,W,C] -> p [B,H/2,W/2,C]; writes mean/invstd [C] (saved for backward) and updates the running statistics.
// work: dl_bn_work_bytes(C) bytes of device scratch (no initialisation needed).
extern "C" int dl_bn_relu_pool_fwd(const float* y, float* p, const float* gamma, const float* beta, float* mean, float* invstd,
                                   float* running_mean, float* running_var, int B, int H, int W, int C, float eps, float momentum,
                                   void* work, void* stream) {
  DL_REQUIRE(y && p && gamma && beta && mean && invstd && work, "dl_bn_relu_pool_fwd: null pointer");
  DL_REQUIRE(C % 4 == 0 && C / 4 <= BN_THREADS && H % 2 == 0 && W % 2 == 0, "dl_bn_relu_pool_fwd: need C %% 4 == 0, C <= %d, even H and W",
             BN_THREADS * 4);
  DL_REQUIRE(reinterpret_cast<uintptr_t>(work) % 16 == 0, "dl_bn_relu_pool_fwd: work must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long M = static_cast<long long>(B) * H * W;
  const int C4 = C / 4;
  float* partial = static_cast<float*>(work);
  const int rows_per_pass = BN_THREADS / C4;
  const int nb = bn_reduce_grid(M, rows_per_pass);
  bn_stats_kernel<<<nb, BN_THREADS, 2 * BN_THREADS * sizeof(float4), st>>>(reinterpret_cast<const float4*>(y), M, C4,
                                                                         reinterpret_cast<float4*>(partial));
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  bn_finalize_kernel<<<(C + FIN_X - 1) / FIN_X, dim3(FIN_X, FIN_Y), 0, st>>>(partial, nb, M, C, eps, momentum, mean, invstd, running_mean, running_var);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  const long long items = M / 4 * C4;
  bn_relu_pool_fwd_kernel<<<bn_grid(items, BN_THREADS), BN_THREADS, 0, st>>>(reinterpret_cast<const float4*>(y),
                                                                              reinterpret_cast<float4*>(p), gamma, beta, mean,
                                                                              invstd, B, H, W, C4);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}

// dp [B,H/2,W/2,C] -> dy [B,H,W,C], dgamma [C], dbeta [C]; y, mean, invstd as saved by the forward call.
extern "C" int dl_bn_relu_pool_bwd(const float* y, const float* dp, float* dy, const float* gamma, const float* beta,
                                   const float* mean, const float* invstd, float* dgamma, float* dbeta, int B, int H, int W, int C,
                                   int accumulate, void* work, void* stream) {
  return dl_bn_relu_pool_bwd_bias(y, dp, dy, gamma, beta, mean, invstd, dgamma, dbeta, nullptr, B, H, W, C, accumulate, work, stream);
}

// Same, and with dconvbias != NULL the column sums of dy — the bias gradient of the convolution that produced y — come
// out of the same pass (added into dconvbias when accumulate != 0), instead of a separate dl_colsum over dy.
extern "C" int dl_bn_relu_pool_bwd_bias(const float* y, const float* dp, float* dy, const float* gamma, const float* beta,
                                        const float* mean, const float* invstd, float* dgamma, float* dbeta, float* dconvbias, int B,
                                        int H, int W, int C, int accumulate, void* work, void* stream) {
  DL_REQUIRE(y && dp && dy && gamma && beta && mean && invstd && dgamma && dbeta && work, "dl_bn_relu_pool_bwd: null pointer");
  DL_REQUIRE(C % 4 == 0 && C / 4 <= BN_THREADS && H % 2 == 0 && W % 2 == 0, "dl_bn_relu_pool_bwd: need C %% 4 == 0, C <= %d, even H and W",
             BN_THREADS * 4);
  DL_REQUIRE(reinterpret_cast<uintptr_t>(work) % 16 == 0, "dl_bn_relu_pool_bwd: work must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int C4 = C / 4;
  const long long windows = static_cast<long long>(B) * (H / 2) * (W / 2);
  float* partial = static_cast<float*>(work);
  double* sums = reinterpret_cast<double*>(partial + static_cast<size_t>(BN_MAX_REDUCE_BLOCKS) * 2 * C);
  const int rows_per_pass = BN_THREADS / C4;
  const int nb = bn_reduce_grid(windows, rows_per_pass);
  bn_relu_pool_bwd_reduce_kernel<<<nb, BN_THREADS, 2 * BN_THREADS * sizeof(float4), st>>>(
      reinterpret_cast<const float4*>(y), reinterpret_cast<const float4*>(dp), gamma, beta, mean, invstd, B, H, W, C4,
      reinterpret_cast<float4*>(partial));
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  bn_bwd_finalize_kernel<<<(C + FIN_X - 1) / FIN_X, dim3(FIN_X, FIN_Y), 0, st>>>(partial, nb, C, sums, dgamma, dbeta, accumulate);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  float* colpart = reinterpret_cast<float*>(sums + 2 * static_cast<size_t>(C));
  int grid = bn_grid(windows * C4, BN_THREADS);
  if (dconvbias != nullptr && grid > BN_MAX_REDUCE_BLOCKS) grid = BN_MAX_REDUCE_BLOCKS;    // one partial row per block
  bn_relu_pool_bwd_apply_kernel<<<grid, BN_THREADS, dconvbias ? BN_THREADS * sizeof(float4) : 0, st>>>(
      reinterpret_cast<const float4*>(y), reinterpret_cast<const float4*>(dp), reinterpret_cast<float4*>(dy), gamma, beta, mean, invstd,
      B, H, W, C4, sums, sums + C, dconvbias ? reinterpret_cast<float4*>(colpart) : nullptr);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  if (dconvbias != nullptr) {
    colsum_finalize_kernel<<<(C + FIN_X - 1) / FIN_X, dim3(FIN_X, FIN_Y), 0, st>>>(colpart, grid, C, dconvbias, accumulate);
    dl_note_launch();
    DL_CUDA(cudaGetLastError());
  }
  return DL_OK;
}

// Column sums of a row-major [M, C] fp32 matrix (the convolution's bias gradient, sum over pixels of dY):
// same two-stage reduction.  work: dl_bn_work_bytes(C) bytes.
__global__ void __launch_bounds__(BN_THREADS) colsum_kernel(const float4* __restrict__ x, long long M, int C4, float4* __restrict__ partial) {
  extern __shared__ float4 sh[];
  const int cg = threadIdx.x % C4;
  const int lane_row = threadIdx.x / C4;
  const int rows_per_pass = BN_THREADS / C4;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (lane_row < rows_per_pass) {
    for (long long r = static_cast<long long>(blockIdx.x) * rows_per_pass + lane_row; r < M;
         r += static_cast<long long>(gridDim.x) * rows_per_pass) {
      const float4 v = x[r * C4 + cg];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
  }
  sh[threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.x < C4) {
    float4 ds = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < rows_per_pass; ++k) {
      const float4 a = sh[k * C4 + threadIdx.x];
      ds.x += a.x; ds.y += a.y; ds.z += a.z; ds.w += a.w;
    }
    partial[static_cast<long long>(blockIdx.x) * C4 + threadIdx.x] = ds;
  }
}

__global__ void colsum_finalize_kernel(const float* partial, int nblocks, int C, float* out, int accumulate) {
  __shared__ double sh[FIN_Y][FIN_X];
  const int c = blockIdx.x * FIN_X + threadIdx.x;
  const int cc = c < C ? c : C - 1;
  const double a = fin_sum(partial, nblocks, C, cc, sh);
  if (c >= C || threadIdx.y != 0) return;
  out[c] = (accumulate ? out[c] : 0.f) + static_cast<float>(a);
}

extern "C" int dl_colsum(const float* x, float* out, long long M, int C, int accumulate, void* work, void* stream) {
  DL_REQUIRE(x && out && work && M > 0, "dl_colsum: bad arguments");
  DL_REQUIRE(C % 4 == 0 && C / 4 <= BN_THREADS, "dl_colsum: need C %% 4 == 0 and C <= %d", BN_THREADS * 4);
  DL_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0 && reinterpret_cast<uintptr_t>(work) % 16 == 0, "dl_colsum: 16-byte alignment required");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int C4 = C / 4;
  const int rows_per_pass = BN_THREADS / C4;
  const int nb = bn_reduce_grid(M, rows_per_pass);
  colsum_kernel<<<nb, BN_THREADS, BN_THREADS * sizeof(float4), st>>>(reinterpret_cast<const float4*>(x), M, C4,
                                                                   reinterpret_cast<float4*>(work));
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  colsum_finalize_kernel<<<(C + FIN_X - 1) / FIN_X, dim3(FIN_X, FIN_Y), 0, st>>>(static_cast<const float*>(work), nb, C, out, accumulate);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}
