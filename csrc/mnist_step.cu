// One training step of the reference's MNIST network as ONE kernel — SURVEY C6/C7, §2.7.
//
// The model of examples/mnist.lua:53-81 is tiny (10,842 parameters, 2.5 MFLOP per sample) and the reference trains
// it at batch size 1 (:33): forward + backward + SGD is ~25 library kernels of a few microseconds each, 95 us per step
// even inside a CUDA graph, and that is what every MNIST configuration really measures.  Here the whole step
//
//     Reshape(1,32,32) -> conv 1->16 5x5 -> tanh -> maxpool 2 -> conv 16->16 5x5 -> tanh -> maxpool 2
//     -> Linear 400->10 -> logSoftMax -> NLL,  its backward pass,  and  p -= lr * g            (:53-66, :79, :86, :112-116)
//
// runs in one launch of a thread-block CLUSTER of 8 CTAs.  CTA r owns feature channels 2r and 2r+1 of both convolution
// layers end to end — their filters, activations, gradients and the matching 50 columns of the Linear layer live in
// its shared memory — and the three places where channels meet go through distributed shared memory:
//   * after conv1 every CTA writes its two pooled maps into the shared memory of all eight (all-gather);
//   * the Linear layer's ten partial sums of each CTA are written to all eight and added in rank order (all-reduce);
//   * conv2's input gradient: CTA r forms the contribution of ITS two output channels to all 16 input channels and
//     writes channel pair j's part into CTA j's slot r; CTA j adds its eight slots in rank order (reduce-scatter).
// Three cluster barriers per sample, no global memory traffic besides the 4 KB input and (once per launch) the
// parameters.  Every sum has a fixed order: two runs give the same bits.
//
// Batches: the cluster walks its samples one after another and accumulates the gradients in shared memory; with more
// than one cluster (B > 8 or so) each takes every n-th sample and the gradients meet in global memory with atomics, the
// update then being a separate dl_local_sgd.  With ONE cluster and apply_sgd != 0 the update is fused: one launch per step.
#include <cooperative_groups.h>
#include <cstdlib>

#include "dl_common.cuh"
#include "dl_host.h"

namespace cg = cooperative_groups;

#define MN_CL 8      // CTAs per cluster: 2 of the 16 channels each
#define MN_T 320     // threads per CTA = 2 output channels x 10 output rows x 16 input channels (conv2's natural shape)
#define MN_T2 416    // second form: 13 warps — conv1 in one round (392 windows), conv2's two backward passes side by side (7 + 5 warps)

struct MnistArgs {
  const float* x;            // [B, 1024]
  const long long* y;        // [B]
  float* p[6];               // conv1.weight [16,1,5,5], conv1.bias [16], conv2.weight [16,16,5,5], conv2.bias [16],
  float* g[6];               //   linear.weight [10,400], linear.bias [10]  — and their gradients (may alias a flat vector)
  float* loss;               // mean NLL of the batch
  float* logp;               // [B, 10] or NULL
  int B;
  float lr;
  int apply_sgd;             // 1: p -= lr * g inside the kernel (single cluster only); 0: g += gradient
};

struct MnistSmem {
  float x[1024];
  float w1[2][25], b1[2];
  float a1[2][28 * 28];        // tanh(conv1); overwritten by the gradient w.r.t. conv1's pre-activation in the backward pass
  unsigned char i1[2][196];    // arg-max position (0..3) inside each pooling window
  alignas(16) float p1[16][196];   // pooled conv1 maps of ALL channels (gathered through DSMEM)
  float w2[2][16][25], b2[2];
  float a2[2][100];            // tanh(conv2), later its pre-activation gradient
  unsigned char i2[2][25];
  float p2[2][25];             // this CTA's 50 features
  float w3[10][50], b3[10];
  float lpart[MN_CL][10];      // partial logits of every CTA (all-reduce slots)
  float dl[10];                // dlogits
  float dfeat[2][25];
  float dpart[MN_CL][2][196];  // conv2 dgrad contributions for MY two channels, one slot per source CTA (second form: the
                               // same words hold MY contributions to all 16 channels, [16][196], and the owners fetch them)
  float gw1[2][25], gb1[2], gw2[2][16][25], gb2[2], gw3[10][50], gb3[10];   // gradient accumulators over the batch
  float loss_acc;
};

__device__ __forceinline__ float warp_sum16(float v) {  // sum over the 16 lanes of a half warp (lanes differing in bits 0..3)
  v += __shfl_xor_sync(0xffffffffu, v, 8);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v;
}

// kV2 = false: the first form (320 threads).  kV2 = true: the form written after its ncu capture
// (profiles/ncu_tails_r2.md: 15 % of the issue slots used, 34 % of the samples in conv1's filter gradient, a single
// chain of 196 dependent FMAs per thread on shared-memory reads that collide 6-fold):
//   * conv1's filter gradient: warp = (channel, filter row), lane = output column — contiguous conflict-free reads,
//     five independent sums per lane, the bias gradient falls out of the same loop;
//   * conv2's weight gradient and input gradient on different warps (416 threads) instead of one after the other;
//   * one tanh per thread after conv2 instead of ten under a predicate; the label is fetched at the top of the sample,
//     not inside the softmax; the update reads the parameters from shared memory, where they already are.
// Same sums up to the order of the additions inside conv1's filter gradient.
template <bool kV2>
__global__ void __cluster_dims__(MN_CL, 1, 1) __launch_bounds__(kV2 ? MN_T2 : MN_T, 1) mnist_step_kernel(MnistArgs a) {
  constexpr int T = kV2 ? MN_T2 : MN_T;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  MnistSmem& s = *reinterpret_cast<MnistSmem*>(smem_raw);
  cg::cluster_group cluster = cg::this_cluster();
  const int r = static_cast<int>(cluster.block_rank());          // owns channels 2r, 2r+1
  const int cid = static_cast<int>(blockIdx.x) / MN_CL;           // which cluster
  const int ncl = static_cast<int>(gridDim.x) / MN_CL;
  const int t = threadIdx.x;
  MnistSmem* peer[MN_CL];
#pragma unroll
  for (int k = 0; k < MN_CL; ++k) peer[k] = cluster.map_shared_rank(&s, k);

  // ---- parameters of my channels into shared memory, gradient accumulators to zero ---------------------------
  for (int i = t; i < 50; i += T) (&s.w1[0][0])[i] = a.p[0][(2 * r) * 25 + i];
  for (int i = t; i < 800; i += T) (&s.w2[0][0][0])[i] = a.p[2][(2 * r) * 400 + i];
  for (int i = t; i < 500; i += T) {
    const int j = i / 50, f = i - j * 50;
    s.w3[j][f] = a.p[4][j * 400 + 2 * r * 25 + f];
  }
  if (t < 2) { s.b1[t] = a.p[1][2 * r + t]; s.b2[t] = a.p[3][2 * r + t]; }
  if (t < 10) s.b3[t] = a.p[5][t];
  for (int i = t; i < 50; i += T) (&s.gw1[0][0])[i] = 0.f;
  for (int i = t; i < 800; i += T) (&s.gw2[0][0][0])[i] = 0.f;
  for (int i = t; i < 500; i += T) (&s.gw3[0][0])[i] = 0.f;
  if (t < 2) { s.gb1[t] = 0.f; s.gb2[t] = 0.f; }
  if (t < 10) s.gb3[t] = 0.f;
  if (t == 0) s.loss_acc = 0.f;
  const float inv_b = 1.f / static_cast<float>(a.B);
  cluster.sync();   // everybody's shared memory exists and is initialised before anyone writes into it remotely

  for (int b = cid; b < a.B; b += ncl) {
    // ---- input -------------------------------------------------------------------------------------------------
    const float4* xin = reinterpret_cast<const float4*>(a.x + static_cast<long long>(b) * 1024);
    long long y_early = 0;
    if (kV2 && t < 32) y_early = a.y[b];      // in flight while the convolutions run
    for (int i = t; i < 256; i += T) reinterpret_cast<float4*>(s.x)[i] = xin[i];
    __syncthreads();

    // ---- conv1 + tanh + 2x2 max-pool for my two channels: one item = one pooling window -----------------------------
    for (int it = t; it < 2 * 196; it += T) {
      const int c = it / 196, pos = it - c * 196, py = pos / 14, px = pos - py * 14;
      float in[6][6];
#pragma unroll
      for (int dy = 0; dy < 6; ++dy)
#pragma unroll
        for (int dx = 0; dx < 6; ++dx) in[dy][dx] = s.x[(2 * py + dy) * 32 + 2 * px + dx];
      float acc[4] = {s.b1[c], s.b1[c], s.b1[c], s.b1[c]};
#pragma unroll
      for (int ky = 0; ky < 5; ++ky)
#pragma unroll
        for (int kx = 0; kx < 5; ++kx) {
          const float w = s.w1[c][ky * 5 + kx];
          acc[0] = fmaf(w, in[ky][kx], acc[0]);
          acc[1] = fmaf(w, in[ky][kx + 1], acc[1]);
          acc[2] = fmaf(w, in[ky + 1][kx], acc[2]);
          acc[3] = fmaf(w, in[ky + 1][kx + 1], acc[3]);
        }
      float best = -2.f;
      int arg = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float v = tanhf(acc[k]);
        s.a1[c][(2 * py + (k >> 1)) * 28 + 2 * px + (k & 1)] = v;
        if (v > best) { best = v; arg = k; }      // first maximum in row-major order, like max_pool2d
      }
      s.i1[c][pos] = static_cast<unsigned char>(arg);
      if (kV2) {
        s.p1[2 * r + c][pos] = best;             // the peers fetch it after the barrier (below)
      } else {
#pragma unroll
        for (int k = 0; k < MN_CL; ++k) peer[k]->p1[2 * r + c][pos] = best;   // all-gather through distributed shared memory
      }
    }
    cluster.sync();   // [1] first form: every CTA holds all 16 pooled maps; second form: every CTA's two maps are final
    if (kV2) {
      // Pull instead of push: a barrier with remote stores in flight waits until the last of them has landed (ERRBAR, 20 % of the
      // samples of the first capture of this form); remote LOADS are 16 bytes wide, issued back to back, and nothing waits
      // for them but their user.  98 float4 per peer, 7 peers.
      for (int i = t; i < (MN_CL - 1) * 98; i += T) {
        const int k0 = i / 98, q = i - k0 * 98, k = k0 + (k0 >= r ? 1 : 0);
        reinterpret_cast<float4*>(&s.p1[2 * k][0])[q] = reinterpret_cast<const float4*>(&peer[k]->p1[2 * k][0])[q];
      }
      __syncthreads();
    }

    // ---- conv2 + tanh for my two output channels: thread = (co, output row, input channel) -------------------------
    if (!kV2 || t < MN_T) {   // (whole warps: 320 = 10 x 32)
      const int ci = t & 15, oy = (t >> 4) % 10, co = t / 160;
      float acc[10];
#pragma unroll
      for (int ox = 0; ox < 10; ++ox) acc[ox] = 0.f;
#pragma unroll
      for (int ky = 0; ky < 5; ++ky) {
        float row[14];
#pragma unroll
        for (int xx = 0; xx < 14; ++xx) row[xx] = s.p1[ci][(oy + ky) * 14 + xx];
#pragma unroll
        for (int kx = 0; kx < 5; ++kx) {
          const float w = s.w2[co][ci][ky * 5 + kx];
#pragma unroll
          for (int ox = 0; ox < 10; ++ox) acc[ox] = fmaf(w, row[ox + kx], acc[ox]);
        }
      }
      if (kV2) {
        float mine = 0.f;
#pragma unroll
        for (int ox = 0; ox < 10; ++ox) {
          const float v = warp_sum16(acc[ox]);    // every lane of the half warp ends up with the sum
          if (ci == ox) mine = v;
        }
        if (ci < 10) s.a2[co][oy * 10 + ci] = tanhf(mine + s.b2[co]);
      } else {
#pragma unroll
        for (int ox = 0; ox < 10; ++ox) {
          const float v = warp_sum16(acc[ox]);      // over the 16 input channels, fixed butterfly order
          if (ci == 0) s.a2[co][oy * 10 + ox] = tanhf(v + s.b2[co]);
        }
      }
    }
    __syncthreads();
    if (t < 50) {  // 2x2 max-pool -> my 50 features
      const int c = t / 25, pos = t - c * 25, py = pos / 5, px = pos - py * 5;
      float best = -2.f;
      int arg = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float v = s.a2[c][(2 * py + (k >> 1)) * 10 + 2 * px + (k & 1)];
        if (v > best) { best = v; arg = k; }
      }
      s.p2[c][pos] = best;
      s.i2[c][pos] = static_cast<unsigned char>(arg);
    }
    __syncthreads();
    // ---- Linear: warp j forms my 50-feature part of logit j and hands it to every CTA ----------------------------
    if (!kV2 || t < MN_T) {
      const int j = t >> 5, lane = t & 31;
      const float* f = &s.p2[0][0];
      float v = s.w3[j][lane] * f[lane] + (lane < 18 ? s.w3[j][32 + lane] * f[32 + lane] : 0.f);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane < MN_CL) peer[lane]->lpart[r][j] = v;
    }
    cluster.sync();   // [2] every CTA holds all partial logits

    // ---- log-softmax, loss, dlogits: every CTA does the same sums in the same order (lane j of warp 0 owns class j) ------
    if (t < 32) {
      float z = -INFINITY;
      if (t < 10) {
        z = s.b3[t];
#pragma unroll
        for (int k = 0; k < MN_CL; ++k) z += s.lpart[k][t];
      }
      float m = z;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
      float sum = t < 10 ? expf(z - m) : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
      const float lse = m + logf(sum);
      const long long yy = kV2 ? y_early : a.y[b];
      if (t < 10) {
        const float lp = z - lse;
        s.dl[t] = (expf(lp) - (t == yy ? 1.f : 0.f)) * inv_b;       // d(mean NLL)/dlogit
        if (r == 0 && a.logp != nullptr) a.logp[static_cast<long long>(b) * 10 + t] = lp;
        if (t == yy) s.loss_acc -= lp;
      }
    }
    __syncthreads();

    // ---- Linear backward for my 50 features; un-pool and tanh' into conv2's pre-activation gradient ----------------
    for (int i = t; i < 500; i += T) {
      const int j = i / 50, f = i - j * 50;
      s.gw3[j][f] = fmaf(s.dl[j], (&s.p2[0][0])[f], s.gw3[j][f]);
    }
    if (t < 10) s.gb3[t] += s.dl[t];
    if (t >= 32 && t < 82) {
      const int f = t - 32;
      float v = 0.f;
#pragma unroll
      for (int j = 0; j < 10; ++j) v = fmaf(s.dl[j], s.w3[j][f], v);
      (&s.dfeat[0][0])[f] = v;
    }
    __syncthreads();
    if (t < 200) {
      const int c = t / 100, pos = t - c * 100, oy = pos / 10, ox = pos - oy * 10;
      const int win = (oy >> 1) * 5 + (ox >> 1), k = ((oy & 1) << 1) | (ox & 1);
      const float av = s.a2[c][pos];
      const float up = s.i2[c][win] == k ? s.dfeat[c][win] : 0.f;
      s.a2[c][pos] = up * (1.f - av * av);                            // now: gradient w.r.t. conv2's pre-activation
    }
    __syncthreads();

    // ---- conv2 weight / bias gradients: thread = (co, ci, filter row), five taps each ---------------------------------
    for (int it = kV2 ? t - 224 : t; it >= 0 && it < 2 * 16 * 5; it += T) {   // (second form: warps 7..11, beside the input gradient)
      const int co = it / 80, rem = it - co * 80, ci = rem / 5, ky = rem - ci * 5;
      float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
      for (int oy = 0; oy < 10; ++oy) {
        float row[14];
#pragma unroll
        for (int xx = 0; xx < 14; ++xx) row[xx] = s.p1[ci][(oy + ky) * 14 + xx];
#pragma unroll
        for (int ox = 0; ox < 10; ++ox) {
          const float d = s.a2[co][oy * 10 + ox];
#pragma unroll
          for (int kx = 0; kx < 5; ++kx) acc[kx] = fmaf(d, row[ox + kx], acc[kx]);
        }
      }
#pragma unroll
      for (int kx = 0; kx < 5; ++kx) s.gw2[co][ci][ky * 5 + kx] += acc[kx];
    }
    if (kV2 ? t < 32 : (t >= 160 && t < 192)) {  // first form: warp 5 has no filter rows; it sums the two bias gradients, 16 lanes each
      const int lane = t & 31, co = lane >> 4;
      float v = 0.f;
      for (int i = lane & 15; i < 100; i += 16) v += s.a2[co][i];
      v = warp_sum16(v);
      if ((lane & 15) == 0) s.gb2[co] += v;
    }
    // ---- conv2 input gradient: my two output channels' contribution to ALL 16 input channels, row by row ---------
    for (int it = t; it < 16 * 14; it += T) {
      const int ci = it / 14, yy = it - ci * 14;
      float acc[14];
#pragma unroll
      for (int xx = 0; xx < 14; ++xx) acc[xx] = 0.f;
#pragma unroll
      for (int co = 0; co < 2; ++co) {
#pragma unroll
        for (int ky = 0; ky < 5; ++ky) {
          const int oy = yy - ky;
          if (oy < 0 || oy > 9) continue;
          float d[10];
#pragma unroll
          for (int ox = 0; ox < 10; ++ox) d[ox] = s.a2[co][oy * 10 + ox];
#pragma unroll
          for (int kx = 0; kx < 5; ++kx) {
            const float w = s.w2[co][ci][ky * 5 + kx];
#pragma unroll
            for (int ox = 0; ox < 10; ++ox) acc[ox + kx] = fmaf(w, d[ox], acc[ox + kx]);
          }
        }
      }
      // reduce-scatter: first form — into slot r of the channel's owner; second form — kept here, the owner fetches it
      float* dst = kV2 ? &(&s.dpart[0][0][0])[ci * 196 + yy * 14] : &peer[ci >> 1]->dpart[r][ci & 1][yy * 14];
#pragma unroll
      for (int xx = 0; xx < 14; ++xx) dst[xx] = acc[xx];
    }
    cluster.sync();   // [3] first form: my eight slots are filled; second form: every CTA's contributions are final

    // ---- conv1 backward for my two channels: sum the slots, un-pool, tanh', then the filter gradients ----------------
    for (int it = t; it < 2 * 196; it += T) {
      const int c = it / 196, pos = it - c * 196, py = pos / 14, px = pos - py * 14;
      float dp = 0.f;
#pragma unroll
      for (int k = 0; k < MN_CL; ++k)                                  // rank order in both forms: the same bits
        dp += kV2 ? (&peer[k]->dpart[0][0][0])[(2 * r + c) * 196 + pos] : s.dpart[k][c][pos];
      const int arg = s.i1[c][pos];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float* av = &s.a1[c][(2 * py + (k >> 1)) * 28 + 2 * px + (k & 1)];
        *av = (k == arg) ? dp * (1.f - *av * *av) : 0.f;              // now: gradient w.r.t. conv1's pre-activation
      }
    }
    __syncthreads();
    if (kV2) {
      // warp = (channel, filter row ky), lane = output column: lane ox adds d[oy][ox] * x[oy+ky][ox+kx] over the 28 rows into
      // five sums (one per kx) — both reads walk consecutive words across the lanes — then the lanes meet by butterfly.  The
      // bias gradient (sum of d) rides along in the two ky == 0 warps.
      const int w = t >> 5, lane = t & 31;
      if (w < 10) {
        const int c = w / 5, ky = w - c * 5;
        float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
        float bs = 0.f;
        if (lane < 28) {
#pragma unroll 4
          for (int oy = 0; oy < 28; ++oy) {
            const float d = s.a1[c][oy * 28 + lane];
            const float* xr = &s.x[(oy + ky) * 32 + lane];
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) acc[kx] = fmaf(d, xr[kx], acc[kx]);
            bs += d;
          }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
          for (int kx = 0; kx < 5; ++kx) acc[kx] += __shfl_xor_sync(0xffffffffu, acc[kx], o);
          bs += __shfl_xor_sync(0xffffffffu, bs, o);
        }
        if (lane == 0) {
#pragma unroll
          for (int kx = 0; kx < 5; ++kx) s.gw1[c][ky * 5 + kx] += acc[kx];
          if (ky == 0) s.gb1[c] += bs;
        }
      }
    } else {
      // thread = (channel, filter tap, one of 4 row groups); the groups meet in fixed order through shared scratch (lpart
      // and dfeat are free by now)
      {
        const int c = (t / 100) & 1, rem = t % 100, tap = rem >> 2, grp = rem & 3, ky = tap / 5, kx = tap - ky * 5;
        float v = 0.f;
        if (t < 200) {
          for (int oy = grp * 7; oy < grp * 7 + 7; ++oy)
            for (int ox = 0; ox < 28; ++ox) v = fmaf(s.a1[c][oy * 28 + ox], s.x[(oy + ky) * 32 + ox + kx], v);
        }
        v += __shfl_xor_sync(0xffffffffu, v, 1);     // whole warps take part (320 threads = 10 warps), idle lanes carry zeros
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        if (t < 200 && grp == 0) s.gw1[c][tap] += v;
        // bias gradients of my two channels: every thread adds its share, the ten warps meet in lpart (free until the next
        // sample's Linear layer: no peer can pass barrier [1] of the next sample before this CTA has arrived there)
        float b0 = 0.f, b1 = 0.f;
        for (int i = t; i < 784; i += T) { b0 += s.a1[0][i]; b1 += s.a1[1][i]; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          b0 += __shfl_xor_sync(0xffffffffu, b0, o);
          b1 += __shfl_xor_sync(0xffffffffu, b1, o);
        }
        float* scratch = &s.lpart[0][0];
        if ((t & 31) == 0) { scratch[(t >> 5) * 2] = b0; scratch[(t >> 5) * 2 + 1] = b1; }
      }
      __syncthreads();
      if (t < 2) {
        const float* scratch = &s.lpart[0][0];
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < MN_T / 32; ++w) v += scratch[w * 2 + t];
        s.gb1[t] += v;
      }
    }
    // No barrier here: the next sample's remote writes are ordered behind this sample's reads by the barriers themselves
    // (p1 is last read before [3]; lpart right after [2]; dpart before a CTA reaches the next [1], and nobody writes
    // dpart again before everybody has passed that [1] and the [2] after it).
    __syncthreads();   // s.x and a1 are rewritten by the next sample
  }

  // ---- gradients out (or the update itself) ----------------------------------------------------------------------
  __syncthreads();
  const bool fused = a.apply_sgd != 0 && ncl == 1;
  auto emit = [&](float* p, float* g, float grad, float mine) {   // `mine`: this CTA's shared-memory copy of *p
    if (fused) {
      *p = fmaf(-a.lr, grad, kV2 ? mine : *p);   // same rounding as the flat-vector SGD kernels (csrc/fused_sgd.cu)
    } else if (ncl == 1) {
      *g += grad;                        // accumulate into .grad like autograd would
    } else {
      atomicAdd(g, grad);
    }
  };
  for (int i = t; i < 50; i += T) emit(a.p[0] + 2 * r * 25 + i, a.g[0] + 2 * r * 25 + i, (&s.gw1[0][0])[i], (&s.w1[0][0])[i]);
  for (int i = t; i < 800; i += T) emit(a.p[2] + 2 * r * 400 + i, a.g[2] + 2 * r * 400 + i, (&s.gw2[0][0][0])[i], (&s.w2[0][0][0])[i]);
  for (int i = t; i < 500; i += T) {
    const int j = i / 50, f = i - j * 50, at = j * 400 + 2 * r * 25 + f;
    emit(a.p[4] + at, a.g[4] + at, s.gw3[j][f], s.w3[j][f]);
  }
  if (t < 2) {
    emit(a.p[1] + 2 * r + t, a.g[1] + 2 * r + t, s.gb1[t], s.b1[t]);
    emit(a.p[3] + 2 * r + t, a.g[3] + 2 * r + t, s.gb2[t], s.b2[t]);
  }
  if (r == 0 && t < 10) emit(a.p[5] + t, a.g[5] + t, s.gb3[t], s.b3[t]);
  if (r == 0 && t == 0) {
    if (ncl == 1) *a.loss = s.loss_acc * inv_b;
    else atomicAdd(a.loss, s.loss_acc * inv_b);
  }
  cluster.sync();   // no CTA may leave while a peer could still touch its shared memory
}

// Which form of the kernel runs: 2 (default) or 1 — DL_B200_MNIST_FORM at load time, dl_mnist_form() afterwards (returns
// the previous value; anything but 1 or 2 only reads it).
static int mnist_form_from_env() {
  const char* e = getenv("DL_B200_MNIST_FORM");
  return (e != nullptr && e[0] == '1') ? 1 : 2;
}
static int g_mnist_form = mnist_form_from_env();
extern "C" int dl_mnist_form(int form) {
  const int before = g_mnist_form;
  if (form == 1 || form == 2) g_mnist_form = form;
  return before;
}

// x [B,1024] fp32, y [B] int64; params / grads: the six tensors of the model in module order.  apply_sgd != 0 (and a
// batch small enough for one cluster, B <= DL_MNIST_FUSED_MAX_BATCH) fuses the update p -= lr * g into the launch and
// leaves grads untouched; otherwise the gradient of the MEAN loss is ADDED into grads.  loss: 1 float (overwritten).
extern "C" int dl_mnist_step(const float* x, const int64_t* y, float* const* params, float* const* grads, float* loss, float* logp,
                             int B, float lr, int apply_sgd, void* stream) {
  DL_REQUIRE(x && y && params && grads && loss && B > 0, "dl_mnist_step: bad arguments");
  DL_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0, "dl_mnist_step: x must be 16-byte aligned");
  MnistArgs a;
  a.x = x; a.y = reinterpret_cast<const long long*>(y); a.loss = loss; a.logp = logp; a.B = B; a.lr = lr;
  for (int i = 0; i < 6; ++i) {
    DL_REQUIRE(params[i] && grads[i], "dl_mnist_step: null parameter / gradient pointer");
    a.p[i] = params[i];
    a.g[i] = grads[i];
  }
  int ncl = B <= DL_MNIST_FUSED_MAX_BATCH ? 1 : (B + 3) / 4;       // a cluster walks ~4 samples when the batch is larger
  const int cap = dl_sm_count() / MN_CL;
  if (ncl > cap) ncl = cap;
  a.apply_sgd = (apply_sgd && ncl == 1) ? 1 : 0;
  DL_REQUIRE(!apply_sgd || ncl == 1, "dl_mnist_step: the fused update needs B <= %d (one cluster)", DL_MNIST_FUSED_MAX_BATCH);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  static bool configured = false;
  if (!configured) {
    DL_CUDA(cudaFuncSetAttribute(mnist_step_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sizeof(MnistSmem))));
    DL_CUDA(cudaFuncSetAttribute(mnist_step_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(sizeof(MnistSmem))));
    configured = true;
  }
  if (ncl > 1) DL_CUDA(cudaMemsetAsync(loss, 0, sizeof(float), st));
  if (g_mnist_form == 1) mnist_step_kernel<false><<<ncl * MN_CL, MN_T, sizeof(MnistSmem), st>>>(a);
  else mnist_step_kernel<true><<<ncl * MN_CL, MN_T2, sizeof(MnistSmem), st>>>(a);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}
