// fused_elastic_average — SURVEY K7/K8/W7.
//
// One launch replaces lua/AllReduceEA.lua:34-45
//     delta = (x - c) * alpha ; x -= delta ; allReduce(delta) ; c += sum(delta)
// without materialising delta:
//
//   arrival barrier
//   phase 1 : the owner of a slice loads that slice of x from EVERY rank over NVLink, its own
//             (replicated) center slice, forms  c_new = c + sum_j alpha (x_j - c)  in rank order
//             and pushes c_new into every rank's *next* center buffer (double buffered, so
//             nobody's c_old is overwritten while it is still needed)
//   barrier : every peer has finished reading our x, every c_new slice has landed
//   phase 2 : x -= alpha (x - c_old), locally over the whole vector.
//
// Center replicas stay bit-identical (owner computes, everyone receives the same bits), which
// removes the floating-point drift the reference documents at lua/AllReduceEA.lua:76.
// A round in which no rank is active (the terminating drain round) moves nothing.
#include "dl_common.cuh"
#include "dl_host.h"
#include "dl_pipe.cuh"

struct EaArgs {
  uint64_t x_off, c_off, cnew_off;
  unsigned long long n4;
  float alpha;
  uint32_t active, payload;
  int channel;
};

template <int kWorld, bool kNvls, bool kEmu>
__global__ void __launch_bounds__(DL_THREADS, 1) fused_ea_kernel(dl_comm_t c, EaArgs a) {
  __shared__ uint32_t pay[DL_MAX_RANKS];
  constexpr int W = kWorld ? kWorld : DL_MAX_RANKS;
  constexpr int U = kNvls ? 8 : kWorld == 1 ? 8 : kWorld == 2 ? 8 : kWorld == 4 ? 4 : 2;
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = kWorld ? kWorld : c.world;
  const int act = kEmu ? ((a.active >> rank) & 1u) : (a.active & 3u);
  const uint32_t user = kEmu ? a.payload + rank : a.payload;
  const bool ok = dl_block_barrier<false>(c, rank, world, a.channel, dl_pack_payload(act, user), pay);
  int n_act;
  dl_active_mask(c, rank, world, a.channel, pay, &n_act);
  const DlSlice s = dl_slice(a.n4, rank, world);
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
  const unsigned long long tid0 = static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x;
  const float alpha = a.alpha;
  const bool go = ok && n_act > 0;
  const float4* c_old = dl_at<float4>(c, rank, a.c_off);

  if (go) {
    for (unsigned long long i0 = s.begin + tid0; i0 < s.end; i0 += stride * U) {
      float4 sum[U];
      float4 cc[U];
      if (kNvls) {
        // switch-side sum of x over all ranks: sum_j a(x_j - c) = a (sum_j x_j - N c)
        const float4* mx = dl_mc_at<float4>(c, a.x_off);
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned long long idx = i0 + u * stride;
          const bool in = idx < s.end;
          const float4 sx = in ? dl_mc_ld_reduce_f4(mx + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
          cc[u] = in ? c_old[idx] : make_float4(0.f, 0.f, 0.f, 0.f);
          const float nw = static_cast<float>(world);
          sum[u] = make_float4(alpha * (sx.x - nw * cc[u].x), alpha * (sx.y - nw * cc[u].y),
                               alpha * (sx.z - nw * cc[u].z), alpha * (sx.w - nw * cc[u].w));
        }
      } else {
        float4 v[U][W];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned long long idx = i0 + u * stride;
          const bool in = idx < s.end;
          cc[u] = in ? c_old[idx] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int j = 0; j < W; ++j)
            v[u][j] = (in && j < world) ? dl_ld_f4(dl_at<float4>(c, j, a.x_off) + idx) : cc[u];
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          sum[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int j = 0; j < W; ++j) {
            if (j < world) {
              // same expression as the reference: delta_j = (x_j - c) * alpha, summed in rank order
              sum[u].x += (v[u][j].x - cc[u].x) * alpha; sum[u].y += (v[u][j].y - cc[u].y) * alpha;
              sum[u].z += (v[u][j].z - cc[u].z) * alpha; sum[u].w += (v[u][j].w - cc[u].w) * alpha;
            }
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const unsigned long long idx = i0 + u * stride;
        if (idx < s.end) {
          const float4 cn = cc[u] + sum[u];
          if (kNvls) {
            dl_mc_st_f4(dl_mc_at<float4>(c, a.cnew_off) + idx, cn);
          } else {
#pragma unroll
            for (int j = 0; j < W; ++j)
              if (j < world) dl_st_f4(dl_at<float4>(c, j, a.cnew_off) + idx, cn);
          }
        }
      }
    }
  }
  const bool ok2 = dl_block_barrier(c, rank, world, a.channel, 0u, pay);
  if (go && ok2) {
    // Block b of each peer read offsets {b-pattern} of its slice of OUR x: after the barrier those
    // are free, so this block moves exactly those elements (all slices, same pattern).
    float4* x = dl_at<float4>(c, rank, a.x_off);
    for (unsigned long long k = tid0; k < s.slice4; k += stride) {
#pragma unroll
      for (int j = 0; j < W; ++j) {
        const unsigned long long idx = s.slice4 * j + k;
        if (j < world && idx < a.n4) {
          float4 xv = x[idx];
          const float4 cv = c_old[idx];
          xv.x -= (xv.x - cv.x) * alpha; xv.y -= (xv.y - cv.y) * alpha;
          xv.z -= (xv.z - cv.z) * alpha; xv.w -= (xv.w - cv.w) * alpha;
          x[idx] = xv;
        }
      }
    }
  } else if (ok && ok2 && a.cnew_off != a.c_off) {
    // nothing moved this round: keep the double buffer coherent (c_new = c_old) so the caller can flip
    float4* cn = dl_at<float4>(c, rank, a.cnew_off);
    for (unsigned long long k = tid0; k < a.n4; k += stride) cn[k] = c_old[k];
  }
}

// ---- DL_VARIANT_HYBRID: software-pipelined slice pass, switch + P2P in one launch (dl_pipe.cuh) ------------
template <int W>
struct EaEpi {
  const float4* c_old;
  float alpha;
  int world;
  __device__ __forceinline__ float4 p2p(const float4 (&v)[W], unsigned long long idx) {
    const float4 cc = c_old[idx];
    float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int j = 0; j < W; ++j) {
      if (j < world) {  // same expression as the reference: delta_j = (x_j - c) * alpha, summed in rank order
        sum.x += (v[j].x - cc.x) * alpha; sum.y += (v[j].y - cc.y) * alpha;
        sum.z += (v[j].z - cc.z) * alpha; sum.w += (v[j].w - cc.w) * alpha;
      }
    }
    return cc + sum;
  }
  __device__ __forceinline__ float4 mc(const float4& sx, unsigned long long idx) {
    // switch-side sum of x over all ranks: sum_j a(x_j - c) = a (sum_j x_j - N c)
    const float4 cc = c_old[idx];
    const float nw = static_cast<float>(world);
    return cc + make_float4(alpha * (sx.x - nw * cc.x), alpha * (sx.y - nw * cc.y), alpha * (sx.z - nw * cc.z), alpha * (sx.w - nw * cc.w));
  }
};

template <int kWorld, bool kEmu>
__global__ void __launch_bounds__(DL_PIPE_THREADS, 1) fused_ea_pipe_kernel(dl_comm_t c, EaArgs a) {
  __shared__ uint32_t pay[DL_MAX_RANKS];
  constexpr int W = kWorld ? kWorld : DL_MAX_RANKS;
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = kWorld ? kWorld : c.world;
  const int act = kEmu ? ((a.active >> rank) & 1u) : (a.active & 3u);
  const uint32_t user = kEmu ? a.payload + rank : a.payload;
  const bool ok = dl_block_barrier<false>(c, rank, world, a.channel, dl_pack_payload(act, user), pay);
  int n_act;
  dl_active_mask(c, rank, world, a.channel, pay, &n_act);
  const DlSlice s = dl_slice_even(a.n4, rank, world);
  const bool go = ok && n_act > 0;
  const float4* c_old = dl_at<float4>(c, rank, a.c_off);
  const float alpha = a.alpha;
  if (go) {
    // every rank's x takes part whoever is "active" (a finished node still moves, lua/AllReduceEA.lua:58-68),
    // so the switch may always add all copies
    EaEpi<W> epi{c_old, alpha, world};
    dl_pipe_pass<W>(c, rank, world, (1u << world) - 1u, true, c.nvls_permille, a.x_off, a.cnew_off, 0, s, epi);
  }
  const bool ok2 = dl_block_barrier(c, rank, world, a.channel, 0u, pay);
  if (go && ok2) {
    float4* x = dl_at<float4>(c, rank, a.x_off);
    dl_pipe_foreach_released(a.n4, world, true, c.mc != nullptr, c.nvls_permille, s, [&](unsigned long long k) {
      float4 xv = x[k];
      const float4 cv = c_old[k];
      xv.x -= (xv.x - cv.x) * alpha; xv.y -= (xv.y - cv.y) * alpha;
      xv.z -= (xv.z - cv.z) * alpha; xv.w -= (xv.w - cv.w) * alpha;
      x[k] = xv;
    });
  } else if (ok && ok2) {
    float4* cn = dl_at<float4>(c, rank, a.cnew_off);
    const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_PIPE_THREADS;
    for (unsigned long long k = static_cast<unsigned long long>(blockIdx.x) * DL_PIPE_THREADS + threadIdx.x; k < a.n4; k += stride)
      cn[k] = c_old[k];
  }
}

// single-rank elastic step (world == 1 / oracle for tests):  d = a(x-c); x -= d; c += d
__global__ void __launch_bounds__(DL_THREADS) local_ea_kernel(float* x, float* c, unsigned long long n, float alpha) {
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
  for (unsigned long long i = static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x; i < n; i += stride) {
    const float d = (x[i] - c[i]) * alpha;
    x[i] -= d;
    c[i] += d;
  }
}

#define DL_EA_DISPATCH_WORLD(...)                                                             \
  switch (c->world) {                                                                         \
    case 1: rc = dl_launch(fused_ea_kernel<1, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
    case 2: rc = dl_launch(fused_ea_kernel<2, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
    case 4: rc = dl_launch(fused_ea_kernel<4, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
    case 8: rc = dl_launch(fused_ea_kernel<8, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
    default: rc = dl_launch(fused_ea_kernel<0, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
  }

extern "C" int dl_fused_ea(const dl_comm_t* c, uint64_t x_off, uint64_t c_off, uint64_t cnew_off, uint64_t n,
                           float alpha, int variant, int active, uint32_t payload, int nblocks, int channel,
                           void* stream) {
  DL_REQUIRE(c != nullptr, "dl_fused_ea: null communicator");
  DL_REQUIRE(n % 4 == 0, "dl_fused_ea: n must be a multiple of 4 floats");
  DL_REQUIRE(x_off % 16 == 0 && c_off % 16 == 0 && cnew_off % 16 == 0, "dl_fused_ea: offsets must be 16-byte aligned");
  DL_REQUIRE(x_off >= DL_PAD_BYTES && c_off >= DL_PAD_BYTES && cnew_off >= DL_PAD_BYTES, "dl_fused_ea: offsets overlap the signal pad");
  DL_REQUIRE(cnew_off != c_off, "dl_fused_ea: the next-center buffer must differ from the current one");
  DL_REQUIRE(payload < (1u << 30), "dl_fused_ea: payload must fit 30 bits");
  EaArgs a;
  a.x_off = x_off; a.c_off = c_off; a.cnew_off = cnew_off; a.n4 = n / 4; a.alpha = alpha;
  a.active = static_cast<uint32_t>(active); a.payload = payload; a.channel = channel;
  if (variant == DL_VARIANT_LL)
    return dl_ll_ea(c, x_off, c_off, cnew_off, n, alpha, active, payload, nblocks, channel, static_cast<cudaStream_t>(stream));
  int nb = dl_check_comm(c, channel, dl_pick_blocks((a.n4 + c->world - 1) / (c->world > 0 ? c->world : 1), 2, nblocks));
  if (nb < 0) return nb;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = DL_OK;
  if (variant == DL_VARIANT_HYBRID) {
    DL_REQUIRE(n % 8 == 0 && x_off % 32 == 0 && c_off % 32 == 0 && cnew_off % 32 == 0,
               "dl_fused_ea: the hybrid variant needs n %% 8 == 0 and 32-byte aligned offsets");
    DL_REQUIRE(c->nvls_permille >= 0 && c->nvls_permille <= 1000, "dl_fused_ea: nvls_permille out of range");
    DL_REQUIRE(c->nvls_permille == 0 || (c->mc && !c->emulate), "dl_fused_ea: nvls_permille > 0 needs a multicast mapping");
    nb = dl_check_comm(c, channel, dl_pick_blocks((a.n4 + c->world - 1) / (c->world > 0 ? c->world : 1) / 2 + 1, 1, nblocks));
    if (nb < 0) return nb;
#define DL_EA_PIPE(WORLD)                                                                                        \
    rc = c->emulate ? dl_launch(fused_ea_pipe_kernel<WORLD, true>, c, nb, DL_PIPE_THREADS, st, *c, a)            \
                    : dl_launch(fused_ea_pipe_kernel<WORLD, false>, c, nb, DL_PIPE_THREADS, st, *c, a)
    switch (c->world) {
      case 1: DL_EA_PIPE(1); break;
      case 2: DL_EA_PIPE(2); break;
      case 4: DL_EA_PIPE(4); break;
      case 8: DL_EA_PIPE(8); break;
      default: DL_EA_PIPE(0); break;
    }
#undef DL_EA_PIPE
    return rc;
  }
  if (variant == DL_VARIANT_NVLS) {
    DL_REQUIRE(c->mc && !c->emulate, "dl_fused_ea: NVLS variant needs a multicast mapping");
    DL_EA_DISPATCH_WORLD(true, false)
  } else {
    if (c->emulate) { DL_EA_DISPATCH_WORLD(false, true) } else { DL_EA_DISPATCH_WORLD(false, false) }
  }
  return rc;
}

extern "C" int dl_local_ea(float* x, float* c, uint64_t n, float alpha, void* stream) {
  DL_REQUIRE(x && c, "dl_local_ea: null pointer");
  if (n == 0) return DL_OK;
  local_ea_kernel<<<dl_pick_blocks(n, 4, 0), DL_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(x, c, n, alpha);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}
