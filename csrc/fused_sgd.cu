// fused_allreduce_normalize_sgd — SURVEY K2/K3/W5.
//
// One launch replaces  tree.allReduce(grads) ; grad:mul(1/n) ; p:add(-lr, g)
// (lua/AllReduceSGD.lua:18-30 + examples/cifar10.lua:187-191):
//
//   arrival barrier  (publishes `active` + step count; yields n_active)
//   two-shot : the owner of a slice loads that slice of every ACTIVE peer's gradient over
//              NVLink (or one multimem.ld_reduce through the switch), sums in rank order,
//              scales by 1/n_active, applies weight decay / momentum / SGD to its parameter
//              slice and pushes the new slice into every rank's parameter buffer
//              (P2P stores or one multimem.st)
//   one-shot : every rank loads the whole gradient of every active peer and updates its own
//              parameters redundantly (bit-identical: same order, same inputs)
//   departure barrier, then the local gradient buffer is zeroed (fused zero_grad).
//
// Parameters stay bit-identical on all ranks because only the owner computes a slice.
#include "dl_common.cuh"
#include "dl_host.h"
#include "dl_pipe.cuh"

struct SgdArgs {
  uint64_t g_off, p_off, m_off;
  unsigned long long n4;
  float lr, momentum, wd;
  int nesterov;
  uint32_t active;  // real: 0/1, emulation: bit r = rank r active
  uint32_t payload;
  int zero_grads;
  int channel;
};

// The update rule with its rounding pinned down (explicit fmaf, no compiler-chosen contraction): every variant of
// the kernel — one-shot, two-shot, NVLS, any bucket geometry — then produces the same bits from the same sums.
__device__ __forceinline__ float sgd_update1(float g, float p, float* m, const SgdArgs& a) {
  if (a.wd != 0.f) g = fmaf(a.wd, p, g);
  if (m != nullptr) {
    const float mv = fmaf(a.momentum, *m, g);
    *m = mv;
    g = a.nesterov ? fmaf(a.momentum, mv, g) : mv;
  }
  return fmaf(-a.lr, g, p);
}
__device__ __forceinline__ float4 sgd_update(float4 g, float4 p, float4* m_ptr, const SgdArgs& a) {
  float* m = reinterpret_cast<float*>(m_ptr);
  float4 r;
  r.x = sgd_update1(g.x, p.x, m ? m + 0 : nullptr, a);
  r.y = sgd_update1(g.y, p.y, m ? m + 1 : nullptr, a);
  r.z = sgd_update1(g.z, p.z, m ? m + 2 : nullptr, a);
  r.w = sgd_update1(g.w, p.w, m ? m + 3 : nullptr, a);
  return r;
}

// One pass over the owner's slice [s.begin, s.end): reduce the gradients of the active ranks
// (kMc: one multimem.ld_reduce through the NVSwitch; else P2P loads summed in rank order), update
// the owner's parameters and distribute them (kMc: one multimem.st; else P2P stores).
// U = 16-byte vectors a thread keeps in flight per source.
template <int W, int U, bool kMc>
__device__ __forceinline__ void sgd_slice_pass(const dl_comm_t& c, const SgdArgs& a, const DlSlice& s, int rank, int world,
                                               uint32_t mask, float inv, unsigned long long tid0,
                                               unsigned long long stride) {
  float4* p_loc = dl_at<float4>(c, rank, a.p_off);
  float4* m_loc = a.m_off ? dl_at<float4>(c, rank, a.m_off) : nullptr;
  for (unsigned long long i0 = s.begin + tid0; i0 < s.end; i0 += stride * U) {
    float4 acc[U];
    if (kMc) {
      const float4* mg = dl_mc_at<float4>(c, a.g_off);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const unsigned long long idx = i0 + u * stride;
        acc[u] = idx < s.end ? dl_mc_ld_reduce_f4(mg + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    } else {
      float4 v[U][W];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const unsigned long long idx = i0 + u * stride;
#pragma unroll
        for (int j = 0; j < W; ++j) {
          v[u][j] = (idx < s.end && j < world && ((mask >> j) & 1u))
                        ? dl_ld_f4(dl_at<float4>(c, j, a.g_off) + idx)
                        : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        acc[u] = v[u][0];
#pragma unroll
        for (int j = 1; j < W; ++j) acc[u] = acc[u] + v[u][j];
      }
    }
    float4 pv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned long long idx = i0 + u * stride;
      if (idx < s.end) pv[u] = p_loc[idx];
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned long long idx = i0 + u * stride;
      if (idx < s.end) {
        const float4 pn = sgd_update(dl_scale(acc[u], inv), pv[u], m_loc ? m_loc + idx : nullptr, a);
        if (kMc) {
          dl_mc_st_f4(dl_mc_at<float4>(c, a.p_off) + idx, pn);
        } else {
#pragma unroll
          for (int j = 0; j < W; ++j)
            if (j < world) dl_st_f4(dl_at<float4>(c, j, a.p_off) + idx, pn);
        }
      }
    }
  }
}

#define DL_NVLS_U 8  // multimem.ld_reduce round trips through the switch are long: keep 128 B per thread in flight

template <int kWorld, bool kNvls, bool kEmu>
__global__ void __launch_bounds__(DL_THREADS, 1) sgd_twoshot_kernel(dl_comm_t c, SgdArgs a) {
  __shared__ uint32_t pay[DL_MAX_RANKS];
  constexpr int W = kWorld ? kWorld : DL_MAX_RANKS;
  constexpr int U = kWorld == 1 ? 8 : kWorld == 2 ? 8 : kWorld == 4 ? 4 : 2;
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = kWorld ? kWorld : c.world;
  const int act = kEmu ? ((a.active >> rank) & 1u) : (a.active & 3u);
  const uint32_t user = kEmu ? a.payload + rank : a.payload;

  __shared__ uint32_t auxs[DL_MAX_RANKS];
  const bool ok = dl_block_barrier<false>(c, rank, world, a.channel, dl_pack_payload(act, user), pay, __float_as_uint(a.lr), auxs);
  int n_act;
  const uint32_t mask = dl_active_mask(c, rank, world, a.channel, pay, &n_act);
  // the learning rate of the lowest active rank wins (identical on all ranks, so owners agree)
  a.lr = __uint_as_float(dl_aux_of_first_active(mask, auxs, rank));
  const DlSlice s = dl_slice(a.n4, rank, world);
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
  const unsigned long long tid0 = static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x;

  if (ok && n_act > 0) {
    const float inv = n_act > 1 ? 1.f / static_cast<float>(n_act) : 1.f;
    if (kNvls && n_act == world) {
      sgd_slice_pass<W, DL_NVLS_U, true>(c, a, s, rank, world, mask, inv, tid0, stride);
    } else if (kNvls) {
      // some ranks are draining: the switch would add their stale gradients too, so this (rare)
      // round reads the active ranks over P2P and pushes the parameters with P2P stores
      sgd_slice_pass<W, 1, false>(c, a, s, rank, world, mask, inv, tid0, stride);
    } else {
      sgd_slice_pass<W, U, false>(c, a, s, rank, world, mask, inv, tid0, stride);
    }
  }
  const bool ok2 = dl_block_barrier(c, rank, world, a.channel, 0u, pay);
  if (ok && ok2 && a.zero_grads) {
    // Block b of every peer has finished reading offsets {b-pattern} of its slice of OUR gradient.
    float4* g_loc = dl_at<float4>(c, rank, a.g_off);
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (unsigned long long k = tid0; k < s.slice4; k += stride) {
#pragma unroll
      for (int j = 0; j < W; ++j) {
        const unsigned long long idx = s.slice4 * j + k;
        if (j < world && idx < a.n4) g_loc[idx] = z;
      }
    }
  }
}

// ---- DL_VARIANT_HYBRID: software-pipelined slice pass, switch + P2P in one launch (dl_pipe.cuh) --------
template <int W>
struct SgdEpi {
  const SgdArgs& a;
  float4* p_loc;
  float4* m_loc;
  float inv;
  __device__ __forceinline__ float4 upd(const float4& sum, unsigned long long idx) {
    return sgd_update(dl_scale(sum, inv), p_loc[idx], m_loc ? m_loc + idx : nullptr, a);
  }
  __device__ __forceinline__ float4 p2p(const float4 (&v)[W], unsigned long long idx) {
    float4 acc = v[0];
#pragma unroll
    for (int j = 1; j < W; ++j) acc = acc + v[j];  // rank order, zeros for ranks that sit this round out
    return upd(acc, idx);
  }
  __device__ __forceinline__ float4 mc(const float4& sum, unsigned long long idx) { return upd(sum, idx); }
};

template <int kWorld, bool kEmu>
__global__ void __launch_bounds__(DL_PIPE_THREADS, 1) sgd_pipe_kernel(dl_comm_t c, SgdArgs a) {
  __shared__ uint32_t pay[DL_MAX_RANKS];
  __shared__ uint32_t auxs[DL_MAX_RANKS];
  constexpr int W = kWorld ? kWorld : DL_MAX_RANKS;
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = kWorld ? kWorld : c.world;
  const int act = kEmu ? ((a.active >> rank) & 1u) : (a.active & 3u);
  const uint32_t user = kEmu ? a.payload + rank : a.payload;
  const bool ok = dl_block_barrier<false>(c, rank, world, a.channel, dl_pack_payload(act, user), pay, __float_as_uint(a.lr), auxs);
  int n_act;
  const uint32_t mask = dl_active_mask(c, rank, world, a.channel, pay, &n_act);
  a.lr = __uint_as_float(dl_aux_of_first_active(mask, auxs, rank));
  const DlSlice s = dl_slice_even(a.n4, rank, world);
  const bool full = n_act == world;
  if (ok && n_act > 0) {
    SgdEpi<W> epi{a, dl_at<float4>(c, rank, a.p_off), a.m_off ? dl_at<float4>(c, rank, a.m_off) : nullptr,
                  n_act > 1 ? 1.f / static_cast<float>(n_act) : 1.f};
    dl_pipe_pass<W>(c, rank, world, mask, full, c.nvls_permille, a.g_off, a.p_off, 0, s, epi);
  }
  const bool ok2 = dl_block_barrier(c, rank, world, a.channel, 0u, pay);
  if (ok && ok2 && a.zero_grads) {
    float4* g_loc = dl_at<float4>(c, rank, a.g_off);
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    // a round nobody contributed to ran no data phase: its geometry is the all-P2P one
    dl_pipe_foreach_released(a.n4, world, full, c.mc != nullptr, c.nvls_permille, s, [&](unsigned long long k) { g_loc[k] = z; });
  }
}

// ---- DL_VARIANT_LIGHT: the per-bucket kernel that runs beside the convolutions (dl_pipe.cuh, "light slice pass") ----
template <int kWorld, bool kEmu>
__global__ void __launch_bounds__(DL_LIGHT_THREADS, 6) sgd_light_kernel(dl_comm_t c, SgdArgs a) {
  __shared__ uint32_t pay[DL_MAX_RANKS];
  __shared__ uint32_t auxs[DL_MAX_RANKS];
  constexpr int W = kWorld ? kWorld : DL_MAX_RANKS;
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = kWorld ? kWorld : c.world;
  const int act = kEmu ? ((a.active >> rank) & 1u) : (a.active & 3u);
  const uint32_t user = kEmu ? a.payload + rank : a.payload;
  const bool ok = dl_block_barrier<false>(c, rank, world, a.channel, dl_pack_payload(act, user), pay, __float_as_uint(a.lr), auxs);
  int n_act;
  const uint32_t mask = dl_active_mask(c, rank, world, a.channel, pay, &n_act);
  a.lr = __uint_as_float(dl_aux_of_first_active(mask, auxs, rank));
  const DlSlice s = dl_slice_even(a.n4, rank, world);
  const bool full = n_act == world;
  if (ok && n_act > 0) {
    SgdEpi<W> epi{a, dl_at<float4>(c, rank, a.p_off), a.m_off ? dl_at<float4>(c, rank, a.m_off) : nullptr,
                  n_act > 1 ? 1.f / static_cast<float>(n_act) : 1.f};
    dl_light_pass<W>(c, world, mask, full, c.nvls_permille, a.g_off, a.p_off, s, epi);
  }
  const bool ok2 = dl_block_barrier(c, rank, world, a.channel, 0u, pay);
  if (ok && ok2 && a.zero_grads) {
    float4* g_loc = dl_at<float4>(c, rank, a.g_off);
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    dl_light_foreach_released(a.n4, world, full, c.mc != nullptr, c.nvls_permille, s, [&](unsigned long long k) { g_loc[k] = z; });
  }
}

template <int kWorld, bool kEmu>
__global__ void __launch_bounds__(DL_THREADS, 1) sgd_oneshot_kernel(dl_comm_t c, SgdArgs a) {
  __shared__ uint32_t pay[DL_MAX_RANKS];
  constexpr int W = kWorld ? kWorld : DL_MAX_RANKS;
  constexpr int U = kWorld == 2 ? 4 : 2;
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = kWorld ? kWorld : c.world;
  const int act = kEmu ? ((a.active >> rank) & 1u) : (a.active & 3u);
  const uint32_t user = kEmu ? a.payload + rank : a.payload;

  __shared__ uint32_t auxs[DL_MAX_RANKS];
  const bool ok = dl_block_barrier<false>(c, rank, world, a.channel, dl_pack_payload(act, user), pay, __float_as_uint(a.lr), auxs);
  int n_act;
  const uint32_t mask = dl_active_mask(c, rank, world, a.channel, pay, &n_act);
  // the learning rate of the lowest active rank wins (identical on all ranks, so owners agree)
  a.lr = __uint_as_float(dl_aux_of_first_active(mask, auxs, rank));
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
  const unsigned long long tid0 = static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x;

  if (ok && n_act > 0) {
    const float inv = n_act > 1 ? 1.f / static_cast<float>(n_act) : 1.f;
    float4* p_loc = dl_at<float4>(c, rank, a.p_off);
    float4* m_loc = a.m_off ? dl_at<float4>(c, rank, a.m_off) : nullptr;
    for (unsigned long long i0 = tid0; i0 < a.n4; i0 += stride * U) {
      float4 v[U][W];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const unsigned long long idx = i0 + u * stride;
#pragma unroll
        for (int j = 0; j < W; ++j) {
          v[u][j] = (idx < a.n4 && j < world && ((mask >> j) & 1u))
                        ? dl_ld_f4(dl_at<float4>(c, j, a.g_off) + idx)
                        : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const unsigned long long idx = i0 + u * stride;
        float4 acc = v[u][0];
#pragma unroll
        for (int j = 1; j < W; ++j) acc = acc + v[u][j];
        if (idx < a.n4)
          p_loc[idx] = sgd_update(dl_scale(acc, inv), p_loc[idx], m_loc ? m_loc + idx : nullptr, a);
      }
    }
  }
  const bool ok2 = dl_block_barrier(c, rank, world, a.channel, 0u, pay);
  if (ok && ok2 && a.zero_grads) {
    float4* g_loc = dl_at<float4>(c, rank, a.g_off);
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (unsigned long long k = tid0; k < a.n4; k += stride) g_loc[k] = z;
  }
}

// ---- world == 1 and helper ops: no communication ---------------------------------------------
struct LocalSgdArgs {
  float* p; const float* p_in; float* g; float* m;
  unsigned long long n;
  float lr, momentum, wd, scale;
  int nesterov, zero_grads;
};

__global__ void __launch_bounds__(DL_THREADS) local_sgd_kernel(LocalSgdArgs a) {
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
  const unsigned long long n4 = a.n / 4;
  SgdArgs s{};
  s.lr = a.lr; s.momentum = a.momentum; s.wd = a.wd; s.nesterov = a.nesterov;
  float4* p4 = reinterpret_cast<float4*>(a.p);
  const float4* pi4 = reinterpret_cast<const float4*>(a.p_in);
  float4* g4 = reinterpret_cast<float4*>(a.g);
  float4* m4 = reinterpret_cast<float4*>(a.m);
  for (unsigned long long i = static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x; i < n4;
       i += stride) {
    const float4 g = dl_scale(g4[i], a.scale);
    p4[i] = sgd_update(g, pi4[i], m4 ? m4 + i : nullptr, s);
    if (a.zero_grads) g4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  // scalar tail (n not a multiple of 4)
  for (unsigned long long i = n4 * 4 + static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x;
       i < a.n; i += stride) {
    a.p[i] = sgd_update1(__fmul_rn(a.g[i], a.scale), a.p_in[i], a.m ? a.m + i : nullptr, s);
    if (a.zero_grads) a.g[i] = 0.f;
  }
}

#define DL_SGD_DISPATCH_WORLD(KERNEL, ...)                                  \
  switch (c->world) {                                                       \
    case 1: rc = dl_launch(KERNEL<1, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
    case 2: rc = dl_launch(KERNEL<2, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
    case 4: rc = dl_launch(KERNEL<4, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
    case 8: rc = dl_launch(KERNEL<8, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
    default: rc = dl_launch(KERNEL<0, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
  }

extern "C" int dl_fused_sgd(const dl_comm_t* c, uint64_t g_off, uint64_t p_off, uint64_t m_off,
                            uint64_t n, float lr, float momentum, float weight_decay, int nesterov,
                            int variant, int active, uint32_t payload, int zero_grads, int nblocks,
                            int channel, void* stream) {
  DL_REQUIRE(c != nullptr, "dl_fused_sgd: null communicator");
  DL_REQUIRE(n % 4 == 0, "dl_fused_sgd: n (%llu) must be a multiple of 4 floats", (unsigned long long)n);
  DL_REQUIRE(g_off % 16 == 0 && p_off % 16 == 0 && m_off % 16 == 0, "dl_fused_sgd: offsets must be 16-byte aligned");
  DL_REQUIRE(g_off >= DL_PAD_BYTES && p_off >= DL_PAD_BYTES, "dl_fused_sgd: offsets overlap the signal pad");
  DL_REQUIRE(payload < (1u << 30), "dl_fused_sgd: payload must fit 30 bits");
  if (variant == DL_VARIANT_NVLS) DL_REQUIRE(c->mc && !c->emulate, "dl_fused_sgd: NVLS variant needs a multicast mapping");
  SgdArgs a;
  a.g_off = g_off; a.p_off = p_off; a.m_off = m_off; a.n4 = n / 4;
  a.lr = lr; a.momentum = momentum; a.wd = weight_decay; a.nesterov = nesterov;
  a.active = static_cast<uint32_t>(active); a.payload = payload; a.zero_grads = zero_grads; a.channel = channel;
  if (variant == DL_VARIANT_LL)
    return dl_ll_sgd(c, g_off, p_off, m_off, n, lr, momentum, weight_decay, nesterov, active, payload, zero_grads, nblocks, channel,
                     static_cast<cudaStream_t>(stream));
  const unsigned long long work = variant == DL_VARIANT_ONESHOT ? a.n4 : (a.n4 + c->world - 1) / (c->world > 0 ? c->world : 1);
  int nb = dl_check_comm(c, channel, dl_pick_blocks(work, 2, nblocks));
  if (nb < 0) return nb;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = DL_OK;
  if (variant == DL_VARIANT_LIGHT) {
    DL_REQUIRE(n % 8 == 0, "dl_fused_sgd: the light variant needs n %% 8 == 0 (even slices)");
    DL_REQUIRE(c->nvls_permille >= 0 && c->nvls_permille <= 1000, "dl_fused_sgd: nvls_permille out of range");
    DL_REQUIRE(c->nvls_permille == 0 || (c->mc && !c->emulate), "dl_fused_sgd: nvls_permille > 0 needs a multicast mapping");
    // one small CTA per SM at most: they move in beside the convolution's CTAs; ~4 vectors per thread at least
    nb = dl_check_comm(c, channel, dl_pick_blocks(work / 4 + 1, 1, nblocks));
    if (nb < 0) return nb;
#define DL_SGD_LIGHT(WORLD)                                                                                     \
    rc = c->emulate ? dl_launch(sgd_light_kernel<WORLD, true>, c, nb, DL_LIGHT_THREADS, st, *c, a)                \
                    : dl_launch(sgd_light_kernel<WORLD, false>, c, nb, DL_LIGHT_THREADS, st, *c, a)
    switch (c->world) {
      case 1: DL_SGD_LIGHT(1); break;
      case 2: DL_SGD_LIGHT(2); break;
      case 4: DL_SGD_LIGHT(4); break;
      case 8: DL_SGD_LIGHT(8); break;
      default: DL_SGD_LIGHT(0); break;
    }
#undef DL_SGD_LIGHT
    return rc;
  }
  if (variant == DL_VARIANT_HYBRID) {
    DL_REQUIRE(n % 8 == 0 && g_off % 32 == 0 && p_off % 32 == 0 && m_off % 32 == 0,
               "dl_fused_sgd: the hybrid variant moves 32 bytes at a time: n %% 8 == 0 and 32-byte aligned offsets");
    DL_REQUIRE(c->nvls_permille >= 0 && c->nvls_permille <= 1000, "dl_fused_sgd: nvls_permille out of range");
    DL_REQUIRE(c->nvls_permille == 0 || (c->mc && !c->emulate), "dl_fused_sgd: nvls_permille > 0 needs a multicast mapping");
    // 256 threads per CTA, one 32-byte item per thread and stage: enough CTAs for ~2 items per thread, all SMs at most
    nb = dl_check_comm(c, channel, dl_pick_blocks(work / 2 + 1, 2 * DL_PIPE_THREADS / DL_THREADS, nblocks));
    if (nb < 0) return nb;
#define DL_SGD_PIPE(WORLD)                                                                                      \
    rc = c->emulate ? dl_launch(sgd_pipe_kernel<WORLD, true>, c, nb, DL_PIPE_THREADS, st, *c, a)                  \
                    : dl_launch(sgd_pipe_kernel<WORLD, false>, c, nb, DL_PIPE_THREADS, st, *c, a)
    switch (c->world) {
      case 1: DL_SGD_PIPE(1); break;
      case 2: DL_SGD_PIPE(2); break;
      case 4: DL_SGD_PIPE(4); break;
      case 8: DL_SGD_PIPE(8); break;
      default: DL_SGD_PIPE(0); break;
    }
#undef DL_SGD_PIPE
    return rc;
  }
  if (variant == DL_VARIANT_ONESHOT) {
    if (c->emulate) { DL_SGD_DISPATCH_WORLD(sgd_oneshot_kernel, true) } else { DL_SGD_DISPATCH_WORLD(sgd_oneshot_kernel, false) }
  } else if (variant == DL_VARIANT_NVLS) {
    DL_SGD_DISPATCH_WORLD(sgd_twoshot_kernel, true, false)
  } else if (variant == DL_VARIANT_TWOSHOT) {
    if (c->emulate) { DL_SGD_DISPATCH_WORLD(sgd_twoshot_kernel, false, true) } else { DL_SGD_DISPATCH_WORLD(sgd_twoshot_kernel, false, false) }
  } else {
    dl_set_error("dl_fused_sgd: unknown variant %d", variant);
    return DL_EINVAL;
  }
  return rc;
}

extern "C" int dl_local_sgd(float* p, float* g, float* m, uint64_t n, float lr, float momentum,
                            float weight_decay, int nesterov, float grad_scale, int zero_grads,
                            void* stream) {
  return dl_local_sgd_from(p, p, g, m, n, lr, momentum, weight_decay, nesterov, grad_scale, zero_grads, stream);
}

extern "C" int dl_local_sgd_from(float* p, const float* p_in, float* g, float* m, uint64_t n, float lr, float momentum,
                                 float weight_decay, int nesterov, float grad_scale, int zero_grads, void* stream) {
  DL_REQUIRE(p && g && p_in, "dl_local_sgd: null pointer");
  DL_REQUIRE((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(p_in) | reinterpret_cast<uintptr_t>(g) |
              reinterpret_cast<uintptr_t>(m)) % 16 == 0,
             "dl_local_sgd: pointers must be 16-byte aligned");
  if (n == 0) return DL_OK;
  LocalSgdArgs a;
  a.p = p; a.p_in = p_in; a.g = g; a.m = m; a.n = n; a.lr = lr; a.momentum = momentum; a.wd = weight_decay;
  a.scale = grad_scale; a.nesterov = nesterov; a.zero_grads = zero_grads;
  const int nb = dl_pick_blocks(n / 4 + 1, 2, 0);
  local_sgd_kernel<<<nb, DL_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(a);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}
