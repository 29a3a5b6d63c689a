// In-place all-reduce(sum) over the symmetric heap — SURVEY K1/K9/W4.
// Replaces tree.allReduce(value, add) (lua/AllReduceSGD.lua:12,20,39; examples/mnist.lua:122).
//
//   one-shot : every rank loads the vector from all (active) peers, sums in rank order into
//              registers, waits until everybody has finished reading, then overwrites its own
//              copy.  Bit-identical on all ranks; used for small messages and for the integer
//              confusion-matrix reduce.
//   two-shot : slice owner reduces over P2P loads and pushes the result slice to all ranks.
//   NVLS     : multimem.ld_reduce + multimem.st (the NVSwitch adds and replicates).
#include "dl_common.cuh"
#include "dl_host.h"
#include "dl_pipe.cuh"

struct ArArgs {
  uint64_t off;
  unsigned long long nvec;  // number of 16-byte vectors
  uint32_t active, payload;
  int scale, channel;
};

template <int kDtype> struct VecOps;
template <> struct VecOps<DL_F32> {
  __device__ static uint4 add(uint4 a, uint4 b) {
    return make_uint4(__float_as_uint(__uint_as_float(a.x) + __uint_as_float(b.x)), __float_as_uint(__uint_as_float(a.y) + __uint_as_float(b.y)),
                      __float_as_uint(__uint_as_float(a.z) + __uint_as_float(b.z)), __float_as_uint(__uint_as_float(a.w) + __uint_as_float(b.w)));
  }
  __device__ static uint4 scale(uint4 a, float s) {
    return make_uint4(__float_as_uint(__uint_as_float(a.x) * s), __float_as_uint(__uint_as_float(a.y) * s),
                      __float_as_uint(__uint_as_float(a.z) * s), __float_as_uint(__uint_as_float(a.w) * s));
  }
};
template <> struct VecOps<DL_I32> {
  __device__ static uint4 add(uint4 a, uint4 b) { return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
  __device__ static uint4 scale(uint4 a, float) { return a; }
};
template <> struct VecOps<DL_I64> {
  __device__ static uint4 add(uint4 a, uint4 b) {
    unsigned long long a0 = (static_cast<unsigned long long>(a.y) << 32) | a.x, a1 = (static_cast<unsigned long long>(a.w) << 32) | a.z;
    unsigned long long b0 = (static_cast<unsigned long long>(b.y) << 32) | b.x, b1 = (static_cast<unsigned long long>(b.w) << 32) | b.z;
    a0 += b0; a1 += b1;
    return make_uint4(static_cast<unsigned>(a0), static_cast<unsigned>(a0 >> 32), static_cast<unsigned>(a1), static_cast<unsigned>(a1 >> 32));
  }
  __device__ static uint4 scale(uint4 a, float) { return a; }
};
template <> struct VecOps<DL_F64> {
  __device__ static uint4 add(uint4 a, uint4 b) {
    double a0 = __hiloint2double(a.y, a.x), a1 = __hiloint2double(a.w, a.z);
    double b0 = __hiloint2double(b.y, b.x), b1 = __hiloint2double(b.w, b.z);
    a0 += b0; a1 += b1;
    return make_uint4(__double2loint(a0), __double2hiint(a0), __double2loint(a1), __double2hiint(a1));
  }
  __device__ static uint4 scale(uint4 a, float s) {
    double a0 = __hiloint2double(a.y, a.x) * s, a1 = __hiloint2double(a.w, a.z) * s;
    return make_uint4(__double2loint(a0), __double2hiint(a0), __double2loint(a1), __double2hiint(a1));
  }
};

#define DL_ONESHOT_U 4  // vectors a thread keeps in registers across the middle barrier

template <int kDtype, bool kEmu>
__global__ void __launch_bounds__(DL_THREADS, 1) allreduce_oneshot_kernel(dl_comm_t c, ArArgs a) {
  __shared__ uint32_t pay[DL_MAX_RANKS];
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = c.world;
  const int act = kEmu ? ((a.active >> rank) & 1u) : (a.active & 3u);
  const uint32_t user = kEmu ? a.payload + rank : a.payload;
  __shared__ uint32_t auxs[DL_MAX_RANKS];
  const bool ok = dl_block_barrier<false>(c, rank, world, a.channel, dl_pack_payload(act, user), pay, a.scale != 0 ? 1u : 0u, auxs);
  int n_act;
  const uint32_t mask = dl_active_mask(c, rank, world, a.channel, pay, &n_act);
  // whether to normalise is dictated by the lowest active rank (draining ranks just follow)
  a.scale = static_cast<int>(dl_aux_of_first_active(mask, auxs, rank) & 1u);
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
  const unsigned long long tid0 = static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x;

  uint4 acc[DL_ONESHOT_U];
#pragma unroll
  for (int u = 0; u < DL_ONESHOT_U; ++u) {
    const unsigned long long idx = tid0 + u * stride;
    acc[u] = make_uint4(0, 0, 0, 0);
    if (ok && idx < a.nvec) {
      uint4 v[DL_MAX_RANKS];
#pragma unroll
      for (int j = 0; j < DL_MAX_RANKS; ++j)
        v[j] = (j < world && ((mask >> j) & 1u)) ? dl_ld_u4(dl_at<uint4>(c, j, a.off) + idx) : make_uint4(0, 0, 0, 0);
      acc[u] = v[0];
#pragma unroll
      for (int j = 1; j < DL_MAX_RANKS; ++j) acc[u] = VecOps<kDtype>::add(acc[u], v[j]);
    }
  }
  // everybody has finished reading -> safe to overwrite our own copy
  const bool ok2 = dl_block_barrier(c, rank, world, a.channel, 0u, pay);
  if (ok && ok2) {
    const float inv = (a.scale && n_act > 1) ? 1.f / static_cast<float>(n_act) : 1.f;
    uint4* mine = dl_at<uint4>(c, rank, a.off);
#pragma unroll
    for (int u = 0; u < DL_ONESHOT_U; ++u) {
      const unsigned long long idx = tid0 + u * stride;
      if (idx < a.nvec) mine[idx] = (a.scale && n_act > 1) ? VecOps<kDtype>::scale(acc[u], inv) : acc[u];
    }
  }
}

// One pass over the owner's slice: reduce (kMc: in the switch; else P2P loads in rank order), scale,
// and hand the result to every rank (kMc: multimem.st; else P2P stores).
template <int W, int U, bool kMc>
__device__ __forceinline__ void allreduce_slice_pass(const dl_comm_t& c, const ArArgs& a, const DlSlice& s, int world,
                                                     uint32_t mask, float inv, unsigned long long tid0,
                                                     unsigned long long stride) {
  for (unsigned long long i0 = s.begin + tid0; i0 < s.end; i0 += stride * U) {
    float4 acc[U];
    if (kMc) {
      const float4* mg = dl_mc_at<float4>(c, a.off);
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
        for (int j = 0; j < W; ++j)
          v[u][j] = (idx < s.end && j < world && ((mask >> j) & 1u)) ? dl_ld_f4(dl_at<float4>(c, j, a.off) + idx)
                                                                    : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        acc[u] = v[u][0];
#pragma unroll
        for (int j = 1; j < W; ++j) acc[u] = acc[u] + v[u][j];
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned long long idx = i0 + u * stride;
      if (idx < s.end) {
        const float4 r = dl_scale(acc[u], inv);
        if (kMc) {
          dl_mc_st_f4(dl_mc_at<float4>(c, a.off) + idx, r);
        } else {
#pragma unroll
          for (int j = 0; j < W; ++j)
            if (j < world) dl_st_f4(dl_at<float4>(c, j, a.off) + idx, r);
        }
      }
    }
  }
}

#define DL_NVLS_U 8  // vectors in flight per thread on the multimem path (long switch round trip)

template <int kWorld, bool kNvls, bool kEmu>
__global__ void __launch_bounds__(DL_THREADS, 1) allreduce_twoshot_kernel(dl_comm_t c, ArArgs a) {
  __shared__ uint32_t pay[DL_MAX_RANKS];
  constexpr int W = kWorld ? kWorld : DL_MAX_RANKS;
  constexpr int U = kWorld == 2 ? 8 : kWorld == 4 ? 4 : 2;
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = kWorld ? kWorld : c.world;
  const int act = kEmu ? ((a.active >> rank) & 1u) : (a.active & 3u);
  const uint32_t user = kEmu ? a.payload + rank : a.payload;
  __shared__ uint32_t auxs[DL_MAX_RANKS];
  const bool ok = dl_block_barrier<false>(c, rank, world, a.channel, dl_pack_payload(act, user), pay, a.scale != 0 ? 1u : 0u, auxs);
  int n_act;
  const uint32_t mask = dl_active_mask(c, rank, world, a.channel, pay, &n_act);
  // whether to normalise is dictated by the lowest active rank (draining ranks just follow)
  a.scale = static_cast<int>(dl_aux_of_first_active(mask, auxs, rank) & 1u);
  const DlSlice s = dl_slice(a.nvec, rank, world);
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
  const unsigned long long tid0 = static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x;

  if (ok) {
    const float inv = (a.scale && n_act > 1) ? 1.f / static_cast<float>(n_act) : 1.f;
    if (kNvls && n_act == world) {
      allreduce_slice_pass<W, DL_NVLS_U, true>(c, a, s, world, mask, inv, tid0, stride);
    } else if (kNvls) {  // ranks are draining: leave their stale data out (P2P, rank order)
      allreduce_slice_pass<W, 1, false>(c, a, s, world, mask, inv, tid0, stride);
    } else {
      allreduce_slice_pass<W, U, false>(c, a, s, world, mask, inv, tid0, stride);
    }
  }
  dl_block_barrier(c, rank, world, a.channel, 0u, pay);
}

// ---- DL_VARIANT_HYBRID (float32): software-pipelined slice pass, switch + P2P in one launch (dl_pipe.cuh) ----
template <int W>
struct ArEpi {
  float inv;
  __device__ __forceinline__ float4 p2p(const float4 (&v)[W], unsigned long long) {
    float4 acc = v[0];
#pragma unroll
    for (int j = 1; j < W; ++j) acc = acc + v[j];
    return dl_scale(acc, inv);
  }
  __device__ __forceinline__ float4 mc(const float4& sum, unsigned long long) { return dl_scale(sum, inv); }
};

template <int kWorld, bool kEmu>
__global__ void __launch_bounds__(DL_PIPE_THREADS, 1) allreduce_pipe_kernel(dl_comm_t c, ArArgs a) {
  __shared__ uint32_t pay[DL_MAX_RANKS];
  __shared__ uint32_t auxs[DL_MAX_RANKS];
  constexpr int W = kWorld ? kWorld : DL_MAX_RANKS;
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = kWorld ? kWorld : c.world;
  const int act = kEmu ? ((a.active >> rank) & 1u) : (a.active & 3u);
  const uint32_t user = kEmu ? a.payload + rank : a.payload;
  const bool ok = dl_block_barrier<false>(c, rank, world, a.channel, dl_pack_payload(act, user), pay, a.scale != 0 ? 1u : 0u, auxs);
  int n_act;
  const uint32_t mask = dl_active_mask(c, rank, world, a.channel, pay, &n_act);
  a.scale = static_cast<int>(dl_aux_of_first_active(mask, auxs, rank) & 1u);
  const DlSlice s = dl_slice_even(a.nvec, rank, world);
  if (ok) {
    ArEpi<W> epi{(a.scale && n_act > 1) ? 1.f / static_cast<float>(n_act) : 1.f};
    dl_pipe_pass<W>(c, rank, world, mask, n_act == world, c.nvls_permille, a.off, a.off, 0, s, epi);
  }
  dl_block_barrier(c, rank, world, a.channel, 0u, pay);
}

#define DL_AR_DISPATCH_WORLD(...)                                                                      \
  switch (c->world) {                                                                                  \
    case 2: rc = dl_launch(allreduce_twoshot_kernel<2, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
    case 4: rc = dl_launch(allreduce_twoshot_kernel<4, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
    case 8: rc = dl_launch(allreduce_twoshot_kernel<8, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
    default: rc = dl_launch(allreduce_twoshot_kernel<0, __VA_ARGS__>, c, nb, DL_THREADS, st, *c, a); break; \
  }

#define DL_AR_ONESHOT(DT)                                                                                   \
  rc = c->emulate ? dl_launch(allreduce_oneshot_kernel<DT, true>, c, nb, DL_THREADS, st, *c, a)             \
                  : dl_launch(allreduce_oneshot_kernel<DT, false>, c, nb, DL_THREADS, st, *c, a)

extern "C" int dl_allreduce(const dl_comm_t* c, int dtype, uint64_t off, uint64_t n, int variant, int active,
                            uint32_t payload, int scale_by_inv_active, int nblocks, int channel, void* stream) {
  DL_REQUIRE(c != nullptr, "dl_allreduce: null communicator");
  const int esz = (dtype == DL_F32 || dtype == DL_I32) ? 4 : 8;
  DL_REQUIRE(dtype >= DL_F32 && dtype <= DL_F64, "dl_allreduce: bad dtype %d", dtype);
  DL_REQUIRE((n * esz) % 16 == 0, "dl_allreduce: byte length must be a multiple of 16 (pad the buffer)");
  DL_REQUIRE(off % 16 == 0 && off >= DL_PAD_BYTES, "dl_allreduce: bad offset");
  DL_REQUIRE(payload < (1u << 30), "dl_allreduce: payload must fit 30 bits");
  ArArgs a;
  a.off = off; a.nvec = n * esz / 16; a.active = static_cast<uint32_t>(active); a.payload = payload;
  a.scale = scale_by_inv_active; a.channel = channel;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = DL_OK;
  if (variant == DL_VARIANT_LL)
    return dl_ll_allreduce(c, dtype, off, n * esz, active, payload, scale_by_inv_active, nblocks, channel, st);
  if (variant == DL_VARIANT_ONESHOT) {
    int nb = dl_check_comm(c, channel, dl_pick_blocks(a.nvec, 1, nblocks));
    if (nb < 0) return nb;
    // capacity: every vector must fit in the registers of the grid
    while (static_cast<unsigned long long>(nb) * DL_THREADS * DL_ONESHOT_U < a.nvec) {
      const int cap = c->emulate ? dl_sm_count() / c->world : (dl_sm_count() < DL_MAX_BLOCKS ? dl_sm_count() : DL_MAX_BLOCKS);
      DL_REQUIRE(nb < cap, "dl_allreduce: %llu bytes exceed the one-shot capacity; use the two-shot variant",
                 (unsigned long long)(n * esz));
      nb = (nb * 2 < cap) ? nb * 2 : cap;
    }
    switch (dtype) {
      case DL_F32: DL_AR_ONESHOT(DL_F32); break;
      case DL_I32: DL_AR_ONESHOT(DL_I32); break;
      case DL_I64: DL_AR_ONESHOT(DL_I64); break;
      default: DL_AR_ONESHOT(DL_F64); break;
    }
    return rc;
  }
  DL_REQUIRE(dtype == DL_F32, "dl_allreduce: two-shot / NVLS / hybrid variants are float32 only");
  int nb = dl_check_comm(c, channel, dl_pick_blocks((a.nvec + c->world - 1) / c->world, 2, nblocks));
  if (nb < 0) return nb;
  if (variant == DL_VARIANT_HYBRID) {
    DL_REQUIRE(n % 8 == 0 && off % 32 == 0, "dl_allreduce: the hybrid variant needs n %% 8 == 0 and a 32-byte aligned offset");
    DL_REQUIRE(c->nvls_permille >= 0 && c->nvls_permille <= 1000, "dl_allreduce: nvls_permille out of range");
    DL_REQUIRE(c->nvls_permille == 0 || (c->mc && !c->emulate), "dl_allreduce: nvls_permille > 0 needs a multicast mapping");
    nb = dl_check_comm(c, channel, dl_pick_blocks((a.nvec + c->world - 1) / c->world / 2 + 1, 1, nblocks));
    if (nb < 0) return nb;
#define DL_AR_PIPE(WORLD)                                                                                         \
    rc = c->emulate ? dl_launch(allreduce_pipe_kernel<WORLD, true>, c, nb, DL_PIPE_THREADS, st, *c, a)            \
                    : dl_launch(allreduce_pipe_kernel<WORLD, false>, c, nb, DL_PIPE_THREADS, st, *c, a)
    switch (c->world) {
      case 1: DL_AR_PIPE(1); break;
      case 2: DL_AR_PIPE(2); break;
      case 4: DL_AR_PIPE(4); break;
      case 8: DL_AR_PIPE(8); break;
      default: DL_AR_PIPE(0); break;
    }
#undef DL_AR_PIPE
    return rc;
  }
  if (variant == DL_VARIANT_NVLS) {
    DL_REQUIRE(c->mc && !c->emulate, "dl_allreduce: NVLS variant needs a multicast mapping");
    DL_AR_DISPATCH_WORLD(true, false)
  } else if (variant == DL_VARIANT_TWOSHOT) {
    if (c->emulate) { DL_AR_DISPATCH_WORLD(false, true) } else { DL_AR_DISPATCH_WORLD(false, false) }
  } else {
    dl_set_error("dl_allreduce: unknown variant %d", variant);
    return DL_EINVAL;
  }
  return rc;
}
