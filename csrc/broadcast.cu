// broadcast_multicast — SURVEY K5/K6/K10/W6.
// Replaces tree.scatter(params) (lua/AllReduceSGD.lua:52; lua/AllReduceEA.lua:83,93) and the
// "zero the losers + all-reduce" winner broadcast (lua/AllReduceSGD.lua:43-47).
//
//   one-shot : every rank pulls the whole vector from the root (small messages)
//   two-shot : every non-root rank pulls one of N-1 slices from the root and pushes it to the other
//              ranks in the same pass, so the root's NVLink egress is S instead of (N-1) S
//   NVLS     : the same N-1 owners pull their slice from the root and hand it to the switch with one
//              multimem.st per 16 bytes, which replicates it to every rank: root egress S, owner egress
//              S/(N-1) instead of S (N-2)/(N-1).  (A root that multicasts the whole vector alone is bounded by
//              one GPU's multicast issue rate: 405 GB/s at 1 GiB on 8 GPUs, against 647 for NCCL.)
//   LL       : small messages, csrc/ll_ops.cu
// The two-shot and NVLS passes are software pipelined (the pull of item k+1 is in flight while item k is pushed) and
// move 32 bytes per access when the buffers allow it.
// With off2 != 0 every rank also receives the values at off2 (EA: center <- params,
// lua/AllReduceEA.lua:93-99) without a second pass.
#include "dl_common.cuh"
#include "dl_host.h"

struct BcArgs {
  uint64_t off, off2;
  unsigned long long n4;
  int root;
  uint32_t payload;
  int channel;
};

// Wide (32-byte) pipelined form of the owner pass: pull from the root, push to everybody.  kMc: the push is a multicast store.
template <bool kMc, bool kEmu>
__global__ void __launch_bounds__(DL_THREADS, 1) broadcast_wide_kernel(dl_comm_t c, BcArgs a) {
  __shared__ uint32_t pay[DL_MAX_RANKS];
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = c.world;
  const uint32_t user = kEmu ? a.payload + rank : a.payload;
  const bool ok = dl_block_barrier<false>(c, rank, world, a.channel, dl_pack_payload(1, user), pay);
  int n_act;
  dl_active_mask(c, rank, world, a.channel, pay, &n_act);
  if (ok && rank != a.root) {
    const int owners = world - 1;
    const int me = rank < a.root ? rank : rank - 1;
    const DlSlice s = dl_slice(a.n4 / 2, me, owners);                 // in 32-byte items
    const float4* src = dl_at<float4>(c, a.root, a.off);
    const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
    unsigned long long it = s.begin + static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x;
    dl_f8 cur, nxt;
    if (it < s.end) cur = dl_ld_f8(src + 2 * it);
    for (; it < s.end; it += stride) {
      const unsigned long long nx = it + stride;
      if (nx < s.end) nxt = dl_ld_f8(src + 2 * nx);
      if (kMc) {
        // the root's own copy is rewritten with identical bits: harmless
        dl_mc_st_f4(dl_mc_at<float4>(c, a.off) + 2 * it, cur.lo);
        dl_mc_st_f4(dl_mc_at<float4>(c, a.off) + 2 * it + 1, cur.hi);
        if (a.off2) {
          dl_mc_st_f4(dl_mc_at<float4>(c, a.off2) + 2 * it, cur.lo);
          dl_mc_st_f4(dl_mc_at<float4>(c, a.off2) + 2 * it + 1, cur.hi);
        }
      } else {
#pragma unroll
        for (int j = 0; j < DL_MAX_RANKS; ++j) {
          if (j < world) {
            if (j != a.root) dl_st_f8(dl_at<float4>(c, j, a.off) + 2 * it, cur);
            if (a.off2) dl_st_f8(dl_at<float4>(c, j, a.off2) + 2 * it, cur);
          }
        }
      }
      cur = nxt;
    }
  }
  dl_block_barrier(c, rank, world, a.channel, 0u, pay);
}

template <int kVariant, bool kEmu>
__global__ void __launch_bounds__(DL_THREADS, 1) broadcast_kernel(dl_comm_t c, BcArgs a) {
  __shared__ uint32_t pay[DL_MAX_RANKS];
  constexpr int U = 4;
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = c.world;
  const uint32_t user = kEmu ? a.payload + rank : a.payload;
  const bool ok = dl_block_barrier<false>(c, rank, world, a.channel, dl_pack_payload(1, user), pay);
  int n_act;
  dl_active_mask(c, rank, world, a.channel, pay, &n_act);
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
  const unsigned long long tid0 = static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x;
  const float4* src = dl_at<float4>(c, a.root, a.off);

  if (ok) {
    if (kVariant == DL_VARIANT_ONESHOT) {
      float4* dst = dl_at<float4>(c, rank, a.off);
      float4* dst2 = a.off2 ? dl_at<float4>(c, rank, a.off2) : nullptr;
      for (unsigned long long i0 = tid0; i0 < a.n4; i0 += stride * U) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned long long idx = i0 + u * stride;
          if (idx < a.n4) v[u] = dl_ld_f4(src + idx);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned long long idx = i0 + u * stride;
          if (idx < a.n4) {
            if (rank != a.root) dst[idx] = v[u];
            if (dst2) dst2[idx] = v[u];
          }
        }
      }
    } else if (kVariant == DL_VARIANT_TWOSHOT) {
      // Slices are owned by the NON-root ranks only (world > 1): the root's NVLink egress is then S — every byte leaves
      // it exactly once — instead of S (pulls) + S (N-1)/N (pushes of a slice of its own), which is what bounded the
      // large-message broadcast (1 GiB, 8 GPUs: 2.48 ms = 1.75 S at 770 GB/s).  The root only takes part in the barriers.
      const int owners = world > 1 ? world - 1 : 1;
      const int me = world > 1 ? (rank < a.root ? rank : rank - 1) : 0;
      DlSlice s = dl_slice(a.n4, me, owners);
      if (world > 1 && rank == a.root) s.begin = s.end = 0;
      for (unsigned long long i0 = s.begin + tid0; i0 < s.end; i0 += stride * U) {
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned long long idx = i0 + u * stride;
          if (idx < s.end) v[u] = dl_ld_f4(src + idx);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned long long idx = i0 + u * stride;
          if (idx < s.end) {
#pragma unroll
            for (int j = 0; j < DL_MAX_RANKS; ++j) {
              if (j < world) {
                if (j != a.root) dl_st_f4(dl_at<float4>(c, j, a.off) + idx, v[u]);
                if (a.off2) dl_st_f4(dl_at<float4>(c, j, a.off2) + idx, v[u]);
              }
            }
          }
        }
      }
    } else {  // NVLS
      if (rank == a.root) {
        for (unsigned long long i0 = tid0; i0 < a.n4; i0 += stride * U) {
          float4 v[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const unsigned long long idx = i0 + u * stride;
            if (idx < a.n4) v[u] = src[idx];
          }
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const unsigned long long idx = i0 + u * stride;
            if (idx < a.n4) {
              // the root's own copy is rewritten with identical bits: harmless
              dl_mc_st_f4(dl_mc_at<float4>(c, a.off) + idx, v[u]);
              if (a.off2) dl_mc_st_f4(dl_mc_at<float4>(c, a.off2) + idx, v[u]);
            }
          }
        }
      }
    }
  }
  dl_block_barrier(c, rank, world, a.channel, 0u, pay);
}

// Plain P2P pull: dst[i] = peer heap[off + i]  (initClient / startTest center fetch).
__global__ void __launch_bounds__(DL_THREADS) peer_pull_kernel(const float4* src, float4* dst, unsigned long long n4,
                                                             const float* src_tail, float* dst_tail, int tail) {
  constexpr int U = 4;
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
  for (unsigned long long i0 = static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x; i0 < n4;
       i0 += stride * U) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i0 + u * stride < n4) v[u] = dl_ld_f4(src + i0 + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i0 + u * stride < n4) dst[i0 + u * stride] = v[u];
  }
  if (blockIdx.x == 0 && threadIdx.x < tail) dst_tail[threadIdx.x] = src_tail[threadIdx.x];
}

#define DL_BC_LAUNCH(V)                                                                         \
  rc = c->emulate ? dl_launch(broadcast_kernel<V, true>, c, nb, DL_THREADS, st, *c, a)          \
                  : dl_launch(broadcast_kernel<V, false>, c, nb, DL_THREADS, st, *c, a)

extern "C" int dl_broadcast(const dl_comm_t* c, uint64_t off, uint64_t off2, uint64_t n, int root, int variant,
                            uint32_t payload, int nblocks, int channel, void* stream) {
  DL_REQUIRE(c != nullptr, "dl_broadcast: null communicator");
  DL_REQUIRE(n % 4 == 0, "dl_broadcast: n must be a multiple of 4 floats");
  DL_REQUIRE(off % 16 == 0 && off2 % 16 == 0 && off >= DL_PAD_BYTES && (off2 == 0 || off2 >= DL_PAD_BYTES),
             "dl_broadcast: bad offsets");
  DL_REQUIRE(root >= 0 && root < c->world, "dl_broadcast: bad root %d", root);
  DL_REQUIRE(payload < (1u << 30), "dl_broadcast: payload must fit 30 bits");
  BcArgs a;
  a.off = off; a.off2 = off2; a.n4 = n / 4; a.root = root; a.payload = payload; a.channel = channel;
  if (variant == DL_VARIANT_LL)
    return dl_ll_broadcast(c, off, off2, n, root, payload, nblocks, channel, static_cast<cudaStream_t>(stream));
  const int owners = c->world > 1 ? c->world - 1 : 1;
  const unsigned long long work = variant != DL_VARIANT_ONESHOT ? (a.n4 + owners - 1) / owners : a.n4;
  int nb = dl_check_comm(c, channel, dl_pick_blocks(work, 4, nblocks));
  if (nb < 0) return nb;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = DL_OK;
  const bool wide = c->world > 1 && n % 8 == 0 && off % 32 == 0 && off2 % 32 == 0;
  if (variant == DL_VARIANT_ONESHOT) {
    DL_BC_LAUNCH(DL_VARIANT_ONESHOT);
  } else if (variant == DL_VARIANT_TWOSHOT && wide) {
    rc = c->emulate ? dl_launch(broadcast_wide_kernel<false, true>, c, nb, DL_THREADS, st, *c, a)
                    : dl_launch(broadcast_wide_kernel<false, false>, c, nb, DL_THREADS, st, *c, a);
  } else if (variant == DL_VARIANT_TWOSHOT) {
    DL_BC_LAUNCH(DL_VARIANT_TWOSHOT);
  } else if (variant == DL_VARIANT_NVLS) {
    DL_REQUIRE(c->mc && !c->emulate, "dl_broadcast: NVLS variant needs a multicast mapping");
    if (wide) {
      rc = dl_launch(broadcast_wide_kernel<true, false>, c, nb, DL_THREADS, st, *c, a);
    } else {
      rc = dl_launch(broadcast_kernel<DL_VARIANT_NVLS, false>, c, nb, DL_THREADS, st, *c, a);
    }
  } else {
    dl_set_error("dl_broadcast: unknown variant %d", variant);
    return DL_EINVAL;
  }
  return rc;
}

extern "C" int dl_peer_pull(const dl_comm_t* c, int src_rank, uint64_t off, float* dst, uint64_t n, int nblocks,
                            void* stream) {
  int nb = dl_check_comm(c, 0, dl_pick_blocks(n / 4 + 1, 4, nblocks));
  if (nb < 0) return nb;
  DL_REQUIRE(src_rank >= 0 && src_rank < c->world, "dl_peer_pull: bad source rank %d", src_rank);
  DL_REQUIRE(off % 16 == 0 && reinterpret_cast<uintptr_t>(dst) % 16 == 0, "dl_peer_pull: 16-byte alignment required");
  const float* src = reinterpret_cast<const float*>(reinterpret_cast<const char*>(c->base[src_rank]) + off);
  const unsigned long long n4 = n / 4;
  const int tail = static_cast<int>(n - n4 * 4);
  peer_pull_kernel<<<nb, DL_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float4*>(src), reinterpret_cast<float4*>(dst), n4, src + n4 * 4, dst + n4 * 4, tail);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}
