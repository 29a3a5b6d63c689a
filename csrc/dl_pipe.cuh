// distlearn_b200 — the software-pipelined slice pass shared by the large-message collectives
// (DL_VARIANT_HYBRID of fused_sgd / allreduce / fused_ea).
//
// Why: in the first-generation kernels every thread issued all its loads, waited one NVLink round
// trip, then issued all its stores, and a 16 MiB vector on 8 GPUs is less than one such round per
// thread — so the whole grid read in lockstep (ingress busy, egress idle) and then wrote in lockstep
// (egress busy, ingress idle).  NVLink is full duplex; here each thread walks its part of the slice in
// small items and issues the loads of item k+1 BEFORE it reduces and stores item k, so in the steady
// state the reduce-scatter traffic of one item overlaps the all-gather traffic of the previous one.
//
// Two data paths run side by side in ONE launch (the "hybrid"):
//   * blocks [0, nb_n): the first `permille` thousandths of every slice go through the NVSwitch —
//     multimem.ld_reduce (the switch adds the N copies) and multimem.st (the switch replicates),
//     (1 + 1/N) S bytes per direction per GPU, but bounded by the switch's reduction engines
//     (measured ~0.70 of the link rate on this pool, NCCL's NVLS all-reduce sits at the same level);
//   * blocks [nb_n, nb): the rest of the slice over plain P2P — 256-bit loads from every peer
//     (LDG.E.ENL2.256), rank-order sum, 256-bit stores to every peer, 2 (N-1)/N S bytes per direction.
// The two paths are limited by different resources, so together they move more than either alone.
//
// Geometry is slice-relative and a pure function of (slice length, grid, permille, block), hence identical on
// every rank: block b of rank r and block b of rank s touch the same offsets of their respective slices,
// which is what the block-level departure barrier certifies ("block b of every peer is done with offsets
// R_b of its slice of MY buffer").
#pragma once
#include "dl_common.cuh"

#define DL_PIPE_THREADS 256
// multimem.ld_reduce vectors per thread per stage (2 stages in flight).  Measured on 8 GPUs, graph-timed, all-switch pass
// (profiles/r2/r2_sweep_8_graph.json with 4, r2z_sweep_8_graph.json with 8): 1 MiB 15.3 -> 13.9 us, 4 MiB 22.7 -> 21.3,
// 16 MiB 60.9 -> 59.7, but 64 MiB 179.4 -> 198.8: short passes want more loads in flight, long ones saturate the switch
// and then queue.  So: 8 while a rank's slice is below DL_PIPE_MC_BIG float4 (4 MiB), 4 from there on.
#define DL_PIPE_MC_U 8
#define DL_PIPE_MC_U_BIG 4
#define DL_PIPE_MC_BIG (1ull << 18)

struct DlPipeRange {
  unsigned long long r0, r1;  // slice-relative range of this block in float4 units (both even)
  bool nvls;
};

// slice4 must be even.  permille: share of the slice that goes through the switch.
__host__ __device__ __forceinline__ DlPipeRange dl_pipe_range(unsigned long long slice4, int nb, int b, int permille) {
  const unsigned long long pairs = slice4 / 2;
  unsigned long long pn = pairs * static_cast<unsigned long long>(permille) / 1000ull;
  int nbn = 0;
  if (pn > 0) {
    nbn = static_cast<int>(static_cast<long long>(nb) * permille / 1000);
    if (nbn < 1) nbn = 1;
    if (pn < pairs && nbn > nb - 1) nbn = nb - 1;
    if (pn == pairs) nbn = nb;
  }
  if (nbn == 0) pn = 0;
  DlPipeRange g;
  if (b < nbn) {
    const unsigned long long per = (pn + nbn - 1) / nbn;
    unsigned long long lo = per * static_cast<unsigned long long>(b);
    if (lo > pn) lo = pn;
    unsigned long long hi = lo + per;
    if (hi > pn) hi = pn;
    g.r0 = 2 * lo; g.r1 = 2 * hi; g.nvls = true;
  } else {
    const unsigned long long q = pairs - pn;
    const int nbp = nb - nbn;
    const unsigned long long per = (q + nbp - 1) / (nbp > 0 ? nbp : 1);
    unsigned long long lo = per * static_cast<unsigned long long>(b - nbn);
    if (lo > q) lo = q;
    unsigned long long hi = lo + per;
    if (hi > q) hi = q;
    g.r0 = 2 * (pn + lo); g.r1 = 2 * (pn + hi); g.nvls = false;
  }
  return g;
}

// even-length slices (the last one may be shorter or empty): 256-bit accesses never straddle a slice boundary
__device__ __forceinline__ DlSlice dl_slice_even(unsigned long long n4, int rank, int world) {
  DlSlice s;
  s.slice4 = (n4 + world - 1) / world;
  s.slice4 += s.slice4 & 1ull;
  s.begin = s.slice4 * static_cast<unsigned long long>(rank);
  s.end = s.begin + s.slice4;
  if (s.begin > n4) s.begin = n4;
  if (s.end > n4) s.end = n4;
  return s;
}

// ---- P2P path: items of 32 bytes ---------------------------------------------------------------------
// Epi::p2p(const float4 (&v)[W], idx) -> float4 to distribute (v[j] = source j's vector, zeros when j is not in
// `lmask`); called for both halves of an item.  dst_off (and dst2_off if non-zero) receive the result on every rank.
template <int W, typename Epi>
__device__ __forceinline__ void dl_pipe_p2p(const dl_comm_t& c, int world, uint32_t lmask, uint64_t src_off, uint64_t dst_off,
                                            uint64_t dst2_off, unsigned long long begin4, unsigned long long end4, Epi& epi) {
  // begin4/end4: absolute float4 range of this block (even).  A stage is U items per thread, U x W = 8 source vectors of 32
  // bytes: the same 256 bytes per thread and stage in flight whatever the world size (with two ranks one item per stage
  // leaves the link idle most of a round trip: measured 487 vs 417 us at 256 MiB against the unpipelined kernel).
  constexpr int U = W >= 8 ? 1 : W >= 4 ? 2 : 4;
  constexpr unsigned long long T = DL_PIPE_THREADS;
  const unsigned long long pb = begin4 / 2, pe = end4 / 2;
  dl_f8 cur[U][W], nxt[U][W];
  const dl_f8 zero = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
  unsigned long long it = pb + threadIdx.x;
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const unsigned long long idx = it + u * T;
#pragma unroll
    for (int j = 0; j < W; ++j)
      cur[u][j] = (idx < pe && j < world && ((lmask >> j) & 1u)) ? dl_ld_f8(dl_at<float4>(c, j, src_off) + 2 * idx) : zero;
  }
  for (; it < pe; it += U * T) {
    const unsigned long long nx = it + U * T;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned long long idx = nx + u * T;
#pragma unroll
      for (int j = 0; j < W; ++j)
        nxt[u][j] = (idx < pe && j < world && ((lmask >> j) & 1u)) ? dl_ld_f8(dl_at<float4>(c, j, src_off) + 2 * idx) : zero;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned long long idx = it + u * T;
      if (idx < pe) {
        float4 lo[W], hi[W];
#pragma unroll
        for (int j = 0; j < W; ++j) { lo[j] = cur[u][j].lo; hi[j] = cur[u][j].hi; }
        dl_f8 out;
        out.lo = epi.p2p(lo, 2 * idx);
        out.hi = epi.p2p(hi, 2 * idx + 1);
#pragma unroll
        for (int j = 0; j < W; ++j) {
          if (j < world) {
            dl_st_f8(dl_at<float4>(c, j, dst_off) + 2 * idx, out);
            if (dst2_off) dl_st_f8(dl_at<float4>(c, j, dst2_off) + 2 * idx, out);
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int j = 0; j < W; ++j) cur[u][j] = nxt[u][j];
  }
}

// ---- switch path: items of DL_PIPE_MC_U x 16 bytes ------------------------------------------------------
// Epi::mc(float4 sum_over_all_ranks, idx) -> float4 to distribute.
template <int U, typename Epi>
__device__ __forceinline__ void dl_pipe_mc(const dl_comm_t& c, uint64_t src_off, uint64_t dst_off, uint64_t dst2_off,
                                           unsigned long long begin4, unsigned long long end4, Epi& epi) {
  const float4* msrc = dl_mc_at<float4>(c, src_off);
  float4* mdst = dl_mc_at<float4>(c, dst_off);
  float4* mdst2 = dst2_off ? dl_mc_at<float4>(c, dst2_off) : nullptr;
  float4 cur[U], nxt[U];
  unsigned long long it = begin4 + threadIdx.x;
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const unsigned long long idx = it + static_cast<unsigned long long>(u) * DL_PIPE_THREADS;
    cur[u] = idx < end4 ? dl_mc_ld_reduce_f4(msrc + idx) : z;
  }
  for (; it < end4; it += static_cast<unsigned long long>(U) * DL_PIPE_THREADS) {
    const unsigned long long nx = it + static_cast<unsigned long long>(U) * DL_PIPE_THREADS;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned long long idx = nx + static_cast<unsigned long long>(u) * DL_PIPE_THREADS;
      nxt[u] = idx < end4 ? dl_mc_ld_reduce_f4(msrc + idx) : z;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned long long idx = it + static_cast<unsigned long long>(u) * DL_PIPE_THREADS;
      if (idx < end4) {
        const float4 out = epi.mc(cur[u], idx);
        dl_mc_st_f4(mdst + idx, out);
        if (mdst2) dl_mc_st_f4(mdst2 + idx, out);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) cur[u] = nxt[u];
  }
}

// The whole data phase of a hybrid collective for block blockIdx.x: pick the path and the range.
// `full` = every rank is active (the switch adds ALL copies, so a round with draining ranks runs P2P only).
template <int W, typename Epi>
__device__ __forceinline__ void dl_pipe_pass(const dl_comm_t& c, int rank, int world, uint32_t lmask, bool full, int permille,
                                             uint64_t src_off, uint64_t dst_off, uint64_t dst2_off, const DlSlice& s, Epi& epi) {
  const DlPipeRange g = dl_pipe_range(s.slice4, static_cast<int>(gridDim.x), static_cast<int>(blockIdx.x),
                                      (full && c.mc != nullptr) ? permille : 0);
  unsigned long long b4 = s.begin + g.r0, e4 = s.begin + g.r1;
  if (b4 > s.end) b4 = s.end;
  if (e4 > s.end) e4 = s.end;
  if (g.nvls) {
    if (s.slice4 >= DL_PIPE_MC_BIG) dl_pipe_mc<DL_PIPE_MC_U_BIG>(c, src_off, dst_off, dst2_off, b4, e4, epi);
    else dl_pipe_mc<DL_PIPE_MC_U>(c, src_off, dst_off, dst2_off, b4, e4, epi);
  } else {
    dl_pipe_p2p<W>(c, world, lmask, src_off, dst_off, dst2_off, b4, e4, epi);
  }
}

// After the departure barrier: the offsets of MY buffer that block b of every peer has finished with — the same
// slice-relative range in each of the `world` slices.  fn(idx) is called once per float4 of that set.
template <typename Fn>
__device__ __forceinline__ void dl_pipe_foreach_released(unsigned long long n4, int world, bool full, bool has_mc, int permille,
                                                         const DlSlice& s, Fn fn) {
  const DlPipeRange g = dl_pipe_range(s.slice4, static_cast<int>(gridDim.x), static_cast<int>(blockIdx.x),
                                      (full && has_mc) ? permille : 0);
  for (int j = 0; j < world; ++j) {
    const unsigned long long base = s.slice4 * static_cast<unsigned long long>(j);
    unsigned long long b4 = base + g.r0, e4 = base + g.r1;
    if (b4 > n4) b4 = n4;
    if (e4 > n4) e4 = n4;
    for (unsigned long long k = b4 + threadIdx.x; k < e4; k += DL_PIPE_THREADS) fn(k);
  }
}

// ---- the LIGHT slice pass: made to run UNDER the convolutions -------------------------------------------------------
// A gradient bucket that is reduced while backward is still running shares the SMs with the tensor-core GEMM, whose
// persistent CTA (576 threads x 96 registers, ~200 KB of shared memory) leaves 10,240 registers and ~25 KB of shared
// memory per SM.  The kernels above (256 threads x 192 registers) cannot move in beside it: they wait for whole SMs to
// drain, and then keep the next GEMM's CTAs out in turn — which is why per-bucket launches gained 1 % in round 1.
// The light pass is sized to fit the gap instead: DL_LIGHT_THREADS threads of at most 80 registers, one 16-byte vector
// per source in flight per thread and no double buffering — the bytes in flight come from running on ALL SMs at once
// (148 x 128 threads x 8 sources x 16 B = 2.4 MB, the bandwidth-delay product of the links).
#define DL_LIGHT_THREADS 128
#define DL_LIGHT_MC_U 8

template <int W, typename Epi>
__device__ __forceinline__ void dl_light_p2p(const dl_comm_t& c, int world, uint32_t lmask, uint64_t src_off, uint64_t dst_off,
                                             unsigned long long begin4, unsigned long long end4, Epi& epi) {
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  for (unsigned long long i = begin4 + threadIdx.x; i < end4; i += DL_LIGHT_THREADS) {
    float4 v[W];
#pragma unroll
    for (int j = 0; j < W; ++j) v[j] = (j < world && ((lmask >> j) & 1u)) ? dl_ld_f4(dl_at<float4>(c, j, src_off) + i) : z;
    const float4 out = epi.p2p(v, i);
#pragma unroll
    for (int j = 0; j < W; ++j)
      if (j < world) dl_st_f4(dl_at<float4>(c, j, dst_off) + i, out);
  }
}

template <typename Epi>
__device__ __forceinline__ void dl_light_mc(const dl_comm_t& c, uint64_t src_off, uint64_t dst_off, unsigned long long begin4,
                                            unsigned long long end4, Epi& epi) {
  constexpr int U = DL_LIGHT_MC_U;
  const float4* msrc = dl_mc_at<float4>(c, src_off);
  float4* mdst = dl_mc_at<float4>(c, dst_off);
  for (unsigned long long i0 = begin4 + threadIdx.x; i0 < end4; i0 += static_cast<unsigned long long>(U) * DL_LIGHT_THREADS) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned long long idx = i0 + static_cast<unsigned long long>(u) * DL_LIGHT_THREADS;
      if (idx < end4) v[u] = dl_mc_ld_reduce_f4(msrc + idx);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned long long idx = i0 + static_cast<unsigned long long>(u) * DL_LIGHT_THREADS;
      if (idx < end4) dl_mc_st_f4(mdst + idx, epi.mc(v[u], idx));
    }
  }
}

template <int W, typename Epi>
__device__ __forceinline__ void dl_light_pass(const dl_comm_t& c, int world, uint32_t lmask, bool full, int permille, uint64_t src_off,
                                              uint64_t dst_off, const DlSlice& s, Epi& epi) {
  const DlPipeRange g = dl_pipe_range(s.slice4, static_cast<int>(gridDim.x), static_cast<int>(blockIdx.x),
                                      (full && c.mc != nullptr) ? permille : 0);
  unsigned long long b4 = s.begin + g.r0, e4 = s.begin + g.r1;
  if (b4 > s.end) b4 = s.end;
  if (e4 > s.end) e4 = s.end;
  if (g.nvls) dl_light_mc(c, src_off, dst_off, b4, e4, epi);
  else dl_light_p2p<W>(c, world, lmask, src_off, dst_off, b4, e4, epi);
}

template <typename Fn>
__device__ __forceinline__ void dl_light_foreach_released(unsigned long long n4, int world, bool full, bool has_mc, int permille,
                                                          const DlSlice& s, Fn fn) {
  const DlPipeRange g = dl_pipe_range(s.slice4, static_cast<int>(gridDim.x), static_cast<int>(blockIdx.x),
                                      (full && has_mc) ? permille : 0);
  for (int j = 0; j < world; ++j) {
    const unsigned long long base = s.slice4 * static_cast<unsigned long long>(j);
    unsigned long long b4 = base + g.r0, e4 = base + g.r1;
    if (b4 > n4) b4 = n4;
    if (e4 > n4) e4 = n4;
    for (unsigned long long k = b4 + threadIdx.x; k < e4; k += DL_LIGHT_THREADS) fn(k);
  }
}
