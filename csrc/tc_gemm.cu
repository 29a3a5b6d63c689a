// tc_gemm — fp32-accurate GEMM on the 5th-generation tensor cores (tcgen05 + TMEM + TMA).
//
//     C[M,N] (+)= A[M,K] * B[N,K]^T (+ bias[N])        A, B, C row-major fp32, K contiguous in A and B
//
// The reference trains in strict fp32 (`grad.util.cast(params, 'float')`, examples/cifar10.lua:136) and
// its convolutions are im2col + SGEMM (SpatialConvolutionMM, examples/Model.lua:20).  On B200 an fp32
// FFMA GEMM tops out near 70 TFLOP/s while the tensor cores take TF32 inputs at >1 PFLOP/s, so this
// kernel keeps fp32 accuracy with the 3xTF32 split:  a = a_hi + a_lo  (a_hi = a rounded to TF32,
// a_lo = a - a_hi exactly),  a*b ~= a_hi*b_hi + a_hi*b_lo + a_lo*b_hi  accumulated in fp32 in TMEM
// (the dropped a_lo*b_lo term is ~2^-22 relative).  `terms = 1` runs plain TF32 (a_hi*b_hi only).
//
// One CTA computes one 128 x BN output tile (optionally one K-split of it), warp-specialised:
//   warp 0      TMA producer: cp.async.bulk.tensor 128x32 (A) and BNx32 (B) fp32 boxes, 128B swizzle,
//               into a 3-stage ring, completion on mbarriers
//   warps 2-5   converters: split the raw fp32 tile in place into hi (same buffer) and lo (sibling
//               buffer) — element-wise at identical byte offsets, so the TMA swizzle is preserved —
//               then fence.proxy.async and arrive; after the main loop the same warps are the epilogue
//   warp 1      allocates TMEM; one elected lane issues tcgen05.mma.kind::tf32 (M=128, N=BN, K=8):
//               3 products x 4 K-steps per stage; tcgen05.commit frees the stage / publishes the tile
//   warps 6-9   accumulator drain + epilogue.  The tensor core adds into TMEM with truncation, which over
//               thousands of K steps would cost ~1e-5 relative — more than an fp32 FFMA chain.  So the
//               accumulator is double-buffered in TMEM and flushed every CHUNK K-blocks (K = 128):
//               tcgen05.ld 32 lanes x 32 columns per warp, added into fp32 registers (round-to-nearest)
//               while the MMAs of the next chunk run into the other buffer.  At the end: (+bias) ->
//               128-byte row segments to global memory (st, or red.add for split-K / accumulate)
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <mutex>

#include "dl_host.h"

namespace {

constexpr int BM = 128;      // UMMA_M
constexpr int BK = 32;       // fp32 elements per 128-byte swizzle row
constexpr int UMMA_K = 8;    // K per tcgen05.mma.kind::tf32
constexpr int STAGES = 3;
constexpr int NUM_THREADS = 320;
constexpr int CHUNK = 4;     // K-blocks accumulated inside TMEM before the fp32 register flush

struct GemmArgs {
  float* C;
  const float* bias;
  int M, N, K;
  long long ldc;
  int kblocks_per_split;  // K blocks (of BK) each blockIdx.z handles
  int atomic_out;         // 1: red.add into C (split-K or accumulate), 0: plain store
  int terms;              // 3 = 3xTF32 (fp32-accurate), 1 = TF32
  int a_mn, b_mn;         // operand stored with its M / N dimension contiguous ([K, MN] row-major) instead of K
  int vec_out;            // rows of C are 16-byte aligned: float4 stores / red.v4
  // implicit-GEMM convolution (5x5, stride 1, pad 2, NHWC): the operand marked (*) is read straight from the
  // activation tensor with 4-D TMA boxes shifted by the filter tap — out-of-image elements arrive as zeros,
  // so no im2col / col2im buffer exists.
  //   mode 1 forward : C[pixels, Cout] = sum_{tap,c}  X(*)[pixel + tap, c]   * W[cout, (tap, c)]
  //   mode 2 dgrad   : C[pixels, Cin]  = sum_{tap,co} dY(*)[pixel - tap, co] * W[co, (tap, cin)]   (W read MN-major)
  //   mode 3 wgrad   : C[Cout, 25 Cin] = sum_pixels   dY[pixel, cout]        * X(*)[pixel + tap, c]  (both MN-major)
  int mode;
  int cH, cW;             // image height / width
  int cA;                 // channels of the 4-D tensor on the A side (mode 1: Cin, mode 2: Cout)
  int cIn;                // Cin (modes 2 and 3)
  int b_split;            // stream-K kernel only: the B operand arrives already split (map_b = hi, map_b2 = lo, made once
                          // per step by dl_split_tf32): the converter warps leave B alone
};

// ---- PTX wrappers ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a mis-programmed pipeline must fail the launch, not hang the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t i = 0; i < (1u << 26); ++i)
    if (mbar_try_wait(bar, parity)) return;
  __trap();
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
// explicit shared-space 128-bit accesses (a generic pointer derived from the aligned dynamic-smem base makes the
// compiler emit generic LD/ST, which are slower than LDS/STS)
__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, const float4& v) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float tf32_round(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

// K-major operand tile with 128-byte swizzle: rows of 128 bytes, 8-row groups 1024 bytes apart.
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);  // start address, 16-byte units
  d |= static_cast<uint64_t>(1) << 16;                       // leading byte offset (unused with swizzle)
  d |= static_cast<uint64_t>(1024 >> 4) << 32;               // stride byte offset: next 8-row group
  d |= static_cast<uint64_t>(1) << 46;                       // descriptor version (Blackwell)
  d |= static_cast<uint64_t>(2) << 61;                       // SWIZZLE_128B
  return d;
}

// MN-major operand tile, as TMA writes it from a [K, MN] row-major matrix in boxes of 32 K-rows x 32 MN-elements.
// For 32-bit (TF32) operands the tensor core accepts exactly one MN-major shared-memory layout: 128-byte rows
// swizzled in 32-byte chunks (UMMA layout type 1 = SWIZZLE_128B_BASE32B, TMA CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B),
// whose atom is 4 K-rows x 128 bytes.  Inside a box the 4-row atoms are 512 bytes apart (SBO); boxes (the next
// 32 MN elements) are 4096 bytes apart (LBO).  One tcgen05.mma (K = 8) consumes two atoms = 1024 bytes.
__device__ __forceinline__ uint64_t make_desc_mn(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(4096 >> 4) << 16;               // leading byte offset: next 32 MN elements
  d |= static_cast<uint64_t>(512 >> 4) << 32;                // stride byte offset: next 4 K rows
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(1) << 61;                       // SWIZZLE_128B_BASE32B
  return d;
}

template <int BN>
struct Smem {
  // [stage][0 = hi (TMA destination), 1 = lo]; every tile is 1024-byte aligned (swizzle atom)
  float a[STAGES][2][BM * BK];
  float b[STAGES][2][BN * BK];
  uint64_t full[STAGES], conv[STAGES], empty[STAGES], acc_full[2], acc_empty[2];
  uint32_t tmem_base;
};

template <int BN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, GemmArgs g) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  Smem<BN>& s = *reinterpret_cast<Smem<BN>*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int total_kb = (g.K + BK - 1) / BK;
  const int kb_begin = blockIdx.z * g.kblocks_per_split;
  int kb_end = kb_begin + g.kblocks_per_split;
  if (kb_end > total_kb) kb_end = total_kb;
  const int nkb = kb_end - kb_begin;

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&s.full[i], 1);
      mbar_init(&s.conv[i], 128);
      mbar_init(&s.empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s.acc_full[i], 1);
      mbar_init(&s.acc_empty[i], 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // TMEM: two accumulator buffers of BN fp32 columns each (a power of two >= 32 in total)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s.tmem_base)), "n"(2 * BN) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s.tmem_base;

  if (nkb > 0) {
    if (warp == 0) {
      // ===== TMA producer =====
      if (lane == 0) {
        for (int i = 0; i < nkb; ++i) {
          const int st = i % STAGES, it = i / STAGES;
          mbar_wait(&s.empty[st], (it & 1) ^ 1);
          mbar_expect_tx(&s.full[st], (BM + BN) * BK * 4);
          const int kb = kb_begin + i;
          const int k0 = kb * BK;
          if (g.mode == 1 || g.mode == 2) {
            // A tile = 128 consecutive pixels of the NHWC tensor, shifted by the filter tap of this K block
            const int hw = g.cH * g.cW;
            const int b0 = m0 / hw, h0 = (m0 - b0 * hw) / g.cW;
            const int per_tap = g.cA / BK;
            const int tap = kb / per_tap, cb = kb - tap * per_tap;
            int dh = tap / 5 - 2, dw = tap % 5 - 2;
            if (g.mode == 2) { dh = -dh; dw = -dw; }
            tma_load_4d(&s.a[st][0][0], &map_a, &s.full[st], cb * BK, dw, h0 + dh, b0);
            if (g.mode == 1) {
              tma_load_2d(&s.b[st][0][0], &map_b, &s.full[st], k0, n0);
            } else {
              // W[co, (tap, cin)] as an MN-major operand: rows = this block of 32 output channels, columns = tap*Cin + n
              for (int bx = 0; bx < BN / 32; ++bx)
                tma_load_2d(&s.b[st][0][bx * 32 * BK], &map_b, &s.full[st], tap * g.cIn + n0 + 32 * bx, cb * BK);
            }
          } else if (g.mode == 3) {
            for (int bx = 0; bx < BM / 32; ++bx) tma_load_2d(&s.a[st][0][bx * 32 * BK], &map_a, &s.full[st], m0 + 32 * bx, k0);
            // B boxes = 32 consecutive pixels (the K block) x 32 channels of one tap, straight from X
            const int hw = g.cH * g.cW;
            const int b0 = k0 / hw, h0 = (k0 - b0 * hw) / g.cW;
            for (int bx = 0; bx < BN / 32; ++bx) {
              const int col = n0 + 32 * bx;
              const int tap = col / g.cIn, c0 = col - tap * g.cIn;
              if (tap < 25) tma_load_4d(&s.b[st][0][bx * 32 * BK], &map_b, &s.full[st], c0, tap % 5 - 2, h0 + tap / 5 - 2, b0);
              else tma_load_4d(&s.b[st][0][bx * 32 * BK], &map_b, &s.full[st], g.cIn, 0, 0, 0);   // past the last tap: zeros
            }
          } else {
            if (!g.a_mn) {
              tma_load_2d(&s.a[st][0][0], &map_a, &s.full[st], k0, m0);
            } else {
              for (int bx = 0; bx < BM / 32; ++bx) tma_load_2d(&s.a[st][0][bx * 32 * BK], &map_a, &s.full[st], m0 + 32 * bx, k0);
            }
            if (!g.b_mn) {
              tma_load_2d(&s.b[st][0][0], &map_b, &s.full[st], k0, n0);
            } else {
              for (int bx = 0; bx < BN / 32; ++bx) tma_load_2d(&s.b[st][0][bx * 32 * BK], &map_b, &s.full[st], n0 + 32 * bx, k0);
            }
          }
        }
      }
    } else if (warp == 1) {
      // ===== MMA issuer =====
      // instruction descriptor: D = fp32, A = B = TF32, both K-major, N = BN, M = 128
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(BN >> 3) << 17) |
                             (static_cast<uint32_t>(BM >> 4) << 24) | (g.a_mn ? (1u << 15) : 0u) | (g.b_mn ? (1u << 16) : 0u);
      // Everything below is computed by all 32 lanes with warp-uniform values, so descriptors live in uniform
      // registers; only the tcgen05 instructions themselves are issued by one elected lane.  (Computing them
      // under `if (lane == 0)` costs a register->uniform-register move per operand per MMA and made the issue
      // loop, not the tensor core, the bottleneck.)
      const uint32_t a_base = smem_u32(&s.a[0][0][0]), b_base = smem_u32(&s.b[0][0][0]);
      constexpr uint32_t A_TILE = BM * BK * 4, B_TILE = BN * BK * 4;   // bytes of one hi (or lo) tile
      const uint64_t a_step = g.a_mn ? (1024 >> 4) : ((UMMA_K * 4) >> 4);  // one K step (8 elements) in 16-byte units
      const uint64_t b_step = g.b_mn ? (1024 >> 4) : ((UMMA_K * 4) >> 4);
      const bool three = g.terms == 3;
      int st = 0, it = 0;
      for (int i = 0; i < nkb; ++i) {
        const int chunk = i / CHUNK, buf = chunk & 1, use = chunk >> 1;
        const bool first_in_chunk = (i % CHUNK) == 0, last_in_chunk = (i % CHUNK) == CHUNK - 1 || i == nkb - 1;
        if (first_in_chunk) mbar_wait(&s.acc_empty[buf], (use & 1) ^ 1);   // drain warps are done with this buffer
        mbar_wait(&s.conv[st], it & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t acc = tmem + static_cast<uint32_t>(buf * BN);
        const uint32_t pa_hi = a_base + static_cast<uint32_t>(st) * 2 * A_TILE, pa_lo = pa_hi + A_TILE;
        const uint32_t pb_hi = b_base + static_cast<uint32_t>(st) * 2 * B_TILE, pb_lo = pb_hi + B_TILE;
        const uint64_t a_hi = g.a_mn ? make_desc_mn(pa_hi) : make_desc(pa_hi), a_lo = g.a_mn ? make_desc_mn(pa_lo) : make_desc(pa_lo);
        const uint64_t b_hi = g.b_mn ? make_desc_mn(pb_hi) : make_desc(pb_hi), b_lo = g.b_mn ? make_desc_mn(pb_lo) : make_desc(pb_lo);
        const uint32_t keep0 = first_in_chunk ? 0u : 1u;                    // a new chunk overwrites its buffer
        if (elect_one()) {
          if (three) {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              // small terms first, the dominant product last
              umma_tf32(acc, a_lo + a_step * k, b_hi + b_step * k, idesc, k == 0 ? keep0 : 1u);
              umma_tf32(acc, a_hi + a_step * k, b_lo + b_step * k, idesc, 1u);
              umma_tf32(acc, a_hi + a_step * k, b_hi + b_step * k, idesc, 1u);
            }
          } else {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32(acc, a_hi + a_step * k, b_hi + b_step * k, idesc, k == 0 ? keep0 : 1u);
          }
          umma_commit(&s.empty[st]);                          // stage reusable once these MMAs have read it
          if (last_in_chunk) umma_commit(&s.acc_full[buf]);   // chunk complete: hand the buffer to the drain warps
        }
        __syncwarp();
        if (++st == STAGES) { st = 0; ++it; }
      }
    } else if (warp < 6) {
      // ===== converters (warps 2..5, 128 threads) =====
      const int t = threadIdx.x - 64;
      const uint32_t a_base = smem_u32(&s.a[0][0][0]), b_base = smem_u32(&s.b[0][0][0]);
      constexpr uint32_t A_TILE = BM * BK * 4, B_TILE = BN * BK * 4;
      const bool three = g.terms == 3;
      int st = 0, it = 0;
      for (int i = 0; i < nkb; ++i) {
        mbar_wait(&s.full[st], it & 1);
        const uint32_t a_hi = a_base + static_cast<uint32_t>(st) * 2 * A_TILE + t * 16, a_lo = a_hi + A_TILE;
        const uint32_t b_hi = b_base + static_cast<uint32_t>(st) * 2 * B_TILE + t * 16, b_lo = b_hi + B_TILE;
#pragma unroll
        for (int j = 0; j < BM * BK / 4 / 128; ++j) {
          const float4 v = lds128(a_hi + j * 2048);
          const float4 h = make_float4(tf32_round(v.x), tf32_round(v.y), tf32_round(v.z), tf32_round(v.w));
          sts128(a_hi + j * 2048, h);
          if (three) sts128(a_lo + j * 2048, make_float4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w));
        }
#pragma unroll
        for (int j = 0; j < BN * BK / 4 / 128; ++j) {
          const float4 v = lds128(b_hi + j * 2048);
          const float4 h = make_float4(tf32_round(v.x), tf32_round(v.y), tf32_round(v.z), tf32_round(v.w));
          sts128(b_hi + j * 2048, h);
          if (three) sts128(b_lo + j * 2048, make_float4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w));
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the tensor core
        mbar_arrive(&s.conv[st]);
        if (++st == STAGES) { st = 0; ++it; }
      }
    } else {
      // ===== drain + epilogue (warps 6..9): this warp owns TMEM lanes [32*(warp%4), +32) = output rows =====
      const int quarter = warp & 3;
      const int row = m0 + quarter * 32 + lane;
      float acc[BN];
#pragma unroll
      for (int j = 0; j < BN; ++j) acc[j] = 0.f;
      const int nchunks = (nkb + CHUNK - 1) / CHUNK;
      for (int c = 0; c < nchunks; ++c) {
        const int buf = c & 1, use = c >> 1;
        mbar_wait(&s.acc_full[buf], use & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
        for (int c0 = 0; c0 < BN; c0 += 32) {
          uint32_t r[32];
          const uint32_t taddr = tmem + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(buf * BN + c0);
          asm volatile(
              "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
              "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
              "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
              : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
                "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
                "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
              : "r"(taddr)
              : "memory");
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int q = 0; q < 32; ++q) acc[c0 + q] += __uint_as_float(r[q]);
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        mbar_arrive(&s.acc_empty[buf]);                      // the MMA warp may overwrite this buffer
      }
      if (row < g.M) {
        float* crow = g.C + static_cast<long long>(row) * g.ldc + n0;
        const bool add_bias = g.bias != nullptr && blockIdx.z == 0;
#pragma unroll
        for (int q = 0; q < BN; q += 4) {
          const int n = n0 + q;
          if (n + 3 < g.N && g.vec_out) {
            float4 v = make_float4(acc[q], acc[q + 1], acc[q + 2], acc[q + 3]);
            if (add_bias) {
              const float4 bv = *reinterpret_cast<const float4*>(g.bias + n);
              v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
            }
            if (g.atomic_out) {
              asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(crow + q), "f"(v.x), "f"(v.y),
                           "f"(v.z), "f"(v.w)
                           : "memory");
            } else {
              *reinterpret_cast<float4*>(crow + q) = v;
            }
          } else {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              if (n + e < g.N) {
                float v = acc[q + e];
                if (add_bias) v += g.bias[n + e];
                if (g.atomic_out) atomicAdd(crow + q + e, v);
                else crow[q + e] = v;
              }
            }
          }
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(2 * BN) : "memory");
  }
}


// tcgen05.mma with the A operand in tensor memory (lanes = rows of the tile, one 32-bit column per K element)
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ float lds32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
// 16 consecutive 32-bit columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
// "TS" variant of the kernel: the A operand never returns to shared memory.  The converter warps read each raw A
// row once from shared memory and write its TF32 hi / lo parts into TENSOR memory (tcgen05.st), from where the
// MMA takes them; only B still makes the smem -> smem round trip.  Per K block that is 144 KB of shared-memory
// traffic instead of 224 KB — and the kernel is shared-memory-bandwidth bound — plus a 4-deep ring, because a
// stage no longer holds an A-lo tile.
constexpr int TS_STAGES = 4;
constexpr int TS_THREADS = 576;   // producer, MMA, 8 converter warps, 8 drain warps
template <int BN>
struct SmemTS {
  float a[TS_STAGES][BM * BK];        // raw fp32 A tile (TMA destination)
  float b[TS_STAGES][2][BN * BK];     // [0] = hi (TMA destination), [1] = lo
  uint64_t full[TS_STAGES], conv[TS_STAGES], empty[TS_STAGES], acc_full[2], acc_empty[2];
  uint32_t tmem_base;
};

template <int BN>
__global__ void __launch_bounds__(TS_THREADS, 1)
tc_gemm_ts_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, GemmArgs g) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  SmemTS<BN>& s = *reinterpret_cast<SmemTS<BN>*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int total_kb = (g.K + BK - 1) / BK;
  const int kb_begin = blockIdx.z * g.kblocks_per_split;
  int kb_end = kb_begin + g.kblocks_per_split;
  if (kb_end > total_kb) kb_end = total_kb;
  const int nkb = kb_end - kb_begin;

  if (threadIdx.x == 0) {
    for (int i = 0; i < TS_STAGES; ++i) {
      mbar_init(&s.full[i], 1);
      mbar_init(&s.conv[i], 256);
      mbar_init(&s.empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s.acc_full[i], 1);
      mbar_init(&s.acc_empty[i], 256);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // TMEM: two accumulator buffers of BN columns + TS_STAGES x (hi, lo) A tiles of 32 columns: all 512 columns
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s.tmem_base)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s.tmem_base;

  if (nkb > 0) {
    if (warp == 0) {
      // ===== TMA producer =====
      if (lane == 0) {
        for (int i = 0; i < nkb; ++i) {
          const int st = i % TS_STAGES, it = i / TS_STAGES;
          mbar_wait(&s.empty[st], (it & 1) ^ 1);
          mbar_expect_tx(&s.full[st], (BM + BN) * BK * 4);
          const int kb = kb_begin + i;
          const int k0 = kb * BK;
          if (g.mode == 1 || g.mode == 2) {
            // A tile = 128 consecutive pixels of the NHWC tensor, shifted by the filter tap of this K block
            const int hw = g.cH * g.cW;
            const int b0 = m0 / hw, h0 = (m0 - b0 * hw) / g.cW;
            const int per_tap = g.cA / BK;
            const int tap = kb / per_tap, cb = kb - tap * per_tap;
            int dh = tap / 5 - 2, dw = tap % 5 - 2;
            if (g.mode == 2) { dh = -dh; dw = -dw; }
            tma_load_4d(&s.a[st][0], &map_a, &s.full[st], cb * BK, dw, h0 + dh, b0);
            if (g.mode == 1) {
              tma_load_2d(&s.b[st][0][0], &map_b, &s.full[st], k0, n0);
            } else {
              // W[co, (tap, cin)] as an MN-major operand: rows = this block of 32 output channels, columns = tap*Cin + n
              for (int bx = 0; bx < BN / 32; ++bx)
                tma_load_2d(&s.b[st][0][bx * 32 * BK], &map_b, &s.full[st], tap * g.cIn + n0 + 32 * bx, cb * BK);
            }
          } else if (g.mode == 3) {
            for (int bx = 0; bx < BM / 32; ++bx) tma_load_2d(&s.a[st][bx * 32 * BK], &map_a, &s.full[st], m0 + 32 * bx, k0);
            // B boxes = 32 consecutive pixels (the K block) x 32 channels of one tap, straight from X
            const int hw = g.cH * g.cW;
            const int b0 = k0 / hw, h0 = (k0 - b0 * hw) / g.cW;
            for (int bx = 0; bx < BN / 32; ++bx) {
              const int col = n0 + 32 * bx;
              const int tap = col / g.cIn, c0 = col - tap * g.cIn;
              if (tap < 25) tma_load_4d(&s.b[st][0][bx * 32 * BK], &map_b, &s.full[st], c0, tap % 5 - 2, h0 + tap / 5 - 2, b0);
              else tma_load_4d(&s.b[st][0][bx * 32 * BK], &map_b, &s.full[st], g.cIn, 0, 0, 0);   // past the last tap: zeros
            }
          } else {
            if (!g.a_mn) {
              tma_load_2d(&s.a[st][0], &map_a, &s.full[st], k0, m0);
            } else {
              for (int bx = 0; bx < BM / 32; ++bx) tma_load_2d(&s.a[st][bx * 32 * BK], &map_a, &s.full[st], m0 + 32 * bx, k0);
            }
            if (!g.b_mn) {
              tma_load_2d(&s.b[st][0][0], &map_b, &s.full[st], k0, n0);
            } else {
              for (int bx = 0; bx < BN / 32; ++bx) tma_load_2d(&s.b[st][0][bx * 32 * BK], &map_b, &s.full[st], n0 + 32 * bx, k0);
            }
          }
        }
      }
    } else if (warp == 1) {
      // ===== MMA issuer =====
      // instruction descriptor: D = fp32, A = B = TF32, both K-major, N = BN, M = 128
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(BN >> 3) << 17) |
                             (static_cast<uint32_t>(BM >> 4) << 24) | (g.b_mn ? (1u << 16) : 0u);   // A comes from TMEM (K-major)
      // Everything below is computed by all 32 lanes with warp-uniform values, so descriptors live in uniform
      // registers; only the tcgen05 instructions themselves are issued by one elected lane.  (Computing them
      // under `if (lane == 0)` costs a register->uniform-register move per operand per MMA and made the issue
      // loop, not the tensor core, the bottleneck.)
      const uint32_t b_base = smem_u32(&s.b[0][0][0]);
      constexpr uint32_t B_TILE = BN * BK * 4;                                // bytes of one hi (or lo) B tile
      const uint64_t b_step = g.b_mn ? (1024 >> 4) : ((UMMA_K * 4) >> 4);      // one K step (8 elements) in 16-byte units
      const bool three = g.terms == 3;
      int st = 0, it = 0;
      for (int i = 0; i < nkb; ++i) {
        const int chunk = i / CHUNK, buf = chunk & 1, use = chunk >> 1;
        const bool first_in_chunk = (i % CHUNK) == 0, last_in_chunk = (i % CHUNK) == CHUNK - 1 || i == nkb - 1;
        if (first_in_chunk) mbar_wait(&s.acc_empty[buf], (use & 1) ^ 1);   // drain warps are done with this buffer
        mbar_wait(&s.conv[st], it & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t acc = tmem + static_cast<uint32_t>(buf * BN);
        const uint32_t ta_hi = tmem + static_cast<uint32_t>(2 * BN + st * 64), ta_lo = ta_hi + 32;   // A(hi|lo): 128 lanes x 32 K columns
        const uint32_t pb_hi = b_base + static_cast<uint32_t>(st) * 2 * B_TILE, pb_lo = pb_hi + B_TILE;
        const uint64_t b_hi = g.b_mn ? make_desc_mn(pb_hi) : make_desc(pb_hi), b_lo = g.b_mn ? make_desc_mn(pb_lo) : make_desc(pb_lo);
        const uint32_t keep0 = first_in_chunk ? 0u : 1u;                    // a new chunk overwrites its buffer
        if (elect_one()) {
          if (three) {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              // small terms first, the dominant product last
              umma_tf32_ts(acc, ta_lo + k * UMMA_K, b_hi + b_step * k, idesc, k == 0 ? keep0 : 1u);
              umma_tf32_ts(acc, ta_hi + k * UMMA_K, b_lo + b_step * k, idesc, 1u);
              umma_tf32_ts(acc, ta_hi + k * UMMA_K, b_hi + b_step * k, idesc, 1u);
            }
          } else {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) umma_tf32_ts(acc, ta_hi + k * UMMA_K, b_hi + b_step * k, idesc, k == 0 ? keep0 : 1u);
          }
          umma_commit(&s.empty[st]);                          // stage (smem B, TMEM A, smem raw A) reusable once these MMAs are done
          if (last_in_chunk) umma_commit(&s.acc_full[buf]);   // chunk complete: hand the buffer to the drain warps
        }
        __syncwarp();
        if (++st == TS_STAGES) { st = 0; ++it; }
      }
    } else if (warp < 10) {
      // ===== converters (warps 2..9, 256 threads: two warps per TMEM lane quarter, each taking half of the K block) =====
      const int t = threadIdx.x - 64;                     // 0..255
      const uint32_t a_base = smem_u32(&s.a[0][0]), b_base = smem_u32(&s.b[0][0][0]);
      constexpr uint32_t A_TILE = BM * BK * 4, B_TILE = BN * BK * 4;
      const bool three = g.terms == 3;
      const int quarter = warp & 3;                       // the TMEM lanes this warp may touch: [32*quarter, +32)
      const int half = (warp - 2) >> 2;                   // which 16 of the 32 K values of the row this warp converts
      const int row = quarter * 32 + lane;                // = row of the A tile held by this thread
      int st = 0, it = 0;
      for (int i = 0; i < nkb; ++i) {
        mbar_wait(&s.full[st], it & 1);
        // All shared-memory loads of this K block are issued first (A: 4 x 16 bytes or 16 x 4 bytes of this thread's
        // row; B: 16 bytes per 4 KB of tile) so that their latencies overlap; conversion and stores follow.
        const uint32_t a_tile = a_base + static_cast<uint32_t>(st) * A_TILE;
        const uint32_t b_hi = b_base + static_cast<uint32_t>(st) * 2 * B_TILE + t * 16, b_lo = b_hi + B_TILE;
        constexpr int NB = BN * BK / 4 / 256;
        float4 bv[NB];
        uint32_t hi[16], lo[16];
        if (!g.a_mn) {
          // K-major tile, 128-byte rows, 16-byte chunks XOR-swizzled with (row % 8) by the TMA
          float4 av[4];
#pragma unroll
          for (int c = 0; c < 4; ++c) av[c] = lds128(a_tile + row * 128 + (((half * 4 + c) ^ (row & 7)) << 4));
#pragma unroll
          for (int j = 0; j < NB; ++j) bv[j] = lds128(b_hi + j * 4096);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const float e[4] = {av[c].x, av[c].y, av[c].z, av[c].w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float h = tf32_round(e[q]);
              hi[c * 4 + q] = __float_as_uint(h);
              lo[c * 4 + q] = __float_as_uint(e[q] - h);
            }
          }
        } else {
          // MN-major boxes (32 K-rows x 32 rows-of-A, unswizzled): element (k, row) of box row/32
          const uint32_t colp = a_tile + (row >> 5) * 4096 + (row & 31) * 4 + half * 16 * 128;
          float e[16];
#pragma unroll
          for (int k = 0; k < 16; ++k) e[k] = lds32(colp + k * 128);
#pragma unroll
          for (int j = 0; j < NB; ++j) bv[j] = lds128(b_hi + j * 4096);
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            const float h = tf32_round(e[k]);
            hi[k] = __float_as_uint(h);
            lo[k] = __float_as_uint(e[k] - h);
          }
        }
        // ---- A -> TMEM lane `row`: columns [16*half, +16) of the stage's hi block, same of its lo block (+32)
        const uint32_t ta = tmem + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(2 * BN + st * 64 + half * 16);
        tmem_st16(ta, hi);
        if (three) tmem_st16(ta + 32, lo);
        // ---- B: hi in place, lo into the sibling tile
#pragma unroll
        for (int j = 0; j < NB; ++j) {
          const float4 v = bv[j];
          const float4 h = make_float4(tf32_round(v.x), tf32_round(v.y), tf32_round(v.z), tf32_round(v.w));
          sts128(b_hi + j * 4096, h);
          if (three) sts128(b_lo + j * 4096, make_float4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w));
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");   // TMEM stores of this thread are complete
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic-proxy smem writes -> visible to the tensor core
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        mbar_arrive(&s.conv[st]);
        if (++st == TS_STAGES) { st = 0; ++it; }
      }
    } else {
      // ===== drain + epilogue (warps 10..17): warp owns TMEM lanes [32*(warp%4), +32) = output rows and one half of the columns =====
      const int quarter = warp & 3;
      const int chalf = (warp - 10) >> 2;
      constexpr int HN = BN / 2;                           // columns per drain warp
      const int row = m0 + quarter * 32 + lane;
      float acc[HN];
#pragma unroll
      for (int j = 0; j < HN; ++j) acc[j] = 0.f;
      const int nchunks = (nkb + CHUNK - 1) / CHUNK;
      for (int c = 0; c < nchunks; ++c) {
        const int buf = c & 1, use = c >> 1;
        mbar_wait(&s.acc_full[buf], use & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
        for (int c0 = 0; c0 < HN; c0 += 16) {
          uint32_t r[16];
          const uint32_t taddr = tmem + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(buf * BN + chalf * HN + c0);
          asm volatile(
              "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
              "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
              : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
              : "r"(taddr)
              : "memory");
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int q = 0; q < 16; ++q) acc[c0 + q] += __uint_as_float(r[q]);
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        mbar_arrive(&s.acc_empty[buf]);                      // the MMA warp may overwrite this buffer
      }
      if (row < g.M) {
        float* crow = g.C + static_cast<long long>(row) * g.ldc + n0 + chalf * HN;
        const bool add_bias = g.bias != nullptr && blockIdx.z == 0;
#pragma unroll
        for (int q = 0; q < HN; q += 4) {
          const int n = n0 + chalf * HN + q;
          if (n + 3 < g.N && g.vec_out) {
            float4 v = make_float4(acc[q], acc[q + 1], acc[q + 2], acc[q + 3]);
            if (add_bias) {
              const float4 bv = *reinterpret_cast<const float4*>(g.bias + n);
              v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
            }
            if (g.atomic_out) {
              asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(crow + q), "f"(v.x), "f"(v.y),
                           "f"(v.z), "f"(v.w)
                           : "memory");
            } else {
              *reinterpret_cast<float4*>(crow + q) = v;
            }
          } else {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              if (n + e < g.N) {
                float v = acc[q + e];
                if (add_bias) v += g.bias[n + e];
                if (g.atomic_out) atomicAdd(crow + q + e, v);
                else crow[q + e] = v;
              }
            }
          }
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512) : "memory");
  }
}

// ---- host: tensor maps ----------------------------------------------------------------------------
typedef CUresult (*EncodeTiled_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiled_t encode_fn() {
  static EncodeTiled_t fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiled_t>(p);
  });
  return fn;
}

int ensure_context() {
  typedef CUresult (*GetCurrent_t)(CUcontext*);
  typedef CUresult (*SetCurrent_t)(CUcontext);
  typedef CUresult (*Retain_t)(CUcontext*, CUdevice);
  static GetCurrent_t get_cur = nullptr;
  static SetCurrent_t set_cur = nullptr;
  static Retain_t retain = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuCtxGetCurrent", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      get_cur = reinterpret_cast<GetCurrent_t>(p);
    if (cudaGetDriverEntryPoint("cuCtxSetCurrent", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      set_cur = reinterpret_cast<SetCurrent_t>(p);
    if (cudaGetDriverEntryPoint("cuDevicePrimaryCtxRetain", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      retain = reinterpret_cast<Retain_t>(p);
  });
  DL_REQUIRE(get_cur && set_cur && retain, "tc_gemm: driver context entry points are not available");
  thread_local bool bound = false;
  if (bound) return DL_OK;
  CUcontext ctx = nullptr;
  if (get_cur(&ctx) == CUDA_SUCCESS && ctx != nullptr) {
    bound = true;
    return DL_OK;
  }
  int dev = 0;
  DL_CUDA(cudaGetDevice(&dev));
  DL_REQUIRE(retain(&ctx, dev) == CUDA_SUCCESS && ctx != nullptr, "tc_gemm: cannot retain the primary context of device %d", dev);
  DL_REQUIRE(set_cur(ctx) == CUDA_SUCCESS, "tc_gemm: cannot bind the primary context of device %d to this thread", dev);
  bound = true;
  return DL_OK;
}

// 2-D fp32 row-major matrix [rows, cols] with row pitch ld (elements); box = box_rows x 32 columns, 128B swizzle
// swz: 0 = 128B swizzle (K-major UMMA operand), 1 = 128B swizzle in 32-byte atoms (MN-major TF32 UMMA operand),
//      2 = none (tile is consumed by threads, not by the tensor core)
int make_map(CUtensorMap* map, const float* ptr, long long rows, long long cols, long long ld, int box_rows, int swz) {
  DL_REQUIRE(reinterpret_cast<uintptr_t>(ptr) % 16 == 0 && ld % 4 == 0, "tc_gemm: operands must be 16-byte aligned with a row pitch that is a multiple of 4 floats");
  EncodeTiled_t enc = encode_fn();
  DL_REQUIRE(enc != nullptr, "tc_gemm: cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 4};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(BK), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swz == 1 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : swz == 2 ? CU_TENSOR_MAP_SWIZZLE_NONE : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DL_REQUIRE(r == CUDA_SUCCESS, "tc_gemm: cuTensorMapEncodeTiled failed with CUresult %d (ptr %p rows %lld cols %lld ld %lld)",
             static_cast<int>(r), static_cast<const void*>(ptr), rows, cols, ld);
  return DL_OK;
}

// 4-D map over an NHWC fp32 tensor [B, H, W, C]: box = 32 channels x (pixels) pixels laid out as whole image rows
// (and whole images when an image has fewer pixels than the box).  Out-of-range coordinates, negative ones
// included, read as zeros: that is the convolution's padding.
int make_map_nhwc(CUtensorMap* map, const float* ptr, int B, int H, int W, int C, int pixels, bool mn_major) {
  EncodeTiled_t enc = encode_fn();
  DL_REQUIRE(enc != nullptr, "tc_conv: cuTensorMapEncodeTiled is not available from this driver");
  DL_REQUIRE(reinterpret_cast<uintptr_t>(ptr) % 16 == 0 && C % 4 == 0, "tc_conv: NHWC tensors must be 16-byte aligned with C %% 4 == 0");
  DL_REQUIRE(pixels % W == 0, "tc_conv: image width %d must divide the %d-pixel box", W, pixels);
  const int rows = pixels / W;
  const int bh = rows < H ? rows : H;
  DL_REQUIRE(H % bh == 0 && rows % bh == 0, "tc_conv: a %d-pixel box does not tile %dx%d images", pixels, H, W);
  const int bb = rows / bh;
  cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H), static_cast<cuuint64_t>(B)};
  cuuint64_t strides[3] = {static_cast<cuuint64_t>(C) * 4, static_cast<cuuint64_t>(W) * C * 4, static_cast<cuuint64_t>(H) * W * C * 4};
  cuuint32_t box[4] = {static_cast<cuuint32_t>(BK), static_cast<cuuint32_t>(W), static_cast<cuuint32_t>(bh), static_cast<cuuint32_t>(bb)};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DL_REQUIRE(r == CUDA_SUCCESS, "tc_conv: cuTensorMapEncodeTiled (NHWC %dx%dx%dx%d, box %d px) failed with CUresult %d", B, H, W, C,
             pixels, static_cast<int>(r));
  return DL_OK;
}

// A-in-TMEM ("TS") kernel unless DL_B200_TC_TS=0
bool use_ts() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("DL_B200_TC_TS");
    v = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
}

#include "tc_gemm_streamk.inl"

template <int BN>
int launch(const CUtensorMap& ma, const CUtensorMap& mb, const GemmArgs& g, int splits, cudaStream_t st) {
  static bool configured[2][64] = {{false}};
  int dev = 0;
  DL_CUDA(cudaGetDevice(&dev));
  const bool ts = use_ts();
  const size_t smem = (ts ? sizeof(SmemTS<BN>) : sizeof(Smem<BN>)) + 1024;
  if (dev >= 0 && dev < 64 && !configured[ts][dev]) {
    if (ts) DL_CUDA(cudaFuncSetAttribute(tc_gemm_ts_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    else DL_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    configured[ts][dev] = true;
  }
  dim3 grid((g.M + BM - 1) / BM, (g.N + BN - 1) / BN, splits);
  if (ts) tc_gemm_ts_kernel<BN><<<grid, TS_THREADS, smem, st>>>(ma, mb, g);
  else tc_gemm_kernel<BN><<<grid, NUM_THREADS, smem, st>>>(ma, mb, g);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}

}  // namespace

// Switch the experimental stream-K kernel on (1) or off (0) for later calls; returns the previous setting.
extern "C" int dl_tc_gemm_streamk(int on) {
  const int before = use_streamk() ? 1 : 0;
  g_streamk = on ? 1 : 0;
  return before;
}

// The stream-K work list of CTA `cta` for an M x N x K problem on `sms` SMs, computed by the same functions the kernel
// uses (host-side check of the decomposition; no GPU needed).  Writes up to `max_segments` rows of
// (m0, n0, first K block, K blocks) into `segments`; returns the number of segments of that CTA, *grid = CTAs launched.
extern "C" int dl_tc_streamk_plan(int M, int N, int K, int sms, int cta, int* grid, int* segments, int max_segments) {
  DL_REQUIRE(M > 0 && N > 0 && K > 0 && sms > 0 && grid != nullptr, "dl_tc_streamk_plan: bad arguments");
  const int bn = N <= 64 ? 64 : 128;
  long long gsz = 0;
  const SkArgs k = sk_plan(M, N, K, bn, sms, &gsz);
  *grid = static_cast<int>(gsz);
  DL_REQUIRE(cta >= 0 && cta < gsz, "dl_tc_streamk_plan: CTA %d outside the grid of %lld", cta, gsz);
  const long long u_begin = k.units * cta / gsz, u_end = k.units * (cta + 1) / gsz;
  int n = 0;
  for (long long u = u_begin; u < u_end;) {
    const SkSegment sg = sk_segment(k, u, u_end, bn);
    if (segments != nullptr && n < max_segments) {
      segments[4 * n + 0] = sg.m0; segments[4 * n + 1] = sg.n0; segments[4 * n + 2] = sg.kb0; segments[4 * n + 3] = sg.nkb;
    }
    ++n;
    u += sg.nkb;
  }
  return n;
}

extern "C" int dl_tc_gemm_ex(const float* A, const float* B, float* C, const float* bias, int M, int N, int K, long long lda,
                             long long ldb, long long ldc, int a_mn, int b_mn, int split_k, int accumulate, int terms,
                             void* stream) {
  DL_REQUIRE(A && B && C, "dl_tc_gemm: null pointer");
  DL_REQUIRE(M > 0 && N > 0 && K > 0, "dl_tc_gemm: empty problem %d x %d x %d", M, N, K);
  DL_REQUIRE(terms == 1 || terms == 3, "dl_tc_gemm: terms must be 1 (TF32) or 3 (3xTF32, fp32-accurate)");
  DL_REQUIRE(reinterpret_cast<uintptr_t>(C) % 4 == 0, "dl_tc_gemm: C must be 4-byte aligned");
  const int total_kb = (K + BK - 1) / BK;
  if (split_k < 1) split_k = 1;
  if (split_k > total_kb) split_k = total_kb;
  GemmArgs g;
  g.C = C; g.bias = bias; g.M = M; g.N = N; g.K = K; g.ldc = ldc;
  g.kblocks_per_split = (total_kb + split_k - 1) / split_k;
  const int splits = (total_kb + g.kblocks_per_split - 1) / g.kblocks_per_split;
  g.atomic_out = (splits > 1 || accumulate) ? 1 : 0;
  g.terms = terms;
  g.a_mn = a_mn ? 1 : 0;
  g.b_mn = b_mn ? 1 : 0;
  g.mode = 0; g.cH = g.cW = g.cA = g.cIn = 0; g.b_split = 0;
  g.vec_out = (reinterpret_cast<uintptr_t>(C) % 16 == 0 && ldc % 4 == 0 &&
               (bias == nullptr || reinterpret_cast<uintptr_t>(bias) % 16 == 0)) ? 1 : 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int bn = N <= 64 ? 64 : 128;
  // The tensor-map encoder is a raw driver entry point: it needs a context current on THIS thread (autograd
  // runs backward on worker threads that may only have used the runtime API lazily).  Binding the primary
  // context through the driver API is legal during stream capture, unlike cudaFree(0).
  {
    const int rc_ctx = ensure_context();
    if (rc_ctx != DL_OK) return rc_ctx;
  }
  CUtensorMap ma, mb;
  // K-major: [MN rows, K cols], box = (tile rows) x 32 K;  MN-major: [K rows, MN cols], box = 32 K rows x 32 MN
  // an MN-major A goes to TMEM through the converter threads in the TS kernel (no swizzle); the smem-operand kernel needs the UMMA layout
  int rc = a_mn ? make_map(&ma, A, K, M, lda, BK, use_ts() ? 2 : 1) : make_map(&ma, A, M, K, lda, BM, 0);
  if (rc != DL_OK) return rc;
  rc = b_mn ? make_map(&mb, B, K, N, ldb, BK, 1) : make_map(&mb, B, N, K, ldb, bn, 0);
  if (rc != DL_OK) return rc;
  if (use_ts() && use_streamk()) return bn == 64 ? launch_sk<64>(ma, mb, mb, g, accumulate, st) : launch_sk<128>(ma, mb, mb, g, accumulate, st);
  if (splits > 1 && !accumulate) {
    // split-K partial sums are added with atomics: start from zero (row by row: C may be a strided view)
    if (ldc == N) {
      DL_CUDA(cudaMemsetAsync(C, 0, static_cast<size_t>(M) * N * 4, st));
    } else {
      DL_CUDA(cudaMemset2DAsync(C, static_cast<size_t>(ldc) * 4, 0, static_cast<size_t>(N) * 4, M, st));
    }
  }
  return bn == 64 ? launch<64>(ma, mb, g, splits, st) : launch<128>(ma, mb, g, splits, st);
}

extern "C" int dl_tc_gemm(const float* A, const float* B, float* C, const float* bias, int M, int N, int K, long long lda,
                          long long ldb, long long ldc, int split_k, int accumulate, int terms, void* stream) {
  return dl_tc_gemm_ex(A, B, C, bias, M, N, K, lda, ldb, ldc, 0, 0, split_k, accumulate, terms, stream);
}

// hi = w rounded to TF32 (the converter warps' rounding), lo = w - hi: the two B operands of the 3xTF32 product, made
// once per step and layer instead of once per K block and CTA.  n % 4 == 0, 16-byte aligned pointers.
namespace {
__global__ void split_tf32_kernel(const float4* __restrict__ w, float4* __restrict__ hi, float4* __restrict__ lo, long long n4) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 v = w[i];
    const float4 h = make_float4(tf32_round(v.x), tf32_round(v.y), tf32_round(v.z), tf32_round(v.w));
    hi[i] = h;
    lo[i] = make_float4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w);
  }
}
}  // namespace

extern "C" int dl_split_tf32(const float* w, float* hi, float* lo, long long n, void* stream) {
  DL_REQUIRE(w && hi && lo && n > 0 && n % 4 == 0, "dl_split_tf32: bad arguments (n must be a positive multiple of 4)");
  DL_REQUIRE((reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(hi) | reinterpret_cast<uintptr_t>(lo)) % 16 == 0,
             "dl_split_tf32: pointers must be 16-byte aligned");
  const long long n4 = n / 4;
  long long blocks = (n4 + 255) / 256;
  if (blocks > 1184) blocks = 1184;
  split_tf32_kernel<<<static_cast<unsigned>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float4*>(w), reinterpret_cast<float4*>(hi), reinterpret_cast<float4*>(lo), n4);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}

// Implicit-GEMM 5x5 / stride 1 / pad 2 convolution pieces on NHWC tensors (see GemmArgs::mode).
//   mode 1: act = x  [B,H,W,Cin],  w = W[Cout, 25*Cin], out = y  [B*H*W, Cout] (+ bias)
//   mode 2: act = dy [B,H,W,Cout], w = W,               out = dx [B*H*W, Cin]
//   mode 3: act = x  [B,H,W,Cin],  w = dy [B*H*W, Cout] (the other operand), out = dW [Cout, 25*Cin]
//
// dl_tc_conv5x5_ws: the same with the weights ALSO given as their TF32 split (w_hi = w rounded to TF32, w_lo = w - w_hi,
// both [Cout, 25*Cin], made by dl_split_tf32).  Modes 1 and 2 of the 3xTF32 stream-K kernel then load both halves by
// TMA and the converter warps only handle the activations: per K block 48 KB less shared-memory traffic (B read, B hi
// and lo written back) for 16 KB more TMA — the kernel is bound by the shared-memory port (profiles/ncu_gemm_r2.md).
// Same products, same order, same bits.  Anything else (mode 3, terms 1, the one-tile-per-CTA kernels) ignores the split.
extern "C" int dl_tc_conv5x5_ws(int mode, const float* act, const float* w, const float* w_hi, const float* w_lo, float* out,
                                const float* bias, int B, int H, int W, int Cin, int Cout, int split_k, int accumulate, int terms,
                                void* stream);

extern "C" int dl_tc_conv5x5(int mode, const float* act, const float* w, float* out, const float* bias, int B, int H, int W,
                             int Cin, int Cout, int split_k, int accumulate, int terms, void* stream) {
  return dl_tc_conv5x5_ws(mode, act, w, nullptr, nullptr, out, bias, B, H, W, Cin, Cout, split_k, accumulate, terms, stream);
}

extern "C" int dl_tc_conv5x5_ws(int mode, const float* act, const float* w, const float* w_hi, const float* w_lo, float* out,
                                const float* bias, int B, int H, int W, int Cin, int Cout, int split_k, int accumulate, int terms,
                                void* stream) {
  DL_REQUIRE(mode >= 1 && mode <= 3, "dl_tc_conv5x5: mode must be 1 (forward), 2 (dgrad) or 3 (wgrad)");
  DL_REQUIRE((w_hi == nullptr) == (w_lo == nullptr), "dl_tc_conv5x5_ws: give both halves of the split or neither");
  DL_REQUIRE(act && w && out && B > 0 && H > 0 && W > 0, "dl_tc_conv5x5: bad arguments");
  DL_REQUIRE(Cin % BK == 0 && Cout % BK == 0, "dl_tc_conv5x5: Cin (%d) and Cout (%d) must be multiples of %d", Cin, Cout, BK);
  DL_REQUIRE(terms == 1 || terms == 3, "dl_tc_conv5x5: terms must be 1 or 3");
  const long long pixels = static_cast<long long>(B) * H * W;
  const int kc = 25 * Cin;
  {
    const int rc_ctx = ensure_context();
    if (rc_ctx != DL_OK) return rc_ctx;
  }
  GemmArgs g;
  g.C = out; g.bias = mode == 1 ? bias : nullptr; g.terms = terms;
  g.mode = mode; g.cH = H; g.cW = W; g.cIn = Cin;
  g.a_mn = mode == 3; g.b_mn = mode != 1;
  g.b_split = (w_hi != nullptr && mode != 3 && terms == 3 && use_ts() && use_streamk()) ? 1 : 0;
  CUtensorMap ma, mb, mb2;
  int rc;
  if (mode == 1) {
    g.M = static_cast<int>(pixels); g.N = Cout; g.K = kc; g.cA = Cin;
    rc = make_map_nhwc(&ma, act, B, H, W, Cin, BM, false);
    if (rc != DL_OK) return rc;
  } else if (mode == 2) {
    g.M = static_cast<int>(pixels); g.N = Cin; g.K = 25 * Cout; g.cA = Cout;
    rc = make_map_nhwc(&ma, act, B, H, W, Cout, BM, false);
    if (rc != DL_OK) return rc;
  } else {
    g.M = Cout; g.N = kc; g.K = static_cast<int>(pixels); g.cA = 0;
    DL_REQUIRE(pixels % BK == 0, "dl_tc_conv5x5: wgrad needs B*H*W (%lld) to be a multiple of %d", pixels, BK);
    rc = make_map(&ma, w, pixels, Cout, Cout, BK, use_ts() ? 2 : 1);    // dY [pixels, Cout] as MN-major A
    if (rc != DL_OK) return rc;
  }
  const int bn = g.N <= 64 ? 64 : 128;
  const float* wb = g.b_split ? w_hi : w;
  if (mode == 1) rc = make_map(&mb, wb, Cout, kc, kc, bn, 0);            // W K-major
  else if (mode == 2) rc = make_map(&mb, wb, Cout, kc, kc, BK, 1);       // W MN-major: rows = cout (K), cols = (tap, cin)
  else rc = make_map_nhwc(&mb, act, B, H, W, Cin, BK, true);             // X as MN-major B, 32-pixel boxes
  if (rc != DL_OK) return rc;
  mb2 = mb;
  if (g.b_split) {
    rc = mode == 1 ? make_map(&mb2, w_lo, Cout, kc, kc, bn, 0) : make_map(&mb2, w_lo, Cout, kc, kc, BK, 1);
    if (rc != DL_OK) return rc;
  }
  g.ldc = g.N;
  const int total_kb = (g.K + BK - 1) / BK;
  if (split_k < 1) split_k = 1;
  if (split_k > total_kb) split_k = total_kb;
  g.kblocks_per_split = (total_kb + split_k - 1) / split_k;
  const int splits = (total_kb + g.kblocks_per_split - 1) / g.kblocks_per_split;
  g.atomic_out = (splits > 1 || accumulate) ? 1 : 0;
  g.vec_out = (reinterpret_cast<uintptr_t>(out) % 16 == 0 && g.ldc % 4 == 0 &&
               (g.bias == nullptr || reinterpret_cast<uintptr_t>(g.bias) % 16 == 0)) ? 1 : 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (use_ts() && use_streamk()) return bn == 64 ? launch_sk<64>(ma, mb, mb2, g, accumulate, st) : launch_sk<128>(ma, mb, mb2, g, accumulate, st);
  if (splits > 1 && !accumulate) DL_CUDA(cudaMemsetAsync(out, 0, static_cast<size_t>(g.M) * g.N * 4, st));
  return bn == 64 ? launch<64>(ma, mb, g, splits, st) : launch<128>(ma, mb, g, splits, st);
}
