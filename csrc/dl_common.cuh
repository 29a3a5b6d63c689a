// distlearn_b200 — device-side primitives shared by every collective kernel.
//
// Signalling model (replaces blocking socket recv): every rank owns
// a signal pad at the start of its symmetric heap.  Block b of rank r and block b
// of every peer meet at a *block-level* barrier: r stores (epoch<<32 | payload)
// into slot [epoch&1][b][r] of every peer's pad with st.release.sys and spins on
// its own slots with ld.acquire.sys.  Epochs are monotonic device-side counters,
// so a launch needs no host-side sequence number and can sit in a CUDA graph.
// Two parities are enough: a peer cannot reach barrier e+2 before we have
// signalled e+1, which we only do after we finished reading the e slots.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/distlearn_b200.h"

#define DL_THREADS 512
#define DL_DEFAULT_TIMEOUT_NS (20ull * 1000ull * 1000ull * 1000ull)

struct DlPad {
  unsigned long long flags[2][DL_MAX_BLOCKS][DL_MAX_RANKS];  // [parity][block][source rank]
  unsigned long long aux[2][DL_MAX_BLOCKS][DL_MAX_RANKS];    // 64 extra bits published with each flag
  unsigned int epoch[DL_MAX_BLOCKS];                         // this rank's barrier count per block
  dl_result_t result;
  unsigned int reserved[22];
};
static_assert(sizeof(DlPad) % 16 == 0, "pad alignment");

// Async-EASGD control block (lives in the server's pad; clients reach it over NVLink).
#define DL_MAILBOX_WORDS 32
struct DlAsyncCtl {
  unsigned int ticket_next;     // ticket lock: next ticket to hand out      (K11)
  unsigned int ticket_serving;  // ticket currently allowed in
  unsigned long long sync_count;  // completed client syncs                  (K13)
  unsigned long long mailbox[DL_MAILBOX_WORDS];  // host-visible control words (test request/ack, ready, done)
  // client-local words (each client uses the ones in its OWN pad)
  unsigned int local_go;    // block 0 holds the lock: other blocks may start
  unsigned int local_done;  // blocks finished in the current launch
  unsigned int local_ticket;
  unsigned int reserved;
};

#define DL_ASYNC_CTL_OFFSET (DL_NUM_CHANNELS * sizeof(DlPad))
static_assert(DL_ASYNC_CTL_OFFSET + sizeof(DlAsyncCtl) <= DL_PAD_BYTES, "pad too small");

__device__ __forceinline__ DlPad* dl_pad(const dl_comm_t& c, int rank, int ch) {
  return reinterpret_cast<DlPad*>(c.base[rank]) + ch;
}
__device__ __forceinline__ DlAsyncCtl* dl_async_ctl(const dl_comm_t& c, int rank) {
  return reinterpret_cast<DlAsyncCtl*>(reinterpret_cast<char*>(c.base[rank]) + DL_ASYNC_CTL_OFFSET);
}

// ---- memory-model primitives -------------------------------------------------------------
__device__ __forceinline__ void dl_st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long dl_ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void dl_st_release_sys_u32(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int dl_ld_acquire_sys_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned int dl_ld_acquire_gpu_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void dl_st_release_gpu_u32(unsigned int* p, unsigned int v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long dl_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// ---- 128-bit data movement ---------------------------------------------------------------
// Peer addresses bypass the local L2 and are served through L1 (B300_MICROARCH "NVLink"), so
// streaming loads use no_allocate; every element is read exactly once per launch.
__device__ __forceinline__ float4 dl_ld_f4(const float4* p) {
  float4 v;
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void dl_st_f4(float4* p, const float4& v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ uint4 dl_ld_u4(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void dl_st_u4(uint4* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}
// NVLS: the switch reads the same offset on every GPU bound to the multicast object and
// returns the sum; multimem.st replicates one store to all of them.
__device__ __forceinline__ float4 dl_mc_ld_reduce_f4(const float4* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void dl_mc_st_f4(float4* p, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
// Vector reduction into (possibly remote) memory: one REDG.E.ADD.F32x4 per 16 bytes.
__device__ __forceinline__ void dl_red_add_f4_sys(float4* p, const float4& v) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

__device__ __forceinline__ float4 operator+(const float4& a, const float4& b) {
  return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}
__device__ __forceinline__ float4 dl_scale(const float4& a, float s) {
  // __fmul_rn: never fused into a following add (the result must not depend on which kernel inlined this)
  return make_float4(__fmul_rn(a.x, s), __fmul_rn(a.y, s), __fmul_rn(a.z, s), __fmul_rn(a.w, s));
}

// ---- the block-level cross-rank barrier ---------------------------------------------------
// All threads of the block must call it.  `pay_out` is a shared array of DL_MAX_RANKS words
// that receives every rank's payload.  Returns false if the bounded spin timed out (the error
// word of the local pad is then set and callers skip their data phase).
__device__ __forceinline__ bool dl_block_barrier(const dl_comm_t& c, int rank, int world, int ch,
                                                 uint32_t payload, uint32_t* pay_out,
                                                 unsigned long long aux = 0ull,
                                                 unsigned long long* aux_out = nullptr) {
  __shared__ int s_ok;
  __syncthreads();  // everything this block wrote so far is ordered before the release below
  if (threadIdx.x < 32) {
    DlPad* mine = dl_pad(c, rank, ch);
    const int lane = threadIdx.x;
    uint32_t e = 0;
    if (lane == 0) e = mine->epoch[blockIdx.x] + 1u;
    e = __shfl_sync(0xffffffffu, e, 0);
    bool ok = true;
    if (lane < world) {
      DlPad* theirs = dl_pad(c, lane, ch);
      if (aux_out != nullptr)  // ordered before the flag by the release below
        *reinterpret_cast<volatile unsigned long long*>(&theirs->aux[e & 1u][blockIdx.x][rank]) = aux;
      dl_st_release_sys(&theirs->flags[e & 1u][blockIdx.x][rank],
                        (static_cast<unsigned long long>(e) << 32) | payload);
      const unsigned long long* slot = &mine->flags[e & 1u][blockIdx.x][lane];
      const unsigned long long limit = c.timeout_ns ? c.timeout_ns : DL_DEFAULT_TIMEOUT_NS;
      unsigned long long t0 = 0;
      unsigned int polls = 0;
      unsigned long long w;
      while (true) {
        w = dl_ld_acquire_sys(slot);
        if (static_cast<uint32_t>(w >> 32) == e) break;
        if ((++polls & 1023u) == 0u) {
          const unsigned long long now = dl_globaltimer();
          if (t0 == 0) t0 = now;
          if (now - t0 > limit) {
            ok = false;
            break;
          }
        }
      }
      pay_out[lane] = static_cast<uint32_t>(w);
      if (aux_out != nullptr)
        aux_out[lane] = *reinterpret_cast<volatile unsigned long long*>(&mine->aux[e & 1u][blockIdx.x][lane]);
    }
    ok = __all_sync(0xffffffffu, ok);
    if (lane == 0) {
      mine->epoch[blockIdx.x] = e;
      if (!ok) mine->result.error = static_cast<uint32_t>(-DL_ETIMEOUT);
      s_ok = ok ? 1 : 0;
    }
  }
  __syncthreads();
  return s_ok != 0;
}

// payload convention on arrival barriers: bit 0 = active, bits 1..31 = user payload
__device__ __forceinline__ uint32_t dl_pack_payload(int active, uint32_t user) {
  // bit 0: counted in n_active AND data included; bit 1: data included but not counted (a
  // finished node that still contributes, lua/AllReduceEA.lua:58-68); bits 2..31: user payload
  return (user << 2) | (static_cast<uint32_t>(active) & 3u);
}

// Decode the arrival payloads: active mask, count, and (block 0) publish the result words.
__device__ __forceinline__ uint32_t dl_active_mask(const dl_comm_t& c, int rank, int world, int ch,
                                                   const uint32_t* pay, int* n_active) {
  uint32_t mask = 0;
  int n = 0;
#pragma unroll
  for (int j = 0; j < DL_MAX_RANKS; ++j) {
    if (j < world && (pay[j] & 3u)) mask |= 1u << j;  // data included
    if (j < world && (pay[j] & 1u)) ++n;              // counted as a contributor
  }
  *n_active = n;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    DlPad* mine = dl_pad(c, rank, ch);
    for (int j = 0; j < DL_MAX_RANKS; ++j) mine->result.payload[j] = j < world ? (pay[j] >> 2) : 0u;
    mine->result.n_active = static_cast<uint32_t>(n);
  }
  return mask;
}

// Hyper-parameters travel in the aux word so that ACTIVE ranks dictate them: a draining rank
// (lua/AllReduceSGD.lua:37) only needs to know the buffer geometry, not the current lr.
__device__ __forceinline__ unsigned long long dl_aux_of_first_active(uint32_t mask, const unsigned long long* aux, int rank) {
  const int first = mask ? (__ffs(static_cast<int>(mask)) - 1) : rank;
  return aux[first];
}

// Slice geometry: the flat vector of n4 float4 is cut into `world` equal slices; rank r owns
// [r*slice4, min((r+1)*slice4, n4)).  Every rank derives the same numbers.
struct DlSlice {
  unsigned long long slice4, begin, end;
};
__device__ __forceinline__ DlSlice dl_slice(unsigned long long n4, int rank, int world) {
  DlSlice s;
  s.slice4 = (n4 + world - 1) / world;
  s.begin = s.slice4 * static_cast<unsigned long long>(rank);
  s.end = s.begin + s.slice4;
  if (s.begin > n4) s.begin = n4;
  if (s.end > n4) s.end = n4;
  return s;
}

template <typename T>
__device__ __forceinline__ T* dl_at(const dl_comm_t& c, int rank, uint64_t off) {
  return reinterpret_cast<T*>(reinterpret_cast<char*>(c.base[rank]) + off);
}
template <typename T>
__device__ __forceinline__ T* dl_mc_at(const dl_comm_t& c, uint64_t off) {
  return reinterpret_cast<T*>(reinterpret_cast<char*>(c.mc) + off);
}
