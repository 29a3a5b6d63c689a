// distlearn_b200 — device-side primitives shared by every collective kernel.
//
// Signalling model (replaces blocking socket recv): every rank owns
// a signal pad at the start of its symmetric heap.  Block b of rank r and block b
// of every peer meet at a *block-level* barrier: r stores the packet {epoch, payload,
// epoch, aux} into slot [epoch&1][b][r] of every peer's pad (after a system-scope
// release fence when the block has data to publish) and polls its own slots, then
// fences (acquire).  Epochs are monotonic device-side counters,
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
  // [parity][block][source rank]: one 16-byte packet {epoch, payload, epoch, aux}.  The epoch rides in both 8-byte
  // halves (each half is written atomically), so payload and aux need no ordering between them.
  uint4 slots[2][DL_MAX_BLOCKS][DL_MAX_RANKS];
  unsigned int epoch[DL_MAX_BLOCKS];                         // this rank's barrier count per block
  dl_result_t result;
  unsigned int ll_epoch;                                     // completed DL_VARIANT_LL calls on this channel
  unsigned int ll_done;                                      // blocks of the running LL call that have finished
  unsigned int reserved[20];
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
__device__ __forceinline__ void dl_st_relaxed_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long dl_ld_relaxed_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void dl_fence_acquire_sys() { asm volatile("fence.acquire.sys;" ::: "memory"); }
__device__ __forceinline__ void dl_fence_release_sys() { asm volatile("fence.release.sys;" ::: "memory"); }
__device__ __forceinline__ void dl_fence_sc_sys() { asm volatile("fence.sc.sys;" ::: "memory"); }
// Sticky error word in pinned host memory (the host polls it without touching the device).
__device__ __forceinline__ void dl_host_error(const dl_comm_t& c, int channel, int code) {
  if (c.host_words != nullptr) {
    dl_host_words_t* h = static_cast<dl_host_words_t*>(c.host_words);
    if (atomicCAS_system(&h->error, 0u, static_cast<unsigned int>(-code)) == 0u) h->error_channel = static_cast<unsigned int>(channel);
  }
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
// 256-bit accesses (sm_100: LDG.E.ENL2.256 / STG.E.ENL2.256): half the memory instructions per byte on the P2P paths.
struct dl_f8 {
  float4 lo, hi;
};
__device__ __forceinline__ dl_f8 dl_ld_f8(const float4* p) {  // p must be 32-byte aligned
  dl_f8 v;
  asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(v.lo.x), "=f"(v.lo.y), "=f"(v.lo.z), "=f"(v.lo.w), "=f"(v.hi.x), "=f"(v.hi.y), "=f"(v.hi.z), "=f"(v.hi.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void dl_st_f8(float4* p, const dl_f8& v) {
  asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v.lo.x), "f"(v.lo.y),
               "f"(v.lo.z), "f"(v.lo.w), "f"(v.hi.x), "f"(v.hi.y), "f"(v.hi.z), "f"(v.hi.w)
               : "memory");
}
// LL packets: 16 bytes = {data0, flag, data1, flag}; each 8-byte half is written and read atomically, so a
// reader that sees the flag it waits for in both halves has the data (the protocol NCCL calls LL).
__device__ __forceinline__ void dl_ll_store(uint4* p, uint32_t d0, uint32_t d1, uint32_t flag) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(d0), "r"(flag), "r"(d1), "r"(flag) : "memory");
}
__device__ __forceinline__ uint4 dl_ll_load(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
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
// that receives every rank's payload (aux_out, if given, every rank's 32-bit aux word).  Returns
// false if the bounded spin timed out (the error word of the local pad and the host's sticky error
// word are then set and callers skip their data phase).
//
// kPublish = true : release pattern — everything this block wrote so far (P2P stores into peers'
//                   heaps included) is visible to whoever sees the flag (fence.release.sys = MEMBAR.SYS
//                   waits for those writes to be acknowledged, which costs an NVLink round trip).
// kPublish = false: arrival only — "my inputs are final and your stores may land in my outputs".  The inputs
//                   were written by earlier kernels on this stream, so they already sit at this GPU's point
//                   of coherence, where peer loads and the switch are served; no fence, one NVLink crossing.
// Both sides poll with relaxed loads and fence once (acquire) after the last flag arrived.
template <bool kPublish = true>
__device__ __forceinline__ bool dl_block_barrier(const dl_comm_t& c, int rank, int world, int ch,
                                                 uint32_t payload, uint32_t* pay_out,
                                                 uint32_t aux = 0u, uint32_t* aux_out = nullptr) {
  __shared__ int s_ok;
  __syncthreads();  // everything this block wrote so far is ordered before the fence below
  if (threadIdx.x < 32) {
    DlPad* mine = dl_pad(c, rank, ch);
    const int lane = threadIdx.x;
    uint32_t e = 0;
    if (lane == 0) e = mine->epoch[blockIdx.x] + 1u;
    e = __shfl_sync(0xffffffffu, e, 0);
    int err = 0;
    if (lane < world) {
      DlPad* theirs = dl_pad(c, lane, ch);
      if (kPublish) dl_fence_release_sys();
      // packet {epoch, payload, epoch, aux} in one 16-byte store
      asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(&theirs->slots[e & 1u][blockIdx.x][rank]),
                   "r"(e), "r"(payload), "r"(e), "r"(aux)
                   : "memory");
      const uint4* slot = &mine->slots[e & 1u][blockIdx.x][lane];
      const unsigned long long limit = c.timeout_ns ? c.timeout_ns : DL_DEFAULT_TIMEOUT_NS;
      unsigned long long t0 = 0;
      unsigned int polls = 0;
      uint4 w;
      while (true) {
        w = dl_ll_load(slot);
        if (w.x == e && w.z == e) break;
        if (c.flags & DL_FLAG_DEBUG) {
          // a slot of parity e&1 can only hold e-2 (its previous use; 0 before the first) or e: anything else
          // means a rank ran ahead, a launch went missing, or somebody else wrote into the pad
          const uint32_t prev = e - 2u;
          const bool okx = w.x == e || w.x == prev || (e <= 2u && w.x == 0u);
          const bool okz = w.z == e || w.z == prev || (e <= 2u && w.z == 0u);
          if (!okx || !okz) { err = DL_EPROTO; break; }
        }
        if ((++polls & 1023u) == 0u) {
          const unsigned long long now = dl_globaltimer();
          if (t0 == 0) t0 = now;
          if (now - t0 > limit) { err = DL_ETIMEOUT; break; }
        }
      }
      dl_fence_acquire_sys();
      pay_out[lane] = w.y;
      if (aux_out != nullptr) aux_out[lane] = w.w;
    }
    const bool ok = __all_sync(0xffffffffu, err == 0);
    if (err != 0) {
      mine->result.error = static_cast<uint32_t>(-err);
      dl_host_error(c, ch, err);
    }
    if (lane == 0) {
      mine->epoch[blockIdx.x] = e;
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
    if (c.host_words != nullptr && rank == c.rank) {
      // the contributor count, readable by the host without a device-to-host copy (lua/AllReduceSGD.lua:20: `_, n = allReduce`)
      dl_host_words_t* h = static_cast<dl_host_words_t*>(c.host_words);
      // ONE 8-byte store carries {ticket, n}: it cannot tear, so no fence is needed between the two words — a system-scope
      // fence here would also wait for this thread's flag / packet stores to the peers to be acknowledged (an NVLink round
      // trip on the critical path of every small collective).  The host reads the slot after the call's event anyway.
      *reinterpret_cast<volatile unsigned long long*>(&h->ring[ch][c.ticket % DL_HOST_RING]) =
          static_cast<unsigned long long>(c.ticket) | (static_cast<unsigned long long>(static_cast<uint32_t>(n)) << 32);
    }
  }
  return mask;
}

// Hyper-parameters travel in the aux word so that ACTIVE ranks dictate them: a draining rank
// (lua/AllReduceSGD.lua:37) only needs to know the buffer geometry, not the current lr.
__device__ __forceinline__ uint32_t dl_aux_of_first_active(uint32_t mask, const uint32_t* aux, int rank) {
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
