// Memory-model litmus tests for the signalling protocol, on the real fabric (SURVEY §5 "race detection": the
// substitute for compute-sanitizer's racecheck, which cannot see across GPUs anyway and is closed on this pool).
//
// Every collective here rests on two patterns across GPUs:
//
//   MP (message passing)  writer: data stores into the peer's heap ; fence.release.sys ; flag store
//                         reader: poll the flag (relaxed)          ; fence.acquire.sys ; data loads
//                         forbidden: the reader sees the flag of iteration i and data older than i.
//                         This is dl_block_barrier<true> (departure barriers) and the LL packets' ordering.
//   SB (store buffering)  A: store X (in B's heap) ; fence.sc.sys ; load Y (own heap)
//                         B: store Y (in A's heap) ; fence.sc.sys ; load X (own heap)
//                         forbidden: both load the old value.  Nothing in the library needs SB to hold, but it is the
//                         shape that tells whether the fences really are system-scope fences on this fabric.
//
// Block b of rank r plays against block b of its ring neighbours, every iteration starts after a random
// __nanosleep skew (per rank, block and iteration), and `fenced = 0` runs the same shapes WITHOUT the fences as a
// control: SB then shows the forbidden outcome readily, which is what proves the test can see a reordering at all.
#include "dl_common.cuh"
#include "dl_host.h"

#define LIT_WORDS 32  // data words per message (one per lane)

struct LitArgs {
  uint64_t off;         // scratch offset in every heap
  int shape;            // 0 = MP, 1 = SB
  int fenced;
  int iterations;
  uint32_t seed;
};

// scratch layout per rank (bytes from off):
//   [0, 64)                      counts: u64 checked, u64 violations
//   per block b (stride LIT_BLOCK_BYTES): flag u64 @0, ack u64 @64, progress u64 @128, data[LIT_WORDS] u32 @256,
//                                          then SB arrays X[iter], R[iter] (u32 each)
__host__ __device__ inline size_t lit_block_bytes(int iterations) { return 512 + static_cast<size_t>(iterations) * 8; }

__device__ __forceinline__ uint32_t lit_hash(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

template <bool kEmu>
__global__ void __launch_bounds__(32) litmus_kernel(dl_comm_t c, LitArgs a) {
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = c.world;
  const int lane = threadIdx.x;
  const int next = (rank + 1) % world, prev = (rank + world - 1) % world;
  const size_t bb = lit_block_bytes(a.iterations);
  auto area = [&](int r) { return dl_at<char>(c, r, a.off) + 64 + bb * blockIdx.x; };
  char* mine = area(rank);
  unsigned long long checked = 0, bad = 0;
  const unsigned long long limit = c.timeout_ns ? c.timeout_ns : DL_DEFAULT_TIMEOUT_NS;
  bool alive = true;

  auto wait_ge = [&](const unsigned long long* p, unsigned long long v) {
    unsigned long long t0 = 0;
    unsigned int polls = 0;
    while (dl_ld_relaxed_sys(p) < v) {
      if ((++polls & 1023u) == 0u) {
        const unsigned long long now = dl_globaltimer();
        if (t0 == 0) t0 = now;
        if (now - t0 > limit) return false;
      }
    }
    return true;
  };

  if (a.shape == 0) {
    // MP around the ring: I write to `next`, I read what `prev` wrote to me.
    unsigned long long* flag_next = reinterpret_cast<unsigned long long*>(area(next));
    uint32_t* data_next = reinterpret_cast<uint32_t*>(area(next) + 256);
    unsigned long long* ack_prev = reinterpret_cast<unsigned long long*>(area(prev) + 64);
    const unsigned long long* flag_mine = reinterpret_cast<const unsigned long long*>(mine);
    const unsigned long long* ack_mine = reinterpret_cast<const unsigned long long*>(mine + 64);
    const volatile uint32_t* data_mine = reinterpret_cast<const volatile uint32_t*>(mine + 256);
    for (int i = 1; i <= a.iterations && alive; ++i) {
      __nanosleep(lit_hash(a.seed ^ (rank * 7919u) ^ (blockIdx.x * 104729u) ^ static_cast<uint32_t>(i)) & 2047u);
      // writer half: the previous message must have been consumed (ack) before its data is overwritten
      if (lane == 0) alive = wait_ge(ack_mine, static_cast<unsigned long long>(i - 1));
      alive = __shfl_sync(0xffffffffu, alive, 0);
      if (!alive) break;
      data_next[lane] = static_cast<uint32_t>(i) * 64u + lane;          // plain (weak) stores into the peer's heap
      __syncwarp();
      if (lane == 0) {
        if (a.fenced) dl_fence_release_sys();
        dl_st_relaxed_sys(flag_next, static_cast<unsigned long long>(i));
      }
      // reader half
      if (lane == 0) alive = wait_ge(flag_mine, static_cast<unsigned long long>(i));
      alive = __shfl_sync(0xffffffffu, alive, 0);
      if (!alive) break;
      if (a.fenced) dl_fence_acquire_sys();
      __syncwarp();
      const uint32_t got = data_mine[lane];
      if (got != static_cast<uint32_t>(i) * 64u + lane) ++bad;
      ++checked;
      __syncwarp();
      if (lane == 0) {
        dl_fence_release_sys();   // my loads are done before the writer may overwrite
        dl_st_relaxed_sys(ack_prev, static_cast<unsigned long long>(i));
      }
    }
  } else {
    // SB between ring neighbours: I store X_i into next's heap and load what prev stored into mine.
    // Results: R[i] = value I loaded; the host pairs R_r[i] (loaded from prev's store) with R_prev[i]... for a 2-cycle
    // (world == 2) that is the classic two-thread SB; for longer rings the forbidden outcome is "every rank of the
    // ring loaded 0 in iteration i", checked on the host.
    uint32_t* x_next = reinterpret_cast<uint32_t*>(area(next) + 512);
    const volatile uint32_t* x_mine = reinterpret_cast<const volatile uint32_t*>(mine + 512);
    uint32_t* r_mine = reinterpret_cast<uint32_t*>(mine + 512) + a.iterations;
    unsigned long long* prog_next = reinterpret_cast<unsigned long long*>(area(next) + 128);
    const unsigned long long* prog_mine = reinterpret_cast<const unsigned long long*>(mine + 128);
    for (int i0 = 0; i0 < a.iterations && alive; i0 += 32) {
      // loose rendezvous every 32 iterations so that the racing accesses of both sides overlap in time
      if (lane == 0) {
        dl_st_relaxed_sys(prog_next, static_cast<unsigned long long>(i0 + 1));
        alive = wait_ge(prog_mine, static_cast<unsigned long long>(i0 + 1));
      }
      alive = __shfl_sync(0xffffffffu, alive, 0);
      if (!alive) break;
      __nanosleep(lit_hash(a.seed ^ (rank * 7919u) ^ (blockIdx.x * 104729u) ^ static_cast<uint32_t>(i0)) & 255u);
      const int i = i0 + lane;
      if (i < a.iterations) {
        *reinterpret_cast<volatile uint32_t*>(x_next + i) = 1u;
        if (a.fenced) dl_fence_sc_sys();
        r_mine[i] = x_mine[i];
        ++checked;
      }
    }
  }
  // totals
  for (int o = 16; o > 0; o >>= 1) {
    checked += __shfl_down_sync(0xffffffffu, checked, o);
    bad += __shfl_down_sync(0xffffffffu, bad, o);
  }
  if (lane == 0) {
    unsigned long long* counts = dl_at<unsigned long long>(c, rank, a.off);
    atomicAdd(counts + 0, checked);
    atomicAdd(counts + 1, bad);
    if (!alive) atomicAdd(counts + 2, 1ull);
  }
}

extern "C" size_t dl_litmus_scratch_bytes(int iterations, int nblocks) {
  return 64 + lit_block_bytes(iterations) * static_cast<size_t>(nblocks);
}

// The scratch region must be zeroed on every rank (and a host barrier passed) before the call.
extern "C" int dl_litmus(const dl_comm_t* c, int shape, int fenced, uint64_t scratch_off, int iterations, uint32_t seed,
                         int nblocks, void* stream) {
  int nb = dl_check_comm(c, 0, nblocks);
  if (nb < 0) return nb;
  DL_REQUIRE(shape == 0 || shape == 1, "dl_litmus: shape must be 0 (MP) or 1 (SB)");
  DL_REQUIRE(iterations > 0 && scratch_off % 64 == 0 && scratch_off >= DL_PAD_BYTES, "dl_litmus: bad arguments");
  DL_REQUIRE(c->world >= 2, "dl_litmus: needs at least two ranks");
  LitArgs a;
  a.off = scratch_off; a.shape = shape; a.fenced = fenced; a.iterations = iterations; a.seed = seed;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (c->emulate) return dl_launch(litmus_kernel<true>, c, nb, 32, st, *c, a);
  return dl_launch(litmus_kernel<false>, c, nb, 32, st, *c, a);
}
