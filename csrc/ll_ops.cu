// DL_VARIANT_LL — the small-message form of every collective (SURVEY K9 and the latency-bound regime of
// §2.7: the 10x10 confusion matrix of examples/cifar10.lua:203, MNIST's 42 KiB of gradients, CIFAR at the
// reference's global batch of 32).
//
// The barrier-based kernels cost two flag barriers plus a P2P read round trip: four NVLink crossings in
// series, 12-17 us on 8 GPUs whatever the size.  Here NOTHING waits for a flag before the data moves:
//
//   push   : every contributing rank writes its whole vector straight into a staging slot of every peer as
//            16-byte packets {data, flag, data, flag} (flag = call number on this channel, never zero);
//   reduce : every rank polls the packets of all sources in its OWN memory until both flags show this call,
//            sums them in rank order and finishes the operation locally — redundantly, but from identical inputs in
//            an identical order, so results stay bit-identical across ranks (test/test_AllReduceSGD.lua:38).
//
// One NVLink crossing on the critical path.  Why no barrier is needed:
//   * staging slots are double buffered by call parity.  A peer can only push call e+2 (same parity as e) after it
//     finished e+1, which needs MY e+1 packets, which I only send after I consumed everything of call e;
//   * flags carry the call number, so packets of call e-2 left in a slot are never mistaken for call e and
//     nothing has to be cleared;
//   * nobody writes into anybody else's operands: each rank reads only its own input and writes only its own output.
// What each rank contributes besides data (active bits, step count, learning rate — the words the arrival barrier
// carries in the other variants) travels in a header packet per source.
//
// Cost: 2x the payload on the wire, N-1 copies of it — which is why this is the path for messages up to
// ll_slot_bytes/2 (64 KiB by default) only.
#include "dl_common.cuh"
#include "dl_host.h"

#define LL_OP_ALLREDUCE 0
#define LL_OP_SGD 1
#define LL_OP_EA 2
#define LL_OP_BROADCAST 3

struct LlArgs {
  uint64_t off_a;  // allreduce: buffer | sgd: grads | ea: x        | broadcast: buffer
  uint64_t off_b;  //                   | sgd: params| ea: c (cur)  | broadcast: second destination or 0
  uint64_t off_c;  //                   | sgd: momentum or 0 | ea: c (next)
  unsigned long long nvec;  // 16-byte vectors
  float lr, momentum, wd, alpha;
  int nesterov, scale, zero_grads, root;
  uint32_t active, payload;
  int channel;
};

template <int kDtype> struct LlAdd;
template <> struct LlAdd<DL_F32> {
  __device__ static uint4 add(uint4 a, uint4 b) {
    return make_uint4(__float_as_uint(__uint_as_float(a.x) + __uint_as_float(b.x)), __float_as_uint(__uint_as_float(a.y) + __uint_as_float(b.y)),
                      __float_as_uint(__uint_as_float(a.z) + __uint_as_float(b.z)), __float_as_uint(__uint_as_float(a.w) + __uint_as_float(b.w)));
  }
  __device__ static uint4 scale(uint4 a, float s) {
    return make_uint4(__float_as_uint(__uint_as_float(a.x) * s), __float_as_uint(__uint_as_float(a.y) * s),
                      __float_as_uint(__uint_as_float(a.z) * s), __float_as_uint(__uint_as_float(a.w) * s));
  }
};
template <> struct LlAdd<DL_I32> {
  __device__ static uint4 add(uint4 a, uint4 b) { return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
  __device__ static uint4 scale(uint4 a, float) { return a; }
};
template <> struct LlAdd<DL_I64> {
  __device__ static uint4 add(uint4 a, uint4 b) {
    unsigned long long a0 = (static_cast<unsigned long long>(a.y) << 32) | a.x, a1 = (static_cast<unsigned long long>(a.w) << 32) | a.z;
    const unsigned long long b0 = (static_cast<unsigned long long>(b.y) << 32) | b.x, b1 = (static_cast<unsigned long long>(b.w) << 32) | b.z;
    a0 += b0; a1 += b1;
    return make_uint4(static_cast<unsigned>(a0), static_cast<unsigned>(a0 >> 32), static_cast<unsigned>(a1), static_cast<unsigned>(a1 >> 32));
  }
  __device__ static uint4 scale(uint4 a, float) { return a; }
};
template <> struct LlAdd<DL_F64> {
  __device__ static uint4 add(uint4 a, uint4 b) {
    double a0 = __hiloint2double(a.y, a.x), a1 = __hiloint2double(a.w, a.z);
    a0 += __hiloint2double(b.y, b.x); a1 += __hiloint2double(b.w, b.z);
    return make_uint4(__double2loint(a0), __double2hiint(a0), __double2loint(a1), __double2hiint(a1));
  }
  __device__ static uint4 scale(uint4 a, float s) {
    const double a0 = __hiloint2double(a.y, a.x) * s, a1 = __hiloint2double(a.w, a.z) * s;
    return make_uint4(__double2loint(a0), __double2hiint(a0), __double2loint(a1), __double2hiint(a1));
  }
};

// slot of source `src` in rank `dst`'s heap for (channel, parity)
__device__ __forceinline__ uint4* ll_slot(const dl_comm_t& c, int dst, int channel, uint32_t parity, int src) {
  const uint64_t index = (static_cast<uint64_t>(channel) * 2u + parity) * static_cast<uint64_t>(c.world) + static_cast<uint64_t>(src);
  return reinterpret_cast<uint4*>(reinterpret_cast<char*>(c.base[dst]) + c.ll_off + index * c.ll_slot_bytes);
}

// Poll one packet until both halves carry `flag`; false on timeout.
__device__ __forceinline__ bool ll_wait(const uint4* p, uint32_t flag, unsigned long long limit, uint4* out) {
  unsigned long long t0 = 0;
  unsigned int polls = 0;
  while (true) {
    const uint4 w = dl_ll_load(p);
    if (w.y == flag && w.w == flag) {
      *out = w;
      return true;
    }
    if ((++polls & 1023u) == 0u) {
      const unsigned long long now = dl_globaltimer();
      if (t0 == 0) t0 = now;
      if (now - t0 > limit) return false;
    }
  }
}

__device__ __forceinline__ float ll_sgd1(float g, float p, float* m, const LlArgs& a) {
  // identical, operation for operation, to sgd_update1 in fused_sgd.cu: every variant produces the same bits
  if (a.wd != 0.f) g = fmaf(a.wd, p, g);
  if (m != nullptr) {
    const float mv = fmaf(a.momentum, *m, g);
    *m = mv;
    g = a.nesterov ? fmaf(a.momentum, mv, g) : mv;
  }
  return fmaf(-a.lr, g, p);
}

template <int kOp, int kDtype, bool kEmu>
__global__ void __launch_bounds__(DL_THREADS) ll_kernel(dl_comm_t c, LlArgs a) {
  __shared__ uint32_t s_call;
  __shared__ uint32_t pay[DL_MAX_RANKS];
  __shared__ uint32_t auxs[DL_MAX_RANKS];
  __shared__ int s_ok;
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int world = c.world;
  const int act = kOp == LL_OP_BROADCAST ? 1 : (kEmu ? ((a.active >> rank) & 1u) : (a.active & 3u));
  const uint32_t user = kEmu ? a.payload + rank : a.payload;
  DlPad* mine = dl_pad(c, rank, a.channel);
  if (threadIdx.x == 0) s_call = *reinterpret_cast<volatile unsigned int*>(&mine->ll_epoch) + 1u;
  __syncthreads();
  const uint32_t call = s_call;
  const uint32_t flag = call | 0x80000000u;  // never zero: a fresh (zeroed) slot is never mistaken for a packet
  const uint32_t parity = call & 1u;
  const unsigned long long limit = c.timeout_ns ? c.timeout_ns : DL_DEFAULT_TIMEOUT_NS;
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
  const unsigned long long tid0 = static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x;

  // which ranks push data: EA moves on every rank's x (a finished node still takes part, lua/AllReduceEA.lua:58-68)
  const bool sends = kOp == LL_OP_BROADCAST ? (rank == a.root) : kOp == LL_OP_EA ? true : (act != 0);

  // ---- push -----------------------------------------------------------------------------------------
  if (blockIdx.x == 0 && threadIdx.x < world) {
    const uint32_t aux = kOp == LL_OP_SGD ? __float_as_uint(a.lr) : (a.scale != 0 ? 1u : 0u);
    dl_ll_store(ll_slot(c, threadIdx.x, a.channel, parity, rank), dl_pack_payload(act, user), aux, flag);
  }
  const uint4* in = dl_at<uint4>(c, rank, a.off_a);
  // every thread keeps the vectors it pushed: its own contribution never takes the detour through memory
  constexpr int kMaxOwn = 4;
  uint4 own[kMaxOwn];
#pragma unroll
  for (int u = 0; u < kMaxOwn; ++u) {
    const unsigned long long idx = tid0 + u * stride;
    own[u] = make_uint4(0u, 0u, 0u, 0u);
    if (idx < a.nvec && sends) {
      own[u] = in[idx];
      for (int j = 0; j < world; ++j) {
        if (j == rank) continue;
        // the two packets of vector idx sit nvec apart, not side by side: a warp's store then covers 512 contiguous bytes
        // (four full lines on the link) instead of 16 bytes out of every 32
        uint4* slot = ll_slot(c, j, a.channel, parity, rank) + 1 + idx;
        dl_ll_store(slot, own[u].x, own[u].y, flag);
        dl_ll_store(slot + a.nvec, own[u].z, own[u].w, flag);
      }
    }
  }

  // ---- headers: who is active, step counts, the hyper-parameters of the first active rank --------------
  if (threadIdx.x < 32) {
    bool ok = true;
    if (threadIdx.x < world) {
      uint4 w;
      if (threadIdx.x == rank) {
        w.x = dl_pack_payload(act, user);
        w.z = kOp == LL_OP_SGD ? __float_as_uint(a.lr) : (a.scale != 0 ? 1u : 0u);
      } else {
        ok = ll_wait(ll_slot(c, rank, a.channel, parity, threadIdx.x), flag, limit, &w);
      }
      pay[threadIdx.x] = w.x;
      auxs[threadIdx.x] = w.z;
    }
    ok = __all_sync(0xffffffffu, ok);
    if (threadIdx.x == 0) s_ok = ok ? 1 : 0;
  }
  __syncthreads();
  bool ok = s_ok != 0;
  int n_act;
  const uint32_t mask = dl_active_mask(c, rank, world, a.channel, pay, &n_act);
  const uint32_t aux0 = dl_aux_of_first_active(mask, auxs, rank);
  const uint32_t srcmask = kOp == LL_OP_BROADCAST ? (1u << a.root) : kOp == LL_OP_EA ? ((1u << world) - 1u) : mask;

  // ---- reduce + local epilogue --------------------------------------------------------------------------
#pragma unroll
  for (int u = 0; u < kMaxOwn; ++u) {
    const unsigned long long idx = tid0 + u * stride;
    if (!ok || idx >= a.nvec) continue;
    uint4 v[DL_MAX_RANKS];
    // Every source's two packets are requested before the first flag is looked at: the 2 x (N-1) loads are independent, so
    // a round of polling costs ONE trip to the L2 instead of 14 in a row (8 GPUs: that chain alone was ~8 us, which is why
    // the push path lost to the barrier kernels above 4 KiB in the first 8-GPU sweep).  Sources that have not arrived yet
    // are asked again, all of them together.
    uint32_t pending = srcmask & ~(1u << rank) & ((1u << world) - 1u);
    unsigned long long t0 = 0;
    unsigned int polls = 0;
#pragma unroll
    for (int j = 0; j < DL_MAX_RANKS; ++j)      // (unrolled, static indices: v stays in registers)
      v[j] = (j == rank && ((srcmask >> j) & 1u)) ? own[u] : make_uint4(0u, 0u, 0u, 0u);
    while (pending != 0u) {
      uint4 p0[DL_MAX_RANKS], p1[DL_MAX_RANKS];
#pragma unroll
      for (int j = 0; j < DL_MAX_RANKS; ++j) {
        if ((pending >> j) & 1u) {
          const uint4* slot = ll_slot(c, rank, a.channel, parity, j) + 1 + idx;
          p0[j] = dl_ll_load(slot);
          p1[j] = dl_ll_load(slot + a.nvec);
        }
      }
#pragma unroll
      for (int j = 0; j < DL_MAX_RANKS; ++j) {
        if (((pending >> j) & 1u) && p0[j].y == flag && p0[j].w == flag && p1[j].y == flag && p1[j].w == flag) {
          v[j] = make_uint4(p0[j].x, p0[j].z, p1[j].x, p1[j].z);
          pending &= ~(1u << j);
        }
      }
      if (pending != 0u && (++polls & 255u) == 0u) {
        const unsigned long long now = dl_globaltimer();
        if (t0 == 0) t0 = now;
        if (now - t0 > limit) { ok = false; break; }
      }
    }
    if (!ok) break;
    if (kOp == LL_OP_ALLREDUCE) {
      uint4 acc = v[0];
#pragma unroll
      for (int j = 1; j < DL_MAX_RANKS; ++j) acc = LlAdd<kDtype>::add(acc, v[j]);
      const bool sc = (aux0 & 1u) && n_act > 1;
      dl_at<uint4>(c, rank, a.off_a)[idx] = sc ? LlAdd<kDtype>::scale(acc, 1.f / static_cast<float>(n_act)) : acc;
    } else if (kOp == LL_OP_SGD) {
      float* g_loc = reinterpret_cast<float*>(dl_at<uint4>(c, rank, a.off_a) + idx);
      if (n_act > 0) {
        uint4 acc = v[0];
#pragma unroll
        for (int j = 1; j < DL_MAX_RANKS; ++j) acc = LlAdd<DL_F32>::add(acc, v[j]);
        const float inv = n_act > 1 ? 1.f / static_cast<float>(n_act) : 1.f;
        LlArgs h = a;
        h.lr = __uint_as_float(aux0);
        float* p = reinterpret_cast<float*>(dl_at<uint4>(c, rank, a.off_b) + idx);
        float* m = a.off_c ? reinterpret_cast<float*>(dl_at<uint4>(c, rank, a.off_c) + idx) : nullptr;
        const float g4[4] = {__uint_as_float(acc.x), __uint_as_float(acc.y), __uint_as_float(acc.z), __uint_as_float(acc.w)};
        float4 pv = *reinterpret_cast<float4*>(p);
        float pn[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) pn[e] = ll_sgd1(__fmul_rn(g4[e], inv), pn[e], m ? m + e : nullptr, h);
        *reinterpret_cast<float4*>(p) = make_float4(pn[0], pn[1], pn[2], pn[3]);
      }
      if (a.zero_grads) *reinterpret_cast<float4*>(g_loc) = make_float4(0.f, 0.f, 0.f, 0.f);
    } else if (kOp == LL_OP_EA) {
      const float4 cc = dl_at<float4>(c, rank, a.off_b)[idx];
      float4* cn = dl_at<float4>(c, rank, a.off_c) + idx;
      if (n_act > 0) {
        float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int j = 0; j < DL_MAX_RANKS; ++j) {
          if (j < world) {
            // same expression and order as fused_ea.cu: delta_j = (x_j - c) * alpha, summed in rank order
            sum.x += (__uint_as_float(v[j].x) - cc.x) * a.alpha; sum.y += (__uint_as_float(v[j].y) - cc.y) * a.alpha;
            sum.z += (__uint_as_float(v[j].z) - cc.z) * a.alpha; sum.w += (__uint_as_float(v[j].w) - cc.w) * a.alpha;
          }
        }
        *cn = make_float4(cc.x + sum.x, cc.y + sum.y, cc.z + sum.z, cc.w + sum.w);
        float4 xv = make_float4(__uint_as_float(own[u].x), __uint_as_float(own[u].y), __uint_as_float(own[u].z), __uint_as_float(own[u].w));
        xv.x -= (xv.x - cc.x) * a.alpha; xv.y -= (xv.y - cc.y) * a.alpha;
        xv.z -= (xv.z - cc.z) * a.alpha; xv.w -= (xv.w - cc.w) * a.alpha;
        dl_at<float4>(c, rank, a.off_a)[idx] = xv;
      } else {
        *cn = cc;  // nothing moved this round: keep the double buffer coherent so the caller can flip
      }
    } else {  // broadcast
      const uint4 r = v[a.root];
      if (rank != a.root) dl_at<uint4>(c, rank, a.off_a)[idx] = r;
      if (a.off_b) dl_at<uint4>(c, rank, a.off_b)[idx] = r;
    }
  }

  // ---- completion: the last block to finish counts the call -------------------------------------------
  if (!ok && threadIdx.x == 0) {
    mine->result.error = static_cast<uint32_t>(-DL_ETIMEOUT);
    dl_host_error(c, a.channel, DL_ETIMEOUT);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // (no fence: the counter orders nothing but itself, the next call on this channel is stream-ordered behind this kernel)
    const unsigned int done = gridDim.x == 1 ? 1u : atomicAdd(&mine->ll_done, 1u) + 1u;
    if (done == gridDim.x) {
      if (gridDim.x != 1) mine->ll_done = 0u;
      *reinterpret_cast<volatile unsigned int*>(&mine->ll_epoch) = call;
    }
  }
}

static int ll_check(const dl_comm_t* c, const char* who, uint64_t nbytes, int channel, int nblocks_req, unsigned long long nvec, int* nb_out) {
  DL_REQUIRE(c->ll_off != 0 && c->ll_slot_bytes >= 48, "%s: the communicator has no small-message staging area (ll_off)", who);
  DL_REQUIRE(c->ll_off % 16 == 0 && c->ll_slot_bytes % 16 == 0, "%s: staging area must be 16-byte aligned", who);
  DL_REQUIRE(16 + 2 * nbytes <= c->ll_slot_bytes, "%s: %llu bytes exceed the small-message limit of %llu", who,
             (unsigned long long)nbytes, (unsigned long long)((c->ll_slot_bytes - 16) / 2));
  int nb = dl_check_comm(c, channel, dl_pick_blocks(nvec, 1, nblocks_req));
  if (nb < 0) return nb;
  // every vector must fit the registers of the grid (4 per thread)
  while (static_cast<unsigned long long>(nb) * DL_THREADS * 4 < nvec) {
    const int cap = c->emulate ? dl_sm_count() / c->world : (dl_sm_count() < DL_MAX_BLOCKS ? dl_sm_count() : DL_MAX_BLOCKS);
    DL_REQUIRE(nb < cap, "%s: %llu bytes exceed the capacity of the small-message kernel", who, (unsigned long long)nbytes);
    nb = nb * 2 < cap ? nb * 2 : cap;
  }
  *nb_out = nb;
  return DL_OK;
}

#define LL_LAUNCH(OP, DT)                                                                          \
  (c->emulate ? dl_launch(ll_kernel<OP, DT, true>, c, nb, DL_THREADS, st, *c, a)                  \
              : dl_launch(ll_kernel<OP, DT, false>, c, nb, DL_THREADS, st, *c, a))

// Entry points used by dl_allreduce / dl_fused_sgd / dl_fused_ea / dl_broadcast when variant == DL_VARIANT_LL.
int dl_ll_allreduce(const dl_comm_t* c, int dtype, uint64_t off, uint64_t nbytes, int active, uint32_t payload, int scale,
                    int nblocks, int channel, cudaStream_t st) {
  LlArgs a{};
  a.off_a = off; a.nvec = nbytes / 16; a.scale = scale; a.active = static_cast<uint32_t>(active); a.payload = payload; a.channel = channel;
  int nb = 0;
  const int rc = ll_check(c, "dl_allreduce", nbytes, channel, nblocks, a.nvec, &nb);
  if (rc != DL_OK) return rc;
  switch (dtype) {
    case DL_F32: return LL_LAUNCH(LL_OP_ALLREDUCE, DL_F32);
    case DL_I32: return LL_LAUNCH(LL_OP_ALLREDUCE, DL_I32);
    case DL_I64: return LL_LAUNCH(LL_OP_ALLREDUCE, DL_I64);
    default: return LL_LAUNCH(LL_OP_ALLREDUCE, DL_F64);
  }
}

int dl_ll_sgd(const dl_comm_t* c, uint64_t g_off, uint64_t p_off, uint64_t m_off, uint64_t n, float lr, float momentum, float wd,
              int nesterov, int active, uint32_t payload, int zero_grads, int nblocks, int channel, cudaStream_t st) {
  LlArgs a{};
  a.off_a = g_off; a.off_b = p_off; a.off_c = m_off; a.nvec = n / 4; a.lr = lr; a.momentum = momentum; a.wd = wd; a.nesterov = nesterov;
  a.zero_grads = zero_grads; a.active = static_cast<uint32_t>(active); a.payload = payload; a.channel = channel;
  int nb = 0;
  const int rc = ll_check(c, "dl_fused_sgd", n * 4, channel, nblocks, a.nvec, &nb);
  if (rc != DL_OK) return rc;
  return LL_LAUNCH(LL_OP_SGD, DL_F32);
}

int dl_ll_ea(const dl_comm_t* c, uint64_t x_off, uint64_t c_off, uint64_t cnew_off, uint64_t n, float alpha, int active,
             uint32_t payload, int nblocks, int channel, cudaStream_t st) {
  LlArgs a{};
  a.off_a = x_off; a.off_b = c_off; a.off_c = cnew_off; a.nvec = n / 4; a.alpha = alpha;
  a.active = static_cast<uint32_t>(active); a.payload = payload; a.channel = channel;
  int nb = 0;
  const int rc = ll_check(c, "dl_fused_ea", n * 4, channel, nblocks, a.nvec, &nb);
  if (rc != DL_OK) return rc;
  return LL_LAUNCH(LL_OP_EA, DL_F32);
}

int dl_ll_broadcast(const dl_comm_t* c, uint64_t off, uint64_t off2, uint64_t n, int root, uint32_t payload, int nblocks, int channel,
                    cudaStream_t st) {
  LlArgs a{};
  a.off_a = off; a.off_b = off2; a.nvec = n / 4; a.root = root; a.active = 1u; a.payload = payload; a.channel = channel;
  int nb = 0;
  const int rc = ll_check(c, "dl_broadcast", n * 4, channel, nblocks, a.nvec, &nb);
  if (rc != DL_OK) return rc;
  return LL_LAUNCH(LL_OP_BROADCAST, DL_F32);
}
