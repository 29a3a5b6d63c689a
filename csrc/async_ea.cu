// async_ea_client_sync — SURVEY K11-K14/W8.
//
// The reference's asynchronous EASGD (lua/AsyncEA.lua) keeps the center on a server PROCESS and
// moves 2x|params| per sync through TCP with a six-message handshake.  Here the center is a
// buffer in the server rank's symmetric heap and one client-side kernel does the whole sync:
//
//     c  = center[server]            (P2P loads over NVLink)        lua/AsyncEA.lua:95-106
//     d  = alpha * (x - c) ; x -= d  (local)                        lua/AsyncEA.lua:109-119
//     center[server] += d            (red.global.add.v4.f32 .sys)   lua/AsyncEA.lua:198-216
//
// with no host round trip and no kernel on the server.  `exclusive` reproduces the server's
// one-client-at-a-time mutex (recvAny, lua/AsyncEA.lua:163-177) with a ticket lock in the
// server's signal pad; the default lock-free mode lets syncs overlap (element-wise atomic adds
// keep the center exact; a client may read a center that is mid-update by another client, which
// EASGD tolerates by design).
#include "dl_common.cuh"
#include "dl_host.h"

struct AsyncArgs {
  float* x;
  float* x_out;  // where the moved parameters go: x itself, or a second buffer when x must stay intact
  uint64_t c_off;
  unsigned long long n;
  float alpha;
  int server;
  int exclusive;
};

__global__ void __launch_bounds__(DL_THREADS, 1) async_ea_client_kernel(dl_comm_t c, AsyncArgs a) {
  DlAsyncCtl* srv = dl_async_ctl(c, a.server);
  DlAsyncCtl* loc = dl_async_ctl(c, c.rank);
  __shared__ int s_ok;
  if (a.exclusive) {
    // block 0 takes the ticket and waits its turn; the other blocks wait for the local go word
    if (threadIdx.x == 0) {
      const unsigned long long limit = c.timeout_ns ? c.timeout_ns : DL_DEFAULT_TIMEOUT_NS;
      unsigned long long t0 = 0;
      unsigned int polls = 0;
      int ok = 1;
      if (blockIdx.x == 0) {
        unsigned int ticket;
        asm volatile("atom.relaxed.sys.global.add.u32 %0, [%1], 1;" : "=r"(ticket) : "l"(&srv->ticket_next) : "memory");
        // Nobody ever walks away from the queue, and nobody can block it for good.  If the ticket being served does
        // not change for a whole timeout its holder is taken for dead and ONE waiter passes that turn on (CAS); a
        // waiter that finds itself overtaken by such a recovery draws a new ticket.  A holder that was only slow then
        // shares the center with its successor for a moment — the lock-free mode's semantics, exact all the same
        // because the center is only ever changed by atomic adds.  Releases never move the counter backwards (max).
        unsigned int last = 0xffffffffu;
        int recoveries = 0;
        while (true) {
          const unsigned int cur = dl_ld_acquire_sys_u32(&srv->ticket_serving);
          if (cur == ticket) break;
          if (static_cast<int>(cur - ticket) > 0) {  // overtaken: queue again
            asm volatile("atom.relaxed.sys.global.add.u32 %0, [%1], 1;" : "=r"(ticket) : "l"(&srv->ticket_next) : "memory");
            last = 0xffffffffu;
            continue;
          }
          if (cur != last) { last = cur; t0 = 0; polls = 0; }
          if ((++polls & 1023u) == 0u) {
            const unsigned long long now = dl_globaltimer();
            if (t0 == 0) t0 = now;
            if (now - t0 > limit) {
              if (++recoveries > 8) { ok = 0; break; }   // the server's memory itself is not answering
              unsigned int seen;
              asm volatile("atom.relaxed.sys.global.cas.b32 %0, [%1], %2, %3;" : "=r"(seen) : "l"(&srv->ticket_serving), "r"(cur), "r"(cur + 1u) : "memory");
              t0 = 0;
            }
          }
        }
        loc->local_ticket = ticket;
        dl_st_release_gpu_u32(&loc->local_go, ok ? 1u : 2u);
      } else {
        unsigned int g;
        while ((g = dl_ld_acquire_gpu_u32(&loc->local_go)) == 0u) {
          if ((++polls & 1023u) == 0u) {
            const unsigned long long now = dl_globaltimer();
            if (t0 == 0) t0 = now;
            if (now - t0 > 2 * limit) { g = 2u; break; }
          }
        }
        ok = (g == 1u);
      }
      s_ok = ok;
    }
    __syncthreads();
  } else if (threadIdx.x == 0) {
    s_ok = 1;
  }
  __syncthreads();
  const bool ok = s_ok != 0;

  if (ok) {
    constexpr int U = 4;
    const unsigned long long n4 = a.n / 4;
    const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
    float4* center = dl_at<float4>(c, a.server, a.c_off);
    float4* x4 = reinterpret_cast<float4*>(a.x);
    float4* xo4 = reinterpret_cast<float4*>(a.x_out);
    const float alpha = a.alpha;
    for (unsigned long long i0 = static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x; i0 < n4;
         i0 += stride * U) {
      float4 cv[U], xv[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const unsigned long long idx = i0 + u * stride;
        if (idx < n4) {
          cv[u] = dl_ld_f4(center + idx);
          xv[u] = x4[idx];
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const unsigned long long idx = i0 + u * stride;
        if (idx < n4) {
          float4 d;
          d.x = (xv[u].x - cv[u].x) * alpha; d.y = (xv[u].y - cv[u].y) * alpha;
          d.z = (xv[u].z - cv[u].z) * alpha; d.w = (xv[u].w - cv[u].w) * alpha;
          xv[u].x -= d.x; xv[u].y -= d.y; xv[u].z -= d.z; xv[u].w -= d.w;
          xo4[idx] = xv[u];
          dl_red_add_f4_sys(center + idx, d);
        }
      }
    }
    // scalar tail
    float* cs = dl_at<float>(c, a.server, a.c_off);
    for (unsigned long long i = n4 * 4 + static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x; i < a.n;
         i += stride) {
      const float cvs = *reinterpret_cast<volatile float*>(cs + i);
      const float d = (a.x[i] - cvs) * alpha;
      a.x_out[i] = a.x[i] - d;
      asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(cs + i), "f"(d) : "memory");
    }
  }
  // completion: the last block to finish publishes the sync (and releases the lock)
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    const unsigned int done = atomicAdd(&loc->local_done, 1u) + 1u;
    if (done == gridDim.x) {
      __threadfence_system();
      if (a.exclusive && ok) {
        const unsigned int ticket = loc->local_ticket;
        unsigned int prev_serving;
        asm volatile("atom.release.sys.global.max.u32 %0, [%1], %2;" : "=r"(prev_serving) : "l"(&srv->ticket_serving), "r"(ticket + 1u) : "memory");
      }
      if (ok) {
        asm volatile("red.release.sys.global.add.u64 [%0], 1;" ::"l"(&srv->sync_count) : "memory");
      } else {
        dl_pad(c, c.rank, 0)->result.error = static_cast<uint32_t>(-DL_ETIMEOUT);
      }
      loc->local_done = 0u;
      dl_st_release_gpu_u32(&loc->local_go, 0u);
    }
  }
}

extern "C" int dl_async_ea_client(const dl_comm_t* c, int server, float* x, uint64_t c_off, uint64_t n, float alpha,
                                  int exclusive, int nblocks, void* stream) {
  return dl_async_ea_client_out(c, server, x, x, c_off, n, alpha, exclusive, nblocks, stream);
}

extern "C" int dl_async_ea_client_out(const dl_comm_t* c, int server, float* x, float* x_out, uint64_t c_off, uint64_t n,
                                      float alpha, int exclusive, int nblocks, void* stream) {
  int nb = dl_check_comm(c, 0, dl_pick_blocks(n / 4 + 1, 4, nblocks));
  if (nb < 0) return nb;
  DL_REQUIRE(server >= 0 && server < c->world, "dl_async_ea_client: bad server rank %d", server);
  DL_REQUIRE(x != nullptr && reinterpret_cast<uintptr_t>(x) % 16 == 0, "dl_async_ea_client: x must be 16-byte aligned");
  DL_REQUIRE(x_out != nullptr && reinterpret_cast<uintptr_t>(x_out) % 16 == 0, "dl_async_ea_client: x_out must be 16-byte aligned");
  DL_REQUIRE(c_off % 16 == 0 && c_off >= DL_PAD_BYTES, "dl_async_ea_client: bad center offset");
  // exclusive mode spins between blocks of one grid: they must all be resident
  if (exclusive && nb > dl_sm_count()) nb = dl_sm_count();
  AsyncArgs a;
  a.x = x; a.x_out = x_out; a.c_off = c_off; a.n = n; a.alpha = alpha; a.server = server; a.exclusive = exclusive;
  async_ea_client_kernel<<<nb, DL_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(*c, a);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}

extern "C" int dl_async_sync_count(const dl_comm_t* c, int server, uint64_t* count, void* stream) {
  int nb = dl_check_comm(c, 0, 1);
  if (nb < 0) return nb;
  DL_REQUIRE(server >= 0 && server < c->world && count, "dl_async_sync_count: bad arguments");
  const DlAsyncCtl* ctl = reinterpret_cast<const DlAsyncCtl*>(reinterpret_cast<const char*>(c->base[server]) + DL_ASYNC_CTL_OFFSET);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned long long v = 0;
  DL_CUDA(cudaMemcpyAsync(&v, &ctl->sync_count, sizeof(v), cudaMemcpyDeviceToHost, st));
  DL_CUDA(cudaStreamSynchronize(st));
  *count = v;
  return DL_OK;
}

// Everything a polling host wants from rank `rank`'s control block in ONE copy: out[0] = sync counter, out[1..32] = the
// mailbox words.  (The server's wait loop used to issue one blocking copy per word, up to nine per tick.)
extern "C" int dl_async_ctl_snapshot(const dl_comm_t* c, int rank, uint64_t* out, void* stream) {
  int nb = dl_check_comm(c, 0, 1);
  if (nb < 0) return nb;
  DL_REQUIRE(rank >= 0 && rank < c->world && out, "dl_async_ctl_snapshot: bad arguments");
  static_assert(offsetof(DlAsyncCtl, mailbox) == offsetof(DlAsyncCtl, sync_count) + sizeof(unsigned long long), "layout");
  const DlAsyncCtl* ctl = reinterpret_cast<const DlAsyncCtl*>(reinterpret_cast<const char*>(c->base[rank]) + DL_ASYNC_CTL_OFFSET);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DL_CUDA(cudaMemcpyAsync(out, &ctl->sync_count, (1 + DL_MAILBOX_WORDS) * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
  DL_CUDA(cudaStreamSynchronize(st));
  return DL_OK;
}

// ---- host-visible mailbox words in a rank's pad (tester handshake, ready/done flags) ----------
extern "C" int dl_mailbox_write(const dl_comm_t* c, int rank, int index, uint64_t value, void* stream) {
  int nb = dl_check_comm(c, 0, 1);
  if (nb < 0) return nb;
  DL_REQUIRE(rank >= 0 && rank < c->world && index >= 0 && index < DL_MAILBOX_WORDS, "dl_mailbox_write: bad rank/index");
  DlAsyncCtl* ctl = reinterpret_cast<DlAsyncCtl*>(reinterpret_cast<char*>(c->base[rank]) + DL_ASYNC_CTL_OFFSET);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned long long v = value;
  DL_CUDA(cudaMemcpyAsync(&ctl->mailbox[index], &v, sizeof(v), cudaMemcpyHostToDevice, st));
  DL_CUDA(cudaStreamSynchronize(st));
  return DL_OK;
}

extern "C" int dl_mailbox_read(const dl_comm_t* c, int rank, int index, uint64_t* value, void* stream) {
  int nb = dl_check_comm(c, 0, 1);
  if (nb < 0) return nb;
  DL_REQUIRE(rank >= 0 && rank < c->world && index >= 0 && index < DL_MAILBOX_WORDS && value, "dl_mailbox_read: bad arguments");
  const DlAsyncCtl* ctl = reinterpret_cast<const DlAsyncCtl*>(reinterpret_cast<const char*>(c->base[rank]) + DL_ASYNC_CTL_OFFSET);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned long long v = 0;
  DL_CUDA(cudaMemcpyAsync(&v, &ctl->mailbox[index], sizeof(v), cudaMemcpyDeviceToHost, st));
  DL_CUDA(cudaStreamSynchronize(st));
  *value = v;
  return DL_OK;
}
