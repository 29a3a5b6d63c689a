// Host-side common code: error text, launch geometry, pad reset / result read-back, plain barrier.
#include <string.h>

#include "dl_common.cuh"
#include "dl_host.h"

static thread_local char g_err[1024] = "";

void dl_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* dl_last_error(void) { return g_err; }

#include <atomic>
static std::atomic<unsigned long long> g_launches{0};
void dl_note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
extern "C" unsigned long long dl_kernel_launches(void) { return g_launches.load(std::memory_order_relaxed); }
extern "C" int dl_version(void) { return 100; }

int dl_sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

int dl_pick_blocks(unsigned long long work_items, int per_thread, int requested) {
  const int sms = dl_sm_count();
  int cap = sms < DL_MAX_BLOCKS ? sms : DL_MAX_BLOCKS;
  if (requested > 0) return requested < cap ? requested : cap;
  const unsigned long long per_block = static_cast<unsigned long long>(DL_THREADS) * (per_thread > 0 ? per_thread : 1);
  unsigned long long nb = (work_items + per_block - 1) / per_block;
  if (nb < 1) nb = 1;
  if (nb > static_cast<unsigned long long>(cap)) nb = cap;
  return static_cast<int>(nb);
}

int dl_check_comm(const dl_comm_t* c, int channel, int nblocks) {
  if (!c) {
    dl_set_error("null communicator");
    return DL_EINVAL;
  }
  if (c->world < 1 || c->world > DL_MAX_RANKS || c->rank < 0 || c->rank >= c->world) {
    dl_set_error("bad rank/world %d/%d (max world %d)", c->rank, c->world, DL_MAX_RANKS);
    return DL_EINVAL;
  }
  if (channel < 0 || channel >= DL_NUM_CHANNELS) {
    dl_set_error("bad channel %d", channel);
    return DL_EINVAL;
  }
  for (int r = 0; r < c->world; ++r) {
    if (!c->base[r]) {
      dl_set_error("heap of rank %d is not mapped", r);
      return DL_EINVAL;
    }
  }
  if (nblocks < 1) nblocks = 1;
  if (nblocks > DL_MAX_BLOCKS) nblocks = DL_MAX_BLOCKS;
  if (c->emulate && c->world > 1) {
    // cooperative launch: all emulated ranks' blocks must be co-resident (1 CTA of 512 threads per SM)
    const int cap = dl_sm_count() / c->world;
    if (cap < 1) {
      dl_set_error("emulated world %d does not fit on this GPU", c->world);
      return DL_EINVAL;
    }
    if (nblocks > cap) nblocks = cap;
  }
  return nblocks;
}

// ---- plain barrier ---------------------------------------------------------------------------
template <bool kEmu>
__global__ void __launch_bounds__(32) barrier_kernel(dl_comm_t c, uint32_t active, uint32_t payload, int channel) {
  __shared__ uint32_t pay[DL_MAX_RANKS];
  const int rank = kEmu ? static_cast<int>(blockIdx.y) : c.rank;
  const int act = kEmu ? ((active >> rank) & 1u) : (active & 3u);
  const uint32_t user = kEmu ? payload + rank : payload;
  dl_block_barrier(c, rank, c.world, channel, dl_pack_payload(act, user), pay);
  int n_act;
  dl_active_mask(c, rank, c.world, channel, pay, &n_act);
}

extern "C" int dl_barrier(const dl_comm_t* c, int active, uint32_t payload, int channel, void* stream) {
  int nb = dl_check_comm(c, channel, 1);
  if (nb < 0) return nb;
  DL_REQUIRE(payload < (1u << 30), "dl_barrier: payload must fit 30 bits");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (c->emulate)
    return dl_launch(barrier_kernel<true>, c, 1, 32, st, *c, static_cast<uint32_t>(active), payload, channel);
  return dl_launch(barrier_kernel<false>, c, 1, 32, st, *c, static_cast<uint32_t>(active), payload, channel);
}

extern "C" int dl_read_result(const dl_comm_t* c, int channel, dl_result_t* out, void* stream) {
  int nb = dl_check_comm(c, channel, 1);
  if (nb < 0) return nb;
  DL_REQUIRE(out != nullptr, "dl_read_result: null out");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const DlPad* pad = reinterpret_cast<const DlPad*>(c->base[c->rank]) + channel;
  DL_CUDA(cudaMemcpyAsync(out, &pad->result, sizeof(dl_result_t), cudaMemcpyDeviceToHost, st));
  DL_CUDA(cudaStreamSynchronize(st));
  if (out->error) {
    // reported once: clear the device word so that later calls are judged on their own
    DL_CUDA(cudaMemsetAsync(const_cast<uint32_t*>(&pad->result.error), 0, sizeof(uint32_t), st));
    DL_CUDA(cudaStreamSynchronize(st));
    if (static_cast<int>(out->error) == -DL_EPROTO) {
      dl_set_error("collective on channel %d found a barrier slot the protocol cannot produce (rank %d of %d): ranks are not "
                   "launching the same sequence of collectives, or something else wrote into the signal pad", channel, c->rank, c->world);
      return DL_EPROTO;
    }
    dl_set_error("collective on channel %d timed out waiting for a peer (rank %d of %d)", channel, c->rank, c->world);
    return DL_ETIMEOUT;
  }
  return DL_OK;
}

extern "C" int dl_reset_pad(const dl_comm_t* c) {
  int nb = dl_check_comm(c, 0, 1);
  if (nb < 0) return nb;
  if (c->emulate) {
    for (int r = 0; r < c->world; ++r) DL_CUDA(cudaMemset(c->base[r], 0, DL_PAD_BYTES));
  } else {
    DL_CUDA(cudaMemset(c->base[c->rank], 0, DL_PAD_BYTES));
  }
  DL_CUDA(cudaDeviceSynchronize());
  return DL_OK;
}

// Pinned, device-mapped host memory for dl_comm_t.host_words: kernels write the contributor count and the sticky
// error word there, the host reads them with plain loads (no copy, no stream synchronisation of its own).
extern "C" int dl_host_words_alloc(void** host_ptr, void** dev_ptr) {
  DL_REQUIRE(host_ptr && dev_ptr, "dl_host_words_alloc: null pointer");
  void* h = nullptr;
  DL_CUDA(cudaHostAlloc(&h, sizeof(dl_host_words_t), cudaHostAllocMapped | cudaHostAllocPortable));
  memset(h, 0, sizeof(dl_host_words_t));
  void* d = nullptr;
  DL_CUDA(cudaHostGetDevicePointer(&d, h, 0));
  *host_ptr = h;
  *dev_ptr = d;
  return DL_OK;
}
extern "C" int dl_host_words_free(void* host_ptr) {
  if (host_ptr) DL_CUDA(cudaFreeHost(host_ptr));
  return DL_OK;
}

extern "C" int dl_memset_sync(int dev, void* ptr, int value, size_t bytes) {
  DL_CUDA(cudaSetDevice(dev));
  DL_CUDA(cudaMemset(ptr, value, bytes));
  DL_CUDA(cudaDeviceSynchronize());
  return DL_OK;
}
