// Host (CPU) transport — SURVEY W9.
//
// The reference runs CPU nodes over libipc TCP sockets (examples/mnist.sh: four CPU processes).
// Here same-host CPU ranks share one POSIX shared-memory segment that holds every rank's
// "symmetric heap", and run the SAME algorithms as the GPU kernels (owner-computes two-shot
// reduce + push, double-buffered elastic averaging, slice-relay broadcast, center-resident async
// EASGD) with C++11 atomics as the release/acquire flag protocol.  It backs BASELINE config #1
// (2 CPU processes, MNIST AllReduceSGD) and lets the multi-rank logic be tested without a GPU.
//
// Layout of the segment:  [header 4 KiB][heap of rank 0][heap of rank 1]...   and every heap
// starts with a HostPad (DL_PAD_BYTES reserved, as on the GPU).
#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <cerrno>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>
#if defined(__x86_64__)
#include <emmintrin.h>
#endif

#include "../include/distlearn_b200.h"

namespace {

thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

constexpr uint64_t kMagic = 0x444c42323030484full;  // "DLB200HO"
constexpr size_t kHeaderBytes = 4096;

struct Header {
  std::atomic<uint64_t> magic;
  uint32_t world;
  uint32_t reserved;
  uint64_t heap_bytes;
  std::atomic<uint32_t> attached;
};

struct HostPad {
  std::atomic<uint64_t> flags[DL_NUM_CHANNELS][2][DL_MAX_RANKS];
  uint64_t aux[DL_NUM_CHANNELS][2][DL_MAX_RANKS];  // published before the flag (release), read after (acquire)
  uint32_t epoch[DL_NUM_CHANNELS];
  // async EASGD control (server side)
  std::atomic<uint32_t> ticket_next;
  std::atomic<uint32_t> ticket_serving;
  std::atomic<uint64_t> sync_count;
  std::atomic<uint64_t> mailbox[32];
};
static_assert(sizeof(HostPad) <= DL_PAD_BYTES, "host pad too large");

struct Handle {
  int rank, world;
  size_t heap_bytes, total_bytes;
  char* map;
  char name[128];
  double timeout_s;
  char* heap(int r) const { return map + kHeaderBytes + static_cast<size_t>(r) * heap_bytes; }
  HostPad* pad(int r) const { return reinterpret_cast<HostPad*>(heap(r)); }
  template <typename T>
  T* at(int r, uint64_t off) const {
    return reinterpret_cast<T*>(heap(r) + off);
  }
};

double now_s() {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec + ts.tv_nsec * 1e-9;
}

// bounded spin with back-off: yields quickly because CI machines have fewer cores than ranks
struct Spinner {
  unsigned polls = 0;
  double t0 = 0.0;
  double limit;
  explicit Spinner(double l) : limit(l) {}
  bool wait_more() {
    ++polls;
    if (polls < 64) return true;
    if (polls < 4096) {
      sched_yield();
    } else {
      struct timespec ts = {0, 50000};
      nanosleep(&ts, nullptr);
    }
    if ((polls & 255u) == 0) {
      const double t = now_s();
      if (t0 == 0.0) t0 = t;
      if (t - t0 > limit) return false;
    }
    return true;
  }
};

// Cross-rank barrier; returns 0 or DL_ETIMEOUT.  pay_out[r] receives rank r's payload word.
int barrier(Handle* h, int ch, uint32_t payload, uint32_t* pay_out, uint64_t aux = 0, uint64_t* aux_out = nullptr) {
  HostPad* mine = h->pad(h->rank);
  const uint32_t e = mine->epoch[ch] + 1u;
  const uint64_t word = (static_cast<uint64_t>(e) << 32) | payload;
  for (int r = 0; r < h->world; ++r) {
    h->pad(r)->aux[ch][e & 1u][h->rank] = aux;
    h->pad(r)->flags[ch][e & 1u][h->rank].store(word, std::memory_order_release);
  }
  for (int r = 0; r < h->world; ++r) {
    Spinner sp(h->timeout_s);
    uint64_t w;
    while (static_cast<uint32_t>((w = mine->flags[ch][e & 1u][r].load(std::memory_order_acquire)) >> 32) != e) {
      if (!sp.wait_more()) {
        set_error("rank %d timed out waiting for rank %d (channel %d, epoch %u)", h->rank, r, ch, e);
        mine->epoch[ch] = e;
        return DL_ETIMEOUT;
      }
    }
    if (pay_out) pay_out[r] = static_cast<uint32_t>(w);
    if (aux_out) aux_out[r] = mine->aux[ch][e & 1u][r];
  }
  mine->epoch[ch] = e;
  return DL_OK;
}

uint32_t decode(const Handle* h, const uint32_t* pay, dl_result_t* res, int* n_active) {
  uint32_t mask = 0;
  int n = 0;
  for (int r = 0; r < h->world; ++r) {
    if (pay[r] & 3u) mask |= 1u << r;  // data included
    if (pay[r] & 1u) ++n;              // counted as a contributor
  }
  if (res) {
    memset(res, 0, sizeof(*res));
    for (int r = 0; r < h->world; ++r) res->payload[r] = pay[r] >> 2;
    res->n_active = static_cast<uint32_t>(n);
  }
  *n_active = n;
  return mask;
}

struct Slice {
  uint64_t len, begin, end;
};
Slice slice_of(uint64_t n, int rank, int world) {
  Slice s;
  s.len = (n + world - 1) / world;
  s.begin = s.len * rank;
  s.end = s.begin + s.len;
  if (s.begin > n) s.begin = n;
  if (s.end > n) s.end = n;
  return s;
}

int check(Handle* h, int ch) {
  if (!h) {
    set_error("null handle");
    return DL_EINVAL;
  }
  if (ch < 0 || ch >= DL_NUM_CHANNELS) {
    set_error("bad channel %d", ch);
    return DL_EINVAL;
  }
  return DL_OK;
}

// ---- blocked, branch-free inner loops for the float paths -----------------------------------------------------------
// The slice is walked in blocks that stay in L1: the active ranks' vectors are summed into a local buffer IN RANK ORDER
// (the same order as the scalar code they replace and as the GPU kernels, so results do not depend on which path ran),
// the update runs on the buffer, and the result is copied out to every rank.  Each loop is a plain a[i] op b[i] over
// restrict pointers, which g++ vectorises; DLH_CLONES adds an AVX2 copy picked at load time.  Built with
// -ffp-contract=off: no fused multiply-adds, the rounding is the one written here.
// (DLH_NO_CLONES: the sanitizer builds switch the ifunc-based dispatch off - its resolvers run before TSan is up)
#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__) && !defined(DLH_NO_CLONES)
#define DLH_CLONES __attribute__((target_clones("avx2", "default")))
#else
#define DLH_CLONES
#endif
constexpr uint64_t kBlock = 2048;

DLH_CLONES void blk_add(float* __restrict acc, const float* __restrict src, uint64_t len) {
  for (uint64_t i = 0; i < len; ++i) acc[i] += src[i];
}
DLH_CLONES void blk_scale(float* __restrict acc, float f, uint64_t len) {
  for (uint64_t i = 0; i < len; ++i) acc[i] *= f;
}
// acc <- sum over the active ranks (ascending) of src[r][0..len)
void blk_sum_active(float* acc, const float* const* src, int world, uint32_t mask, uint64_t len) {
  if (mask & 1u) memcpy(acc, src[0], len * sizeof(float));
  else memset(acc, 0, len * sizeof(float));
  for (int r = 1; r < world; ++r)
    if ((mask >> r) & 1u) blk_add(acc, src[r], len);
}
// g (summed gradients, in) -> new parameters (out), exactly: g*=inv; g+=wd*p; m=mom*m+g; g=nesterov? g+mom*m : m; p-=lr*g
DLH_CLONES void blk_sgd_plain(float* __restrict g, const float* __restrict p, float inv, float wd, float lr, uint64_t len) {
  if (wd != 0.f) {
    for (uint64_t i = 0; i < len; ++i) {
      const float gi = g[i] * inv + wd * p[i];
      g[i] = p[i] - lr * gi;
    }
  } else {
    for (uint64_t i = 0; i < len; ++i) g[i] = p[i] - lr * (g[i] * inv);
  }
}
DLH_CLONES void blk_sgd_momentum(float* __restrict g, const float* __restrict p, float* __restrict m, float inv, float wd,
                                 float lr, float momentum, int nesterov, uint64_t len) {
  for (uint64_t i = 0; i < len; ++i) {
    float gi = g[i] * inv;
    if (wd != 0.f) gi += wd * p[i];
    const float mi = momentum * m[i] + gi;
    m[i] = mi;
    gi = nesterov ? gi + momentum * mi : mi;
    g[i] = p[i] - lr * gi;
  }
}
// acc <- c + sum_r (x_r - c) * alpha   (ranks ascending, starting from 0.f like the scalar form)
DLH_CLONES void blk_ea_accumulate(float* __restrict sum, const float* __restrict x, const float* __restrict c, float alpha,
                                  uint64_t len) {
  for (uint64_t i = 0; i < len; ++i) sum[i] += (x[i] - c[i]) * alpha;
}
DLH_CLONES void blk_add_to(float* __restrict out, const float* __restrict a, const float* __restrict b, uint64_t len) {
  for (uint64_t i = 0; i < len; ++i) out[i] = a[i] + b[i];
}
DLH_CLONES void blk_ea_local(float* __restrict x, const float* __restrict c, float alpha, uint64_t len) {
  for (uint64_t i = 0; i < len; ++i) x[i] -= (x[i] - c[i]) * alpha;
}

// Copy a finished block to every rank.  The destinations are written once and next read by OTHER cores, so on x86 they
// go out with non-temporal stores (no read-for-ownership of lines that are overwritten in full); stream_fence() must
// run before the barrier that publishes them, because a release store does not order non-temporal stores.  Only for
// vectors that do not fit the last-level cache anyway (kStreamFloats): below that the readers find the lines in cache and
// streaming them out to memory is a loss (measured: 20 -> 35 us at 10,848 floats, 6.5 -> 6.0 ms at 4.3 M).
constexpr uint64_t kStreamFloats = 1ull << 21;
#if defined(__x86_64__)
inline void blk_publish(float* dst, const float* src, uint64_t len, bool stream) {
  if (!stream || (reinterpret_cast<uintptr_t>(dst) & 15u) != 0 || len < 16) {
    memcpy(dst, src, len * sizeof(float));
    return;
  }
  uint64_t i = 0;
  for (; i + 4 <= len; i += 4) _mm_stream_ps(dst + i, _mm_load_ps(src + i));   // src is the 64-byte aligned block buffer
  for (; i < len; ++i) dst[i] = src[i];
}
inline void stream_fence() { _mm_sfence(); }
#else
inline void blk_publish(float* dst, const float* src, uint64_t len, bool) { memcpy(dst, src, len * sizeof(float)); }
inline void stream_fence() {}
#endif

template <typename T>
void reduce_typed(Handle* h, uint64_t off, uint64_t n, uint32_t mask, bool scale, int n_act) {
  // owner-computes over this rank's slice, push to everyone (same as the two-shot GPU kernel)
  const Slice s = slice_of(n, h->rank, h->world);
  const T* src[DL_MAX_RANKS];
  T* dst[DL_MAX_RANKS];
  for (int r = 0; r < h->world; ++r) {
    src[r] = h->at<T>(r, off);
    dst[r] = h->at<T>(r, off);
  }
  for (uint64_t i = s.begin; i < s.end; ++i) {
    T acc = (mask & 1u) ? src[0][i] : T(0);
    for (int r = 1; r < h->world; ++r) acc += ((mask >> r) & 1u) ? src[r][i] : T(0);
    if (scale && n_act > 1) acc = static_cast<T>(acc * (T(1) / static_cast<T>(n_act)));
    for (int r = 0; r < h->world; ++r) dst[r][i] = acc;
  }
}
template <>
void reduce_typed<float>(Handle* h, uint64_t off, uint64_t n, uint32_t mask, bool scale, int n_act) {
  const Slice s = slice_of(n, h->rank, h->world);
  const float inv = (scale && n_act > 1) ? 1.f / static_cast<float>(n_act) : 1.f;
  const float* src[DL_MAX_RANKS];
  float* dst[DL_MAX_RANKS];
  for (int r = 0; r < h->world; ++r) src[r] = dst[r] = h->at<float>(r, off);
  alignas(64) float acc[kBlock];
  for (uint64_t b0 = s.begin; b0 < s.end; b0 += kBlock) {
    const uint64_t len = s.end - b0 < kBlock ? s.end - b0 : kBlock;
    const float* blk[DL_MAX_RANKS];
    for (int r = 0; r < h->world; ++r) blk[r] = src[r] + b0;
    blk_sum_active(acc, blk, h->world, mask, len);
    blk_scale(acc, inv, len);
    for (int r = 0; r < h->world; ++r) blk_publish(dst[r] + b0, acc, len, n >= kStreamFloats);
  }
  stream_fence();
}

// Ranks that live in ONE process (threads, the way the reference's own tests run their nodes: ipc.map,
// test/test_AllReduceSGD.lua:26-35) share a single mapping of the segment: one mmap per (process, segment), counted.
// Besides saving address space this is what lets ThreadSanitizer see the cross-rank protocol at all — it tracks
// accesses by virtual address, and two mappings of the same pages would hide every race from it.
struct Mapping {
  char* map;
  size_t bytes;
  int refs;
};
std::mutex g_maps_mu;
std::map<std::string, Mapping> g_maps;

char* acquire_existing(const char* name, size_t total) {
  std::lock_guard<std::mutex> lk(g_maps_mu);
  auto it = g_maps.find(name);
  if (it == g_maps.end() || it->second.bytes != total) return nullptr;
  ++it->second.refs;
  return it->second.map;
}
// registers a fresh mapping, or - if another thread got there first - drops ours and returns theirs
char* publish_mapping(const char* name, char* map, size_t total) {
  std::lock_guard<std::mutex> lk(g_maps_mu);
  auto it = g_maps.find(name);
  if (it != g_maps.end() && it->second.bytes == total) {
    munmap(map, total);
    ++it->second.refs;
    return it->second.map;
  }
  g_maps[name] = Mapping{map, total, 1};
  return map;
}
void release_mapping(const char* name, char* map, size_t total) {
  std::lock_guard<std::mutex> lk(g_maps_mu);
  auto it = g_maps.find(name);
  if (it != g_maps.end() && it->second.map == map) {
    if (--it->second.refs > 0) return;
    g_maps.erase(it);
  }
  munmap(map, total);
}

}  // namespace

extern "C" {

const char* dlh_last_error(void) { return g_err; }

int dlh_open(const char* name, int rank, int world, uint64_t heap_bytes, double timeout_s, void** out) {
  if (world < 1 || world > DL_MAX_RANKS || rank < 0 || rank >= world || heap_bytes < DL_PAD_BYTES || heap_bytes % 4096) {
    set_error("dlh_open: bad arguments (rank %d world %d heap %llu)", rank, world, (unsigned long long)heap_bytes);
    return DL_EINVAL;
  }
  const size_t total = kHeaderBytes + static_cast<size_t>(world) * heap_bytes;
  const double limit = timeout_s > 0 ? timeout_s : 60.0;
  int fd = -1;
  void* map = nullptr;
  if (rank == 0) {
    shm_unlink(name);
    fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0 || ftruncate(fd, static_cast<off_t>(total)) != 0) {
      set_error("shm_open/ftruncate '%s': %s", name, strerror(errno));
      if (fd >= 0) close(fd);
      return DL_ESYS;
    }
  } else {
    const double t0 = now_s();
    while (true) {
      map = acquire_existing(name, total);       // a rank of this very process has mapped the segment already
      if (map) break;
      fd = shm_open(name, O_RDWR, 0600);
      if (fd >= 0) {
        struct stat st;
        if (fstat(fd, &st) == 0 && static_cast<size_t>(st.st_size) >= total) break;
        close(fd);
        fd = -1;
      }
      if (now_s() - t0 > limit) {
        set_error("timed out waiting for rank 0 to create '%s'", name);
        return DL_ETIMEOUT;
      }
      usleep(1000);
    }
  }
  if (!map) {
    map = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (map == MAP_FAILED) {
      set_error("mmap '%s': %s", name, strerror(errno));
      return DL_ESYS;
    }
    map = publish_mapping(name, static_cast<char*>(map), total);
  }
  Handle* h = new Handle();
  h->rank = rank;
  h->world = world;
  h->heap_bytes = heap_bytes;
  h->total_bytes = total;
  h->map = static_cast<char*>(map);
  h->timeout_s = limit;
  snprintf(h->name, sizeof(h->name), "%s", name);
  Header* hdr = reinterpret_cast<Header*>(h->map);
  if (rank == 0) {
    // a fresh segment is zero-filled by the kernel: pads start at epoch 0
    hdr->world = static_cast<uint32_t>(world);
    hdr->heap_bytes = heap_bytes;
    hdr->attached.store(1, std::memory_order_relaxed);
    hdr->magic.store(kMagic, std::memory_order_release);
  } else {
    const double t0 = now_s();
    while (hdr->magic.load(std::memory_order_acquire) != kMagic) {
      if (now_s() - t0 > limit) {
        set_error("timed out waiting for rank 0 to initialise '%s'", name);
        release_mapping(name, static_cast<char*>(map), total);
        delete h;
        return DL_ETIMEOUT;
      }
      usleep(500);
    }
    if (hdr->world != static_cast<uint32_t>(world) || hdr->heap_bytes != heap_bytes) {
      set_error("segment '%s' was created with a different geometry", name);
      release_mapping(name, static_cast<char*>(map), total);
      delete h;
      return DL_EINVAL;
    }
    hdr->attached.fetch_add(1, std::memory_order_acq_rel);
  }
  // everybody waits until all ranks are attached, then rank 0 may unlink the name
  {
    const double t0 = now_s();
    while (hdr->attached.load(std::memory_order_acquire) < static_cast<uint32_t>(world)) {
      if (now_s() - t0 > limit) {
        set_error("timed out waiting for all %d ranks to attach to '%s'", world, name);
        if (rank == 0) shm_unlink(name);
        release_mapping(name, static_cast<char*>(map), total);
        delete h;
        return DL_ETIMEOUT;
      }
      usleep(500);
    }
  }
  // Every rank holds a mapping now, so the NAME is no longer needed: drop it here rather than at close, and a job that
  // is killed later leaves nothing behind in /dev/shm.
  if (rank == 0) shm_unlink(name);
  *out = h;
  return DL_OK;
}

int dlh_close(void* handle, int unlink_name) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h) return DL_OK;
  if (unlink_name) shm_unlink(h->name);
  release_mapping(h->name, h->map, h->total_bytes);
  delete h;
  return DL_OK;
}

void* dlh_base(void* handle, int rank) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || rank < 0 || rank >= h->world) return nullptr;
  return h->heap(rank);
}

int dlh_set_timeout(void* handle, double seconds) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h) return DL_EINVAL;
  h->timeout_s = seconds;
  return DL_OK;
}

int dlh_barrier(void* handle, int active, uint32_t payload, int channel, dl_result_t* res) {
  Handle* h = static_cast<Handle*>(handle);
  int rc = check(h, channel);
  if (rc) return rc;
  uint32_t pay[DL_MAX_RANKS] = {0};
  rc = barrier(h, channel, (payload << 2) | (static_cast<uint32_t>(active) & 3u), pay);
  if (rc) return rc;
  int n;
  decode(h, pay, res, &n);
  return DL_OK;
}

int dlh_allreduce(void* handle, int dtype, uint64_t off, uint64_t n, int active, uint32_t payload, int scale,
                  int channel, dl_result_t* res) {
  Handle* h = static_cast<Handle*>(handle);
  int rc = check(h, channel);
  if (rc) return rc;
  uint32_t pay[DL_MAX_RANKS] = {0};
  uint64_t auxs[DL_MAX_RANKS] = {0};
  rc = barrier(h, channel, (payload << 2) | (static_cast<uint32_t>(active) & 3u), pay, scale ? 1u : 0u, auxs);
  if (rc) return rc;
  int n_act;
  const uint32_t mask = decode(h, pay, res, &n_act);
  if (mask) scale = static_cast<int>(auxs[__builtin_ctz(mask)] & 1u);  // the lowest active rank dictates
  switch (dtype) {
    case DL_F32: reduce_typed<float>(h, off, n, mask, scale != 0, n_act); break;
    case DL_I32: reduce_typed<int32_t>(h, off, n, mask, false, n_act); break;
    case DL_I64: reduce_typed<int64_t>(h, off, n, mask, false, n_act); break;
    case DL_F64: reduce_typed<double>(h, off, n, mask, scale != 0, n_act); break;
    default: set_error("bad dtype %d", dtype); return DL_EINVAL;
  }
  return barrier(h, channel, 0u, nullptr);
}

int dlh_fused_sgd(void* handle, uint64_t g_off, uint64_t p_off, uint64_t m_off, uint64_t n, float lr, float momentum,
                  float wd, int nesterov, int active, uint32_t payload, int zero_grads, int channel, dl_result_t* res) {
  Handle* h = static_cast<Handle*>(handle);
  int rc = check(h, channel);
  if (rc) return rc;
  uint32_t pay[DL_MAX_RANKS] = {0};
  uint64_t auxs[DL_MAX_RANKS] = {0};
  uint32_t lr_bits;
  memcpy(&lr_bits, &lr, 4);
  rc = barrier(h, channel, (payload << 2) | (static_cast<uint32_t>(active) & 3u), pay, lr_bits, auxs);
  if (rc) return rc;
  int n_act;
  const uint32_t mask = decode(h, pay, res, &n_act);
  if (mask) {  // the learning rate of the lowest active rank wins (draining ranks just follow)
    lr_bits = static_cast<uint32_t>(auxs[__builtin_ctz(mask)]);
    memcpy(&lr, &lr_bits, 4);
  }
  if (n_act > 0) {
    const Slice s = slice_of(n, h->rank, h->world);
    const float inv = n_act > 1 ? 1.f / static_cast<float>(n_act) : 1.f;
    float* p_loc = h->at<float>(h->rank, p_off);
    float* m_loc = m_off ? h->at<float>(h->rank, m_off) : nullptr;
    const float* g_src[DL_MAX_RANKS];
    float* p_dst[DL_MAX_RANKS];
    for (int r = 0; r < h->world; ++r) {
      g_src[r] = h->at<float>(r, g_off);
      p_dst[r] = h->at<float>(r, p_off);
    }
    alignas(64) float acc[kBlock];
    for (uint64_t b0 = s.begin; b0 < s.end; b0 += kBlock) {
      const uint64_t len = s.end - b0 < kBlock ? s.end - b0 : kBlock;
      const float* blk[DL_MAX_RANKS];
      for (int r = 0; r < h->world; ++r) blk[r] = g_src[r] + b0;
      blk_sum_active(acc, blk, h->world, mask, len);
      // zero_grad by the owner while the lines are still in cache (nobody touches gradients before the call returns)
      if (zero_grads)
        for (int r = 0; r < h->world; ++r) memset(const_cast<float*>(blk[r]), 0, len * sizeof(float));
      if (m_loc) blk_sgd_momentum(acc, p_loc + b0, m_loc + b0, inv, wd, lr, momentum, nesterov, len);
      else blk_sgd_plain(acc, p_loc + b0, inv, wd, lr, len);
      for (int r = 0; r < h->world; ++r) blk_publish(p_dst[r] + b0, acc, len, n >= kStreamFloats);
    }
    stream_fence();
  }
  rc = barrier(h, channel, 0u, nullptr);
  if (rc) return rc;
  if (zero_grads && n_act == 0) memset(h->at<float>(h->rank, g_off), 0, n * sizeof(float));   // nobody owned a pass
  return DL_OK;
}

int dlh_fused_ea(void* handle, uint64_t x_off, uint64_t c_off, uint64_t cnew_off, uint64_t n, float alpha, int active,
                 uint32_t payload, int channel, dl_result_t* res) {
  Handle* h = static_cast<Handle*>(handle);
  int rc = check(h, channel);
  if (rc) return rc;
  if (c_off == cnew_off) {
    set_error("dlh_fused_ea: next-center buffer must differ from the current one");
    return DL_EINVAL;
  }
  uint32_t pay[DL_MAX_RANKS] = {0};
  rc = barrier(h, channel, (payload << 2) | (static_cast<uint32_t>(active) & 3u), pay);
  if (rc) return rc;
  int n_act;
  decode(h, pay, res, &n_act);
  const float* c_old = h->at<float>(h->rank, c_off);
  if (n_act > 0) {
    const Slice s = slice_of(n, h->rank, h->world);
    alignas(64) float sum[kBlock];
    alignas(64) float cn[kBlock];
    for (uint64_t b0 = s.begin; b0 < s.end; b0 += kBlock) {
      const uint64_t len = s.end - b0 < kBlock ? s.end - b0 : kBlock;
      memset(sum, 0, len * sizeof(float));
      for (int r = 0; r < h->world; ++r) blk_ea_accumulate(sum, h->at<float>(r, x_off) + b0, c_old + b0, alpha, len);
      blk_add_to(cn, c_old + b0, sum, len);
      for (int r = 0; r < h->world; ++r) blk_publish(h->at<float>(r, cnew_off) + b0, cn, len, n >= kStreamFloats);
    }
    stream_fence();
  }
  rc = barrier(h, channel, 0u, nullptr);
  if (rc) return rc;
  if (n_act > 0) {
    float* x = h->at<float>(h->rank, x_off);
    blk_ea_local(x, c_old, alpha, n);
  } else {
    memcpy(h->at<float>(h->rank, cnew_off), c_old, n * sizeof(float));
  }
  return DL_OK;
}

int dlh_broadcast(void* handle, uint64_t off, uint64_t off2, uint64_t n, int root, uint32_t payload, int channel,
                  dl_result_t* res) {
  Handle* h = static_cast<Handle*>(handle);
  int rc = check(h, channel);
  if (rc) return rc;
  if (root < 0 || root >= h->world) {
    set_error("bad root %d", root);
    return DL_EINVAL;
  }
  uint32_t pay[DL_MAX_RANKS] = {0};
  rc = barrier(h, channel, (payload << 2) | 1u, pay);
  if (rc) return rc;
  int n_act;
  decode(h, pay, res, &n_act);
  const Slice s = slice_of(n, h->rank, h->world);
  const float* src = h->at<float>(root, off);
  for (int r = 0; r < h->world; ++r) {
    if (r != root) memcpy(h->at<float>(r, off) + s.begin, src + s.begin, (s.end - s.begin) * sizeof(float));
    if (off2) memcpy(h->at<float>(r, off2) + s.begin, src + s.begin, (s.end - s.begin) * sizeof(float));
  }
  return barrier(h, channel, 0u, nullptr);
}

// Async EASGD client sync against a center resident in `server`'s heap (lua/AsyncEA.lua:134-146).
int dlh_async_ea_client(void* handle, int server, float* x, uint64_t c_off, uint64_t n, float alpha, int exclusive) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || server < 0 || server >= h->world) {
    set_error("bad server rank");
    return DL_EINVAL;
  }
  HostPad* srv = h->pad(server);
  uint32_t ticket = 0;
  if (exclusive) {
    // Same queue discipline as the CUDA kernel (csrc/async_ea.cu): nobody leaves the queue; a ticket that is not
    // released for a whole timeout is passed on by one of the waiters (its holder is taken for dead), a waiter that
    // was overtaken queues again, releases never move the counter backwards.
    ticket = srv->ticket_next.fetch_add(1, std::memory_order_relaxed);
    uint32_t last = 0xffffffffu;
    int recoveries = 0;
    Spinner sp(h->timeout_s);
    while (true) {
      uint32_t cur = srv->ticket_serving.load(std::memory_order_acquire);
      if (cur == ticket) break;
      if (static_cast<int32_t>(cur - ticket) > 0) {
        ticket = srv->ticket_next.fetch_add(1, std::memory_order_relaxed);
        last = 0xffffffffu;
        continue;
      }
      if (cur != last) {
        last = cur;
        sp = Spinner(h->timeout_s);
      }
      if (!sp.wait_more()) {
        if (++recoveries > 8) {
          set_error("timed out waiting for the center lock (ticket %u)", ticket);
          return DL_ETIMEOUT;
        }
        srv->ticket_serving.compare_exchange_strong(cur, cur + 1u, std::memory_order_acq_rel);
        sp = Spinner(h->timeout_s);
      }
    }
  }
  float* center = h->at<float>(server, c_off);
  {
    // the center is only ever changed by atomic adds, with or without the lock: a recovery that lets a slow holder and
    // its successor overlap cannot lose an update
    for (uint64_t i = 0; i < n; ++i) {
      std::atomic<uint32_t>* cell = reinterpret_cast<std::atomic<uint32_t>*>(center + i);
      uint32_t old_bits = cell->load(std::memory_order_relaxed);
      float cv;
      memcpy(&cv, &old_bits, 4);
      const float d = (x[i] - cv) * alpha;
      x[i] -= d;
      // atomic float add (CAS loop), the CPU counterpart of red.global.add.f32
      while (true) {
        float cur;
        memcpy(&cur, &old_bits, 4);
        const float nv = cur + d;
        uint32_t new_bits;
        memcpy(&new_bits, &nv, 4);
        if (cell->compare_exchange_weak(old_bits, new_bits, std::memory_order_relaxed)) break;
      }
    }
  }
  if (exclusive) {  // release: serving = max(serving, ticket + 1)
    uint32_t cur = srv->ticket_serving.load(std::memory_order_relaxed);
    while (static_cast<int32_t>(ticket + 1u - cur) > 0 &&
           !srv->ticket_serving.compare_exchange_weak(cur, ticket + 1u, std::memory_order_release, std::memory_order_relaxed)) {
    }
  }
  srv->sync_count.fetch_add(1, std::memory_order_release);
  return DL_OK;
}

// Fault injection for the tests: draw a ticket of the exclusive-mode queue and never release it (a client that died
// while holding the center lock).
int dlh_debug_take_ticket(void* handle, int server, uint32_t* ticket) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || server < 0 || server >= h->world || !ticket) return DL_EINVAL;
  *ticket = h->pad(server)->ticket_next.fetch_add(1, std::memory_order_relaxed);
  return DL_OK;
}

int dlh_async_sync_count(void* handle, int server, uint64_t* count) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || server < 0 || server >= h->world || !count) return DL_EINVAL;
  *count = h->pad(server)->sync_count.load(std::memory_order_acquire);
  return DL_OK;
}

int dlh_mailbox_write(void* handle, int rank, int index, uint64_t value) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || rank < 0 || rank >= h->world || index < 0 || index >= 32) return DL_EINVAL;
  h->pad(rank)->mailbox[index].store(value, std::memory_order_release);
  return DL_OK;
}

int dlh_mailbox_read(void* handle, int rank, int index, uint64_t* value) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || rank < 0 || rank >= h->world || index < 0 || index >= 32 || !value) return DL_EINVAL;
  *value = h->pad(rank)->mailbox[index].load(std::memory_order_acquire);
  return DL_OK;
}

int dlh_peer_pull(void* handle, int src_rank, uint64_t off, float* dst, uint64_t n) {
  Handle* h = static_cast<Handle*>(handle);
  if (!h || src_rank < 0 || src_rank >= h->world) return DL_EINVAL;
  memcpy(dst, h->at<float>(src_rank, off), n * sizeof(float));
  return DL_OK;
}

}  // extern "C"
