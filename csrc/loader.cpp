// Native batch loader: worker threads assemble training batches into caller-provided (pinned) host buffers ahead of the
// consumer.  Counterpart of the reference's `dataset` package as it is used by the examples
// (`Dataset(path, {partition, partitions})` + `sampledBatcher{samplerKind = 'permutation' | 'label-uniform', batchSize,
// inputDims, processor}` -> `getBatch()`, `numBatches()`; examples/mnist.lua:26-40, examples/cifar10.lua:41-71,
// examples/Data.lua:22-61), which is Lua over C.
//
// * The dataset stays where the caller put it: [n_total, sample_elems] uint8 (CIFAR-style bytes) or float32, labels int64.
//   Node `partition` of `partitions` sees samples partition-1, partition-1+partitions, ... (the reference's partitioning).
// * A batch is a pure function of (seed, batch number): the sampler uses a counter-based generator, so the stream of
//   batches does not depend on the number of worker threads or on their timing.
//     permutation   : epoch e = batch / ceil(n / B) uses the Fisher-Yates shuffle seeded by (seed, e); the last batch of
//                     an epoch is short when B does not divide n
//     label-uniform : each sample = uniform class (among the classes this partition holds), then uniform member of it
// * uint8 samples are converted to float32 as  x = u8 * scale[c] + shift[c]  per channel (mean/std normalisation folded
//   into one FMA); float32 samples go through the same FMA (scale 1, shift 0 = copy).
// * Ring of `depth` slots.  Workers fill batch b into slot b % depth as soon as the consumer has released batch
//   b - depth; dll_next() hands out batches strictly in order; dll_release() recycles the oldest outstanding one.
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <map>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

namespace {

thread_local char g_err[256] = "";

constexpr int kMaxChannels = 16;

inline uint64_t mix(uint64_t x) {  // splitmix64 finaliser
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
inline uint64_t rnd(uint64_t seed, uint64_t a, uint64_t b, uint64_t c) { return mix(mix(mix(seed ^ mix(a)) ^ b) ^ c); }
// unbiased enough for sampling: 64-bit value scaled to [0, n)
inline uint64_t below(uint64_t r, uint64_t n) { return static_cast<uint64_t>((static_cast<unsigned __int128>(r) * n) >> 64); }

struct Loader {
  const uint8_t* u8 = nullptr;
  const float* f32 = nullptr;
  const int64_t* labels = nullptr;
  int64_t n_total = 0, first = 0, stride = 1, n = 0;
  int sample_elems = 0, channels = 1, chan_elems = 0, num_classes = 0;
  float scale[kMaxChannels], shift[kMaxChannels];
  int kind = 0, batch = 1, depth = 2;
  uint64_t seed = 0;
  int64_t batches_per_epoch = 0;
  std::vector<std::vector<int64_t>> by_class;   // local sample numbers (0..n) per class
  std::vector<int> nonempty;
  std::vector<float*> xs;
  std::vector<int64_t*> ys;
  std::vector<int> sizes;
  std::vector<int64_t> ready;                    // batch number sitting in each slot, -1 = none
  // permutations of the epochs in flight
  std::mutex perm_mu;
  std::map<int64_t, std::shared_ptr<const std::vector<int64_t>>> perms;
  // ring state
  std::mutex mu;
  std::condition_variable cv_work, cv_ready;
  int64_t next_fill = 0, next_consume = 0, released = 0;
  bool stop = false;
  std::vector<std::thread> workers;
  std::atomic<uint64_t> samples_done{0};

  int64_t global_index(int64_t local) const { return first + local * stride; }

  // shared ownership: a worker may still be reading epoch e when the cache drops it
  std::shared_ptr<const std::vector<int64_t>> permutation(int64_t epoch) {
    std::lock_guard<std::mutex> lk(perm_mu);
    auto it = perms.find(epoch);
    if (it != perms.end()) return it->second;
    auto p = std::make_shared<std::vector<int64_t>>(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i) (*p)[static_cast<size_t>(i)] = i;
    for (int64_t i = n - 1; i > 0; --i) {
      const int64_t j = static_cast<int64_t>(below(rnd(seed, 0x70, static_cast<uint64_t>(epoch), static_cast<uint64_t>(i)), static_cast<uint64_t>(i + 1)));
      std::swap((*p)[static_cast<size_t>(i)], (*p)[static_cast<size_t>(j)]);
    }
    while (perms.size() >= 3) perms.erase(perms.begin());   // batches are handed out in order: older epochs are (nearly) done
    perms.emplace(epoch, p);
    return p;
  }

  // local sample numbers of batch b; returns the batch size
  int indices(int64_t b, int64_t* out) {
    if (kind == 0) {
      const int64_t epoch = b / batches_per_epoch, pos = (b % batches_per_epoch) * batch;
      const std::shared_ptr<const std::vector<int64_t>> p = permutation(epoch);
      const int64_t cnt = std::min<int64_t>(batch, n - pos);
      for (int64_t i = 0; i < cnt; ++i) out[i] = (*p)[static_cast<size_t>(pos + i)];
      return static_cast<int>(cnt);
    }
    for (int i = 0; i < batch; ++i) {
      const int cls = nonempty[below(rnd(seed, 0x4c, static_cast<uint64_t>(b), 2 * static_cast<uint64_t>(i)), nonempty.size())];
      const std::vector<int64_t>& pool = by_class[static_cast<size_t>(cls)];
      out[i] = pool[below(rnd(seed, 0x4c, static_cast<uint64_t>(b), 2 * static_cast<uint64_t>(i) + 1), pool.size())];
    }
    return batch;
  }

  void fill(int64_t b, int slot) {
    std::vector<int64_t> idx(static_cast<size_t>(batch));
    const int cnt = indices(b, idx.data());
    float* x = xs[static_cast<size_t>(slot)];
    int64_t* y = ys[static_cast<size_t>(slot)];
    for (int i = 0; i < cnt; ++i) {
      const int64_t g = global_index(idx[static_cast<size_t>(i)]);
      y[i] = labels[g];
      float* dst = x + static_cast<size_t>(i) * sample_elems;
      if (u8 != nullptr) {
        const uint8_t* src = u8 + static_cast<size_t>(g) * sample_elems;
        for (int c = 0; c < channels; ++c) {
          const float sc = scale[c], sh = shift[c];
          const uint8_t* s = src + static_cast<size_t>(c) * chan_elems;
          float* d = dst + static_cast<size_t>(c) * chan_elems;
          for (int e = 0; e < chan_elems; ++e) d[e] = static_cast<float>(s[e]) * sc + sh;
        }
      } else {
        const float* src = f32 + static_cast<size_t>(g) * sample_elems;
        for (int c = 0; c < channels; ++c) {
          const float sc = scale[c], sh = shift[c];
          const float* s = src + static_cast<size_t>(c) * chan_elems;
          float* d = dst + static_cast<size_t>(c) * chan_elems;
          if (sc == 1.f && sh == 0.f) memcpy(d, s, sizeof(float) * static_cast<size_t>(chan_elems));
          else for (int e = 0; e < chan_elems; ++e) d[e] = s[e] * sc + sh;
        }
      }
    }
    sizes[static_cast<size_t>(slot)] = cnt;
    samples_done.fetch_add(static_cast<uint64_t>(cnt), std::memory_order_relaxed);
  }

  void work() {
    for (;;) {
      int64_t b;
      {
        std::unique_lock<std::mutex> lk(mu);
        cv_work.wait(lk, [&] { return stop || next_fill < released + depth; });
        if (stop) return;
        b = next_fill++;
      }
      const int slot = static_cast<int>(b % depth);
      fill(b, slot);
      {
        std::lock_guard<std::mutex> lk(mu);
        ready[static_cast<size_t>(slot)] = b;
      }
      cv_ready.notify_all();
    }
  }
};

int fail(const char* msg) {
  strncpy(g_err, msg, sizeof(g_err) - 1);
  g_err[sizeof(g_err) - 1] = 0;
  return -1;
}

}  // namespace

extern "C" {

const char* dll_last_error(void) { return g_err; }

// data: uint8 (is_u8 = 1) or float32 [n_total, sample_elems]; labels int64 [n_total] with values in [0, num_classes).
// channels divides sample_elems; scale / shift have `channels` entries (null = 1 / 0).
// kind: 0 = permutation, 1 = label-uniform.  x_slots / y_slots: `depth` caller-owned buffers of batch * sample_elems
// floats / batch int64 each (pinned host memory when the batches go to a GPU).
int dll_create(const void* data, int is_u8, const int64_t* labels, int64_t n_total, int sample_elems, int channels,
               const float* scale, const float* shift, int num_classes, int partition, int partitions, int kind, int batch,
               uint64_t seed, int threads, int depth, float* const* x_slots, int64_t* const* y_slots, void** out) {
  if (!data || !labels || !out || !x_slots || !y_slots) return fail("dll_create: null argument");
  if (n_total <= 0 || sample_elems <= 0 || batch <= 0 || depth < 2 || threads < 1 || threads > 64 || num_classes <= 0)
    return fail("dll_create: sizes must be positive (depth >= 2, 1 <= threads <= 64)");
  if (channels < 1 || channels > kMaxChannels || sample_elems % channels) return fail("dll_create: channels must divide sample_elems (max 16)");
  if (partitions < 1 || partition < 1 || partition > partitions) return fail("dll_create: partition must be in 1..partitions");
  if (kind != 0 && kind != 1) return fail("dll_create: kind must be 0 (permutation) or 1 (label-uniform)");
  Loader* L = new Loader();
  if (is_u8) L->u8 = static_cast<const uint8_t*>(data); else L->f32 = static_cast<const float*>(data);
  L->labels = labels;
  L->n_total = n_total;
  L->first = partition - 1;
  L->stride = partitions;
  L->n = L->first < n_total ? (n_total - L->first + partitions - 1) / partitions : 0;
  if (L->n <= 0) { delete L; return fail("dll_create: this partition is empty"); }
  L->sample_elems = sample_elems;
  L->channels = channels;
  L->chan_elems = sample_elems / channels;
  L->num_classes = num_classes;
  for (int c = 0; c < channels; ++c) {
    L->scale[c] = scale ? scale[c] : 1.f;
    L->shift[c] = shift ? shift[c] : 0.f;
  }
  L->kind = kind; L->batch = batch; L->depth = depth; L->seed = seed;
  L->batches_per_epoch = (L->n + batch - 1) / batch;
  L->by_class.resize(static_cast<size_t>(num_classes));
  for (int64_t i = 0; i < L->n; ++i) {
    const int64_t y = labels[L->global_index(i)];
    if (y < 0 || y >= num_classes) { delete L; return fail("dll_create: label out of range"); }
    L->by_class[static_cast<size_t>(y)].push_back(i);
  }
  for (int c = 0; c < num_classes; ++c) if (!L->by_class[static_cast<size_t>(c)].empty()) L->nonempty.push_back(c);
  L->xs.assign(x_slots, x_slots + depth);
  L->ys.assign(y_slots, y_slots + depth);
  L->sizes.assign(static_cast<size_t>(depth), 0);
  L->ready.assign(static_cast<size_t>(depth), -1);
  for (int t = 0; t < threads; ++t) L->workers.emplace_back([L] { L->work(); });
  *out = L;
  return 0;
}

// Next batch in order: its slot and size.  Returns 0, or -2 on timeout (timeout_s <= 0: wait for ever).
int dll_next(void* h, int* slot, int* size, int64_t* batch_no, double timeout_s) {
  Loader* L = static_cast<Loader*>(h);
  if (!L || !slot || !size) return fail("dll_next: null argument");
  std::unique_lock<std::mutex> lk(L->mu);
  const int64_t b = L->next_consume;
  const int s = static_cast<int>(b % L->depth);
  auto have = [&] { return L->ready[static_cast<size_t>(s)] == b; };
  if (b >= L->released + L->depth) return fail("dll_next: every slot is outstanding, release one first");
  if (timeout_s > 0) {
    if (!L->cv_ready.wait_for(lk, std::chrono::duration<double>(timeout_s), have)) {
      fail("dll_next: timed out");
      return -2;
    }
  } else {
    L->cv_ready.wait(lk, have);
  }
  L->next_consume = b + 1;
  *slot = s;
  *size = L->sizes[static_cast<size_t>(s)];
  if (batch_no) *batch_no = b;
  return 0;
}

// Give the OLDEST outstanding batch back (batches are released in the order they were handed out).
int dll_release(void* h) {
  Loader* L = static_cast<Loader*>(h);
  if (!L) return fail("dll_release: null handle");
  {
    std::lock_guard<std::mutex> lk(L->mu);
    if (L->released >= L->next_consume) return fail("dll_release: nothing is outstanding");
    L->released += 1;
  }
  L->cv_work.notify_all();
  return 0;
}

int64_t dll_num_batches(void* h) { return h ? static_cast<Loader*>(h)->batches_per_epoch : 0; }
int64_t dll_num_samples(void* h) { return h ? static_cast<Loader*>(h)->n : 0; }
uint64_t dll_samples_done(void* h) { return h ? static_cast<Loader*>(h)->samples_done.load() : 0; }

int dll_destroy(void* h) {
  Loader* L = static_cast<Loader*>(h);
  if (!L) return 0;
  {
    std::lock_guard<std::mutex> lk(L->mu);
    L->stop = true;
  }
  L->cv_work.notify_all();
  for (std::thread& t : L->workers) t.join();
  delete L;
  return 0;
}

// CIFAR-10 binary batches (data_batch_N.bin / test_batch.bin): records of 1 label byte + 3072 pixel bytes (3 planes of
// 32x32).  Fills data[n, 3072] and labels[n]; returns the number of records read (<= max_records) or -1.
int64_t dll_read_cifar_bin(const char* path, uint8_t* data, int64_t* labels, int64_t max_records) {
  FILE* f = fopen(path, "rb");
  if (!f) return fail("dll_read_cifar_bin: cannot open file");
  uint8_t rec[3073];
  int64_t n = 0;
  while (n < max_records && fread(rec, 1, sizeof(rec), f) == sizeof(rec)) {
    labels[n] = rec[0];
    memcpy(data + static_cast<size_t>(n) * 3072, rec + 1, 3072);
    ++n;
  }
  fclose(f);
  return n;
}

}  // extern "C"
