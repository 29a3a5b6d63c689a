// ThreadSanitizer harness for the CPU-side native code (SURVEY §5 "race detection / sanitizers").
//
// Built with -fsanitize=thread together with csrc/host_transport.cpp and csrc/loader.cpp and run by
// tests/test_tsan_cpu.py.  The ranks of the shared-memory transport run as THREADS of this one process (the way the
// reference's own tests run their nodes, ipc.map in test/test_AllReduceSGD.lua:26-35) and share one mapping of the
// segment, so every cross-rank access goes through the same virtual addresses and TSan checks the flag protocol
// (release/acquire epochs, owner-computes slices, double-buffered center, the ticket queue) for real.
// Exit code 0 = every collective gave the right numbers; TSan itself reports races on stderr and makes the exit code 66.
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <unistd.h>
#include <vector>

#include "../../include/distlearn_b200.h"
#include "../../include/distlearn_loader.h"

extern "C" {
const char* dlh_last_error(void);
int dlh_open(const char* name, int rank, int world, uint64_t heap_bytes, double timeout_s, void** out);
int dlh_close(void* handle, int unlink_name);
void* dlh_base(void* handle, int rank);
int dlh_barrier(void* handle, int active, uint32_t payload, int channel, dl_result_t* res);
int dlh_allreduce(void* handle, int dtype, uint64_t off, uint64_t n, int active, uint32_t payload, int scale, int channel, dl_result_t* res);
int dlh_fused_sgd(void* handle, uint64_t g_off, uint64_t p_off, uint64_t m_off, uint64_t n, float lr, float momentum, float wd,
                  int nesterov, int active, uint32_t payload, int zero_grads, int channel, dl_result_t* res);
int dlh_fused_ea(void* handle, uint64_t x_off, uint64_t c_off, uint64_t cnew_off, uint64_t n, float alpha, int active, uint32_t payload,
                 int channel, dl_result_t* res);
int dlh_broadcast(void* handle, uint64_t off, uint64_t off2, uint64_t n, int root, uint32_t payload, int channel, dl_result_t* res);
int dlh_async_ea_client(void* handle, int server, float* x, uint64_t c_off, uint64_t n, float alpha, int exclusive);
int dlh_async_sync_count(void* handle, int server, uint64_t* count);
}

static std::atomic<int> g_fail{0};
#define CHECK(cond, ...)                      \
  do {                                        \
    if (!(cond)) {                            \
      fprintf(stderr, "FAIL rank %d: ", rank); \
      fprintf(stderr, __VA_ARGS__);           \
      fprintf(stderr, "\n");                  \
      g_fail.fetch_add(1);                    \
    }                                         \
  } while (0)
#define OK(call) CHECK((call) == 0, "%s -> %s", #call, dlh_last_error())

static void rank_main(const char* name, int rank, int world, int rounds) {
  const uint64_t heap = 4u << 20;
  void* h = nullptr;
  if (dlh_open(name, rank, world, heap, 30.0, &h) != 0) {
    fprintf(stderr, "rank %d: open failed: %s\n", rank, dlh_last_error());
    g_fail.fetch_add(1);
    return;
  }
  char* base = static_cast<char*>(dlh_base(h, rank));
  const uint64_t n = 6000;  // floats; not a multiple of the block size or of the world
  const uint64_t off_g = DL_PAD_BYTES, off_p = off_g + 32768, off_m = off_p + 32768, off_c0 = off_m + 32768, off_c1 = off_c0 + 32768,
                 off_ctr = off_c1 + 32768;
  float* g = reinterpret_cast<float*>(base + off_g);
  float* p = reinterpret_cast<float*>(base + off_p);
  float* m = reinterpret_cast<float*>(base + off_m);
  float* c0 = reinterpret_cast<float*>(base + off_c0);
  dl_result_t res;
  for (uint64_t i = 0; i < n; ++i) { p[i] = 1.0f; m[i] = 0.f; c0[i] = 0.5f; }
  OK(dlh_barrier(h, 1, rank, 0, &res));
  for (int it = 0; it < rounds; ++it) {
    // all-reduce with a rotating inactive rank, normalised
    const int idle = it % world;
    for (uint64_t i = 0; i < n; ++i) g[i] = static_cast<float>(rank + 1);
    OK(dlh_allreduce(h, DL_F32, off_g, n, rank == idle ? 0 : 1, it, 1, 0, &res));
    float want = 0.f;
    for (int r = 0; r < world; ++r) want += r == idle ? 0.f : static_cast<float>(r + 1);
    want /= static_cast<float>(world - 1 > 1 ? world - 1 : 1);
    CHECK(std::fabs(g[0] - want) < 1e-5f && std::fabs(g[n - 1] - want) < 1e-5f, "allreduce round %d: %f != %f", it, g[0], want);
    CHECK(static_cast<int>(res.n_active) == world - 1, "n_active %u", res.n_active);
    // fused step (momentum), everybody active
    for (uint64_t i = 0; i < n; ++i) g[i] = 0.01f * static_cast<float>(rank + 1);
    OK(dlh_fused_sgd(h, off_g, off_p, off_m, n, 0.1f, 0.5f, 0.f, 0, 1, it, 1, 0, &res));
    CHECK(g[0] == 0.f && g[n - 1] == 0.f, "grads not zeroed");
    // elastic average, double-buffered center
    const uint64_t cur = (it & 1) ? off_c1 : off_c0, nxt = (it & 1) ? off_c0 : off_c1;
    OK(dlh_fused_ea(h, off_p, cur, nxt, n, 0.2f, 1, it, 0, &res));
    // broadcast from a rotating root into two buffers
    OK(dlh_broadcast(h, off_p, nxt, n, it % world, it, 0, &res));
  }
  // parameters are bit-identical on all ranks after the broadcast: compare with rank 0's copy
  OK(dlh_barrier(h, 1, 0, 1, &res));
  const float* p0 = reinterpret_cast<const float*>(static_cast<char*>(dlh_base(h, 0)) + off_p);
  CHECK(memcmp(p, p0, n * sizeof(float)) == 0, "parameters differ from rank 0");
  OK(dlh_barrier(h, 1, 0, 1, &res));
  // asynchronous EASGD: rank 0 holds the center, the others sync concurrently (lock-free, then exclusive)
  float* ctr = reinterpret_cast<float*>(static_cast<char*>(dlh_base(h, 0)) + off_ctr);
  if (rank == 0)
    for (uint64_t i = 0; i < 512; ++i) ctr[i] = 100.f;
  OK(dlh_barrier(h, 1, 0, 1, &res));
  if (rank != 0) {
    std::vector<float> x(512, static_cast<float>(rank));
    for (int k = 0; k < 20; ++k) OK(dlh_async_ea_client(h, 0, x.data(), off_ctr, 512, 0.1f, k & 1));
  }
  OK(dlh_barrier(h, 1, 0, 1, &res));
  if (rank == 0) {
    uint64_t cnt = 0;
    OK(dlh_async_sync_count(h, 0, &cnt));
    CHECK(cnt == static_cast<uint64_t>(20 * (world - 1)), "sync count %llu", (unsigned long long)cnt);
  }
  OK(dlh_barrier(h, 1, 0, 1, &res));
  dlh_close(h, 0);
}

static int loader_main() {
  // two worker threads assembling batches into a ring the consumer drains: samplers, uint8 -> float, slot hand-over
  const int n = 1000, elems = 48, batch = 32, depth = 3;
  std::vector<uint8_t> data(static_cast<size_t>(n) * elems);
  std::vector<int64_t> labels(n);
  for (int i = 0; i < n; ++i) {
    labels[i] = i % 10;
    for (int e = 0; e < elems; ++e) data[static_cast<size_t>(i) * elems + e] = static_cast<uint8_t>((i + e) & 255);
  }
  std::vector<std::vector<float>> xs(depth, std::vector<float>(static_cast<size_t>(batch) * elems));
  std::vector<std::vector<int64_t>> ys(depth, std::vector<int64_t>(batch));
  float* xp[depth];
  int64_t* yp[depth];
  for (int d = 0; d < depth; ++d) { xp[d] = xs[d].data(); yp[d] = ys[d].data(); }
  int bad = 0;
  for (int kind = 0; kind < 2; ++kind) {
    void* L = nullptr;
    if (dll_create(data.data(), 1, labels.data(), n, elems, 3, nullptr, nullptr, 10, 1, 2, kind, batch, 7, 2, depth, xp, yp, &L) != 0) {
      fprintf(stderr, "dll_create: %s\n", dll_last_error());
      return 1;
    }
    for (int b = 0; b < 60; ++b) {
      int slot = -1, size = 0;
      int64_t no = -1;
      if (dll_next(L, &slot, &size, &no, 10.0) != 0 || no != b || size <= 0) { ++bad; break; }
      double s = 0;
      for (int i = 0; i < size * elems; ++i) s += xs[slot][i];
      for (int i = 0; i < size; ++i) bad += (ys[slot][i] < 0 || ys[slot][i] > 9);
      if (!(s >= 0)) ++bad;
      dll_release(L);
    }
    dll_destroy(L);
  }
  return bad;
}

int main(int argc, char** argv) {
  const int world = argc > 1 ? atoi(argv[1]) : 4;
  const int rounds = argc > 2 ? atoi(argv[2]) : 25;
  char name[64];
  snprintf(name, sizeof(name), "/dl-tsan-%d", static_cast<int>(getpid()));
  std::vector<std::thread> th;
  for (int r = 0; r < world; ++r) th.emplace_back(rank_main, name, r, world, rounds);
  for (auto& t : th) t.join();
  const int lb = loader_main();
  if (lb) fprintf(stderr, "loader: %d bad batches\n", lb);
  const int fails = g_fail.load() + lb;
  printf("tsan_host: world %d, %d rounds, %d failures\n", world, rounds, fails);
  return fails ? 1 : 0;
}
