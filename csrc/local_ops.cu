// Single-GPU helper kernels used by metrics/logging: deterministic squared distance between two
// flat vectors (|x - center|^2 is the elastic-averaging "how far did we drift" metric, |g|^2 the
// gradient norm) and confusion-matrix accumulation (optim.ConfusionMatrix:add in
// examples/cifar10.lua:194-196, examples/mnist.lua:119).
#include "dl_common.cuh"
#include "dl_host.h"

#define DL_SUMSQ_BLOCKS 296  // 2 CTAs per SM on 148 SMs
#define DL_SUMSQ_THREADS 256

struct SumsqScratch {
  double partial[DL_SUMSQ_BLOCKS];
  unsigned int arrived;  // must be zero before a launch; the last block leaves it zero again
  unsigned int pad[3];
};

__global__ void __launch_bounds__(DL_SUMSQ_THREADS) sumsq_diff_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                                    unsigned long long n, double* out, SumsqScratch* s) {
  __shared__ double warp_sum[DL_SUMSQ_THREADS / 32];
  __shared__ bool last;
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_SUMSQ_THREADS;
  double acc = 0.0;
  for (unsigned long long i = static_cast<unsigned long long>(blockIdx.x) * DL_SUMSQ_THREADS + threadIdx.x; i < n; i += stride) {
    const float d = b ? a[i] - b[i] : a[i];
    acc += static_cast<double>(d) * static_cast<double>(d);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < DL_SUMSQ_THREADS / 32; ++w) t += warp_sum[w];
    s->partial[blockIdx.x] = t;
    __threadfence();
    last = (atomicAdd(&s->arrived, 1u) + 1u == gridDim.x);
  }
  __syncthreads();
  if (last && threadIdx.x == 0) {
    __threadfence();
    double t = 0.0;
    for (unsigned int k = 0; k < gridDim.x; ++k) t += reinterpret_cast<volatile double*>(s->partial)[k];  // fixed order
    *out = t;
    s->arrived = 0u;
  }
}

extern "C" size_t dl_sumsq_scratch_bytes(void) { return sizeof(SumsqScratch); }

static int sumsq_blocks(uint64_t n) {
  unsigned long long nb = (n + DL_SUMSQ_THREADS * 8ull - 1) / (DL_SUMSQ_THREADS * 8ull);
  if (nb < 1) nb = 1;
  if (nb > DL_SUMSQ_BLOCKS) nb = DL_SUMSQ_BLOCKS;
  return static_cast<int>(nb);
}

extern "C" int dl_sumsq_diff(const float* a, const float* b, uint64_t n, double* out, void* scratch, void* stream) {
  DL_REQUIRE(a && out && scratch, "dl_sumsq_diff: null pointer");
  DL_REQUIRE(reinterpret_cast<uintptr_t>(scratch) % 8 == 0, "dl_sumsq_diff: scratch must be 8-byte aligned");
  sumsq_diff_kernel<<<sumsq_blocks(n), DL_SUMSQ_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
      a, b, n, out, static_cast<SumsqScratch*>(scratch));
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}

// mat[target][argmax(logits)] += 1, one thread per sample
__global__ void confusion_add_kernel(const float* __restrict__ logits, const long long* __restrict__ target, int batch,
                                     int classes, unsigned long long* mat) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= batch) return;
  const float* row = logits + static_cast<long long>(i) * classes;
  int best = 0;
  float bv = row[0];
  for (int k = 1; k < classes; ++k) {
    const float v = row[k];
    if (v > bv) { bv = v; best = k; }
  }
  const long long t = target[i];
  if (t >= 0 && t < classes) atomicAdd(&mat[t * classes + best], 1ull);
}

extern "C" int dl_confusion_add(const float* logits, const int64_t* target, int batch, int classes, int64_t* mat,
                                void* stream) {
  DL_REQUIRE(logits && target && mat && batch >= 0 && classes > 0, "dl_confusion_add: bad arguments");
  if (batch == 0) return DL_OK;
  confusion_add_kernel<<<(batch + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      logits, reinterpret_cast<const long long*>(target), batch, classes, reinterpret_cast<unsigned long long*>(mat));
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}

// ---- dl_sumsq_diff_static: the same reduction for callers that cannot allocate -----------------
// Some hosts of this library have no way to hand dl_sumsq_diff() a scratch buffer: a job that
// gave every byte of HBM to its own pool at start-up and treats any later driver allocation as a
// bug (cudaMalloc can fail there, and it synchronises the whole device when it succeeds); a
// metrics or watchdog thread that must not touch the allocator while the training thread owns
// it; a plain-C trainer that links this .so without torch and so has no caching allocator to
// borrow from.  For them the scratch lives in the library image itself: one fixed array per GPU,
// resident from the moment the library is loaded on that GPU until the process exits.
//
// DL_SUMSQ_STATIC_SCRATCH_BYTES (2,560 bytes per GPU) is all there is.  It holds the partial sums
// of the largest grid this reduction ever launches, and the entry point refuses vectors longer
// than DL_SUMSQ_STATIC_MAX_N elements (include/distlearn_b200.h) instead of growing anything.
#define DL_SUMSQ_STATIC_SCRATCH_BYTES 2560  // >= sizeof(SumsqScratch), multiple of 256
static_assert(sizeof(SumsqScratch) <= DL_SUMSQ_STATIC_SCRATCH_BYTES, "static scratch too small");

static __device__ __align__(256) unsigned char g_sumsq_static_scratch[DL_SUMSQ_STATIC_SCRATCH_BYTES];

#include <mutex>

// In host code `g_sumsq_static_scratch` is only a handle for the runtime, not an address the GPU
// understands; the device address differs per GPU and is looked up once per device.
static void* g_sumsq_static_ptr[64] = {nullptr};
static std::mutex g_sumsq_static_mu;

static int sumsq_static_scratch(void** out) {
  int dev = -1;
  DL_CUDA(cudaGetDevice(&dev));
  DL_REQUIRE(dev >= 0 && dev < 64, "dl_sumsq_diff_static: device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_sumsq_static_mu);
  if (g_sumsq_static_ptr[dev] == nullptr) {
    void* p = nullptr;
    DL_CUDA(cudaGetSymbolAddress(&p, g_sumsq_static_scratch));
    DL_REQUIRE(p != nullptr, "dl_sumsq_diff_static: no device address for the static scratch on device %d", dev);
    g_sumsq_static_ptr[dev] = p;
  }
  *out = g_sumsq_static_ptr[dev];
  return DL_OK;
}

extern "C" size_t dl_sumsq_static_scratch_bytes(void) { return DL_SUMSQ_STATIC_SCRATCH_BYTES; }
extern "C" uint64_t dl_sumsq_static_max_n(void) { return DL_SUMSQ_STATIC_MAX_N; }

extern "C" int dl_sumsq_diff_static(const float* a, const float* b, uint64_t n, double* out, void* stream) {
  DL_REQUIRE(a && out, "dl_sumsq_diff_static: null pointer");
  DL_REQUIRE(n <= DL_SUMSQ_STATIC_MAX_N,
             "dl_sumsq_diff_static: n = %llu exceeds the supported maximum of %llu elements; "
             "use dl_sumsq_diff with a caller-provided scratch buffer",
             (unsigned long long)n, (unsigned long long)DL_SUMSQ_STATIC_MAX_N);
  const int nb = sumsq_blocks(n);  // <= DL_SUMSQ_BLOCKS, whose partial sums the array holds (static_assert above)
  void* scratch = nullptr;
  const int rc = sumsq_static_scratch(&scratch);
  if (rc != DL_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // The array keeps whatever the previous call (or an aborted one) left behind.  The partial sums
  // are overwritten before they are read; the arrival counter is the one word that must be zero.
  SumsqScratch* s = static_cast<SumsqScratch*>(scratch);
  DL_CUDA(cudaMemsetAsync(&s->arrived, 0, sizeof(s->arrived), st));
  sumsq_diff_kernel<<<nb, DL_SUMSQ_THREADS, 0, st>>>(a, b, n, out, s);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}
