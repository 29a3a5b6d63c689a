// Measured NVLink peaks ("measure a P2P copy / multicast peak on the box and use the measured figure").
// Plain streaming kernels over the symmetric heap with nothing else in them, so that the roofline fractions of the
// collectives are taken against what THIS fabric delivers to THIS kind of access, not against a nominal link rate:
//   mode 0  pull      local[dst_off] = peer[src_off]           256-bit loads over NVLink, local stores
//   mode 1  push      peer[dst_off]  = local[src_off]          local loads, 256-bit stores over NVLink
//   mode 2  mc store  multicast[dst_off] = local[src_off]      one multimem.st per 16 bytes, replicated by the switch
//   mode 3  mc reduce local[dst_off] = sum over ranks [src_off] one multimem.ld_reduce per 16 bytes
// plus the copy engine (cudaMemcpyAsync between the local heap and a peer's mapping).  Timed by bench/peak.py.
#include "dl_common.cuh"
#include "dl_host.h"

__global__ void __launch_bounds__(DL_THREADS) peak_kernel(dl_comm_t c, int mode, int peer, uint64_t src_off, uint64_t dst_off,
                                                          unsigned long long n4) {
  constexpr int U = 4;
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * DL_THREADS;
  const unsigned long long tid = static_cast<unsigned long long>(blockIdx.x) * DL_THREADS + threadIdx.x;
  if (mode == 0 || mode == 1) {
    const float4* src = dl_at<float4>(c, mode == 0 ? peer : c.rank, src_off);
    float4* dst = dl_at<float4>(c, mode == 0 ? c.rank : peer, dst_off);
    const unsigned long long np = n4 / 2;
    for (unsigned long long i0 = tid; i0 < np; i0 += stride * U) {
      dl_f8 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u * stride < np) v[u] = dl_ld_f8(src + 2 * (i0 + u * stride));
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u * stride < np) dl_st_f8(dst + 2 * (i0 + u * stride), v[u]);
    }
  } else if (mode == 2) {
    const float4* src = dl_at<float4>(c, c.rank, src_off);
    float4* dst = dl_mc_at<float4>(c, dst_off);
    for (unsigned long long i0 = tid; i0 < n4; i0 += stride * U) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u * stride < n4) v[u] = src[i0 + u * stride];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u * stride < n4) dl_mc_st_f4(dst + i0 + u * stride, v[u]);
    }
  } else {
    const float4* src = dl_mc_at<float4>(c, src_off);
    float4* dst = dl_at<float4>(c, c.rank, dst_off);
    for (unsigned long long i0 = tid; i0 < n4; i0 += stride * 2 * U) {
      float4 v[2 * U];
#pragma unroll
      for (int u = 0; u < 2 * U; ++u)
        if (i0 + u * stride < n4) v[u] = dl_mc_ld_reduce_f4(src + i0 + u * stride);
#pragma unroll
      for (int u = 0; u < 2 * U; ++u)
        if (i0 + u * stride < n4) dst[i0 + u * stride] = v[u];
    }
  }
}

// [src_off, src_off + n floats) -> [dst_off, ...) as described above; n % 8 == 0, offsets 32-byte aligned.
extern "C" int dl_peak_stream(const dl_comm_t* c, int mode, int peer, uint64_t src_off, uint64_t dst_off, uint64_t n, int nblocks,
                              void* stream) {
  int nb = dl_check_comm(c, 0, dl_pick_blocks(n / 8 + 1, 4, nblocks));
  if (nb < 0) return nb;
  DL_REQUIRE(mode >= 0 && mode <= 3 && peer >= 0 && peer < c->world, "dl_peak_stream: bad mode / peer");
  DL_REQUIRE(n % 8 == 0 && src_off % 32 == 0 && dst_off % 32 == 0 && src_off >= DL_PAD_BYTES && dst_off >= DL_PAD_BYTES,
             "dl_peak_stream: n %% 8 == 0 and 32-byte aligned offsets outside the pad required");
  DL_REQUIRE(mode < 2 || (c->mc != nullptr && !c->emulate), "dl_peak_stream: multicast modes need a multicast mapping");
  peak_kernel<<<nb, DL_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(*c, mode, peer, src_off, dst_off, n / 4);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}

// The copy engine: local[dst_off] <- peer[src_off] (pull != 0) or peer[dst_off] <- local[src_off].
extern "C" int dl_peak_memcpy(const dl_comm_t* c, int peer, int pull, uint64_t src_off, uint64_t dst_off, uint64_t bytes, void* stream) {
  int nb = dl_check_comm(c, 0, 1);
  if (nb < 0) return nb;
  DL_REQUIRE(peer >= 0 && peer < c->world, "dl_peak_memcpy: bad peer");
  const char* src = static_cast<const char*>(c->base[pull ? peer : c->rank]) + src_off;
  char* dst = static_cast<char*>(c->base[pull ? c->rank : peer]) + dst_off;
  DL_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream)));
  return DL_OK;
}
