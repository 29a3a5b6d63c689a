// Host-side helpers shared by the launchers: error text, CUDA error mapping, launch geometry.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

#include "../include/distlearn_b200.h"

void dl_set_error(const char* fmt, ...);
// Every kernel launch of the library is counted (host side) so that callers can report how many of OUR kernels ran.
void dl_note_launch();

#define DL_CUDA(expr)                                                                      \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess) {                                                               \
      dl_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));   \
      return DL_ECUDA;                                                                     \
    }                                                                                      \
  } while (0)

#define DL_REQUIRE(cond, ...)      \
  do {                             \
    if (!(cond)) {                 \
      dl_set_error(__VA_ARGS__);   \
      return DL_EINVAL;            \
    }                              \
  } while (0)

// Validate a communicator + channel + block count; returns the clamped block count or <0.
int dl_check_comm(const dl_comm_t* c, int channel, int nblocks);
// Number of SMs of the current device (cached per device).
int dl_sm_count();
// Default block count for `work_items` items of per-thread work.
int dl_pick_blocks(unsigned long long work_items, int per_thread, int requested);

// DL_VARIANT_LL forms of the collectives (csrc/ll_ops.cu); reached through the public entry points.
int dl_ll_allreduce(const dl_comm_t* c, int dtype, uint64_t off, uint64_t nbytes, int active, uint32_t payload, int scale,
                    int nblocks, int channel, cudaStream_t st);
int dl_ll_sgd(const dl_comm_t* c, uint64_t g_off, uint64_t p_off, uint64_t m_off, uint64_t n, float lr, float momentum, float wd,
              int nesterov, int active, uint32_t payload, int zero_grads, int nblocks, int channel, cudaStream_t st);
int dl_ll_ea(const dl_comm_t* c, uint64_t x_off, uint64_t c_off, uint64_t cnew_off, uint64_t n, float alpha, int active,
             uint32_t payload, int nblocks, int channel, cudaStream_t st);
int dl_ll_broadcast(const dl_comm_t* c, uint64_t off, uint64_t off2, uint64_t n, int root, uint32_t payload, int nblocks, int channel,
                    cudaStream_t st);

// Launch `kernel` normally, or cooperatively with gridDim.y = world when emulating ranks.
template <typename... Args>
int dl_launch(void (*kernel)(Args...), const dl_comm_t* c, int nblocks, int threads,
              cudaStream_t stream, Args... args) {
  dim3 grid(nblocks, c->emulate ? c->world : 1, 1);
  if (c->emulate && c->world > 1) {
    void* argv[] = {(void*)&args...};
    DL_CUDA(cudaLaunchCooperativeKernel((const void*)kernel, grid, dim3(threads), argv, 0, stream));
  } else {
    kernel<<<grid, threads, 0, stream>>>(args...);
    DL_CUDA(cudaGetLastError());
  }
  dl_note_launch();
  return DL_OK;
}
