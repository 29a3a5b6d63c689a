/*
 * distlearn_b200 — C ABI of the native library (libdistlearn_b200.so).
 *
 * Everything the Python package does on a GPU goes through these entry points
 * (ctypes), so a C/C++ trainer can use them too.  All functions return 0 on
 * success and a negative DL_E* code on failure; dl_last_error() gives the text.
 *
 * Reference parity is cited as path:line into shanlior/torch-distlearn.
 */
#ifndef DISTLEARN_B200_H
#define DISTLEARN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DL_MAX_RANKS 8      /* one NVSwitch domain (HGX B200) */
#define DL_MAX_BLOCKS 160   /* >= 148 SMs: one persistent CTA per SM */
#define DL_NUM_CHANNELS 4   /* independent signal pads (one per concurrent stream) */
#define DL_PAD_BYTES (256u * 1024u) /* bytes reserved at the start of every heap for signalling */

#define DL_OK 0
#define DL_EINVAL (-1)
#define DL_ECUDA (-2)
#define DL_EUNSUPPORTED (-3)
#define DL_ESYS (-4)
#define DL_ETIMEOUT (-5)

#define DL_EPROTO (-6) /* DL_FLAG_DEBUG: a barrier slot held an epoch the protocol cannot produce */

/* collective algorithm variants (chosen per message size by the variant table) */
#define DL_VARIANT_ONESHOT 0 /* every rank reads all peers, reduces in rank order          */
#define DL_VARIANT_TWOSHOT 1 /* owner-computes reduce-scatter + push all-gather over P2P   */
#define DL_VARIANT_NVLS 2    /* multimem.ld_reduce + multimem.st through the NVSwitch      */
#define DL_VARIANT_LL 3      /* small messages: every rank pushes {data, flag} packets into its peers' staging
                                slots and reduces what arrives - one NVLink crossing, no barrier at all */
#define DL_VARIANT_HYBRID 4  /* large messages: software-pipelined slice pass, part of every slice reduced in the
                                switch (multimem) and the rest over P2P loads in the same launch; nvls_permille
                                in dl_comm_t sets the split (0 = all P2P, 1000 = all NVLS) */

#define DL_VARIANT_LIGHT 5   /* fused_sgd only: the hybrid pass in small CTAs (128 threads, <= 80 registers) that fit on an SM
                                beside the tensor-core GEMM's CTA - the variant gradient buckets use under backward */

/* dl_comm_t.flags */
#define DL_FLAG_DEBUG 1u     /* barriers check epoch monotonicity / slot contents and report DL_EPROTO */

/* data types of the plain all-reduce */
#define DL_F32 0
#define DL_I32 1
#define DL_I64 2
#define DL_F64 3

/*
 * A communicator as the kernels see it: the same heap mapped from every rank.
 * base[r] is rank r's heap as addressable from THIS process (base[rank] is the
 * local allocation); mc is the multicast (NVLS) mapping of the same heap or
 * NULL.  The first DL_PAD_BYTES of every heap are the signal pad.
 * With emulate != 0 all `world` heaps live on one GPU and one cooperative
 * launch plays every rank (gridDim.y = world) — the single-GPU test mode.
 */
typedef struct dl_comm {
  int32_t rank;
  int32_t world;
  int32_t emulate;
  uint32_t flags;      /* DL_FLAG_* */
  void* base[DL_MAX_RANKS];
  void* mc;
  uint64_t timeout_ns; /* bounded spin: 0 = default 20 s */
  /* Small-message (DL_VARIANT_LL) staging area inside every heap: DL_NUM_CHANNELS x 2 parities x world source
   * slots of ll_slot_bytes each, starting at heap offset ll_off (0 = not set up).  A slot holds one 16-byte
   * header packet plus two 16-byte packets per 16 bytes of payload; it must start out zeroed. */
  uint64_t ll_off;
  uint64_t ll_slot_bytes;
  /* Device-visible address of pinned host memory (dl_host_words_t) or NULL: kernels leave n_active there under the
   * caller's ticket, and the sticky error word, so the host can read both without a device-to-host copy. */
  void* host_words;
  uint32_t ticket;       /* caller's sequence number for this call (host_words ring index) */
  int32_t nvls_permille; /* DL_VARIANT_HYBRID: share of every slice that goes through the switch */
} dl_comm_t;

/* What kernels write into pinned host memory (one per communicator). */
#define DL_HOST_RING 64
typedef struct dl_host_words {
  uint32_t error;                            /* sticky: 0, or the positive DL_E* code of the first failure */
  uint32_t error_channel;
  uint32_t reserved[2];
  struct { uint32_t ticket, n_active; } ring[DL_NUM_CHANNELS][DL_HOST_RING];
} dl_host_words_t;

/* Per-call result words the kernels leave in the local signal pad. */
typedef struct dl_result {
  uint32_t payload[DL_MAX_RANKS]; /* what each rank published at the arrival barrier */
  uint32_t n_active;              /* number of ranks that contributed               */
  uint32_t error;                 /* 0, or DL_ETIMEOUT (positive) if a spin gave up */
} dl_result_t;

const char* dl_last_error(void);
int dl_version(void);
/* Number of kernels this library has launched so far in this process (host-side count). */
unsigned long long dl_kernel_launches(void);

/* ---- symmetric heap: CUDA VMM + POSIX-fd export (+ multicast), cudaIpc fallback ---- */
/* Replaces libipc's socket transport + CUDA-IPC hint (examples/mnist.lua:93). */
int dl_device_caps(int dev, int* vmm, int* fd_handles, int* multicast);
int dl_vmm_granularity(int dev, size_t* gran);
int dl_vmm_alloc(int dev, size_t bytes, uint64_t* handle, int* fd);
int dl_vmm_import(int fd, uint64_t* handle);
int dl_vmm_map(int dev, uint64_t handle, size_t bytes, void** ptr);
int dl_vmm_unmap(void* ptr, size_t bytes);
int dl_vmm_release(uint64_t handle);
int dl_mc_granularity(int ndev, size_t bytes, size_t* gran);
int dl_mc_create(int ndev, size_t bytes, uint64_t* handle, int* fd);
int dl_mc_add_device(uint64_t handle, int dev);
int dl_mc_bind(uint64_t mc_handle, uint64_t mem_handle, size_t bytes);
int dl_mc_unbind(uint64_t mc_handle, int dev, size_t bytes);
int dl_ipc_alloc(int dev, size_t bytes, void** ptr, void* handle64);
int dl_ipc_open(int dev, const void* handle64, void** ptr);
int dl_ipc_close(void* ptr);
int dl_ipc_free(void* ptr);
int dl_enable_peer(int dev, int peer);
int dl_can_access_peer(int dev, int peer, int* can);
int dl_memset_sync(int dev, void* ptr, int value, size_t bytes);
/* unix-domain-socket fd passing (SCM_RIGHTS) used to move VMM handles between local ranks */
int dl_uds_listen(const char* name, int* sock);
int dl_uds_send_fd(const char* name, int fd, int tag, int timeout_ms);
int dl_uds_recv_fd(int sock, int* fd, int* tag, int timeout_ms);
int dl_uds_close(int sock);
int dl_close_fd(int fd);

/* ---- collectives (all offsets are byte offsets into the symmetric heap) ---- */

/* In-place all-reduce(sum) of n elements at heap offset `off`.  With
 * scale_by_inv_active != 0 the f32 result is divided by the number of ranks
 * that passed active != 0 (lua/AllReduceSGD.lua:20-27); inactive ranks
 * contribute zeros (the drain all-reduce of lua/AllReduceSGD.lua:37).
 * `payload` (31 bits) is published to every peer together with `active`. */
int dl_allreduce(const dl_comm_t* c, int dtype, uint64_t off, uint64_t n, int variant, int active,
                 uint32_t payload, int scale_by_inv_active, int nblocks, int channel, void* stream);

/* Fused gradient all-reduce + 1/n_active + (momentum, weight decay) SGD update
 * + parameter all-gather: grads at g_off are summed over active ranks, the
 * owner of each slice updates params at p_off and pushes the new slice to every
 * rank; grads are zeroed afterwards when zero_grads != 0.
 * Replaces lua/AllReduceSGD.lua:18-30 followed by examples/cifar10.lua:187-191. */
int dl_fused_sgd(const dl_comm_t* c, uint64_t g_off, uint64_t p_off, uint64_t m_off, uint64_t n,
                 float lr, float momentum, float weight_decay, int nesterov, int variant,
                 int active, uint32_t payload, int zero_grads, int nblocks, int channel,
                 void* stream);

/* Fused elastic averaging (lua/AllReduceEA.lua:34-45): c_new = c + sum_i a(x_i - c) is computed by
 * the slice owner from peers' x and written to every replica at cnew_off, then
 * x -= a (x - c_old) locally.  One launch. */
int dl_fused_ea(const dl_comm_t* c, uint64_t x_off, uint64_t c_off, uint64_t cnew_off, uint64_t n,
                float alpha, int variant, int active, uint32_t payload, int nblocks, int channel,
                void* stream);

/* Broadcast n floats at `off` from `root` to every rank (tree.scatter,
 * lua/AllReduceSGD.lua:52, and the winner broadcast :43-47); when off2 != 0 the
 * same values are also written at off2 (EA: center <- params,
 * lua/AllReduceEA.lua:93-99). */
int dl_broadcast(const dl_comm_t* c, uint64_t off, uint64_t off2, uint64_t n, int root, int variant,
                 uint32_t payload, int nblocks, int channel, void* stream);

/* Pinned, device-mapped host memory for dl_comm_t.host_words (one dl_host_words_t, zeroed).  With it set, every
 * collective leaves its contributor count under dl_comm_t.ticket in ring[channel][ticket % DL_HOST_RING] and a
 * failed spin sets the sticky `error` word, both readable by the host with plain loads: the `n` that
 * lua/AllReduceSGD.lua:20-29 returns costs no device-to-host copy. */
int dl_host_words_alloc(void** host_ptr, void** dev_ptr);
int dl_host_words_free(void* host_ptr);

/* Bandwidth probes behind the roofline denominators (csrc/peak.cu, bench/peak.py; no counterpart in the reference, which
 * publishes no numbers): stream n floats between the local heap and `peer`'s (mode 0 pull, 1 push) or through the
 * multicast mapping (mode 2 multimem.st, 3 multimem.ld_reduce); dl_peak_memcpy is the same transfer on the copy engine. */
int dl_peak_stream(const dl_comm_t* c, int mode, int peer, uint64_t src_off, uint64_t dst_off, uint64_t n, int nblocks,
                   void* stream);
int dl_peak_memcpy(const dl_comm_t* c, int peer, int pull, uint64_t src_off, uint64_t dst_off, uint64_t bytes, void* stream);

/* Memory-model litmus tests of the signalling patterns on the real fabric (csrc/litmus.cu; the reference has no
 * counterpart - it synchronises with blocking sockets).  shape 0 = message passing around the ring of ranks (data
 * stores, release fence, flag | poll, acquire fence, data loads), shape 1 = store buffering with fence.sc.sys;
 * fenced = 0 drops the fences (control run).  Block b of a rank plays against block b of its neighbours with a random
 * __nanosleep skew per iteration.  scratch: dl_litmus_scratch_bytes(iterations, nblocks) zeroed bytes at scratch_off on
 * every rank; afterwards its first words hold {checked, violations (MP), blocks that timed out} as u64, and for SB the
 * per-iteration loaded values follow per block (see csrc/litmus.cu). */
size_t dl_litmus_scratch_bytes(int iterations, int nblocks);
int dl_litmus(const dl_comm_t* c, int shape, int fenced, uint64_t scratch_off, int iterations, uint32_t seed, int nblocks,
              void* stream);

/* A barrier only: publishes payload/active, fills dl_result. */
int dl_barrier(const dl_comm_t* c, int active, uint32_t payload, int channel, void* stream);

/* Copy the dl_result of `channel` to host (synchronises `stream`). */
int dl_read_result(const dl_comm_t* c, int channel, dl_result_t* out, void* stream);
/* Zero this rank's signal pad (call on every rank, then barrier on the host). */
int dl_reset_pad(const dl_comm_t* c);

/* ---- asynchronous EASGD (lua/AsyncEA.lua) ---- */
/* Client side of one sync (lua/AsyncEA.lua:134-146 and :163-228 on the server):
 * c = center on `server`'s heap (P2P loads), d = alpha (x - c), x -= d (local),
 * center += d (red.add over NVLink).  exclusive != 0 serialises clients with a
 * ticket lock in the server's pad (the recvAny mutex, lua/AsyncEA.lua:163-177).
 * x is a plain device pointer; the center lives at c_off in the server heap. */
int dl_async_ea_client(const dl_comm_t* c, int server, float* x, uint64_t c_off, uint64_t n,
                       float alpha, int exclusive, int nblocks, void* stream);
/* Same, but the moved parameters x - d are written to x_out and x is left untouched.  d depends only on x and the
 * center, so a trainer can run the sync on a side stream UNDER the forward/backward pass that still reads x, and apply
 * the gradient afterwards from x_out (dl_local_sgd_from): the reference's order "gradient at x, elastic move, apply
 * the gradient" (examples/EASGD_client.lua:106-117) with the exchange hidden behind the compute. */
int dl_async_ea_client_out(const dl_comm_t* c, int server, float* x, float* x_out, uint64_t c_off, uint64_t n,
                           float alpha, int exclusive, int nblocks, void* stream);
/* Copy n floats from `src_rank`'s heap at off into dst (P2P loads): initClient /
 * startTest center pull (lua/AsyncEA.lua:64-78, :268-285). */
int dl_peer_pull(const dl_comm_t* c, int src_rank, uint64_t off, float* dst, uint64_t n,
                 int nblocks, void* stream);
/* Read the server's sync counter (number of completed client syncs). */
int dl_async_sync_count(const dl_comm_t* c, int server, uint64_t* count, void* stream);

/* Sync counter (out[0]) and all 32 mailbox words (out[1..32]) of rank `rank` in one copy: what a server or tester
 * polls while it waits (lua/AsyncEA.lua:163-177 blocks in recvAny instead). */
int dl_async_ctl_snapshot(const dl_comm_t* c, int rank, uint64_t* out, void* stream);
/* Host-visible 64-bit mailbox words in rank `rank`'s pad (index < 32): the tester handshake
 * "Test?"/"Ack" (lua/AsyncEA.lua:239-258,268-292) and ready/done flags travel here. */
int dl_mailbox_write(const dl_comm_t* c, int rank, int index, uint64_t value, void* stream);
int dl_mailbox_read(const dl_comm_t* c, int rank, int index, uint64_t* value, void* stream);

/* ---- tensor-core GEMM (tcgen05 + TMEM + TMA), fp32 in / fp32 out ---- */
/* C[M,N] (+)= A[M,K] * B[N,K]^T (+ bias[N]); A, B, C row-major fp32 with K contiguous in A and B
 * (leading dimensions in elements, multiples of 4).  terms = 3: 3xTF32 split with fp32-level accuracy
 * (the reference is strict fp32, examples/cifar10.lua:136; its convolutions are im2col + SGEMM,
 * examples/Model.lua:20); terms = 1: plain TF32.  split_k > 1 splits K over blockIdx.z and adds the
 * partial tiles with red.add (C is zeroed first unless accumulate != 0, which adds into C as is). */
int dl_tc_gemm(const float* A, const float* B, float* C, const float* bias, int M, int N, int K, long long lda,
               long long ldb, long long ldc, int split_k, int accumulate, int terms, void* stream);
/* Same, with either operand optionally "MN-major": a_mn != 0 means A is stored as [K, M] row-major
 * (M contiguous), b_mn != 0 means B is stored as [K, N] row-major.  This is what the weight-gradient
 * GEMM  dW[Cout, Kc] = dY[M, Cout]^T * col[M, Kc]  and the data-gradient GEMM  dcol = dY * W  need,
 * without materialising any transpose. */
int dl_tc_gemm_ex(const float* A, const float* B, float* C, const float* bias, int M, int N, int K, long long lda,
                  long long ldb, long long ldc, int a_mn, int b_mn, int split_k, int accumulate, int terms,
                  void* stream);

/* Implicit-GEMM 5x5 / stride 1 / pad 2 convolution on NHWC fp32 tensors (Cin, Cout multiples of 32): the
 * tcgen05 GEMM above with its TMA producer reading the activation tensor in 4-D boxes shifted by the filter
 * tap (zero fill = padding), so no im2col / col2im buffer exists.  Replaces SpatialConvolutionMM forward and
 * both of its backward passes (examples/Model.lua:26,32,38).
 *   mode 1 forward: act = x [B,H,W,Cin],  w = W [Cout, 25*Cin] ((kh,kw,cin) columns), out = y [B*H*W, Cout] (+bias)
 *   mode 2 dgrad  : act = dy [B,H,W,Cout], w = W,                                      out = dx [B*H*W, Cin]
 *   mode 3 wgrad  : act = x [B,H,W,Cin],  w = dy [B*H*W, Cout],                        out = dW [Cout, 25*Cin] */
int dl_tc_conv5x5(int mode, const float* act, const float* w, float* out, const float* bias, int B, int H, int W,
                  int Cin, int Cout, int split_k, int accumulate, int terms, void* stream);
/* The same with the weights also given as their TF32 split (w_hi = w rounded to TF32, w_lo = w - w_hi, both
 * [Cout, 25*Cin], from dl_split_tf32): forward and dgrad of the 3xTF32 stream-K kernel then load both halves by TMA and
 * its converter warps only handle the activations (less shared-memory traffic; same bits).  w_hi = w_lo = NULL, mode 3 and
 * terms 1 behave exactly like dl_tc_conv5x5. */
int dl_tc_conv5x5_ws(int mode, const float* act, const float* w, const float* w_hi, const float* w_lo, float* out,
                     const float* bias, int B, int H, int W, int Cin, int Cout, int split_k, int accumulate, int terms,
                     void* stream);
/* hi[i] = w[i] rounded to TF32, lo[i] = w[i] - hi[i]; n a multiple of 4, 16-byte aligned pointers. */
int dl_split_tf32(const float* w, float* hi, float* lo, long long n, void* stream);

/* Stream-K form of the tensor-core GEMM — one persistent CTA per SM takes a contiguous range of (tile, K-block) units,
 * so 256 tiles on 148 SMs cost 86.5 K-block times instead of 100 (csrc/tc_gemm_streamk.inl).  on = 1 / 0 switches it
 * for later dl_tc_gemm* / dl_tc_conv5x5 calls (initial value: 1 unless the environment says DL_B200_TC_STREAMK=0);
 * returns the previous setting. */
int dl_tc_gemm_streamk(int on);
/* Host-side view of the stream-K decomposition (same code as the kernel): the segments (m0, n0, first K block, K blocks)
 * of CTA `cta`; returns their number, *grid = CTAs the launch would use on `sms` SMs. */
int dl_tc_streamk_plan(int M, int N, int K, int sms, int cta, int* grid, int* segments, int max_segments);

/* 5x5 / stride 1 / pad 2 convolution pieces on NHWC activations (SpatialConvolutionMM = im2col + GEMM,
 * examples/Model.lua:20-38): col[B*H*W, ld] with columns ordered (kh, kw, cin); col2im sums the 25 taps
 * back into dx[B,H,W,C]. */
int dl_im2col5x5(const float* x, float* col, int B, int H, int W, int C, long long ld, void* stream);
int dl_col2im5x5(const float* dcol, float* dx, int B, int H, int W, int C, long long ld, void* stream);

/* Fused training-mode batch-norm + ReLU + 2x2/2 max-pool on NHWC fp32 activations, forward and backward
 * (SpatialBatchNormalization(eps 1e-3) -> ReLU -> SpatialMaxPooling(2,2,2,2), examples/Model.lua:21-24).
 * y [B,H,W,C] -> p [B,H/2,W/2,C]; mean/invstd [C] are outputs of the forward call and inputs of the backward
 * one; running_mean/var (may be NULL) are updated like nn.BatchNorm2d.  work: dl_bn_work_bytes(C) bytes. */
size_t dl_bn_work_bytes(int C);
int dl_bn_relu_pool_fwd(const float* y, float* p, const float* gamma, const float* beta, float* mean, float* invstd,
                        float* running_mean, float* running_var, int B, int H, int W, int C, float eps, float momentum,
                        void* work, void* stream);
int dl_bn_relu_pool_bwd(const float* y, const float* dp, float* dy, const float* gamma, const float* beta, const float* mean,
                        const float* invstd, float* dgamma, float* dbeta, int B, int H, int W, int C, int accumulate,
                        void* work, void* stream);
/* dl_bn_relu_pool_bwd that also produces the bias gradient of the convolution in front of the block (dconvbias[C] (+)=
 * column sums of dy; NULL: skip) in the same pass over dy — examples/Model.lua:20-24: conv -> BN -> ReLU -> pool. */
int dl_bn_relu_pool_bwd_bias(const float* y, const float* dp, float* dy, const float* gamma, const float* beta, const float* mean,
                             const float* invstd, float* dgamma, float* dbeta, float* dconvbias, int B, int H, int W, int C,
                             int accumulate, void* work, void* stream);
/* out[C] (+)= column sums of x[M, C] (a convolution's bias gradient); work: dl_bn_work_bytes(C) bytes.
 * accumulate != 0 (here and in dl_bn_relu_pool_bwd for dgamma/dbeta) adds into the buffer: the autograd functions
 * pass the parameters' .grad views so that no separate gradient-accumulation kernel runs. */
int dl_colsum(const float* x, float* out, long long M, int C, int accumulate, void* work, void* stream);

/* Classifier head: Reshape -> Linear(HW*C, J) -> LogSoftMax -> ClassNLLCriterion (examples/Model.lua:44-52) on an NHWC
 * feature map feat[B, HW, C]; w[J, HW*C] keeps the reference's (c, h, w) column order.  J = 10 or 16.
 * forward: logp[B, J], loss[1] = mean NLL (rowloss: B floats of scratch).  backward: gloss = device pointer to the
 * upstream gradient of the loss; dfeat[B, HW, C], dw[J, HW*C], db[J] (accumulate != 0: added into dw / db). */
int dl_head_fwd(const float* feat, const float* w, const float* bias, const int64_t* target, float* logp, float* loss,
                float* rowloss, int B, int HW, int C, int J, void* stream);
int dl_head_bwd(const float* feat, const float* w, const float* logp, const int64_t* target, const float* gloss, float* dfeat,
                float* dw, float* db, int B, int HW, int C, int J, int accumulate, void* stream);

/* One training step of the reference's MNIST network (examples/mnist.lua:53-81: conv 1->16 5x5, tanh, maxpool, conv 16->16
 * 5x5, tanh, maxpool, Linear 400->10, logSoftMax + NLL) as ONE launch of an 8-CTA thread-block cluster (csrc/mnist_step.cu):
 * forward, backward and - with apply_sgd != 0 and B <= DL_MNIST_FUSED_MAX_BATCH - the update p -= lr * g of
 * examples/mnist.lua:112-116.  x [B,1024] fp32, y [B] int64; params / grads: the six tensors in module order (conv1.weight,
 * conv1.bias, conv2.weight, conv2.bias, linear.weight, linear.bias).  Without the fused update the gradient of the mean
 * loss is ADDED into grads.  loss: 1 float; logp: [B,10] log-probabilities or NULL. */
#define DL_MNIST_FUSED_MAX_BATCH 8
int dl_mnist_step(const float* x, const int64_t* y, float* const* params, float* const* grads, float* loss, float* logp, int B,
                  float lr, int apply_sgd, void* stream);
/* Which form of the one-launch MNIST kernel runs: 2 (default: 416 threads, csrc/mnist_step.cu) or 1 (the first form, kept as
 * the reference for it); returns the previous value, any other argument only reads it.  DL_B200_MNIST_FORM=1 sets it at load. */
int dl_mnist_form(int form);

/* ---- local flat-vector ops (single GPU, no communication) ---- */
/* p -= lr * (g*scale [+ wd*p] [momentum]); optionally zero g. world==1 fast path. */
int dl_local_sgd(float* p, float* g, float* m, uint64_t n, float lr, float momentum,
                 float weight_decay, int nesterov, float grad_scale, int zero_grads, void* stream);
/* Same update, reading the old parameters from p_in and writing the new ones to p (p_in may equal p). */
int dl_local_sgd_from(float* p, const float* p_in, float* g, float* m, uint64_t n, float lr, float momentum,
                      float weight_decay, int nesterov, float grad_scale, int zero_grads, void* stream);
/* x -= alpha (x - c); c += world_scale * alpha (x_old - c)  (single-rank EA step) */
int dl_local_ea(float* x, float* c, uint64_t n, float alpha, void* stream);
/* Deterministic sum of (a[i]-b[i])^2 (b may be NULL) into out[0] (double).
 * scratch: caller-provided device buffer of dl_sumsq_scratch_bytes() bytes,
 * zero-initialised once; the kernel leaves it ready for the next call. */
size_t dl_sumsq_scratch_bytes(void);
int dl_sumsq_diff(const float* a, const float* b, uint64_t n, double* out, void* scratch,
                  void* stream);
/* dl_sumsq_diff for callers that cannot allocate device memory.  (No counterpart in the
 * reference: it has no metrics kernels and no allocation-free mode.)
 * Meant for hosts that cannot supply dl_sumsq_diff's scratch buffer: jobs that hand all of HBM to
 * their own pool at start-up and forbid driver allocations afterwards, monitoring threads that
 * must stay out of the allocator, C trainers that link this library without torch.  The scratch
 * is a fixed dl_sumsq_static_scratch_bytes() = 2,560-byte array inside the library, held on every
 * GPU that loads it until the process exits; nothing is allocated or freed by this call, by the
 * CUDA runtime on its behalf, or by torch.  a, b (may be NULL) and out are the caller's.
 *  - n may not exceed DL_SUMSQ_STATIC_MAX_N elements; larger calls fail with DL_EINVAL and a
 *    message, and should use dl_sumsq_diff with their own scratch.
 *  - The first call on each device (with that device current) looks up the array's device
 *    address; make that first call before starting any stream capture.
 *  - All calls on one GPU share the array.  Each call re-zeroes the one word that has to start at
 *    zero and overwrites the rest before reading it, so stale contents are harmless; but two
 *    calls that may run at the same time on the same GPU (different streams, different threads)
 *    race on it and produce garbage.  Keep them on one stream or order them with events.
 *  - Same arithmetic and block order as dl_sumsq_diff: results agree bit for bit. */
#define DL_SUMSQ_STATIC_MAX_N (1ull << 31)
size_t dl_sumsq_static_scratch_bytes(void);
uint64_t dl_sumsq_static_max_n(void);
int dl_sumsq_diff_static(const float* a, const float* b, uint64_t n, double* out, void* stream);
/* 10x10-style confusion matrix accumulation: mat[target][argmax(logits)] += 1 */
int dl_confusion_add(const float* logits, const int64_t* target, int batch, int classes,
                     int64_t* mat, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DISTLEARN_B200_H */
