// Stream-K ("persistent") form of the TS GEMM kernel — the default since it was validated on a B200 (round 2:
// tests/test_tc_streamk_gpu.py, 24 cases against the one-tile-per-CTA kernel and float64; flagship step 1.155 -> 1.132 ms,
// profiles/bench_r2.md).  DL_B200_TC_STREAMK=0 or dl_tc_gemm_streamk(0) goes back to one tile per CTA.
//
// Why: one CTA per SM and 128 x BN tiles leave a tail.  The layer-2 forward GEMM has 256 tiles: 148 SMs run them as
// two waves of 50 K-blocks, 100 K-block times for 86.5 K-block times of work (profiles/ncu_tc_gemm.md: "27 % second-
// wave tail"); every layer of the CIFAR net has a tile count like that.  Here the work is cut in units of one K block
// (all tiles x all K blocks, tile-major) and CTA c of G = #SMs takes the contiguous range [c U / G, (c+1) U / G).  A
// range crosses tile boundaries: the CTA runs its segments (tile, K-block range) back to back through the SAME
// pipeline — ring stages, mbarrier phases and the two TMEM accumulator buffers simply keep counting — so the epilogue
// of one segment overlaps the main loop of the next.  A segment that covers its whole tile stores it; a partial one
// adds it with red.add into C, which the host zeroed first (or which is being accumulated into anyway: wgrad -> .grad).
//
// Included inside the anonymous namespace of tc_gemm.cu, after tc_gemm_ts_kernel (it uses that kernel's shared-memory
// layout, thread roles and PTX wrappers unchanged).

struct SkArgs {
  int m_tiles, n_tiles;   // output tiles (BM x BN)
  int total_kb;           // K blocks per tile
  long long units;        // m_tiles * n_tiles * total_kb
};

struct SkSegment {
  int m0, n0, kb0, nkb;
};

// the segment that starts at unit u of a range ending at u_end
__host__ __device__ __forceinline__ SkSegment sk_segment(const SkArgs& k, long long u, long long u_end, int bn) {
  SkSegment s;
  const long long tile = u / k.total_kb;
  s.kb0 = static_cast<int>(u - tile * k.total_kb);
  const long long left = u_end - u;
  const int in_tile = k.total_kb - s.kb0;
  s.nkb = left < in_tile ? static_cast<int>(left) : in_tile;
  // consecutive tiles share their A rows (m) and walk over n: the A tile stays in L2 for its n_tiles readers
  s.m0 = static_cast<int>(tile / k.n_tiles) * BM;
  s.n0 = static_cast<int>(tile % k.n_tiles) * bn;
  return s;
}

template <int BN>
__global__ void __launch_bounds__(TS_THREADS, 1)
tc_gemm_sk_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                  const __grid_constant__ CUtensorMap map_b2, GemmArgs g, SkArgs k) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  SmemTS<BN>& s = *reinterpret_cast<SmemTS<BN>*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long u_begin = k.units * blockIdx.x / gridDim.x;
  const long long u_end = k.units * (blockIdx.x + 1) / gridDim.x;

  if (threadIdx.x == 0) {
    for (int i = 0; i < TS_STAGES; ++i) {
      mbar_init(&s.full[i], 1);
      mbar_init(&s.conv[i], 256);
      mbar_init(&s.empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s.acc_full[i], 1);
      mbar_init(&s.acc_empty[i], 256);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s.tmem_base)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s.tmem_base;

  if (u_end > u_begin) {
    if (warp == 0) {
      // ===== TMA producer: same boxes as tc_gemm_ts_kernel, tile by tile =====
      if (lane == 0) {
        int st = 0, it = 0;
        for (long long u = u_begin; u < u_end;) {
          const SkSegment sg = sk_segment(k, u, u_end, BN);
          const int m0 = sg.m0, n0 = sg.n0;
          for (int i = 0; i < sg.nkb; ++i) {
            mbar_wait(&s.empty[st], (it & 1) ^ 1);
            mbar_expect_tx(&s.full[st], (BM + (g.b_split ? 2 * BN : BN)) * BK * 4);
            const int kb = sg.kb0 + i;
            const int k0 = kb * BK;
            if (g.mode == 1 || g.mode == 2) {
              const int hw = g.cH * g.cW;
              const int b0 = m0 / hw, h0 = (m0 - b0 * hw) / g.cW;
              const int per_tap = g.cA / BK;
              const int tap = kb / per_tap, cb = kb - tap * per_tap;
              int dh = tap / 5 - 2, dw = tap % 5 - 2;
              if (g.mode == 2) { dh = -dh; dw = -dw; }
              tma_load_4d(&s.a[st][0], &map_a, &s.full[st], cb * BK, dw, h0 + dh, b0);
              if (g.mode == 1) {
                tma_load_2d(&s.b[st][0][0], &map_b, &s.full[st], k0, n0);
                if (g.b_split) tma_load_2d(&s.b[st][1][0], &map_b2, &s.full[st], k0, n0);
              } else {
                for (int bx = 0; bx < BN / 32; ++bx)
                  tma_load_2d(&s.b[st][0][bx * 32 * BK], &map_b, &s.full[st], tap * g.cIn + n0 + 32 * bx, cb * BK);
                if (g.b_split)
                  for (int bx = 0; bx < BN / 32; ++bx)
                    tma_load_2d(&s.b[st][1][bx * 32 * BK], &map_b2, &s.full[st], tap * g.cIn + n0 + 32 * bx, cb * BK);
              }
            } else if (g.mode == 3) {
              for (int bx = 0; bx < BM / 32; ++bx) tma_load_2d(&s.a[st][bx * 32 * BK], &map_a, &s.full[st], m0 + 32 * bx, k0);
              const int hw = g.cH * g.cW;
              const int b0 = k0 / hw, h0 = (k0 - b0 * hw) / g.cW;
              for (int bx = 0; bx < BN / 32; ++bx) {
                const int col = n0 + 32 * bx;
                const int tap = col / g.cIn, c0 = col - tap * g.cIn;
                if (tap < 25) tma_load_4d(&s.b[st][0][bx * 32 * BK], &map_b, &s.full[st], c0, tap % 5 - 2, h0 + tap / 5 - 2, b0);
                else tma_load_4d(&s.b[st][0][bx * 32 * BK], &map_b, &s.full[st], g.cIn, 0, 0, 0);
              }
            } else {
              if (!g.a_mn) {
                tma_load_2d(&s.a[st][0], &map_a, &s.full[st], k0, m0);
              } else {
                for (int bx = 0; bx < BM / 32; ++bx) tma_load_2d(&s.a[st][bx * 32 * BK], &map_a, &s.full[st], m0 + 32 * bx, k0);
              }
              if (!g.b_mn) {
                tma_load_2d(&s.b[st][0][0], &map_b, &s.full[st], k0, n0);
              } else {
                for (int bx = 0; bx < BN / 32; ++bx) tma_load_2d(&s.b[st][0][bx * 32 * BK], &map_b, &s.full[st], n0 + 32 * bx, k0);
              }
            }
            if (++st == TS_STAGES) { st = 0; ++it; }
          }
          u += sg.nkb;
        }
      }
    } else if (warp == 1) {
      // ===== MMA issuer: accumulator chunks restart at every segment, the two TMEM buffers alternate over ALL chunks =====
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (static_cast<uint32_t>(BN >> 3) << 17) |
                             (static_cast<uint32_t>(BM >> 4) << 24) | (g.b_mn ? (1u << 16) : 0u);
      const uint32_t b_base = smem_u32(&s.b[0][0][0]);
      constexpr uint32_t B_TILE = BN * BK * 4;
      const uint64_t b_step = g.b_mn ? (1024 >> 4) : ((UMMA_K * 4) >> 4);
      const bool three = g.terms == 3;
      int st = 0, it = 0;
      int gc = 0;                                            // chunks started so far by this CTA (all segments)
      for (long long u = u_begin; u < u_end;) {
        const SkSegment sg = sk_segment(k, u, u_end, BN);
        for (int i = 0; i < sg.nkb; ++i) {
          const int buf = gc & 1, use = gc >> 1;
          const bool first_in_chunk = (i % CHUNK) == 0, last_in_chunk = (i % CHUNK) == CHUNK - 1 || i == sg.nkb - 1;
          if (first_in_chunk) mbar_wait(&s.acc_empty[buf], (use & 1) ^ 1);
          mbar_wait(&s.conv[st], it & 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t acc = tmem + static_cast<uint32_t>(buf * BN);
          const uint32_t ta_hi = tmem + static_cast<uint32_t>(2 * BN + st * 64), ta_lo = ta_hi + 32;
          const uint32_t pb_hi = b_base + static_cast<uint32_t>(st) * 2 * B_TILE, pb_lo = pb_hi + B_TILE;
          const uint64_t b_hi = g.b_mn ? make_desc_mn(pb_hi) : make_desc(pb_hi), b_lo = g.b_mn ? make_desc_mn(pb_lo) : make_desc(pb_lo);
          const uint32_t keep0 = first_in_chunk ? 0u : 1u;
          if (elect_one()) {
            if (three) {
#pragma unroll
              for (int kk = 0; kk < BK / UMMA_K; ++kk) {
                umma_tf32_ts(acc, ta_lo + kk * UMMA_K, b_hi + b_step * kk, idesc, kk == 0 ? keep0 : 1u);
                umma_tf32_ts(acc, ta_hi + kk * UMMA_K, b_lo + b_step * kk, idesc, 1u);
                umma_tf32_ts(acc, ta_hi + kk * UMMA_K, b_hi + b_step * kk, idesc, 1u);
              }
            } else {
#pragma unroll
              for (int kk = 0; kk < BK / UMMA_K; ++kk) umma_tf32_ts(acc, ta_hi + kk * UMMA_K, b_hi + b_step * kk, idesc, kk == 0 ? keep0 : 1u);
            }
            umma_commit(&s.empty[st]);
            if (last_in_chunk) umma_commit(&s.acc_full[buf]);
          }
          __syncwarp();
          if (last_in_chunk) ++gc;
          if (++st == TS_STAGES) { st = 0; ++it; }
        }
        u += sg.nkb;
      }
    } else if (warp < 10) {
      // ===== converters: tile-agnostic, they just follow the ring for every K block of the range =====
      const int t = threadIdx.x - 64;
      const uint32_t a_base = smem_u32(&s.a[0][0]), b_base = smem_u32(&s.b[0][0][0]);
      constexpr uint32_t A_TILE = BM * BK * 4, B_TILE = BN * BK * 4;
      const bool three = g.terms == 3;
      const bool conv_b = g.b_split == 0;                    // B arrives split (hi and lo by TMA): nothing to do for it here
      const int quarter = warp & 3;
      const int half = (warp - 2) >> 2;
      const int row = quarter * 32 + lane;
      const long long nkb_all = u_end - u_begin;
      int st = 0, it = 0;
      for (long long i = 0; i < nkb_all; ++i) {
        mbar_wait(&s.full[st], it & 1);
        const uint32_t a_tile = a_base + static_cast<uint32_t>(st) * A_TILE;
        const uint32_t b_hi = b_base + static_cast<uint32_t>(st) * 2 * B_TILE + t * 16, b_lo = b_hi + B_TILE;
        constexpr int NB = BN * BK / 4 / 256;
        float4 bv[NB];
        uint32_t hi[16], lo[16];
        if (!g.a_mn) {
          float4 av[4];
#pragma unroll
          for (int c = 0; c < 4; ++c) av[c] = lds128(a_tile + row * 128 + (((half * 4 + c) ^ (row & 7)) << 4));
          if (conv_b) {
#pragma unroll
            for (int j = 0; j < NB; ++j) bv[j] = lds128(b_hi + j * 4096);
          }
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const float e[4] = {av[c].x, av[c].y, av[c].z, av[c].w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float h = tf32_round(e[q]);
              hi[c * 4 + q] = __float_as_uint(h);
              lo[c * 4 + q] = __float_as_uint(e[q] - h);
            }
          }
        } else {
          const uint32_t colp = a_tile + (row >> 5) * 4096 + (row & 31) * 4 + half * 16 * 128;
          float e[16];
#pragma unroll
          for (int kk = 0; kk < 16; ++kk) e[kk] = lds32(colp + kk * 128);
          if (conv_b) {
#pragma unroll
            for (int j = 0; j < NB; ++j) bv[j] = lds128(b_hi + j * 4096);
          }
#pragma unroll
          for (int kk = 0; kk < 16; ++kk) {
            const float h = tf32_round(e[kk]);
            hi[kk] = __float_as_uint(h);
            lo[kk] = __float_as_uint(e[kk] - h);
          }
        }
        const uint32_t ta = tmem + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(2 * BN + st * 64 + half * 16);
        tmem_st16(ta, hi);
        if (three) tmem_st16(ta + 32, lo);
        if (conv_b) {
#pragma unroll
          for (int j = 0; j < NB; ++j) {
            const float4 v = bv[j];
            const float4 h = make_float4(tf32_round(v.x), tf32_round(v.y), tf32_round(v.z), tf32_round(v.w));
            sts128(b_hi + j * 4096, h);
            if (three) sts128(b_lo + j * 4096, make_float4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w));
          }
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        if (conv_b) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // (no generic-proxy writes to shared memory otherwise)
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        mbar_arrive(&s.conv[st]);
        if (++st == TS_STAGES) { st = 0; ++it; }
      }
    } else {
      // ===== drain + epilogue: one register accumulator per segment, written (or added) when the segment ends =====
      const int quarter = warp & 3;
      const int chalf = (warp - 10) >> 2;
      constexpr int HN = BN / 2;
      int gc = 0;                                            // mirrors the MMA warp's chunk counter
      for (long long u = u_begin; u < u_end;) {
        const SkSegment sg = sk_segment(k, u, u_end, BN);
        float acc[HN];
#pragma unroll
        for (int j = 0; j < HN; ++j) acc[j] = 0.f;
        const int nchunks = (sg.nkb + CHUNK - 1) / CHUNK;
        for (int c = 0; c < nchunks; ++c, ++gc) {
          const int buf = gc & 1, use = gc >> 1;
          mbar_wait(&s.acc_full[buf], use & 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
          for (int c0 = 0; c0 < HN; c0 += 16) {
            uint32_t r[16];
            const uint32_t taddr = tmem + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(buf * BN + chalf * HN + c0);
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                  "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                : "r"(taddr)
                : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int q = 0; q < 16; ++q) acc[c0 + q] += __uint_as_float(r[q]);
          }
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          mbar_arrive(&s.acc_empty[buf]);
        }
        // the MMA warp is already filling the buffers with the next segment while this one is written out
        const int row = sg.m0 + quarter * 32 + lane;
        if (row < g.M) {
          float* crow = g.C + static_cast<long long>(row) * g.ldc + sg.n0 + chalf * HN;
          const bool add_bias = g.bias != nullptr && sg.kb0 == 0;
          const bool atomic = g.atomic_out != 0 || sg.nkb != k.total_kb;     // a partial tile is added to its siblings
#pragma unroll
          for (int q = 0; q < HN; q += 4) {
            const int n = sg.n0 + chalf * HN + q;
            if (n + 3 < g.N && g.vec_out) {
              float4 v = make_float4(acc[q], acc[q + 1], acc[q + 2], acc[q + 3]);
              if (add_bias) {
                const float4 bv = *reinterpret_cast<const float4*>(g.bias + n);
                v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
              }
              if (atomic) {
                asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(crow + q), "f"(v.x), "f"(v.y),
                             "f"(v.z), "f"(v.w)
                             : "memory");
              } else {
                *reinterpret_cast<float4*>(crow + q) = v;
              }
            } else {
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                if (n + e < g.N) {
                  float v = acc[q + e];
                  if (add_bias) v += g.bias[n + e];
                  if (atomic) atomicAdd(crow + q + e, v);
                  else crow[q + e] = v;
                }
              }
            }
          }
        }
        u += sg.nkb;
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512) : "memory");
  }
}

// On by default; DL_B200_TC_STREAMK=0 deselects this kernel (it needs the TS layout, i.e. DL_B200_TC_TS unset or 1)
int g_streamk = -1;   // -1: not decided yet (environment), 0 / 1: off / on (dl_tc_gemm_streamk)
bool use_streamk() {
  if (g_streamk < 0) {
    const char* e = getenv("DL_B200_TC_STREAMK");
    g_streamk = (e != nullptr && e[0] == '0') ? 0 : 1;   // on unless DL_B200_TC_STREAMK=0
  }
  return g_streamk != 0;
}

// Work decomposition for an M x N x K problem on `sms` SMs: one CTA per SM (512 TMEM columns and ~200 KB of shared
// memory each), never fewer than 4 K blocks per CTA.
SkArgs sk_plan(int M, int N, int K, int bn, int sms, long long* grid) {
  SkArgs k;
  k.m_tiles = (M + BM - 1) / BM;
  k.n_tiles = (N + bn - 1) / bn;
  k.total_kb = (K + BK - 1) / BK;
  k.units = static_cast<long long>(k.m_tiles) * k.n_tiles * k.total_kb;
  long long gsz = sms;
  if (gsz > k.units / 4) gsz = k.units / 4 > 0 ? k.units / 4 : 1;
  *grid = gsz;
  return k;
}

// `accumulate`: C already holds values to add to (every segment uses red.add, nothing is zeroed)
template <int BN>
int launch_sk(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mb2, GemmArgs g, int accumulate, cudaStream_t st) {
  static bool configured[64] = {false};
  static int sm_count[64] = {0};
  int dev = 0;
  DL_CUDA(cudaGetDevice(&dev));
  DL_REQUIRE(dev >= 0 && dev < 64, "tc_gemm: device index %d out of range", dev);
  const size_t smem = sizeof(SmemTS<BN>) + 1024;
  if (!configured[dev]) {
    DL_CUDA(cudaFuncSetAttribute(tc_gemm_sk_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    DL_CUDA(cudaDeviceGetAttribute(&sm_count[dev], cudaDevAttrMultiProcessorCount, dev));
    configured[dev] = true;
  }
  long long grid = 0;
  const SkArgs k = sk_plan(g.M, g.N, g.K, BN, sm_count[dev] > 0 ? sm_count[dev] : 148, &grid);
  const long long tiles = static_cast<long long>(k.m_tiles) * k.n_tiles;
  g.atomic_out = accumulate ? 1 : 0;
  g.kblocks_per_split = k.total_kb;
  if (!accumulate && tiles % grid != 0) {
    // some tile is shared between CTAs and arrives as a sum of red.adds: start from zero
    if (g.ldc == g.N) {
      DL_CUDA(cudaMemsetAsync(g.C, 0, static_cast<size_t>(g.M) * g.N * 4, st));
    } else {
      DL_CUDA(cudaMemset2DAsync(g.C, static_cast<size_t>(g.ldc) * 4, 0, static_cast<size_t>(g.N) * 4, g.M, st));
    }
  }
  tc_gemm_sk_kernel<BN><<<static_cast<unsigned>(grid), TS_THREADS, smem, st>>>(ma, mb, mb2, g, k);
  dl_note_launch();
  DL_CUDA(cudaGetLastError());
  return DL_OK;
}
