// tcgen05 / TMEM / TMA implicit-GEMM kernels for the ResBlock convolution (sm_100a only).
//
//   k_tc_conv<C, MODE>   y[p][co] = sum_{tap,ci} x[p+tap][ci] * W[co][tap][ci]      (fwd, MODE 0: + BN partial sums)
//                        dx[p][ci] = sum_{tap,co} dy[p-tap][co] * W[co][tap][ci] + g (dgrad, MODE 1: same kernel on the
//                        tap-flipped / transposed bf16 weight copy, residual add in the epilogue)
//   k_tc_wgrad<C>        dW[tap][ci][co] (+)= sum_p x[p+tap][ci] * dy[p][co]        (per-CTA partial, fixed-order reduce later)
//
// Data movement: one CTA tile = 8 rows x 16 cols of one image = 128 pixels = the UMMA M (fwd) or K (wgrad) extent.
// TMA loads the NHWC bf16 tile of each of the 9 taps as a 4-D box {C,16,8,1} at coordinates shifted by the tap;
// out-of-bounds rows/cols are zero-filled by the TMA unit, which IS the conv padding.  The inner extent is C*2 bytes
// (64 B for C=32 -> SWIZZLE_64B, 128 B for C=64 -> SWIZZLE_128B), the layout the UMMA smem descriptors name.
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM alloc + MMA issuer (one lane), warps 2..5 = epilogue
// (TMEM -> registers -> bf16 -> global), double-buffered TMEM accumulators so the epilogue of tile i overlaps the MMAs
// of tile i+1.  Every mbarrier wait has a clock64() timeout that traps instead of hanging the GPU.
#include "nrd_common.cuh"

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  long long t0 = 0;
  int spins = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if (++spins > 4096) {
      if (t0 == 0) t0 = clock64();
      else if (clock64() - t0 > 4000000000LL) __trap();   // ~2 s: a protocol bug must not hang the box
    }
  }
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tm) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]; bf16 operands, fp32 accumulate
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (thread = its own TMEM lane)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// ---- descriptors --------------------------------------------------------------------------
// smem matrix descriptor (cute::UMMA::SmemDescriptor): start>>4 [0,14) | LBO>>4 [16,30) | SBO>>4 [32,46) | version=1 [46,48)
// | layout [61,64): 2 = SWIZZLE_128B, 4 = SWIZZLE_64B
__device__ __forceinline__ uint64_t make_sdesc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout) {
  return (uint64_t)((addr & 0x3FFFF) >> 4) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46) | ((uint64_t)layout << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor), kind::f16: c_format F32 [4,6)=1, a/b format BF16 [7,10)/[10,13)=1,
// a_major [15], b_major [16] (0 = K-major, 1 = MN-major), N>>3 [17,23), M>>4 [24,29)
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, int a_mn, int b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}

template <int C> struct Cfg {
  static constexpr int ROW_BYTES = C * 2;                       // one pixel's channels
  static constexpr int TILE_BYTES = 128 * ROW_BYTES;            // 128-pixel activation tile
  static constexpr int WTAP_BYTES = C * ROW_BYTES;              // one tap of the weights: [C rows][C k]
  static constexpr uint32_t LAYOUT = (C == 32) ? 4u : 2u;       // SWIZZLE_64B : SWIZZLE_128B
  static constexpr int SBO = 8 * ROW_BYTES;                     // 8-row core-matrix group stride
  static constexpr int NST = (C == 32) ? 8 : 6;                 // activation stages (conv kernel)
  static constexpr int KSTEPS = C / 16;                         // UMMA K=16 steps per tap
};

// recursive-halving warp transpose-reduce: on return lane l holds sum over the 32 lanes of v[l]
__device__ __forceinline__ float warp_col_sum(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool upper = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      float send = upper ? v[i] : v[i + off];
      float keep = upper ? v[i + off] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

// ============================================================================================
// conv fwd / dgrad
// ============================================================================================
template <int C, int MODE>
__global__ void __launch_bounds__(192, 1)
k_tc_conv(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, bf16* __restrict__ out,
          const bf16* __restrict__ residual, float* __restrict__ stats_partial, int ntiles) {
  using K = Cfg<C>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t a_smem = base;                                   // NST stages
  const uint32_t w_smem = a_smem + K::NST * K::TILE_BYTES;       // 9 taps
  const uint32_t bar0 = w_smem + 9 * K::WTAP_BYTES;              // barriers (8 B each)
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 8u * (K::NST + s); };
  const uint32_t wfull = bar0 + 8u * (2 * K::NST);
  auto tfull = [&](int s) { return bar0 + 8u * (2 * K::NST + 1 + s); };
  auto tempty = [&](int s) { return bar0 + 8u * (2 * K::NST + 3 + s); };
  const uint32_t tmem_slot = bar0 + 8u * (2 * K::NST + 5);
  const uint32_t red_off = 9 * K::WTAP_BYTES + K::NST * K::TILE_BYTES + 8 * (2 * K::NST + 6);
  float* red = reinterpret_cast<float*>(gen_base + red_off);      // [4][2][C]
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t TMEM_COLS = (2 * C <= 64) ? 64 : 128;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmW);
    for (int s = 0; s < K::NST; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 1); }
    mbar_init(wfull, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(tfull(s), 1); mbar_init(tempty(s), 4); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer =====
      mbar_expect_tx(wfull, 9 * K::WTAP_BYTES);
      for (int t = 0; t < 9; ++t) tma_load_2d(w_smem + t * K::WTAP_BYTES, &tmW, wfull, t * C, 0);
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int b = tile >> 1, r0 = (tile & 1) * 8;
        for (int t = 0; t < 9; ++t) {
          mbar_wait(empty(stage), phase ^ 1);
          mbar_expect_tx(full(stage), K::TILE_BYTES);
          tma_load_4d(a_smem + stage * K::TILE_BYTES, &tmA, full(stage), 0, t % 3 - 1, r0 + t / 3 - 1, b);
          if (++stage == K::NST) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer =====
      constexpr uint32_t idesc = make_idesc(128, C, 0, 0);
      mbar_wait(wfull, 0);
      int stage = 0; uint32_t phase = 0; int it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(tempty(as), aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * C;
        for (int t = 0; t < 9; ++t) {
          mbar_wait(full(stage), phase);
          tc_fence_after();
          const uint64_t ad = make_sdesc(a_smem + stage * K::TILE_BYTES, 16, K::SBO, K::LAYOUT);
          const uint64_t bd = make_sdesc(w_smem + t * K::WTAP_BYTES, 16, K::SBO, K::LAYOUT);
#pragma unroll
          for (int k = 0; k < K::KSTEPS; ++k)   // +32 B along K inside the swizzled row = +2 in the (addr>>4) field
            umma_bf16(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (t | k) != 0);
          umma_commit(empty(stage));
          if (t == 8) umma_commit(tfull(as));
          if (++stage == K::NST) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===== epilogue warps 2..5: TMEM lane quarter q = warp % 4 =====
    const int q = warp & 3;
    float ssum[C / 32], qsum[C / 32];
#pragma unroll
    for (int h = 0; h < C / 32; ++h) ssum[h] = qsum[h] = 0.f;
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      const int b = tile >> 1, r0 = (tile & 1) * 8;
      mbar_wait(tfull(as), aphase);
      tc_fence_after();
      const size_t gp = (size_t)b * NRD_PIX + r0 * 16 + q * 32 + lane;   // this thread's pixel
      bf16* orow = out + gp * C;
#pragma unroll
      for (int h = 0; h < C / 32; ++h) {
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + as * C + h * 32, v);
        if (h == C / 32 - 1) {               // accumulator is in registers: hand the TMEM stage back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty(as));
        }
        if (MODE == 1) {
          const uint4* rrow = reinterpret_cast<const uint4*>(residual + gp * C + h * 32);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 rv = rrow[j];
            const __nv_bfloat162* rh = reinterpret_cast<const __nv_bfloat162*>(&rv);
#pragma unroll
            for (int e = 0; e < 4; ++e) { float2 f = __bfloat1622float2(rh[e]); v[j * 8 + 2 * e] += f.x; v[j * 8 + 2 * e + 1] += f.y; }
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 pk; __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
          for (int e = 0; e < 4; ++e) ph[e] = __floats2bfloat162_rn(v[j * 8 + 2 * e], v[j * 8 + 2 * e + 1]);
          reinterpret_cast<uint4*>(orow + h * 32)[j] = pk;
        }
        if (MODE == 0) {
          float sq[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) sq[i] = v[i] * v[i];
          ssum[h] += warp_col_sum(v, lane);
          qsum[h] += warp_col_sum(sq, lane);
        }
      }
    }
    if (MODE == 0) {
#pragma unroll
      for (int h = 0; h < C / 32; ++h) { red[(q * 2 + 0) * C + h * 32 + lane] = ssum[h]; red[(q * 2 + 1) * C + h * 32 + lane] = qsum[h]; }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      const int t = threadIdx.x - 64;
      if (t < 2 * C) {
        // fixed order over the 4 lane quarters; red index: [q][2][C]
        float s = red[t] + red[2 * C + t] + red[4 * C + t] + red[6 * C + t];
        stats_partial[(size_t)blockIdx.x * 2 * C + t] = s;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
}

// ============================================================================================
// wgrad: both operands MN-major (channels contiguous, pixels = K).  A = up to 128/C taps of x stacked along M
// (the tap tiles are consecutive in shared memory: LBO = TILE_BYTES jumps from one tap to the next), B = dy tile.
// ============================================================================================
template <int C> struct WCfg {
  static constexpr int TAPS_PER_MMA = 128 / C;                         // 4 (C=32) or 2 (C=64)
  static constexpr int GROUPS = (9 + TAPS_PER_MMA - 1) / TAPS_PER_MMA; // 3 or 5 accumulators of C columns
  static constexpr int SLOTS = GROUPS * TAPS_PER_MMA;                  // tile slots the A descriptors may touch (12 / 10)
  static constexpr int DY_SLOT = 9;                                    // dy sits right after the 9 taps (and doubles as junk rows)
  static constexpr int SS_TILES = SLOTS > 10 ? SLOTS : 10;             // slots 10.. (C=32 only) are zero-filled junk
  static constexpr int NSS = (C == 32) ? 2 : 1;                        // super-stages that fit in 227 KB
  static constexpr uint32_t TMEM_COLS = (GROUPS * C <= 128) ? 128 : 512;
};

template <int C>
__global__ void __launch_bounds__(192, 1)
k_tc_wgrad(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmDY, float* __restrict__ partial,
           int ntiles, int first) {
  using K = Cfg<C>; using W = WCfg<C>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
  constexpr int SS_BYTES = W::SS_TILES * K::TILE_BYTES;
  const uint32_t bar0 = base + W::NSS * SS_BYTES;
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 8u * (W::NSS + s); };
  const uint32_t tfull = bar0 + 8u * (2 * W::NSS);
  const uint32_t tmem_slot = bar0 + 8u * (2 * W::NSS + 1);
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_slot - base));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // junk tile slots (10..SS_TILES-1) are read by the last MMA group (their accumulator rows are discarded): zero them once
  if constexpr (W::SS_TILES > 10) {
    constexpr int per = (W::SS_TILES - 10) * K::TILE_BYTES / 16;
    for (int i = threadIdx.x; i < W::NSS * per; i += blockDim.x) {
      const int ss = i / per, o = i % per;
      reinterpret_cast<uint4*>(gen_base + ss * SS_BYTES + 10 * K::TILE_BYTES)[o] = make_uint4(0, 0, 0, 0);
    }
    fence_proxy_async();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX); tma_prefetch_desc(&tmDY);
    for (int s = 0; s < W::NSS; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 1); }
    mbar_init(tfull, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, W::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;

  if (warp == 0) {
    if (lane == 0) {
      int ss = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int b = tile >> 1, r0 = (tile & 1) * 8;
        mbar_wait(empty(ss), phase ^ 1);
        mbar_expect_tx(full(ss), 10 * K::TILE_BYTES);
        const uint32_t sbase = base + ss * SS_BYTES;
        for (int t = 0; t < 9; ++t) tma_load_4d(sbase + t * K::TILE_BYTES, &tmX, full(ss), 0, t % 3 - 1, r0 + t / 3 - 1, b);
        tma_load_4d(sbase + W::DY_SLOT * K::TILE_BYTES, &tmDY, full(ss), 0, 0, r0, b);
        if (++ss == W::NSS) { ss = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(128, C, 1, 1);
      int ss = 0; uint32_t phase = 0; bool any = false;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        mbar_wait(full(ss), phase);
        tc_fence_after();
        const uint32_t sbase = base + ss * SS_BYTES;
#pragma unroll
        for (int g = 0; g < W::GROUPS; ++g) {
#pragma unroll
          for (int k = 0; k < 8; ++k) {          // 16 pixels (two 8-row swizzle atoms) per MMA
            const uint64_t ad = make_sdesc(sbase + g * W::TAPS_PER_MMA * K::TILE_BYTES + k * 16 * K::ROW_BYTES, K::TILE_BYTES, K::SBO, K::LAYOUT);
            const uint64_t bd = make_sdesc(sbase + W::DY_SLOT * K::TILE_BYTES + k * 16 * K::ROW_BYTES, K::TILE_BYTES, K::SBO, K::LAYOUT);
            umma_bf16(tmem_base + g * C, ad, bd, idesc, (any || k > 0) ? 1u : 0u);
          }
        }
        any = true;
        umma_commit(empty(ss));
        if (++ss == W::NSS) { ss = 0; phase ^= 1; }
      }
      umma_commit(tfull);
    }
  } else {
    const int q = warp & 3;
    mbar_wait(tfull, 0);
    tc_fence_after();
    float* outp = partial + (size_t)blockIdx.x * 9 * C * C;
    const int m = q * 32 + lane;                       // accumulator row = (tap within group)*C + ci
#pragma unroll
    for (int g = 0; g < W::GROUPS; ++g) {
      const int tap = g * W::TAPS_PER_MMA + m / C, ci = m % C;
#pragma unroll
      for (int h = 0; h < C / 32; ++h) {
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + g * C + h * 32, v);
        if (tap < 9) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const size_t o = ((size_t)tap * C + h * 32 + j) * C + ci;      // [tap][co][ci]
            outp[o] = first ? v[j] : outp[o] + v[j];
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, W::TMEM_COLS); }
}

// ---- host ---------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  }
  return fn;
}

static int make_act_map(CUtensorMap* tm, void* ptr, int B, int C) {
  PFN_encodeTiled enc = get_encode();
  NRD_REQUIRE(enc, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[4] = {(cuuint64_t)C, 16, 16, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)16 * C * 2, (cuuint64_t)256 * C * 2};
  cuuint32_t box[4] = {(cuuint32_t)C, 16, 8, 1};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, ptr, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   C == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { nrd_set_error("cuTensorMapEncodeTiled(activation) failed: %d", (int)r); return NRD_ECUDA; }
  return NRD_OK;
}
static int make_w_map(CUtensorMap* tm, void* ptr, int C) {
  PFN_encodeTiled enc = get_encode();
  NRD_REQUIRE(enc, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[2] = {(cuuint64_t)9 * C, (cuuint64_t)C};
  cuuint64_t strides[1] = {(cuuint64_t)9 * C * 2};
  cuuint32_t box[2] = {(cuuint32_t)C, (cuuint32_t)C};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   C == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { nrd_set_error("cuTensorMapEncodeTiled(weights) failed: %d", (int)r); return NRD_ECUDA; }
  return NRD_OK;
}

template <int C> constexpr size_t conv_smem() {
  return 1024 + Cfg<C>::NST * Cfg<C>::TILE_BYTES + 9 * Cfg<C>::WTAP_BYTES + 8 * (2 * Cfg<C>::NST + 6) + 4 * 2 * C * sizeof(float) + 64;
}
template <int C> constexpr size_t wgrad_smem() {
  return 1024 + WCfg<C>::NSS * WCfg<C>::SS_TILES * Cfg<C>::TILE_BYTES + 8 * (2 * WCfg<C>::NSS + 2) + 64;
}

static int conv_grid(const nrd_plan* p) { int t = p->B * 2; return t < NRD_NUM_SMS ? t : NRD_NUM_SMS; }

template <typename F> static int set_attr(F* fn, size_t smem) {
  cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) { nrd_set_error("cudaFuncSetAttribute(%zu B smem): %s", smem, cudaGetErrorString(e)); return NRD_ECUDA; }
  return NRD_OK;
}

}  // namespace

void nrd_count_launch();

int nrd_tc_plan_init(nrd_plan* p) {
  p->tc_ready = 0;
  if (p->dtype != NRD_BF16) return NRD_OK;
  int major = 0;
  NRD_CHECK_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, p->device));
  NRD_REQUIRE(major == 10, "the bf16 path needs an sm_100a (B200) device: tcgen05/TMEM kernels, no fallback (found sm_%d*)", major * 10);
  const size_t act = p->act_elems * p->esz;
  for (int k = 0; k <= p->NB; ++k) NRD_TRY(make_act_map(&p->tm_x[k], p->xs + act * k, p->B, p->C));
  NRD_TRY(make_act_map(&p->tm_dy, p->dy, p->B, p->C));
  NRD_TRY(make_w_map(&p->tm_wf, p->wtc_fwd, p->C));
  NRD_TRY(make_w_map(&p->tm_wd, p->wtc_dgrad, p->C));
  if (p->C == 32) {
    NRD_TRY(set_attr(k_tc_conv<32, 0>, conv_smem<32>())); NRD_TRY(set_attr(k_tc_conv<32, 1>, conv_smem<32>()));
    NRD_TRY(set_attr(k_tc_wgrad<32>, wgrad_smem<32>()));
  } else {
    NRD_TRY(set_attr(k_tc_conv<64, 0>, conv_smem<64>())); NRD_TRY(set_attr(k_tc_conv<64, 1>, conv_smem<64>()));
    NRD_TRY(set_attr(k_tc_wgrad<64>, wgrad_smem<64>()));
  }
  p->tc_ready = 1;
  return NRD_OK;
}

int nrd_tc_stats_blocks(const nrd_plan* p) { return conv_grid(p); }

int nrd_tc_conv_fwd(nrd_plan* p, int k, cudaStream_t st) {
  NRD_REQUIRE(p->tc_ready, "tcgen05 path not initialised");
  const size_t act = p->act_elems * p->esz;
  bf16* y = (bf16*)(p->ys + act * k);
  const int grid = conv_grid(p), nt = p->B * 2;
  if (p->C == 32) k_tc_conv<32, 0><<<grid, 192, conv_smem<32>(), st>>>(p->tm_x[k], p->tm_wf, y, nullptr, p->stats_partial, nt);
  else k_tc_conv<64, 0><<<grid, 192, conv_smem<64>(), st>>>(p->tm_x[k], p->tm_wf, y, nullptr, p->stats_partial, nt);
  nrd_count_launch();
  NRD_CHECK_CUDA(cudaGetLastError());
  return NRD_OK;
}

int nrd_tc_conv_dgrad(nrd_plan* p, const void* g, void* dx_out, cudaStream_t st) {
  NRD_REQUIRE(p->tc_ready, "tcgen05 path not initialised");
  const int grid = conv_grid(p), nt = p->B * 2;
  if (p->C == 32) k_tc_conv<32, 1><<<grid, 192, conv_smem<32>(), st>>>(p->tm_dy, p->tm_wd, (bf16*)dx_out, (const bf16*)g, nullptr, nt);
  else k_tc_conv<64, 1><<<grid, 192, conv_smem<64>(), st>>>(p->tm_dy, p->tm_wd, (bf16*)dx_out, (const bf16*)g, nullptr, nt);
  nrd_count_launch();
  NRD_CHECK_CUDA(cudaGetLastError());
  return NRD_OK;
}

int nrd_tc_conv_wgrad(nrd_plan* p, int k, int first, cudaStream_t st) {
  NRD_REQUIRE(p->tc_ready, "tcgen05 path not initialised");
  const int grid = conv_grid(p), nt = p->B * 2;   // same CTA count as the CUDA-core wgrad => same partial buffer / reduce
  if (p->C == 32) k_tc_wgrad<32><<<grid, 192, wgrad_smem<32>(), st>>>(p->tm_x[k], p->tm_dy, p->wgrad_partial, nt, first);
  else k_tc_wgrad<64><<<grid, 192, wgrad_smem<64>(), st>>>(p->tm_x[k], p->tm_dy, p->wgrad_partial, nt, first);
  nrd_count_launch();
  NRD_CHECK_CUDA(cudaGetLastError());
  return NRD_OK;
}
