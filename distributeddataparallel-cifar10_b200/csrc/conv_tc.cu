// tcgen05 / TMEM / TMA implicit-GEMM kernels for the ResBlock convolution (sm_100a only).
//
//   k_tc_conv<C, MODE>   y[p][co] = sum_{tap,ci} x[p+tap][ci] * W[co][tap][ci]      (fwd, MODE 0: + BN partial sums)
//                        dx[p][ci] = sum_{tap,co} dy[p-tap][co] * W[co][tap][ci] + g (dgrad, MODE 1: same kernel on the
//                        tap-flipped / transposed bf16 weight copy, residual add in the epilogue)
//   k_tc_wgrad<C>        dW[tap][ci][co] (+)= sum_p x[p+tap][ci] * dy[p][co]        (per-CTA partial, fixed-order reduce later)
//
// Data movement (v2): TMA issue rate, not bytes, bounded v1 (one box per tap = 9 x 128 rows of 64 B per 128 pixels,
// ~4 cycles/row; profiles/r01_launches_bench_nograph_v1.md).  Now a work item is ROWS image rows (a whole 16x16 image
// for C=32, half for C=64) and TMA loads it THREE times, once per horizontal tap kx, as a 4-D box {C,16,ROWS+2,1} at
// w-coordinate kx-1 and h-coordinate r0-1: out-of-bounds rows/cols are zero-filled by the TMA unit, which IS the conv
// padding.  The vertical taps ky are then just +ky*16 pixel rows = +ky*1024 B (C=32) on the UMMA descriptor start
// address, which keeps every operand start aligned to the swizzle atom.  Inner extent = C*2 bytes (64 B -> SWIZZLE_64B,
// 128 B -> SWIZZLE_128B), the layout the UMMA smem descriptors name.
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM alloc + MMA issuer (one lane), warps 2..5 = epilogue
// (TMEM -> registers -> bf16 -> global), double-buffered TMEM accumulators so the epilogue of tile i overlaps the MMAs
// of tile i+1.  Every mbarrier wait has a clock64() timeout that traps instead of hanging the GPU.
#include <type_traits>
#include "tc_common.cuh"

namespace {
template <int C> struct Cfg {
  static constexpr int ROW_BYTES = C * 2;                       // one pixel's channels
  static constexpr int TILE_BYTES = 128 * ROW_BYTES;            // 128-pixel activation tile
  static constexpr int WTAP_BYTES = C * ROW_BYTES;              // one tap of the weights: [C rows][C k]
  static constexpr uint32_t LAYOUT = (C == 32) ? 4u : 2u;       // SWIZZLE_64B : SWIZZLE_128B
  static constexpr int SBO = 8 * ROW_BYTES;                     // 8-row core-matrix group stride
  static constexpr int KSTEPS = C / 16;                         // UMMA K=16 steps per tap
  static constexpr int ROWS = (C == 32) ? 16 : 8;               // image rows per work item
  static constexpr int MT = ROWS / 8;                           // 128-pixel M tiles per work item
  static constexpr int ITEMS_PER_IMG = 16 / ROWS;
  static constexpr int IMGROW_BYTES = 16 * ROW_BYTES;           // one image row (16 pixels)
  static constexpr int BUF_BYTES = (ROWS + 2) * IMGROW_BYTES;   // one kx-shifted, row-haloed copy of the item
  static constexpr int NST = 2;                                 // work-item stages
};

// 16 values per lane -> after the call lanes 2c and 2c+1 both hold the sum over the 32 lanes of v[c] (c = lane >> 1)
__device__ __forceinline__ float warp_col_sum16(float (&v)[16], int lane) {
#pragma unroll
  for (int off = 16; off >= 2; off >>= 1) {
    const bool upper = (lane & off) != 0;
    const int n = off >> 1;
#pragma unroll
    for (int i = 0; i < n; ++i) {
      float send = upper ? v[i] : v[i + n];
      float keep = upper ? v[i + n] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 1);
}

// 1/sqrt(x) to fp32 accuracy without the fp64 sqrt/div software sequences (hundreds of cycles on the per-block critical path):
// hardware rsqrt + one Newton step
__device__ __forceinline__ float fast_invstd(double var_plus_eps) {
  const float x = (float)var_plus_eps;
  float y = rsqrtf(x);
  y = y * (1.5f - 0.5f * x * y * y);
  return y;
}

// ============================================================================================
// conv fwd / dgrad / fused fwd block
//   MODE 0: y = conv(x), per-CTA BN partial sums                       (two-kernel forward, any batch)
//   MODE 1: dx = conv(dy, flipped W) + g                               (dgrad)
//   MODE 2: whole ResBlock forward: accumulators of ALL the CTA's tiles stay in TMEM; pass 1 takes the BN partial sums,
//           a grid barrier makes every CTA's partials visible, each CTA combines them in the same fixed order,
//           pass 2 re-reads TMEM and writes y (bf16) and x_next = relu(bn(y)) + x.   (cooperative launch)
// ============================================================================================
struct BlockFwdArgs {
  const float* gamma; const float* beta;     // [C]
  float* saved;                              // [2C] mean | invstd of this application (for backward)
  float* scale_shift;                        // [2C]
  float* running;                            // [2C] running_mean | running_var (EMA, CTA 0)
  long long* nbt;                            // num_batches_tracked
  const bf16* x;                             // block input (residual)
  bf16* xnext;                               // block output
  unsigned long long* gbar;                  // grid barrier counter
  int count;                                 // B*256
};

template <int C, int MODE>
__global__ void __launch_bounds__(192, 1)
k_tc_conv(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, bf16* __restrict__ out,
          const bf16* __restrict__ residual, float* __restrict__ stats_partial, int nitems, BlockFwdArgs fa) {
  using K = Cfg<C>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
  constexpr int STAGE_BYTES = 3 * K::BUF_BYTES;
  constexpr int NACC = (MODE == 2) ? (512 / C) : 2;               // accumulator slots of C columns
  const uint32_t a_smem = base;                                   // NST stages x 3 kx-buffers
  const uint32_t w_smem = a_smem + K::NST * STAGE_BYTES;         // 9 taps
  const uint32_t bar0 = w_smem + 9 * K::WTAP_BYTES;              // barriers (8 B each)
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 8u * (K::NST + s); };
  const uint32_t wfull = bar0 + 8u * (2 * K::NST);
  auto tfull = [&](int s) { return bar0 + 8u * (2 * K::NST + 1 + s); };
  auto tempty = [&](int s) { return bar0 + 8u * (2 * K::NST + 1 + NACC + s); };
  const uint32_t tmem_slot = bar0 + 8u * (2 * K::NST + 1 + 2 * NACC);
  constexpr int BAR_BYTES = 8 * (2 * K::NST + 2 + 2 * NACC);
  float* red = reinterpret_cast<float*>(gen_base + K::NST * STAGE_BYTES + 9 * K::WTAP_BYTES + BAR_BYTES);   // [4][2][C] (+ [2C] scale/shift)
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t TMEM_COLS = (MODE == 2) ? 512 : ((2 * C <= 64) ? 64 : 128);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmW);
    for (int s = 0; s < K::NST; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 1); }
    mbar_init(wfull, 1);
    for (int s = 0; s < NACC; ++s) { mbar_init(tfull(s), 1); mbar_init(tempty(s), 4); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer: 3 loads (kx = 0,1,2) per work item =====
      mbar_expect_tx(wfull, 9 * K::WTAP_BYTES);
      for (int t = 0; t < 9; ++t) tma_load_2d(w_smem + t * K::WTAP_BYTES, &tmW, wfull, t * C, 0);
      int stage = 0; uint32_t phase = 0;
      for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
        const int b = item / K::ITEMS_PER_IMG, r0 = (item % K::ITEMS_PER_IMG) * K::ROWS;
        mbar_wait(empty(stage), phase ^ 1);
        mbar_expect_tx(full(stage), STAGE_BYTES);
        for (int kx = 0; kx < 3; ++kx)
          tma_load_4d(a_smem + stage * STAGE_BYTES + kx * K::BUF_BYTES, &tmA, full(stage), 0, kx - 1, r0 - 1, b);
        if (++stage == K::NST) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer =====
      constexpr uint32_t idesc = make_idesc(128, C, 0, 0);
      mbar_wait(wfull, 0);
      int stage = 0; uint32_t phase = 0; int tile = 0;
      for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
        mbar_wait(full(stage), phase);
        tc_fence_after();
        for (int m = 0; m < K::MT; ++m, ++tile) {
          const int as = (MODE == 2) ? tile : (tile & 1);
          if (MODE != 2) { mbar_wait(tempty(as), ((tile >> 1) & 1) ^ 1); tc_fence_after(); }
          const uint32_t d_tmem = tmem_base + as * C;
#pragma unroll
          for (int t = 0; t < 9; ++t) {
            const uint32_t a_addr = a_smem + stage * STAGE_BYTES + (t % 3) * K::BUF_BYTES + (m * 8 + t / 3) * K::IMGROW_BYTES;
            const uint64_t ad = make_sdesc(a_addr, 16, K::SBO, K::LAYOUT);
            const uint64_t bd = make_sdesc(w_smem + t * K::WTAP_BYTES, 16, K::SBO, K::LAYOUT);
#pragma unroll
            for (int k = 0; k < K::KSTEPS; ++k)   // +32 B along K inside the swizzled row = +2 in the (addr>>4) field
              umma_bf16(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (t | k) != 0);
          }
          umma_commit(tfull(as));
        }
        umma_commit(empty(stage));
        if (++stage == K::NST) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===== epilogue warps 2..5: TMEM lane quarter q = warp % 4 =====
    const int q = warp & 3;
    unsigned long long gbar_base = 0;
    if (MODE == 2 && threadIdx.x == 64) gbar_base = nrd_grid_barrier_base(fa.gbar, gridDim.x);   // before this CTA's own arrival
    float ssum[C / 32], qsum[C / 32];
#pragma unroll
    for (int h = 0; h < C / 32; ++h) ssum[h] = qsum[h] = 0.f;
    int tile = 0;
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
      const int b = item / K::ITEMS_PER_IMG, r0 = (item % K::ITEMS_PER_IMG) * K::ROWS;
      for (int m = 0; m < K::MT; ++m, ++tile) {
        const int as = (MODE == 2) ? tile : (tile & 1);
        mbar_wait(tfull(as), (MODE == 2) ? 0u : (uint32_t)((tile >> 1) & 1));
        tc_fence_after();
        const size_t gp = (size_t)b * NRD_PIX + (r0 + m * 8) * 16 + q * 32 + lane;   // this thread's pixel
#pragma unroll
        for (int h = 0; h < C / 32; ++h) {
          float v[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + as * C + h * 32, v);
          if (MODE != 2 && h == C / 32 - 1) {      // accumulator is in registers: hand the TMEM slot back
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
          if (MODE != 2) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 pk; __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
              for (int e = 0; e < 4; ++e) ph[e] = __floats2bfloat162_rn(v[j * 8 + 2 * e], v[j * 8 + 2 * e + 1]);
              reinterpret_cast<uint4*>(out + gp * C + h * 32)[j] = pk;
            }
          }
          if (MODE != 1) {
            float sq[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) sq[i] = v[i] * v[i];
            ssum[h] += warp_col_sum(v, lane);
            qsum[h] += warp_col_sum(sq, lane);
          }
        }
      }
    }
    if (MODE != 1) {
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
    if (MODE == 2) {
      // ---- every CTA's partials -> global; grid barrier; identical fixed-order combine in every CTA ----
      const int t = threadIdx.x - 64;                 // 0..127
      float* ss_sm = red + 8 * C;                     // [2C] scale | shift
      double* dred = reinterpret_cast<double*>(red + 10 * C);   // [RP][2C], RP*2C = 512
      asm volatile("bar.sync 1, 128;" ::: "memory");  // partial row written
      if (t == 0) nrd_grid_barrier(fa.gbar, gbar_base + gridDim.x);
      asm volatile("bar.sync 1, 128;" ::: "memory");
      {
        // partial matrix [gridDim.x][2C] fp32: every thread owns one float4 column chunk and every RP-th row, all loads
        // independent (one L2 round trip), fp64 accumulation, then a fixed-order combine over the RP row classes
        constexpr int NC4 = 2 * C / 4, RP = 128 / NC4;
        const int c4 = t % NC4, r = t / NC4;
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
        for (int row0 = r; row0 < (int)gridDim.x; row0 += 8 * RP) {       // batches of 8 independent, predicated loads
          float4 v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int row = row0 + u * RP;
            v[u] = (row < (int)gridDim.x) ? __ldcg(reinterpret_cast<const float4*>(stats_partial + (size_t)row * 2 * C) + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) { a0 += (double)v[u].x; a1 += (double)v[u].y; a2 += (double)v[u].z; a3 += (double)v[u].w; }
        }
        double* d = dred + (size_t)r * 2 * C + c4 * 4;
        d[0] = a0; d[1] = a1; d[2] = a2; d[3] = a3;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (t < C) {
        constexpr int RP = 128 / (2 * C / 4);
        double S = 0.0, Q = 0.0;
#pragma unroll
        for (int k = 0; k < RP; ++k) { S += dred[k * 2 * C + t]; Q += dred[k * 2 * C + C + t]; }
        const double mean = S / fa.count;
        double var = Q / fa.count - mean * mean;
        if (var < 0.0) var = 0.0;
        const float invstd = (float)(1.0 / sqrt(var + (double)NRD_BN_EPS));
        const float sc = fa.gamma[t] * invstd, sh = fa.beta[t] - (float)mean * sc;
        ss_sm[t] = sc; ss_sm[C + t] = sh;
        if (blockIdx.x == 0) {
          fa.saved[t] = (float)mean; fa.saved[C + t] = invstd;
          fa.scale_shift[t] = sc; fa.scale_shift[C + t] = sh;
          const double unbiased = var * ((double)fa.count / (double)(fa.count - 1));
          fa.running[t] = (1.f - NRD_BN_MOMENTUM) * fa.running[t] + NRD_BN_MOMENTUM * (float)mean;
          fa.running[C + t] = (1.f - NRD_BN_MOMENTUM) * fa.running[C + t] + NRD_BN_MOMENTUM * (float)unbiased;
          if (t == 0) *fa.nbt += 1;
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      // ---- pass 2: y (bf16) and x_next = relu(y_bf16*scale + shift) + x ----
      tile = 0;
      for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
        const int b = item / K::ITEMS_PER_IMG, r0 = (item % K::ITEMS_PER_IMG) * K::ROWS;
        for (int m = 0; m < K::MT; ++m, ++tile) {
          const size_t gp = (size_t)b * NRD_PIX + (r0 + m * 8) * 16 + q * 32 + lane;
#pragma unroll
          for (int h = 0; h < C / 32; ++h) {
            float v[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + tile * C + h * 32, v);
            const uint4* xrow = reinterpret_cast<const uint4*>(fa.x + gp * C + h * 32);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 yk, xk = xrow[j], ok;
              __nv_bfloat162* yh = reinterpret_cast<__nv_bfloat162*>(&yk);
              const __nv_bfloat162* xh = reinterpret_cast<const __nv_bfloat162*>(&xk);
              __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&ok);
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int c = h * 32 + j * 8 + 2 * e;
                yh[e] = __floats2bfloat162_rn(v[j * 8 + 2 * e], v[j * 8 + 2 * e + 1]);
                const float2 yr = __bfloat1622float2(yh[e]), xr = __bfloat1622float2(xh[e]);   // BN sees the STORED y (as backward will)
                oh[e] = __floats2bfloat162_rn(fmaxf(fmaf(yr.x, ss_sm[c], ss_sm[C + c]), 0.f) + xr.x,
                                              fmaxf(fmaf(yr.y, ss_sm[c + 1], ss_sm[C + c + 1]), 0.f) + xr.y);
              }
              reinterpret_cast<uint4*>(out + gp * C + h * 32)[j] = yk;
              reinterpret_cast<uint4*>(fa.xnext + gp * C + h * 32)[j] = ok;
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
}

// ============================================================================================
// Persistent forward over ALL residual blocks (C = 32, at most 2 images per CTA, cooperative launch).
// The block's convolution of an image depends only on that image, so each CTA keeps its images' block input on chip for
// the whole stack: x_0 comes in by TMA (one haloed copy, or 3 kx-shifted row-haloed copies: template parameter ONE), and per block
//     MMA warp     36 tcgen05.mma per image from the shared-memory copy/copies -> 2 accumulators (TMEM) per image
//     pass 1       BN sums straight from TMEM (fp32) -> CTA sum -> returning atomics on count-carrying words (nrd_common.cuh);
//                  y_k (bf16) staged in shared memory for a bulk tensor store
//     exchange     one warp polls its channels' words until every CTA has contributed, derives scale / shift
//     pass 2       x_{k+1} = relu(bn(y_k)) + x_k written (swizzled) into the operand copy/copies = the next block's A operand
//                  (and, ONE, into a staging tile); the store thread sends it to global memory (saved for backward)
// so between blocks nothing is re-read from global memory and there is one launch instead of NB (or 3*NB).
// 576 threads: warp 0 TMA loads + bulk stores, warp 1 MMA, 16 epilogue warps = (channel half, pixel tile, TMEM lane quarter):
// one pixel x 16 channels per thread, image slot 0 then slot 1.
// ============================================================================================
struct PersistFwdArgs {
  const float* gamma; const float* beta;
  float* saved;            // [NB][2C]
  float* scale_shift;      // [NB][2C]
  float* running;          // [2C]
  long long* nbt;
  bf16* xs;                // x_0 .. x_NB   (x_0 is read through the tensor map)
  bf16* ys;                // y_0 .. y_NB-1
  float* partial;          // (unused by the persistent kernel; kept for the per-block kernels)
  unsigned long long* acc; // [NB][2C] 64-bit fixed-point accumulators, zeroed before the launch
  unsigned long long* gbar;
  int B, NB;
  size_t act_elems;        // B*256*C
  double inv_count, unbias; // 1/(B*256), n/(n-1)
  long long* trace;        // optional [gridDim.x][NB][8] clock64 stamps of epilogue thread 0 (profiling aid), or NULL
  int dbg_skip_stores;     // profiling experiment only: drop the global y/x stores (results are then garbage)
  bf16* pooled;            // [B,8,8,C] maxpool2 of x_NB and its arg-max (position 0..3 in the window), written by the last
  uint8_t* idx2;           // block's pass 2; NULL = leave the pooling to the head kernel
  int poll_ns;
  // fused stem (NRD_OPT_FUSE_STEM): x_0 = maxpool2(relu(conv1(images) + b)) computed here instead of loaded; NULL = off
  const float* images;     // [B,3,32,32] fp32
  const float* conv1_wp;   // [27][C] fp32
  const float* conv1_b;    // [C]
  uint8_t* idx0;           // [B,16,16,C] stem pool arg-max
  nrd_bn_sync_args bs;     // DDP buffer sync that the stem kernel would have applied
  unsigned int* status;    // the plan's sticky status word (nrd_acc_add)
};

// ONE = 1: ONE operand copy per image, [18 rows][18 pixels][64 B] with a row AND column halo, instead of three kx-shifted
// copies: SWIZZLE_64B is a function of the shared-memory address, so a tap is just a descriptor start (ky*18 + kx pixels
// further, tools/umma_shift_test.cu) and a pixel tile is 16 image rows x 8 columns (8-row groups one 18-pixel pitch apart).
// A third of the operand writes in pass 2, a third of the shared memory.
template <int ONE>
__global__ void __launch_bounds__(576, 1)
k_tc_fwd_persist(const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmXst,
                 const __grid_constant__ CUtensorMap tmYst, PersistFwdArgs a) {
  constexpr int C = 32;
  using K = Cfg<C>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
  constexpr int PITCH = 18;                                       // ONE: pixels per operand row (16 + column halo)
  constexpr int SLOT_BYTES = ONE ? 21 * 1024 : 3 * K::BUF_BYTES;  // 18*18*64 = 20,736 B rounded up / 55,296 B: three kx-copies
  const uint32_t a_smem = base;                                   // 2 image slots
  const uint32_t w_smem = a_smem + 2 * SLOT_BYTES;               // 9 taps
  constexpr int YST_BYTES = NRD_PIX * C * 2;                      // 16,384 B: y_k of one image, staged for its TMA store
  const uint32_t y_smem = w_smem + 9 * K::WTAP_BYTES;             // 2 image slots (1024-B aligned)
  // ONE: x_{k+1} is staged like y_k for its bulk store (a haloed 18 x 18 box cannot be the source of a tensor store)
  const uint32_t xst_smem = y_smem + 2 * YST_BYTES;
  const uint32_t bar0 = xst_smem + (ONE ? 2 * YST_BYTES : 0);
  const uint32_t wfull = bar0;
  auto xfull = [&](int s) { return bar0 + 8u * (1 + s); };
  auto opready = [&](int s) { return bar0 + 8u * (3 + s); };
  auto tfull = [&](int s, int m) { return bar0 + 8u * (5 + s * 2 + m); };
  const uint32_t tmem_slot = bar0 + 8u * 9;
  auto yready = [&](int s) { return bar0 + 8u * (10 + s); };
  const uint32_t yfree = bar0 + 8u * 12, xfree = bar0 + 8u * 13;
  const uint32_t st_afull = bar0 + 8u * 14, st_tfull = bar0 + 8u * 15;   // fused stem: im2col tiles written / accumulators ready
  const uint32_t st_x0free = bar0 + 8u * 16;                             // ... x_0's bulk store has read the centre copies
  constexpr int BAR_BYTES = 8 * 18;
  constexpr int MISC_OFF = 2 * SLOT_BYTES + 9 * K::WTAP_BYTES + (ONE ? 4 : 2) * YST_BYTES + BAR_BYTES;
  float* red = reinterpret_cast<float*>(gen_base + MISC_OFF);            // [16 warps][sum | sum of squares][16 channels]
  float* ss_sm = red + 1024;                                             // [2C] scale | shift of the current block
  // fused stem scratch: W1 tile (1024-B aligned), bias, two zero-padded fp32 images; the four im2col tiles of a half image
  // alias the y staging tiles (free until block 0's pass 1)
  constexpr int STEM_OFF = (MISC_OFF + (1024 + 64) * 4 + 1023) & ~1023;
  uint8_t* st_w_gen = gen_base + STEM_OFF;                               // [32 co][32 taps] bf16, SWIZZLE_64B rows
  float* st_bias = reinterpret_cast<float*>(st_w_gen + 2048);
  float* st_img = st_bias + 32;                                          // [2][3][34][40] fp32
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int img0 = blockIdx.x, img1 = blockIdx.x + gridDim.x;
  const int nslots = (img1 < a.B) ? 2 : 1;
  constexpr uint32_t TMEM_COLS = 256;                              // 2 slots x 2 tiles x 32 columns + 4 stem window accumulators
  const bool fuse_stem = a.images != nullptr;
  if (a.trace && threadIdx.x == 64) a.trace[(size_t)blockIdx.x * a.NB * 8 + 7] = clock64();      // kernel entry
  nrd_pdl_launch();                                                // the head kernels behind us may be scheduled (they wait for our exit)

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX0); tma_prefetch_desc(&tmW);
    mbar_init(wfull, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(xfull(s), 1); mbar_init(opready(s), 512); mbar_init(tfull(s, 0), 1); mbar_init(tfull(s, 1), 1); mbar_init(yready(s), 512);
    }
    mbar_init(yfree, 1); mbar_init(xfree, 1);
    mbar_init(st_afull, 512); mbar_init(st_tfull, 1); mbar_init(st_x0free, 1);
    if (fuse_stem) { mbar_init(xfull(0), 512); mbar_init(xfull(1), 512); }      // x_0 is written by the epilogue threads, not by TMA
    tma_prefetch_desc(&tmXst); tma_prefetch_desc(&tmYst);
    fence_barrier_init();
  }
  if (fuse_stem) {
    // the operand copies' halo rows / edge columns are the convolution padding (TMA zero-fills them on the loading path) and
    // the padded images' borders are conv1's: zero everything once
    for (int i = threadIdx.x; i < 2 * SLOT_BYTES / 16; i += blockDim.x) reinterpret_cast<uint4*>(gen_base)[i] = make_uint4(0, 0, 0, 0);
    // images first (the longest latency of the prologue): interior by cp.async, borders zeroed by plain stores (disjoint)
    if (threadIdx.x >= 64) {
      const uint32_t im_s = smem_u32(st_img);
      for (int c = threadIdx.x - 64; c < nslots * 768; c += 512) {                 // 768 16-byte chunks per image
        const int s_ = c / 768, qq = c % 768, ci = qq >> 8, r = (qq >> 3) & 31, c4 = qq & 7;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(im_s + (uint32_t)((s_ * 4080 + (ci * 34 + r + 1) * 40 + 4 + c4 * 4) * 4)),
                     "l"(a.images + (size_t)(s_ == 0 ? img0 : img1) * 3072 + ci * 1024 + r * 32 + c4 * 4) : "memory");
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
    for (int i = threadIdx.x; i < 2 * 3 * 34 * 40; i += blockDim.x) {
      const int r = (i / 40) % 34, c = i % 40;
      if (r == 0 || r == 33 || c < 4 || c >= 36) st_img[i] = 0.f;
    }
    if (threadIdx.x < 128) {                                       // W1 -> bf16 K-major rows; row co, chunk ch = taps ch*8..+7
      const int co = threadIdx.x >> 2, ch = threadIdx.x & 3;
      uint32_t pk[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int t0 = ch * 8 + 2 * e;
        const __nv_bfloat162 hh = __floats2bfloat162_rn(t0 < 27 ? a.conv1_wp[t0 * C + co] : 0.f, t0 + 1 < 27 ? a.conv1_wp[(t0 + 1) * C + co] : 0.f);
        pk[e] = *reinterpret_cast<const uint32_t*>(&hh);
      }
      *reinterpret_cast<uint4*>(st_w_gen + co * 64 + ((ch ^ ((co >> 1) & 3)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
    }
    if (threadIdx.x < 32) st_bias[threadIdx.x] = a.conv1_b[threadIdx.x];
    fence_proxy_async();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(wfull, 9 * K::WTAP_BYTES);
      for (int t = 0; t < 9; ++t) tma_load_2d(w_smem + t * K::WTAP_BYTES, &tmW, wfull, t * C, 0);
      if (!fuse_stem) {
        for (int s = 0; s < nslots; ++s) {
          if (ONE) {
            mbar_expect_tx(xfull(s), PITCH * PITCH * 64);
            tma_load_4d(a_smem + s * SLOT_BYTES, &tmX0, xfull(s), 0, -1, -1, s == 0 ? img0 : img1);     // box 18 x 18: halo zero-filled
          } else {
            mbar_expect_tx(xfull(s), SLOT_BYTES);
            for (int kx = 0; kx < 3; ++kx)
              tma_load_4d(a_smem + s * SLOT_BYTES + kx * K::BUF_BYTES, &tmX0, xfull(s), 0, kx - 1, -1, s == 0 ? img0 : img1);
          }
        }
      } else {
        // fused stem: x_0 was produced on chip; it still goes to global memory (block 0's weight gradient and the stem
        // backward read it), straight from the centre copies
        for (int s = 0; s < nslots; ++s) {
          mbar_wait(xfull(s), 0);
          tma_store_5d(&tmXst, ONE ? xst_smem + s * YST_BYTES : a_smem + s * SLOT_BYTES + K::BUF_BYTES + K::IMGROW_BYTES, 0, 0, 0, s == 0 ? img0 : img1, 0);
        }
        bulk_commit(); bulk_wait_read0();
        mbar_arrive(st_x0free);
      }
      // ... then this thread is the store engine: y_k from its staging tile and x_{k+1} from the CENTRE operand copy (the same
      // swizzled bytes the tensor map expects) go to global memory as two bulk tensor stores per image, instead of eight
      // 16-byte stores per thread
      for (int k = 0; k < a.NB; ++k) {
        for (int s = 0; s < nslots; ++s) {
          mbar_wait(yready(s), (uint32_t)(k & 1));
          if (!a.dbg_skip_stores) tma_store_5d(&tmYst, y_smem + s * YST_BYTES, 0, 0, 0, s == 0 ? img0 : img1, k);
        }
        bulk_commit(); bulk_wait_read0();
        mbar_arrive(yfree);
        for (int s = 0; s < nslots; ++s) {
          mbar_wait(opready(s), (uint32_t)(k & 1));
          if (!a.dbg_skip_stores)
            tma_store_5d(&tmXst, ONE ? xst_smem + s * YST_BYTES : a_smem + s * SLOT_BYTES + K::BUF_BYTES + K::IMGROW_BYTES, 0, 0, 0, s == 0 ? img0 : img1, k + 1);
        }
        bulk_commit(); bulk_wait_read0();
        mbar_arrive(xfree);
      }
      bulk_wait0();
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(128, C, 0, 0);
      if (fuse_stem) {
        // stem: per half image, D_w[pooled pixel][co] = im2col_w[pooled pixel][27 taps -> 32] x W1, one accumulator per
        // position w of the 2x2 pooling window (the epilogue takes the maximum over the four)
        const uint32_t tiles = y_smem, w1 = base + STEM_OFF;
        const uint64_t bd = make_sdesc(w1, 16, K::SBO, K::LAYOUT);
        for (int i = 0; i < 2 * nslots; ++i) {
          mbar_wait(st_afull, (uint32_t)(i & 1));
          tc_fence_after();
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            const uint64_t ad = make_sdesc(tiles + w * 8192, 16, K::SBO, K::LAYOUT);
            umma_bf16(tmem_base + 128 + w * C, ad, bd, idesc, 0);
            umma_bf16(tmem_base + 128 + w * C, ad + 2, bd + 2, idesc, 1);
          }
          umma_commit(st_tfull);
        }
      }
      mbar_wait(wfull, 0);
      for (int k = 0; k < a.NB; ++k) {
        for (int s = 0; s < nslots; ++s) {        // slot-major: slot 0's MMAs start as soon as ITS operands are rewritten
          if (k == 0) mbar_wait(xfull(s), 0);
          else mbar_wait(opready(s), (uint32_t)((k - 1) & 1));
          tc_fence_after();
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            const uint32_t d_tmem = tmem_base + (s * 2 + m) * C;
#pragma unroll
            for (int t = 0; t < 9; ++t) {
              // ONE: tile m = columns 8m..8m+7 of all 16 rows; tap (ky, kx) starts ky rows and kx pixels into the haloed copy
              const uint32_t a_addr = ONE ? a_smem + s * SLOT_BYTES + ((t / 3) * PITCH + m * 8 + (t % 3)) * 64
                                          : a_smem + s * SLOT_BYTES + (t % 3) * K::BUF_BYTES + (m * 8 + t / 3) * K::IMGROW_BYTES;
              const uint64_t ad = make_sdesc(a_addr, 16, ONE ? PITCH * 64 : K::SBO, K::LAYOUT);
              const uint64_t bd = make_sdesc(w_smem + t * K::WTAP_BYTES, 16, K::SBO, K::LAYOUT);
#pragma unroll
              for (int kk = 0; kk < K::KSTEPS; ++kk) umma_bf16(d_tmem, ad + 2 * kk, bd + 2 * kk, idesc, (t | kk) != 0);
            }
            umma_commit(tfull(s, m));
          }
        }
      }
    }
  } else {
    // ===== epilogue: 16 warps; warp = (channel half h, tile m, TMEM lane quarter q); thread = one pixel x 16 channels, of =====
    // ===== image slot 0 and then of slot 1: slot 0's operands are ready (and its MMAs run) while slot 1 is still in pass 2 =====
    const int ew = warp - 2, h = ew >> 3, m = (ew >> 2) & 1, q = warp & 3, et = threadIdx.x - 64;    // et: 0..511
    // pixel of this thread = row of its TMEM lane in tile m: 8 rows x 16 columns per tile, or (ONE) 16 rows x 8 columns
    const int p = q * 32 + lane, pi = ONE ? q * 4 + (lane >> 3) : m * 8 + (p >> 4), pj = ONE ? m * 8 + (lane & 7) : (p & 15);
    const int P1 = ONE ? (pi + 1) * PITCH + pj + 1 : (pi + 1) * 16 + pj;       // its index inside the (centre) operand copy
    constexpr int CEN = ONE ? 0 : K::BUF_BYTES;                                // byte offset of the centre copy in a slot
    const int P0y = pi * 16 + pj;                                              // ... and inside the y staging tile
    const uint32_t my_tmem = tmem_base + ((uint32_t)(q * 32) << 16) + m * C + h * 16;   // + s * 2C
    const float* ss_w = ss_sm + h * 16;                                        // this thread's channels: scale at [c], shift at [C + c]
    const float gamma_r = a.gamma[lane], beta_r = a.beta[lane];             // off the per-block critical path (used by warp ew == 0)
    // the LAST row of the previous launch could not be cleaned inside that launch (its readers were still polling): do it now,
    // long before anyone adds to it -- every CTA reaches that row only after seeing this CTA's contribution to row 0, which
    // is issued after the fence
    if (blockIdx.x == 0 && ew == 0) {
      unsigned long long* last = a.acc + (size_t)(a.NB - 1) * NRD_ACC_ROW;
#pragma unroll
      for (int r = 0; r < NRD_ACC_REPL; ++r) { last[NRD_ACC_IDX(r, lane)] = 0ull; last[NRD_ACC_IDX(r, C + lane)] = 0ull; }
      __threadfence();
    }
    if (fuse_stem) {
      // ===== fused stem: x_0 = maxpool2(relu(conv1(image) + b)) of this CTA's images, straight into the operand copies =====
      constexpr int IMG_ROW = 40, IMG_FLOATS = 3 * 34 * IMG_ROW;     // zero-padded rows -1..32, data column c at index 4 + c
      if (blockIdx.x == gridDim.x - 1 && ew == 0 && a.bs.slot) {     // DDP buffer sync, by the thread that does the EMA later
        a.bs.running[lane] = a.bs.slot[lane] * a.bs.inv_world; a.bs.running[C + lane] = a.bs.slot[C + lane] * a.bs.inv_world;
      }
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      asm volatile("bar.sync 1, 512;" ::: "memory");
      long long* trs = (a.trace && et == 0) ? a.trace + (size_t)gridDim.x * a.NB * 8 + (size_t)blockIdx.x * 16 : nullptr;   // stem stamps
      if (trs) trs[0] = clock64();
      uint8_t* tiles = gen_base + (y_smem - base);
      for (int i = 0; i < 2 * nslots; ++i) {
        const int s_ = i >> 1, hh = i & 1;
        const float* im = st_img + s_ * IMG_FLOATS;
        if (i > 0) mbar_wait(st_tfull, (uint32_t)((i - 1) & 1));     // the previous half's MMAs have read the tiles
        {
          // builder: thread = (pooled pixel bp of the half, window position w): one 64-byte im2col row (27 taps, bf16)
          const int bp = et & 127, w = et >> 7, pr = 8 * hh + (bp >> 4), pc = bp & 15, dy = w >> 1, dx = w & 1;
          float kv[32];
#pragma unroll
          for (int ci = 0; ci < 3; ++ci)
#pragma unroll
            for (int ky = 0; ky < 3; ++ky)
#pragma unroll
              for (int kx = 0; kx < 3; ++kx) kv[ci * 9 + ky * 3 + kx] = im[(ci * 34 + 2 * pr + dy + ky) * IMG_ROW + 3 + 2 * pc + dx + kx];
          kv[27] = kv[28] = kv[29] = kv[30] = kv[31] = 0.f;
          uint8_t* row = tiles + w * 8192 + bp * 64;
          const int sw = (bp >> 1) & 3;
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            uint32_t pk[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const __nv_bfloat162 hv = __floats2bfloat162_rn(kv[ch * 8 + 2 * e], kv[ch * 8 + 2 * e + 1]);
              pk[e] = *reinterpret_cast<const uint32_t*>(&hv);
            }
            *reinterpret_cast<uint4*>(row + ((ch ^ sw) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          }
        }
        fence_proxy_async();
        mbar_arrive(st_afull);
        if (trs) trs[1 + i * 3] = clock64();
        if (m == hh) {
          // epilogue of the half (the warps whose pixel tile it is; the others already build the next half's tiles): this
          // thread's pooled pixel x 16 channels; maximum over the four accumulators, first one wins
          mbar_wait(st_tfull, (uint32_t)(i & 1));
          tc_fence_after();
          if (trs) trs[2 + i * 3] = clock64();
          float mx[16]; uint32_t mi[4] = {0u, 0u, 0u, 0u};
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            float v[16];
            tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + 128 + w * C + h * 16, v);
#pragma unroll
            for (int c = 0; c < 16; ++c) {
              if (w == 0) mx[c] = v[c];
              else if (v[c] > mx[c]) { mx[c] = v[c]; mi[c >> 2] = (mi[c >> 2] & ~(0xFFu << (8 * (c & 3)))) | ((uint32_t)w << (8 * (c & 3))); }
            }
          }
          tc_fence_before();
          uint8_t* slot_gen = gen_base + s_ * SLOT_BYTES;
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int ch = h * 2 + j;
            uint32_t pk[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int c = j * 8 + 2 * e;
              const __nv_bfloat162 hv = __floats2bfloat162_rn(fmaxf(mx[c] + st_bias[h * 16 + c], 0.f), fmaxf(mx[c + 1] + st_bias[h * 16 + c + 1], 0.f));
              pk[e] = *reinterpret_cast<const uint32_t*>(&hv);
            }
            const uint4 ok = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            // the stem's accumulator rows are pooled pixels in row-major order of the half image (8 rows x 16 columns)
            const int sj = p & 15, P1s = ONE ? (hh * 8 + (p >> 4) + 1) * PITCH + sj + 1 : (hh * 8 + (p >> 4) + 1) * 16 + sj;
            *reinterpret_cast<uint4*>(slot_gen + CEN + P1s * 64 + ((ch ^ ((P1s >> 1) & 3)) << 4)) = ok;
            if (ONE) { const int P0s = (hh * 8 + (p >> 4)) * 16 + sj; *reinterpret_cast<uint4*>(gen_base + (xst_smem - base) + s_ * YST_BYTES + P0s * 64 + ((ch ^ ((P0s >> 1) & 3)) << 4)) = ok; }
            if (!ONE && sj < 15) { const int P0 = P1s + 1; *reinterpret_cast<uint4*>(slot_gen + P0 * 64 + ((ch ^ ((P0 >> 1) & 3)) << 4)) = ok; }
            if (!ONE && sj > 0) { const int P2 = P1s - 1; *reinterpret_cast<uint4*>(slot_gen + 2 * K::BUF_BYTES + P2 * 64 + ((ch ^ ((P2 >> 1) & 3)) << 4)) = ok; }
          }
          *reinterpret_cast<uint4*>(a.idx0 + ((size_t)(s_ == 0 ? img0 : img1) * NRD_PIX + hh * 128 + p) * C + h * 16) = make_uint4(mi[0], mi[1], mi[2], mi[3]);
          if (trs) trs[3 + i * 3] = clock64();
        }
      }
      fence_proxy_async();
      for (int s_ = 0; s_ < nslots; ++s_) mbar_arrive(xfull(s_));
    }
    for (int k = 0; k < a.NB; ++k) {
      long long* tr = (a.trace && et == 0) ? a.trace + ((size_t)blockIdx.x * a.NB + k) * 8 : nullptr;
      if (tr) tr[0] = clock64();
      // ---- pass 1: y_k store + per-channel sum / sum of squares of this warp's 32 pixels x 16 channels ----
      float ssum = 0.f, qsum = 0.f;
      uint4 ypk[2];
      auto stage_y = [&](int s) {      // ypk -> slot s's staging tile (swizzled as the tensor map expects); the store thread sends it off
        if (s == 0 && k > 0) mbar_wait(yfree, (uint32_t)((k - 1) & 1));     // block k-1's y stores have read the tiles
        uint8_t* yst = gen_base + (y_smem - base) + s * YST_BYTES;
#pragma unroll
        for (int j = 0; j < 2; ++j) *reinterpret_cast<uint4*>(yst + P0y * 64 + (((h * 2 + j) ^ ((P0y >> 1) & 3)) << 4)) = ypk[j];
        fence_proxy_async();
        mbar_arrive(yready(s));
      };
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        if (s >= nslots) break;
        mbar_wait(tfull(s, m), (uint32_t)(k & 1));
        tc_fence_after();
        if (tr && s == 0) tr[1] = clock64();
        float v[16];
        tmem_ld16(my_tmem + s * 2 * C, v);
        // y_k (bf16) for the staging tile; the LAST slot's goes there only after the statistics have left (off the critical path)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&ypk[j]);
#pragma unroll
          for (int e = 0; e < 4; ++e) ph[e] = __floats2bfloat162_rn(v[j * 8 + 2 * e], v[j * 8 + 2 * e + 1]);
        }
        if (s + 1 < nslots) stage_y(s);
        {
          float sq[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) sq[i] = v[i] * v[i];
          qsum += warp_col_sum16(sq, lane);
        }
        ssum += warp_col_sum16(v, lane);       // lanes 2c, 2c+1: channel h*16 + c
      }
      if (tr) tr[2] = clock64();
      red[ew * 32 + (lane & 1) * 16 + (lane >> 1)] = (lane & 1) ? qsum : ssum;
      asm volatile("bar.sync 1, 512;" ::: "memory");
      // cross-CTA sum WITHOUT a partial matrix and WITHOUT a separate barrier: every CTA adds its 2C per-channel sums as fixed
      // point (2^-24) into one word per statistic that also counts the contributors (nrd_common.cuh); integer addition is
      // associative, so the result is bit-identical whatever order the CTAs arrive in
      unsigned long long* acc = a.acc + (size_t)k * NRD_ACC_ROW;      // [replica][2C] words, NRD_ACC_WS apart
      if (et < 2 * C) {
        const int which = et >> 5, ch = et & 31;
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[((ch >> 4) * 8 + w) * 32 + which * 16 + (ch & 15)];   // fixed order over the 8 warps of the half
        nrd_acc_check(nrd_acc_add(acc + NRD_ACC_IDX(blockIdx.x % NRD_ACC_REPL, et), t, 16777216.0, a.status), gridDim.x);   // wait for the return: measured faster
      }
      stage_y(nslots - 1);
      if (tr) tr[3] = clock64();
      // one warp (lane = channel) polls its channels' words until every CTA has contributed, derives the coefficients (ONE
      // warp's worth of the slow fp64 pipe) and publishes them with the only CTA-wide barrier of the wait
      if (ew == 0) {
        unsigned long long ws, wq;
        nrd_acc_wait2(acc + NRD_ACC_IDX(0, lane), acc + NRD_ACC_IDX(0, C + lane), gridDim.x, ws, wq, a.poll_ns);
        if (tr) tr[4] = clock64();
        // every CTA has contributed to this row, so every CTA is done reading the previous one: clean it for the next launch
        if (blockIdx.x == 0 && k > 0) {
#pragma unroll
          for (int r = 0; r < NRD_ACC_REPL; ++r) { (acc - NRD_ACC_ROW)[NRD_ACC_IDX(r, lane)] = 0ull; (acc - NRD_ACC_ROW)[NRD_ACC_IDX(r, C + lane)] = 0ull; }
        }
        const double S = nrd_acc_value(ws, gridDim.x, 1.0 / 16777216.0), Q = nrd_acc_value(wq, gridDim.x, 1.0 / 16777216.0);
        const double mean = S * a.inv_count;
        double var = Q * a.inv_count - mean * mean;
        if (var < 0.0) var = 0.0;
        const float invstd = fast_invstd(var + (double)NRD_BN_EPS);
        const float sc = gamma_r * invstd, sh = beta_r - (float)mean * sc;
        ss_sm[lane] = sc; ss_sm[C + lane] = sh;
        if (blockIdx.x == gridDim.x - 1) {      // the last CTA has the fewest images: the bookkeeping is off the grid's critical path
          a.saved[(size_t)k * 2 * C + lane] = (float)mean; a.saved[(size_t)k * 2 * C + C + lane] = invstd;
          a.scale_shift[(size_t)k * 2 * C + lane] = sc; a.scale_shift[(size_t)k * 2 * C + C + lane] = sh;
          const double unbiased = var * a.unbias;
          a.running[lane] = (1.f - NRD_BN_MOMENTUM) * a.running[lane] + NRD_BN_MOMENTUM * (float)mean;
          a.running[C + lane] = (1.f - NRD_BN_MOMENTUM) * a.running[C + lane] + NRD_BN_MOMENTUM * (float)unbiased;
          if (lane == 0) *a.nbt += 1;
        }
      }
      asm volatile("bar.sync 1, 512;" ::: "memory");
      if (tr) tr[5] = clock64();
      // ---- pass 2: x_{k+1} = relu(bn(y_k)) + x_k -> the three operand copies in shared memory, then global ----
      if (k > 0) mbar_wait(xfree, (uint32_t)((k - 1) & 1));       // block k-1's x stores have read the centre copies
      else if (fuse_stem) mbar_wait(st_x0free, 0);
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        if (s >= nslots) break;
        if (k == 0) mbar_wait(xfull(s), 0);      // the TMA-written x_0 copy is read below (residual) by this thread itself
        uint8_t* slot_gen = gen_base + s * SLOT_BYTES;
        float v[16];
        tmem_ld16(my_tmem + s * 2 * C, v);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          // residual x_k from the centre copy (kx = 1); 16-B chunk ch of pixel P sits at chunk ch ^ ((P>>1)&3) (SWIZZLE_64B)
          const int ch = h * 2 + j;
          const uint4 xk = *reinterpret_cast<const uint4*>(slot_gen + CEN + P1 * 64 + ((ch ^ ((P1 >> 1) & 3)) << 4));
          const __nv_bfloat162* xh = reinterpret_cast<const __nv_bfloat162*>(&xk);
          uint4 ok;
          __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&ok);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int c = j * 8 + 2 * e;
            const float2 yr = __bfloat1622float2(__floats2bfloat162_rn(v[c], v[c + 1]));   // BN sees the STORED y (as backward will)
            const float2 xr = __bfloat1622float2(xh[e]);
            const float2 sc2 = *reinterpret_cast<const float2*>(ss_w + c), sh2 = *reinterpret_cast<const float2*>(ss_w + C + c);
            oh[e] = __floats2bfloat162_rn(fmaxf(fmaf(yr.x, sc2.x, sh2.x), 0.f) + xr.x, fmaxf(fmaf(yr.y, sc2.y, sh2.y), 0.f) + xr.y);
          }
          *reinterpret_cast<uint4*>(slot_gen + CEN + P1 * 64 + ((ch ^ ((P1 >> 1) & 3)) << 4)) = ok;   // (!ONE: also the store's source)
          if (ONE) *reinterpret_cast<uint4*>(gen_base + (xst_smem - base) + s * YST_BYTES + P0y * 64 + ((ch ^ ((P0y >> 1) & 3)) << 4)) = ok;
          if (!ONE && k + 1 < a.NB) {
            if (pj < 15) { const int P0 = P1 + 1; *reinterpret_cast<uint4*>(slot_gen + P0 * 64 + ((ch ^ ((P0 >> 1) & 3)) << 4)) = ok; }
            if (pj > 0) { const int P2 = P1 - 1; *reinterpret_cast<uint4*>(slot_gen + 2 * K::BUF_BYTES + P2 * 64 + ((ch ^ ((P2 >> 1) & 3)) << 4)) = ok; }
          } else if (k + 1 == a.NB && a.pooled) {
            // last block: max-pool 2x2 of the network body's output right here.  A warp holds two image rows (lanes 0..15 and
            // 16..31), so the window of the even/even pixel is lanes l, l^1, l^16, l^17 = positions 0..3; first maximum wins
            // (k_pool_fc1_fwd_part's rule).  The even/even lane writes 8 pooled channels and their arg-max bytes.
            const uint32_t own[4] = {ok.x, ok.y, ok.z, ok.w};
            uint32_t pk[4]; uint32_t ib[2] = {0u, 0u};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const uint32_t o1 = __shfl_xor_sync(0xffffffffu, own[e], 1), o2 = __shfl_xor_sync(0xffffffffu, own[e], ONE ? 8 : 16),
                             o3 = __shfl_xor_sync(0xffffffffu, own[e], ONE ? 9 : 17);      // next row: 8 lanes on (ONE), else 16
              const uint32_t w4[4] = {own[e], o1, o2, o3};
              float2 mx = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w4[0]));
              uint32_t i0 = 0, i1 = 0;
#pragma unroll
              for (int qq = 1; qq < 4; ++qq) {
                const float2 vv = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w4[qq]));
                if (vv.x > mx.x) { mx.x = vv.x; i0 = qq; }
                if (vv.y > mx.y) { mx.y = vv.y; i1 = qq; }
              }
              const __nv_bfloat162 pm = __floats2bfloat162_rn(mx.x, mx.y);      // exact: the maxima are bf16 values
              pk[e] = *reinterpret_cast<const uint32_t*>(&pm);
              ib[e >> 1] |= (i0 | (i1 << 8)) << ((e & 1) * 16);
            }
            if (((pi | pj) & 1) == 0) {
              const size_t po = ((size_t)(s == 0 ? img0 : img1) * 64 + (pi >> 1) * 8 + (pj >> 1)) * C + ch * 8;
              *reinterpret_cast<uint4*>(a.pooled + po) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
              *reinterpret_cast<uint2*>(a.idx2 + po) = make_uint2(ib[0], ib[1]);
            }
          }
        }
        // TMEM reads and shared-memory operand writes of this thread for slot s are done: release the slot to the MMA warp
        // (and the centre copy to the store thread)
        tc_fence_before();
        fence_proxy_async();
        mbar_arrive(opready(s));
      }
      if (tr) tr[6] = clock64();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
  if (a.trace && threadIdx.x == 64) a.trace[((size_t)blockIdx.x * a.NB + a.NB - 1) * 8 + 7] = clock64();   // kernel exit
}

// ============================================================================================
// Streaming persistent forward: ALL residual blocks, ANY batch size (C = 32, cooperative launch) -- the kernel for batches
// beyond the two images per CTA that k_tc_fwd_persist keeps on chip.  A CTA owns images j = 0 .. n-1 (blockIdx.x + j * grid) and
// streams them through two haloed operand slots per block:
//   phase 1, per image   DMA thread: x_k[img] by ONE TMA load (18 x 18 box, halo zero-filled) into operand slot j & 1;
//                        MMA warp: 36 tcgen05.mma into TMEM slot j & 3;  epilogue: BN sums straight from TMEM (accumulated over
//                        the CTA's images), y_k staged and bulk-stored
//   exchange             the count-carrying fixed-point words of k_tc_fwd_persist
//   phase 2, per image   epilogue: x_{k+1} = relu(bn(y_k)) + x_k from the STORED y_k and x_k (this thread's own 32 B of each, L2
//                        hits: written / loaded a few microseconds earlier), staged and bulk-stored: the next block's TMA source
// so every tensor crosses L2 once per use and HBM at most once, with ONE launch for the whole stack and no per-block kernel
// boundary (the per-block path it replaces: one fused launch per block, accumulators of <= 8 images per CTA).
// 576 threads: warp 0 lane 0 = DMA (TMA loads and bulk stores), warp 1 lane 0 = MMA issue, 16 epilogue warps as in
// k_tc_fwd_persist<1> (thread = one pixel x 16 channels; pixel tile m = columns 8m .. 8m+7 of all 16 rows).
// ============================================================================================
constexpr int STREAM_NSLOT = 4;          // haloed operand slots of the streaming forward (21 KB each)
constexpr int STREAM_NT = 4;             // its TMEM image slots (64 columns each).  Measured: 6 operand / 8 TMEM slots (more of the next block's
//                                          MMAs under pass 2) shorten pass 1 by 4k cycles per block and lengthen pass 2 by 8k (B = 2048): rejected
constexpr int STREAM_BAR_BYTES = 8 * 48;
struct StreamFwdArgs {
  const float* gamma; const float* beta;
  float* saved; float* scale_shift; float* running; long long* nbt;
  const bf16* xs; const bf16* ys;             // the x_k / y_k stacks (phase 2 reads this thread's own pixels back)
  unsigned long long* acc;
  int B, NB;
  size_t act_elems;
  double inv_count, unbias;
  int poll_ns;
  unsigned int* status;
  long long* trace;           // optional clock64 stamps [CTA][block][8] (profiling aid), or NULL
};

__global__ void __launch_bounds__(576, 1)
k_tc_fwd_stream(const __grid_constant__ CUtensorMap tmX18, const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmXst,
                const __grid_constant__ CUtensorMap tmYst, StreamFwdArgs a) {
  constexpr int C = 32;
  using K = Cfg<C>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
  constexpr int PITCH = 18, SLOT_BYTES = 21 * 1024, YST_BYTES = NRD_PIX * C * 2, NSLOT = STREAM_NSLOT, NT = STREAM_NT;
  const uint32_t a_smem = base;                                    // NSLOT operand slots
  const uint32_t w_smem = a_smem + NSLOT * SLOT_BYTES;             // 9 taps
  const uint32_t y_smem = w_smem + 9 * K::WTAP_BYTES;              // 2 y staging tiles
  const uint32_t xst_smem = y_smem + 2 * YST_BYTES;                // 2 x staging tiles
  const uint32_t bar0 = xst_smem + 2 * YST_BYTES;
  const uint32_t wfull = bar0;
  auto xfull = [&](int s) { return bar0 + 8u * (1 + s); };         // operand slot loaded (TMA bytes)
  auto opfree = [&](int s) { return bar0 + 8u * (1 + NSLOT + s); };        // the MMAs that read operand slot s are done
  auto tfull = [&](int t, int m) { return bar0 + 8u * (1 + 2 * NSLOT + t * 2 + m); };   // accumulator (TMEM slot t, tile m) ready
  auto tfree = [&](int t) { return bar0 + 8u * (1 + 2 * NSLOT + 2 * NT + t); };          // pass 1 has read TMEM slot t (512 arrivals)
  constexpr int BX = 1 + 2 * NSLOT + 3 * NT;
  auto yready = [&](int s) { return bar0 + 8u * (BX + s); };       // y staged (512)
  auto xready = [&](int s) { return bar0 + 8u * (BX + 2 + s); };   // x staged (512)
  auto yfree = [&](int s) { return bar0 + 8u * (BX + 4 + s); };    // the bulk store has read y staging tile s
  auto xfree = [&](int s) { return bar0 + 8u * (BX + 6 + s); };
  const uint32_t p1done = bar0 + 8u * (BX + 8);                    // every y store of this block's phase 1 is complete
  const uint32_t tmem_slot = bar0 + 8u * (BX + 9);
  constexpr int BAR_BYTES = STREAM_BAR_BYTES;
  static_assert(8 * (BX + 10) <= BAR_BYTES, "barrier area");
  constexpr int MISC_OFF = NSLOT * SLOT_BYTES + 9 * K::WTAP_BYTES + 4 * YST_BYTES + BAR_BYTES;
  float* red = reinterpret_cast<float*>(gen_base + MISC_OFF);            // [16 warps][sum | sum of squares][16 channels]
  float* ss_sm = red + 1024;                                             // [2C] scale | shift of the current block
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nimg = (a.B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;      // images of this CTA (>= 1)
  auto img_of = [&](int j) { return (int)blockIdx.x + j * (int)gridDim.x; };
  constexpr uint32_t TMEM_COLS = NT * 2 * C;                       // NT image slots x 2 tiles x 32 columns
  // Slots rotate over the GLOBAL image counter g = k * nimg + j (operand slot g % NSLOT, TMEM slot g % NT, staging tile g & 1), so
  // the n-th use of a slot is g / slots (operand slot g % NSLOT, TMEM slot g % NT) and every mbarrier parity follows from g alone.

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX18); tma_prefetch_desc(&tmW); tma_prefetch_desc(&tmXst); tma_prefetch_desc(&tmYst);
    mbar_init(wfull, 1); mbar_init(p1done, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(yready(s), 512); mbar_init(xready(s), 512); mbar_init(yfree(s), 1); mbar_init(xfree(s), 1); }
    for (int s = 0; s < NSLOT; ++s) { mbar_init(xfull(s), 1); mbar_init(opfree(s), 1); }
    for (int t = 0; t < NT; ++t) { mbar_init(tfull(t, 0), 1); mbar_init(tfull(t, 1), 1); mbar_init(tfree(t), 512); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;

  if (warp == 0) {
    if (lane == 0) {
      // ===== DMA thread =====
      mbar_expect_tx(wfull, 9 * K::WTAP_BYTES);
      for (int t = 0; t < 9; ++t) tma_load_2d(w_smem + t * K::WTAP_BYTES, &tmW, wfull, t * C, 0);
      auto load = [&](int k, int j) {
        const int g = k * nimg + j, s = g % NSLOT, u = g / NSLOT;
        if (u > 0) mbar_wait(opfree(s), (u - 1) & 1);              // the previous occupant's MMAs are done
        mbar_expect_tx(xfull(s), PITCH * PITCH * 64);
        tma_load_5d(a_smem + s * SLOT_BYTES, &tmX18, xfull(s), 0, -1, -1, img_of(j), k);   // box 18 x 18: halo zero-filled
      };
      const int npre = nimg < NSLOT ? nimg : NSLOT;                // loads issued ahead of a block's first MMA
      for (int j = 0; j < npre; ++j) load(0, j);
      for (int k = 0; k < a.NB; ++k) {
        // phase 1: NSLOT loads ahead, y stores behind
        for (int j = 0; j < nimg; ++j) {
          const int g = k * nimg + j, s = g & 1;
          if (j + NSLOT < nimg) load(k, j + NSLOT);                // as soon as image j's MMAs have read the slot
          mbar_wait(yready(s), (g >> 1) & 1);
          tma_store_5d(&tmYst, y_smem + s * YST_BYTES, 0, 0, 0, img_of(j), k);
          bulk_commit(); bulk_wait_read0();
          mbar_arrive(yfree(s));
        }
        bulk_wait0();                                              // y_k is in memory: phase 2 reads it back
        fence_proxy_async();
        mbar_arrive(p1done);
        // phase 2: x_{k+1} stores; the next block's first operands are requested as soon as their stores are complete
        const bool more = k + 1 < a.NB;
        for (int j = 0; j < nimg; ++j) {
          const int g = k * nimg + j, s = g & 1;
          mbar_wait(xready(s), (g >> 1) & 1);
          tma_store_5d(&tmXst, xst_smem + s * YST_BYTES, 0, 0, 0, img_of(j), k + 1);
          bulk_commit(); bulk_wait_read0();
          mbar_arrive(xfree(s));
          if (more && j >= 1 && j - 1 < npre) { bulk_wait1(); load(k + 1, j - 1); }
        }
        bulk_wait0();                                              // x_{k+1} is in memory
        fence_proxy_async();
        if (more && nimg - 1 < npre) load(k + 1, nimg - 1);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer =====
      constexpr uint32_t idesc = make_idesc(128, C, 0, 0);
      mbar_wait(wfull, 0);
      const int total = a.NB * nimg;
      for (int g = 0; g < total; ++g) {
        const int s = g % NSLOT, t = g % NT;
        mbar_wait(xfull(s), (g / NSLOT) & 1);
        if (g >= NT) mbar_wait(tfree(t), ((g / NT) - 1) & 1);      // pass 1 has read this TMEM slot's previous accumulators
        tc_fence_after();
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          const uint32_t d_tmem = tmem_base + (t * 2 + m) * C;
#pragma unroll
          for (int tp = 0; tp < 9; ++tp) {
            const uint32_t a_addr = a_smem + s * SLOT_BYTES + ((tp / 3) * PITCH + m * 8 + (tp % 3)) * 64;
            const uint64_t ad = make_sdesc(a_addr, 16, PITCH * 64, K::LAYOUT);
            const uint64_t bd = make_sdesc(w_smem + tp * K::WTAP_BYTES, 16, K::SBO, K::LAYOUT);
#pragma unroll
            for (int kk = 0; kk < K::KSTEPS; ++kk) umma_bf16(d_tmem, ad + 2 * kk, bd + 2 * kk, idesc, (tp | kk) != 0);
          }
          umma_commit(tfull(t, m));
        }
        umma_commit(opfree(s));
      }
    }
  } else {
    // ===== epilogue: 16 warps; warp = (channel half h, tile m, TMEM lane quarter q); thread = one pixel x 16 channels =====
    const int ew = warp - 2, h = ew >> 3, m = (ew >> 2) & 1, q = warp & 3, et = threadIdx.x - 64;    // et: 0..511
    const int pi = q * 4 + (lane >> 3), pj = m * 8 + (lane & 7);
    const int P0y = pi * 16 + pj;                                                    // pixel index inside a staging tile
    const size_t pix_off = (size_t)P0y * C + h * 16;                                 // this thread's 16 channels inside an image
    const uint32_t my_tmem = tmem_base + ((uint32_t)(q * 32) << 16) + m * C + h * 16;   // + t * 2C
    const uint32_t st_off = P0y * 64;                                                // this pixel's 64-byte row of a staging tile
    const uint32_t st_c0 = ((h * 2 + 0) ^ ((P0y >> 1) & 3)) << 4, st_c1 = ((h * 2 + 1) ^ ((P0y >> 1) & 3)) << 4;
    const float* ss_w = ss_sm + h * 16;
    const float gamma_r = a.gamma[lane], beta_r = a.beta[lane];
    if (blockIdx.x == 0 && ew == 0) {            // the last row of the previous launch (see k_tc_fwd_persist)
      unsigned long long* last = a.acc + (size_t)(a.NB - 1) * NRD_ACC_ROW;
#pragma unroll
      for (int r = 0; r < NRD_ACC_REPL; ++r) { last[NRD_ACC_IDX(r, lane)] = 0ull; last[NRD_ACC_IDX(r, C + lane)] = 0ull; }
      __threadfence();
    }
    for (int k = 0; k < a.NB; ++k) {
      long long* tr = (a.trace && et == 0) ? a.trace + ((size_t)blockIdx.x * a.NB + k) * 8 : nullptr;
      long long w_mma = 0, w_st = 0;
      if (tr) tr[0] = clock64();
      // ---- phase 1: statistics of every image's accumulators, y_k staged for its bulk store ----
      float ssum = 0.f, qsum = 0.f;           // (summing the images per thread first and reducing once per block: 3 % faster at
      //                                         B = 2048, but no longer the resident kernel's order of additions: not taken)
      for (int j = 0; j < nimg; ++j) {
        const int g = k * nimg + j, s = g & 1, t = g % NT;
        const long long c0 = tr ? clock64() : 0;
        mbar_wait(tfull(t, m), (g / NT) & 1);
        if (tr) w_mma += clock64() - c0;
        tc_fence_after();
        float v[16];
        tmem_ld16(my_tmem + t * 2 * C, v);
        tc_fence_before();
        mbar_arrive(tfree(t));                                     // the accumulators are in registers
        uint4 ypk[2];
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
          __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&ypk[jj]);
#pragma unroll
          for (int e = 0; e < 4; ++e) ph[e] = __floats2bfloat162_rn(v[jj * 8 + 2 * e], v[jj * 8 + 2 * e + 1]);
        }
        const long long c1 = tr ? clock64() : 0;
        if (g >= 2) mbar_wait(yfree(s), ((g >> 1) - 1) & 1);       // the previous store has read this staging tile
        if (tr) w_st += clock64() - c1;
        uint8_t* yst = gen_base + (y_smem - base) + s * YST_BYTES + st_off;
        *reinterpret_cast<uint4*>(yst + st_c0) = ypk[0]; *reinterpret_cast<uint4*>(yst + st_c1) = ypk[1];
        fence_proxy_async();
        mbar_arrive(yready(s));
        {
          float sq[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) sq[i] = v[i] * v[i];
          qsum += warp_col_sum16(sq, lane);
        }
        ssum += warp_col_sum16(v, lane);
      }
      if (tr) tr[1] = clock64();
      red[ew * 32 + (lane & 1) * 16 + (lane >> 1)] = (lane & 1) ? qsum : ssum;
      asm volatile("bar.sync 1, 512;" ::: "memory");
      unsigned long long* acc = a.acc + (size_t)k * NRD_ACC_ROW;
      if (et < 2 * C) {
        const int which = et >> 5, ch = et & 31;
        float tsum = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) tsum += red[((ch >> 4) * 8 + w) * 32 + which * 16 + (ch & 15)];
        nrd_acc_check(nrd_acc_add(acc + NRD_ACC_IDX(blockIdx.x % NRD_ACC_REPL, et), tsum, 16777216.0, a.status), gridDim.x);
      }
      if (tr) tr[2] = clock64();
      // phase 2's first operands travel during the exchange: x_k is final since the previous block; y_k needs p1done
      const bf16* yk = a.ys + (size_t)k * a.act_elems + pix_off;
      const bf16* xk = a.xs + (size_t)k * a.act_elems + pix_off;
      uint4 yA[2], xA[2], yB[2], xB[2];
      auto fetch_x = [&](uint4 (&xv)[2], int j) {
        const uint4* px = reinterpret_cast<const uint4*>(xk + (size_t)img_of(j) * NRD_PIX * C);
        xv[0] = __ldcg(px); xv[1] = __ldcg(px + 1);
      };
      auto fetch_y = [&](uint4 (&yv)[2], int j) {
        const uint4* py = reinterpret_cast<const uint4*>(yk + (size_t)img_of(j) * NRD_PIX * C);
        yv[0] = __ldcg(py); yv[1] = __ldcg(py + 1);
      };
      fetch_x(xA, 0);
      if (nimg > 1) fetch_x(xB, 1);
      if (ew == 0) {
        unsigned long long ws, wq;
        nrd_acc_wait2(acc + NRD_ACC_IDX(0, lane), acc + NRD_ACC_IDX(0, C + lane), gridDim.x, ws, wq, a.poll_ns);
        if (blockIdx.x == 0 && k > 0) {
#pragma unroll
          for (int r = 0; r < NRD_ACC_REPL; ++r) { (acc - NRD_ACC_ROW)[NRD_ACC_IDX(r, lane)] = 0ull; (acc - NRD_ACC_ROW)[NRD_ACC_IDX(r, C + lane)] = 0ull; }
        }
        const double S = nrd_acc_value(ws, gridDim.x, 1.0 / 16777216.0), Q = nrd_acc_value(wq, gridDim.x, 1.0 / 16777216.0);
        const double mean = S * a.inv_count;
        double var = Q * a.inv_count - mean * mean;
        if (var < 0.0) var = 0.0;
        const float invstd = fast_invstd(var + (double)NRD_BN_EPS);
        const float sc = gamma_r * invstd, sh = beta_r - (float)mean * sc;
        ss_sm[lane] = sc; ss_sm[C + lane] = sh;
        if (blockIdx.x == gridDim.x - 1) {
          a.saved[(size_t)k * 2 * C + lane] = (float)mean; a.saved[(size_t)k * 2 * C + C + lane] = invstd;
          a.scale_shift[(size_t)k * 2 * C + lane] = sc; a.scale_shift[(size_t)k * 2 * C + C + lane] = sh;
          const double unbiased = var * a.unbias;
          a.running[lane] = (1.f - NRD_BN_MOMENTUM) * a.running[lane] + NRD_BN_MOMENTUM * (float)mean;
          a.running[C + lane] = (1.f - NRD_BN_MOMENTUM) * a.running[C + lane] + NRD_BN_MOMENTUM * (float)unbiased;
          if (lane == 0) *a.nbt += 1;
        }
      }
      if (tr) tr[3] = clock64();
      mbar_wait(p1done, (uint32_t)(k & 1));                        // this block's y_k stores are complete
      fetch_y(yA, 0);
      if (nimg > 1) fetch_y(yB, 1);
      asm volatile("bar.sync 1, 512;" ::: "memory");               // scale | shift of this block are in shared memory
      if (tr) tr[4] = clock64();
      // ---- phase 2: x_{k+1} = relu(bn(y_k)) + x_k from the stored tensors (this thread's own 32 B of each, two images in flight) ----
      auto emit = [&](uint4 (&yv)[2], uint4 (&xv)[2], int j) {
        const int g = k * nimg + j, s = g & 1;
        uint4 ok[2];
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
          const __nv_bfloat162* yh = reinterpret_cast<const __nv_bfloat162*>(&yv[jj]);
          const __nv_bfloat162* xh = reinterpret_cast<const __nv_bfloat162*>(&xv[jj]);
          __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&ok[jj]);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int c = jj * 8 + 2 * e;
            const float2 yr = __bfloat1622float2(yh[e]), xr = __bfloat1622float2(xh[e]);
            const float2 sc2 = *reinterpret_cast<const float2*>(ss_w + c), sh2 = *reinterpret_cast<const float2*>(ss_w + C + c);
            oh[e] = __floats2bfloat162_rn(fmaxf(fmaf(yr.x, sc2.x, sh2.x), 0.f) + xr.x, fmaxf(fmaf(yr.y, sc2.y, sh2.y), 0.f) + xr.y);
          }
        }
        if (j + 2 < nimg) { fetch_y(yv, j + 2); fetch_x(xv, j + 2); }     // two images ahead, while this one is staged
        const long long c1 = tr ? clock64() : 0;
        if (g >= 2) mbar_wait(xfree(s), ((g >> 1) - 1) & 1);
        if (tr) w_st += clock64() - c1;
        uint8_t* xst = gen_base + (xst_smem - base) + s * YST_BYTES + st_off;
        *reinterpret_cast<uint4*>(xst + st_c0) = ok[0]; *reinterpret_cast<uint4*>(xst + st_c1) = ok[1];
        fence_proxy_async();
        mbar_arrive(xready(s));
      };
      for (int j = 0; j < nimg; j += 2) {
        emit(yA, xA, j);
        if (j + 1 < nimg) emit(yB, xB, j + 1);
      }
      if (tr) { tr[5] = clock64(); tr[6] = w_mma; tr[7] = w_st; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
}

constexpr size_t stream_fwd_smem() { return 1024 + STREAM_NSLOT * 21 * 1024 + 9 * Cfg<32>::WTAP_BYTES + 4 * NRD_PIX * 32 * 2 + STREAM_BAR_BYTES + (1024 + 64) * sizeof(float) + 64; }

// ============================================================================================
// Persistent backward over ALL residual blocks (C = 32, at most 2 images per CTA, cooperative launch).
// dL/dx of an image never leaves the SM: it lives in the epilogue threads' registers (bf16 pairs, the same rounding the
// per-block kernels apply when they store it).  Per block k = NB-1 .. 0 and image:
//   phase A   y_k of the thread's pixel (cp.async, one block ahead); dz = g*[bn(y)>0]; per-channel sums of dz and dz*y
//             (warp transpose-reduce) -> CTA sum -> returning atomics on count-carrying words -> one warp polls, derives
//             the dy coefficients (no separate grid barrier)
//   phase B   dy = gamma*invstd*(dz - mean(dz) - xhat*mean(dz*xhat)) written (bf16, swizzled) into three kx-shifted,
//             row-haloed copies in shared memory (or ONE haloed copy: template parameter) = operand of BOTH gradients
//   MMA       dgrad: 36 tcgen05.mma per image (K-major dy x flipped weights) -> 2 accumulators per image
//             wgrad: 16 tcgen05.mma per image, both operands MN-major: A = 4 vertical taps of the row-haloed x_k tile (TMA)
//             stacked along M, B = the three kx-shifted dy rows stacked along N = 96; ONE accumulator set that persists over
//             all blocks and both images: one partial store per CTA per step, issued under the last block's dgrad
//   phase C   dx_k = dgrad + g -> registers (-> global only for k = 0), then straight into phase A of the next block
// dL/dx_NB itself is built from dL/d(pooled) through pool-2's arg-max while the registers are loaded.
// HBM traffic per block: y_k and x_k read once; nothing written.
// ============================================================================================
struct PersistBwdArgs {
  const float* gamma;
  const float* scale_shift;   // [NB][2C]
  const float* saved;         // [NB][2C] mean | invstd
  float* dgamma; float* dbeta;
  const bf16* ys;             // y_0 .. y_NB-1
  const bf16* g_pool;         // dL/d(pooled) [B,8,8,C], expanded through pool-2's argmax while loading (dL/dx_NB never exists)
  const uint8_t* idx2;        // pool-2 argmax [B,8,8,C] (position 0..3 inside the 2x2 window)
  bf16* dx0_out;              // dL/dx_0
  unsigned long long* acc;    // [NB][2C] 64-bit fixed-point accumulators, zeroed before the launch
  float* wpartial;            // [gridDim.x][9][C][C]
  unsigned long long* gbar;
  int B, NB;
  size_t act_elems;
  double inv_count;
  int dbg_skip;               // profiling experiment only (results garbage)
  long long* trace;           // optional clock64 stamps (profiling aid), or NULL
  int poll_ns;
  unsigned int* status;       // the plan's sticky status word (nrd_acc_add)
  // fused stem backward (NRD_OPT_FUSE_STEM_BWD, ONE = 0 only): conv1's weight / bias gradient of this CTA's images right here, from
  // dL/dx_0 in the epilogue registers and x_0 still in shared memory; NULL images = off (the stand-alone stem kernel follows)
  const float* images;        // [B,3,32,32] fp32
  const uint8_t* idx0;        // [B,16,16,C] stem pool arg-max
  float* stem_partial;        // [gridDim.x][28][C]
};

// ONE = 1: dy lives in ONE row- and column-haloed copy per image (see k_tc_fwd_persist): dgrad taps are pixel-granular
// descriptor starts, the weight gradient's three kx taps are N groups ONE pixel apart (LBO = 64 B)
template <int ONE>
__global__ void __launch_bounds__(576, 1)
k_tc_bwd_persist(const __grid_constant__ CUtensorMap tmXall, const __grid_constant__ CUtensorMap tmWd, PersistBwdArgs a) {
  constexpr int C = 32;
  using K = Cfg<C>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
  constexpr int PITCH = 18;
  constexpr int DY_BYTES = ONE ? 21 * 1024 : 3 * K::BUF_BYTES;    // one haloed copy (18*18*64 B rounded up) / three kx-copies of dy
  constexpr int SLOT_BYTES = DY_BYTES + K::BUF_BYTES;             // + row-haloed x_k tile
  const uint32_t a_smem = base;
  const uint32_t w_smem = a_smem + 2 * SLOT_BYTES;
  const uint32_t bar0 = w_smem + 9 * K::WTAP_BYTES;
  const uint32_t wfull = bar0;
  auto xready = [&](int s) { return bar0 + 8u * (1 + s); };
  auto dyready = [&](int s) { return bar0 + 8u * (3 + s); };
  auto wfree = [&](int s) { return bar0 + 8u * (5 + s); };
  auto dxfull = [&](int s, int m) { return bar0 + 8u * (7 + s * 2 + m); };
  const uint32_t wdone = bar0 + 8u * 11;
  const uint32_t stfull = bar0 + 8u * 12, stdone = bar0 + 8u * 14;      // fused stem backward: tiles of an image built / its MMAs done
  const uint32_t tmem_slot = bar0 + 8u * 13;
  constexpr int BAR_BYTES = 8 * 16;
  constexpr int MISC_OFF = 2 * SLOT_BYTES + 9 * K::WTAP_BYTES + BAR_BYTES;
  float* red = reinterpret_cast<float*>(gen_base + MISC_OFF);            // [16 warps][2][C] = 1024 floats
  float* coef = red + 1024;                                               // [3C] = gamma*invstd | mean(dz) | mean(dz*xhat)
  float* bnp2 = coef + 96;                                               // 2 x 64 floats (scale/shift pairs, double-buffered) + 128 floats (dy coefficients)
  double* dsum = reinterpret_cast<double*>(bnp2 + 256);                  // [32 row classes][2C] = 2048
  double* dsum2 = dsum + 2048;                                           // [4][2C]
  uint4* ybuf = reinterpret_cast<uint4*>(dsum2 + 256);                   // [4 chunks][512 threads]: each thread's y_k row (cp.async)                   // [1024]
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int img0 = blockIdx.x, img1 = blockIdx.x + gridDim.x;
  const int nslots = (img1 < a.B) ? 2 : 1;
  constexpr uint32_t TMEM_COLS = 256;                              // dgrad 2x2x32 columns + wgrad 3x32 columns
  constexpr uint32_t ACCW = 128;
  if (a.trace && threadIdx.x == 64) a.trace[(size_t)blockIdx.x * a.NB * 16 + 14] = clock64();     // kernel entry
  nrd_pdl_launch();                                                // the stem backward behind us may be scheduled (it waits for our exit)

  // the halo rows / columns of the dy copies are the convolution padding: zero everything once, never write them again
  for (int i = threadIdx.x; i < 2 * DY_BYTES / 16; i += blockDim.x) {
    const int s = i / (DY_BYTES / 16), o = i % (DY_BYTES / 16);
    reinterpret_cast<uint4*>(gen_base + s * SLOT_BYTES)[o] = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async();
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmXall); tma_prefetch_desc(&tmWd);
    mbar_init(wfull, 1); mbar_init(wdone, 1); mbar_init(stfull, 512); mbar_init(stdone, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(xready(s), 1); mbar_init(dyready(s), 512); mbar_init(wfree(s), 1);
      mbar_init(dxfull(s, 0), 1); mbar_init(dxfull(s, 1), 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer: flipped weights once; x_k tile per block and image slot =====
      mbar_expect_tx(wfull, 9 * K::WTAP_BYTES);
      for (int t = 0; t < 9; ++t) tma_load_2d(w_smem + t * K::WTAP_BYTES, &tmWd, wfull, t * C, 0);
      for (int it = 0; it < a.NB; ++it) {
        const int k = a.NB - 1 - it;
        for (int s = 0; s < nslots; ++s) {
          if (it > 0) mbar_wait(wfree(s), (uint32_t)((it - 1) & 1));      // the wgrad MMAs that read the old tile are done
          mbar_expect_tx(xready(s), K::BUF_BYTES);
          tma_load_5d(a_smem + s * SLOT_BYTES + DY_BYTES, &tmXall, xready(s), 0, 0, -1, s == 0 ? img0 : img1, k);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer =====
      constexpr uint32_t idesc_d = make_idesc(128, C, 0, 0);      // dgrad: both K-major
      constexpr uint32_t idesc_w = make_idesc(128, 3 * C, 1, 1);  // wgrad: both MN-major, N = three kx taps x C
      mbar_wait(wfull, 0);
      auto dgrad = [&](int it) {
        for (int s = 0; s < nslots; ++s) {
          mbar_wait(dyready(s), (uint32_t)(it & 1));
          tc_fence_after();
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            const uint32_t d_tmem = tmem_base + (s * 2 + m) * C;
#pragma unroll
            for (int t = 0; t < 9; ++t) {
              const uint32_t a_addr = ONE ? a_smem + s * SLOT_BYTES + ((t / 3) * PITCH + m * 8 + (t % 3)) * 64
                                          : a_smem + s * SLOT_BYTES + (t % 3) * K::BUF_BYTES + (m * 8 + t / 3) * K::IMGROW_BYTES;
              const uint64_t ad = make_sdesc(a_addr, 16, ONE ? PITCH * 64 : K::SBO, K::LAYOUT);
              const uint64_t bd = make_sdesc(w_smem + t * K::WTAP_BYTES, 16, K::SBO, K::LAYOUT);
#pragma unroll
              for (int kk = 0; kk < K::KSTEPS; ++kk) umma_bf16(d_tmem, ad + 2 * kk, bd + 2 * kk, idesc_d, (t | kk) != 0);
            }
            umma_commit(dxfull(s, m));
          }
        }
      };
      auto wgrad = [&](int it) {
        for (int s = 0; s < nslots; ++s) {
          mbar_wait(dyready(s), (uint32_t)(it & 1));
          mbar_wait(xready(s), (uint32_t)(it & 1));
          tc_fence_after();
          const uint32_t sb = a_smem + s * SLOT_BYTES;
          for (int yy = 0; yy < 16; ++yy) {              // K step = image row yy = 16 pixels; ONE MMA per row
            // A (MN-major): x_k rows yy .. yy+3 of the row-haloed tile stacked along M = (j, ci): vertical tap ky = j (j = 3 junk)
            // B (MN-major): row yy of the three kx-shifted dy copies stacked along N = (kxb, co): horizontal tap kx = 2 - kxb
            const uint64_t ad = make_sdesc(sb + DY_BYTES + yy * K::IMGROW_BYTES, K::IMGROW_BYTES, K::SBO, K::LAYOUT);
            const uint64_t bd = ONE ? make_sdesc(sb + (yy + 1) * PITCH * 64, 64, K::SBO, K::LAYOUT)          // N groups one pixel apart
                                    : make_sdesc(sb + (yy + 1) * K::IMGROW_BYTES, K::BUF_BYTES, K::SBO, K::LAYOUT);
            umma_bf16(tmem_base + ACCW, ad, bd, idesc_w, (it > 0 || s > 0 || yy > 0) ? 1u : 0u);
          }
          umma_commit(wfree(s));
        }
      };
      for (int it = 0; it + 1 < a.NB; ++it) { dgrad(it); wgrad(it); }
      // last block: weight gradient first, so the partial store runs under the last dgrad instead of after it
      wgrad(a.NB - 1);
      umma_commit(wdone);
      dgrad(a.NB - 1);
      if (!ONE && a.images) {
        // fused stem backward: D[tap (M rows 0..31; 32..127 junk M-groups)][co] += A_w[pixel][tap]^T * G_w[pixel][co], both MN-major,
        // per image 2 halves x 4 window positions x 8 k-steps of 16 pooled pixels (the order of k_stem_tc_bwd: bit-identical partials)
        constexpr uint32_t idesc_s = make_idesc(128, C, 1, 1);
        for (int s = 0; s < nslots; ++s) {
          mbar_wait(stfull, (uint32_t)(s & 1));
          tc_fence_after();
#pragma unroll 1
          for (int hw = 0; hw < 8; ++hw) {              // hw = half * 4 + window
            const uint32_t at = a_smem + ((hw >> 2) * 8 + (hw & 3)) * 8192, gt = at + 4 * 8192;
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
              umma_bf16(tmem_base + ACCW + 3 * C, make_sdesc(at + ks * 1024, 8192, 512, 4), make_sdesc(gt + ks * 1024, 8192, 512, 4), idesc_s,
                        (s | hw | ks) != 0);
          }
          umma_commit(stdone);
        }
      }
    }
  } else {
    // ===== epilogue: 16 warps; warp = (channel half h, tile m, TMEM lane quarter q); thread = one pixel x 16 channels, of =====
    // ===== image slot 0 and then of slot 1 (slot 0's dgrad MMAs run while slot 1's dy is still being written) =====
    const int ew = warp - 2, h = ew >> 3, m = (ew >> 2) & 1, q = warp & 3, et = threadIdx.x - 64;    // et: 0..511
    // pixel of this thread = row of its TMEM lane in tile m: 8 rows x 16 columns per tile, or (ONE) 16 rows x 8 columns
    const int pi = ONE ? q * 4 + (lane >> 3) : m * 8 + ((q * 32 + lane) >> 4), pj = ONE ? m * 8 + (lane & 7) : (lane & 15);
    const size_t pix_off = (size_t)(pi * 16 + pj) * C + h * 16;               // this thread's 16 channels inside an image
    const int P1 = ONE ? (pi + 1) * PITCH + pj + 1 : (pi + 1) * 16 + pj;       // its index inside the (centre) dy copy
    const uint32_t my_tmem = tmem_base + ((uint32_t)(q * 32) << 16) + m * C + h * 16;   // + s * 2C
    float dg_acc = 0.f, db_acc = 0.f;                 // CTA 0, warp ew == 0 (lane = channel): running dgamma / dbeta over the blocks
    // y_k of this thread's pixel: prefetched one block ahead with cp.async into this thread's own shared-memory slots,
    // one commit group per image slot
    const uint32_t ybuf_s = smem_u32(ybuf);
    auto load_y = [&](int k) {
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        if (s < nslots) {
          const bf16* yk = a.ys + (size_t)k * a.act_elems + (size_t)(s == 0 ? img0 : img1) * NRD_PIX * C + pix_off;
#pragma unroll
          for (int j = 0; j < 2; ++j)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ybuf_s + (uint32_t)(((s * 2 + j) * 512 + et) * 16)),
                         "l"(reinterpret_cast<const uint4*>(yk) + j) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
      }
    };
    // y of 8 channels (chunk j of slot s) of this thread's pixel: one conflict-free LDS.128 from the thread's own slots
    auto y_chunk = [&](int s, int j) -> uint4 { return ybuf[(s * 2 + j) * 512 + et]; };
    load_y(a.NB - 1);                                 // the longest latency of the prologue (DRAM) goes first
    uint32_t gpk[2][8];                               // dL/d(out) of this thread's pixel and channels, per image slot, bf16 pairs
    uint4 st_idx[2] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};   // fused stem backward: this thread's 16 arg-max bytes per slot
#pragma unroll
    for (int s = 0; s < 2; ++s) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (s < nslots) {
          {
            // max-pool backward on the fly: the pooled gradient goes to the window position the forward picked
            const size_t po = ((size_t)(s == 0 ? img0 : img1) * 64 + (pi >> 1) * 8 + (pj >> 1)) * C + h * 16 + j * 8;
            const uint4 gp4 = *reinterpret_cast<const uint4*>(a.g_pool + po);
            const uint2 id = *reinterpret_cast<const uint2*>(a.idx2 + po);
            const uint32_t sel = (uint32_t)((pi & 1) * 2 + (pj & 1));
            const uint32_t gw[4] = {gp4.x, gp4.y, gp4.z, gp4.w};
            uint32_t o[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const uint32_t ib = (e < 2 ? id.x : id.y) >> ((e & 1) * 16);
              const uint32_t lo = ((ib & 0xffu) == sel) ? 0x0000ffffu : 0u, hi = (((ib >> 8) & 0xffu) == sel) ? 0xffff0000u : 0u;
              o[e] = gw[e] & (lo | hi);
            }
            v = make_uint4(o[0], o[1], o[2], o[3]);
          }
        }
        gpk[s][j * 4 + 0] = v.x; gpk[s][j * 4 + 1] = v.y; gpk[s][j * 4 + 2] = v.z; gpk[s][j * 4 + 3] = v.w;
      }
    }
    auto write_ssp = [&](int it2) {                  // scale/shift pairs of block NB-1-it2 -> buffer it2 & 1 (warp ew == 0, lane = channel)
      const int k2 = a.NB - 1 - it2;
      float* d = bnp2 + (it2 & 1) * 64 + (lane >> 1) * 4 + (lane & 1) * 2;
      d[0] = a.scale_shift[(size_t)k2 * 2 * C + lane]; d[1] = a.scale_shift[(size_t)k2 * 2 * C + C + lane];
    };
    if (ew == 0) write_ssp(0);
    const float gamma_r = a.gamma[lane];
    // the LAST row of the previous launch could not be cleaned inside that launch (its readers were still polling): do it now,
    // long before anyone adds to it -- every CTA reaches that row only after seeing this CTA's contribution to row 0, which
    // is issued after the fence
    if (blockIdx.x == 0 && ew == 0) {
      unsigned long long* last = a.acc + (size_t)(a.NB - 1) * NRD_ACC_ROW;
#pragma unroll
      for (int r = 0; r < NRD_ACC_REPL; ++r) { last[NRD_ACC_IDX(r, lane)] = 0ull; last[NRD_ACC_IDX(r, C + lane)] = 0ull; }
      __threadfence();
    }
    asm volatile("bar.sync 1, 512;" ::: "memory");
    float4* cfp = reinterpret_cast<float4*>(bnp2 + 128);   // per channel pair: {A1, A2, A3, -, A1', A2', A3', -}, dy = dz*A1 + y*A2 + A3
    float s1 = 0.f, s2 = 0.f;                              // lanes 2c, 2c+1: sums of dz / dz*y over this warp's pixels, channel 16h + c
    // ---- phase A of block it for image slot s: per-channel sums of dz and dz*y over this warp's 32 pixels ----
    auto phase_a = [&](int s, int it) {
      const float4* ssp = reinterpret_cast<const float4*>(bnp2 + (it & 1) * 64);
      float t1[16], t2[16];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const uint4 yv = y_chunk(s, j);
        const uint32_t* yw = reinterpret_cast<const uint32_t*>(&yv);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int i = j * 8 + 2 * e;
          const float2 y2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&yw[e]));
          const float2 g2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gpk[s][j * 4 + e]));
          const float4 ss = ssp[h * 8 + j * 4 + e];
          const float d0 = fmaf(y2.x, ss.x, ss.y) > 0.f ? g2.x : 0.f, d1 = fmaf(y2.y, ss.z, ss.w) > 0.f ? g2.y : 0.f;
          t1[i] = d0; t1[i + 1] = d1; t2[i] = d0 * y2.x; t2[i + 1] = d1 * y2.y;
        }
      }
      s1 += warp_col_sum16(t1, lane);
      s2 += warp_col_sum16(t2, lane);
    };
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    phase_a(0, 0);
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    if (nslots > 1) phase_a(1, 0);
    for (int it = 0; it < a.NB; ++it) {
      const int k = a.NB - 1 - it;
      long long* tr = (a.trace && et == 0) ? a.trace + ((size_t)blockIdx.x * a.NB + it) * 16 : nullptr;
      if (tr) { const long long t = clock64(); tr[0] = t; tr[7] = t; tr[1] = t; for (int i = 8; i < 14; ++i) tr[i] = t; }
      // saved statistics of this block for the combine below: fetched now, far ahead of their use on the critical path
      const float mean_r = (ew == 0) ? a.saved[(size_t)k * 2 * C + lane] : 0.f, is_r = (ew == 0) ? a.saved[(size_t)k * 2 * C + C + lane] : 0.f;
      red[ew * 32 + (lane & 1) * 16 + (lane >> 1)] = (lane & 1) ? s2 : s1;
      s1 = 0.f; s2 = 0.f;
      asm volatile("bar.sync 1, 512;" ::: "memory");
      // cross-CTA sum: fixed point (2^-36) + contributor count in one word per statistic (nrd_common.cuh): order-independent,
      // bit-exact, and no separate grid barrier
      unsigned long long* acc = a.acc + (size_t)it * NRD_ACC_ROW;     // [replica][2C] words, NRD_ACC_WS apart
      if (et < 2 * C) {
        const int which = et >> 5, ch = et & 31;
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[((ch >> 4) * 8 + w) * 32 + which * 16 + (ch & 15)];
        nrd_acc_check(nrd_acc_add(acc + NRD_ACC_IDX(blockIdx.x % NRD_ACC_REPL, et), t, 68719476736.0, a.status), gridDim.x);   // wait for the return: measured faster
      }
      if (tr) tr[2] = clock64();
      // one warp (lane = channel) polls its channels' words, derives the dy coefficients and publishes them (and the NEXT
      // block's scale/shift pairs: every thread is past phase B of the previous block) with the one CTA barrier that follows
      if (ew == 0) {
        unsigned long long wb, wg;
        nrd_acc_wait2(acc + NRD_ACC_IDX(0, lane), acc + NRD_ACC_IDX(0, C + lane), gridDim.x, wb, wg, a.poll_ns);
        if (tr) tr[3] = clock64();
        if (blockIdx.x == 0 && it > 0) {       // the previous row is dead (see the forward kernel): clean it for the next launch
#pragma unroll
          for (int r = 0; r < NRD_ACC_REPL; ++r) { (acc - NRD_ACC_ROW)[NRD_ACC_IDX(r, lane)] = 0ull; (acc - NRD_ACC_ROW)[NRD_ACC_IDX(r, C + lane)] = 0ull; }
        }
        const double db = nrd_acc_value(wb, gridDim.x, 1.0 / 68719476736.0), dgy = nrd_acc_value(wg, gridDim.x, 1.0 / 68719476736.0);
        if (it + 1 < a.NB) write_ssp(it + 1);
        const double mean = (double)mean_r, is = (double)is_r;
        const double dg = is * (dgy - mean * db);       // sum dz*xhat = invstd * (sum dz*y - mean * sum dz)
        db_acc += (float)db; dg_acc += (float)dg;       // same order as the per-block path: k = NB-1 first
        const double a_ = (double)gamma_r * is, b_ = db * a.inv_count, c_ = dg * a.inv_count;
        float* d = reinterpret_cast<float*>(cfp) + (lane >> 1) * 8 + (lane & 1) * 4;
        d[0] = (float)a_; d[1] = (float)(-a_ * c_ * is); d[2] = (float)(a_ * c_ * is * mean - a_ * b_); d[3] = 0.f;
      }
      asm volatile("bar.sync 1, 512;" ::: "memory");
      if (tr) tr[4] = clock64();
      // ---- phase B: dy -> the three operand copies (once the previous block's wgrad MMAs no longer read them) ----
      {
        const float4* ssp = reinterpret_cast<const float4*>(bnp2 + (it & 1) * 64);
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          if (s >= nslots) break;
          if (it > 0) mbar_wait(wfree(s), (uint32_t)((it - 1) & 1));
          uint8_t* slot_gen = gen_base + s * SLOT_BYTES;
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int ch = h * 2 + j;
            uint4 ok;
            __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&ok);
            const uint4 yv = y_chunk(s, j);
            const uint32_t* yw = reinterpret_cast<const uint32_t*>(&yv);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int pe = ch * 4 + e;
              const float2 y2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&yw[e]));
              const float2 g2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gpk[s][j * 4 + e]));
              const float4 ss = ssp[pe], c0 = cfp[pe * 2], c1 = cfp[pe * 2 + 1];
              const float d0 = fmaf(y2.x, ss.x, ss.y) > 0.f ? g2.x : 0.f, d1 = fmaf(y2.y, ss.z, ss.w) > 0.f ? g2.y : 0.f;
              oh[e] = __floats2bfloat162_rn(fmaf(d0, c0.x, fmaf(y2.x, c0.y, c0.z)), fmaf(d1, c1.x, fmaf(y2.y, c1.y, c1.z)));
            }
            *reinterpret_cast<uint4*>(slot_gen + (ONE ? 0 : K::BUF_BYTES) + P1 * 64 + ((ch ^ ((P1 >> 1) & 3)) << 4)) = ok;
            if (!ONE && pj < 15) { const int P0 = P1 + 1; *reinterpret_cast<uint4*>(slot_gen + P0 * 64 + ((ch ^ ((P0 >> 1) & 3)) << 4)) = ok; }
            if (!ONE && pj > 0) { const int P2 = P1 - 1; *reinterpret_cast<uint4*>(slot_gen + 2 * K::BUF_BYTES + P2 * 64 + ((ch ^ ((P2 >> 1) & 3)) << 4)) = ok; }
          }
          fence_proxy_async();
          mbar_arrive(dyready(s));
        }
      }
      if (tr) tr[5] = clock64();
      if (k > 0 && !a.dbg_skip) load_y(k - 1);          // y_k is dead: fetch the next block's rows while the dgrad MMAs run
      if (!ONE && k == 0 && a.images) {
        // fused stem backward: the y prefetch buffers are dead from here on -- this CTA's (zero-padded, fp32) images go there while
        // the last dgrad runs, and this thread's arg-max bytes into registers
        asm volatile("bar.sync 1, 512;" ::: "memory");                           // every thread is past its last read of ybuf
        float* im_gen = reinterpret_cast<float*>(ybuf);
        const uint32_t im_s = smem_u32(im_gen);
        for (int c = et; c < nslots * 768; c += 512) {                           // 768 16-byte chunks per image
          const int s_ = c / 768, qq = c % 768, ci = qq >> 8, r = (qq >> 3) & 31, c4 = qq & 7;
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(im_s + (uint32_t)((s_ * 4080 + (ci * 34 + r + 1) * 40 + 4 + c4 * 4) * 4)),
                       "l"(a.images + (size_t)(s_ == 0 ? img0 : img1) * 3072 + ci * 1024 + r * 32 + c4 * 4) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        for (int i = et; i < 2016; i += 512) {                                    // halo of both slots: rows 0 / 33, columns 0..3 / 36..39
          int off;
          if (i < 480) { const int s_ = i / 240, rem = i % 240, ci = rem / 80, rr = (rem % 80) / 40, c = rem % 40; off = s_ * 4080 + (ci * 34 + rr * 33) * 40 + c; }
          else { const int j = i - 480, s_ = j / 768, rem = j % 768, ci = rem / 256, r = 1 + (rem % 256) / 8, cc = rem % 8; off = s_ * 4080 + (ci * 34 + r) * 40 + (cc < 4 ? cc : 32 + cc); }
          im_gen[off] = 0.f;
        }
#pragma unroll
        for (int s = 0; s < 2; ++s)
          if (s < nslots) st_idx[s] = *reinterpret_cast<const uint4*>(a.idx0 + (size_t)(s == 0 ? img0 : img1) * NRD_PIX * C + pix_off);
      }
      // ---- weight-gradient partial of this CTA: accumulator lanes (j, ci), columns (kxb, co) ----
      if (it == a.NB - 1 && h == 0 && m == 0) {
        mbar_wait(wdone, 0);
        tc_fence_after();
        float* outp = a.wpartial + (size_t)blockIdx.x * 9 * C * C;
        const int j = q, ci = lane;
  #pragma unroll
        for (int kxb = 0; kxb < 3; ++kxb) {
          float v[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + ACCW + kxb * C, v);
          if (j < 3) {
            float* dst = outp + (size_t)(j * 3 + (2 - kxb)) * C * C + ci;                     // [tap = (ky, kx)][co][ci]
  #pragma unroll
            for (int co = 0; co < 32; ++co) dst[co * C] = v[co];
          }
        }
      }
      // ---- phase C: dx_k = dgrad + g, then straight on to phase A of the next block for the same image ----
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        if (s >= nslots) break;
        mbar_wait(dxfull(s, m), (uint32_t)(it & 1));
        tc_fence_after();
        float v[16];
        tmem_ld16(my_tmem + s * 2 * C, v);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float2 g2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gpk[s][e]));
          const __nv_bfloat162 o = __floats2bfloat162_rn(v[2 * e] + g2.x, v[2 * e + 1] + g2.y);
          gpk[s][e] = *reinterpret_cast<const uint32_t*>(&o);
        }
        tc_fence_before();      // the accumulator is in registers: the next block's dgrad may overwrite it (ordered by dyready)
        if (k == 0) {
          uint4* dst = reinterpret_cast<uint4*>(a.dx0_out + (size_t)(s == 0 ? img0 : img1) * NRD_PIX * C + pix_off);
#pragma unroll
          for (int j = 0; j < 2; ++j) dst[j] = make_uint4(gpk[s][j * 4], gpk[s][j * 4 + 1], gpk[s][j * 4 + 2], gpk[s][j * 4 + 3]);
        } else if (!a.dbg_skip) {
          if (s == 0) asm volatile("cp.async.wait_group 1;" ::: "memory");
          else asm volatile("cp.async.wait_group 0;" ::: "memory");
          phase_a(s, it + 1);
        }
      }
      if (tr) tr[6] = clock64();
    }
    if (blockIdx.x == 0 && ew == 0) { a.dgamma[lane] = dg_acc; a.dbeta[lane] = db_acc; }
    if (!ONE && a.images) {
      // ===== fused stem backward: conv1's weight / bias gradient of this CTA's images (model/resnet.py:16 under main.py:38) =====
      // dL/dx_0 is in gpk (this thread's pixel x 16 channels), x_0 still in the x tile of the last block (k = 0), the images
      // in the y buffers.  Per image: 4 im2col tiles A_w [128 pooled pixels][27 taps + 1] and 4 masked-gradient tiles G_w
      // [128 pooled pixels][C] per half image, as in k_stem_tc_bwd; the tiles take over the (dead) dy / x slots.
      uint32_t x0r[2][8];
#pragma unroll
      for (int s = 0; s < 2; ++s)
        if (s < nslots) {
          const uint8_t* xt = gen_base + s * SLOT_BYTES + DY_BYTES;
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int ch = h * 2 + j;
            const uint4 v = *reinterpret_cast<const uint4*>(xt + P1 * 64 + ((ch ^ ((P1 >> 1) & 3)) << 4));
            x0r[s][j * 4] = v.x; x0r[s][j * 4 + 1] = v.y; x0r[s][j * 4 + 2] = v.z; x0r[s][j * 4 + 3] = v.w;
          }
        }
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      asm volatile("bar.sync 1, 512;" ::: "memory");     // x_0 read, images landed, every dgrad MMA consumed (each thread waited for its tile)
      const float* im_gen = reinterpret_cast<const float*>(ybuf);
      const int hh = pi >> 3, pl = (pi & 7) * 16 + pj, sw = (pl >> 1) & 3;        // half image, row inside its tiles, its swizzle
#pragma unroll
      for (int s = 0; s < 2; ++s) {                      // (unrolled: gpk / x0r / st_idx stay in registers)
        if (s >= nslots) break;
        if (s > 0) mbar_wait(stdone, 0u);                // the previous image's MMAs have read the tiles
        const float* im = im_gen + s * 4080;
        // A rows of window positions w = 2h, 2h + 1 (the other channel-half thread of this pixel builds the other two)
        {
          float patch[3][4][4];
#pragma unroll
          for (int ci = 0; ci < 3; ++ci)
#pragma unroll
            for (int rr = 0; rr < 4; ++rr)
#pragma unroll
              for (int cc = 0; cc < 4; ++cc) patch[ci][rr][cc] = im[(ci * 34 + 2 * pi + rr) * 40 + 3 + 2 * pj + cc];
          auto rows = [&](auto HC) {                     // windows 2H, 2H + 1: static indices into the patch (h is warp-uniform)
            constexpr int H = decltype(HC)::value;
#pragma unroll
          for (int wi = 0; wi < 2; ++wi) {
            constexpr int dy_ = H;
            const int w = H * 2 + wi, dx_ = wi;
            float kv[32];
#pragma unroll
            for (int ci = 0; ci < 3; ++ci)
#pragma unroll
              for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) kv[ci * 9 + ky * 3 + kx] = patch[ci][dy_ + ky][dx_ + kx];
            kv[27] = 1.0f; kv[28] = kv[29] = kv[30] = kv[31] = 0.f;             // tap 27 = 1: accumulator row 27 is the bias gradient
            uint8_t* row = gen_base + (hh * 8 + w) * 8192 + pl * 64;
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
              uint32_t pk[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const __nv_bfloat162 hv = __floats2bfloat162_rn(kv[ch * 8 + 2 * e], kv[ch * 8 + 2 * e + 1]);
                pk[e] = *reinterpret_cast<const uint32_t*>(&hv);
              }
              *reinterpret_cast<uint4*>(row + ((ch ^ sw) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            }
          }
          };
          if (h == 0) rows(std::integral_constant<int, 0>{}); else rows(std::integral_constant<int, 1>{});
        }
        // G rows: this thread's 16 channels of its pixel, for each window position: dL/dx_0 where that position was the arg-max and x_0 > 0
        {
          const uint8_t* ib = reinterpret_cast<const uint8_t*>(&st_idx[s]);
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            uint8_t* row = gen_base + (hh * 8 + 4 + w) * 8192 + pl * 64;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              uint32_t o4[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int c = j * 8 + 2 * e;
                const uint32_t g2 = gpk[s][j * 4 + e], x2 = x0r[s][j * 4 + e];
                const bool k0 = ib[c] == w && (x2 & 0x7FFFu) != 0, k1 = ib[c + 1] == w && (x2 & 0x7FFF0000u) != 0;
                o4[e] = (k0 ? (g2 & 0xFFFFu) : 0u) | (k1 ? (g2 & 0xFFFF0000u) : 0u);
              }
              *reinterpret_cast<uint4*>(row + (((h * 2 + j) ^ sw) << 4)) = make_uint4(o4[0], o4[1], o4[2], o4[3]);
            }
          }
        }
        fence_proxy_async();
        mbar_arrive(stfull);
      }
      if (ew == 2) {                                     // warp 4: the TMEM lane quarter 0 (warp % 4 == 0) = lanes 0..31 = taps (27 + bias row)
        mbar_wait(stdone, (uint32_t)((nslots - 1) & 1));
        tc_fence_after();
        float v[32];
        tmem_ld32(tmem_base + ACCW + 3 * C, v);
        if (lane < 28) {
          float4* dst = reinterpret_cast<float4*>(a.stem_partial + ((size_t)blockIdx.x * 28 + lane) * C);
#pragma unroll
          for (int c4 = 0; c4 < 8; ++c4) dst[c4] = make_float4(v[c4 * 4], v[c4 * 4 + 1], v[c4 * 4 + 2], v[c4 * 4 + 3]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
  if (a.trace && threadIdx.x == 64) a.trace[((size_t)blockIdx.x * a.NB + a.NB - 1) * 16 + 15] = clock64();   // kernel exit
}

// ============================================================================================
// Streaming persistent backward: ALL residual blocks, ANY batch size (C = 32, cooperative launch) -- k_tc_bwd_persist for batches
// beyond two images per CTA.  dL/dx lives in ONE global buffer G [B,16,16,C] (bf16, the rounding k_tc_bwd_persist applies to its
// registers), updated in place: a thread only ever reads and writes its own pixel's 16 channels of it, and an image's new value is
// written after the dgrad MMAs that needed the old one.  A CTA owns images j = 0 .. n-1 (blockIdx.x + j * grid).  Per block
// k = NB-1 .. 0:
//   exchange   per-channel sums of dz and dz*y over the whole batch (count-carrying fixed-point words) -> dy coefficients
//   B(j)       dy of image j from G and y_k (both prefetched one image ahead: G into registers, y_k by cp.async into a 3-deep ring
//              of per-thread shared-memory slots) into the haloed dy copy of operand slot gi % 3
//   MMA        dgrad (36 tcgen05.mma, TMEM slot gi & 1) and wgrad (16 tcgen05.mma against the x_k tile the DMA thread loaded into
//              the same operand slot; ONE accumulator set over all blocks and images, one partial store per CTA)
//   C(j)       dx = dgrad + g -> G, and phase A of block k-1 for the same image (y_{k-1} prefetched into the ring slot y_k left)
// in the order B(0) B(1) C(0) B(2) C(1) ..., so image j's MMAs run under B(j+1).  Only block NB-1 has a stand-alone phase A: it
// also builds dL/dx_NB from dL/d(pooled) through pool-2's arg-max and stores it to G.
// L2 / HBM traffic per image and block: G read + written, y_k read twice (phase A one block earlier, then B), x_k read once.
// ============================================================================================
struct StreamBwdArgs {
  const float* gamma;
  const float* scale_shift;   // [NB][2C]
  const float* saved;         // [NB][2C] mean | invstd
  float* dgamma; float* dbeta;
  const bf16* ys;             // y_0 .. y_NB-1
  const bf16* g_pool;         // dL/d(pooled) [B,8,8,C]
  const uint8_t* idx2;        // pool-2 argmax [B,8,8,C]
  bf16* G;                    // dL/dx of the current block, in place; dL/dx_0 at exit
  unsigned long long* acc;    // [NB] rows of count-carrying words
  float* wpartial;            // [gridDim.x][9][C][C]
  int B, NB;
  size_t act_elems;
  double inv_count;
  int poll_ns;
  unsigned int* status;
  long long* trace;           // optional clock64 stamps [CTA][block][16] (profiling aid), or NULL
};

constexpr int STREAM_BWD_NS = 3;          // operand slots (dy copy + x_k tile)
constexpr int STREAM_BWD_NY = 3;          // y ring depth (per-thread slots)

__global__ void __launch_bounds__(576, 1)
k_tc_bwd_stream(const __grid_constant__ CUtensorMap tmXall, const __grid_constant__ CUtensorMap tmWd, StreamBwdArgs a) {
  constexpr int C = 32;
  using K = Cfg<C>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
  constexpr int PITCH = 18, NS = STREAM_BWD_NS, NY = STREAM_BWD_NY;
  constexpr int DY_BYTES = 21 * 1024;                             // one haloed copy (18*18*64 B rounded up)
  constexpr int SLOT_BYTES = DY_BYTES + K::BUF_BYTES;             // + row-haloed x_k tile
  const uint32_t a_smem = base;
  const uint32_t w_smem = a_smem + NS * SLOT_BYTES;
  const uint32_t bar0 = w_smem + 9 * K::WTAP_BYTES;
  const uint32_t wfull = bar0, wdone = bar0 + 8u;
  auto xready = [&](int s) { return bar0 + 8u * (2 + s); };       // x_k tile landed (TMA bytes)
  auto dyready = [&](int s) { return bar0 + 8u * (6 + s); };      // dy written (512 arrivals)
  auto wfree = [&](int s) { return bar0 + 8u * (10 + s); };       // the MMAs that read operand slot s are done
  auto dxfull = [&](int t, int m) { return bar0 + 8u * (14 + t * 2 + m); };
  const uint32_t tmem_slot = bar0 + 8u * 18;
  constexpr int BAR_BYTES = 8 * 20;
  constexpr int MISC_OFF = NS * SLOT_BYTES + 9 * K::WTAP_BYTES + BAR_BYTES;
  float* red = reinterpret_cast<float*>(gen_base + MISC_OFF);            // [16 warps][2][16] = 512 floats (1024 reserved)
  float* bnp2 = red + 1024;                                              // 2 x 64 floats (scale/shift pairs, double-buffered) + 128 floats (dy coefficients)
  uint4* ybuf = reinterpret_cast<uint4*>(bnp2 + 256);                    // [NY][2 chunks][512 threads]
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_slot - base));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nimg = (a.B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;      // images of this CTA (>= 1)
  auto img_of = [&](int j) { return (int)blockIdx.x + j * (int)gridDim.x; };
  const int total = a.NB * nimg;                                   // gi = it * nimg + j: operand slot gi % NS, TMEM slot gi & 1, y ring gi % NY
  constexpr uint32_t TMEM_COLS = 256;                              // dgrad 2 x 2 x 32 columns + wgrad 3 x 32 columns
  constexpr uint32_t ACCW = 128;
  nrd_pdl_launch();

  // the halo rows / columns of the dy copies are the convolution padding: zero everything once, never write them again
  for (int i = threadIdx.x; i < NS * DY_BYTES / 16; i += blockDim.x) {
    const int s = i / (DY_BYTES / 16), o = i % (DY_BYTES / 16);
    reinterpret_cast<uint4*>(gen_base + s * SLOT_BYTES)[o] = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async();
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmXall); tma_prefetch_desc(&tmWd);
    mbar_init(wfull, 1); mbar_init(wdone, 1);
    for (int s = 0; s < NS; ++s) { mbar_init(xready(s), 1); mbar_init(dyready(s), 512); mbar_init(wfree(s), 1); }
    for (int t = 0; t < 2; ++t) { mbar_init(dxfull(t, 0), 1); mbar_init(dxfull(t, 1), 1); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer: flipped weights once; the x_k tile of every (block, image), NS ahead of the MMAs =====
      mbar_expect_tx(wfull, 9 * K::WTAP_BYTES);
      for (int t = 0; t < 9; ++t) tma_load_2d(w_smem + t * K::WTAP_BYTES, &tmWd, wfull, t * C, 0);
      int it = 0, j = 0;
      for (int gi = 0; gi < total; ++gi) {
        const int s = gi % NS, u = gi / NS;
        if (u > 0) mbar_wait(wfree(s), (uint32_t)((u - 1) & 1));  // the wgrad MMAs that read the old tile are done
        mbar_expect_tx(xready(s), K::BUF_BYTES);
        tma_load_5d(a_smem + s * SLOT_BYTES + DY_BYTES, &tmXall, xready(s), 0, 0, -1, img_of(j), a.NB - 1 - it);
        if (++j == nimg) { j = 0; ++it; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer =====
      constexpr uint32_t idesc_d = make_idesc(128, C, 0, 0);      // dgrad: both K-major
      constexpr uint32_t idesc_w = make_idesc(128, 3 * C, 1, 1);  // wgrad: both MN-major, N = three kx taps x C
      mbar_wait(wfull, 0);
      for (int gi = 0; gi < total; ++gi) {
        const int s = gi % NS, t = gi & 1;
        const uint32_t par = (uint32_t)((gi / NS) & 1);
        const uint32_t sb = a_smem + s * SLOT_BYTES;
        mbar_wait(dyready(s), par);          // (also: every thread has read TMEM slot t's previous accumulators, see C below)
        tc_fence_after();
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          const uint32_t d_tmem = tmem_base + (t * 2 + m) * C;
#pragma unroll
          for (int tp = 0; tp < 9; ++tp) {
            const uint64_t ad = make_sdesc(sb + ((tp / 3) * PITCH + m * 8 + (tp % 3)) * 64, 16, PITCH * 64, K::LAYOUT);
            const uint64_t bd = make_sdesc(w_smem + tp * K::WTAP_BYTES, 16, K::SBO, K::LAYOUT);
#pragma unroll
            for (int kk = 0; kk < K::KSTEPS; ++kk) umma_bf16(d_tmem, ad + 2 * kk, bd + 2 * kk, idesc_d, (tp | kk) != 0);
          }
          umma_commit(dxfull(t, m));
        }
        mbar_wait(xready(s), par);
        tc_fence_after();
        for (int yy = 0; yy < 16; ++yy) {              // K step = image row yy = 16 pixels; ONE MMA per row (see k_tc_bwd_persist<1>)
          const uint64_t ad = make_sdesc(sb + DY_BYTES + yy * K::IMGROW_BYTES, K::IMGROW_BYTES, K::SBO, K::LAYOUT);
          const uint64_t bd = make_sdesc(sb + (yy + 1) * PITCH * 64, 64, K::SBO, K::LAYOUT);          // N groups one pixel apart
          umma_bf16(tmem_base + ACCW, ad, bd, idesc_w, (gi > 0 || yy > 0) ? 1u : 0u);
        }
        umma_commit(wfree(s));
      }
      umma_commit(wdone);
    }
  } else {
    // ===== epilogue: 16 warps; warp = (channel half h, tile m, TMEM lane quarter q); thread = one pixel x 16 channels =====
    const int ew = warp - 2, h = ew >> 3, m = (ew >> 2) & 1, q = warp & 3, et = threadIdx.x - 64;    // et: 0..511
    const int pi = q * 4 + (lane >> 3), pj = m * 8 + (lane & 7);
    const size_t pix_off = (size_t)(pi * 16 + pj) * C + h * 16;               // this thread's 16 channels inside an image
    const int P1 = (pi + 1) * PITCH + pj + 1;                                 // its index inside the haloed dy copy
    const uint32_t dy_c0 = P1 * 64 + (((h * 2 + 0) ^ ((P1 >> 1) & 3)) << 4), dy_c1 = P1 * 64 + (((h * 2 + 1) ^ ((P1 >> 1) & 3)) << 4);
    const uint32_t my_tmem = tmem_base + ((uint32_t)(q * 32) << 16) + m * C + h * 16;   // + t * 2C
    float dg_acc = 0.f, db_acc = 0.f;                 // CTA 0, warp ew == 0 (lane = channel): running dgamma / dbeta over the blocks
    const uint32_t ybuf_s = smem_u32(ybuf);
    // y of block k, image j -> ring slot r of this thread (2 x 16 B), one cp.async group
    auto load_y = [&](int k, int j, int r) {
      const bf16* yk = a.ys + (size_t)k * a.act_elems + (size_t)img_of(j) * NRD_PIX * C + pix_off;
#pragma unroll
      for (int c = 0; c < 2; ++c)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ybuf_s + (uint32_t)(((r * 2 + c) * 512 + et) * 16)),
                     "l"(reinterpret_cast<const uint4*>(yk) + c) : "memory");
    };
    auto commit = [&]() { asm volatile("cp.async.commit_group;" ::: "memory"); };
    auto y_chunk = [&](int r, int c) -> uint4 { return ybuf[(r * 2 + c) * 512 + et]; };
    auto g_ptr = [&](int j) { return reinterpret_cast<uint4*>(a.G + (size_t)img_of(j) * NRD_PIX * C + pix_off); };
    auto write_ssp = [&](int it2) {                  // scale/shift pairs of block NB-1-it2 -> buffer it2 & 1 (warp ew == 0, lane = channel)
      const int k2 = a.NB - 1 - it2;
      float* d = bnp2 + (it2 & 1) * 64 + (lane >> 1) * 4 + (lane & 1) * 2;
      d[0] = a.scale_shift[(size_t)k2 * 2 * C + lane]; d[1] = a.scale_shift[(size_t)k2 * 2 * C + C + lane];
    };
    if (ew == 0) write_ssp(0);
    const float gamma_r = a.gamma[lane];
    if (blockIdx.x == 0 && ew == 0) {                // the last row of the previous launch (see k_tc_bwd_persist)
      unsigned long long* last = a.acc + (size_t)(a.NB - 1) * NRD_ACC_ROW;
#pragma unroll
      for (int r = 0; r < NRD_ACC_REPL; ++r) { last[NRD_ACC_IDX(r, lane)] = 0ull; last[NRD_ACC_IDX(r, C + lane)] = 0ull; }
      __threadfence();
    }
    asm volatile("bar.sync 1, 512;" ::: "memory");
    float4* cfp = reinterpret_cast<float4*>(bnp2 + 128);   // per channel pair: {A1, A2, A3, -, A1', A2', A3', -}, dy = dz*A1 + y*A2 + A3
    float s1 = 0.f, s2 = 0.f;                              // lanes 2c, 2c+1: sums of dz / dz*y over this warp's pixels, channel 16h + c
    // ---- phase A of block it for one image: per-channel sums of dz and dz*y over this warp's 32 pixels ----
    auto phase_a = [&](const uint32_t (&g)[8], int r, int it) {
      const float4* ssp = reinterpret_cast<const float4*>(bnp2 + (it & 1) * 64);
      float t1[16], t2[16];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const uint4 yv = y_chunk(r, c);
        const uint32_t* yw = reinterpret_cast<const uint32_t*>(&yv);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int i = c * 8 + 2 * e;
          const float2 y2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&yw[e]));
          const float2 g2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&g[c * 4 + e]));
          const float4 ss = ssp[h * 8 + c * 4 + e];
          const float d0 = fmaf(y2.x, ss.x, ss.y) > 0.f ? g2.x : 0.f, d1 = fmaf(y2.y, ss.z, ss.w) > 0.f ? g2.y : 0.f;
          t1[i] = d0; t1[i + 1] = d1; t2[i] = d0 * y2.x; t2[i + 1] = d1 * y2.y;
        }
      }
      s1 += warp_col_sum16(t1, lane);
      s2 += warp_col_sum16(t2, lane);
    };
    // ---- block NB-1: dL/dx_NB from dL/d(pooled) through pool-2's arg-max -> G, and its phase A ----
    {
      load_y(a.NB - 1, 0, 0); commit();
      for (int j = 0; j < nimg; ++j) {
        if (j + 1 < nimg) load_y(a.NB - 1, j + 1, (j + 1) & 1);
        commit();
        uint32_t g[8];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const size_t po = ((size_t)img_of(j) * 64 + (pi >> 1) * 8 + (pj >> 1)) * C + h * 16 + c * 8;
          const uint4 gp4 = *reinterpret_cast<const uint4*>(a.g_pool + po);
          const uint2 id = *reinterpret_cast<const uint2*>(a.idx2 + po);
          const uint32_t sel = (uint32_t)((pi & 1) * 2 + (pj & 1));
          const uint32_t gw[4] = {gp4.x, gp4.y, gp4.z, gp4.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint32_t ib = (e < 2 ? id.x : id.y) >> ((e & 1) * 16);
            const uint32_t lo = ((ib & 0xffu) == sel) ? 0x0000ffffu : 0u, hi = (((ib >> 8) & 0xffu) == sel) ? 0xffff0000u : 0u;
            g[c * 4 + e] = gw[e] & (lo | hi);
          }
        }
        uint4* dst = g_ptr(j);
        dst[0] = make_uint4(g[0], g[1], g[2], g[3]); dst[1] = make_uint4(g[4], g[5], g[6], g[7]);
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        phase_a(g, j & 1, 0);
      }
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    uint32_t gA[8], gB[8], pg[8];                    // dL/dx of the even / odd image in flight; the next image's (prefetched)
    // The prefetch loads write pg DIRECTLY (inline asm outputs): written as uint4 temporaries copied into pg, the compiler put
    // the copies right behind the loads and every thread waited out the L2 latency there (ncu: 14 % of all stall samples).
    // (Prefetching two images ahead -- a second register set, a 4-deep y ring -- measured 4 % SLOWER: the kernel is bound by
    // shared-memory bandwidth and issue slots, not by this latency.)
    auto prefetch_g = [&](int j) {
      const uint4* gp = g_ptr(j);
      asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(pg[0]), "=r"(pg[1]), "=r"(pg[2]), "=r"(pg[3]) : "l"(gp));
      asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(pg[4]), "=r"(pg[5]), "=r"(pg[6]), "=r"(pg[7]) : "l"(gp + 1));
    };
    for (int it = 0; it < a.NB; ++it) {
      const int k = a.NB - 1 - it;
      long long* tr = (a.trace && et == 0) ? a.trace + ((size_t)blockIdx.x * a.NB + it) * 16 : nullptr;
      long long w_ld = 0, w_slot = 0, w_mma = 0, t_b = 0, t_c = 0;
      if (tr) tr[0] = clock64();
      const float mean_r = (ew == 0) ? a.saved[(size_t)k * 2 * C + lane] : 0.f, is_r = (ew == 0) ? a.saved[(size_t)k * 2 * C + C + lane] : 0.f;
      red[ew * 32 + (lane & 1) * 16 + (lane >> 1)] = (lane & 1) ? s2 : s1;
      s1 = 0.f; s2 = 0.f;
      asm volatile("bar.sync 1, 512;" ::: "memory");
      unsigned long long* acc = a.acc + (size_t)it * NRD_ACC_ROW;
      if (et < 2 * C) {
        const int which = et >> 5, ch = et & 31;
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[((ch >> 4) * 8 + w) * 32 + which * 16 + (ch & 15)];
        nrd_acc_check(nrd_acc_add(acc + NRD_ACC_IDX(blockIdx.x % NRD_ACC_REPL, et), t, 68719476736.0, a.status), gridDim.x);
      }
      if (tr) tr[1] = clock64();
      {   // the first image's operands travel during the exchange
        prefetch_g(0);
        load_y(k, 0, (it * nimg) % NY); commit();
      }
      if (ew == 0) {
        unsigned long long wb, wg;
        nrd_acc_wait2(acc + NRD_ACC_IDX(0, lane), acc + NRD_ACC_IDX(0, C + lane), gridDim.x, wb, wg, a.poll_ns);
        if (blockIdx.x == 0 && it > 0) {
#pragma unroll
          for (int r = 0; r < NRD_ACC_REPL; ++r) { (acc - NRD_ACC_ROW)[NRD_ACC_IDX(r, lane)] = 0ull; (acc - NRD_ACC_ROW)[NRD_ACC_IDX(r, C + lane)] = 0ull; }
        }
        const double db = nrd_acc_value(wb, gridDim.x, 1.0 / 68719476736.0), dgy = nrd_acc_value(wg, gridDim.x, 1.0 / 68719476736.0);
        if (it + 1 < a.NB) write_ssp(it + 1);
        const double mean = (double)mean_r, is = (double)is_r;
        const double dg = is * (dgy - mean * db);       // sum dz*xhat = invstd * (sum dz*y - mean * sum dz)
        db_acc += (float)db; dg_acc += (float)dg;
        const double a_ = (double)gamma_r * is, b_ = db * a.inv_count, c_ = dg * a.inv_count;
        float* d = reinterpret_cast<float*>(cfp) + (lane >> 1) * 8 + (lane & 1) * 4;
        d[0] = (float)a_; d[1] = (float)(-a_ * c_ * is); d[2] = (float)(a_ * c_ * is * mean - a_ * b_); d[3] = 0.f;
      }
      asm volatile("bar.sync 1, 512;" ::: "memory");
      if (tr) tr[2] = clock64();
      const float4* ssp = reinterpret_cast<const float4*>(bnp2 + (it & 1) * 64);
      // ---- B(i): dy of image i -> operand slot; then y_{k-1}[i] into the ring slot y_k[i] leaves, and image i+1's operands ----
      auto phase_b = [&](uint32_t (&g)[8], int i) {
        const int gi = it * nimg + i, s = gi % NS, u = gi / NS, r = gi % NY;
        const long long c0 = tr ? clock64() : 0;
        asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
        for (int e = 0; e < 8; ++e) g[e] = pg[e];
        const long long c1 = tr ? clock64() : 0;
        if (u > 0) mbar_wait(wfree(s), (uint32_t)((u - 1) & 1));
        if (tr) { const long long c2 = clock64(); w_ld += c1 - c0; w_slot += c2 - c1; }
        uint8_t* slot_gen = gen_base + s * SLOT_BYTES;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const int ch = h * 2 + c;
          uint4 ok;
          __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&ok);
          const uint4 yv = y_chunk(r, c);
          const uint32_t* yw = reinterpret_cast<const uint32_t*>(&yv);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int pe = ch * 4 + e;
            const float2 y2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&yw[e]));
            const float2 g2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&g[c * 4 + e]));
            const float4 ss = ssp[pe], c0 = cfp[pe * 2], c1 = cfp[pe * 2 + 1];
            const float d0 = fmaf(y2.x, ss.x, ss.y) > 0.f ? g2.x : 0.f, d1 = fmaf(y2.y, ss.z, ss.w) > 0.f ? g2.y : 0.f;
            oh[e] = __floats2bfloat162_rn(fmaf(d0, c0.x, fmaf(y2.x, c0.y, c0.z)), fmaf(d1, c1.x, fmaf(y2.y, c1.y, c1.z)));
          }
          *reinterpret_cast<uint4*>(slot_gen + (c == 0 ? dy_c0 : dy_c1)) = ok;
        }
        fence_proxy_async();
        mbar_arrive(dyready(s));
        if (k > 0) load_y(k - 1, i, r);
        commit();
        if (i + 1 < nimg) {
          prefetch_g(i + 1);
          load_y(k, i + 1, (gi + 1) % NY);
        }
        commit();
        if (tr) t_b += clock64() - c0;
      };
      // ---- C(i): dx_k = dgrad + g -> G, then phase A of block k-1 for the same image ----
      auto phase_c = [&](uint32_t (&g)[8], int i) {
        const int gi = it * nimg + i, t = gi & 1, r = gi % NY;
        const long long c0 = tr ? clock64() : 0;
        mbar_wait(dxfull(t, m), (uint32_t)((gi >> 1) & 1));
        if (tr) w_mma += clock64() - c0;
        tc_fence_after();
        float v[16];
        tmem_ld16(my_tmem + t * 2 * C, v);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float2 g2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&g[e]));
          const __nv_bfloat162 o = __floats2bfloat162_rn(v[2 * e] + g2.x, v[2 * e + 1] + g2.y);
          g[e] = *reinterpret_cast<const uint32_t*>(&o);
        }
        tc_fence_before();      // the accumulator is in registers: image i+2's dgrad may overwrite it (ordered by its dyready)
        uint4* dst = g_ptr(i);
        dst[0] = make_uint4(g[0], g[1], g[2], g[3]); dst[1] = make_uint4(g[4], g[5], g[6], g[7]);
        if (k > 0) {
          // y_{k-1}[i] was committed in B(i); behind it at most the two groups of B(i+1)
          if (i + 1 < nimg) asm volatile("cp.async.wait_group 2;" ::: "memory");
          else asm volatile("cp.async.wait_group 0;" ::: "memory");
          phase_a(g, r, it + 1);
        }
        if (tr) t_c += clock64() - c0;
      };
      for (int i0 = 0; i0 <= nimg; i0 += 2) {
        if (i0 < nimg) phase_b(gA, i0);
        if (i0 >= 1) phase_c(gB, i0 - 1);
        if (i0 + 1 < nimg) phase_b(gB, i0 + 1);
        if (i0 + 1 <= nimg) phase_c(gA, i0);
      }
      if (tr) { tr[3] = clock64(); tr[4] = w_ld; tr[5] = w_slot; tr[6] = w_mma; tr[7] = t_b; tr[8] = t_c; }
    }
    if (blockIdx.x == 0 && ew == 0) { a.dgamma[lane] = dg_acc; a.dbeta[lane] = db_acc; }
    // ---- weight-gradient partial of this CTA: accumulator lanes (j, ci), columns (kxb, co) ----
    if (h == 0 && m == 0) {
      mbar_wait(wdone, 0);
      tc_fence_after();
      float* outp = a.wpartial + (size_t)blockIdx.x * 9 * C * C;
      const int j = q, ci = lane;
#pragma unroll
      for (int kxb = 0; kxb < 3; ++kxb) {
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + ACCW + kxb * C, v);
        if (j < 3) {
          float* dst = outp + (size_t)(j * 3 + (2 - kxb)) * C * C + ci;                     // [tap = (ky, kx)][co][ci]
#pragma unroll
          for (int co = 0; co < 32; ++co) dst[co * C] = v[co];
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
}

constexpr size_t stream_bwd_smem() {
  return 1024 + STREAM_BWD_NS * (21 * 1024 + Cfg<32>::BUF_BYTES) + 9 * Cfg<32>::WTAP_BYTES + 8 * 20 + (1024 + 256) * sizeof(float) +
         (size_t)STREAM_BWD_NY * 2 * 512 * 16 + 64;
}

// ============================================================================================
// wgrad: both operands MN-major (channels contiguous, pixels = K).  Per work item: the 3 kx-shifted row-haloed copies of x
// and the dy rows.  One MMA stacks 128/C vertical taps along M: tap ky+1 is one image row further = LBO.
// ============================================================================================
template <int C> struct WCfg {
  static constexpr int TAPS_PER_MMA = 128 / C;                         // 4 (C=32) or 2 (C=64)
  static constexpr int GPK = (3 + TAPS_PER_MMA - 1) / TAPS_PER_MMA;    // accumulator groups per kx: 1 or 2
  static constexpr int GROUPS = 3 * GPK;                               // 3 or 6 accumulators of C columns
  static constexpr uint32_t TMEM_COLS = (GROUPS * C <= 128) ? 128 : 512;
  static constexpr int DY_BYTES = Cfg<C>::ROWS * Cfg<C>::IMGROW_BYTES;
  static constexpr int STAGE_BYTES = 3 * Cfg<C>::BUF_BYTES + DY_BYTES;
  static constexpr int NSS = 2;
  static constexpr int SLACK = 3 * Cfg<C>::IMGROW_BYTES;               // junk taps (ky >= 3) read up to 2 rows past the last stage
};

template <int C>
__global__ void __launch_bounds__(192, 1)
k_tc_wgrad(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmDY, float* __restrict__ partial,
           int nitems, int first) {
  using K = Cfg<C>; using W = WCfg<C>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen_base = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bar0 = base + W::NSS * W::STAGE_BYTES + W::SLACK;
  auto full = [&](int s) { return bar0 + 8u * s; };
  auto empty = [&](int s) { return bar0 + 8u * (W::NSS + s); };
  const uint32_t tfull = bar0 + 8u * (2 * W::NSS);
  const uint32_t tmem_slot = bar0 + 8u * (2 * W::NSS + 1);
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen_base + (tmem_slot - base));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

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
      for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
        const int b = item / K::ITEMS_PER_IMG, r0 = (item % K::ITEMS_PER_IMG) * K::ROWS;
        mbar_wait(empty(ss), phase ^ 1);
        mbar_expect_tx(full(ss), W::STAGE_BYTES);
        const uint32_t sbase = base + ss * W::STAGE_BYTES;
        for (int kx = 0; kx < 3; ++kx) tma_load_4d(sbase + kx * K::BUF_BYTES, &tmX, full(ss), 0, kx - 1, r0 - 1, b);
        tma_load_4d(sbase + 3 * K::BUF_BYTES, &tmDY, full(ss), 0, 0, r0, b);
        if (++ss == W::NSS) { ss = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(128, C, 1, 1);
      int ss = 0; uint32_t phase = 0; bool any = false;
      for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
        mbar_wait(full(ss), phase);
        tc_fence_after();
        const uint32_t sbase = base + ss * W::STAGE_BYTES;
        for (int y = 0; y < K::ROWS; ++y) {            // K step = one image row = 16 pixels
          const uint64_t bd = make_sdesc(sbase + 3 * K::BUF_BYTES + y * K::IMGROW_BYTES, K::IMGROW_BYTES, K::SBO, K::LAYOUT);
#pragma unroll
          for (int kx = 0; kx < 3; ++kx)
#pragma unroll
            for (int j = 0; j < W::GPK; ++j) {
              // rows of this accumulator: (tap ky = j*TAPS_PER_MMA + m/C, ci = m%C); tap ky reads image row y+ky of the haloed copy
              const uint64_t ad = make_sdesc(sbase + kx * K::BUF_BYTES + (y + j * W::TAPS_PER_MMA) * K::IMGROW_BYTES, K::IMGROW_BYTES, K::SBO, K::LAYOUT);
              umma_bf16(tmem_base + (kx * W::GPK + j) * C, ad, bd, idesc, (any || y > 0) ? 1u : 0u);
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
    for (int kx = 0; kx < 3; ++kx)
#pragma unroll
      for (int j = 0; j < W::GPK; ++j) {
        const int ky = j * W::TAPS_PER_MMA + m / C, ci = m % C, tap = ky * 3 + kx;
#pragma unroll
        for (int h = 0; h < C / 32; ++h) {
          float v[32];
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (kx * W::GPK + j) * C + h * 32, v);
          if (ky < 3) {
            float* dst = outp + ((size_t)tap * C + h * 32) * C + ci;          // [tap][co][ci], co stride = C floats
            if (!first) {     // all 32 loads first: a load/store chain through possibly-aliasing pointers would serialise
              float old[32];
#pragma unroll
              for (int jj = 0; jj < 32; ++jj) old[jj] = __ldcg(dst + (size_t)jj * C);
#pragma unroll
              for (int jj = 0; jj < 32; ++jj) v[jj] += old[jj];
            }
#pragma unroll
            for (int jj = 0; jj < 32; ++jj) dst[(size_t)jj * C] = v[jj];
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

// activation map: dims {C, W=16, H=16, N=B}; box {C, 16, box_rows, 1}
static int make_act_map(CUtensorMap* tm, void* ptr, int B, int C, int box_rows) {
  PFN_encodeTiled enc = get_encode();
  NRD_REQUIRE(enc, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[4] = {(cuuint64_t)C, 16, 16, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)16 * C * 2, (cuuint64_t)256 * C * 2};
  cuuint32_t box[4] = {(cuuint32_t)C, 16, (cuuint32_t)box_rows, 1};
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

template <int C, int MODE> constexpr size_t conv_smem() {
  constexpr int NACC = (MODE == 2) ? (512 / C) : 2;
  return 1024 + Cfg<C>::NST * 3 * Cfg<C>::BUF_BYTES + 9 * Cfg<C>::WTAP_BYTES + 8 * (2 * Cfg<C>::NST + 2 + 2 * NACC) +
         (10 * C) * sizeof(float) + 512 * sizeof(double) + 64;
}
template <int C> constexpr size_t wgrad_smem() {
  return 1024 + WCfg<C>::NSS * WCfg<C>::STAGE_BYTES + WCfg<C>::SLACK + 8 * (2 * WCfg<C>::NSS + 2) + 64;
}

constexpr size_t persist_smem(int one) {
  return 1024 + 2 * (one ? 21 * 1024 : 3 * Cfg<32>::BUF_BYTES) + 9 * Cfg<32>::WTAP_BYTES + (one ? 4 : 2) * NRD_PIX * 32 * 2 + 144 + (1024 + 64) * sizeof(float) + 64 +
         1024 + 2048 + 128 + 2 * 3 * 34 * 40 * sizeof(float);     // + fused stem: alignment, W1 tile, bias, two padded images
}

constexpr size_t persist_bwd_smem(int one) {
  return 1024 + 2 * ((one ? 21 * 1024 : 3 * Cfg<32>::BUF_BYTES) + Cfg<32>::BUF_BYTES) + 9 * Cfg<32>::WTAP_BYTES + 128 + (1024 + 96 + 256) * sizeof(float) + (2048 + 256) * sizeof(double) + 4 * 512 * 16 + 64;
}

static int items(const nrd_plan* p) { return p->B * (p->C == 32 ? 1 : 2); }
static int conv_grid(const nrd_plan* p) { int t = items(p), sms = NRD_NUM_SMS - p->reserve_sms; return t < sms ? t : sms; }

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
  const int rows = (p->C == 32 ? 16 : 8);
  for (int k = 0; k <= p->NB; ++k) NRD_TRY(make_act_map(&p->tm_x[k], p->xs + act * k, p->B, p->C, rows + 2));
  NRD_TRY(make_act_map(&p->tm_dy, p->dy, p->B, p->C, rows + 2));      // dgrad: haloed
  NRD_TRY(make_act_map(&p->tm_dy0, p->dy, p->B, p->C, rows));         // wgrad: the rows themselves
  NRD_TRY(make_w_map(&p->tm_wf, p->wtc_fwd, p->C));
  NRD_TRY(make_w_map(&p->tm_wd, p->wtc_dgrad, p->C));
  if (p->C == 32) {
    NRD_TRY(set_attr(k_tc_conv<32, 0>, conv_smem<32, 0>())); NRD_TRY(set_attr(k_tc_conv<32, 1>, conv_smem<32, 1>()));
    NRD_TRY(set_attr(k_tc_conv<32, 2>, conv_smem<32, 2>())); NRD_TRY(set_attr(k_tc_wgrad<32>, wgrad_smem<32>()));
  } else {
    NRD_TRY(set_attr(k_tc_conv<64, 0>, conv_smem<64, 0>())); NRD_TRY(set_attr(k_tc_conv<64, 1>, conv_smem<64, 1>()));
    NRD_TRY(set_attr(k_tc_conv<64, 2>, conv_smem<64, 2>())); NRD_TRY(set_attr(k_tc_wgrad<64>, wgrad_smem<64>()));
  }
  // fused block forward: every CTA keeps all its accumulators in TMEM (512 columns) and all CTAs must be co-resident
  nrd_tc_refresh_limits(p);
  if (p->C == 32) {
    NRD_TRY(set_attr(k_tc_fwd_persist<0>, persist_smem(0)));
    NRD_TRY(set_attr(k_tc_fwd_persist<1>, persist_smem(1)));
    NRD_TRY(set_attr(k_tc_fwd_stream, stream_fwd_smem()));
    NRD_TRY(set_attr(k_tc_bwd_stream, stream_bwd_smem()));
    NRD_TRY(set_attr(k_tc_bwd_persist<0>, persist_bwd_smem(0)));
    NRD_TRY(set_attr(k_tc_bwd_persist<1>, persist_bwd_smem(1)));
    // all saved x_k as ONE 5-D tensor {C, W, H, N, k}: the block index is just the last TMA coordinate
    PFN_encodeTiled enc = get_encode();
    cuuint64_t dims[5] = {32, 16, 16, (cuuint64_t)p->B, (cuuint64_t)(p->NB + 1)};
    cuuint64_t strides[4] = {64, 1024, 16384, (cuuint64_t)p->act_elems * 2};
    cuuint32_t box[5] = {32, 16, 18, 1, 1};
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = enc(&p->tm_xall, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, p->xs, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { nrd_set_error("cuTensorMapEncodeTiled(x stack) failed: %d", (int)r); return NRD_ECUDA; }
    // store maps of the persistent forward: one whole image {C, 16, 16} of x_k / y_k per bulk tensor store
    cuuint32_t sbox[5] = {32, 16, 16, 1, 1};
    r = enc(&p->tm_xst, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, p->xs, dims, strides, sbox, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { nrd_set_error("cuTensorMapEncodeTiled(x store) failed: %d", (int)r); return NRD_ECUDA; }
    {   // streaming forward: any x_k WITH its halo
      cuuint32_t b5[5] = {32, 18, 18, 1, 1};
      r = enc(&p->tm_x18all, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, p->xs, dims, strides, b5, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) { nrd_set_error("cuTensorMapEncodeTiled(x stack, haloed) failed: %d", (int)r); return NRD_ECUDA; }
    }
    {   // one-copy layout: x_0 WITH its halo, 18 x 18 pixels (zero-filled by the load)
      cuuint64_t d4[4] = {32, 16, 16, (cuuint64_t)p->B};
      cuuint64_t s4[3] = {64, 1024, 16384};
      cuuint32_t b4[4] = {32, 18, 18, 1}, e4[4] = {1, 1, 1, 1};
      r = enc(&p->tm_x18, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, p->xs, d4, s4, b4, e4, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) { nrd_set_error("cuTensorMapEncodeTiled(x_0, haloed) failed: %d", (int)r); return NRD_ECUDA; }
    }
    cuuint64_t ydims[5] = {32, 16, 16, (cuuint64_t)p->B, (cuuint64_t)p->NB};
    r = enc(&p->tm_yst, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, p->ys, ydims, strides, sbox, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { nrd_set_error("cuTensorMapEncodeTiled(y store) failed: %d", (int)r); return NRD_ECUDA; }
  }
  p->tc_ready = 1;
  return NRD_OK;
}

// which cooperative kernels fit, given the grid (148 SMs minus the ones left free for a concurrent collective)
void nrd_tc_refresh_limits(nrd_plan* p) {
  const int rows = p->C == 32 ? 16 : 8;
  const int grid = conv_grid(p), per_cta = (items(p) + grid - 1) / grid * (rows / 8);
  p->tc_fused_fwd = (per_cta * p->C <= 512) ? 1 : 0;
  const bool fits2 = p->B <= 2 * grid;
  // NB = 1: the row cleaned at kernel start would be the live row 0
  p->tc_persist_fwd = (p->C == 32 && p->NB >= 2 && (fits2 || p->stream_opt != 2)) ? 1 : 0;
  p->tc_stream = (p->tc_persist_fwd && (!fits2 || p->stream_opt == 1)) ? 1 : 0;
}

int nrd_tc_stats_blocks(const nrd_plan* p) { return conv_grid(p); }
int nrd_tc_wgrad_blocks(const nrd_plan* p) { return conv_grid(p); }

int nrd_tc_conv_fwd(nrd_plan* p, int k, cudaStream_t st) {
  NRD_REQUIRE(p->tc_ready, "tcgen05 path not initialised");
  const size_t act = p->act_elems * p->esz;
  bf16* y = (bf16*)(p->ys + act * k);
  const int grid = conv_grid(p), nt = items(p);
  BlockFwdArgs fa{};
  if (p->C == 32) k_tc_conv<32, 0><<<grid, 192, conv_smem<32, 0>(), st>>>(p->tm_x[k], p->tm_wf, y, nullptr, p->stats_partial, nt, fa);
  else k_tc_conv<64, 0><<<grid, 192, conv_smem<64, 0>(), st>>>(p->tm_x[k], p->tm_wf, y, nullptr, p->stats_partial, nt, fa);
  nrd_count_launch();
  NRD_CHECK_CUDA(cudaGetLastError());
  return NRD_OK;
}

// whole ResBlock forward k: y_k, x_{k+1}, saved stats, running-stat EMA in one cooperative launch
int nrd_tc_block_fwd(nrd_plan* p, int k, cudaStream_t st) {
  NRD_REQUIRE(p->tc_ready && p->tc_fused_fwd, "fused block forward not available for this plan");
  const size_t act = p->act_elems * p->esz;
  const int C = p->C;
  bf16* y = (bf16*)(p->ys + act * k);
  const bf16* res = nullptr;
  int nt = items(p);
  BlockFwdArgs fa;
  fa.gamma = p->params + p->off[3]; fa.beta = p->params + p->off[4];
  fa.saved = p->saved_stats + (size_t)k * 2 * C; fa.scale_shift = p->scale_shift + (size_t)k * 2 * C;
  fa.running = p->bn_running; fa.nbt = (long long*)p->nbt;
  fa.x = (const bf16*)(p->xs + act * k); fa.xnext = (bf16*)(p->xs + act * (k + 1));
  fa.gbar = p->gbar + NRD_GBAR_BLOCK_FWD; fa.count = p->B * NRD_PIX;
  float* sp = p->stats_partial;
  void* args[] = {(void*)&p->tm_x[k], (void*)&p->tm_wf, (void*)&y, (void*)&res, (void*)&sp, (void*)&nt, (void*)&fa};
  const dim3 grid(conv_grid(p)), block(192);
  if (C == 32) NRD_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)k_tc_conv<32, 2>, grid, block, args, conv_smem<32, 2>(), st));
  else NRD_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)k_tc_conv<64, 2>, grid, block, args, conv_smem<64, 2>(), st));
  nrd_count_launch();
  return NRD_OK;
}

// all NB residual blocks in ONE cooperative launch (C = 32, <= 2 images per CTA)
int nrd_tc_forward_blocks(nrd_plan* p, int pool, const float* stem_images, int training, cudaStream_t st) {
  NRD_REQUIRE(p->tc_ready && p->tc_persist_fwd, "persistent block-stack forward not available for this plan");
  if (p->tc_stream) {
    NRD_REQUIRE(!pool && !stem_images, "the streaming forward neither pools nor runs the stem");
    StreamFwdArgs sa;
    sa.gamma = p->params + p->off[3]; sa.beta = p->params + p->off[4];
    sa.saved = p->saved_stats; sa.scale_shift = p->scale_shift; sa.running = p->bn_running; sa.nbt = (long long*)p->nbt;
    sa.xs = (const bf16*)p->xs; sa.ys = (const bf16*)p->ys; sa.acc = p->bn_acc; sa.B = p->B; sa.NB = p->NB; sa.act_elems = p->act_elems;
    sa.inv_count = 1.0 / ((double)p->B * NRD_PIX); sa.unbias = ((double)p->B * NRD_PIX) / ((double)p->B * NRD_PIX - 1.0);
    sa.poll_ns = p->poll_ns; sa.status = reinterpret_cast<unsigned int*>(p->gbar + NRD_GBAR_STATUS); sa.trace = (long long*)p->trace;
    void* sargs[] = {(void*)&p->tm_x18all, (void*)&p->tm_wf, (void*)&p->tm_xst, (void*)&p->tm_yst, (void*)&sa};
    NRD_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)k_tc_fwd_stream, dim3(conv_grid(p)), dim3(576), sargs, stream_fwd_smem(), st));
    nrd_count_launch();
    return NRD_OK;
  }
  PersistFwdArgs a;
  a.gamma = p->params + p->off[3]; a.beta = p->params + p->off[4];
  a.saved = p->saved_stats; a.scale_shift = p->scale_shift; a.running = p->bn_running; a.nbt = (long long*)p->nbt;
  a.xs = (bf16*)p->xs; a.ys = (bf16*)p->ys; a.partial = p->stats_partial; a.acc = p->bn_acc; a.gbar = p->gbar + NRD_GBAR_PERSIST_FWD;
  a.B = p->B; a.NB = p->NB; a.act_elems = p->act_elems; a.trace = (long long*)p->trace; a.dbg_skip_stores = p->dbg_skip_stores;
  a.pooled = pool ? (bf16*)p->pooled : nullptr; a.idx2 = p->idx2; a.poll_ns = p->poll_ns;
  a.status = reinterpret_cast<unsigned int*>(p->gbar + NRD_GBAR_STATUS);
  a.images = stem_images; a.conv1_wp = p->conv1_wp; a.conv1_b = p->params + p->off[1]; a.idx0 = p->idx0; a.bs = nrd_bn_sync_of(p, training);
  a.inv_count = 1.0 / ((double)p->B * NRD_PIX); a.unbias = ((double)p->B * NRD_PIX) / ((double)p->B * NRD_PIX - 1.0);
  void* args[] = {(void*)((p->one_copy & 1) ? &p->tm_x18 : &p->tm_x[0]), (void*)&p->tm_wf, (void*)&p->tm_xst, (void*)&p->tm_yst, (void*)&a};
  if (p->one_copy & 1) NRD_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)k_tc_fwd_persist<1>, dim3(conv_grid(p)), dim3(576), args, persist_smem(1), st));
  else NRD_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)k_tc_fwd_persist<0>, dim3(conv_grid(p)), dim3(576), args, persist_smem(0), st));
  nrd_count_launch();
  return NRD_OK;
}

// all NB residual blocks of the backward in ONE cooperative launch: g_in = dL/dx_NB -> dx0_out = dL/dx_0, dgamma, dbeta,
// per-CTA weight-gradient partials (reduced by k_wgrad_reduce)
int nrd_tc_backward_blocks(nrd_plan* p, const void* g_pooled, void* dx0_out, const float* stem_images, cudaStream_t st) {
  NRD_REQUIRE(p->tc_ready && nrd_persist_bwd_ok(p), "persistent block-stack backward not available for this plan");
  if (p->tc_stream) {
    NRD_REQUIRE(!stem_images, "the streaming backward does not run the stem");
    StreamBwdArgs sa;
    sa.gamma = p->params + p->off[3]; sa.scale_shift = p->scale_shift; sa.saved = p->saved_stats;
    sa.dgamma = p->grads + p->off[3]; sa.dbeta = p->grads + p->off[4];
    sa.ys = (const bf16*)p->ys; sa.g_pool = (const bf16*)g_pooled; sa.idx2 = p->idx2; sa.G = (bf16*)dx0_out;
    sa.acc = p->bn_acc + (size_t)32 * NRD_ACC_ROW; sa.wpartial = p->wgrad_partial;
    sa.B = p->B; sa.NB = p->NB; sa.act_elems = p->act_elems; sa.inv_count = 1.0 / ((double)p->B * NRD_PIX); sa.poll_ns = p->poll_ns;
    sa.status = reinterpret_cast<unsigned int*>(p->gbar + NRD_GBAR_STATUS); sa.trace = (long long*)p->trace_bwd;
    void* sargs[] = {(void*)&p->tm_xall, (void*)&p->tm_wd, (void*)&sa};
    NRD_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)k_tc_bwd_stream, dim3(conv_grid(p)), dim3(576), sargs, stream_bwd_smem(), st));
    nrd_count_launch();
    return NRD_OK;
  }
  PersistBwdArgs a;
  a.gamma = p->params + p->off[3]; a.scale_shift = p->scale_shift; a.saved = p->saved_stats;
  a.dgamma = p->grads + p->off[3]; a.dbeta = p->grads + p->off[4];
  a.ys = (const bf16*)p->ys; a.g_pool = (const bf16*)g_pooled; a.idx2 = p->idx2; a.poll_ns = p->poll_ns; a.dx0_out = (bf16*)dx0_out;
  a.acc = p->bn_acc + (size_t)32 * NRD_ACC_ROW; a.wpartial = p->wgrad_partial; a.gbar = p->gbar + NRD_GBAR_PERSIST_BWD;
  a.B = p->B; a.NB = p->NB; a.act_elems = p->act_elems; a.trace = (long long*)p->trace_bwd; a.dbg_skip = p->dbg_skip_stores; a.inv_count = 1.0 / ((double)p->B * NRD_PIX);
  a.status = reinterpret_cast<unsigned int*>(p->gbar + NRD_GBAR_STATUS);
  const bool stem_here = stem_images != nullptr && !(p->one_copy & 2);
  a.images = stem_here ? stem_images : nullptr; a.idx0 = p->idx0; a.stem_partial = p->stem_partial;
  void* args[] = {(void*)&p->tm_xall, (void*)&p->tm_wd, (void*)&a};
  if (p->one_copy & 2) NRD_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)k_tc_bwd_persist<1>, dim3(conv_grid(p)), dim3(576), args, persist_bwd_smem(1), st));
  else NRD_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)k_tc_bwd_persist<0>, dim3(conv_grid(p)), dim3(576), args, persist_bwd_smem(0), st));
  nrd_count_launch();
  return NRD_OK;
}

int nrd_tc_conv_dgrad(nrd_plan* p, const void* g, void* dx_out, cudaStream_t st) {
  NRD_REQUIRE(p->tc_ready, "tcgen05 path not initialised");
  const int grid = conv_grid(p), nt = items(p);
  BlockFwdArgs fa{};
  if (p->C == 32) k_tc_conv<32, 1><<<grid, 192, conv_smem<32, 1>(), st>>>(p->tm_dy, p->tm_wd, (bf16*)dx_out, (const bf16*)g, nullptr, nt, fa);
  else k_tc_conv<64, 1><<<grid, 192, conv_smem<64, 1>(), st>>>(p->tm_dy, p->tm_wd, (bf16*)dx_out, (const bf16*)g, nullptr, nt, fa);
  nrd_count_launch();
  NRD_CHECK_CUDA(cudaGetLastError());
  return NRD_OK;
}

int nrd_tc_conv_wgrad(nrd_plan* p, int k, int first, cudaStream_t st) {
  NRD_REQUIRE(p->tc_ready, "tcgen05 path not initialised");
  const int grid = conv_grid(p), nt = items(p);
  if (p->C == 32) k_tc_wgrad<32><<<grid, 192, wgrad_smem<32>(), st>>>(p->tm_x[k], p->tm_dy0, p->wgrad_partial, nt, first);
  else k_tc_wgrad<64><<<grid, 192, wgrad_smem<64>(), st>>>(p->tm_x[k], p->tm_dy0, p->wgrad_partial, nt, first);
  nrd_count_launch();
  NRD_CHECK_CUDA(cudaGetLastError());
  return NRD_OK;
}
