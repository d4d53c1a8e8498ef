// Stem on the tensor cores (C = 32, bf16 plans): x0 = maxpool2(relu(conv1(img) + b))  (model/resnet.py:16) and its weight
// gradient, as implicit GEMMs whose im2col operand is built in shared memory by the CTA itself.
//
// The 3-channel 3x3 stem has K = 27 (padded to 32): far too thin for a TMA-fed pipeline, but 226 M MACs per batch of 256 are
// too many for the CUDA cores next to a ~70 us step.  One work item = half an image = 128 POOLED pixels.  For each of the
// four positions (dy,dx) of the 2x2 pooling window the builder warps write one im2col tile A_w [128 pooled pixels][32 taps]
// (bf16, SWIZZLE_64B K-major rows of 64 B) from the zero-padded fp32 image staged in shared memory:
//   forward   D_w[pixel][co] = A_w * W1^T  (4 x 2 tcgen05.mma, M=128, N=32, K=16): the epilogue thread of a pooled pixel reads
//             the four accumulators of ITS OWN TMEM lane and takes the running max -> no cross-lane traffic for the pool,
//             first-max arg index for backward, + bias, ReLU, bf16 store.  The [B,32,32,32] conv output never exists.
//   backward  dW1[tap][co] = sum_w sum_pixels A_w[pixel][tap] * G_w[pixel][co] with G_w = g0 masked to the pixels whose arg-max is
//             window position w (and x0 > 0): the same A tiles as MN-major operand, G_w tiles built beside them, 4 x 8
//             tcgen05.mma per item into ONE accumulator that lives in TMEM for the whole kernel; tap 27 of every A row is 1.0,
//             so row 27 of the accumulator is the bias gradient.
#include <stdlib.h>
#include "tc_common.cuh"

namespace {

constexpr int SC = 32;                        // stem output channels handled here
constexpr int IMG_ROW = 40;                   // padded row: data column c at index 4 + c (16-B aligned for cp.async)
constexpr int IMG_FLOATS = 3 * 34 * IMG_ROW;  // zero-padded image, rows -1..32
constexpr int TILE_BYTES = 128 * 64;          // one im2col / gradient tile: 128 rows x 32 bf16

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// cp.async one fp32 image [3][32][32] into its padded shared-memory slot (128 threads)
__device__ __forceinline__ void load_image_async(const float* __restrict__ img, uint32_t dst_s, int t) {
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    const int q = t + i * 128, ci = q >> 8, r = (q >> 3) & 31, c4 = q & 7;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_s + (uint32_t)(((ci * 34 + r + 1) * IMG_ROW + 4 + c4 * 4) * 4)),
                 "l"(img + ci * 1024 + r * 32 + c4 * 4) : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}

// the four im2col tiles of one item (half image h) for pooled pixel p of that half; `one` = value of tap 27
__device__ __forceinline__ void build_a_tiles(const float* __restrict__ im, uint8_t* __restrict__ tiles, int h, int p, float one) {
  const int pr = 8 * h + (p >> 4), pc = p & 15;
  float patch[3][4][4];
#pragma unroll
  for (int ci = 0; ci < 3; ++ci)
#pragma unroll
    for (int rr = 0; rr < 4; ++rr)
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) patch[ci][rr][cc] = im[(ci * 34 + 2 * pr + rr) * IMG_ROW + 3 + 2 * pc + cc];
  const int sw = (p >> 1) & 3;
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    const int dy = w >> 1, dx = w & 1;
    float k[32];
#pragma unroll
    for (int ci = 0; ci < 3; ++ci)
#pragma unroll
      for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) k[ci * 9 + ky * 3 + kx] = patch[ci][dy + ky][dx + kx];
    k[27] = one; k[28] = k[29] = k[30] = k[31] = 0.f;
    uint8_t* row = tiles + w * TILE_BYTES + p * 64;
#pragma unroll
    for (int ch = 0; ch < 4; ++ch)
      *reinterpret_cast<uint4*>(row + ((ch ^ sw) << 4)) = make_uint4(pack_bf16(k[ch * 8], k[ch * 8 + 1]), pack_bf16(k[ch * 8 + 2], k[ch * 8 + 3]),
                                                                     pack_bf16(k[ch * 8 + 4], k[ch * 8 + 5]), pack_bf16(k[ch * 8 + 6], k[ch * 8 + 7]));
  }
}

// zero the halo of the two padded image slots (rows 0 and 33, columns 0..3 and 36..39 of rows 1..32): 2016 floats, disjoint
// from the bytes cp.async fills
__device__ __forceinline__ void zero_image_halos(float* im_gen, int tid, int nthreads) {
  for (int i = tid; i < 2016; i += nthreads) {
    int off;
    if (i < 480) { const int s_ = i / 240, rem = i % 240, ci = rem / 80, rr = (rem % 80) / 40, c = rem % 40; off = s_ * IMG_FLOATS + (ci * 34 + rr * 33) * IMG_ROW + c; }
    else { const int j = i - 480, s_ = j / 768, rem = j % 768, ci = rem / 256, r = 1 + (rem % 256) / 8, cc = rem % 8; off = s_ * IMG_FLOATS + (ci * 34 + r) * IMG_ROW + (cc < 4 ? cc : 32 + cc); }
    im_gen[off] = 0.f;
  }
}

constexpr int STAGE_F = 4 * TILE_BYTES;                 // forward stage: 4 A tiles
constexpr int STAGE_B = 8 * TILE_BYTES;                 // backward stage: 4 A tiles + 4 gradient tiles

// ============================================================================================
// forward.  288 threads: warp 0 MMA, warps 1..4 image loader + im2col builders, warps 5..8 epilogue
// ============================================================================================
__global__ void __launch_bounds__(288, 2)
k_stem_tc_fwd(const float* __restrict__ img, const float* __restrict__ wp /*[27][32] fp32*/, const float* __restrict__ bias,
              bf16* __restrict__ x0, uint8_t* __restrict__ idx0, int B, nrd_bn_sync_args bs, long long* __restrict__ dbg) {
  extern __shared__ uint8_t smem_raw[];
  long long* dbgt = (dbg && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) ? dbg : nullptr;     // stamp rows 16 / 17
  const int drow = blockIdx.x == 0 ? 16 : 17;
  if (threadIdx.x == 0) nrd_dbg_stamp(dbgt, drow, 0);
  nrd_bn_sync_apply(bs);
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  uint8_t* a_gen = gen;                                            // 2 stages x 4 tiles
  uint8_t* w_gen = gen + 2 * STAGE_F;                              // W1 as [32 co][32 taps] bf16, SWIZZLE_64B rows
  float* im_gen = reinterpret_cast<float*>(w_gen + 2048);          // 2 padded images
  float* bias_s = im_gen + 2 * IMG_FLOATS;
  const uint32_t bar0 = smem_u32(bias_s + 32);
  auto afull = [&](int s) { return bar0 + 8u * s; };
  auto aempty = [&](int s) { return bar0 + 8u * (2 + s); };
  auto tfull = [&](int s) { return bar0 + 8u * (4 + s); };
  auto tempty = [&](int s) { return bar0 + 8u * (6 + s); };
  const uint32_t tmem_slot = bar0 + 64;
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - base));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t TMEM_COLS = 256;                              // 2 buffers x 4 window positions x 32 columns

  // image 0 first (the longest latency of the prologue: DRAM): its interior goes in by cp.async, the halo is zeroed below by
  // plain stores -- disjoint bytes, so the two need no ordering
  const int nimg = (B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // images of this CTA
  if (warp >= 1 && warp <= 4 && nimg > 0) load_image_async(img + (size_t)blockIdx.x * 3072, smem_u32(im_gen), threadIdx.x - 32);
  if (warp >= 1 && warp <= 4 && nimg > 1) load_image_async(img + (size_t)(blockIdx.x + gridDim.x) * 3072, smem_u32(im_gen) + IMG_FLOATS * 4, threadIdx.x - 32);
  zero_image_halos(im_gen, threadIdx.x, blockDim.x);                                    // halos stay zero for good
  if (threadIdx.x < 128) {                                         // W1 -> bf16 K-major rows; row co, chunk ch = taps ch*8..+7
    const int co = threadIdx.x >> 2, ch = threadIdx.x & 3;
    float k[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) { const int t = ch * 8 + e; k[e] = t < 27 ? wp[t * SC + co] : 0.f; }
    *reinterpret_cast<uint4*>(w_gen + co * 64 + ((ch ^ ((co >> 1) & 3)) << 4)) =
        make_uint4(pack_bf16(k[0], k[1]), pack_bf16(k[2], k[3]), pack_bf16(k[4], k[5]), pack_bf16(k[6], k[7]));
  }
  if (threadIdx.x < 32) bias_s[threadIdx.x] = bias[threadIdx.x];
  fence_proxy_async();
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < 2; ++s) { mbar_init(afull(s), 128); mbar_init(aempty(s), 1); mbar_init(tfull(s), 1); mbar_init(tempty(s), 4); }
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;
  if (threadIdx.x == 0) nrd_dbg_stamp(dbgt, drow, 1);

  if (warp == 0) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(128, SC, 0, 0);
      const uint32_t a_s = base, w_s = base + 2 * STAGE_F;
      for (int it = 0; it < 2 * nimg; ++it) {
        const int st = it & 1; const uint32_t ph = (it >> 1) & 1;
        mbar_wait(afull(st), ph);
        mbar_wait(tempty(st), ph ^ 1);
        tc_fence_after();
        const uint64_t bd = make_sdesc(w_s, 16, 512, 4);
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          const uint64_t ad = make_sdesc(a_s + st * STAGE_F + w * TILE_BYTES, 16, 512, 4);
          umma_bf16(tmem_base + (st * 4 + w) * SC, ad, bd, idesc, 0);
          umma_bf16(tmem_base + (st * 4 + w) * SC, ad + 2, bd + 2, idesc, 1);
        }
        umma_commit(aempty(st));
        umma_commit(tfull(st));
      }
    }
  } else if (warp <= 4) {
    // ===== image loader + im2col builders =====
    const int t = threadIdx.x - 32;
    const uint32_t im_s = smem_u32(im_gen);
    for (int i = 0; i < nimg; ++i) {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      asm volatile("bar.sync 2, 128;" ::: "memory");               // image i complete and visible; image i-1's buffer is free
      if (t == 0 && i == 0) nrd_dbg_stamp(dbgt, drow, 2);
      if (i >= 1 && i + 1 < nimg) load_image_async(img + (size_t)(blockIdx.x + (i + 1) * gridDim.x) * 3072, im_s + ((i + 1) & 1) * IMG_FLOATS * 4, t);
      const float* im = im_gen + (i & 1) * IMG_FLOATS;
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        const int it = i * 2 + h, st = it & 1;
        mbar_wait(aempty(st), ((it >> 1) & 1) ^ 1);
        build_a_tiles(im, a_gen + st * STAGE_F, h, t, 1.0f);
        fence_proxy_async();
        mbar_arrive(afull(st));
        if (t == 0 && it < 2) nrd_dbg_stamp(dbgt, drow, 3 + it);
      }
    }
  } else {
    // ===== epilogue: thread = pooled pixel (TMEM lane); running max over the 4 window positions =====
    const int q = warp & 3, p = q * 32 + lane;
    for (int it = 0; it < 2 * nimg; ++it) {
      const int st = it & 1, i = it >> 1, h = it & 1;
      const int b = blockIdx.x + i * gridDim.x;
      mbar_wait(tfull(st), (it >> 1) & 1);
      tc_fence_after();
      if (threadIdx.x == 160 && it < 2) nrd_dbg_stamp(dbgt, drow, 5 + 2 * it);
      float m[32]; uint32_t mi[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) mi[j] = 0;
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (st * 4 + w) * SC, v);
        if (w == 3) { tc_fence_before(); __syncwarp(); if (lane == 0) mbar_arrive(tempty(st)); }
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          if (w == 0) m[c] = v[c];
          else if (v[c] > m[c]) { m[c] = v[c]; mi[c >> 2] = (mi[c >> 2] & ~(0xFFu << (8 * (c & 3)))) | ((uint32_t)w << (8 * (c & 3))); }   // first max wins
        }
      }
      const size_t o = ((size_t)b * 256 + h * 128 + p) * SC;
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        uint4 pk;
        pk.x = pack_bf16(fmaxf(m[ch * 8 + 0] + bias_s[ch * 8 + 0], 0.f), fmaxf(m[ch * 8 + 1] + bias_s[ch * 8 + 1], 0.f));
        pk.y = pack_bf16(fmaxf(m[ch * 8 + 2] + bias_s[ch * 8 + 2], 0.f), fmaxf(m[ch * 8 + 3] + bias_s[ch * 8 + 3], 0.f));
        pk.z = pack_bf16(fmaxf(m[ch * 8 + 4] + bias_s[ch * 8 + 4], 0.f), fmaxf(m[ch * 8 + 5] + bias_s[ch * 8 + 5], 0.f));
        pk.w = pack_bf16(fmaxf(m[ch * 8 + 6] + bias_s[ch * 8 + 6], 0.f), fmaxf(m[ch * 8 + 7] + bias_s[ch * 8 + 7], 0.f));
        reinterpret_cast<uint4*>(x0 + o)[ch] = pk;
      }
      reinterpret_cast<uint4*>(idx0 + o)[0] = make_uint4(mi[0], mi[1], mi[2], mi[3]);
      reinterpret_cast<uint4*>(idx0 + o)[1] = make_uint4(mi[4], mi[5], mi[6], mi[7]);
      if (threadIdx.x == 160 && it < 2) nrd_dbg_stamp(dbgt, drow, 6 + 2 * it);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
  if (threadIdx.x == 0) nrd_dbg_stamp(dbgt, drow, 9);
}

// ============================================================================================
// backward (weight + bias gradient).  288 threads: warp 0 MMA, warps 1..4 im2col builders, warps 5..8 gradient-tile builders
// ============================================================================================
__global__ void __launch_bounds__(288, 1)
k_stem_tc_bwd(const float* __restrict__ img, const bf16* __restrict__ g0, const bf16* __restrict__ x0, const uint8_t* __restrict__ idx0,
              float* __restrict__ partial /*[grid][28][32]*/, int B, long long* __restrict__ dbg) {
  extern __shared__ uint8_t smem_raw[];
  long long* dbgt = (dbg && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) ? dbg : nullptr;     // stamp rows 24 / 25
  const int drow = blockIdx.x == 0 ? 24 : 25;
  if (threadIdx.x == 0) nrd_dbg_stamp(dbgt, drow, 0);
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  uint8_t* a_gen = gen;                                            // 2 stages x (4 A tiles + 4 G tiles)
  float* im_gen = reinterpret_cast<float*>(gen + 2 * STAGE_B + 4 * TILE_BYTES);   // (+ slack: M-groups 1..3 of the last tile read past it)
  const uint32_t bar0 = smem_u32(im_gen + 2 * IMG_FLOATS);
  auto afull = [&](int s) { return bar0 + 8u * s; };
  auto aempty = [&](int s) { return bar0 + 8u * (2 + s); };
  const uint32_t done = bar0 + 32;
  const uint32_t tmem_slot = bar0 + 40;
  volatile uint32_t* tmem_slot_gen = reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - base));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t TMEM_COLS = 32;

  // image 0 first (it does not depend on the predecessor kernel): interior by cp.async, halo zeroed by plain stores (disjoint bytes)
  const int nimg = (B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  if (warp >= 1 && warp <= 4 && nimg > 0) load_image_async(img + (size_t)blockIdx.x * 3072, smem_u32(im_gen), threadIdx.x - 32);
  if (warp >= 1 && warp <= 4 && nimg > 1) load_image_async(img + (size_t)(blockIdx.x + gridDim.x) * 3072, smem_u32(im_gen) + IMG_FLOATS * 4, threadIdx.x - 32);
  zero_image_halos(im_gen, threadIdx.x, blockDim.x);
  // (the slack behind the last stage, which the junk M-groups 1..3 of the last tiles read, stays uninitialised: a junk row of A
  // only ever reaches the same junk row of the accumulator, which nobody reads)
  fence_proxy_async();
  if (warp == 0 && lane == 0) {
    for (int s = 0; s < 2; ++s) { mbar_init(afull(s), 256); mbar_init(aempty(s), 1); }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (threadIdx.x == 0) nrd_dbg_stamp(dbgt, drow, 1);
  nrd_pdl_launch();
  nrd_pdl_wait();          // set-up above overlaps the predecessor's tail; dL/dx_0 is read below
  if (threadIdx.x == 0) nrd_dbg_stamp(dbgt, drow, 2);
  const uint32_t tmem_base = *tmem_slot_gen;

  if (warp == 0) {
    if (lane == 0) {
      // D[tap (M rows 0..31; rows 32..127 are junk M-groups)][co] += A_w[pixel][tap]^T * G_w[pixel][co]; both operands MN-major
      constexpr uint32_t idesc = make_idesc(128, SC, 1, 1);
      for (int it = 0; it < 2 * nimg; ++it) {
        const int st = it & 1;
        mbar_wait(afull(st), (it >> 1) & 1);
        tc_fence_after();
        const uint32_t sb = base + st * STAGE_B;
#pragma unroll
        for (int w = 0; w < 4; ++w)
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {     // 16 pooled pixels per MMA
            const uint64_t ad = make_sdesc(sb + w * TILE_BYTES + ks * 1024, TILE_BYTES, 512, 4);
            const uint64_t bd = make_sdesc(sb + (4 + w) * TILE_BYTES + ks * 1024, TILE_BYTES, 512, 4);
            umma_bf16(tmem_base, ad, bd, idesc, (it | w | ks) != 0);
          }
        umma_commit(aempty(st));
      }
      umma_commit(done);
    }
  } else if (warp <= 4) {
    const int t = threadIdx.x - 32;
    const uint32_t im_s = smem_u32(im_gen);
    for (int i = 0; i < nimg; ++i) {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      asm volatile("bar.sync 2, 128;" ::: "memory");
      if (t == 0 && i == 0) nrd_dbg_stamp(dbgt, drow, 3);
      // images 0 and 1 were requested in the prologue; image i+1 (i >= 1) goes into the slot image i-1 has left (the barrier above)
      if (i >= 1 && i + 1 < nimg) load_image_async(img + (size_t)(blockIdx.x + (i + 1) * gridDim.x) * 3072, im_s + ((i + 1) & 1) * IMG_FLOATS * 4, t);
      const float* im = im_gen + (i & 1) * IMG_FLOATS;
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        const int it = i * 2 + h, st = it & 1;
        mbar_wait(aempty(st), ((it >> 1) & 1) ^ 1);
        build_a_tiles(im, a_gen + st * STAGE_B, h, t, 1.0f);       // tap 27 = 1 -> accumulator row 27 = bias gradient
        fence_proxy_async();
        mbar_arrive(afull(st));
        if (t == 0 && it < 4) nrd_dbg_stamp(dbgt, drow, 4 + it);
      }
    }
  } else {
    // ===== gradient tiles: G_w[pixel][co] = g0 where (argmax == w and x0 > 0) else 0 =====
    // the next half image's g0 / x0 / arg-max rows are fetched while the current one is turned into tiles (L2 latency off the
    // per-half critical path)
    const int t = threadIdx.x - 160, q = warp & 3;
    uint4 gv[4], xv[4], iv[2], gn[4], xn[4], in_[2];
    auto fetch = [&](int it, uint4 (&g)[4], uint4 (&x)[4], uint4 (&ii)[2]) {
      const int i = it >> 1, h = it & 1, b = blockIdx.x + i * gridDim.x;
      const size_t o = ((size_t)b * 256 + h * 128 + t) * SC;
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) { g[ch] = reinterpret_cast<const uint4*>(g0 + o)[ch]; x[ch] = reinterpret_cast<const uint4*>(x0 + o)[ch]; }
      ii[0] = reinterpret_cast<const uint4*>(idx0 + o)[0]; ii[1] = reinterpret_cast<const uint4*>(idx0 + o)[1];
    };
    if (nimg > 0) fetch(0, gv, xv, iv);
    for (int it = 0; it < 2 * nimg; ++it) {
      const int st = it & 1;
      if (it + 1 < 2 * nimg) fetch(it + 1, gn, xn, in_);
      mbar_wait(aempty(st), ((it >> 1) & 1) ^ 1);
      const uint32_t* gw = reinterpret_cast<const uint32_t*>(gv); const uint32_t* xw = reinterpret_cast<const uint32_t*>(xv);
      const uint8_t* ib = reinterpret_cast<const uint8_t*>(iv);
      const int sw = (t >> 1) & 3;
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        uint8_t* row = a_gen + st * STAGE_B + (4 + w) * TILE_BYTES + t * 64;
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          uint32_t o4[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int c = ch * 8 + 2 * e;
            const uint32_t g2 = gw[ch * 4 + e], x2 = xw[ch * 4 + e];
            // bf16 pair: low half = channel c, high half = channel c+1; x0 >= 0 always, so "x0 > 0" is "bits != 0 (ignoring -0)"
            const bool k0 = ib[c] == w && (x2 & 0x7FFFu) != 0, k1 = ib[c + 1] == w && (x2 & 0x7FFF0000u) != 0;
            o4[e] = (k0 ? (g2 & 0xFFFFu) : 0u) | (k1 ? (g2 & 0xFFFF0000u) : 0u);
          }
          *reinterpret_cast<uint4*>(row + ((ch ^ sw) << 4)) = make_uint4(o4[0], o4[1], o4[2], o4[3]);
        }
      }
      fence_proxy_async();
      mbar_arrive(afull(st));
      if (t == 0 && it < 4) nrd_dbg_stamp(dbgt, drow, 8 + it);
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) { gv[ch] = gn[ch]; xv[ch] = xn[ch]; }
      iv[0] = in_[0]; iv[1] = in_[1];
    }
    if (q == 0) {                                                  // TMEM lanes 0..31 = taps
      mbar_wait(done, 0);
      tc_fence_after();
      if (t == 0) nrd_dbg_stamp(dbgt, drow, 12);
      float v[32];
      tmem_ld32(tmem_base, v);
      if (nimg == 0) {
#pragma unroll
        for (int c = 0; c < 32; ++c) v[c] = 0.f;
      }
      if (lane < 28) {
        float4* dst = reinterpret_cast<float4*>(partial + ((size_t)blockIdx.x * 28 + lane) * SC);
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) dst[c4] = make_float4(v[c4 * 4], v[c4 * 4 + 1], v[c4 * 4 + 2], v[c4 * 4 + 3]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
  if (threadIdx.x == 0) nrd_dbg_stamp(dbgt, drow, 13);
}

constexpr size_t stem_fwd_smem() { return 1024 + 2 * STAGE_F + 2048 + 2 * IMG_FLOATS * 4 + 128 + 128 + 64; }
constexpr size_t stem_bwd_smem() { return 1024 + 2 * STAGE_B + 4 * TILE_BYTES + 2 * IMG_FLOATS * 4 + 128 + 64; }

}  // namespace

void nrd_count_launch();

int nrd_tc_stem_init(nrd_plan* p) {
  p->tc_stem = 0;
  if (p->dtype != NRD_BF16 || p->C != 32 || !p->tc_ready) return NRD_OK;
  NRD_CHECK_CUDA(cudaFuncSetAttribute(k_stem_tc_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stem_fwd_smem()));
  // two CTAs per SM (2 x 101 KB of shared memory, 2 x 256 TMEM columns): ask for the full shared-memory carve-out
  NRD_CHECK_CUDA(cudaFuncSetAttribute(k_stem_tc_fwd, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  NRD_CHECK_CUDA(cudaFuncSetAttribute(k_stem_tc_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stem_bwd_smem()));
  p->tc_stem = 1;
  return NRD_OK;
}

static int stem_grid(const nrd_plan* p) { return p->B < NRD_NUM_SMS ? p->B : NRD_NUM_SMS; }
// forward: the kernel is a latency chain per image (load, im2col build, 8 small MMAs, pooled epilogue), so two CTAs share an SM
// and at B = 256 every CTA has ONE image instead of 108 CTAs having two in a row (NRD_STEM_CTAS_PER_SM=1 restores one per SM)
static int stem_fwd_grid(const nrd_plan* p) {
  static int per_sm = 0;
  if (!per_sm) { const char* e = getenv("NRD_STEM_CTAS_PER_SM"); per_sm = (e && atoi(e) == 1) ? 1 : 2; }
  const int cap = per_sm * NRD_NUM_SMS;
  return p->B < cap ? p->B : cap;
}

int nrd_tc_stem_fwd(nrd_plan* p, const float* images, int training, cudaStream_t st) {
  NRD_REQUIRE(p->tc_stem, "tensor-core stem not available for this plan");
  k_stem_tc_fwd<<<stem_fwd_grid(p), 288, stem_fwd_smem(), st>>>(images, p->conv1_wp, p->params + p->off[1], (bf16*)p->xs, p->idx0, p->B, nrd_bn_sync_of(p, training), p->dbg);
  nrd_count_launch();
  NRD_CHECK_CUDA(cudaGetLastError());
  return NRD_OK;
}

// per-CTA partials [stem_grid][28][32] into p->stem_partial (same layout the CUDA-core path writes; reduced by the caller)
int nrd_tc_stem_bwd(nrd_plan* p, const float* images, const void* g0, cudaStream_t st) {
  NRD_REQUIRE(p->tc_stem, "tensor-core stem not available for this plan");
  NRD_CHECK_CUDA(nrd_launch_pdl(k_stem_tc_bwd, dim3(stem_grid(p)), dim3(288), stem_bwd_smem(), st, (p->pdl >> 1) & 1, images, (const bf16*)g0, (const bf16*)p->xs,
                                (const uint8_t*)p->idx0, p->stem_partial, p->B, p->dbg));
  nrd_count_launch();
  NRD_CHECK_CUDA(cudaGetLastError());
  return NRD_OK;
}
