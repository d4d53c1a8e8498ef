// CUDA-core kernels of the NetResDeep training step (validation-mode fp32 path and every
// non-GEMM op of the bf16 path).  Activations are NHWC, T = float | bf16; parameters, gradients,
// BN statistics and all accumulation are fp32 (fp64 for the final cross-CTA combines).
// Every cross-CTA reduction is two-stage with a fixed summation order => bitwise reproducible.
#include <stdlib.h>
#include "nrd_common.cuh"
#include <stdarg.h>

// ------------------------------------------------------------------------------------------
// element helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

// 8 contiguous elements <-> 8 floats (16 B for bf16, 32 B for fp32); pointer must be 16-B aligned
__device__ __forceinline__ void ld8(const float* p, float (&v)[8]) {
  float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void ld8(const bf16* p, float (&v)[8]) {
  uint4 r = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) { float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
}
__device__ __forceinline__ void st8(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void st8(bf16* p, const float (&v)[8]) {
  uint4 r; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = r;
}
__device__ __forceinline__ void ld4(const float* p, float (&v)[4]) {
  float4 a = *reinterpret_cast<const float4*>(p); v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
__device__ __forceinline__ void ld4(const bf16* p, float (&v)[4]) {
  uint2 r = *reinterpret_cast<const uint2*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
  float2 a = __bfloat1622float2(h[0]), b = __bfloat1622float2(h[1]);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}
__device__ __forceinline__ void st4(float* p, const float (&v)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void st4(bf16* p, const float (&v)[4]) {
  uint2 r; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
  h[0] = __floats2bfloat162_rn(v[0], v[1]); h[1] = __floats2bfloat162_rn(v[2], v[3]);
  *reinterpret_cast<uint2*>(p) = r;
}

// ------------------------------------------------------------------------------------------
// weight packing (runs after every parameter change)
// ------------------------------------------------------------------------------------------
// w_fwd[t][ci][co]   = W[co][ci][ky][kx],            t = ky*3+kx
// w_dgrad[t][co][ci] = W[co][ci][2-ky][2-kx]  (dgrad is a forward conv of dy with flipped taps)
// wtc_fwd[co][t*C+ci]  (bf16, K-major B operand), wtc_dgrad[ci][t*C+co] (bf16, flipped taps)
// w1p[hw*C+c][j]     = fc1.weight[j][c*64+hw]        (k-major transpose; flatten order c,h,w -> NHWC order)
// conv1_wp[k][co]    = conv1.weight[co][k],          k = ci*9+ky*3+kx
__global__ void k_pack_weights(const float* __restrict__ rb_w, const float* __restrict__ fc1_w,
                               const float* __restrict__ conv1_w, int C, float* __restrict__ w_fwd,
                               float* __restrict__ w_dgrad, bf16* __restrict__ wtc_fwd,
                               bf16* __restrict__ wtc_dgrad, float* __restrict__ w1p,
                               float* __restrict__ conv1_wp) {
  const int n_rb = C * C * 9, n_fc1 = NRD_FC1_OUT * 64 * C, n_c1 = C * 27;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_rb + n_fc1 + n_c1; i += gridDim.x * blockDim.x) {
    if (i < n_rb) {
      int co = i / (C * 9), r = i % (C * 9), ci = r / 9, t = r % 9;
      float w = rb_w[i];
      if (w_fwd) w_fwd[(t * C + ci) * C + co] = w;
      if (w_dgrad) w_dgrad[((8 - t) * C + co) * C + ci] = w;
      if (wtc_fwd) wtc_fwd[(size_t)co * 9 * C + t * C + ci] = __float2bfloat16_rn(w);
      if (wtc_dgrad) wtc_dgrad[(size_t)ci * 9 * C + (8 - t) * C + co] = __float2bfloat16_rn(w);
    } else if (i < n_rb + n_fc1) {
      int q = i - n_rb;
      int j = q / (64 * C), r = q % (64 * C), c = r / 64, hw = r % 64;
      if (w1p) w1p[((size_t)hw * C + c) * NRD_FC1_OUT + j] = fc1_w[q];   // w1t[k][j], k = NHWC column
    } else {
      int q = i - n_rb - n_fc1;
      int co = q / 27, k = q % 27;
      if (conv1_wp) conv1_wp[k * C + co] = conv1_w[q];
    }
  }
}

// ------------------------------------------------------------------------------------------
// stem: x0 = maxpool2(relu(conv1(x) + b))            (model/resnet.py:16)
// grid (4 bands, B); block 64*(C/8) threads: thread = (pooled pixel of the band, 8-channel group)
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void k_stem_fwd(const float* __restrict__ img, const float* __restrict__ wp /*[27][C]*/,
                           const float* __restrict__ bias, T* __restrict__ x0, uint8_t* __restrict__ idx0, int C, nrd_bn_sync_args bs) {
  extern __shared__ float sm[];
  nrd_bn_sync_apply(bs);
  float* tile = sm;                 // [3][10][35], padded to 1056 floats so wsm stays 16-B aligned
  float* wsm = sm + 1056;           // [27][C]
  float* bsm = wsm + 27 * C;        // [C]
  const int band = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, nt = blockDim.x;
  const float* im = img + (size_t)b * 3 * 1024;
  for (int i = tid; i < 3 * 10 * 34; i += nt) {
    int ci = i / 340, r = (i % 340) / 34, c = i % 34;
    int gr = band * 8 - 1 + r, gc = c - 1;
    float v = 0.f;
    if (gr >= 0 && gr < 32 && gc >= 0 && gc < 32) v = im[ci * 1024 + gr * 32 + gc];
    tile[(ci * 10 + r) * 35 + c] = v;
  }
  for (int i = tid; i < 27 * C; i += nt) wsm[i] = wp[i];
  for (int i = tid; i < C; i += nt) bsm[i] = bias[i];
  __syncthreads();
  const int ncg = C / 8, cg = tid % ncg, px = tid / ncg, ppy = px / 16, ppx = px % 16;
  float acc[4][8];
#pragma unroll
  for (int p = 0; p < 4; ++p)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[p][j] = 0.f;
#pragma unroll
  for (int ci = 0; ci < 3; ++ci) {
    float patch[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) patch[r][c] = tile[(ci * 10 + 2 * ppy + r) * 35 + 2 * ppx + c];
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const float* w = wsm + (ci * 9 + ky * 3 + kx) * C + cg * 8;
        float4 w0 = *reinterpret_cast<const float4*>(w), w1 = *reinterpret_cast<const float4*>(w + 4);
        float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          float v = patch[(p >> 1) + ky][(p & 1) + kx];
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[p][j] = fmaf(v, wv[j], acc[p][j]);
        }
      }
  }
  float out[8]; uint8_t id[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float bj = bsm[cg * 8 + j];
    float m = acc[0][j] + bj; int mi = 0;
#pragma unroll
    for (int p = 1; p < 4; ++p) { float v = acc[p][j] + bj; if (v > m) { m = v; mi = p; } }  // first max wins
    out[j] = fmaxf(m, 0.f); id[j] = (uint8_t)mi;
  }
  size_t o = (((size_t)b * 16 + band * 4 + ppy) * 16 + ppx) * C + cg * 8;
  st8(x0 + o, out);
  uint2 pk; uint8_t* pb = reinterpret_cast<uint8_t*>(&pk);
#pragma unroll
  for (int j = 0; j < 8; ++j) pb[j] = id[j];
  *reinterpret_cast<uint2*>(idx0 + o) = pk;
}

// stem backward: dconv1.weight / dconv1.bias only (the input needs no grad).  One CTA loops over images; per image the
// masked pooled gradient (relu'(0)=0, pool routes to the first max) and the arg-max indices are staged in shared memory
// with coalesced loads, then thread = (co, pixel-group) accumulates its 27 taps + bias from shared memory only.
// Per-CTA partial [28][C], reduced afterwards in fixed order.
template <typename T>
__global__ void __launch_bounds__(256) k_stem_bwd(const float* __restrict__ img, const T* __restrict__ g0, const T* __restrict__ x0,
                           const uint8_t* __restrict__ idx0, float* __restrict__ partial, int B, int C) {
  extern __shared__ float sm[];
  float* tile = sm;                       // [3][34][35] zero-padded image
  float* gm = sm + 3 * 34 * 35 + 2;       // [256][C] masked gradient (16-B aligned: 3570+2 floats)
  uint8_t* im_idx = reinterpret_cast<uint8_t*>(gm + 256 * C);   // [256][C]
  float* red = gm;                        // reused at the end: [ngroups][28][C]  (ngroups*28 <= 256)
  const int tid = threadIdx.x, nt = blockDim.x, co = tid % C, grp = tid / C, ngrp = nt / C;
  float acc[28];
#pragma unroll
  for (int k = 0; k < 28; ++k) acc[k] = 0.f;
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    __syncthreads();
    const float* im = img + (size_t)b * 3072;
    for (int i = tid; i < 3 * 34 * 34; i += nt) {
      int ci = i / 1156, r = (i % 1156) / 34, c = i % 34;
      float v = 0.f;
      if (r >= 1 && r <= 32 && c >= 1 && c <= 32) v = im[ci * 1024 + (r - 1) * 32 + (c - 1)];
      tile[(ci * 34 + r) * 35 + c] = v;
    }
    const size_t base = (size_t)b * 256 * C;
    for (int i = tid; i < 256 * C / 8; i += nt) {
      float gv[8], xv[8];
      ld8(g0 + base + (size_t)i * 8, gv); ld8(x0 + base + (size_t)i * 8, xv);
#pragma unroll
      for (int j = 0; j < 8; ++j) gm[i * 8 + j] = xv[j] > 0.f ? gv[j] : 0.f;
      *reinterpret_cast<uint2*>(im_idx + (size_t)i * 8) = *reinterpret_cast<const uint2*>(idx0 + base + (size_t)i * 8);
    }
    __syncthreads();
    for (int px = grp; px < 256; px += ngrp) {
      const float g = gm[px * C + co];
      const int id = im_idx[px * C + co];
      const int h = 2 * (px / 16) + (id >> 1), w = 2 * (px % 16) + (id & 1);   // conv-output position
      acc[27] += g;
#pragma unroll
      for (int ci = 0; ci < 3; ++ci)
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
          for (int kx = 0; kx < 3; ++kx)
            acc[ci * 9 + ky * 3 + kx] = fmaf(g, tile[(ci * 34 + h + ky) * 35 + w + kx], acc[ci * 9 + ky * 3 + kx]);
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 28; ++k) red[(grp * 28 + k) * C + co] = acc[k];
  __syncthreads();
  for (int i = tid; i < 28 * C; i += nt) {
    float s = 0.f;
    for (int g = 0; g < ngrp; ++g) s += red[g * 28 * C + i];
    partial[(size_t)blockIdx.x * 28 * C + i] = s;
  }
}

// Where a finished gradient element goes when the optimizer step rides the kernel that produces it (nrd_plan_set_fused_update):
// the master parameter and, for the three tensors that have them, the packed copies the kernels read (k_pack_weights' layouts).
struct nrd_upd_args {
  float* params;            // flat master parameters, or NULL: no update here (a separate nrd_sgd_update_pack follows)
  float step_scale;         // lr * inv_world
  float *w_fwd, *w_dgrad; bf16 *wtc_fwd, *wtc_dgrad; float* conv1_wp;
};
// flat index pi (conv weight [co][ci][tap] at off_rbw, conv1 weight [co][27] at off_c1w): master + packed copies
static __device__ __forceinline__ void nrd_apply_update(const nrd_upd_args& u, long long pi, float g, float p_old, int C, long long off_rbw, long long off_c1w) {
  const float w = fmaf(-u.step_scale, g, p_old);
  u.params[pi] = w;
  const long long n_rb = (long long)C * C * 9, n_c1 = (long long)C * 27;
  if (pi >= off_rbw && pi < off_rbw + n_rb) {
    const int q = (int)(pi - off_rbw), co = q / (C * 9), r9 = q % (C * 9), ci = r9 / 9, t = r9 % 9;
    u.w_fwd[(t * C + ci) * C + co] = w;
    u.w_dgrad[((8 - t) * C + co) * C + ci] = w;
    u.wtc_fwd[(size_t)co * 9 * C + t * C + ci] = __float2bfloat16_rn(w);
    u.wtc_dgrad[(size_t)ci * 9 * C + (8 - t) * C + co] = __float2bfloat16_rn(w);
  } else if (pi >= off_c1w && pi < off_c1w + n_c1) {
    const int q = (int)(pi - off_c1w), co = q / 27, k = q % 27;
    u.conv1_wp[k * C + co] = w;
  }
}

// Body gradients, last kernel of the step's critical path: the per-CTA partials of the block-conv weight gradient ([ctas][9][C][C])
// and of the stem's ([ctas][28][C]) are summed in a fixed order -- CTA = 16 float4 columns x 16 row classes, every thread's
// <= 10 loads independent and issued together (one L2 round trip), row classes combined in order -- and each finished element
//   * goes to the flat gradient vector (PyTorch layouts),
//   * peer-memory exchange on (x.world > 0): is stored into this rank's slot of EVERY rank's exchange buffer (the kernel's own
//     order: a thread's 4 values are 16 contiguous bytes per peer); the last CTA adds the BN gradients (written by the
//     persistent backward) and the BN running-stat slot, and whichever CTA finishes last publishes the epoch in every rank's
//     flag (nrd_common.cuh),
//   * fused update on (u.params != NULL): takes the optimizer step right here -- alone (world 1), or after waiting for the
//     world's flags and summing the world's slots in rank order (the same order on every rank => identical weights) --
//     so no optimizer launch follows the backward.
__global__ void __launch_bounds__(256) k_body_reduce(const float* __restrict__ wpartial, int wblk, const float* __restrict__ spartial, int sblk,
                                                     int C, float* __restrict__ grads_base, float* __restrict__ bn_sync,
                                                     const float* __restrict__ running, float contrib, nrd_xchg_args x,
                                                     int off_bng, int off_rbw, int off_c1w, int off_c1b, nrd_upd_args u, long long* __restrict__ dbg) {
  __shared__ float4 red[16][16];
  long long* dbgt = (dbg && threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) ? dbg : nullptr;
  const int drow = blockIdx.x == 0 ? 8 : 9;
  nrd_dbg_stamp(dbgt, drow, 0);
  nrd_pdl_launch();
  nrd_pdl_wait();
  nrd_dbg_stamp(dbgt, drow, 1);
  const bool px = x.world > 0, upd = u.params != nullptr;
  const bool ll = px && upd;                           // fused update: flag-in-data slots (nrd_common.cuh); else flags published per bucket
  // the step's epoch: read by every CTA at entry, bumped by whichever CTA arrives last at a counter (after every CTA has read it)
  __shared__ unsigned long long s_epoch;
  if (px && threadIdx.x == 0) {
    unsigned long long* ctrl = nrd_xchg_flags(x.peer[x.rank], x.world, x.nx, 0);
    s_epoch = ctrl[NRD_XCHG_EPOCH] + 1;
    if (ll) {                                          // (the flag protocol bumps it in nrd_xchg_publish)
      unsigned int* arrive = reinterpret_cast<unsigned int*>(ctrl + NRD_XCHG_ARRIVE);
      if (atomicAdd(arrive, 1u) == gridDim.x - 1) { *arrive = 0u; ctrl[NRD_XCHG_EPOCH] = s_epoch; }
    }
  }
  if (px) __syncthreads();
  const unsigned long long epoch = px ? s_epoch : 0;
  const size_t slot = px ? ((size_t)(epoch & 1) * x.world + x.rank) * x.nx : 0;
  const bool last_cta = blockIdx.x == gridDim.x - 1;
  // DDP buffer sync riding the body bucket: rank 0 contributes world_size * running, the others exact zeros; + dgamma | dbeta
  if (last_cta && (bn_sync || px)) {
    for (int j = threadIdx.x; j < 2 * C; j += blockDim.x) {
      const float v = contrib != 0.f ? running[j] * contrib : 0.f;
      if (bn_sync && !px) bn_sync[j] = v;
      if (px) {
        const float gj = grads_base[off_bng + j];
        if (ll) { nrd_ll_push1(x, epoch, j, v); nrd_ll_push1(x, epoch, 2 * C + off_bng + j, gj); }
        else for (int p = 0; p < x.world; ++p) { x.peer[p][slot + j] = v; x.peer[p][slot + 2 * C + off_bng + j] = gj; }
      }
    }
  }
  const int c4 = threadIdx.x & 15, rc = threadIdx.x >> 4, nconv4 = 9 * C * C / 4;
  const bool conv = (int)blockIdx.x * 16 < nconv4;     // CTA-uniform: both column ranges are multiples of 16 float4
  const int n = conv ? 9 * C * C : 28 * C, nblk = conv ? wblk : sblk;
  const int ibase = (conv ? (int)blockIdx.x * 16 : (int)blockIdx.x * 16 - nconv4) * 4;       // first element of this CTA's 64
  // output role: thread t < 64 finishes element ibase + t; its destination (and, fused update, its old parameter) is known now
  const int t = threadIdx.x;
  long long pi = 0; size_t sidx = 0; float p_old = 0.f;
  if (t < 64) {
    const int i = ibase + t;
    if (conv) { const int tap = i / (C * C), co = (i / C) % C, ci = i % C; pi = off_rbw + ((long long)co * C + ci) * 9 + tap; }
    else { const int k = i / C, co = i % C; pi = k < 27 ? off_c1w + co * 27 + k : off_c1b + co; }
    // the exchange slot keeps THIS kernel's order (conv [tap][co][ci], conv1 [tap][co], then the bias row): a warp's 32 values are
    // 128 contiguous bytes per peer; the consumer un-permutes locally
    sidx = 2 * C + (conv ? off_rbw + i : (i < 27 * C ? off_c1w + i : off_c1b + (i - 27 * C)));
    if (upd) p_old = u.params[pi];
  }
  const float* part = (conv ? wpartial : spartial) + ibase + c4 * 4;
  float4 v[10];
#pragma unroll
  for (int q = 0; q < 10; ++q) {                       // nblk <= 160: rows rc, rc + 16, ...; all loads in flight together
    const int b = rc + 16 * q;
    v[q] = b < nblk ? __ldcg(reinterpret_cast<const float4*>(part + (size_t)b * n)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  float4 s = v[0];
#pragma unroll
  for (int q = 1; q < 10; ++q) { s.x += v[q].x; s.y += v[q].y; s.z += v[q].z; s.w += v[q].w; }
  red[rc][c4] = s;
  __syncthreads();
  nrd_dbg_stamp(dbgt, drow, 2);
  float tot = 0.f;
  if (t < 64) {
    const float* rf = reinterpret_cast<const float*>(&red[0][0]) + t;      // element t of row class k: red[k][t / 4].{x,y,z,w}
#pragma unroll
    for (int k = 0; k < 16; ++k) tot += rf[k * 64];                        // row classes in order
    grads_base[pi] = tot;
    if (px) {
      // 16-byte peer stores (warps 0 and 1 are fully inside t < 64): flag-in-data: lanes 2q, 2q+1 -> {v, epoch, v', epoch};
      // flag protocol: lanes 4q .. 4q+3 -> 4 consecutive slot elements
      const float v1 = __shfl_down_sync(0xffffffffu, tot, 1), v2 = __shfl_down_sync(0xffffffffu, tot, 2), v3 = __shfl_down_sync(0xffffffffu, tot, 3);
      if (ll) { if ((t & 1) == 0) nrd_ll_push2(x, epoch, sidx, tot, v1); }
      else if ((t & 3) == 0) for (int p = 0; p < x.world; ++p) *reinterpret_cast<float4*>(x.peer[p] + slot + sidx) = make_float4(tot, v1, v2, v3);
    }
  }
  if (px && !ll) nrd_xchg_publish(x, 0, epoch);
  nrd_dbg_stamp(dbgt, drow, 4);
  if (!upd) return;
  if (px) {
    // ---- every element is the rank-ordered sum of the world's slots in OUR buffer; each word is waited for by itself ----
    if (t < 64) {
      tot = nrd_ll_sum(x, epoch, sidx);
      grads_base[pi] = tot;                            // .grad keeps holding the world's sum, as after an all-reduce
    }
    nrd_dbg_stamp(dbgt, drow, 5);
    if (last_cta)
      for (int j = threadIdx.x; j < 2 * C; j += blockDim.x) {
        const float sres = nrd_ll_sum(x, epoch, j), gj = nrd_ll_sum(x, epoch, 2 * C + off_bng + j);
        if (bn_sync) bn_sync[j] = sres;
        grads_base[off_bng + j] = gj;
      }
  }
  if (t < 64) nrd_apply_update(u, pi, tot, p_old, C, off_rbw, off_c1w);
  if (last_cta)
    for (int j = threadIdx.x; j < 2 * C; j += blockDim.x) u.params[off_bng + j] = fmaf(-u.step_scale, grads_base[off_bng + j], u.params[off_bng + j]);
  nrd_dbg_stamp(dbgt, drow, 3);
}

// head bucket of the peer-memory exchange: this rank's fc1 / fc2 gradients (flat range [lo, hi)) into its slot of every
// rank's buffer; runs on the side stream behind the head weight-gradient kernels, i.e. under the body backward
__global__ void __launch_bounds__(256) k_xchg_push_head(nrd_xchg_args x, const float* __restrict__ grads, int lo, int hi, int C) {
  const unsigned long long epoch = nrd_xchg_flags(x.peer[x.rank], x.world, x.nx, 1)[NRD_XCHG_EPOCH] + 1;
  const size_t slot = ((size_t)(epoch & 1) * x.world + x.rank) * x.nx + 2 * C;
  const int n4 = (hi - lo) / 4;                       // lo, 2C and nx are multiples of 4: 16-byte stores
  for (int v4 = blockIdx.x * blockDim.x + threadIdx.x; v4 < n4; v4 += gridDim.x * blockDim.x) {
    const float4 v = *reinterpret_cast<const float4*>(grads + lo + v4 * 4);
    for (int p = 0; p < x.world; ++p) *reinterpret_cast<float4*>(x.peer[p] + slot + lo + v4 * 4) = v;
  }
  if (blockIdx.x == 0 && threadIdx.x < (hi - lo) - n4 * 4) {
    const int i = lo + n4 * 4 + threadIdx.x;
    for (int p = 0; p < x.world; ++p) x.peer[p][slot + i] = grads[i];
  }
  nrd_xchg_publish(x, 1, epoch);
}

// CTA = 32 outputs x 8 row classes, fixed-order combine
__global__ void __launch_bounds__(256) k_stem_bwd_reduce(const float* __restrict__ partial, int nblk, int C, float* __restrict__ dw,
                                  float* __restrict__ db) {
  __shared__ double red[8][33];
  const int o = threadIdx.x & 31, r = threadIdx.x >> 5, i = blockIdx.x * 32 + o, n = 28 * C;
  double s = 0.0;
#pragma unroll 8
  for (int b = r; b < nblk; b += 8) s += (double)__ldcg(partial + (size_t)b * n + i);
  red[r][o] = s;
  __syncthreads();
  if (r == 0) {
    double tot = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) tot += red[k][o];
    int k = i / C, co = i % C;
    if (k < 27) dw[co * 27 + k] = (float)tot; else db[co] = (float)tot;
  }
}

// ------------------------------------------------------------------------------------------
// conv3x3 pad 1, C -> C, NHWC, direct on CUDA cores.  CTA = 8x16 pixel tile of one image.
// w: [9][Cin][Cout] fp32.  Optional: per-CTA BN partial sums of the fp32 results; residual add.
// ------------------------------------------------------------------------------------------
template <typename T, int C>
__global__ void __launch_bounds__(256) k_conv3x3(const T* __restrict__ x, const float* __restrict__ w,
                                                 T* __restrict__ y, float* __restrict__ stats_partial,
                                                 const T* __restrict__ residual) {
  constexpr int CP = C + 1;            // padded channel stride (bank-conflict free)
  constexpr int NCOG = C / 4;          // threads along output channels (4 each)
  constexpr int NPG = 256 / NCOG;      // pixel groups
  constexpr int PPT = 128 / NPG;       // pixels per thread (consecutive in a row)
  constexpr int GPR = 16 / PPT;        // groups per row
  extern __shared__ float sm[];
  float* xs = sm;                      // [10][18][CP]
  float* ws = sm + 10 * 18 * CP;       // [9][C][C]
  const int tid = threadIdx.x, blk = blockIdx.x, b = blk >> 1, r0 = (blk & 1) * 8;
  const T* xb = x + (size_t)b * NRD_PIX * C;
  for (int i = tid; i < 10 * 18 * (C / 4); i += 256) {
    int c4 = i % (C / 4), pc = (i / (C / 4)) % 18, pr = i / (C / 4) / 18;
    int gr = r0 - 1 + pr, gc = pc - 1;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (gr >= 0 && gr < 16 && gc >= 0 && gc < 16) ld4(xb + ((size_t)gr * 16 + gc) * C + c4 * 4, v);
    float* d = xs + (pr * 18 + pc) * CP + c4 * 4;
    d[0] = v[0]; d[1] = v[1]; d[2] = v[2]; d[3] = v[3];
  }
  for (int i = tid; i < 9 * C * C / 4; i += 256)
    reinterpret_cast<float4*>(ws)[i] = reinterpret_cast<const float4*>(w)[i];
  __syncthreads();
  const int cog = tid % NCOG, pg = tid / NCOG, row = pg / GPR, col0 = (pg % GPR) * PPT;
  float acc[PPT][4];
#pragma unroll
  for (int j = 0; j < PPT; ++j) { acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f; }
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    const float* xr = xs + ((row + ky) * 18 + col0) * CP;
#pragma unroll 4
    for (int ci = 0; ci < C; ++ci) {
      float in[PPT + 2];
#pragma unroll
      for (int j = 0; j < PPT + 2; ++j) in[j] = xr[j * CP + ci];
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        float4 wv = *reinterpret_cast<const float4*>(ws + ((ky * 3 + kx) * C + ci) * C + cog * 4);
#pragma unroll
        for (int j = 0; j < PPT; ++j) {
          acc[j][0] = fmaf(in[j + kx], wv.x, acc[j][0]); acc[j][1] = fmaf(in[j + kx], wv.y, acc[j][1]);
          acc[j][2] = fmaf(in[j + kx], wv.z, acc[j][2]); acc[j][3] = fmaf(in[j + kx], wv.w, acc[j][3]);
        }
      }
    }
  }
  // epilogue
  float s[4] = {0.f, 0.f, 0.f, 0.f}, q[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int j = 0; j < PPT; ++j) {
#pragma unroll
    for (int e = 0; e < 4; ++e) { s[e] += acc[j][e]; q[e] = fmaf(acc[j][e], acc[j][e], q[e]); }
    size_t o = ((size_t)b * NRD_PIX + (r0 + row) * 16 + col0 + j) * C + cog * 4;
    float v[4] = {acc[j][0], acc[j][1], acc[j][2], acc[j][3]};
    if (residual) { float rr[4]; ld4(residual + o, rr); v[0] += rr[0]; v[1] += rr[1]; v[2] += rr[2]; v[3] += rr[3]; }
    st4(y + o, v);
  }
  if (stats_partial) {
    __syncthreads();                   // everyone is done with xs; reuse it as [NPG][2][C]
    float* red = sm;
#pragma unroll
    for (int e = 0; e < 4; ++e) { red[(pg * 2 + 0) * C + cog * 4 + e] = s[e]; red[(pg * 2 + 1) * C + cog * 4 + e] = q[e]; }
    __syncthreads();
    if (tid < 2 * C) {
      float t = 0.f;
      for (int g = 0; g < NPG; ++g) t += red[g * 2 * C + tid];
      stats_partial[(size_t)blk * 2 * C + tid] = t;
    }
  }
}

// wgrad: partial[blk][ky*3+kx][co][ci] (+)= sum over the CTA's tiles of dy[p][co] * x[p+tap][ci].
// grid (nblk, 3): blockIdx.y = ky.  thread = (ci, group of CO_PT output channels).
template <typename T, int C>
__global__ void __launch_bounds__(256) k_wgrad(const T* __restrict__ x, const T* __restrict__ dy,
                                               float* __restrict__ partial, int ntiles, int first) {
  constexpr int CP = C + 1;
  constexpr int NCOG = 256 / C;
  constexpr int CO_PT = C / NCOG;
  extern __shared__ float sm[];
  float* xs = sm;                 // [3 rows... we keep 10][18][CP]
  float* ds = sm + 10 * 18 * CP;  // [128][C]
  const int tid = threadIdx.x, ky = blockIdx.y, ci = tid % C, cog = tid / C;
  float acc[3][CO_PT];
#pragma unroll
  for (int k = 0; k < 3; ++k)
#pragma unroll
    for (int e = 0; e < CO_PT; ++e) acc[k][e] = 0.f;
  for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int b = t >> 1, r0 = (t & 1) * 8;
    const T* xb = x + (size_t)b * NRD_PIX * C;
    __syncthreads();
    for (int i = tid; i < 10 * 18 * (C / 4); i += 256) {
      int c4 = i % (C / 4), pc = (i / (C / 4)) % 18, pr = i / (C / 4) / 18;
      int gr = r0 - 1 + pr, gc = pc - 1;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      if (gr >= 0 && gr < 16 && gc >= 0 && gc < 16) ld4(xb + ((size_t)gr * 16 + gc) * C + c4 * 4, v);
      float* d = xs + (pr * 18 + pc) * CP + c4 * 4;
      d[0] = v[0]; d[1] = v[1]; d[2] = v[2]; d[3] = v[3];
    }
    const T* db = dy + ((size_t)b * NRD_PIX + r0 * 16) * C;
    for (int i = tid; i < 128 * C / 4; i += 256) {
      float v[4]; ld4(db + (size_t)i * 4, v);
      *reinterpret_cast<float4*>(ds + i * 4) = make_float4(v[0], v[1], v[2], v[3]);
    }
    __syncthreads();
    for (int p = 0; p < 128; ++p) {
      const int r = p >> 4, c = p & 15;
      const float* xr = xs + ((r + ky) * 18 + c) * CP + ci;
      float x0 = xr[0], x1 = xr[CP], x2 = xr[2 * CP];
      const float* dp = ds + p * C + cog * CO_PT;
#pragma unroll
      for (int e4 = 0; e4 < CO_PT / 4; ++e4) {
        float4 d4 = *reinterpret_cast<const float4*>(dp + e4 * 4);
        float dv[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          acc[0][e4 * 4 + e] = fmaf(x0, dv[e], acc[0][e4 * 4 + e]);
          acc[1][e4 * 4 + e] = fmaf(x1, dv[e], acc[1][e4 * 4 + e]);
          acc[2][e4 * 4 + e] = fmaf(x2, dv[e], acc[2][e4 * 4 + e]);
        }
      }
    }
  }
  float* out = partial + (size_t)blockIdx.x * 9 * C * C;
#pragma unroll
  for (int kx = 0; kx < 3; ++kx)
#pragma unroll
    for (int e = 0; e < CO_PT; ++e) {
      size_t o = ((size_t)(ky * 3 + kx) * C + cog * CO_PT + e) * C + ci;
      out[o] = first ? acc[kx][e] : out[o] + acc[kx][e];
    }
}

// dW[co][ci][t] (OIHW) = sum_blk partial[blk][t][co][ci].  CTA = 32 outputs x 8 row classes, fixed-order combine.
__global__ void __launch_bounds__(256) k_wgrad_reduce(const float* __restrict__ partial, int nblk, int C, float* __restrict__ dw) {
  __shared__ double red[8][33];
  const int o = threadIdx.x & 31, r = threadIdx.x >> 5, i = blockIdx.x * 32 + o, n = 9 * C * C;
  double s = 0.0;
#pragma unroll 8
  for (int b = r; b < nblk; b += 8) s += (double)__ldcg(partial + (size_t)b * n + i);
  red[r][o] = s;
  __syncthreads();
  if (r == 0) {
    double tot = 0.0;
#pragma unroll
    for (int k = 0; k < 8; ++k) tot += red[k][o];
    int t = i / (C * C), co = (i / C) % C, ci = i % C;
    dw[((size_t)co * C + ci) * 9 + t] = (float)tot;
  }
}

// ------------------------------------------------------------------------------------------
// BatchNorm2d (training) forward pieces                                 (model/resnet.py:35-37)
// ------------------------------------------------------------------------------------------
// Column sums of a per-CTA partial matrix [rows][2C] (fp32) in fp64, fixed order, by one CTA of 256 threads: every
// thread owns one float4 column chunk and every RP-th row (all loads independent: one L2 round trip), then the RP row
// classes are combined in order.  On return dsum[0..2C) (shared) holds the sums.
template <int NT>
__device__ __forceinline__ void combine_partials(const float* __restrict__ partial, int rows, int C, double* dsum /*[NT*4]*/) {
  const int t = threadIdx.x, NC4 = 2 * C / 4, RP = NT / NC4, c4 = t % NC4, r = t / NC4;
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  for (int row0 = r; row0 < rows; row0 += 8 * RP) {       // batches of 8 independent, predicated loads (one L2 round trip each)
    float4 v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int row = row0 + u * RP;
      v[u] = (row < rows) ? __ldcg(reinterpret_cast<const float4*>(partial + (size_t)row * 2 * C) + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) { a0 += (double)v[u].x; a1 += (double)v[u].y; a2 += (double)v[u].z; a3 += (double)v[u].w; }
  }
  double* d = dsum + (size_t)r * 2 * C + c4 * 4;
  d[0] = a0; d[1] = a1; d[2] = a2; d[3] = a3;
  __syncthreads();
  double tot = 0.0;
  if (t < 2 * C) {
    for (int k = 0; k < RP; ++k) tot += dsum[(size_t)k * 2 * C + t];
  }
  __syncthreads();
  if (t < 2 * C) dsum[t] = tot;
  __syncthreads();
}

// finalize: partial [blocks][2C] -> mean, invstd (saved), scale/shift, running-stat EMA, nbt += 1
__global__ void __launch_bounds__(256) k_bn_finalize(const float* __restrict__ partial, int blocks, int count, int C,
                              const float* __restrict__ gamma, const float* __restrict__ beta,
                              float* __restrict__ saved, float* __restrict__ scale_shift,
                              float* __restrict__ running, int64_t* __restrict__ nbt) {
  __shared__ double dsum[1024];
  combine_partials<256>(partial, blocks, C, dsum);
  const int tid = threadIdx.x;
  if (tid < C) {
    const double S = dsum[tid], Q = dsum[C + tid];
    double mean = S / count, var = Q / count - mean * mean;
    if (var < 0.0) var = 0.0;
    float invstd = (float)(1.0 / sqrt(var + (double)NRD_BN_EPS));
    float sc = gamma[tid] * invstd;
    saved[tid] = (float)mean; saved[C + tid] = invstd;
    scale_shift[tid] = sc; scale_shift[C + tid] = beta[tid] - (float)mean * sc;
    if (running) {
      double unbiased = var * ((double)count / (double)(count - 1));
      running[tid] = (1.f - NRD_BN_MOMENTUM) * running[tid] + NRD_BN_MOMENTUM * (float)mean;
      running[C + tid] = (1.f - NRD_BN_MOMENTUM) * running[C + tid] + NRD_BN_MOMENTUM * (float)unbiased;
    }
  }
  if (tid == 0 && nbt) *nbt += 1;
}

// eval mode: scale/shift from the running statistics
__global__ void k_bn_eval_coeffs(const float* __restrict__ running, const float* __restrict__ gamma,
                                 const float* __restrict__ beta, int C, float* __restrict__ scale_shift) {
  int c = threadIdx.x;
  if (c < C) {
    float invstd = (float)(1.0 / sqrt((double)running[C + c] + (double)NRD_BN_EPS));
    float sc = gamma[c] * invstd;
    scale_shift[c] = sc; scale_shift[C + c] = beta[c] - running[c] * sc;
  }
}

// out = relu(y*scale + shift) + x
template <typename T>
__global__ void k_bn_relu_add(const T* __restrict__ y, const T* __restrict__ x, const float* __restrict__ scale_shift,
                              T* __restrict__ out, size_t n8, int C) {
  __shared__ float ss[128];
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) ss[i] = scale_shift[i];
  __syncthreads();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
    int c0 = (int)((i * 8) % C);
    float yv[8], xv[8], o[8];
    ld8(y + i * 8, yv); ld8(x + i * 8, xv);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = fmaxf(fmaf(yv[j], ss[c0 + j], ss[C + c0 + j]), 0.f) + xv[j];
    st8(out + i * 8, o);
  }
}

// ------------------------------------------------------------------------------------------
// BN + ReLU backward.  g = dL/d(out) (the residual branch passes g through unchanged).
//   z = y*scale+shift, dz = g*[z>0], xhat = (y-mean)*invstd
//   dbeta = sum dz, dgamma = sum dz*xhat, dy = gamma*invstd*(dz - dbeta/n - xhat*dgamma/n)
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_bn_bwd_reduce(const T* __restrict__ g, const T* __restrict__ y,
                                                       const float* __restrict__ scale_shift,
                                                       const float* __restrict__ saved, float* __restrict__ partial,
                                                       int C) {
  __shared__ float red[4096];  // [rows][2][C], rows = 256/(C/8) => rows*2*C = 4096 floats
  const int tid = threadIdx.x, ncg = C / 8, cg = tid % ncg, pr = tid / ncg, nrow = 256 / ncg;
  float sc[8], sh[8], mu[8], is[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = scale_shift[cg * 8 + j]; sh[j] = scale_shift[C + cg * 8 + j];
    mu[j] = saved[cg * 8 + j]; is[j] = saved[C + cg * 8 + j];
  }
  float s1[8], s2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s1[j] = s2[j] = 0.f;
  const size_t base = (size_t)blockIdx.x * NRD_PIX * C;   // one image per CTA
  for (int px = pr; px < NRD_PIX; px += nrow) {
    float gv[8], yv[8];
    ld8(g + base + (size_t)px * C + cg * 8, gv); ld8(y + base + (size_t)px * C + cg * 8, yv);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float z = fmaf(yv[j], sc[j], sh[j]);
      float dz = z > 0.f ? gv[j] : 0.f;
      s1[j] += dz; s2[j] = fmaf(dz, (yv[j] - mu[j]) * is[j], s2[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) { red[(pr * 2 + 0) * C + cg * 8 + j] = s1[j]; red[(pr * 2 + 1) * C + cg * 8 + j] = s2[j]; }
  __syncthreads();
  if (tid < 2 * C) {
    float t = 0.f;
    for (int r = 0; r < nrow; ++r) t += red[r * 2 * C + tid];
    partial[(size_t)blockIdx.x * 2 * C + tid] = t;
  }
}

// partial [blocks][2C] -> dbeta/dgamma (set or accumulate) + coefficients coef[3][C] = {gamma*invstd, dbeta/n, dgamma/n}
__global__ void __launch_bounds__(256) k_bn_bwd_finalize(const float* __restrict__ partial, int blocks, int count, int C,
                                  const float* __restrict__ gamma, const float* __restrict__ saved,
                                  float* __restrict__ dgamma, float* __restrict__ dbeta, int accumulate,
                                  float* __restrict__ coef) {
  __shared__ double dsum[1024];
  combine_partials<256>(partial, blocks, C, dsum);
  const int tid = threadIdx.x;
  if (tid < C) {
    const double db = dsum[tid], dg = dsum[C + tid];
    if (accumulate) { dbeta[tid] += (float)db; dgamma[tid] += (float)dg; }
    else { dbeta[tid] = (float)db; dgamma[tid] = (float)dg; }
    coef[tid] = gamma[tid] * saved[C + tid];
    coef[C + tid] = (float)(db / count);
    coef[2 * C + tid] = (float)(dg / count);
  }
}

template <typename T>
__global__ void k_bn_bwd_apply(const T* __restrict__ g, const T* __restrict__ y, const float* __restrict__ scale_shift,
                               const float* __restrict__ saved, const float* __restrict__ coef, T* __restrict__ dy,
                               size_t n8, int C) {
  __shared__ float p[7 * 64];
  for (int i = threadIdx.x; i < C; i += blockDim.x) {
    p[i] = scale_shift[i]; p[C + i] = scale_shift[C + i]; p[2 * C + i] = saved[i]; p[3 * C + i] = saved[C + i];
    p[4 * C + i] = coef[i]; p[5 * C + i] = coef[C + i]; p[6 * C + i] = coef[2 * C + i];
  }
  __syncthreads();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
    int c0 = (int)((i * 8) % C);
    float gv[8], yv[8], o[8];
    ld8(g + i * 8, gv); ld8(y + i * 8, yv);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int c = c0 + j;
      float z = fmaf(yv[j], p[c], p[C + c]);
      float dz = z > 0.f ? gv[j] : 0.f;
      float xh = (yv[j] - p[2 * C + c]) * p[3 * C + c];
      o[j] = p[4 * C + c] * (dz - p[5 * C + c] - xh * p[6 * C + c]);
    }
    st8(dy + i * 8, o);
  }
}

// Fused BN + ReLU backward (cooperative launch, grid <= #SMs, every CTA owns <= IPC whole images, the first IPC_S of which it
// keeps in shared memory between the two phases): phase 1 stages g and y and takes the per-CTA sums, a grid barrier publishes
// them, every CTA combines all partial rows in the same fixed order, phase 2 produces dy from shared memory -- and, for the
// images beyond IPC_S, from g and y read again (they were read microseconds earlier: L2 hits while 2T fits there).  One launch
// and 3T of HBM traffic instead of three launches and 5T.
template <typename T>
__global__ void __launch_bounds__(256, 1) k_bn_bwd_fused(const T* __restrict__ g, const T* __restrict__ y,
                                                         const float* __restrict__ scale_shift, const float* __restrict__ saved,
                                                         const float* __restrict__ gamma, T* __restrict__ dy,
                                                         float* __restrict__ dgamma, float* __restrict__ dbeta, int accumulate,
                                                         float* __restrict__ partial, unsigned long long* gbar, int B, int C, int ipc, int ipc_s) {
  extern __shared__ __align__(16) uint8_t smraw[];
  const int img_elems = NRD_PIX * C;
  T* gs = reinterpret_cast<T*>(smraw);
  T* ys = gs + (size_t)ipc_s * img_elems;
  double* dsum = reinterpret_cast<double*>(ys + (size_t)ipc_s * img_elems);   // [1024]
  float* red = reinterpret_cast<float*>(dsum + 1024);                       // [4096]
  float* coef = red + 4096;                                                 // [3C]
  const int tid = threadIdx.x, ncg = C / 8, cg = tid % ncg, pr = tid / ncg, nrow = 256 / ncg;
  float sc[8], sh[8], mu[8], is[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = scale_shift[cg * 8 + j]; sh[j] = scale_shift[C + cg * 8 + j];
    mu[j] = saved[cg * 8 + j]; is[j] = saved[C + cg * 8 + j];
  }
  unsigned long long gbar_target = 0;
  if (tid == 0) gbar_target = nrd_grid_barrier_base(gbar, gridDim.x) + gridDim.x;   // read before this CTA's own arrival
  float s1[8], s2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s1[j] = s2[j] = 0.f;
  const int chunks = img_elems / 8;          // 16-B (bf16) / 32-B (fp32) chunks per image; chunk % ncg == cg for this thread
  for (int i = 0; i < ipc; ++i) {
    const int b = blockIdx.x + i * gridDim.x;
    if (b >= B) break;
    const size_t base = (size_t)b * img_elems;
#pragma unroll 4
    for (int ch = tid; ch < chunks; ch += 256) {
      float gv[8], yv[8];
      ld8(g + base + (size_t)ch * 8, gv); ld8(y + base + (size_t)ch * 8, yv);
      if (i < ipc_s) { st8(gs + (size_t)i * img_elems + (size_t)ch * 8, gv); st8(ys + (size_t)i * img_elems + (size_t)ch * 8, yv); }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float z = fmaf(yv[j], sc[j], sh[j]);
        float dz = z > 0.f ? gv[j] : 0.f;
        s1[j] += dz; s2[j] = fmaf(dz, (yv[j] - mu[j]) * is[j], s2[j]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) { red[(pr * 2 + 0) * C + cg * 8 + j] = s1[j]; red[(pr * 2 + 1) * C + cg * 8 + j] = s2[j]; }
  __syncthreads();
  if (tid < 2 * C) {
    float t = 0.f;
    for (int r = 0; r < nrow; ++r) t += red[r * 2 * C + tid];
    partial[(size_t)blockIdx.x * 2 * C + tid] = t;
  }
  __syncthreads();
  if (tid == 0) nrd_grid_barrier(gbar, gbar_target);
  __syncthreads();
  combine_partials<256>(partial, gridDim.x, C, dsum);
  const int count = B * NRD_PIX;
  if (tid < C) {
    const double db = dsum[tid], dg = dsum[C + tid];
    if (blockIdx.x == 0) {
      if (accumulate) { dbeta[tid] += (float)db; dgamma[tid] += (float)dg; }
      else { dbeta[tid] = (float)db; dgamma[tid] = (float)dg; }
    }
    coef[tid] = gamma[tid] * saved[C + tid];
    coef[C + tid] = (float)(db / count);
    coef[2 * C + tid] = (float)(dg / count);
  }
  __syncthreads();
  float ca[8], cb[8], cc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { ca[j] = coef[cg * 8 + j]; cb[j] = coef[C + cg * 8 + j]; cc[j] = coef[2 * C + cg * 8 + j]; }
  for (int i = 0; i < ipc; ++i) {
    const int b = blockIdx.x + i * gridDim.x;
    if (b >= B) break;
    const size_t base = (size_t)b * img_elems;
#pragma unroll 4
    for (int ch = tid; ch < chunks; ch += 256) {
      float gv[8], yv[8], o[8];
      if (i < ipc_s) { ld8(gs + (size_t)i * img_elems + (size_t)ch * 8, gv); ld8(ys + (size_t)i * img_elems + (size_t)ch * 8, yv); }
      else { ld8(g + base + (size_t)ch * 8, gv); ld8(y + base + (size_t)ch * 8, yv); }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float z = fmaf(yv[j], sc[j], sh[j]);
        float dz = z > 0.f ? gv[j] : 0.f;
        float xh = (yv[j] - mu[j]) * is[j];
        o[j] = ca[j] * (dz - cb[j] - xh * cc[j]);
      }
      st8(dy + base + (size_t)ch * 8, o);
    }
  }
}

// ------------------------------------------------------------------------------------------
// 2x2/2 max pool at 16x16 -> 8x8 (model/resnet.py:18); gradient goes to the FIRST max
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void k_maxpool_fwd(const T* __restrict__ x, T* __restrict__ p, uint8_t* __restrict__ idx, int B, int C) {
  const int ncg = C / 8;
  size_t n = (size_t)B * 64 * ncg;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    int cg = (int)(i % ncg); size_t pp = i / ncg; int pw = (int)(pp % 8), ph = (int)((pp / 8) % 8); size_t b = pp / 64;
    float m[8]; int mi[8];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float v[8];
      ld8(x + ((b * 16 + 2 * ph + (q >> 1)) * 16 + 2 * pw + (q & 1)) * C + cg * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) if (q == 0 || v[j] > m[j]) { m[j] = v[j]; mi[j] = q; }
    }
    st8(p + pp * C + cg * 8, m);
    uint2 pk; uint8_t* pb = reinterpret_cast<uint8_t*>(&pk);
#pragma unroll
    for (int j = 0; j < 8; ++j) pb[j] = (uint8_t)mi[j];
    *reinterpret_cast<uint2*>(idx + pp * C + cg * 8) = pk;
  }
}

template <typename T>
__global__ void k_maxpool_bwd(const T* __restrict__ gp, const uint8_t* __restrict__ idx, T* __restrict__ gx, int B, int C) {
  const int ncg = C / 8;
  size_t n = (size_t)B * 64 * ncg;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    int cg = (int)(i % ncg); size_t pp = i / ncg; int pw = (int)(pp % 8), ph = (int)((pp / 8) % 8); size_t b = pp / 64;
    float g[8]; ld8(gp + pp * C + cg * 8, g);
    uint2 pk = *reinterpret_cast<const uint2*>(idx + pp * C + cg * 8);
    const uint8_t* pb = reinterpret_cast<const uint8_t*>(&pk);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = (pb[j] == q) ? g[j] : 0.f;
      st8(gx + ((b * 16 + 2 * ph + (q >> 1)) * 16 + 2 * pw + (q & 1)) * C + cg * 8, v);
    }
  }
}

// ------------------------------------------------------------------------------------------
// head: fc1 + relu, fc2, softmax cross-entropy                     (model/resnet.py:19-21, main.py:28)
// ------------------------------------------------------------------------------------------
// softmax cross-entropy, forward and gradient, one CTA: per-row loss (kept in loss_rows) and dlogits =
// (softmax - onehot) * grad_scale / B, then the mean loss by a fixed-shape tree over the CTA (deterministic).
__global__ void __launch_bounds__(256) k_ce(const float* __restrict__ logits, const int64_t* __restrict__ labels, int B, float gscale,
                                            float* __restrict__ loss_rows, float* __restrict__ dlogits, float* __restrict__ loss) {
  __shared__ double red[256];
  double acc = 0.0;
  for (int b = threadIdx.x; b < B; b += 256) {
    float v[NRD_NCLS], m = -INFINITY;
#pragma unroll
    for (int c = 0; c < NRD_NCLS; ++c) { v[c] = logits[b * NRD_NCLS + c]; m = fmaxf(m, v[c]); }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < NRD_NCLS; ++c) { v[c] = expf(v[c] - m); s += v[c]; }
    const long long lab64 = labels[b];
    const bool lab_ok = lab64 >= 0 && lab64 < NRD_NCLS;            // an out-of-range label poisons the loss, never the address
    const int lab = lab_ok ? (int)lab64 : 0;
    const float inv = 1.f / s, sc = gscale / (float)B;
    const float l = lab_ok ? logf(s) + m - logits[b * NRD_NCLS + lab] : NAN;
    loss_rows[b] = l;
    acc += (double)l;
    if (dlogits) {
#pragma unroll
      for (int c = 0; c < NRD_NCLS; ++c) dlogits[b * NRD_NCLS + c] = lab_ok ? (v[c] * inv - (c == lab ? 1.f : 0.f)) * sc : NAN;
    }
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0 && loss) *loss = (float)(red[0] / B);
}

// dh1[b][j] = sum_c dlogits[b][c] * w2[c][j] * [h1[b][j] > 0] is 10 MACs per element: every consumer below recomputes it
// while staging its rows instead of reading it from a kernel of its own, which also makes the three consumers independent
// of one another (head weight gradients can run beside the body backward).
__device__ __forceinline__ float dh1_elem(const float* __restrict__ dl_row /*[10]*/, const float* __restrict__ w2s /*[10][32]*/, int j, float h) {
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < NRD_NCLS; ++c) s = fmaf(dl_row[c], w2s[c * NRD_FC1_OUT + j], s);
  return h > 0.f ? s : 0.f;
}

// dW2[c][j], db2[c], dbfc1[j].  grid = 12 CTAs of 256 threads: CTA c < 10 -> dW2[c][:], CTA 10 -> db2, CTA 11 -> dbfc1; thread =
// (output, one of 8 batch-row classes); every class sums its rows in order and the 8 classes are combined in order.
__global__ void __launch_bounds__(256) k_head_wgrad(const float* __restrict__ dlogits, const float* __restrict__ h1,
                             const float* __restrict__ w2, int B, float* __restrict__ dw2,
                             float* __restrict__ db2, float* __restrict__ dbfc1) {
  __shared__ float red[8][33];
  __shared__ float w2s[NRD_NCLS * NRD_FC1_OUT];
  const int t = threadIdx.x, o = t & 31, rg = t >> 5, role = blockIdx.x;
  if (role == 11) { for (int i = t; i < NRD_NCLS * NRD_FC1_OUT; i += 256) w2s[i] = w2[i]; __syncthreads(); }
  float s = 0.f;
  if (role < 10) {
#pragma unroll 4
    for (int b = rg; b < B; b += 8) s = fmaf(dlogits[(size_t)b * NRD_NCLS + role], h1[(size_t)b * NRD_FC1_OUT + o], s);
  } else if (role == 10) {
    if (o < NRD_NCLS)
#pragma unroll 4
      for (int b = rg; b < B; b += 8) s += dlogits[(size_t)b * NRD_NCLS + o];
  } else {
#pragma unroll 2
    for (int b = rg; b < B; b += 8) s += dh1_elem(dlogits + (size_t)b * NRD_NCLS, w2s, o, h1[(size_t)b * NRD_FC1_OUT + o]);
  }
  red[rg][o] = s;
  __syncthreads();
  if (rg == 0) {
    float tot = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) tot += red[k][o];
    if (role < 10) dw2[role * NRD_FC1_OUT + o] = tot;
    else if (role == 10) { if (o < NRD_NCLS) db2[o] = tot; }
    else dbfc1[o] = tot;
  }
}

// dWfc1[j][c*64+hw] = sum_b dh1[b][j] * p[b][hw*C+c].  CTA (x, y) = 64 consecutive NHWC columns x 32 rows over the batch rows
// b == y*64 .. (64-row chunks strided by gridDim.y); partial[y][j][k] summed in order by k_fc1_wgrad_reduce.
template <typename T>
__global__ void __launch_bounds__(256) k_fc1_wgrad(const T* __restrict__ p, const float* __restrict__ dlogits, const float* __restrict__ h1,
                                                   const float* __restrict__ w2, float* __restrict__ partial, int B, int C) {
  __shared__ __align__(16) float ds[64 * NRD_FC1_OUT];
  __shared__ float ps[64 * 64];
  __shared__ float dl[64 * NRD_NCLS], w2s[NRD_NCLS * NRD_FC1_OUT];
  const int tid = threadIdx.x, il = tid % 64, jg = tid / 64, K = 64 * C, i0 = blockIdx.x * 64;
  for (int i = tid; i < NRD_NCLS * NRD_FC1_OUT; i += 256) w2s[i] = w2[i];
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  for (int bb = blockIdx.y * 64; bb < B; bb += gridDim.y * 64) {
    const int nb = min(64, B - bb);
    __syncthreads();
    for (int q = tid; q < nb * NRD_NCLS; q += 256) dl[q] = dlogits[(size_t)bb * NRD_NCLS + q];
    for (int q = tid; q < nb * 16; q += 256) {          // 64 rows x 16 four-element chunks
      int r = q >> 4, c4 = q & 15;
      float v[4]; ld4(p + (size_t)(bb + r) * K + i0 + c4 * 4, v);
      float* d = ps + r * 64 + c4 * 4;
      d[0] = v[0]; d[1] = v[1]; d[2] = v[2]; d[3] = v[3];
    }
    __syncthreads();
    for (int q = tid; q < nb * NRD_FC1_OUT; q += 256)
      ds[q] = dh1_elem(dl + (q >> 5) * NRD_NCLS, w2s, q & 31, h1[(size_t)bb * NRD_FC1_OUT + q]);
    __syncthreads();
#pragma unroll 4
    for (int b = 0; b < nb; ++b) {
      const float pv = ps[b * 64 + il];
      const float4* d4 = reinterpret_cast<const float4*>(ds + b * NRD_FC1_OUT + jg * 8);
      const float4 a = d4[0], c4 = d4[1];
      acc[0] = fmaf(pv, a.x, acc[0]); acc[1] = fmaf(pv, a.y, acc[1]); acc[2] = fmaf(pv, a.z, acc[2]); acc[3] = fmaf(pv, a.w, acc[3]);
      acc[4] = fmaf(pv, c4.x, acc[4]); acc[5] = fmaf(pv, c4.y, acc[5]); acc[6] = fmaf(pv, c4.z, acc[6]); acc[7] = fmaf(pv, c4.w, acc[7]);
    }
  }
  float* out = partial + (size_t)blockIdx.y * NRD_FC1_OUT * K;
#pragma unroll
  for (int e = 0; e < 8; ++e) out[(size_t)(jg * 8 + e) * K + i0 + il] = acc[e];     // [split][j][NHWC column]: coalesced
}

// dWfc1[j][c*64+hw] = sum_split partial[split][j][hw*C+c]
__global__ void k_fc1_wgrad_reduce(const float* __restrict__ partial, int nsplit, int C, float* __restrict__ dw1) {
  const int K = 64 * C, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= NRD_FC1_OUT * K) return;
  float s = 0.f;
  for (int q = 0; q < nsplit; ++q) s += partial[(size_t)q * NRD_FC1_OUT * K + i];
  const int j = i / K, k = i % K, hw = k / C, c = k % C;
  dw1[(size_t)j * K + c * 64 + hw] = s;
}

// d(pooled)[b][hw][c] = sum_j dh1[b][j] * w1t[hw*C+c][j], scattered through pool-2's argmax into gx [B,16,16,C].
// CTA = 8 batch rows x 256 columns: each weight row is read once per 8 rows.
template <typename T>
__global__ void __launch_bounds__(256) k_fc1_dgrad_unpool(const float* __restrict__ dlogits, const float* __restrict__ h1,
                                                          const float* __restrict__ w2, const float* __restrict__ w1t,
                                                          const uint8_t* __restrict__ idx2, T* __restrict__ gx,
                                                          int B, int C) {
  __shared__ float d[8][NRD_FC1_OUT];
  __shared__ float dl[8 * NRD_NCLS], w2s[NRD_NCLS * NRD_FC1_OUT];
  const int K = 64 * C, per_b = K / 256, b0 = (blockIdx.x / per_b) * 8, i = (blockIdx.x % per_b) * 256 + threadIdx.x;
  nrd_pdl_launch();
  for (int q = threadIdx.x; q < NRD_NCLS * NRD_FC1_OUT; q += 256) w2s[q] = w2[q];
  nrd_pdl_wait();
  if (threadIdx.x < 8 * NRD_NCLS) {
    const int r = threadIdx.x / NRD_NCLS;
    dl[threadIdx.x] = (b0 + r < B) ? dlogits[(size_t)b0 * NRD_NCLS + threadIdx.x] : 0.f;
  }
  __syncthreads();
  {
    const int r = threadIdx.x >> 5, j = threadIdx.x & 31;
    d[r][j] = (b0 + r < B) ? dh1_elem(dl + r * NRD_NCLS, w2s, j, h1[(size_t)(b0 + r) * NRD_FC1_OUT + j]) : 0.f;
  }
  __syncthreads();
  float s[8];
#pragma unroll
  for (int r = 0; r < 8; ++r) s[r] = 0.f;
  const float4* wrow = reinterpret_cast<const float4*>(w1t + (size_t)i * NRD_FC1_OUT);
#pragma unroll
  for (int j4 = 0; j4 < 8; ++j4) {
    const float4 w = wrow[j4];
#pragma unroll
    for (int r = 0; r < 8; ++r)
      s[r] += d[r][j4 * 4] * w.x + d[r][j4 * 4 + 1] * w.y + d[r][j4 * 4 + 2] * w.z + d[r][j4 * 4 + 3] * w.w;
  }
  const int hw = i / C, c = i % C, ph = hw / 8, pw = hw % 8;
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    const int b = b0 + r;
    if (b < B) {
      const int id = idx2[(size_t)b * K + i];
#pragma unroll
      for (int q = 0; q < 4; ++q)
        gx[(((size_t)b * 16 + 2 * ph + (q >> 1)) * 16 + 2 * pw + (q & 1)) * C + c] = from_f<T>(q == id ? s[r] : 0.f);
    }
  }
}

// The same without the scatter: d(pooled) [B,8,8,C] only.  The persistent backward expands it through pool-2's argmax while it
// loads its gradient registers, so the 4x larger (and 3/4 zero) dL/dx_NB never exists in memory.
template <typename T>
__global__ void __launch_bounds__(256) k_fc1_dgrad_pooled(const float* __restrict__ dlogits, const float* __restrict__ h1,
                                                          const float* __restrict__ w2, const float* __restrict__ w1t,
                                                          T* __restrict__ dp, int B, int C) {
  __shared__ float d[8][NRD_FC1_OUT];
  __shared__ float dl[8 * NRD_NCLS], w2s[NRD_NCLS * NRD_FC1_OUT];
  const int K = 64 * C, per_b = K / 256, b0 = (blockIdx.x / per_b) * 8, i = (blockIdx.x % per_b) * 256 + threadIdx.x;
  nrd_pdl_launch();
  for (int q = threadIdx.x; q < NRD_NCLS * NRD_FC1_OUT; q += 256) w2s[q] = w2[q];
  nrd_pdl_wait();
  if (threadIdx.x < 8 * NRD_NCLS) {
    const int r = threadIdx.x / NRD_NCLS;
    dl[threadIdx.x] = (b0 + r < B) ? dlogits[(size_t)b0 * NRD_NCLS + threadIdx.x] : 0.f;
  }
  __syncthreads();
  {
    const int r = threadIdx.x >> 5, j = threadIdx.x & 31;
    d[r][j] = (b0 + r < B) ? dh1_elem(dl + r * NRD_NCLS, w2s, j, h1[(size_t)(b0 + r) * NRD_FC1_OUT + j]) : 0.f;
  }
  __syncthreads();
  float s[8];
#pragma unroll
  for (int r = 0; r < 8; ++r) s[r] = 0.f;
  const float4* wrow = reinterpret_cast<const float4*>(w1t + (size_t)i * NRD_FC1_OUT);
  float su[8];
#pragma unroll
  for (int r = 0; r < 8; ++r) su[r] = 0.f;
#pragma unroll
  for (int j4 = 0; j4 < 8; ++j4) {
    const float4 w = wrow[j4];
#pragma unroll
    for (int r = 0; r < 8; ++r) {      // k_head_fused's order (bit-identical): one fma chain over j = 0..15, one over 16..31, then lower + upper
      float& a = j4 < 4 ? s[r] : su[r];
      a = fmaf(d[r][j4 * 4], w.x, a); a = fmaf(d[r][j4 * 4 + 1], w.y, a);
      a = fmaf(d[r][j4 * 4 + 2], w.z, a); a = fmaf(d[r][j4 * 4 + 3], w.w, a);
    }
  }
#pragma unroll
  for (int r = 0; r < 8; ++r)
    if (b0 + r < B) dp[(size_t)(b0 + r) * K + i] = from_f<T>(s[r] + su[r]);
}

// ------------------------------------------------------------------------------------------
// The whole head in ONE cooperative launch (model/resnet.py:18-21, main.py:35 and the first products of main.py:38):
//   max-pool 2x2 -> fc1 (split-K over the CTAs of a row group) -> +bias, ReLU -> fc2 -> softmax cross-entropy, dlogits
//   -> dh1 -> dL/d(pooled) (each CTA its own K slice)
// CTA (x, ks) serves row groups rg = x, x + gridDim.x, ... (32 batch rows each) for the K slice [ks*128, ks*128+128) of the
// pooled NHWC vector; its 16 KB slice of the fc1 weight is fetched ONCE and serves both the forward partial product and
// the backward slice.  The only cross-CTA dependency -- fc1's split-K sum -- is a counter per row group: its KS CTAs arrive
// after storing their partials, wait for one another (cooperative launch: all CTAs are co-resident), and then EACH redoes the
// tiny tail of its 32 rows (sum of the KS partials in slice order, fc2, softmax: ~1.4 k MACs per row) instead of waiting for
// a second launch or for one CTA to publish it.  The ks == 0 CTAs publish h1 / logits / loss rows / dlogits, and the last of
// them sums the loss rows in k_ce's fixed order.  Everything is fixed-order: bit-reproducible.
// ------------------------------------------------------------------------------------------
#define NRD_HEAD_THREADS 512
// shared memory (floats): weight slice k-major [128][32] | the same transposed [32][132] | pooled slice, k-major [128][33] |
// dL/dh1, j-major [32][32] | cross-thread partial sums [8 k-groups | 2 j-halves][...] (32 KB) | fc2 weight + bias [320 + 16]
#define NRD_HEAD_SMEM_FLOATS (128 * 32 + 32 * 132 + 128 * 33 + 32 * 32 + 8 * 64 * 16 + 336)
static constexpr size_t head_smem_bytes() { return (size_t)NRD_HEAD_SMEM_FLOATS * 4 + 256 * 8 + 16; }

// 512 threads.  The two GEMM-shaped phases are register-tiled 4 x 4 (16 FMAs per 2 shared-memory reads; the first version's
// 1 x 4 tiles were bound by shared-memory bandwidth, ~3 us per phase) with the reduction dimension split over thread groups
// and combined in a fixed order; the tail of a row (fc2, softmax, dh1) runs in the registers of its 8 consecutive lanes.
template <typename T>
__global__ void __launch_bounds__(NRD_HEAD_THREADS, 1)
k_head_fused(const T* __restrict__ x, T* __restrict__ pooled, uint8_t* __restrict__ idx2, int pooled_in,
             const float* __restrict__ w1t, const float* __restrict__ b1, const float* __restrict__ w2, const float* __restrict__ b2,
             float* __restrict__ part, float* __restrict__ h1, float* __restrict__ logits, int B, int C,
             const int64_t* __restrict__ labels, float gscale, float* __restrict__ loss_rows, float* __restrict__ dlogits,
             float* __restrict__ loss, unsigned long long* __restrict__ rg_ctr, unsigned int* __restrict__ counter,
             T* __restrict__ dp, long long* __restrict__ dbg, volatile float* loss_mirror, volatile unsigned int* seq_mirror,
             unsigned int* __restrict__ seq_dev) {
  extern __shared__ __align__(16) float hsm[];
  float* ws = hsm;                    // [128 k][32 j]      fc1 weight slice, k-major (forward partial product)
  float* wsT = ws + 128 * 32;         // [32 j][132]        the same slice transposed (backward slice), rows padded
  float* psT = wsT + 32 * 132;        // [128 k][33]        pooled slice of the current row group, k-major
  float* ddT = psT + 128 * 33;        // [32 j][32 rows]    dL/dh1 of the row group
  float* xr = ddT + 32 * 32;          // [8][64 tiles][16]  partial sums of the k-groups (forward) / [256 tiles][16] upper j half (backward)
  float* w2s = xr + 8 * 64 * 16;      // [10][32] fc2 weight, then [10] fc2 bias
  float* b2s = w2s + 320;
  double* red = reinterpret_cast<double*>((reinterpret_cast<uintptr_t>(b2s + 16) + 15) & ~(uintptr_t)15);   // [256]
  __shared__ int last;
  const int tid = threadIdx.x, lane = tid & 31, K = 64 * C, KS = gridDim.y, ks = blockIdx.y, k0 = ks * 128, nrg = (B + 31) / 32;
  const int kh = tid >> 8, r = (tid >> 3) & 31, jq = tid & 7;      // tail role: (copy kh of) row r, outputs jq*4 .. +3
  // stamps: rows 0..3 = CTAs (0,0), (0,KS-1), (last,0), (last,KS-1)
  const int drow = (blockIdx.x == 0 ? 0 : (blockIdx.x == gridDim.x - 1 ? 2 : -1)) + (blockIdx.y == 0 ? 0 : (blockIdx.y == gridDim.y - 1 ? 1 : -100));
  long long* dbgt = (dbg && tid == 0 && drow >= 0) ? dbg : nullptr;
  nrd_dbg_stamp(dbgt, drow, 0);
  // ---- everything that does not depend on the row group is requested first (weights are untouched since the last optimizer
  // ---- step): this CTA's fc1 weight slice, fc1 bias, fc2 weight and bias ----
  float4 wreg[2];
#pragma unroll
  for (int it = 0; it < 2; ++it) wreg[it] = reinterpret_cast<const float4*>(w1t + (size_t)k0 * 32)[tid + it * NRD_HEAD_THREADS];
  const float4 bb4 = *reinterpret_cast<const float4*>(b1 + jq * 4);
  for (int q = tid; q < NRD_NCLS * NRD_FC1_OUT + NRD_NCLS; q += NRD_HEAD_THREADS) w2s[q] = q < NRD_NCLS * NRD_FC1_OUT ? w2[q] : b2[q - NRD_NCLS * NRD_FC1_OUT];
  bool w_stored = false;
  for (int rg = blockIdx.x; rg < nrg; rg += gridDim.x) {
    const int b0 = rg * 32, b = b0 + r;
    nrd_dbg_stamp(dbgt, drow, 1);
    long long lab64 = 0;
    if (labels != nullptr && b < B) lab64 = labels[b];             // needed only after the barrier: in flight until then
    // ---- pool: 32 rows x 16 eight-channel chunks of the slice, one per thread; all four window loads in flight together ----
    {
      const int rr = tid >> 4, c8 = tid & 15, bb = b0 + rr, kk = k0 + c8 * 8;
      float v[4][8];
      if (bb < B) {
        if (pooled_in) ld8(pooled + (size_t)bb * K + kk, v[0]);
        else {
          const int hw = kk / C, c = kk % C, ph_ = hw >> 3, pw = hw & 7;
#pragma unroll
          for (int q = 0; q < 4; ++q) ld8(x + (((size_t)bb * 16 + 2 * ph_ + (q >> 1)) * 16 + 2 * pw + (q & 1)) * C + c, v[q]);
        }
      }
      __syncthreads();                                 // previous row group's readers of psT / ddT / xr are done
      if (!w_stored) {
#pragma unroll
        for (int it = 0; it < 2; ++it) {
          const int i = tid + it * NRD_HEAD_THREADS, k = i >> 3, j = (i & 7) * 4;
          const float4 w = wreg[it];
          reinterpret_cast<float4*>(ws)[i] = w;
          wsT[(j + 0) * 132 + k] = w.x; wsT[(j + 1) * 132 + k] = w.y; wsT[(j + 2) * 132 + k] = w.z; wsT[(j + 3) * 132 + k] = w.w;
        }
        w_stored = true;
      }
      float m[8]; int mi[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) { m[j] = 0.f; mi[j] = 0; }
      if (bb < B) {
        if (pooled_in) {
#pragma unroll
          for (int j = 0; j < 8; ++j) m[j] = v[0][j];
        } else {
#pragma unroll
          for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int j = 0; j < 8; ++j) if (q == 0 || v[q][j] > m[j]) { m[j] = v[q][j]; mi[j] = q; }     // first max wins
          st8(pooled + (size_t)bb * K + kk, m);
          uint2 pk; uint8_t* pb = reinterpret_cast<uint8_t*>(&pk);
#pragma unroll
          for (int j = 0; j < 8; ++j) pb[j] = (uint8_t)mi[j];
          *reinterpret_cast<uint2*>(idx2 + (size_t)bb * K + kk) = pk;
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) psT[(c8 * 8 + j) * 33 + rr] = m[j];
    }
    __syncthreads();
    nrd_dbg_stamp(dbgt, drow, 2);
    // ---- fc1 partial product: thread = (k-group kg of 16 k, tile of 4 rows x 4 outputs); the 8 k-groups are combined in order ----
    {
      const int kg = tid >> 6, tile = tid & 63, rt = tile >> 3, jt = tile & 7;
      float acc[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[i][e] = 0.f;
      const float* pk_ = psT + (kg * 16) * 33 + rt * 4;
      const float* wk_ = ws + (kg * 16) * 32 + jt * 4;
#pragma unroll 4
      for (int k = 0; k < 16; ++k) {
        const float4 w = *reinterpret_cast<const float4*>(wk_ + k * 32);
        const float p0 = pk_[k * 33], p1 = pk_[k * 33 + 1], p2 = pk_[k * 33 + 2], p3 = pk_[k * 33 + 3];
        acc[0][0] = fmaf(p0, w.x, acc[0][0]); acc[0][1] = fmaf(p0, w.y, acc[0][1]); acc[0][2] = fmaf(p0, w.z, acc[0][2]); acc[0][3] = fmaf(p0, w.w, acc[0][3]);
        acc[1][0] = fmaf(p1, w.x, acc[1][0]); acc[1][1] = fmaf(p1, w.y, acc[1][1]); acc[1][2] = fmaf(p1, w.z, acc[1][2]); acc[1][3] = fmaf(p1, w.w, acc[1][3]);
        acc[2][0] = fmaf(p2, w.x, acc[2][0]); acc[2][1] = fmaf(p2, w.y, acc[2][1]); acc[2][2] = fmaf(p2, w.z, acc[2][2]); acc[2][3] = fmaf(p2, w.w, acc[2][3]);
        acc[3][0] = fmaf(p3, w.x, acc[3][0]); acc[3][1] = fmaf(p3, w.y, acc[3][1]); acc[3][2] = fmaf(p3, w.z, acc[3][2]); acc[3][3] = fmaf(p3, w.w, acc[3][3]);
      }
      float4* dst = reinterpret_cast<float4*>(xr + (kg * 64 + tile) * 16);
#pragma unroll
      for (int i = 0; i < 4; ++i) dst[i] = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
      __syncthreads();
      // thread -> 2 outputs of one tile: (row rt*4 + i, outputs jt*4 + e, e = 2*(tid&1) .. +1), k-groups in order
      const int tl = tid >> 3, i2 = (tid >> 1) & 3, e2 = (tid & 1) * 2;
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int g = 0; g < 8; ++g) { const float2 t2 = *reinterpret_cast<const float2*>(xr + (g * 64 + tl) * 16 + i2 * 4 + e2); s0 += t2.x; s1 += t2.y; }
      const int orow = b0 + (tl >> 3) * 4 + i2, oj = (tl & 7) * 4 + e2;
      if (orow < B) *reinterpret_cast<float2*>(part + ((size_t)ks * B + orow) * NRD_FC1_OUT + oj) = make_float2(s0, s1);
    }
    // ---- the KS CTAs of this row group wait for one another (monotonic counter per row group, nrd_common.cuh).  The CTA barrier
    // ---- orders every thread's partial store before thread 0's release-add, which is cumulative at gpu scope: the grid-sync
    // ---- pattern, one release per CTA ----
    __syncthreads();
    nrd_dbg_stamp(dbgt, drow, 3);
    if (tid == 0) {
      const unsigned long long base = nrd_grid_barrier_base(rg_ctr + rg, (unsigned)KS);
      nrd_grid_barrier(rg_ctr + rg, base + (unsigned)KS);
    }
    __syncthreads();
    nrd_dbg_stamp(dbgt, drow, 4);
    // ---- tail of the 32 rows, in registers, redone by every CTA of the group (threads 0..255: the other half would only double
    // ---- the L2 traffic of the partials): thread (r, jq) holds h1[r][jq*4 .. +3]; the row's 8 threads are 8 consecutive lanes ----
    float h[4] = {0.f, 0.f, 0.f, 0.f};
    if (kh == 0 && b < B) {
      float4 hv[16];
#pragma unroll 1
      for (int s0 = 0; s0 < KS; s0 += 16) {            // KS = 16 (C = 32) or 32 (C = 64): all loads of a batch in flight together
#pragma unroll
        for (int s2 = 0; s2 < 16; ++s2) hv[s2] = __ldcg(reinterpret_cast<const float4*>(part + ((size_t)(s0 + s2) * B + b) * NRD_FC1_OUT) + jq);
#pragma unroll
        for (int s2 = 0; s2 < 16; ++s2) { h[0] += hv[s2].x; h[1] += hv[s2].y; h[2] += hv[s2].z; h[3] += hv[s2].w; }     // slice order: fixed
      }
      h[0] = fmaxf(h[0] + bb4.x, 0.f); h[1] = fmaxf(h[1] + bb4.y, 0.f); h[2] = fmaxf(h[2] + bb4.z, 0.f); h[3] = fmaxf(h[3] + bb4.w, 0.f);
      if (ks == 0 && kh == 0) *reinterpret_cast<float4*>(h1 + (size_t)b * NRD_FC1_OUT + jq * 4) = make_float4(h[0], h[1], h[2], h[3]);
    }
    // logits: each thread's 4-term partial of the 32-term products, butterfly over the row's 8 lanes (xor 1, 2, 4: a fixed tree)
    float lg[NRD_NCLS];
#pragma unroll
    for (int c = 0; c < NRD_NCLS; ++c) lg[c] = 0.f;
    if (kh == 0)                                       // (warp-uniform: warps 0..7)
#pragma unroll
    for (int c = 0; c < NRD_NCLS; ++c) {
      const float4 w = *reinterpret_cast<const float4*>(w2s + c * NRD_FC1_OUT + jq * 4);
      float t = fmaf(h[3], w.w, fmaf(h[2], w.z, fmaf(h[1], w.y, h[0] * w.x)));
      t += __shfl_xor_sync(0xffffffffu, t, 1); t += __shfl_xor_sync(0xffffffffu, t, 2); t += __shfl_xor_sync(0xffffffffu, t, 4);
      lg[c] = t + b2s[c];
    }
    const bool pub = ks == 0 && kh == 0 && jq == 0 && b < B;      // one thread per row publishes
    if (pub) {
#pragma unroll
      for (int c = 0; c < NRD_NCLS; ++c) logits[(size_t)b * NRD_NCLS + c] = lg[c];
    }
    if (labels == nullptr) continue;                   // forward only (uniform over the grid)
    nrd_dbg_stamp(dbgt, drow, 5);
    // softmax cross-entropy of the row, by each of its threads (same arithmetic, same order as k_ce)
    float dlr[NRD_NCLS];
#pragma unroll
    for (int c = 0; c < NRD_NCLS; ++c) dlr[c] = 0.f;
    if (kh == 0 && b < B) {
      float v[NRD_NCLS], m = -INFINITY;
#pragma unroll
      for (int c = 0; c < NRD_NCLS; ++c) { v[c] = lg[c]; m = fmaxf(m, v[c]); }
      float s2 = 0.f;
#pragma unroll
      for (int c = 0; c < NRD_NCLS; ++c) { v[c] = expf(v[c] - m); s2 += v[c]; }
      const bool lab_ok = lab64 >= 0 && lab64 < NRD_NCLS;          // an out-of-range label poisons the loss, never the address
      const int lab = lab_ok ? (int)lab64 : 0;
      const float inv = 1.f / s2, sc = gscale / (float)B;
      float llab = lg[0];
#pragma unroll
      for (int c = 1; c < NRD_NCLS; ++c) llab = (c == lab) ? lg[c] : llab;
#pragma unroll
      for (int c = 0; c < NRD_NCLS; ++c) dlr[c] = lab_ok ? (v[c] * inv - (c == lab ? 1.f : 0.f)) * sc : NAN;
      if (pub) {
        loss_rows[b] = lab_ok ? logf(s2) + m - llab : NAN;
#pragma unroll
        for (int c = 0; c < NRD_NCLS; ++c) dlogits[(size_t)b * NRD_NCLS + c] = dlr[c];
      }
    }
    if (dp != nullptr) {
      // dL/dh1 of this thread's 4 outputs (c ascending, one fma chain: dh1_elem's order) -> ddT[j][row]
      if (kh == 0) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float sacc = 0.f;
#pragma unroll
          for (int c = 0; c < NRD_NCLS; ++c) sacc = fmaf(dlr[c], w2s[c * NRD_FC1_OUT + jq * 4 + e], sacc);
          ddT[(jq * 4 + e) * 32 + r] = (b < B && h[e] > 0.f) ? sacc : 0.f;
        }
      }
      __syncthreads();
      nrd_dbg_stamp(dbgt, drow, 6);
      // ---- dL/d(pooled) for this slice: thread = (j half jh, tile of 4 rows x 4 columns); per output one fma chain over its 16 j,
      // ---- ascending; lower half + upper half, in that order ----
      const int jh = tid >> 8, tile = tid & 255, rt = tile >> 5, kq = tile & 31;
      float acc[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[i][e] = 0.f;
#pragma unroll 4
      for (int jj = 0; jj < 16; ++jj) {
        const int j = jh * 16 + jj;
        const float4 w = *reinterpret_cast<const float4*>(wsT + j * 132 + kq * 4);
        const float4 d = *reinterpret_cast<const float4*>(ddT + j * 32 + rt * 4);
        acc[0][0] = fmaf(d.x, w.x, acc[0][0]); acc[0][1] = fmaf(d.x, w.y, acc[0][1]); acc[0][2] = fmaf(d.x, w.z, acc[0][2]); acc[0][3] = fmaf(d.x, w.w, acc[0][3]);
        acc[1][0] = fmaf(d.y, w.x, acc[1][0]); acc[1][1] = fmaf(d.y, w.y, acc[1][1]); acc[1][2] = fmaf(d.y, w.z, acc[1][2]); acc[1][3] = fmaf(d.y, w.w, acc[1][3]);
        acc[2][0] = fmaf(d.z, w.x, acc[2][0]); acc[2][1] = fmaf(d.z, w.y, acc[2][1]); acc[2][2] = fmaf(d.z, w.z, acc[2][2]); acc[2][3] = fmaf(d.z, w.w, acc[2][3]);
        acc[3][0] = fmaf(d.w, w.x, acc[3][0]); acc[3][1] = fmaf(d.w, w.y, acc[3][1]); acc[3][2] = fmaf(d.w, w.z, acc[3][2]); acc[3][3] = fmaf(d.w, w.w, acc[3][3]);
      }
      float4* up = reinterpret_cast<float4*>(xr + tile * 16);
      if (jh == 1) {
#pragma unroll
        for (int i = 0; i < 4; ++i) up[i] = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
      }
      __syncthreads();
      if (jh == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 u4 = up[i];
          const float o4[4] = {acc[i][0] + u4.x, acc[i][1] + u4.y, acc[i][2] + u4.z, acc[i][3] + u4.w};
          const int orow = b0 + rt * 4 + i;
          if (orow < B) st4(dp + (size_t)orow * K + k0 + kq * 4, o4);
        }
      }
    }
    nrd_dbg_stamp(dbgt, drow, 7);
    // mean loss: the last row group to get here sums the B loss rows in k_ce's fixed order
    if (ks == 0) {
      __syncthreads();                                 // every publishing thread of this CTA has stored its rows
      if (tid == 0) { __threadfence(); last = (atomicAdd(counter, 1u) == (unsigned)nrg - 1) ? 1 : 0; }
      __syncthreads();
      if (last) {
        __threadfence();
        double acc = 0.0;
        if (tid < 256) { for (int bb = tid; bb < B; bb += 256) acc += (double)__ldcg(loss_rows + bb); red[tid] = acc; }
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) { if (tid < o) red[tid] += red[tid + o]; __syncthreads(); }
        if (tid == 0) {
          const float lv = (float)(red[0] / B);
          *loss = lv; *counter = 0u;
          if (loss_mirror) {
            // main.py:41 `loss.item()` without a copy or a sync in the stream: the loss lands in (mapped, pinned) HOST memory by
            // itself, then the step's sequence number; the host reads the value once it sees the number (nrd_plan_set_loss_mirror)
            const unsigned int sq = *seq_dev + 1u;
            *seq_dev = sq;
            loss_mirror[(sq - 1u) % NRD_LOSS_RING] = lv;
            __threadfence_system();
            *seq_mirror = sq;
          }
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// fused DDP mean + SGD step: p <- p - lr * inv_world * g                      (main.py:27,39,63)
// ------------------------------------------------------------------------------------------
__global__ void k_sgd(float* __restrict__ p, const float* __restrict__ g, size_t n, float step_scale) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    p[i] = fmaf(-step_scale, g[i], p[i]);
}

// SGD with momentum as torch.optim.SGD(lr, momentum) runs it (ppe_main_ddp.py:133; dampening 0, no nesterov, no weight decay):
// buf <- g on the first step, buf <- mu * buf + g afterwards; p <- p - lr * buf
__global__ void k_sgd_momentum(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ buf, size_t n, float lr, float mu, int first) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float gi = g[i];
    const float b = first ? gi : fmaf(mu, buf[i], gi);
    buf[i] = b;
    p[i] = fmaf(-lr, b, p[i]);
  }
}

// binary_cross_entropy_with_logits, mean reduction (ppe_main_ddp.py:147), forward and gradient in one pass:
// loss_i = max(x,0) - x*t + log1p(exp(-|x|));  dL/dx_i = (sigmoid(x) - t) * grad_scale / n.  One CTA, fixed-shape tree.
__global__ void __launch_bounds__(256) k_bce_logits(const float* __restrict__ x, const float* __restrict__ t, size_t n, float gscale,
                                                    float* __restrict__ dx, float* __restrict__ loss) {
  __shared__ double red[256];
  double acc = 0.0;
  const float sc = gscale / (float)n;
  for (size_t i = threadIdx.x; i < n; i += 256) {
    const float xi = x[i], ti = t[i];
    const float l = fmaxf(xi, 0.f) - xi * ti + log1pf(expf(-fabsf(xi)));
    acc += (double)l;
    if (dx) dx[i] = (1.f / (1.f + expf(-xi)) - ti) * sc;
  }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0 && loss) *loss = (float)(red[0] / (double)n);
}

// p <- p - step_scale * g over the whole flat vector AND the packed copies of the three tensors that have them, one launch
// (off[] = flat offsets of the 9 tensors; index order as k_pack_weights)
// [lo, hi): the part of the flat vector this launch updates (whole vector, or one bucket when the update rides the backward);
// wait_mask: which exchange buckets it needs (bit 0 body, bit 1 head)
__global__ void k_sgd_pack(float* __restrict__ p, float* g, float step_scale, int C, long long lo, long long hi, int wait_mask,
                           long long o_c1w, long long o_rbw, long long o_fc1w, long long n,
                           float* __restrict__ w_fwd, float* __restrict__ w_dgrad, bf16* __restrict__ wtc_fwd, bf16* __restrict__ wtc_dgrad,
                           float* __restrict__ w1p, float* __restrict__ conv1_wp, nrd_xchg_args x, float* bn_sync) {
  (void)n;
  nrd_pdl_wait();
  // peer-memory exchange: every gradient is the rank-ordered sum of the world's slots in OUR buffer (body bucket [0, o_fc1w)
  // and head bucket [o_fc1w, n) arrive separately)
  const bool px = x.world > 0;
  const float* slots = nullptr;
  const float* slots_head = nullptr;
  if (px) {
    float* own = x.peer[x.rank];
    const unsigned long long e_body = (wait_mask & 1) ? nrd_xchg_wait(x, 0, (int)threadIdx.x) : 0ull;
    const unsigned long long e_head = (wait_mask & 2) ? nrd_xchg_wait(x, 1, (int)threadIdx.x - 32) : 0ull;
    __syncthreads();
    slots = own + (size_t)(e_body & 1) * x.world * x.nx;
    slots_head = own + (size_t)(e_head & 1) * x.world * x.nx;
    if (blockIdx.x == 0 && bn_sync && (wait_mask & 1))
      for (int j = threadIdx.x; j < 2 * C; j += blockDim.x) {
        float sres = 0.f;
        for (int r = 0; r < x.world; ++r) sres += __ldcg(slots + (size_t)r * x.nx + j);
        bn_sync[j] = sres;
      }
  }
  const long long n_rb = (long long)C * C * 9, n_fc1 = (long long)NRD_FC1_OUT * 64 * C, n_c1 = (long long)C * 27;
  for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (long long)gridDim.x * blockDim.x) {
    float gi;
    if (px) {
      const float* sl = i < o_fc1w ? slots : slots_head;
      long long si = i;                            // slot index: k_body_reduce's order for the two conv weights
      if (i >= o_rbw && i < o_rbw + n_rb) { const int q = (int)(i - o_rbw), co = q / (C * 9), r9 = q % (C * 9), ci = r9 / 9, t = r9 % 9; si = o_rbw + ((long long)t * C + co) * C + ci; }
      else if (i >= o_c1w && i < o_c1w + n_c1) { const int q = (int)(i - o_c1w), co = q / 27, k = q % 27; si = o_c1w + (long long)k * C + co; }
      gi = 0.f;
      for (int r = 0; r < x.world; ++r) gi += __ldcg(sl + (size_t)r * x.nx + 2 * C + si);
      g[i] = gi;                                   // .grad keeps holding the world's sum, as after the all-reduce
    } else {
      gi = g[i];
    }
    const float w = fmaf(-step_scale, gi, p[i]);
    p[i] = w;
    if (i >= o_rbw && i < o_rbw + n_rb) {
      const int q = (int)(i - o_rbw), co = q / (C * 9), r = q % (C * 9), ci = r / 9, t = r % 9;
      w_fwd[(t * C + ci) * C + co] = w;
      w_dgrad[((8 - t) * C + co) * C + ci] = w;
      wtc_fwd[(size_t)co * 9 * C + t * C + ci] = __float2bfloat16_rn(w);
      wtc_dgrad[(size_t)ci * 9 * C + (8 - t) * C + co] = __float2bfloat16_rn(w);
    } else if (i >= o_fc1w && i < o_fc1w + n_fc1) {
      const int q = (int)(i - o_fc1w), j = q / (64 * C), r = q % (64 * C), c = r / 64, hw = r % 64;
      w1p[((size_t)hw * C + c) * NRD_FC1_OUT + j] = w;
    } else if (i >= o_c1w && i < o_c1w + n_c1) {
      const int q = (int)(i - o_c1w), co = q / 27, k = q % 27;
      conv1_wp[k * C + co] = w;
    }
  }
}

// ------------------------------------------------------------------------------------------
// Real input path (main.py:53-61): the uint8 HWC dataset stays resident in HBM; a batch is a gather by sample index fused with
// ToTensor (u8 -> [0,1] fp32, HWC -> CHW) and Normalize((x - mean[c]) / std[c]) -- 4x fewer bytes read than an fp32 dataset,
// one pass.  Thread = one pixel of one sample: reads its 3 bytes, writes one float into each of the 3 channel planes (the
// warp's 32 pixels are 128 contiguous bytes per plane).  The arithmetic is torchvision's, in its order:
// (float(u8) / 255 - mean) / std with fp32 division, so the batch equals the reference DataLoader's bit for bit.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_gather_normalize(const uint8_t* __restrict__ data /*[N,32,32,3]*/, const int64_t* __restrict__ labels_all,
                                                          const int64_t* __restrict__ idx /*[B]*/, int B, long long N, float m0, float m1, float m2,
                                                          float s0, float s1, float s2, float* __restrict__ out /*[B,3,32,32]*/,
                                                          int64_t* __restrict__ labels_out) {
  const int b = blockIdx.x >> 2, p = (blockIdx.x & 3) * 256 + threadIdx.x;       // 4 CTAs per sample, 1024 pixels
  if (b >= B) return;
  long long src = idx[b];
  if (src < 0 || src >= N) src = 0;                    // never read outside the dataset; the host validates the index list
  const uint8_t* px = data + ((size_t)src * 1024 + p) * 3;
  const float r = (float)px[0], g = (float)px[1], bl = (float)px[2];
  float* o = out + (size_t)b * 3072 + p;
  o[0] = (r / 255.f - m0) / s0; o[1024] = (g / 255.f - m1) / s1; o[2048] = (bl / 255.f - m2) / s2;
  if (p == 0 && labels_out) labels_out[b] = labels_all[src];
}

// ------------------------------------------------------------------------------------------
// synthetic CIFAR-shaped batch: Philox-4x32-10, counter = (element/4, rank, step lo, step hi)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t (&out)[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

__global__ void k_gen_batch(float* __restrict__ img, int64_t* __restrict__ labels, int B, uint64_t seed, uint32_t rank,
                            uint64_t step) {
  const size_t n4 = (size_t)B * 3072 / 4;
  const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4 + (size_t)B; i += (size_t)gridDim.x * blockDim.x) {
    uint32_t r[4];
    if (i < n4) {
      philox4x32_10((uint32_t)i, rank, (uint32_t)step, (uint32_t)(step >> 32), k0, k1, r);
      // Box-Muller on two pairs of 32-bit uniforms; u in (0,1]
      float u0 = ((float)(r[0] >> 8) + 1.0f) * (1.0f / 16777216.0f), u1 = (float)(r[1] >> 8) * (1.0f / 16777216.0f);
      float u2 = ((float)(r[2] >> 8) + 1.0f) * (1.0f / 16777216.0f), u3 = (float)(r[3] >> 8) * (1.0f / 16777216.0f);
      float ra = sqrtf(-2.0f * logf(u0)), rb = sqrtf(-2.0f * logf(u2));
      float s0, c0, s1, c1;
      sincospif(2.0f * u1, &s0, &c0); sincospif(2.0f * u3, &s1, &c1);
      reinterpret_cast<float4*>(img)[i] = make_float4(ra * c0, ra * s0, rb * c1, rb * s1);
    } else {
      size_t b = i - n4;
      philox4x32_10((uint32_t)b, rank | 0x80000000u, (uint32_t)step, (uint32_t)(step >> 32), k0, k1, r);
      labels[b] = (int64_t)(r[0] % 10u);
    }
  }
}

// ==========================================================================================
// host side
// ==========================================================================================
static thread_local char g_err[512] = "";
void nrd_set_error(const char* fmt, ...) {
  va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
}
extern "C" const char* nrd_last_error(void) { return g_err; }
extern "C" int nrd_version(void) { return 100; }

#include <atomic>
#include <mutex>
static std::atomic<unsigned long long> g_launches{0};
void nrd_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
extern "C" unsigned long long nrd_launch_count(void) { return g_launches.load(); }
#define LAUNCH_CHECK() do { nrd_count_launch(); NRD_CHECK_CUDA(cudaGetLastError()); } while (0)

static inline int grid_for(size_t n, int block, int cap = NRD_NUM_SMS * 8) {
  size_t g = (n + block - 1) / block;
  if (g > (size_t)cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}
static bool valid_C(int C) { return C == 32 || C == 64; }
static bool valid_dt(int dt) { return dt == NRD_F32 || dt == NRD_BF16; }

extern "C" int nrd_param_offsets(int C, int64_t off[10]) {
  NRD_REQUIRE(valid_C(C), "nrd_param_offsets: C must be 32 or 64 (got %d)", C);
  int64_t sizes[9] = {(int64_t)C * 27, C, (int64_t)C * C * 9, C, C, (int64_t)NRD_FC1_OUT * 64 * C, NRD_FC1_OUT,
                      NRD_NCLS * NRD_FC1_OUT, NRD_NCLS};
  off[0] = 0;
  for (int i = 0; i < 9; ++i) off[i + 1] = off[i] + sizes[i];
  return NRD_OK;
}

// ---- per-op launchers -------------------------------------------------------------------
static int fc1_wgrad_splits(int B) { int n = (B + 63) / 64; return n < 8 ? n : 8; }
static int wgrad_blocks(int B) { int t = B * 2; return t < NRD_NUM_SMS ? t : NRD_NUM_SMS; }
static int stem_bwd_blocks(int B) { return B < NRD_NUM_SMS ? B : NRD_NUM_SMS; }

// opt-in dynamic shared memory is a per-device function attribute: set it once per (kernel, device),
// outside any stream capture (the first, eager call does it)
template <typename F>
static int set_smem_once(F* fn, size_t smem, bool (&done)[64]) {
  int dev = 0;
  NRD_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 64 && done[dev]) return NRD_OK;
  NRD_CHECK_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (dev < 64) done[dev] = true;
  return NRD_OK;
}

template <typename T>
static int launch_conv(const void* x, const float* w, void* y, int B, int C, float* stats, const void* res, cudaStream_t st) {
  if (C == 32) {
    size_t smem = (10 * 18 * 33 + 9 * 32 * 32) * sizeof(float);
    static bool done[64]; NRD_TRY(set_smem_once(k_conv3x3<T, 32>, smem, done));
    k_conv3x3<T, 32><<<B * 2, 256, smem, st>>>((const T*)x, w, (T*)y, stats, (const T*)res);
  } else {
    size_t smem = (10 * 18 * 65 + 9 * 64 * 64) * sizeof(float);
    static bool done[64]; NRD_TRY(set_smem_once(k_conv3x3<T, 64>, smem, done));
    k_conv3x3<T, 64><<<B * 2, 256, smem, st>>>((const T*)x, w, (T*)y, stats, (const T*)res);
  }
  LAUNCH_CHECK();
  return NRD_OK;
}
static int conv_fwd_any(const void* x, const float* w, void* y, int B, int C, int dt, float* stats, const void* res, cudaStream_t st) {
  return dt == NRD_F32 ? launch_conv<float>(x, w, y, B, C, stats, res, st) : launch_conv<bf16>(x, w, y, B, C, stats, res, st);
}

template <typename T>
static int launch_wgrad(const void* x, const void* dy, float* partial, int B, int C, int first, cudaStream_t st) {
  dim3 grid(wgrad_blocks(B), 3);
  if (C == 32) {
    size_t smem = (10 * 18 * 33 + 128 * 32) * sizeof(float);
    static bool done[64]; NRD_TRY(set_smem_once(k_wgrad<T, 32>, smem, done));
    k_wgrad<T, 32><<<grid, 256, smem, st>>>((const T*)x, (const T*)dy, partial, B * 2, first);
  } else {
    size_t smem = (10 * 18 * 65 + 128 * 64) * sizeof(float);
    static bool done[64]; NRD_TRY(set_smem_once(k_wgrad<T, 64>, smem, done));
    k_wgrad<T, 64><<<grid, 256, smem, st>>>((const T*)x, (const T*)dy, partial, B * 2, first);
  }
  LAUNCH_CHECK();
  return NRD_OK;
}
static int wgrad_any(const void* x, const void* dy, float* partial, int B, int C, int dt, int first, cudaStream_t st) {
  return dt == NRD_F32 ? launch_wgrad<float>(x, dy, partial, B, C, first, st) : launch_wgrad<bf16>(x, dy, partial, B, C, first, st);
}
static int wgrad_reduce(const float* partial, int nblk, int C, float* dw, cudaStream_t st) {
  k_wgrad_reduce<<<9 * C * C / 32, 256, 0, st>>>(partial, nblk, C, dw);
  LAUNCH_CHECK();
  return NRD_OK;
}

static int bn_relu_add_any(const void* y, const void* x, const float* ss, void* out, int B, int C, int dt, cudaStream_t st) {
  size_t n8 = (size_t)B * NRD_PIX * C / 8;
  int g = grid_for(n8, 256);
  if (dt == NRD_F32) k_bn_relu_add<float><<<g, 256, 0, st>>>((const float*)y, (const float*)x, ss, (float*)out, n8, C);
  else k_bn_relu_add<bf16><<<g, 256, 0, st>>>((const bf16*)y, (const bf16*)x, ss, (bf16*)out, n8, C);
  LAUNCH_CHECK();
  return NRD_OK;
}

// grid, images per CTA, and how many of those stay in shared memory (200 KB budget); returns the dynamic shared-memory size
static size_t bn_bwd_fused_smem(int B, int C, int dt, int* grid_out, int* ipc_out, int* ipcs_out = nullptr) {
  const int grid = B < NRD_NUM_SMS ? B : NRD_NUM_SMS, ipc = (B + grid - 1) / grid;
  const size_t fixed = 1024 * sizeof(double) + (4096 + 3 * 64) * sizeof(float), per_img = (size_t)2 * NRD_PIX * C * (dt == NRD_F32 ? 4 : 2);
  const int fit = (int)((200 * 1024 - fixed) / per_img), ipc_s = ipc < fit ? ipc : fit;
  if (grid_out) *grid_out = grid;
  if (ipc_out) *ipc_out = ipc;
  if (ipcs_out) *ipcs_out = ipc_s;
  return (size_t)ipc_s * per_img + fixed;
}

// scratch layout: [B][2C] partials | coef [3C].  gbar != NULL selects the one-launch cooperative kernel when it fits.
static int bn_relu_bwd_any(const void* g, const void* y, const float* ss, const float* saved, const float* gamma, void* dy,
                           float* dgamma, float* dbeta, int accumulate, int B, int C, int dt, float* scratch, cudaStream_t st,
                           unsigned long long* gbar = nullptr) {
  float* partial = scratch; float* coef = scratch + (size_t)B * 2 * C;
  int grid = 0, ipc = 0, ipc_s = 0;
  const size_t fsmem = bn_bwd_fused_smem(B, C, dt, &grid, &ipc, &ipc_s);
  if (gbar) {
    void* args[] = {(void*)&g, (void*)&y, (void*)&ss, (void*)&saved, (void*)&gamma, (void*)&dy, (void*)&dgamma, (void*)&dbeta,
                    (void*)&accumulate, (void*)&partial, (void*)&gbar, (void*)&B, (void*)&C, (void*)&ipc, (void*)&ipc_s};
    static bool done32[64], done16[64];
    if (dt == NRD_F32) {
      NRD_TRY(set_smem_once(k_bn_bwd_fused<float>, 200 * 1024, done32));
      NRD_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)k_bn_bwd_fused<float>, dim3(grid), dim3(256), args, fsmem, st));
    } else {
      NRD_TRY(set_smem_once(k_bn_bwd_fused<bf16>, 200 * 1024, done16));
      NRD_CHECK_CUDA(cudaLaunchCooperativeKernel((const void*)k_bn_bwd_fused<bf16>, dim3(grid), dim3(256), args, fsmem, st));
    }
    nrd_count_launch();
    return NRD_OK;
  }
  if (dt == NRD_F32) k_bn_bwd_reduce<float><<<B, 256, 0, st>>>((const float*)g, (const float*)y, ss, saved, partial, C);
  else k_bn_bwd_reduce<bf16><<<B, 256, 0, st>>>((const bf16*)g, (const bf16*)y, ss, saved, partial, C);
  LAUNCH_CHECK();
  k_bn_bwd_finalize<<<1, 256, 0, st>>>(partial, B, B * NRD_PIX, C, gamma, saved, dgamma, dbeta, accumulate, coef);
  LAUNCH_CHECK();
  size_t n8 = (size_t)B * NRD_PIX * C / 8;
  int gr = grid_for(n8, 256);
  if (dt == NRD_F32) k_bn_bwd_apply<float><<<gr, 256, 0, st>>>((const float*)g, (const float*)y, ss, saved, coef, (float*)dy, n8, C);
  else k_bn_bwd_apply<bf16><<<gr, 256, 0, st>>>((const bf16*)g, (const bf16*)y, ss, saved, coef, (bf16*)dy, n8, C);
  LAUNCH_CHECK();
  return NRD_OK;
}

static int stem_fwd_any(const float* img, const float* wp, const float* bias, void* x0, uint8_t* idx0, int B, int C, int dt, cudaStream_t st,
                        nrd_bn_sync_args bs = nrd_bn_sync_args{nullptr, nullptr, 0.f, 0}) {
  dim3 grid(4, B); int threads = 64 * (C / 8);
  size_t smem = (1056 + 27 * C + C) * sizeof(float);
  if (dt == NRD_F32) k_stem_fwd<float><<<grid, threads, smem, st>>>(img, wp, bias, (float*)x0, idx0, C, bs);
  else k_stem_fwd<bf16><<<grid, threads, smem, st>>>(img, wp, bias, (bf16*)x0, idx0, C, bs);
  LAUNCH_CHECK();
  return NRD_OK;
}
static int stem_bwd_any(const float* img, const void* g0, const void* x0, const uint8_t* idx0, float* dw, float* db, int B, int C,
                        int dt, float* partial, cudaStream_t st, bool do_reduce = true) {
  int nblk = stem_bwd_blocks(B);
  size_t smem = (3 * 34 * 35 + 2 + 256 * C) * sizeof(float) + 256 * C;
  {
    static bool done32[64], done16[64];
    if (dt == NRD_F32) NRD_TRY(set_smem_once(k_stem_bwd<float>, (3 * 34 * 35 + 2 + 256 * 64) * sizeof(float) + 256 * 64, done32));
    else NRD_TRY(set_smem_once(k_stem_bwd<bf16>, (3 * 34 * 35 + 2 + 256 * 64) * sizeof(float) + 256 * 64, done16));
  }
  if (dt == NRD_F32) k_stem_bwd<float><<<nblk, 256, smem, st>>>(img, (const float*)g0, (const float*)x0, idx0, partial, B, C);
  else k_stem_bwd<bf16><<<nblk, 256, smem, st>>>(img, (const bf16*)g0, (const bf16*)x0, idx0, partial, B, C);
  LAUNCH_CHECK();
  if (do_reduce) {
    k_stem_bwd_reduce<<<28 * C / 32, 256, 0, st>>>(partial, nblk, C, dw, db);
    LAUNCH_CHECK();
  }
  return NRD_OK;
}

// ---- per-op C entries -------------------------------------------------------------------
extern "C" int nrd_pack_conv_weight(const float* w, int C, float* w_fwd, float* w_dgrad, void* stream) {
  NRD_REQUIRE(valid_C(C) && w, "nrd_pack_conv_weight: bad arguments");
  k_pack_weights<<<grid_for((size_t)C * C * 9, 256), 256, 0, (cudaStream_t)stream>>>(w, nullptr, nullptr, C, w_fwd, w_dgrad,
                                                                                     nullptr, nullptr, nullptr, nullptr);
  // the kernel's index space also covers fc1/conv1; with null outputs those iterations are no-ops
  LAUNCH_CHECK();
  return NRD_OK;
}

extern "C" int nrd_conv3x3_fwd(const void* x, const float* w_packed, void* y, int B, int C, int dtype, float* stats_partial,
                               int* stats_blocks, const void* residual, void* stream) {
  NRD_REQUIRE(valid_C(C) && valid_dt(dtype) && B > 0 && x && w_packed && y, "nrd_conv3x3_fwd: bad arguments (B=%d C=%d dtype=%d)", B, C, dtype);
  if (stats_blocks) *stats_blocks = B * 2;
  return conv_fwd_any(x, w_packed, y, B, C, dtype, stats_partial, residual, (cudaStream_t)stream);
}

extern "C" size_t nrd_wgrad_partial_bytes(int B, int C) { return (size_t)wgrad_blocks(B) * 9 * C * C * sizeof(float); }

extern "C" int nrd_conv3x3_wgrad(const void* x, const void* dy, float* dw, int B, int C, int dtype, float* partial, void* stream) {
  NRD_REQUIRE(valid_C(C) && valid_dt(dtype) && B > 0 && x && dy && dw && partial, "nrd_conv3x3_wgrad: bad arguments");
  NRD_TRY(wgrad_any(x, dy, partial, B, C, dtype, 1, (cudaStream_t)stream));
  return wgrad_reduce(partial, wgrad_blocks(B), C, dw, (cudaStream_t)stream);
}

extern "C" int nrd_bn_finalize(const float* partial, int blocks, int count, int C, const float* gamma, const float* beta,
                               float* saved, float* scale_shift, float* bn_running, int64_t* nbt, void* stream) {
  NRD_REQUIRE(valid_C(C) && blocks > 0 && count > 1 && partial && gamma && beta && saved && scale_shift, "nrd_bn_finalize: bad arguments");
  k_bn_finalize<<<1, 256, 0, (cudaStream_t)stream>>>(partial, blocks, count, C, gamma, beta, saved, scale_shift, bn_running, nbt);
  LAUNCH_CHECK();
  return NRD_OK;
}

extern "C" int nrd_bn_relu_add(const void* y, const void* x, const float* scale_shift, void* out, int B, int C, int dtype, void* stream) {
  NRD_REQUIRE(valid_C(C) && valid_dt(dtype) && B > 0 && y && x && out && scale_shift, "nrd_bn_relu_add: bad arguments");
  return bn_relu_add_any(y, x, scale_shift, out, B, C, dtype, (cudaStream_t)stream);
}

extern "C" size_t nrd_bn_bwd_scratch_bytes(int B, int C) { return ((size_t)B * 2 * C + 3 * C) * sizeof(float); }

extern "C" int nrd_bn_relu_bwd(const void* g, const void* y, const float* scale_shift, const float* saved, const float* gamma, void* dy,
                               float* dgamma_dbeta, int accumulate, int B, int C, int dtype, float* scratch, void* stream) {
  NRD_REQUIRE(valid_C(C) && valid_dt(dtype) && B > 0 && g && y && dy && scratch && dgamma_dbeta, "nrd_bn_relu_bwd: bad arguments");
  return bn_relu_bwd_any(g, y, scale_shift, saved, gamma, dy, dgamma_dbeta, dgamma_dbeta + C, accumulate, B, C, dtype, scratch,
                         (cudaStream_t)stream);
}

extern "C" int nrd_stem_fwd(const float* images, const float* conv1_w, const float* conv1_b, void* x0, uint8_t* idx0, int B, int C,
                            int dtype, void* stream) {
  NRD_REQUIRE(valid_C(C) && valid_dt(dtype) && B > 0 && images && conv1_w && conv1_b && x0 && idx0, "nrd_stem_fwd: bad arguments");
  // conv1_w here is the PACKED [27][C] copy (see nrd_plan / k_pack_weights)
  return stem_fwd_any(images, conv1_w, conv1_b, x0, idx0, B, C, dtype, (cudaStream_t)stream);
}
extern "C" size_t nrd_stem_bwd_partial_bytes(int B, int C) { return (size_t)stem_bwd_blocks(B) * 28 * C * sizeof(float); }
extern "C" int nrd_stem_bwd(const float* images, const void* g0, const void* x0, const uint8_t* idx0, float* dw, float* db, int B, int C,
                            int dtype, float* partial, void* stream) {
  NRD_REQUIRE(valid_C(C) && valid_dt(dtype) && B > 0 && images && g0 && x0 && idx0 && dw && db && partial, "nrd_stem_bwd: bad arguments");
  return stem_bwd_any(images, g0, x0, idx0, dw, db, B, C, dtype, partial, (cudaStream_t)stream);
}

extern "C" int nrd_maxpool_fwd(const void* x, void* pooled, uint8_t* idx, int B, int C, int dtype, void* stream) {
  NRD_REQUIRE(valid_C(C) && valid_dt(dtype) && B > 0 && x && pooled && idx, "nrd_maxpool_fwd: bad arguments");
  size_t n = (size_t)B * 64 * (C / 8);
  if (dtype == NRD_F32) k_maxpool_fwd<float><<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>((const float*)x, (float*)pooled, idx, B, C);
  else k_maxpool_fwd<bf16><<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)x, (bf16*)pooled, idx, B, C);
  LAUNCH_CHECK();
  return NRD_OK;
}
extern "C" int nrd_maxpool_bwd(const void* gp, const uint8_t* idx, void* gx, int B, int C, int dtype, void* stream) {
  NRD_REQUIRE(valid_C(C) && valid_dt(dtype) && B > 0 && gp && gx && idx, "nrd_maxpool_bwd: bad arguments");
  size_t n = (size_t)B * 64 * (C / 8);
  if (dtype == NRD_F32) k_maxpool_bwd<float><<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>((const float*)gp, idx, (float*)gx, B, C);
  else k_maxpool_bwd<bf16><<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>((const bf16*)gp, idx, (bf16*)gx, B, C);
  LAUNCH_CHECK();
  return NRD_OK;
}

extern "C" int nrd_cross_entropy(const float* logits, const int64_t* labels, int B, float grad_scale, float* loss_out, float* dlogits,
                                 float* row_scratch, void* stream) {
  NRD_REQUIRE(B > 0 && logits && labels && row_scratch, "nrd_cross_entropy: bad arguments");
  k_ce<<<1, 256, 0, (cudaStream_t)stream>>>(logits, labels, B, grad_scale, row_scratch, dlogits, loss_out);
  LAUNCH_CHECK();
  return NRD_OK;
}

// ---- allocation-free cross-entropy ---------------------------------------------------------
// Some callers cannot hand us scratch memory because they cannot allocate any: code that runs after
// the application has given the whole device to another allocator or a fixed arena, plug-in hooks
// that only receive input/output pointers, and callers already inside a region where cudaMalloc is
// illegal.  For them the per-row loss scratch of nrd_cross_entropy lives in this file-scope device
// array instead (32 KiB, resident on every GPU that loads the library until the process exits).
#define NRD_CE_STATIC_MAX_ROWS 8192
static __device__ __align__(256) float g_ce_static_rows[NRD_CE_STATIC_MAX_ROWS];
static float* g_ce_static_ptr[64];       // device address of the array, looked up once per device
static std::mutex g_ce_static_mu;

extern "C" int nrd_cross_entropy_static(const float* logits, const int64_t* labels, int B, float grad_scale, float* loss_out,
                                        float* dlogits, void* stream) {
  NRD_REQUIRE(B > 0 && logits && labels && loss_out, "nrd_cross_entropy_static: bad arguments");
  NRD_REQUIRE(B <= NRD_CE_STATIC_MAX_ROWS,
              "nrd_cross_entropy_static: B=%d is beyond the built-in scratch (%d rows); call nrd_cross_entropy with your own row_scratch",
              B, NRD_CE_STATIC_MAX_ROWS);
  int dev = -1;
  NRD_CHECK_CUDA(cudaGetDevice(&dev));
  NRD_REQUIRE(dev >= 0 && dev < 64, "nrd_cross_entropy_static: device index %d not supported", dev);
  float* rows = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_ce_static_mu);
    if (!g_ce_static_ptr[dev]) {
      // In host code the array's name is only a handle; the address a kernel can use comes from the runtime,
      // per device, with that device current.  This is the one-time set-up of the first call on a device.
      void* p = nullptr;
      NRD_CHECK_CUDA(cudaGetSymbolAddress(&p, g_ce_static_rows));
      g_ce_static_ptr[dev] = static_cast<float*>(p);
    }
    rows = g_ce_static_ptr[dev];
  }
  // rows [0,B) are fully overwritten before they are read, so nothing has to start at zero
  k_ce<<<1, 256, 0, (cudaStream_t)stream>>>(logits, labels, B, grad_scale, rows, dlogits, loss_out);
  LAUNCH_CHECK();
  return NRD_OK;
}

extern "C" int nrd_sgd_update(float* params, const float* grads, size_t n, float lr, float inv_world, void* stream) {
  NRD_REQUIRE(params && grads, "nrd_sgd_update: null pointer");
  if (n == 0) return NRD_OK;
  k_sgd<<<grid_for(n, 256, NRD_NUM_SMS * 2), 256, 0, (cudaStream_t)stream>>>(params, grads, n, lr * inv_world);
  LAUNCH_CHECK();
  return NRD_OK;
}

extern "C" int nrd_sgd_momentum_update(float* params, const float* grads, float* momentum_buf, size_t n, float lr, float momentum, int first_step,
                                       void* stream) {
  NRD_REQUIRE(params && grads && momentum_buf && n > 0, "nrd_sgd_momentum_update: bad arguments");
  k_sgd_momentum<<<grid_for(n, 256, NRD_NUM_SMS * 2), 256, 0, (cudaStream_t)stream>>>(params, grads, momentum_buf, n, lr, momentum, first_step ? 1 : 0);
  LAUNCH_CHECK();
  return NRD_OK;
}

extern "C" int nrd_bce_with_logits(const float* logits, const float* targets, size_t n, float grad_scale, float* loss_out, float* dlogits, void* stream) {
  NRD_REQUIRE(logits && targets && n > 0 && (loss_out || dlogits), "nrd_bce_with_logits: bad arguments");
  k_bce_logits<<<1, 256, 0, (cudaStream_t)stream>>>(logits, targets, n, grad_scale, dlogits, loss_out);
  LAUNCH_CHECK();
  return NRD_OK;
}

extern "C" int nrd_gather_normalize(const uint8_t* data_u8_nhwc, const int64_t* labels_all, long long n_samples, const int64_t* indices, int B,
                                    const float mean[3], const float std[3], float* images, int64_t* labels, void* stream) {
  NRD_REQUIRE(data_u8_nhwc && labels_all && indices && images && mean && std && B > 0 && n_samples > 0, "nrd_gather_normalize: bad arguments");
  NRD_REQUIRE(std[0] != 0.f && std[1] != 0.f && std[2] != 0.f, "nrd_gather_normalize: zero std");
  k_gather_normalize<<<B * 4, 256, 0, (cudaStream_t)stream>>>(data_u8_nhwc, labels_all, indices, B, n_samples, mean[0], mean[1], mean[2], std[0], std[1], std[2],
                                                              images, labels);
  LAUNCH_CHECK();
  return NRD_OK;
}

extern "C" int nrd_gen_batch(float* images, int64_t* labels, int B, uint64_t seed, uint32_t rank, uint64_t step, void* stream) {
  NRD_REQUIRE(B > 0 && images && labels, "nrd_gen_batch: bad arguments");
  size_t n = (size_t)B * 768 + B;
  k_gen_batch<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(images, labels, B, seed, rank, step);
  LAUNCH_CHECK();
  return NRD_OK;
}

// ---- head (one cooperative launch) -----------------------------------------------------------
static int head_init(nrd_plan* p) {
  int occ = 0, sms = 0;
  NRD_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, p->device));
  if (p->dtype == NRD_F32) {
    NRD_CHECK_CUDA(cudaFuncSetAttribute(k_head_fused<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)head_smem_bytes()));
    NRD_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_head_fused<float>, NRD_HEAD_THREADS, head_smem_bytes()));
  } else {
    NRD_CHECK_CUDA(cudaFuncSetAttribute(k_head_fused<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)head_smem_bytes()));
    NRD_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_head_fused<bf16>, NRD_HEAD_THREADS, head_smem_bytes()));
  }
  p->head_ctas = occ * sms;
  NRD_REQUIRE(p->head_ctas >= 64 * p->C / 128, "the head kernel does not fit this device (%d co-resident CTAs)", p->head_ctas);
  return NRD_OK;
}

// pool + fc1 + fc2 (+ loss, dlogits when labels != NULL) (+ dL/d(pooled) into dp when dp != NULL), one launch
static int launch_head(nrd_plan* p, const void* x_last, int pooled_in, float* logits, const int64_t* labels, float gscale, float* loss_rows,
                       float* dlogits, float* loss, void* dp, cudaStream_t st) {
  const int B = p->B, C = p->C, KS = 64 * C / 128, nrg = (B + 31) / 32;
  int gx = p->head_ctas / KS; if (gx > nrg) gx = nrg;
  const float* w1t = p->w1p; const float* b1 = p->params + p->off[6]; const float* w2 = p->params + p->off[7]; const float* b2 = p->params + p->off[8];
  unsigned int* counter = reinterpret_cast<unsigned int*>(p->gbar + NRD_GBAR_HEAD_COUNTER);
  unsigned int* seq_dev = reinterpret_cast<unsigned int*>(p->gbar + NRD_GBAR_LOSS_SEQ);
  void* pooled = p->pooled; uint8_t* idx2 = p->idx2; float* part = p->fc1_part; float* h1 = p->h1; unsigned long long* ctr = p->head_ctr;
  int Bv = B, Cv = C;
  void* args[] = {(void*)&x_last, (void*)&pooled, (void*)&idx2, (void*)&pooled_in, (void*)&w1t, (void*)&b1, (void*)&w2, (void*)&b2, (void*)&part, (void*)&h1,
                  (void*)&logits, (void*)&Bv, (void*)&Cv, (void*)&labels, (void*)&gscale, (void*)&loss_rows, (void*)&dlogits, (void*)&loss, (void*)&ctr,
                  (void*)&counter, (void*)&dp, (void*)&p->dbg, (void*)&p->loss_mirror, (void*)&p->seq_mirror, (void*)&seq_dev};
  const void* fn = p->dtype == NRD_F32 ? (const void*)k_head_fused<float> : (const void*)k_head_fused<bf16>;
  NRD_CHECK_CUDA(cudaLaunchCooperativeKernel(fn, dim3(gx, KS), dim3(NRD_HEAD_THREADS), args, head_smem_bytes(), st));
  nrd_count_launch();
  return NRD_OK;
}

// ---- plan ---------------------------------------------------------------------------------
static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

struct Carver {
  char* base; size_t off;
  template <typename U> U* take(size_t bytes) { U* p = base ? reinterpret_cast<U*>(base + off) : nullptr; off += align256(bytes); return p; }
};

static void carve(nrd_plan* p, char* base, size_t* total) {
  const int B = p->B, C = p->C, NB = p->NB;
  const size_t act = p->act_elems * p->esz;
  Carver cv{base, 0};
  p->xs = cv.take<char>(act * (NB + 1));
  p->ys = cv.take<char>(act * NB);
  p->dy = cv.take<char>(act);
  p->dxa = cv.take<char>(act);
  p->dxb = cv.take<char>(act);
  p->idx0 = cv.take<uint8_t>(p->act_elems);
  p->pooled = cv.take<char>(act / 4);
  p->idx2 = cv.take<uint8_t>(p->act_elems / 4);
  p->h1 = cv.take<float>((size_t)B * NRD_FC1_OUT * 4);
  p->dh1 = cv.take<float>((size_t)B * NRD_FC1_OUT * 4);
  p->fc1_part = cv.take<float>((size_t)(64 * C / 128) * B * NRD_FC1_OUT * 4);
  p->fc1w_partial = cv.take<float>((size_t)fc1_wgrad_splits(B) * NRD_FC1_OUT * 64 * C * 4);
  p->dlogits = cv.take<float>((size_t)B * NRD_NCLS * 4);
  p->loss_rows = cv.take<float>((size_t)B * 4);
  p->saved_stats = cv.take<float>((size_t)NB * 2 * C * 4);
  p->scale_shift = cv.take<float>((size_t)NB * 2 * C * 4);
  p->stats_partial = cv.take<float>((size_t)(B * 2 > 1024 ? B * 2 : 1024) * 2 * C * 4);
  p->bnbwd_scratch = cv.take<float>(nrd_bn_bwd_scratch_bytes(B, C));
  p->wgrad_partial = cv.take<float>(nrd_wgrad_partial_bytes(B, C));
  p->stem_partial = cv.take<float>(nrd_stem_bwd_partial_bytes(B, C));
  p->w_fwd = cv.take<float>((size_t)9 * C * C * 4);
  p->w_dgrad = cv.take<float>((size_t)9 * C * C * 4);
  p->w1p = cv.take<float>((size_t)NRD_FC1_OUT * 64 * C * 4);
  p->conv1_wp = cv.take<float>((size_t)27 * C * 4);
  p->wtc_fwd = cv.take<bf16>((size_t)9 * C * C * 2);
  p->wtc_dgrad = cv.take<bf16>((size_t)9 * C * C * 2);
  p->gbar = cv.take<unsigned long long>(256);
  p->head_ctr = cv.take<unsigned long long>((size_t)((B + 31) / 32) * sizeof(unsigned long long));
  p->dbg_buf = cv.take<long long>(64 * 16 * 2 * sizeof(long long));
  p->bn_acc = cv.take<unsigned long long>((size_t)2 * 32 * NRD_ACC_ROW * sizeof(unsigned long long));
  *total = cv.off;
}

static int fill_shape(nrd_plan* p, int B, int C, int NB, int dtype) {
  NRD_REQUIRE(valid_C(C), "C (n_chans1) must be 32 or 64, got %d", C);
  NRD_REQUIRE(valid_dt(dtype), "dtype must be NRD_F32 or NRD_BF16, got %d", dtype);
  NRD_REQUIRE(B >= 1 && B <= 8192, "per-rank batch must be in [1, 8192], got %d", B);
  NRD_REQUIRE(NB >= 1 && NB <= 31, "n_blocks must be in [1, 31], got %d", NB);
  p->B = B; p->C = C; p->NB = NB; p->dtype = dtype;
  p->esz = dtype == NRD_F32 ? 4 : 2;
  p->act_elems = (size_t)B * NRD_PIX * C;
  return nrd_param_offsets(C, p->off);
}

extern "C" size_t nrd_workspace_bytes(int B, int C, int NB, int dtype) {
  nrd_plan tmp; memset(&tmp, 0, sizeof(tmp));
  if (fill_shape(&tmp, B, C, NB, dtype) != NRD_OK) return 0;
  size_t total = 0; carve(&tmp, nullptr, &total);
  return total;
}

extern "C" int nrd_plan_create(nrd_plan** out, int device, int B, int C, int NB, int dtype, void* workspace, size_t workspace_bytes,
                               float* params, float* grads, float* bn_running, int64_t* nbt) {
  NRD_REQUIRE(out && workspace && params && grads && bn_running && nbt, "nrd_plan_create: null pointer");
  NRD_REQUIRE(((uintptr_t)workspace & 255) == 0, "nrd_plan_create: workspace must be 256-byte aligned");
  nrd_plan* p = new nrd_plan; memset(p, 0, sizeof(*p));
  int r = fill_shape(p, B, C, NB, dtype);
  if (r != NRD_OK) { delete p; return r; }
  size_t total = 0; carve(p, (char*)workspace, &total);
  if (total > workspace_bytes) { delete p; nrd_set_error("workspace too small: need %zu bytes, got %zu", total, workspace_bytes); return NRD_EINVAL; }
  p->device = device; p->params = params; p->grads = grads; p->bn_running = bn_running; p->nbt = nbt;
  p->conv_impl = NRD_CONV_CUDA_CORE;
  p->fuse_bn_bwd = 1;
  p->persist_fwd = 1;
  p->persist_bwd = 1;
  p->use_tc_stem = 1;
  p->fuse_head_tail = 1;
  p->one_copy = 1;           // forward only: in the backward the one-pixel-apart N groups of the weight gradient measured 2 us slower
  p->fuse_stem_bwd = 0;      // conv1's weight-gradient partials in the persistent backward's tail: bit-identical, measured +-0 (the tail
                             // costs the 2-image CTAs ~8 us, what the stand-alone kernel costs behind its programmatic launch)
  p->fuse_stem = 0;          // measured: as long as the stem kernel it replaces (its four half images run one after the other)
  p->fuse_pool = 0;          // measured 1 us slower per step than pooling in the fc1 kernel (it sits on the blocks' critical path)
  { const char* e = getenv("NRD_POLL_NS"); p->poll_ns = e ? atoi(e) : 0; }
  p->pdl = 5;     // measured: the consumers of the two persistent kernels (bit 1) lose 13 us when made programmatic dependents
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) { delete p; nrd_set_error("cudaSetDevice(%d): %s", device, cudaGetErrorString(e)); return NRD_ECUDA; }
  e = cudaMemset(p->gbar, 0, 256);
  if (e != cudaSuccess) { delete p; nrd_set_error("cudaMemset(grid barrier): %s", cudaGetErrorString(e)); return NRD_ECUDA; }
  e = cudaMemset(p->bn_acc, 0, (size_t)2 * 32 * NRD_ACC_ROW * sizeof(unsigned long long));   // the persistent kernels keep it zero
  if (e != cudaSuccess) { delete p; nrd_set_error("cudaMemset(BN accumulators): %s", cudaGetErrorString(e)); return NRD_ECUDA; }
  e = cudaMemset(p->head_ctr, 0, (size_t)((B + 31) / 32) * sizeof(unsigned long long));
  if (e != cudaSuccess) { delete p; nrd_set_error("cudaMemset(head counters): %s", cudaGetErrorString(e)); return NRD_ECUDA; }
  r = head_init(p);
  if (r != NRD_OK) { delete p; return r; }
  if (dtype == NRD_BF16) {
    r = nrd_tc_plan_init(p);
    if (r != NRD_OK) { delete p; return r; }   // no silent fallback: the bf16 product path IS the tcgen05 path
    if (p->tc_ready) p->conv_impl = NRD_CONV_TCGEN05;
    r = nrd_tc_stem_init(p);
    if (r != NRD_OK) { delete p; return r; }
  }
  *out = p;
  return NRD_OK;
}

extern "C" int nrd_plan_destroy(nrd_plan* p) { delete p; return NRD_OK; }

extern "C" int nrd_plan_set_conv_impl(nrd_plan* p, int impl) {
  NRD_REQUIRE(p, "null plan");
  NRD_REQUIRE(impl == NRD_CONV_CUDA_CORE || ((impl == NRD_CONV_TCGEN05 || impl == NRD_CONV_TCGEN05_UNFUSED) && p->dtype == NRD_BF16 && p->tc_ready),
              "conv impl %d not available for this plan (dtype=%d)", impl, p->dtype);
  p->conv_impl = impl;
  return NRD_OK;
}

extern "C" int nrd_plan_set_option(nrd_plan* p, int option, int value) {
  NRD_REQUIRE(p, "null plan");
  switch (option) {
    case NRD_OPT_FUSE_BN_BWD: p->fuse_bn_bwd = value ? 1 : 0; return NRD_OK;
    case NRD_OPT_PERSIST_FWD: p->persist_fwd = value ? 1 : 0; return NRD_OK;
    case NRD_OPT_PERSIST_BWD: p->persist_bwd = value ? 1 : 0; return NRD_OK;
    case NRD_OPT_TC_STEM: p->use_tc_stem = value ? 1 : 0; return NRD_OK;
    case NRD_OPT_PDL: p->pdl = value; return NRD_OK;
    case NRD_OPT_HEAD_TAIL: p->fuse_head_tail = value ? 1 : 0; return NRD_OK;
    case NRD_OPT_FUSE_POOL: p->fuse_pool = value ? 1 : 0; return NRD_OK;
    case NRD_OPT_FUSE_STEM: p->fuse_stem = value ? 1 : 0; return NRD_OK;
    case NRD_OPT_FUSE_STEM_BWD: p->fuse_stem_bwd = value ? 1 : 0; return NRD_OK;
    case NRD_OPT_STREAM:
      NRD_REQUIRE(value >= 0 && value <= 2, "NRD_OPT_STREAM: 0 (auto), 1 (always) or 2 (never), got %d", value);
      p->stream_opt = value;
      if (p->tc_ready) nrd_tc_refresh_limits(p);
      return NRD_OK;
    case NRD_OPT_ONE_COPY: p->one_copy = value & 3; return NRD_OK;
    case NRD_OPT_RESERVE_SMS:
      NRD_REQUIRE(value >= 0 && value <= NRD_NUM_SMS / 2, "NRD_OPT_RESERVE_SMS: %d out of range [0, %d]", value, NRD_NUM_SMS / 2);
      if (value != p->reserve_sms) {
        // the grid of the cooperative kernels changes: their monotonic barrier counters recover the launch's base as
        // v - v % grid, which only holds while the grid never changes -- start them from zero again (nothing of this plan
        // may be in flight or being captured when an option is set)
        NRD_CHECK_CUDA(cudaSetDevice(p->device));
        NRD_CHECK_CUDA(cudaDeviceSynchronize());
        NRD_CHECK_CUDA(cudaMemset(p->gbar, 0, 4 * sizeof(unsigned long long)));
      }
      p->reserve_sms = value;
      if (p->tc_ready) nrd_tc_refresh_limits(p);
      return NRD_OK;
    case NRD_OPT_TIMELINE:
      if (value && !p->tl_ev[0]) {
        NRD_CHECK_CUDA(cudaSetDevice(p->device));
        for (int i = 0; i < 40; ++i) NRD_CHECK_CUDA(cudaEventCreate(&p->tl_ev[i]));
      }
      p->tl_on = value ? 1 : 0; p->tl_n = 0;
      return NRD_OK;
    case NRD_OPT_TRACE: p->dbg_skip_stores = (value == 2);
      p->trace = value ? (void*)p->dxb : nullptr; p->trace_bwd = value ? (void*)p->dy : nullptr; p->dbg = value ? p->dbg_buf : nullptr; return NRD_OK;   // stamps land in the dx pong / dy scratch
    default: nrd_set_error("nrd_plan_set_option: unknown option %d", option); return NRD_EINVAL;
  }
}

extern "C" int nrd_plan_set_bn_sync(nrd_plan* p, float* slot, float contrib, float inv_world) {
  NRD_REQUIRE(p, "null plan");
  NRD_REQUIRE(slot == nullptr || slot + 2 * p->C == p->grads, "nrd_plan_set_bn_sync: the slot must be the 2C floats right in front of the gradient vector");
  p->bn_sync = slot;
  p->bn_sync_contrib = contrib;
  p->bn_sync_inv_world = inv_world;
  return NRD_OK;
}

// ---- peer-memory gradient exchange: buffers and wiring -------------------------------------------------------------------
extern "C" size_t nrd_exchange_bytes(int C, int world) {
  int64_t off[10];
  if (world < 2 || world > NRD_XCHG_MAX_WORLD || nrd_param_offsets(C, off) != NRD_OK) return 0;
  return nrd_xchg_bytes(world, (2 * C + (int)off[9] + 3) & ~3, (2 * C + (int)off[5] + 3) & ~3);
}

// The exchange buffer is the one piece of device memory the library allocates itself: it must be a whole cudaMalloc
// allocation to be exportable with cudaIpcGetMemHandle (a sub-range of a framework's caching allocator is not).
extern "C" int nrd_exchange_alloc(int device, size_t bytes, void** ptr, void* ipc_handle_64) {
  NRD_REQUIRE(ptr && ipc_handle_64 && bytes > 0, "nrd_exchange_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  NRD_CHECK_CUDA(cudaSetDevice(device));
  NRD_CHECK_CUDA(cudaMalloc(ptr, bytes));
  NRD_CHECK_CUDA(cudaMemset(*ptr, 0, bytes));
  NRD_CHECK_CUDA(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, *ptr);
  if (e != cudaSuccess) { cudaFree(*ptr); *ptr = nullptr; nrd_set_error("cudaIpcGetMemHandle: %s", cudaGetErrorString(e)); return NRD_ECUDA; }
  memcpy(ipc_handle_64, &h, 64);
  return NRD_OK;
}

extern "C" int nrd_exchange_open(int device, const void* ipc_handle_64, void** ptr) {
  NRD_REQUIRE(ptr && ipc_handle_64, "nrd_exchange_open: null pointer");
  NRD_CHECK_CUDA(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle_64, 64);
  NRD_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return NRD_OK;
}

extern "C" int nrd_exchange_release(void* ptr, int opened) {
  if (!ptr) return NRD_OK;
  if (opened) NRD_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  else NRD_CHECK_CUDA(cudaFree(ptr));
  return NRD_OK;
}

extern "C" int nrd_plan_set_exchange(nrd_plan* p, int rank, int world, void* const* bufs) {
  NRD_REQUIRE(p, "null plan");
  if (world == 0 || !bufs) { memset(&p->xchg, 0, sizeof(p->xchg)); return NRD_OK; }
  NRD_REQUIRE(world >= 2 && world <= NRD_XCHG_MAX_WORLD && rank >= 0 && rank < world, "nrd_plan_set_exchange: world %d / rank %d out of range", world, rank);
  for (int r = 0; r < world; ++r) { NRD_REQUIRE(bufs[r], "nrd_plan_set_exchange: null buffer for rank %d", r); p->xchg.peer[r] = (float*)bufs[r]; }
  p->xchg.world = world; p->xchg.rank = rank; p->xchg.nx = (2 * p->C + (int)p->off[9] + 3) & ~3;
  p->xchg.nll = (2 * p->C + (int)p->off[5] + 3) & ~3;
  {
    // how long the optimizer kernel waits for a peer's gradients before it traps the context: 10 minutes by default (the
    // default of the NCCL watchdog this exchange replaces), NRD_XCHG_TIMEOUT_S overrides; converted with the SM clock rate
    const char* e = getenv("NRD_XCHG_TIMEOUT_S");
    double secs = e ? atof(e) : 600.0;
    if (!(secs > 0.0)) secs = 600.0;
    int khz = 0;
    if (cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, p->device) != cudaSuccess || khz <= 0) khz = 2000000;
    p->xchg.timeout_cycles = (long long)(secs * 1e3 * (double)khz);
  }
  return NRD_OK;
}

extern "C" int nrd_plan_conv_impl(const nrd_plan* p) { return p ? p->conv_impl : NRD_EINVAL; }

// Sticky status of the plan's kernels (NRD_STATUS_* bits): synchronises `stream`, reads the word, optionally clears it.
extern "C" int nrd_plan_status(nrd_plan* p, unsigned int* flags_out, int clear, void* stream) {
  NRD_REQUIRE(p && flags_out, "nrd_plan_status: null pointer");
  NRD_CHECK_CUDA(cudaSetDevice(p->device));
  cudaStream_t st = (cudaStream_t)stream;
  unsigned int* w = reinterpret_cast<unsigned int*>(p->gbar + NRD_GBAR_STATUS);
  NRD_CHECK_CUDA(cudaMemcpyAsync(flags_out, w, sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
  if (clear) NRD_CHECK_CUDA(cudaMemsetAsync(w, 0, sizeof(unsigned int), st));
  NRD_CHECK_CUDA(cudaStreamSynchronize(st));
  if (*flags_out & NRD_STATUS_BN_NONFINITE) { nrd_set_error("BatchNorm statistics: a per-CTA sum was NaN/Inf (the activations or gradients diverged)"); return NRD_ERANGE; }
  if (*flags_out & NRD_STATUS_BN_SATURATED) { nrd_set_error("BatchNorm statistics: a per-CTA sum left the fixed-point range of the persistent kernels (forward |sum| >= 2^22, backward >= 2^10): statistics of that step were clamped"); return NRD_ERANGE; }
  return NRD_OK;
}

extern "C" int nrd_plan_tensor(nrd_plan* p, int what, int k, void** ptr, size_t* bytes) {
  NRD_REQUIRE(p && ptr && bytes, "null pointer");
  const size_t act = p->act_elems * p->esz;
  switch (what) {
    case 0: NRD_REQUIRE(k >= 0 && k <= p->NB, "k out of range"); *ptr = p->xs + act * k; *bytes = act; break;
    case 1: NRD_REQUIRE(k >= 0 && k < p->NB, "k out of range"); *ptr = p->ys + act * k; *bytes = act; break;
    case 2: *ptr = p->dy; *bytes = act; break;
    case 3: *ptr = p->pooled; *bytes = act / 4; break;
    case 4: NRD_REQUIRE(k >= 0 && k < p->NB, "k out of range"); *ptr = p->saved_stats + (size_t)k * 2 * p->C; *bytes = 2 * p->C * 4; break;
    case 5: *ptr = p->dxa; *bytes = act; break;
    case 6: *ptr = p->dxb; *bytes = act; break;
    case 7: *ptr = p->wgrad_partial; *bytes = nrd_wgrad_partial_bytes(p->B, p->C); break;
    case 9: *ptr = p->stem_partial; *bytes = nrd_stem_bwd_partial_bytes(p->B, p->C); break;
    case 11: *ptr = p->dbg_buf; *bytes = 64 * 16 * 2 * sizeof(long long); break;
    case 10: *ptr = p->h1; *bytes = (size_t)p->B * NRD_FC1_OUT * 4; break;
    case 8: *ptr = p->stats_partial; *bytes = (size_t)(p->B * 2 > 1024 ? p->B * 2 : 1024) * 2 * p->C * 4; break;
    default: nrd_set_error("nrd_plan_tensor: unknown selector %d", what); return NRD_EINVAL;
  }
  return NRD_OK;
}

#define PRM(i) (p->params + p->off[i])
#define GRD(i) (p->grads + p->off[i])
enum { P_C1W = 0, P_C1B, P_RBW, P_BNG, P_BNB, P_FC1W, P_FC1B, P_FC2W, P_FC2B };

extern "C" int nrd_pack_weights(nrd_plan* p, void* stream) {
  NRD_REQUIRE(p, "null plan");
  NRD_CHECK_CUDA(cudaSetDevice(p->device));
  const int C = p->C;
  size_t n = (size_t)C * C * 9 + (size_t)NRD_FC1_OUT * 64 * C + C * 27;
  k_pack_weights<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(PRM(P_RBW), PRM(P_FC1W), PRM(P_C1W), C, p->w_fwd, p->w_dgrad,
                                                                     p->wtc_fwd, p->wtc_dgrad, p->w1p, p->conv1_wp);
  LAUNCH_CHECK();
  return NRD_OK;
}

// optimizer step over [lo, hi) of the flat vector (+ the packed copies of what it touches); wait_mask: exchange buckets it needs
static int sgd_pack_range(nrd_plan* p, float step_scale, long long lo, long long hi, int wait_mask, int pdl, cudaStream_t st) {
  NRD_CHECK_CUDA(nrd_launch_pdl(k_sgd_pack, dim3(grid_for((size_t)(hi - lo), 256, NRD_NUM_SMS * 2)), dim3(256), 0, st, pdl,
                                p->params, p->grads, step_scale, p->C, lo, hi, wait_mask, p->off[P_C1W], p->off[P_RBW], p->off[P_FC1W], (long long)p->off[9], p->w_fwd,
                                p->w_dgrad, p->wtc_fwd, p->wtc_dgrad, p->w1p, p->conv1_wp, p->xchg, p->bn_sync));
  LAUNCH_CHECK();
  return NRD_OK;
}

// optimizer.step() (+ DDP mean) and the weight re-pack in ONE launch over the flat vector
extern "C" int nrd_sgd_update_pack(nrd_plan* p, float lr, float inv_world, void* stream) {
  NRD_REQUIRE(p, "null plan");
  NRD_CHECK_CUDA(cudaSetDevice(p->device));
  const int r = sgd_pack_range(p, lr * inv_world, 0, p->off[9], 3, (p->pdl >> 2) & 1, (cudaStream_t)stream);
  nrd_tl_mark(p, (cudaStream_t)stream, "join+sgd_pack");
  return r;
}

// Read the timeline of the last step run with NRD_OPT_TIMELINE on (synchronise first): us[i] = microseconds between mark i and
// mark i+1, names[i] = what ran in between (the kernel launched before mark i+1).  Returns the number of intervals, < 0 on error.
extern "C" int nrd_plan_timeline(nrd_plan* p, int max_n, float* us, const char** names) {
  NRD_REQUIRE(p && us && names, "nrd_plan_timeline: null pointer");
  NRD_REQUIRE(p->tl_on, "nrd_plan_timeline: NRD_OPT_TIMELINE is off");
  NRD_CHECK_CUDA(cudaSetDevice(p->device));
  int n = 0;
  for (int i = 0; i + 1 < p->tl_n && n < max_n; ++i, ++n) {
    float ms = 0.f;
    NRD_CHECK_CUDA(cudaEventElapsedTime(&ms, p->tl_ev[i], p->tl_ev[i + 1]));
    us[n] = ms * 1e3f; names[n] = p->tl_name[i + 1];
  }
  return n;
}

extern "C" int nrd_plan_set_loss_mirror(nrd_plan* p, float* host_loss, unsigned int* host_seq) {
  NRD_REQUIRE(p, "null plan");
  NRD_REQUIRE((host_loss == nullptr) == (host_seq == nullptr), "nrd_plan_set_loss_mirror: give both pointers or neither");
  if (host_loss) {
    NRD_CHECK_CUDA(cudaSetDevice(p->device));
    // must be memory the device can write: pinned host memory is mapped at its own address under unified addressing
    cudaPointerAttributes a;
    NRD_CHECK_CUDA(cudaPointerGetAttributes(&a, host_loss));
    NRD_REQUIRE(a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged || a.type == cudaMemoryTypeDevice,
                "nrd_plan_set_loss_mirror: the mirror must be pinned (page-locked) host memory");
    NRD_CHECK_CUDA(cudaMemset(p->gbar + NRD_GBAR_LOSS_SEQ, 0, sizeof(unsigned long long)));
    *host_seq = 0u;
  }
  p->loss_mirror = host_loss; p->seq_mirror = host_seq;
  return NRD_OK;
}

extern "C" int nrd_plan_set_fused_update(nrd_plan* p, float lr, float inv_world) {
  NRD_REQUIRE(p, "null plan");
  NRD_REQUIRE(!(lr > 0.f) || inv_world > 0.f, "nrd_plan_set_fused_update: inv_world must be positive (got %g)", (double)inv_world);
  p->upd_lr = lr > 0.f ? lr : 0.f;
  p->upd_inv_world = inv_world;
  return NRD_OK;
}

static int forward_impl(nrd_plan* p, const float* images, float* logits, int training, void* stream, const int64_t* labels,
                        float grad_scale, float* loss_out, float* dlogits, float* loss_rows);

extern "C" int nrd_forward(nrd_plan* p, const float* images, float* logits, int training, void* stream) {
  return forward_impl(p, images, logits, training, stream, nullptr, 0.f, nullptr, nullptr, nullptr);
}

// training forward + CrossEntropyLoss forward/backward: the loss rides the last kernel of the forward (one launch less)
extern "C" int nrd_forward_loss(nrd_plan* p, const float* images, const int64_t* labels, float grad_scale, float* logits,
                                float* loss_out, float* dlogits, float* row_scratch, void* stream) {
  NRD_REQUIRE(labels && loss_out && dlogits && row_scratch, "nrd_forward_loss: null pointer");
  return forward_impl(p, images, logits, 1, stream, labels, grad_scale, loss_out, dlogits, row_scratch);
}

static int forward_impl(nrd_plan* p, const float* images, float* logits, int training, void* stream, const int64_t* labels,
                        float grad_scale, float* loss_out, float* dlogits, float* loss_rows) {
  NRD_REQUIRE(p && images && logits, "nrd_forward: null pointer");
  NRD_CHECK_CUDA(cudaSetDevice(p->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int B = p->B, C = p->C, NB = p->NB, dt = p->dtype;
  const size_t act = p->act_elems * p->esz;
  p->images_last = images;
  nrd_tl_mark(p, st, "start", true);
  const bool persist = training && p->conv_impl == NRD_CONV_TCGEN05 && nrd_persist_fwd_ok(p) && p->persist_fwd;
  const bool stem_in_blocks = persist && !p->tc_stream && p->fuse_stem && p->tc_stem && p->use_tc_stem;
  if (stem_in_blocks) {}                                         // x_0 is produced inside the persistent kernel
  else if (p->conv_impl != NRD_CONV_CUDA_CORE && p->tc_stem && p->use_tc_stem) NRD_TRY(nrd_tc_stem_fwd(p, images, training, st));
  else NRD_TRY(stem_fwd_any(images, p->conv1_wp, PRM(P_C1B), p->xs, p->idx0, B, C, dt, st, nrd_bn_sync_of(p, training)));
  nrd_tl_mark(p, st, "stem_fwd");
  if (!training) { k_bn_eval_coeffs<<<1, 64, 0, st>>>(p->bn_running, PRM(P_BNG), PRM(P_BNB), C, p->scale_shift); LAUNCH_CHECK(); }
  const bool pool_in_blocks = persist && !p->tc_stream && p->fuse_pool;
  if (persist) NRD_TRY(nrd_tc_forward_blocks(p, pool_in_blocks ? 1 : 0, stem_in_blocks ? images : nullptr, training, st));       // ... and the max-pool behind the last block
  if (persist) nrd_tl_mark(p, st, "fwd_blocks");
  for (int k = 0; k < (persist ? 0 : NB); ++k) {
    char* xk = p->xs + act * k; char* yk = p->ys + act * k; char* xn = p->xs + act * (k + 1);
    float* ss = training ? p->scale_shift + (size_t)k * 2 * C : p->scale_shift;
    if (training && p->conv_impl == NRD_CONV_TCGEN05 && p->tc_fused_fwd) {   // conv + BN stats + grid barrier + apply: one launch
      NRD_TRY(nrd_tc_block_fwd(p, k, st));
      continue;
    }
    int blocks = B * 2;
    if (p->conv_impl != NRD_CONV_CUDA_CORE) { NRD_TRY(nrd_tc_conv_fwd(p, k, st)); blocks = nrd_tc_stats_blocks(p); }
    else NRD_TRY(conv_fwd_any(xk, p->w_fwd, yk, B, C, dt, training ? p->stats_partial : nullptr, nullptr, st));
    if (training) {
      k_bn_finalize<<<1, 256, 0, st>>>(p->stats_partial, blocks, B * NRD_PIX, C, PRM(P_BNG), PRM(P_BNB),
                                       p->saved_stats + (size_t)k * 2 * C, ss, p->bn_running, p->nbt);
      LAUNCH_CHECK();
    }
    NRD_TRY(bn_relu_add_any(yk, xk, ss, xn, B, C, dt, st));
  }
  // head: pool + fc1 + fc2 (+ loss and dlogits, + the backward's first product dL/d(pooled)) in ONE cooperative launch
  p->dp_for = nullptr;
  const bool tail = labels && dt == NRD_BF16 && p->conv_impl == NRD_CONV_TCGEN05 && nrd_persist_bwd_ok(p) && p->persist_bwd && p->fuse_head_tail;
  NRD_TRY(launch_head(p, p->xs + act * NB, pool_in_blocks ? 1 : 0, logits, labels, grad_scale, loss_rows, dlogits, loss_out, tail ? (void*)p->dxa : nullptr, st));
  nrd_tl_mark(p, st, "head");
  if (tail) p->dp_for = dlogits;       // d(pooled) is ready when nrd_backward_body(dlogits) is called
  if (labels) { p->labels_last = labels; p->logits_last = logits; p->loss_last = loss_out; p->dlogits_last = dlogits; p->loss_rows_last = loss_rows; }
  return NRD_OK;
}

// head weight gradients (fc1.weight/bias, fc2.weight/bias): independent of the body backward given dlogits
extern "C" int nrd_backward_head(nrd_plan* p, const float* dlogits, void* stream) {
  NRD_REQUIRE(p && dlogits, "nrd_backward_head: null pointer");
  NRD_CHECK_CUDA(cudaSetDevice(p->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int B = p->B, C = p->C, dt = p->dtype;
  k_head_wgrad<<<12, 256, 0, st>>>(dlogits, p->h1, PRM(P_FC2W), B, GRD(P_FC2W), GRD(P_FC2B), GRD(P_FC1B));
  LAUNCH_CHECK();
  const int nsplit = fc1_wgrad_splits(B);
  if (dt == NRD_F32) k_fc1_wgrad<float><<<dim3(C, nsplit), 256, 0, st>>>((const float*)p->pooled, dlogits, p->h1, PRM(P_FC2W), p->fc1w_partial, B, C);
  else k_fc1_wgrad<bf16><<<dim3(C, nsplit), 256, 0, st>>>((const bf16*)p->pooled, dlogits, p->h1, PRM(P_FC2W), p->fc1w_partial, B, C);
  LAUNCH_CHECK();
  k_fc1_wgrad_reduce<<<(NRD_FC1_OUT * 64 * C + 255) / 256, 256, 0, st>>>(p->fc1w_partial, nsplit, C, GRD(P_FC1W));
  LAUNCH_CHECK();
  if (p->xchg.world > 0) {      // peer-memory exchange: the head bucket leaves now, under the body backward
    k_xchg_push_head<<<32, 256, 0, st>>>(p->xchg, p->grads, (int)p->off[P_FC1W], (int)p->off[9], C);
    LAUNCH_CHECK();
  }
  // fused update: the head bucket (fc1 / fc2, 87 % of the parameters) takes its optimizer step here, on the caller's side stream,
  // under the body backward -- nothing after this point reads these weights or their packed copy before the next forward
  if (p->upd_lr > 0.f) NRD_TRY(sgd_pack_range(p, p->upd_lr * p->upd_inv_world, p->off[P_FC1W], p->off[9], 2, 0, st));
  return NRD_OK;
}

extern "C" int nrd_backward(nrd_plan* p, const float* dlogits, void* stream) {
  NRD_TRY(nrd_backward_head(p, dlogits, stream));
  return nrd_backward_body(p, dlogits, stream);
}

// fixed-order reduce of the conv / stem partials into the flat gradient vector (+ exchange push, + fused optimizer step)
static int launch_body_reduce(nrd_plan* p, int stem_rows, int after_persistent, cudaStream_t st) {
  const int B = p->B, C = p->C;
  nrd_upd_args u = {};
  if (p->upd_lr > 0.f) {
    u.params = p->params; u.step_scale = p->upd_lr * p->upd_inv_world;
    u.w_fwd = p->w_fwd; u.w_dgrad = p->w_dgrad; u.wtc_fwd = p->wtc_fwd; u.wtc_dgrad = p->wtc_dgrad; u.conv1_wp = p->conv1_wp;
  }
  // directly behind a persistent kernel a programmatic dependent would only sit on the SMs and wait (measured: slower), see NRD_OPT_PDL
  NRD_CHECK_CUDA(nrd_launch_pdl(k_body_reduce, dim3((9 * C * C + 28 * C) / 64), dim3(256), 0, st, after_persistent ? (p->pdl >> 1) & 1 : (p->pdl >> 2) & 1, (const float*)p->wgrad_partial,
                                p->conv_impl != NRD_CONV_CUDA_CORE ? nrd_tc_wgrad_blocks(p) : wgrad_blocks(B), (const float*)p->stem_partial, stem_rows, C,
                                p->grads, p->bn_sync, (const float*)p->bn_running, p->bn_sync_contrib, p->xchg,
                                (int)p->off[P_BNG], (int)p->off[P_RBW], (int)p->off[P_C1W], (int)p->off[P_C1B], u, p->dbg));
  LAUNCH_CHECK();
  return NRD_OK;
}

// everything below the head: d(pooled) -> residual blocks -> stem; fills the body gradients (conv1, block conv, gamma, beta)
extern "C" int nrd_backward_body(nrd_plan* p, const float* dlogits, void* stream) {
  NRD_REQUIRE(p && dlogits, "nrd_backward_body: null pointer");
  NRD_CHECK_CUDA(cudaSetDevice(p->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int B = p->B, C = p->C, NB = p->NB, dt = p->dtype;
  const size_t act = p->act_elems * p->esz;
  char* g = p->dxa; char* gn = p->dxb;
  const int K = 64 * C;
  const bool persist = p->conv_impl == NRD_CONV_TCGEN05 && nrd_persist_bwd_ok(p) && p->persist_bwd;
  const bool stem_in_blocks = persist && !p->tc_stream && p->fuse_stem_bwd && p->tc_stem && p->use_tc_stem && !(p->one_copy & 2) && p->images_last != nullptr;
  if (persist && p->dp_for == dlogits) p->dp_for = nullptr;      // nrd_forward_loss already left d(pooled) in the dx scratch
  else if (persist) NRD_CHECK_CUDA(nrd_launch_pdl(k_fc1_dgrad_pooled<bf16>, dim3(((B + 7) / 8) * (K / 256)), dim3(256), 0, st, p->pdl & 1, dlogits, p->h1,
                                             PRM(P_FC2W), p->w1p, (bf16*)g, B, C));
  else if (dt == NRD_F32) k_fc1_dgrad_unpool<float><<<((B + 7) / 8) * (K / 256), 256, 0, st>>>(dlogits, p->h1, PRM(P_FC2W), p->w1p, p->idx2, (float*)g, B, C);
  else k_fc1_dgrad_unpool<bf16><<<((B + 7) / 8) * (K / 256), 256, 0, st>>>(dlogits, p->h1, PRM(P_FC2W), p->w1p, p->idx2, (bf16*)g, B, C);
  LAUNCH_CHECK();
  // residual blocks, last to first; the shared block's grads accumulate over the NB applications
  nrd_tl_mark(p, st, "fork/dgrad_fc1");
  if (persist) {
    NRD_TRY(nrd_tc_backward_blocks(p, g, gn, stem_in_blocks ? p->images_last : nullptr, st));
    char* t = g; g = gn; gn = t;
    nrd_tl_mark(p, st, "bwd_blocks");
  }
  for (int k = NB - 1; k >= (persist ? NB : 0); --k) {
    const int first = (k == NB - 1);
    char* xk = p->xs + act * k; char* yk = p->ys + act * k;
    NRD_TRY(bn_relu_bwd_any(g, yk, p->scale_shift + (size_t)k * 2 * C, p->saved_stats + (size_t)k * 2 * C, PRM(P_BNG), p->dy,
                            GRD(P_BNG), GRD(P_BNB), !first, B, C, dt, p->bnbwd_scratch, st,
                            p->fuse_bn_bwd ? p->gbar + NRD_GBAR_BN_BWD : nullptr));
    if (p->conv_impl != NRD_CONV_CUDA_CORE) {
      NRD_TRY(nrd_tc_conv_wgrad(p, k, first, st));
      NRD_TRY(nrd_tc_conv_dgrad(p, g, gn, st));
    } else {
      NRD_TRY(wgrad_any(xk, p->dy, p->wgrad_partial, B, C, dt, first, st));
      NRD_TRY(conv_fwd_any(p->dy, p->w_dgrad, gn, B, C, dt, nullptr, g, st));
    }
    char* t = g; g = gn; gn = t;
  }
  p->dx0_last = g;
  if (stem_in_blocks) {}                                         // conv1's partials came out of the persistent backward
  else if (p->conv_impl != NRD_CONV_CUDA_CORE && p->tc_stem && p->use_tc_stem) NRD_TRY(nrd_tc_stem_bwd(p, p->images_last, g, st));
  else NRD_TRY(stem_bwd_any(p->images_last, g, p->xs, p->idx0, GRD(P_C1W), GRD(P_C1B), B, C, dt, p->stem_partial, st, false));
  nrd_tl_mark(p, st, "stem_bwd");
  NRD_TRY(launch_body_reduce(p, stem_in_blocks ? nrd_tc_wgrad_blocks(p) : stem_bwd_blocks(p->B), stem_in_blocks ? 1 : 0, st));
  nrd_tl_mark(p, st, "body_reduce(+sgd)");
  return NRD_OK;
}

extern "C" int nrd_train_step(nrd_plan* p, const float* images, const int64_t* labels, float lr, float* logits, float* loss_out,
                              void* stream) {
  NRD_REQUIRE(p && images && labels && logits && loss_out, "nrd_train_step: null pointer");
  NRD_TRY(nrd_forward_loss(p, images, labels, 1.0f, logits, loss_out, p->dlogits, p->loss_rows, stream));
  NRD_TRY(nrd_backward(p, p->dlogits, stream));
  return nrd_sgd_update_pack(p, lr, 1.0f, stream);
}

// Run ONE internal op of the step on the plan's buffers (bench.py times each kernel class with CUDA
// events on the launching stream; ncu captures use the same entry).  Inputs are whatever the last
// nrd_forward / nrd_backward left in the workspace.
extern "C" int nrd_plan_run_op(nrd_plan* p, int op, int k, void* stream) {
  NRD_REQUIRE(p && k >= 0 && k < p->NB, "nrd_plan_run_op: bad arguments");
  NRD_CHECK_CUDA(cudaSetDevice(p->device));
  cudaStream_t st = (cudaStream_t)stream;
  const int B = p->B, C = p->C, dt = p->dtype;
  const size_t act = p->act_elems * p->esz;
  char* xk = p->xs + act * k; char* yk = p->ys + act * k;
  float* ss = p->scale_shift + (size_t)k * 2 * C; float* sv = p->saved_stats + (size_t)k * 2 * C;
  switch (op) {
    case NRD_OP_CONV_FWD:
      if (p->conv_impl != NRD_CONV_CUDA_CORE) return nrd_tc_conv_fwd(p, k, st);
      return conv_fwd_any(xk, p->w_fwd, yk, B, C, dt, p->stats_partial, nullptr, st);
    case NRD_OP_BN_APPLY: return bn_relu_add_any(yk, xk, ss, p->dxb, B, C, dt, st);
    case NRD_OP_BN_BWD: {
      // dgamma/dbeta go to a sink (h1 is dead after backward and holds >= 2C floats) so profiling leaves the grads alone
      return bn_relu_bwd_any(p->dxa, yk, ss, sv, PRM(P_BNG), p->dy, p->h1, p->h1 + C, 0, B, C, dt, p->bnbwd_scratch, st,
                             p->fuse_bn_bwd ? p->gbar + NRD_GBAR_BN_BWD : nullptr);
    }
    case NRD_OP_WGRAD:
      if (p->conv_impl != NRD_CONV_CUDA_CORE) return nrd_tc_conv_wgrad(p, k, 1, st);
      return wgrad_any(xk, p->dy, p->wgrad_partial, B, C, dt, 1, st);
    case NRD_OP_DGRAD:
      if (p->conv_impl != NRD_CONV_CUDA_CORE) return nrd_tc_conv_dgrad(p, p->dxa, p->dxb, st);
      return conv_fwd_any(p->dy, p->w_dgrad, p->dxb, B, C, dt, nullptr, p->dxa, st);
    case NRD_OP_STEM_FWD:
      NRD_REQUIRE(p->images_last, "run a forward first");
      if (p->conv_impl != NRD_CONV_CUDA_CORE && p->tc_stem && p->use_tc_stem) return nrd_tc_stem_fwd(p, p->images_last, 0, st);   // rewrites x_0 in place (same values)
      return stem_fwd_any(p->images_last, p->conv1_wp, PRM(P_C1B), p->dxb, p->idx0, B, C, dt, st);
    case NRD_OP_STEM_BWD:
      NRD_REQUIRE(p->images_last, "run a forward first");
      NRD_REQUIRE(p->dx0_last, "run a backward first");
      if (p->conv_impl != NRD_CONV_CUDA_CORE && p->tc_stem && p->use_tc_stem) return nrd_tc_stem_bwd(p, p->images_last, p->dx0_last, st);
      return stem_bwd_any(p->images_last, p->dx0_last, p->xs, p->idx0, GRD(P_C1W), GRD(P_C1B), B, C, dt, p->stem_partial, st);
    case NRD_OP_FWD_BLOCKS:
      NRD_REQUIRE(p->conv_impl == NRD_CONV_TCGEN05 && p->tc_persist_fwd, "persistent forward not available for this plan");
      return nrd_tc_forward_blocks(p, 0, nullptr, 0, st);
    case NRD_OP_BWD_BLOCKS:
      NRD_REQUIRE(p->conv_impl == NRD_CONV_TCGEN05 && nrd_persist_bwd_ok(p), "persistent backward not available for this plan");
      return nrd_tc_backward_blocks(p, p->dxa, p->dxb, nullptr, st);
    case NRD_OP_BLOCK_FWD:
      NRD_REQUIRE(p->conv_impl != NRD_CONV_CUDA_CORE && p->tc_fused_fwd, "fused block forward not available for this plan");
      return nrd_tc_block_fwd(p, k, st);
    case NRD_OP_HEAD: {
      NRD_REQUIRE(p->labels_last, "run nrd_forward_loss first");
      const bool tail = dt == NRD_BF16 && p->conv_impl == NRD_CONV_TCGEN05 && nrd_persist_bwd_ok(p) && p->persist_bwd && p->fuse_head_tail;
      return launch_head(p, p->xs + act * p->NB, 0, p->logits_last, p->labels_last, 1.0f, p->loss_rows_last, p->dlogits_last, p->loss_last,
                         tail ? (void*)p->dxa : nullptr, st);     // rewrites the same values
    }
    case NRD_OP_BODY_REDUCE: {
      // timing only: no exchange (the peers are not running this) and no optimizer step
      const float lr = p->upd_lr; const int world = p->xchg.world;
      p->upd_lr = 0.f; p->xchg.world = 0;
      const int r = launch_body_reduce(p, stem_bwd_blocks(p->B), 0, st);
      p->upd_lr = lr; p->xchg.world = world;
      return r;
    }
    case NRD_OP_SGD: {
      NRD_REQUIRE(p->xchg.world == 0, "NRD_OP_SGD would wait for peers that are not running it");
      return sgd_pack_range(p, 0.f, 0, p->off[9], 0, 0, st);      // step 0: same memory traffic, weights unchanged
    }
    default: nrd_set_error("nrd_plan_run_op: unknown op %d", op); return NRD_EINVAL;
  }
}
