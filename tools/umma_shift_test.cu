// Feasibility test for a ONE-copy activation layout (round-2 idea): does a K-major SWIZZLE_64B shared-memory descriptor whose
// start address is NOT aligned to the 512-byte swizzle atom (start = base + s*64 B, i.e. shifted by s pixels) read the rows
// base+s, base+s+1, ... correctly, given that the data was written with the address-based swizzle?  If yes, the three
// kx-shifted operand copies of the conv kernels can become one copy with a column halo.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/umma_shift_test tools/umma_shift_test.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../distributeddataparallel-cifar10_b200/csrc/tc_common.cuh"

void nrd_set_error(const char*, ...) {}

__global__ void __launch_bounds__(128, 1) k_test(int shift, int sbo, float* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t a_s = base, b_s = base + 32768, bar = base + 40960, tslot = bar + 8;
  volatile uint32_t* tslot_gen = reinterpret_cast<volatile uint32_t*>(gen + (tslot - base));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, t = threadIdx.x;
  // A region: 512 "pixels" of 64 B; pixel P channel c holds value (P + c/64): written with the address-based SWIZZLE_64B
  for (int i = t; i < 512 * 4; i += 128) {
    const int P = i >> 2, ch = i & 3;
    __nv_bfloat16 v[8];
    for (int e = 0; e < 8; ++e) v[e] = __float2bfloat16((float)(P % 256) + (float)(ch * 8 + e) / 64.f);
    const uint32_t addr = (uint32_t)(P * 64);
    *reinterpret_cast<uint4*>(gen + addr - 0 + (((ch ^ ((addr >> 7) & 3))) << 4) - 0 + 0) = *reinterpret_cast<uint4*>(v);
  }
  // B: 32x32 identity, K-major rows of 64 B, swizzled
  for (int i = t; i < 32 * 4; i += 128) {
    const int n = i >> 2, ch = i & 3;
    __nv_bfloat16 v[8];
    for (int e = 0; e < 8; ++e) v[e] = __float2bfloat16((ch * 8 + e) == n ? 1.f : 0.f);
    *reinterpret_cast<uint4*>(gen + 32768 + n * 64 + ((ch ^ ((n >> 1) & 3)) << 4)) = *reinterpret_cast<uint4*>(v);
  }
  fence_proxy_async();
  if (t == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(tslot, 32);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = *tslot_gen;
  if (t == 0) {
    const uint32_t idesc = make_idesc(128, 32, 0, 0);
    const uint64_t ad = make_sdesc(a_s + shift * 64, 16, sbo, 4), bd = make_sdesc(b_s, 16, 512, 4);
    umma_bf16(tmem, ad, bd, idesc, 0);
    umma_bf16(tmem, ad + 2, bd + 2, idesc, 1);
    umma_commit(bar);
  }
  mbar_wait(bar, 0);
  tc_fence_after();
  float v[32];
  tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16), v);
  for (int c = 0; c < 32; ++c) out[(warp * 32 + lane) * 32 + c] = v[c];
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 32); }
}

// MN-major B with N = 96 = three 32-channel groups whose LBO is ONE pixel (64 B): group j reads pixels shifted by j, i.e. the
// three horizontal taps of the weight gradient out of ONE dy copy.  A (MN-major, M = 128) is a selector: A[m][k] = (m == k).
__global__ void __launch_bounds__(128, 1) k_test_mn(int p0, float* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t b_s = base, a_s = base + 32768, bar = base + 45056, tslot = bar + 8;
  volatile uint32_t* tslot_gen = reinterpret_cast<volatile uint32_t*>(gen + (tslot - base));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, t = threadIdx.x;
  for (int i = t; i < 512 * 4; i += 128) {            // pixels: value(P, c) = P + c/64
    const int P = i >> 2, ch = i & 3;
    __nv_bfloat16 v[8];
    for (int e = 0; e < 8; ++e) v[e] = __float2bfloat16((float)(P % 256) + (float)(ch * 8 + e) / 64.f);
    const uint32_t addr = (uint32_t)(P * 64);
    *reinterpret_cast<uint4*>(gen + addr + ((ch ^ ((addr >> 7) & 3)) << 4)) = *reinterpret_cast<uint4*>(v);
  }
  // A: element (m, k) at (m/32)*1024 + k*64 + (m%32)*2 (K = 16 rows of 64 B, two 8-row atoms 512 B apart = contiguous), swizzled
  for (int i = t; i < 4 * 16 * 4; i += 128) {
    const int g = i / 64, k = (i / 4) % 16, ch = i & 3;
    __nv_bfloat16 v[8];
    for (int e = 0; e < 8; ++e) { const int m = g * 32 + ch * 8 + e; v[e] = __float2bfloat16(m == k ? 1.f : 0.f); }
    const uint32_t addr = (uint32_t)(g * 1024 + k * 64);
    *reinterpret_cast<uint4*>(gen + 32768 + addr + ((ch ^ ((addr >> 7) & 3)) << 4)) = *reinterpret_cast<uint4*>(v);
  }
  fence_proxy_async();
  if (t == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(tslot, 128);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = *tslot_gen;
  if (t == 0) {
    const uint32_t idesc = make_idesc(128, 96, 1, 1);
    const uint64_t ad = make_sdesc(a_s, 1024, 512, 4), bd = make_sdesc(b_s + p0 * 64, 64, 512, 4);
    umma_bf16(tmem, ad, bd, idesc, 0);
    umma_commit(bar);
  }
  mbar_wait(bar, 0);
  tc_fence_after();
  for (int j = 0; j < 3; ++j) {
    float v[32];
    tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + j * 32, v);
    for (int c = 0; c < 32; ++c) out[(warp * 32 + lane) * 96 + j * 32 + c] = v[c];
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 128); }
}

int main() {
  float* d; cudaMalloc(&d, 128 * 32 * 4);
  cudaFuncSetAttribute(k_test, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024);
  static float h[128 * 32];
  // (shift in pixels, SBO): SBO 512 = contiguous pixels; SBO 1152 = rows of 18 pixels (8-pixel groups 18 pixels apart)
  const int cases[][2] = {{0, 512}, {8, 512}, {1, 512}, {3, 512}, {17, 512}, {0, 1152}, {1, 1152}, {19, 1152}};
  for (auto& cs : cases) {
    k_test<<<1, 128, 48 * 1024>>>(cs[0], cs[1], d);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("shift %d sbo %d: launch failed\n", cs[0], cs[1]); return 1; }
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    int bad = 0; float first_bad = 0, first_exp = 0;
    for (int r = 0; r < 128; ++r) {
      const int P = cs[0] + (r / 8) * (cs[1] / 64) + (r % 8);
      for (int c = 0; c < 32; ++c) {
        const float e = (float)(P % 256) + (float)c / 64.f;
        // bf16 rounding of the stored value
        const float eb = __bfloat162float(__float2bfloat16(e));
        if (h[r * 32 + c] != eb) { if (!bad) { first_bad = h[r * 32 + c]; first_exp = eb; } ++bad; }
      }
    }
    printf("shift %2d pixels, SBO %4d: %s (%d mismatches; first got %.4f expected %.4f)\n", cs[0], cs[1], bad ? "WRONG" : "exact", bad, first_bad, first_exp);
    fflush(stdout);
  }
  {
    float* d2; cudaMalloc(&d2, 128 * 96 * 4);
    cudaFuncSetAttribute(k_test_mn, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024);
    static float h2[128 * 96];
    for (int p0 : {0, 1, 19}) {
      k_test_mn<<<1, 128, 48 * 1024>>>(p0, d2);
      if (cudaDeviceSynchronize() != cudaSuccess) { printf("MN test p0 %d: launch failed\n", p0); return 1; }
      cudaMemcpy(h2, d2, sizeof(h2), cudaMemcpyDeviceToHost);
      int bad = 0; float fb = 0, fe = 0;
      for (int m = 0; m < 16; ++m)
        for (int j = 0; j < 3; ++j)
          for (int c = 0; c < 32; ++c) {
            const float eb = __bfloat162float(__float2bfloat16((float)((p0 + j + m) % 256) + (float)c / 64.f));
            const float got = h2[m * 96 + j * 32 + c];
            if (got != eb) { if (!bad) { fb = got; fe = eb; } ++bad; }
          }
      printf("MN-major B, N = 96, LBO = 64 B (one pixel), first pixel %2d: %s (%d mismatches; first got %.4f expected %.4f)\n", p0, bad ? "WRONG" : "exact", bad, fb, fe);
    }
  }
  return 0;
}
