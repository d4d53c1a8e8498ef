// Micro-benchmark: issue rate of tcgen05.mma (kind::f16, M=128, K=16, SS mode) as a function of N and of the shared-memory
// layout of the A operand.  Timing only (operands are whatever is in shared memory).  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/mma_rate tools/mma_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../distributeddataparallel-cifar10_b200/csrc/tc_common.cuh"

void nrd_set_error(const char*, ...) {}

struct Variant { int layout; int row_bytes; int sbo; int kstep_off; int N; int a_mn; int lbo; int b_mn; int b_lbo; };

__global__ void __launch_bounds__(128, 1) k_rate(Variant v, int nmma, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + 200 * 1024, tslot = bar + 8;
  volatile uint32_t* tslot_gen = reinterpret_cast<volatile uint32_t*>(smem_raw + (tslot - smem_u32(smem_raw)));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0x3c003c00u;
  fence_proxy_async();
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(tslot, 256);
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = *tslot_gen;
  if (warp == 1 && lane == 0) {
    const uint32_t idesc = make_idesc(128, v.N, v.a_mn, v.b_mn);
    const uint32_t a0 = base, b0 = base + 160 * 1024;
    // descriptors precomputed: the timed loop is nothing but tcgen05.mma instructions (18 per round, as one conv tile)
    uint64_t ad[18], bd[18];
#pragma unroll
    for (int i = 0; i < 18; ++i) {
      const int t = i >> 1, kk = i & 1;
      const uint32_t a_addr = a0 + (t % 3) * (18 * 16 * v.row_bytes) + (t / 3) * 16 * v.row_bytes + kk * v.kstep_off;
      ad[i] = make_sdesc(a_addr, v.lbo, v.sbo, v.layout);
      bd[i] = v.b_mn ? make_sdesc(base + 96 * 1024 + i * 1152, v.b_lbo, 512, 4) : make_sdesc(b0 + t * 2048 + kk * 32, 16, 512, 4);
    }
    long long t0 = clock64();
    for (int r = 0; r < nmma / 18; ++r) {
#pragma unroll
      for (int i = 0; i < 18; ++i) umma_bf16(tmem, ad[i], bd[i], idesc, 1);
    }
    umma_commit(bar);
    mbar_wait(bar, 0);
    long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 256); }
}

int main() {
  long long* d; cudaMalloc(&d, 148 * 8);
  cudaFuncSetAttribute(k_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 210 * 1024);
  struct { const char* name; Variant v; } vs[] = {
      {"SW64  K-major rows 64B  N=32 (current C=32)", {4, 64, 512, 32, 32, 0, 16}},
      {"SW32  K-major rows 32B  N=32 (planar halves)", {6, 32, 256, 18 * 16 * 32 * 3, 32, 0, 16}},
      {"SW128 K-major rows 128B N=32", {2, 128, 1024, 32, 32, 0, 16}},
      {"NONE  K-major 8x16B cores N=32 (LBO=128,SBO=256)", {0, 32, 256, 18 * 16 * 32 * 3, 32, 0, 128}},
      {"SW64  K-major N=64", {4, 64, 512, 32, 64, 0, 16}},
      {"SW64  K-major N=96", {4, 64, 512, 32, 96, 0, 16}},
      {"SW64  K-major N=128", {4, 64, 512, 32, 128, 0, 16}},
      {"SW64  K-major N=256", {4, 64, 512, 32, 256, 0, 16}},
      {"SW128 K-major N=64 (current C=64)", {2, 128, 1024, 32, 64, 0, 16}},
      {"SW128 K-major N=256", {2, 128, 1024, 32, 256, 0, 16}},
      {"SW32  K-major N=64", {6, 32, 256, 18 * 16 * 32 * 3, 64, 0, 16}},
      {"SW32  K-major N=128", {6, 32, 256, 18 * 16 * 32 * 3, 128, 0, 16}},
      {"SW64  MN-major A (wgrad, LBO=1024) N=32", {4, 64, 512, 1024, 32, 1, 1024}},
      {"wgrad: MN-major A, MN-major B N=96, groups 18432 B apart", {4, 64, 512, 1024, 96, 1, 1024, 1, 18432}},
      {"wgrad: MN-major A, MN-major B N=96, groups 64 B apart", {4, 64, 512, 1024, 96, 1, 1024, 1, 64}},
      {"wgrad: MN-major A, MN-major B N=96, groups 1152 B apart", {4, 64, 512, 1024, 96, 1, 1024, 1, 1152}},
  };
  const int nmma = 1800;
  for (auto& e : vs) {
    for (int grid : {1, 148}) {
      k_rate<<<grid, 128, 210 * 1024>>>(e.v, nmma, d);
      cudaError_t err = cudaDeviceSynchronize();
      if (err != cudaSuccess) { printf("%s: %s\n", e.name, cudaGetErrorString(err)); return 1; }
      long long h[148]; cudaMemcpy(h, d, grid * 8, cudaMemcpyDeviceToHost);
      long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
      printf("%-52s grid %3d: %7.1f cycles / MMA\n", e.name, grid, (double)mx / nmma);
    }
  }
  return 0;
}
