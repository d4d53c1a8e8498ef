// Micro-benchmark: round time of "every CTA adds to shared words, then waits until all CTAs have" (the BN-sum exchange of
// the persistent kernels) for different numbers of words per CTA and word spacings.  One CTA per SM (cooperative-like:
// 148 CTAs of 64 threads are always co-resident).  Build:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/barrier_rate tools/barrier_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long peek(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// mode 0: nwords atomics per CTA (threads 0..nwords-1), lane l of warp 0 polls words l and l+32 (if nwords == 64) or word l
// mode 1: ONE atomic per CTA (arrival counter only), one polling thread: the floor of any atomic+poll exchange
__global__ void __launch_bounds__(64, 1) k_round(unsigned long long* acc, int rounds, int nwords, int ws, int mode, long long* out) {
  const int t = threadIdx.x;
  __shared__ int dummy;
  long long t0 = clock64();
  for (int r = 0; r < rounds; ++r) {
    unsigned long long* row = acc + (size_t)r * 64 * ws;
    if (mode == 0) {
      if (t < nwords) atomicAdd(row + t * ws, (1ull << 56) + 12345ull);
      if (t < 32) {
        while (true) {
          const int w0 = t < nwords ? t : 0;
          unsigned long long a = peek(row + w0 * ws), b = nwords == 64 ? peek(row + (t + 32) * ws) : a;
          const bool done = (a >> 56) == gridDim.x && (b >> 56) == gridDim.x;
          if (__all_sync(0xffffffffu, done)) break;
        }
      }
    } else if (mode == 1) {
      if (t == 0) {
        atomicAdd(row, 1ull);
        while (peek(row) < gridDim.x) {}
      }
    } else if (mode == 3) {
      // data + count in every word, but only ONE lane polls (word 0) while waiting; then every lane fetches its words
      if (t < nwords) atomicAdd(row + t * ws, (1ull << 56) + 12345ull);
      if (t < 32) {
        if (t == 0) while ((peek(row) >> 56) != gridDim.x) {}
        __syncwarp();
        while (true) {
          unsigned long long a = peek(row + t * ws), b = peek(row + (t + 32) * ws);
          const bool done = (a >> 56) == gridDim.x && (b >> 56) == gridDim.x;
          if (__all_sync(0xffffffffu, done)) break;
        }
      }
    } else if (mode == 4) {
      // 64 polling threads, one word each
      atomicAdd(row + t * ws, (1ull << 56) + 12345ull);
      while ((peek(row + t * ws) >> 56) != gridDim.x) {}
    } else if (mode == 5) {
      // returning atomics (ATOM) and the returned count tells the LAST contributor immediately; everyone still polls
      const unsigned long long old = atomicAdd(row + t * ws, (1ull << 56) + 12345ull);
      if ((old >> 56) != gridDim.x - 1) while ((peek(row + t * ws) >> 56) != gridDim.x) {}
    }
    __syncthreads();
  }
  if (t == 0) { out[blockIdx.x] = clock64() - t0; dummy = 0; }
}

int main() {
  const int rounds = 200;
  unsigned long long* acc; long long* out;
  const size_t bytes = (size_t)rounds * 64 * 16 * 8;
  cudaMalloc(&acc, bytes); cudaMalloc(&out, 148 * 8);
  struct { const char* name; int nwords, ws, mode, grid; } vs[] = {
      {"64 words/CTA, packed (8 B apart), 148 CTAs", 64, 1, 0, 148},
      {"64 words/CTA, one line each (128 B apart), 148 CTAs", 64, 16, 0, 148},
      {"32 words/CTA, one line each, 148 CTAs", 32, 16, 0, 148},
      {"1 word/CTA (counter only), 148 CTAs", 1, 16, 1, 148},
      {"64 words/CTA, one line each, 16 CTAs", 64, 16, 0, 16},
      {"mode 3: one lane polls word 0, then all fetch, 148 CTAs", 64, 16, 3, 148},
      {"mode 4: 64 polling threads, one word each, 148 CTAs", 64, 16, 4, 148},
      {"mode 5: returning atomics + 64 pollers, 148 CTAs", 64, 16, 5, 148},
      {"mode 4, 2 CTAs", 64, 16, 4, 2},
      {"1 word/CTA (counter only), 2 CTAs", 1, 16, 1, 2},
  };
  for (auto& v : vs) {
    cudaMemset(acc, 0, bytes);
    void* args[] = {&acc, (void*)&rounds, &v.nwords, &v.ws, &v.mode, &out};
    cudaError_t e = cudaLaunchCooperativeKernel((const void*)k_round, dim3(v.grid), dim3(64), args, 0, 0);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", v.name, cudaGetErrorString(e)); return 1; }
    long long h[148]; cudaMemcpy(h, out, v.grid * 8, cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < v.grid; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("%-56s %7.0f cycles / round\n", v.name, (double)mx / rounds);
    fflush(stdout);
  }
  return 0;
}
