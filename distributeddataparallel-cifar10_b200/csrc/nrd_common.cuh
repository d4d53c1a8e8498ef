// Shared declarations for libnrd.so (B200 / sm_100a NetResDeep training step).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/nrd.h"

typedef __nv_bfloat16 bf16;

#define NRD_BN_EPS 1e-5f      // nn.BatchNorm2d default (model/resnet.py:28)
#define NRD_BN_MOMENTUM 0.1f  // nn.BatchNorm2d default (model/resnet.py:28)
#define NRD_HW 16             // ResBlock resolution
#define NRD_PIX 256           // 16*16
#define NRD_TILE_PIX 128      // one conv CTA tile: 8 rows x 16 cols of one image
#define NRD_FC1_OUT 32
#define NRD_NCLS 10
#define NRD_NUM_SMS 148

void nrd_set_error(const char* fmt, ...);

#define NRD_CHECK_CUDA(call)                                                                   \
  do {                                                                                         \
    cudaError_t e__ = (call);                                                                  \
    if (e__ != cudaSuccess) {                                                                  \
      nrd_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));     \
      return NRD_ECUDA;                                                                        \
    }                                                                                          \
  } while (0)

#define NRD_REQUIRE(cond, ...)                                                                 \
  do {                                                                                         \
    if (!(cond)) {                                                                             \
      nrd_set_error(__VA_ARGS__);                                                              \
      return NRD_EINVAL;                                                                       \
    }                                                                                          \
  } while (0)

#define NRD_TRY(call)                                                                          \
  do {                                                                                         \
    int r__ = (call);                                                                          \
    if (r__ != NRD_OK) return r__;                                                             \
  } while (0)

// ---- gradient exchange over NVLink peer memory (graph path of DDP, one node, <= 8 ranks) ----------------------------------
// Every rank owns one exchange buffer, mapped into every other rank's address space (CUDA IPC):
//   data   [2 parities][world][nx] floats   slot (parity, r) = what rank r produced in a step of that parity
//   ctrl   u64 words (NRD_XCHG_*): per bucket (body, head) flags [2][NRD_XCHG_MAX_WORLD] = epoch of the data rank r put in
//          slot (parity, r), and the producing kernel's local epoch and arrival counter
// nx = 2C (BN running-stat slot) + all gradients in flat order.  Body bucket: the kernel that finishes the body gradients
// (k_body_reduce) stores every value straight into slot (parity, own rank) of ALL ranks' buffers and its last CTA publishes
// the epoch in every rank's flag.  Head bucket: k_xchg_push_head does the same for fc1/fc2 gradients on the side stream,
// under the body backward.  The optimizer kernel (k_sgd_pack) waits for the world's flags in its OWN buffer and sums the
// slots in rank order: one NVLink write latency instead of a collective launch, no NCCL kernel competing for SMs, and a sum
// order that is the same on every rank.  Two parities: a rank one step ahead never overwrites data a peer is still reading.
#define NRD_XCHG_MAX_WORLD 8
#define NRD_XCHG_BUCKET_WORDS (2 * NRD_XCHG_MAX_WORLD + 2)      // flags[2][8], epoch, arrive
#define NRD_XCHG_EPOCH (2 * NRD_XCHG_MAX_WORLD)
#define NRD_XCHG_ARRIVE (2 * NRD_XCHG_MAX_WORLD + 1)
#define NRD_XCHG_CTRL_WORDS (2 * NRD_XCHG_BUCKET_WORDS)         // bucket 0 = body, bucket 1 = head
struct nrd_xchg_args {
  float* peer[NRD_XCHG_MAX_WORLD];
  int world, rank, nx;          // world == 0: exchange off
  int nll;                      // elements of the body bucket's flag-in-data slots (BN slot + body gradients, rounded up to 4)
  long long timeout_cycles;     // how long a consumer spins for a peer's flag before it gives up (nrd_plan_set_exchange)
};
static __host__ __device__ __forceinline__ size_t nrd_xchg_flags_off(int world, int nx) {
  return (((size_t)2 * world * nx * sizeof(float)) + 127) & ~(size_t)127;
}
// Body bucket with the optimizer step fused into the producing kernel: "flag in data" slots behind the control words --
//   ll   [2 parities][world][nll] x {float value, u32 epoch}
// every 8-byte word carries its own arrival flag, so the producer needs no fence, no arrival counter and no flag store, and
// the consumer's wait is ONE NVLink crossing: it polls the words it sums until each shows this step's epoch.
static __host__ __device__ __forceinline__ size_t nrd_xchg_ll_off(int world, int nx) {
  return (nrd_xchg_flags_off(world, nx) + NRD_XCHG_CTRL_WORDS * sizeof(unsigned long long) + 127) & ~(size_t)127;
}
static __host__ __device__ __forceinline__ size_t nrd_xchg_bytes(int world, int nx, int nll) {
  return nrd_xchg_ll_off(world, nx) + (size_t)2 * world * nll * sizeof(uint2);
}
static __device__ __forceinline__ uint2* nrd_xchg_ll(float* base, const nrd_xchg_args& x, unsigned long long epoch, int from_rank) {
  return reinterpret_cast<uint2*>(reinterpret_cast<char*>(base) + nrd_xchg_ll_off(x.world, x.nx)) + ((size_t)(epoch & 1) * x.world + from_rank) * x.nll;
}
// two consecutive elements {v0, v1} of this rank's slot into every rank's buffer: one 16-byte store per peer, each 8-byte half
// carries the epoch
static __device__ __forceinline__ void nrd_ll_push2(const nrd_xchg_args& x, unsigned long long epoch, size_t idx, float v0, float v1) {
  const unsigned int e = (unsigned int)epoch;
  for (int p = 0; p < x.world; ++p) {
    uint2* d = nrd_xchg_ll(x.peer[p], x, epoch, x.rank) + idx;
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(d), "r"(__float_as_uint(v0)), "r"(e), "r"(__float_as_uint(v1)), "r"(e) : "memory");
  }
}
static __device__ __forceinline__ void nrd_ll_push1(const nrd_xchg_args& x, unsigned long long epoch, size_t idx, float v) {
  const unsigned int e = (unsigned int)epoch;
  for (int p = 0; p < x.world; ++p) {
    uint2* d = nrd_xchg_ll(x.peer[p], x, epoch, x.rank) + idx;
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(d), "r"(__float_as_uint(v)), "r"(e) : "memory");
  }
}
// rank-ordered sum of element idx over the world's slots in OUR buffer, waiting for each word's epoch (all loads of a round in flight)
static __device__ __forceinline__ float nrd_ll_sum(const nrd_xchg_args& x, unsigned long long epoch, size_t idx) {
  const unsigned int e = (unsigned int)epoch;
  const uint2* base = nrd_xchg_ll(x.peer[x.rank], x, epoch, 0) + idx;
  uint2 v[NRD_XCHG_MAX_WORLD];
  long long t0 = 0;
  int spins = 0;
  while (true) {
    bool ok = true;
#pragma unroll
    for (int r = 0; r < NRD_XCHG_MAX_WORLD; ++r)
      if (r < x.world) {
        asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v[r].x), "=r"(v[r].y) : "l"(base + (size_t)r * x.nll) : "memory");
      }
#pragma unroll
    for (int r = 0; r < NRD_XCHG_MAX_WORLD; ++r) if (r < x.world) ok = ok && (v[r].y == e);
    if (ok) break;
    if (++spins > 4096) {
      if (t0 == 0) t0 = clock64();
      else if (clock64() - t0 > x.timeout_cycles) __trap();     // a peer is gone
    }
  }
  float g = 0.f;
#pragma unroll
  for (int r = 0; r < NRD_XCHG_MAX_WORLD; ++r) if (r < x.world) g += __uint_as_float(v[r].x);      // rank order: the same on every rank
  return g;
}
static __device__ __forceinline__ unsigned long long* nrd_xchg_flags(float* base, int world, int nx, int bucket = 0) {
  return reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(base) + nrd_xchg_flags_off(world, nx)) + bucket * NRD_XCHG_BUCKET_WORDS;
}
// last step of a producing kernel (every thread calls it after its peer stores): the CTA that arrives last publishes `epoch` in
// every rank's flag of this bucket and records it locally.  Ordering, one fence per CTA instead of one per thread and one per
// flag: the CTA barrier orders every thread's peer stores before thread 0's system-scope fence (fences are cumulative), which
// precedes the CTA's arrival; the last arriver fences once more (acquire side of the arrival counter) and then writes the
// world's flags as RELAXED stores from `world` lanes at once -- data, fence, flag is the release pattern; a release store per
// flag would put a full system-scope fence (an NVLink round trip, ~2 us) in front of every one of them, one after the other:
// measured 12-18 us between a rank's publish and the arrival of the last peer's flag at 8 ranks (profiles/r02_ddp_tail.md).
static __device__ __forceinline__ void nrd_xchg_publish(const nrd_xchg_args& x, int bucket, unsigned long long epoch) {
  __shared__ int s_last;
  __syncthreads();
  unsigned long long* ctrl = nrd_xchg_flags(x.peer[x.rank], x.world, x.nx, bucket);
  if (threadIdx.x == 0) {
    __threadfence_system();                   // this CTA's peer stores are performed system-wide before it reports in
    unsigned int* arrive = reinterpret_cast<unsigned int*>(ctrl + NRD_XCHG_ARRIVE);
    s_last = (atomicAdd(arrive, 1u) == gridDim.x * gridDim.y - 1) ? 1 : 0;
    if (s_last) { *arrive = 0u; __threadfence_system(); }
  }
  __syncthreads();
  if (s_last) {
    if ((int)threadIdx.x < x.world) {
      unsigned long long* f = nrd_xchg_flags(x.peer[threadIdx.x], x.world, x.nx, bucket) + (epoch & 1) * NRD_XCHG_MAX_WORLD + x.rank;
      asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(f), "l"(epoch) : "memory");
    }
    if (threadIdx.x == 0) ctrl[NRD_XCHG_EPOCH] = epoch;
  }
}
// consumer: threads 0 .. world-1 of the CTA wait for the world's flags of a bucket in the own buffer; returns the epoch
static __device__ __forceinline__ unsigned long long nrd_xchg_wait(const nrd_xchg_args& x, int bucket, int t) {
  unsigned long long* ctrl = nrd_xchg_flags(x.peer[x.rank], x.world, x.nx, bucket);
  const unsigned long long epoch = ctrl[NRD_XCHG_EPOCH];       // set by this step's producer (ordered before us by the stream)
  if (t >= 0 && t < x.world) {
    const unsigned long long* f = ctrl + (epoch & 1) * NRD_XCHG_MAX_WORLD + t;
    unsigned long long v;
    long long t0 = 0;
    int spins = 0;
    while (true) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
      if (v >= epoch) break;
      if (++spins > 4096) {
        if (t0 == 0) t0 = clock64();
        else if (clock64() - t0 > x.timeout_cycles) __trap();   // default ~10 min (NCCL's default watchdog): a peer is gone
      }
    }
  }
  return epoch;
}

// the same wait when the caller knows the epoch itself (a kernel that is both producer and consumer of the bucket)
static __device__ __forceinline__ void nrd_xchg_wait_epoch(const nrd_xchg_args& x, int bucket, int t, unsigned long long epoch) {
  if (t >= 0 && t < x.world) {
    const unsigned long long* f = nrd_xchg_flags(x.peer[x.rank], x.world, x.nx, bucket) + (epoch & 1) * NRD_XCHG_MAX_WORLD + t;
    unsigned long long v;
    long long t0 = 0;
    int spins = 0;
    while (true) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
      if (v >= epoch) break;
      if (++spins > 4096) {
        if (t0 == 0) t0 = clock64();
        else if (clock64() - t0 > x.timeout_cycles) __trap();
      }
    }
  }
}

struct nrd_plan {
  int device, B, C, NB, dtype, conv_impl;
  size_t esz;       // activation element size
  size_t act_elems; // B*256*C
  float *params, *grads, *bn_running;
  int64_t* nbt;
  int64_t off[10];
  const float* images_last;  // images of the last nrd_forward (stem backward reads them)
  const float* dp_for;       // dlogits pointer whose d(pooled) nrd_forward_loss already computed (consumed by nrd_backward_body)
  int fuse_stem_bwd;         // option: the persistent backward also computes conv1's weight / bias gradient partials (no stem backward kernel)
  int fuse_stem;             // option: the persistent forward computes x_0 itself (no stem kernel, no x_0 round trip)
  int fuse_pool;             // option: the persistent forward's last block also does the max-pool in front of fc1
  int fuse_head_tail;        // option: fc2 + loss + fc1 dgrad in one launch inside nrd_forward_loss
  const char* dx0_last;      // dL/dx_0 of the last nrd_backward_body (which of the two dx scratch buffers holds it)
  // workspace carve-up
  char* xs;            // (NB+1) saved block inputs x_k, NHWC T
  char* ys;            // NB saved conv outputs y_k, NHWC T
  char* dy;            // dL/dy_k scratch
  char* dxa;           // dL/dx ping
  char* dxb;           // dL/dx pong
  uint8_t* idx0;       // stem pool argmax [B,16,16,C]
  char* pooled;        // [B,8,8,C] T
  uint8_t* idx2;       // [B,8,8,C]
  float* h1;           // [B,32] relu(fc1)
  float* dh1;          // [B,32]
  float* fc1_part;     // [K/128][B][32] split-K partial products of fc1
  float* fc1w_partial; // [splits][32][64*C] batch-split partials of the fc1 weight gradient
  float* dlogits;      // [B,10]  (train_step path)
  float* loss_rows;    // [B]
  float* saved_stats;  // [NB][2C] mean | invstd
  float* scale_shift;  // [NB][2C]
  float* stats_partial;// [B*2][2C]
  float* bnbwd_scratch;
  float* wgrad_partial;
  float* stem_partial;
  float* w_fwd;        // [9][C][C] tap-major fp32 (ci rows, co contiguous)
  float* w_dgrad;      // [9][C][C] flipped taps, (co rows, ci contiguous)
  float* w1p;          // fc1 weight, k-major transpose in NHWC column order: w1t[64*C][32]
  float* conv1_wp;     // [27][C]
  bf16* wtc_fwd;       // tcgen05 B operand: [Cout][9*Cin] bf16, K-major
  bf16* wtc_dgrad;     // same for dgrad (transposed + flipped)
  // TMA descriptors for the tcgen05 path (filled by nrd_tc_plan_init)
  CUtensorMap tm_x[32];   // per saved x_k (k <= NB), box [1,8,16,C]
  CUtensorMap tm_dy;      // dy scratch, row-haloed box (dgrad)
  CUtensorMap tm_dy0;     // dy scratch, plain rows (wgrad)
  CUtensorMap tm_wf, tm_wd;
  CUtensorMap tm_xall;    // all x_k as one 5-D tensor (persistent backward)
  CUtensorMap tm_x18;     // x_0 with an 18 x 18 (haloed) box: the one-copy operand layout
  int one_copy;           // option (bit 0 forward, bit 1 backward): ONE haloed operand copy per image instead of three shifted ones
  CUtensorMap tm_xst, tm_yst; // x_k / y_k stacks, whole-image boxes: bulk tensor STORES of the persistent forward
  int tc_ready;
  nrd_xchg_args xchg;     // peer-memory gradient exchange (nrd_plan_set_exchange); world == 0: off
  int poll_ns;            // experiment knob (env NRD_POLL_NS at plan creation): back-off between polls of the BN exchange
  int pdl;                // option: programmatic dependent launch between the small kernels of the step
  int reserve_sms;        // option: SMs the cooperative kernels leave free (a concurrent NCCL kernel needs somewhere to run)
  // NRD_OPT_TIMELINE: an event after every kernel of the step's main stream (event-record NODES when the step is captured), so
  // that the replayed step can be read kernel by kernel in its real context (warm L2, neighbours running)
  int tl_on, tl_n;
  cudaEvent_t tl_ev[40];
  const char* tl_name[40];
  float upd_lr, upd_inv_world; // nrd_plan_set_fused_update: the optimizer step rides nrd_backward_head / _body (upd_lr > 0)
  unsigned long long* head_ctr; // per-row-group arrival counters of the fused head kernel
  float* loss_mirror; unsigned int* seq_mirror;   // pinned host words the head kernel also writes the loss / its sequence number to
  const int64_t* labels_last;   // labels of the last nrd_forward_loss (nrd_plan_run_op replays the head with them)
  float *logits_last, *loss_last, *dlogits_last, *loss_rows_last;
  int head_ctas;          // how many CTAs of the head kernel fit on the device at once (cooperative launch)
  float* bn_sync;         // optional [2C] slot in front of the gradient vector: BN running stats ride the body bucket's all-reduce
  float bn_sync_contrib;  // what this rank adds to the slot: world_size on rank 0, 0 elsewhere
  float bn_sync_inv_world;
  int tc_stem;            // tensor-core stem kernels available (bf16, C = 32)
  int use_tc_stem;        // option
  int persist_fwd;        // option: use the persistent block-stack forward when available
  int persist_bwd;        // option: use the persistent block-stack backward when available
  int fuse_bn_bwd;        // BN+ReLU backward as one cooperative kernel when it fits shared memory
  int tc_persist_fwd;     // all NB blocks of the forward / backward in ONE cooperative kernel each (C = 32, NB >= 2)
  int tc_stream;          // ... as the STREAMING kernels (any batch size) instead of the resident ones (<= 2 images per CTA)
  int stream_opt;         // NRD_OPT_STREAM: 0 = streaming kernels only beyond 2 images per CTA, 1 = always, 2 = never (per-block kernels there)
  CUtensorMap tm_x18all;  // all x_k, 18 x 18 (haloed) boxes: the streaming forward's operand loads
  int tc_fused_fwd;       // whole ResBlock forward in one cooperative kernel (accumulators fit TMEM)
  unsigned long long* bn_acc; // BN-sum words of the persistent kernels: fwd [32 blocks][NRD_ACC_REPL][64] then bwd, same shape;
                              // zero between launches (the kernels clean up behind themselves)
  unsigned long long* gbar; // grid-barrier counters (one per cooperative kernel type)
  void* trace;            // optional device buffer for in-kernel clock64 stamps (profiling aid; NULL = off)
  int dbg_skip_stores;    // profiling experiment (NRD_OPT_TRACE value 2)
  long long* dbg_buf;     // [64 rows][16 stamps][clock64, globaltimer]
  long long* dbg;         // in-kernel stamps of the small kernels (nrd_dbg_stamp; NULL = off; what = 11 in nrd_plan_tensor)
  void* trace_bwd;        // same for the persistent backward (lives in the pooled-gradient-free part of the dy scratch)
};

// Programmatic dependent launch: a kernel launched with nrd_launch_pdl() may start (be scheduled, run its prologue) while its
// predecessor on the stream is still running; it must call nrd_pdl_wait() before touching anything the predecessor (or any
// kernel before that) writes.  Producers call nrd_pdl_launch() as early as they like: it only lets the dependent grid be
// scheduled.  Without the launch attribute both are no-ops.  Stream capture keeps the programmatic edge in the graph.
static __device__ __forceinline__ void nrd_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
static __device__ __forceinline__ void nrd_pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <typename... KArgs, typename... Args>
static inline cudaError_t nrd_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// Grid-wide barrier for cooperative launches (all CTAs co-resident).  ONE 64-bit monotonic counter per kernel type and plan:
// an arrival is a release-add of 1, waiting is an acquire-poll until the counter reaches this episode's target.  The counter
// is never reset (a given kernel type always runs with the same grid, so it is a multiple of the grid size between launches);
// nrd_grid_barrier_base() recovers the launch's base value even if faster CTAs have already arrived at the first barrier.
// Call from ONE thread per CTA, bracketed by CTA-level barriers.  A spin that lasts ~2 s traps instead of hanging the GPU.
static __device__ __forceinline__ unsigned long long nrd_grid_barrier_base(const unsigned long long* ctr, unsigned int nblocks) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(ctr) : "memory");
  return v - v % nblocks;
}
static __device__ __forceinline__ void nrd_grid_barrier(unsigned long long* ctr, unsigned long long target) {
  asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(ctr) : "memory");
  unsigned long long v;
  long long t0 = 0;
  int spins = 0;
  while (true) {
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(ctr) : "memory");
    if (v >= target) break;
    if (++spins > 1024) {
      if (t0 == 0) t0 = clock64();
      else if (clock64() - t0 > 4000000000LL) __trap();
    }
  }
}
// DDP buffer sync folded into the gradient exchange (nrd_plan_set_bn_sync): the stem forward, the first kernel of a training
// step, overwrites the running stats with rank 0's values of the previous step (slot = world_size * value after the
// all-reduce), exactly where DDP's broadcast in front of the forward would.
struct nrd_bn_sync_args { const float* slot; float* running; float inv_world; int n; };
static __device__ __forceinline__ void nrd_bn_sync_apply(const nrd_bn_sync_args& a) {
  if (a.slot && blockIdx.x == 0 && blockIdx.y == 0)
    for (int j = threadIdx.x; j < a.n; j += blockDim.x) a.running[j] = a.slot[j] * a.inv_world;
}
static inline nrd_bn_sync_args nrd_bn_sync_of(const nrd_plan* p, int training) {
  nrd_bn_sync_args a = {nullptr, nullptr, 0.f, 0};
  if (p && training && p->bn_sync) { a.slot = p->bn_sync; a.running = p->bn_running; a.inv_world = p->bn_sync_inv_world; a.n = 2 * p->C; }
  return a;
}
// BN sums of the persistent kernels: ONE 64-bit word per channel statistic carries both the cross-CTA sum and the arrival
// count, so waiting for "all CTAs have contributed" and fetching the sum are the same load, and no CTA waits for its own
// atomics to be acknowledged before it signals (that is what a separate barrier counter costs).
//   bits 56..63  number of CTAs that have added (grid <= 255)
//   bits  0..55  sum of (fixed-point contribution + NRD_ACC_BIAS): the bias keeps every contribution positive so the low
//                field never borrows from / carries into the count; integer addition = order-independent, bit-exact.
// A contribution is clamped to +-(2^46 - 1) fixed-point units (see the scale at each use).
#define NRD_ACC_CNT1 (1ull << 56)
#define NRD_ACC_BIAS (1ull << 46)
// The atomic RETURNS the old word on purpose: measured (tools/barrier_rate.cu), a round of returning atomics completes in
// ~2050 cycles on 148 CTAs against ~3000 for fire-and-forget reductions (REDs seem to drain behind the polling loads).  The
// returned count doubles as a check that the row was clean (a dirty row would otherwise make every CTA spin to the trap).
// `status` is the plan's sticky status word (NRD_STATUS_*, read by nrd_plan_status): a contribution that does not fit (or is
// not finite) is still clamped -- the count field must stay intact or every CTA would spin to the trap -- but the step is
// marked, so wrong statistics never pass silently.
#define NRD_STATUS_BN_SATURATED 1u     // a per-CTA BN sum left the fixed-point range (forward: |sum| >= 2^22, backward: >= 2^10)
#define NRD_STATUS_BN_NONFINITE 2u     // a per-CTA BN sum was NaN / Inf
static __device__ __forceinline__ unsigned long long nrd_acc_add(unsigned long long* word, float v, double scale, unsigned int* status) {
  const long long lim = (long long)NRD_ACC_BIAS - 1;
  long long f;
  if (!isfinite(v)) { f = 0; atomicOr(status, NRD_STATUS_BN_NONFINITE); }
  else {
    f = __double2ll_rn((double)v * scale);
    if (f > lim || f < -lim) { f = f > lim ? lim : -lim; atomicOr(status, NRD_STATUS_BN_SATURATED); }
  }
  return atomicAdd(word, NRD_ACC_CNT1 + NRD_ACC_BIAS + (unsigned long long)f);
}
// checking right away, i.e. waiting for the atomic's return before the first poll, measured 2.5 us per kernel FASTER than
// polling beside the outstanding atomic
static __device__ __forceinline__ void nrd_acc_check(unsigned long long old, unsigned int nblocks) {
  if ((old >> 56) >= nblocks) __trap();
}
static __device__ __forceinline__ unsigned long long nrd_acc_peek(const unsigned long long* word) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(word) : "memory");
  return v;
}
static __device__ __forceinline__ double nrd_acc_value(unsigned long long w, unsigned int nblocks, double inv_scale) {
  return (double)((long long)w - (long long)((unsigned long long)nblocks * NRD_ACC_BIAS)) * inv_scale;   // w: summed low fields
}
// Atomics on ONE address retire at ~27 cycles each in the L2 slice: 108 CTAs arriving together would queue for ~3000 cycles.
// Every statistic therefore has NRD_ACC_REPL replica words (CTA i adds to replica i % NRD_ACC_REPL, rows of `stride` words
// apart); the waiting warp reads all replicas, adds the counts and the low fields (still pure integer arithmetic).
#define NRD_ACC_REPL 1
// words of one row are NRD_ACC_WS words (128 B) apart: every statistic has its own L2 line, so the ~9500 atomics that
// arrive within a few hundred cycles spread over 64 lines / L2 slices instead of queueing on 4
#define NRD_ACC_WS 16
#define NRD_ACC_IDX(r, e) (((r) * 64 + (e)) * NRD_ACC_WS)
#define NRD_ACC_ROW (NRD_ACC_REPL * 64 * NRD_ACC_WS)          // words per block row
// one warp waits until both of each lane's statistics are complete (all lanes leave together); a ~2 s spin traps.
// v0 / v1 return the summed low fields (feed them to nrd_acc_value).
static __device__ __forceinline__ void nrd_acc_wait2(const unsigned long long* w0, const unsigned long long* w1,
                                                     unsigned int nblocks, unsigned long long& v0, unsigned long long& v1, int poll_ns = 0) {
  long long t0 = 0;
  int spins = 0;
  while (true) {
    unsigned long long r0[NRD_ACC_REPL], r1[NRD_ACC_REPL];
#pragma unroll
    for (int r = 0; r < NRD_ACC_REPL; ++r) { r0[r] = nrd_acc_peek(w0 + NRD_ACC_IDX(r, 0)); r1[r] = nrd_acc_peek(w1 + NRD_ACC_IDX(r, 0)); }
    unsigned long long c0 = 0, c1 = 0;
    v0 = 0; v1 = 0;
#pragma unroll
    for (int r = 0; r < NRD_ACC_REPL; ++r) {
      c0 += r0[r] >> 56; c1 += r1[r] >> 56;
      v0 += r0[r] & (NRD_ACC_CNT1 - 1); v1 += r1[r] & (NRD_ACC_CNT1 - 1);
    }
    const bool done = c0 == nblocks && c1 == nblocks;
    if (__all_sync(0xffffffffu, done)) break;
    if (poll_ns) __nanosleep(poll_ns);
    if (++spins > 1024) {
      if (t0 == 0) t0 = clock64();
      else if (clock64() - t0 > 4000000000LL) __trap();
    }
  }
}
#define NRD_GBAR_BLOCK_FWD 0
#define NRD_GBAR_PERSIST_FWD 1
#define NRD_GBAR_BN_BWD 2
#define NRD_GBAR_PERSIST_BWD 3
#define NRD_GBAR_HEAD_COUNTER 16       // word of the gbar block used as the head kernels' arrival counter
#define NRD_GBAR_STATUS 24             // word of the gbar block that holds the sticky status flags
#define NRD_GBAR_LOSS_SEQ 25           // ... the number of losses written to the host mirror so far (nrd_plan_set_loss_mirror)

// in-kernel stamps (profiling aid, NRD_OPT_TRACE): dbg = NULL turns them off.  Row `row` of the buffer = 16 x {clock64, globaltimer}.
static __device__ __forceinline__ void nrd_dbg_stamp(long long* dbg, int row, int slot) {
  if (dbg) {
    unsigned long long gt;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
    dbg[(row * 16 + slot) * 2] = clock64(); dbg[(row * 16 + slot) * 2 + 1] = (long long)gt;
  }
}

// timeline mark: record the next event of the plan's timeline on `st` behind the kernel that was just launched
static inline void nrd_tl_mark(nrd_plan* p, cudaStream_t st, const char* name, bool restart = false) {
  if (!p->tl_on) return;
  if (restart) p->tl_n = 0;
  if (p->tl_n >= 40) return;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(st, &cs);
  cudaEventRecordWithFlags(p->tl_ev[p->tl_n], st, cs == cudaStreamCaptureStatusActive ? cudaEventRecordExternal : cudaEventRecordDefault);
  p->tl_name[p->tl_n++] = name;
}

// tcgen05 path (conv_tc.cu)
int nrd_tc_plan_init(nrd_plan* p);
void nrd_tc_refresh_limits(nrd_plan* p);
int nrd_tc_conv_fwd(nrd_plan* p, int k, cudaStream_t st);                       // y_k = conv(x_k) + stats partials
int nrd_tc_conv_dgrad(nrd_plan* p, const void* g, void* dx_out, cudaStream_t st); // dx = dgrad(dy) + g
int nrd_tc_conv_wgrad(nrd_plan* p, int k, int first, cudaStream_t st);          // partial (+)= wgrad(x_k, dy)
int nrd_tc_stem_init(nrd_plan* p);
int nrd_tc_stem_fwd(nrd_plan* p, const float* images, int training, cudaStream_t st);
int nrd_tc_stem_bwd(nrd_plan* p, const float* images, const void* g0, cudaStream_t st);   // per-CTA partials only
int nrd_tc_stats_blocks(const nrd_plan* p);
int nrd_tc_wgrad_blocks(const nrd_plan* p);
int nrd_tc_forward_blocks(nrd_plan* p, int pool, const float* stem_images, int training, cudaStream_t st);   // x_0 -> all y_k, x_{k+1}
                                                                                // (stem_images: compute x_0 from the images first)
                                                                                // (pool: also maxpool2(x_NB) -> pooled, idx2)
int nrd_tc_backward_blocks(nrd_plan* p, const void* g_pooled, void* dx0_out, const float* stem_images, cudaStream_t st);   // all blocks' backward, one launch
                                                                            // (stem_images: also conv1's weight-gradient partials, no stem kernel);
                                                                            // g_pooled = dL/d(pooled) [B,8,8,C]
int nrd_tc_block_fwd(nrd_plan* p, int k, cudaStream_t st);                      // y_k, x_{k+1}, BN stats + EMA, one launch
// persistent stacks usable for this plan?  (the streaming backward exists from NRD_HAVE_BWD_STREAM on)
#define NRD_HAVE_BWD_STREAM 1
static inline bool nrd_persist_fwd_ok(const nrd_plan* p) { return p->tc_persist_fwd != 0; }
static inline bool nrd_persist_bwd_ok(const nrd_plan* p) { return p->tc_persist_fwd != 0 && (!p->tc_stream || NRD_HAVE_BWD_STREAM); }
