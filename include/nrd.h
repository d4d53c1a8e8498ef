/* nrd.h -- C ABI of libnrd.so, the B200 (sm_100a) NetResDeep data-parallel training step.
 *
 * The reference (BaamPark/DistributedDataParallel-Cifar10) has no FFI: its hot path is a
 * chain of PyTorch calls.  This header is the boundary a maintainer would bind instead
 * (ctypes stub in INTEGRATION.md); each entry names the reference call site it replaces
 * (file:line relative to the reference tree).
 *
 * Conventions
 *  - every entry returns 0 on success or a negative NRD_E* code; the message is in
 *    nrd_last_error() (thread-local);
 *  - all pointers are DEVICE pointers unless said otherwise; the library never allocates
 *    or frees caller tensors -- scratch comes from a caller-provided workspace whose size
 *    nrd_workspace_bytes() reports;
 *  - every launch goes to the `stream` argument (a cudaStream_t passed as void*), is
 *    asynchronous to the host and capturable in a CUDA graph; entries may be called from
 *    any host thread (autograd runs backward on its own thread) -- they cudaSetDevice()
 *    to the plan's device first;
 *  - activations are NHWC (pixels x channels); `dtype` is the activation storage type and
 *    the tensor-core operand type (NRD_F32: validation mode, fp32 everywhere, CUDA cores;
 *    NRD_BF16: bf16 storage + tcgen05 tensor cores, fp32 accumulate).  Parameters,
 *    gradients, BN statistics and the loss are always fp32 in PyTorch's own layouts
 *    (conv OIHW, fc1 columns in (c,h,w) order) so they alias the nn.Module's tensors.
 */
#ifndef NRD_H_
#define NRD_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NRD_F32 0
#define NRD_BF16 1

#define NRD_OK 0
#define NRD_EINVAL (-1)   /* bad argument (shape, dtype, null pointer, workspace too small) */
#define NRD_ECUDA (-2)    /* a CUDA runtime/driver call failed */
#define NRD_EUNSUPPORTED (-3)
#define NRD_ERANGE (-4)    /* nrd_plan_status: a kernel of the plan clamped or saw non-finite BatchNorm sums */

/* conv implementation selector for NRD_BF16 plans (nrd_plan_set_conv_impl) */
#define NRD_CONV_CUDA_CORE 0  /* direct conv on CUDA cores (also the only NRD_F32 path) */
#define NRD_CONV_TCGEN05 1    /* implicit GEMM on tcgen05 tensor cores, TMA-staged NHWC tiles; ResBlock forward fused
                                 (conv + BN statistics + grid barrier + BN apply + ReLU + residual) when it fits TMEM */
#define NRD_CONV_TCGEN05_UNFUSED 2 /* same tensor-core kernels, BN as separate kernels (any batch size; cross-check) */

typedef struct nrd_plan nrd_plan; /* opaque: shapes + carved workspace + TMA descriptors */

const char* nrd_last_error(void);
int nrd_version(void);

/* ---- flat parameter vector --------------------------------------------------------------
 * The 9 unique parameter tensors of NetResDeep (model/resnet.py:9-13,27-28) live in ONE flat
 * fp32 vector (gradients in a second one of the same shape) so that the bucketed allreduce
 * and the fused update need no gather.  Order: conv1.weight, conv1.bias, resblock conv.weight,
 * bn.weight, bn.bias  [= bucket "body"]  then fc1.weight, fc1.bias, fc2.weight, fc2.bias
 * [= bucket "head"].  offsets[i] is the first float of tensor i, offsets[9] the total. */
int nrd_param_offsets(int C, int64_t offsets[10]);

/* ---- plan -------------------------------------------------------------------------------- */
size_t nrd_workspace_bytes(int B, int C, int NB, int dtype);
/* params/grads: flat fp32 vectors (above).  bn_running: [2*C] fp32 = running_mean|running_var,
 * nbt: int64 scalar (BatchNorm2d buffers, model/resnet.py:28).  workspace must be 256-B aligned. */
int nrd_plan_create(nrd_plan** out, int device, int B, int C, int NB, int dtype,
                    void* workspace, size_t workspace_bytes,
                    float* params, float* grads, float* bn_running, int64_t* nbt);
int nrd_plan_destroy(nrd_plan* plan);
int nrd_plan_set_conv_impl(nrd_plan* plan, int impl);
int nrd_plan_conv_impl(const nrd_plan* plan); /* current selector */
/* options: NRD_OPT_FUSE_BN_BWD (default 1): BN+ReLU backward as ONE cooperative kernel (sums, grid barrier, dy) when the
 * CTA's images fit shared memory; 0 = the three-kernel path (any batch size; cross-check). */
#define NRD_OPT_FUSE_BN_BWD 0
/* NRD_OPT_PERSIST_FWD (default 1): run ALL residual blocks of the forward in one persistent cooperative kernel that keeps
 * each image's activations in shared memory / TMEM between blocks (C = 32, B <= 2 x (148 - NRD_OPT_RESERVE_SMS)); 0 = one fused kernel per block. */
#define NRD_OPT_PERSIST_FWD 1
/* NRD_OPT_TRACE (default 0): profiling aid -- the persistent kernels write clock64 stamps of their phases into the dx pong
 * scratch (forward; nrd_plan_tensor what=6) and the dy scratch (backward; what=2): [cta][block][8] int64.  Never enable in
 * a real step when B*NB*64 bytes could overlap live data (per-block kernels use those buffers). */
#define NRD_OPT_TRACE 2
/* NRD_OPT_PERSIST_BWD (default 1): the same for the backward of the residual blocks: dL/dx stays in registers, dy in
 * shared memory, the weight-gradient accumulators in TMEM for the whole stack; 0 = per-block kernels. */
#define NRD_OPT_PERSIST_BWD 3
/* NRD_OPT_TC_STEM (default 1): stem forward / weight gradient as tcgen05 implicit GEMMs with the im2col operand built in
 * shared memory (bf16 plans, C = 32); 0 = the CUDA-core stem kernels. */
#define NRD_OPT_TC_STEM 4
/* NRD_OPT_RESERVE_SMS (default 0): the cooperative kernels use 148 - value SMs.  They fill an SM's shared memory, so anything
 * launched beside them (side-stream kernels: head weight gradients, the head bucket's all-reduce or push; programmatically
 * launched successors) can only run on SMs they leave free.  A caller that overlaps such work sets 8 (graph.py does, on one GPU
 * and on DDP ranks: 195 -> 184 us per step at B = 256); a caller that runs nrd_train_step on ONE stream gains nothing from it. */
#define NRD_OPT_RESERVE_SMS 5
/* NRD_OPT_PDL (bit mask, default 5): kernels launched with programmatic stream serialization, i.e. already resident with
 * their prologue done when the predecessor's last CTA retires.  bit 0: loss and fc1-dgrad kernels, bit 2: body gradient
 * reduce and optimizer kernels, bit 1: the kernels right behind the two persistent kernels (measured slower: off). */
#define NRD_OPT_PDL 6
/* NRD_OPT_HEAD_TAIL (default 1): nrd_forward_loss also computes the backward's first product, dL/d(pooled), in the same
 * launch as fc2 and the loss; nrd_backward_body(dlogits) with the pointer nrd_forward_loss filled then starts at the blocks. */
#define NRD_OPT_HEAD_TAIL 7
/* NRD_OPT_FUSE_POOL (default 0: measured 1 us per step slower): the persistent forward max-pools its last block's output itself (warp shuffles over the
 * 2x2 window) and fc1 reads the pooled tensor, instead of a head kernel re-reading x_NB. */
#define NRD_OPT_FUSE_POOL 8
/* NRD_OPT_FUSE_STEM (default 0: bit-identical, measured no faster than the stem kernel it replaces): the persistent forward computes x_0 = maxpool2(relu(conv1(images)+b)) itself (im2col tiles
 * built by its 512 epilogue threads, 8 tcgen05.mma per half image, result written straight into the operand copies):
 * no stem launch, no x_0 load. */
#define NRD_OPT_FUSE_STEM 9
/* NRD_OPT_ONE_COPY (bit mask, default 1): bit 0 = the persistent forward, bit 1 = the persistent backward keep ONE row- and
 * column-haloed operand copy per image (a tap is a pixel-granular descriptor start) instead of three kx-shifted copies.
 * Measured: forward -1 us; backward +2 us (its weight-gradient MMAs read N groups one pixel apart), hence off there. */
#define NRD_OPT_ONE_COPY 10
/* NRD_OPT_FUSE_STEM_BWD (default 0: bit-identical, measured no faster -- 195.1 vs 194.4 us per step): the persistent backward also computes conv1's weight / bias gradient (model/resnet.py:16 under
 * main.py:38) of its images in its tail -- dL/dx_0 is in its registers, x_0 in its shared memory -- as the same tcgen05 GEMM
 * over the same tiles, in the same order as the stem backward kernel it replaces (bit-identical partials at equal grids);
 * 0 = the stand-alone stem backward kernel. */
#define NRD_OPT_FUSE_STEM_BWD 12
/* NRD_OPT_STREAM (default 0): which kernels run the residual blocks of a bf16, C = 32 plan.  0: the resident persistent kernels up to
 * 2 images per CTA (B <= 296 on 148 SMs), the STREAMING persistent kernels beyond (any batch size: every image passes through two
 * operand slots per block, one launch for the whole stack); 1: the streaming kernels always (cross-check: bit-identical to the
 * resident ones); 2: never (per-block kernels beyond 2 images per CTA, as in round 1). */
#define NRD_OPT_STREAM 13
/* NRD_OPT_TIMELINE (default 0): measurement aid.  An event is recorded behind every kernel of the step's main stream -- an
 * event-record NODE when the step is being captured -- and nrd_plan_timeline() reads the intervals after a (replayed) step:
 * the step kernel by kernel in its real context (warm L2, side stream running).  The marks serialise neighbouring kernels
 * (no programmatic overlap across a mark), so the marked step is a little slower than the unmarked one. */
#define NRD_OPT_TIMELINE 11
int nrd_plan_set_option(nrd_plan* plan, int option, int value);
/* Sticky status of the plan's kernels.  The persistent kernels add their per-CTA BatchNorm sums (model/resnet.py:35 and its
 * backward) as 64-bit fixed point; a contribution outside the representable range (forward |sum| >= 2^22, backward >= 2^10)
 * or a NaN/Inf is clamped so that the grid keeps running, and marked here: NRD_STATUS bit 0 = clamped, bit 1 = non-finite.
 * nrd_plan_status synchronises `stream`, copies the word to *flags_out (HOST pointer), clears it when clear != 0 and returns
 * NRD_ERANGE (message in nrd_last_error) if any bit was set.  The reference has no counterpart (PyTorch would propagate
 * Inf/NaN into the loss); call it where the reference reads the loss on the host (main.py:41,44). */
int nrd_plan_status(nrd_plan* plan, unsigned int* flags_out, int clear, void* stream);
/* DDP's per-forward buffer broadcast (torch DDP `broadcast_buffers`, reached from main.py:63 and run in front of every
 * forward, main.py:33) folded into the gradient exchange: `slot` = the 2C floats directly in front of `grads` (so slot +
 * body bucket are ONE contiguous all-reduce), `contrib` = world_size on rank 0 and 0 elsewhere, `inv_world` = 1/world_size.
 * nrd_backward_body() writes running_stats * contrib into the slot; after the SUM all-reduce it holds world_size * (rank
 * 0's stats) on every rank, and the next training nrd_forward() starts by storing slot * inv_world into the running
 * stats -- what the broadcast would have delivered, bit-exact for a power-of-two world size.  Between steps every rank
 * still sees its OWN running stats, as with the reference.  The caller initialises the slot (world_size * current stats)
 * before the first step.  slot = NULL (default) turns it off. */
int nrd_plan_set_bn_sync(nrd_plan* plan, float* slot, float contrib, float inv_world);
/* Gradient exchange over NVLink peer memory for the body bucket (one node, 2..8 ranks), replacing the all-reduce that DDP's
 * reducer runs after the last gradient (reached from main.py:38 through the hooks main.py:63 installs).  Every rank
 * allocates one exchange buffer (nrd_exchange_bytes / nrd_exchange_alloc: a cudaMalloc allocation, exported as a 64-byte CUDA
 * IPC handle), the ranks swap handles with any transport they like, open each other's buffers (nrd_exchange_open) and hand
 * the table (own pointer at index `rank`) to nrd_plan_set_exchange.  From then on nrd_backward_body() stores the body
 * gradients (and the BN slot of nrd_plan_set_bn_sync, which it then needs contrib = 1 on rank 0 and inv_world = 1 for) into
 * every rank's buffer and nrd_sgd_update_pack() sums the world's contributions in rank order before the update, so the
 * caller all-reduces nothing (the head bucket leaves the same way from nrd_backward_head).  The optimizer kernel waits for a
 * peer's flag for NRD_XCHG_TIMEOUT_S seconds (environment, read here; default 600 like NCCL's watchdog) before it traps.  Every rank must run the same sequence of steps.  world = 0 turns it off. */
size_t nrd_exchange_bytes(int C, int world);
int nrd_exchange_alloc(int device, size_t bytes, void** ptr, void* ipc_handle_64);
int nrd_exchange_open(int device, const void* ipc_handle_64, void** ptr);
int nrd_exchange_release(void* ptr, int opened);
int nrd_plan_set_exchange(nrd_plan* plan, int rank, int world, void* const* bufs);
/* debugging/tests: device pointer to a saved tensor.  what: 0 = x_k (k in 0..NB), 1 = y_k,
 * 2 = dy scratch, 3 = pooled, 4 = saved BN mean|invstd of block k ([2C] fp32), 5 = dx scratch (ping),
 * 6 = dx scratch (pong), 7 = wgrad per-CTA partials [ctas][9][C][C], 8 = BN-stat per-CTA partials [ctas][2][C],
 * 9 = stem weight-gradient per-CTA partials [ctas][28][C] (27 taps + bias), 10 = h1 = relu(fc1) [B,32] fp32,
 * 11 = in-kernel stamps of the small kernels (NRD_OPT_TRACE): [64 rows][16][clock64, globaltimer] int64 */
int nrd_plan_tensor(nrd_plan* plan, int what, int k, void** ptr, size_t* bytes);

/* Re-derive the packed weight copies (tap-major / transposed / bf16 / NHWC-ordered fc1) from the
 * fp32 master parameters.  Must run after anything that changed the parameters
 * (optimizer.step() main.py:39, load_state_dict, the DDP ctor broadcast main.py:63). */
int nrd_pack_weights(nrd_plan* plan, void* stream);

/* ---- whole-model entries ------------------------------------------------------------------
 * nrd_forward  = NetResDeep.forward (model/resnet.py:15-22) in training mode: batch-stat BN,
 *   running-stat EMA applied NB times, nbt += NB; keeps x_k, y_k, pool indices for backward.
 *   images: fp32 NCHW [B,3,32,32] (what main.py:33 moves to the device); logits: fp32 [B,10].
 *   training=0 gives eval-mode BN (running statistics, no side effects).
 * nrd_cross_entropy = nn.CrossEntropyLoss() forward AND its gradient (main.py:28,35,38):
 *   loss = mean_b(-log softmax(logits)[label]); dlogits = (softmax - onehot)/B * grad_scale.
 * nrd_backward = loss.backward() below the logits (main.py:38): overwrites the flat grads.
 * nrd_sgd_update = the DDP mean (main.py:63; c10d reducer divides by world size) fused with
 *   optimizer.step() of optim.SGD(lr) (main.py:27,39): p <- p - lr * inv_world * g, n floats.
 * nrd_train_step = main.py:34-39 for one rank without the allreduce: forward, loss, backward,
 *   update (inv_world = 1) and weight re-pack; loss_out: fp32 scalar on the device. */
int nrd_forward(nrd_plan* plan, const float* images, float* logits, int training, void* stream);
/* main.py:34-35 in one call: training forward, then loss and dlogits exactly as nrd_cross_entropy computes them (bit
 * for bit), with the loss folded into the forward's last kernel.  row_scratch: B floats. */
int nrd_forward_loss(nrd_plan* plan, const float* images, const int64_t* labels, float grad_scale, float* logits,
                     float* loss_out, float* dlogits, float* row_scratch, void* stream);
int nrd_cross_entropy(const float* logits, const int64_t* labels, int B, float grad_scale,
                      float* loss_out, float* dlogits, float* row_scratch /* [B] */, void* stream);
/* nrd_cross_entropy_static: the same computation for callers that cannot allocate device memory and
 * therefore cannot supply row_scratch (the device was handed to another allocator or a fixed arena,
 * a hook that only gets input/output pointers, a context where cudaMalloc is not allowed).  It has no
 * counterpart in the reference.  The per-row scratch is a 32 KiB array built into the library (held on
 * every GPU that loads it until the process exits), so B must be <= 8192 or the call fails with
 * NRD_EINVAL; nothing is allocated or freed.  The first call on each device looks up the array's
 * device address (one-time set-up): make that first call before starting any stream capture.  All
 * calls on a device share the one array: calls that may overlap on the GPU (different streams) race,
 * so keep them on one stream or order them with events.  No state carries from call to call.
 * Precondition: the CURRENT device of the calling thread must be the device that holds `logits` (this entry has no plan and
 * therefore cannot cudaSetDevice itself; every plan-based entry does): the scratch address is looked up per current device. */
int nrd_cross_entropy_static(const float* logits, const int64_t* labels, int B, float grad_scale,
                             float* loss_out, float* dlogits, void* stream);
int nrd_backward(nrd_plan* plan, const float* dlogits, void* stream);
/* The two independent halves of nrd_backward, so a caller can run them on different streams and start the all-reduce of
 * the head bucket (86.6 % of the gradient bytes) while the body backward is still running:
 * _head fills fc1/fc2 weight+bias gradients, _body everything else (d(pooled) -> residual blocks -> stem). */
int nrd_backward_head(nrd_plan* plan, const float* dlogits, void* stream);
int nrd_backward_body(nrd_plan* plan, const float* dlogits, void* stream);
int nrd_sgd_update(float* params, const float* grads, size_t n, float lr, float inv_world, void* stream);
/* nrd_sgd_update_pack = nrd_sgd_update over the plan's whole flat vector + nrd_pack_weights, one launch. */
int nrd_sgd_update_pack(nrd_plan* plan, float lr, float inv_world, void* stream);
/* optimizer.step() (main.py:39) riding the backward: after nrd_plan_set_fused_update(plan, lr, inv_world) with lr > 0,
 * nrd_backward_head() ends with the update of the head bucket (fc1/fc2: on the caller's side stream, under the body backward)
 * and nrd_backward_body()'s last kernel -- the one that finishes the body gradients -- applies p <- p - lr * inv_world * g to the
 * body bucket (after summing the world's slots when the peer-memory exchange is on) and re-derives the packed weight copies, so
 * NO nrd_sgd_update_pack() follows (calling it as well would step twice).  The flat gradient vector still holds the step's
 * gradients (the world's sum under the exchange).  lr <= 0 turns it off (default). */
int nrd_plan_set_fused_update(nrd_plan* plan, float lr, float inv_world);
/* `loss.item()` (main.py:41) without a copy node or a stream synchronisation: nrd_forward_loss() additionally writes the mean
 * loss of its n-th call (n = 1, 2, ...) to host_loss[(n - 1) % NRD_LOSS_RING] and then, behind a system-scope fence, n to
 * *host_seq.  Both must be page-locked HOST memory (device-accessible at the same address under unified addressing, e.g.
 * cudaHostAlloc / torch pin_memory; host_loss: NRD_LOSS_RING floats); the host polls *host_seq (volatile) and reads the ring
 * slot of the step it waits for once the count has reached it -- the ring keeps a value valid for NRD_LOSS_RING - 1 further
 * steps.  NULL, NULL turns it off (default). */
#define NRD_LOSS_RING 8
int nrd_plan_set_loss_mirror(nrd_plan* plan, float* host_loss, unsigned int* host_seq);
int nrd_train_step(nrd_plan* plan, const float* images, const int64_t* labels, float lr,
                   float* logits, float* loss_out, void* stream);

/* ---- synthetic CIFAR-shaped batch (replaces the DataLoader of main.py:53-61) ---------------
 * images[b,c,h,w] ~ N(0,1) fp32 NCHW, labels ~ U{0..9} int64; counter-based (Philox-4x32-10
 * keyed by seed, counter = (element, rank, step)), reproducible on the CPU. */
int nrd_gen_batch(float* images, int64_t* labels, int B, uint64_t seed, uint32_t rank, uint64_t step,
                  void* stream);

/* ---- optimizer / loss variants of the reference's second script (ppe_main_ddp.py:133,147) -------------------
 * nrd_sgd_momentum_update = optim.SGD(lr, momentum).step() with torch's defaults (dampening 0, nesterov off, weight decay 0):
 *   buf <- g (first_step != 0) or momentum * buf + g; p <- p - lr * buf, over n floats of the flat vectors; the caller re-packs
 *   the weight copies afterwards (nrd_pack_weights).
 * nrd_bce_with_logits = F.binary_cross_entropy_with_logits(logits, targets) (mean over all n elements) AND its gradient
 *   dlogits = (sigmoid(x) - t) * grad_scale / n, one pass; either output may be NULL. */
int nrd_sgd_momentum_update(float* params, const float* grads, float* momentum_buf, size_t n, float lr, float momentum, int first_step,
                            void* stream);
int nrd_bce_with_logits(const float* logits, const float* targets, size_t n, float grad_scale, float* loss_out, float* dlogits, void* stream);

/* ---- the reference's real input path on the device (main.py:53-61; main_no_ddp.py:22-34) ---------
 * The CIFAR-10 training set as torchvision holds it -- uint8 [N,32,32,3] (HWC) plus int64 labels [N] -- stays resident in HBM.
 * One call builds a batch: gather by sample index (the DistributedSampler order of main.py:60, computed on the host), ToTensor
 * (u8 -> [0,1] fp32, HWC -> CHW) and Normalize(mean, std) (main.py:55-57) in one pass, into the fp32 NCHW batch nrd_forward
 * takes, and the batch's labels.  (float(u8)/255 - mean)/std in fp32, torchvision's order: equal to the DataLoader's batch bit
 * for bit.  mean/std are HOST pointers to 3 floats; everything else is on the device.  B may be any size (the reference's
 * drop_last=False leaves a ragged last batch: 10 samples at 8 ranks). */
int nrd_gather_normalize(const uint8_t* data_u8_nhwc, const int64_t* labels_all, long long n_samples, const int64_t* indices, int B,
                         const float mean[3], const float std[3], float* images, int64_t* labels, void* stream);

/* ---- per-op entries (unit parity tests; the Python autograd Functions of single ops) ------
 * T = plan dtype is replaced by an explicit `dtype`; tensors NHWC.  H = W = 16 for the block
 * convolution (the only resolution the ResBlock runs at, model/resnet.py:16-17). */
/* conv3x3 pad 1, no bias (ResBlock.conv, model/resnet.py:34).  w_packed: fp32 [9][Cin=C][Cout=C]
 * (tap-major, from nrd_pack_conv_weight).  stats_partial (optional): [blocks][2][C] per-CTA sum
 * and sum of squares of the fp32 outputs; *stats_blocks receives the CTA count (host pointer). */
int nrd_pack_conv_weight(const float* w_oihw, int C, float* w_fwd /*[9][C][C]*/, float* w_dgrad /*[9][C][C]*/, void* stream);
int nrd_conv3x3_fwd(const void* x, const float* w_packed, void* y, int B, int C, int dtype,
                    float* stats_partial, int* stats_blocks, const void* residual, void* stream);
/* wgrad: dW[co][ci][ky][kx] (OIHW fp32, OVERWRITTEN) = sum_p dy[p][co] * x[p+tap][ci].
 * partial: scratch of nrd_wgrad_partial_bytes(B, C). */
size_t nrd_wgrad_partial_bytes(int B, int C);
int nrd_conv3x3_wgrad(const void* x, const void* dy, float* dw_oihw, int B, int C, int dtype,
                      float* partial, void* stream);
/* BatchNorm2d training forward pieces (model/resnet.py:35-37): finalize turns the per-CTA
 * partials into mean/invstd (saved: [2C]) and scale/shift ([2C]), and applies the running-stat
 * EMA (momentum 0.1, unbiased variance) when bn_running != NULL; apply computes
 * out = relu(y*scale+shift) + x. */
int nrd_bn_finalize(const float* stats_partial, int blocks, int count, int C, const float* gamma,
                    const float* beta, float* saved_mean_invstd, float* scale_shift,
                    float* bn_running, int64_t* nbt, void* stream);
int nrd_bn_relu_add(const void* y, const void* x, const float* scale_shift, void* out, int B, int C,
                    int dtype, void* stream);
/* BatchNorm2d + ReLU + residual backward (native_batch_norm_backward + threshold_backward):
 * given g = dL/d(out), y, scale_shift and saved stats: dgamma/dbeta (ACCUMULATED into
 * dgamma_dbeta [2C] when accumulate != 0) and dy.  scratch: [blocks][2][C] + [2C] floats. */
size_t nrd_bn_bwd_scratch_bytes(int B, int C);
int nrd_bn_relu_bwd(const void* g, const void* y, const float* scale_shift, const float* saved_mean_invstd,
                    const float* gamma, void* dy, float* dgamma_dbeta, int accumulate, int B, int C,
                    int dtype, float* scratch, void* stream);
/* stem: max_pool2d(relu(conv1(x)),2) (model/resnet.py:16).  x0: NHWC [B,16,16,C], idx0: u8. */
int nrd_stem_fwd(const float* images, const float* conv1_w, const float* conv1_b, void* x0, uint8_t* idx0,
                 int B, int C, int dtype, void* stream);
size_t nrd_stem_bwd_partial_bytes(int B, int C);
int nrd_stem_bwd(const float* images, const void* g0, const void* x0, const uint8_t* idx0, float* dconv1_w,
                 float* dconv1_b, int B, int C, int dtype, float* partial, void* stream);
/* max_pool2d(out,2) at 16x16 -> 8x8 (model/resnet.py:18) and its gradient scatter */
int nrd_maxpool_fwd(const void* x, void* pooled, uint8_t* idx, int B, int C, int dtype, void* stream);
int nrd_maxpool_bwd(const void* gpooled, const uint8_t* idx, void* gx, int B, int C, int dtype, void* stream);

/* ---- measurement hooks (bench.py / ncu; no reference counterpart: the reference has no profiler) ---
 * nrd_launch_count: kernels launched by this library in this process so far (graph replays do not
 * pass through the library and are counted by the caller as captured launches x replays).
 * nrd_plan_run_op: launch ONE kernel class of the step on the plan's buffers, for CUDA-event timing
 * on the launching stream.  Needs a preceding nrd_forward + nrd_backward on the plan. */
#define NRD_OP_CONV_FWD 0  /* y_k = conv3x3(x_k) + BN partial sums      (model/resnet.py:34) */
#define NRD_OP_BN_APPLY 1  /* relu(bn(y_k)) + x_k                        (model/resnet.py:35-37) */
#define NRD_OP_BN_BWD 2    /* BN+ReLU backward: reduce, finalize, apply  (loss.backward, main.py:38) */
#define NRD_OP_WGRAD 3     /* conv weight gradient partials */
#define NRD_OP_DGRAD 4     /* conv input gradient + residual */
#define NRD_OP_STEM_FWD 5  /* conv1 + relu + pool                        (model/resnet.py:16) */
#define NRD_OP_STEM_BWD 6
#define NRD_OP_BLOCK_FWD 7 /* whole ResBlock forward k in one cooperative launch (tcgen05 plans): conv + BN + ReLU + add;
                              NOTE: applies the running-stat EMA and nbt += 1 like a real forward */
#define NRD_OP_FWD_BLOCKS 8 /* all residual blocks' forward in the persistent kernel (same side effects as a real forward) */
#define NRD_OP_BWD_BLOCKS 9 /* all residual blocks' backward in the persistent kernel (overwrites the block gradients) */
#define NRD_OP_HEAD 10       /* pool + fc1 + fc2 + loss + dlogits + dL/d(pooled), the one head launch (model/resnet.py:18-21, main.py:35);
                                needs a preceding nrd_forward_loss (its labels and outputs are reused) */
#define NRD_OP_BODY_REDUCE 11 /* fixed-order reduce of the body-gradient partials (no exchange, no optimizer step) */
#define NRD_OP_SGD 12        /* the optimizer kernel over the whole flat vector with step 0 (same traffic, weights unchanged) */
unsigned long long nrd_launch_count(void);
/* us[i] = microseconds the i-th kernel of the main stream took in the last step run with NRD_OPT_TIMELINE on (synchronise the
 * device first), names[i] = a static string naming it; returns the number of entries (< 0: error).  HOST pointers. */
int nrd_plan_timeline(nrd_plan* plan, int max_n, float* us, const char** names);
int nrd_plan_run_op(nrd_plan* plan, int op, int k, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NRD_H_ */
