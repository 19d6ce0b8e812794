/* libfedb200 - C ABI of the fedb200 sm_100a kernels and the peer-memory manager.
 *
 * Conventions: every function returns 0 on success, a non-zero code otherwise, and leaves a
 * human-readable message in fedb_last_error().  `stream` is a cudaStream_t passed as void*.
 * All pointers are device pointers unless the name says host.  Device memory is allocated only by the
 * fedb_mem_* / fedb_mc_* family (peer-shared arenas) and by fedb_cross_entropy (a per-device row scratch
 * that grows with cudaMalloc/cudaFree when the batch grows; fedb_cross_entropy_static never allocates).
 *
 * "reference:" lines cite the file:line of amolahinge/739-839-federated-learning-using-grpc whose
 * behaviour the entry point replaces.
 */
#ifndef FEDB200_H
#define FEDB200_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FEDB_MAX_PEERS 16
#define FEDB_MAX_DST 20

const char* fedb_last_error(void);
int fedb_version(void);

/* ------------------------------------------------------------------ control block (device flags)
 * One per peer-shared allocation.  arrive[j]/done[j] are written by rank j with st.release.sys and
 * polled locally with ld.acquire.sys (bounded spin).  Replaces the thread joins of the reference
 * round loop.  reference: src/server.py:132-135,150-153 */
typedef struct {
  uint32_t arrive[FEDB_MAX_PEERS];
  uint32_t done[FEDB_MAX_PEERS];
  uint32_t epoch;        /* rounds completed by this rank's K-AGG */
  uint32_t status;       /* 0 ok | 1 timeout in arrive wait | 2 timeout in done wait | 3 aborted */
  uint32_t blocks_done;  /* grid completion counter, self-resetting */
  uint32_t abort_flag;   /* host writes 1 to break every spin */
  uint32_t version;      /* model version held by the arena this block guards */
  uint32_t reserved[27];
} fedb_ctrl_t;

/* ------------------------------------------------------------------ K-AGG / K-BCAST
 * out[i] = mean_k src[k][i] over [begin, end) of the flat fp32 arena, in the reference's order
 * ((s0+s1)+s2...)/N with the division last; int64 counters are summed in int64, true-divided in
 * fp32 and truncated.  The result is stored to every dst[] (peer or local pointers; optional bf16
 * shadow copies) - that fan-out IS the global-model broadcast.  With sync_mode=1 the kernel first
 * tells every peer "my arena is final" and waits for theirs, and afterwards tells every peer "my
 * shard is written" and waits for theirs, so one launch per rank is a complete
 * reduce-scatter + all-gather over NVLink.
 * reference: src/server.py:155-179 (allreduce), :144-153 (SendModel fan-out), :141-142 (backup) */
typedef struct {
  const float* src[FEDB_MAX_PEERS];       /* client fp32 arenas */
  const int64_t* src_i64[FEDB_MAX_PEERS]; /* client int64 counter segments (may be NULL) */
  float* dst[FEDB_MAX_DST];               /* destinations: clients, server global, backup replica */
  int64_t* dst_i64[FEDB_MAX_DST];
  uint16_t* dst_bf16[FEDB_MAX_DST];       /* optional bf16 shadow per destination (NULL = none) */
  float weight[FEDB_MAX_PEERS];           /* sample counts as floats; ignored when uniform */
  int32_t n_src, n_dst;                   /* n_dst counts dst[] (fp32) */
  int32_t n_dst_i64;                      /* entries of dst_i64[]: may exceed n_dst when the fp32 fan-out
                                             goes through mc_dst but the counters go out as unicast stores */
  int32_t uniform;                        /* 1: plain mean (N1) */
  int32_t do_i64;                         /* 1: this launch also reduces the int64 segment */
  int64_t begin, end;                     /* fp32 element range, multiples of 4 */
  int64_t n_i64;
  /* compressed uploads (K-CODEC decode fused): src_q[k] fp8-e4m3 deltas, src_scale[k] one fp32
   * scale per 32 elements, base = the model the deltas are relative to.  NULL src_q = raw fp32. */
  const uint8_t* src_q[FEDB_MAX_PEERS];
  const float* src_scale[FEDB_MAX_PEERS];
  const float* base;
  /* NVLS: when mc_src != NULL the sum comes from one multimem.ld_reduce on the multicast alias
   * of the client arenas (in-switch reduction, order not defined); when mc_dst != NULL the result
   * goes out with one multimem.st instead of n unicast stores (dst[] may still list extras). */
  const float* mc_src;
  float* mc_dst;
  /* cross-rank synchronisation */
  int32_t sync_mode;                      /* 0 none | 1 device flags */
  int32_t rank, world;
  fedb_ctrl_t* ctrl_local;
  fedb_ctrl_t* ctrl_peer[FEDB_MAX_PEERS];
  uint64_t timeout_ns;
  uint32_t version;                       /* stamped into ctrl of every destination when nonzero */
  fedb_ctrl_t* dst_ctrl[FEDB_MAX_DST];    /* optional: where to stamp the version */
  /* optional gate: when non-NULL and *gate != 0 at run time the launch does nothing (no stores, no version stamp).
   * The follow-up copies of a round (global model, replica) pass the status word of the round's K-AGG here, so a
   * round whose flag wait failed cannot leak un-averaged weights into the server's copies. */
  const uint32_t* gate;
} fedb_agg_desc_t;

int fedb_fedavg(const fedb_agg_desc_t* desc, int max_blocks, void* stream);

/* K-REPL: versioned copy of a flat arena (fp32 + int64) into the backup's replica, meant for a
 * low-priority side stream so it overlaps the next round.  reference: src/server.py:141-142,236-242 */
int fedb_replicate(float* dst, const float* src, int64_t n_f32, int64_t* dst_i64, const int64_t* src_i64,
                   int64_t n_i64, fedb_ctrl_t* dst_ctrl, uint32_t version, void* stream);

/* K-PUSH: bit-exact fan-out of `bytes` (multiple of 16) from local memory into n_dst peer buffers, or through the NVLS
 * multicast alias mc_dst when it is non-NULL (one multimem.st per 16 bytes).  No arithmetic: carries any data type.
 * Used to all-gather per-round inputs that every client needs (the evaluation set) over NVLink instead of N PCIe
 * uploads.  No counterpart in the reference (every worker loads its own copy of CIFAR-10, src/main.py:43-56). */
int fedb_push_copy(const void* src, void* const* dst, int n_dst, void* mc_dst, uint64_t bytes, void* stream);

/* Barrier on device flags only (bounded spin). */
int fedb_flag_barrier(fedb_ctrl_t* local, fedb_ctrl_t* const* peers, int rank, int world,
                      uint64_t timeout_ns, void* stream);

/* ------------------------------------------------------------------ K-CODEC (the -c switch)
 * Lossy opt-in mode: per 32-element block, delta = cur - base + residual; scale = amax/448;
 * q = e4m3(delta/scale); residual = delta - q*scale (error feedback).  Lossless mode needs no
 * kernel on the GPU plane (weights cross NVLink raw).  reference: src/server.py:103-107,
 * src/client.py:39-43 */
int fedb_codec_encode(const float* cur, const float* base, float* residual, uint8_t* q, float* scale,
                      int64_t n, void* stream);
int fedb_codec_decode(float* out, const float* base, const uint8_t* q, const float* scale, int64_t n,
                      void* stream);

/* ------------------------------------------------------------------ K10 fused SGD / Adam
 * torch.optim.SGD semantics over the flat parameter range: g += wd*p; v = mu*v + g; p -= lr*v
 * (v starts at zero, which makes the first step v = g exactly).  Optionally emits a bf16 shadow.
 * reference: src/main.py:99-100,151 */
int fedb_sgd_step(float* param, const float* grad, float* momentum, uint16_t* shadow_bf16, int64_t n,
                  float lr, float mu, float weight_decay, void* stream);
/* End-of-step form: same update, and in the same pass the gradient is cleared (the next backward pass accumulates into
 * zeros) and every int64 counter of the arena (BatchNorm's num_batches_tracked) advances by one. */
int fedb_sgd_step_end(float* param, float* grad, float* momentum, int64_t n, float lr, float mu, float weight_decay,
                      int64_t* counters, int64_t n_counters, void* stream);
int fedb_adam_step(float* param, const float* grad, float* m, float* v, int64_t n, float lr, float beta1,
                   float beta2, float eps, float weight_decay, const int32_t* step_dev, void* stream);

/* ------------------------------------------------------------------ K9/K13 loss + counters
 * logits [B, C] -> mean CE loss; dlogits = (softmax - onehot)/B; stats[0] += loss_sum (fp32),
 * stats[1] += correct (as float), stats[2] += B.  A label outside [0, C) is not dereferenced; it turns
 * that row's loss (and so the batch loss) into NaN.  reference: src/main.py:98,149,153-156 */
int fedb_cross_entropy(const float* logits, const int64_t* target, float* dlogits, float* loss_out,
                       float* stats, int B, int C, void* stream);

/* Same computation and same arguments as fedb_cross_entropy, for callers that cannot allocate device
 * memory when they call: a process whose whole GPU already belongs to its own pool (torch's caching
 * allocator at its high-water mark, a serving runtime with one fixed arena), where a library's
 * cudaMalloc fails or fragments, and latency-critical loops, where cudaMalloc/cudaFree would
 * synchronise the device.  This entry point calls no cudaMalloc/cudaFree and no allocating routine:
 * its per-row scratch is a fixed 128 KiB array inside the library image (held by every GPU that loads
 * the library until exit).  Limits and rules: B <= FEDB_CE_STATIC_MAX_ROWS (16384), larger batches are
 * refused with an error; all calls on a GPU share that array, so calls that may overlap on the GPU
 * race - keep them on one stream or order them with an event; nothing in the array needs clearing
 * between calls (each call writes the rows it later reads); the first call on each device looks up
 * the array's address and should be made before any stream capture on that device.
 * This entry point has no counterpart in the reference. */
#define FEDB_CE_STATIC_MAX_ROWS 16384
int fedb_cross_entropy_static(const float* logits, const int64_t* target, float* dlogits, float* loss_out,
                              float* stats, int B, int C, void* stream);

/* ------------------------------------------------------------------ K14 input pipeline
 * uint8 NHWC [B,32,32,3] -> fp32 NCHW (nhwc_out=0) or NHWC (1), ToTensor+Normalize, optional
 * RandomCrop(32,4)+flip drawn from (seed, sample index).  reference: src/main.py:36-46 */
int fedb_prepare_batch(const uint8_t* src_u8, float* dst, int B, int nhwc_out, int augment,
                       uint64_t seed, void* stream);
/* Same, with a device-side seed word mixed into `seed` at run time: a step captured in a CUDA graph bakes `seed`
 * in, the word (bumped by the host once per pass) keeps every replayed pass drawing fresh crops and flips, as the
 * reference does every epoch (src/main.py:36-41). */
int fedb_prepare_batch_dev(const uint8_t* src_u8, float* dst, int B, int nhwc_out, int augment,
                           uint64_t seed, const uint64_t* seed_dev, void* stream);

/* ------------------------------------------------------------------ LeNet fused training step
 * Whole step in 4 kernels: features fwd, classifier fwd+loss+bwd, features bwd, reduce+SGD.
 * `p` points at the flat fp32 arena; off[] are the element offsets of conv1.w, conv1.b, conv2.w,
 * conv2.b, fc1.w, fc1.b, fc2.w, fc2.b, fc3.w, fc3.b.  Workspace sizes come from
 * fedb_lenet_workspace_bytes(B).  reference: src/models/lenet.py:5-23, src/main.py:142-151 */
typedef struct {
  float* param;            /* arena base */
  float* momentum;         /* same layout as param */
  int64_t off[10];
  float lr, mu, weight_decay;
  int32_t augment;
  uint64_t seed;
  const uint64_t* seed_dev; /* optional device word XOR-ed into every step seed: lets a replayed CUDA
                               graph draw fresh crops each round (bump it between replays) */
} fedb_lenet_t;

size_t fedb_lenet_workspace_bytes(int B);
/* debug: clock64() stamps of CTA 0 at the 15 phase boundaries of the sample kernel (tools/prof_lenet.py --phases) */
int fedb_lenet_phase_clocks(const fedb_lenet_t* net, const uint8_t* images_u8, const int64_t* labels, int B,
                            void* workspace, long long* clk_dev, void* stream);
size_t fedb_lenet_eval_bytes(int B);   /* workspace of fedb_lenet_eval: two rows of B floats (always <= workspace_bytes(B)) */
int fedb_lenet_train_step(const fedb_lenet_t* net, const uint8_t* images_u8, const int64_t* labels,
                          int B, void* workspace, float* stats, uint64_t step_seed, void* stream);
int fedb_lenet_eval(const fedb_lenet_t* net, const uint8_t* images_u8, const int64_t* labels, int B,
                    void* workspace, float* stats, void* stream);
/* The same evaluation through a WORK QUEUE, so that it can run off the round's critical path: persistent CTAs pull chunks
 * of samples from `counter` (one device word, zeroed by the caller before the first launch of an evaluation).  A small
 * background launch (ctas = idle SMs) scores the global model next to the next round's training kernels; a full-size
 * launch on the training stream takes whatever is left when the result is needed; fedb_lenet_eval_stats then reduces
 * the per-sample rows in a fixed order (the numbers do not depend on who scored which chunk).
 * reference: src/client.py:30 + src/main.py:167-191 evaluate inside SendModel, on the critical path. */
int fedb_lenet_eval_queue(const fedb_lenet_t* net, const uint8_t* images_u8, const int64_t* labels, int B, void* workspace,
                          uint32_t* counter, int ctas, void* stream);
int fedb_lenet_eval_stats(void* workspace, int B, float* stats, void* stream);
/* forward only, logits out (tests) */
int fedb_lenet_forward(const fedb_lenet_t* net, const uint8_t* images_u8, const float* images_f32,
                       int B, void* workspace, float* logits, void* stream);
/* gradient only (no update): writes dparam with the arena layout (tests) */
int fedb_lenet_grad(const fedb_lenet_t* net, const float* images_f32, const int64_t* labels, int B,
                    void* workspace, float* dparam, float* loss_out, void* stream);

/* ------------------------------------------------------------------ peer memory manager
 * CUDA VMM allocations exported as POSIX file descriptors, imported and mapped by peers, plus
 * NVLS multicast objects when the fabric supports them.  Replaces checkpoint files and base64
 * strings as the IPC medium.  reference: src/client.py:19-23, src/server.py:55-58 */
typedef struct {
  uint64_t handle;     /* CUmemGenericAllocationHandle */
  uint64_t ptr;        /* mapped device VA in this process */
  uint64_t bytes;      /* padded to the allocation granularity */
  int32_t device;      /* device the physical memory lives on (-1: imported, unknown) */
  int32_t fd;          /* exported fd (-1 if not exported) */
} fedb_mem_t;

int fedb_mem_supported(int device, int* vmm, int* posix_fd, int* multicast);
int fedb_mem_create(uint64_t bytes, int device, fedb_mem_t* out);
int fedb_mem_export_fd(fedb_mem_t* mem);
int fedb_mem_import_fd(int fd, uint64_t bytes, int device, fedb_mem_t* out);
int fedb_mem_release(fedb_mem_t* mem);
int fedb_mc_create(uint64_t bytes, int n_devices, fedb_mem_t* out_mc);          /* rank 0 */
int fedb_mc_import_fd(int fd, uint64_t bytes, fedb_mem_t* out_mc);              /* other ranks */
int fedb_mc_add_device(fedb_mem_t* mc, int device);
int fedb_mc_bind_and_map(fedb_mem_t* mc, const fedb_mem_t* local_mem, int device); /* sets mc->ptr */
/* fd exchange over a unix domain socket (SCM_RIGHTS) */
int fedb_fd_serve(const char* path, const int* fds, int n_fds, int n_clients, int timeout_ms);
int fedb_fd_fetch(const char* path, int* out_fds, int n_fds, int timeout_ms);
/* cudaMemcpyAsync(cudaMemcpyDefault) on `stream`: pinned host <-> device <-> peer; capturable as a memcpy node */
int fedb_copy_async(void* dst, const void* src, uint64_t bytes, void* stream);
/* Zero-fill of `bytes` at dst as a memset on the stream (inside a captured graph: a memset node, no kernel). */
int fedb_zero_async(void* dst, uint64_t bytes, void* stream);

/* ------------------------------------------------------------------ tensor-core GEMM / convolution (csrc/gemm_tc.cu)
 * tcgen05.mma with TMEM accumulators, operands staged by TMA through a 4-stage mbarrier ring, 128x128 output tiles,
 * optional split-K (fp32 red.add into D).  Everything is row-major fp32 in HBM unless noted; leading dimensions in
 * elements.  These carry the model zoo's convolutions (reference: src/models/*.py nn.Conv2d / nn.Linear, which the
 * reference leaves to ATen on the CPU). */
/* D[M,N] (+)= A[M,K] * B[N,K]^T (+ bias[N]) (relu).  TF32 inputs, fp32 accumulate.  K % 32 == 0, N % 4 == 0. */
int fedb_gemm_tf32(const float* A, const float* B, float* D, const float* bias, int M, int N, int K, long long lda,
                   long long ldb, long long ldd, int relu, int splits, int accumulate, void* stream);
/* D[M,N] (+)= At[K,M]^T * Bt[K,N] with bf16 inputs stored reduction-major (the weight-gradient shape: both operands
 * are activations [pixels, channels]).  M % 64 == 0, N % 64 == 0, K % 64 == 0. */
int fedb_gemm_bf16_tn(const uint16_t* At, const uint16_t* Bt, float* D, int M, int N, int K, long long lda, long long ldb,
                      long long ldd, int splits, int accumulate, void* stream);
/* Implicit-GEMM convolution on NHWC fp32: Y[N,H/stride,W/stride,Cout] = conv(X[N,H,W,Cin], Wk) (+bias)(relu).
 * fs/pad: 3/1 or 1/0; stride 1 or 2; output maps must be powers of two with width 2..32; Cin % 32 == 0.
 * Wk is the filter packed by fedb_pack_filter mode 0 ([Cout][tap][Cin]).  The data gradient of a stride-1 layer is the
 * same call on dY with the mode-1 packing (flipped taps, channels swapped). */
int fedb_conv_tf32(const float* X, const float* Wk, float* Y, const float* bias, int N, int H, int W, int Cin, int Cout, int fs,
                   int stride, int pad, int relu, void* stream);
int fedb_conv3x3_tf32(const float* X, const float* Wk, float* Y, const float* bias, int N, int H, int W, int Cin, int Cout,
                      int relu, void* stream);
/* Weight gradient: dWk[Cout][tap][Cin] += dY^T * shifted(X), bf16 inputs (fedb_cast_bf16), fp32 accumulate, split over
 * pixels.  dWk must be zeroed by the caller; fedb_pack_filter mode 2 turns it back into OIHW.  Channels % 64 == 0. */
int fedb_conv_wgrad_bf16(const uint16_t* dY, const uint16_t* X, float* dWk, int N, int H, int W, int Cin, int Cout, int fs,
                         int stride, int pad, int splits, void* stream);
/* Data gradient of a stride-2 3x3/pad-1 or 1x1 convolution, decomposed by the parity of the dX pixel into four stride-1
 * implicit GEMMs whose epilogues write the interleaved positions of dX [N,2P,2Q,Cin] directly.  dY is [N,P,Q,Cout].
 * fs == 3: Wcls from fedb_pack_dgrad_s2 (9*Cin*Cout floats).  fs == 1: Wcls = W^T and dX must be zero-filled first. */
int fedb_pack_dgrad_s2(const float* w_oihw, float* out, int K, int C, void* stream);
int fedb_conv_dgrad_s2_tf32(const float* dY, const float* Wcls, float* dX, int N, int P, int Q, int Cout, int Cin, int fs,
                            void* stream);
/* One group of a grouped convolution per call: pointers address the group's first channel, the pitches are the channel
 * counts of the whole NHWC tensors (0 = dense).  Same kernels, same constraints on the per-group channel counts. */
int fedb_conv_tf32_g(const float* X, const float* Wk, float* Y, const float* bias, int N, int H, int W, int Cin, int Cout, int fs,
                     int stride, int pad, int relu, long long xpitch, long long ypitch, void* stream);
int fedb_conv_wgrad_bf16_g(const uint16_t* dY, const uint16_t* X, float* dWk, int N, int H, int W, int Cin, int Cout, int fs,
                           int stride, int pad, int splits, long long dypitch, long long xpitch, void* stream);
int fedb_conv_dgrad_s2_tf32_g(const float* dY, const float* Wcls, float* dX, int N, int P, int Q, int Cout, int Cin, int fs,
                              long long dypitch, long long dxpitch, void* stream);
/* All groups of a grouped convolution in ONE launch (blockIdx.z / blockIdx.y carries the group): C and K are the channel
 * counts of the whole tensors, the filter is the grouped OIHW tensor [K][C/groups][fs][fs].  Packing: forward = plain
 * fedb_pack_filter mode 0 with C/groups input channels; stride-1 data gradient = fedb_pack_filter_grouped mode 1 (one row
 * per input channel of the whole tensor; a 1x1 filter gives the unpadded per-group transposes); stride-2 data gradient =
 * fedb_pack_dgrad_s2_grouped.  Per-group channel counts: % 4 (forward / data gradient), % 8 (weight gradient). */
int fedb_conv_tf32_grouped(const float* X, const float* Wk, float* Y, const float* bias, int N, int H, int W, int C, int K,
                           int groups, int fs, int stride, int pad, int relu, void* stream);
int fedb_conv_dgrad_s2_tf32_grouped(const float* dY, const float* Wcls, float* dX, int N, int P, int Q, int K, int C, int groups,
                                    int fs, void* stream);
int fedb_conv_wgrad_bf16_grouped(const uint16_t* dY, const uint16_t* X, float* dWk, int N, int H, int W, int C, int K, int groups,
                                 int fs, int stride, int pad, int splits, void* stream);
int fedb_pack_filter_grouped(const float* w_oihw, float* out, int K, int C_per_group, int groups, int mode, int taps, void* stream);
int fedb_pack_dgrad_s2_grouped(const float* w_oihw, float* out, int K, int C_per_group, int groups, void* stream);
int fedb_conv3x3_wgrad_bf16(const uint16_t* dY, const uint16_t* X, float* dWk, int N, int H, int W, int Cin, int Cout,
                            int splits, void* stream);
/* mode 0: OIHW -> [K][tap][C]; mode 1: OIHW -> [C][flipped tap][K] (data gradient); mode 2: [K][tap][C] -> OIHW */
int fedb_pack_filter(const float* w_oihw, float* out, int K, int C, int mode, int taps, void* stream);
int fedb_pack3x3(const float* w_oihw, float* out, int K, int C, int mode, void* stream);
/* All filter (un)packings of a model in ONE launch.  jobs_dev: device array of njobs 56-byte records
 *   { const float* src; float* dst; int64 total; int32 kind, K, C, T, Rp, groups, chunk0, pad; }
 * chunk0 = sum over the preceding jobs of ceil(total / FEDB_PACK_CHUNK); total_chunks = that sum over all jobs (one CTA
 * per chunk, so every job gets CTAs in proportion to its size).
 * kind 0 / 1 = fedb_pack_filter modes 0 / 1 (Rp: padded reduction width), 2 = dst(OIHW) += src([K][tap][C]) and src is
 * cleared (weight gradients after the backward pass), 3 = fedb_pack_dgrad_s2 classes, 4 = transpose [K][C] -> [C][K];
 * 5 / 6 / 7 = kinds 0 / 1 / 2 for 3x3 filters with K % 32 == 0 and C % 32 == 0, moved as 32 x 32 x 9 tiles through shared
 * memory (total must then be K * C * 9, i.e. no padding).
 * The reference has no counterpart (its convolutions are torch.nn.Conv2d, src/models). */
#define FEDB_PACK_CHUNK 9216
int fedb_pack_multi(const void* jobs_dev, int njobs, int total_chunks, void* stream);
/* im2col for tiny reductions (Cin*fs*fs <= 32, the 3-channel stems): NHWC x -> [pixels][32] fp32 and/or [pixels][64] bf16,
 * column = OIHW flattening of the filter, zero padded; the convolution and its weight gradient become plain GEMMs */
int fedb_im2col_small(const float* x, float* out_f32, uint16_t* out_bf16, int N, int H, int W, int C, int fs, int stride,
                      int pad, void* stream);
int fedb_cast_bf16(const float* src, uint16_t* dst, long long n, void* stream);          /* n % 4 == 0 */
/* fp32 [rows][C] -> bf16 [rows][Cpad], zero columns appended (Cpad % 8 == 0): 16-byte pitch for any channel count */
int fedb_cast_bf16_pad(const float* src, uint16_t* dst, long long rows, int C, int Cpad, void* stream);
/* the same from rows that are src_pitch floats apart: extracts one group's channels of an NHWC tensor (src at its first channel) */
int fedb_cast_bf16_pad_p(const float* src, uint16_t* dst, long long rows, int C, int Cpad, long long src_pitch, void* stream);
int fedb_transpose(const float* in, float* out, int rows, int cols, void* stream);        /* out[cols][rows] */

/* ------------------------------------------------------------------ generic NHWC layer kernels (csrc/nn_generic.cu)
 * Every layer type the model zoo uses (reference: src/models/*.py), for the shapes the tensor-core path does not
 * take: grouped / depthwise / odd-sized convolutions, batch norm, activations, pooling, channel cat/slice/shuffle,
 * squeeze-excite gates, drop-connect masks, the classifier.
 * shape14 = {N,H,W,C,K,R,S,P,Q,stride_h,stride_w,pad_h,pad_w,groups}; shape12 = {N,H,W,C,P,Q,kh,kw,sh,sw,ph,pw}.
 * act: 0 none, 1 relu, 2 swish, 3 sigmoid. */
int fedb_nn_weight_transpose(const float* w, float* wt, float* wt2, int K, int Cg, int R, int S, void* stream);
int fedb_nn_conv_fwd(const float* x, const float* wt, const float* bias, float* y, const int* shape14, void* stream);
int fedb_nn_conv_dgrad(const float* dy, const float* wt2, float* dx, const int* shape14, void* stream);
int fedb_nn_conv_wgrad(const float* x, const float* dy, float* dw, float* dbias, const int* shape14, void* stream);
/* Batch norm over rows = N*H*W of an NHWC tensor.  scratch: caller-owned floats used for the chunked two-pass
 * reduction (at least 2*C; more chunks = more parallelism).  Training mode updates the running statistics (N4). */
int fedb_nn_bn_fwd(const float* x, float* y, float* mean, float* invstd, float* run_mean, float* run_var, const float* gamma,
                   const float* beta, long long rows, int C, float eps, float momentum, int training, int act, float* scratch,
                   long long scratch_floats, void* stream);
int fedb_nn_bn_bwd(const float* x, const float* dy, float* dx, const float* mean, const float* invstd, const float* gamma,
                   const float* beta, float* dgamma, float* dbeta, long long rows, int C, int act, int training, float* scratch,
                   long long scratch_floats, void* stream);
/* Same, plus an optional bf16 copy of the result (y / dx) written by the same kernel: the operand of the tensor-core
 * weight-gradient GEMM of the neighbouring convolution, so no separate cast pass re-reads the tensor.  C % 4 == 0. */
int fedb_nn_bn_fwd16(const float* x, float* y, float* mean, float* invstd, float* run_mean, float* run_var, const float* gamma,
                     const float* beta, long long rows, int C, float eps, float momentum, int training, int act, float* scratch,
                     long long scratch_floats, uint16_t* y_bf16, void* stream);
int fedb_nn_bn_bwd16(const float* x, const float* dy, float* dx, const float* mean, const float* invstd, const float* gamma,
                     const float* beta, float* dgamma, float* dbeta, long long rows, int C, int act, int training, float* scratch,
                     long long scratch_floats, uint16_t* dx_bf16, void* stream);
int fedb_nn_add_act16(const float* a, const float* b, float* y, long long n, int act, uint16_t* y_bf16, void* stream);
int fedb_nn_act_fwd(const float* x, float* y, long long n, int act, void* stream);
int fedb_nn_act_bwd(const float* x, const float* dy, float* dx, long long n, int act, void* stream);
int fedb_nn_add_act(const float* a, const float* b, float* y, long long n, int act, void* stream);
/* mode 0: y = x * s[n][c] (squeeze-excite gate); mode 1: y = x * s[n] (drop-connect / per-sample mask) */
int fedb_nn_scale(const float* x, const float* s, float* y, long long total, long long per_sample, int C, int mode, void* stream);
int fedb_nn_scale_bwd_gate(const float* x, const float* dy, float* ds, int HW, int C, int N, void* stream);
int fedb_nn_bernoulli_mask(float* m, long long n, float keep, uint64_t seed, const uint64_t* ctr_dev, void* stream);
int fedb_nn_pool_fwd(const float* x, float* y, int* arg, const int* shape12, int is_max, void* stream);
int fedb_nn_pool_bwd(const float* dy, const int* arg, float* dx, const int* shape12, int is_max, void* stream);
/* out[row][dst_off + j] (+)= in[row][src(j)]: cat, slice and (shuffle_groups > 1) channel shuffle in one kernel */
int fedb_nn_channel_copy(const float* in, float* out, long long rows, int Cin, int Cout, int src_off, int dst_off, int width,
                         int shuffle_groups, int accumulate, void* stream);
int fedb_nn_linear_fwd(const float* x, const float* W, const float* bias, float* y, int B, int K, int J, int act, void* stream);
int fedb_nn_linear_bwd(const float* dy, const float* x, const float* W, float* dx, float* dW, float* db, int B, int K, int J,
                       void* stream);

/* ------------------------------------------------------------------ depthwise 3x3 convolution (csrc/depthwise.cu)
 * groups == Cin == Cout, NHWC fp32, any stride / padding, C % 4 == 0; w is OIHW = [C][1][3][3].
 * shape10 = {N, H, W, C, P, Q, stride_h, stride_w, pad_h, pad_w}.  Streaming kernels: four channels per thread, taps in
 * registers; the weight gradient is a two-pass deterministic reduction through caller-owned scratch.
 * reference: src/models/mobilenet.py:14-17, mobilenetv2.py:20-22, efficientnet.py:60-67, shufflenetv2.py:33-36 */
int fedb_dw_conv_fwd(const float* x, const float* w, const float* bias, float* y, const int* shape10, void* stream);
int fedb_dw_conv_dgrad(const float* dy, const float* w, float* dx, const int* shape10, void* stream);
long long fedb_dw_conv_wgrad_scratch(const int* shape10);          /* floats */
int fedb_dw_conv_wgrad(const float* x, const float* dy, float* dw, float* scratch, const int* shape10, void* stream);
/* the same with fs x fs taps, fs in {3, 5, 7} (PNASNet's and EfficientNet's separable convolutions) */
int fedb_dw_conv_fwd_k(const float* x, const float* w, const float* bias, float* y, const int* shape10, int fs, void* stream);
int fedb_dw_conv_dgrad_k(const float* dy, const float* w, float* dx, const int* shape10, int fs, void* stream);
long long fedb_dw_conv_wgrad_scratch_k(const int* shape10, int fs);
int fedb_dw_conv_wgrad_k(const float* x, const float* dy, float* dw, float* scratch, const int* shape10, int fs, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FEDB200_H */
