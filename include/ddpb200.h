/* ddpb200.h -- C ABI of libddpb200.so, the sm_100a kernel library of the ddpb200 framework.
 *
 * Every entry point returns 0 on success; on failure it returns non-zero and
 * ddpb_last_error() describes the problem (thread local).  `stream` is a cudaStream_t passed
 * as void*.  Activations are NHWC bf16 matrices [N*H*W][C]; 3x3 filters are bf16
 * [C_out][9][C_in] (the fp32 masters have the same order); accumulation is fp32.
 *
 * Each comment names the operation of the reference it replaces (file:line under
 * /root/reference, op numbers O1..O16 and kernel numbers K1..K16 from SURVEY.md).  The
 * reference itself ships no native code (SURVEY.md par. 2.5): these are new implementations.
 */
#ifndef DDPB200_H_
#define DDPB200_H_

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ------------------------------------------------------------------------------ */
const char* ddpb_last_error(void);
int ddpb_abi_version(void);          /* bumped when a signature changes */
int ddpb_sm_count(void);             /* SMs of the current device (148 without a device) */
/* Programmatic dependent launch between the library's consecutive kernels (default on, env
 * DDPB_PDL=0): each kernel's prologue overlaps the previous kernel's tail. */
void ddpb_set_pdl(int on);
int ddpb_get_pdl(void);
/* Diagnostics: a one-thread kernel writes %globaltimer (ns) to *slot at this point of the stream. */
int ddpb_stamp(unsigned long long* slot, void* stream);

/* Loss read-back without a copy call (part1/main.py:80 reads loss.item() every iteration, SURVEY.md
 * Q8): a one-thread kernel bumps *dev_seq and stores (sequence << 32 | bits of *value) as one 8-byte
 * word into pinned host memory; the host polls the word for the sequence number it expects. */
int ddpb_publish_scalar(const float* value, unsigned int* dev_seq, unsigned long long* host_slot, void* stream);

/* ---- convolution (K1, K2, K5, K6) ----------------------------------------------------------- */
/* O1: nn.Conv2d(3x3, stride 1, pad 1, bias) forward, part1/model.py:18-23.
 * y[N*H*W][C_out] = conv(x[N][H][W][C_in], w[C_out][taps*C_in]) (+bias), bf16 out.
 * taps = 9, or 1 for the first layer's pre-gathered operand (see ddpb_first_layer_operand_*).
 * stats (fp32 [2][C_out], may be null) is ACCUMULATED with per-channel sum / sum of squares of
 * y (the BatchNorm statistics, O4).  block_n: 0 = automatic.  scratch: fp32 buffer of
 * scratch_elems floats for split-K (ddpb_conv_fprop_scratch_elems() says how many the planner
 * would like for this shape), or null to forbid splitting.  A split-K plan that does not fit in
 * scratch_elems is replaced by one without split-K; nothing is ever written past the buffer. */
int ddpb_conv_fprop(const void* x, const void* w, const float* bias, void* y, float* stats, int N, int H, int W,
                    int C_in, int C_out, int taps, int block_n, float* scratch, long long scratch_elems,
                    unsigned int* split_sync, void* stream);
long long ddpb_conv_fprop_scratch_elems(int N, int H, int W, int C_in, int C_out, int taps, int block_n);

/* O1+O4+O6+O7 in one call: convolution, then BatchNorm(batch statistics) + ReLU (+2x2 max-pool),
 * part1/model.py:15-25.  conv_out ([N*H*W][C_out] bf16) is still written (the backward pass reads
 * it); act is the block's output; saved_mean_rstd fp32 [2][C_out].  stats: fp32 scratch of
 * 2*C_out + 32 floats, ZERO on entry.  The persistent kernel normalises its own tiles after a
 * grid-wide barrier; split-K shapes run the separate finalize and BatchNorm kernels.
 * scratch / scratch_elems as for ddpb_conv_fprop. */
int ddpb_conv_bn_relu_pool_fwd(const void* x, const void* w, const float* bias, void* conv_out, float* stats,
                               const float* gamma, const float* beta, void* act, float* saved_mean_rstd, int N, int H,
                               int W, int C_in, int C_out, int taps, int pool, float eps, float* scratch,
                               long long scratch_elems, unsigned int* split_sync, void* stream);

/* O2: data gradient of the same convolution (autograd of part1/model.py:18-23).
 * dx[N*H*W][C_in] from dy[..][C_out] and the SAME filter pack fprop reads (no transposed copy).
 * scratch sized with ddpb_conv_fprop_scratch_elems(N, H, W, C_out, C_in, 9, block_n). */
int ddpb_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int C_in, int C_out, int block_n,
                    float* scratch, long long scratch_elems, unsigned int* split_sync, void* stream);

/* O3: weight gradient.  dw (fp32 [C_out][dw_ld]) gets its first dw_cols_valid (<= taps*C_in)
 * columns per row.  scratch: scratch_elems floats of device memory for the split reduction over
 * pixels; ddpb_conv_wgrad_scratch_elems_ld() gives the need for this shape AND dw layout (a dw
 * with padded rows or fewer valid columns goes through scratch even without a split),
 * ddpb_conv_wgrad_scratch_elems() the need for a dense dw (dw_cols_valid == dw_ld == taps*C_in).
 * Null / 0 are allowed when the need is 0; a buffer that is too small is reported as an error. */
int ddpb_conv_wgrad(const void* dy, const void* x, float* dw, float* scratch, long long scratch_elems, int N, int H,
                    int W, int C_in, int C_out, int taps, int dw_cols_valid, int dw_ld, void* stream);
long long ddpb_conv_wgrad_scratch_elems(int N, int H, int W, int C_in, int C_out, int taps);
long long ddpb_conv_wgrad_scratch_elems_ld(int N, int H, int W, int C_in, int C_out, int taps, int dw_cols_valid,
                                           int dw_ld);
int ddpb_conv_wgrad_splits(int N, int H, int W, int C_in, int C_out, int taps);
/* No counterpart in the reference.  ddpb_conv_wgrad for callers that must not allocate device
 * memory (a loop that cannot take cudaMalloc's device-wide synchronisation, a process whose
 * memory was handed to another allocator, a host application that forbids allocator calls):
 * same arguments minus `scratch`; the split-reduction scratch is a fixed array of
 * ddpb_conv_wgrad_static_max_elems() floats (4 Mi = 16 MiB) that every GPU loading the library
 * holds until exit.  Larger requests (ddpb_conv_wgrad_scratch_elems_ld() above that size) fail
 * with an error.  The first call on a device resolves the array's address: make it with that
 * device current, before any stream capture, and from one thread (the per-device cache is not
 * locked; later calls may come from any thread).  All calls share the array; calls that overlap
 * on the GPU race. */
int ddpb_conv_wgrad_static(const void* dy, const void* x, float* dw, int N, int H, int W, int C_in, int C_out,
                           int taps, int dw_cols_valid, int dw_ld, void* stream);
long long ddpb_conv_wgrad_static_max_elems(void);

/* Tuning / profiling hooks for the conv launch planner (kernel: 0 one tile per CTA, 2 persistent,
 * 3 persistent CTA pair with cta_group::2 -- no split-K, no 64-wide tile: such shapes fall back
 * to 2; -1 = automatic = 2). */
void ddpb_conv_force(int kernel, int block_n, int splits);
void ddpb_conv_plan(int N, int H, int W, int C_in, int C_out, int taps, int block_n, int* out3);
/* plan for a given call situation: with_tail = ddpb_conv_bn_relu_pool_fwd, with_sync = split_sync passed;
 * out4 = {kernel, block_n, splits, 1 if the split-K sum happens inside the kernel} */
void ddpb_conv_plan_ex(int N, int H, int W, int C_in, int C_out, int taps, int block_n, int with_tail, int with_sync,
                       int* out4);
void ddpb_conv_debug_timing(unsigned long long* per_cta_stamps);   /* 8 slots per CTA, see bench/conv_probe2.py */
/* split_sync (optional, may be null): ddpb_conv_split_sync_words() uint32 words, zeroed ONCE by the caller and left
 * zero by every launch.  With it, split-K partial sums are added up inside the persistent kernel (the warp that
 * draws the last ticket of a tile sums the slices in split order: deterministic) and the BatchNorm tail stays
 * fused; without it a second kernel sums them.  Launches that can run concurrently need separate buffers. */
int ddpb_conv_split_sync_words(void);
void ddpb_conv_fuse_split(int on);       /* tuning: 0 = never sum split-K inside the kernel */
void ddpb_conv_tail_from_tmem(int on);   /* fused BatchNorm tail re-reads the accumulators from TMEM: 1 (default) with one
                                            tile per CTA, 2 wherever the shape allows, 0 never (stored tiles from L2).
                                            Same output. */
void ddpb_conv_debug_mode(int mode);   /* timing probes, wrong results: bit 0 skip B loads, bit 1 skip A loads */

/* ---- BatchNorm + ReLU (+ MaxPool) (K3, K4) -------------------------------------------------- */
/* O4+O6+O7: BatchNorm2d(track_running_stats=False) -> ReLU -> [MaxPool2d(2,2)],
 * part1/model.py:15-16,24,25.  Batch statistics always (also in eval). */
int ddpb_channel_stats(const void* x, float* stats, long long P, int C, void* stream);
int ddpb_bn_relu_pool_fwd(const void* x, const float* stats, const float* gamma, const float* beta, void* y,
                          float* saved_mean_rstd, int N, int H, int W, int C, int pool, float eps, void* stream);
/* O5+O6+O7 backward.  sums: fp32 scratch of 2*C + 32 floats that must be ZERO on entry.  partials:
 * ddpb_bn_bwd_partials_elems() floats of scratch (contents irrelevant) for the default path: two
 * ordinary kernels, every sum taken in a fixed order (dgamma / dbeta bitwise reproducible, no float
 * atomics), dbias evaluated from the sums.  partials == null: the cooperative one-kernel variant.
 * ddpb_bn_bwd_set_fused: 3 default, 2 / 1 cooperative (two / one CTA per SM), 0 two-pass with atomics. */
int ddpb_bn_relu_pool_bwd(const void* x, const void* dy, const float* saved_mean_rstd, const float* gamma,
                          const float* beta, float* sums, void* dx, float* dgamma, float* dbeta, float* dbias,
                          float* partials, long long partial_elems, int N, int H, int W, int C, int pool, float eps,
                          void* stream);
long long ddpb_bn_bwd_partials_elems(int N, int H, int W, int C, int pool);
void ddpb_bn_bwd_set_fused(int mode);

/* ---- classifier + loss (K7, K8, K15) --------------------------------------------------------- */
/* O9+O10+O15: Linear(512,10) + CrossEntropyLoss(mean) + top-1, part1/model.py:40,45,
 * part1/main.py:75,105-106. */
int ddpb_fc_ce_fwd(const void* x, const float* w, const float* b, const long long* labels, float* logits,
                   float* row_loss, float* dlogits, int* row_correct, float* loss, int* correct, int B, int F, int K,
                   void* stream);
/* row_loss / row_correct / loss / correct (may all be null): also reduce the per-row losses of a
 * preceding ddpb_fc_ce_fwd(loss = correct = null) to the mean loss / correct count. */
int ddpb_fc_bwd(const void* x, const float* w, const float* dlogits, float* dw, float* db, void* dx, int B, int F,
                int K, const float* row_loss, const int* row_correct, float* loss, int* correct, void* stream);

/* ---- optimizer (K9) -------------------------------------------------------------------------- */
/* O11+O12: optim.SGD(lr, momentum, weight_decay).step() over the flat arena, part1/main.py:77,
 * 124-125.  shadow (bf16, may be null) receives the rounded parameters = the conv operands. */
int ddpb_sgd_step(float* p, const float* g, float* m, void* shadow, long long n, float lr, float momentum,
                  float weight_decay, float grad_scale, void* stream);
int ddpb_pack_weights(int count, const float* const* master, void* const* fprop, void* const* dgrad, const int* K,
                      const int* C, const int* taps, const int* fprop_ld, const int* flat, void* stream);

/* ---- input pipeline (K16) -------------------------------------------------------------------- */
/* O16: RandomCrop(32,4) / RandomHorizontalFlip / ToTensor / Normalize, part1/main.py:20-28, fused
 * with the gather that builds the first layer's [N*H*W][64] operand.
 * w0_master / w0_packed (both or neither null): fp32 filters [c_out0][27] of the first layer and
 * their zero-padded bf16 operand [c_out0][64], refreshed by one extra CTA of the same launch. */
int ddpb_first_layer_operand_f32(const float* src_nchw, void* out, int N, int H, int W, const float* w0_master,
                                 void* w0_packed, int c_out0, void* stream);
int ddpb_first_layer_operand_u8(const unsigned char* src_nhwc, const signed char* aug, void* out, int N, int H,
                                int W, const float* mean3, const float* std3, int pad, const float* w0_master,
                                void* w0_packed, int c_out0, void* stream);

/* Device-side random numbers for the input pipeline (Philox4x32-10 keyed by seed, counter = (*step, image)).
 * aug: int8 [N][4] = {crop y in [0, 2 pad], crop x in [0, 2 pad], flip, 0}, the layout
 * ddpb_first_layer_operand_u8 reads.  *step (device counter) is read and then incremented by the launch, so a
 * CUDA graph that contains it draws a fresh set on every replay. */
int ddpb_draw_augmentation(signed char* aug, int N, int pad, unsigned long long seed, unsigned int* step, void* stream);
/* Synthetic CIFAR-shaped batch generated on the device: uint8 NHWC pixels (per-class smooth colour template +
 * Gaussian noise) and int64 labels; increments *step. */
int ddpb_synthetic_batch(unsigned char* img_nhwc, long long* labels, int N, int H, int W, int classes,
                         unsigned long long seed, unsigned int* step, void* stream);

/* ---- gradient synchronisation over NVLink peer memory (K10-K14) ------------------------------ */
/* bufs: `world` base pointers of the symmetric buffers ([64 KiB signal pad | fp32 payload]) as
 * mapped in this process; mc: NVLS multicast mapping of the same buffer or null; off/n: payload
 * range in floats (multiples of 4); blocks/slot_base: CTAs and their signal-pad slots;
 * p/m/shadow: local parameter / momentum / bf16 arenas for the fused SGD step (p null = none);
 * sgd_off/sgd_n: range that update covers (sgd_n < 0: the synchronised range). */
int ddpb_comm_signal_pad_bytes(void);
int ddpb_comm_max_ranks(void);
/* part2b dist.all_reduce(SUM) + divide (part2/part2b/main.py:97-103) and part3's bucketed
 * all-reduce (part3/main.py:174): in-place two-shot all-reduce, mean, optional SGD. */
int ddpb_allreduce_avg(const void* const* bufs, const void* mc, int rank, int world, long long off, long long n,
                       int blocks, int slot_base, int use_mc, float* p, float* m, void* shadow, float lr,
                       float momentum, float weight_decay, long long sgd_off, long long sgd_n, int threads,
                       void* stream);
/* Many ranges, one launch: the reference's 34 per-parameter all-reduces (part2/part2b/main.py:101-103) or a set of
 * small buckets.  offs / ns: `count` (<= 64) pairwise disjoint ranges in floats (multiples of 4).  Ranges of at most
 * small_max floats are reduced one-shot (every rank reduces every element itself and keeps the mean in registers
 * across the closing barrier) while they fit in the threads' registers; the others two-shot.  p / m / shadow as
 * above: the SGD update covers exactly the listed ranges. */
int ddpb_allreduce_multi(const void* const* bufs, const void* mc, int rank, int world, const long long* offs,
                         const long long* ns, int count, long long small_max, int blocks, int slot_base, int use_mc,
                         float* p, float* m, void* shadow, float lr, float momentum, float weight_decay,
                         void* stream);
/* part2a gather -> mean on rank 0 -> scatter (part2/part2a/main.py:97-115). */
int ddpb_reduce_root_bcast(const void* const* bufs, const void* mc, int rank, int world, int root, long long off,
                           long long n, int blocks, int slot_base, int use_mc, float* p, float* m, void* shadow,
                           float lr, float momentum, float weight_decay, long long sgd_off, long long sgd_n,
                           void* stream);
/* DDP's constructor-time parameter broadcast. */
int ddpb_broadcast(const void* const* bufs, const void* mc, int rank, int world, int root, long long off,
                   long long n, int blocks, int slot_base, int use_mc, void* stream);
int ddpb_peer_barrier(const void* const* bufs, int rank, int world, int blocks, int slot_base, void* stream);
/* Own VMM + NVLS multicast bootstrap of the symmetric buffer (csrc/vmm.cu; the counterpart of the Gloo
 * rendezvous at part2b/main.py:52-58).  Handles cross processes as POSIX file descriptors; sizes must be
 * multiples of ddpb_vmm_granularity().  All calls act on the current device. */
int ddpb_vmm_granularity(int ndev, long long* bytes_out);
int ddpb_vmm_create(long long bytes, unsigned long long* handle, int* fd);   /* cuMemCreate + export */
int ddpb_vmm_import(int fd, unsigned long long* handle);                     /* allocation or multicast object */
int ddpb_vmm_map(unsigned long long handle, long long bytes, void** ptr);    /* reserve + map + read-write access */
int ddpb_vmm_unmap(void* ptr, long long bytes);
int ddpb_vmm_release(unsigned long long handle);
int ddpb_vmm_close_fd(int fd);
int ddpb_mc_create(int ndev, long long bytes, unsigned long long* handle, int* fd);   /* one rank */
int ddpb_mc_add_device(unsigned long long mc_handle);                                  /* every rank, before any bind */
int ddpb_mc_bind(unsigned long long mc_handle, unsigned long long mem_handle, long long bytes);
/* CUDA IPC bootstrap of the symmetric buffer (last fallback: peer mappings only, no multicast). */
int ddpb_ipc_alloc(long long bytes, void** ptr, void* handle64);
int ddpb_ipc_open(const void* handle64, void** ptr);
int ddpb_ipc_close(void* ptr);
int ddpb_ipc_free(void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* DDPB200_H_ */
