/* b200seg -- C ABI of the sm_100a kernels (lib/libb200seg.so).
 *
 * Every function launches on the given stream and returns a cudaError_t value as int (0 = ok).
 * Activations are bf16 NHWC; `ld_*` is the pixel stride in elements (>= C: tensors may be
 * channel slices of a wider buffer).  Statistics, BN vectors, parameters, gradients are fp32.
 * "R:" cites /root/reference/"Vaihingen PyTorch 2 (кластер).py", whose stock-PyTorch op each
 * kernel replaces (the reference itself ships no device code).
 */
#ifndef B200SEG_H_
#define B200SEG_H_

#include <cuda_runtime.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- BatchNorm(train) + ReLU (+ 2x2 max-pool): R:580-581, R:583-584, R:595 ------------------- */
int b200seg_stats_rows(void);
int b200seg_channel_stats(const void* y, long long P, int C, long long ld, float* partial, int rows,
                          cudaStream_t st);
/* conv_bias: bias of the producing conv (may be NULL); the conv kernels never add it, see bn_act.cu */
int b200seg_bn_finalize(const float* partial, int rows, long long count, const float* gamma,
                        const float* beta, const float* conv_bias, float* running_mean,
                        float* running_var, long long* nbt, float* ss, int C, float momentum, float eps,
                        int training, cudaStream_t st);
int b200seg_bn_relu_pool_fwd(const void* y, long long ld_y, const float* ss, void* out, long long ld_o,
                             void* pool, long long ld_p, int N, int H, int W, int C, cudaStream_t st);
/* workspace: float[b200seg_stats_rows()*2*C + 3*C] */
int b200seg_bn_relu_pool_bwd(const void* d_out, long long ld_do, const void* d_pool, long long ld_dp,
                             const void* y, long long ld_y, const float* ss, const float* gamma,
                             float* dgamma, float* dbeta, void* dy, long long ld_dy, int N, int H, int W,
                             int C, int pooled, float* workspace, cudaStream_t st);

/* BN backward of a stage whose two per-channel sums came out of the kernel that produced d_out
 * (b200seg_head_tail): finalize + apply only.  partial: float[rows][2][C]; coef: float[3*C] scratch. */
int b200seg_bn_relu_bwd_from_partial(const float* partial, int rows, const void* d_out, long long ld_do,
                                     const void* y, long long ld_y, const float* ss, const float* gamma,
                                     float* dgamma, float* dbeta, void* dy, long long ld_dy, int N, int H,
                                     int W, int C, float* coef, cudaStream_t st);

/* ---- 1x1 head + softmax cross-entropy + accuracy + head backward: R:641, R:703, R:755, R:775 -- */
long long b200seg_head_ce_workspace_bytes(int C, int K);
/* workspace: the last 256 bytes hold a CTA counter that must be zero before the first launch. */
int b200seg_head_ce(const void* a, long long ld_a, const float* w, const float* bias,
                    const unsigned char* labels, long long P, int C, int K, float grad_scale, void* da,
                    long long ld_da, float* dw, float* db, unsigned char* pred, float* loss_acc,
                    void* workspace, cudaStream_t st);

/* The whole tail in one pass over the last conv's RAW output y: BatchNorm(train)+ReLU (R:583-584) applied
 * on load, head + loss + accuracy + head backward as b200seg_head_ce, and the two per-channel sums of that
 * BatchNorm's backward (sum g*[a>0], sum g*[a>0]*y) into bn_partial[rows][2][C], rows =
 * b200seg_head_tail_rows(P, C).  The activated tensor never reaches HBM; g_out receives d(loss)/d(a).
 * C in {32, 64, 128}, K <= 8 (b200seg_head_tail_supported).  ss: float[4][C] from b200seg_bn_finalize.
 * g_out == NULL: forward only (loss, accuracy, pred).  Workspace: as b200seg_head_ce. */
int b200seg_head_tail_supported(int C, int K);
int b200seg_head_tail_rows(long long P, int C);
int b200seg_head_tail(const void* y, long long ld_y, const float* ss, const float* w, const float* bias,
                      const unsigned char* labels, long long P, int C, int K, float grad_scale, void* g_out,
                      long long ld_g, float* dw, float* db, unsigned char* pred, float* loss_acc,
                      float* bn_partial, void* workspace, cudaStream_t st);

/* Same computation with NO workspace argument and no allocation of any kind.  (No counterpart in the
 * reference: it has no device code and no notion of a workspace.)  For callers that must not take
 * device memory at call time -- hosts that give the whole GPU to their own pool and forbid driver
 * allocations afterwards, latency-critical loops (cudaMalloc / cudaFree serialise the device), plain C
 * programs without a framework allocator.  The per-CTA partial sums live in an array built into the
 * library (b200seg_head_ce_static_scratch_bytes(), ~0.6 MB, resident on every GPU that loads the
 * library).  Limits: C <= 128, K <= 8; anything larger returns cudaErrorInvalidValue with a message on
 * stderr and launches nothing.  All calls on a device share that array: calls that may overlap on the
 * GPU race.  The first call on each device resolves the array's device address; make it before any
 * stream capture begins. */
long long b200seg_head_ce_static_scratch_bytes(void);
int b200seg_head_ce_static(const void* a, long long ld_a, const float* w, const float* bias,
                           const unsigned char* labels, long long P, int C, int K, float grad_scale,
                           void* da, long long ld_da, float* dw, float* db, unsigned char* pred,
                           float* loss_acc, cudaStream_t st);

/* ---- optimizer over the flat buffers: R:704, R:437-438, R:552-553 ------------------------------ */
int b200seg_fused_adam(float* p, float* g, float* m, float* v, long long n, float lr, float b1, float b2,
                       float eps, float wd, float gscale, float bc1, float bc2, void* p_lowp,
                       int zero_grad, cudaStream_t st);
int b200seg_fused_sgd(float* p, float* g, float* buf, long long n, float lr, float momentum, float wd,
                      float gscale, void* p_lowp, int zero_grad, int first_step, cudaStream_t st);

/* ---- gradient codecs: R:451-496 (encode), R:513-545 (decode) -----------------------------------
 * codec: 0 fp32, 1 bf16, 2 fp16, 3 q201, 4 q21, 5 int8s, 6 fp8 */
int b200seg_amax(const float* x, long long n, float* out, cudaStream_t st);
int b200seg_codec_encode(const float* x, long long n, const float* scale, int codec, void* out,
                         cudaStream_t st);
int b200seg_codec_decode(const void* in, long long n, const float* scale, int codec, float* out,
                         int accumulate, cudaStream_t st);

/* ---- input pipeline and bilinear up-sampling: R:737-741, R:754; R:609 --------------------------- */
int b200seg_prep_input(const unsigned char* img, long long P, int Cin, void* out, int Cpad,
                       cudaStream_t st);
int b200seg_upsample2x_fwd(const void* x, long long ld_x, void* out, long long ld_o, int N, int h, int w,
                           int C, cudaStream_t st);
int b200seg_upsample2x_bwd(const void* g, long long ld_g, void* dx, long long ld_x, int N, int h, int w,
                           int C, cudaStream_t st);

/* glue that keeps eager framework ops out of the captured step (no reference counterpart: autograd does this
 * bookkeeping there): pad bf16 rows cin -> cpad with zeros; dst[r][c<cin] += src[r][c], src := 0;
 * out[c] += sum_r part[r * row_stride + c]. */
int b200seg_pad_rows(const void* src, int rows, int cin, void* dst, int cpad, cudaStream_t st);
int b200seg_unpad_add_zero(float* src, int rows, int cpad, float* dst, int cin, cudaStream_t st);
int b200seg_rows_sum_add(const float* part, int rows, long long row_stride, int C, float* out, cudaStream_t st);

/* ---- gradient all-reduce over NVLink peer memory: R:255-440, R:443-556 -------------------------
 * Pointer arrays are indexed by rank and hold addresses valid in THIS process (symmetric memory).
 * wire_kind: 0 fp32, 1 bf16, 2 fp16 (also q201), 3 int8 (q21, int8s), 4 fp8.  n_elems % 16 == 0.
 * oneshot_out == NULL: in-place two-shot into every rank's grad buffer; else one-shot into
 * oneshot_out.  Two epochs are consumed per call: pass epoch = 2 * call_index + 1. */
/* Every device-side wait for a peer is bounded (default 600 s; the reference blocks forever in recv, R:98).
 * On expiry the waiting rank stores 0x80000000 | rank into word b200seg_allreduce_abort_word() of EVERY
 * rank's signal pad and the kernels return without reducing; the host polls that word. */
void b200seg_set_comm_timeout(double seconds);
int b200seg_allreduce_abort_word(void);
int b200seg_allreduce_max_blocks(void);
long long b200seg_allreduce_signal_bytes(void);
int b200seg_write_mult(const float* scale, float levels, float* out, cudaStream_t st);
int b200seg_allreduce(const void* const* wire_ptrs, float* const* grad_ptrs,
                      unsigned int* const* signal_ptrs, const float* const* mult_ptrs, int rank, int world,
                      long long n_elems, int wire_kind, float post_scale, unsigned int epoch,
                      float* oneshot_out, int blocks, cudaStream_t st);

/* Quantize-on-send two-shot for the lossy codecs (1 bf16, 2 fp16, 3 q201, 4 q21, 5 int8s, 6 fp8): R:451-496 and
 * R:513-545 without an encode pass -- the sender quantizes in registers and stores into the owner's inbox over
 * NVLink, the owner dequantizes, adds in rank order in fp32 and stores fp32 into every rank's gradient buffer.
 * scale: this rank's max-abs on the device (b200seg_amax), NULL for bf16 / fp16.  n_elems % 64 == 0. */
int b200seg_allreduce_push(void* const* inbox_ptrs, float* const* grad_ptrs, unsigned int* const* signal_ptrs,
                           float* const* mult_ptrs, const float* scale, int codec, int rank, int world,
                           long long n_elems, float post_scale, unsigned int epoch, int blocks, cudaStream_t st);

/* NVLS: the NVSwitch performs the addition (multimem.ld_reduce / multimem.st on the multicast mapping of
 * the symmetric fp32 buffer).  multicast_ptr == NULL -> cudaErrorInvalidValue (use b200seg_allreduce). */
int b200seg_allreduce_nvls(unsigned int* const* signal_ptrs, float* multicast_ptr, int rank, int world,
                           long long n_elems, float post_scale, unsigned int epoch, int blocks,
                           cudaStream_t st);

/* Error record of the tcgen05 kernels (no reference counterpart): >= 8 words of pinned, device-mapped host memory,
 * registered per device.  A bounded mbarrier wait that times out writes {0xB200DEAD, barrier address, parity, CTA,
 * thread} there and then traps; the record outlives the CUDA context. */
int b200seg_set_error_record(unsigned int* host_mapped);

/* ---- tcgen05 / TMEM / TMA convolutions (csrc/conv): R:579, R:582 (3x3 conv), their autograd
 * backward (dgrad, wgrad), R:607 (ConvTranspose2d k=2 s=2 fused with the concat placement, R:616).
 * Weights are bf16 in the flat-buffer order (Cout,3,3,Cin); dgrad reads them un-transposed as an
 * MN-major UMMA operand.  `partials` (fprop): float[rows][2][Cout] per-CTA sum / sum of squares
 * of the bf16-rounded output, rows = b200conv_fprop_stat_rows(...). */
int b200conv_supported(int cin, int cout);
/* The kernel variant and tiling a conv call of a given shape takes (host-only: no launch, no GPU needed).
 * op: 0 fprop, 1 dgrad, 2 wgrad, 3 transposed-conv forward, 4 transposed-conv dgrad.
 * kernel: 0 tap GEMM on square pixel boxes, 1 tap GEMM on halo strips (whole 128-pixel row segments);
 *         for op 2: 2 split-K wgrad, 3 128-column strip wgrad, 4 row-pair wgrad (other fields unused). */
typedef struct B200ConvPlan {
  int kernel, block_k, block_n, n_tiles, ocb;
  int rows_per_unit;      /* output rows per work unit: 1, 2 (streamed weights) or 4 (vertical taps merged along N) */
  int weights_resident;   /* the CTA's slice of the filter bank stays in shared memory for its whole life */
  int private_epilogue;   /* per-warp epilogue pipelines (no CTA-wide barrier per tile) */
  int staging_bufs, strip_slots, stages, grid, cluster, m_tiles;
  long long smem_bytes;
} B200ConvPlan;
int b200conv_plan(int op, int N, int H, int W, int Cin, int Cout, B200ConvPlan* out);
int b200conv_fprop_stat_rows(int N, int H, int W, int Cin, int Cout);
int b200conv_fprop(const void* x, long long ld_x, const void* w, const float* bias, void* y,
                   long long ld_y, int N, int H, int W, int Cin, int Cout, float* partials,
                   cudaStream_t st);
int b200conv_dgrad(const void* dy, long long ld_dy, const void* w, void* dx, long long ld_dx, int N,
                   int H, int W, int Cin, int Cout, cudaStream_t st);
/* dgrad + per-CTA column sums of dx: partials = float[rows][2][Cin], rows = b200conv_dgrad_stat_rows(...) */
int b200conv_dgrad_stat_rows(int N, int H, int W, int Cin, int Cout);
int b200conv_dgrad_stats(const void* dy, long long ld_dy, const void* w, void* dx, long long ld_dx, int N,
                         int H, int W, int Cin, int Cout, float* partials, cudaStream_t st);
int b200conv_wgrad(const void* x, long long ld_x, const void* dy, long long ld_dy, float* dw, int N,
                   int H, int W, int Cin, int Cout, cudaStream_t st);
/* the same for filter taps tap0 .. tap0+ntaps-1 only (tap = 3*r + s), split-K kernel */
int b200conv_wgrad_taps(const void* x, long long ld_x, const void* dy, long long ld_dy, float* dw, int N,
                        int H, int W, int Cin, int Cout, int tap0, int ntaps, cudaStream_t st);
int b200conv_convT_fwd(const void* x, long long ld_x, const void* w, const float* bias, void* out,
                       long long ld_o, int N, int h, int w_, int Cin, int Cout, cudaStream_t st);
/* backward of the transposed conv: d_up (N,2h,2w,Cout) slice; w, dw laid out [(dy,dx,co)][ci] */
int b200conv_convT_dgrad(const void* d_up, long long ld_g, const void* w, void* dx, long long ld_dx, int N,
                         int h, int w_, int Cin, int Cout, cudaStream_t st);
int b200conv_convT_wgrad(const void* x, long long ld_x, const void* d_up, long long ld_g, float* dw, int N,
                         int h, int w_, int Cin, int Cout, cudaStream_t st);

/* One cross-GPU barrier over the all-reduce signal pads (one epoch of the shared counter). */
int b200seg_barrier(unsigned int* const* signal_ptrs, int rank, int world, unsigned int epoch, cudaStream_t st);
/* Fused weight-gradient + all-reduce (no counterpart in the reference: its exchange is a separate TCP
 * round, R:255-556).  Registers the multicast alias of the gradient buffer; weight-gradient launchers
 * whose destination lies inside [local_base, local_base + bytes) then add with multimem.red through
 * the alias, i.e. into every replica at once.  multicast_base == NULL unregisters. */
int b200seg_set_grad_multicast(const void* local_base, long long bytes, const void* multicast_base);
long long b200seg_grad_multicast_delta(const void* p);

/* ---- gradient exchange overlapped with backward (csrc/comm/overlap.cu): replaces the strictly-after-
 * backward round of R:756-763 -> R:255-440 / R:443-556.  The flat gradient buffer is reduced bucket by
 * bucket while later layers still compute.
 *  b200seg_arm_next_wgrad: the LAST kernel launched by the next b200conv_wgrad / b200conv_convT_wgrad
 *    call becomes the bucket's trigger: its last CTA stores the exchange number into flags[bucket][rank]
 *    of every rank's pad.  `arm` is a device pointer to {u32* counter; const u32* epoch;
 *    u32* peer_flag[16]; int world; int pad;} (all device addresses); NULL disarms.
 *  b200seg_bucket_reduce: waits until every rank has published the bucket, then reduces it in place
 *    (rank r owns slice r of every range): NVLS through `multicast_ptr`, or peer loads / stores over
 *    `grad_ptrs` when multicast_ptr is NULL.  offs / counts are in floats (multiples of 4), up to 4
 *    ranges.  `state` = local device u32[2] {exchange number (start at 1), CTA counter (0)}.
 *    self_publish: the kernel publishes the bucket flag itself (no arming wgrad launch).  final: last
 *    bucket of the exchange: ends with a cross-rank "done" exchange and advances the exchange number.
 *    Waits are bounded by timeout_s; on expiry the abort word (b200seg_overlap_abort_index()) is raised in
 *    every rank's pad and the kernels return without reducing. */
void b200seg_arm_next_wgrad(const void* arm);
int b200seg_overlap_pad_words(void);
int b200seg_overlap_max_buckets(void);
int b200seg_overlap_flag_index(int bucket, int rank);
int b200seg_overlap_abort_index(void);
int b200seg_bucket_reduce(float* multicast_ptr, float* const* grad_ptrs, unsigned int* const* pad_ptrs,
                          unsigned int* state, int rank, int world, int bucket, int self_publish, int final,
                          int n_ranges, const long long* offs, const long long* counts, float post_scale,
                          double timeout_s, int blocks, cudaStream_t st);

#ifdef __cplusplus
}
#endif
#endif /* B200SEG_H_ */
