// torch.ops.b200seg.* : thin tensor-checking wrappers over the C ABI in include/b200seg.h.
// All launches go to the current CUDA stream, so they are capturable in CUDA graphs.
#include <torch/library.h>
#include <ATen/ATen.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <optional>
#include <vector>
#include <string>
#include <map>
#include "b200seg.h"

namespace {

using at::Tensor;
using OptT = std::optional<Tensor>;

inline cudaStream_t cur() { return at::cuda::getCurrentCUDAStream(); }

inline void ok(int err, const char* what) {
  TORCH_CHECK(err == 0, "b200seg::", what, " failed: ", cudaGetErrorString((cudaError_t)err));
}

// (N,H,W,C) view with contiguous channels and "dense pixels": returns the pixel stride.
inline int64_t nhwc_ld(const Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda() && t.dim() == 4, name, ": expected a CUDA (N,H,W,C) tensor");
  TORCH_CHECK(t.stride(3) == 1, name, ": channels must be contiguous");
  const int64_t ld = t.stride(2);
  TORCH_CHECK(t.stride(1) == ld * t.size(2) && t.stride(0) == ld * t.size(2) * t.size(1), name,
              ": pixels must be densely strided");
  TORCH_CHECK(ld % 8 == 0 && (reinterpret_cast<uintptr_t>(t.data_ptr()) % 16) == 0, name,
              ": needs 16-byte aligned pixels");
  return ld;
}
inline void* ptr_or_null(const OptT& t) { return t.has_value() ? t->data_ptr() : nullptr; }
inline float* f32(const Tensor& t) {
  TORCH_CHECK(t.scalar_type() == at::kFloat && t.is_cuda(), "expected a CUDA float32 tensor");
  return t.data_ptr<float>();
}
inline void bf16(const Tensor& t, const char* n) {
  TORCH_CHECK(t.scalar_type() == at::kBFloat16, n, ": expected bfloat16");
}

int64_t stats_rows() { return b200seg_stats_rows(); }

void channel_stats(const Tensor& y, Tensor partial) {
  c10::cuda::CUDAGuard g(y.device());
  bf16(y, "y");
  int64_t ld = nhwc_ld(y, "y");
  ok(b200seg_channel_stats(y.data_ptr(), y.size(0) * y.size(1) * y.size(2), (int)y.size(3), ld, f32(partial),
                           (int)partial.size(0), cur()), "channel_stats");
}

void bn_finalize(const OptT& partial, int64_t count, const Tensor& gamma, const Tensor& beta,
                 const OptT& conv_bias, const Tensor& rmean, const Tensor& rvar, const Tensor& nbt, Tensor ss, double momentum,
                 double eps, bool training) {
  c10::cuda::CUDAGuard g(ss.device());
  const int C = (int)gamma.numel();
  TORCH_CHECK(!training || partial.has_value(), "training-mode bn_finalize needs the partial sums");
  TORCH_CHECK(nbt.scalar_type() == at::kLong, "num_batches_tracked must be int64");
  ok(b200seg_bn_finalize(partial.has_value() ? f32(*partial) : nullptr,
                         partial.has_value() ? (int)partial->size(0) : 0, count, f32(gamma), f32(beta),
                         conv_bias.has_value() ? f32(*conv_bias) : nullptr, f32(rmean), f32(rvar), (long long*)nbt.data_ptr<int64_t>(), f32(ss), C,
                         (float)momentum, (float)eps, training ? 1 : 0, cur()), "bn_finalize");
}

void bn_relu_pool_fwd(const Tensor& y, const Tensor& ss, Tensor out, OptT pool) {
  c10::cuda::CUDAGuard g(y.device());
  bf16(y, "y"); bf16(out, "out");
  int64_t ldy = nhwc_ld(y, "y"), ldo = nhwc_ld(out, "out"), ldp = 0;
  if (pool.has_value()) { bf16(*pool, "pool"); ldp = nhwc_ld(*pool, "pool"); }
  ok(b200seg_bn_relu_pool_fwd(y.data_ptr(), ldy, f32(ss), out.data_ptr(), ldo, ptr_or_null(pool), ldp,
                              (int)y.size(0), (int)y.size(1), (int)y.size(2), (int)y.size(3), cur()),
     "bn_relu_pool_fwd");
}

void bn_relu_pool_bwd(const OptT& d_out, const OptT& d_pool, const Tensor& y, const Tensor& ss,
                      const Tensor& gamma, Tensor dgamma, Tensor dbeta, const OptT& dbias, Tensor dy) {
  c10::cuda::CUDAGuard g(y.device());
  (void)dbias;   // a conv bias that feeds a train-mode BN has an exactly zero gradient
  bf16(y, "y"); bf16(dy, "dy");
  const int C = (int)y.size(3);
  int64_t ld_do = 0, ld_dp = 0;
  if (d_out.has_value()) { bf16(*d_out, "d_out"); ld_do = nhwc_ld(*d_out, "d_out"); }
  if (d_pool.has_value()) { bf16(*d_pool, "d_pool"); ld_dp = nhwc_ld(*d_pool, "d_pool"); }
  auto ws = at::empty({(int64_t)b200seg_stats_rows() * 2 * C + 3 * C}, y.options().dtype(at::kFloat));
  ok(b200seg_bn_relu_pool_bwd(ptr_or_null(d_out), ld_do, ptr_or_null(d_pool), ld_dp, y.data_ptr(),
                              nhwc_ld(y, "y"), f32(ss), f32(gamma), f32(dgamma), f32(dbeta), dy.data_ptr(),
                              nhwc_ld(dy, "dy"), (int)y.size(0), (int)y.size(1), (int)y.size(2), C,
                              d_pool.has_value() ? 1 : 0, ws.data_ptr<float>(), cur()),
     "bn_relu_pool_bwd");
}

// One zeroed workspace per (device, size) for the head kernel's CTA partials + counter.
Tensor& head_workspace(const Tensor& like, int64_t bytes) {
  static std::map<std::pair<int, int64_t>, Tensor> cache;
  auto key = std::make_pair((int)like.get_device(), bytes);
  auto it = cache.find(key);
  if (it == cache.end()) {
    it = cache.emplace(key, at::zeros({bytes}, like.options().dtype(at::kByte))).first;
  }
  return it->second;
}

void head_ce_fwd_bwd(const Tensor& a, const Tensor& w, const Tensor& b, const Tensor& labels,
                     double grad_scale, OptT da, OptT dw, OptT db, OptT pred, Tensor loss_acc) {
  c10::cuda::CUDAGuard g(a.device());
  bf16(a, "a");
  const int C = (int)a.size(3), K = (int)w.size(0);
  TORCH_CHECK(w.size(1) == C && w.is_contiguous(), "head weight must be (classes, C) contiguous");
  TORCH_CHECK(labels.scalar_type() == at::kByte && labels.is_contiguous(), "labels must be uint8");
  int64_t ld_da = 0;
  if (da.has_value()) { bf16(*da, "da"); ld_da = nhwc_ld(*da, "da"); }
  Tensor& ws = head_workspace(a, b200seg_head_ce_workspace_bytes(C, K));
  ok(b200seg_head_ce(a.data_ptr(), nhwc_ld(a, "a"), f32(w), f32(b), labels.data_ptr<uint8_t>(),
                     a.size(0) * a.size(1) * a.size(2), C, K, (float)grad_scale, ptr_or_null(da), ld_da,
                     dw.has_value() ? f32(*dw) : nullptr, db.has_value() ? f32(*db) : nullptr,
                     pred.has_value() ? pred->data_ptr<uint8_t>() : nullptr, f32(loss_acc), ws.data_ptr(),
                     cur()), "head_ce");
}

// BN(train)+ReLU on load + head + loss + head backward + the BN-backward sums, one pass over the raw conv output
int64_t head_tail_rows(int64_t P, int64_t C) { return b200seg_head_tail_rows(P, (int)C); }
bool head_tail_supported(int64_t C, int64_t K) { return b200seg_head_tail_supported((int)C, (int)K) != 0; }

void head_tail_fwd_bwd(const Tensor& y, const Tensor& ss, const Tensor& w, const Tensor& b, const Tensor& labels,
                       double grad_scale, OptT g_out, OptT dw, OptT db, OptT pred, Tensor loss_acc, OptT bn_partial) {
  c10::cuda::CUDAGuard g(y.device());
  bf16(y, "y");
  const int C = (int)y.size(3), K = (int)w.size(0);
  const int64_t P = y.size(0) * y.size(1) * y.size(2);
  TORCH_CHECK(b200seg_head_tail_supported(C, K), "head_tail: unsupported (C, classes) = (", C, ", ", K, ")");
  TORCH_CHECK(w.size(1) == C && w.is_contiguous(), "head weight must be (classes, C) contiguous");
  TORCH_CHECK(ss.is_contiguous() && ss.size(0) == 4 && ss.size(1) == C, "ss must be (4, C)");
  TORCH_CHECK(labels.scalar_type() == at::kByte && labels.is_contiguous() && labels.numel() == P, "labels must be uint8, one per pixel");
  int64_t ld_g = 0;
  if (g_out.has_value()) {
    bf16(*g_out, "g_out");
    ld_g = nhwc_ld(*g_out, "g_out");
    TORCH_CHECK(dw.has_value() && db.has_value() && bn_partial.has_value(), "head_tail backward needs dw, db and bn_partial");
    TORCH_CHECK(bn_partial->is_contiguous() && bn_partial->numel() >= b200seg_head_tail_rows(P, C) * 2 * (int64_t)C,
                "bn_partial too small");
  }
  Tensor& ws = head_workspace(y, b200seg_head_ce_workspace_bytes(C, K));
  ok(b200seg_head_tail(y.data_ptr(), nhwc_ld(y, "y"), f32(ss), f32(w), f32(b), labels.data_ptr<uint8_t>(), P, C, K,
                       (float)grad_scale, ptr_or_null(g_out), ld_g, dw.has_value() ? f32(*dw) : nullptr,
                       db.has_value() ? f32(*db) : nullptr, pred.has_value() ? pred->data_ptr<uint8_t>() : nullptr,
                       f32(loss_acc), bn_partial.has_value() ? f32(*bn_partial) : nullptr, ws.data_ptr(), cur()),
     "head_tail");
}

void bn_relu_bwd_from_partial(const Tensor& partial, int64_t rows, const Tensor& d_out, const Tensor& y, const Tensor& ss,
                              const Tensor& gamma, Tensor dgamma, Tensor dbeta, Tensor dy) {
  c10::cuda::CUDAGuard g(y.device());
  bf16(y, "y"); bf16(dy, "dy"); bf16(d_out, "d_out");
  const int C = (int)y.size(3);
  TORCH_CHECK(partial.is_contiguous() && partial.numel() >= rows * 2 * (int64_t)C, "partial too small");
  auto coef = at::empty({3 * (int64_t)C}, y.options().dtype(at::kFloat));
  ok(b200seg_bn_relu_bwd_from_partial(f32(partial), (int)rows, d_out.data_ptr(), nhwc_ld(d_out, "d_out"), y.data_ptr(),
                                      nhwc_ld(y, "y"), f32(ss), f32(gamma), f32(dgamma), f32(dbeta), dy.data_ptr(),
                                      nhwc_ld(dy, "dy"), (int)y.size(0), (int)y.size(1), (int)y.size(2), C,
                                      coef.data_ptr<float>(), cur()),
     "bn_relu_bwd_from_partial");
}

// host-only: which kernel variant / tiling a conv of this shape takes (see B200ConvPlan in b200seg.h)
std::vector<int64_t> conv_plan(int64_t op, int64_t N, int64_t H, int64_t W, int64_t Cin, int64_t Cout) {
  B200ConvPlan q{};
  ok(b200conv_plan((int)op, (int)N, (int)H, (int)W, (int)Cin, (int)Cout, &q), "conv_plan");
  return {q.kernel, q.block_k, q.block_n, q.n_tiles, q.ocb, q.rows_per_unit, q.weights_resident, q.private_epilogue,
          q.staging_bufs, q.strip_slots, q.stages, q.grid, q.cluster, q.m_tiles, q.smem_bytes};
}

void fused_adam(Tensor p, Tensor gr, Tensor m, Tensor v, double lr, double b1, double b2, double eps,
                double wd, double gscale, double bc1, double bc2, OptT p_lowp, bool zero_grad) {
  c10::cuda::CUDAGuard g(p.device());
  if (p_lowp.has_value()) bf16(*p_lowp, "p_lowp");
  ok(b200seg_fused_adam(f32(p), f32(gr), f32(m), f32(v), p.numel(), (float)lr, (float)b1, (float)b2,
                        (float)eps, (float)wd, (float)gscale, (float)bc1, (float)bc2, ptr_or_null(p_lowp),
                        zero_grad ? 1 : 0, cur()), "fused_adam");
}

void fused_sgd(Tensor p, Tensor gr, OptT buf, double lr, double momentum, double wd, double gscale,
               OptT p_lowp, bool zero_grad, bool first_step) {
  c10::cuda::CUDAGuard g(p.device());
  ok(b200seg_fused_sgd(f32(p), f32(gr), buf.has_value() ? f32(*buf) : nullptr, p.numel(), (float)lr,
                       (float)momentum, (float)wd, (float)gscale, ptr_or_null(p_lowp), zero_grad ? 1 : 0,
                       first_step ? 1 : 0, cur()), "fused_sgd");
}

int codec_id(const std::string& name) {
  static const char* names[] = {"fp32", "bf16", "fp16", "q201", "q21", "int8s", "fp8"};
  for (int i = 0; i < 7; ++i) if (name == names[i]) return i;
  TORCH_CHECK(false, "unknown codec ", name);
  return -1;
}

void amax(const Tensor& x, Tensor out) {
  c10::cuda::CUDAGuard g(x.device());
  ok(b200seg_amax(f32(x), x.numel(), f32(out), cur()), "amax");
}
void codec_encode(const Tensor& x, const Tensor& scale, Tensor out_bytes, std::string codec) {
  c10::cuda::CUDAGuard g(x.device());
  ok(b200seg_codec_encode(f32(x), x.numel(), f32(scale), codec_id(codec), out_bytes.data_ptr(), cur()),
     "codec_encode");
}
void codec_decode(const Tensor& in_bytes, const Tensor& scale, Tensor out, std::string codec,
                  bool accumulate) {
  c10::cuda::CUDAGuard g(out.device());
  const int id = codec_id(codec);
  if (id == 0) {
    auto src = at::from_blob(in_bytes.data_ptr(), {out.numel()}, out.options());
    if (accumulate) out.add_(src); else out.copy_(src);
    return;
  }
  ok(b200seg_codec_decode(in_bytes.data_ptr(), out.numel(), f32(scale), id, f32(out), accumulate ? 1 : 0,
                          cur()), "codec_decode");
}

void prep_input(const Tensor& img, Tensor out) {
  c10::cuda::CUDAGuard g(out.device());
  TORCH_CHECK(img.scalar_type() == at::kByte && img.is_contiguous() && out.is_contiguous(), "prep_input");
  bf16(out, "out");
  ok(b200seg_prep_input(img.data_ptr<uint8_t>(), img.numel() / img.size(-1), (int)img.size(-1), out.data_ptr(),
                        (int)out.size(-1), cur()), "prep_input");
}
void pad_rows(const Tensor& src, Tensor dst) {
  c10::cuda::CUDAGuard g(dst.device());
  bf16(src, "src"); bf16(dst, "dst");
  TORCH_CHECK(src.is_contiguous() && dst.is_contiguous() && src.numel() / src.size(-1) == dst.numel() / dst.size(-1),
              "pad_rows: contiguous tensors with the same number of rows");
  ok(b200seg_pad_rows(src.data_ptr(), (int)(src.numel() / src.size(-1)), (int)src.size(-1), dst.data_ptr(),
                      (int)dst.size(-1), cur()), "pad_rows");
}
void unpad_add_zero(Tensor src, Tensor dst) {
  c10::cuda::CUDAGuard g(dst.device());
  TORCH_CHECK(src.is_contiguous() && dst.is_contiguous() && src.scalar_type() == at::kFloat &&
              dst.scalar_type() == at::kFloat && src.numel() / src.size(-1) == dst.numel() / dst.size(-1),
              "unpad_add_zero: contiguous fp32 tensors with the same number of rows");
  ok(b200seg_unpad_add_zero(src.data_ptr<float>(), (int)(src.numel() / src.size(-1)), (int)src.size(-1),
                            dst.data_ptr<float>(), (int)dst.size(-1), cur()), "unpad_add_zero");
}
void rows_sum_add(const Tensor& part, int64_t C, Tensor out) {
  c10::cuda::CUDAGuard g(out.device());
  TORCH_CHECK(part.dim() >= 2 && part.is_contiguous() && part.scalar_type() == at::kFloat && out.is_contiguous() &&
              out.scalar_type() == at::kFloat && out.numel() >= C && part.size(-1) >= C, "rows_sum_add");
  const long long row_stride = part.numel() / part.size(0);
  ok(b200seg_rows_sum_add(part.data_ptr<float>(), (int)part.size(0), row_stride, (int)C, out.data_ptr<float>(), cur()),
     "rows_sum_add");
}

void set_error_record(const Tensor& pinned) {
  TORCH_CHECK(pinned.is_pinned() && pinned.scalar_type() == at::kInt && pinned.numel() >= 8 && pinned.is_contiguous(),
              "set_error_record: a pinned int32 tensor of >= 8 elements");
  ok(b200seg_set_error_record(reinterpret_cast<unsigned int*>(pinned.data_ptr<int>())), "set_error_record");
}

void upsample_fwd(const Tensor& x, Tensor out) {
  c10::cuda::CUDAGuard g(x.device());
  bf16(x, "x"); bf16(out, "out");
  ok(b200seg_upsample2x_fwd(x.data_ptr(), nhwc_ld(x, "x"), out.data_ptr(), nhwc_ld(out, "out"), (int)x.size(0),
                            (int)x.size(1), (int)x.size(2), (int)x.size(3), cur()), "upsample2x_fwd");
}
void upsample_bwd(const Tensor& gup, Tensor dx) {
  c10::cuda::CUDAGuard g(dx.device());
  bf16(gup, "d_up"); bf16(dx, "dx");
  ok(b200seg_upsample2x_bwd(gup.data_ptr(), nhwc_ld(gup, "d_up"), dx.data_ptr(), nhwc_ld(dx, "dx"),
                            (int)dx.size(0), (int)dx.size(1), (int)dx.size(2), (int)dx.size(3), cur()),
     "upsample2x_bwd");
}

// ---- tensor-core convolutions (csrc/conv) ---------------------------------------------------------
Tensor conv3x3_fwd(const Tensor& x, const Tensor& w, const OptT& bias, Tensor y, bool want_stats) {
  c10::cuda::CUDAGuard g(x.device());
  bf16(x, "x"); bf16(w, "w"); bf16(y, "y");
  const int N = (int)x.size(0), H = (int)x.size(1), W = (int)x.size(2);
  const int Cout = (int)w.size(0), Cin = (int)w.size(3);
  TORCH_CHECK(w.is_contiguous() && w.size(1) == 3 && w.size(2) == 3, "w must be (Cout,3,3,Cin) contiguous");
  TORCH_CHECK(x.size(3) >= Cin && y.size(3) == Cout, "channel mismatch");
  int rows = b200conv_fprop_stat_rows(N, H, W, Cin, Cout);
  Tensor part = want_stats ? at::empty({rows, 2, Cout}, x.options().dtype(at::kFloat)) : Tensor();
  ok(b200conv_fprop(x.data_ptr(), nhwc_ld(x, "x"), w.data_ptr(), bias.has_value() ? f32(*bias) : nullptr,
                    y.data_ptr(), nhwc_ld(y, "y"), N, H, W, Cin, Cout, want_stats ? part.data_ptr<float>() : nullptr,
                    cur()), "conv3x3_fwd");
  return part;
}

void conv3x3_dgrad(const Tensor& dy, const Tensor& w, Tensor dx) {
  c10::cuda::CUDAGuard g(dy.device());
  bf16(dy, "dy"); bf16(w, "w"); bf16(dx, "dx");
  const int Cout = (int)w.size(0), Cin = (int)w.size(3);
  TORCH_CHECK(dx.size(3) == Cin && dy.size(3) == Cout, "channel mismatch");
  ok(b200conv_dgrad(dy.data_ptr(), nhwc_ld(dy, "dy"), w.data_ptr(), dx.data_ptr(), nhwc_ld(dx, "dx"),
                    (int)dy.size(0), (int)dy.size(1), (int)dy.size(2), Cin, Cout, cur()), "conv3x3_dgrad");
}

Tensor conv3x3_dgrad_stats(const Tensor& dy, const Tensor& w, Tensor dx) {
  c10::cuda::CUDAGuard g(dy.device());
  bf16(dy, "dy"); bf16(w, "w"); bf16(dx, "dx");
  const int Cout = (int)w.size(0), Cin = (int)w.size(3);
  TORCH_CHECK(dx.size(3) == Cin && dy.size(3) == Cout, "channel mismatch");
  const int N = (int)dy.size(0), H = (int)dy.size(1), W = (int)dy.size(2);
  const int rows = b200conv_dgrad_stat_rows(N, H, W, Cin, Cout);
  TORCH_CHECK(rows > 0, "conv3x3_dgrad_stats: unsupported shape");
  Tensor part = at::empty({rows, 2, Cin}, dy.options().dtype(at::kFloat));
  ok(b200conv_dgrad_stats(dy.data_ptr(), nhwc_ld(dy, "dy"), w.data_ptr(), dx.data_ptr(), nhwc_ld(dx, "dx"), N, H, W,
                          Cin, Cout, part.data_ptr<float>(), cur()), "conv3x3_dgrad_stats");
  return part;
}

void conv3x3_wgrad(const Tensor& x, const Tensor& dy, Tensor dw) {
  c10::cuda::CUDAGuard g(x.device());
  bf16(x, "x"); bf16(dy, "dy");
  const int Cout = (int)dw.size(0), Cin = (int)dw.size(3);
  TORCH_CHECK(dw.is_contiguous() && dw.scalar_type() == at::kFloat, "dw must be fp32 (Cout,3,3,Cin) contiguous");
  ok(b200conv_wgrad(x.data_ptr(), nhwc_ld(x, "x"), dy.data_ptr(), nhwc_ld(dy, "dy"), dw.data_ptr<float>(),
                    (int)x.size(0), (int)x.size(1), (int)x.size(2), Cin, Cout, cur()), "conv3x3_wgrad");
}

void convT2x2_fwd(const Tensor& x, const Tensor& w, const Tensor& bias, Tensor out) {
  c10::cuda::CUDAGuard g(x.device());
  bf16(x, "x"); bf16(w, "w"); bf16(out, "out");
  const int Cin = (int)w.size(1), Cout = (int)w.size(0) / 4;
  TORCH_CHECK(w.is_contiguous() && x.size(3) == Cin && out.size(3) == Cout, "convT2x2_fwd shapes");
  ok(b200conv_convT_fwd(x.data_ptr(), nhwc_ld(x, "x"), w.data_ptr(), f32(bias), out.data_ptr(),
                        nhwc_ld(out, "out"), (int)x.size(0), (int)x.size(1), (int)x.size(2), Cin, Cout, cur()),
     "convT2x2_fwd");
}

void allreduce(at::IntArrayRef wire_ptrs, at::IntArrayRef grad_ptrs, at::IntArrayRef signal_ptrs,
               at::IntArrayRef mult_ptrs, int64_t rank, int64_t n_elems, int64_t wire_kind, double post_scale,
               int64_t epoch, OptT oneshot_out, int64_t blocks) {
  const int world = (int)wire_ptrs.size();
  TORCH_CHECK(world >= 1 && world <= 16 && (int)signal_ptrs.size() == world, "allreduce: bad pointer lists");
  const void* wp[16]; float* gp[16]; unsigned int* sp[16]; const float* mp[16];
  for (int r = 0; r < world; ++r) {
    wp[r] = reinterpret_cast<const void*>(wire_ptrs[r]);
    gp[r] = grad_ptrs.size() ? reinterpret_cast<float*>(grad_ptrs[r]) : nullptr;
    sp[r] = reinterpret_cast<unsigned int*>(signal_ptrs[r]);
    mp[r] = mult_ptrs.size() ? reinterpret_cast<const float*>(mult_ptrs[r]) : nullptr;
  }
  ok(b200seg_allreduce(wp, gp, sp, mp, (int)rank, world, n_elems, (int)wire_kind, (float)post_scale,
                       (unsigned int)epoch, oneshot_out.has_value() ? f32(*oneshot_out) : nullptr, (int)blocks,
                       cur()), "allreduce");
}

void allreduce_push(at::IntArrayRef inbox_ptrs, at::IntArrayRef grad_ptrs, at::IntArrayRef signal_ptrs,
                    at::IntArrayRef mult_ptrs, OptT scale, std::string codec, int64_t rank, int64_t n_elems,
                    double post_scale, int64_t epoch, int64_t blocks) {
  const int world = (int)inbox_ptrs.size();
  TORCH_CHECK(world >= 2 && world <= 16 && (int)grad_ptrs.size() == world && (int)signal_ptrs.size() == world &&
              (mult_ptrs.empty() || (int)mult_ptrs.size() == world), "allreduce_push: bad pointer lists");
  void* ip[16]; float* gp[16]; unsigned int* sp[16]; float* mp[16];
  for (int r = 0; r < world; ++r) {
    ip[r] = reinterpret_cast<void*>(inbox_ptrs[r]);
    gp[r] = reinterpret_cast<float*>(grad_ptrs[r]);
    sp[r] = reinterpret_cast<unsigned int*>(signal_ptrs[r]);
    mp[r] = mult_ptrs.empty() ? nullptr : reinterpret_cast<float*>(mult_ptrs[r]);
  }
  ok(b200seg_allreduce_push(ip, gp, sp, mult_ptrs.empty() ? nullptr : mp, scale.has_value() ? f32(*scale) : nullptr,
                            codec_id(codec), (int)rank, world, n_elems, (float)post_scale, (unsigned int)epoch,
                            (int)blocks, cur()), "allreduce_push");
}

void allreduce_nvls(at::IntArrayRef signal_ptrs, int64_t multicast_ptr, int64_t rank, int64_t n_elems,
                    double post_scale, int64_t epoch, int64_t blocks) {
  const int world = (int)signal_ptrs.size();
  TORCH_CHECK(world >= 2 && world <= 16, "allreduce_nvls: bad pointer list");
  unsigned int* sp[16];
  for (int r = 0; r < world; ++r) sp[r] = reinterpret_cast<unsigned int*>(signal_ptrs[r]);
  ok(b200seg_allreduce_nvls(sp, reinterpret_cast<float*>(multicast_ptr), (int)rank, world, n_elems,
                            (float)post_scale, (unsigned int)epoch, (int)blocks, cur()), "allreduce_nvls");
}

void peer_barrier(at::IntArrayRef signal_ptrs, int64_t rank, int64_t epoch) {
  const int world = (int)signal_ptrs.size();
  TORCH_CHECK(world >= 2 && world <= 16, "peer_barrier: bad pointer list");
  unsigned int* sp[16];
  for (int r = 0; r < world; ++r) sp[r] = reinterpret_cast<unsigned int*>(signal_ptrs[r]);
  ok(b200seg_barrier(sp, (int)rank, world, (unsigned int)epoch, cur()), "peer_barrier");
}

// multicast_ptr == 0 unregisters.  n_elems floats starting at buf are covered.
void set_grad_multicast(const Tensor& buf, int64_t n_elems, int64_t multicast_ptr) {
  ok(b200seg_set_grad_multicast(buf.data_ptr(), n_elems * 4, reinterpret_cast<const void*>(multicast_ptr)),
     "set_grad_multicast");
}

void write_mult(const Tensor& scale, double levels, Tensor out) {
  c10::cuda::CUDAGuard g(out.device());
  ok(b200seg_write_mult(f32(scale), (float)levels, f32(out), cur()), "write_mult");
}

void arm_next_wgrad(int64_t arm_ptr) { b200seg_arm_next_wgrad(reinterpret_cast<const void*>(arm_ptr)); }

void bucket_reduce(int64_t multicast_ptr, at::IntArrayRef grad_ptrs, at::IntArrayRef pad_ptrs, Tensor state,
                   int64_t rank, int64_t bucket, bool self_publish, bool final_bucket, at::IntArrayRef offs,
                   at::IntArrayRef counts, double post_scale, double timeout_s, int64_t blocks) {
  const int world = (int)pad_ptrs.size();
  TORCH_CHECK(world >= 2 && world <= 16 && (grad_ptrs.empty() || (int)grad_ptrs.size() == world),
              "bucket_reduce: bad pointer lists");
  TORCH_CHECK(state.is_cuda() && state.scalar_type() == at::kInt && state.numel() >= 2 && state.is_contiguous(),
              "bucket_reduce: state must be a CUDA int32[2]");
  TORCH_CHECK(offs.size() == counts.size() && offs.size() >= 1 && offs.size() <= 4, "bucket_reduce: 1..4 ranges");
  c10::cuda::CUDAGuard g(state.device());
  float* gp[16]; unsigned int* pp[16];
  for (int r = 0; r < world; ++r) {
    gp[r] = grad_ptrs.empty() ? nullptr : reinterpret_cast<float*>(grad_ptrs[r]);
    pp[r] = reinterpret_cast<unsigned int*>(pad_ptrs[r]);
  }
  long long o[4], c[4];
  for (size_t i = 0; i < offs.size(); ++i) { o[i] = offs[i]; c[i] = counts[i]; }
  ok(b200seg_bucket_reduce(reinterpret_cast<float*>(multicast_ptr), grad_ptrs.empty() ? nullptr : gp, pp,
                           reinterpret_cast<unsigned int*>(state.data_ptr<int>()), (int)rank, world, (int)bucket,
                           self_publish ? 1 : 0, final_bucket ? 1 : 0, (int)offs.size(), o, c, (float)post_scale,
                           timeout_s, (int)blocks, cur()), "bucket_reduce");
}
int64_t overlap_pad_words() { return b200seg_overlap_pad_words(); }
int64_t overlap_max_buckets() { return b200seg_overlap_max_buckets(); }
int64_t overlap_flag_index(int64_t bucket, int64_t rank) { return b200seg_overlap_flag_index((int)bucket, (int)rank); }
int64_t overlap_abort_index() { return b200seg_overlap_abort_index(); }

void set_comm_timeout(double seconds) { b200seg_set_comm_timeout(seconds); }
int64_t allreduce_abort_word() { return b200seg_allreduce_abort_word(); }
int64_t allreduce_signal_bytes() { return b200seg_allreduce_signal_bytes(); }
int64_t allreduce_max_blocks() { return b200seg_allreduce_max_blocks(); }

void convT2x2_bwd(const Tensor& x, const Tensor& d_up, const Tensor& w, Tensor dw, OptT dx) {
  c10::cuda::CUDAGuard g(x.device());
  bf16(x, "x"); bf16(d_up, "d_up"); bf16(w, "w");
  const int Cin = (int)w.size(1), Cout = (int)w.size(0) / 4;
  TORCH_CHECK(dw.is_contiguous() && dw.scalar_type() == at::kFloat && dw.size(0) == 4 * Cout && dw.size(1) == Cin,
              "dw must be fp32 (4*Cout, Cin) contiguous");
  const int N = (int)x.size(0), h = (int)x.size(1), w_ = (int)x.size(2);
  ok(b200conv_convT_wgrad(x.data_ptr(), nhwc_ld(x, "x"), d_up.data_ptr(), nhwc_ld(d_up, "d_up"), dw.data_ptr<float>(),
                          N, h, w_, Cin, Cout, cur()), "convT2x2_wgrad");
  if (dx.has_value()) {
    bf16(*dx, "dx");
    ok(b200conv_convT_dgrad(d_up.data_ptr(), nhwc_ld(d_up, "d_up"), w.data_ptr(), dx->data_ptr(), nhwc_ld(*dx, "dx"),
                            N, h, w_, Cin, Cout, cur()), "convT2x2_dgrad");
  }
}

bool conv_supported(int64_t cin, int64_t cout) { return b200conv_supported((int)cin, (int)cout) != 0; }

}  // namespace

TORCH_LIBRARY(b200seg, m) {
  m.def("stats_rows() -> int", &stats_rows);
  m.def("conv_supported(int cin, int cout) -> bool", &conv_supported);
  m.def("channel_stats(Tensor y, Tensor(a!) partial) -> ()", &channel_stats);
  m.def("bn_finalize(Tensor? partial, int count, Tensor gamma, Tensor beta, Tensor? conv_bias, Tensor(a!) running_mean, "
        "Tensor(b!) running_var, Tensor(c!) num_batches_tracked, Tensor(d!) ss, float momentum, float eps, "
        "bool training) -> ()", &bn_finalize);
  m.def("bn_relu_pool_fwd(Tensor y, Tensor ss, Tensor(a!) out, Tensor(b!)? pool) -> ()", &bn_relu_pool_fwd);
  m.def("bn_relu_pool_bwd(Tensor? d_out, Tensor? d_pool, Tensor y, Tensor ss, Tensor gamma, Tensor(a!) dgamma, "
        "Tensor(b!) dbeta, Tensor? dbias, Tensor(c!) dy) -> ()", &bn_relu_pool_bwd);
  m.def("conv_plan(int op, int N, int H, int W, int Cin, int Cout) -> int[]", &conv_plan);
  m.def("head_tail_rows(int P, int C) -> int", &head_tail_rows);
  m.def("head_tail_supported(int C, int K) -> bool", &head_tail_supported);
  m.def("head_tail_fwd_bwd(Tensor y, Tensor ss, Tensor w, Tensor b, Tensor labels, float grad_scale, Tensor(a!)? g_out, "
        "Tensor(b!)? dw, Tensor(c!)? db, Tensor(d!)? pred, Tensor(e!) loss_acc, Tensor(f!)? bn_partial) -> ()",
        &head_tail_fwd_bwd);
  m.def("bn_relu_bwd_from_partial(Tensor partial, int rows, Tensor d_out, Tensor y, Tensor ss, Tensor gamma, "
        "Tensor(a!) dgamma, Tensor(b!) dbeta, Tensor(c!) dy) -> ()", &bn_relu_bwd_from_partial);
  m.def("head_ce_fwd_bwd(Tensor a, Tensor w, Tensor b, Tensor labels, float grad_scale, Tensor(a!)? da, "
        "Tensor(b!)? dw, Tensor(c!)? db, Tensor(d!)? pred, Tensor(e!) loss_acc) -> ()", &head_ce_fwd_bwd);
  m.def("fused_adam(Tensor(a!) p, Tensor(b!) g, Tensor(c!) m, Tensor(d!) v, float lr, float beta1, float beta2, "
        "float eps, float weight_decay, float grad_scale, float bc1, float bc2, Tensor(e!)? p_lowp, "
        "bool zero_grad) -> ()", &fused_adam);
  m.def("fused_sgd(Tensor(a!) p, Tensor(b!) g, Tensor(c!)? buf, float lr, float momentum, float weight_decay, "
        "float grad_scale, Tensor(d!)? p_lowp, bool zero_grad, bool first_step) -> ()", &fused_sgd);
  m.def("amax(Tensor x, Tensor(a!) out) -> ()", &amax);
  m.def("codec_encode(Tensor x, Tensor scale, Tensor(a!) out_bytes, str codec) -> ()", &codec_encode);
  m.def("codec_decode(Tensor in_bytes, Tensor scale, Tensor(a!) out, str codec, bool accumulate) -> ()",
        &codec_decode);
  m.def("prep_input(Tensor img, Tensor(a!) out) -> ()", &prep_input);
  m.def("set_error_record(Tensor pinned) -> ()", &set_error_record);
  m.def("pad_rows(Tensor src, Tensor(a!) dst) -> ()", &pad_rows);
  m.def("unpad_add_zero(Tensor(a!) src, Tensor(b!) dst) -> ()", &unpad_add_zero);
  m.def("rows_sum_add(Tensor part, int C, Tensor(a!) out) -> ()", &rows_sum_add);
  m.def("upsample_bilinear2x_fwd(Tensor x, Tensor(a!) out) -> ()", &upsample_fwd);
  m.def("upsample_bilinear2x_bwd(Tensor d_up, Tensor(a!) dx) -> ()", &upsample_bwd);
  m.def("allreduce(int[] wire_ptrs, int[] grad_ptrs, int[] signal_ptrs, int[] mult_ptrs, int rank, int n_elems, "
        "int wire_kind, float post_scale, int epoch, Tensor(a!)? oneshot_out, int blocks) -> ()", &allreduce);
  m.def("write_mult(Tensor scale, float levels, Tensor(a!) out) -> ()", &write_mult);
  m.def("allreduce_push(int[] inbox_ptrs, int[] grad_ptrs, int[] signal_ptrs, int[] mult_ptrs, Tensor? scale, str codec, "
        "int rank, int n_elems, float post_scale, int epoch, int blocks) -> ()", &allreduce_push);
  m.def("allreduce_nvls(int[] signal_ptrs, int multicast_ptr, int rank, int n_elems, float post_scale, int epoch, "
        "int blocks) -> ()", &allreduce_nvls);
  m.def("peer_barrier(int[] signal_ptrs, int rank, int epoch) -> ()", &peer_barrier);
  m.def("set_grad_multicast(Tensor buf, int n_elems, int multicast_ptr) -> ()", &set_grad_multicast);
  m.def("arm_next_wgrad(int arm_ptr) -> ()", &arm_next_wgrad);
  m.def("bucket_reduce(int multicast_ptr, int[] grad_ptrs, int[] pad_ptrs, Tensor(a!) state, int rank, int bucket, "
        "bool self_publish, bool final_bucket, int[] offs, int[] counts, float post_scale, float timeout_s, "
        "int blocks) -> ()", &bucket_reduce);
  m.def("overlap_pad_words() -> int", &overlap_pad_words);
  m.def("overlap_max_buckets() -> int", &overlap_max_buckets);
  m.def("overlap_flag_index(int bucket, int rank) -> int", &overlap_flag_index);
  m.def("overlap_abort_index() -> int", &overlap_abort_index);
  m.def("set_comm_timeout(float seconds) -> ()", &set_comm_timeout);
  m.def("allreduce_abort_word() -> int", &allreduce_abort_word);
  m.def("allreduce_signal_bytes() -> int", &allreduce_signal_bytes);
  m.def("allreduce_max_blocks() -> int", &allreduce_max_blocks);
  m.def("conv3x3_fwd(Tensor x, Tensor w, Tensor? bias, Tensor(a!) y, bool want_stats) -> Tensor", &conv3x3_fwd);
  m.def("conv3x3_dgrad(Tensor dy, Tensor w, Tensor(a!) dx) -> ()", &conv3x3_dgrad);
  m.def("conv3x3_dgrad_stats(Tensor dy, Tensor w, Tensor(a!) dx) -> Tensor", &conv3x3_dgrad_stats);
  m.def("conv3x3_wgrad(Tensor x, Tensor dy, Tensor(a!) dw) -> ()", &conv3x3_wgrad);
  m.def("convT2x2_fwd(Tensor x, Tensor w, Tensor bias, Tensor(a!) out) -> ()", &convT2x2_fwd);
  m.def("convT2x2_bwd(Tensor x, Tensor d_up, Tensor w, Tensor(a!) dw, Tensor(b!)? dx) -> ()", &convT2x2_bwd);
}
