// C-ABI entry points of the convolution group: pick the tcgen05/TMEM kernel
// when the shape qualifies, otherwise the SIMT tiled kernel.  Both are CUDA;
// there is no CPU path.
#include "common.cuh"

namespace mk {
int conv_fwd_simt_dispatch(int dtype, const void *x, const void *w,
                           int transpose_w, int kflip, const int32_t *nbr,
                           int64_t ld, int64_t n_out, int K, int c_in,
                           int c_out, void *y, cudaStream_t st);
int conv_wgrad_simt_dispatch(int dtype, const void *x, const void *dy,
                             const int32_t *nbr, int64_t ld, int64_t n_out,
                             int K, int c_in, int c_out, void *dw,
                             void *workspace, cudaStream_t st, int dw_acc);
// conv_tc.cu
bool conv_fwd_tc_eligible(int dtype, int64_t n_out, int K, int c_in, int c_out);
int64_t conv_fwd_tc_workspace(int64_t n_out, int K, int c_in, int c_out);
int conv_fwd_tc(int dtype, const void *x, const void *w, int transpose_w,
                int kflip, const int32_t *nbr, const uint32_t *tile_mask,
                int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in,
                int c_out, void *y, const int32_t *out_rows, void *workspace,
                cudaStream_t st, int out_f32);
// conv_wgrad_tc.cu
bool conv_wgrad_tc_eligible(int dtype, int64_t n_out, int K, int c_in, int c_out);
int64_t conv_wgrad_tc_workspace(int64_t n_out, int K, int c_in, int c_out);
int conv_wgrad_tc(int dtype, const void *x, const void *dy,
                  const int32_t *dy_rows, const int32_t *nbr,
                  const uint32_t *tile_mask, int64_t ld, int64_t n_out,
                  int64_t n_in, int K, int c_in, int c_out, void *dw,
                  void *workspace, cudaStream_t st, int dw_f32);
// conv_f32.cu: fp32 on the tensor cores through bf16 operand splitting
bool conv_fwd_f32_eligible(int64_t n_out, int K, int c_in, int c_out);
bool conv_wgrad_f32_eligible(int64_t n_out, int K, int c_in, int c_out);
int64_t conv_fwd_f32_workspace(int64_t n_out, int64_t n_in, int K, int c_in, int c_out);
int64_t conv_wgrad_f32_workspace(int64_t n_out, int64_t n_in, int K, int c_in, int c_out);
int conv_fwd_f32(const float *x, const float *w, int transpose_w, int kflip, const int32_t *nbr,
                 const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                 float *y, const int32_t *out_rows, void *workspace, cudaStream_t st);
int conv_wgrad_f32(const float *x, const float *dy, const int32_t *dy_rows, const int32_t *nbr,
                   const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                   float *dw, void *workspace, cudaStream_t st);

// ... and zero-padding of narrow 16-bit layers to 16 channels
bool conv_fwd_pad16_eligible(int dtype, int64_t n_out, int K, int c_in, int c_out);
bool conv_wgrad_pad16_eligible(int dtype, int64_t n_out, int K, int c_in, int c_out);
int64_t conv_fwd_pad16_workspace(int64_t n_out, int64_t n_in, int K, int c_in, int c_out);
int64_t conv_wgrad_pad16_workspace(int64_t n_out, int64_t n_in, int K, int c_in, int c_out);
int conv_fwd_pad16(int dtype, const void *x, const void *w, int transpose_w, int kflip, const int32_t *nbr,
                   const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                   void *y, const int32_t *out_rows, void *workspace, cudaStream_t st);
int conv_wgrad_pad16(int dtype, const void *x, const void *dy, const int32_t *dy_rows, const int32_t *nbr,
                     const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                     void *dw, int dw_f32, void *workspace, cudaStream_t st);

// conv_simt2.cu: tuned fp32 FFMA kernels (c_in, c_out multiples of 4)
bool conv_f32_simt2_eligible(int c_in, int c_out, int K);
int64_t conv_f32_simt2_workspace(int K, int c_in, int c_out);
int conv_fwd_f32_simt2(const float *x, const float *w, int transpose_w, int kflip, const int32_t *nbr,
                       const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                       float *y, const int32_t *out_rows, void *workspace, cudaStream_t st);
int conv_wgrad_f32_simt2(const float *x, const float *dy, const int32_t *dy_rows, const int32_t *nbr,
                         const uint32_t *tile_mask, int64_t ld, int64_t n_out, int K, int c_in, int c_out, float *dw,
                         cudaStream_t st);

static thread_local int g_last_path = 0;
static int g_force_simt = 0;
}  // namespace mk

using namespace mk;

extern "C" {

int mk_conv_last_path(void) { return g_last_path; }

int mk_conv_force_simt(int on) {
  const int prev = g_force_simt;
  g_force_simt = on ? 1 : 0;
  return prev;
}

int64_t mk_conv_workspace_bytes(int dtype, int64_t n_rows, int K, int c_in,
                                int c_out) {
  int64_t b = 256;
  // SIMT wgrad: fp32 accumulation buffer for 16-bit dtypes
  if (dtype == MK_BF16 || dtype == MK_F16) b += (int64_t)K * c_in * c_out * 4;
  b = std::max<int64_t>(b, conv_fwd_tc_workspace(n_rows, K, c_in, c_out));
  b = std::max<int64_t>(b, conv_wgrad_tc_workspace(n_rows, K, c_in, c_out));
  return b;
}

static bool tc_layout_ok(const void *x, const void *other, const int32_t *nbr,
                         int64_t ld) {
  return ld % 128 == 0 && ((uintptr_t)x & 15) == 0 &&
         ((uintptr_t)other & 15) == 0 && ((uintptr_t)nbr & 15) == 0;
}

static int check_conv_args(int64_t ld, int64_t n_out, int K, int c_in,
                           int c_out) {
  MK_CHECK_ARG(K >= 1, "kernel volume must be >= 1");
  MK_CHECK_ARG(c_in >= 1 && c_out >= 1, "channel counts must be >= 1");
  MK_CHECK_ARG(n_out >= 0 && ld >= n_out, "ld must be >= n_out");
  return MK_OK;
}

int mk_conv_tc_eligible(int dtype, int64_t n_rows, int K, int c_in,
                        int c_out) {
  if (g_force_simt || n_rows <= 0) return 0;
  if (dtype == MK_F32)
    return (conv_fwd_f32_eligible(n_rows, K, c_in, c_out) && conv_wgrad_f32_eligible(n_rows, K, c_in, c_out)) ? 1 : 0;
  if ((c_in % 16) || (c_out % 16))
    return (conv_fwd_pad16_eligible(dtype, n_rows, K, c_in, c_out) &&
            conv_wgrad_pad16_eligible(dtype, n_rows, K, c_in, c_out)) ? 1 : 0;
  return (conv_fwd_tc_eligible(dtype, n_rows, K, c_in, c_out) &&
          conv_wgrad_tc_eligible(dtype, n_rows, K, c_in, c_out))
             ? 1
             : 0;
}

int mk_conv_rows_eligible(int dtype, int64_t n_rows, int K, int c_in, int c_out) {
  if (mk_conv_tc_eligible(dtype, n_rows, K, c_in, c_out)) return 1;
  return (!g_force_simt && n_rows > 0 && dtype == MK_F32 && conv_f32_simt2_eligible(c_in, c_out, K)) ? 1 : 0;
}

int64_t mk_conv_workspace_bytes2(int dtype, int64_t n_out, int64_t n_in, int K, int c_in, int c_out) {
  int64_t b = mk_conv_workspace_bytes(dtype, n_out, K, c_in, c_out);
  if (dtype == MK_F32 && conv_f32_simt2_eligible(c_in, c_out, K))
    b = std::max<int64_t>(b, conv_f32_simt2_workspace(K, c_in, c_out));
  if (dtype == MK_F32 && n_out > 0 && n_in > 0) {
    if (conv_fwd_f32_eligible(n_out, K, c_in, c_out))
      b = std::max<int64_t>(b, conv_fwd_f32_workspace(n_out, n_in, K, c_in, c_out));
    if (conv_wgrad_f32_eligible(n_out, K, c_in, c_out))
      b = std::max<int64_t>(b, conv_wgrad_f32_workspace(n_out, n_in, K, c_in, c_out));
  }
  if ((dtype == MK_BF16 || dtype == MK_F16) && n_out > 0 && n_in > 0) {
    if (conv_fwd_pad16_eligible(dtype, n_out, K, c_in, c_out))
      b = std::max<int64_t>(b, conv_fwd_pad16_workspace(n_out, n_in, K, c_in, c_out));
    if (conv_wgrad_pad16_eligible(dtype, n_out, K, c_in, c_out))
      b = std::max<int64_t>(b, conv_wgrad_pad16_workspace(n_out, n_in, K, c_in, c_out));
  }
  return b;
}

int mk_conv_fwd(int dtype, const void *x, const void *w, int transpose_w,
                int kflip, const int32_t *nbr, const uint32_t *tile_mask,
                int64_t ld, int64_t n_out, int64_t n_in,
                int K, int c_in, int c_out, void *y, void *workspace,
                void *stream) {
  return mk_conv_fwd_ex(dtype, x, w, transpose_w, kflip, nbr, tile_mask, ld,
                        n_out, n_in, K, c_in, c_out, y, nullptr, workspace, -1,
                        stream);
}

int mk_conv_fwd_rows(int dtype, const void *x, const void *w, int transpose_w,
                     int kflip, const int32_t *nbr, const uint32_t *tile_mask,
                     int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in,
                     int c_out, void *y, const int32_t *out_rows,
                     void *workspace, void *stream) {
  return mk_conv_fwd_ex(dtype, x, w, transpose_w, kflip, nbr, tile_mask, ld,
                        n_out, n_in, K, c_in, c_out, y, out_rows, workspace, -1,
                        stream);
}

int mk_conv_fwd_ex(int dtype, const void *x, const void *w, int transpose_w,
                   int kflip, const int32_t *nbr, const uint32_t *tile_mask,
                   int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in,
                   int c_out, void *y, const int32_t *out_rows,
                   void *workspace, int64_t workspace_bytes, void *stream) {
  int rc = check_conv_args(ld, n_out, K, c_in, c_out);
  if (rc) return rc;
  MK_CHECK_ARG(n_in >= 0, "n_in must be >= 0");
  if (n_out == 0) return MK_OK;
  MK_CHECK_ARG(w && nbr && y && (x || n_in == 0), "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const bool tc_ok = !g_force_simt && workspace && n_in > 0 && tc_layout_ok(x, y, nbr, ld);
  if (tc_ok && dtype != MK_F32 && conv_fwd_tc_eligible(dtype, n_out, K, c_in, c_out)) {
    g_last_path = 1;
    return conv_fwd_tc(dtype, x, w, transpose_w, kflip, nbr, tile_mask, ld,
                       n_out, n_in, K, c_in, c_out, y, out_rows, workspace, st, 0);
  }
  if (tc_ok && conv_fwd_pad16_eligible(dtype, n_out, K, c_in, c_out) &&
      workspace_bytes >= conv_fwd_pad16_workspace(n_out, n_in, K, c_in, c_out)) {
    g_last_path = 1;
    return conv_fwd_pad16(dtype, x, w, transpose_w, kflip, nbr, tile_mask, ld, n_out, n_in, K, c_in, c_out, y, out_rows,
                          workspace, st);
  }
  // fp32: bf16 operand splitting on the tensor cores.  Needs the workspace size (it depends
  // on n_in), so only the _ex entry point can take it.
  if (tc_ok && dtype == MK_F32 && ((uintptr_t)w & 3) == 0 && conv_fwd_f32_eligible(n_out, K, c_in, c_out) &&
      workspace_bytes >= conv_fwd_f32_workspace(n_out, n_in, K, c_in, c_out)) {
    g_last_path = 1;
    return conv_fwd_f32((const float *)x, (const float *)w, transpose_w, kflip, nbr, tile_mask, ld, n_out, n_in, K,
                        c_in, c_out, (float *)y, out_rows, workspace, st);
  }
  // fp32 beyond the split path's accumulation limit: the tuned FFMA kernel (also honours out_rows)
  if (!g_force_simt && dtype == MK_F32 && n_in > 0 && conv_f32_simt2_eligible(c_in, c_out, K) &&
      (((uintptr_t)x | (uintptr_t)y | (uintptr_t)w) & 15) == 0 &&
      (!transpose_w || (workspace && workspace_bytes >= conv_f32_simt2_workspace(K, c_in, c_out)))) {
    g_last_path = 0;
    return conv_fwd_f32_simt2((const float *)x, (const float *)w, transpose_w, kflip, nbr, tile_mask, ld, n_out, n_in, K,
                              c_in, c_out, (float *)y, out_rows, workspace, st);
  }
  if (out_rows) {
    set_error("mk_conv_fwd_rows: an output-row permutation is not served for this dtype / shape "
              "(ask mk_conv_rows_eligible first)");
    return MK_EUNSUPPORTED;
  }
  g_last_path = 0;
  return conv_fwd_simt_dispatch(dtype, x, w, transpose_w, kflip, nbr, ld,
                                n_out, K, c_in, c_out, y, st);
}

int mk_conv_wgrad_ex(int dtype, const void *x, const void *dy,
                     const int32_t *dy_rows, const int32_t *nbr,
                     const uint32_t *tile_mask, int64_t ld, int64_t n_out,
                     int64_t n_in, int K, int c_in, int c_out, void *dw,
                     int dw_f32, void *workspace, int64_t workspace_bytes,
                     void *stream) {
  int rc = check_conv_args(ld, n_out, K, c_in, c_out);
  if (rc) return rc;
  MK_CHECK_ARG(n_in >= 0, "n_in must be >= 0");
  MK_CHECK_ARG(dw, "null pointer");
  MK_CHECK_ARG(n_out == 0 || (x && dy && nbr), "null pointer");
  MK_CHECK_ARG(workspace || dtype == MK_F32 || dtype == MK_F64 || dw_f32,
               "workspace required for 16-bit dtypes");
  MK_CHECK_ARG(!dw_f32 || dtype != MK_F64, "dw_f32 is for 16-bit and fp32 features");
  if (dtype == MK_F32) dw_f32 = 0;   // already fp32
  cudaStream_t st = (cudaStream_t)stream;
  const bool tc_ok = !g_force_simt && workspace && n_in > 0 && n_out > 0 && tc_layout_ok(x, dy, nbr, ld) &&
                     (dy_rows == nullptr || ((uintptr_t)dy_rows & 15) == 0);
  if (tc_ok && dtype != MK_F32 && conv_wgrad_tc_eligible(dtype, n_out, K, c_in, c_out)) {
    g_last_path = 1;
    return conv_wgrad_tc(dtype, x, dy, dy_rows, nbr, tile_mask, ld, n_out,
                         n_in, K, c_in, c_out, dw, workspace, st, dw_f32);
  }
  if (tc_ok && conv_wgrad_pad16_eligible(dtype, n_out, K, c_in, c_out) &&
      workspace_bytes >= conv_wgrad_pad16_workspace(n_out, n_in, K, c_in, c_out)) {
    g_last_path = 1;
    return conv_wgrad_pad16(dtype, x, dy, dy_rows, nbr, tile_mask, ld, n_out, n_in, K, c_in, c_out, dw, dw_f32,
                            workspace, st);
  }
  if (tc_ok && dtype == MK_F32 && conv_wgrad_f32_eligible(n_out, K, c_in, c_out) &&
      workspace_bytes >= conv_wgrad_f32_workspace(n_out, n_in, K, c_in, c_out)) {
    g_last_path = 1;
    return conv_wgrad_f32((const float *)x, (const float *)dy, dy_rows, nbr, tile_mask, ld, n_out, n_in, K, c_in,
                          c_out, (float *)dw, workspace, st);
  }
  if (!g_force_simt && dtype == MK_F32 && n_in > 0 && n_out > 0 && conv_f32_simt2_eligible(c_in, c_out, K) &&
      (((uintptr_t)x | (uintptr_t)dy | (uintptr_t)dw) & 15) == 0) {
    g_last_path = 0;
    return conv_wgrad_f32_simt2((const float *)x, (const float *)dy, dy_rows, nbr, tile_mask, ld, n_out, K, c_in, c_out,
                                (float *)dw, st);
  }
  if (dy_rows) {
    set_error("mk_conv_wgrad_rows: dy_rows is not served for this dtype / shape "
              "(ask mk_conv_rows_eligible first)");
    return MK_EUNSUPPORTED;
  }
  g_last_path = 0;
  return conv_wgrad_simt_dispatch(dtype, x, dy, nbr, ld, n_out, K, c_in,
                                  c_out, dw, workspace, st, dw_f32);
}

int mk_conv_wgrad_rows(int dtype, const void *x, const void *dy,
                       const int32_t *dy_rows, const int32_t *nbr,
                       const uint32_t *tile_mask, int64_t ld, int64_t n_out,
                       int64_t n_in, int K, int c_in, int c_out, void *dw,
                       void *workspace, void *stream) {
  return mk_conv_wgrad_ex(dtype, x, dy, dy_rows, nbr, tile_mask, ld, n_out,
                          n_in, K, c_in, c_out, dw, 0, workspace, -1, stream);
}

int mk_conv_wgrad(int dtype, const void *x, const void *dy,
                  const int32_t *nbr, const uint32_t *tile_mask, int64_t ld,
                  int64_t n_out, int64_t n_in, int K,
                  int c_in, int c_out, void *dw, void *workspace,
                  void *stream) {
  return mk_conv_wgrad_ex(dtype, x, dy, nullptr, nbr, tile_mask, ld, n_out,
                          n_in, K, c_in, c_out, dw, 0, workspace, -1, stream);
}

}  // extern "C"
