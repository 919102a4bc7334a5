// fp32 convolution on the tensor cores: bf16 operand splitting.
//
// The reference engine computes in fp32 (MinkowskiConvolution.py:265,285 FloatTensor
// kernels; cuBLAS SGEMM / its own SIMT GEMM) and the example networks run
// in fp32 unless cast.  tcgen05 has no fp32 MMA and kind::tf32 keeps 10 mantissa bits
// (1e-3, two orders above north_star's 1e-5), so an fp32 operand is split into bf16
// pieces, x = x0 + x1 + x2 with x0 = bf16(x), x1 = bf16(x - x0), x2 = bf16(x - x0 - x1)
// (each difference is exact in fp32), and the product is expanded:
//
//   x w  =  x0 w0 + (x0 w1 + x1 w0) + (x0 w2 + x1 w1 + x2 w0) + O(2^-24)
//
// Six bf16 MMAs per fp32 MMA (terms = 6, the default: every product of pieces whose
// exponents sum to less than 2^-24 relative is dropped, nothing else) or three
// (terms = 3: error 2^-16 per product, for callers that accept it: MK_F32_TERMS=3).
// All pieces accumulate in ONE fp32 TMEM accumulator, so the sum is formed at fp32.
//
// The expansion is laid out as extra reduction channels, so the bf16 kernels of
// conv_tc.cu / conv_wgrad_tc.cu run unchanged (their fp32 accumulators are stored as
// they are):
//   forward / dgrad:  A' = [x0|x0|x0|x1|x1|x2]   B' rows = [w0;w1;w2;w0;w1;w0]
//   weight gradient:  X' = [x0|x1|x2], DY' = [dy0|dy1|dy2]; the kernel produces the
//                     3x3 block matrix X'^T DY', k_fold_dw adds the blocks a + b <= 2.
// Channel counts are padded to multiples of 16 with zeros, which also serves the
// narrow layers of the example networks (Cin or Cout in {1, 3, 4}).
#include <algorithm>
#include <cstdlib>

#include "common.cuh"

namespace mk {
// conv_tc.cu / conv_wgrad_tc.cu
bool conv_fwd_tc_eligible(int dtype, int64_t n_out, int K, int c_in, int c_out);
void conv_fwd_tc_set_kper_max(int kper_max);
int64_t conv_fwd_tc_workspace(int64_t n_out, int K, int c_in, int c_out);
int conv_fwd_tc(int dtype, const void *x, const void *w, int transpose_w, int kflip, const int32_t *nbr,
                const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                void *y, const int32_t *out_rows, void *workspace, cudaStream_t st, int out_f32);
bool conv_wgrad_tc_eligible(int dtype, int64_t n_out, int K, int c_in, int c_out);
int64_t conv_wgrad_tc_workspace(int64_t n_out, int K, int c_in, int c_out);
int conv_wgrad_tc(int dtype, const void *x, const void *dy, const int32_t *dy_rows, const int32_t *nbr,
                  const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                  void *dw, void *workspace, cudaStream_t st, int dw_f32);
int64_t conv_wgrad_tc_rows_per_cta(int64_t n_out, int K, int c_in, int c_out);
void conv_wgrad_tc_set_rows_max(int64_t rows_max);

namespace {

inline int pad16(int c) { return (c + 15) & ~15; }
inline size_t al256(size_t b) { return (b + 255) & ~(size_t)255; }

// piece index of a (forward) reduction block t: A side / B side
struct Terms {
  int n;            // reduction blocks of the forward / dgrad expansion (6 or 3)
  int a[6], b[6];
  int nw;           // pieces per operand of the weight-gradient expansion (3 or 2)
};

inline Terms terms() {
  const char *e = getenv("MK_F32_TERMS");
  if (e && atoi(e) == 3) return Terms{3, {0, 0, 1, 0, 0, 0}, {0, 1, 0, 0, 0, 0}, 2};
  return Terms{6, {0, 0, 0, 1, 1, 2}, {0, 1, 2, 0, 1, 0}, 3};
}

__device__ __forceinline__ void split3(float v, __nv_bfloat16 (&p)[3]) {
  p[0] = __float2bfloat16_rn(v);
  const float r1 = v - __bfloat162float(p[0]);
  p[1] = __float2bfloat16_rn(r1);
  p[2] = __float2bfloat16_rn(r1 - __bfloat162float(p[1]));
}

struct PieceMap { int piece[6]; };

// out[r, t * cp + c] = piece[t] of x[r, c]  (zero for c >= c);  8 channels (16 B) per thread
__global__ void k_split_rows(const float *__restrict__ x, int64_t n, int c, int cp, int nblk, PieceMap pm,
                             __nv_bfloat16 *__restrict__ out) {
  const int groups = cp >> 3;
  const int64_t total = n * groups;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = e / groups;
    const int g = (int)(e - r * groups);
    __nv_bfloat16 pc[3][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int ch = g * 8 + j;
      __nv_bfloat16 p3[3];
      split3(ch < c ? __ldg(x + r * c + ch) : 0.f, p3);
      pc[0][j] = p3[0]; pc[1][j] = p3[1]; pc[2][j] = p3[2];
    }
    __nv_bfloat16 *orow = out + r * (int64_t)nblk * cp + g * 8;
    for (int t = 0; t < nblk; ++t)
      *reinterpret_cast<uint4 *>(orow + (int64_t)t * cp) = *reinterpret_cast<const uint4 *>(pc[pm.piece[t]]);
  }
}

// Weight pieces for the forward / dgrad expansion.
//   transpose == 0: w [K, c_red, c_n] -> out [K, nblk * crp, cnp],  out[k][t*crp + c][n] = piece[t](w[k][c][n])
//   transpose == 1: w [K, c_n, c_red] -> out [K, cnp, nblk * crp],  out[k][n][t*crp + c] = piece[t](w[k][n][c])
__global__ void k_split_w(const float *__restrict__ w, int transpose, int K, int c_red, int c_n, int crp, int cnp,
                          int nblk, PieceMap pm, __nv_bfloat16 *__restrict__ out) {
  const int64_t total = (int64_t)K * crp * cnp;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = e;
    int c, n;
    if (!transpose) { n = (int)(r % cnp); r /= cnp; c = (int)(r % crp); r /= crp; }
    else            { c = (int)(r % crp); r /= crp; n = (int)(r % cnp); r /= cnp; }
    const int k = (int)r;
    float v = 0.f;
    if (c < c_red && n < c_n) v = transpose ? w[((size_t)k * c_n + n) * c_red + c] : w[((size_t)k * c_red + c) * c_n + n];
    __nv_bfloat16 p3[3];
    split3(v, p3);
    for (int t = 0; t < nblk; ++t) {
      const size_t o = transpose ? ((size_t)k * cnp + n) * ((size_t)nblk * crp) + (size_t)t * crp + c
                                 : ((size_t)k * nblk * crp + (size_t)t * crp + c) * cnp + n;
      out[o] = p3[pm.piece[t]];
    }
  }
}

// y [n, c] = ypad [n, cp][:, :c]
__global__ void k_unpad_cols(const float *__restrict__ yp, int64_t n, int c, int cp, float *__restrict__ y) {
  const int64_t total = n * c;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = e / c;
    y[e] = yp[r * cp + (e - r * c)];
  }
}

// dw [K, c_in, c_out] = sum over blocks (a, b), a + b < nw, of dw9 [K, nw*cip, nw*cop][a*cip + i][b*cop + j]
__global__ void k_fold_dw(const float *__restrict__ dw9, int K, int c_in, int c_out, int cip, int cop, int nw,
                          float *__restrict__ dw) {
  const int64_t total = (int64_t)K * c_in * c_out;
  const size_t ldw = (size_t)nw * cop;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = e;
    const int j = (int)(r % c_out); r /= c_out;
    const int i = (int)(r % c_in);
    const int k = (int)(r / c_in);
    const float *base = dw9 + (size_t)k * nw * cip * ldw;
    // smallest terms first
    float acc = 0.f;
    for (int s = nw - 1; s >= 0; --s)
      for (int a = 0; a <= s; ++a) acc += base[((size_t)a * cip + i) * ldw + (size_t)(s - a) * cop + j];
    dw[e] = acc;
  }
}

inline int grid_for(int64_t total, int threads) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div(total, threads), (int64_t)num_sms() * 16));
}

}  // namespace

// MK_F32_TC: 0 = never, 1 (default) = where the accumulation chain is short enough for
// north_star's 1e-5, 2 = always ("fast fp32", ~1e-4 on 512-channel layers).
//
// Why a limit: tcgen05 adds every MMA's 16 products into the fp32 TMEM accumulator with
// truncation (round toward zero, as on earlier tensor-core generations), so an accumulator
// drifts toward zero by ~3e-8 of its magnitude per MMA in the chain.  Measured here
// (tests/test_gpu_conv.py, 3^3 kernel): 648 chained MMAs 7e-6, 972 MMAs 1.07e-5 of the
// output's maximum.  An SIMT FFMA accumulation rounds to nearest and has no such drift, so
// beyond kChainLimit MMAs per accumulator fp32 stays on the SIMT kernels.
static int f32_mode() {
  const char *e = getenv("MK_F32_TC");
  return e ? atoi(e) : 1;
}
static int64_t chain_limit() {
  const char *e = getenv("MK_F32_TC_CHAIN");
  return e ? atoll(e) : 400;
}

// Offsets one accumulator may see under the chain limit (0 = no limit: "fast fp32"; -1 = not even one).
// Beyond it the tensor-core kernel splits the offsets over grid.z and the partial outputs are added in
// fp32 with round to nearest (conv_tc.cu, ksplit).
static int kper_max_for(int c_in) {
  const Terms t = terms();
  if (f32_mode() >= 2) return 0;
  const int64_t per_offset = (int64_t)(pad16(c_in) / 16) * t.n;      // MMAs one offset adds to the chain
  const int64_t k = chain_limit() / per_offset;
  return k >= 1 ? (int)std::min<int64_t>(k, 1 << 20) : -1;
}
struct KperGuard {
  explicit KperGuard(int v) { conv_fwd_tc_set_kper_max(v); }
  ~KperGuard() { conv_fwd_tc_set_kper_max(0); }
};

// the weight gradient's chain runs over the rows a CTA reduces (16 per MMA): bound them the same way
struct RowsGuard {
  RowsGuard() { conv_wgrad_tc_set_rows_max(f32_mode() >= 2 ? 0 : chain_limit() * 16); }
  ~RowsGuard() { conv_wgrad_tc_set_rows_max(0); }
};

bool conv_fwd_f32_eligible(int64_t n_out, int K, int c_in, int c_out) {
  const Terms t = terms();
  const int mode = f32_mode();
  if (mode <= 0) return false;
  const int kmax = kper_max_for(c_in);
  if (kmax < 0) return false;
  KperGuard g(kmax);
  return conv_fwd_tc_eligible(MK_BF16, n_out, K, t.n * pad16(c_in), pad16(c_out));
}

bool conv_wgrad_f32_eligible(int64_t n_out, int K, int c_in, int c_out) {
  const Terms t = terms();
  const int mode = f32_mode();
  if (mode <= 0) return false;
  const int cip = t.nw * pad16(c_in), cop = t.nw * pad16(c_out);
  RowsGuard g;
  if (!conv_wgrad_tc_eligible(MK_BF16, n_out, K, cip, cop)) return false;
  // the reduction runs over the rows a CTA walks: 16 rows per MMA
  const int64_t chain = conv_wgrad_tc_rows_per_cta(n_out, K, cip, cop) / 16;
  return mode >= 2 || chain <= chain_limit();
}

int64_t conv_fwd_f32_workspace(int64_t n_out, int64_t n_in, int K, int c_in, int c_out) {
  const Terms t = terms();
  const int cip = pad16(c_in), cop = pad16(c_out);
  KperGuard g(std::max(0, kper_max_for(c_in)));
  return (int64_t)(al256((size_t)n_in * t.n * cip * 2) + al256((size_t)K * t.n * cip * cop * 2) +
                   (cop != c_out ? al256((size_t)n_out * cop * 4) : 0) +
                   (size_t)conv_fwd_tc_workspace(n_out, K, t.n * cip, cop) + 256);
}

int64_t conv_wgrad_f32_workspace(int64_t n_out, int64_t n_in, int K, int c_in, int c_out) {
  const Terms t = terms();
  const int cip = pad16(c_in), cop = pad16(c_out);
  RowsGuard g;
  return (int64_t)(al256((size_t)n_in * t.nw * cip * 2) + al256((size_t)n_out * t.nw * cop * 2) +
                   al256((size_t)K * t.nw * cip * t.nw * cop * 4) +
                   (size_t)conv_wgrad_tc_workspace(n_out, K, t.nw * cip, t.nw * cop) + 256);
}

// ---------------------------------------------------------------------------
// 16-bit features with channel counts that are not multiples of 16 (the 1 -> 32, 512 -> 4,
// 4 -> 4, 4 -> 512 layers of the octree VAE, examples/ae_res.py:973; the 3 -> 32 stem of
// MinkUNet): operands are zero-padded to 16 channels in the workspace and run on the
// tensor-core kernels; the padded output columns are dropped.
// ---------------------------------------------------------------------------
namespace {
template <typename T>
__global__ void k_pad_rows(const T *__restrict__ x, int64_t n, int c, int cp, T *__restrict__ out) {
  const int64_t total = n * cp;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = e / cp;
    const int ch = (int)(e - r * cp);
    out[e] = ch < c ? x[r * c + ch] : T(0.f);
  }
}

// w [K, a, b] -> out [K, ap, bp] (zero padded)
template <typename T>
__global__ void k_pad_w(const T *__restrict__ w, int K, int a, int b, int ap, int bp, T *__restrict__ out) {
  const int64_t total = (int64_t)K * ap * bp;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = e;
    const int j = (int)(r % bp); r /= bp;
    const int i = (int)(r % ap);
    const int k = (int)(r / ap);
    out[e] = (i < a && j < b) ? w[((size_t)k * a + i) * b + j] : T(0.f);
  }
}

// y [.., a, b] = yp [.., ap, bp][:a, :b]   (rows = K * a for weights, n for features with a = 1 ... handled by caller)
template <typename T>
__global__ void k_unpad(const T *__restrict__ yp, int64_t outer, int a, int b, int ap, int bp, T *__restrict__ y) {
  const int64_t total = outer * a * b;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = e;
    const int j = (int)(r % b); r /= b;
    const int i = (int)(r % a);
    const int64_t o = r / a;
    y[e] = yp[((size_t)o * ap + i) * bp + j];
  }
}
}  // namespace

bool conv_pad16_needed(int c_in, int c_out) { return (c_in % 16) != 0 || (c_out % 16) != 0; }

bool conv_fwd_pad16_eligible(int dtype, int64_t n_out, int K, int c_in, int c_out) {
  return (dtype == MK_BF16 || dtype == MK_F16) && conv_pad16_needed(c_in, c_out) &&
         conv_fwd_tc_eligible(dtype, n_out, K, pad16(c_in), pad16(c_out));
}

bool conv_wgrad_pad16_eligible(int dtype, int64_t n_out, int K, int c_in, int c_out) {
  return (dtype == MK_BF16 || dtype == MK_F16) && conv_pad16_needed(c_in, c_out) &&
         conv_wgrad_tc_eligible(dtype, n_out, K, pad16(c_in), pad16(c_out));
}

int64_t conv_fwd_pad16_workspace(int64_t n_out, int64_t n_in, int K, int c_in, int c_out) {
  const int cip = pad16(c_in), cop = pad16(c_out);
  return (int64_t)((cip != c_in ? al256((size_t)n_in * cip * 2) : 0) + al256((size_t)K * cip * cop * 2) +
                   (cop != c_out ? al256((size_t)n_out * cop * 2) : 0) +
                   (size_t)conv_fwd_tc_workspace(n_out, K, cip, cop) + 256);
}

int64_t conv_wgrad_pad16_workspace(int64_t n_out, int64_t n_in, int K, int c_in, int c_out) {
  const int cip = pad16(c_in), cop = pad16(c_out);
  return (int64_t)((cip != c_in ? al256((size_t)n_in * cip * 2) : 0) + (cop != c_out ? al256((size_t)n_out * cop * 2) : 0) +
                   al256((size_t)K * cip * cop * 4) + (size_t)conv_wgrad_tc_workspace(n_out, K, cip, cop) + 256);
}

template <typename T>
static int conv_fwd_pad16_t(int dtype, const T *x, const T *w, int transpose_w, int kflip, const int32_t *nbr,
                            const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in,
                            int c_out, T *y, const int32_t *out_rows, void *workspace, cudaStream_t st) {
  const int cip = pad16(c_in), cop = pad16(c_out);
  uint8_t *ws = (uint8_t *)workspace;
  const T *xp = x;
  if (cip != c_in) {
    T *buf = (T *)ws; ws += al256((size_t)n_in * cip * 2);
    k_pad_rows<T><<<grid_for(n_in * cip, 256), 256, 0, st>>>(x, n_in, c_in, cip, buf);
    MK_LAUNCH_CHECK();
    xp = buf;
  }
  T *wp = (T *)ws; ws += al256((size_t)K * cip * cop * 2);
  if (!transpose_w) k_pad_w<T><<<grid_for((int64_t)K * cip * cop, 256), 256, 0, st>>>(w, K, c_in, c_out, cip, cop, wp);
  else              k_pad_w<T><<<grid_for((int64_t)K * cip * cop, 256), 256, 0, st>>>(w, K, c_out, c_in, cop, cip, wp);
  MK_LAUNCH_CHECK();
  T *yp = y;
  if (cop != c_out) { yp = (T *)ws; ws += al256((size_t)n_out * cop * 2); }
  int rc = conv_fwd_tc(dtype, xp, wp, transpose_w, kflip, nbr, tile_mask, ld, n_out, n_in, K, cip, cop, yp, out_rows, ws,
                       st, 0);
  if (rc) return rc;
  if (yp != y) {
    k_unpad<T><<<grid_for(n_out * c_out, 256), 256, 0, st>>>(yp, n_out, 1, c_out, 1, cop, y);
    MK_LAUNCH_CHECK();
  }
  return MK_OK;
}

int conv_fwd_pad16(int dtype, const void *x, const void *w, int transpose_w, int kflip, const int32_t *nbr,
                   const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                   void *y, const int32_t *out_rows, void *workspace, cudaStream_t st) {
  if (dtype == MK_BF16)
    return conv_fwd_pad16_t<__nv_bfloat16>(dtype, (const __nv_bfloat16 *)x, (const __nv_bfloat16 *)w, transpose_w, kflip,
                                           nbr, tile_mask, ld, n_out, n_in, K, c_in, c_out, (__nv_bfloat16 *)y, out_rows,
                                           workspace, st);
  return conv_fwd_pad16_t<__half>(dtype, (const __half *)x, (const __half *)w, transpose_w, kflip, nbr, tile_mask, ld,
                                  n_out, n_in, K, c_in, c_out, (__half *)y, out_rows, workspace, st);
}

template <typename T>
static int conv_wgrad_pad16_t(int dtype, const T *x, const T *dy, const int32_t *dy_rows, const int32_t *nbr,
                              const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in,
                              int c_out, void *dw, int dw_f32, void *workspace, cudaStream_t st) {
  const int cip = pad16(c_in), cop = pad16(c_out);
  uint8_t *ws = (uint8_t *)workspace;
  const T *xp = x, *dyp = dy;
  if (cip != c_in) {
    T *buf = (T *)ws; ws += al256((size_t)n_in * cip * 2);
    k_pad_rows<T><<<grid_for(n_in * cip, 256), 256, 0, st>>>(x, n_in, c_in, cip, buf);
    MK_LAUNCH_CHECK();
    xp = buf;
  }
  if (cop != c_out) {
    T *buf = (T *)ws; ws += al256((size_t)n_out * cop * 2);
    k_pad_rows<T><<<grid_for(n_out * cop, 256), 256, 0, st>>>(dy, n_out, c_out, cop, buf);
    MK_LAUNCH_CHECK();
    dyp = buf;
  }
  float *dwp = (float *)ws; ws += al256((size_t)K * cip * cop * 4);
  int rc = conv_wgrad_tc(dtype, xp, dyp, dy_rows, nbr, tile_mask, ld, n_out, n_in, K, cip, cop, dwp, ws, st, dw_f32);
  if (rc) return rc;
  const int64_t total = (int64_t)K * c_in * c_out;
  if (dw_f32) k_unpad<float><<<grid_for(total, 256), 256, 0, st>>>(dwp, K, c_in, c_out, cip, cop, (float *)dw);
  else        k_unpad<T><<<grid_for(total, 256), 256, 0, st>>>((const T *)dwp, K, c_in, c_out, cip, cop, (T *)dw);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int conv_wgrad_pad16(int dtype, const void *x, const void *dy, const int32_t *dy_rows, const int32_t *nbr,
                     const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                     void *dw, int dw_f32, void *workspace, cudaStream_t st) {
  if (dtype == MK_BF16)
    return conv_wgrad_pad16_t<__nv_bfloat16>(dtype, (const __nv_bfloat16 *)x, (const __nv_bfloat16 *)dy, dy_rows, nbr,
                                             tile_mask, ld, n_out, n_in, K, c_in, c_out, dw, dw_f32, workspace, st);
  return conv_wgrad_pad16_t<__half>(dtype, (const __half *)x, (const __half *)dy, dy_rows, nbr, tile_mask, ld, n_out,
                                    n_in, K, c_in, c_out, dw, dw_f32, workspace, st);
}

// y (fp32) = conv(x (fp32), w (fp32)) through the bf16 tensor-core kernel
int conv_fwd_f32(const float *x, const float *w, int transpose_w, int kflip, const int32_t *nbr,
                 const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                 float *y, const int32_t *out_rows, void *workspace, cudaStream_t st) {
  const Terms t = terms();
  const int cip = pad16(c_in), cop = pad16(c_out);
  uint8_t *ws = (uint8_t *)workspace;
  __nv_bfloat16 *a = (__nv_bfloat16 *)ws;           ws += al256((size_t)n_in * t.n * cip * 2);
  __nv_bfloat16 *b = (__nv_bfloat16 *)ws;           ws += al256((size_t)K * t.n * cip * cop * 2);
  float *ypad = y;
  if (cop != c_out) { ypad = (float *)ws;           ws += al256((size_t)n_out * cop * 4); }
  PieceMap pa, pb;
  for (int i = 0; i < 6; ++i) { pa.piece[i] = t.a[i]; pb.piece[i] = t.b[i]; }
  k_split_rows<<<grid_for(n_in * (cip >> 3), 256), 256, 0, st>>>(x, n_in, c_in, cip, t.n, pa, a);
  MK_LAUNCH_CHECK();
  // W: [K, c_red = c_in, c_n = c_out] (transpose_w == 0) or [K, c_n = c_out, c_red = c_in]
  k_split_w<<<grid_for((int64_t)K * cip * cop, 256), 256, 0, st>>>(w, transpose_w, K, c_in, c_out, cip, cop, t.n, pb, b);
  MK_LAUNCH_CHECK();
  int rc;
  {
    KperGuard g(std::max(0, kper_max_for(c_in)));
    rc = conv_fwd_tc(MK_BF16, a, b, transpose_w, kflip, nbr, tile_mask, ld, n_out, n_in, K, t.n * cip, cop, ypad,
                     out_rows, ws, st, 1);
  }
  if (rc) return rc;
  if (ypad != y) {
    k_unpad_cols<<<grid_for(n_out * c_out, 256), 256, 0, st>>>(ypad, n_out, c_out, cop, y);
    MK_LAUNCH_CHECK();
  }
  return MK_OK;
}

// dw (fp32) [K, c_in, c_out] = sum_o x[nbr[k,o]]^T dy[o]
int conv_wgrad_f32(const float *x, const float *dy, const int32_t *dy_rows, const int32_t *nbr,
                   const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                   float *dw, void *workspace, cudaStream_t st) {
  const Terms t = terms();
  const int cip = pad16(c_in), cop = pad16(c_out);
  uint8_t *ws = (uint8_t *)workspace;
  __nv_bfloat16 *xs = (__nv_bfloat16 *)ws;          ws += al256((size_t)n_in * t.nw * cip * 2);
  __nv_bfloat16 *dys = (__nv_bfloat16 *)ws;         ws += al256((size_t)n_out * t.nw * cop * 2);
  float *dw9 = (float *)ws;                         ws += al256((size_t)K * t.nw * cip * t.nw * cop * 4);
  PieceMap id;
  for (int i = 0; i < 6; ++i) id.piece[i] = i < 3 ? i : 0;
  k_split_rows<<<grid_for(n_in * (cip >> 3), 256), 256, 0, st>>>(x, n_in, c_in, cip, t.nw, id, xs);
  MK_LAUNCH_CHECK();
  k_split_rows<<<grid_for(n_out * (cop >> 3), 256), 256, 0, st>>>(dy, n_out, c_out, cop, t.nw, id, dys);
  MK_LAUNCH_CHECK();
  RowsGuard g;
  int rc = conv_wgrad_tc(MK_BF16, xs, dys, dy_rows, nbr, tile_mask, ld, n_out, n_in, K, t.nw * cip, t.nw * cop, dw9, ws,
                         st, 1);
  if (rc) return rc;
  k_fold_dw<<<grid_for((int64_t)K * c_in * c_out, 256), 256, 0, st>>>(dw9, K, c_in, c_out, cip, cop, t.nw, dw);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

}  // namespace mk
