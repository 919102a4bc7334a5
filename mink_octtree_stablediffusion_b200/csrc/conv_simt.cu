// Generic (any dtype, any channel count) sparse convolution kernels on the
// neighbour table: output-stationary forward / data-gradient and split-row
// weight-gradient.  SIMT shared-memory tiled GEMMs with fp32 (fp64 for double)
// accumulation.  These serve fp32/fp64 (gradcheck, config 1) and every shape the
// tcgen05 path (conv_tc.cu) does not take.
//
// Replaces matmul<T,uint,TILE> / matmul2<T,uint,TILE> behind
// ConvolutionForwardGPU / ConvolutionBackwardGPU (pybind/extern.hpp:84-114).
// Unlike the reference there is no atomicAdd scatter in forward/dgrad: a CTA
// owns its output rows and loops over the kernel offsets.
#include <algorithm>

#include "common.cuh"

namespace mk {

constexpr int BM = 64;   // output rows per CTA
constexpr int BN = 64;   // output channels per CTA
constexpr int BK = 16;   // reduction chunk
constexpr int kThreads = 256;

// y[o, n] = sum_k sum_c x[nbr[kk(k), o], c] * w(k, c, n)
//   w(k,c,n) = transpose_w ? W[k][n][c] : W[k][c][n]
template <typename T>
__global__ void __launch_bounds__(kThreads)
k_conv_fwd_simt(const T *__restrict__ x, const T *__restrict__ w,
                int transpose_w, int kflip, const int32_t *__restrict__ nbr,
                int64_t ld, int64_t n_out, int K, int c_in, int c_out,
                T *__restrict__ y) {
  using A = typename Acc<T>::type;
  __shared__ A As[BK][BM + 4];
  __shared__ A Bs[BK][BN + 4];
  __shared__ int32_t rows[BM];
  __shared__ int any_valid;

  const int tid = threadIdx.x;
  const int64_t o0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int tx = tid & 15;  // -> 4 output channels
  const int ty = tid >> 4;  // -> 4 output rows

  A acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = (A)0;

  for (int k = 0; k < K; ++k) {
    const int kk = kflip ? K - 1 - k : k;
    __syncthreads();  // previous iteration done with rows/As/Bs
    if (tid == 0) any_valid = 0;
    __syncthreads();
    if (tid < BM) {
      const int64_t o = o0 + tid;
      const int32_t r = o < n_out ? nbr[(int64_t)kk * ld + o] : -1;
      rows[tid] = r;
      if (r >= 0) any_valid = 1;
    }
    __syncthreads();
    if (!any_valid) continue;
    const T *wk = w + (size_t)k * c_in * c_out;
    for (int c0 = 0; c0 < c_in; c0 += BK) {
      // A tile: BM rows x BK channels (stored transposed)
      for (int e = tid; e < BM * BK; e += kThreads) {
        const int m = e / BK, c = e % BK;
        const int32_t r = rows[m];
        A v = (A)0;
        if (r >= 0 && c0 + c < c_in) v = to_acc<T>(x[(size_t)r * c_in + c0 + c]);
        As[c][m] = v;
      }
      // B tile: BK channels x BN outputs
      for (int e = tid; e < BK * BN; e += kThreads) {
        int c, n;
        if (transpose_w) { c = e % BK; n = e / BK; } else { c = e / BN; n = e % BN; }
        A v = (A)0;
        if (c0 + c < c_in && n0 + n < c_out)
          v = to_acc<T>(transpose_w ? wk[(size_t)(n0 + n) * c_in + c0 + c]
                                    : wk[(size_t)(c0 + c) * c_out + n0 + n]);
        Bs[c][n] = v;
      }
      __syncthreads();
#pragma unroll
      for (int c = 0; c < BK; ++c) {
        A a[4], b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = As[c][ty * 4 + i];
#pragma unroll
        for (int j = 0; j < 4; ++j) b[j] = Bs[c][tx * 4 + j];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] += a[i] * b[j];
      }
      __syncthreads();
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t o = o0 + ty * 4 + i;
    if (o >= n_out) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n < c_out) y[(size_t)o * c_out + n] = from_acc<T, A>(acc[i][j]);
    }
  }
}

// dW[k][ci][co] += sum_{o in chunk} x[nbr[k,o]][ci] * dy[o][co]
template <typename T>
__global__ void __launch_bounds__(kThreads)
k_conv_wgrad_simt(const T *__restrict__ x, const T *__restrict__ dy,
                  const int32_t *__restrict__ nbr, int64_t ld, int64_t n_out,
                  int K, int c_in, int c_out, int64_t rows_per_split,
                  int tiles_n, typename Acc<T>::type *__restrict__ dw_acc) {
  using A = typename Acc<T>::type;
  __shared__ A As[BK][BM + 4];  // [row in chunk][ci]
  __shared__ A Bs[BK][BN + 4];  // [row in chunk][co]
  __shared__ int32_t rows[BK];
  __shared__ int any_valid;
  const int tid = threadIdx.x;
  const int k = blockIdx.z;
  const int tile = blockIdx.y;
  const int m0 = (tile / tiles_n) * BM;  // ci base
  const int n0 = (tile % tiles_n) * BN;  // co base
  const int64_t o_begin = (int64_t)blockIdx.x * rows_per_split;
  const int64_t o_end = (o_begin + rows_per_split < n_out) ? o_begin + rows_per_split : n_out;
  const int tx = tid & 15, ty = tid >> 4;

  A acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = (A)0;

  for (int64_t ob = o_begin; ob < o_end; ob += BK) {
    __syncthreads();
    if (tid == 0) any_valid = 0;
    __syncthreads();
    if (tid < BK) {
      const int64_t o = ob + tid;
      const int32_t r = o < o_end ? nbr[(int64_t)k * ld + o] : -1;
      rows[tid] = r;
      if (r >= 0) any_valid = 1;
    }
    __syncthreads();
    if (!any_valid) continue;
    for (int e = tid; e < BK * BM; e += kThreads) {
      const int rr = e / BM, c = e % BM;
      const int32_t r = rows[rr];
      A v = (A)0;
      if (r >= 0 && m0 + c < c_in) v = to_acc<T>(x[(size_t)r * c_in + m0 + c]);
      As[rr][c] = v;
    }
    for (int e = tid; e < BK * BN; e += kThreads) {
      const int rr = e / BN, c = e % BN;
      A v = (A)0;
      if (rows[rr] >= 0 && n0 + c < c_out)
        v = to_acc<T>(dy[(size_t)(ob + rr) * c_out + n0 + c]);
      Bs[rr][c] = v;
    }
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < BK; ++rr) {
      A a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[rr][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[rr][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] += a[i] * b[j];
    }
  }
  A *dwk = dw_acc + (size_t)k * c_in * c_out;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int ci = m0 + ty * 4 + i;
    if (ci >= c_in) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = n0 + tx * 4 + j;
      if (co < c_out && acc[i][j] != (A)0)
        atomicAdd(&dwk[(size_t)ci * c_out + co], acc[i][j]);
    }
  }
}

template <typename T, typename A>
__global__ void k_convert(const A *__restrict__ src, T *__restrict__ dst,
                          int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = from_acc<T, A>(src[i]);
}

template <typename T>
int conv_fwd_simt(const void *x, const void *w, int transpose_w, int kflip,
                  const int32_t *nbr, int64_t ld, int64_t n_out, int K,
                  int c_in, int c_out, void *y, cudaStream_t st) {
  dim3 grid((unsigned)ceil_div(n_out, BM), (unsigned)ceil_div(c_out, BN));
  k_conv_fwd_simt<T><<<grid, kThreads, 0, st>>>(
      (const T *)x, (const T *)w, transpose_w, kflip, nbr, ld, n_out, K, c_in,
      c_out, (T *)y);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

template <typename T>
int conv_wgrad_simt(const void *x, const void *dy, const int32_t *nbr,
                    int64_t ld, int64_t n_out, int K, int c_in, int c_out,
                    void *dw, void *workspace, cudaStream_t st, int dw_acc) {
  using A = typename Acc<T>::type;
  const size_t n_w = (size_t)K * c_in * c_out;
  // dw_acc: dw is of the accumulator type (fp32 for 16-bit features): no conversion pass
  const bool direct = std::is_same<T, A>::value || dw_acc;
  A *acc = direct ? (A *)dw : (A *)workspace;
  MK_CUDA(cudaMemsetAsync(acc, 0, n_w * sizeof(A), st));
  if (n_out > 0) {
    const int tiles_m = (int)ceil_div(c_in, BM), tiles_n = (int)ceil_div(c_out, BN);
    const int64_t ctas_per_split = (int64_t)K * tiles_m * tiles_n;
    int64_t splits = std::max<int64_t>(1, (int64_t)num_sms() * 8 / ctas_per_split);
    splits = std::min<int64_t>(splits, ceil_div(n_out, 256));
    int64_t rows_per_split = ceil_div(n_out, splits);
    rows_per_split = ceil_div(rows_per_split, BK) * BK;
    splits = ceil_div(n_out, rows_per_split);
    dim3 grid((unsigned)splits, (unsigned)(tiles_m * tiles_n), (unsigned)K);
    k_conv_wgrad_simt<T><<<grid, kThreads, 0, st>>>(
        (const T *)x, (const T *)dy, nbr, ld, n_out, K, c_in, c_out,
        rows_per_split, tiles_n, acc);
    MK_LAUNCH_CHECK();
  }
  if (!direct) {
    k_convert<T, A><<<(unsigned)ceil_div((int64_t)n_w, 256), 256, 0, st>>>(
        acc, (T *)dw, (int64_t)n_w);
    MK_LAUNCH_CHECK();
  }
  return MK_OK;
}

int conv_fwd_simt_dispatch(int dtype, const void *x, const void *w,
                           int transpose_w, int kflip, const int32_t *nbr,
                           int64_t ld, int64_t n_out, int K, int c_in,
                           int c_out, void *y, cudaStream_t st) {
  MK_DISPATCH_DTYPE(dtype, return conv_fwd_simt<T>(x, w, transpose_w, kflip,
                                                   nbr, ld, n_out, K, c_in,
                                                   c_out, y, st));
  return MK_OK;
}

int conv_wgrad_simt_dispatch(int dtype, const void *x, const void *dy,
                             const int32_t *nbr, int64_t ld, int64_t n_out,
                             int K, int c_in, int c_out, void *dw,
                             void *workspace, cudaStream_t st, int dw_acc) {
  MK_DISPATCH_DTYPE(dtype, return conv_wgrad_simt<T>(x, dy, nbr, ld, n_out, K,
                                                     c_in, c_out, dw,
                                                     workspace, st, dw_acc));
  return MK_OK;
}

}  // namespace mk
