// Pooling and broadcast kernels on the neighbour table / batch segments.
//
// Replaces LocalPooling{,Transpose}{Forward,Backward}GPU,
// GlobalPooling{Forward,Backward}GPU and Broadcast{Forward,Backward}GPU
// (pybind/extern.hpp:215-392), which in the reference are cuSPARSE SpMM +
// layout transposes.  Here: one thread per (row, channel), rows
// of the table read as a broadcast, features coalesced along the channel axis;
// backward passes walk the inverse table, so nothing needs atomics except the
// per-batch segment sums (fp32 atomics on a [B, C] buffer, one flush per CTA).
#include <algorithm>

#include "common.cuh"

namespace mk {

template <typename T>
__global__ void k_pool_fwd(int mode, const T *__restrict__ x,
                           const int32_t *__restrict__ nbr, int64_t ld,
                           int64_t n_out, int K, int c, T *__restrict__ y,
                           T *__restrict__ nnz, int32_t *__restrict__ argmax) {
  using A = typename Acc<T>::type;
  const int64_t total = n_out * c;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t o = t / c;
    const int ch = (int)(t - o * c);
    A acc = (A)0;
    int cnt = 0;
    int32_t best = -1;
    for (int k = 0; k < K; ++k) {
      const int32_t r = __ldg(nbr + (int64_t)k * ld + o);
      if (r < 0) continue;
      const A v = to_acc<T>(x[(size_t)r * c + ch]);
      if (mode == MK_POOL_MAX) {
        if (best < 0 || v > acc) { acc = v; best = r; }
      } else {
        acc += v;
      }
      ++cnt;
    }
    if (mode == MK_POOL_AVG && cnt > 0) acc /= (A)cnt;
    y[t] = from_acc<T, A>(acc);
    if (argmax && mode == MK_POOL_MAX) argmax[t] = best;
    if (nnz && ch == 0) nnz[o] = from_acc<T, A>((A)cnt);
  }
}

template <typename T>
__global__ void k_pool_bwd(int mode, const T *__restrict__ dy,
                           const T *__restrict__ nnz,
                           const int32_t *__restrict__ inv, int64_t ld_in,
                           int64_t n_in, int K, int c,
                           const int32_t *__restrict__ argmax,
                           T *__restrict__ dx) {
  using A = typename Acc<T>::type;
  const int64_t total = n_in * c;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = t / c;
    const int ch = (int)(t - i * c);
    A acc = (A)0;
    for (int k = 0; k < K; ++k) {
      const int32_t o = __ldg(inv + (int64_t)k * ld_in + i);
      if (o < 0) continue;
      if (mode == MK_POOL_MAX) {
        if (argmax[(size_t)o * c + ch] != (int32_t)i) continue;
        acc += to_acc<T>(dy[(size_t)o * c + ch]);
      } else if (mode == MK_POOL_AVG) {
        acc += to_acc<T>(dy[(size_t)o * c + ch]) / to_acc<T>(nnz[o]);
      } else {
        acc += to_acc<T>(dy[(size_t)o * c + ch]);
      }
    }
    dx[t] = from_acc<T, A>(acc);
  }
}

// ---- batch-segment sums ------------------------------------------------------
// SRC: 0 -> x[r,ch]; 1 -> x[r,ch]*z[r,ch]
constexpr int kSegRows = 256;

template <typename T, int SRC>
__global__ void k_segment_sum(const T *__restrict__ x, const T *__restrict__ z,
                              const int32_t *__restrict__ batch_of_row,
                              int64_t n, int c, int64_t n_batch,
                              typename Acc<T>::type *__restrict__ acc_out,
                              typename Acc<T>::type *__restrict__ cnt_out) {
  using A = typename Acc<T>::type;
  const int ch = blockIdx.y * blockDim.x + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.x * kSegRows;
  const int64_t r1 = (r0 + kSegRows < n) ? r0 + kSegRows : n;
  A acc = (A)0;
  int cnt = 0;
  int32_t cur = -1;
  for (int64_t r = r0; r < r1; ++r) {
    const int32_t b = __ldg(batch_of_row + r);
    if (b != cur) {
      if (cur >= 0 && cur < n_batch) {
        if (ch < c) atomicAdd(&acc_out[(size_t)cur * c + ch], acc);
        if (cnt_out && ch == 0) atomicAdd(&cnt_out[cur], (A)cnt);
      }
      acc = (A)0; cnt = 0; cur = b;
    }
    if (ch < c) {
      A v = to_acc<T>(x[(size_t)r * c + ch]);
      if (SRC == 1) v *= to_acc<T>(z[(size_t)r * c + ch]);
      acc += v;
    }
    ++cnt;
  }
  if (cur >= 0 && cur < n_batch) {
    if (ch < c) atomicAdd(&acc_out[(size_t)cur * c + ch], acc);
    if (cnt_out && ch == 0) atomicAdd(&cnt_out[cur], (A)cnt);
  }
}

template <typename T>
__global__ void k_global_finalize(int mode, const typename Acc<T>::type *acc,
                                  const typename Acc<T>::type *cnt, int c,
                                  int64_t n_batch, T *__restrict__ y,
                                  T *__restrict__ cnt_out) {
  using A = typename Acc<T>::type;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_batch * c) return;
  const int64_t b = t / c;
  A v = acc[t];
  if (mode == MK_POOL_AVG && cnt[b] > (A)0) v /= cnt[b];
  y[t] = from_acc<T, A>(v);
  if (cnt_out && t % c == 0) cnt_out[b] = from_acc<T, A>(cnt[b]);
}

// max: one CTA per batch index (rare path; O(n) scan of batch_of_row per CTA)
template <typename T>
__global__ void k_global_max(const T *__restrict__ x,
                             const int32_t *__restrict__ batch_of_row,
                             int64_t n, int c, T *__restrict__ y,
                             T *__restrict__ cnt_out,
                             int32_t *__restrict__ argmax) {
  using A = typename Acc<T>::type;
  const int b = blockIdx.x;
  for (int ch = threadIdx.x; ch < c; ch += blockDim.x) {
    A best = (A)0;
    int32_t arg = -1;
    int cnt = 0;
    for (int64_t r = 0; r < n; ++r) {
      if (__ldg(batch_of_row + r) != b) continue;
      const A v = to_acc<T>(x[(size_t)r * c + ch]);
      if (arg < 0 || v > best) { best = v; arg = (int32_t)r; }
      ++cnt;
    }
    y[(size_t)b * c + ch] = from_acc<T, A>(best);
    if (argmax) argmax[(size_t)b * c + ch] = arg;
    if (cnt_out && ch == 0) cnt_out[b] = from_acc<T, A>((A)cnt);
  }
}

template <typename T>
__global__ void k_global_bwd(int mode, const T *__restrict__ dy,
                             const T *__restrict__ cnt,
                             const int32_t *__restrict__ batch_of_row,
                             int64_t n, int c,
                             const int32_t *__restrict__ argmax,
                             T *__restrict__ dx) {
  using A = typename Acc<T>::type;
  const int64_t total = n * c;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = t / c;
    const int ch = (int)(t - r * c);
    const int32_t b = __ldg(batch_of_row + r);
    if (b < 0) { dx[t] = from_acc<T, A>((A)0); continue; }   // row whose batch index has no origin row
    A v = to_acc<T>(dy[(size_t)b * c + ch]);
    if (mode == MK_POOL_AVG) v /= to_acc<T>(cnt[b]);
    if (mode == MK_POOL_MAX && argmax[(size_t)b * c + ch] != (int32_t)r) v = (A)0;
    dx[t] = from_acc<T, A>(v);
  }
}

template <typename T>
__global__ void k_broadcast_fwd(int mode, const T *__restrict__ x,
                                const T *__restrict__ g,
                                const int32_t *__restrict__ batch_of_row,
                                int64_t n, int c, T *__restrict__ y) {
  using A = typename Acc<T>::type;
  const int64_t total = n * c;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = t / c;
    const int ch = (int)(t - r * c);
    const int32_t b = __ldg(batch_of_row + r);
    // a row whose batch index has no origin row takes the operation's neutral element
    const A gv = b >= 0 ? to_acc<T>(g[(size_t)b * c + ch]) : (mode == MK_BCAST_ADD ? (A)0 : (A)1);
    const A xv = to_acc<T>(x[t]);
    y[t] = from_acc<T, A>(mode == MK_BCAST_ADD ? xv + gv : xv * gv);
  }
}

template <typename T>
__global__ void k_broadcast_bwd_dx(int mode, const T *__restrict__ g,
                                   const T *__restrict__ dy,
                                   const int32_t *__restrict__ batch_of_row,
                                   int64_t n, int c, T *__restrict__ dx) {
  using A = typename Acc<T>::type;
  const int64_t total = n * c;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = t / c;
    const int ch = (int)(t - r * c);
    A v = to_acc<T>(dy[t]);
    if (mode == MK_BCAST_MUL) {
      const int32_t b = __ldg(batch_of_row + r);
      if (b >= 0) v *= to_acc<T>(g[(size_t)b * c + ch]);
    }
    dx[t] = from_acc<T, A>(v);
  }
}

static inline unsigned ew_blocks(int64_t total) {
  return (unsigned)std::min<int64_t>(ceil_div(total, 256), (int64_t)num_sms() * 16);
}

template <typename T>
int segment_sum_launch(const T *x, const T *z, const int32_t *batch_of_row,
                       int64_t n, int c, int64_t n_batch,
                       typename Acc<T>::type *acc, typename Acc<T>::type *cnt,
                       cudaStream_t st) {
  const int threads = c >= 256 ? 256 : (c >= 128 ? 128 : (c >= 64 ? 64 : 32));
  dim3 grid((unsigned)ceil_div(n, kSegRows), (unsigned)ceil_div(c, threads));
  if (z)
    k_segment_sum<T, 1><<<grid, threads, 0, st>>>(x, z, batch_of_row, n, c,
                                                  n_batch, acc, cnt);
  else
    k_segment_sum<T, 0><<<grid, threads, 0, st>>>(x, z, batch_of_row, n, c,
                                                  n_batch, acc, cnt);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

}  // namespace mk

using namespace mk;

extern "C" {

int mk_pool_fwd(int dtype, int mode, const void *x, const int32_t *nbr,
                int64_t ld, int64_t n_out, int K, int c, void *y, void *nnz,
                int32_t *argmax, void *stream) {
  MK_CHECK_ARG(mode >= MK_POOL_SUM && mode <= MK_POOL_MAX, "bad pooling mode");
  MK_CHECK_ARG(K >= 1 && c >= 1 && ld >= n_out, "bad shape");
  if (n_out == 0) return MK_OK;
  MK_CHECK_ARG(x && nbr && y, "null pointer");
  MK_CHECK_ARG(mode != MK_POOL_MAX || argmax, "max pooling needs argmax");
  cudaStream_t st = (cudaStream_t)stream;
  MK_DISPATCH_DTYPE(dtype, k_pool_fwd<T><<<ew_blocks(n_out * c), 256, 0, st>>>(
      mode, (const T *)x, nbr, ld, n_out, K, c, (T *)y, (T *)nnz, argmax));
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int mk_pool_bwd(int dtype, int mode, const void *dy, const void *nnz,
                const int32_t *inv, int64_t ld_in, int64_t n_in, int K, int c,
                const int32_t *argmax, int64_t n_out, void *dx, void *stream) {
  (void)n_out;
  MK_CHECK_ARG(mode >= MK_POOL_SUM && mode <= MK_POOL_MAX, "bad pooling mode");
  MK_CHECK_ARG(K >= 1 && c >= 1 && ld_in >= n_in, "bad shape");
  if (n_in == 0) return MK_OK;
  MK_CHECK_ARG(dy && inv && dx, "null pointer");
  MK_CHECK_ARG(mode != MK_POOL_AVG || nnz, "avg pooling needs nnz");
  MK_CHECK_ARG(mode != MK_POOL_MAX || argmax, "max pooling needs argmax");
  cudaStream_t st = (cudaStream_t)stream;
  MK_DISPATCH_DTYPE(dtype, k_pool_bwd<T><<<ew_blocks(n_in * c), 256, 0, st>>>(
      mode, (const T *)dy, (const T *)nnz, inv, ld_in, n_in, K, c, argmax,
      (T *)dx));
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int64_t mk_segment_workspace_bytes(int dtype, int c, int64_t n_batch) {
  const int64_t a = dtype == MK_F64 ? 8 : 4;
  return (n_batch * c + n_batch) * a + 256;
}

int mk_global_pool_fwd(int dtype, int mode, const void *x,
                       const int32_t *batch_of_row, int64_t n, int c,
                       int64_t n_batch, void *y, void *cnt, int32_t *argmax,
                       void *workspace, void *stream) {
  MK_CHECK_ARG(mode >= MK_POOL_SUM && mode <= MK_POOL_MAX, "bad pooling mode");
  MK_CHECK_ARG(c >= 1 && n_batch >= 0, "bad shape");
  if (n_batch == 0) return MK_OK;
  MK_CHECK_ARG(y && (n == 0 || (x && batch_of_row)), "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == MK_POOL_MAX) {
    MK_DISPATCH_DTYPE(dtype, k_global_max<T><<<(unsigned)n_batch, 128, 0, st>>>(
        (const T *)x, batch_of_row, n, c, (T *)y, (T *)cnt, argmax));
    MK_LAUNCH_CHECK();
    return MK_OK;
  }
  MK_CHECK_ARG(workspace, "workspace required");
  MK_DISPATCH_DTYPE(dtype, {
    using A = typename Acc<T>::type;
    A *acc = (A *)workspace;
    A *cn = acc + (size_t)n_batch * c;
    MK_CUDA(cudaMemsetAsync(acc, 0, sizeof(A) * (size_t)(n_batch * c + n_batch), st));
    if (n > 0) {
      int rc = segment_sum_launch<T>((const T *)x, nullptr, batch_of_row, n, c,
                                     n_batch, acc, cn, st);
      if (rc) return rc;
    }
    k_global_finalize<T><<<(unsigned)ceil_div(n_batch * c, 256), 256, 0, st>>>(
        mode, acc, cn, c, n_batch, (T *)y, (T *)cnt);
  });
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int mk_global_pool_bwd(int dtype, int mode, const void *dy, const void *cnt,
                       const int32_t *batch_of_row, int64_t n, int c,
                       int64_t n_batch, const int32_t *argmax, void *dx,
                       void *stream) {
  (void)n_batch;
  MK_CHECK_ARG(mode >= MK_POOL_SUM && mode <= MK_POOL_MAX, "bad pooling mode");
  if (n == 0) return MK_OK;
  MK_CHECK_ARG(dy && batch_of_row && dx, "null pointer");
  MK_CHECK_ARG(mode != MK_POOL_AVG || cnt, "avg pooling needs counts");
  MK_CHECK_ARG(mode != MK_POOL_MAX || argmax, "max pooling needs argmax");
  cudaStream_t st = (cudaStream_t)stream;
  MK_DISPATCH_DTYPE(dtype, k_global_bwd<T><<<ew_blocks(n * c), 256, 0, st>>>(
      mode, (const T *)dy, (const T *)cnt, batch_of_row, n, c, argmax, (T *)dx));
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int mk_broadcast_fwd(int dtype, int mode, const void *x, const void *g,
                     const int32_t *batch_of_row, int64_t n, int c, void *y,
                     void *stream) {
  MK_CHECK_ARG(mode == MK_BCAST_ADD || mode == MK_BCAST_MUL, "bad broadcast mode");
  if (n == 0) return MK_OK;
  MK_CHECK_ARG(x && g && batch_of_row && y, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  MK_DISPATCH_DTYPE(dtype, k_broadcast_fwd<T><<<ew_blocks(n * c), 256, 0, st>>>(
      mode, (const T *)x, (const T *)g, batch_of_row, n, c, (T *)y));
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int mk_broadcast_bwd(int dtype, int mode, const void *x, const void *g,
                     const void *dy, const int32_t *batch_of_row, int64_t n,
                     int c, int64_t n_batch, void *dx, void *dg,
                     void *workspace, void *stream) {
  MK_CHECK_ARG(mode == MK_BCAST_ADD || mode == MK_BCAST_MUL, "bad broadcast mode");
  MK_CHECK_ARG(workspace || n_batch == 0, "workspace required");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_batch == 0) return MK_OK;
  MK_DISPATCH_DTYPE(dtype, {
    using A = typename Acc<T>::type;
    A *acc = (A *)workspace;
    A *cn = acc + (size_t)n_batch * c;
    MK_CUDA(cudaMemsetAsync(acc, 0, sizeof(A) * (size_t)(n_batch * c + n_batch), st));
    if (n > 0) {
      k_broadcast_bwd_dx<T><<<ew_blocks(n * c), 256, 0, st>>>(
          mode, (const T *)g, (const T *)dy, batch_of_row, n, c, (T *)dx);
      int rc = segment_sum_launch<T>((const T *)dy,
                                     mode == MK_BCAST_MUL ? (const T *)x : nullptr,
                                     batch_of_row, n, c, n_batch, acc, cn, st);
      if (rc) return rc;
    }
    k_global_finalize<T><<<(unsigned)ceil_div(n_batch * c, 256), 256, 0, st>>>(
        MK_POOL_SUM, acc, cn, c, n_batch, (T *)dg, (T *)nullptr);
  });
  MK_LAUNCH_CHECK();
  return MK_OK;
}

}  // extern "C"
