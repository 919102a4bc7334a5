// Segment statistics and per-segment affine maps: the two passes every
// normalisation layer of the example networks is made of.
//
// Replaces, for MinkowskiBatchNorm / MinkowskiSyncBatchNorm (MinkowskiNormalization.py:
// 55-195, torch.nn.BatchNorm1d on the feature matrix) and for the instance-norm family
// (MinkowskiInstanceNorm :198-314, the fork's MinkowskiStableInstanceNorm :317-402 and
// AdaStableInstanceNorm :546-622: two GlobalPooling + three Broadcast launches + ~10
// elementwise torch ops per call, extern.hpp:332-392), the chain of full passes over
// [N, C] by one read (statistics) and one read + write (normalise + affine (+ AdaIN)):
//
//   mk_seg_sums    out[s][0][c] = sum_{r in s} (a[r,c] - sa[s,c])
//                  out[s][1][c] = sum_{r in s} (a[r,c] - sa[s,c]) * (b[r,c] - sb[s,c])
//       forward statistics: a = b = x, sa = sb = K (any value near the mean: the shifted
//       sums do not cancel); backward: a = dy, b = x, sb = mean.
//   mk_seg_affine  out[r,c] = A[s,c] * a[r,c] + B[s,c] * b[r,c] + C[s,c],  s = seg[r]
//       forward: y = A x + C; backward: dx = A dy + B x + C.
// A segment is a batch item (instance norms; seg = batch row of every voxel) or the whole
// tensor (batch norm; seg == NULL).  Sums are fp32 (fp64 for fp64 features), flushed with
// one atomic per (CTA, segment change, channel); both kernels stream rows with 16-byte
// accesses, grid = a multiple of the SM count, and are HBM-bound (algorithmic bytes:
// N*C*b read for the sums, 2*N*C*b for the affine map with one input).
#include <algorithm>

#include "common.cuh"

namespace mk {
namespace {

constexpr int kNormThreads = 256;

template <typename T> struct Vec16 { static constexpr int n = 16 / sizeof(T); };

template <typename T, int V>
__device__ __forceinline__ void load_vec(const T *p, typename Acc<T>::type (&v)[V]) {
  union { uint4 u; T t[V]; } raw;
  raw.u = __ldg(reinterpret_cast<const uint4 *>(p));
#pragma unroll
  for (int i = 0; i < V; ++i) v[i] = to_acc<T>(raw.t[i]);
}

template <typename T, int V>
__device__ __forceinline__ void store_vec(T *p, const typename Acc<T>::type (&v)[V]) {
  union { uint4 u; T t[V]; } raw;
#pragma unroll
  for (int i = 0; i < V; ++i) raw.t[i] = from_acc<T, typename Acc<T>::type>(v[i]);
  *reinterpret_cast<uint4 *>(p) = raw.u;
}

// Vector path: c % V == 0, 16-byte aligned rows.  Thread (tx, ty): channel vector tx of rows
// row0 + ty, row0 + ty + TY, ...; a CTA owns `rows_per_cta` consecutive rows.
template <typename T>
__global__ void __launch_bounds__(kNormThreads)
k_seg_sums_vec(const T *__restrict__ a, const T *__restrict__ b, const typename Acc<T>::type *__restrict__ sa,
               const typename Acc<T>::type *__restrict__ sb, const int32_t *__restrict__ seg, int64_t n, int c,
               int tx_n, int64_t rows_per_cta, typename Acc<T>::type *__restrict__ out) {
  using A = typename Acc<T>::type;
  constexpr int V = Vec16<T>::n;
  const int tx = threadIdx.x % tx_n, ty = threadIdx.x / tx_n, ty_n = kNormThreads / tx_n;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t r1 = min(n, r0 + rows_per_cta);
  const int ch = tx * V;
  const bool active = ch < c;
  __shared__ A red[2][kNormThreads * V];
  A s0[V], s1[V], va[V], vb[V], ka[V], kb[V];
#pragma unroll
  for (int i = 0; i < V; ++i) { s0[i] = 0; s1[i] = 0; ka[i] = 0; kb[i] = 0; }
  int cur = -1;
  auto flush = [&]() {
    if (cur < 0) return;
    A *o = out + (size_t)cur * 2 * c + ch;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      atomicAdd(o + i, s0[i]);
      atomicAdd(o + c + i, s1[i]);
      s0[i] = 0; s1[i] = 0;
    }
  };
  if (active) {
    int64_t r = r0 + ty;
    // Four rows per iteration with their loads issued together (one 16-byte load in flight per thread
    // left the pass latency-bound: 1.4 TB/s), taken while the four rows stay in the current segment;
    // otherwise one row at a time (segment boundary, start and tail of the range).
    constexpr int U = 4;
    while (r < r1) {
      bool fast = cur >= 0 && r + (int64_t)(U - 1) * ty_n < r1;
      if (fast && seg) {
#pragma unroll
        for (int u = 0; u < U; ++u) fast = fast && __ldg(seg + r + (int64_t)u * ty_n) == cur;
      }
      if (fast) {
        A xa[U][V], xb[U][V];
#pragma unroll
        for (int u = 0; u < U; ++u) load_vec<T, V>(a + (size_t)(r + (int64_t)u * ty_n) * c + ch, xa[u]);
        if (b != a) {
#pragma unroll
          for (int u = 0; u < U; ++u) load_vec<T, V>(b + (size_t)(r + (int64_t)u * ty_n) * c + ch, xb[u]);
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int i = 0; i < V; ++i) {
            const A x = xa[u][i] - ka[i];
            s0[i] += x;
            s1[i] += x * ((b != a ? xb[u][i] : xa[u][i]) - kb[i]);
          }
        r += (int64_t)U * ty_n;
        continue;
      }
      const int s = seg ? __ldg(seg + r) : 0;
      if (s != cur) {
        flush();   // a segment boundary inside the CTA's range (rare: rows of a batch item are contiguous)
        cur = s;
        if (s >= 0) {
#pragma unroll
          for (int i = 0; i < V; ++i) {
            ka[i] = sa ? sa[(size_t)s * c + ch + i] : (A)0;
            kb[i] = sb ? sb[(size_t)s * c + ch + i] : (A)0;
          }
        }
      }
      if (s >= 0) {
        load_vec<T, V>(a + (size_t)r * c + ch, va);
        if (b == a) {
#pragma unroll
          for (int i = 0; i < V; ++i) vb[i] = va[i];
        } else {
          load_vec<T, V>(b + (size_t)r * c + ch, vb);
        }
#pragma unroll
        for (int i = 0; i < V; ++i) {
          const A x = va[i] - ka[i];
          s0[i] += x;
          s1[i] += x * (vb[i] - kb[i]);
        }
      }
      r += ty_n;
    }
  }
  // the CTA's partial sums of its last segment are added up in shared memory first: one
  // atomic per (CTA, channel) instead of one per thread
  const int last = r1 > r0 ? (seg ? __ldg(seg + (r1 - 1)) : 0) : -1;
  if (active && cur != last) { flush(); cur = -1; }
#pragma unroll
  for (int i = 0; i < V; ++i) {
    red[0][threadIdx.x * V + i] = (active && cur == last) ? s0[i] : (A)0;
    red[1][threadIdx.x * V + i] = (active && cur == last) ? s1[i] : (A)0;
  }
  __syncthreads();
  if (ty == 0 && active && last >= 0) {
    A *o = out + (size_t)last * 2 * c + ch;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      A t0 = 0, t1 = 0;
      for (int y = 0; y < ty_n; ++y) {
        t0 += red[0][(y * tx_n + tx) * V + i];
        t1 += red[1][(y * tx_n + tx) * V + i];
      }
      atomicAdd(o + i, t0);
      atomicAdd(o + c + i, t1);
    }
  }
}

// Scalar path (any c, any alignment): one thread per (row slot, channel)
template <typename T>
__global__ void __launch_bounds__(kNormThreads)
k_seg_sums_scalar(const T *__restrict__ a, const T *__restrict__ b, const typename Acc<T>::type *__restrict__ sa,
                  const typename Acc<T>::type *__restrict__ sb, const int32_t *__restrict__ seg, int64_t n, int c,
                  int64_t rows_per_cta, typename Acc<T>::type *__restrict__ out) {
  using A = typename Acc<T>::type;
  const int tx_n = min(c, kNormThreads);
  const int ty_n = kNormThreads / tx_n;
  const int tx = threadIdx.x % tx_n, ty = threadIdx.x / tx_n;
  if (ty >= ty_n) return;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t r1 = min(n, r0 + rows_per_cta);
  for (int ch = tx; ch < c; ch += tx_n) {
    A s0 = 0, s1 = 0, ka = 0, kb = 0;
    int cur = -1;
    for (int64_t r = r0 + ty; r < r1; r += ty_n) {
      const int s = seg ? __ldg(seg + r) : 0;
      if (s != cur) {
        if (cur >= 0) { atomicAdd(out + (size_t)cur * 2 * c + ch, s0); atomicAdd(out + (size_t)cur * 2 * c + c + ch, s1); }
        s0 = 0; s1 = 0; cur = s;
        if (s >= 0) { ka = sa ? sa[(size_t)s * c + ch] : (A)0; kb = sb ? sb[(size_t)s * c + ch] : (A)0; }
      }
      if (s < 0) continue;
      const A x = to_acc<T>(a[(size_t)r * c + ch]) - ka;
      s0 += x;
      s1 += x * (to_acc<T>(b[(size_t)r * c + ch]) - kb);
    }
    if (cur >= 0) { atomicAdd(out + (size_t)cur * 2 * c + ch, s0); atomicAdd(out + (size_t)cur * 2 * c + c + ch, s1); }
  }
}

template <typename T>
__global__ void __launch_bounds__(kNormThreads)
k_seg_affine_vec(const T *__restrict__ a, const T *__restrict__ b, const int32_t *__restrict__ seg, int64_t n, int c,
                 const typename Acc<T>::type *__restrict__ ca, const typename Acc<T>::type *__restrict__ cb,
                 const typename Acc<T>::type *__restrict__ cc, T *__restrict__ out) {
  using A = typename Acc<T>::type;
  constexpr int V = Vec16<T>::n;
  const int vec_per_row = c / V;
  const int64_t total = n * vec_per_row;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = e / vec_per_row;
    const int ch = (int)(e - r * vec_per_row) * V;
    const int s = seg ? __ldg(seg + r) : 0;
    A va[V], vb[V], o[V];
    load_vec<T, V>(a + (size_t)r * c + ch, va);
    if (cb) load_vec<T, V>(b + (size_t)r * c + ch, vb);
    if (s < 0) {
#pragma unroll
      for (int i = 0; i < V; ++i) o[i] = 0;
    } else {
      const A *pa = ca + (size_t)s * c + ch, *pc = cc + (size_t)s * c + ch;
#pragma unroll
      for (int i = 0; i < V; ++i) o[i] = __ldg(pa + i) * va[i] + __ldg(pc + i);
      if (cb) {
        const A *pb = cb + (size_t)s * c + ch;
#pragma unroll
        for (int i = 0; i < V; ++i) o[i] += __ldg(pb + i) * vb[i];
      }
    }
    store_vec<T, V>(out + (size_t)r * c + ch, o);
  }
}

template <typename T>
__global__ void __launch_bounds__(kNormThreads)
k_seg_affine_scalar(const T *__restrict__ a, const T *__restrict__ b, const int32_t *__restrict__ seg, int64_t n,
                    int c, const typename Acc<T>::type *__restrict__ ca, const typename Acc<T>::type *__restrict__ cb,
                    const typename Acc<T>::type *__restrict__ cc, T *__restrict__ out) {
  using A = typename Acc<T>::type;
  const int64_t total = n * c;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = e / c;
    const int ch = (int)(e - r * c);
    const int s = seg ? __ldg(seg + r) : 0;
    A o = 0;
    if (s >= 0) {
      o = ca[(size_t)s * c + ch] * to_acc<T>(a[e]) + cc[(size_t)s * c + ch];
      if (cb) o += cb[(size_t)s * c + ch] * to_acc<T>(b[e]);
    }
    out[e] = from_acc<T, A>(o);
  }
}

// Batch-norm coefficient kernels: the [C]-sized algebra between the two passes as ONE launch each
// (as torch ops it is ~40 tiny launches per layer and dominated the host time of the VAE step).
//   forward : S = sums of (x - K), (x - K)^2 over m rows  ->  mean, invstd, y = A x + C coefficients,
//             running statistics (momentum f, unbiased variance) as torch.nn.BatchNorm1d keeps them
//   backward: S = sums of dy, dy (x - mean)  ->  dx = cA dy + cB x + cC coefficients, dweight, dbias
template <typename A>
__global__ void k_bn_coef_fwd(const A *__restrict__ S, const A *__restrict__ shift, const A *__restrict__ count_dev,
                              double count_host, const A *__restrict__ weight, const A *__restrict__ bias, double eps,
                              int c, A *__restrict__ coef_a, A *__restrict__ coef_c, A *__restrict__ mean_out,
                              A *__restrict__ invstd_out, A *__restrict__ running_mean, A *__restrict__ running_var,
                              double momentum) {
  const int ch = blockIdx.x * blockDim.x + threadIdx.x;
  if (ch >= c) return;
  const A m = count_dev ? max(count_dev[0], (A)1) : (A)count_host;
  const A d = S[ch] / m;
  const A mean = (shift ? shift[ch] : (A)0) + d;
  A var = S[c + ch] / m - d * d;
  var = var > (A)0 ? var : (A)0;
  const A inv = (A)1 / sqrt(var + (A)eps);
  const A a = weight ? weight[ch] * inv : inv;
  coef_a[ch] = a;
  coef_c[ch] = (bias ? bias[ch] : (A)0) - mean * a;
  mean_out[ch] = mean;
  invstd_out[ch] = inv;
  if (running_mean) {
    const A f = (A)momentum;
    const A unbiased = var * (m / max(m - (A)1, (A)1));
    running_mean[ch] = ((A)1 - f) * running_mean[ch] + f * mean;
    running_var[ch] = ((A)1 - f) * running_var[ch] + f * unbiased;
  }
}

template <typename A>
__global__ void k_bn_coef_bwd(const A *__restrict__ S, const A *__restrict__ mean, const A *__restrict__ invstd,
                              const A *__restrict__ count_dev, double count_host, const A *__restrict__ weight, int c,
                              A *__restrict__ coef_a, A *__restrict__ coef_b, A *__restrict__ coef_c,
                              A *__restrict__ dweight, A *__restrict__ dbias) {
  const int ch = blockIdx.x * blockDim.x + threadIdx.x;
  if (ch >= c) return;
  const A m = count_dev ? max(count_dev[0], (A)1) : (A)count_host;
  const A s0 = S[ch], s1 = S[c + ch], inv = invstd[ch], w = weight ? weight[ch] : (A)1;
  const A a = w * inv;
  const A b = -w * inv * inv * inv * s1 / m;
  coef_a[ch] = a;
  coef_b[ch] = b;
  coef_c[ch] = -a * s0 / m - b * mean[ch];
  if (dweight) dweight[ch] = inv * s1;
  if (dbias) dbias[ch] = s0;
}

template <typename T>
int seg_sums_t(const void *a, const void *b, const void *sa, const void *sb, const int32_t *seg, int64_t n, int c,
               int64_t n_seg, void *out, cudaStream_t st) {
  using A = typename Acc<T>::type;
  MK_CUDA(cudaMemsetAsync(out, 0, (size_t)n_seg * 2 * c * sizeof(A), st));
  if (n == 0) return MK_OK;
  constexpr int V = Vec16<T>::n;
  // enough CTAs to fill the machine, each with a long enough row range to amortise its flush
  // (two CTAs per SM: every CTA ends with one atomic per channel on the SAME 2C addresses, and 1184 CTAs
  // serialised there for ~60 us; 296 keep 4.8 MB of loads in flight, which is enough)
  const int64_t ctas = std::max<int64_t>(1, std::min<int64_t>((int64_t)num_sms() * 2, ceil_div(n, 64)));
  const int64_t rows_per_cta = ceil_div(n, ctas);
  const unsigned grid = (unsigned)ceil_div(n, rows_per_cta);
  const bool vec = c % V == 0 && c / V <= kNormThreads && ((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0;
  if (vec) {
    int tx_n = 1;
    while (tx_n < c / V) tx_n <<= 1;
    k_seg_sums_vec<T><<<grid, kNormThreads, 0, st>>>((const T *)a, (const T *)b, (const A *)sa, (const A *)sb, seg, n, c,
                                                     tx_n, rows_per_cta, (A *)out);
  } else {
    k_seg_sums_scalar<T><<<grid, kNormThreads, 0, st>>>((const T *)a, (const T *)b, (const A *)sa, (const A *)sb, seg, n,
                                                        c, rows_per_cta, (A *)out);
  }
  MK_LAUNCH_CHECK();
  return MK_OK;
}

template <typename T>
int seg_affine_t(const void *a, const void *b, const int32_t *seg, int64_t n, int c, const void *ca, const void *cb,
                 const void *cc, void *out, cudaStream_t st) {
  using A = typename Acc<T>::type;
  if (n == 0) return MK_OK;
  constexpr int V = Vec16<T>::n;
  const bool vec = c % V == 0 && ((uintptr_t)a & 15) == 0 && ((uintptr_t)out & 15) == 0 && (!cb || ((uintptr_t)b & 15) == 0);
  const int64_t total = vec ? n * (c / V) : n * c;
  const unsigned grid = (unsigned)std::max<int64_t>(1, std::min<int64_t>(ceil_div(total, kNormThreads), (int64_t)num_sms() * 16));
  if (vec)
    k_seg_affine_vec<T><<<grid, kNormThreads, 0, st>>>((const T *)a, (const T *)b, seg, n, c, (const A *)ca, (const A *)cb,
                                                       (const A *)cc, (T *)out);
  else
    k_seg_affine_scalar<T><<<grid, kNormThreads, 0, st>>>((const T *)a, (const T *)b, seg, n, c, (const A *)ca,
                                                          (const A *)cb, (const A *)cc, (T *)out);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

}  // namespace
}  // namespace mk

using namespace mk;

// ---------------------------------------------------------------------------
// Instance / group norm coefficients (MinkowskiInstanceNorm, MinkowskiStableInstanceNorm,
// MinkowskiStableGroupNorm: MinkowskiNormalization.py:198-478 of the reference).  Per batch item s
// the statistics are taken over the item's rows and the `gs` adjacent channels of a statistics
// group G; the affine parameters are shared by `gw` adjacent channels (weight / bias [C / gw]).
//   forward:  per channel mean_c, var_c from S = sums of (x - K), (x - K)^2;
//             mu_G = avg_c mean_c,  v_G = avg_c (var_c + (mean_c - mu_G)^2),  inv = rsqrt(v_G + eps)
//             y = A x + C,  A = inv w,  C = b - mu_G A
//   backward: S = sums of dy, dy (x - mean_c);  P0 = sum_c w S0_c,
//             P1 = sum_c w inv (S1_c + (mean_c - mu_G) S0_c),  M = m gs
//             dx = A dy + cB x + cC,  cB = -inv^2 P1 / M,  cC = -inv P0 / M - cB mu_G
//             dweight_j = sum_s sum_{c in j} inv (S1_c + (mean_c - mu) S0_c),  dbias_j = sum_s sum_{c in j} S0_c
// One thread per (batch item, statistics group); these are [n_seg, C]-sized problems - the point is
// ONE launch instead of the ~40 tiny torch kernels of the same algebra (a diffusion-UNet step holds
// 102 such norms and was bound by their launches).
template <typename A>
__global__ void k_gn_coef_fwd(const A *__restrict__ S, const A *__restrict__ shift, const A *__restrict__ counts,
                              const A *__restrict__ weight, const A *__restrict__ bias, double eps, int c, int gs,
                              int gw, int n_seg, A *__restrict__ coef_a, A *__restrict__ coef_c,
                              A *__restrict__ mean_out, A *__restrict__ mu_out, A *__restrict__ inv_out) {
  const int ng = c / gs;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_seg * ng) return;
  const int s = t / ng, g = t - s * ng;
  const A m = max(counts[s], (A)1);
  const A *s0 = S + (size_t)s * 2 * c, *s1 = s0 + c;
  A mu = 0;
  for (int i = 0; i < gs; ++i) {
    const int ch = g * gs + i;
    const A mean = (shift ? shift[(size_t)s * c + ch] : (A)0) + s0[ch] / m;
    mean_out[(size_t)s * c + ch] = mean;
    mu += mean;
  }
  mu /= (A)gs;
  A v = 0;
  for (int i = 0; i < gs; ++i) {
    const int ch = g * gs + i;
    const A d = s0[ch] / m;
    A var = s1[ch] / m - d * d;
    var = var > (A)0 ? var : (A)0;
    const A dm = (shift ? shift[(size_t)s * c + ch] : (A)0) + d - mu;
    v += var + dm * dm;
  }
  v /= (A)gs;
  const A inv = (A)1 / sqrt(v + (A)eps);
  mu_out[t] = mu;
  inv_out[t] = inv;
  for (int i = 0; i < gs; ++i) {
    const int ch = g * gs + i;
    const A a = inv * (weight ? weight[ch / gw] : (A)1);
    coef_a[(size_t)s * c + ch] = a;
    coef_c[(size_t)s * c + ch] = (bias ? bias[ch / gw] : (A)0) - mu * a;
  }
}

template <typename A>
__global__ void k_gn_coef_bwd(const A *__restrict__ S, const A *__restrict__ mean, const A *__restrict__ mu_in,
                              const A *__restrict__ inv_in, const A *__restrict__ counts, const A *__restrict__ weight,
                              int c, int gs, int gw, int n_seg, A *__restrict__ coef_a, A *__restrict__ coef_b,
                              A *__restrict__ coef_c) {
  const int ng = c / gs;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_seg * ng) return;
  const int s = t / ng, g = t - s * ng;
  const A M = max(counts[s], (A)1) * (A)gs;
  const A mu = mu_in[t], inv = inv_in[t];
  const A *s0 = S + (size_t)s * 2 * c, *s1 = s0 + c;
  A p0 = 0, p1 = 0;
  for (int i = 0; i < gs; ++i) {
    const int ch = g * gs + i;
    const A w = weight ? weight[ch / gw] : (A)1;
    p0 += w * s0[ch];
    p1 += w * inv * (s1[ch] + (mean[(size_t)s * c + ch] - mu) * s0[ch]);
  }
  const A cb = -inv * inv * p1 / M;
  const A cc = -inv * p0 / M - cb * mu;
  for (int i = 0; i < gs; ++i) {
    const int ch = g * gs + i;
    coef_a[(size_t)s * c + ch] = inv * (weight ? weight[ch / gw] : (A)1);
    coef_b[(size_t)s * c + ch] = cb;
    coef_c[(size_t)s * c + ch] = cc;
  }
}

// parameter gradients: one thread per weight index j (the channels [j gw, (j+1) gw) of every batch item)
template <typename A>
__global__ void k_gn_param_grads(const A *__restrict__ S, const A *__restrict__ mean, const A *__restrict__ mu_in,
                                 const A *__restrict__ inv_in, int c, int gs, int gw, int n_seg,
                                 A *__restrict__ dweight, A *__restrict__ dbias) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= c / gw) return;
  const int ng = c / gs;
  A dw = 0, db = 0;
  for (int s = 0; s < n_seg; ++s) {
    const A *s0 = S + (size_t)s * 2 * c, *s1 = s0 + c;
    for (int i = 0; i < gw; ++i) {
      const int ch = j * gw + i;
      const int t = s * ng + ch / gs;
      dw += inv_in[t] * (s1[ch] + (mean[(size_t)s * c + ch] - mu_in[t]) * s0[ch]);
      db += s0[ch];
    }
  }
  if (dweight) dweight[j] = dw;
  if (dbias) dbias[j] = db;
}

extern "C" {

int mk_seg_sums(int dtype, const void *a, const void *b, const void *shift_a, const void *shift_b,
                const int32_t *seg_of_row, int64_t n, int c, int64_t n_seg, void *out, void *stream) {
  MK_CHECK_ARG(n >= 0 && c >= 1 && n_seg >= 1, "bad sizes");
  MK_CHECK_ARG(out && (n == 0 || (a && b)), "null pointer");
  MK_DISPATCH_DTYPE(dtype, return seg_sums_t<T>(a, b, shift_a, shift_b, seg_of_row, n, c, n_seg, out,
                                                (cudaStream_t)stream));
  return MK_OK;
}

int mk_seg_affine(int dtype, const void *a, const void *b, const int32_t *seg_of_row, int64_t n, int c,
                  const void *coef_a, const void *coef_b, const void *coef_c, void *out, void *stream) {
  MK_CHECK_ARG(n >= 0 && c >= 1, "bad sizes");
  MK_CHECK_ARG(n == 0 || (a && out && coef_a && coef_c && (!coef_b || b)), "null pointer");
  MK_DISPATCH_DTYPE(dtype, return seg_affine_t<T>(a, b, seg_of_row, n, c, coef_a, coef_b, coef_c, out,
                                                  (cudaStream_t)stream));
  return MK_OK;
}

int mk_bn_coef_fwd(int acc_f64, const void *sums, const void *shift, const void *count_dev, double count_host,
                   const void *weight, const void *bias, double eps, int c, void *coef_a, void *coef_c,
                   void *mean_out, void *invstd_out, void *running_mean, void *running_var, double momentum,
                   void *stream) {
  MK_CHECK_ARG(c >= 1 && sums && coef_a && coef_c && mean_out && invstd_out, "bad arguments");
  MK_CHECK_ARG((running_mean == nullptr) == (running_var == nullptr), "running statistics come in pairs");
  const unsigned grid = (unsigned)ceil_div(c, 128);
  if (acc_f64)
    k_bn_coef_fwd<double><<<grid, 128, 0, (cudaStream_t)stream>>>(
        (const double *)sums, (const double *)shift, (const double *)count_dev, count_host, (const double *)weight,
        (const double *)bias, eps, c, (double *)coef_a, (double *)coef_c, (double *)mean_out, (double *)invstd_out,
        (double *)running_mean, (double *)running_var, momentum);
  else
    k_bn_coef_fwd<float><<<grid, 128, 0, (cudaStream_t)stream>>>(
        (const float *)sums, (const float *)shift, (const float *)count_dev, count_host, (const float *)weight,
        (const float *)bias, eps, c, (float *)coef_a, (float *)coef_c, (float *)mean_out, (float *)invstd_out,
        (float *)running_mean, (float *)running_var, momentum);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int mk_bn_coef_bwd(int acc_f64, const void *sums, const void *mean, const void *invstd, const void *count_dev,
                   double count_host, const void *weight, int c, void *coef_a, void *coef_b, void *coef_c,
                   void *dweight, void *dbias, void *stream) {
  MK_CHECK_ARG(c >= 1 && sums && mean && invstd && coef_a && coef_b && coef_c, "bad arguments");
  const unsigned grid = (unsigned)ceil_div(c, 128);
  if (acc_f64)
    k_bn_coef_bwd<double><<<grid, 128, 0, (cudaStream_t)stream>>>(
        (const double *)sums, (const double *)mean, (const double *)invstd, (const double *)count_dev, count_host,
        (const double *)weight, c, (double *)coef_a, (double *)coef_b, (double *)coef_c, (double *)dweight,
        (double *)dbias);
  else
    k_bn_coef_bwd<float><<<grid, 128, 0, (cudaStream_t)stream>>>(
        (const float *)sums, (const float *)mean, (const float *)invstd, (const float *)count_dev, count_host,
        (const float *)weight, c, (float *)coef_a, (float *)coef_b, (float *)coef_c, (float *)dweight, (float *)dbias);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int mk_gn_coef_fwd(int acc_f64, const void *sums, const void *shift, const void *counts, const void *weight,
                   const void *bias, double eps, int c, int gs, int gw, int64_t n_seg, void *coef_a, void *coef_c,
                   void *mean_out, void *mu_out, void *inv_out, void *stream) {
  MK_CHECK_ARG(c >= 1 && gs >= 1 && gw >= 1 && c % gs == 0 && c % gw == 0 && n_seg >= 0, "bad channel grouping");
  MK_CHECK_ARG(n_seg <= (1 << 20), "too many segments");
  if (n_seg == 0) return MK_OK;
  MK_CHECK_ARG(sums && counts && coef_a && coef_c && mean_out && mu_out && inv_out, "null pointer");
  const unsigned grid = (unsigned)ceil_div((int64_t)n_seg * (c / gs), 128);
  if (acc_f64)
    k_gn_coef_fwd<double><<<grid, 128, 0, (cudaStream_t)stream>>>(
        (const double *)sums, (const double *)shift, (const double *)counts, (const double *)weight,
        (const double *)bias, eps, c, gs, gw, (int)n_seg, (double *)coef_a, (double *)coef_c, (double *)mean_out,
        (double *)mu_out, (double *)inv_out);
  else
    k_gn_coef_fwd<float><<<grid, 128, 0, (cudaStream_t)stream>>>(
        (const float *)sums, (const float *)shift, (const float *)counts, (const float *)weight, (const float *)bias,
        eps, c, gs, gw, (int)n_seg, (float *)coef_a, (float *)coef_c, (float *)mean_out, (float *)mu_out,
        (float *)inv_out);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int mk_gn_coef_bwd(int acc_f64, const void *sums, const void *mean, const void *mu, const void *inv,
                   const void *counts, const void *weight, int c, int gs, int gw, int64_t n_seg, void *coef_a,
                   void *coef_b, void *coef_c, void *dweight, void *dbias, void *stream) {
  MK_CHECK_ARG(c >= 1 && gs >= 1 && gw >= 1 && c % gs == 0 && c % gw == 0 && n_seg >= 0, "bad channel grouping");
  MK_CHECK_ARG(n_seg <= (1 << 20), "too many segments");
  MK_CHECK_ARG(sums && mean && mu && inv && counts && coef_a && coef_b && coef_c, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  const unsigned grid = (unsigned)std::max<int64_t>(1, ceil_div((int64_t)n_seg * (c / gs), 128));
  const unsigned gridp = (unsigned)ceil_div(c / gw, 128);
  if (acc_f64) {
    if (n_seg)
      k_gn_coef_bwd<double><<<grid, 128, 0, st>>>((const double *)sums, (const double *)mean, (const double *)mu,
                                                  (const double *)inv, (const double *)counts, (const double *)weight,
                                                  c, gs, gw, (int)n_seg, (double *)coef_a, (double *)coef_b,
                                                  (double *)coef_c);
    if (dweight || dbias)
      k_gn_param_grads<double><<<gridp, 128, 0, st>>>((const double *)sums, (const double *)mean, (const double *)mu,
                                                      (const double *)inv, c, gs, gw, (int)n_seg, (double *)dweight,
                                                      (double *)dbias);
  } else {
    if (n_seg)
      k_gn_coef_bwd<float><<<grid, 128, 0, st>>>((const float *)sums, (const float *)mean, (const float *)mu,
                                                 (const float *)inv, (const float *)counts, (const float *)weight, c,
                                                 gs, gw, (int)n_seg, (float *)coef_a, (float *)coef_b, (float *)coef_c);
    if (dweight || dbias)
      k_gn_param_grads<float><<<gridp, 128, 0, st>>>((const float *)sums, (const float *)mean, (const float *)mu,
                                                     (const float *)inv, c, gs, gw, (int)n_seg, (float *)dweight,
                                                     (float *)dbias);
  }
  MK_LAUNCH_CHECK();
  return MK_OK;
}

}  // extern "C"
