// fp32 sparse convolution on the CUDA cores: register-tiled, cp.async double-buffered.
//
// The reference computes fp32 (MinkowskiConvolution.py:265,285; cuBLAS SGEMM / matmul<T>
// behind extern.hpp:84-114) and fp32 is what the example networks run in by default.  The
// tensor cores serve fp32 only while the accumulation chain is short (conv_f32.cu: tcgen05
// accumulates with truncation), so wide fp32 layers need FFMA kernels that run near the FP32
// pipe's rate.  The generic kernels of conv_simt.cu (64 x 64 tile, scalar loads, no overlap of
// loads and math) reach ~10 TFLOP/s; these are the tuned fp32 versions:
//
//   forward / dgrad   CTA = 128 table columns x 128 (or 64) output channels, 256 threads, 8 x 8
//                     (8 x 4) accumulators per thread (rows ty + 16 i, columns tx*4 + 64 j: both
//                     interleaved so that the LDS.128 of a warp are bank-conflict free); per step one offset k and 16 input channels:
//                     the gathered A chunk [128 x 16] (cp.async 16 B, zero-fill for missing
//                     neighbours, row stride 20 floats = conflict-free LDS.128 along the channel)
//                     and the weight chunk [16 x BN]; two buffers, the copies of step s+1 in flight
//                     under the math of step s.  Offsets without any neighbour in the tile are
//                     skipped (tile mask, or a scan of the table when none is given).  With
//                     `out_rows` the table is tile-ordered and column o is output row out_rows[o].
//   weight gradient   CTA = (offset k, 128 x 128 block of dW, range of 128-column tiles); valid
//                     rows of a tile are compacted (ballot scan) so no FFMA is spent on missing
//                     neighbours; x / dy chunks of 16 rows lie in shared memory as they lie in
//                     global memory; partial sums are added to dW with fp32 atomics (as the
//                     reference's matmul2 does).
// Requirements: c_in % 4 == 0 and c_out % 4 == 0 (16-byte row pieces); other shapes stay on
// conv_simt.cu.  For dgrad the weights are transposed into the workspace first.
#include <algorithm>

#include "common.cuh"
#include "sm100.cuh"

namespace mk {
using namespace sm100;

namespace s2 {
constexpr int kThreads = 256;
constexpr int BM = 128;
constexpr int BK = 16;
constexpr int AS = BK + 4;        // A row stride (floats): 20 -> the 8 rows a thread reads hit distinct banks
constexpr int kMaxK = 128;

struct FwdParams {
  const float *x, *w;             // w: [K, c_in, c_out] (already transposed for dgrad)
  const int32_t *nbr;
  const uint32_t *mask;           // [ntiles, W] or nullptr
  const int32_t *out_rows;        // optional
  float *y;
  int64_t ld, n_out, n_in;
  int K, W, kflip, c_in, c_out;
};

template <int BN>
__global__ void __launch_bounds__(kThreads, 2)
k_conv_fwd_f32(const FwdParams p) {
  constexpr int TN = BN / 16;                       // columns per thread: 8 or 4
  __shared__ __align__(16) float As[2][BM][AS];
  __shared__ __align__(16) float Bs[2][BK][BN];
  __shared__ int s_active[kMaxK];
  __shared__ int s_nactive;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t tile = blockIdx.x;
  const int64_t o0 = tile * BM;
  const int n0 = blockIdx.y * BN;
  const int tx = tid & 15, ty = tid >> 4;

  // ---- offsets with any neighbour in this tile ----
  if (tid == 0) s_nactive = 0;
  __syncthreads();
  if (p.mask) {
    if (tid == 0) {
      int n = 0;
      for (int k = 0; k < p.K; ++k) {
        const int kk = p.kflip ? p.K - 1 - k : k;
        if ((__ldg(p.mask + tile * p.W + (kk >> 5)) >> (kk & 31)) & 1u) s_active[n++] = k;
      }
      s_nactive = n;
    }
  } else {
    for (int k = warp; k < p.K; k += kThreads / 32) {
      const int kk = p.kflip ? p.K - 1 - k : k;
      bool any = false;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t o = o0 + j * 32 + lane;
        any |= o < p.n_out && __ldg(p.nbr + (int64_t)kk * p.ld + o) >= 0;
      }
      if (__any_sync(0xffffffffu, any) && lane == 0) s_active[atomicAdd(&s_nactive, 1)] = k;
    }
  }
  __syncthreads();
  const int nactive = s_nactive;
  if (!p.mask && tid == 0) {      // the scan filled the list in arrival order: sort it (K <= 128, once per CTA)
    for (int i = 1; i < nactive; ++i) {
      const int v = s_active[i];
      int j = i - 1;
      for (; j >= 0 && s_active[j] > v; --j) s_active[j + 1] = s_active[j];
      s_active[j + 1] = v;
    }
  }
  __syncthreads();
  const int nchunk = (p.c_in + BK - 1) / BK;
  const int nsteps = nactive * nchunk;

  float acc[8][TN];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  // copies of one step: A 128 rows x 4 pieces (2 per thread), B 16 rows x BN/4 pieces
  const uint32_t as_base = smem_u32(&As[0][0][0]), bs_base = smem_u32(&Bs[0][0][0]);
  auto issue = [&](int s, int buf) {
    const int k = s_active[s / nchunk];
    const int ch = s - (s / nchunk) * nchunk;
    const int kk = p.kflip ? p.K - 1 - k : k;
    const int c0 = ch * BK;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int piece = tid + h * kThreads;        // 0..511
      const int row = piece >> 2, q = piece & 3;
      const int64_t o = o0 + row;
      const int32_t r = o < p.n_out ? __ldg(p.nbr + (int64_t)kk * p.ld + o) : -1;
      const bool ok = r >= 0 && c0 + q * 4 < p.c_in;
      const float *src = p.x + (size_t)(ok ? r : 0) * p.c_in + (ok ? c0 + q * 4 : 0);
      cp_async16(as_base + (uint32_t)(((buf * BM + row) * AS + q * 4) * 4), src, ok ? 16u : 0u);
    }
    const float *wk = p.w + (size_t)k * p.c_in * p.c_out;
    constexpr int BP = BK * BN / 4;                 // B pieces
#pragma unroll
    for (int h = 0; h < BP / kThreads; ++h) {
      const int piece = tid + h * kThreads;
      const int row = piece / (BN / 4), q = piece - row * (BN / 4);
      const bool ok = c0 + row < p.c_in && n0 + q * 4 < p.c_out;
      const float *src = wk + (size_t)(ok ? c0 + row : 0) * p.c_out + (ok ? n0 + q * 4 : 0);
      cp_async16(bs_base + (uint32_t)(((buf * BK + row) * BN + q * 4) * 4), src, ok ? 16u : 0u);
    }
    cp_async_commit();
  };

  if (nsteps > 0) issue(0, 0);
  for (int s = 0; s < nsteps; ++s) {
    const int buf = s & 1;
    if (s + 1 < nsteps) {
      issue(s + 1, buf ^ 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
#pragma unroll
    for (int k4 = 0; k4 < BK / 4; ++k4) {
      float4 a[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = *reinterpret_cast<const float4 *>(&As[buf][ty + 16 * i][k4 * 4]);
#pragma unroll
      for (int kq = 0; kq < 4; ++kq) {
        float b[TN];
#pragma unroll
        for (int j4 = 0; j4 < TN / 4; ++j4) {
          const float4 v = *reinterpret_cast<const float4 *>(&Bs[buf][k4 * 4 + kq][tx * 4 + j4 * 64]);
          b[j4 * 4 + 0] = v.x; b[j4 * 4 + 1] = v.y; b[j4 * 4 + 2] = v.z; b[j4 * 4 + 3] = v.w;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float av = kq == 0 ? a[i].x : kq == 1 ? a[i].y : kq == 2 ? a[i].z : a[i].w;
#pragma unroll
          for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(av, b[j], acc[i][j]);
        }
      }
    }
    __syncthreads();     // the buffer is refilled by the copies issued in the next iteration
  }

#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t o = o0 + ty + 16 * i;
    if (o >= p.n_out) continue;
    const int64_t orow = p.out_rows ? (int64_t)__ldg(p.out_rows + o) : o;
    float *yr = p.y + (size_t)orow * p.c_out + n0 + tx * 4;
#pragma unroll
    for (int j4 = 0; j4 < TN / 4; ++j4)
      if (n0 + tx * 4 + j4 * 64 < p.c_out)
        *reinterpret_cast<float4 *>(yr + j4 * 64) = make_float4(acc[i][j4 * 4], acc[i][j4 * 4 + 1], acc[i][j4 * 4 + 2],
                                                                acc[i][j4 * 4 + 3]);
  }
}

// wt[k][b][a] = w[k][a][b]
__global__ void k_transpose_w(const float *__restrict__ w, int K, int a, int b, float *__restrict__ wt) {
  const int64_t total = (int64_t)K * a * b;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t r = e;
    const int j = (int)(r % a); r /= a;
    const int i = (int)(r % b);
    const int k = (int)(r / b);
    wt[e] = w[((size_t)k * a + j) * b + i];
  }
}

// ---------------------------------------------------------------------------
// weight gradient
// ---------------------------------------------------------------------------
struct WgParams {
  const float *x, *dy;
  const int32_t *nbr, *dy_rows;
  const uint32_t *mask;           // [ntiles, W] or nullptr
  float *dw;                      // [K, c_in, c_out], zeroed before the launch
  int64_t ld, n_out;
  int K, W, c_in, c_out;
  int tiles_per_cta, ntiles, tiles_n;
};

__global__ void __launch_bounds__(kThreads, 2)
k_conv_wgrad_f32(const WgParams p) {
  constexpr int BN = 128;
  __shared__ __align__(16) float As[2][BK][BM];     // [row in chunk][ci]
  __shared__ __align__(16) float Bs[2][BK][BN];     // [row in chunk][co]
  __shared__ int32_t s_in[BM + BK], s_out[BM + BK];   // compacted (x row, dy row) of the current tile
  __shared__ int s_wcnt[kThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int k = blockIdx.z;
  const int m0 = (blockIdx.y / p.tiles_n) * BM, n0 = (blockIdx.y % p.tiles_n) * BN;
  const int tx = tid & 15, ty = tid >> 4;
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  const uint32_t as_base = smem_u32(&As[0][0][0]), bs_base = smem_u32(&Bs[0][0][0]);

  auto issue = [&](int chunk, int buf) {
    // 16 rows x 32 pieces for each of A (x rows) and B (dy rows): 2 + 2 pieces per thread
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int piece = tid + h * kThreads;
      const int rr = piece >> 5, q = piece & 31;
      const int32_t ri = s_in[chunk * BK + rr], ro = s_out[chunk * BK + rr];
      const bool oka = ri >= 0 && m0 + q * 4 < p.c_in;
      cp_async16(as_base + (uint32_t)(((buf * BK + rr) * BM + q * 4) * 4),
                 p.x + (size_t)(oka ? ri : 0) * p.c_in + (oka ? m0 + q * 4 : 0), oka ? 16u : 0u);
      const bool okb = ri >= 0 && n0 + q * 4 < p.c_out;
      cp_async16(bs_base + (uint32_t)(((buf * BK + rr) * BN + q * 4) * 4),
                 p.dy + (size_t)(okb ? ro : 0) * p.c_out + (okb ? n0 + q * 4 : 0), okb ? 16u : 0u);
    }
    cp_async_commit();
  };

  const int t_begin = blockIdx.x * p.tiles_per_cta;
  const int t_end = min(p.ntiles, t_begin + p.tiles_per_cta);
  for (int t = t_begin; t < t_end; ++t) {
    if (p.mask && !((__ldg(p.mask + (int64_t)t * p.W + (k >> 5)) >> (k & 31)) & 1u)) continue;
    // ---- compact the valid rows of the tile (ordered) ----
    __syncthreads();                                // previous tile's lists are no longer read
    int32_t r = -1;
    int64_t o = 0;
    if (tid < BM) {
      o = (int64_t)t * BM + tid;
      r = o < p.n_out ? __ldg(p.nbr + (int64_t)k * p.ld + o) : -1;
    }
    const unsigned bal = __ballot_sync(0xffffffffu, r >= 0);
    if (lane == 0) s_wcnt[warp] = __popc(bal);
    for (int i = tid; i < BM + BK; i += kThreads) s_in[i] = -1;
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const int c = s_wcnt[w];
      if (w < warp) before += c;
      total += c;
    }
    if (tid < BM && r >= 0) {
      const int pos = before + __popc(bal & ((1u << lane) - 1u));
      s_in[pos] = r;
      s_out[pos] = (int32_t)(p.dy_rows ? __ldg(p.dy_rows + o) : o);
    }
    __syncthreads();
    if (total == 0) continue;
    const int nchunks = (total + BK - 1) / BK;
    issue(0, 0);
    for (int c = 0; c < nchunks; ++c) {
      const int buf = c & 1;
      if (c + 1 < nchunks) {
        issue(c + 1, buf ^ 1);
        cp_async_wait<1>();
      } else {
        cp_async_wait<0>();
      }
      __syncthreads();
#pragma unroll
      for (int rr = 0; rr < BK; ++rr) {
        const float4 a0 = *reinterpret_cast<const float4 *>(&As[buf][rr][ty * 4]);
        const float4 a1 = *reinterpret_cast<const float4 *>(&As[buf][rr][ty * 4 + 64]);
        const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][rr][tx * 4]);
        const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][rr][tx * 4 + 64]);
        const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
      }
      __syncthreads();
    }
  }
  float *dwk = p.dw + (size_t)k * p.c_in * p.c_out;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int ci = m0 + ty * 4 + (i & 3) + 64 * (i >> 2);
    if (ci >= p.c_in) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int co = n0 + tx * 4 + (j & 3) + 64 * (j >> 2);
      if (co < p.c_out && acc[i][j] != 0.f) atomicAdd(&dwk[(size_t)ci * p.c_out + co], acc[i][j]);
    }
  }
}
}  // namespace s2

bool conv_f32_simt2_eligible(int c_in, int c_out, int K) {
  return c_in % 4 == 0 && c_out % 4 == 0 && K <= s2::kMaxK && c_in >= 4 && c_out >= 4;
}

int64_t conv_f32_simt2_workspace(int K, int c_in, int c_out) { return (int64_t)K * c_in * c_out * 4 + 256; }

int conv_fwd_f32_simt2(const float *x, const float *w, int transpose_w, int kflip, const int32_t *nbr,
                       const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                       float *y, const int32_t *out_rows, void *workspace, cudaStream_t st) {
  using namespace s2;
  const float *wk = w;
  if (transpose_w) {      // caller's w is [K, c_out, c_in]: the kernel wants [K, c_in (reduction), c_out]
    float *wt = (float *)workspace;
    const int64_t total = (int64_t)K * c_in * c_out;
    k_transpose_w<<<(unsigned)std::min<int64_t>(ceil_div(total, 256), (int64_t)num_sms() * 8), 256, 0, st>>>(
        w, K, c_out, c_in, wt);
    MK_LAUNCH_CHECK();
    wk = wt;
  }
  FwdParams p;
  p.x = x; p.w = wk; p.nbr = nbr; p.mask = tile_mask; p.out_rows = out_rows; p.y = y;
  p.ld = ld; p.n_out = n_out; p.n_in = n_in; p.K = K; p.W = (K + 31) / 32; p.kflip = kflip; p.c_in = c_in; p.c_out = c_out;
  const unsigned tiles = (unsigned)ceil_div(n_out, BM);
  if (c_out > 64) {
    dim3 grid(tiles, (unsigned)ceil_div(c_out, 128));
    k_conv_fwd_f32<128><<<grid, kThreads, 0, st>>>(p);
  } else {
    dim3 grid(tiles, (unsigned)ceil_div(c_out, 64));
    k_conv_fwd_f32<64><<<grid, kThreads, 0, st>>>(p);
  }
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int conv_wgrad_f32_simt2(const float *x, const float *dy, const int32_t *dy_rows, const int32_t *nbr,
                         const uint32_t *tile_mask, int64_t ld, int64_t n_out, int K, int c_in, int c_out, float *dw,
                         cudaStream_t st) {
  using namespace s2;
  MK_CUDA(cudaMemsetAsync(dw, 0, (size_t)K * c_in * c_out * sizeof(float), st));
  if (n_out == 0) return MK_OK;
  WgParams p;
  p.x = x; p.dy = dy; p.nbr = nbr; p.dy_rows = dy_rows; p.mask = tile_mask; p.dw = dw;
  p.ld = ld; p.n_out = n_out; p.K = K; p.W = (K + 31) / 32; p.c_in = c_in; p.c_out = c_out;
  p.ntiles = (int)ceil_div(n_out, BM);
  const int tiles_m = (int)ceil_div(c_in, BM);
  p.tiles_n = (int)ceil_div(c_out, 128);
  const int64_t per_split = (int64_t)K * tiles_m * p.tiles_n;
  int64_t splits = std::max<int64_t>(1, (int64_t)num_sms() * 4 / per_split);
  splits = std::min<int64_t>(splits, p.ntiles);
  p.tiles_per_cta = (int)ceil_div(p.ntiles, splits);
  splits = ceil_div(p.ntiles, p.tiles_per_cta);
  dim3 grid((unsigned)splits, (unsigned)(tiles_m * p.tiles_n), (unsigned)K);
  k_conv_wgrad_f32<<<grid, kThreads, 0, st>>>(p);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

}  // namespace mk
