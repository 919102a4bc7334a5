// Kernel-map construction.
//
// Replaces count_kernel / preallocated_kernel_map_iteration / direct_in_out_map
// / stride_map_kernel behind CoordinateMapManager::kernel_map
// (pybind/extern.hpp:804; MinkowskiCoordinateManager.py:377-421).
//
// Native form: the output-stationary neighbour table nbr[k*ld + o] (k-major).
// One thread owns one output voxel and probes all K offsets; probes are issued
// in batches of G independent 16-byte loads so that the L2 latency of the hash
// table is overlapped, and each k-row of the table is written fully coalesced.
// Per-offset pair counts fall out of a warp ballot + popc.  The (in,out) pair
// lists the Python API exports are an ordered ballot-scan compaction of the
// table (pairs of an offset ascend in the output row => deterministic).
#include <string.h>

#include <algorithm>

#include "common.cuh"

namespace mk {

constexpr int kMaxK = 512;
struct KOffsets { int32_t v[kMaxK * 3]; };

constexpr int kBuildThreads = 128;

template <int G>
__global__ void __launch_bounds__(kBuildThreads)
k_kmap_build(const int32_t *__restrict__ out_coords, int64_t n_out, int ncols,
             const Slot *__restrict__ table, uint32_t mask,
             const __grid_constant__ KOffsets offs, int K,
             int32_t *__restrict__ nbr, int64_t ld,
             unsigned long long *__restrict__ counts) {
  extern __shared__ int s_cnt[];  // [K] per-block pair counts
  const bool do_count = counts != nullptr;
  if (do_count) {
    for (int k = threadIdx.x; k < K; k += blockDim.x) s_cnt[k] = 0;
    __syncthreads();
  }
  const int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = o < n_out;
  int b = 0, x = 0, y = 0, z = 0;
  if (live) load_coord(out_coords, o, ncols, b, x, y, z);
  const int nd = ncols - 1;

  for (int k0 = 0; k0 < K; k0 += G) {
    uint64_t key[G];
    uint32_t h[G];
    uint4 s[G];
    bool ok[G];
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const int k = k0 + g;
      ok[g] = false;
      if (k < K && live) {
        const int32_t *d = offs.v + k * nd;
        const int nx = x + d[0];
        const int ny = nd > 1 ? y + d[1] : 0;
        const int nz = nd > 2 ? z + d[2] : 0;
        if (coord_in_range(b, nx, ny, nz)) {
          ok[g] = true;
          key[g] = pack_key(b, nx, ny, nz);
          h[g] = hash_key(key[g]) & mask;
          s[g] = __ldg(reinterpret_cast<const uint4 *>(table + h[g]));
        }
      }
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const int k = k0 + g;
      if (k >= K) break;
      int32_t r = -1;
      if (ok[g]) {
        uint4 sl = s[g];
        uint32_t hh = h[g];
        while (true) {
          const uint64_t kk = ((uint64_t)sl.y << 32) | sl.x;
          if (kk == key[g]) { r = (int32_t)sl.z; break; }
          if (kk == kEmptyKey) break;
          hh = (hh + 1) & mask;
          sl = __ldg(reinterpret_cast<const uint4 *>(table + hh));
        }
      }
      if (o < ld) nbr[(int64_t)k * ld + o] = r;   // padding rows [n_out, ld) read as missing
      if (do_count) {
        const unsigned bal = __ballot_sync(0xffffffffu, r >= 0);
        if ((threadIdx.x & 31) == 0 && bal) atomicAdd(&s_cnt[k], __popc(bal));
      }
    }
  }
  if (do_count) {
    __syncthreads();
    for (int k = threadIdx.x; k < K; k += blockDim.x)
      if (s_cnt[k]) atomicAdd(&counts[k], (unsigned long long)s_cnt[k]);
  }
}

__global__ void k_kmap_invert(const int32_t *__restrict__ nbr, int64_t n_out,
                              int64_t ld, int K, int32_t *__restrict__ inv,
                              int64_t n_in, int64_t ld_in) {
  const int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int k = blockIdx.y;
  if (o >= n_out) return;
  const int32_t i = nbr[(int64_t)k * ld + o];
  if (i >= 0 && i < n_in) inv[(int64_t)k * ld_in + i] = (int32_t)o;
}

constexpr int kPairThreads = 256;
constexpr int kPairItems = 8;
constexpr int kPairTile = kPairThreads * kPairItems;

__global__ void __launch_bounds__(kPairThreads)
k_pairs_count(const int32_t *__restrict__ nbr, int64_t n_out, int64_t ld,
              int64_t nblk, int32_t *__restrict__ block_sums) {
  __shared__ int warp_cnt[kPairThreads / 32];
  const int k = blockIdx.y;
  const int64_t base = (int64_t)blockIdx.x * kPairTile;
  int cnt = 0;
#pragma unroll
  for (int r = 0; r < kPairItems; ++r) {
    const int64_t o = base + r * kPairThreads + threadIdx.x;
    const bool f = o < n_out && nbr[(int64_t)k * ld + o] >= 0;
    cnt += __popc(__ballot_sync(0xffffffffu, f));
  }
  if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < kPairThreads / 32; ++w) t += warp_cnt[w];
    block_sums[(int64_t)k * nblk + blockIdx.x] = t;
  }
}

// starts[0]=0, starts[k+1]=starts[k]+totals[k]   (K is small)
__global__ void k_pairs_starts(const int64_t *__restrict__ totals, int K,
                               int64_t *__restrict__ starts) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    int64_t run = 0;
    for (int k = 0; k < K; ++k) {
      starts[k] = run;
      run += totals[k];
    }
    starts[K] = run;
  }
}

__global__ void __launch_bounds__(kPairThreads)
k_pairs_fill(const int32_t *__restrict__ nbr, int64_t n_out, int64_t ld,
             int64_t nblk, const int32_t *__restrict__ block_offsets,
             const int64_t *__restrict__ starts, int32_t *__restrict__ in_rows,
             int32_t *__restrict__ out_rows) {
  __shared__ int warp_cnt[kPairThreads / 32];
  __shared__ long long running_s;
  const int k = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0)
    running_s = starts[k] + block_offsets[(int64_t)k * nblk + blockIdx.x];
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * kPairTile;
#pragma unroll 1
  for (int r = 0; r < kPairItems; ++r) {
    const int64_t o = base + r * kPairThreads + threadIdx.x;
    const int32_t i = o < n_out ? nbr[(int64_t)k * ld + o] : -1;
    const bool f = i >= 0;
    const unsigned bal = __ballot_sync(0xffffffffu, f);
    if (lane == 0) warp_cnt[warp] = __popc(bal);
    __syncthreads();
    int woff = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kPairThreads / 32; ++w) {
      const int c = warp_cnt[w];
      if (w < warp) woff += c;
      tot += c;
    }
    const long long running = running_s;
    if (f) {
      const long long p = running + woff + __popc(bal & ((1u << lane) - 1u));
      in_rows[p] = i;
      out_rows[p] = (int32_t)o;
    }
    __syncthreads();
    if (threadIdx.x == 0) running_s = running + tot;
    __syncthreads();
  }
}

}  // namespace mk

using namespace mk;

extern "C" {

int mk_kmap_build(const int32_t *out_coords, int64_t n_out, int ncols,
                  const void *in_table, int64_t in_capacity,
                  const int32_t *offsets_host, int K, int32_t *nbr, int64_t ld,
                  int64_t *counts_dev, void *stream) {
  MK_CHECK_ARG(ncols >= 2 && ncols <= 4, "ncols must be in [2,4]");
  MK_CHECK_ARG(K >= 1 && K <= kMaxK, "kernel volume must be in [1,512]");
  MK_CHECK_ARG(ld >= n_out, "ld must be >= n_out");
  MK_CHECK_ARG(in_capacity > 0 && (in_capacity & (in_capacity - 1)) == 0,
               "capacity must be a power of two");
  cudaStream_t st = (cudaStream_t)stream;
  if (counts_dev)
    MK_CUDA(cudaMemsetAsync(counts_dev, 0, sizeof(int64_t) * K, st));
  if (n_out == 0) return MK_OK;
  MK_CHECK_ARG(out_coords && in_table && nbr && offsets_host, "null pointer");
  KOffsets offs;
  memcpy(offs.v, offsets_host, sizeof(int32_t) * K * (ncols - 1));
  const unsigned blocks = (unsigned)ceil_div(ld, kBuildThreads);
  const size_t smem = counts_dev ? sizeof(int) * K : 0;
  const uint32_t mask = (uint32_t)(in_capacity - 1);
  if (K % 9 == 0)
    k_kmap_build<9><<<blocks, kBuildThreads, smem, st>>>(
        out_coords, n_out, ncols, (const Slot *)in_table, mask, offs, K, nbr,
        ld, (unsigned long long *)counts_dev);
  else
    k_kmap_build<8><<<blocks, kBuildThreads, smem, st>>>(
        out_coords, n_out, ncols, (const Slot *)in_table, mask, offs, K, nbr,
        ld, (unsigned long long *)counts_dev);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int mk_kmap_invert(const int32_t *nbr, int64_t n_out, int64_t ld, int K,
                   int32_t *inv, int64_t n_in, int64_t ld_in, void *stream) {
  MK_CHECK_ARG(K >= 1 && K <= kMaxK, "kernel volume must be in [1,512]");
  MK_CHECK_ARG(ld >= n_out && ld_in >= n_in, "bad leading dimension");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_in > 0)
    MK_CUDA(cudaMemsetAsync(inv, 0xFF, sizeof(int32_t) * (size_t)K * ld_in, st));
  if (n_out == 0 || n_in == 0) return MK_OK;
  dim3 grid((unsigned)ceil_div(n_out, 256), K);
  k_kmap_invert<<<grid, 256, 0, st>>>(nbr, n_out, ld, K, inv, n_in, ld_in);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

static inline int64_t align256k(int64_t b) { return (b + 255) & ~(int64_t)255; }

int64_t mk_kmap_pairs_workspace_bytes(int64_t n_out, int K) {
  const int64_t nblk = ceil_div(n_out > 0 ? n_out : 1, kPairTile);
  return align256k(nblk * K * 4) + align256k((int64_t)K * 8) + 256;
}

int mk_kmap_pairs(const int32_t *nbr, int64_t n_out, int64_t ld, int K,
                  int64_t *starts_dev, int32_t *in_rows, int32_t *out_rows,
                  void *workspace, void *stream) {
  MK_CHECK_ARG(K >= 1 && K <= kMaxK, "kernel volume must be in [1,512]");
  MK_CHECK_ARG(starts_dev, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_out == 0) {
    MK_CUDA(cudaMemsetAsync(starts_dev, 0, sizeof(int64_t) * (K + 1), st));
    return MK_OK;
  }
  MK_CHECK_ARG(nbr && workspace, "null pointer");
  const int64_t nblk = ceil_div(n_out, kPairTile);
  int32_t *block_sums = (int32_t *)workspace;
  int64_t *totals =
      (int64_t *)((char *)workspace + align256k(nblk * K * 4));
  dim3 grid((unsigned)nblk, K);
  if (in_rows == nullptr) {  // count-only call: starts[] gives the sizes
    k_pairs_count<<<grid, kPairThreads, 0, st>>>(nbr, n_out, ld, nblk,
                                                 block_sums);
    MK_LAUNCH_CHECK();
    int rc = scan_block_sums(block_sums, nblk, K, totals, st);
    if (rc) return rc;
    k_pairs_starts<<<1, 32, 0, st>>>(totals, K, starts_dev);
    MK_LAUNCH_CHECK();
    return MK_OK;
  }
  // fill call: workspace still holds the scanned block offsets of the count call
  k_pairs_fill<<<grid, kPairThreads, 0, st>>>(nbr, n_out, ld, nblk, block_sums,
                                              starts_dev, in_rows, out_rows);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

}  // extern "C"
