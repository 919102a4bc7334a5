// Kernel-map construction.
//
// Replaces count_kernel / preallocated_kernel_map_iteration / direct_in_out_map
// / stride_map_kernel behind CoordinateMapManager::kernel_map
// (pybind/extern.hpp:804; MinkowskiCoordinateManager.py:377-421).
//
// Native form: the output-stationary neighbour table nbr[k*ld + o] (k-major).
// One thread owns one output voxel and probes all K offsets; probes are issued
// in batches of G offsets, one 32-byte sector load per distinct RUN of 4 cells
// (common.cuh: a 3-wide kernel row is 1 sector, 2 across a run boundary), so
// that the L2 latency of the hash table is overlapped; each k-row of the table
// is written fully coalesced.
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
             const Bucket *__restrict__ table, uint32_t mask,
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
    BucketRegs s[G];
    int sub[G];
    uint32_t pend = 0, dup = 0, hit = 0;
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const int k = k0 + g;
      key[g] = kEmptyKey;
      sub[g] = 0;
      if (k < K && live) {
        const int32_t *d = offs.v + k * nd;
        const int nx = x + d[0];
        const int ny = nd > 1 ? y + d[1] : 0;
        const int nz = nd > 2 ? z + d[2] : 0;
        if (coord_in_range(b, nx, ny, nz)) {
          key[g] = run_key(b, nx, ny, nz);
          sub[g] = nx & 3;
          // offsets enumerate x fastest: the previous offset is usually the
          // same run, whose bucket is already on its way
          if (g > 0 && key[g] == key[g - 1]) {
            dup |= 1u << g;
          } else {
            pend |= 1u << g;
            h[g] = bucket_home(key[g], mask);
            s[g] = ldg_bucket(table + h[g]);
          }
        }
      }
    }
    // all chains of the batch advance together: the (rare) second sector of
    // several runs is in flight at once instead of one after the other
    while (pend) {
#pragma unroll
      for (int g = 0; g < G; ++g) {
        if (pend & (1u << g)) {
          const uint64_t kk = bucket_key(s[g]);
          if (kk == key[g]) {
            hit |= 1u << g;
            pend &= ~(1u << g);
          } else if (kk == kEmptyKey) {
            pend &= ~(1u << g);
          } else {
            h[g] = (h[g] + 1) & mask;
            s[g] = ldg_bucket(table + h[g]);
          }
        }
      }
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const int k = k0 + g;
      if (k >= K) break;
      if (g > 0 && (dup & (1u << g))) {
        s[g] = s[g - 1];
        hit |= ((hit >> (g - 1)) & 1u) << g;
      }
      const int32_t r = (hit & (1u << g)) ? bucket_row(s[g], sub[g]) : -1;
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

// ---- row-structured kernels (hyper-cubes, dilation 1 along x) ----------------
// Offsets k = j*KX + i with dx = dx0 + i and (dy, dz) = column j: the KX cells of
// one kernel row lie in at most two consecutive runs, i.e. two buckets A and B
// whose 8 row ids form a window the cells index as (cx0 & 3) + i.  A probe ends
// on the bucket with the run's key or on an empty bucket, and an empty bucket's
// rows already read -1, so no hit flag is carried.  Out-of-range neighbours use
// a key with the reserved batch 65535, which matches nothing.
struct KRows {
  int32_t dx0, J;
  int32_t dyz[2 * kMaxK];  // (dy, dz) of column j
};

constexpr uint32_t kNoKeyHi = 0xFFFF0000u;  // batch 65535: never inserted, and != empty with lo = 0

template <int KX, int CB, bool kCount>
__global__ void __launch_bounds__(kBuildThreads)
k_kmap_build_rows(const int32_t *__restrict__ out_coords, int64_t n_out,
                  int ncols, const Bucket *__restrict__ table, uint32_t mask,
                  const __grid_constant__ KRows rows, int32_t *__restrict__ nbr,
                  int64_t ld, unsigned long long *__restrict__ counts) {
  extern __shared__ int s_cnt[];  // [K] per-block pair counts
  const int J = rows.J;
  if (kCount) {
    for (int k = threadIdx.x; k < J * KX; k += blockDim.x) s_cnt[k] = 0;
    __syncthreads();
  }
  const int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = o < n_out;
  int b = 65535, x = 0, y = 0, z = 0;
  if (live) load_coord(out_coords, o, ncols, b, x, y, z);
  const bool b_ok = (unsigned)b < 65535u;
  const int cx0 = x + rows.dx0 + 32768;         // biased x of cell 0
  const int cx1 = cx0 + KX - 1;                 // ... of the last cell
  const uint32_t hiA = (b_ok && (unsigned)cx0 < 65536u) ? ((uint32_t)b << 16) | (uint32_t)(cx0 >> 2) : kNoKeyHi;
  const uint32_t hiB = (b_ok && (unsigned)cx1 < 65536u) ? ((uint32_t)b << 16) | (uint32_t)(cx1 >> 2) : kNoKeyHi;
  const bool two = (cx1 >> 2) != (cx0 >> 2);
  const uint32_t pA = hiA * kHashC1, pB = hiB * kHashC1;
  const int s0 = cx0 & 3;
  const int shift = __clz(mask);
  const int by = y + 32768, bz = z + 32768;

  for (int j0 = 0; j0 < J; j0 += CB) {
    uint32_t lo[CB], kA[CB], kB[CB], hA[CB], hB[CB];
    BucketRegs sA[CB], sB[CB];
    uint32_t pend = 0;
#pragma unroll
    for (int c = 0; c < CB; ++c) {
      const int j = j0 + c;
      if (j < J) {
        const int ny = by + rows.dyz[2 * j], nz = bz + rows.dyz[2 * j + 1];
        const bool ok = (unsigned)ny < 65536u && (unsigned)nz < 65536u;
        lo[c] = ok ? ((uint32_t)ny << 16) | (uint32_t)nz : 0u;
        kA[c] = ok ? hiA : kNoKeyHi;
        kB[c] = ok ? hiB : kNoKeyHi;
        const uint32_t pl = lo[c] * kHashC2;
        hA[c] = hash_mix((ok ? pA : kNoKeyHi * kHashC1) + pl) >> shift;
        sA[c] = ldg_bucket(table + hA[c]);
        pend |= 1u << (2 * c);
        if (two) {
          hB[c] = hash_mix((ok ? pB : kNoKeyHi * kHashC1) + pl) >> shift;
          sB[c] = ldg_bucket(table + hB[c]);
          pend |= 2u << (2 * c);
        }
      }
    }
    // all chains of the batch advance together
    while (pend) {
#pragma unroll
      for (int c = 0; c < CB; ++c) {
        if (pend & (1u << (2 * c))) {
          const bool fin = (sA[c].v[0] == lo[c] && sA[c].v[1] == kA[c]) ||
                           (sA[c].v[0] & sA[c].v[1]) == 0xFFFFFFFFu;
          if (fin) {
            pend &= ~(1u << (2 * c));
          } else {
            hA[c] = (hA[c] + 1) & mask;
            sA[c] = ldg_bucket(table + hA[c]);
          }
        }
        if (pend & (2u << (2 * c))) {
          const bool fin = (sB[c].v[0] == lo[c] && sB[c].v[1] == kB[c]) ||
                           (sB[c].v[0] & sB[c].v[1]) == 0xFFFFFFFFu;
          if (fin) {
            pend &= ~(2u << (2 * c));
          } else {
            hB[c] = (hB[c] + 1) & mask;
            sB[c] = ldg_bucket(table + hB[c]);
          }
        }
      }
    }
#pragma unroll
    for (int c = 0; c < CB; ++c) {
      const int j = j0 + c;
      if (j >= J) break;
      // window of 8 row ids (B = A when the row sits inside one run: then
      // s0 + KX <= 4 and B is never indexed), shifted down by s0
      uint32_t w[8];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        w[t] = sA[c].v[2 + t];
        w[4 + t] = two ? sB[c].v[2 + t] : kNoRow;
      }
#pragma unroll
      for (int t = 0; t < KX + 1 && t + 2 < 8; ++t) w[t] = (s0 & 2) ? w[t + 2] : w[t];
#pragma unroll
      for (int t = 0; t < KX; ++t) w[t] = (s0 & 1) ? w[t + 1] : w[t];
#pragma unroll
      for (int i = 0; i < KX; ++i) {
        const int k = j * KX + i;
        const int32_t r = (int32_t)w[i];
        if (o < ld) nbr[(int64_t)k * ld + o] = r;   // padding rows [n_out, ld) read as missing
        if (kCount) {
          const unsigned bal = __ballot_sync(0xffffffffu, r >= 0);
          if ((threadIdx.x & 31) == 0 && bal) atomicAdd(&s_cnt[k], __popc(bal));
        }
      }
    }
  }
  if (kCount) {
    __syncthreads();
    for (int k = threadIdx.x; k < J * KX; k += blockDim.x)
      if (s_cnt[k]) atomicAdd(&counts[k], (unsigned long long)s_cnt[k]);
  }
}

// offsets k = j*KX + i, dx = dx0 + i, (dy,dz) constant per j?  KX in [2,5]
static bool row_structure(const int32_t *offs, int K, int nd, KRows *out, int *kx_out) {
  if (nd < 1) return false;
  for (int KX = 5; KX >= 2; --KX) {
    if (K % KX) continue;
    const int J = K / KX;
    bool ok = true;
    for (int j = 0; j < J && ok; ++j)
      for (int i = 0; i < KX && ok; ++i) {
        const int32_t *d = offs + (size_t)(j * KX + i) * nd;
        const int32_t *d0 = offs + (size_t)(j * KX) * nd;
        ok = d[0] == offs[0] + i && (nd < 2 || d[1] == d0[1]) && (nd < 3 || d[2] == d0[2]);
      }
    if (!ok) continue;
    out->dx0 = offs[0];
    out->J = J;
    for (int j = 0; j < J; ++j) {
      const int32_t *d0 = offs + (size_t)(j * KX) * nd;
      out->dyz[2 * j] = nd > 1 ? d0[1] : 0;
      out->dyz[2 * j + 1] = nd > 2 ? d0[2] : 0;
    }
    *kx_out = KX;
    return true;
  }
  return false;
}

template <int KX>
static void launch_rows(const int32_t *out_coords, int64_t n_out, int ncols, const Bucket *table, uint32_t mask,
                        const KRows &rows, int32_t *nbr, int64_t ld, unsigned long long *counts, cudaStream_t st) {
  const unsigned blocks = (unsigned)ceil_div(ld, kBuildThreads);
  constexpr int CB = 3;
  if (counts)
    k_kmap_build_rows<KX, CB, true><<<blocks, kBuildThreads, sizeof(int) * rows.J * KX, st>>>(
        out_coords, n_out, ncols, table, mask, rows, nbr, ld, counts);
  else
    k_kmap_build_rows<KX, CB, false><<<blocks, kBuildThreads, 0, st>>>(
        out_coords, n_out, ncols, table, mask, rows, nbr, ld, counts);
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
  MK_CHECK_ARG(in_capacity >= 1024 && (in_capacity & (in_capacity - 1)) == 0,
               "capacity must be a power of two >= 1024");
  cudaStream_t st = (cudaStream_t)stream;
  if (counts_dev)
    MK_CUDA(cudaMemsetAsync(counts_dev, 0, sizeof(int64_t) * K, st));
  if (n_out == 0) return MK_OK;
  MK_CHECK_ARG(out_coords && in_table && nbr && offsets_host, "null pointer");
  KOffsets offs;
  memcpy(offs.v, offsets_host, sizeof(int32_t) * K * (ncols - 1));
  const unsigned blocks = (unsigned)ceil_div(ld, kBuildThreads);
  const size_t smem = counts_dev ? sizeof(int) * K : 0;
  const uint32_t mask = (uint32_t)(in_capacity / 2 - 1);  // buckets - 1
  static thread_local KRows rows;
  int kx = 0;
  if (row_structure(offsets_host, K, ncols - 1, &rows, &kx)) {
    const Bucket *tb = (const Bucket *)in_table;
    unsigned long long *cd = (unsigned long long *)counts_dev;
    switch (kx) {
      case 2: launch_rows<2>(out_coords, n_out, ncols, tb, mask, rows, nbr, ld, cd, st); break;
      case 3: launch_rows<3>(out_coords, n_out, ncols, tb, mask, rows, nbr, ld, cd, st); break;
      case 4: launch_rows<4>(out_coords, n_out, ncols, tb, mask, rows, nbr, ld, cd, st); break;
      default: launch_rows<5>(out_coords, n_out, ncols, tb, mask, rows, nbr, ld, cd, st); break;
    }
    MK_LAUNCH_CHECK();
    return MK_OK;
  }
  if (K % 9 == 0)
    k_kmap_build<9><<<blocks, kBuildThreads, smem, st>>>(
        out_coords, n_out, ncols, (const Bucket *)in_table, mask, offs, K, nbr,
        ld, (unsigned long long *)counts_dev);
  else
    k_kmap_build<8><<<blocks, kBuildThreads, smem, st>>>(
        out_coords, n_out, ncols, (const Bucket *)in_table, mask, offs, K, nbr,
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


// ---------------------------------------------------------------------------
// Octree split (generative transposed convolution with kernel == stride, extern.hpp:151-165
// with expand_coordinates; modules/vae_block.py:279-296): the K children of a parent are new
// by construction and the child at row i*K + k hangs under offset k of parent i, so the
// kernel map needs no probe: every form the convolution kernels read is written directly.
//   nbr         [K, ld_out]  nbr[k][o] = o / K if o % K == k else -1        (fine row -> parent)
//   inv         [K, ld_in]   inv[k][i] = i * K + k                          (parent -> child: dense)
//   order       [n_out]      tile order: column s = child (s % n_in) under offset s / n_in,
//   nbr_ordered [K, ld_out]  so a 128-row tile holds ONE offset (two at a segment boundary)
//                            instead of all K at 1/K fill
__global__ void k_kmap_children(int64_t n_in, int K, int64_t ld_out, int64_t ld_in, int32_t *__restrict__ nbr,
                                int32_t *__restrict__ inv, int32_t *__restrict__ order,
                                int32_t *__restrict__ nbr_ordered) {
  const int64_t n_out = n_in * K;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (int64_t e = t0; e < (int64_t)K * ld_out; e += stride) {
    const int k = (int)(e / ld_out);
    const int64_t o = e - (int64_t)k * ld_out;
    if (nbr) nbr[e] = (o < n_out && (int)(o % K) == k) ? (int32_t)(o / K) : -1;
    if (nbr_ordered) nbr_ordered[e] = (o < n_out && (int)(o / n_in) == k) ? (int32_t)(o % n_in) : -1;
  }
  if (inv)
    for (int64_t e = t0; e < (int64_t)K * ld_in; e += stride) {
      const int k = (int)(e / ld_in);
      const int64_t i = e - (int64_t)k * ld_in;
      inv[e] = i < n_in ? (int32_t)(i * K + k) : -1;
    }
  if (order)
    for (int64_t s = t0; s < n_out; s += stride) order[s] = (int32_t)((s % n_in) * K + s / n_in);
}

int mk_kmap_children(int64_t n_in, int K, int64_t ld_out, int64_t ld_in, int32_t *nbr, int32_t *inv, int32_t *order,
                     int32_t *nbr_ordered, void *stream) {
  MK_CHECK_ARG(n_in >= 0 && K >= 1 && K <= 128, "bad sizes");
  MK_CHECK_ARG(n_in * K < (int64_t)0x7fffffff, "too many children");
  MK_CHECK_ARG(ld_out >= n_in * K && ld_in >= n_in, "leading dimensions too small");
  const int64_t total = std::max((int64_t)K * ld_out, (int64_t)K * ld_in);
  if (total == 0) return MK_OK;
  const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(total, 256), (int64_t)num_sms() * 16);
  k_kmap_children<<<grid, 256, 0, (cudaStream_t)stream>>>(n_in, K, ld_out, ld_in, nbr, inv, order, nbr_ordered);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

}  // extern "C"
