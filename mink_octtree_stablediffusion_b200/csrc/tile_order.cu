// Tile ordering of a kernel map for the tensor-core convolution kernels.
//
// The tcgen05 kernels work on tiles of 128 consecutive table columns and skip a
// (tile, offset) pair only when NONE of the 128 rows has a neighbour under the
// offset; every other pair costs a full 16 KB gathered stage however few of its
// rows are real.  On surface-like data in the row order of the coordinate map
// (first occurrence in the input, ME's convention) a tile touches ~26.5 of the
// 27 offsets of a 3^3 kernel with 42 % of the gathered rows real.  Rows with the
// SAME neighbour pattern grouped into the same tiles touch ~16 offsets at 70 %.
//
// mk_kmap_tile_order computes, from the finished table, a permutation `order` of
// the table's columns (slot -> row) sorted by (hash of batch and 2^b-voxel block,
// neighbour pattern) and the table re-laid in slot order.  Blocks keep the gathers of
// concurrently processed tiles inside a few MB of the feature matrix (L2).  The
// engine-visible row order of the coordinate map is NOT changed: the convolution
// writes output row order[slot] (mk_conv_fwd_rows) and the weight gradient reads
// its dy rows through the same permutation.
//
// The sort itself is cub::DeviceRadixSort (the toolkit's library; a helper stage
// of the map build, not a hot kernel).
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace mk {

constexpr int kBlockBits = 13;

// key = [hash13(batch, x>>b, y>>b, z>>b) | pattern : min(K, 32)]: 40 significant bits at
// K = 27 = five 8-bit passes of the radix sort instead of eight.  Rows of one block always share
// the hash; distinct blocks that collide merely share tiles where their patterns
// agree, which costs a little locality and nothing else.  pattern = bit k set iff
// nbr[k, row] >= 0 (first 32 offsets), sorted by its rank in the Gray sequence.
__global__ void k_order_key(const int32_t *__restrict__ nbr, int64_t n, int64_t ld, int K,
                            const int32_t *__restrict__ coords, int ncols, int bl,
                            unsigned long long *__restrict__ keys, int32_t *__restrict__ rows) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  uint32_t pat = 0;
  const int kk = K < 32 ? K : 32;
  for (int k = 0; k < kk; ++k) pat |= (uint32_t)(__ldg(nbr + (int64_t)k * ld + r) >= 0) << k;
  uint32_t blk = 0;
  if (coords) {
    const int32_t *c = coords + r * ncols;
    blk = (uint32_t)__ldg(c) * 0x9E3779B1u;
    for (int d = 1; d < ncols; ++d) blk = (blk ^ (uint32_t)(__ldg(c + d) >> bl)) * 0x85EBCA6Bu;
    blk = (blk ^ (blk >> 15)) & ((1u << kBlockBits) - 1u);
  }
  // rank of the pattern along the reflected Gray sequence: neighbours in the sorted order
  // differ in one offset, so a tile that straddles several patterns touches fewer offsets
  // than in plain binary order (126 k -> 122.5 k non-empty (tile, offset) pairs at 1 M voxels)
  uint32_t gr = pat;
  gr ^= gr >> 1; gr ^= gr >> 2; gr ^= gr >> 4; gr ^= gr >> 8; gr ^= gr >> 16;
  keys[r] = ((unsigned long long)blk << kk) | gr;
  rows[r] = (int32_t)r;
}

// dst[k, slot] = src[k, order[slot]] for slot < n, -1 for the padding columns
__global__ void k_permute_table(const int32_t *__restrict__ src, const int32_t *__restrict__ order, int64_t n,
                                int64_t ld, int K, int32_t *__restrict__ dst) {
  const int64_t slot = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= ld) return;
  const int64_t r = slot < n ? (int64_t)__ldg(order + slot) : -1;
  // nine independent loads in flight per thread (each is a 32-byte sector for 4 useful bytes:
  // the pass is bound by L2 sector throughput and by latency, not by HBM)
  for (int k0 = 0; k0 < K; k0 += 9) {
    int32_t v[9];
#pragma unroll
    for (int j = 0; j < 9; ++j) v[j] = (r >= 0 && k0 + j < K) ? __ldg(src + (int64_t)(k0 + j) * ld + r) : -1;
#pragma unroll
    for (int j = 0; j < 9; ++j)
      if (k0 + j < K) dst[(int64_t)(k0 + j) * ld + slot] = v[j];
  }
}


static inline size_t align256o(size_t b) { return (b + 255) & ~(size_t)255; }

static size_t sort_temp_bytes(int64_t n) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const unsigned long long *)nullptr, (unsigned long long *)nullptr,
                                  (const int32_t *)nullptr, (int32_t *)nullptr, (int)n);
  return bytes;
}

}  // namespace mk

using namespace mk;

extern "C" int64_t mk_kmap_tile_order_workspace_bytes(int64_t n_rows) {
  if (n_rows <= 0) return 256;
  const size_t n = (size_t)n_rows;
  return (int64_t)(2 * align256o(n * 8) + align256o(n * 4) + align256o(sort_temp_bytes(n_rows)) + 256);
}

extern "C" int mk_kmap_tile_order(const int32_t *nbr, int64_t n_rows, int64_t ld, int K, const int32_t *coords,
                                  int ncols, int block_log2, int32_t *order, int32_t *nbr_ordered,
                                  void *workspace, void *stream) {
  MK_CHECK_ARG(K >= 1 && ld >= n_rows && n_rows >= 0, "bad K / ld / n_rows");
  MK_CHECK_ARG(n_rows < (int64_t)1 << 31, "too many rows");
  MK_CHECK_ARG(block_log2 >= 0 && block_log2 <= 15, "block_log2 out of range");
  MK_CHECK_ARG(coords == nullptr || ncols >= 1, "bad ncols");
  if (ld == 0) return MK_OK;
  MK_CHECK_ARG(nbr && nbr_ordered && (order || n_rows == 0) && workspace, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_rows > 0) {
    const size_t n = (size_t)n_rows;
    uint8_t *ws = (uint8_t *)workspace;
    unsigned long long *keys_in = (unsigned long long *)ws;
    unsigned long long *keys_out = (unsigned long long *)(ws + align256o(n * 8));
    int32_t *rows_in = (int32_t *)(ws + 2 * align256o(n * 8));
    void *temp = ws + 2 * align256o(n * 8) + align256o(n * 4);
    size_t temp_bytes = sort_temp_bytes(n_rows);
    k_order_key<<<(unsigned)((n_rows + 255) / 256), 256, 0, st>>>(nbr, n_rows, ld, K, coords, ncols, block_log2,
                                                                   keys_in, rows_in);
    MK_LAUNCH_CHECK();
    MK_CUDA(cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, rows_in, order, (int)n_rows, 0, (K < 32 ? K : 32) + kBlockBits,
                                            st));
  }
  k_permute_table<<<(unsigned)((ld + 255) / 256), 256, 0, st>>>(nbr, order, n_rows, ld, K, nbr_ordered);
  MK_LAUNCH_CHECK();
  return MK_OK;
}
