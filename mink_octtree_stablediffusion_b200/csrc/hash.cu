// Coordinate hash: insert / dedup / compact / lookup, coordinate generation,
// flagged selection and row gather/scatter.
//
// Replaces (reference, sources absent — behaviour from pybind/extern.hpp:779-806
// and the kernel inventory in SURVEY.md §2.3): init_hashtbl,
// insert_and_map_kernel, remap / remap_inverse_map, thrust copy_if + cub sort,
// find_coordinate, stride_copy, kernel_region_insert, origin_map_kernel,
// prune_copy_and_insert.
//
// Design: 32-byte buckets {u64 run key, u32 row[4]} (common.cuh), one aligned
// sector each, linear probing over buckets, <= 50 % bucket load.  A "slot" is
// bucket*4 + (x&3).  Duplicates are resolved deterministically (atomicMin on the
// row => first occurrence wins), and compaction is an ordered block scan built
// from warp ballots, so unique_index is ascending without any sort.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace mk {

static thread_local char g_err[512] = "";
unsigned long long g_launch_count = 0;

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static unsigned int *g_hang_host = nullptr, *g_hang_dev = nullptr;

unsigned int *hang_report_device_ptr() {
  static bool tried = false;
  if (!tried) {
    tried = true;
    void *h = nullptr, *d = nullptr;
    if (cudaHostAlloc(&h, 16, cudaHostAllocMapped) == cudaSuccess &&
        cudaHostGetDevicePointer(&d, h, 0) == cudaSuccess) {
      g_hang_host = (unsigned int *)h;
      g_hang_dev = (unsigned int *)d;
      memset(h, 0, 16);
    } else {
      (void)cudaGetLastError();
    }
  }
  return g_hang_dev;
}

}  // namespace mk
// development aid: the watchdog's last report {magic, block, warp, (smem addr << 1) | parity}
extern "C" int mk_debug_hang_words(unsigned int *out4) {
  if (!mk::g_hang_host) return 1;
  for (int i = 0; i < 4; ++i) out4[i] = mk::g_hang_host[i];
  return 0;
}
namespace mk {
int cuda_fail(cudaError_t e, const char *what) {
  if (g_hang_host && g_hang_host[0] == 0xDEADBA44u)
    set_error("CUDA error %d (%s) at %s; pipeline wait timed out in block (%u,%u) warp %u on mbarrier smem 0x%x parity %u",
              (int)e, cudaGetErrorString(e), what, g_hang_host[1] & 0xFFFFu, g_hang_host[1] >> 16, g_hang_host[2],
              g_hang_host[3] >> 1, g_hang_host[3] & 1u);
  else
    set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return MK_ECUDA;
}

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;  // rows per block

// ---------------------------------------------------------------------------
__global__ void k_table_clear(uint4 *__restrict__ table, int64_t n16) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16;
       i += stride)
    table[i] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
}

__device__ __forceinline__ uint32_t *slot_row(Bucket *table, int32_t slot) {
  return &table[slot >> 2].row[slot & 3];
}

__global__ void k_insert(const int32_t *__restrict__ coords, int64_t n,
                         int ncols, Bucket *__restrict__ table, uint32_t mask,
                         int32_t *__restrict__ slot_of_row,
                         int32_t *__restrict__ status) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b, x, y, z;
  load_coord(coords, i, ncols, b, x, y, z);
  if (!coord_in_range(b, x, y, z)) {
    atomicMax(status, MK_ERANGE);
    slot_of_row[i] = -1;
    return;
  }
  const unsigned long long key = run_key(b, x, y, z);
  uint32_t h = bucket_home(key, mask);
  for (uint32_t probes = 0; probes <= mask; ++probes) {
    // cheap pre-check avoids most CAS (a bucket's key never changes once set,
    // so a stale read can only show "empty", which the CAS then corrects)
    unsigned long long old = *(volatile unsigned long long *)&table[h].key;
    if (old == kEmptyKey) old = atomicCAS(&table[h].key, kEmptyKey, key);
    if (old == kEmptyKey || old == key) {
      atomicMin(&table[h].row[x & 3], (uint32_t)i);
      slot_of_row[i] = (int32_t)(h * 4u + (uint32_t)(x & 3));
      return;
    }
    h = (h + 1) & mask;
  }
  atomicMax(status, MK_EFULL);
  slot_of_row[i] = -1;
}

// first_of_row[i] = original row that owns i's key; block_sums[b] = #firsts
__global__ void __launch_bounds__(kScanThreads)
k_first_count(Bucket *__restrict__ table,
              const int32_t *__restrict__ slot_of_row, int64_t n,
              int32_t *__restrict__ first_of_row,
              int32_t *__restrict__ block_sums) {
  __shared__ int warp_cnt[kScanThreads / 32];
  const int64_t base = (int64_t)blockIdx.x * kScanTile;
  int cnt = 0;
#pragma unroll
  for (int r = 0; r < kScanItems; ++r) {
    const int64_t i = base + r * kScanThreads + threadIdx.x;
    bool first = false;
    if (i < n) {
      const int32_t s = slot_of_row[i];
      const int32_t f = s >= 0 ? (int32_t)*slot_row(table, s) : -1;
      first_of_row[i] = f;
      first = (f == (int32_t)i);
    }
    cnt += __popc(__ballot_sync(0xffffffffu, first));
  }
  if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < kScanThreads / 32; ++w) t += warp_cnt[w];
    block_sums[blockIdx.x] = t;
  }
}

// exclusive scan of block_sums (single block), total -> *total_dev (int64)
__global__ void __launch_bounds__(1024)
k_scan_block_sums(int32_t *__restrict__ block_sums_all, int64_t nblocks,
                  int64_t *__restrict__ total_dev) {
  // one CTA per batch entry (blockIdx.x); each scans its own row of sums
  int32_t *__restrict__ block_sums = block_sums_all + (int64_t)blockIdx.x * nblocks;
  __shared__ int warp_tot[32];
  __shared__ int chunk_total;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long carry = 0;  // identical in every thread
  for (int64_t base = 0; base < nblocks; base += 1024) {
    const int64_t i = base + threadIdx.x;
    const int v = i < nblocks ? block_sums[i] : 0;
    int incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      const int w = warp_tot[lane];
      int wincl = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, wincl, d);
        if (lane >= d) wincl += t;
      }
      warp_tot[lane] = wincl - w;  // exclusive warp offsets
      if (lane == 31) chunk_total = wincl;
    }
    __syncthreads();
    if (i < nblocks) block_sums[i] = (int32_t)(carry + incl - v + warp_tot[warp]);
    carry += chunk_total;
    __syncthreads();
  }
  if (threadIdx.x == 0) total_dev[blockIdx.x] = carry;
}

int scan_block_sums(int32_t *block_sums, int64_t nblocks, int batch,
                    int64_t *total_dev, cudaStream_t st) {
  k_scan_block_sums<<<batch, 1024, 0, st>>>(block_sums, nblocks, total_dev);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

// ordered compaction of first rows
__global__ void __launch_bounds__(kScanThreads)
k_compact_first(const int32_t *__restrict__ coords, int ncols,
                const int32_t *__restrict__ first_of_row, int64_t n,
                const int32_t *__restrict__ block_offsets,
                int32_t *__restrict__ newid_of_first,
                int32_t *__restrict__ unique_index,
                int32_t *__restrict__ coords_out) {
  __shared__ int warp_cnt[kScanThreads / 32];
  __shared__ int running_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) running_s = block_offsets[blockIdx.x];
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * kScanTile;
#pragma unroll 1
  for (int r = 0; r < kScanItems; ++r) {
    const int64_t i = base + r * kScanThreads + threadIdx.x;
    const bool first = i < n && first_of_row[i] == (int32_t)i;
    const unsigned bal = __ballot_sync(0xffffffffu, first);
    if (lane == 0) warp_cnt[warp] = __popc(bal);
    __syncthreads();
    int woff = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; ++w) {
      const int c = warp_cnt[w];
      if (w < warp) woff += c;
      tot += c;
    }
    const int running = running_s;
    if (first) {
      const int id = running + woff + __popc(bal & ((1u << lane) - 1u));
      newid_of_first[i] = id;
      if (unique_index) unique_index[id] = (int32_t)i;
      if (coords_out) {
        int b, x, y, z;
        load_coord(coords, i, ncols, b, x, y, z);
        store_coord(coords_out, id, ncols, b, x, y, z);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) running_s = running + tot;
    __syncthreads();
  }
}

__global__ void k_finalize(Bucket *__restrict__ table,
                           const int32_t *__restrict__ slot_of_row,
                           const int32_t *__restrict__ first_of_row,
                           const int32_t *__restrict__ newid_of_first,
                           int64_t n, int32_t *__restrict__ inverse_map) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t f = first_of_row[i];
  if (f < 0) {
    if (inverse_map) inverse_map[i] = -1;
    return;
  }
  const int32_t id = newid_of_first[f];
  if (inverse_map) inverse_map[i] = id;
  if (f == (int32_t)i) *slot_row(table, slot_of_row[i]) = (uint32_t)id;
}

__global__ void k_lookup(const Bucket *__restrict__ table, uint32_t mask,
                         const int32_t *__restrict__ query, int64_t nq,
                         int ncols, int32_t *__restrict__ rows) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  int b, x, y, z;
  load_coord(query, i, ncols, b, x, y, z);
  rows[i] = coord_in_range(b, x, y, z)
                ? table_find(table, mask, b, x, y, z)
                : -1;
}

// ---------------------------------------------------------------------------
struct Int3 { int v[3]; };

__device__ __forceinline__ int floor_div_mul(int c, int s) {
  int q = c / s;
  if ((c % s != 0) && ((c < 0) != (s < 0))) --q;
  return q * s;
}

__global__ void k_coords_stride(const int32_t *__restrict__ in, int64_t n,
                                int ncols, Int3 ts, int32_t *__restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b, x, y, z;
  load_coord(in, i, ncols, b, x, y, z);
  x = floor_div_mul(x, ts.v[0]);
  if (ncols > 2) y = floor_div_mul(y, ts.v[1]);
  if (ncols > 3) z = floor_div_mul(z, ts.v[2]);
  store_coord(out, i, ncols, b, x, y, z);
}

__global__ void k_coords_origin(const int32_t *__restrict__ in, int64_t n,
                                int ncols, int32_t *__restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  store_coord(out, i, ncols, in[i * ncols], 0, 0, 0);
}

// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kScanThreads)
k_flag_count(const uint8_t *__restrict__ keep, int64_t n,
             int32_t *__restrict__ block_sums) {
  __shared__ int warp_cnt[kScanThreads / 32];
  const int64_t base = (int64_t)blockIdx.x * kScanTile;
  int cnt = 0;
#pragma unroll
  for (int r = 0; r < kScanItems; ++r) {
    const int64_t i = base + r * kScanThreads + threadIdx.x;
    cnt += __popc(__ballot_sync(0xffffffffu, i < n && keep[i] != 0));
  }
  if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < kScanThreads / 32; ++w) t += warp_cnt[w];
    block_sums[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(kScanThreads)
k_flag_compact(const uint8_t *__restrict__ keep, int64_t n,
               const int32_t *__restrict__ block_offsets,
               int32_t *__restrict__ kept) {
  __shared__ int warp_cnt[kScanThreads / 32];
  __shared__ int running_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) running_s = block_offsets[blockIdx.x];
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * kScanTile;
#pragma unroll 1
  for (int r = 0; r < kScanItems; ++r) {
    const int64_t i = base + r * kScanThreads + threadIdx.x;
    const bool f = i < n && keep[i] != 0;
    const unsigned bal = __ballot_sync(0xffffffffu, f);
    if (lane == 0) warp_cnt[warp] = __popc(bal);
    __syncthreads();
    int woff = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; ++w) {
      const int c = warp_cnt[w];
      if (w < warp) woff += c;
      tot += c;
    }
    const int running = running_s;
    if (f) kept[running + woff + __popc(bal & ((1u << lane) - 1u))] = (int32_t)i;
    __syncthreads();
    if (threadIdx.x == 0) running_s = running + tot;
    __syncthreads();
  }
}

// rows as 16-byte / 4-byte / 1-byte units; one warp-slice per row chunk
template <typename V>
__global__ void k_rows_gather(const V *__restrict__ src,
                              const int32_t *__restrict__ idx, int64_t n_idx,
                              int units, V *__restrict__ dst, bool scatter) {
  const int64_t total = n_idx * units;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += stride) {
    const int64_t j = t / units;
    const int u = (int)(t - j * units);
    const int64_t r = idx[j];
    if (scatter)
      dst[r * units + u] = src[j * units + u];
    else
      dst[j * units + u] = src[r * units + u];
  }
}

static int rows_copy(const void *src, const int32_t *idx, int64_t n_idx,
                     int64_t row_bytes, void *dst, bool scatter,
                     cudaStream_t st) {
  if (n_idx == 0 || row_bytes == 0) return MK_OK;
  const bool a16 = row_bytes % 16 == 0 && ((uintptr_t)src % 16 == 0) &&
                   ((uintptr_t)dst % 16 == 0);
  const bool a4 = row_bytes % 4 == 0 && ((uintptr_t)src % 4 == 0) &&
                  ((uintptr_t)dst % 4 == 0);
  const int unit = a16 ? 16 : (a4 ? 4 : 1);
  const int units = (int)(row_bytes / unit);
  const int64_t total = n_idx * units;
  const int threads = 256;
  const int blocks = (int)std::min<int64_t>(ceil_div(total, threads),
                                            (int64_t)num_sms() * 32);
  if (unit == 16)
    k_rows_gather<uint4><<<blocks, threads, 0, st>>>(
        (const uint4 *)src, idx, n_idx, units, (uint4 *)dst, scatter);
  else if (unit == 4)
    k_rows_gather<uint32_t><<<blocks, threads, 0, st>>>(
        (const uint32_t *)src, idx, n_idx, units, (uint32_t *)dst, scatter);
  else
    k_rows_gather<uint8_t><<<blocks, threads, 0, st>>>(
        (const uint8_t *)src, idx, n_idx, units, (uint8_t *)dst, scatter);
  MK_LAUNCH_CHECK();
  return MK_OK;
}


// ---------------------------------------------------------------------------
// mk_map_insert2: the same result as mk_map_insert in three launches (four with the
// inverse map) and no memset:
//   k_table_clear2   table + scan state + status words
//   k_insert         (as above)
//   k_dedup_compact  first-occurrence flags, ONE-pass ordered compaction (decoupled look-back
//                    over 2048-row blocks handed out by a ticket), coordinates / unique index
//                    out, compact ids written into the table, unique count by the last block
//   k_inverse_map    inverse_map[i] = id in the slot of row i (only when asked for)
// Why the slot update is race-free inside the compaction pass: a slot holds the first row f
// of its coordinate until f's own thread replaces it by f's compact id, and id(f) <= f < i
// for every other row i of that coordinate, so "slot == i" stays false for them whichever
// value they read.  With MK_INSERT_UNIQUE the caller guarantees distinct rows (children of
// distinct octree parents, pruned rows of a map): ids are the row numbers, no dedup pass.
constexpr unsigned kLbAggregate = 1u << 30, kLbPrefix = 2u << 30, kLbMask = (1u << 30) - 1u;

__global__ void k_table_clear2(uint4 *__restrict__ table, int64_t n16, unsigned *__restrict__ state, int64_t nstate,
                               int32_t *__restrict__ status, int64_t *__restrict__ n_unique) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (int64_t i = t0; i < n16; i += stride)
    table[i] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
  for (int64_t i = t0; i < nstate; i += stride) state[i] = 0u;
  if (t0 == 0) { *status = 0; *n_unique = 0; }
}

// rows distinct by contract: slot row = row number
__global__ void k_insert_unique(const int32_t *__restrict__ coords, int64_t n, int ncols, Bucket *__restrict__ table,
                                uint32_t mask, int32_t *__restrict__ status) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b, x, y, z;
  load_coord(coords, i, ncols, b, x, y, z);
  if (!coord_in_range(b, x, y, z)) { atomicMax(status, MK_ERANGE); return; }
  const unsigned long long key = run_key(b, x, y, z);
  uint32_t h = bucket_home(key, mask);
  for (uint32_t probes = 0; probes <= mask; ++probes) {
    unsigned long long old = *(volatile unsigned long long *)&table[h].key;
    if (old == kEmptyKey) old = atomicCAS(&table[h].key, kEmptyKey, key);
    if (old == kEmptyKey || old == key) {
      table[h].row[x & 3] = (uint32_t)i;
      return;
    }
    h = (h + 1) & mask;
  }
  atomicMax(status, MK_EFULL);
}

template <typename IdxT>
__global__ void __launch_bounds__(kScanThreads)
k_dedup_compact(const int32_t *__restrict__ coords, int ncols, Bucket *__restrict__ table,
                const int32_t *__restrict__ slot_of_row, int64_t n, unsigned *__restrict__ state /* [0] = ticket */,
                IdxT *__restrict__ unique_index, int32_t *__restrict__ coords_out, int64_t *__restrict__ n_unique) {
  __shared__ int warp_cnt[kScanItems][kScanThreads / 32];
  __shared__ unsigned s_block, s_prefix;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_block = atomicAdd(&state[0], 1u);      // blocks enter the scan in ticket order
  __syncthreads();
  const unsigned blk = s_block;
  unsigned *lb = state + 1;
  const int64_t base = (int64_t)blk * kScanTile;
  bool first[kScanItems];
  unsigned bal[kScanItems];
#pragma unroll
  for (int r = 0; r < kScanItems; ++r) {
    const int64_t i = base + r * kScanThreads + threadIdx.x;
    first[r] = false;
    if (i < n) {
      const int32_t s = slot_of_row[i];
      first[r] = s >= 0 && *(volatile uint32_t *)slot_row(table, s) == (uint32_t)i;
    }
    bal[r] = __ballot_sync(0xffffffffu, first[r]);
    if (lane == 0) warp_cnt[r][warp] = __popc(bal[r]);
  }
  __syncthreads();
  // block aggregate -> published; exclusive prefix by looking back over the predecessors
  if (warp == 0) {
    unsigned tot = 0;
    for (int e = lane; e < kScanItems * (kScanThreads / 32); e += 32) tot += (unsigned)(&warp_cnt[0][0])[e];
    tot = __reduce_add_sync(0xffffffffu, tot);
    if (lane == 0) {
      __threadfence();
      atomicExch(&lb[blk], (blk == 0 ? kLbPrefix : kLbAggregate) | tot);
    }
    unsigned excl = 0;
    int64_t look = (int64_t)blk - 1;
    while (look >= 0) {
      const int64_t j = look - lane;
      unsigned v = kLbPrefix;                       // lanes before block 0: a zero prefix
      if (j >= 0) {
        do { v = *(volatile unsigned *)&lb[j]; } while ((v & ~kLbMask) == 0u);
      }
      const unsigned is_prefix = __ballot_sync(0xffffffffu, (v & ~kLbMask) == kLbPrefix);
      const int stop = is_prefix ? __ffs((int)is_prefix) - 1 : 32;    // nearest predecessor holding a prefix
      const unsigned contrib = lane <= stop ? (v & kLbMask) : 0u;
      excl += __reduce_add_sync(0xffffffffu, contrib);
      if (is_prefix) break;
      look -= 32;
    }
    if (lane == 0) {
      if (blk != 0) { __threadfence(); atomicExch(&lb[blk], kLbPrefix | (excl + tot)); }
      s_prefix = excl;
      if ((int64_t)(blk + 1) * kScanTile >= n) *n_unique = (int64_t)(excl + tot);
    }
  }
  __syncthreads();
  unsigned running = s_prefix;
#pragma unroll
  for (int r = 0; r < kScanItems; ++r) {
    int woff = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; ++w) {
      const int c = warp_cnt[r][w];
      if (w < warp) woff += c;
      tot += c;
    }
    if (first[r]) {
      const int64_t i = base + r * kScanThreads + threadIdx.x;
      const unsigned id = running + (unsigned)woff + (unsigned)__popc(bal[r] & ((1u << lane) - 1u));
      if (unique_index) unique_index[id] = (IdxT)i;
      if (coords_out) {
        int b, x, y, z;
        load_coord(coords, i, ncols, b, x, y, z);
        store_coord(coords_out, id, ncols, b, x, y, z);
      }
      *slot_row(table, slot_of_row[i]) = id;
    }
    running += (unsigned)tot;
  }
}

template <typename IdxT>
__global__ void k_inverse_map(const Bucket *__restrict__ table, const int32_t *__restrict__ slot_of_row, int64_t n,
                              IdxT *__restrict__ inverse_map) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t s = slot_of_row[i];
  inverse_map[i] = s >= 0 ? (IdxT)table[s >> 2].row[s & 3] : (IdxT)-1;
}

__global__ void k_set_count(int64_t *p, int64_t v) { *p = v; }

template <typename IdxT>
__global__ void k_iota(IdxT *__restrict__ a, IdxT *__restrict__ b, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (a) a[i] = (IdxT)i;
  if (b) b[i] = (IdxT)i;
}
}  // namespace mk

using namespace mk;

extern "C" {

const char *mk_last_error(void) { return g_err; }
int mk_version(void) { return 100; }
int64_t mk_launch_count(void) { return (int64_t)g_launch_count; }

// capacity is counted in 16-byte units (table bytes = 16 * capacity): two per
// bucket, and >= 2n buckets so that even a cloud with one cell per run stays at
// <= 50 % bucket load
int64_t mk_hash_capacity(int64_t n) {
  int64_t cap = 1024;
  while (cap < 4 * n) cap <<= 1;
  return cap;
}

static inline int64_t align256(int64_t b) { return (b + 255) & ~(int64_t)255; }

int64_t mk_insert_workspace_bytes(int64_t n) {
  const int64_t nblk = ceil_div(n > 0 ? n : 1, kScanTile);
  return 3 * align256(n * 4) + align256((nblk + 1) * 4) + 256;
}

int mk_map_insert(const int32_t *coords, int64_t n, int ncols, void *table,
                  int64_t capacity, int32_t *unique_index,
                  int32_t *inverse_map, int32_t *coords_out, void *workspace,
                  int64_t *n_unique_dev, int32_t *status_dev, void *stream) {
  MK_CHECK_ARG(ncols >= 2 && ncols <= 4, "ncols must be in [2,4] (D <= 3)");
  MK_CHECK_ARG(n >= 0 && n < (int64_t)0x7fffffff, "row count out of range");
  MK_CHECK_ARG(capacity >= 4 * n && capacity >= 1024 &&
                   (capacity & (capacity - 1)) == 0 &&
                   capacity <= ((int64_t)1 << 30),
               "capacity must be a power of two >= max(4n, 1024)");
  MK_CHECK_ARG(table && n_unique_dev && status_dev && (workspace || n == 0),
               "null pointer");
  MK_CHECK_ARG((uintptr_t)table % 32 == 0, "table must be 32-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  MK_CUDA(cudaMemsetAsync(status_dev, 0, sizeof(int32_t), st));
  {
    const int threads = 256;
    const int blocks = (int)std::min<int64_t>(ceil_div(capacity, threads),
                                              (int64_t)num_sms() * 16);
    k_table_clear<<<blocks, threads, 0, st>>>((uint4 *)table, capacity);
    MK_LAUNCH_CHECK();
  }
  if (n == 0) {
    MK_CUDA(cudaMemsetAsync(n_unique_dev, 0, sizeof(int64_t), st));
    return MK_OK;
  }
  char *ws = (char *)workspace;
  int32_t *slot_of_row = (int32_t *)ws; ws += align256(n * 4);
  int32_t *first_of_row = (int32_t *)ws; ws += align256(n * 4);
  int32_t *newid = (int32_t *)ws; ws += align256(n * 4);
  int32_t *block_sums = (int32_t *)ws;
  const int64_t nblk = ceil_div(n, kScanTile);
  const uint32_t mask = (uint32_t)(capacity / 2 - 1);  // buckets - 1
  k_insert<<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(
      coords, n, ncols, (Bucket *)table, mask, slot_of_row, status_dev);
  MK_LAUNCH_CHECK();
  k_first_count<<<(unsigned)nblk, kScanThreads, 0, st>>>(
      (Bucket *)table, slot_of_row, n, first_of_row, block_sums);
  MK_LAUNCH_CHECK();
  k_scan_block_sums<<<1, 1024, 0, st>>>(block_sums, nblk, n_unique_dev);
  MK_LAUNCH_CHECK();
  k_compact_first<<<(unsigned)nblk, kScanThreads, 0, st>>>(
      coords, ncols, first_of_row, n, block_sums, newid, unique_index,
      coords_out);
  MK_LAUNCH_CHECK();
  k_finalize<<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(
      (Bucket *)table, slot_of_row, first_of_row, newid, n, inverse_map);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int64_t mk_insert2_workspace_bytes(int64_t n) {
  const int64_t nblk = ceil_div(n > 0 ? n : 1, kScanTile);
  return align256(n * 4) + align256((nblk + 2) * 4) + 256;
}

int mk_map_insert2(const int32_t *coords, int64_t n, int ncols, void *table, int64_t capacity, int flags,
                   void *unique_index, void *inverse_map, int32_t *coords_out, void *workspace,
                   int64_t *n_unique_dev, int32_t *status_dev, void *stream) {
  MK_CHECK_ARG(ncols >= 2 && ncols <= 4, "ncols must be in [2,4] (D <= 3)");
  MK_CHECK_ARG(n >= 0 && n < (int64_t)(1 << 30), "row count out of range");
  MK_CHECK_ARG(capacity >= 4 * n && capacity >= 1024 && (capacity & (capacity - 1)) == 0 &&
                   capacity <= ((int64_t)1 << 29),
               "capacity must be a power of two in [max(4n, 1024), 2^29]");
  MK_CHECK_ARG(table && n_unique_dev && status_dev && (workspace || n == 0), "null pointer");
  MK_CHECK_ARG((uintptr_t)table % 32 == 0, "table must be 32-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const bool unique = (flags & MK_INSERT_UNIQUE) != 0, idx64 = (flags & MK_INSERT_INT64_MAPS) != 0;
  char *ws = (char *)workspace;
  int32_t *slot_of_row = (int32_t *)ws;
  const int64_t nblk = ceil_div(n > 0 ? n : 1, kScanTile);
  unsigned *state = n > 0 ? (unsigned *)(ws + align256(n * 4)) : nullptr;
  {
    const int threads = 256;
    const int blocks = (int)std::min<int64_t>(ceil_div(capacity, threads), (int64_t)num_sms() * 16);
    k_table_clear2<<<blocks, threads, 0, st>>>((uint4 *)table, capacity, state, n > 0 ? nblk + 1 : 0, status_dev,
                                               n_unique_dev);
    MK_LAUNCH_CHECK();
  }
  if (n == 0) return MK_OK;
  const uint32_t mask = (uint32_t)(capacity / 2 - 1);
  const unsigned g256 = (unsigned)ceil_div(n, 256);
  if (unique) {
    k_insert_unique<<<g256, 256, 0, st>>>(coords, n, ncols, (Bucket *)table, mask, status_dev);
    MK_LAUNCH_CHECK();
    if (coords_out && coords_out != coords)
      MK_CUDA(cudaMemcpyAsync(coords_out, coords, (size_t)n * ncols * 4, cudaMemcpyDeviceToDevice, st));
    if (unique_index || inverse_map) {
      if (idx64) k_iota<int64_t><<<g256, 256, 0, st>>>((int64_t *)unique_index, (int64_t *)inverse_map, n);
      else k_iota<int32_t><<<g256, 256, 0, st>>>((int32_t *)unique_index, (int32_t *)inverse_map, n);
      MK_LAUNCH_CHECK();
    }
    k_set_count<<<1, 1, 0, st>>>(n_unique_dev, n);
    MK_LAUNCH_CHECK();
    return MK_OK;
  }
  k_insert<<<g256, 256, 0, st>>>(coords, n, ncols, (Bucket *)table, mask, slot_of_row, status_dev);
  MK_LAUNCH_CHECK();
  if (idx64)
    k_dedup_compact<int64_t><<<(unsigned)nblk, kScanThreads, 0, st>>>(coords, ncols, (Bucket *)table, slot_of_row, n, state,
                                                                      (int64_t *)unique_index, coords_out, n_unique_dev);
  else
    k_dedup_compact<int32_t><<<(unsigned)nblk, kScanThreads, 0, st>>>(coords, ncols, (Bucket *)table, slot_of_row, n, state,
                                                                      (int32_t *)unique_index, coords_out, n_unique_dev);
  MK_LAUNCH_CHECK();
  if (inverse_map) {
    if (idx64) k_inverse_map<int64_t><<<g256, 256, 0, st>>>((const Bucket *)table, slot_of_row, n, (int64_t *)inverse_map);
    else k_inverse_map<int32_t><<<g256, 256, 0, st>>>((const Bucket *)table, slot_of_row, n, (int32_t *)inverse_map);
    MK_LAUNCH_CHECK();
  }
  return MK_OK;
}

int mk_map_lookup(const void *table, int64_t capacity, const int32_t *query,
                  int64_t nq, int ncols, int32_t *rows, void *stream) {
  MK_CHECK_ARG(ncols >= 2 && ncols <= 4, "ncols must be in [2,4]");
  MK_CHECK_ARG(capacity >= 1024 && (capacity & (capacity - 1)) == 0,
               "capacity must be a power of two >= 1024");
  if (nq == 0) return MK_OK;
  MK_CHECK_ARG(table && query && rows, "null pointer");
  k_lookup<<<(unsigned)ceil_div(nq, 256), 256, 0, (cudaStream_t)stream>>>(
      (const Bucket *)table, (uint32_t)(capacity / 2 - 1), query, nq, ncols,
      rows);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int mk_coords_stride(const int32_t *in, int64_t n, int ncols,
                     const int32_t *new_stride_host, int32_t *out,
                     void *stream) {
  MK_CHECK_ARG(ncols >= 2 && ncols <= 4, "ncols must be in [2,4]");
  if (n == 0) return MK_OK;
  Int3 ts = {{1, 1, 1}};
  for (int d = 0; d < ncols - 1; ++d) {
    MK_CHECK_ARG(new_stride_host[d] > 0, "stride must be positive");
    ts.v[d] = new_stride_host[d];
  }
  k_coords_stride<<<(unsigned)ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(
      in, n, ncols, ts, out);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

// small per-call offset tables live in constant-sized device staging memory
// owned by the caller-visible API: we pass them by kernel argument instead.
struct OffsArg { int32_t v[3 * 128]; };

__global__ void k_coords_expand_arg(const int32_t *__restrict__ in, int64_t n,
                                    int ncols, OffsArg offs, int K,
                                    int32_t *__restrict__ out) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * K) return;
  const int64_t i = t / K;
  const int k = (int)(t - i * K);
  int b, x, y, z;
  load_coord(in, i, ncols, b, x, y, z);
  const int32_t *o = offs.v + k * (ncols - 1);
  x += o[0];
  if (ncols > 2) y += o[1];
  if (ncols > 3) z += o[2];
  store_coord(out, t, ncols, b, x, y, z);
}

int mk_coords_expand(const int32_t *in, int64_t n, int ncols,
                     const int32_t *offsets_host, int K, int32_t *out,
                     void *stream) {
  MK_CHECK_ARG(ncols >= 2 && ncols <= 4, "ncols must be in [2,4]");
  MK_CHECK_ARG(K >= 1 && K <= 128, "K must be in [1,128] for expansion");
  if (n == 0) return MK_OK;
  OffsArg offs;
  memcpy(offs.v, offsets_host, sizeof(int32_t) * K * (ncols - 1));
  k_coords_expand_arg<<<(unsigned)ceil_div(n * K, 256), 256, 0,
                        (cudaStream_t)stream>>>(in, n, ncols, offs, K, out);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int mk_coords_origin(const int32_t *in, int64_t n, int ncols, int32_t *out,
                     void *stream) {
  MK_CHECK_ARG(ncols >= 2 && ncols <= 4, "ncols must be in [2,4]");
  if (n == 0) return MK_OK;
  k_coords_origin<<<(unsigned)ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(
      in, n, ncols, out);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int64_t mk_select_workspace_bytes(int64_t n) {
  return align256((ceil_div(n > 0 ? n : 1, kScanTile) + 1) * 4);
}

int mk_select_flagged(const uint8_t *keep, int64_t n, int32_t *kept,
                      void *workspace, int64_t *n_kept_dev, void *stream) {
  MK_CHECK_ARG(n >= 0 && n < (int64_t)0x7fffffff, "row count out of range");
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    MK_CUDA(cudaMemsetAsync(n_kept_dev, 0, sizeof(int64_t), st));
    return MK_OK;
  }
  MK_CHECK_ARG(keep && kept && workspace && n_kept_dev, "null pointer");
  int32_t *block_sums = (int32_t *)workspace;
  const int64_t nblk = ceil_div(n, kScanTile);
  k_flag_count<<<(unsigned)nblk, kScanThreads, 0, st>>>(keep, n, block_sums);
  MK_LAUNCH_CHECK();
  k_scan_block_sums<<<1, 1024, 0, st>>>(block_sums, nblk, n_kept_dev);
  MK_LAUNCH_CHECK();
  k_flag_compact<<<(unsigned)nblk, kScanThreads, 0, st>>>(keep, n, block_sums,
                                                          kept);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

// ---------------------------------------------------------------------------
// mk_select_flagged_static: mk_select_flagged without a caller workspace and
// without any allocation.  The block-sum scratch is the file-scope device array
// below (one copy per GPU that loads the library, held until exit).  The array's
// NAME is only a host-side stand-in in host code; the device address is fetched
// once per device with cudaGetSymbolAddress (that device current) and cached.
// Nothing in the scratch has to start at zero: k_flag_count writes every entry
// [0, nblk) before k_scan_block_sums / k_flag_compact read it, so the previous
// call's contents are simply overwritten.  All calls share the array: two calls
// that overlap on the GPU (different streams) race on it.
constexpr int64_t kStaticSelectMaxBlocks = 8192;
constexpr int64_t kStaticSelectMaxRows = kStaticSelectMaxBlocks * kScanTile;  // 16 777 216
static __device__ __align__(256) int32_t g_select_scratch[kStaticSelectMaxBlocks + 64];

constexpr int kMaxDevices = 64;
static int32_t *g_select_scratch_ptr[kMaxDevices];  // device address per device, or null

int64_t mk_select_static_max_rows(void) { return kStaticSelectMaxRows; }

int mk_select_flagged_static(const uint8_t *keep, int64_t n, int32_t *kept,
                             int64_t *n_kept_dev, void *stream) {
  MK_CHECK_ARG(n >= 0, "row count out of range");
  if (n > kStaticSelectMaxRows) {
    set_error("mk_select_flagged_static: n = %lld exceeds the static scratch limit of %lld rows; "
              "use mk_select_flagged with a caller workspace",
              (long long)n, (long long)kStaticSelectMaxRows);
    return MK_EINVAL;
  }
  MK_CHECK_ARG(n_kept_dev, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    MK_CUDA(cudaMemsetAsync(n_kept_dev, 0, sizeof(int64_t), st));
    return MK_OK;
  }
  MK_CHECK_ARG(keep && kept, "null pointer");
  int dev = -1;
  MK_CUDA(cudaGetDevice(&dev));
  MK_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "device ordinal out of range");
  int32_t *block_sums = __atomic_load_n(&g_select_scratch_ptr[dev], __ATOMIC_ACQUIRE);
  if (!block_sums) {
    void *p = nullptr;
    MK_CUDA(cudaGetSymbolAddress(&p, g_select_scratch));
    MK_CHECK_ARG(p != nullptr, "cudaGetSymbolAddress returned null");
    block_sums = (int32_t *)p;
    __atomic_store_n(&g_select_scratch_ptr[dev], block_sums, __ATOMIC_RELEASE);
  }
  const int64_t nblk = ceil_div(n, kScanTile);
  k_flag_count<<<(unsigned)nblk, kScanThreads, 0, st>>>(keep, n, block_sums);
  MK_LAUNCH_CHECK();
  k_scan_block_sums<<<1, 1024, 0, st>>>(block_sums, nblk, n_kept_dev);
  MK_LAUNCH_CHECK();
  k_flag_compact<<<(unsigned)nblk, kScanThreads, 0, st>>>(keep, n, block_sums,
                                                          kept);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int mk_rows_gather(const void *src, const int32_t *idx, int64_t n_idx,
                   int64_t row_bytes, void *dst, void *stream) {
  return rows_copy(src, idx, n_idx, row_bytes, dst, false,
                   (cudaStream_t)stream);
}

int mk_rows_scatter(const void *src, const int32_t *idx, int64_t n_idx,
                    int64_t row_bytes, void *dst, void *stream) {
  return rows_copy(src, idx, n_idx, row_bytes, dst, true,
                   (cudaStream_t)stream);
}

}  // extern "C"
