// Tensor-core (tcgen05 / TMEM) sparse convolution, forward and data gradient.
//
// Replaces the matmul<T,uint,TILE> gather-GEMM-scatter kernels behind
// ConvolutionForwardGPU / ConvolutionBackwardGPU (pybind/extern.hpp:84-114;
// SURVEY §2.3): all kernel offsets in ONE persistent launch, output-stationary,
// accumulators in TMEM, no atomics.
//
//   y[o, n0:n0+NT] = sum_k  x[nbr[kk(k), o], :] @ B_k          (fp32 accumulate)
//
// CTA roles (736 threads, 1 CTA / SM, all 512 TMEM columns):
//   warps 0-3   epilogue : tcgen05.ld accumulators -> bf16/f16 -> global
//   warp  4     indices  : one lane bulk-copies, per offset, the T*128 neighbour-
//                          table entries of the group into a ring in shared memory
//   warp  5     weights  : one lane bulk-copies the pre-swizzled B image of each
//                          (offset, channel chunk) into a ring
//   warp  6     MMA      : driven by tickets; one elected lane issues tcgen05.mma
//                          (M=128, N=NT, K=16) from uniform registers
//   warps 7-22  gather   : four teams of four warps; LDGSTS 16-byte chunks of the
//                          neighbour rows into the A stage (128B swizzle), missing
//                          neighbours zero-filled
// A CTA owns "groups" of T 128-row output tiles whose accumulators live side by
// side in one half of TMEM; the two halves ping-pong so the epilogue of one
// group overlaps the main loop of the next.  One A stage holds the T gathered
// tiles of the group under one (offset, channel chunk) - a coarse stage keeps the
// number of barrier hand-offs per MMA low, which is what bounds this kernel.
// Which (tile, offset) pairs have any neighbour at all is known up front from
// the kernel map's tile mask, so every role walks the same static schedule and
// empty pairs are skipped.  The gather teams, which walk the schedule anyway,
// hand the MMA warp a "ticket" per stage (tiles present, accumulate flags, group
// begin / end) so that the serial MMA warp does no bookkeeping.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "common.cuh"
#include "sm100.cuh"
#include "tmap.cuh"

namespace mk {
using namespace sm100;

constexpr int kTileM = 128;
constexpr int kTileBytes = kTileM * 128;   // one gathered tile: 128 rows x 64 16-bit channels
// one A stage = the T tiles of a group under one (offset, channel chunk): T * kTileBytes
constexpr int kGatherTeams = 4;              // stages in flight on the producer side
constexpr int kGatherWarps = 4 * kGatherTeams;   // 4 warps (32 rows each) per stage
#ifndef MK_TC_MMA_WARPS
#define MK_TC_MMA_WARPS 2
#endif
constexpr int kMmaWarps = MK_TC_MMA_WARPS;   // tile t of a stage belongs to MMA warp t % kMmaWarps.  Every MMA warp waits on
                               // every stage and commits to its release barrier (count kMmaWarps), so stages are
                               // still released in order: the producers' parity waits stay alias-free
constexpr int kWarpIdx = 4, kWarpWeights = 5, kWarpMma0 = 6, kWarpGather0 = kWarpMma0 + kMmaWarps;
constexpr int kNumThreadsTC = 32 * (kWarpGather0 + kGatherWarps);
constexpr int kSchedCopies = 2 + kGatherWarps;
constexpr int kMaxT = 4;                   // tiles per group (256 TMEM columns / NT)
constexpr int kMaxIdxSlots = 8;            // ring of neighbour-index rows: one slot = the T*128 entries of a group under one offset
constexpr int kMaxBStages = 8;             // weight-image ring
constexpr int kMaxK = 128;

// -DMK_TC_PROF: per-CTA cycle counters of where each role waits (development aid;
// read back with mk_debug_prof_read).  Compiled out of the product build.
#ifdef MK_TC_PROF
__device__ unsigned long long g_prof[256 * 32];
#define PROF_DECL(...) unsigned long long __VA_ARGS__
#define PROF_NOW() clock64()
#define PROF_T0(v) v = clock64()
#define PROF_ADD(acc, t0) acc += clock64() - (t0)
#define PROF_PUT(slot, v) g_prof[blockIdx.x * 32 + (slot)] = (v)
#define PROF_INC(acc) ++acc
#else
#define PROF_DECL(...)
#define PROF_NOW() 0
#define PROF_T0(v)
#define PROF_ADD(acc, t0)
#define PROF_PUT(slot, v)
#define PROF_INC(acc)
#endif

struct ConvTCParams {
  const void *x;          // [n_in, c_in] 16-bit
  const void *bpack;      // [nsplit][K][nchunk][NT rows x 128 B] pre-swizzled
  const int32_t *nbr;     // [K, ld]
  const uint32_t *mask;   // [ntiles, W]
  void *y;                // [n_out, c_out]
  const int32_t *out_rows; // optional [n_out]: table column o is output row out_rows[o] (tile-ordered maps)
  int64_t ld, n_out, n_in;
  int K, W, kflip;
  int *group_counter;     // optional [gridDim.y], zeroed before the launch: groups are handed out in order, on demand
  int b_merged;           // weight ring = A ring (same depth): image s travels with stage s on the stage's barriers
  int c_in, c_out, nchunk, NT, T;
  int a_stages, b_stages, i_slots;
  int ksplit, kper;       // offsets split over blockIdx.z (small maps): CTA z serves schedule positions [z kper, (z+1) kper)
  int lean;               // gather warps: the lean loop (T == 4, one channel chunk, one mask word per tile)
  int half_cols;          // TMEM columns between the two accumulator halves: 256, or 0 = one accumulator set only
                          // (wide outputs: twice the tiles per group, i.e. per weight image, instead of an
                          // epilogue that overlaps the next group)
  int ngroups;
  uint32_t idesc;
};

// ring position: stage index + phase parity, advanced without division
struct Ring {
  uint32_t stage = 0, phase = 0;
  __device__ __forceinline__ void advance(uint32_t n) {
    if (++stage == n) { stage = 0; phase ^= 1u; }
  }
  __device__ __forceinline__ void skip(uint32_t by, uint32_t n) {
    stage += by;
    while (stage >= n) { stage -= n; phase ^= 1u; }
  }
};

// B image of W for the MMA: per (split, k, chunk) NT rows (n) x 64 reduction
// channels (128 B, K-major), 16-byte chunks XOR-swizzled with (n & 7).
//   transpose == 0: B(n, c) = W[k][c][n0 + n]        W: [K, c_red, c_n]
//   transpose == 1: B(n, c) = W[k][n0 + n][c]        W: [K, c_n, c_red]
template <typename T>
__global__ void k_pack_b(const T *__restrict__ w, int transpose, int K, int c_red, int c_n, int nchunk, int NT,
                         int nsplit, T *__restrict__ out, int *__restrict__ group_counter) {
  // (also resets the forward kernel's group counters: it runs right before it on the stream)
  if (group_counter && blockIdx.x == 0 && (int)threadIdx.x < nsplit) group_counter[threadIdx.x] = 0;
  const int64_t total = (int64_t)nsplit * K * nchunk * NT * 64;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (int64_t)gridDim.x * blockDim.x) {
    const int cc = (int)(e & 63);
    int64_t r = e >> 6;
    const int n = (int)(r % NT); r /= NT;
    const int ch = (int)(r % nchunk); r /= nchunk;
    const int k = (int)(r % K);
    const int sp = (int)(r / K);
    const int c = ch * 64 + cc;
    const int ng = sp * NT + n;
    T v = T(0.f);
    if (c < c_red && ng < c_n)
      v = transpose ? w[((size_t)k * c_n + ng) * c_red + c] : w[((size_t)k * c_red + c) * c_n + ng];
    const size_t base = ((((size_t)sp * K + k) * nchunk + ch) * NT + n) * 64;
    const int chunk16 = (cc >> 3) ^ (n & 7);
    out[base + chunk16 * 8 + (cc & 7)] = v;
  }
}

// transpose == 1 (data gradient: B(n, c) = W[k][n0 + n][c], contiguous along c like the image): one thread
// per 16-byte chunk, 32-bit index arithmetic (the element-wise kernel spends ~100 instructions of 64-bit
// division per 2-byte element: 64 us for a 27 x 960 x 960 kernel).
template <typename T>
__global__ void __launch_bounds__(256)
k_pack_b_rows(const T *__restrict__ w, int K, int c_red, int c_n, int nchunk, int NT, int nsplit, T *__restrict__ out,
              int *__restrict__ group_counter) {
  if (group_counter && blockIdx.x == 0 && (int)threadIdx.x < nsplit) group_counter[threadIdx.x] = 0;
  const uint32_t total = (uint32_t)nsplit * K * nchunk * NT * 8u;
  for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const uint32_t chunk = e & 7u;
    uint32_t r = e >> 3;
    const uint32_t n = r % (uint32_t)NT; r /= (uint32_t)NT;
    const uint32_t ch = r % (uint32_t)nchunk; r /= (uint32_t)nchunk;
    const uint32_t k = r % (uint32_t)K;
    const uint32_t sp = r / (uint32_t)K;
    const uint32_t c0 = ch * 64u + chunk * 8u, ng = sp * (uint32_t)NT + n;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (c0 < (uint32_t)c_red && ng < (uint32_t)c_n)   // c_red is a multiple of 16: a chunk is whole or absent
      v = __ldg(reinterpret_cast<const uint4 *>(w + ((size_t)k * c_n + ng) * c_red + c0));
    const size_t base = ((((size_t)sp * K + k) * nchunk + ch) * NT + n) * 64;
    *reinterpret_cast<uint4 *>(out + base + ((chunk ^ (n & 7u)) << 3)) = v;
  }
}

// The same image for transpose == 0 through a shared-memory tile: W[k][c][n] is contiguous along n, the
// image along c, so the element-wise kernel above reads 2 bytes per 32-byte sector (64 us for the 50 MB of
// a 27 x 960 x 960 kernel, 13 % of the kernel time of a diffusion-UNet step).  One block = 64 channels x
// 64 output columns of one offset: coalesced reads along n, 16-byte writes along c.
template <typename T>
__global__ void __launch_bounds__(256)
k_pack_b_tiled(const T *__restrict__ w, int K, int c_red, int c_n, int nchunk, int NT, int nsplit,
               T *__restrict__ out, int *__restrict__ group_counter) {
  __shared__ T tile[64][66];
  if (group_counter && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (int)threadIdx.x < nsplit)
    group_counter[threadIdx.x] = 0;
  const int ng0 = blockIdx.x * 64, ch = blockIdx.y, k = blockIdx.z;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
#pragma unroll
  for (int r = 0; r < 16; ++r) {
    const int cl = r * 4 + ty;
    const int c = ch * 64 + cl, ng = ng0 + tx;
    tile[cl][tx] = (c < c_red && ng < c_n) ? w[((size_t)k * c_red + c) * c_n + ng] : T(0.f);
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int idx = (int)threadIdx.x + j * 256;
    const int nl = idx >> 3, chunk = idx & 7;
    const int ng = ng0 + nl;
    if (ng >= nsplit * NT) continue;
    const int sp = ng / NT, n = ng - sp * NT;
    alignas(16) T v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = tile[chunk * 8 + i][nl];
    const size_t base = ((((size_t)sp * K + k) * nchunk + ch) * NT + n) * 64;
    *reinterpret_cast<uint4 *>(out + base + ((chunk ^ (n & 7)) << 3)) = *reinterpret_cast<const uint4 *>(v);
  }
}

// mask[tile*W + k/32] bit k%32 = any(nbr[k, 128*tile .. 128*tile+127] >= 0)
__global__ void k_tile_mask(const int32_t *__restrict__ nbr, int64_t n_rows, int64_t ld, int K, int W,
                            uint32_t *__restrict__ mask) {
  const int64_t tile = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
  extern __shared__ uint32_t s_mask[];
  for (int i = threadIdx.x; i < W; i += blockDim.x) s_mask[i] = 0;
  __syncthreads();
  for (int k = warp; k < K; k += nwarp) {
    const int32_t *row = nbr + (int64_t)k * ld + tile * kTileM;
    bool any = false;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t o = tile * kTileM + j * 32 + lane;
      any |= (o < n_rows) && (__ldg(row + j * 32 + lane) >= 0);
    }
    if (__any_sync(0xffffffffu, any) && lane == 0) atomicOr(&s_mask[k >> 5], 1u << (k & 31));
  }
  __syncthreads();
  for (int i = threadIdx.x; i < W; i += blockDim.x) mask[tile * W + i] = s_mask[i];
}

template <typename T> __device__ __forceinline__ uint32_t pack2(float a, float b);
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t *>(&v);
}
template <> __device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t *>(&v);
}

// Per-role copy of a group's schedule: sched[k] = T-bit mask of the group's
// tiles that have any neighbour under the table offset used by weight k.
// Called by a whole (converged) warp.
__device__ __forceinline__ uint32_t fill_sched(uint8_t *sched, const ConvTCParams &p, int64_t tile0,
                                               int64_t ntiles, int lane) {
  __syncwarp();
  uint32_t cnt = 0;
  const int kz0 = (int)blockIdx.z * p.kper, kz1 = min(p.K, kz0 + p.kper);
  for (int k = lane; k < p.K; k += 32) {
    const int kk = p.kflip ? p.K - 1 - k : k;
    uint32_t bits = 0;
    for (int t = 0; t < p.T; ++t) {
      const int64_t tile = tile0 + t;
      if (tile < ntiles && k >= kz0 && k < kz1) bits |= ((__ldg(p.mask + tile * p.W + (kk >> 5)) >> (kk & 31)) & 1u) << t;
    }
    sched[k] = (uint8_t)bits;
    cnt += bits ? 1u : 0u;
  }
  __syncwarp();
  return __reduce_add_sync(0xffffffffu, cnt);   // number of offsets with any neighbour in the group
}

// Bits of mask word `wi` (table offsets 32 wi .. 32 wi + 31) that this CTA's share of the offsets covers
__device__ __forceinline__ uint32_t subset_word(const ConvTCParams &p, int wi) {
  const int kz0 = (int)blockIdx.z * p.kper, kz1 = min(p.K, kz0 + p.kper);
  const int lo = (p.kflip ? p.K - kz1 : kz0) - 32 * wi, hi = (p.kflip ? p.K - kz0 : kz1) - 32 * wi;
  const int a = max(lo, 0), b = min(hi, 32);
  if (b <= a) return 0u;
  return (b == 32 ? 0xffffffffu : (1u << b) - 1u) & ~((1u << a) - 1u);
}

// Queue through which the index warp tells the other roles which group comes next (dynamic
// group scheduling).  One claim per group: the cost is a handful of barrier operations per
// ~40 k cycles of work.  Every consumer warp waits on every phase of `full`, the producer on
// every phase of `empty` (parity-wait soundness).
constexpr int kGroupQ = 4;
constexpr int kGroupConsumers = 4 /*epilogue*/ + 1 /*weights*/ + kGatherWarps;
struct GroupQueue {
  int *q;
  uint64_t *full, *empty;
  int *counter;   // nullptr: static schedule (every gridDim.x-th group)
  int ngroups;
  uint32_t slot = 0, phase = 0;
  __device__ __forceinline__ int fetch(bool producer, int lane) {
    int g = 0;
    if (producer) {
      mbar_wait_warp(&empty[slot], phase ^ 1u);
      if (lane == 0) {
        g = atomicAdd(counter, 1);
        if (g >= ngroups) g = -1;
        *(volatile int *)&q[slot] = g;
        mbar_arrive(&full[slot]);   // release: orders the entry before the phase completes
      }
      g = __shfl_sync(0xffffffffu, g, 0);
    } else {
      if (lane == 0) mbar_wait_shared(&full[slot], phase);   // two dozen waiters: short polls (no sleeping wait)
      __syncwarp();
      g = *(volatile int *)&q[slot];
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
    }
    if (++slot == kGroupQ) { slot = 0; phase ^= 1u; }
    return g;
  }
  __device__ __forceinline__ int first(bool producer, int lane) {
    if (counter) return fetch(producer, lane);
    return (int)blockIdx.x < ngroups ? (int)blockIdx.x : -1;
  }
  __device__ __forceinline__ int next(bool producer, int lane, int g) {
    if (counter) return fetch(producer, lane);
    g += (int)gridDim.x;
    return g < ngroups ? g : -1;
  }
};

// Ticket: what the MMA warp has to do with one A stage (= the T gathered tiles of
// one offset and channel chunk), written by a gather team (the teams walk the
// schedule anyway; the MMA warp is the serial resource of the CTA and must not).
//   bits 0-3 tiles present, bits 4-7 tiles whose accumulator already holds data,
//   bit 8 TMEM half, bits 16-18 K steps
constexpr uint32_t kTkGroupBegin = 1u << 12, kTkGroupEnd = 1u << 13, kTkEmptyParity = 1u << 14, kTkEnd = 1u << 15;
// single accumulator set (half_cols == 0): the group must also wait for the epilogue of the PREVIOUS group
constexpr uint32_t kTkPrevParity = 1u << 9, kTkHasPrev = 1u << 10;

// 16 accumulator columns -> 16 outputs of type TOut (T: 2 x 16-byte stores; float: 4)
template <typename T, typename TOut>
__device__ __forceinline__ void store16(TOut *dst, const uint32_t (&v)[16]) {
  if constexpr (sizeof(TOut) == 4) {
    uint4 *d = reinterpret_cast<uint4 *>(dst);
    d[0] = make_uint4(v[0], v[1], v[2], v[3]);
    d[1] = make_uint4(v[4], v[5], v[6], v[7]);
    d[2] = make_uint4(v[8], v[9], v[10], v[11]);
    d[3] = make_uint4(v[12], v[13], v[14], v[15]);
  } else {
    uint4 o0, o1;
    o0.x = pack2<T>(__uint_as_float(v[0]), __uint_as_float(v[1]));
    o0.y = pack2<T>(__uint_as_float(v[2]), __uint_as_float(v[3]));
    o0.z = pack2<T>(__uint_as_float(v[4]), __uint_as_float(v[5]));
    o0.w = pack2<T>(__uint_as_float(v[6]), __uint_as_float(v[7]));
    o1.x = pack2<T>(__uint_as_float(v[8]), __uint_as_float(v[9]));
    o1.y = pack2<T>(__uint_as_float(v[10]), __uint_as_float(v[11]));
    o1.z = pack2<T>(__uint_as_float(v[12]), __uint_as_float(v[13]));
    o1.w = pack2<T>(__uint_as_float(v[14]), __uint_as_float(v[15]));
    reinterpret_cast<uint4 *>(dst)[0] = o0;
    reinterpret_cast<uint4 *>(dst)[1] = o1;
  }
}

// TOut = T (16-bit features in, 16-bit out) or float (the fp32 split path of conv_f32.cu:
// bf16 operand pieces in, the fp32 accumulator out as it is)
template <typename T, bool kFull, typename TOut>
__global__ void __launch_bounds__(kNumThreadsTC, 1)
k_conv_tc(const __grid_constant__ ConvTCParams p, const __grid_constant__ CUtensorMap tmap_x) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment for the 128B-swizzle atoms
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t S = (uint32_t)p.a_stages;
  const uint32_t b_bytes = (uint32_t)p.NT * 128u;
  uint8_t *a_ring = smem;
  const uint32_t stage_bytes = (uint32_t)p.T * kTileBytes;
  uint8_t *b_ring = a_ring + (size_t)S * stage_bytes;
  const uint32_t kBStages = (uint32_t)p.b_stages, kIdxSlots = (uint32_t)p.i_slots;
  const uint32_t idx_slot = (uint32_t)p.T * kTileM;   // entries per slot
  int32_t *idx_ring = (int32_t *)(b_ring + kBStages * b_bytes);
  const uint32_t sched_stride = ((uint32_t)p.K + 31u) & ~31u;            // bytes per role copy of the schedule
  uint8_t *sched_all = (uint8_t *)(idx_ring + kIdxSlots * idx_slot);   // per-role copies x sched_stride
  uint32_t *tickets = (uint32_t *)(sched_all + kSchedCopies * sched_stride);   // [S]
  uint64_t *bars = (uint64_t *)(tickets + 16);
  uint64_t *a_full = bars, *a_empty = a_full + S;
  uint64_t *b_full = a_empty + S, *b_empty = b_full + kBStages;
  uint64_t *i_full = b_empty + kBStages, *i_empty = i_full + kIdxSlots;
  uint64_t *t_full = i_empty + kIdxSlots, *t_empty = t_full + 2;
  uint64_t *gq_full = t_empty + 2, *gq_empty = gq_full + kGroupQ;
  uint32_t *tmem_slot = (uint32_t *)(gq_empty + kGroupQ);
  int *gq = (int *)(tmem_slot + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t ntiles = (p.n_out + kTileM - 1) / kTileM;
  const bool role_is_idx = warp == kWarpIdx;
  const int sp = blockIdx.y;
  const int acc_stride = (p.NT + 31) & ~31;
  // Which groups this CTA works on.  Static: every gridDim.x-th group.  Dynamic (group_counter):
  // the index warp claims the next group of the layer when its prefetch gets there and
  // publishes it to the other roles through a small queue (gq); groups are still handed
  // out in table order, so concurrently processed groups stay neighbours in L2, but a CTA
  // that drew cheap groups simply takes more of them (static: the slowest CTA of the
  // north-star layer carried 9 % more stages than the mean).
  GroupQueue gq_ctl{gq, gq_full, gq_empty,
                    p.group_counter ? p.group_counter + blockIdx.z * gridDim.y + blockIdx.y : nullptr, p.ngroups};

  if (threadIdx.x == 0) {
    for (uint32_t s = 0; s < S; ++s) { mbar_init(&a_full[s], (uint32_t)p.T * (p.lean == 2 ? 132u : 128u) + 1u + (p.b_merged ? 1u : 0u)); mbar_init(&a_empty[s], kMmaWarps); }
    for (uint32_t s = 0; s < kBStages; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], kMmaWarps); }
    for (uint32_t s = 0; s < kIdxSlots; ++s) { mbar_init(&i_full[s], 1); mbar_init(&i_empty[s], (uint32_t)(p.nchunk * p.T) * 4u); }
    for (int s = 0; s < 2; ++s) { mbar_init(&t_full[s], kMmaWarps); mbar_init(&t_empty[s], 4); }
    for (int s = 0; s < kGroupQ; ++s) { mbar_init(&gq_full[s], 1); mbar_init(&gq_empty[s], kGroupConsumers); }
    fence_barrier_init();
  }
  if (warp == kWarpGather0 && lane == 0 && p.lean == 2) tma_prefetch_desc(&tmap_x);
  if (warp == kWarpMma0) tmem_alloc_512(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    // ===================== epilogue =====================
    uint32_t it = 0;   // counts the groups that have any work (the others never reach the MMA thread)
    PROF_DECL(pe_wait = 0, pe_work = 0, pe_t0 = 0, pe_t1 = 0;)
    for (int g = gq_ctl.first(role_is_idx, lane); g >= 0; g = gq_ctl.next(role_is_idx, lane, g)) {
      const uint32_t h = it & 1u;
      const int64_t tile0 = (int64_t)g * p.T;
      uint32_t any = 0;
      for (int t = 0; t < p.T; ++t)
        if (tile0 + t < ntiles)
          for (int wi = 0; wi < p.W; ++wi) any |= __ldg(p.mask + (tile0 + t) * p.W + wi) & subset_word(p, wi);
      if (any) {
        PROF_T0(pe_t0);
        mbar_wait_warp(&t_full[h], (it >> 1) & 1u, 3, 400);   // outermost wait: report inner stalls first
        tc_fence_after();
        PROF_ADD(pe_wait, pe_t0);
      }
      PROF_T0(pe_t1);
      for (int t = 0; t < p.T; ++t) {
        const int64_t tile = tile0 + t;
        if (tile >= ntiles) break;
        uint32_t touched = 0;
        for (int wi = 0; wi < p.W; ++wi) touched |= __ldg(p.mask + tile * p.W + wi) & subset_word(p, wi);
        const int64_t row = tile * kTileM + warp * 32 + lane;   // table column
        const int64_t orow = (p.out_rows && row < p.n_out) ? (int64_t)__ldg(p.out_rows + row) : row;
        TOut *yrow = (TOut *)p.y + ((size_t)blockIdx.z * p.n_out + (size_t)orow) * p.c_out + (size_t)sp * p.NT;
        const uint32_t tbase = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(h * p.half_cols + t * acc_stride);
        for (int c0 = 0; c0 < p.NT; c0 += 16) {
          uint32_t v[16];
          if (touched) {
            tmem_ld16(tbase + (uint32_t)c0, v);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = 0u;
          }
          if (row < p.n_out) store16<T, TOut>(yrow + c0, v);
        }
      }
      PROF_ADD(pe_work, pe_t1);
      if (any) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&t_empty[h]);
        ++it;
      }
    }
    if (warp == 0 && lane == 0) { PROF_PUT(0, pe_wait); PROF_PUT(1, pe_work); }
  } else if (warp == kWarpIdx) {
    // ===================== neighbour-index prefetch =====================
    // one slot of the idx ring per offset: the T*128 table entries of the group; the
    // channel chunks of an offset reuse the slot
    uint8_t *sched = sched_all + 0 * sched_stride;
    const uint32_t i_base = smem_u32(idx_ring);
    Ring ir;
    PROF_DECL(pi_wait = 0, pi_t0 = 0;)
    for (int g = gq_ctl.first(role_is_idx, lane); g >= 0; g = gq_ctl.next(role_is_idx, lane, g)) {
      const int64_t tile0 = (int64_t)g * p.T;
      fill_sched(sched, p, tile0, ntiles, lane);
      for (int k = 0; k < p.K; ++k) {
        const uint32_t bits = sched[k];
        if (!bits) continue;
        const int kk = p.kflip ? p.K - 1 - k : k;
        PROF_T0(pi_t0);
        mbar_wait_warp(&i_empty[ir.stage], ir.phase ^ 1u);
        PROF_ADD(pi_wait, pi_t0);
        if (lane == 0) {
          // the T tiles of a group are consecutive rows of the table: one copy
          const int64_t first = tile0 * kTileM;
          const uint32_t bytes = (uint32_t)min((int64_t)p.T * kTileM, p.ld - first) * 4u;
          const int32_t *src = p.nbr + (int64_t)kk * p.ld + first;
          mbar_arrive_expect_tx(&i_full[ir.stage], bytes);
          bulk_g2s(i_base + ir.stage * (idx_slot * 4u), src, bytes, &i_full[ir.stage]);
        }
        __syncwarp();
        ir.advance(kIdxSlots);
      }
    }
    if (lane == 0) PROF_PUT(2, pi_wait);
  } else if (warp >= kWarpGather0) {
    // ===================== gather producers (LDGSTS) =====================
    // Four teams of four warps.  Tile slot t of stage number n (stages are counted
    // over the whole life of the CTA) belongs to team (n*T + t) & 3, so the stages
    // of a team are at most 4/T < S apart, which keeps the parity waits on a_empty
    // alias-free as long as the MMA warp releases stages in order.  Warp q of a
    // team moves tile rows [32q, 32q+32); lane = (rsub, c16): instruction i moves
    // 16-byte chunk c16 of rows 32q + 8*rsub + i (four 128-byte feature rows per
    // warp instruction).  Completion is tracked by the hardware
    // (cp.async.mbarrier.arrive.noinc); a team whose tile has no neighbour under
    // the offset just arrives; one extra arrive publishes the ticket.
    // The loop is bound by instruction issue (16 warps x a few hundred instructions
    // per stage, measured with ncu's source view), so everything that does not change
    // is computed once, shared memory is addressed in its own space (a generic
    // pointer costs generic LD/ST plus a cvta per barrier operation) and the
    // schedule is walked on a register bit mask.
    const int gw = warp - kWarpGather0;
    const uint32_t team = (uint32_t)gw >> 2;
    const int q = gw & 3;
    uint8_t *sched = sched_all + (2 + gw) * sched_stride;
    const uint32_t sched_s = smem_u32(sched);
    const uint32_t row_bytes = (uint32_t)p.c_in * 2u;
    const int c16 = lane & 7, rsub = lane >> 3;
    const int row0 = 32 * q + 8 * rsub;            // first of this lane's 8 rows (multiple of 8)
    const uint32_t T = (uint32_t)p.T;
    // lane-constant parts of the addresses.  Destination of instruction i: bits 4-6 =
    // c16 ^ i (swizzle), bits 7-9 = i (row in the 8-row atom): both an XOR with i * 144,
    // as the atom base has those bits clear.
    const uint32_t dst_lane = smem_u32(a_ring) + (uint32_t)row0 * 128u + ((uint32_t)c16 << 4);
    const uint32_t idx_lane = smem_u32(idx_ring) + (uint32_t)row0 * 4u;
    const uint8_t *x_lane = (const uint8_t *)p.x + (size_t)c16 * 16;
    asm volatile("" : "+l"(x_lane));   // keep the lane's base in a register pair: one IMAD.WIDE per row instead of
                                       // IMAD.WIDE + a 64-bit add of the parameter re-read from the constant bank
    const uint32_t b_a_full = smem_u32(a_full), b_a_empty = smem_u32(a_empty);
    const uint32_t b_i_full = smem_u32(i_full), b_i_empty = smem_u32(i_empty);
    const uint32_t tickets_s = smem_u32(tickets);
    const bool ticket_lane = q == 0 && lane == 0;
    const uint32_t nchunk = (uint32_t)p.nchunk;
    Ring ir, ar;                                   // idx slot of the row's first tile / current A stage
    uint32_t nst = 0;                              // stages so far
    uint32_t it = 0;                               // groups with work so far
    PROF_DECL(pg_we = 0, pg_wi = 0, pg_iss = 0, pg_t0 = 0, pg_n = 0, pg_start = PROF_NOW(), pg_grp = 0, pg_g0 = PROF_NOW(), pg_abs = 0;)
    if (kFull && p.lean) {
      // Lean form of the loop below for the layers the ingest side bounds (Cin = 64, Cout <= 64, K <= 32:
      // T = 4, one channel chunk, one mask word per tile).  Tile slot t of every stage then belongs to team
      // t for good, a stage is an offset, and the group's schedule is the four tiles' mask words in
      // registers: no schedule copy in shared memory, no ballots, no nested loops.  (The general loop
      // spends ~350 instructions per warp and stage at one instruction per ~6 cycles, of which 8 are
      // copies: the gather warps' own chain, not the copy rate, set the stage period - DESIGN 4.9.)
      const int kz0 = (int)blockIdx.z * p.kper, kz1 = min(p.K, kz0 + p.kper);
      const uint32_t kz_mask = (kz1 >= 32 ? 0xffffffffu : (1u << kz1) - 1u) & ~((1u << kz0) - 1u);
      for (int g = gq_ctl.first(role_is_idx, lane); g >= 0; g = gq_ctl.next(role_is_idx, lane, g)) {
        const int64_t tile0 = (int64_t)g * 4;
        uint32_t mm[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t v = tile0 + j < ntiles ? __ldg(p.mask + tile0 + j) : 0u;
          if (p.kflip) v = __brev(v) >> (32 - p.K);     // schedule position k <-> table offset K-1-k
          mm[j] = v & kz_mask;
        }
        const uint32_t all = mm[0] | mm[1] | mm[2] | mm[3];
        PROF_ADD(pg_grp, pg_g0);
        if (!all) { PROF_T0(pg_g0); continue; }
        const uint32_t mine = team == 0 ? mm[0] : team == 1 ? mm[1] : team == 2 ? mm[2] : mm[3];
        uint32_t tk_group = (it & 1u) << 8 | ((((it >> 1) & 1u) ^ 1u) ? kTkEmptyParity : 0u) | (4u << 16);
        if (p.half_cols == 0 && it > 0)
          tk_group |= kTkHasPrev | ((((it - 1u) >> 1) & 1u) ? kTkPrevParity : 0u);
        uint32_t seen = 0;
#pragma unroll 1
        for (uint32_t m = all; m; m &= m - 1u, ++nst) {
          const uint32_t k = (uint32_t)__ffs((int)m) - 1u;
          const uint32_t a_full_s = b_a_full + ar.stage * 8u;
          const uint32_t i_full_k = b_i_full + ir.stage * 8u, i_empty_k = b_i_empty + ir.stage * 8u;
          PROF_T0(pg_t0);
          mbar_wait_warp_a(b_a_empty + ar.stage * 8u, ar.phase ^ 1u);
          PROF_ADD(pg_we, pg_t0);
          if (team == 0 && ticket_lane) {
            const uint32_t bits = ((mm[0] >> k) & 1u) | ((mm[1] >> k) & 1u) << 1 | ((mm[2] >> k) & 1u) << 2 |
                                  ((mm[3] >> k) & 1u) << 3;
            uint32_t tk = tk_group | bits | (seen << 4);
            if (m == all) tk |= kTkGroupBegin;
            if (!(m & (m - 1u))) tk |= kTkGroupEnd;
            sts_u32(tickets_s + ar.stage * 4u, tk);
            mbar_arrive_a(a_full_s);                 // release: orders the ticket before the phase completes
            seen |= bits;
          }
          PROF_T0(pg_t0);
          mbar_wait_warp_a(i_full_k, ir.phase);      // (also when the tile is absent: see the general loop)
          PROF_ADD(pg_wi, pg_t0);
          if (((mine >> k) & 1u) && p.lean == 2) {
            // Hybrid ingest.  On a tile-ordered table 71 % of the 4-row groups of a non-empty (tile,
            // offset) pair are fully valid (96 % of the valid rows), 6 % partly, 23 % empty
            // (scripts/group_stats.py).  Full groups go through TMA tile::gather4 - one instruction of
            // one lane per 512 bytes, written by the async proxy with the map's 128-byte swizzle, bytes
            // counted on the stage's barrier - and only the rest through LDGSTS with zero fill (one warp
            // instruction per group): the LSU path, which drains a 16 KB tile in ~400 cycles, carries 29 %
            // of the groups.  gather4 alone is unusable here: a missing row costs it ~10 cycles (4.5).
            PROF_T0(pg_t0);
            const uint32_t idx_w = smem_u32(idx_ring) + ir.stage * (4u * kTileM * 4u) + team * (kTileM * 4u) +
                                   (uint32_t)q * 128u;                    // this warp's 32 table entries
            int4 gi = make_int4(-1, -1, -1, -1);
            if (lane < 8) gi = lds_v4(idx_w + (uint32_t)lane * 16u);       // lane j < 8: rows 4j .. 4j+3
            const bool fullg = (gi.x | gi.y | gi.z | gi.w) >= 0;
            const uint32_t fm = __ballot_sync(0xffffffffu, fullg);
            const uint32_t tile_s = smem_u32(a_ring) + ar.stage * (4u * kTileBytes) + team * kTileBytes +
                                    (uint32_t)q * 4096u;
            if (lane == 0) mbar_arrive_expect_tx_a(a_full_s, 512u * (uint32_t)__popc(fm));
            if (fullg) tma_gather4_a(tile_s + (uint32_t)lane * 512u, &tmap_x, 0, gi, a_full_s);
            // lane (r4, c16) moves chunk c16 of row 4j + r4: bits 4-6 of the address = c16 ^ ((4j + r4) & 7)
            const uint32_t r4 = (uint32_t)lane >> 3;
            const uint32_t dst_l = tile_s + r4 * 128u + (((uint32_t)c16 ^ r4) << 4);
#pragma unroll 1
            for (uint32_t rest = ~fm & 0xFFu; rest; rest &= rest - 1u) {
              const uint32_t j = (uint32_t)__ffs((int)rest) - 1u;
              const int32_t r = (int32_t)lds_u32(idx_w + (4u * j + r4) * 4u);
              cp_async16((dst_l + j * 512u) ^ ((j & 1u) << 6), x_lane + (size_t)(uint32_t)max(r, 0) * row_bytes,
                         r >= 0 ? 16u : 0u);
            }
            cp_async_arrive_noinc_a(a_full_s);
            __syncwarp();                            // every index has been consumed: now the slot may go
            if (lane == 0) mbar_arrive_a(i_empty_k);
            PROF_ADD(pg_iss, pg_t0);
            PROF_INC(pg_n);
          } else if ((mine >> k) & 1u) {
            PROF_T0(pg_t0);
            const uint32_t idx_k = idx_lane + ir.stage * (4u * kTileM * 4u) + team * (kTileM * 4u);
            const int4 i0 = lds_v4(idx_k), i1 = lds_v4(idx_k + 16u);
            const uint32_t sb = dst_lane + ar.stage * (4u * kTileBytes) + team * kTileBytes;
            const int32_t ri[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
              cp_async16(sb ^ (uint32_t)(i * 144), x_lane + (size_t)(uint32_t)max(ri[i], 0) * row_bytes,
                         ri[i] >= 0 ? 16u : 0u);
            cp_async_arrive_noinc_a(a_full_s);
            __syncwarp();                            // the copies consumed the indices: now the slot may go
            if (lane == 0) mbar_arrive_a(i_empty_k);
            PROF_ADD(pg_iss, pg_t0);
            PROF_INC(pg_n);
          } else {
            mbar_arrive_a(a_full_s);                 // nothing to gather: stand in for the cp.async arrivals
            if (lane == 0) {
              if (p.lean == 2) mbar_arrive_a(a_full_s);   // ... and for the warp's expect_tx arrival
              mbar_arrive_a(i_empty_k);
            }
          }
          ar.advance(S);
          ir.advance(kIdxSlots);
        }
        ++it;
        PROF_T0(pg_g0);
      }
    } else
    for (int g = gq_ctl.first(role_is_idx, lane); g >= 0; g = gq_ctl.next(role_is_idx, lane, g)) {
      const int64_t tile0 = (int64_t)g * p.T;
      const uint32_t n_stages = fill_sched(sched, p, tile0, ntiles, lane) * nchunk;
      PROF_ADD(pg_grp, pg_g0);
      if (!n_stages) { PROF_T0(pg_g0); continue; }
      // offsets with work, as a bit mask in registers (sched[] itself is in shared memory)
      uint32_t any0, any1 = 0, any2 = 0, any3 = 0;
      any0 = __ballot_sync(0xffffffffu, lane < p.K && lds_u8(sched_s + (uint32_t)lane) != 0u);
      if (p.K > 32) {
        any1 = __ballot_sync(0xffffffffu, 32 + lane < p.K && lds_u8(sched_s + 32u + (uint32_t)lane) != 0u);
        any2 = __ballot_sync(0xffffffffu, 64 + lane < p.K && lds_u8(sched_s + 64u + (uint32_t)lane) != 0u);
        any3 = __ballot_sync(0xffffffffu, 96 + lane < p.K && lds_u8(sched_s + 96u + (uint32_t)lane) != 0u);
      }
      uint32_t tk_group = (it & 1u) << 8 | ((((it >> 1) & 1u) ^ 1u) ? kTkEmptyParity : 0u);
      if (p.half_cols == 0 && it > 0)     // completion ((it-1)>>1)+1 of the other barrier = epilogue of group it-1
        tk_group |= kTkHasPrev | ((((it - 1u) >> 1) & 1u) ? kTkPrevParity : 0u);
      uint32_t seen = 0, done = 0;                 // tiles already accumulated into / stages of the group so far
#pragma unroll 1
      for (int j = 0; j < 4; ++j) {
#pragma unroll 1
        for (uint32_t m = j == 0 ? any0 : j == 1 ? any1 : j == 2 ? any2 : any3; m; m &= m - 1u) {
          const uint32_t k = 32u * (uint32_t)j + (uint32_t)__ffs((int)m) - 1u;
          const uint32_t bits = lds_u8(sched_s + k);
          const uint32_t i_full_k = b_i_full + ir.stage * 8u, i_empty_k = b_i_empty + ir.stage * 8u;
          const uint32_t idx_k = idx_lane + ir.stage * (idx_slot * 4u);
#pragma unroll 1
          for (uint32_t ch = 0; ch < nchunk; ++ch, ++nst, ++done) {
            const uint32_t a_full_s = b_a_full + ar.stage * 8u;
            const uint32_t dst_s = dst_lane + ar.stage * stage_bytes;
            bool waited = false;
#pragma unroll 1
            for (uint32_t t = (team - nst * T) & 3u; t < T; t += 4) {
              if (!waited) {
                PROF_T0(pg_t0);
                mbar_wait_warp_a(b_a_empty + ar.stage * 8u, ar.phase ^ 1u);
                PROF_ADD(pg_we, pg_t0);
                waited = true;
              }
              if (t == 0 && ticket_lane) {
                const uint32_t nks = kFull ? 4u : (uint32_t)min(4, (p.c_in - (int)ch * 64) >> 4);
                uint32_t tk = tk_group | bits | ((ch > 0 ? 0xFu : seen) << 4) | (nks << 16);
                if (done == 0) tk |= kTkGroupBegin;
                if (done == n_stages - 1) tk |= kTkGroupEnd;
                sts_u32(tickets_s + ar.stage * 4u, tk);
                mbar_arrive_a(a_full_s);             // release: orders the ticket before the phase completes
              }
              if (!((bits >> t) & 1u)) {
                PROF_T0(pg_t0);
                mbar_arrive_a(a_full_s);             // nothing to gather: stand in for the cp.async arrivals
                // The indices are not needed, but the team must still OBSERVE this phase of the
                // slot's barrier: a parity wait is only sound for a waiter that has seen every
                // earlier phase.  A team that skipped here could run a whole idx ring ahead (the A
                // ring is deeper than the idx ring when T < 4) and find the slot's next wait
                // satisfied by THIS phase still in flight - it then gathered through stale
                // indices (seen as an illegal address with the T = 2 plan).
                mbar_wait_warp_a(i_full_k, ir.phase);
                if (lane == 0) mbar_arrive_a(i_empty_k);
                PROF_ADD(pg_abs, pg_t0);
                continue;
              }
              PROF_T0(pg_t0);
              mbar_wait_warp_a(i_full_k, ir.phase);
              PROF_ADD(pg_wi, pg_t0);
              PROF_T0(pg_t0);
              const int4 i0 = lds_v4(idx_k + t * (kTileM * 4u)), i1 = lds_v4(idx_k + t * (kTileM * 4u) + 16u);
              const uint32_t szok = (kFull || (ch * 64 + c16 * 8) < (uint32_t)p.c_in) ? 16u : 0u;
              const uint8_t *xb = x_lane + (size_t)ch * 128;
              asm volatile("" : "+l"(xb));   // one register pair: row address = a single IMAD.WIDE
              const uint32_t sb = dst_s + t * kTileBytes;
              const int32_t ri[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
#ifdef MK_TC_CHECK
              // development check: a gathered index outside the feature matrix (a pipeline bug:
              // the slot was read before it was loaded or after it was overwritten)
              for (int i = 0; i < 8; ++i)
                if (ri[i] >= (int32_t)p.n_in || ri[i] < -1) {
                  unsigned int *hh = mk_hang_slot;
                  if (hh) {
                    hh[1] = blockIdx.x | (blockIdx.y << 16);
                    hh[2] = (uint32_t)warp | (k << 8) | (t << 16) | (ir.stage << 20) | (ir.phase << 24) | (ar.stage << 25);
                    hh[3] = (uint32_t)ri[i];
                    hh[0] = 0xBAD1D000u | (uint32_t)i;
                    __threadfence_system();
                  }
                  asm volatile("trap;");
                }
#endif
              // table entries are -1 for a missing neighbour and for rows past n_out
#pragma unroll
              for (int i = 0; i < 8; ++i)
                cp_async16(sb ^ (uint32_t)(i * 144), xb + (size_t)(uint32_t)max(ri[i], 0) * row_bytes,
                           ri[i] >= 0 ? szok : 0u);
              cp_async_arrive_noinc_a(a_full_s);
              // Release the index slot only now: the copies above consumed every loaded
              // index, so the loads have completed.  (An asm LDS followed directly by the
              // asm arrive carries no dependence; the arrive then overtook the load and the
              // index warp overwrote the slot under it - a few wrong rows, run to run.)
              __syncwarp();
              if (lane == 0) mbar_arrive_a(i_empty_k);
              PROF_ADD(pg_iss, pg_t0);
              PROF_INC(pg_n);
            }
            ar.advance(S);
          }
          seen |= bits;
          ir.advance(kIdxSlots);
        }
      }
      ++it;
      PROF_T0(pg_g0);
    }
    if (lane == 0 && gw == 0) { PROF_PUT(28, PROF_NOW() - pg_start); PROF_PUT(29, pg_grp); PROF_PUT(30, pg_abs); }
    if (lane == 0 && q == 0) { PROF_PUT(8 + 4 * team, pg_we); PROF_PUT(9 + 4 * team, pg_wi); PROF_PUT(10 + 4 * team, pg_iss); PROF_PUT(11 + 4 * team, pg_n); }
    // the stage after the last one carries the END ticket
    for (uint32_t t = (team - nst * T) & 3u; t < T; t += 4) {
      if (t == (team - nst * T & 3u)) mbar_wait_warp(&a_empty[ar.stage], ar.phase ^ 1u);
      if (t == 0 && q == 0 && lane == 0) {
        tickets[ar.stage] = kTkEnd;
        mbar_arrive(&a_full[ar.stage]);
      }
      mbar_arrive(&a_full[ar.stage]);
      if (p.lean == 2 && lane == 0) mbar_arrive(&a_full[ar.stage]);
    }
  } else if (warp >= kWarpMma0 && warp < kWarpMma0 + kMmaWarps) {
    // ===================== MMA issuers: driven by tickets =====================
    // The whole warp runs the loop converged on warp-uniform values (the ticket is
    // broadcast with a shuffle), so that the descriptors live in uniform registers
    // and one elected lane issues back-to-back UTCHMMA without per-operand R2UR.
    {
      const uint32_t a_base = smem_u32(a_ring), b_base = smem_u32(b_ring);
      const uint32_t tm_base = __shfl_sync(0xffffffffu, tmem_base, 0);
      // descriptor template: SBO = 1024 B (8-row atoms), LBO = 16 B (unused for one
      // 128-byte K atom), version 1, SWIZZLE_128B; the start address is added per MMA
      const uint64_t desc0 = smem_desc_sw128(0, 16, 1024);
      const uint32_t mw = (uint32_t)(warp - kWarpMma0);
      uint32_t s = 0, ph = 0, bs = 0, bph = 0;
      PROF_DECL(pm_wa = 0, pm_wt = 0, pm_wb = 0, pm_iss = 0, pm_n = 0, pm_t0 = 0, pm_start = PROF_NOW(), pm_f = 0, pm_m = 0, pm_c = 0;)
      for (;;) {
        PROF_T0(pm_t0);
        mbar_wait_shared(&a_full[s], ph);
        PROF_ADD(pm_wa, pm_t0);
        const uint32_t tk = __shfl_sync(0xffffffffu, *(volatile uint32_t *)&tickets[s], 0);
        if (tk & kTkEnd) break;
        const uint32_t h = (tk >> 8) & 1u;
        PROF_T0(pm_t0);
        if (tk & kTkGroupBegin) {
          mbar_wait_shared(&t_empty[h], (tk & kTkEmptyParity) ? 1u : 0u);
          if (tk & kTkHasPrev) mbar_wait_shared(&t_empty[h ^ 1u], (tk & kTkPrevParity) ? 1u : 0u);
        }
        PROF_ADD(pm_wt, pm_t0);
        PROF_T0(pm_t0);
        if (!p.b_merged) mbar_wait_shared(&b_full[bs], bph);
        PROF_ADD(pm_wb, pm_t0);
        PROF_T0(pm_t0);
        tc_fence_after();
        const uint64_t bd0 = desc0 | (uint64_t)(((b_base + (p.b_merged ? s : bs) * b_bytes) & 0x3FFFFu) >> 4);
        const uint32_t d0 = tm_base + h * (uint32_t)p.half_cols;
        const uint32_t a0 = a_base + s * stage_bytes;
        if (elect_one()) {
          PROF_DECL(pq_t0 = clock64();)
          fence_proxy_async_smem();   // cp.async (generic proxy) writes -> tensor-core reads
          PROF_ADD(pm_f, pq_t0);
          PROF_T0(pq_t0);
#pragma unroll
          for (int t = 0; t < kMaxT; ++t) {
            if ((uint32_t)(t % kMmaWarps) != mw || !((tk >> t) & 1u)) continue;
            const uint32_t d = d0 + (uint32_t)(t * acc_stride);
            const uint64_t ad0 = desc0 | (uint64_t)(((a0 + (uint32_t)t * kTileBytes) & 0x3FFFFu) >> 4);
            const uint32_t acc0 = (tk >> (4 + t)) & 1u;
            if (kFull) {
              umma_f16(d, ad0, bd0, p.idesc, acc0);
              umma_f16(d, ad0 + 2, bd0 + 2, p.idesc, 1u);
              umma_f16(d, ad0 + 4, bd0 + 4, p.idesc, 1u);
              umma_f16(d, ad0 + 6, bd0 + 6, p.idesc, 1u);
            } else {
              const int nks = (int)((tk >> 16) & 7u);
              uint32_t acc = acc0;
              for (int ks = 0; ks < nks; ++ks) {
                umma_f16(d, ad0 + (uint64_t)(ks * 2), bd0 + (uint64_t)(ks * 2), p.idesc, acc);
                acc = 1u;
              }
            }
          }
          PROF_ADD(pm_m, pq_t0);
          PROF_T0(pq_t0);
          umma_commit(&a_empty[s]);
          if (!p.b_merged) umma_commit(&b_empty[bs]);
          if (tk & kTkGroupEnd) umma_commit(&t_full[h]);
          PROF_ADD(pm_c, pq_t0);
        }
        __syncwarp();
        PROF_ADD(pm_iss, pm_t0);
        PROF_INC(pm_n);
        if (++bs == kBStages) { bs = 0; bph ^= 1u; }
        if (++s == S) { s = 0; ph ^= 1u; }
      }
      if (lane == 0 && mw == 0) {
        PROF_PUT(3, pm_wa); PROF_PUT(4, pm_wt); PROF_PUT(5, pm_wb); PROF_PUT(6, pm_iss); PROF_PUT(7, pm_n);
        PROF_PUT(24, PROF_NOW() - pm_start);
      }
      if (mw == 0) {
        // the elected lane's private counters (max over lanes: the others hold 0)
        PROF_PUT(25, __reduce_max_sync(0xffffffffu, (unsigned)pm_f)); PROF_PUT(26, __reduce_max_sync(0xffffffffu, (unsigned)pm_m));
        PROF_PUT(27, __reduce_max_sync(0xffffffffu, (unsigned)pm_c));
      }
    }
    __syncwarp();
  } else if (warp == kWarpWeights) {
    // ===================== weight loader =====================
    uint8_t *sched = sched_all + 1 * sched_stride;
    const uint32_t b_base = smem_u32(b_ring);
    Ring br;
    for (int g = gq_ctl.first(role_is_idx, lane); g >= 0; g = gq_ctl.next(role_is_idx, lane, g)) {
      const int64_t tile0 = (int64_t)g * p.T;
      fill_sched(sched, p, tile0, ntiles, lane);
      for (int k = 0; k < p.K; ++k) {
        if (!sched[k]) continue;
        for (int ch = 0; ch < p.nchunk; ++ch) {
          // merged ring: the image rides on the stage's own barriers (one wait and one
          // commit less per stage on the MMA warps, whose dependent chain bounds the kernel)
          uint64_t *empty = p.b_merged ? &a_empty[br.stage] : &b_empty[br.stage];
          uint64_t *full = p.b_merged ? &a_full[br.stage] : &b_full[br.stage];
          mbar_wait_warp(empty, br.phase ^ 1u);
          if (lane == 0) {
            const uint8_t *src = (const uint8_t *)p.bpack + (((size_t)sp * p.K + k) * p.nchunk + ch) * b_bytes;
            mbar_arrive_expect_tx(full, b_bytes);
            bulk_g2s(b_base + br.stage * b_bytes, src, b_bytes, full);
          }
          __syncwarp();
          br.advance(p.b_merged ? S : kBStages);
        }
      }
    }
    if (p.b_merged) {
      // the stage that carries the END ticket counts an arrival of this warp as well
      mbar_wait_warp(&a_empty[br.stage], br.phase ^ 1u);
      if (lane == 0) mbar_arrive(&a_full[br.stage]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kWarpMma0) {
    tc_fence_after();
    tmem_dealloc_512(tmem_base);
  }
}

// y[i] = sum_z part[z][i] in the fixed order z = 0, 1, ... (n is a multiple of 16: c_out is)
template <typename TOut>
__global__ void __launch_bounds__(256)
k_ksplit_reduce(const float *__restrict__ part, int ks, int64_t n, TOut *__restrict__ y) {
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += (int64_t)gridDim.x * blockDim.x * 4) {
    float4 a = __ldg(reinterpret_cast<const float4 *>(part + i));
    for (int z = 1; z < ks; ++z) {
      const float4 v = __ldg(reinterpret_cast<const float4 *>(part + (size_t)z * n + i));
      a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
    }
    if constexpr (sizeof(TOut) == 4) {
      *reinterpret_cast<float4 *>(y + i) = a;
    } else {
      uint2 o;
      o.x = pack2<TOut>(a.x, a.y);
      o.y = pack2<TOut>(a.z, a.w);
      *reinterpret_cast<uint2 *>(y + i) = o;
    }
  }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
static int arm_hang_report() {
  // per device: the symbol lives in each device's copy of the module
  static bool done[64] = {};
  int dev = 0;
  MK_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) dev = 0;
  if (!done[dev]) {
    unsigned int *d = hang_report_device_ptr();
    MK_CUDA(cudaMemcpyToSymbol(sm100::mk_hang_slot, &d, sizeof(d)));
    done[dev] = true;
  }
  return MK_OK;
}

struct TCPlan {
  int nsplit, NT, nchunk, T, W, a_stages, b_stages, i_slots, b_merged, half_cols;
  int ksplit, kper;       // offsets split over grid.z with fp32 partial outputs (small maps), else 1 / K
  size_t smem, bpack_bytes, mask_bytes, range_bytes, part_bytes;
};

static inline size_t align256s(size_t b) { return (b + 255) & ~(size_t)255; }

// Upper bound on the offsets one accumulator may see (0 = none): set by the fp32 split path around its
// calls (conv_f32.cu), whose bf16-piece products must not chain more than ~400 MMAs into one fp32
// TMEM accumulator (tcgen05 accumulates with truncation); the plan then splits the offsets
// over grid.z however many CTAs the layer has, and the partial outputs are added in fp32 (round to nearest).
static thread_local int g_kper_max = 0;
void conv_fwd_tc_set_kper_max(int kper_max) { g_kper_max = kper_max; }

// One candidate: T tiles per group on accumulator halves `half_cols` apart (0 = a single set).
static bool plan_with(int dtype, int64_t n_out, int K, int c_in, int c_out, int T, int half_cols, TCPlan &pl) {
  int nsplit = (c_out + 255) / 256;
  while (c_out % (16 * nsplit)) ++nsplit;
  pl.nsplit = nsplit;
  pl.NT = c_out / nsplit;
  pl.nchunk = (c_in + 63) / 64;
  pl.T = T;
  pl.half_cols = half_cols;
  pl.b_stages = 2;
  pl.i_slots = 4;
  // development knobs (tuning sweeps): MK_TC_BSTAGES / MK_TC_ISLOTS
  if (const char *e = getenv("MK_TC_BSTAGES")) pl.b_stages = std::max(2, std::min(kMaxBStages, atoi(e)));
  // (a multiple of 4: the teams that use an idx slot must be the same every time round the ring)
  if (const char *e = getenv("MK_TC_ISLOTS")) pl.i_slots = std::max(4, std::min(kMaxIdxSlots, atoi(e) & ~3));
  pl.W = (K + 31) / 32;
  const size_t b_bytes = (size_t)pl.NT * 128;
  size_t fixed = 1024 /*align slack*/ + pl.b_stages * b_bytes + (size_t)pl.i_slots * pl.T * kTileM * 4 +
                 kSchedCopies * (size_t)((K + 31) & ~31) + 64 + 1024 /*barriers*/;
  const size_t budget = 227 * 1024;
  const size_t stage_bytes = (size_t)pl.T * kTileBytes;
  // Four tiles on one accumulator set may run on a ring of TWO 64 KB stages (128 output columns: three do
  // not fit beside the weight images).  What bounds these kernels is the MMA warps' fixed chain per stage
  // (~1400 cycles), so a stage with twice the MMA work beats a deeper ring of small stages:
  // 128 -> 128 forward 0.795 -> 0.645 ms against two tiles on each of two sets and five stages.
  const int s_min = (T == 4 && half_cols == 0 && !getenv("MK_TC_S3")) ? 2 : 3;
  if (fixed + s_min * stage_bytes > budget) return false;
  int S = (int)((budget - fixed) / stage_bytes);
  S = std::min(S, 12);   // tickets[] holds 16
  if (S < s_min || S * pl.T <= 4) return false;   // producer teams: consecutive stages of a team must be < S apart
  pl.a_stages = S;
  // Whatever is left after the A ring goes to the weight ring: with two images the copy of
  // the image for stage n+2 starts when stage n retires, one stage period before it is
  // needed, and the MMA warps wait for it (10 % of their time at 64 -> 64, -DMK_TC_PROF).
  // If there is room for one weight image per A stage, the weight ring is merged into the A
  // ring: image and gathered tiles of a stage share the stage's full / empty barriers.
  pl.b_merged = 0;
  if (!getenv("MK_TC_BSTAGES")) {
    while (pl.b_stages < S && pl.b_stages < kMaxBStages && fixed + b_bytes + (size_t)S * stage_bytes <= budget) {
      ++pl.b_stages;
      fixed += b_bytes;
    }
    pl.b_merged = (pl.b_stages == S && !getenv("MK_TC_NOMERGE")) ? 1 : 0;
  }
  pl.smem = fixed + (size_t)S * stage_bytes;
  pl.bpack_bytes = align256s((size_t)nsplit * K * pl.nchunk * b_bytes);
  pl.mask_bytes = align256s((size_t)((n_out + kTileM - 1) / kTileM) * pl.W * 4);
  pl.range_bytes = 256;   // group counters, one int per output split (and offset split)
  // Small maps: a layer with fewer groups than SMs leaves most of the GPU idle while a handful of CTAs
  // stream ALL K weight images (27 x 960 x 960 on a 206-row latent map: 8 CTAs, 335 us).  The offsets
  // are then split over grid.z; every CTA writes an fp32 partial output and a second kernel adds them in
  // a fixed order.  Partials are bounded by the SM count: <= 148 x T x 128 rows x NT columns x 4 B.
  pl.ksplit = 1;
  pl.kper = K;
  pl.part_bytes = 0;
  const int64_t ngroups = ((n_out + kTileM - 1) / kTileM + pl.T - 1) / pl.T;
  const int64_t ctas = ngroups * pl.nsplit;
  if (dtype == MK_BF16 && K >= 2 && ctas * 2 <= num_sms() && !getenv("MK_TC_NOKSPLIT")) {
    int ks = (int)std::min<int64_t>(K, num_sms() / ctas);
    ks = std::min(ks, 64 / pl.nsplit);
    if (ks >= 2) {
      pl.kper = (K + ks - 1) / ks;
      pl.ksplit = (K + pl.kper - 1) / pl.kper;
    }
  }
  if (dtype == MK_BF16 && g_kper_max > 0 && pl.kper > g_kper_max) {
    pl.kper = g_kper_max;
    pl.ksplit = (K + pl.kper - 1) / pl.kper;
    if (pl.nsplit * pl.ksplit > 64) return false;   // one group counter per (offset split, output split)
  }
  if (pl.ksplit > 1) pl.part_bytes = align256s((size_t)pl.ksplit * n_out * c_out * 4);
  return true;
}

static bool make_plan(int dtype, int64_t n_out, int K, int c_in, int c_out, TCPlan &pl) {
  if (dtype != MK_BF16 && dtype != MK_F16) return false;
  if (c_in < 16 || c_in % 16 || c_out < 16 || c_out % 16 || K < 1 || K > kMaxK || n_out < 1) return false;
  int nsplit = (c_out + 255) / 256;
  while (c_out % (16 * nsplit)) ++nsplit;
  const int acc_stride = (c_out / nsplit + 31) & ~31;
  int t_two = std::min(kMaxT, 256 / acc_stride);
  int t_one = std::min(kMaxT, 512 / acc_stride);
  if (const char *e = getenv("MK_TC_T")) {   // development knob
    t_two = std::max(1, std::min(t_two, atoi(e)));
    t_one = std::max(1, std::min(t_one, atoi(e)));
  }
  // Wide outputs: with two accumulator sets only 256 / NT tiles share a weight image (ONE at NT = 256: 32 KB
  // of weights re-read per 16 KB of gathered rows, and the SM ingests ~64 B/clk whatever the source).  One
  // set of up to 512 columns lets more tiles share an image; the price is that a group's first MMA waits for
  // the previous group's epilogue (a few thousand cycles per ~100 k cycles of MMA work at these widths).
  // Taken when it at least doubles the tiles per image and the A ring still holds three stages (two for
  // four tiles, see plan_with; three tiles on one set measured slower than two on each of two sets).
  if (!getenv("MK_TC_TWO_SETS"))
    for (int t = t_one; t >= 2 * t_two; --t)
      if (plan_with(dtype, n_out, K, c_in, c_out, t, 0, pl)) return true;
  return plan_with(dtype, n_out, K, c_in, c_out, t_two, 256, pl);
}

bool conv_fwd_tc_eligible(int dtype, int64_t n_out, int K, int c_in, int c_out) {
  TCPlan pl;
  return make_plan(dtype, n_out, K, c_in, c_out, pl);
}

int64_t conv_fwd_tc_workspace(int64_t n_out, int K, int c_in, int c_out) {
  TCPlan pl;
  if (!make_plan(MK_BF16, std::max<int64_t>(n_out, 1), K, c_in, c_out, pl)) return 0;
  return (int64_t)(pl.bpack_bytes + pl.mask_bytes + pl.range_bytes + pl.part_bytes + 256);
}

int tile_mask_launch(const int32_t *nbr, int64_t n_rows, int64_t ld, int K, uint32_t *mask, cudaStream_t st) {
  const int W = (K + 31) / 32;
  const int64_t ntiles = (n_rows + kTileM - 1) / kTileM;
  if (ntiles == 0) return MK_OK;
  k_tile_mask<<<(unsigned)ntiles, 256, W * sizeof(uint32_t), st>>>(nbr, n_rows, ld, K, W, mask);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

template <typename T, typename TOut>
static int conv_fwd_tc_typed(const TCPlan &pl, int fmt, const void *x, const void *w, int transpose_w, int kflip,
                             const int32_t *nbr, const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in,
                             int K, int c_in, int c_out, void *y, const int32_t *out_rows, void *workspace,
                             cudaStream_t st) {
  uint8_t *ws = (uint8_t *)workspace;
  T *bpack = (T *)ws;
  uint32_t *mask = (uint32_t *)(ws + pl.bpack_bytes);
  // group counters (one per output split) behind the mask; development switch MK_TC_STATIC
  int *group_counter = getenv("MK_TC_STATIC") ? nullptr : (int *)(ws + pl.bpack_bytes + pl.mask_bytes);
  {
    const int64_t total = (int64_t)pl.nsplit * K * pl.nchunk * pl.NT * 64;
    if (!transpose_w && !getenv("MK_TC_PACK_FLAT")) {
      dim3 pg((unsigned)ceil_div(pl.nsplit * pl.NT, 64), (unsigned)pl.nchunk, (unsigned)K);
      k_pack_b_tiled<T><<<pg, 256, 0, st>>>((const T *)w, K, c_in, c_out, pl.nchunk, pl.NT, pl.nsplit, bpack,
                                            group_counter);
    } else if (transpose_w && (c_in % 8) == 0 && ((uintptr_t)w & 15) == 0 && total / 8 < (1ll << 31) &&
               !getenv("MK_TC_PACK_FLAT")) {
      const int blocks = (int)std::min<int64_t>(ceil_div(total / 8, 256), (int64_t)num_sms() * 8);
      k_pack_b_rows<T><<<blocks, 256, 0, st>>>((const T *)w, K, c_in, c_out, pl.nchunk, pl.NT, pl.nsplit, bpack,
                                               group_counter);
    } else {
      const int blocks = (int)std::min<int64_t>(ceil_div(total, 256), (int64_t)num_sms() * 8);
      k_pack_b<T><<<blocks, 256, 0, st>>>((const T *)w, transpose_w, K, c_in, c_out, pl.nchunk, pl.NT, pl.nsplit,
                                          bpack, group_counter);
    }
    MK_LAUNCH_CHECK();
  }
  if (tile_mask == nullptr) {
    int rc = tile_mask_launch(nbr, n_out, ld, K, mask, st);
    if (rc) return rc;
    tile_mask = mask;
  }
  ConvTCParams p;
  p.x = x;
  p.bpack = bpack; p.nbr = nbr; p.mask = tile_mask; p.y = y; p.out_rows = out_rows;
  p.ld = ld; p.n_out = n_out; p.n_in = n_in; p.K = K; p.W = pl.W; p.kflip = kflip;
  p.c_in = c_in; p.c_out = c_out; p.nchunk = pl.nchunk; p.NT = pl.NT; p.T = pl.T;
  p.a_stages = pl.a_stages; p.b_stages = pl.b_stages; p.i_slots = pl.i_slots; p.b_merged = pl.b_merged; p.group_counter = group_counter;
  p.half_cols = pl.half_cols;
  p.ksplit = pl.ksplit;
  p.kper = pl.kper;
  if (pl.ksplit > 1 && group_counter)   // one counter per (offset split, output split); k_pack_b zeroed the first nsplit
    MK_CUDA(cudaMemsetAsync(group_counter, 0, sizeof(int) * pl.nsplit * pl.ksplit, st));
  p.lean = (pl.T == 4 && pl.nchunk == 1 && pl.W == 1 && (c_in % 64) == 0 && !getenv("MK_TC_NOLEAN")) ? 1 : 0;
  if (const char *e = getenv("MK_TC_LEAN")) p.lean = p.lean ? std::max(1, std::min(2, atoi(e))) : 0;   // development knob
  CUtensorMap tmap_x;
  memset(&tmap_x, 0, sizeof(tmap_x));
  if (p.lean == 2) {      // the feature matrix as [n_in rows x 64 channels], one-row box: the tile::gather4 form
    int rc = make_tmap_rows16(&tmap_x, std::is_same<T, __nv_bfloat16>::value ? MK_BF16 : MK_F16, x, n_in, c_in, 64, 1);
    if (rc) return rc;
  }
  const int64_t ntiles = (n_out + kTileM - 1) / kTileM;
  p.ngroups = (int)((ntiles + pl.T - 1) / pl.T);
  p.idesc = idesc_f16(fmt, kTileM, pl.NT, 0, 0);
  { int rc = arm_hang_report(); if (rc) return rc; }
  const bool full = (c_in % 64) == 0;
  // (set on every call: the attribute is per device and the call costs about a microsecond)
  if (full)
    MK_CUDA(cudaFuncSetAttribute(k_conv_tc<T, true, TOut>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  else
    MK_CUDA(cudaFuncSetAttribute(k_conv_tc<T, false, TOut>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  const int gx = (int)std::min<int64_t>((int64_t)p.ngroups, std::max(1, num_sms() / (pl.nsplit * pl.ksplit)));
  dim3 grid(gx, pl.nsplit, pl.ksplit);
  if (full)
    k_conv_tc<T, true, TOut><<<grid, kNumThreadsTC, pl.smem, st>>>(p, tmap_x);
  else
    k_conv_tc<T, false, TOut><<<grid, kNumThreadsTC, pl.smem, st>>>(p, tmap_x);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int conv_fwd_tc(int dtype, const void *x, const void *w, int transpose_w, int kflip, const int32_t *nbr,
                const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                void *y, const int32_t *out_rows, void *workspace, cudaStream_t st, int out_f32) {
  TCPlan pl;
  if (!make_plan(dtype, n_out, K, c_in, c_out, pl)) {
    set_error("conv_fwd_tc: shape not eligible");
    return MK_EUNSUPPORTED;
  }
  if (dtype == MK_BF16 && pl.ksplit > 1) {
    float *part = (float *)((uint8_t *)workspace + pl.bpack_bytes + pl.mask_bytes + pl.range_bytes);
    int rc = conv_fwd_tc_typed<__nv_bfloat16, float>(pl, 1, x, w, transpose_w, kflip, nbr, tile_mask, ld, n_out, n_in,
                                                     K, c_in, c_out, part, out_rows, workspace, st);
    if (rc) return rc;
    const int64_t n_el = n_out * (int64_t)c_out;
    const unsigned blocks = (unsigned)std::min<int64_t>(ceil_div(n_el / 4, 256), (int64_t)num_sms() * 16);
    if (out_f32)
      k_ksplit_reduce<float><<<blocks, 256, 0, st>>>(part, pl.ksplit, n_el, (float *)y);
    else
      k_ksplit_reduce<__nv_bfloat16><<<blocks, 256, 0, st>>>(part, pl.ksplit, n_el, (__nv_bfloat16 *)y);
    MK_LAUNCH_CHECK();
    return MK_OK;
  }
  if (dtype == MK_BF16) {
    if (out_f32)
      return conv_fwd_tc_typed<__nv_bfloat16, float>(pl, 1, x, w, transpose_w, kflip, nbr, tile_mask, ld, n_out, n_in,
                                                     K, c_in, c_out, y, out_rows, workspace, st);
    return conv_fwd_tc_typed<__nv_bfloat16, __nv_bfloat16>(pl, 1, x, w, transpose_w, kflip, nbr, tile_mask, ld, n_out,
                                                           n_in, K, c_in, c_out, y, out_rows, workspace, st);
  }
  if (out_f32) {
    set_error("conv_fwd_tc: fp32 output is served for bf16 operands only");
    return MK_EUNSUPPORTED;
  }
  return conv_fwd_tc_typed<__half, __half>(pl, 0, x, w, transpose_w, kflip, nbr, tile_mask, ld, n_out, n_in, K, c_in,
                                           c_out, y, out_rows, workspace, st);
}

// weight gradient on tensor cores: conv_wgrad_tc.cu
}  // namespace mk

extern "C" int mk_has_tcgen05(void) { return 1; }

#ifdef MK_TC_PROF
extern "C" int mk_debug_prof_read(unsigned long long *host, int n) {
  return (int)cudaMemcpyFromSymbol(host, mk::g_prof, sizeof(unsigned long long) * n);
}
#endif

extern "C" int64_t mk_kmap_tile_mask_words(int K) { return (K + 31) / 32; }

extern "C" int mk_kmap_tile_mask(const int32_t *nbr, int64_t n_rows, int64_t ld, int K, uint32_t *mask,
                                 void *stream) {
  MK_CHECK_ARG(K >= 1 && ld >= n_rows, "bad K / ld");
  if (n_rows == 0) return MK_OK;
  MK_CHECK_ARG(nbr && mask, "null pointer");
  return mk::tile_mask_launch(nbr, n_rows, ld, K, mask, (cudaStream_t)stream);
}
