// Tensor-core (tcgen05 / TMEM) sparse convolution, forward and data gradient.
//
// Replaces the matmul<T,uint,TILE> gather-GEMM-scatter kernels behind
// ConvolutionForwardGPU / ConvolutionBackwardGPU (pybind/extern.hpp:84-114;
// SURVEY §2.3): all kernel offsets in ONE persistent launch, output-stationary,
// accumulators in TMEM, no atomics.
//
//   y[o, n0:n0+NT] = sum_k  x[nbr[kk(k), o], :] @ B_k          (fp32 accumulate)
//
// CTA roles (352 threads, 1 CTA / SM, all 512 TMEM columns):
//   warps 0-3  epilogue : tcgen05.ld accumulators -> bf16/f16 -> global
//   warp  4    indices  : one lane bulk-copies the 128 neighbour-table entries
//                         of every (tile, offset) into a ring in shared memory
//   warps 5-8  gather   : LDGSTS 16-byte chunks of the neighbour rows into a
//                         ring of 128-row x 128-byte A stages (128B swizzle),
//                         missing neighbours zero-filled; stages are dealt
//                         round-robin to the four warps
//   warp  9    MMA      : one lane issues tcgen05.mma (M=128, N=NT, K=16)
//   warp  10   weights  : one lane bulk-copies the pre-swizzled B image of each
//                         (offset, channel chunk) into a ring
// A CTA owns "groups" of T 128-row output tiles whose accumulators live side by
// side in one half of TMEM; the two halves ping-pong so the epilogue of one
// group overlaps the main loop of the next.  Which (tile, offset) pairs have
// any neighbour at all is known up front from the kernel map's tile mask, so
// every role walks the same static schedule and empty pairs are skipped.
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "sm100.cuh"
#include "tmap.cuh"

namespace mk {
using namespace sm100;

constexpr int kTileM = 128;
constexpr int kStageBytes = kTileM * 128;  // one A stage: 128 rows x 64 16-bit channels
constexpr int kNumThreadsTC = 352;
constexpr int kGatherWarps = 4;
constexpr int kMaxT = 4;                   // tiles per group (256 TMEM columns / NT)
constexpr int kIdxSlots = 32;              // ring of 512-byte neighbour-index rows
constexpr int kBStages = 4;                // weight-image ring
constexpr int kMaxK = 128;

struct ConvTCParams {
  const void *x;          // [n_in, c_in] 16-bit
  const void *bpack;      // [nsplit][K][nchunk][NT rows x 128 B] pre-swizzled
  const int32_t *nbr;     // [K, ld]
  const uint32_t *mask;   // [ntiles, W]
  void *y;                // [n_out, c_out]
  int64_t ld, n_out, n_in;
  int K, W, kflip;
  int c_in, c_out, nchunk, NT, T;
  int a_stages;
  int ngroups;
  uint32_t idesc;
};

// ring position: stage index + phase parity, advanced without division
struct Ring {
  uint32_t stage = 0, phase = 0;
  __device__ __forceinline__ void advance(uint32_t n) {
    if (++stage == n) { stage = 0; phase ^= 1u; }
  }
};

// B image of W for the MMA: per (split, k, chunk) NT rows (n) x 64 reduction
// channels (128 B, K-major), 16-byte chunks XOR-swizzled with (n & 7).
//   transpose == 0: B(n, c) = W[k][c][n0 + n]        W: [K, c_red, c_n]
//   transpose == 1: B(n, c) = W[k][n0 + n][c]        W: [K, c_n, c_red]
template <typename T>
__global__ void k_pack_b(const T *__restrict__ w, int transpose, int K, int c_red, int c_n, int nchunk, int NT,
                         int nsplit, T *__restrict__ out) {
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
__device__ __forceinline__ void fill_sched(uint8_t *sched, const ConvTCParams &p, int64_t tile0, int64_t ntiles,
                                           int lane) {
  __syncwarp();
  for (int k = lane; k < p.K; k += 32) {
    const int kk = p.kflip ? p.K - 1 - k : k;
    uint32_t bits = 0;
    for (int t = 0; t < p.T; ++t) {
      const int64_t tile = tile0 + t;
      if (tile < ntiles) bits |= ((__ldg(p.mask + tile * p.W + (kk >> 5)) >> (kk & 31)) & 1u) << t;
    }
    sched[k] = (uint8_t)bits;
  }
  __syncwarp();
}

template <typename T>
__global__ void __launch_bounds__(kNumThreadsTC, 1)
k_conv_tc(const __grid_constant__ ConvTCParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment for the 128B-swizzle atoms
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t S = (uint32_t)p.a_stages;
  const uint32_t b_bytes = (uint32_t)p.NT * 128u;
  uint8_t *a_ring = smem;
  uint8_t *b_ring = a_ring + (size_t)S * kStageBytes;
  int32_t *idx_ring = (int32_t *)(b_ring + kBStages * b_bytes);
  uint8_t *sched_all = (uint8_t *)(idx_ring + kIdxSlots * kTileM);   // 4 role copies x kMaxK
  uint64_t *bars = (uint64_t *)(sched_all + 7 * kMaxK);
  uint64_t *a_full = bars, *a_empty = a_full + S;
  uint64_t *b_full = a_empty + S, *b_empty = b_full + kBStages;
  uint64_t *i_full = b_empty + kBStages, *i_empty = i_full + kIdxSlots;
  uint64_t *t_full = i_empty + kIdxSlots, *t_empty = t_full + 2;
  uint32_t *tmem_slot = (uint32_t *)(t_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t ntiles = (p.n_out + kTileM - 1) / kTileM;
  const int sp = blockIdx.y;
  const int acc_stride = (p.NT + 31) & ~31;

  if (threadIdx.x == 0) {
    for (uint32_t s = 0; s < S; ++s) { mbar_init(&a_full[s], 32); mbar_init(&a_empty[s], 1); }
    for (int s = 0; s < kBStages; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
    for (int s = 0; s < kIdxSlots; ++s) { mbar_init(&i_full[s], 1); mbar_init(&i_empty[s], (uint32_t)p.nchunk); }
    for (int s = 0; s < 2; ++s) { mbar_init(&t_full[s], 1); mbar_init(&t_empty[s], 4); }
    fence_barrier_init();
  }
  if (warp == 9) tmem_alloc_512(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    // ===================== epilogue =====================
    uint32_t it = 0;
    for (int g = blockIdx.x; g < p.ngroups; g += gridDim.x, ++it) {
      const uint32_t h = it & 1u;
      const int64_t tile0 = (int64_t)g * p.T;
      mbar_wait(&t_full[h], (it >> 1) & 1u);
      tc_fence_after();
      for (int t = 0; t < p.T; ++t) {
        const int64_t tile = tile0 + t;
        if (tile >= ntiles) break;
        uint32_t touched = 0;
        for (int wi = 0; wi < p.W; ++wi) touched |= __ldg(p.mask + tile * p.W + wi);
        const int64_t row = tile * kTileM + warp * 32 + lane;
        T *yrow = (T *)p.y + (size_t)row * p.c_out + (size_t)sp * p.NT;
        const uint32_t tbase = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(h * 256 + t * acc_stride);
        for (int c0 = 0; c0 < p.NT; c0 += 16) {
          uint32_t v[16];
          if (touched) {
            tmem_ld16(tbase + (uint32_t)c0, v);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = 0u;
          }
          if (row < p.n_out) {
            uint4 o0, o1;
            o0.x = pack2<T>(__uint_as_float(v[0]), __uint_as_float(v[1]));
            o0.y = pack2<T>(__uint_as_float(v[2]), __uint_as_float(v[3]));
            o0.z = pack2<T>(__uint_as_float(v[4]), __uint_as_float(v[5]));
            o0.w = pack2<T>(__uint_as_float(v[6]), __uint_as_float(v[7]));
            o1.x = pack2<T>(__uint_as_float(v[8]), __uint_as_float(v[9]));
            o1.y = pack2<T>(__uint_as_float(v[10]), __uint_as_float(v[11]));
            o1.z = pack2<T>(__uint_as_float(v[12]), __uint_as_float(v[13]));
            o1.w = pack2<T>(__uint_as_float(v[14]), __uint_as_float(v[15]));
            *reinterpret_cast<uint4 *>(yrow + c0) = o0;
            *reinterpret_cast<uint4 *>(yrow + c0 + 8) = o1;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&t_empty[h]);
    }
  } else if (warp == 4) {
    // ===================== neighbour-index prefetch =====================
    // one 512-byte row of the table per (offset, tile) -> idx ring; the channel
    // chunks of a stage reuse the same indices, so the gather warp re-reads them
    uint8_t *sched = sched_all + 0 * kMaxK;
    const uint32_t i_base = smem_u32(idx_ring);
    Ring ir;
    for (int g = blockIdx.x; g < p.ngroups; g += gridDim.x) {
      const int64_t tile0 = (int64_t)g * p.T;
      fill_sched(sched, p, tile0, ntiles, lane);
      for (int k = 0; k < p.K; ++k) {
        const uint32_t bits = sched[k];
        if (!bits) continue;
        const int kk = p.kflip ? p.K - 1 - k : k;
        for (int t = 0; t < p.T; ++t) {
          if (!((bits >> t) & 1u)) continue;
          mbar_wait(&i_empty[ir.stage], ir.phase ^ 1u);
          if (lane == 0) {
            const int32_t *src = p.nbr + (int64_t)kk * p.ld + (tile0 + t) * kTileM;
            mbar_arrive_expect_tx(&i_full[ir.stage], kTileM * 4);
            bulk_g2s(i_base + ir.stage * (kTileM * 4), src, kTileM * 4, &i_full[ir.stage]);
          }
          __syncwarp();
          ir.advance(kIdxSlots);
        }
      }
    }
  } else if (warp < 5 + kGatherWarps) {
    // ===================== gather producers (LDGSTS) =====================
    // lane = (rsub, c16): instruction i of a stage moves 16-byte chunk c16 of tile
    // rows rsub*32 + i, i.e. four 128-byte feature rows per warp instruction
    const int w = warp - 5;
    uint8_t *sched = sched_all + (1 + w) * kMaxK;
    const uint32_t a_base = smem_u32(a_ring);
    const uint32_t row_bytes = (uint32_t)p.c_in * 2u;
    const int c16 = lane & 7, rsub = lane >> 3;
    Ring ir, ar;
    uint32_t e = 0;
    for (int g = blockIdx.x; g < p.ngroups; g += gridDim.x) {
      const int64_t tile0 = (int64_t)g * p.T;
      fill_sched(sched, p, tile0, ntiles, lane);
      for (int k = 0; k < p.K; ++k) {
        const uint32_t bits = sched[k];
        if (!bits) continue;
        Ring irt = ir;
        for (int ch = 0; ch < p.nchunk; ++ch) {
          irt = ir;
          const uint32_t szok = (ch * 64 + c16 * 8) < p.c_in ? 16u : 0u;
          const uint8_t *xb = (const uint8_t *)p.x + (size_t)ch * 128 + (size_t)c16 * 16;
          for (int t = 0; t < p.T; ++t) {
            if (!((bits >> t) & 1u)) continue;
            if ((e & (kGatherWarps - 1)) == (uint32_t)w) {
              mbar_wait(&i_full[irt.stage], irt.phase);
              const int4 *ip = reinterpret_cast<const int4 *>(idx_ring + irt.stage * kTileM + rsub * 32);
              int4 iv[8];
#pragma unroll
              for (int q = 0; q < 8; ++q) iv[q] = ip[q];
              __syncwarp();
              if (lane == 0) mbar_arrive(&i_empty[irt.stage]);
              mbar_wait(&a_empty[ar.stage], ar.phase ^ 1u);
              // the 8 swizzled chunk positions of this lane inside the stage
              const uint32_t sb = a_base + ar.stage * kStageBytes + (uint32_t)rsub * 4096u;
              uint32_t sbx[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) sbx[j] = sb + (uint32_t)((c16 ^ j) << 4);
              // table entries are -1 for a missing neighbour and for rows past n_out
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                const int32_t ri = (i & 3) == 0 ? iv[i >> 2].x : (i & 3) == 1 ? iv[i >> 2].y
                                   : (i & 3) == 2 ? iv[i >> 2].z : iv[i >> 2].w;
                cp_async16(sbx[i & 7] + (uint32_t)i * 128u, xb + (size_t)(uint32_t)max(ri, 0) * row_bytes,
                           ri >= 0 ? szok : 0u);
              }
              cp_async_arrive_noinc(&a_full[ar.stage]);
            }
            irt.advance(kIdxSlots);
            ar.advance(S);
            ++e;
          }
        }
        ir = irt;
      }
    }
  } else if (warp == 9) {
    // ===================== MMA issuer =====================
    uint8_t *sched = sched_all + 5 * kMaxK;
    const uint32_t a_base = smem_u32(a_ring), b_base = smem_u32(b_ring);
    // descriptor template: SBO = 1024 B (8-row atoms), LBO = 16 B (unused for one
    // 128-byte K atom), version 1, SWIZZLE_128B; the start address is added per MMA
    const uint64_t desc0 = smem_desc_sw128(0, 16, 1024);
    Ring ar, br;
    uint32_t it = 0;
    for (int g = blockIdx.x; g < p.ngroups; g += gridDim.x, ++it) {
      const uint32_t h = it & 1u;
      const int64_t tile0 = (int64_t)g * p.T;
      fill_sched(sched, p, tile0, ntiles, lane);
      mbar_wait(&t_empty[h], ((it >> 1) & 1u) ^ 1u);
      tc_fence_after();
      uint32_t fresh = 0xFu;   // tiles whose accumulator has not been written in this group
      for (int k = 0; k < p.K; ++k) {
        const uint32_t bits = sched[k];
        if (!bits) continue;
        for (int ch = 0; ch < p.nchunk; ++ch) {
          mbar_wait(&b_full[br.stage], br.phase);
          const int nks = min(4, (p.c_in - ch * 64) >> 4);
          const uint64_t bd0 = desc0 | (uint64_t)(((b_base + br.stage * b_bytes) & 0x3FFFFu) >> 4);
#pragma unroll
          for (int t = 0; t < kMaxT; ++t) {
            if (t < p.T && ((bits >> t) & 1u)) {
              mbar_wait(&a_full[ar.stage], ar.phase);
              fence_proxy_async_smem();
              tc_fence_after();
              if (elect_one()) {
                const uint32_t d = tmem_base + (uint32_t)(h * 256 + t * acc_stride);
                const uint64_t ad0 = desc0 | (uint64_t)(((a_base + ar.stage * kStageBytes) & 0x3FFFFu) >> 4);
                uint32_t acc = ((fresh >> t) & 1u) ? 0u : 1u;
                for (int ks = 0; ks < nks; ++ks) {
                  umma_f16(d, ad0 + (uint64_t)(ks * 2), bd0 + (uint64_t)(ks * 2), p.idesc, acc);
                  acc = 1u;
                }
                umma_commit(&a_empty[ar.stage]);
              }
              __syncwarp();
              fresh &= ~(1u << t);
              ar.advance(S);
            }
          }
          if (elect_one()) umma_commit(&b_empty[br.stage]);
          __syncwarp();
          br.advance(kBStages);
        }
      }
      if (elect_one()) umma_commit(&t_full[h]);
      __syncwarp();
    }
  } else {
    // ===================== weight loader =====================
    uint8_t *sched = sched_all + 6 * kMaxK;
    const uint32_t b_base = smem_u32(b_ring);
    Ring br;
    for (int g = blockIdx.x; g < p.ngroups; g += gridDim.x) {
      const int64_t tile0 = (int64_t)g * p.T;
      fill_sched(sched, p, tile0, ntiles, lane);
      for (int k = 0; k < p.K; ++k) {
        if (!sched[k]) continue;
        for (int ch = 0; ch < p.nchunk; ++ch) {
          mbar_wait(&b_empty[br.stage], br.phase ^ 1u);
          if (lane == 0) {
            const uint8_t *src = (const uint8_t *)p.bpack + (((size_t)sp * p.K + k) * p.nchunk + ch) * b_bytes;
            mbar_arrive_expect_tx(&b_full[br.stage], b_bytes);
            bulk_g2s(b_base + br.stage * b_bytes, src, b_bytes, &b_full[br.stage]);
          }
          __syncwarp();
          br.advance(kBStages);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc_512(tmem_base);
  }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
struct TCPlan {
  int nsplit, NT, nchunk, T, W, a_stages;
  size_t smem, bpack_bytes, mask_bytes;
};

static inline size_t align256s(size_t b) { return (b + 255) & ~(size_t)255; }

static bool make_plan(int dtype, int64_t n_out, int K, int c_in, int c_out, TCPlan &pl) {
  if (dtype != MK_BF16 && dtype != MK_F16) return false;
  if (c_in < 16 || c_in % 16 || c_out < 16 || c_out % 16 || K < 1 || K > kMaxK || n_out < 1) return false;
  int nsplit = (c_out + 255) / 256;
  while (c_out % (16 * nsplit)) ++nsplit;
  pl.nsplit = nsplit;
  pl.NT = c_out / nsplit;
  pl.nchunk = (c_in + 63) / 64;
  const int acc_stride = (pl.NT + 31) & ~31;
  pl.T = std::min(kMaxT, 256 / acc_stride);
  pl.W = (K + 31) / 32;
  const size_t b_bytes = (size_t)pl.NT * 128;
  const size_t fixed = 1024 /*align slack*/ + kBStages * b_bytes + kIdxSlots * kTileM * 4 + 7 * kMaxK +
                       1024 /*barriers*/;
  const size_t budget = 227 * 1024;
  int S = (int)((budget - fixed) / kStageBytes);
  S = std::min(S, 12);
  if (S < 3) return false;
  pl.a_stages = S;
  pl.smem = fixed + (size_t)S * kStageBytes;
  pl.bpack_bytes = align256s((size_t)nsplit * K * pl.nchunk * b_bytes);
  pl.mask_bytes = align256s((size_t)((n_out + kTileM - 1) / kTileM) * pl.W * 4);
  return true;
}

bool conv_fwd_tc_eligible(int dtype, int64_t n_out, int K, int c_in, int c_out) {
  TCPlan pl;
  return make_plan(dtype, n_out, K, c_in, c_out, pl);
}

int64_t conv_fwd_tc_workspace(int64_t n_out, int K, int c_in, int c_out) {
  TCPlan pl;
  if (!make_plan(MK_BF16, std::max<int64_t>(n_out, 1), K, c_in, c_out, pl)) return 0;
  return (int64_t)(pl.bpack_bytes + pl.mask_bytes + 256);
}

int tile_mask_launch(const int32_t *nbr, int64_t n_rows, int64_t ld, int K, uint32_t *mask, cudaStream_t st) {
  const int W = (K + 31) / 32;
  const int64_t ntiles = (n_rows + kTileM - 1) / kTileM;
  if (ntiles == 0) return MK_OK;
  k_tile_mask<<<(unsigned)ntiles, 256, W * sizeof(uint32_t), st>>>(nbr, n_rows, ld, K, W, mask);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

template <typename T>
static int conv_fwd_tc_typed(const TCPlan &pl, int fmt, const void *x, const void *w, int transpose_w, int kflip,
                             const int32_t *nbr, const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in,
                             int K, int c_in, int c_out, void *y, void *workspace, cudaStream_t st) {
  uint8_t *ws = (uint8_t *)workspace;
  T *bpack = (T *)ws;
  uint32_t *mask = (uint32_t *)(ws + pl.bpack_bytes);
  {
    const int64_t total = (int64_t)pl.nsplit * K * pl.nchunk * pl.NT * 64;
    const int blocks = (int)std::min<int64_t>(ceil_div(total, 256), (int64_t)num_sms() * 8);
    k_pack_b<T><<<blocks, 256, 0, st>>>((const T *)w, transpose_w, K, c_in, c_out, pl.nchunk, pl.NT, pl.nsplit,
                                        bpack);
    MK_LAUNCH_CHECK();
  }
  if (tile_mask == nullptr) {
    int rc = tile_mask_launch(nbr, n_out, ld, K, mask, st);
    if (rc) return rc;
    tile_mask = mask;
  }
  ConvTCParams p;
  p.x = x;
  p.bpack = bpack; p.nbr = nbr; p.mask = tile_mask; p.y = y;
  p.ld = ld; p.n_out = n_out; p.n_in = n_in; p.K = K; p.W = pl.W; p.kflip = kflip;
  p.c_in = c_in; p.c_out = c_out; p.nchunk = pl.nchunk; p.NT = pl.NT; p.T = pl.T;
  p.a_stages = pl.a_stages;
  const int64_t ntiles = (n_out + kTileM - 1) / kTileM;
  p.ngroups = (int)((ntiles + pl.T - 1) / pl.T);
  p.idesc = idesc_f16(fmt, kTileM, pl.NT, 0, 0);
  static bool attr_set[2] = {false, false};
  if (!attr_set[fmt]) {
    MK_CUDA(cudaFuncSetAttribute(k_conv_tc<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set[fmt] = true;
  }
  const int gx = (int)std::min<int64_t>((int64_t)p.ngroups, std::max(1, num_sms() / pl.nsplit));
  dim3 grid(gx, pl.nsplit);
  k_conv_tc<T><<<grid, kNumThreadsTC, pl.smem, st>>>(p);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

int conv_fwd_tc(int dtype, const void *x, const void *w, int transpose_w, int kflip, const int32_t *nbr,
                const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                void *y, void *workspace, cudaStream_t st) {
  TCPlan pl;
  if (!make_plan(dtype, n_out, K, c_in, c_out, pl)) {
    set_error("conv_fwd_tc: shape not eligible");
    return MK_EUNSUPPORTED;
  }
  if (dtype == MK_BF16)
    return conv_fwd_tc_typed<__nv_bfloat16>(pl, 1, x, w, transpose_w, kflip, nbr, tile_mask, ld, n_out, n_in, K,
                                            c_in, c_out, y, workspace, st);
  return conv_fwd_tc_typed<__half>(pl, 0, x, w, transpose_w, kflip, nbr, tile_mask, ld, n_out, n_in, K, c_in, c_out,
                                   y, workspace, st);
}

// weight gradient on tensor cores: conv_wgrad_tc.cu
}  // namespace mk

extern "C" int mk_has_tcgen05(void) { return 1; }

extern "C" int64_t mk_kmap_tile_mask_words(int K) { return (K + 31) / 32; }

extern "C" int mk_kmap_tile_mask(const int32_t *nbr, int64_t n_rows, int64_t ld, int K, uint32_t *mask,
                                 void *stream) {
  MK_CHECK_ARG(K >= 1 && ld >= n_rows, "bad K / ld");
  if (n_rows == 0) return MK_OK;
  MK_CHECK_ARG(nbr && mask, "null pointer");
  return mk::tile_mask_launch(nbr, n_rows, ld, K, mask, (cudaStream_t)stream);
}
