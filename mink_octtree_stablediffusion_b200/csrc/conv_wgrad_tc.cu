// Tensor-core (tcgen05 / TMEM) weight gradient of the sparse convolution.
//
// Replaces the dB part of matmul2<T,uint,TILE> behind ConvolutionBackwardGPU
// (pybind/extern.hpp:100-114), which accumulates with atomicAdd;
// here the result is deterministic.
//
//   dW[k][ci][co] = sum_o  x[nbr[k,o], ci] * dy[o, co]
//
// As an MMA: D[M x N] += A[M x 16] B[16 x N] with the 16 reduced "k" being
// output rows o.  A = gathered x rows, B = dy rows; both sit in shared memory
// exactly as they sit in global memory (row = voxel, 128 contiguous bytes = 64
// channels, 128B swizzle), which for UMMA is the MN-major canonical layout.
//
// Work decomposition
//   panel  = (offset k, 64-channel chunk of c_in): one gathered [128 x 64] tile
//   pair   = two consecutive panels = the M=128 rows of one accumulator
//   unit   = (pair, n-split): a [128 x NB] fp32 accumulator = NB TMEM columns
// A CTA owns up to U = 512/NB units for its whole life (accumulators never
// leave TMEM until the end) and walks the 128-row tiles xs, xs+X, ...; for each
// tile it loads the dy tile once (TMA) and, per pair, gathers the two panels
// (LDGSTS, zero-fill for missing neighbours).  Pairs whose offsets have no
// neighbour in the tile are skipped via the kernel map's tile mask.  At the end
// each CTA writes its fp32 partial to the workspace and a second kernel sums
// the X partials in a fixed order.
//
// CTA roles (640 threads, 1 CTA / SM):
//   warps 0-15  gather producers (four teams of four warps, one panel per team
//               at a time); warps 0-3 then run the epilogue (TMEM -> fp32 partial)
//   warp  16    neighbour-index prefetch (bulk copies into a smem ring)
//   warps 17-18 MMA issuers (unit u of the CTA -> warp u mod 2)
//   warp  19    dy tile loader (TMA)
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "sm100.cuh"
#include "tmap.cuh"

namespace mk {
using namespace sm100;

namespace wg {
// -DMK_TC_PROF: per-CTA cycle counters of where each role waits (development aid;
// read back with mk_debug_wg_prof_read).  Compiled out of the product build.
#ifdef MK_TC_PROF
__device__ unsigned long long g_wprof[256 * 32];
#define WP_DECL(...) unsigned long long __VA_ARGS__
#define WP_T0(v) v = clock64()
#define WP_ADD(acc, t0) acc += clock64() - (t0)
#define WP_ELAPSED(v) v = clock64() - (v)
#define WP_PUT(slot, v) g_wprof[(blockIdx.y * gridDim.x + blockIdx.x) * 32 + (slot)] = (v)
#else
#define WP_DECL(...)
#define WP_T0(v)
#define WP_ADD(acc, t0)
#define WP_ELAPSED(v)
#define WP_PUT(slot, v)
#endif
constexpr int kTileM = 128;
constexpr int kPanelBytes = kTileM * 128;   // [128 rows x 64 ch] 16-bit
constexpr int kStageBytes = 2 * kPanelBytes;
constexpr int kGatherTeams = 4;
constexpr int kGatherWarps = 4 * kGatherTeams;   // a team of 4 warps fills one panel (32 rows per warp)
#ifndef MK_WG_MMA_WARPS
#define MK_WG_MMA_WARPS 2
#endif
constexpr int kMmaWarps = MK_WG_MMA_WARPS;   // MMA warp m issues for the CTA's units u = m (mod kMmaWarps): accumulators nobody else touches
static_assert(kMmaWarps == 2 || kMmaWarps == 4, "unit masks below are written for 2 or 4 issuing warps");
constexpr int kWarpIdx = kGatherWarps, kWarpMma = kGatherWarps + 1, kWarpDy = kWarpMma + kMmaWarps;
constexpr int kThreads = 32 * (kGatherWarps + 2 + kMmaWarps);
constexpr int kIdxSlots = 32;
constexpr int kMaxUnits = 16;

struct Params {
  const void *x;          // [n_in, c_in]
  const int32_t *nbr;     // [K, ld]
  const uint32_t *mask;   // [ntiles, W]
  const int32_t *dy_rows; // optional [n_out]: table column o pairs with dy row dy_rows[o] (tile-ordered table)
  float *partial;         // [X][K][c_in][c_out]
  int64_t ld, n_out, n_in;
  int K, W, c_in, c_out;
  int pa;                 // 64-channel panels per offset
  int npairs, ppg;        // pairs in total / per CTA group
  int G;                  // pair groups (blockIdx.y = ns * G + g)
  int NB, nbpad, pb;      // accumulator columns, padded stride, dy panels per tile
  int a_stages, b_bufs;
  int ntiles, X;
  uint32_t idesc;
  int shared_idx;         // pa even: the two panels of a pair are channel chunks of ONE offset and share an index slot
  int knock;   // development aid (MK_WG_KNOCK): 1 = no gather copies, 2 = no MMA
};

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

// bit k of a tile's mask words (W <= 4)
__device__ __forceinline__ uint32_t mask_bit(const uint32_t (&w)[4], int k) {
  const uint32_t word = k < 32 ? w[0] : k < 64 ? w[1] : k < 96 ? w[2] : w[3];
  return (word >> (k & 31)) & 1u;
}

// Lane u of every role warp keeps the constants of the CTA's unit u (computed
// once, with the divisions): the offsets / channel panels of its two panels.
struct UnitConst {
  int k0, k1, pi0, pi1;   // k1 == K: the second panel does not exist
  bool valid;
};
// Unit u of pair group g is pair g + u * G: the groups interleave over the offsets.  (With
// contiguous halves the group of the low offsets had 25 % more stages on a tile-ordered
// table - the order sorts by neighbour pattern, whose high bits are the high offsets, so
// those are all-or-nothing inside a tile while the low ones are mixed - and the kernel
// ran at the pace of that group.)
__device__ __forceinline__ UnitConst unit_const(const Params &p, int g, int npair_cta, int lane) {
  UnitConst c;
  const int q0 = 2 * (g + lane * p.G);
  c.k0 = q0 / p.pa; c.pi0 = q0 % p.pa;
  c.k1 = (q0 + 1) / p.pa; c.pi1 = (q0 + 1) % p.pa;
  c.valid = lane < npair_cta;
  if (c.k1 >= p.K) c.k1 = p.K;
  return c;
}

// per-tile view of this CTA's units: bit u of `pairs` = unit u has work; bit u of
// `p0` / `p1` = its first / second panel is real and non-empty.  Whole (converged)
// warp; a few instructions (the role warps run it once per tile on their critical path).
struct TileSched { uint32_t pairs, p0, p1; };

__device__ __forceinline__ void load_mask(const Params &p, int tile, uint32_t (&w)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) w[i] = (i < p.W && (unsigned)tile < (unsigned)p.ntiles) ? __ldg(p.mask + (int64_t)tile * p.W + i) : 0u;
}

__device__ __forceinline__ TileSched tile_sched(const Params &p, const uint32_t (&w)[4], const UnitConst &c) {
  const uint32_t b0 = c.valid ? mask_bit(w, c.k0) : 0u;
  const uint32_t b1 = (c.valid && c.k1 < p.K) ? mask_bit(w, c.k1) : 0u;
  TileSched s;
  s.p0 = __ballot_sync(0xffffffffu, b0 != 0u);
  s.p1 = __ballot_sync(0xffffffffu, b1 != 0u);
  s.pairs = s.p0 | s.p1;
  return s;
}
}  // namespace wg

__global__ void __launch_bounds__(wg::kThreads, 1)
k_conv_wgrad_tc(const __grid_constant__ wg::Params p, const __grid_constant__ CUtensorMap tmap_dy,
                const __grid_constant__ CUtensorMap tmap_dy_row) {
  using namespace wg;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t S = (uint32_t)p.a_stages, NBUF = (uint32_t)p.b_bufs;
  const uint32_t b_bytes = (uint32_t)p.pb * kPanelBytes;
  uint8_t *a_ring = smem;
  uint8_t *b_ring = a_ring + (size_t)S * kStageBytes;
  int32_t *idx_ring = (int32_t *)(b_ring + (size_t)NBUF * b_bytes);
  uint64_t *bars = (uint64_t *)(idx_ring + kIdxSlots * kTileM);
  uint64_t *a_full = bars, *a_empty = a_full + S;
  // b_empty / done: one barrier per MMA warp.  Two tcgen05.commit from different warps
  // onto ONE mbarrier were seen to lose an arrival when they come close together
  // (the kernel hung with the MMAs knocked out, i.e. back-to-back commits), so
  // every barrier here has a single committing thread.
  uint64_t *b_full = a_empty + S, *b_empty = b_full + NBUF;             // b_empty[buf * kMmaWarps + m]
  uint64_t *i_full = b_empty + NBUF * kMmaWarps, *i_empty = i_full + kIdxSlots;
  uint64_t *done = i_empty + kIdxSlots;                                 // done[m]
  uint32_t *tmem_slot = (uint32_t *)(done + kMmaWarps);
  uint32_t *inited_slot = tmem_slot + 1;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = blockIdx.y % p.G, ns = blockIdx.y / p.G;
  const int npair_cta = (p.npairs - g + p.G - 1) / p.G;
  const int xs = blockIdx.x;

  if (threadIdx.x == 0) {
    for (uint32_t s = 0; s < S; ++s) { mbar_init(&a_full[s], 256); mbar_init(&a_empty[s], kMmaWarps); }
    for (uint32_t s = 0; s < NBUF; ++s) mbar_init(&b_full[s], 1);
    for (uint32_t s = 0; s < NBUF * kMmaWarps; ++s) mbar_init(&b_empty[s], 1);
    for (int s = 0; s < kIdxSlots; ++s) { mbar_init(&i_full[s], 1); mbar_init(&i_empty[s], p.shared_idx ? 8 : 4); }
    for (int m = 0; m < kMmaWarps; ++m) mbar_init(&done[m], 2);
    *inited_slot = 0u;
    fence_barrier_init();
  }
  if (warp == kWarpDy && lane == 0) tma_prefetch_desc(p.dy_rows ? &tmap_dy_row : &tmap_dy);
  if (warp == kWarpMma) tmem_alloc_512(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < kGatherWarps) {
    // ===================== gather producers =====================
    // one work item = one panel; items are dealt round-robin to the four teams
    // (a team's consecutive panels are 4 apart = 2 stages < S: alias-free parity
    // waits with in-order release).  Warp q of a team moves panel rows
    // [32q, 32q+32); lane = (rsub, c16): instruction i moves 16-byte chunk c16 of
    // rows 32q + 8*rsub + i (four 128-byte feature rows per warp instruction)
    const uint32_t row_bytes = (uint32_t)p.c_in * 2u;
    const uint32_t team = (uint32_t)warp >> 2;
    const int q = warp & 3;
    const int c16 = lane & 7, rsub = lane >> 3;
    const int row0 = 32 * q + 8 * rsub;
    // The loop is bound by instruction issue: lane-constant address parts are computed
    // once and shared memory is addressed in its own space (see conv_tc.cu).  Destination
    // of copy i: bits 4-6 = c16 ^ i (swizzle), bits 7-9 = i (row in the 8-row atom) = an
    // XOR with i * 144, the atom base having those bits clear.
    const uint32_t dst_lane = smem_u32(a_ring) + (uint32_t)row0 * 128u + ((uint32_t)c16 << 4);
    const uint32_t idx_lane = smem_u32(idx_ring) + (uint32_t)row0 * 4u;
    const uint8_t *x_lane = (const uint8_t *)p.x + (size_t)c16 * 16;
    const uint32_t b_a_full = smem_u32(a_full), b_a_empty = smem_u32(a_empty);
    const uint32_t b_i_full = smem_u32(i_full), b_i_empty = smem_u32(i_empty);
    Ring ir, ar;
    uint32_t nstage = 0;                           // stages so far
    WP_DECL(g_wi = 0, g_wa = 0, g_is = 0, g_t0 = 0, g_start = 0, g_blk = 0, g_b0 = 0, g_sch = 0, g_s0 = 0;)
    WP_T0(g_start);
    const UnitConst uc = unit_const(p, g, npair_cta, lane);
    uint32_t wn[4];
    load_mask(p, xs, wn);
    for (int tile = xs; tile < p.ntiles; tile += p.X) {
      WP_T0(g_s0);
      const uint32_t w[4] = {wn[0], wn[1], wn[2], wn[3]};
      load_mask(p, tile + p.X, wn);   // next tile's mask: its latency hides behind this tile
      const TileSched ts = tile_sched(p, w, uc);
      WP_ADD(g_sch, g_s0);
      // Panel e = 2 * (stage number) + hf goes to team e & 3: a team owns half hf =
      // team & 1 of every second stage, so it walks the set bits of `pairs` and acts
      // on the stages of its parity only; the ring positions of a stage / of a real
      // panel follow from bit counts, nothing is advanced panel by panel.
      const uint32_t hf = team & 1u;
      const uint32_t real_mask = hf ? ts.p1 : ts.p0;
      uint32_t rank = 0;
      for (uint32_t b = ts.pairs; b; b &= b - 1u, ++rank) {
        if (((nstage + rank) & 1u) != (team >> 1)) continue;
        const int u = __ffs((int)b) - 1;
        const uint32_t below = (1u << u) - 1u;
        const bool real = (real_mask >> u) & 1u;
        WP_T0(g_b0);
        int pi = 0;
        if (p.pa > 1) pi = __shfl_sync(0xffffffffu, hf ? uc.pi1 : uc.pi0, u);
        const uint32_t szok = (real && (pi * 64 + c16 * 8) < p.c_in) ? 16u : 0u;
        const uint8_t *xb = x_lane + (size_t)pi * 128;
        asm volatile("" : "+l"(xb));   // one register pair: row address = a single IMAD.WIDE
        Ring as = ar;
        as.skip(rank, S);
        int4 i0 = make_int4(-1, -1, -1, -1), i1 = i0;
        uint32_t i_empty_s = 0;
        if (real) {
          // real panels before this one: both halves of the pairs below u, and the first half of u
          Ring is = ir;
          is.skip(p.shared_idx ? (uint32_t)__popc(ts.pairs & below)
                               : (uint32_t)__popc(ts.p0 & below) + (uint32_t)__popc(ts.p1 & below) +
                                     (hf ? ((ts.p0 >> u) & 1u) : 0u),
                  kIdxSlots);
          WP_T0(g_t0);
          mbar_wait_warp_a(b_i_full + is.stage * 8u, is.phase);
          WP_ADD(g_wi, g_t0);
          i0 = lds_v4(idx_lane + is.stage * (kTileM * 4u));
          i1 = lds_v4(idx_lane + is.stage * (kTileM * 4u) + 16u);
          i_empty_s = b_i_empty + is.stage * 8u;
        }
        WP_T0(g_t0);
        mbar_wait_warp_a(b_a_empty + as.stage * 8u, as.phase ^ 1u);
        WP_ADD(g_wa, g_t0);
        WP_T0(g_t0);
        const uint32_t sb = dst_lane + as.stage * kStageBytes + hf * kPanelBytes;
        const int32_t ri[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
        if (!(p.knock & 1))
#pragma unroll
        for (int i = 0; i < 8; ++i)
          cp_async16(sb ^ (uint32_t)(i * 144), xb + (size_t)(uint32_t)max(ri[i], 0) * row_bytes, ri[i] >= 0 ? szok : 0u);
        cp_async_arrive_noinc_a(b_a_full + as.stage * 8u);
        if (real) {
          // release the index slot only after the copies consumed every loaded index (an
          // asm LDS followed directly by the asm arrive carries no dependence: the arrive
          // can overtake the load and the slot is overwritten under it)
          __syncwarp();
          if (lane == 0) mbar_arrive_a(i_empty_s);
        }
        WP_ADD(g_is, g_t0);
        WP_ADD(g_blk, g_b0);
      }
      const uint32_t n_st = (uint32_t)__popc(ts.pairs);
      nstage += n_st;
      ar.skip(n_st, S);
      ir.skip(p.shared_idx ? n_st : (uint32_t)__popc(ts.p0) + (uint32_t)__popc(ts.p1), kIdxSlots);
    }
    if (lane == 0 && (warp & 3) == 0) {
      WP_PUT(8 + 4 * team, g_wi); WP_PUT(9 + 4 * team, g_wa); WP_PUT(10 + 4 * team, g_is);
      if (team == 0) { WP_PUT(4, g_blk); WP_PUT(5, g_sch); }
      WP_ELAPSED(g_start); WP_PUT(11 + 4 * team, g_start);
    }
    if (warp >= 4) goto done_role;   // only warps 0-3 (TMEM lane quarters) run the epilogue
    // ===================== epilogue: TMEM -> fp32 partial =====================
    for (int m = 0; m < kMmaWarps; ++m) mbar_wait_warp(&done[m], 0u);
    tc_fence_after();
    const uint32_t inited = *(volatile uint32_t *)inited_slot;
    float *part = p.partial + (size_t)xs * p.K * p.c_in * p.c_out;
    for (int u = 0; u < npair_cta; ++u) {
      const int r = warp * 32 + lane;                 // accumulator row = TMEM lane
      const int q = 2 * (g + u * p.G) + (r >> 6);
      const int k = q / p.pa, ci = (q % p.pa) * 64 + (r & 63);
      const bool row_ok = k < p.K && ci < p.c_in;
      float *dst = part + ((size_t)k * p.c_in + ci) * p.c_out + (size_t)ns * p.NB;
      const uint32_t tbase = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(u * p.nbpad);
      for (int c0 = 0; c0 < p.NB; c0 += 16) {
        uint32_t v[16];
        if ((inited >> u) & 1u) {
          tmem_ld16(tbase + (uint32_t)c0, v);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = 0u;
        }
        if (row_ok) {
#pragma unroll
          for (int j = 0; j < 16; j += 4)
            *reinterpret_cast<uint4 *>(dst + c0 + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
      }
    }
    tc_fence_before();
  } else if (warp == kWarpIdx) {
    // ===================== neighbour-index prefetch =====================
    // Lane i moves the tile's i-th real panel: its own slot, its own wait, its own
    // 512-byte bulk copy, so the per-panel barrier latency is paid once per tile
    // instead of once per panel (a tile has up to 2 * kMaxUnits = kIdxSlots panels).
    const uint32_t i_base = smem_u32(idx_ring);
    Ring ir;
    const UnitConst uc = unit_const(p, g, npair_cta, lane);
    uint32_t wn[4];
    load_mask(p, xs, wn);
    WP_DECL(i_w = 0, i_t0 = 0, i_start = clock64(), i_n = 0, i_a = 0, i_b = 0, i_ta = 0, i_c = 0;)
    for (int tile = xs; tile < p.ntiles; tile += p.X) {
      WP_T0(i_ta);
      const uint32_t w[4] = {wn[0], wn[1], wn[2], wn[3]};
      load_mask(p, tile + p.X, wn);
      WP_DECL(i_tb = clock64();)
      const TileSched ts = tile_sched(p, w, uc);
      WP_ADD(i_c, i_tb);
      // Lane q owns panel q = 2u + hf (real iff bit q of pm) and the slot at its rank among the real
      // panels.  (An earlier form gave lane i the i-th real panel through __fns and built pm in a loop
      // over the units: ~3300 cycles per tile on this single warp, which a CTA of a wide layer - few
      // units, so a tile every ~1500 cycles - could not hide: the gather teams waited for indices 41 %
      // of the time at 256 -> 256.)
      // With an even number of channel panels per offset (shared_idx) both panels of a pair read the
      // same table row: one slot and one copy per PAIR, lane u = unit u with its own constants.
      const uint32_t hf = p.shared_idx ? 0u : (uint32_t)lane & 1u;
      const uint32_t pm = p.shared_idx
                              ? ts.pairs
                              : __ballot_sync(0xffffffffu, (((hf ? ts.p1 : ts.p0) >> (lane >> 1)) & 1u) != 0u);
      const uint32_t n = (uint32_t)__popc(pm);
      const bool act = (pm >> lane) & 1u;
      const int src_lane = p.shared_idx ? lane : lane >> 1;
      const int k0 = __shfl_sync(0xffffffffu, uc.k0, src_lane), k1 = __shfl_sync(0xffffffffu, uc.k1, src_lane);
      WP_ADD(i_a, i_ta);
      WP_T0(i_ta);
      if (act) {
        const int k = hf ? k1 : k0;
        uint32_t slot = ir.stage + (uint32_t)__popc(pm & ((1u << lane) - 1u)), ph = ir.phase;
        if (slot >= (uint32_t)kIdxSlots) { slot -= kIdxSlots; ph ^= 1u; }
        WP_T0(i_t0);
        mbar_wait(&i_empty[slot], ph ^ 1u);
        WP_ADD(i_w, i_t0);
        const int32_t *src = p.nbr + (int64_t)k * p.ld + (int64_t)tile * kTileM;
        mbar_arrive_expect_tx(&i_full[slot], kTileM * 4);
        bulk_g2s(i_base + slot * (kTileM * 4), src, kTileM * 4, &i_full[slot]);
      }
      __syncwarp();
      WP_ADD(i_b, i_ta);
      ir.skip(n, kIdxSlots);
    }
    if (lane == 0) { WP_PUT(25, i_w); WP_ELAPSED(i_start); WP_PUT(26, i_start); WP_PUT(27, i_a); WP_PUT(28, i_b); WP_PUT(29, i_c); }
  } else if (warp >= kWarpMma && warp < kWarpMma + kMmaWarps) {
    // ===================== MMA issuers =====================
    // The serial cost of a stage on the issuing warp (barrier wait, proxy fence,
    // eight blocking UMMA issues, commit: ~600 cycles, measured) exceeds the ~390
    // cycles the stage keeps the tensor pipe busy, so one warp cannot keep the pipe
    // fed.  The CTA's units are independent accumulators: two warps share them by
    // unit parity, each walks the set bits of the tile's schedule and touches only
    // the A stages of its own units.  On the dy buffers, shared by both, every warp
    // waits and arrives once per tile.
    const uint32_t mw = (uint32_t)(warp - kWarpMma);
    const uint32_t a_base = smem_u32(a_ring), b_base = smem_u32(b_ring);
    // MN-major SW128 operands: LBO = distance between 64-channel panels,
    // SBO = 1024 B between 8-row groups; 16 rows (one MMA) = 2048 B
    const uint64_t desc0 = smem_desc_sw128(0, kPanelBytes, 1024);
    Ring ar, br;
    uint32_t inited = 0;
    WP_DECL(m_wb = 0, m_wa = 0, m_is = 0, m_t0 = 0, m_start = 0;)
    WP_T0(m_start);
    const UnitConst uc = unit_const(p, g, npair_cta, lane);
    const uint32_t mine_mask = (kMmaWarps == 2 ? 0x55555555u : 0x11111111u) << mw;
    uint32_t wn[4];
    load_mask(p, xs, wn);
    for (int tile = xs; tile < p.ntiles; tile += p.X) {
      const uint32_t w[4] = {wn[0], wn[1], wn[2], wn[3]};
      load_mask(p, tile + p.X, wn);
      const TileSched ts = tile_sched(p, w, uc);
      if (!ts.pairs) continue;
      WP_T0(m_t0);
      mbar_wait_shared(&b_full[br.stage], br.phase);   // kMmaWarps waiters
      __syncwarp();
      WP_ADD(m_wb, m_t0);
      const uint64_t bd0 = desc0 | (uint64_t)(((b_base + br.stage * b_bytes) & 0x3FFFFu) >> 4);
      Ring as = ar;
      for (uint32_t b = ts.pairs; b; b &= b - 1u, as.advance(S)) {
        const int u = __ffs((int)b) - 1;
        // Every warp waits on every stage, also on the other warp's, and the slot is
        // released by both (the owner's commit + the other's plain arrive): a parity
        // wait is only sound for a waiter that sees each phase of its barrier - one
        // that skips uses, or is not part of the release, can be two phases out.
        WP_T0(m_t0);
        mbar_wait_shared(&a_full[as.stage], as.phase);
        __syncwarp();
        WP_ADD(m_wa, m_t0);
        if (!((mine_mask >> u) & 1u)) {
          if (lane == 0) mbar_arrive(&a_empty[as.stage]);
          __syncwarp();
          continue;
        }
        WP_T0(m_t0);
        fence_proxy_async_smem();
        tc_fence_after();
        if (elect_one()) {
          const uint32_t d = tmem_base + (uint32_t)(u * p.nbpad);
          const uint64_t ad0 = desc0 | (uint64_t)(((a_base + as.stage * kStageBytes) & 0x3FFFFu) >> 4);
          uint32_t acc = (inited >> u) & 1u;
          if (!(p.knock & 2))
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            umma_f16(d, ad0 + (uint64_t)(ks * 128), bd0 + (uint64_t)(ks * 128), p.idesc, acc);
            acc = 1u;
          }
          umma_commit(&a_empty[as.stage]);
        }
        __syncwarp();
        WP_ADD(m_is, m_t0);
        inited |= 1u << u;
      }
      ar.skip((uint32_t)__popc(ts.pairs), S);
      if (elect_one()) umma_commit(&b_empty[br.stage * kMmaWarps + mw]);
      __syncwarp();
      br.advance(NBUF);
    }
    if (lane == 0 && mw == 0) {
      WP_PUT(0, m_wb); WP_PUT(1, m_wa); WP_PUT(2, m_is);
      WP_ELAPSED(m_start); WP_PUT(3, m_start);
    }
    if (lane == 0) {
      atomicOr(inited_slot, inited);
      mbar_arrive(&done[mw]);                  // release: orders the flag before the phase completes
    }
    __syncwarp();
    if (elect_one()) umma_commit(&done[mw]);   // ... and the warp's MMAs
    __syncwarp();
  } else {
    // ===================== dy tile loader (TMA) =====================
    const uint32_t b_base = smem_u32(b_ring);
    Ring br;
    const UnitConst uc = unit_const(p, g, npair_cta, lane);
    uint32_t wn[4];
    load_mask(p, xs, wn);
    WP_DECL(d_we = 0, d_ld = 0, d_t0 = 0, d_start = clock64(), d_n = 0;)
    for (int tile = xs; tile < p.ntiles; tile += p.X) {
      const uint32_t w[4] = {wn[0], wn[1], wn[2], wn[3]};
      load_mask(p, tile + p.X, wn);
      const TileSched ts = tile_sched(p, w, uc);
      if (!ts.pairs) continue;
      WP_T0(d_t0);
      for (int m = 0; m < kMmaWarps; ++m) mbar_wait_warp(&b_empty[br.stage * kMmaWarps + m], br.phase ^ 1u);
      WP_ADD(d_we, d_t0);
      WP_T0(d_t0);
      if (p.dy_rows) {
        // rows through the permutation: lane l brings rows 4l..4l+3 of the tile with one
        // tile::gather4 per 64-channel panel (same swizzled image as the tiled load);
        // columns past n_out get an out-of-range row index = zero fill
        const int64_t o = (int64_t)tile * kTileM + 4 * lane;
        int4 r;
        if (o + 3 < p.n_out) {
          r = __ldg(reinterpret_cast<const int4 *>(p.dy_rows + o));
        } else {
          const int32_t oob = (int32_t)p.n_out;
          r.x = o + 0 < p.n_out ? __ldg(p.dy_rows + o + 0) : oob;
          r.y = o + 1 < p.n_out ? __ldg(p.dy_rows + o + 1) : oob;
          r.z = o + 2 < p.n_out ? __ldg(p.dy_rows + o + 2) : oob;
          r.w = oob;
        }
        if (lane == 0) mbar_arrive_expect_tx(&b_full[br.stage], b_bytes);
        __syncwarp();
        for (int pn = 0; pn < p.pb; ++pn)
          tma_gather4(b_base + br.stage * b_bytes + (uint32_t)pn * kPanelBytes + (uint32_t)lane * 512u, &tmap_dy_row,
                      ns * p.NB + pn * 64, r, &b_full[br.stage]);
      } else if (lane == 0) {
        mbar_arrive_expect_tx(&b_full[br.stage], b_bytes);
        for (int pn = 0; pn < p.pb; ++pn)
          tma_load_2d(b_base + br.stage * b_bytes + (uint32_t)pn * kPanelBytes, &tmap_dy, ns * p.NB + pn * 64,
                      tile * kTileM, &b_full[br.stage]);
      }
      __syncwarp();
      WP_ADD(d_ld, d_t0);
      br.advance(NBUF);
    }
    if (lane == 0) { WP_PUT(6, d_we); WP_PUT(7, d_ld); WP_ELAPSED(d_start); WP_PUT(24, d_start); }
  }

done_role:
  tc_fence_before();
  __syncthreads();
  if (warp == kWarpMma) {
    tc_fence_after();
    tmem_dealloc_512(tmem_base);
  }
}

// dw[e] = sum over the X partials in a fixed order (deterministic): warp w of a block sums
// the partials s = w, w + 8, ... of 32 consecutive float4 columns (coalesced 512-byte rows,
// eight independent chains in flight per column instead of one of X dependent loads), then
// warp 0 adds the eight sums in order.
template <typename T>
__global__ void __launch_bounds__(256)
k_wgrad_reduce(const float *__restrict__ partial, int X, int64_t n, T *__restrict__ dw) {
  __shared__ float4 part[8][32];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t i = ((int64_t)blockIdx.x * 32 + lane) * 4;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i < n) {
#pragma unroll 4
    for (int s = w; s < X; s += 8) {
      const float4 v = __ldg(reinterpret_cast<const float4 *>(partial + (size_t)s * n + i));
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  part[w][lane] = acc;
  __syncthreads();
  if (w != 0 || i >= n) return;
#pragma unroll
  for (int j = 1; j < 8; ++j) {
    const float4 v = part[j][lane];
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
  }
  dw[i] = from_acc<T, float>(acc.x);
  dw[i + 1] = from_acc<T, float>(acc.y);
  dw[i + 2] = from_acc<T, float>(acc.z);
  dw[i + 3] = from_acc<T, float>(acc.w);
}

// out[i, :] = dy[rows[i], :] (16-bit rows of `row_bytes`, a multiple of 16), 16 bytes per thread.
// For wide outputs the weight-gradient kernel reads dy tiles with one tiled TMA load per panel from
// this copy instead of 32 tile::gather4 per panel through the permutation: a CTA of a layer wider than 64 columns needs
// a new dy tile every ~1300 cycles (few units per CTA, so few stages per tile), and the single loader
// warp's gather4 stream took ~4500 (measured: 63 % of the MMA warps' time waiting for the dy tile at
// 256 -> 256; 4.68 -> 3.2 ms + 0.2 ms for this pass).
__global__ void __launch_bounds__(256)
k_gather_rows16(const uint8_t *__restrict__ dy, const int32_t *__restrict__ rows, int64_t n, int row_bytes,
                uint8_t *__restrict__ out) {
  const int per_row = row_bytes >> 4;
  const int64_t total = n * per_row;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = t / per_row;
    const int c = (int)(t - r * per_row);
    const int64_t src = (int64_t)__ldg(rows + r);
    reinterpret_cast<uint4 *>(out + r * row_bytes)[c] =
        __ldg(reinterpret_cast<const uint4 *>(dy + src * row_bytes) + c);
  }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
static int arm_hang_report() {
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

// Upper bound on the rows one CTA may reduce over (0 = none): set by the fp32 split path around its calls
// (conv_f32.cu) - its bf16-piece products must not chain more than ~400 MMAs (16 rows each) into one fp32
// TMEM accumulator (tcgen05 accumulates with truncation).  The plan then takes as many tile
// splits as that needs, whatever the SM count (several waves of CTAs), and k_wgrad_reduce adds the
// partials in fp32 (round to nearest).
static thread_local int64_t g_rows_max = 0;
void conv_wgrad_tc_set_rows_max(int64_t rows_max) { g_rows_max = rows_max; }

namespace {
struct WPlan {
  int pa, npairs, nsplit, NB, nbpad, pb, U, G, ppg, a_stages, b_bufs, ntiles, X, W;
  size_t smem, partial_bytes, mask_bytes, dyg_bytes;
};

// dy through a tile order: pre-gathered copy (wide outputs) or tile::gather4 inside the kernel (narrow)
inline bool pregather_dy(int c_out) {
  if (const char *e = getenv("MK_WG_PREGATHER")) return atoi(e) != 0;   // development knob
  return c_out > 64;   // measured: 96 -> 96 1.03 -> 0.72 ms with the copy, 64 -> 64 0.306 -> 0.359 (one panel: gather4 keeps up)
}

inline size_t align256w(size_t b) { return (b + 255) & ~(size_t)255; }

bool make_wplan(int dtype, int64_t n_out, int K, int c_in, int c_out, WPlan &pl) {
  using namespace wg;
  if (dtype != MK_BF16 && dtype != MK_F16) return false;
  if (c_in < 16 || c_in % 16 || c_out < 16 || c_out % 16 || K < 1 || K > 128 || n_out < 1) return false;
  if (n_out > (int64_t)kTileM * 0x7fffff00) return false;
  pl.pa = (c_in + 63) / 64;
  pl.npairs = (K * pl.pa + 1) / 2;
  // 128-column accumulators: dy tile of 32 KB, four A stages in flight.  From 512 output channels on, 256
  // columns (64 KB dy tile, two A stages): every gathered panel is then used for twice the columns
  // (512 -> 512 on a dense 48^3 cube 817 -> 1015 TFLOP/s, on the sphere surface 400 -> 415; at 256 -> 256
  // the sphere loses 5 %).
  int nb_max = c_out >= 512 ? 256 : 128;
  if (const char *e = getenv("MK_WG_NB")) nb_max = std::max(16, std::min(256, atoi(e)));   // development knob
  int nsplit = (c_out + nb_max - 1) / nb_max;
  while (c_out % (16 * nsplit)) ++nsplit;
  pl.nsplit = nsplit;
  pl.NB = c_out / nsplit;
  pl.nbpad = (pl.NB + 31) & ~31;
  pl.pb = (pl.NB + 63) / 64;
  pl.U = std::min(kMaxUnits, 512 / pl.nbpad);
  pl.G = (pl.npairs + pl.U - 1) / pl.U;
  pl.ppg = (pl.npairs + pl.G - 1) / pl.G;
  pl.G = (pl.npairs + pl.ppg - 1) / pl.ppg;
  pl.W = (K + 31) / 32;
  pl.ntiles = (int)((n_out + kTileM - 1) / kTileM);
  const int ctas_y = pl.G * nsplit;
  pl.X = std::max(1, std::min(pl.ntiles, num_sms() / ctas_y));
  if (g_rows_max > 0) {
    const int tiles_max = (int)std::max<int64_t>(1, g_rows_max / kTileM);
    pl.X = std::max(pl.X, (pl.ntiles + tiles_max - 1) / tiles_max);
    if ((size_t)pl.X * K * c_in * c_out * 4 > ((size_t)8 << 30)) return false;   // partial sums beyond 8 GB: not served
    if ((int64_t)pl.X * ctas_y > 0x7fff0000ll || ctas_y > 65535) return false;
  }
  const size_t b_tile = (size_t)pl.pb * kPanelBytes;
  const size_t fixed = 1024 + (size_t)kIdxSlots * kTileM * 4 + 1024;
  const size_t budget = 227 * 1024;
  pl.b_bufs = 2;
  if (fixed + 2 * b_tile + 2 * (size_t)kStageBytes > budget) pl.b_bufs = 1;
  if (fixed + pl.b_bufs * b_tile + 2 * (size_t)kStageBytes > budget) return false;
  pl.a_stages = (int)std::min<size_t>(6, (budget - fixed - pl.b_bufs * b_tile) / kStageBytes);
  // A gather team owns the stages of one parity (of the running stage number).  With an
  // even ring the same teams refill a given slot every time, so a team sees every
  // phase of the a_empty barriers it waits on whatever the order the two MMA warps
  // release the slots in (with an odd ring it would see every other phase: unsound).
  pl.a_stages &= ~1;
  if (pl.a_stages < 2) return false;
  if (fixed + 3 * b_tile + (size_t)pl.a_stages * kStageBytes <= budget && pl.b_bufs == 2) pl.b_bufs = 3;
  pl.smem = fixed + pl.b_bufs * b_tile + (size_t)pl.a_stages * kStageBytes;
  pl.partial_bytes = align256w((size_t)pl.X * K * c_in * c_out * 4);
  pl.mask_bytes = align256w((size_t)pl.ntiles * pl.W * 4);
  pl.dyg_bytes = pregather_dy(c_out) ? align256w((size_t)n_out * c_out * 2) : 0;
  return true;
}
}  // namespace

int tile_mask_launch(const int32_t *nbr, int64_t n_rows, int64_t ld, int K, uint32_t *mask, cudaStream_t st);

bool conv_wgrad_tc_eligible(int dtype, int64_t n_out, int K, int c_in, int c_out) {
  WPlan pl;
  return make_wplan(dtype, n_out, K, c_in, c_out, pl);
}

// rows of the table one CTA reduces over (its share of the tile splits)
int64_t conv_wgrad_tc_rows_per_cta(int64_t n_out, int K, int c_in, int c_out) {
  WPlan pl;
  if (!make_wplan(MK_BF16, std::max<int64_t>(n_out, 1), K, c_in, c_out, pl)) return n_out;
  return ((int64_t)(pl.ntiles + pl.X - 1) / pl.X) * wg::kTileM;
}

int64_t conv_wgrad_tc_workspace(int64_t n_out, int K, int c_in, int c_out) {
  WPlan pl;
  if (!make_wplan(MK_BF16, std::max<int64_t>(n_out, 1), K, c_in, c_out, pl)) return 0;
  return (int64_t)(pl.partial_bytes + pl.mask_bytes + pl.dyg_bytes + 256);
}

int conv_wgrad_tc(int dtype, const void *x, const void *dy, const int32_t *dy_rows, const int32_t *nbr,
                  const uint32_t *tile_mask, int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in, int c_out,
                  void *dw, void *workspace, cudaStream_t st, int dw_f32) {
  using namespace wg;
  WPlan pl;
  if (!make_wplan(dtype, n_out, K, c_in, c_out, pl)) {
    set_error("conv_wgrad_tc: shape not eligible");
    return MK_EUNSUPPORTED;
  }
  uint8_t *ws = (uint8_t *)workspace;
  float *partial = (float *)ws;
  // one tile split and an fp32 result: the only partial IS the result (small maps: the reduction pass of a
  // 27 x 960 x 960 kernel is 200 MB of traffic for nothing)
  const bool direct = pl.X == 1 && dw_f32;
  if (direct) partial = (float *)dw;
  if (tile_mask == nullptr) {
    uint32_t *mask = (uint32_t *)(ws + pl.partial_bytes);
    int rc = tile_mask_launch(nbr, n_out, ld, K, mask, st);
    if (rc) return rc;
    tile_mask = mask;
  }
  if (dy_rows && pl.dyg_bytes) {
    uint8_t *dyg = ws + pl.partial_bytes + pl.mask_bytes;
    const int64_t total = n_out * (int64_t)(c_out * 2 / 16);
    const unsigned blocks = (unsigned)std::min<int64_t>(ceil_div(total, 256), (int64_t)num_sms() * 16);
    k_gather_rows16<<<blocks, 256, 0, st>>>((const uint8_t *)dy, dy_rows, n_out, c_out * 2, dyg);
    MK_LAUNCH_CHECK();
    dy = dyg;
    dy_rows = nullptr;
  }
  CUtensorMap tmap_dy, tmap_dy_row;
  {
    int rc = make_tmap_rows16(&tmap_dy, dtype, dy, n_out, c_out, 64, kTileM);
    if (rc) return rc;
    rc = make_tmap_rows16(&tmap_dy_row, dtype, dy, n_out, c_out, 64, 1);   // tile::gather4 form
    if (rc) return rc;
  }
  Params p;
  p.x = x; p.nbr = nbr; p.mask = tile_mask; p.dy_rows = dy_rows; p.partial = partial;
  p.ld = ld; p.n_out = n_out; p.n_in = n_in;
  p.K = K; p.W = pl.W; p.c_in = c_in; p.c_out = c_out;
  p.pa = pl.pa; p.npairs = pl.npairs; p.ppg = pl.ppg; p.G = pl.G;
  p.NB = pl.NB; p.nbpad = pl.nbpad; p.pb = pl.pb;
  p.a_stages = pl.a_stages; p.b_bufs = pl.b_bufs;
  p.ntiles = pl.ntiles; p.X = pl.X;
  p.idesc = idesc_f16(dtype == MK_BF16 ? 1 : 0, kTileM, pl.NB, 1, 1);
  p.shared_idx = (pl.pa % 2 == 0 && !getenv("MK_WG_NOSHARE")) ? 1 : 0;
  p.knock = getenv("MK_WG_KNOCK") ? atoi(getenv("MK_WG_KNOCK")) : 0;
  { int rc = arm_hang_report(); if (rc) return rc; }
  MK_CUDA(cudaFuncSetAttribute(k_conv_wgrad_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  dim3 grid(pl.X, pl.G * pl.nsplit);
  k_conv_wgrad_tc<<<grid, kThreads, pl.smem, st>>>(p, tmap_dy, tmap_dy_row);
  MK_LAUNCH_CHECK();
  if (direct) return MK_OK;
  const int64_t n_w = (int64_t)K * c_in * c_out;
  const unsigned blocks = (unsigned)ceil_div(n_w / 4, 32);
  if (dw_f32)   // the fp32 sum as it is (fp32 parameters; the fp32 split path folds its blocks from it)
    k_wgrad_reduce<float><<<blocks, 256, 0, st>>>(partial, pl.X, n_w, (float *)dw);
  else if (dtype == MK_BF16)
    k_wgrad_reduce<__nv_bfloat16><<<blocks, 256, 0, st>>>(partial, pl.X, n_w, (__nv_bfloat16 *)dw);
  else
    k_wgrad_reduce<__half><<<blocks, 256, 0, st>>>(partial, pl.X, n_w, (__half *)dw);
  MK_LAUNCH_CHECK();
  return MK_OK;
}

}  // namespace mk

#ifdef MK_TC_PROF
extern "C" int mk_debug_wg_prof_read(unsigned long long *host, int n) {
  return (int)cudaMemcpyFromSymbol(host, mk::wg::g_wprof, sizeof(unsigned long long) * n);
}
#endif
