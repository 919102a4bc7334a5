// Micro-benchmark (development aid): how fast can one SM pull 128-byte feature rows
// into swizzled shared-memory tiles?  LDGSTS (cp.async 16 B) vs TMA tile::gather4.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_bw gather_bw.cu -I../../mink_octtree_stablediffusion_b200/csrc -I../../include
//   ./gather_bw
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include "common.cuh"
#include "sm100.cuh"
#include "tmap.cuh"
namespace mk { void set_error(const char *, ...) {} }
using namespace mk::sm100;

constexpr int kTile = 128, kTileBytes = 128 * 128;

// mode 0: LDGSTS, a warp moves 32 rows per round (4 warps = one tile); mode 1: TMA gather4, a warp moves a whole tile
template <int MODE0>
__global__ void __launch_bounds__(1024, 1)
k_gather(const __grid_constant__ CUtensorMap tm, const uint8_t *x, const int32_t *idx, int tiles_per_cta, int nwarps,
         int depth, unsigned long long *cycles, int zero_row, int n_ldgs_warps) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bars[64];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // MODE 1: slot = warp * depth + j (one tile each), barrier count 1 (+tx)
  // MODE 0: team = warp / 4; slot = team * depth + j; barrier count 128
  if (threadIdx.x == 0) {
    if (MODE0 == 3) { for (int s = 0; s < 64; ++s) mbar_init(&bars[s], s < (n_ldgs_warps / 4) * depth ? 128 : 1); }
    else for (int s = 0; s < 64; ++s) mbar_init(&bars[s], MODE0 >= 1 ? 1 : 128);
    fence_barrier_init();
  }
  __syncthreads();
  const long long t0 = clock64();
  if (warp < nwarps) {
    // mixed mode: warps [0, n_ldgs_warps) are LDGSTS teams, the rest TMA warps; units share the tile list
    const int MODE = MODE0 == 3 ? (warp < n_ldgs_warps ? 0 : 1) : (MODE0 == 2 ? 1 : MODE0);
    const int nteams = MODE0 == 3 ? n_ldgs_warps / 4 : 0;
    const int unit = MODE0 == 3 ? (MODE == 0 ? warp / 4 : nteams + warp - n_ldgs_warps) : (MODE == 1 ? warp : warp / 4);
    const int nunits = MODE0 == 3 ? nteams + nwarps - n_ldgs_warps : (MODE == 1 ? nwarps : nwarps / 4);
    const int32_t *my = idx + (size_t)blockIdx.x * tiles_per_cta * kTile;
    int n = 0;
    for (int t = unit; t < tiles_per_cta; t += nunits, ++n) {
      const int j = n % depth, u = n / depth;
      const int slot = unit * depth + j;
      if (u >= 1) mbar_wait_warp(&bars[slot], (u - 1) & 1);   // previous use of the slot has landed
      const uint32_t dst = smem_u32(smem) + (slot % 12) * kTileBytes;
      if (MODE == 1) {
        int4 r = *reinterpret_cast<const int4 *>(my + (size_t)t * kTile + 4 * lane);
        if (zero_row >= 0) { r.x = r.x < 0 ? zero_row : r.x; r.y = r.y < 0 ? zero_row : r.y; r.z = r.z < 0 ? zero_row : r.z; r.w = r.w < 0 ? zero_row : r.w; }
        if (lane == 0) mbar_arrive_expect_tx(&bars[slot], kTileBytes);
        __syncwarp();
        tma_gather4(dst + lane * 512, &tm, 0, r, &bars[slot]);
      } else {
        const int q = warp & 3, c16 = lane & 7, rsub = lane >> 3, row0 = 32 * q + 8 * rsub;
        const int4 *ip = reinterpret_cast<const int4 *>(my + (size_t)t * kTile + row0);
        const int4 i0 = ip[0], i1 = ip[1];
        const int32_t ri[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
        const uint8_t *xb = x + c16 * 16;
#pragma unroll
        for (int i = 0; i < 8; ++i)
          cp_async16(dst + row0 * 128 + i * 128 + ((c16 ^ i) << 4), xb + (size_t)(uint32_t)max(ri[i], 0) * 128,
                     ri[i] >= 0 ? 16u : 0u);
        cp_async_arrive_noinc(&bars[slot]);
      }
    }
    for (int m = max(0, n - depth); m < n; ++m)   // drain: last use of every slot
      mbar_wait_warp(&bars[unit * depth + m % depth], (m / depth) & 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

int main(int argc, char **argv) {
  const int64_t N = 1 << 20;
  const int tiles_per_cta = argc > 2 ? atoi(argv[2]) : 1404, ncta = 148;
  const int only = argc > 3 ? atoi(argv[3]) : -1;
  const double density = argc > 1 ? atof(argv[1]) : 0.42;
  uint8_t *x;
  cudaMalloc(&x, (N + 1) * 128);
  cudaMemset(x, 1, N * 128);
  cudaMemset(x + N * 128, 0, 128);
  std::vector<int32_t> h((size_t)ncta * tiles_per_cta * kTile);
  srand(1);
  for (int c = 0; c < ncta; ++c)
    for (int t = 0; t < tiles_per_cta; ++t) {
      // 52 tile-offsets per output tile: consecutive output tiles walk through the cloud
      const int64_t tile = ((int64_t)c * tiles_per_cta + t) / 27;
      const int shift = (rand() % 6001) - 3000;
      for (int m = 0; m < kTile; ++m) {
        int64_t r = (tile * kTile + m + shift) % N;
        if (r < 0) r += N;
        h[((size_t)c * tiles_per_cta + t) * kTile + m] = (rand() / (double)RAND_MAX) < density ? (int32_t)r : -1;
      }
    }
  int32_t *idx;
  cudaMalloc(&idx, h.size() * 4);
  cudaMemcpy(idx, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
  unsigned long long *cyc;
  cudaMalloc(&cyc, ncta * 8);
  CUtensorMap tm;
  if (mk::make_tmap_rows16(&tm, MK_BF16, x, N + 1, 64, 64, 1)) { printf("tmap failed\n"); return 1; }
  cudaFuncSetAttribute(k_gather<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(k_gather<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(k_gather<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  // {mode, warps, depth, ldgs warps (mixed mode)}
  const int cfgs[][4] = {{0, 16, 3, 0}, {0, 32, 1, 0}, {1, 12, 1, 0}, {2, 8, 1, 0}, {2, 12, 1, 0}, {2, 16, 1, 0}, {2, 24, 1, 0}, {2, 32, 1, 0},
                         {2, 6, 2, 0}, {2, 12, 2, 0}, {3, 28, 1, 16}, {3, 32, 1, 16}, {3, 24, 1, 8}};
  int ci = -1;
  for (auto &cf : cfgs) {
    if (++ci != only && only >= 0) continue;
    const int mode = cf[0], nw = cf[1], depth = cf[2];
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e9;
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(a);
      if (nw * depth > 12 && mode != 0) { }  // smem: 12 slots
      if (mode == 0) k_gather<0><<<ncta, 1024, 200 * 1024>>>(tm, x, idx, tiles_per_cta, nw, depth, cyc, -1, 0);
      else if (mode == 3) k_gather<3><<<ncta, 1024, 200 * 1024>>>(tm, x, idx, tiles_per_cta, nw, depth, cyc, (int)N, cf[3]);
      else k_gather<1><<<ncta, 1024, 200 * 1024>>>(tm, x, idx, tiles_per_cta, nw, depth, cyc, mode == 2 ? (int)N : -1, 0);
      cudaEventRecord(b);
      if (cudaEventSynchronize(b) != cudaSuccess) { printf("error %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
      float ms; cudaEventElapsedTime(&ms, a, b);
      if (ms < best) best = ms;
    }
    std::vector<unsigned long long> hc(ncta);
    cudaMemcpy(hc.data(), cyc, ncta * 8, cudaMemcpyDeviceToHost);
    double mean = 0; for (auto v : hc) mean += v; mean /= ncta;
    printf("mode %s warps %2d depth %d: %.3f ms, %.0f cycles/tile/SM, nominal %.1f B/clk/SM, real %.2f TB/s\n",
           mode == 0 ? "LDGS" : mode == 1 ? "TMA " : mode == 2 ? "TMAz" : "MIX ", nw, depth, best, mean / tiles_per_cta, kTileBytes / (mean / tiles_per_cta),
           density * ncta * (double)tiles_per_cta * kTileBytes / best / 1e9);
  }
  return 0;
}
