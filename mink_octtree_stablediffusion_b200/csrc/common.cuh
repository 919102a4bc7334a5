// Shared device/host helpers for the mink_b200 engine (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/mink_b200.h"

namespace mk {

// ---- error plumbing -------------------------------------------------------
void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);

#define MK_CHECK_ARG(cond, msg)                                                \
  do {                                                                         \
    if (!(cond)) {                                                             \
      mk::set_error("%s: %s", __func__, msg);                                  \
      return MK_EINVAL;                                                        \
    }                                                                          \
  } while (0)

#define MK_CUDA(call)                                                          \
  do {                                                                         \
    cudaError_t e__ = (call);                                                  \
    if (e__ != cudaSuccess) return mk::cuda_fail(e__, #call);                  \
  } while (0)

// every kernel launch of the library passes through MK_LAUNCH_CHECK, which also
// feeds mk_launch_count() (bench.py's gpu_launches)
extern unsigned long long g_launch_count;

#define MK_LAUNCH_CHECK()                                                      \
  do {                                                                         \
    ++mk::g_launch_count;                                                      \
    cudaError_t e__ = cudaGetLastError();                                      \
    if (e__ != cudaSuccess) return mk::cuda_fail(e__, __func__);               \
  } while (0)

// batched single-CTA exclusive scan (hash.cu): row b of block_sums[batch][nblocks]
// is scanned in place, its total written to total_dev[b]
int scan_block_sums(int32_t *block_sums, int64_t nblocks, int batch,
                    int64_t *total_dev, cudaStream_t st);

// 4 words of mapped host memory a timed-out mbarrier wait reports into
// (sm100.cuh); device pointer or nullptr.  mk_last_error() decodes it.
unsigned int *hang_report_device_ptr();

inline int num_sms() {
  // per device (a process may drive several GPUs, one manager each)
  static int n[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (n[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
    n[dev] = v;
  }
  return n[dev];
}

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---- coordinate keys --------------------------------------------------------
// One coordinate row (b, x, y, z) -> 64-bit key.  Unused trailing axes are 0.
// key = b << 48 | (x+32768) << 32 | (y+32768) << 16 | (z+32768); the hash table
// is keyed by runs of 4 cells along x (run_key below).
constexpr uint64_t kEmptyKey = 0xFFFFFFFFFFFFFFFFull;  // b=65535 is reserved

// A bucket is one aligned 32-byte sector and holds a RUN of 4 cells along the
// first spatial axis: key = (b, x>>2, y, z), row[x&3] = compact row id (or
// 0xFFFFFFFF).  Kernel offsets enumerate that axis fastest, so the 3 (or 2, 4,
// 5 ...) neighbours of a kernel row cost one sector — two when the run
// boundary is crossed — instead of one sector each.
struct __align__(32) Bucket {
  unsigned long long key;
  uint32_t row[4];
  uint32_t pad[2];
};
constexpr uint32_t kNoRow = 0xFFFFFFFFu;

__host__ __device__ __forceinline__ bool coord_in_range(int b, int x, int y,
                                                        int z) {
  return b >= 0 && b < 65535 && x >= -32768 && x <= 32767 && y >= -32768 &&
         y <= 32767 && z >= -32768 && z <= 32767;
}

__host__ __device__ __forceinline__ uint64_t pack_key(int b, int x, int y,
                                                      int z) {
  return ((uint64_t)(uint32_t)(b & 0xFFFF) << 48) |
         ((uint64_t)(uint32_t)((x + 32768) & 0xFFFF) << 32) |
         ((uint64_t)(uint32_t)((y + 32768) & 0xFFFF) << 16) |
         (uint64_t)(uint32_t)((z + 32768) & 0xFFFF);
}

// home bucket of a run key {hi = b<<16 | run, lo = y<<16 | z}: multiply-add of
// the two words, one xorshift-multiply round, TOP bits (mask = buckets-1, a
// power of two minus one, >= 511).  5 integer instructions: the kernel-map
// build evaluates it ~14 times per voxel.
constexpr uint32_t kHashC1 = 0x9E3779B1u, kHashC2 = 0x85EBCA77u, kHashC3 = 0xC2B2AE3Du;

__host__ __device__ __forceinline__ uint32_t hash_mix(uint32_t x) {
  x ^= x >> 16;
  x *= kHashC3;
  return x;
}

__device__ __forceinline__ uint32_t bucket_home(uint32_t hi, uint32_t lo,
                                                uint32_t mask) {
  return hash_mix(hi * kHashC1 + lo * kHashC2) >> __clz(mask);
}

__device__ __forceinline__ uint32_t bucket_home(uint64_t key, uint32_t mask) {
  return bucket_home((uint32_t)(key >> 32), (uint32_t)key, mask);
}

// read one coordinate row of ncols (<=4) int32 into b,x,y,z (missing = 0)
__device__ __forceinline__ void load_coord(const int32_t *__restrict__ c,
                                           int64_t i, int ncols, int &b,
                                           int &x, int &y, int &z) {
  if (ncols == 4) {
    int4 v = __ldg(reinterpret_cast<const int4 *>(c) + i);
    b = v.x; x = v.y; y = v.z; z = v.w;
  } else {
    const int32_t *p = c + i * ncols;
    b = p[0];
    x = ncols > 1 ? p[1] : 0;
    y = ncols > 2 ? p[2] : 0;
    z = 0;
  }
}

__device__ __forceinline__ void store_coord(int32_t *__restrict__ c, int64_t i,
                                            int ncols, int b, int x, int y,
                                            int z) {
  if (ncols == 4) {
    reinterpret_cast<int4 *>(c)[i] = make_int4(b, x, y, z);
  } else {
    int32_t *p = c + i * ncols;
    p[0] = b;
    if (ncols > 1) p[1] = x;
    if (ncols > 2) p[2] = y;
  }
}

// ---- probing -----------------------------------------------------------------
// Linear probing over buckets; a probe step is one LDG.256 of one sector.
struct BucketRegs { uint32_t v[8]; };  // {key lo, key hi, row0..row3, pad, pad}

__device__ __forceinline__ BucketRegs ldg_bucket(const Bucket *p) {
  BucketRegs r;
  asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]),
                 "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7])
               : "l"(p));
  return r;
}

// key of the run that holds (b,x,y,z); the cell is row[x & 3] of its bucket
__host__ __device__ __forceinline__ uint64_t run_key(int b, int x, int y,
                                                     int z) {
  return ((uint64_t)(uint32_t)(b & 0xFFFF) << 48) |
         ((uint64_t)(uint32_t)(((x + 32768) & 0xFFFF) >> 2) << 32) |
         ((uint64_t)(uint32_t)((y + 32768) & 0xFFFF) << 16) |
         (uint64_t)(uint32_t)((z + 32768) & 0xFFFF);
}

__device__ __forceinline__ uint64_t bucket_key(const BucketRegs &s) {
  return ((uint64_t)s.v[1] << 32) | s.v[0];
}

__device__ __forceinline__ int32_t bucket_row(const BucketRegs &s, int sub) {
  const uint32_t lo = (sub & 1) ? s.v[3] : s.v[2];
  const uint32_t hi = (sub & 1) ? s.v[5] : s.v[4];
  return (int32_t)((sub & 2) ? hi : lo);  // kNoRow reads as -1
}

// probe the table for cell (b,x,y,z); returns compact row or -1
__device__ __forceinline__ int32_t table_find(const Bucket *__restrict__ table,
                                              uint32_t mask, int b, int x,
                                              int y, int z) {
  const uint64_t key = run_key(b, x, y, z);
  uint32_t h = bucket_home(key, mask);
  while (true) {
    const BucketRegs s = ldg_bucket(table + h);
    const uint64_t k = bucket_key(s);
    if (k == key) return bucket_row(s, x & 3);
    if (k == kEmptyKey) return -1;
    h = (h + 1) & mask;
  }
}

// ---- feature dtypes -----------------------------------------------------------
template <typename T> struct Acc { using type = float; };
template <> struct Acc<double> { using type = double; };

template <typename T> __device__ __forceinline__ typename Acc<T>::type to_acc(T v) { return (typename Acc<T>::type)v; }
template <> __device__ __forceinline__ float to_acc<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_acc<__half>(__half v) { return __half2float(v); }

template <typename T, typename A> __device__ __forceinline__ T from_acc(A v) { return (T)v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_acc<__nv_bfloat16, float>(float v) { return __float2bfloat16(v); }
template <> __device__ __forceinline__ __half from_acc<__half, float>(float v) { return __float2half(v); }

inline int dtype_size(int dtype) {
  switch (dtype) {
    case MK_F32: return 4;
    case MK_F64: return 8;
    case MK_BF16: return 2;
    case MK_F16: return 2;
  }
  return 0;
}

#define MK_DISPATCH_DTYPE(dtype, ...)                                          \
  switch (dtype) {                                                             \
    case MK_F32: { using T = float; __VA_ARGS__; break; }                      \
    case MK_F64: { using T = double; __VA_ARGS__; break; }                     \
    case MK_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; }             \
    case MK_F16: { using T = __half; __VA_ARGS__; break; }                     \
    default: mk::set_error("%s: unsupported dtype %d", __func__, dtype);       \
             return MK_EUNSUPPORTED;                                           \
  }

}  // namespace mk
