/*
 * mink_b200.h — C-ABI of the B200-native sparse-tensor convolution engine.
 *
 * This is the drop-in boundary for the MinkowskiEngine hot path.  Every entry
 * point takes plain device pointers, sizes and a cudaStream_t (passed as
 * void*), returns 0 on success or a non-zero MK_E* status, and never aborts;
 * the message of the last failure on the calling thread is mk_last_error().
 *
 * Memory is owned by the caller (the Python host layer allocates everything
 * with torch.empty, i.e. PyTorch's caching allocator — the reference default
 * GPUMemoryAllocatorType.PYTORCH, MinkowskiCoordinateManager.py:46).  The
 * library keeps no state besides the per-thread error string.
 *
 * Each group cites the reference interface it replaces (paths relative to the
 * reference repo root).
 *
 * Coordinates are int32 rows [N, ncols] with the batch index first
 * (utils/collation.py:89-91), ncols = D+1 <= 4.  Internally a row is packed
 * into one 64-bit key: 16 bits batch (0..65535) + 3 x 16 bits signed
 * coordinate (-32768..32767).  Out-of-range rows raise MK_ERANGE.
 */
#ifndef MINK_B200_H_
#define MINK_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MK_OK 0
#define MK_EINVAL 1   /* bad argument                                   */
#define MK_ERANGE 2   /* coordinate outside the packable range          */
#define MK_ECUDA 3    /* CUDA runtime error (message in mk_last_error)  */
#define MK_EUNSUPPORTED 4
#define MK_EFULL 5    /* hash table full (capacity too small)           */

/* feature dtypes */
#define MK_F32 0
#define MK_F64 1
#define MK_BF16 2
#define MK_F16 3

/* pooling modes (subset of PoolingMode, pybind/extern.hpp:706-732) */
#define MK_POOL_SUM 0
#define MK_POOL_AVG 1
#define MK_POOL_MAX 2

/* broadcast modes (BroadcastMode, pybind/extern.hpp:734-739) */
#define MK_BCAST_ADD 0
#define MK_BCAST_MUL 1

const char *mk_last_error(void);
int mk_version(void);
/* number of CUDA kernels this library has launched in the process so far */
int64_t mk_launch_count(void);
/* 1 when the library was built with sm_100a tcgen05 kernels */
int mk_has_tcgen05(void);

/* ------------------------------------------------------------------------
 * Coordinate hash.  Replaces CoordinateMapManager::insert_and_map and the
 * cudf concurrent_unordered_map kernels behind it (pybind/extern.hpp:779;
 * MinkowskiCoordinateManager.py:153-179).
 *
 * The table is capacity*16 bytes (32-byte aligned): capacity/2 buckets of one
 * 32-byte sector {uint64 run key, uint32 row[4], pad}, a bucket holding a run of
 * 4 cells along the first spatial axis.  capacity is a power of two from
 * mk_hash_capacity(n) (>= 4n, i.e. <= 50% bucket load in the worst case).
 * ---------------------------------------------------------------------- */
int64_t mk_hash_capacity(int64_t n);
/* bytes of scratch mk_map_insert needs for n rows */
int64_t mk_insert_workspace_bytes(int64_t n);

/*
 * Insert n rows, deduplicate (first occurrence wins), and compact.
 *   table        [capacity*16 B]  out: key -> compact row id
 *   unique_index [n]   out: ascending original row of each unique coordinate
 *   inverse_map  [n]   out: compact row id of every original row
 *   coords_out   [n*ncols] out: coordinates of the unique rows, compact order
 *   n_unique_dev [1]   out (device int64)
 *   status_dev   [1]   out (device int32): 0 ok, MK_ERANGE, MK_EFULL
 * Any of unique_index / inverse_map / coords_out may be NULL.
 */
int mk_map_insert(const int32_t *coords, int64_t n, int ncols, void *table,
                  int64_t capacity, int32_t *unique_index,
                  int32_t *inverse_map, int32_t *coords_out, void *workspace,
                  int64_t *n_unique_dev, int32_t *status_dev, void *stream);

/* mk_map_insert in three launches (four with the inverse map): the table clear also zeroes
 * the scan state, the dedup / ordered compaction / table update is ONE pass (decoupled
 * look-back scan), and the maps can be written as int64 directly (what the Python API
 * returns, MinkowskiCoordinateManager.py:153-179).  flags:
 *   MK_INSERT_UNIQUE      the caller guarantees distinct rows (children of distinct octree
 *                         parents - extern.hpp:151-165 with kernel 2 / stride 2 -, pruned rows
 *                         of a map - extern.hpp:411-419): no dedup pass, ids = row numbers
 *   MK_INSERT_INT64_MAPS  unique_index / inverse_map are int64 [n]
 * Same outputs as mk_map_insert otherwise (tested against it). */
#define MK_INSERT_UNIQUE 1
#define MK_INSERT_INT64_MAPS 2
int64_t mk_insert2_workspace_bytes(int64_t n);
int mk_map_insert2(const int32_t *coords, int64_t n, int ncols, void *table,
                   int64_t capacity, int flags, void *unique_index,
                   void *inverse_map, int32_t *coords_out, void *workspace,
                   int64_t *n_unique_dev, int32_t *status_dev, void *stream);

/* rows[i] = compact row id of query[i] in the table, or -1.
 * Replaces thrust find_coordinate. */
int mk_map_lookup(const void *table, int64_t capacity, const int32_t *query,
                  int64_t nq, int ncols, int32_t *rows, void *stream);

/* ------------------------------------------------------------------------
 * Coordinate generation.  Replaces manager.stride (extern.hpp:789), the
 * expand_coordinates path of ConvolutionTransposeForwardGPU
 * (extern.hpp:151-165) and origin maps (extern.hpp:790).
 * ---------------------------------------------------------------------- */
/* out[i, 1+d] = floor(in[i, 1+d] / new_stride[d]) * new_stride[d]; batch kept */
int mk_coords_stride(const int32_t *in, int64_t n, int ncols,
                     const int32_t *new_stride_host, int32_t *out,
                     void *stream);
/* out[i*K + k, :] = in[i, :] + (0, offsets[k, :]) ; offsets host [K, ncols-1] */
int mk_coords_expand(const int32_t *in, int64_t n, int ncols,
                     const int32_t *offsets_host, int K, int32_t *out,
                     void *stream);
/* out[i, :] = (in[i, 0], 0, 0, 0) */
int mk_coords_origin(const int32_t *in, int64_t n, int ncols, int32_t *out,
                     void *stream);

/* ------------------------------------------------------------------------
 * Pruning.  Replaces PruningForwardGPU / PruningBackwardGPU
 * (extern.hpp:411-428; MinkowskiPruning.py:38-72).
 * ---------------------------------------------------------------------- */
int64_t mk_select_workspace_bytes(int64_t n);
/* kept[j] = index of the j-th non-zero keep byte (order preserved) */
int mk_select_flagged(const uint8_t *keep, int64_t n, int32_t *kept,
                      void *workspace, int64_t *n_kept_dev, void *stream);
/* mk_select_flagged for callers that cannot allocate (CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in
 * the reference.  Same result as
 * mk_select_flagged, but the scan scratch is a fixed static __device__ array
 * inside the library (33 KB per GPU that loads it, held until exit) instead of
 * a caller workspace; the call performs no cudaMalloc / cudaFree and calls no
 * allocating routine.  n > mk_select_static_max_rows() (16 777 216) is rejected
 * with MK_EINVAL.  Every call on a device shares that one array: calls that
 * overlap on the GPU (different streams) race; serialise them on one stream. */
int64_t mk_select_static_max_rows(void);
int mk_select_flagged_static(const uint8_t *keep, int64_t n, int32_t *kept,
                             int64_t *n_kept_dev, void *stream);
/* dst[j, :] = src[idx[j], :]   (row_bytes per row, any dtype) */
int mk_rows_gather(const void *src, const int32_t *idx, int64_t n_idx,
                   int64_t row_bytes, void *dst, void *stream);
/* dst[idx[j], :] = src[j, :]   (dst rows not listed are left untouched) */
int mk_rows_scatter(const void *src, const int32_t *idx, int64_t n_idx,
                    int64_t row_bytes, void *dst, void *stream);

/* ------------------------------------------------------------------------
 * Kernel maps.  Replaces the count_kernel / preallocated_kernel_map_iteration
 * pair behind CoordinateMapManager::kernel_map (extern.hpp:804;
 * MinkowskiCoordinateManager.py:377-421).
 *
 * The engine's native form is the output-stationary neighbour table
 *   nbr[k * ld + o] = input row hit by output row o under kernel offset k,
 *                     or -1
 * (k-major so that a tile of consecutive output rows reads/writes
 * contiguous memory).  The padding entries o in [n_out, ld) of every k are
 * written as -1 by mk_kmap_build / mk_kmap_invert; the convolution kernels
 * rely on every entry being -1 or a valid row.  `offsets_host` is [K, ncols-1] in coordinate units
 * (already multiplied by dilation * tensor_stride).
 * ---------------------------------------------------------------------- */
int mk_kmap_build(const int32_t *out_coords, int64_t n_out, int ncols,
                  const void *in_table, int64_t in_capacity,
                  const int32_t *offsets_host, int K, int32_t *nbr, int64_t ld,
                  int64_t *counts_dev /* [K] pair counts, may be NULL */,
                  void *stream);
/* inverse table over input rows: inv[k*ld_in + nbr[k*ld+o]] = o, else -1 */
int mk_kmap_invert(const int32_t *nbr, int64_t n_out, int64_t ld, int K,
                   int32_t *inv, int64_t n_in, int64_t ld_in, void *stream);
int64_t mk_kmap_pairs_workspace_bytes(int64_t n_out, int K);
/* CSR export: for offset k, pairs [starts[k], starts[k+1]) ascending in o.
 * Two calls with the same workspace: first with in_rows == NULL (counts ->
 * starts_dev [K+1] int64; the caller reads starts_dev[K] and allocates), then
 * with in_rows/out_rows sized >= starts[K] to fill them. */
int mk_kmap_pairs(const int32_t *nbr, int64_t n_out, int64_t ld, int K,
                  int64_t *starts_dev, int32_t *in_rows, int32_t *out_rows,
                  void *workspace, void *stream);

/* Octree split: kernel map of a generative transposed convolution whose kernel equals its
 * stride (extern.hpp:151-165 with expand_coordinates = true; modules/vae_block.py:279-296),
 * where the child rows are i*K + k by construction: the table, its inverse (dense), the tile
 * order and the tile-ordered table are written in closed form, no probes.  Any output may be NULL. */
int mk_kmap_children(int64_t n_in, int K, int64_t ld_out, int64_t ld_in,
                     int32_t *nbr, int32_t *inv, int32_t *order,
                     int32_t *nbr_ordered, void *stream);

/* Per 128-row tile of the table, a K-bit mask of the offsets that have at
 * least one neighbour in the tile: mask[tile * W + k/32] bit (k%32), with
 * W = mk_kmap_tile_mask_words(K).  The tensor-core convolution kernels use it
 * to skip empty (tile, offset) pairs; it depends only on the kernel map, so it
 * is built once and cached beside it. */
int64_t mk_kmap_tile_mask_words(int K);
int mk_kmap_tile_mask(const int32_t *nbr, int64_t n_rows, int64_t ld, int K,
                      uint32_t *mask, void *stream);

/* Tile ordering (an internal layout of the engine, no counterpart in the
 * reference: ME's gather-GEMM-scatter works on per-offset pair lists).  The
 * tensor-core kernels process 128 table columns at a time and skip an offset
 * only when none of the 128 rows has a neighbour under it.  `order` [n_rows]
 * is a permutation of the table's columns sorted by (batch, 2^block_log2
 * block of the row's coordinate, neighbour pattern), nbr_ordered [K, ld] the
 * same table with column s = old column order[s] (padding -1).  Row indices
 * INSIDE the table are untouched.  coords may be NULL (pattern only).  The
 * coordinate map's row order does not change: mk_conv_fwd_rows writes row
 * order[s]; the weight gradient reads dy row order[s] (mk_conv_wgrad_rows). */
int64_t mk_kmap_tile_order_workspace_bytes(int64_t n_rows);
int mk_kmap_tile_order(const int32_t *nbr, int64_t n_rows, int64_t ld, int K,
                       const int32_t *coords, int ncols, int block_log2,
                       int32_t *order, int32_t *nbr_ordered, void *workspace,
                       void *stream);

/* ------------------------------------------------------------------------
 * Convolution.  Replaces ConvolutionForwardGPU / ConvolutionBackwardGPU and
 * the Transpose variants (extern.hpp:84-114,151-181): the matmul / matmul2
 * gather-GEMM-scatter kernels.
 *
 *   y[o, :] = sum_k  x[nbr[k,o], :] @ W[k]          W: [K, Cin, Cout]
 *
 * mk_conv_fwd with (transpose_w = 1) computes x @ W[k]^T (W: [K, Cout_, Cin_]
 * seen from the caller as [K, c_out_of_this_call... ]) and is the data
 * gradient when called with the inverse table.  `kflip` != 0 reads offset
 * K-1-k of the table for weight k (stride-1 same-map shortcut: the inverse
 * table of a symmetric kernel is the table itself with offsets mirrored).
 * bf16/f16 accumulate in fp32; f64 accumulates in f64.
 * `tile_mask` (mk_kmap_tile_mask of the SAME table) may be NULL: the tensor-
 * core path then derives it into the workspace on every call.
 * `n_in` is the number of rows of x (table entries outside [0, n_in) count as
 * missing).  The tensor-core path needs ld % 128 == 0 and 16-byte aligned
 * x / nbr; other layouts are served by the SIMT kernels.
 * ---------------------------------------------------------------------- */
int64_t mk_conv_workspace_bytes(int dtype, int64_t n_rows, int K, int c_in,
                                int c_out);
int mk_conv_fwd(int dtype, const void *x, const void *w, int transpose_w,
                int kflip, const int32_t *nbr, const uint32_t *tile_mask,
                int64_t ld, int64_t n_out, int64_t n_in,
                int K, int c_in, int c_out, void *y, void *workspace,
                void *stream);
/* 1 if the tcgen05 kernels (forward and weight gradient) serve this shape. */
int mk_conv_tc_eligible(int dtype, int64_t n_rows, int K, int c_in, int c_out);
/* mk_conv_fwd on a tile-ordered table (mk_kmap_tile_order): column o of the
 * table is output row out_rows[o], i.e. y[out_rows[o], :] = sum_k ...  Served
 * by the tensor-core path only (MK_EUNSUPPORTED otherwise); out_rows == NULL
 * is mk_conv_fwd. */
int mk_conv_fwd_rows(int dtype, const void *x, const void *w, int transpose_w,
                     int kflip, const int32_t *nbr, const uint32_t *tile_mask,
                     int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in,
                     int c_out, void *y, const int32_t *out_rows,
                     void *workspace, void *stream);
/* dW[k] = sum_o x[nbr[k,o], :]^T  dy[o, :]    dW: [K, Cin, Cout] (dtype) */
int mk_conv_wgrad(int dtype, const void *x, const void *dy,
                  const int32_t *nbr, const uint32_t *tile_mask, int64_t ld,
                  int64_t n_out, int64_t n_in, int K,
                  int c_in, int c_out, void *dw, void *workspace,
                  void *stream);
/* mk_conv_wgrad on a tile-ordered table: table column o pairs with dy row
 * dy_rows[o] (the `order` of mk_kmap_tile_order; 16-byte aligned), read inside
 * the kernel with TMA tile::gather4 - no permuted copy of dy is made.  Served
 * by the tensor-core path only (MK_EUNSUPPORTED otherwise); dy_rows == NULL is
 * mk_conv_wgrad. */
int mk_conv_wgrad_rows(int dtype, const void *x, const void *dy,
                       const int32_t *dy_rows, const int32_t *nbr,
                       const uint32_t *tile_mask, int64_t ld, int64_t n_out,
                       int64_t n_in, int K, int c_in, int c_out, void *dw,
                       void *workspace, void *stream);
/* The same two operations with the workspace SIZE passed in (workspace_bytes, from
 * mk_conv_workspace_bytes2, which also takes n_in).  They serve everything the entry
 * points above serve and additionally
 *  - MK_F32 features on the tensor cores: operands are split into bf16 pieces (x = x0 +
 *    x1 + x2, each difference exact) and the six products above 2^-24 are accumulated in
 *    one fp32 TMEM accumulator; channels are padded to multiples of 16 in the workspace, so
 *    the narrow layers (Cin or Cout of 1, 3, 4) take this path too.  The reference computes
 *    fp32 (MinkowskiConvolution.py:265,285; cuBLAS SGEMM behind extern.hpp:84-114);
 *  - dw_f32 != 0: the weight gradient of 16-bit features is written as float32
 *    [K, Cin, Cout] (the fp32 sum the kernel holds anyway), for fp32 parameters whose
 *    gradients are reduced over ranks in fp32 (examples/multigpu_ddp.py:81-106).
 * workspace_bytes < what the split path needs (or -1) falls back to the SIMT kernels. */
int64_t mk_conv_workspace_bytes2(int dtype, int64_t n_out, int64_t n_in, int K,
                                 int c_in, int c_out);
int mk_conv_fwd_ex(int dtype, const void *x, const void *w, int transpose_w,
                   int kflip, const int32_t *nbr, const uint32_t *tile_mask,
                   int64_t ld, int64_t n_out, int64_t n_in, int K, int c_in,
                   int c_out, void *y, const int32_t *out_rows,
                   void *workspace, int64_t workspace_bytes, void *stream);
int mk_conv_wgrad_ex(int dtype, const void *x, const void *dy,
                     const int32_t *dy_rows, const int32_t *nbr,
                     const uint32_t *tile_mask, int64_t ld, int64_t n_out,
                     int64_t n_in, int K, int c_in, int c_out, void *dw,
                     int dw_f32, void *workspace, int64_t workspace_bytes,
                     void *stream);
/* 1 if the row-permuted forms (out_rows / dy_rows on a tile-ordered table) are served for this
 * dtype and shape: the tensor-core kernels, or for MK_F32 beyond the split path's accumulation
 * limit the tuned FFMA kernels (conv_simt2.cu; c_in and c_out multiples of 4). */
int mk_conv_rows_eligible(int dtype, int64_t n_rows, int K, int c_in, int c_out);
/* which implementation the last mk_conv_* call on this thread used:
 * 0 = SIMT tiled kernel, 1 = tcgen05/TMEM kernel */
int mk_conv_last_path(void);
/* force the SIMT path (testing); returns previous value */
int mk_conv_force_simt(int on);

/* ------------------------------------------------------------------------
 * Pooling / broadcast.  Replaces LocalPooling*GPU, GlobalPooling*GPU,
 * Broadcast*GPU (extern.hpp:215-392).
 * ---------------------------------------------------------------------- */
/* y[o,:] = reduce_k x[nbr[k,o],:];  nnz[o] = #valid k (dtype of x);
 * argmax [n_out*c] int32 (MK_POOL_MAX only, may be NULL otherwise) */
int mk_pool_fwd(int dtype, int mode, const void *x, const int32_t *nbr,
                int64_t ld, int64_t n_out, int K, int c, void *y, void *nnz,
                int32_t *argmax, void *stream);
/* dx (zero-initialised by the callee), via the inverse table (sum/avg) or the
 * arg-max indices (max) */
int mk_pool_bwd(int dtype, int mode, const void *dy, const void *nnz,
                const int32_t *inv, int64_t ld_in, int64_t n_in, int K, int c,
                const int32_t *argmax, int64_t n_out, void *dx, void *stream);
/* scratch for the batch-segment kernels below */
int64_t mk_segment_workspace_bytes(int dtype, int c, int64_t n_batch);
/* batch-segment reduction: y[b,:] = reduce_{r: batch(r)=b} x[r,:].
 * batch_of_row [n] int32, n_batch rows out; cnt [n_batch] (dtype) out;
 * argmax [n_batch*c] for MK_POOL_MAX */
int mk_global_pool_fwd(int dtype, int mode, const void *x,
                       const int32_t *batch_of_row, int64_t n, int c,
                       int64_t n_batch, void *y, void *cnt, int32_t *argmax,
                       void *workspace, void *stream);
int mk_global_pool_bwd(int dtype, int mode, const void *dy, const void *cnt,
                       const int32_t *batch_of_row, int64_t n, int c,
                       int64_t n_batch, const int32_t *argmax, void *dx,
                       void *stream);
/* y[r,:] = x[r,:] (+|*) g[batch(r),:] */
int mk_broadcast_fwd(int dtype, int mode, const void *x, const void *g,
                     const int32_t *batch_of_row, int64_t n, int c, void *y,
                     void *stream);
/* dx [n,c], dg [n_batch,c] */
int mk_broadcast_bwd(int dtype, int mode, const void *x, const void *g,
                     const void *dy, const int32_t *batch_of_row, int64_t n,
                     int c, int64_t n_batch, void *dx, void *dg,
                     void *workspace, void *stream);

/* ------------------------------------------------------------------------
 * Normalisation passes.  Replace, for MinkowskiBatchNorm / SyncBatchNorm
 * (MinkowskiNormalization.py:55-195) and the instance-norm family (:198-622:
 * two GlobalPooling + three Broadcast calls + elementwise torch ops per layer,
 * extern.hpp:332-392), every full pass over [N, C] by one statistics read and
 * one normalise+affine read/write.  A segment is a batch item (seg_of_row, the
 * batch row of every voxel; negative = row ignored) or the whole tensor
 * (seg_of_row == NULL, n_seg = 1).  Sums and coefficients are float32 (float64
 * for MK_F64 features), shaped [n_seg, 2, c] / [n_seg, c].
 *   mk_seg_sums:   out[s][0][c] = sum_r (a - shift_a[s,c]),
 *                  out[s][1][c] = sum_r (a - shift_a[s,c]) * (b - shift_b[s,c])
 *                  (shift_* may be NULL = 0; out is zeroed by the call)
 *   mk_seg_affine: out[r,c] = coef_a[s,c] * a[r,c] + coef_b[s,c] * b[r,c] + coef_c[s,c]
 *                  (coef_b / b may be NULL)
 * ---------------------------------------------------------------------- */
int mk_seg_sums(int dtype, const void *a, const void *b, const void *shift_a,
                const void *shift_b, const int32_t *seg_of_row, int64_t n, int c,
                int64_t n_seg, void *out, void *stream);
int mk_seg_affine(int dtype, const void *a, const void *b,
                  const int32_t *seg_of_row, int64_t n, int c,
                  const void *coef_a, const void *coef_b, const void *coef_c,
                  void *out, void *stream);

/* Batch-norm coefficients between the two passes, one launch each (torch.nn.BatchNorm1d semantics,
 * MinkowskiNormalization.py:55-103).  All arrays are [c] of the accumulator type (float, or double
 * when acc_f64); the row count m is count_dev[0] (SyncBatchNorm: all-reduced on the device) or
 * count_host.
 *   fwd: sums = mk_seg_sums(x, x, shift, shift) -> mean, invstd, y = coef_a x + coef_c, and, when
 *        running_mean / running_var are given, r = (1 - momentum) r + momentum {mean, unbiased var}
 *   bwd: sums = mk_seg_sums(dy, x, NULL, mean)  -> dx = coef_a dy + coef_b x + coef_c, dweight, dbias */
int mk_bn_coef_fwd(int acc_f64, const void *sums, const void *shift, const void *count_dev, double count_host,
                   const void *weight, const void *bias, double eps, int c, void *coef_a, void *coef_c,
                   void *mean_out, void *invstd_out, void *running_mean, void *running_var, double momentum,
                   void *stream);
int mk_bn_coef_bwd(int acc_f64, const void *sums, const void *mean, const void *invstd, const void *count_dev,
                   double count_host, const void *weight, int c, void *coef_a, void *coef_b, void *coef_c,
                   void *dweight, void *dbias, void *stream);

/* Instance / group norm coefficients in one launch each way (MinkowskiInstanceNorm,
 * MinkowskiStableInstanceNorm, MinkowskiStableGroupNorm: MinkowskiNormalization.py:198-478).  Per
 * batch item the statistics are taken over the item's rows and `gs` adjacent channels; weight / bias
 * [c / gw] are shared by `gw` adjacent channels.  Arrays are of the accumulator type; sums [n_seg, 2, c],
 * shift / coefficients / mean [n_seg, c], mu / inv [n_seg, c / gs], counts [n_seg].
 *   fwd: sums = mk_seg_sums(x, x, shift, shift, seg)  -> y = coef_a[seg] x + coef_c[seg]
 *   bwd: sums = mk_seg_sums(dy, x, NULL, mean, seg)   -> dx = coef_a dy + coef_b x + coef_c, dweight, dbias */
int mk_gn_coef_fwd(int acc_f64, const void *sums, const void *shift, const void *counts, const void *weight,
                   const void *bias, double eps, int c, int gs, int gw, int64_t n_seg, void *coef_a, void *coef_c,
                   void *mean_out, void *mu_out, void *inv_out, void *stream);
int mk_gn_coef_bwd(int acc_f64, const void *sums, const void *mean, const void *mu, const void *inv,
                   const void *counts, const void *weight, int c, int gs, int gw, int64_t n_seg, void *coef_a,
                   void *coef_b, void *coef_c, void *dweight, void *dbias, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* MINK_B200_H_ */
