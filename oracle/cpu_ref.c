/*
 * cpu_ref.c — CPU restatement (C + OpenMP) of the MinkowskiEngine CPU backend's
 * algorithm for the hot path.  TEST / BASELINE INFRASTRUCTURE ONLY: used by
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs as the checker and the timed CPU baseline; never by the product.
 *
 * PARITY UNPINNED (see oracle/oracle.py header): the reference ships no engine
 * sources, no loadable binary and no golden vectors; this restates the
 * published semantics of NVIDIA MinkowskiEngine 0.5.4 (CPU path):
 *   - coordinate map  = flat open-addressing hash of the int32 row -> row index
 *                        (reference: coordinate_map_cpu.hpp, robin_hood map;
 *                         entry point pybind/extern.hpp:779 insert_and_map)
 *   - kernel map      = per kernel offset, probe the in-map at out + offset and
 *                        append (in_row, out_row) (extern.hpp:804 kernel_map;
 *                        kernel_region.hpp iterator, first axis fastest)
 *   - convolution     = per offset: gather rows -> GEMM with W[k] -> scatter-add
 *                        (ConvolutionForwardCPU / BackwardCPU, extern.hpp:53-81)
 * Threads: OpenMP, all cores unless OMP_NUM_THREADS says otherwise (ME's own
 * default caps at 16-20: MinkowskiEngine/__init__.py:34-46).
 */
#include <omp.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { int32_t c[4]; int32_t row; } slot_t;

static inline uint32_t hash4(const int32_t *c) {
  /* murmur3-style mix of the four int32 (reference: coordinate_murmur3) */
  uint32_t h = 0x9747b28cu;
  for (int i = 0; i < 4; ++i) {
    uint32_t k = (uint32_t)c[i];
    k *= 0xcc9e2d51u; k = (k << 15) | (k >> 17); k *= 0x1b873593u;
    h ^= k; h = (h << 13) | (h >> 19); h = h * 5u + 0xe6546b64u;
  }
  h ^= 16u; h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}

static inline void load4(const int32_t *coords, int64_t i, int ncols, int32_t *o) {
  o[0] = o[1] = o[2] = o[3] = 0;
  for (int d = 0; d < ncols; ++d) o[d] = coords[i * ncols + d];
}

typedef struct { slot_t *slots; uint64_t mask; } cmap_t;

int cr_num_threads(void) { return omp_get_max_threads(); }
void cr_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }

void *cr_map_create(int64_t n) {
  cmap_t *m = (cmap_t *)malloc(sizeof(cmap_t));
  uint64_t cap = 1024;
  while (cap < (uint64_t)(2 * n)) cap <<= 1;
  m->slots = (slot_t *)malloc(cap * sizeof(slot_t));
  m->mask = cap - 1;
  for (uint64_t i = 0; i < cap; ++i) m->slots[i].row = -1;
  return m;
}

void cr_map_destroy(void *p) {
  cmap_t *m = (cmap_t *)p;
  free(m->slots);
  free(m);
}

static inline int32_t map_find(const cmap_t *m, const int32_t *c) {
  uint64_t h = hash4(c) & m->mask;
  for (;;) {
    const slot_t *s = &m->slots[h];
    if (s->row < 0) return -1;
    if (s->c[0] == c[0] && s->c[1] == c[1] && s->c[2] == c[2] && s->c[3] == c[3]) return s->row;
    h = (h + 1) & m->mask;
  }
}

/* Sequential insert in row order (the ME CPU map is single-writer): the first
 * occurrence of a coordinate gets the next compact row id.  Returns n_unique. */
int64_t cr_insert_and_map(void *p, const int32_t *coords, int64_t n, int ncols,
                          int64_t *unique_index, int64_t *inverse_map) {
  cmap_t *m = (cmap_t *)p;
  int64_t nu = 0;
  for (int64_t i = 0; i < n; ++i) {
    int32_t c[4];
    load4(coords, i, ncols, c);
    uint64_t h = hash4(c) & m->mask;
    for (;;) {
      slot_t *s = &m->slots[h];
      if (s->row < 0) {
        memcpy(s->c, c, sizeof(c));
        s->row = (int32_t)nu;
        if (unique_index) unique_index[nu] = i;
        if (inverse_map) inverse_map[i] = nu;
        ++nu;
        break;
      }
      if (s->c[0] == c[0] && s->c[1] == c[1] && s->c[2] == c[2] && s->c[3] == c[3]) {
        if (inverse_map) inverse_map[i] = s->row;
        break;
      }
      h = (h + 1) & m->mask;
    }
  }
  return nu;
}

/* Kernel map: for offset k, pairs are written to in_rows/out_rows[k*n_out ...]
 * ascending in the out row; counts[k] = number of pairs. */
void cr_kernel_map(const void *in_map, const int32_t *out_coords, int64_t n_out,
                   int ncols, const int32_t *offsets, int K, int64_t *counts,
                   int32_t *in_rows, int32_t *out_rows) {
  const cmap_t *m = (const cmap_t *)in_map;
  const int nd = ncols - 1;
#pragma omp parallel for schedule(dynamic, 1)
  for (int k = 0; k < K; ++k) {
    int64_t cnt = 0;
    int32_t *ir = in_rows + (int64_t)k * n_out, *orr = out_rows + (int64_t)k * n_out;
    for (int64_t o = 0; o < n_out; ++o) {
      int32_t c[4];
      load4(out_coords, o, ncols, c);
      for (int d = 0; d < nd; ++d) c[1 + d] += offsets[k * nd + d];
      const int32_t r = map_find(m, c);
      if (r >= 0) { ir[cnt] = r; orr[cnt] = (int32_t)o; ++cnt; }
    }
    counts[k] = cnt;
  }
}

/* out[o,:] += x[i,:] @ W[k]   for every pair of every offset (out pre-zeroed by
 * the caller).  Within one offset every out row appears at most once, so the
 * pair loop is race-free. */
void cr_conv_forward(const float *x, const float *w, int K, int c_in, int c_out,
                     const int64_t *counts, const int32_t *in_rows,
                     const int32_t *out_rows, int64_t n_out, float *out) {
  for (int k = 0; k < K; ++k) {
    const float *wk = w + (size_t)k * c_in * c_out;
    const int32_t *ir = in_rows + (int64_t)k * n_out, *orr = out_rows + (int64_t)k * n_out;
    const int64_t cnt = counts[k];
#pragma omp parallel for schedule(static)
    for (int64_t p = 0; p < cnt; ++p) {
      const float *xi = x + (size_t)ir[p] * c_in;
      float *yo = out + (size_t)orr[p] * c_out;
      for (int ci = 0; ci < c_in; ++ci) {
        const float a = xi[ci];
        const float *wr = wk + (size_t)ci * c_out;
        for (int co = 0; co < c_out; ++co) yo[co] += a * wr[co];
      }
    }
  }
}

/* grad_in[i,:] += dy[o,:] @ W[k]^T ; grad_w[k] += x[i,:]^T dy[o,:]
 * (grad_in and grad_w pre-zeroed by the caller). */
void cr_conv_backward(const float *x, const float *dy, const float *w, int K,
                      int c_in, int c_out, const int64_t *counts,
                      const int32_t *in_rows, const int32_t *out_rows,
                      int64_t n_out, float *grad_in, float *grad_w) {
  const int nt = omp_get_max_threads();
  const size_t wsz = (size_t)c_in * c_out;
  float *partial = (float *)malloc(sizeof(float) * wsz * nt);
  for (int k = 0; k < K; ++k) {
    const float *wk = w + (size_t)k * wsz;
    const int32_t *ir = in_rows + (int64_t)k * n_out, *orr = out_rows + (int64_t)k * n_out;
    const int64_t cnt = counts[k];
    memset(partial, 0, sizeof(float) * wsz * nt);
#pragma omp parallel
    {
      float *gw = partial + (size_t)omp_get_thread_num() * wsz;
#pragma omp for schedule(static)
      for (int64_t p = 0; p < cnt; ++p) {
        const float *xi = x + (size_t)ir[p] * c_in;
        const float *go = dy + (size_t)orr[p] * c_out;
        float *gi = grad_in + (size_t)ir[p] * c_in;   /* in rows are distinct within an offset */
        for (int ci = 0; ci < c_in; ++ci) {
          const float *wr = wk + (size_t)ci * c_out;
          float *gwr = gw + (size_t)ci * c_out;
          const float a = xi[ci];
          float acc = 0.f;
          for (int co = 0; co < c_out; ++co) {
            acc += go[co] * wr[co];
            gwr[co] += a * go[co];
          }
          gi[ci] += acc;
        }
      }
    }
    float *gwk = grad_w + (size_t)k * wsz;
    for (int t = 0; t < nt; ++t)
      for (size_t e = 0; e < wsz; ++e) gwk[e] += partial[(size_t)t * wsz + e];
  }
  free(partial);
}
