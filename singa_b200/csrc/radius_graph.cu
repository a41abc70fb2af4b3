// radius_graph as a cell-list neighbour search that emits a destination-sorted CSR edge list.
//
// Replaces torch_cluster.radius_graph (torch-cluster 1.6.3, un-vendored; reference imports it at gen.py:17 and the
// dataset's `linked_to` edges are radius graphs).  Semantics restated from the published algorithm:
// edge (j -> i) iff j != i (unless loop), batch[j] == batch[i] and
//     ((xi-xj)^2 + (yi-yj)^2) + (zi-zj)^2 < r*r        evaluated in fp32 with separately rounded operations
// (the brute-force oracle's expression, oracle/ef_oracle.py: radius_graph_bruteforce).  Neighbour cap as upstream
// applies it: radius_graph(loop=False) calls radius(..., max_num_neighbors + 1), whose kernel scans sources in index
// order, the self hit included, and stops at the cap; self loops are removed afterwards (see neighbour_kernel).  Output: edge_index[0] = source j, edge_index[1] = target i, grouped by target in
// ascending i and, inside a target, ascending j - a canonical order, so the edge LIST (not only the set) is
// reproducible bit for bit and is directly the CSR the attention kernels consume.
//
// Pipeline (all HBM-bound integer/byte work; no float atomics; integer atomics only where order is irrelevant):
//   1 bbox      per-batch bounding boxes, order-preserving uint encoding + atomicMin/Max        (host reads B*6 floats)
//   2 bin       cell id per point (cell edge = 1.001 r), histogram                                 \
//   3 scan      cub exclusive sum -> cell_start                                                     |  efl_radius_graph_count
//   4 scatter   points into cell order as float4 (x, y, z, original index)                          |
//   5 count     thread per point scans its 27 cells -> deg[i]; capped and full exclusive scans     /   (host reads 2 totals)
//   6 fill      same scan, writes sources into the full CSR                                        \   efl_radius_graph_fill
//   7 rowsort   warp per target: bitonic sort of its sources in shared memory                      /
//   8 emit      first min(deg, cap) sources of each row -> int64 edge_index [2, E]                     efl_radius_graph_emit
// The scatter order inside a cell depends on atomics, the final result does not (step 7 canonicalises it).
#include <cub/cub.cuh>

#include "efl_common.cuh"

namespace efl {
namespace rg {

struct Grid {
  float cs_inv;
  int nx, ny, nz;   // cells per batch item along each axis
  int nbatch;
};

__device__ __forceinline__ unsigned enc_f(float f) {
  const unsigned b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float dec_f(unsigned u) {
  const unsigned b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  float f;
  memcpy(&f, &b, 4);
  return f;
#endif
}

__global__ void bbox_init_kernel(unsigned* __restrict__ box, int nbatch) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nbatch * 6) box[t] = (t % 6 < 3) ? 0xffffffffu : 0u;
}

// box [B, 6] encoded (min xyz | max xyz).  The batch vector is sorted, so a CTA whose first and last point share a
// batch item reduces in shared memory and issues 6 global atomics instead of 6 per point.
__global__ void __launch_bounds__(256) bbox_kernel(const float* __restrict__ pos, const int* __restrict__ batch, int n,
                                                   unsigned* __restrict__ box) {
  __shared__ unsigned sb[6];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int first = blockIdx.x * blockDim.x, last = min(n, first + (int)blockDim.x) - 1;
  const int bf = batch ? batch[first] : 0, bl = batch ? batch[last] : 0;
  const bool uniform = bf == bl;
  if (threadIdx.x < 6) sb[threadIdx.x] = threadIdx.x < 3 ? 0xffffffffu : 0u;
  __syncthreads();
  if (i < n) {
    const int b = batch ? batch[i] : 0;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      unsigned e = enc_f(pos[3 * (size_t)i + d]);
      if (uniform) {
        const unsigned lo = __reduce_min_sync(__activemask(), e), hi = __reduce_max_sync(__activemask(), e);
        if ((threadIdx.x & 31) == 0 || i == first) {
          atomicMin(sb + d, lo);
          atomicMax(sb + 3 + d, hi);
        }
      } else {
        atomicMin(box + 6 * b + d, e);
        atomicMax(box + 6 * b + 3 + d, e);
      }
    }
  }
  __syncthreads();
  if (uniform && threadIdx.x < 6) {
    if (threadIdx.x < 3) atomicMin(box + 6 * bf + threadIdx.x, sb[threadIdx.x]);
    else atomicMax(box + 6 * bf + threadIdx.x, sb[threadIdx.x]);
  }
}

__global__ void bbox_decode_kernel(const unsigned* __restrict__ box, float* __restrict__ out, int n) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) out[t] = dec_f(box[t]);
}

__device__ __forceinline__ int cell_coord(float p, float origin, float cs_inv, int n) {
  int c = (int)floorf((p - origin) * cs_inv);
  return c < 0 ? 0 : (c >= n ? n - 1 : c);
}

__global__ void bin_kernel(const float* __restrict__ pos, const int* __restrict__ batch, int n,
                           const float* __restrict__ origin /* [B,3] */, Grid g, int* __restrict__ cell_of,
                           int* __restrict__ cell_count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int b = batch ? batch[i] : 0;
  const int cx = cell_coord(pos[3 * (size_t)i + 0], origin[3 * b + 0], g.cs_inv, g.nx);
  const int cy = cell_coord(pos[3 * (size_t)i + 1], origin[3 * b + 1], g.cs_inv, g.ny);
  const int cz = cell_coord(pos[3 * (size_t)i + 2], origin[3 * b + 2], g.cs_inv, g.nz);
  const int c = ((b * g.nz + cz) * g.ny + cy) * g.nx + cx;
  cell_of[i] = c;
  atomicAdd(cell_count + c, 1);
}

__global__ void scatter_kernel(const float* __restrict__ pos, int n, const int* __restrict__ cell_of,
                               const int* __restrict__ cell_start, int* __restrict__ cursor,
                               float4* __restrict__ sorted) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = cell_of[i];
  const int slot = cell_start[c] + atomicAdd(cursor + c, 1);
  sorted[slot] = make_float4(pos[3 * (size_t)i], pos[3 * (size_t)i + 1], pos[3 * (size_t)i + 2], __int_as_float(i));
}

// Visits every point within r of sorted slot s (27-cell stencil).  FILL = false: count; true: write sources.
template <bool FILL>
__global__ void __launch_bounds__(128) neighbour_kernel(const float4* __restrict__ sorted, int n,
                                                        const int* __restrict__ cell_of,
                                                        const int* __restrict__ cell_start, Grid g, float r2, int loop,
                                                        int cap, int* __restrict__ deg, int* __restrict__ deg_cap,
                                                        const int* __restrict__ rowptr, int* __restrict__ col) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const float4 me = sorted[s];
  const int i = __float_as_int(me.w);
  const int c = cell_of[i];
  const int cx = c % g.nx, cy = (c / g.nx) % g.ny, cz = (c / (g.nx * g.ny)) % g.nz, b = c / (g.nx * g.ny * g.nz);
  int cnt = 0, cnt_lo = 0;  // hits; hits whose source index is below the target's own index
  int* out = FILL ? col + rowptr[i] : nullptr;
  for (int dz = -1; dz <= 1; ++dz) {
    const int z = cz + dz;
    if (z < 0 || z >= g.nz) continue;
    for (int dy = -1; dy <= 1; ++dy) {
      const int y = cy + dy;
      if (y < 0 || y >= g.ny) continue;
      // the three x-neighbours are contiguous in cell order: one candidate range per (z, y)
      const int x0 = cx > 0 ? cx - 1 : 0, x1 = cx + 1 < g.nx ? cx + 1 : g.nx - 1;
      const int base = ((b * g.nz + z) * g.ny + y) * g.nx;
      const int q0 = cell_start[base + x0], q1 = cell_start[base + x1 + 1];
      for (int q = q0; q < q1; ++q) {
        const float4 o = __ldg(sorted + q);
        const float ddx = __fsub_rn(me.x, o.x), ddy = __fsub_rn(me.y, o.y), ddz = __fsub_rn(me.z, o.z);
        const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(ddx, ddx), __fmul_rn(ddy, ddy)), __fmul_rn(ddz, ddz));
        if (d2 < r2 && (loop || q != s)) {
          if (FILL) out[cnt] = __float_as_int(o.w);
          else cnt_lo += __float_as_int(o.w) < i;
          ++cnt;
        }
      }
    }
  }
  if (!FILL) {
    deg[i] = cnt;
    // Neighbour cap as torch_cluster.radius_graph applies it (torch-cluster 1.6.x radius_graph -> radius): with
    // loop=False it searches with max_num_neighbors + 1, the self hit included, in ascending source index, and
    // removes self loops afterwards.  A target whose own index comes after its first cap+1 hits keeps cap+1
    // sources; otherwise the self hit used one slot and min(deg, cap) remain.  Either way the kept sources are the
    // smallest indices of the (sorted) row, so emit takes a prefix.
    deg_cap[i] = loop ? min(cnt, cap) : (cnt_lo >= cap + 1 ? cap + 1 : min(cnt, cap));
  }
}

// ascending sort of every CSR row; one warp per row, bitonic network in shared memory (rows <= CAP elements)
template <int CAP>
__global__ void __launch_bounds__(256) rowsort_kernel(const int* __restrict__ rowptr, int n, int* __restrict__ col) {
  __shared__ int buf[8][CAP];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= n) return;
  const int r0 = rowptr[row], len = rowptr[row + 1] - r0;
  if (len <= 1) return;
  if (len > CAP) return;  // sorted by rowsort_long_kernel
  int P = 32;
  while (P < len) P <<= 1;
  int* b = buf[warp];
  for (int t = lane; t < P; t += 32) b[t] = t < len ? col[r0 + t] : 0x7fffffff;
  __syncwarp();
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = lane; t < (P >> 1); t += 32) {
        const int lo = ((t & ~(j - 1)) << 1) | (t & (j - 1));  // index with bit j cleared
        const int hi = lo | j;
        const bool up = (lo & k) == 0;
        const int a = b[lo], c = b[hi];
        if ((a > c) == up) { b[lo] = c; b[hi] = a; }
      }
      __syncwarp();
    }
  }
  for (int t = lane; t < len; t += 32) col[r0 + t] = b[t];
}

// long rows: one CTA per row, bitonic sort in dynamic shared memory (P ints)
__global__ void __launch_bounds__(256) rowsort_long_kernel(const int* __restrict__ rowptr, int n, int* __restrict__ col,
                                                           int min_len, int P) {
  extern __shared__ int lb[];
  for (int row = blockIdx.x; row < n; row += gridDim.x) {
    const int r0 = rowptr[row], len = rowptr[row + 1] - r0;
    if (len <= min_len) continue;  // uniform per CTA
    for (int t = threadIdx.x; t < P; t += blockDim.x) lb[t] = t < len ? col[r0 + t] : 0x7fffffff;
    __syncthreads();
    for (int k = 2; k <= P; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int t = threadIdx.x; t < (P >> 1); t += blockDim.x) {
          const int lo = ((t & ~(j - 1)) << 1) | (t & (j - 1));
          const int hi = lo | j;
          const bool up = (lo & k) == 0;
          const int a = lb[lo], c = lb[hi];
          if ((a > c) == up) { lb[lo] = c; lb[hi] = a; }
        }
        __syncthreads();
      }
    }
    for (int t = threadIdx.x; t < len; t += blockDim.x) col[r0 + t] = lb[t];
    __syncthreads();
  }
}

// edge_index [2, E] int64: row 0 = sources, row 1 = targets; warp per target
__global__ void emit_kernel(const int* __restrict__ rowptr_full, const int* __restrict__ rowptr_cap,
                            const int* __restrict__ col, int n, long long E, long long* __restrict__ edge_index,
                            int flow_target_to_source) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= n) return;
  const int s0 = rowptr_full[row], d0 = rowptr_cap[row], len = rowptr_cap[row + 1] - d0;
  long long* src = edge_index + (flow_target_to_source ? E : 0);
  long long* dst = edge_index + (flow_target_to_source ? 0 : E);
  for (int t = lane; t < len; t += 32) {
    src[d0 + t] = col[s0 + t];
    dst[d0 + t] = row;
  }
}

// ---------------------------------------------------------------------------------------------------
// knn_graph on the same cell list (reference call sites model/CProMG.py:293, 330:
// knn_graph(pos, k, batch=batch, flow='target_to_source'), torch-cluster 1.6.3, un-vendored).
// One warp per query point.  The warp keeps its K best (distance, index) pairs as sorted 64-bit keys
// (fp32 distance bits << 32 | source index: non-negative floats order like their bit patterns, ties go to the
// smaller index) in shared memory, followed by a staging area for new candidates; when the staging area fills, the
// whole buffer is bitonic-sorted and cut back to K.  Cells are visited in shells of growing Chebyshev radius around
// the query's cell; after shell r every point closer than r * cs is known, so the search stops as soon as the K-th
// best distance is inside that radius (or the shell has left the batch item's grid).  Distances use the radius
// kernel's separately rounded fp32 expression.  Output nbr [N, K] (ascending distance, -1 padded) and cnt [N].
// ---------------------------------------------------------------------------------------------------
constexpr int KNN_BUF = 256;           // keys per warp: K best + staging (K <= 64)
constexpr int KNN_WARPS = 4;

__device__ __forceinline__ void knn_sort(unsigned long long* b, int lane) {
  // bitonic sort of KNN_BUF keys by one warp
  for (int k = 2; k <= KNN_BUF; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = lane; t < (KNN_BUF >> 1); t += 32) {
        const int lo = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const int hi = lo | j;
        const bool up = (lo & k) == 0;
        const unsigned long long a = b[lo], c = b[hi];
        if ((a > c) == up) { b[lo] = c; b[hi] = a; }
      }
      __syncwarp();
    }
  }
}

__global__ void __launch_bounds__(32 * KNN_WARPS) knn_kernel(const float4* __restrict__ sorted, int n,
                                                             const int* __restrict__ cell_of,
                                                             const int* __restrict__ cell_start, Grid g, float cs,
                                                             int K, int loop, int* __restrict__ nbr,
                                                             int* __restrict__ cnt) {
  __shared__ unsigned long long buf[KNN_WARPS][KNN_BUF];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int s = blockIdx.x * KNN_WARPS + warp;
  if (s >= n) return;
  unsigned long long* b = buf[warp];
  const unsigned long long EMPTY = ~0ull;
  for (int t = lane; t < KNN_BUF; t += 32) b[t] = EMPTY;
  __syncwarp();
  const float4 me = sorted[s];
  const int i = __float_as_int(me.w);
  const int c = cell_of[i];
  const int cx = c % g.nx, cy = (c / g.nx) % g.ny, cz = (c / (g.nx * g.ny)) % g.nz, bi = c / (g.nx * g.ny * g.nz);
  int fill = K;                       // next free staging slot (warp-uniform)
  unsigned long long kth = EMPTY;     // current K-th best key (warp-uniform): worse candidates are dropped at once
  const int rmax = max(max(max(cx, g.nx - 1 - cx), max(cy, g.ny - 1 - cy)), max(cz, g.nz - 1 - cz));
  for (int r = 0; r <= rmax; ++r) {
    for (int dz = -r; dz <= r; ++dz) {
      const int z = cz + dz;
      if (z < 0 || z >= g.nz) continue;
      for (int dy = -r; dy <= r; ++dy) {
        const int y = cy + dy;
        if (y < 0 || y >= g.ny) continue;
        const bool face = (dz == -r || dz == r || dy == -r || dy == r);
        // rows on the shell's faces take the whole x range; interior rows only the two end cells
        for (int part = 0; part < (face || r == 0 ? 1 : 2); ++part) {
          int x0, x1;
          if (face || r == 0) { x0 = max(cx - r, 0); x1 = min(cx + r, g.nx - 1); }
          else { x0 = x1 = part == 0 ? cx - r : cx + r; if (x0 < 0 || x0 >= g.nx) continue; }
          const int base = ((bi * g.nz + z) * g.ny + y) * g.nx;
          const int q0 = cell_start[base + x0], q1 = cell_start[base + x1 + 1];
          for (int qb = q0; qb < q1; qb += 32) {
            const int q = qb + lane;
            unsigned long long key = EMPTY;
            if (q < q1) {
              const float4 o = __ldg(sorted + q);
              const float ddx = __fsub_rn(me.x, o.x), ddy = __fsub_rn(me.y, o.y), ddz = __fsub_rn(me.z, o.z);
              const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(ddx, ddx), __fmul_rn(ddy, ddy)), __fmul_rn(ddz, ddz));
              if (loop || q != s)
                key = ((unsigned long long)__float_as_uint(d2) << 32) | (unsigned)__float_as_int(o.w);
            }
            const bool take = key < kth;
            const unsigned m = __ballot_sync(0xffffffffu, take);
            const int nnew = __popc(m);
            if (nnew == 0) continue;
            if (fill + nnew > KNN_BUF) {      // staging full: sort, keep the K best, refresh the threshold
              knn_sort(b, lane);
              for (int t = K + lane; t < KNN_BUF; t += 32) b[t] = EMPTY;
              __syncwarp();
              fill = K;
              kth = b[K - 1];
            }
            if (take && key < kth) b[fill + __popc(m & ((1u << lane) - 1))] = key;
            fill += nnew;                      // slots of candidates rejected by the refreshed threshold stay EMPTY
            __syncwarp();
          }
        }
      }
    }
    // shell r done: every point closer than r * cs has been seen
    knn_sort(b, lane);
    for (int t = K + lane; t < KNN_BUF; t += 32) b[t] = EMPTY;
    __syncwarp();
    fill = K;
    kth = b[K - 1];
    if (kth != EMPTY) {
      const float covered = (float)r * cs * 0.999f;   // margin for the rounding of the cell assignment
      const float d2k = __uint_as_float((unsigned)(kth >> 32));
      if (d2k < covered * covered) break;
    }
  }
  int have = 0;
  for (int t = lane; t < K; t += 32) {
    const unsigned long long key = b[t];
    nbr[(size_t)i * K + t] = key == EMPTY ? -1 : (int)(unsigned)(key & 0xffffffffu);
    have += key != EMPTY;
  }
  have = __reduce_add_sync(0xffffffffu, have);
  if (lane == 0) cnt[i] = have;
}

__global__ void knn_rowptr_full_kernel(int n, int K, int* __restrict__ rowptr_full) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i <= n) rowptr_full[i] = i * K;
}

}  // namespace rg
}  // namespace efl

using namespace efl;
using efl::rg::Grid;

static inline unsigned nblk(long long n, int t) { return (unsigned)((n + t - 1) / t); }

// box_ws: B*6 uints (scratch); box_out: B*6 floats (min xyz | max xyz per batch item), read by the host to size the grid
extern "C" int efl_radius_graph_bbox(const float* pos, const int* batch, int N, int nbatch, void* box_ws,
                                     float* box_out, void* stream) {
  EFL_REQUIRE(N > 0 && nbatch > 0, "efl_radius_graph_bbox: empty input");
  cudaStream_t s = (cudaStream_t)stream;
  rg::bbox_init_kernel<<<nblk(nbatch * 6, 128), 128, 0, s>>>((unsigned*)box_ws, nbatch);
  rg::bbox_kernel<<<nblk(N, 256), 256, 0, s>>>(pos, batch, N, (unsigned*)box_ws);
  rg::bbox_decode_kernel<<<nblk(nbatch * 6, 128), 128, 0, s>>>((const unsigned*)box_ws, box_out, nbatch * 6);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

extern "C" long long efl_radius_graph_temp_bytes(long long max_items) {
  size_t bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, bytes, (const int*)nullptr, (int*)nullptr, (int)max_items);
  return (long long)bytes;
}

// Steps 2-5.  origin [B,3] device floats; cell edge cs >= r; ncells = nbatch*nx*ny*nz.
// Buffers (device): cell_of [N], cell_start [ncells+1], scratch [2*(ncells+1)] ints, sorted [N] float4,
// deg [2*(N+1)] ints (full | capped), rowptr_full [N+1], rowptr_cap [N+1], temp (efl_radius_graph_temp_bytes(max(N,ncells)+1)).
extern "C" int efl_radius_graph_count(const float* pos, const int* batch, int N, int nbatch, const float* origin,
                                      float cs, int nx, int ny, int nz, float r, int loop, int cap, int* cell_of,
                                      int* cell_start, int* scratch, void* sorted, int* deg, int* rowptr_full,
                                      int* rowptr_cap, void* temp, long long temp_bytes, void* stream) {
  EFL_REQUIRE(N > 0 && nbatch > 0 && nx > 0 && ny > 0 && nz > 0, "efl_radius_graph_count: empty input");
  EFL_REQUIRE(cs >= r && r > 0.f, "efl_radius_graph_count: cell edge must be >= r > 0");
  const long long ncells_ll = (long long)nbatch * nx * ny * nz;
  EFL_REQUIRE(ncells_ll < (1ll << 30), "efl_radius_graph_count: too many cells");
  const int ncells = (int)ncells_ll;
  cudaStream_t s = (cudaStream_t)stream;
  Grid g{1.0f / cs, nx, ny, nz, nbatch};
  int* cell_count = scratch;
  int* cursor = scratch + (ncells + 1);
  cudaMemsetAsync(scratch, 0, sizeof(int) * 2 * (size_t)(ncells + 1), s);
  rg::bin_kernel<<<nblk(N, 256), 256, 0, s>>>(pos, batch, N, origin, g, cell_of, cell_count);
  size_t tb = (size_t)temp_bytes;
  cub::DeviceScan::ExclusiveSum(temp, tb, cell_count, cell_start, ncells + 1, s);
  rg::scatter_kernel<<<nblk(N, 256), 256, 0, s>>>(pos, N, cell_of, cell_start, cursor, (float4*)sorted);
  int* deg_full = deg;
  int* deg_cap = deg + (N + 1);
  cudaMemsetAsync(deg, 0, sizeof(int) * 2 * (size_t)(N + 1), s);
  rg::neighbour_kernel<false><<<nblk(N, 128), 128, 0, s>>>((const float4*)sorted, N, cell_of, cell_start, g,
                                                           r * r, loop, cap, deg_full, deg_cap, nullptr, nullptr);
  tb = (size_t)temp_bytes;
  cub::DeviceScan::ExclusiveSum(temp, tb, deg_full, rowptr_full, N + 1, s);
  tb = (size_t)temp_bytes;
  cub::DeviceScan::ExclusiveSum(temp, tb, deg_cap, rowptr_cap, N + 1, s);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// Steps 6-7: col [rowptr_full[N]] ints; max_deg = largest full degree (the host read it with the totals).
extern "C" int efl_radius_graph_fill(int N, int nbatch, float cs, int nx, int ny, int nz, float r, int loop,
                                     const int* cell_of, const int* cell_start, const void* sorted,
                                     const int* rowptr_full, int* col, int max_deg, void* stream) {
  EFL_REQUIRE(N > 0, "efl_radius_graph_fill: empty input");
  cudaStream_t s = (cudaStream_t)stream;
  Grid g{1.0f / cs, nx, ny, nz, nbatch};
  rg::neighbour_kernel<true><<<nblk(N, 128), 128, 0, s>>>((const float4*)sorted, N, cell_of, cell_start, g,
                                                          r * r, loop, 0, nullptr, nullptr, rowptr_full, col);
  rg::rowsort_kernel<1024><<<nblk(N, 8), 256, 0, s>>>(rowptr_full, N, col);
  if (max_deg > 1024) {
    int P = 2048;
    while (P < max_deg) P <<= 1;
    EFL_REQUIRE(P <= 32768, "efl_radius_graph_fill: more than 32768 neighbours of one point");
    {  // per call: the attribute is per device and a process may touch several (cheap next to the sort itself)
      cudaError_t e = cudaFuncSetAttribute(rg::rowsort_long_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 * 4);
      if (e != cudaSuccess) {
        efl_set_error(cudaGetErrorString(e));
        return (int)e;
      }
    }
    rg::rowsort_long_kernel<<<N < 1184 ? N : 1184, 256, (size_t)P * 4, s>>>(rowptr_full, N, col, 1024, P);
  }
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// Step 8: edge_index [2, E] int64, E = rowptr_cap[N]; flow 0 = source_to_target (row 0 sources), 1 = target_to_source
extern "C" int efl_radius_graph_emit(const int* rowptr_full, const int* rowptr_cap, const int* col, int N, long long E,
                                     long long* edge_index, int flow_target_to_source, void* stream) {
  if (N <= 0 || E <= 0) return EFL_OK;
  rg::emit_kernel<<<nblk((long long)N * 32, 256), 256, 0, (cudaStream_t)stream>>>(rowptr_full, rowptr_cap, col, N, E,
                                                                               edge_index, flow_target_to_source);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// knn_graph, steps after the bounding boxes (efl_radius_graph_bbox): bin + scan + scatter into the cell list, then
// the K nearest neighbours of every point inside its batch item.  Buffers as efl_radius_graph_count plus
// nbr [N*K] ints, cnt [N+1] ints (cnt[N] is set to 0), rowptr_full [N+1] (= i*K), rowptr [N+1] (exclusive scan of cnt).
// efl_radius_graph_emit(rowptr_full, rowptr, nbr, N, E, edge_index, flow) then writes the int64 edge list.
extern "C" int efl_knn_graph(const float* pos, const int* batch, int N, int nbatch, const float* origin, float cs,
                             int nx, int ny, int nz, int K, int loop, int* cell_of, int* cell_start, int* scratch,
                             void* sorted, int* nbr, int* cnt, int* rowptr_full, int* rowptr, void* temp,
                             long long temp_bytes, void* stream) {
  EFL_REQUIRE(N > 0 && nbatch > 0 && nx > 0 && ny > 0 && nz > 0, "efl_knn_graph: empty input");
  EFL_REQUIRE(K >= 1 && K <= 64, "efl_knn_graph: k must be in [1, 64]");
  EFL_REQUIRE(cs > 0.f, "efl_knn_graph: cell edge must be positive");
  const long long ncells_ll = (long long)nbatch * nx * ny * nz;
  EFL_REQUIRE(ncells_ll < (1ll << 30), "efl_knn_graph: too many cells");
  EFL_REQUIRE((long long)N * K < (1ll << 31), "efl_knn_graph: N * k overflows int");
  const int ncells = (int)ncells_ll;
  cudaStream_t s = (cudaStream_t)stream;
  Grid g{1.0f / cs, nx, ny, nz, nbatch};
  int* cell_count = scratch;
  int* cursor = scratch + (ncells + 1);
  cudaMemsetAsync(scratch, 0, sizeof(int) * 2 * (size_t)(ncells + 1), s);
  rg::bin_kernel<<<nblk(N, 256), 256, 0, s>>>(pos, batch, N, origin, g, cell_of, cell_count);
  size_t tb = (size_t)temp_bytes;
  cub::DeviceScan::ExclusiveSum(temp, tb, cell_count, cell_start, ncells + 1, s);
  rg::scatter_kernel<<<nblk(N, 256), 256, 0, s>>>(pos, N, cell_of, cell_start, cursor, (float4*)sorted);
  cudaMemsetAsync(cnt + N, 0, sizeof(int), s);
  rg::knn_kernel<<<nblk(N, rg::KNN_WARPS), 32 * rg::KNN_WARPS, 0, s>>>((const float4*)sorted, N, cell_of, cell_start, g,
                                                                     cs, K, loop, nbr, cnt);
  rg::knn_rowptr_full_kernel<<<nblk(N + 1, 256), 256, 0, s>>>(N, K, rowptr_full);
  tb = (size_t)temp_bytes;
  cub::DeviceScan::ExclusiveSum(temp, tb, cnt, rowptr, N + 1, s);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
