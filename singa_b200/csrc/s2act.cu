// Separable S2 activation on per-edge (or per-node) coefficient blocks, fp32 SIMT version.
//
// Replaces SeparableS2Activation / S2Activation (reference EF_layers.py:1736-1773):
//   grid = to_grid[G, M] . x[M, CH];  s = SiLU(grid);  y = from_grid[M, G]^T-contract . s;
//   y[row 0] = SiLU(gate).
// The [*, G, CH] grid tensor (32 GB per layer at config 2 in the reference) lives only in
// registers.  One thread per channel; to/from matrices in shared memory, read as float4
// broadcasts.  The coefficient rows are addressed through (pointer, stride) descriptors so the
// same kernel serves the m-primary split edge buffers and the l-primary node tensors.
#include "efl_common.cuh"

namespace efl {

struct RowMap {
  // element (row j, item e, channel c) = base[seg(j)] + e * item_stride[seg] + (j - first[seg]) * CH + c
  // the M rows are covered by up to 5 contiguous segments (m=0 | +1 | -1 | +2 | -2 blocks)
  void* base[5];
  long long item_stride[5];
  int first[6];
  int nseg;
};

// segment of row j (uniform across the block); register indices stay compile-time
__device__ __forceinline__ int seg_of(const RowMap& rm, int j) {
  int s = 0;
#pragma unroll
  for (int t = 1; t < 5; ++t) s += (t < rm.nseg && j >= rm.first[t]) ? 1 : 0;
  return s;
}

template <int M, typename T>
__device__ __forceinline__ void load_rows(const RowMap& rm, long long e, int CH, int c, float* h) {
#pragma unroll
  for (int j = 0; j < M; ++j) {
    const int s = seg_of(rm, j);
    const T* p = reinterpret_cast<const T*>(rm.base[s]) + e * rm.item_stride[s] + (long long)(j - rm.first[s]) * CH + c;
    h[j] = to_f<T>(*p);
  }
}

template <int M, typename T>
__device__ __forceinline__ void store_rows(const RowMap& rm, long long e, int CH, int c, const float* y) {
#pragma unroll
  for (int j = 0; j < M; ++j) {
    const int s = seg_of(rm, j);
    T* p = reinterpret_cast<T*>(rm.base[s]) + e * rm.item_stride[s] + (long long)(j - rm.first[s]) * CH + c;
    *p = from_f<T>(y[j]);
  }
}

// MP = M padded to a multiple of 4 (float4 broadcasts)
template <int M, int MP, typename TIn, typename TOut>
__global__ void __launch_bounds__(128) s2act_fwd_kernel(RowMap in, RowMap out, const float* __restrict__ gate,
                                                        long long gate_ld, const float* __restrict__ to_grid,
                                                        const float* __restrict__ from_grid, int G, int CH,
                                                        long long n_items) {
  extern __shared__ float4 smem4[];
  float* sT = reinterpret_cast<float*>(smem4);  // [G][MP]
  float* sF = sT + G * MP;                      // [G][MP]
  for (int i = threadIdx.x; i < G * MP; i += blockDim.x) {
    const int g = i / MP, j = i % MP;
    sT[i] = j < M ? to_grid[g * M + j] : 0.f;
    sF[i] = j < M ? from_grid[g * M + j] : 0.f;
  }
  __syncthreads();
  const int cpb = blockDim.x;  // channels handled per block pass
  const int chunks = (CH + cpb - 1) / cpb;
  for (long long w = blockIdx.x; w < n_items * chunks; w += gridDim.x) {
    const long long e = w / chunks;
    const int c = (int)(w % chunks) * cpb + threadIdx.x;
    if (c >= CH) continue;
    float h[MP], y[MP];
#pragma unroll
    for (int j = 0; j < MP; ++j) { h[j] = 0.f; y[j] = 0.f; }
    load_rows<M, TIn>(in, e, CH, c, h);
    for (int g = 0; g < G; ++g) {
      const float4* t4 = reinterpret_cast<const float4*>(sT + g * MP);
      const float4* f4 = reinterpret_cast<const float4*>(sF + g * MP);
      float acc = 0.f;
#pragma unroll
      for (int q = 0; q < MP / 4; ++q) {
        const float4 t = t4[q];
        acc = fmaf(t.x, h[4 * q], acc);
        acc = fmaf(t.y, h[4 * q + 1], acc);
        acc = fmaf(t.z, h[4 * q + 2], acc);
        acc = fmaf(t.w, h[4 * q + 3], acc);
      }
      const float s = siluf_(acc);
#pragma unroll
      for (int q = 0; q < MP / 4; ++q) {
        const float4 f = f4[q];
        y[4 * q] = fmaf(f.x, s, y[4 * q]);
        y[4 * q + 1] = fmaf(f.y, s, y[4 * q + 1]);
        y[4 * q + 2] = fmaf(f.z, s, y[4 * q + 2]);
        y[4 * q + 3] = fmaf(f.w, s, y[4 * q + 3]);
      }
    }
    y[0] = siluf_(gate[e * gate_ld + c]);
    store_rows<M, TOut>(out, e, CH, c, y);
  }
}

// backward.  dy rows arrive through `din`; results:
//   dgate[e, c] = dy[0] * SiLU'(gate)
//   dx rows written through `dout`.  When dup != 0 the m>0 segments use the duplicated,
//   column-interleaved "dY" layout of the SO(2) GEMM backward (row 2e: (d re_q, d im_q),
//   row 2e+1: (d im_q, -d re_q)); segment pairs (1,2), (3,4) are the (+m, -m) blocks.
template <int M, int MP, typename TIn, typename TD, typename TOut>
__global__ void __launch_bounds__(128) s2act_bwd_kernel(RowMap in, RowMap din, RowMap dout, int dup,
                                                        const float* __restrict__ gate, long long gate_ld,
                                                        TOut* __restrict__ dgate, long long dgate_ld,
                                                        const float* __restrict__ to_grid,
                                                        const float* __restrict__ from_grid, int G, int CH,
                                                        long long n_items) {
  extern __shared__ float4 smem4[];
  float* sT = reinterpret_cast<float*>(smem4);
  float* sF = sT + G * MP;
  for (int i = threadIdx.x; i < G * MP; i += blockDim.x) {
    const int g = i / MP, j = i % MP;
    sT[i] = j < M ? to_grid[g * M + j] : 0.f;
    sF[i] = (j < M && j > 0) ? from_grid[g * M + j] : 0.f;  // row 0 of the S2 output is discarded
  }
  __syncthreads();
  const int cpb = blockDim.x;
  const int chunks = (CH + cpb - 1) / cpb;
  for (long long w = blockIdx.x; w < n_items * chunks; w += gridDim.x) {
    const long long e = w / chunks;
    const int c = (int)(w % chunks) * cpb + threadIdx.x;
    if (c >= CH) continue;
    float h[MP], dy[MP], dx[MP];
#pragma unroll
    for (int j = 0; j < MP; ++j) { h[j] = 0.f; dy[j] = 0.f; dx[j] = 0.f; }
    load_rows<M, TIn>(in, e, CH, c, h);
    load_rows<M, TD>(din, e, CH, c, dy);
    for (int g = 0; g < G; ++g) {
      const float4* t4 = reinterpret_cast<const float4*>(sT + g * MP);
      const float4* f4 = reinterpret_cast<const float4*>(sF + g * MP);
      float acc = 0.f, ds = 0.f;
#pragma unroll
      for (int q = 0; q < MP / 4; ++q) {
        const float4 t = t4[q];
        const float4 f = f4[q];
        acc = fmaf(t.x, h[4 * q], acc);
        acc = fmaf(t.y, h[4 * q + 1], acc);
        acc = fmaf(t.z, h[4 * q + 2], acc);
        acc = fmaf(t.w, h[4 * q + 3], acc);
        ds = fmaf(f.x, dy[4 * q], ds);
        ds = fmaf(f.y, dy[4 * q + 1], ds);
        ds = fmaf(f.z, dy[4 * q + 2], ds);
        ds = fmaf(f.w, dy[4 * q + 3], ds);
      }
      const float dg = ds * dsiluf_(acc);
#pragma unroll
      for (int q = 0; q < MP / 4; ++q) {
        const float4 t = t4[q];
        dx[4 * q] = fmaf(t.x, dg, dx[4 * q]);
        dx[4 * q + 1] = fmaf(t.y, dg, dx[4 * q + 1]);
        dx[4 * q + 2] = fmaf(t.z, dg, dx[4 * q + 2]);
        dx[4 * q + 3] = fmaf(t.w, dg, dx[4 * q + 3]);
      }
    }
    dgate[e * dgate_ld + c] = from_f<TOut>(dy[0] * dsiluf_(gate[e * gate_ld + c]));
    if (!dup) {
      store_rows<M, TOut>(dout, e, CH, c, dx);
    } else {
      // m = 0 segment plain; segment pairs (1,2), (3,4) duplicated + interleaved
#pragma unroll
      for (int j = 0; j < M; ++j) {
        const int s = seg_of(dout, j);
        if (s == 0) {
          TOut* p = reinterpret_cast<TOut*>(dout.base[0]) + e * dout.item_stride[0] + (long long)j * CH + c;
          *p = from_f<TOut>(dx[j]);
        } else if (s & 1) {
          // j is a "+m" (real) row; its partner "-m" row sits nr rows later
          const int nr = dout.first[s + 1] - dout.first[s];
          const long long half = (long long)nr * CH;
          const float gr = dx[j];
          float gi = 0.f;
#pragma unroll
          for (int jj = 0; jj < M; ++jj) gi = (jj == j + nr) ? dx[jj] : gi;
          TOut* r0 = reinterpret_cast<TOut*>(dout.base[s]) + e * dout.item_stride[s];
          TOut* r1 = r0 + 2 * half;
          const long long q = (long long)(j - dout.first[s]) * CH + c;
          r0[2 * q] = from_f<TOut>(gr);
          r0[2 * q + 1] = from_f<TOut>(gi);
          r1[2 * q] = from_f<TOut>(gi);
          r1[2 * q + 1] = from_f<TOut>(-gr);
        }
      }
    }
  }
}

}  // namespace efl

using namespace efl;

static RowMap make_map(int nseg, void* const* base, const long long* strides, const int* first) {
  RowMap r;
  r.nseg = nseg;
  for (int i = 0; i < 5; ++i) {
    r.base[i] = i < nseg ? base[i] : nullptr;
    r.item_stride[i] = i < nseg ? strides[i] : 0;
  }
  for (int i = 0; i < 6; ++i) r.first[i] = i <= nseg ? first[i] : first[nseg];
  return r;
}

static int pick_grid(long long work) {
  long long g = 148LL * 8;
  return (int)(work < g ? (work > 0 ? work : 1) : g);
}

// Row maps are passed as (nseg, base pointers[nseg], item strides[nseg] in elements, first[nseg+1]).
extern "C" int efl_s2act_fwd(int M, int G, int CH, int bf16_in, int bf16_out, int in_nseg, void* const* in_base,
                             const long long* in_stride, const int* in_first, int out_nseg, void* const* out_base,
                             const long long* out_stride, const int* out_first, const float* gate,
                             long long gate_ld, const float* to_grid, const float* from_grid, long long n_items,
                             void* stream) {
  if (n_items <= 0) return EFL_OK;
  EFL_REQUIRE(in_nseg <= 5 && out_nseg <= 5, "efl_s2act_fwd: at most 5 row segments");
  RowMap in = make_map(in_nseg, in_base, in_stride, in_first);
  RowMap out = make_map(out_nseg, out_base, out_stride, out_first);
  cudaStream_t s = (cudaStream_t)stream;
  const int threads = CH >= 128 ? 128 : ((CH + 31) / 32) * 32;
  const int chunks = (CH + threads - 1) / threads;
  const int grid = pick_grid(n_items * chunks);
#define LAUNCH(MM, MP, TI, TO)                                                                        \
  {                                                                                                   \
    const size_t smem = 2 * (size_t)G * MP * sizeof(float);                                           \
    cudaFuncSetAttribute(s2act_fwd_kernel<MM, MP, TI, TO>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    s2act_fwd_kernel<MM, MP, TI, TO><<<grid, threads, smem, s>>>(in, out, gate, gate_ld, to_grid,     \
                                                                 from_grid, G, CH, n_items);          \
  }
#define BY_TYPE(MM, MP)                                                                               \
  if (bf16_in && bf16_out) LAUNCH(MM, MP, __nv_bfloat16, __nv_bfloat16)                               \
  else if (!bf16_in && !bf16_out) LAUNCH(MM, MP, float, float)                                        \
  else { efl_set_error("efl_s2act_fwd: mixed in/out dtypes unsupported"); return EFL_ERR_UNSUPPORTED; }
  if (M == 29) BY_TYPE(29, 32)
  else if (M == 49) BY_TYPE(49, 52)
  else if (M == 39) BY_TYPE(39, 40)
  else if (M == 81) BY_TYPE(81, 84)
  else { efl_set_error("efl_s2act_fwd: unsupported coefficient count"); return EFL_ERR_UNSUPPORTED; }
#undef BY_TYPE
#undef LAUNCH
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

extern "C" int efl_s2act_bwd(int M, int G, int CH, int bf16_in, int bf16_d, int in_nseg, void* const* in_base,
                             const long long* in_stride, const int* in_first, int din_nseg, void* const* din_base,
                             const long long* din_stride, const int* din_first, int dout_nseg,
                             void* const* dout_base, const long long* dout_stride, const int* dout_first, int dup,
                             const float* gate, long long gate_ld, void* dgate, long long dgate_ld,
                             const float* to_grid, const float* from_grid, long long n_items, void* stream) {
  if (n_items <= 0) return EFL_OK;
  EFL_REQUIRE(in_nseg <= 5 && din_nseg <= 5 && dout_nseg <= 5, "efl_s2act_bwd: at most 5 row segments");
  RowMap in = make_map(in_nseg, in_base, in_stride, in_first);
  RowMap din = make_map(din_nseg, din_base, din_stride, din_first);
  RowMap dout = make_map(dout_nseg, dout_base, dout_stride, dout_first);
  cudaStream_t s = (cudaStream_t)stream;
  const int threads = CH >= 128 ? 128 : ((CH + 31) / 32) * 32;
  const int chunks = (CH + threads - 1) / threads;
  const int grid = pick_grid(n_items * chunks);
#define LAUNCH(MM, MP, TI, TD)                                                                        \
  {                                                                                                   \
    const size_t smem = 2 * (size_t)G * MP * sizeof(float);                                           \
    cudaFuncSetAttribute(s2act_bwd_kernel<MM, MP, TI, TD, TD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    s2act_bwd_kernel<MM, MP, TI, TD, TD><<<grid, threads, smem, s>>>(in, din, dout, dup, gate, gate_ld, (TD*)dgate, \
                                                                     dgate_ld, to_grid, from_grid, G, CH, n_items); \
  }
#define BY_TYPE(MM, MP)                                                                               \
  if (bf16_in && bf16_d) LAUNCH(MM, MP, __nv_bfloat16, __nv_bfloat16)                                 \
  else if (!bf16_in && !bf16_d) LAUNCH(MM, MP, float, float)                                          \
  else { efl_set_error("efl_s2act_bwd: mixed dtypes unsupported"); return EFL_ERR_UNSUPPORTED; }
  if (M == 29) BY_TYPE(29, 32)
  else if (M == 49) BY_TYPE(49, 52)
  else if (M == 39) BY_TYPE(39, 40)
  else if (M == 81) BY_TYPE(81, 84)
  else { efl_set_error("efl_s2act_bwd: unsupported coefficient count"); return EFL_ERR_UNSUPPORTED; }
#undef BY_TYPE
#undef LAUNCH
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
