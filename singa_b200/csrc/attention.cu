// Attention-weight path: alpha logits and the per-destination segment softmax.
//
//  alpha_logits : [E, heads*A] (the first heads*A "extra" m=0 outputs of so2_conv_1)
//                 -> LayerNorm(A) -> SmoothLeakyReLU -> dot with alpha_dot[heads, A] -> [E, heads]
//                 (reference EF_layers.py:1175-1178, 1024, 1669-1677)
//  seg_softmax  : torch_geometric.utils.softmax(alpha, edge_index[1]) over the dst-sorted CSR
//                 (EF_layers.py:1180): exp(x - max_seg) / (sum_seg + 1e-16); one warp per
//                 destination, two passes in registers/L1, no atomics.
#include "efl_common.cuh"

namespace efl {

constexpr float kLnEps = 1e-5f;
constexpr float kSlope = 0.2f;

__device__ __forceinline__ float slrelu(float x) {
  return 0.5f * (1.f + kSlope) * x + 0.5f * (1.f - kSlope) * x * (2.f * sigmoidf_(x) - 1.f);
}
__device__ __forceinline__ float dslrelu(float x) {
  const float s = sigmoidf_(x);
  return 0.5f * (1.f + kSlope) + 0.5f * (1.f - kSlope) * ((2.f * s - 1.f) + 2.f * x * s * (1.f - s));
}

// One warp per edge; lane = (head h = lane >> 2, quarter q = lane & 3) owns channels [8q, 8q+8) of head h, loaded
// as two float4 (the row pitch and the head offset are multiples of 16 bytes).  LayerNorm statistics are reduced
// over the 4 lanes of a head with two shuffles.  heads <= 8; lanes of absent heads idle.
struct Lane8 {
  float v[8];
};
__device__ __forceinline__ Lane8 load8(const float* p) {
  Lane8 r;
  const float4 a = __ldg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  r.v[0] = a.x; r.v[1] = a.y; r.v[2] = a.z; r.v[3] = a.w;
  r.v[4] = b.x; r.v[5] = b.y; r.v[6] = b.z; r.v[7] = b.w;
  return r;
}
__device__ __forceinline__ float quad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  return v;
}
__device__ __forceinline__ void store8(float* p, const float* o) {
  reinterpret_cast<float4*>(p)[0] = make_float4(o[0], o[1], o[2], o[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(o[4], o[5], o[6], o[7]);
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float* o) {
  __nv_bfloat162 h0 = __floats2bfloat162_rn(o[0], o[1]), h1 = __floats2bfloat162_rn(o[2], o[3]);
  __nv_bfloat162 h2 = __floats2bfloat162_rn(o[4], o[5]), h3 = __floats2bfloat162_rn(o[6], o[7]);
  uint4 u;
  u.x = *reinterpret_cast<uint32_t*>(&h0);
  u.y = *reinterpret_cast<uint32_t*>(&h1);
  u.z = *reinterpret_cast<uint32_t*>(&h2);
  u.w = *reinterpret_cast<uint32_t*>(&h3);
  *reinterpret_cast<uint4*>(p) = u;
}

__global__ void __launch_bounds__(256) alpha_logits_fwd_kernel(const float* __restrict__ extra, int ld, int heads,
                                                               const float* __restrict__ ln_w,
                                                               const float* __restrict__ ln_b,
                                                               const float* __restrict__ adot,
                                                               float* __restrict__ logits, int Ec) {
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int h = lane >> 2, q = lane & 3;
  const bool on = h < heads;
  const int hc = on ? h : 0;
  const Lane8 w = load8(ln_w + 8 * q), b = load8(ln_b + 8 * q), ad = load8(adot + hc * 32 + 8 * q);
  for (int e = blockIdx.x * nw + wib; e < Ec; e += gridDim.x * nw) {
    const Lane8 x = load8(extra + (size_t)e * ld + hc * 32 + 8 * q);
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += x.v[j];
    const float mean = quad_sum(s) * (1.f / 32.f);
    float vs = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) vs = fmaf(x.v[j] - mean, x.v[j] - mean, vs);
    const float rstd = rsqrtf(quad_sum(vs) * (1.f / 32.f) + kLnEps);
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc = fmaf(slrelu((x.v[j] - mean) * rstd * w.v[j] + b.v[j]), ad.v[j], acc);
    acc = quad_sum(acc);
    if (on && q == 0) logits[(size_t)e * heads + h] = acc;
  }
}

// backward: d logits -> d extra[:, :heads*32]; parameter-gradient partials per block:
// part[block, 0:heads*32] = d alpha_dot, [heads*32 : +32] = d ln_w, [+32 : +64] = d ln_b
template <typename TOut>
__global__ void __launch_bounds__(256) alpha_logits_bwd_kernel(const float* __restrict__ extra, int ld, int heads,
                                                               const float* __restrict__ ln_w,
                                                               const float* __restrict__ ln_b,
                                                               const float* __restrict__ adot,
                                                               const float* __restrict__ dlogits,
                                                               TOut* __restrict__ dextra, int ldo,
                                                               float* __restrict__ part, int Ec) {
  extern __shared__ float sp[];  // [8 warps][heads*32 + 64]
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int h = lane >> 2, q = lane & 3;
  const bool on = h < heads;
  const int hc = on ? h : 0;
  const int stride = heads * 32 + 64;
  const Lane8 w = load8(ln_w + 8 * q), b = load8(ln_b + 8 * q), ad = load8(adot + hc * 32 + 8 * q);
  float g_w[8], g_b[8], g_dot[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) g_w[j] = g_b[j] = g_dot[j] = 0.f;
  for (int e = blockIdx.x * nw + wib; e < Ec; e += gridDim.x * nw) {
    const Lane8 x = load8(extra + (size_t)e * ld + hc * 32 + 8 * q);
    const float dl = on ? __ldg(dlogits + (size_t)e * heads + h) : 0.f;
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += x.v[j];
    const float mean = quad_sum(s) * (1.f / 32.f);
    float vs = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) vs = fmaf(x.v[j] - mean, x.v[j] - mean, vs);
    const float rstd = rsqrtf(quad_sum(vs) * (1.f / 32.f) + kLnEps);
    float xh[8], dxh[8], m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      xh[j] = (x.v[j] - mean) * rstd;
      const float y = xh[j] * w.v[j] + b.v[j];
      g_dot[j] = fmaf(dl, slrelu(y), g_dot[j]);
      const float dy = dl * ad.v[j] * dslrelu(y);
      g_w[j] = fmaf(dy, xh[j], g_w[j]);
      g_b[j] += dy;
      dxh[j] = dy * w.v[j];
      m1 += dxh[j];
      m2 = fmaf(dxh[j], xh[j], m2);
    }
    m1 = quad_sum(m1) * (1.f / 32.f);
    m2 = quad_sum(m2) * (1.f / 32.f);
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = rstd * (dxh[j] - m1 - xh[j] * m2);
    if (on) store8(dextra + (size_t)e * ldo + h * 32 + 8 * q, o);
  }
  // LayerNorm parameter gradients: sum over the heads (lanes that share q), fixed shuffle order
#pragma unroll
  for (int j = 0; j < 8; ++j) {
#pragma unroll
    for (int o = 4; o < 32; o <<= 1) {
      g_w[j] += __shfl_xor_sync(0xffffffffu, g_w[j], o);
      g_b[j] += __shfl_xor_sync(0xffffffffu, g_b[j], o);
    }
  }
  float* mine = sp + wib * stride;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (on) mine[h * 32 + 8 * q + j] = g_dot[j];
    if (h == 0) {
      mine[heads * 32 + 8 * q + j] = g_w[j];
      mine[heads * 32 + 32 + 8 * q + j] = g_b[j];
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < stride; i += blockDim.x) {
    float acc = 0.f;
    for (int k = 0; k < nw; ++k) acc += sp[k * stride + i];
    part[(size_t)blockIdx.x * stride + i] = acc;
  }
}

// one warp per destination node
__global__ void __launch_bounds__(256) seg_softmax_fwd_kernel(const float* __restrict__ logits, int heads,
                                                              const int* __restrict__ rowptr,
                                                              float* __restrict__ alpha, int n0, int n1) {
  const int n = n0 + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (n >= n1) return;
  const int kb = rowptr[n], ke = rowptr[n + 1];
  for (int h = 0; h < heads; ++h) {
    float mx = -INFINITY;
    for (int k = kb + lane; k < ke; k += 32) mx = fmaxf(mx, logits[(size_t)k * heads + h]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int k = kb + lane; k < ke; k += 32) sum += expf(logits[(size_t)k * heads + h] - mx);
    sum = warp_sum(sum) + 1e-16f;
    for (int k = kb + lane; k < ke; k += 32)
      alpha[(size_t)k * heads + h] = expf(logits[(size_t)k * heads + h] - mx) / sum;
  }
}

// d logits = alpha * (d alpha - sum_seg alpha * d alpha)   (the 1e-16 in the denominator is
// immaterial: the reference's autograd differentiates e/(s+1e-16), identical up to 1e-16/s)
// d alpha arrives as `nparts` partial planes (one per degree group of the rotate-back backward), summed here in
// fixed order.
__global__ void __launch_bounds__(256) seg_softmax_bwd_kernel(const float* __restrict__ alpha,
                                                              const float* __restrict__ dalpha, int nparts,
                                                              long long part_stride, int heads,
                                                              const int* __restrict__ rowptr,
                                                              float* __restrict__ dlogits, int n0, int n1) {
  const int n = n0 + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (n >= n1) return;
  const int kb = rowptr[n], ke = rowptr[n + 1];
  for (int h = 0; h < heads; ++h) {
    float dot = 0.f;
    for (int k = kb + lane; k < ke; k += 32) {
      float da = 0.f;
      for (int p = 0; p < nparts; ++p) da += dalpha[(size_t)p * part_stride + (size_t)k * heads + h];
      dot += alpha[(size_t)k * heads + h] * da;
    }
    dot = warp_sum(dot);
    for (int k = kb + lane; k < ke; k += 32) {
      float da = 0.f;
      for (int p = 0; p < nparts; ++p) da += dalpha[(size_t)p * part_stride + (size_t)k * heads + h];
      const float a = alpha[(size_t)k * heads + h];
      dlogits[(size_t)k * heads + h] = a * (da - dot);
    }
  }
}

}  // namespace efl

using namespace efl;

extern "C" int efl_alpha_logits_fwd(const float* extra, int ld, int heads, int alpha_ch, const float* ln_w,
                                    const float* ln_b, const float* alpha_dot, float* logits, int Ec, void* stream) {
  if (Ec <= 0) return EFL_OK;
  EFL_REQUIRE(alpha_ch == 32 && heads <= 8, "efl_alpha_logits_fwd: attn_alpha_channels must be 32, heads <= 8");
  EFL_REQUIRE((ld % 4) == 0 && ((uintptr_t)extra & 15) == 0, "efl_alpha_logits_fwd: extra must be 16-byte aligned, ld % 4 == 0");
  const int warps_per_block = 8;
  const int want = (Ec + warps_per_block - 1) / warps_per_block;
  const int cap = 148 * 8 * 4;
  alpha_logits_fwd_kernel<<<want < cap ? want : cap, 256, 0, (cudaStream_t)stream>>>(
      extra, ld, heads, ln_w, ln_b, alpha_dot, logits, Ec);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// part must hold nblocks * (heads*32 + 64) floats; returns nblocks through *nblocks_out when part == NULL
extern "C" int efl_alpha_logits_bwd(const float* extra, int ld, int heads, int alpha_ch, const float* ln_w,
                                    const float* ln_b, const float* alpha_dot, const float* dlogits, void* dextra,
                                    int ldo, int bf16_out, float* part, int nblocks, int Ec, void* stream) {
  EFL_REQUIRE(alpha_ch == 32 && heads <= 8, "efl_alpha_logits_bwd: attn_alpha_channels must be 32, heads <= 8");
  EFL_REQUIRE(nblocks > 0, "efl_alpha_logits_bwd: nblocks must be positive");
  EFL_REQUIRE((ld % 4) == 0 && (ldo % 8) == 0 && ((uintptr_t)extra & 15) == 0 && ((uintptr_t)dextra & 15) == 0,
              "efl_alpha_logits_bwd: extra/dextra must be 16-byte aligned (ld % 4 == 0, ldo % 8 == 0)");
  const size_t smem = 8 * (heads * 32 + 64) * sizeof(float);
  if (bf16_out)
    alpha_logits_bwd_kernel<__nv_bfloat16><<<nblocks, 256, smem, (cudaStream_t)stream>>>(
        extra, ld, heads, ln_w, ln_b, alpha_dot, dlogits, (__nv_bfloat16*)dextra, ldo, part, Ec);
  else
    alpha_logits_bwd_kernel<float><<<nblocks, 256, smem, (cudaStream_t)stream>>>(
        extra, ld, heads, ln_w, ln_b, alpha_dot, dlogits, (float*)dextra, ldo, part, Ec);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

extern "C" int efl_segment_softmax_fwd(const float* logits, int heads, const int* rowptr, float* alpha, int n0,
                                       int n1, void* stream) {
  if (n1 <= n0) return EFL_OK;
  const int nodes = n1 - n0;
  seg_softmax_fwd_kernel<<<(nodes + 7) / 8, 256, 0, (cudaStream_t)stream>>>(logits, heads, rowptr, alpha, n0, n1);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

extern "C" int efl_segment_softmax_bwd(const float* alpha, const float* dalpha, int nparts, long long part_stride,
                                       int heads, const int* rowptr, float* dlogits, int n0, int n1, void* stream) {
  EFL_REQUIRE(nparts >= 1, "efl_segment_softmax_bwd: nparts must be >= 1");
  if (n1 <= n0) return EFL_OK;
  const int nodes = n1 - n0;
  seg_softmax_bwd_kernel<<<(nodes + 7) / 8, 256, 0, (cudaStream_t)stream>>>(alpha, dalpha, nparts, part_stride, heads,
                                                                             rowptr, dlogits, n0, n1);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// ---------------------------------------------------------------------------------------------------
// Deterministic gradient of an embedding lookup with few rows (source/target element embeddings of the edge scalars,
// Embedding.py:97-99: 43 rows x 16 channels, E lookups).  torch's embedding_dense_backward sums with float atomics
// (run-to-run different bits); here every thread owns private accumulators and all reductions run in a fixed order.
//   stage 1: CTA b takes a contiguous slice of lookups; thread = (sub-lane s, channel c) adds row e into its private
//            shared-memory table acc[s][id][c]; the sub-lanes are then summed in order -> part[b][rows][C]
//   stage 2: dW[r][c] = sum_b part[b][r][c] in order
// ---------------------------------------------------------------------------------------------------
namespace efl {

__global__ void __launch_bounds__(256) embedding_grad_partial_kernel(const long long* __restrict__ ids,
                                                                     const float* __restrict__ dx, long long ld,
                                                                     int E, int rows, int C, float* __restrict__ part) {
  extern __shared__ float acc[];  // [S][rows][C]
  const int S = 256 / C;
  const int s = threadIdx.x / C, c = threadIdx.x % C;
  for (int i = threadIdx.x; i < S * rows * C; i += blockDim.x) acc[i] = 0.f;
  __syncthreads();
  const int per = (E + gridDim.x - 1) / gridDim.x;
  const int e0 = blockIdx.x * per, e1 = min(E, e0 + per);
  if (s < S) {
    float* mine = acc + (size_t)s * rows * C + c;
    for (int e = e0 + s; e < e1; e += S) {
      const long long id = ids[e];
      if (id >= 0 && id < rows) mine[id * C] += dx[(size_t)e * ld + c];
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < rows * C; i += blockDim.x) {
    float t = 0.f;
    for (int k = 0; k < S; ++k) t += acc[(size_t)k * rows * C + i];
    part[(size_t)blockIdx.x * rows * C + i] = t;
  }
}

__global__ void embedding_grad_reduce_kernel(const float* __restrict__ part, int nb, int n, float* __restrict__ dW) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float t = 0.f;
  for (int b = 0; b < nb; ++b) t += part[(size_t)b * n + i];
  dW[i] = t;
}

}  // namespace efl

// ids [E] int64, dx rows of ld floats (pointer already offset to the embedding's columns), dW [rows, C].
// workspace: nblocks * rows * C floats.  C must divide 256; rows * C * (256 / C) floats of shared memory (<= 200 KB).
extern "C" int efl_embedding_grad(const long long* ids, const float* dx, long long ld, int E, int rows, int C,
                                  float* dW, float* workspace, int nblocks, void* stream) {
  EFL_REQUIRE(C > 0 && 256 % C == 0 && rows > 0 && nblocks > 0, "efl_embedding_grad: C must divide 256");
  const size_t smem = sizeof(float) * (size_t)(256 / C) * rows * C;
  EFL_REQUIRE(smem <= 200 * 1024, "efl_embedding_grad: table too large for the shared-memory accumulators");
  cudaStream_t s = (cudaStream_t)stream;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(efl::embedding_grad_partial_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) {
      efl_set_error(cudaGetErrorString(e));
      return (int)e;
    }
  }
  efl::embedding_grad_partial_kernel<<<nblocks, 256, smem, s>>>(ids, dx, ld, E, rows, C, workspace);
  EFL_CHECK_LAUNCH();
  const int n = rows * C;
  efl::embedding_grad_reduce_kernel<<<(n + 255) / 256, 256, 0, s>>>(workspace, nblocks, n, dW);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
