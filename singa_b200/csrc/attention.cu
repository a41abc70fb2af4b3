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

// one warp per edge, lane = alpha channel (A == 32)
__global__ void __launch_bounds__(256) alpha_logits_fwd_kernel(const float* __restrict__ extra, int ld, int heads,
                                                               const float* __restrict__ ln_w,
                                                               const float* __restrict__ ln_b,
                                                               const float* __restrict__ adot,
                                                               float* __restrict__ logits, int Ec) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= Ec) return;
  const float w = ln_w[lane], b = ln_b[lane];
  for (int h = 0; h < heads; ++h) {
    const float v = extra[(size_t)warp * ld + h * 32 + lane];
    const float mean = warp_sum(v) * (1.f / 32.f);
    const float d = v - mean;
    const float var = warp_sum(d * d) * (1.f / 32.f);
    const float y = d * rsqrtf(var + kLnEps) * w + b;
    const float s = warp_sum(slrelu(y) * adot[h * 32 + lane]);
    if (lane == 0) logits[(size_t)warp * heads + h] = s;
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
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nw = blockDim.x >> 5;
  const int stride = heads * 32 + 64;
  const float w = ln_w[lane], b = ln_b[lane];
  float g_w = 0.f, g_b = 0.f;
  float g_dot[8];
#pragma unroll
  for (int h = 0; h < 8; ++h) g_dot[h] = 0.f;
  for (int e = blockIdx.x * nw + wib; e < Ec; e += gridDim.x * nw) {
#pragma unroll
    for (int h = 0; h < 8; ++h) {
      if (h < heads) {
        const float v = extra[(size_t)e * ld + h * 32 + lane];
        const float mean = warp_sum(v) * (1.f / 32.f);
        const float d = v - mean;
        const float var = warp_sum(d * d) * (1.f / 32.f);
        const float rstd = rsqrtf(var + kLnEps);
        const float xh = d * rstd;
        const float y = xh * w + b;
        const float dl = dlogits[(size_t)e * heads + h];
        g_dot[h] += dl * slrelu(y);
        const float dy = dl * adot[h * 32 + lane] * dslrelu(y);
        g_w += dy * xh;
        g_b += dy;
        const float dxh = dy * w;
        const float m1 = warp_sum(dxh) * (1.f / 32.f);
        const float m2 = warp_sum(dxh * xh) * (1.f / 32.f);
        dextra[(size_t)e * ldo + h * 32 + lane] = from_f<TOut>(rstd * (dxh - m1 - xh * m2));
      }
    }
  }
  float* mine = sp + wib * stride;
#pragma unroll
  for (int h = 0; h < 8; ++h)
    if (h < heads) mine[h * 32 + lane] = g_dot[h];
  mine[heads * 32 + lane] = g_w;
  mine[heads * 32 + 32 + lane] = g_b;
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
__global__ void __launch_bounds__(256) seg_softmax_bwd_kernel(const float* __restrict__ alpha,
                                                              const float* __restrict__ dalpha, int heads,
                                                              const int* __restrict__ rowptr,
                                                              float* __restrict__ dlogits, int n0, int n1) {
  const int n = n0 + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (n >= n1) return;
  const int kb = rowptr[n], ke = rowptr[n + 1];
  for (int h = 0; h < heads; ++h) {
    float dot = 0.f;
    for (int k = kb + lane; k < ke; k += 32) dot += alpha[(size_t)k * heads + h] * dalpha[(size_t)k * heads + h];
    dot = warp_sum(dot);
    for (int k = kb + lane; k < ke; k += 32) {
      const float a = alpha[(size_t)k * heads + h];
      dlogits[(size_t)k * heads + h] = a * (dalpha[(size_t)k * heads + h] - dot);
    }
  }
}

}  // namespace efl

using namespace efl;

extern "C" int efl_alpha_logits_fwd(const float* extra, int ld, int heads, int alpha_ch, const float* ln_w,
                                    const float* ln_b, const float* alpha_dot, float* logits, int Ec, void* stream) {
  if (Ec <= 0) return EFL_OK;
  EFL_REQUIRE(alpha_ch == 32 && heads <= 8, "efl_alpha_logits_fwd: attn_alpha_channels must be 32, heads <= 8");
  const int warps_per_block = 8;
  alpha_logits_fwd_kernel<<<(Ec + warps_per_block - 1) / warps_per_block, 256, 0, (cudaStream_t)stream>>>(
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

extern "C" int efl_segment_softmax_bwd(const float* alpha, const float* dalpha, int heads, const int* rowptr,
                                       float* dlogits, int n0, int n1, void* stream) {
  if (n1 <= n0) return EFL_OK;
  const int nodes = n1 - n0;
  seg_softmax_bwd_kernel<<<(nodes + 7) / 8, 256, 0, (cudaStream_t)stream>>>(alpha, dalpha, heads, rowptr, dlogits,
                                                                             n0, n1);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
