// EquivariantRMSNormArraySphericalHarmonicsV2 as configured by get_normalization_layer
// (reference EF_layers.py:2099-2192, 2254-2273): centre l=0 over channels, one degree-balanced RMS per node,
// per-degree gain, bias on l=0.  One warp per node, C = 16 channels: lane -> channel lane%16, coefficient
// rows 2t + lane/16.  HBM-bound: reads and writes the node tensor once (fwd) / twice (bwd).
#include "efl_common.cuh"

namespace efl {

template <int LMAX>
__device__ __forceinline__ int degree_of(int i) {
  int l = 0;
#pragma unroll
  for (int k = 1; k <= LMAX; ++k) l += (i >= k * k) ? 1 : 0;
  return l;
}

template <int LMAX>
__global__ void __launch_bounds__(256) rmsnorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, float eps,
                                                          float* __restrict__ y, int N) {
  constexpr int L = (LMAX + 1) * (LMAX + 1), C = 16, T = (L + 1) / 2;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= N) return;
  const int c = lane & 15, half = lane >> 4;
  const float* xp = x + (size_t)warp * L * C;
  float v[T];
#pragma unroll
  for (int t = 0; t < T; ++t) {
    const int i = 2 * t + half;
    v[t] = i < L ? xp[i * C + c] : 0.f;
  }
  // centre the l = 0 row (held by lanes 0..15 at t = 0)
  float m = half == 0 ? v[0] : 0.f;
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) m += __shfl_xor_sync(0xffffffffu, m, o);
  m = __shfl_sync(0xffffffffu, m, 0) * (1.f / C);
  if (half == 0) v[0] -= m;
  float p = 0.f;
#pragma unroll
  for (int t = 0; t < T; ++t) {
    const int i = 2 * t + half;
    if (i < L) p += v[t] * v[t] / (float)((2 * degree_of<LMAX>(i) + 1) * (LMAX + 1));
  }
  p = warp_sum(p) * (1.f / C);
  const float s = rsqrtf(p + eps);
  float* yp = y + (size_t)warp * L * C;
#pragma unroll
  for (int t = 0; t < T; ++t) {
    const int i = 2 * t + half;
    if (i < L) {
      float o = v[t] * s * gamma[degree_of<LMAX>(i) * C + c];
      if (i == 0) o += beta[c];
      yp[i * C + c] = o;
    }
  }
}

// part[block, 0:(LMAX+1)*16] = d gamma partials, [(LMAX+1)*16 : +16] = d beta partials
template <int LMAX>
__global__ void __launch_bounds__(256) rmsnorm_bwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                          float eps, const float* __restrict__ dy,
                                                          float* __restrict__ dx, float* __restrict__ part, int N) {
  constexpr int L = (LMAX + 1) * (LMAX + 1), C = 16, T = (L + 1) / 2, NP = (LMAX + 2) * C;
  __shared__ float sp[8][NP];
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c = lane & 15, half = lane >> 4;
  float gg[LMAX + 1], gb = 0.f;
#pragma unroll
  for (int l = 0; l <= LMAX; ++l) gg[l] = 0.f;
  float gam[LMAX + 1];
#pragma unroll
  for (int l = 0; l <= LMAX; ++l) gam[l] = gamma[l * C + c];
  for (int n = blockIdx.x * 8 + wib; n < N; n += gridDim.x * 8) {
    const float* xp = x + (size_t)n * L * C;
    const float* dp = dy + (size_t)n * L * C;
    float v[T], d[T];
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const int i = 2 * t + half;
      v[t] = i < L ? xp[i * C + c] : 0.f;
      d[t] = i < L ? dp[i * C + c] : 0.f;
    }
    float m = half == 0 ? v[0] : 0.f;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) m += __shfl_xor_sync(0xffffffffu, m, o);
    m = __shfl_sync(0xffffffffu, m, 0) * (1.f / C);
    if (half == 0) v[0] -= m;
    float p = 0.f, D = 0.f;
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const int i = 2 * t + half;
      if (i < L) {
        const int l = degree_of<LMAX>(i);
        p += v[t] * v[t] / (float)((2 * l + 1) * (LMAX + 1));
        D += d[t] * gam[l] * v[t];
      }
    }
    p = warp_sum(p) * (1.f / C);
    D = warp_sum(D);
    const float s = rsqrtf(p + eps);
    const float k = s * s * s * D * (1.f / C);
    float g0 = 0.f;  // d x' of the l = 0 row (for the centring backward)
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const int i = 2 * t + half;
      if (i < L) {
        const int l = degree_of<LMAX>(i);
        const float w = 1.f / (float)((2 * l + 1) * (LMAX + 1));
        gg[l] += d[t] * v[t] * s;
        if (i == 0) gb += d[t];
        const float g = s * gam[l] * d[t] - k * w * v[t];
        if (i == 0) g0 = g;
        d[t] = g;
      }
    }
    float gm = half == 0 ? g0 : 0.f;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) gm += __shfl_xor_sync(0xffffffffu, gm, o);
    gm = __shfl_sync(0xffffffffu, gm, 0) * (1.f / C);
    if (half == 0) d[0] -= gm;
    float* op = dx + (size_t)n * L * C;
#pragma unroll
    for (int t = 0; t < T; ++t) {
      const int i = 2 * t + half;
      if (i < L) op[i * C + c] = d[t];
    }
  }
  // combine the two lanes of a channel, then the warps of the block
#pragma unroll
  for (int l = 0; l <= LMAX; ++l) gg[l] += __shfl_xor_sync(0xffffffffu, gg[l], 16);
  gb += __shfl_xor_sync(0xffffffffu, gb, 16);
  if (half == 0) {
#pragma unroll
    for (int l = 0; l <= LMAX; ++l) sp[wib][l * C + c] = gg[l];
    sp[wib][(LMAX + 1) * C + c] = gb;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < NP; i += blockDim.x) {
    float acc = 0.f;
    for (int w = 0; w < 8; ++w) acc += sp[w][i];
    part[(size_t)blockIdx.x * NP + i] = acc;
  }
}

}  // namespace efl

using namespace efl;

extern "C" int efl_rmsnorm_sh_fwd(int lmax, int C, const float* x, const float* gamma, const float* beta, float eps,
                                  float* y, int N, void* stream) {
  if (N <= 0) return EFL_OK;
  EFL_REQUIRE(C == 16, "efl_rmsnorm_sh_fwd: sphere_channels must be 16");
  cudaStream_t s = (cudaStream_t)stream;
  const int blocks = (N + 7) / 8;
  if (lmax == 6) rmsnorm_fwd_kernel<6><<<blocks, 256, 0, s>>>(x, gamma, beta, eps, y, N);
  else if (lmax == 8) rmsnorm_fwd_kernel<8><<<blocks, 256, 0, s>>>(x, gamma, beta, eps, y, N);
  else if (lmax == 4) rmsnorm_fwd_kernel<4><<<blocks, 256, 0, s>>>(x, gamma, beta, eps, y, N);
  else { efl_set_error("efl_rmsnorm_sh_fwd: unsupported lmax"); return EFL_ERR_UNSUPPORTED; }
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// part: nblocks * (lmax + 2) * 16 floats (per-block partials of d gamma [lmax+1,16] | d beta [16])
extern "C" int efl_rmsnorm_sh_bwd(int lmax, int C, const float* x, const float* gamma, float eps, const float* dy,
                                  float* dx, float* part, int nblocks, int N, void* stream) {
  EFL_REQUIRE(C == 16, "efl_rmsnorm_sh_bwd: sphere_channels must be 16");
  EFL_REQUIRE(nblocks > 0, "efl_rmsnorm_sh_bwd: nblocks must be positive");
  cudaStream_t s = (cudaStream_t)stream;
  if (lmax == 6) rmsnorm_bwd_kernel<6><<<nblocks, 256, 0, s>>>(x, gamma, eps, dy, dx, part, N);
  else if (lmax == 8) rmsnorm_bwd_kernel<8><<<nblocks, 256, 0, s>>>(x, gamma, eps, dy, dx, part, N);
  else if (lmax == 4) rmsnorm_bwd_kernel<4><<<nblocks, 256, 0, s>>>(x, gamma, eps, dy, dx, part, N);
  else { efl_set_error("efl_rmsnorm_sh_bwd: unsupported lmax"); return EFL_ERR_UNSUPPORTED; }
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
