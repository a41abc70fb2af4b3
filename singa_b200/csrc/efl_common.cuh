// Shared device helpers and compile-time SO(3) index tables for the EF_layers path.
// No torch headers: this library is a plain C-ABI CUDA library (include/efl.h).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#define EFL_OK 0
#define EFL_ERR_UNSUPPORTED 10001
#define EFL_ERR_ARG 10002

void efl_set_error(const char* msg);

#define EFL_CHECK_LAUNCH()                                      \
  do {                                                          \
    cudaError_t e__ = cudaGetLastError();                       \
    if (e__ != cudaSuccess) {                                   \
      efl_set_error(cudaGetErrorString(e__));                   \
      return (int)e__;                                          \
    }                                                           \
  } while (0)

#define EFL_REQUIRE(cond, msg)                                  \
  do {                                                          \
    if (!(cond)) {                                              \
      efl_set_error(msg);                                       \
      return EFL_ERR_ARG;                                       \
    }                                                           \
  } while (0)

// cudaFuncSetAttribute is per DEVICE: "done once" flags are kept as a bit per device ordinal, so a process that
// touches a second GPU does not launch with an unset dynamic shared-memory limit there.
static inline bool efl_first_use_on_device(unsigned long long* mask) {
  int dev = 0;
  cudaGetDevice(&dev);
  const unsigned long long bit = 1ull << (dev & 63);
  if (*mask & bit) return false;
  *mask |= bit;
  return true;
}

namespace efl {

// Coefficient bookkeeping for one resolution (lmax, mmax); mirrors what
// CoefficientMappingModule builds at run time (reference EF_layers.py:1421-1549).
//   l-primary full index : l*l + l + n,           n = -l..l              (L entries)
//   kept rows            : |m'| <= min(l, MMAX)                          (M entries)
//   m-primary index      : m=0 block (l=0..LMAX), then for m=1..MMAX the +m block
//                          (l=m..LMAX, "real") followed by the -m block ("imag")
//   Wigner store         : for each l the kept rows m' of D^l, each 2l+1 wide and padded to a multiple of 4 floats
//                          (16-byte rows: float4 / cp.async staging), WSZ floats per edge (296 for (6,2))
template <int LMAX, int MMAX>
struct Tab {
  static constexpr int L = (LMAX + 1) * (LMAX + 1);
  __host__ __device__ static constexpr int nrows(int l) { return (l < MMAX) ? 2 * l + 1 : 2 * MMAX + 1; }
  __host__ __device__ static constexpr int mm(int l) { return l < MMAX ? l : MMAX; }
  __host__ __device__ static constexpr int rowlen(int l) { return (2 * l + 1 + 3) & ~3; }
  __host__ __device__ static constexpr int woff(int l) {
    int o = 0;
    for (int k = 0; k < l; ++k) o += nrows(k) * rowlen(k);
    return o;
  }
  static constexpr int WSZ = woff(LMAX + 1);
  __host__ __device__ static constexpr int count_m() {
    int o = 0;
    for (int k = 0; k <= LMAX; ++k) o += nrows(k);
    return o;
  }
  static constexpr int M = count_m();
  // offset of the first (real) block of order am in m-primary order
  __host__ __device__ static constexpr int moff(int am) {
    int o = 0;
    if (am == 0) return 0;
    o = LMAX + 1;
    for (int k = 1; k < am; ++k) o += 2 * (LMAX + 1 - k);
    return o;
  }
  __host__ __device__ static constexpr int mprim(int l, int mp) {
    int am = mp < 0 ? -mp : mp;
    if (am == 0) return l;
    return moff(am) + (mp > 0 ? 0 : (LMAX + 1 - am)) + (l - am);
  }
  // index of row m' of degree l inside the Wigner store
  __host__ __device__ static constexpr int wrow(int l, int mp) { return woff(l) + (mp + mm(l)) * rowlen(l); }
  // rotate-back work split: one warp per degree group; group 0 = degrees 0..3, group g >= 1 = degree 3 + g
  // (FMA work per group for (6,2): 70 | 45 | 55 | 65)
  static constexpr int NG = LMAX > 3 ? LMAX - 2 : 1;
  __host__ __device__ static constexpr int glo(int g) { return g == 0 ? 0 : 3 + g; }
  __host__ __device__ static constexpr int ghi(int g) { return g == 0 ? (LMAX < 3 ? LMAX : 3) : 3 + g; }
  // rotate-back rescale sqrt((2l+1)/(2 MMAX+1)) for l > MMAX (EF_layers.py:1530-1549)
  __host__ __device__ static float rescale(int l) {
    return l > MMAX ? sqrtf((float)(2 * l + 1) / (float)(2 * MMAX + 1)) : 1.0f;
  }
};

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ float siluf_(float x) { return x * sigmoidf_(x); }
// d/dx silu(x) = s + x s (1 - s)
__device__ __forceinline__ float dsiluf_(float x) {
  float s = sigmoidf_(x);
  return s * (1.0f + x * (1.0f - s));
}

}  // namespace efl
