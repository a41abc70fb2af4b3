// bf16x3 operand preparation for the accurate tensor-core mode ("3xTF32 / bf16x3 split as the
// fp32-accurate mode").  An fp32 value x is split into hi = bf16(x) and lo = bf16(x - hi); a product a*b is then
// a_hi*b_hi + a_hi*b_lo + a_lo*b_hi (the dropped a_lo*b_lo term is 2^-16 * 2^-16 relative), accumulated in fp32 by
// the tensor cores.  Instead of three GEMMs the three terms are laid side by side along the REDUCTION dimension, so
// the unchanged tcgen05 GEMM kernels (csrc/tc_gemm.cu) do all three in one launch with a 3x longer K:
//   NT (reduce over columns):  A' = [a_hi | a_hi | a_lo]  [M, 3K],   B' = [b_hi | b_lo | b_hi]  [N, 3K]
//   TN (reduce over rows):     A' = [a_hi ; a_hi ; a_lo]  [3R, N1],  B' = [b_hi ; b_lo ; b_hi]  [3R, N2]
// HBM-bound elementwise kernel: reads 4 B, writes 6 B per element.
#include "efl_common.cuh"

namespace efl {

// pattern[t] = 0 -> hi, 1 -> lo for the three copies
template <bool STACK>
__global__ void __launch_bounds__(256) split3_kernel(const float* __restrict__ x, long long ld, long long rows, int cols,
                                                     __nv_bfloat16* __restrict__ out, long long ldo, int p0, int p1,
                                                     int p2) {
  const int c4 = cols >> 2;   // cols % 4 == 0
  const long long total = rows * c4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / c4;
    const int c = (int)(i % c4) << 2;
    const float4 v = *reinterpret_cast<const float4*>(x + r * ld + c);
    const float f[4] = {v.x, v.y, v.z, v.w};
    __nv_bfloat16 hi[4], lo[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      hi[j] = __float2bfloat16_rn(f[j]);
      lo[j] = __float2bfloat16_rn(f[j] - __bfloat162float(hi[j]));
    }
    const uint2 h = make_uint2(*reinterpret_cast<uint32_t*>(&hi[0]), *reinterpret_cast<uint32_t*>(&hi[2]));
    const uint2 l = make_uint2(*reinterpret_cast<uint32_t*>(&lo[0]), *reinterpret_cast<uint32_t*>(&lo[2]));
    const int pat[3] = {p0, p1, p2};
#pragma unroll
    for (int t = 0; t < 3; ++t) {
      __nv_bfloat16* o = STACK ? out + ((long long)t * rows + r) * ldo + c : out + r * ldo + (long long)t * cols + c;
      *reinterpret_cast<uint2*>(o) = pat[t] ? l : h;
    }
  }
}

}  // namespace efl

// x [rows, cols] fp32 (row pitch ld) -> bf16 out.  stack = 0: out [rows, 3*cols] (pitch ldo >= 3*cols), copies side by
// side; stack = 1: out [3*rows, cols] (pitch ldo), copies one under the other.  second = 0: (hi, hi, lo) - the A
// operand; second = 1: (hi, lo, hi) - the B operand.  cols % 4 == 0, x 16-byte aligned, out 8-byte aligned.
extern "C" int efl_split3(const float* x, long long ld, long long rows, int cols, void* out, long long ldo, int stack,
                          int second, void* stream) {
  if (rows <= 0 || cols <= 0) return EFL_OK;
  EFL_REQUIRE(cols % 4 == 0 && ld % 4 == 0 && ldo % 4 == 0, "efl_split3: cols, ld and ldo must be multiples of 4");
  EFL_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)out & 7) == 0, "efl_split3: misaligned operand");
  const long long total = rows * (cols / 4);
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  const int p1 = second ? 1 : 0, p2 = second ? 0 : 1;
  if (stack)
    efl::split3_kernel<true><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, ld, rows, cols, (__nv_bfloat16*)out,
                                                                               ldo, 0, p1, p2);
  else
    efl::split3_kernel<false><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(x, ld, rows, cols, (__nv_bfloat16*)out,
                                                                                ldo, 0, p1, p2);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
