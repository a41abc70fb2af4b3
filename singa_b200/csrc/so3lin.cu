// Operand layout kernels that put SO3_LinearV2 (reference model/EF_layers.py:624-674) on the tcgen05 GEMM engine.
//
// SO3_LinearV2 is out[n, i, :] = W[l(i)] x[n, i, :] (+ bias on i = 0): seven per-degree weights applied to the
// 49 coefficient rows of each node.  The reference materialises an expanded [49, out, in] weight per call and runs a
// batched matmul.  Here the degree selection moves into the OPERAND instead, so that every use is one plain
// [N*49, K] x [K, N'] GEMM of efl_tc_gemm_nt / efl_tc_gemm_tn:
//   * "expand-in"  (few input channels, FFN 16 -> 512): row (n, i) of the A operand holds x[n, i, :] in column block
//     l(i) of a (lmax+1)*Cin wide row, zeros elsewhere, and a 1 in the bias column when i = 0; B = [Cout, KP] holds
//     W[l] side by side and the bias.  The 7x redundant MACs cost ~0.1 ms on the tensor pipe; the output write is
//     the bound either way.
//   * "select-out" (few output channels, 512 -> 16 and proj 112 -> 16): A is x as bf16, B = all seven W[l] stacked
//     ([(lmax+1)*Cout, Cin]); the select kernel keeps column block l(i) of each fp32 result row and adds the bias.
// Backward uses the same two kernels with the roles swapped (expand the incoming gradient / select the data
// gradient); weight gradients fall out of one TN GEMM as exact per-degree blocks because the expanded operand is
// zero outside its block.  HBM-bound, fully coalesced 16-byte accesses.
#include "efl_common.cuh"

namespace efl {

__device__ __forceinline__ int degree_of_row(int i) {
  int l = (int)sqrtf((float)i);
  while ((l + 1) * (l + 1) <= i) ++l;
  while (l * l > i) --l;
  return l;
}

__device__ __forceinline__ uint32_t pack2_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// xe [rows, KP] bf16 from x [rows, C] fp32 (rows = N * L).  expand = 1: block l(i); expand = 0: plain cast + zero pad.
// One thread per 8 output columns (16-byte store).  C % 8 == 0, KP % 8 == 0.
__global__ void so3_expand_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ xe, long long rows, int L,
                                  int C, int KP, int expand, int bias_col) {
  const int chunks = KP >> 3;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= rows * chunks) return;
  const long long r = t / chunks;
  const int q = (int)(t % chunks);
  const int i = (int)(r % L);
  const int col = q << 3;
  int src = -1;  // source column of x for this chunk, -1: zeros
  if (expand) {
    const int l = degree_of_row(i);
    if (col >= l * C && col < (l + 1) * C) src = col - l * C;
  } else if (col < C) {
    src = col;
  }
  uint4 o = make_uint4(0u, 0u, 0u, 0u);
  if (src >= 0) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(x + r * C + src));
    const float4 b = __ldg(reinterpret_cast<const float4*>(x + r * C + src + 4));
    o = make_uint4(pack2_bf16(a.x, a.y), pack2_bf16(a.z, a.w), pack2_bf16(b.x, b.y), pack2_bf16(b.z, b.w));
  } else if (bias_col >= col && bias_col < col + 8 && i == 0) {
    uint32_t w[4] = {0u, 0u, 0u, 0u};
    const int j = bias_col - col;
    w[j >> 1] = (j & 1) ? 0x3f800000u : 0x00003f80u;  // bf16(1.0) = 0x3f80
    o = make_uint4(w[0], w[1], w[2], w[3]);
  }
  *reinterpret_cast<uint4*>(xe + r * KP + col) = o;
}

// y [rows, C] fp32 = ye[r, l(i)*C + c] (+ bias[c] on i = 0); ye fp32 with row pitch ld.  Thread per 4 columns.
__global__ void so3_select_kernel(const float* __restrict__ ye, long long ld, const float* __restrict__ bias,
                                  float* __restrict__ y, long long rows, int L, int C) {
  const int chunks = C >> 2;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= rows * chunks) return;
  const long long r = t / chunks;
  const int c = (int)(t % chunks) << 2;
  const int i = (int)(r % L);
  const int l = degree_of_row(i);
  float4 v = __ldg(reinterpret_cast<const float4*>(ye + r * ld + l * C + c));
  if (bias != nullptr && i == 0) {
    const float4 b = __ldg(reinterpret_cast<const float4*>(bias + c));
    v.x += b.x; v.y += b.y; v.z += b.z; v.w += b.w;
  }
  *reinterpret_cast<float4*>(y + r * C + c) = v;
}

}  // namespace efl

using namespace efl;

extern "C" int efl_so3_expand(const float* x, void* xe, long long rows, int lmax, int C, int KP, int expand,
                              int bias_col, void* stream) {
  if (rows <= 0) return EFL_OK;
  EFL_REQUIRE(C % 8 == 0 && KP % 8 == 0, "efl_so3_expand: C and KP must be multiples of 8");
  EFL_REQUIRE(KP >= (expand ? (lmax + 1) * C : C), "efl_so3_expand: KP too small");
  EFL_REQUIRE(bias_col < KP && (bias_col < 0 || bias_col >= (expand ? (lmax + 1) * C : C)),
              "efl_so3_expand: bias column must lie in the padding");
  const int L = (lmax + 1) * (lmax + 1);
  const long long n = rows * (KP >> 3);
  so3_expand_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      x, (__nv_bfloat16*)xe, rows, L, C, KP, expand, bias_col);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

extern "C" int efl_so3_select(const float* ye, long long ld, const float* bias, float* y, long long rows, int lmax,
                              int C, void* stream) {
  if (rows <= 0) return EFL_OK;
  EFL_REQUIRE(C % 4 == 0 && ld % 4 == 0, "efl_so3_select: C and ld must be multiples of 4");
  const int L = (lmax + 1) * (lmax + 1);
  const long long n = rows * (C >> 2);
  so3_select_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(ye, ld, bias, y, rows, L, C);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
