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

// ---------------------------------------------------------------------------------------------------
// Operand kernels of the node-level S2 activation on grids the fused tcgen05 node kernel is not built for (lmax = 8:
// 81 coefficients x 342 grid points, BASELINE.json config 5).  The two projections run as plain tcgen05 GEMMs over
// all (node, channel) rows (ops._node_s2_gemm_*); these kernels replace the torch transposes / casts / SiLU calls
// around them (about 20 full passes over [N, L, CH] per layer, 65 ms of a 292 ms step in round 1):
//   pack      x [N, L, CH] fp32 -> bf16 rows [(n, c), LP] (coefficients contiguous, zero padded; optionally l = 0 zeroed)
//   unpack    bf16 rows -> y [N, L, CH] fp32, with the l = 0 row replaced by SiLU(gate) when a gate is given
//   silu      grid <- SiLU(grid) in place (bf16)
//   dsilu_mul dact <- dact * SiLU'(grid) in place (bf16)
// All HBM-bound; pack / unpack transpose through a padded shared-memory tile so both sides stay coalesced.
// ---------------------------------------------------------------------------------------------------
namespace efl {

constexpr int NS2_TC = 64;   // channels per tile

__device__ __forceinline__ float silu_fast_(float x) { return __fdividef(x, 1.0f + __expf(-x)); }
__device__ __forceinline__ float dsilu_fast_(float x) {
  const float s = __fdividef(1.0f, 1.0f + __expf(-x));
  return s * fmaf(x, 1.0f - s, 1.0f);
}

// LP is even; a thread writes / reads two adjacent coefficients of one (node, channel) row as one bf16x2 word
__global__ void __launch_bounds__(256) node_rows_pack_kernel(const float* __restrict__ x, int L, int CH, int LP,
                                                             int zero_row0, __nv_bfloat16* __restrict__ out) {
  extern __shared__ float tile[];   // [L][NS2_TC + 1]
  const int n = blockIdx.x, c0 = blockIdx.y * NS2_TC;
  const float* xn = x + (size_t)n * L * CH;
  for (int i = threadIdx.x; i < L * NS2_TC; i += blockDim.x) {
    const int l = i / NS2_TC, c = i % NS2_TC;
    tile[l * (NS2_TC + 1) + c] = (c0 + c < CH) ? xn[(size_t)l * CH + c0 + c] : 0.f;
  }
  __syncthreads();
  const int half = LP >> 1;
  uint32_t* o = reinterpret_cast<uint32_t*>(out + ((size_t)n * CH + c0) * LP);
  for (int i = threadIdx.x; i < NS2_TC * half; i += blockDim.x) {
    const int c = i / half, l = (i % half) << 1;
    if (c0 + c < CH) {
      const float v0 = (l < L && !(zero_row0 && l == 0)) ? tile[l * (NS2_TC + 1) + c] : 0.f;
      const float v1 = (l + 1 < L) ? tile[(l + 1) * (NS2_TC + 1) + c] : 0.f;
      __nv_bfloat162 h = __floats2bfloat162_rn(v0, v1);
      o[(size_t)c * half + (l >> 1)] = *reinterpret_cast<uint32_t*>(&h);
    }
  }
}

__global__ void __launch_bounds__(256) node_rows_unpack_kernel(const __nv_bfloat16* __restrict__ yt, int L, int CH,
                                                               int LP, const float* __restrict__ gate,
                                                               float* __restrict__ y) {
  extern __shared__ float tile[];   // [L + 1][NS2_TC + 1]
  const int n = blockIdx.x, c0 = blockIdx.y * NS2_TC;
  const int half = LP >> 1;
  const uint32_t* src = reinterpret_cast<const uint32_t*>(yt + ((size_t)n * CH + c0) * LP);
  for (int i = threadIdx.x; i < NS2_TC * half; i += blockDim.x) {
    const int c = i / half, l = (i % half) << 1;
    if (l < L && c0 + c < CH) {
      const uint32_t w = src[(size_t)c * half + (l >> 1)];
      tile[l * (NS2_TC + 1) + c] = __uint_as_float(w << 16);
      tile[(l + 1) * (NS2_TC + 1) + c] = __uint_as_float(w & 0xffff0000u);   // row L is scratch when L is odd
    }
  }
  __syncthreads();
  float* yn = y + (size_t)n * L * CH;
  for (int i = threadIdx.x; i < L * NS2_TC; i += blockDim.x) {
    const int l = i / NS2_TC, c = i % NS2_TC;
    if (c0 + c < CH) {
      float v = tile[l * (NS2_TC + 1) + c];
      if (gate && l == 0) v = silu_fast_(gate[(size_t)n * CH + c0 + c]);
      yn[(size_t)l * CH + c0 + c] = v;
    }
  }
}

// mode 0: a <- SiLU(a);  mode 1: a <- a * SiLU'(b)
__global__ void __launch_bounds__(256) bf16_silu_kernel(__nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                                                        long long n8, int mode) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    uint4 va = reinterpret_cast<uint4*>(a)[i];
    uint4 vb = mode ? reinterpret_cast<const uint4*>(b)[i] : va;
    __nv_bfloat16* pa = reinterpret_cast<__nv_bfloat16*>(&va);
    const __nv_bfloat16* pb = reinterpret_cast<const __nv_bfloat16*>(&vb);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float x = __bfloat162float(pa[j]);
      pa[j] = __float2bfloat16_rn(mode ? x * dsilu_fast_(__bfloat162float(pb[j])) : silu_fast_(x));
    }
    reinterpret_cast<uint4*>(a)[i] = va;
  }
}

}  // namespace efl

extern "C" int efl_node_rows_pack(const float* x, int N, int L, int CH, int LP, int zero_row0, void* out,
                                  void* stream) {
  if (N <= 0) return EFL_OK;
  EFL_REQUIRE(LP >= L && L <= 128 && LP % 2 == 0, "efl_node_rows_pack: LP >= L, LP even, L <= 128");
  dim3 grid(N, (CH + efl::NS2_TC - 1) / efl::NS2_TC);
  efl::node_rows_pack_kernel<<<grid, 256, sizeof(float) * L * (efl::NS2_TC + 1), (cudaStream_t)stream>>>(
      x, L, CH, LP, zero_row0, (__nv_bfloat16*)out);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

extern "C" int efl_node_rows_unpack(const void* yt, int N, int L, int CH, int LP, const float* gate, float* y,
                                    void* stream) {
  if (N <= 0) return EFL_OK;
  EFL_REQUIRE(LP >= L && L <= 128 && LP % 2 == 0, "efl_node_rows_unpack: LP >= L, LP even, L <= 128");
  dim3 grid(N, (CH + efl::NS2_TC - 1) / efl::NS2_TC);
  efl::node_rows_unpack_kernel<<<grid, 256, sizeof(float) * (L + 1) * (efl::NS2_TC + 1), (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)yt, L, CH, LP, gate, y);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// a, b: bf16 arrays of n elements (n % 8 == 0, 16-byte aligned).  mode 0: a <- SiLU(a); mode 1: a <- a * SiLU'(b).
extern "C" int efl_bf16_silu(void* a, const void* b, long long n, int mode, void* stream) {
  if (n <= 0) return EFL_OK;
  EFL_REQUIRE(n % 8 == 0 && ((uintptr_t)a & 15) == 0 && (mode == 0 || ((uintptr_t)b & 15) == 0),
              "efl_bf16_silu: n % 8 == 0 and 16-byte aligned arrays");
  long long blocks = (n / 8 + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  efl::bf16_silu_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)a, (const __nv_bfloat16*)b,
                                                                          n / 8, mode);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
