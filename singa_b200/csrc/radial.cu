// Fused RadialFunction MLP (reference model/EF_layers.py:1634-1657) for the configured shape
//   Linear(K0 -> 16) -> LayerNorm(16) -> SiLU -> Linear(16 -> 16) -> LayerNorm(16) -> SiLU -> Linear(16 -> NO)
// (edge_channels_list [48, 16, 16] + the output width: NO = 576 for so2_conv_1, 112 for EdgeDegreeEmbedding).
// The reference runs it as 3 addmm + 2 LayerNorm + 2 SiLU kernels per call and autograd keeps every intermediate;
// here forward is one kernel that keeps only the two 16-wide pre-norm activations per edge, and backward is one
// kernel that recomputes the rest and emits all parameter gradients as per-CTA partials (reduced in fixed order
// by the caller: deterministic, no atomics).
//
// HBM-bound: per edge forward reads K0*4 B and writes (NO + 32)*4 B; backward reads (NO + K0 + 32)*4 B and
// writes K0*4 B.
#include "efl_common.cuh"

namespace efl {

constexpr int RH = 16;        // hidden width
constexpr int RT = 32;        // edges per tile
constexpr int RTHREADS = 256;
constexpr float kRadLnEps = 1e-5f;

__device__ __forceinline__ float half_sum(float v) {  // sum over the 16 lanes of a half-warp
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct RadialParams {
  const float *W1, *b1, *g1, *be1, *W2, *b2, *g2, *be2, *W3, *b3;
};

// hidden part for one edge on a half-warp (lane j = hidden unit j): x row in shared memory.
// Returns a2[j]; optionally the pre-norm activations and the normalised values needed by backward.
struct Hidden {
  float h1, xh1, r1, y1, a1, h2, xh2, r2, y2, a2;
};
__device__ __forceinline__ Hidden radial_hidden(const float* __restrict__ xs, int K0, const float* __restrict__ sW1,
                                                const float* __restrict__ sW2, const RadialParams& p, int j,
                                                float* __restrict__ a1s /* [16] shared scratch of this half-warp */) {
  Hidden h;
  float acc = __ldg(p.b1 + j);
  for (int i = 0; i < K0; ++i) acc = fmaf(xs[i], sW1[j * K0 + i], acc);
  h.h1 = acc;
  const float m1 = half_sum(acc) * (1.f / RH);
  const float d1 = acc - m1;
  h.r1 = rsqrtf(half_sum(d1 * d1) * (1.f / RH) + kRadLnEps);
  h.xh1 = d1 * h.r1;
  h.y1 = h.xh1 * __ldg(p.g1 + j) + __ldg(p.be1 + j);
  h.a1 = siluf_(h.y1);
  a1s[j] = h.a1;
  __syncwarp();
  float acc2 = __ldg(p.b2 + j);
#pragma unroll
  for (int i = 0; i < RH; ++i) acc2 = fmaf(a1s[i], sW2[j * RH + i], acc2);
  h.h2 = acc2;
  const float m2 = half_sum(acc2) * (1.f / RH);
  const float d2 = acc2 - m2;
  h.r2 = rsqrtf(half_sum(d2 * d2) * (1.f / RH) + kRadLnEps);
  h.xh2 = d2 * h.r2;
  h.y2 = h.xh2 * __ldg(p.g2 + j) + __ldg(p.be2 + j);
  h.a2 = siluf_(h.y2);
  __syncwarp();
  return h;
}

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(RTHREADS) radial_fwd_kernel(const float* __restrict__ x, int K0, RadialParams p,
                                                              int NO, float* __restrict__ out, int E) {
  extern __shared__ float sm[];
  float* sW1 = sm;                    // [16][K0]
  float* sW2 = sW1 + RH * K0;         // [16][16]
  float* sx = sW2 + RH * RH;          // [RT][K0]
  float* sa2 = sx + RT * K0;          // [RT][16]
  float* sa1 = sa2 + RT * RH;         // [16 half-warps][16]
  for (int i = threadIdx.x; i < RH * K0; i += RTHREADS) sW1[i] = p.W1[i];
  for (int i = threadIdx.x; i < RH * RH; i += RTHREADS) sW2[i] = p.W2[i];
  const int hw = threadIdx.x >> 4, j = threadIdx.x & 15;
  const int ntiles = (E + RT - 1) / RT;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int e0 = tile * RT;
    const int ne = min(RT, E - e0);
    __syncthreads();
    for (int i = threadIdx.x; i < ne * K0; i += RTHREADS) sx[i] = x[(size_t)e0 * K0 + i];
    __syncthreads();
    for (int r = 0; r < RT / 16; ++r) {
      const int el = r * 16 + hw;
      const bool ok = el < ne;
      const Hidden h = radial_hidden(sx + (ok ? el : 0) * K0, K0, sW1, sW2, p, j, sa1 + hw * RH);
      sa2[el * RH + j] = h.a2;
    }
    __syncthreads();
    // last layer: thread = output column (strided by the block size); W3 row in registers
    for (int c = threadIdx.x; c < NO; c += RTHREADS) {
      float w[RH];
#pragma unroll
      for (int k = 0; k < RH; k += 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(p.W3 + (size_t)c * RH + k));
        w[k] = t.x; w[k + 1] = t.y; w[k + 2] = t.z; w[k + 3] = t.w;
      }
      const float b = __ldg(p.b3 + c);
      for (int el = 0; el < ne; ++el) {
        float acc = b;
#pragma unroll
        for (int k = 0; k < RH; k += 4) {
          const float4 a = *reinterpret_cast<const float4*>(sa2 + el * RH + k);
          acc = fmaf(a.x, w[k], acc);
          acc = fmaf(a.y, w[k + 1], acc);
          acc = fmaf(a.z, w[k + 2], acc);
          acc = fmaf(a.w, w[k + 3], acc);
        }
        out[(size_t)(e0 + el) * NO + c] = acc;
      }
    }
  }
}

}  // namespace efl

namespace efl {

// ------------------------------------------------------------------------------------------
// backward.  Per-CTA partial layout (floats):
//   dW3 [NO,16] | db3 [NO] | dW2 [16,16] | db2 | dg2 | dbe2 [16 each] | dW1 [16,K0] | db1 | dg1 | dbe1 [16 each]
// ------------------------------------------------------------------------------------------
template <int K0>
__global__ void __launch_bounds__(RTHREADS) radial_bwd_kernel(const float* __restrict__ x, RadialParams p, int NO,
                                                              const float* __restrict__ dout, float* __restrict__ dx,
                                                              float* __restrict__ part, int E) {
  constexpr int MAXS = 3;                 // output columns per thread (NO <= 768)
  constexpr int HWSZ = RH * (RH + K0 + 6);  // per half-warp parameter-gradient floats
  extern __shared__ float sm[];
  const int SDS = NO + 4;                 // padded row pitch of the d(out) tile (breaks the 2-way bank conflict)
  float* sW1 = sm;                        // [16][K0]
  float* sW2 = sW1 + RH * K0;             // [16][16]
  float* sx = sW2 + RH * RH;              // [RT][K0]
  float* sa2 = sx + RT * K0;              // [RT][16]
  float* sa1 = sa2 + RT * RH;             // [RT][16]
  float* sdh = sa1 + RT * RH;             // [16 half-warps][16]
  float* sd = sdh + 16 * RH;              // [RT][SDS]  (re-used for the final reduction)
  for (int i = threadIdx.x; i < RH * K0; i += RTHREADS) sW1[i] = p.W1[i];
  for (int i = threadIdx.x; i < RH * RH; i += RTHREADS) sW2[i] = p.W2[i];
  const int hw = threadIdx.x >> 4, j = threadIdx.x & 15;
  const float g1 = __ldg(p.g1 + j), g2 = __ldg(p.g2 + j);
  float dW3[MAXS][RH], db3[MAXS];
#pragma unroll
  for (int s = 0; s < MAXS; ++s) {
    db3[s] = 0.f;
#pragma unroll
    for (int k = 0; k < RH; ++k) dW3[s][k] = 0.f;
  }
  float dW2[RH], dW1[K0], db2 = 0.f, dg2 = 0.f, dbe2 = 0.f, db1 = 0.f, dg1 = 0.f, dbe1 = 0.f;
#pragma unroll
  for (int i = 0; i < RH; ++i) dW2[i] = 0.f;
#pragma unroll
  for (int i = 0; i < K0; ++i) dW1[i] = 0.f;

  const int ntiles = (E + RT - 1) / RT;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int e0 = tile * RT;
    const int ne = min(RT, E - e0);
    __syncthreads();
    for (int i = threadIdx.x; i < ne * K0; i += RTHREADS) sx[i] = x[(size_t)e0 * K0 + i];
    for (int i = threadIdx.x; i < ne * (NO / 4); i += RTHREADS) {
      const int el = i / (NO / 4), c4 = i % (NO / 4);
      *reinterpret_cast<float4*>(sd + el * SDS + 4 * c4) =
          __ldg(reinterpret_cast<const float4*>(dout + (size_t)(e0 + el) * NO) + c4);
    }
    __syncthreads();
    Hidden hid[RT / 16];
#pragma unroll
    for (int r = 0; r < RT / 16; ++r) {
      const int el = r * 16 + hw;
      hid[r] = radial_hidden(sx + (el < ne ? el : 0) * K0, K0, sW1, sW2, p, j, sa1 + el * RH);
      sa2[el * RH + j] = hid[r].a2;
    }
    __syncthreads();
    // ---- last layer: dW3, db3 (thread = output column)
#pragma unroll
    for (int s = 0; s < MAXS; ++s) {
      const int c = threadIdx.x + s * RTHREADS;
      if (c < NO) {
        for (int el = 0; el < ne; ++el) {
          const float d = sd[el * SDS + c];
          db3[s] += d;
#pragma unroll
          for (int k = 0; k < RH; k += 4) {
            const float4 a = *reinterpret_cast<const float4*>(sa2 + el * RH + k);
            dW3[s][k] = fmaf(d, a.x, dW3[s][k]);
            dW3[s][k + 1] = fmaf(d, a.y, dW3[s][k + 1]);
            dW3[s][k + 2] = fmaf(d, a.z, dW3[s][k + 2]);
            dW3[s][k + 3] = fmaf(d, a.w, dW3[s][k + 3]);
          }
        }
      }
    }
    // ---- hidden layers: half-warp per edge, lane = hidden unit
#pragma unroll
    for (int r = 0; r < RT / 16; ++r) {
      const int el = r * 16 + hw;
      const bool ok = el < ne;
      const Hidden& h = hid[r];
      // d a2[j] = sum_c dout[e, c] W3[c, j]
      float da2 = 0.f;
      {
        const float* drow = sd + (ok ? el : 0) * SDS;
        float t0 = 0.f, t1 = 0.f, t2 = 0.f, t3 = 0.f;
        for (int c = 0; c < NO; c += 4) {
          const float4 d = *reinterpret_cast<const float4*>(drow + c);
          t0 = fmaf(d.x, __ldg(p.W3 + (size_t)(c + 0) * RH + j), t0);
          t1 = fmaf(d.y, __ldg(p.W3 + (size_t)(c + 1) * RH + j), t1);
          t2 = fmaf(d.z, __ldg(p.W3 + (size_t)(c + 2) * RH + j), t2);
          t3 = fmaf(d.w, __ldg(p.W3 + (size_t)(c + 3) * RH + j), t3);
        }
        da2 = ok ? (t0 + t1) + (t2 + t3) : 0.f;
      }
      // SiLU, LayerNorm 2
      const float dz2 = da2 * dsiluf_(h.y2);
      dg2 = fmaf(dz2, h.xh2, dg2);
      dbe2 += dz2;
      const float dxh2 = dz2 * g2;
      const float m21 = half_sum(dxh2) * (1.f / RH);
      const float m22 = half_sum(dxh2 * h.xh2) * (1.f / RH);
      const float dh2 = h.r2 * (dxh2 - m21 - h.xh2 * m22);
      db2 += dh2;
      const float* a1row = sa1 + el * RH;
#pragma unroll
      for (int i = 0; i < RH; ++i) dW2[i] = fmaf(dh2, a1row[i], dW2[i]);
      sdh[hw * RH + j] = dh2;
      __syncwarp();
      float da1 = 0.f;
#pragma unroll
      for (int jj = 0; jj < RH; ++jj) da1 = fmaf(sdh[hw * RH + jj], sW2[jj * RH + j], da1);
      __syncwarp();
      // SiLU, LayerNorm 1
      const float dz1 = da1 * dsiluf_(h.y1);
      dg1 = fmaf(dz1, h.xh1, dg1);
      dbe1 += dz1;
      const float dxh1 = dz1 * g1;
      const float m11 = half_sum(dxh1) * (1.f / RH);
      const float m12 = half_sum(dxh1 * h.xh1) * (1.f / RH);
      const float dh1 = h.r1 * (dxh1 - m11 - h.xh1 * m12);
      db1 += dh1;
      const float* xrow = sx + (ok ? el : 0) * K0;
#pragma unroll
      for (int i = 0; i < K0; ++i) dW1[i] = fmaf(dh1, xrow[i], dW1[i]);
      sdh[hw * RH + j] = dh1;
      __syncwarp();
#pragma unroll
      for (int t = 0; t < K0 / 16; ++t) {
        const int i = j + 16 * t;
        float acc = 0.f;
#pragma unroll
        for (int jj = 0; jj < RH; ++jj) acc = fmaf(sdh[hw * RH + jj], sW1[jj * K0 + i], acc);
        if (ok) dx[(size_t)(e0 + el) * K0 + i] = acc;
      }
      __syncwarp();
    }
  }
  // ---- per-CTA partials
  const int PSZ = NO * (RH + 1) + RH * RH + 3 * RH + RH * K0 + 3 * RH;
  float* my = part + (size_t)blockIdx.x * PSZ;
#pragma unroll
  for (int s = 0; s < MAXS; ++s) {
    const int c = threadIdx.x + s * RTHREADS;
    if (c < NO) {
#pragma unroll
      for (int k = 0; k < RH; k += 4)
        *reinterpret_cast<float4*>(my + (size_t)c * RH + k) = make_float4(dW3[s][k], dW3[s][k + 1], dW3[s][k + 2], dW3[s][k + 3]);
      my[(size_t)NO * RH + c] = db3[s];
    }
  }
  __syncthreads();
  // half-warp accumulators -> shared [16 hw][HWSZ] -> fixed-order sum over the 16 half-warps
  float* red = sd;
  {
    float* m = red + hw * HWSZ;
#pragma unroll
    for (int i = 0; i < RH; ++i) m[j * RH + i] = dW2[i];
    m[RH * RH + j] = db2;
    m[RH * RH + RH + j] = dg2;
    m[RH * RH + 2 * RH + j] = dbe2;
    float* m1 = m + RH * RH + 3 * RH;
#pragma unroll
    for (int i = 0; i < K0; ++i) m1[j * K0 + i] = dW1[i];
    m1[RH * K0 + j] = db1;
    m1[RH * K0 + RH + j] = dg1;
    m1[RH * K0 + 2 * RH + j] = dbe1;
  }
  __syncthreads();
  float* tail = my + (size_t)NO * (RH + 1);
  for (int i = threadIdx.x; i < HWSZ; i += RTHREADS) {
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) acc += red[k * HWSZ + i];
    tail[i] = acc;
  }
}

}  // namespace efl

using namespace efl;

static size_t radial_fwd_smem(int K0) { return sizeof(float) * (RH * K0 + RH * RH + RT * K0 + RT * RH + 16 * RH); }
static size_t radial_bwd_smem(int K0, int NO) {
  size_t sd = (size_t)RT * (NO + 4);
  const size_t red = (size_t)16 * RH * (RH + K0 + 6);
  if (red > sd) sd = red;
  return sizeof(float) * (RH * K0 + RH * RH + RT * K0 + 2 * RT * RH + 16 * RH + sd);
}

extern "C" int efl_radial_mlp_part_size(int K0, int NO) { return NO * (RH + 1) + RH * RH + 3 * RH + RH * K0 + 3 * RH; }

// x [E,K0] -> out [E,NO].  Weights in the nn.Linear / nn.LayerNorm layouts of RadialFunction.net.{0,1,3,4,6}.
extern "C" int efl_radial_mlp_fwd(const float* x, int K0, const float* W1, const float* b1, const float* g1,
                                  const float* be1, const float* W2, const float* b2, const float* g2,
                                  const float* be2, const float* W3, const float* b3, int NO, float* out, int E,
                                  void* stream) {
  if (E <= 0) return EFL_OK;
  EFL_REQUIRE(K0 > 0 && K0 <= 64 && K0 % 4 == 0 && NO % 4 == 0, "efl_radial_mlp_fwd: K0 <= 64, K0 % 4 == 0, NO % 4 == 0");
  RadialParams p{W1, b1, g1, be1, W2, b2, g2, be2, W3, b3};
  const int ntiles = (E + RT - 1) / RT;
  const int grid = ntiles < 148 * 4 ? ntiles : 148 * 4;
  radial_fwd_kernel<<<grid, RTHREADS, radial_fwd_smem(K0), (cudaStream_t)stream>>>(x, K0, p, NO, out, E);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// dout [E,NO] -> dx [E,K0] and part [nblocks, efl_radial_mlp_part_size] (per-CTA parameter-gradient partials).
extern "C" int efl_radial_mlp_bwd(const float* x, int K0, const float* W1, const float* b1, const float* g1,
                                  const float* be1, const float* W2, const float* b2, const float* g2,
                                  const float* be2, const float* W3, const float* b3, int NO, const float* dout,
                                  float* dx, float* part, int nblocks, int E, void* stream) {
  EFL_REQUIRE(nblocks > 0, "efl_radial_mlp_bwd: nblocks must be positive");
  EFL_REQUIRE((K0 == 48 || K0 == 16 || K0 == 32 || K0 == 64) && NO % 4 == 0 && NO <= 768,
              "efl_radial_mlp_bwd: K0 in {16,32,48,64}, NO % 4 == 0, NO <= 768");
  RadialParams p{W1, b1, g1, be1, W2, b2, g2, be2, W3, b3};
  const size_t smem = radial_bwd_smem(K0, NO);
  cudaStream_t s = (cudaStream_t)stream;
#define RLAUNCH(KK)                                                                                         \
  {                                                                                                         \
    static bool done = false;                                                                               \
    if (!done) {                                                                                            \
      cudaError_t e = cudaFuncSetAttribute(radial_bwd_kernel<KK>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                           (int)radial_bwd_smem(KK, 768));                                  \
      if (e != cudaSuccess) {                                                                               \
        efl_set_error(cudaGetErrorString(e));                                                               \
        return (int)e;                                                                                      \
      }                                                                                                     \
      done = true;                                                                                          \
    }                                                                                                       \
    radial_bwd_kernel<KK><<<nblocks, RTHREADS, smem, s>>>(x, p, NO, dout, dx, part, E);                     \
  }
  if (K0 == 48) RLAUNCH(48)
  else if (K0 == 16) RLAUNCH(16)
  else if (K0 == 32) RLAUNCH(32)
  else RLAUNCH(64)
#undef RLAUNCH
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
