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
__device__ __forceinline__ Hidden radial_hidden(const float* __restrict__ xs, int K0, const float* __restrict__ sW1t,
                                                const float* __restrict__ sW2t, const RadialParams& p, int j,
                                                float* __restrict__ a1s /* [16] shared scratch of this half-warp */) {
  Hidden h;
  float acc = __ldg(p.b1 + j);
  for (int i = 0; i < K0; ++i) acc = fmaf(xs[i], sW1t[i * RH + j], acc);  // [K0][16]: conflict-free
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
  for (int i = 0; i < RH; ++i) acc2 = fmaf(a1s[i], sW2t[i * RH + j], acc2);
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
// a2_out != nullptr: hidden layers only, a2 written as bf16 [E,16] (the last layer then runs as a K=16 tcgen05 GEMM)
__global__ void __launch_bounds__(RTHREADS) radial_fwd_kernel(const float* __restrict__ x, int K0, RadialParams p,
                                                              int NO, float* __restrict__ out,
                                                              __nv_bfloat16* __restrict__ a2_out, int E) {
  extern __shared__ float sm[];
  float* sW1 = sm;                    // [K0][16]  (transposed: lane = hidden unit reads consecutive banks)
  float* sW2 = sW1 + RH * K0;         // [16][16]  (transposed)
  float* sx = sW2 + RH * RH;          // [RT][K0]
  float* sa2 = sx + RT * K0;          // [RT][16]
  float* sa1 = sa2 + RT * RH;         // [16 half-warps][16]
  for (int i = threadIdx.x; i < RH * K0; i += RTHREADS) sW1[(i % K0) * RH + i / K0] = p.W1[i];
  for (int i = threadIdx.x; i < RH * RH; i += RTHREADS) sW2[(i % RH) * RH + i / RH] = p.W2[i];
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
      const Hidden h = radial_hidden(sx + (ok ? el : 0) * K0, K0, sW1, sW2, p, j, sa1 + hw * RH);  // sW1/sW2 hold the transposes here
      sa2[el * RH + j] = h.a2;
      if (a2_out != nullptr && ok) a2_out[(size_t)(e0 + el) * RH + j] = __float2bfloat16_rn(h.a2);
    }
    if (a2_out != nullptr) continue;
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
// Per tile of RT edges:
//   A  recompute the hidden activations (half-warp per edge, lane = hidden unit)
//   B  dW3 / db3: thread = output column, a2 row of the edge loaded once for the thread's columns
//   C  d a2 = dout . W3: half-warp owns a chunk of columns with its W3 entries in registers and sweeps the tile's
//      edges (4 FMAs per shared-memory load); chunk partials combined in fixed order
//   D  LayerNorm/SiLU backward per edge; d h2, d h1 parked in shared memory; d x written
//   E  dW2 += dh2^T a1, dW1 += dh1^T x as tile-level outer products (thread = weight element)
// ------------------------------------------------------------------------------------------
template <int K0>
__global__ void __launch_bounds__(RTHREADS, 2) radial_bwd_kernel(const float* __restrict__ x, RadialParams p, int NO,
                                                              const float* __restrict__ dout, float* __restrict__ dx,
                                                              float* __restrict__ part, int E) {
  constexpr int MAXS = 3;                 // output columns per thread (NO <= 768)
  constexpr int CWMAX = 48;               // 768 / 16
  constexpr int NW1 = (RH * K0 + RTHREADS - 1) / RTHREADS;  // dW1 elements per thread
  constexpr int NWARP = RTHREADS / 32;
  extern __shared__ float sm[];
  const int SDS = NO + 4;                 // padded row pitch of the d(out) tile
  float* sW1 = sm;                        // [16][K0]
  float* sW2 = sW1 + RH * K0;             // [16][16]
  float* sx = sW2 + RH * RH;              // [RT][K0]
  float* sa2 = sx + RT * K0;              // [RT][16]
  float* sa1 = sa2 + RT * RH;             // [RT][16]
  float* sdh = sa1 + RT * RH;             // [16 half-warps][16] exchange scratch
  float* sdh2 = sdh + 16 * RH;            // [RT][16]
  float* sdh1 = sdh2 + RT * RH;           // [RT][16]
  float* spart = sdh1 + RT * RH;          // [8 warps][RT/2][16] chunk partials of d a2
  float* sda2 = spart + NWARP * (RT / 2) * RH;  // [RT][16]
  float* sW1t = sda2 + RT * RH;           // [K0][16] transposed copies for the forward recompute
  float* sW2t = sW1t + RH * K0;           // [16][16]
  float* sd = sW2t + RH * RH;             // [RT][SDS]
  const int CW = (((NO + 15) / 16) + 3) & ~3;  // columns per half-warp chunk (multiple of 4, <= CWMAX)
  for (int i = threadIdx.x; i < RH * K0; i += RTHREADS) {
    const float v = p.W1[i];
    sW1[i] = v;
    sW1t[(i % K0) * RH + i / K0] = v;
  }
  for (int i = threadIdx.x; i < RH * RH; i += RTHREADS) {
    const float v = p.W2[i];
    sW2[i] = v;
    sW2t[(i % RH) * RH + i / RH] = v;
  }
  const int hw = threadIdx.x >> 4, j = threadIdx.x & 15;
  const float g1 = __ldg(p.g1 + j), g2 = __ldg(p.g2 + j);
  float dW3[MAXS][RH], db3[MAXS];
#pragma unroll
  for (int s = 0; s < MAXS; ++s) {
    db3[s] = 0.f;
#pragma unroll
    for (int k = 0; k < RH; ++k) dW3[s][k] = 0.f;
  }
  float dW1[NW1], dW2 = 0.f, db2 = 0.f, dg2 = 0.f, dbe2 = 0.f, db1 = 0.f, dg1 = 0.f, dbe1 = 0.f;
#pragma unroll
  for (int q = 0; q < NW1; ++q) dW1[q] = 0.f;

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
    // ---- A
    Hidden hid[RT / 16];
#pragma unroll
    for (int r = 0; r < RT / 16; ++r) {
      const int el = r * 16 + hw;
      hid[r] = radial_hidden(sx + (el < ne ? el : 0) * K0, K0, sW1t, sW2t, p, j, sa1 + el * RH);
      sa2[el * RH + j] = hid[r].a2;
    }
    __syncthreads();
    // ---- B
    for (int el = 0; el < ne; ++el) {
      float a[RH];
#pragma unroll
      for (int k = 0; k < RH; k += 4) {
        const float4 t = *reinterpret_cast<const float4*>(sa2 + el * RH + k);
        a[k] = t.x; a[k + 1] = t.y; a[k + 2] = t.z; a[k + 3] = t.w;
      }
#pragma unroll
      for (int s = 0; s < MAXS; ++s) {
        const int c = threadIdx.x + s * RTHREADS;
        if (c < NO) {
          const float d = sd[el * SDS + c];
          db3[s] += d;
#pragma unroll
          for (int k = 0; k < RH; ++k) dW3[s][k] = fmaf(d, a[k], dW3[s][k]);
        }
      }
    }
    // ---- C
    {
      const int c0 = hw * CW;
      float w3[CWMAX];
#pragma unroll
      for (int cc = 0; cc < CWMAX; ++cc)
        w3[cc] = (cc < CW && c0 + cc < NO) ? __ldg(p.W3 + (size_t)(c0 + cc) * RH + j) : 0.f;
      // two sub-batches of RT/2 edges share one [8 warps][RT/2][16] partial buffer
      for (int sb = 0; sb < 2; ++sb) {
        const int el0 = sb * (RT / 2), el1 = min(ne, el0 + RT / 2);
        for (int el = el0; el < el1; ++el) {
          const float* drow = sd + el * SDS + c0;
          float t0 = 0.f, t1 = 0.f, t2 = 0.f, t3 = 0.f;
#pragma unroll
          for (int cc = 0; cc < CWMAX; cc += 4) {
            if (cc < CW && c0 + cc < NO) {
              const float4 d = *reinterpret_cast<const float4*>(drow + cc);
              t0 = fmaf(d.x, w3[cc], t0);
              t1 = fmaf(d.y, w3[cc + 1], t1);
              t2 = fmaf(d.z, w3[cc + 2], t2);
              t3 = fmaf(d.w, w3[cc + 3], t3);
            }
          }
          float acc = (t0 + t1) + (t2 + t3);
          acc += __shfl_xor_sync(0xffffffffu, acc, 16);
          if ((threadIdx.x & 16) == 0) spart[((threadIdx.x >> 5) * (RT / 2) + (el - el0)) * RH + j] = acc;
        }
        __syncthreads();
        for (int t = threadIdx.x; t < (el1 - el0) * RH; t += RTHREADS) {
          float acc = 0.f;
#pragma unroll
          for (int w = 0; w < NWARP; ++w) acc += spart[w * (RT / 2) * RH + t];
          sda2[el0 * RH + t] = acc;
        }
        __syncthreads();
      }
    }
    // ---- D
#pragma unroll
    for (int r = 0; r < RT / 16; ++r) {
      const int el = r * 16 + hw;
      const bool ok = el < ne;
      const Hidden& h = hid[r];
      const float da2 = ok ? sda2[el * RH + j] : 0.f;
      const float dz2 = da2 * dsiluf_(h.y2);
      dg2 = fmaf(dz2, h.xh2, dg2);
      dbe2 += dz2;
      const float dxh2 = dz2 * g2;
      const float m21 = half_sum(dxh2) * (1.f / RH);
      const float m22 = half_sum(dxh2 * h.xh2) * (1.f / RH);
      const float dh2 = h.r2 * (dxh2 - m21 - h.xh2 * m22);
      db2 += dh2;
      sdh2[el * RH + j] = dh2;
      sdh[hw * RH + j] = dh2;
      __syncwarp();
      float da1 = 0.f;
#pragma unroll
      for (int jj = 0; jj < RH; ++jj) da1 = fmaf(sdh[hw * RH + jj], sW2[jj * RH + j], da1);
      __syncwarp();
      const float dz1 = da1 * dsiluf_(h.y1);
      dg1 = fmaf(dz1, h.xh1, dg1);
      dbe1 += dz1;
      const float dxh1 = dz1 * g1;
      const float m11 = half_sum(dxh1) * (1.f / RH);
      const float m12 = half_sum(dxh1 * h.xh1) * (1.f / RH);
      const float dh1 = h.r1 * (dxh1 - m11 - h.xh1 * m12);
      db1 += dh1;
      sdh1[el * RH + j] = dh1;
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
    __syncthreads();
    // ---- E  (rows el >= ne of sdh2/sdh1 hold zeros: d a2 = 0 there)
    {
      const int jj = threadIdx.x >> 4, i = threadIdx.x & 15;
      float acc = 0.f;
      for (int el = 0; el < RT; ++el) acc = fmaf(sdh2[el * RH + jj], sa1[el * RH + i], acc);
      dW2 += acc;
    }
#pragma unroll
    for (int q = 0; q < NW1; ++q) {
      const int idx = threadIdx.x + q * RTHREADS;
      if (idx < RH * K0) {
        const int jj = idx / K0, i = idx % K0;
        float acc = 0.f;
        for (int el = 0; el < ne; ++el) acc = fmaf(sdh1[el * RH + jj], sx[el * K0 + i], acc);
        dW1[q] += acc;
      }
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
  float* tail = my + (size_t)NO * (RH + 1);
  tail[threadIdx.x] = dW2;                       // [16][16], thread = jj*16 + i
  float* t1 = tail + RH * RH + 3 * RH;
#pragma unroll
  for (int q = 0; q < NW1; ++q) {
    const int idx = threadIdx.x + q * RTHREADS;
    if (idx < RH * K0) t1[idx] = dW1[q];
  }
  // the six 16-vectors: half-warp accumulators -> shared [16 hw][6][16] -> fixed-order sum
  __syncthreads();
  float* red = sd;
  red[(hw * 6 + 0) * RH + j] = db2;
  red[(hw * 6 + 1) * RH + j] = dg2;
  red[(hw * 6 + 2) * RH + j] = dbe2;
  red[(hw * 6 + 3) * RH + j] = db1;
  red[(hw * 6 + 4) * RH + j] = dg1;
  red[(hw * 6 + 5) * RH + j] = dbe1;
  __syncthreads();
  if (threadIdx.x < 6 * RH) {
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) acc += red[k * 6 * RH + threadIdx.x];
    const int v = threadIdx.x / RH, jj = threadIdx.x % RH;
    float* dst = v < 3 ? tail + RH * RH + v * RH : t1 + RH * K0 + (v - 3) * RH;
    dst[jj] = acc;
  }
}

}  // namespace efl


namespace efl {

// ------------------------------------------------------------------------------------------
// hidden-layer backward (bf16 mode): d a2 [E,16] fp32 comes from the tcgen05 GEMM dout . W3; this kernel does phases
// A, D, E of radial_bwd_kernel only.  Per-CTA partial layout (floats):
//   dW2 [16,16] | db2 | dg2 | dbe2 [16 each] | dW1 [16,K0] | db1 | dg1 | dbe1 [16 each]
// ------------------------------------------------------------------------------------------
template <int K0>
__global__ void __launch_bounds__(RTHREADS) radial_hidden_bwd_kernel(const float* __restrict__ x, RadialParams p,
                                                                     const float* __restrict__ da2_in,
                                                                     float* __restrict__ dx, float* __restrict__ part,
                                                                     int E) {
  constexpr int NW1 = (RH * K0 + RTHREADS - 1) / RTHREADS;
  extern __shared__ float sm[];
  float* sW1 = sm;                        // [16][K0]
  float* sW2 = sW1 + RH * K0;             // [16][16]
  float* sW1t = sW2 + RH * RH;            // [K0][16]
  float* sW2t = sW1t + RH * K0;           // [16][16]
  float* sx = sW2t + RH * RH;             // [RT][K0]
  float* sa1 = sx + RT * K0;              // [RT][16]
  float* sdh = sa1 + RT * RH;             // [16 half-warps][16]
  float* sdh2 = sdh + 16 * RH;            // [RT][16]
  float* sdh1 = sdh2 + RT * RH;           // [RT][16]
  float* sda2 = sdh1 + RT * RH;           // [RT][16]  (re-used for the final reduction: 16*6*16 floats fit in sda2..)
  for (int i = threadIdx.x; i < RH * K0; i += RTHREADS) {
    const float v = p.W1[i];
    sW1[i] = v;
    sW1t[(i % K0) * RH + i / K0] = v;
  }
  for (int i = threadIdx.x; i < RH * RH; i += RTHREADS) {
    const float v = p.W2[i];
    sW2[i] = v;
    sW2t[(i % RH) * RH + i / RH] = v;
  }
  const int hw = threadIdx.x >> 4, j = threadIdx.x & 15;
  const float g1 = __ldg(p.g1 + j), g2 = __ldg(p.g2 + j);
  float dW1[NW1], dW2 = 0.f, db2 = 0.f, dg2 = 0.f, dbe2 = 0.f, db1 = 0.f, dg1 = 0.f, dbe1 = 0.f;
#pragma unroll
  for (int q = 0; q < NW1; ++q) dW1[q] = 0.f;
  const int ntiles = (E + RT - 1) / RT;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int e0 = tile * RT;
    const int ne = min(RT, E - e0);
    __syncthreads();
    for (int i = threadIdx.x; i < ne * K0; i += RTHREADS) sx[i] = x[(size_t)e0 * K0 + i];
    for (int i = threadIdx.x; i < RT * RH; i += RTHREADS) sda2[i] = i < ne * RH ? da2_in[(size_t)e0 * RH + i] : 0.f;
    __syncthreads();
#pragma unroll
    for (int r = 0; r < RT / 16; ++r) {
      const int el = r * 16 + hw;
      const bool ok = el < ne;
      const Hidden h = radial_hidden(sx + (ok ? el : 0) * K0, K0, sW1t, sW2t, p, j, sa1 + el * RH);
      const float da2 = ok ? sda2[el * RH + j] : 0.f;
      const float dz2 = da2 * dsiluf_(h.y2);
      dg2 = fmaf(dz2, h.xh2, dg2);
      dbe2 += dz2;
      const float dxh2 = dz2 * g2;
      const float m21 = half_sum(dxh2) * (1.f / RH);
      const float m22 = half_sum(dxh2 * h.xh2) * (1.f / RH);
      const float dh2 = h.r2 * (dxh2 - m21 - h.xh2 * m22);
      db2 += dh2;
      sdh2[el * RH + j] = dh2;
      sdh[hw * RH + j] = dh2;
      __syncwarp();
      float da1 = 0.f;
#pragma unroll
      for (int jj = 0; jj < RH; ++jj) da1 = fmaf(sdh[hw * RH + jj], sW2[jj * RH + j], da1);
      __syncwarp();
      const float dz1 = da1 * dsiluf_(h.y1);
      dg1 = fmaf(dz1, h.xh1, dg1);
      dbe1 += dz1;
      const float dxh1 = dz1 * g1;
      const float m11 = half_sum(dxh1) * (1.f / RH);
      const float m12 = half_sum(dxh1 * h.xh1) * (1.f / RH);
      const float dh1 = h.r1 * (dxh1 - m11 - h.xh1 * m12);
      db1 += dh1;
      sdh1[el * RH + j] = dh1;
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
    __syncthreads();
    {
      const int jj = threadIdx.x >> 4, i = threadIdx.x & 15;
      float acc = 0.f;
      for (int el = 0; el < RT; ++el) acc = fmaf(sdh2[el * RH + jj], sa1[el * RH + i], acc);
      dW2 += acc;
    }
#pragma unroll
    for (int q = 0; q < NW1; ++q) {
      const int idx = threadIdx.x + q * RTHREADS;
      if (idx < RH * K0) {
        const int jj = idx / K0, i = idx % K0;
        float acc = 0.f;
        for (int el = 0; el < ne; ++el) acc = fmaf(sdh1[el * RH + jj], sx[el * K0 + i], acc);
        dW1[q] += acc;
      }
    }
  }
  const int PSZ = RH * RH + 3 * RH + RH * K0 + 3 * RH;
  float* tail = part + (size_t)blockIdx.x * PSZ;
  tail[threadIdx.x] = dW2;
  float* t1 = tail + RH * RH + 3 * RH;
#pragma unroll
  for (int q = 0; q < NW1; ++q) {
    const int idx = threadIdx.x + q * RTHREADS;
    if (idx < RH * K0) t1[idx] = dW1[q];
  }
  __syncthreads();
  float* red = sx;  // RT*K0 >= 16*6*16 for K0 >= 48; smaller K0: host sizes sx accordingly
  red[(hw * 6 + 0) * RH + j] = db2;
  red[(hw * 6 + 1) * RH + j] = dg2;
  red[(hw * 6 + 2) * RH + j] = dbe2;
  red[(hw * 6 + 3) * RH + j] = db1;
  red[(hw * 6 + 4) * RH + j] = dg1;
  red[(hw * 6 + 5) * RH + j] = dbe1;
  __syncthreads();
  if (threadIdx.x < 6 * RH) {
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) acc += red[k * 6 * RH + threadIdx.x];
    const int v = threadIdx.x / RH, jj = threadIdx.x % RH;
    float* dst = v < 3 ? tail + RH * RH + v * RH : t1 + RH * K0 + (v - 3) * RH;
    dst[jj] = acc;
  }
}

}  // namespace efl

using namespace efl;

static size_t radial_fwd_smem(int K0) { return sizeof(float) * (RH * K0 + RH * RH + RT * K0 + RT * RH + 16 * RH); }
static size_t radial_bwd_smem(int K0, int NO) {
  size_t sd = (size_t)RT * (NO + 4);
  if (sd < 16 * 6 * RH) sd = 16 * 6 * RH;
  return sizeof(float) * (2 * (RH * K0 + RH * RH) + RT * K0 + 4 * RT * RH + 16 * RH + (RTHREADS / 64 + 1) * RT * RH + sd);
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
  radial_fwd_kernel<<<grid, RTHREADS, radial_fwd_smem(K0), (cudaStream_t)stream>>>(x, K0, p, NO, out, nullptr, E);
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
    static unsigned long long done = 0;                                                                     \
    if (efl_first_use_on_device(&done)) {                                                                   \
      cudaError_t e = cudaFuncSetAttribute(radial_bwd_kernel<KK>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                           (int)radial_bwd_smem(KK, 768));                                  \
      if (e != cudaSuccess) {                                                                               \
        efl_set_error(cudaGetErrorString(e));                                                               \
        return (int)e;                                                                                      \
      }                                                                                                     \
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

// ---- bf16 mode: hidden layers only; the 16 -> NO layer and its backward run on efl_tc_gemm_nt / efl_tc_gemm_tn ----
extern "C" int efl_radial_hidden_part_size(int K0) { return RH * RH + 3 * RH + RH * K0 + 3 * RH; }

// x [E,K0] fp32 -> a2 [E,16] bf16 (activation entering the last Linear)
extern "C" int efl_radial_hidden_fwd(const float* x, int K0, const float* W1, const float* b1, const float* g1,
                                     const float* be1, const float* W2, const float* b2, const float* g2,
                                     const float* be2, void* a2_out, int E, void* stream) {
  if (E <= 0) return EFL_OK;
  EFL_REQUIRE(K0 > 0 && K0 <= 64 && K0 % 4 == 0, "efl_radial_hidden_fwd: K0 <= 64, K0 % 4 == 0");
  RadialParams p{W1, b1, g1, be1, W2, b2, g2, be2, nullptr, nullptr};
  const int ntiles = (E + RT - 1) / RT;
  const int grid = ntiles < 148 * 4 ? ntiles : 148 * 4;
  radial_fwd_kernel<<<grid, RTHREADS, radial_fwd_smem(K0), (cudaStream_t)stream>>>(x, K0, p, 0, nullptr,
                                                                                  (__nv_bfloat16*)a2_out, E);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// da2 [E,16] fp32 -> dx [E,K0] and part [nblocks, efl_radial_hidden_part_size(K0)]
extern "C" int efl_radial_hidden_bwd(const float* x, int K0, const float* W1, const float* b1, const float* g1,
                                     const float* be1, const float* W2, const float* b2, const float* g2,
                                     const float* be2, const float* da2, float* dx, float* part, int nblocks, int E,
                                     void* stream) {
  EFL_REQUIRE(nblocks > 0, "efl_radial_hidden_bwd: nblocks must be positive");
  EFL_REQUIRE(K0 == 48 || K0 == 16 || K0 == 32 || K0 == 64, "efl_radial_hidden_bwd: K0 in {16,32,48,64}");
  RadialParams p{W1, b1, g1, be1, W2, b2, g2, be2, nullptr, nullptr};
  const int sxf = RT * K0 > 16 * 6 * RH ? RT * K0 : 16 * 6 * RH;
  const size_t smem = sizeof(float) * (2 * (RH * K0 + RH * RH) + sxf + 4 * RT * RH + 16 * RH);
  cudaStream_t s = (cudaStream_t)stream;
  if (K0 == 48) radial_hidden_bwd_kernel<48><<<nblocks, RTHREADS, smem, s>>>(x, p, da2, dx, part, E);
  else if (K0 == 16) radial_hidden_bwd_kernel<16><<<nblocks, RTHREADS, smem, s>>>(x, p, da2, dx, part, E);
  else if (K0 == 32) radial_hidden_bwd_kernel<32><<<nblocks, RTHREADS, smem, s>>>(x, p, da2, dx, part, E);
  else radial_hidden_bwd_kernel<64><<<nblocks, RTHREADS, smem, s>>>(x, p, da2, dx, part, E);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
