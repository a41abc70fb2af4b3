// Edge-frame rotations (HBM-bound kernels).
//
//  rotate_in        : gather x[src] || x[dst], rotate into the edge frame with the kept
//                     Wigner rows, optional per-edge radial scale, emit m-primary GEMM operands.
//                     Replaces SO3_Embedding.clone/_expand_edge, cat, SO3_Rotation.rotate,
//                     _m_primary and the radial multiply of SO2_Convolution
//                     (reference EF_layers.py:1110-1139, 494-497, 354-355, 820-850).
//  rotate_back_agg  : alpha * value, rotate back (rescale folded in) and segment-sum per
//                     destination over the dst-sorted CSR.  Replaces EF_layers.py:1186-1199,
//                     500-505, 342-351.  Deterministic: one CTA owns a destination, fixed order,
//                     no atomics; the [E, L, heads*value] tensor never exists.
//  edge_degree      : EdgeDegreeEmbedding tail (EF_layers.py:116-147): only the m'=0 Wigner row.
//
// Per-edge operand layout ("edge rows, m-primary"): every edge owns ONE row of `pitch` elements; inside the row the
// kept coefficients sit in m-primary order, m = 0 block (LMAX+1 degrees), then for each order m > 0 the +m ("real")
// block followed by the -m ("imag") block, each (LMAX+1-m) * CH wide.  The three base pointers address the start of
// the m = 0, m = 1 and m = 2 blocks of edge 0 (they may carry a column offset: the m = 0 block of the rotate-in
// output is preceded by the bias/ones pad of the SO(2) GEMM).  Because (+m | -m) are adjacent, one edge row is
// directly the [x_re | x_im] operand of the SO(2) linear written as a real GEMM (ops.py: _Weights).
#include "efl_common.cuh"

namespace efl {

template <int LMAX>
struct OpBufs {
  void* p[3];
  long long pitch;   // elements between consecutive edges (the same for the three blocks: one row per edge)
};

// element (l, m') of local edge e, channel c
template <int LMAX, typename T>
__device__ __forceinline__ T* op_ptr(const OpBufs<LMAX>& b, int e, int l, int mp, int CH, int c) {
  const int am = mp < 0 ? -mp : mp;
  const int K = (LMAX + 1 - am) * CH;
  return reinterpret_cast<T*>(b.p[am]) + (size_t)e * b.pitch + (mp < 0 ? K : 0) + (l - am) * CH + c;
}

// ------------------------------------------------------------------------------------------
// rotate-in forward: one warp per edge, lane = channel of the src||dst concat (2C = 32)
// ------------------------------------------------------------------------------------------
template <int LMAX, int MMAX, int C, typename TOut>
__global__ void __launch_bounds__(128) rotate_in_fwd_kernel(
    const float* __restrict__ x_src, const float* __restrict__ x_dst, const int* __restrict__ src,
    const int* __restrict__ dst, const int* __restrict__ perm, const float* __restrict__ wig,
    const float* __restrict__ rad, int rad_ld, OpBufs<LMAX> out, int ones_pad, int e0, int Ec) {
  using T = Tab<LMAX, MMAX>;
  static_assert(2 * C == 32, "lane = channel mapping needs 2C == 32");
  constexpr int CH = 2 * C;
  static_assert(T::WSZ % 4 == 0, "Wigner rows are 16-byte padded");
  // 16-byte aligned rows: the broadcast reads of a row compile to LDS.128 (the LSU pipe, at 77 % busy with scalar
  // LDS, was the limiter of this kernel: ncu r1k)
  __shared__ __align__(16) float sW[4][T::WSZ];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int el = blockIdx.x * 4 + warp;
  if (el >= Ec) return;
  const int e = e0 + el;
  const int eo = perm[e];
  const float4* w4 = reinterpret_cast<const float4*>(wig + (size_t)eo * T::WSZ);
  for (int i = lane; i < T::WSZ / 4; i += 32) reinterpret_cast<float4*>(sW[warp])[i] = __ldg(w4 + i);
  const int node = lane < C ? src[e] : dst[e];
  const float* xp = (lane < C ? x_src : x_dst) + (size_t)node * T::L * C + (lane & (C - 1));
  float x[T::L];
#pragma unroll
  for (int i = 0; i < T::L; ++i) x[i] = __ldg(xp + i * C);
  __syncwarp();
  const float* rd = rad ? rad + (size_t)eo * rad_ld : nullptr;
  // this edge's row: one 64-bit multiply here, compile-time offsets below (per-store 64-bit address arithmetic cost
  // 33 registers and 30 % of this kernel's speed: ncu r2f)
  const size_t roff_e = (size_t)el * (size_t)out.pitch;
  TOut* const row[3] = {reinterpret_cast<TOut*>(out.p[0]) + roff_e + lane,
                        reinterpret_cast<TOut*>(out.p[MMAX >= 1 ? 1 : 0]) + roff_e + lane,
                        reinterpret_cast<TOut*>(out.p[MMAX >= 2 ? 2 : 0]) + roff_e + lane};
#pragma unroll
  for (int l = 0; l <= LMAX; ++l) {
#pragma unroll
    for (int mp = -T::mm(l); mp <= T::mm(l); ++mp) {
      const float* wr = &sW[warp][T::wrow(l, mp)];
      float acc = 0.f;
#pragma unroll
      for (int n = 0; n < 2 * l + 1; ++n) acc = fmaf(wr[n], x[l * l + n], acc);
      const int am = mp < 0 ? -mp : mp;
      if (rd) {
        // radial layout: [m=0 | m=1 | m=2 ...], each (LMAX+1-m)*CH wide (EF_layers.py:823,848)
        int roff = 0;
#pragma unroll
        for (int k = 0; k < am; ++k) roff += (LMAX + 1 - k) * CH;
        acc *= __ldg(rd + roff + (l - am) * CH + lane);
      }
      row[am][(mp < 0 ? (LMAX + 1 - am) * CH : 0) + (l - am) * CH] = from_f<TOut>(acc);
    }
  }
  // bias column of the SO(2) m = 0 GEMM: the ones_pad (<= 64) columns in front of the m = 0 block hold (1, 0, 0, ...),
  // so the bias is a weight column forward and a row of the wgrad product backward (no bias epilogue, no column sum)
  for (int t = lane; t < ones_pad; t += 32) (row[0] - lane - ones_pad)[t] = from_f<TOut>(t == 0 ? 1.f : 0.f);
}

// ------------------------------------------------------------------------------------------
// rotate-in backward: per edge, dA (GEMM dgrad output) -> d rad, per-edge d x (src | dst halves)
// ------------------------------------------------------------------------------------------
template <int LMAX, int MMAX, int C, typename TIn>
__global__ void __launch_bounds__(128) rotate_in_bwd_kernel(
    const float* __restrict__ x_src, const float* __restrict__ x_dst, const int* __restrict__ src,
    const int* __restrict__ dst, const int* __restrict__ perm, const float* __restrict__ wig,
    const float* __restrict__ rad, int rad_ld, OpBufs<LMAX> dA, float* __restrict__ drad,
    float* __restrict__ dxe /* [E, L, 2C] sorted order */, int e0, int Ec) {
  using T = Tab<LMAX, MMAX>;
  constexpr int CH = 2 * C;
  __shared__ float sW[4][T::WSZ + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int el = blockIdx.x * 4 + warp;
  if (el >= Ec) return;
  const int e = e0 + el;
  const int eo = perm[e];
  const float* w = wig + (size_t)eo * T::WSZ;
  for (int i = lane; i < T::WSZ; i += 32) sW[warp][i] = w[i];
  const int node = lane < C ? src[e] : dst[e];
  const float* xp = (lane < C ? x_src : x_dst) + (size_t)node * T::L * C + (lane & (C - 1));
  float x[T::L], dx[T::L];
#pragma unroll
  for (int i = 0; i < T::L; ++i) {
    x[i] = rad ? __ldg(xp + i * C) : 0.f;
    dx[i] = 0.f;
  }
  __syncwarp();
  const float* rd = rad ? rad + (size_t)eo * rad_ld : nullptr;
  float* drd = rad ? drad + (size_t)eo * rad_ld : nullptr;
#pragma unroll
  for (int l = 0; l <= LMAX; ++l) {
#pragma unroll
    for (int am = 0; am <= T::mm(l); ++am) {
      int roff = 0;
#pragma unroll
      for (int k = 0; k < am; ++k) roff += (LMAX + 1 - k) * CH;
      const int rcol = roff + (l - am) * CH + lane;
      const float rv = rd ? __ldg(rd + rcol) : 1.f;
      float dr = 0.f;
#pragma unroll
      for (int sgn = 0; sgn < (am == 0 ? 1 : 2); ++sgn) {
        const int mp = sgn == 0 ? am : -am;
        const float* wr = &sW[warp][T::wrow(l, mp)];
        const float da = to_f<TIn>(*op_ptr<LMAX, TIn>(dA, el, l, mp, CH, lane));
        if (rd) {
          float xr = 0.f;
#pragma unroll
          for (int n = 0; n < 2 * l + 1; ++n) xr = fmaf(wr[n], x[l * l + n], xr);
          dr = fmaf(da, xr, dr);
        }
        const float dxr = da * rv;
#pragma unroll
        for (int n = 0; n < 2 * l + 1; ++n) dx[l * l + n] = fmaf(wr[n], dxr, dx[l * l + n]);
      }
      if (rd) drd[rcol] = dr;
    }
  }
  float* o = dxe + (size_t)e * T::L * CH + lane;
#pragma unroll
  for (int i = 0; i < T::L; ++i) o[i * CH] = dx[i];
}

// ------------------------------------------------------------------------------------------
// rotate-in backward, node-owned: one warp per node (destination side: edges of its CSR row; source side: its edges
// in the source grouping that fall into the chunk), the two half-warps take alternate edges, lane & 15 = channel of
// that side's half of the src||dst concat.  d x of the node accumulates in registers over its edges in a fixed
// order (deterministic, no atomics) and is written once: the per-edge [E, L, 2C] gradient tensor of the gather
// never exists (it cost 6.3 KB/edge written + read back).  x of the node (needed for d rad) is loaded once per node
// instead of once per edge.
// ------------------------------------------------------------------------------------------
template <int LMAX, int MMAX, int C, typename TIn, int WPB>
__global__ void __launch_bounds__(32 * WPB, 16 / WPB) rotate_in_bwd_node_kernel(
    const float* __restrict__ x_node, const int* __restrict__ seg_ptr, const int* __restrict__ seg_idx,
    const int* __restrict__ perm, const float* __restrict__ wig, const float* __restrict__ rad, int rad_ld,
    OpBufs<LMAX> dA, float* __restrict__ drad, float* __restrict__ dx_out, int coff, int e0, int Ec, int n0, int n1,
    int accumulate) {
  using T = Tab<LMAX, MMAX>;
  static_assert(C == 16, "half-warp = channels of one side");
  static_assert(T::WSZ % 4 == 0, "Wigner rows are 16-byte padded");
  constexpr int CH = 2 * C;
  constexpr int NPAIR = (LMAX + 1) * (MMAX + 1) - MMAX * (MMAX + 1) / 2;  // (l, |m'|) pairs
  __shared__ __align__(16) float sW[WPB][2][2][T::WSZ];  // [warp][half][double buffer]
  __shared__ float sX[WPB][T::L][C];  // x of the warp's node (only d rad needs it): shared memory, not registers
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int half = lane >> 4, c = lane & 15;
  const int n = n0 + blockIdx.x * WPB + warp;
  if (n >= n1) return;
  int lo = seg_ptr[n], hi = seg_ptr[n + 1];
  if (seg_idx) {  // source grouping spans all chunks: keep the positions inside [e0, e0 + Ec)
    int a = lo, b = hi;
    while (a < b) { const int m = (a + b) >> 1; if (seg_idx[m] < e0) a = m + 1; else b = m; }
    const int first = a;
    b = hi;
    while (a < b) { const int m = (a + b) >> 1; if (seg_idx[m] < e0 + Ec) a = m + 1; else b = m; }
    lo = first; hi = a;
  }
  float* o = dx_out + (size_t)n * T::L * C + c;
  if (lo >= hi) {
    if (!accumulate && half == 0) {
#pragma unroll
      for (int i = 0; i < T::L; ++i) o[i * C] = 0.f;
    }
    return;
  }
  float dx[T::L];
#pragma unroll
  for (int i = 0; i < T::L; ++i) dx[i] = 0.f;
  if (rad) {
    const float* xp = x_node + (size_t)n * T::L * C;
    for (int i = lane; i < T::L * C; i += 32) (&sX[warp][0][0])[i] = __ldg(xp + i);
  }
  __syncwarp();  // sX is filled by all 32 lanes and read per half-warp: order the stores of the other half first
  const float* x = &sX[warp][0][c];
  // Software pipeline over the half-warp's edges: edge ids (segment position -> original id, two dependent loads) are
  // fetched two iterations ahead and the Wigner rows one iteration ahead (cp.async into the other buffer), so only
  // the d A / radial loads of the current edge are still exposed; before, the Wigner copy alone (load -> shared
  // store behind the id chain) was 44 % of the stall samples (ncu r1k).
  auto edge_pos = [&](int k) { return k < hi ? (seg_idx ? __ldg(seg_idx + k) : k) : -1; };
  auto wig_async = [&](float* dst, int eo_) {
    const float4* w4 = reinterpret_cast<const float4*>(wig + (size_t)eo_ * T::WSZ);
    for (int i = c; i < T::WSZ / 4; i += 16)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst + 4 * i)),
                   "l"(w4 + i)
                   : "memory");
  };
  int pos_c = edge_pos(lo + half), pos_n = edge_pos(lo + 2 + half);
  int eo_c = pos_c >= 0 ? __ldg(perm + pos_c) : 0, eo_n = pos_n >= 0 ? __ldg(perm + pos_n) : 0;
  int buf = 0;
  const uint32_t hmask = half ? 0xffff0000u : 0x0000ffffu;
  if (pos_c >= 0) wig_async(sW[warp][half][0], eo_c);
  asm volatile("cp.async.commit_group;" ::: "memory");
  for (int it = lo; it < hi; it += 2) {
    const int pos_2 = edge_pos(it + 4 + half);
    const int eo_2 = pos_2 >= 0 ? __ldg(perm + pos_2) : 0;
    if (pos_n >= 0) wig_async(sW[warp][half][buf ^ 1], eo_n);
    asm volatile("cp.async.commit_group;" ::: "memory");
    const bool act = pos_c >= 0;
    const int pos = act ? pos_c : 0;
    const int el = pos - e0;
    const int eo = eo_c;
    float* sw = sW[warp][half][buf];
    // every global load of this edge is issued before the first store of the iteration (the d rad stores would
    // otherwise order the loads behind them): d A rows and radial values of the lane's channel into registers
    const float* rd = rad ? rad + (size_t)eo * rad_ld : nullptr;
    float* drd = rad ? drad + (size_t)eo * rad_ld : nullptr;
    float da_r[T::M], rv_r[NPAIR];
    {
      int t = 0, s2 = 0;
#pragma unroll
      for (int l = 0; l <= LMAX; ++l) {
#pragma unroll
        for (int am = 0; am <= T::mm(l); ++am) {
          int roff = 0;
#pragma unroll
          for (int q = 0; q < am; ++q) roff += (LMAX + 1 - q) * CH;
          rv_r[s2++] = (act && rd) ? __ldg(rd + roff + (l - am) * CH + coff + c) : 1.f;
#pragma unroll
          for (int sgn = 0; sgn < (am == 0 ? 1 : 2); ++sgn) {
            const int mp = sgn == 0 ? am : -am;
            da_r[t++] = act ? to_f<TIn>(*op_ptr<LMAX, TIn>(dA, el, l, mp, CH, coff + c)) : 0.f;
          }
        }
      }
    }
    asm volatile("cp.async.wait_group 1;" ::: "memory");  // this edge's Wigner rows have landed (next edge's may not)
    __syncwarp(hmask);   // the two halves own separate buffers: synchronise per half-warp
    pos_c = pos_n; pos_n = pos_2;
    eo_c = eo_n; eo_n = eo_2;
    buf ^= 1;
    if (!act) continue;
    int t = 0, s2 = 0;
#pragma unroll
    for (int l = 0; l <= LMAX; ++l) {
#pragma unroll
      for (int am = 0; am <= T::mm(l); ++am) {
        int roff = 0;
#pragma unroll
        for (int q = 0; q < am; ++q) roff += (LMAX + 1 - q) * CH;
        const int rcol = roff + (l - am) * CH + coff + c;
        const float rv = rv_r[s2++];
        float dr = 0.f;
#pragma unroll
        for (int sgn = 0; sgn < (am == 0 ? 1 : 2); ++sgn) {
          const int mp = sgn == 0 ? am : -am;
          const float* wr = sw + T::wrow(l, mp);
          const float da = da_r[t++];
          if (rd) {
            float xr = 0.f;
#pragma unroll
            for (int q = 0; q < 2 * l + 1; ++q) xr = fmaf(wr[q], x[(l * l + q) * C], xr);
            dr = fmaf(da, xr, dr);
          }
          const float dxr = da * rv;
#pragma unroll
          for (int q = 0; q < 2 * l + 1; ++q) dx[l * l + q] = fmaf(wr[q], dxr, dx[l * l + q]);
        }
        if (rd) drd[rcol] = dr;
      }
    }
    __syncwarp(hmask);   // every lane of the half is done with sw before the next iteration refills it
  }
#pragma unroll
  for (int i = 0; i < T::L; ++i) {
    const float v = dx[i] + __shfl_xor_sync(0xffffffffu, dx[i], 16);
    if (half == 0) o[i * C] = accumulate ? (o[i * C] + v) : v;
  }
}

// out[n, i, c] = sum_{k in [ptr[n], ptr[n+1])} buf[idx ? idx[k] : k, i, coff + c]
__global__ void segment_sum_rows_kernel(const float* __restrict__ buf, int row_elems /* L*ldc */, int ldc, int coff,
                                        int C, int L, const int* __restrict__ ptr, const int* __restrict__ idx,
                                        float* __restrict__ out, int accumulate) {
  const int n = blockIdx.x;
  const int b = ptr[n], e = ptr[n + 1];
  for (int t = threadIdx.x; t < L * C; t += blockDim.x) {
    const int i = t / C, c = t % C;
    float acc = 0.f;
    for (int k = b; k < e; ++k) {
      const int r = idx ? idx[k] : k;
      acc += buf[(size_t)r * row_elems + i * ldc + coff + c];
    }
    float* o = out + (size_t)n * L * C + t;
    *o = accumulate ? (*o + acc) : acc;
  }
}

// ------------------------------------------------------------------------------------------
// rotate-back + aggregate: one CTA per destination node, one WARP per degree group (Tab::NG groups, independent,
// no block-level synchronisation), lane = 4 adjacent value channels (CH / 4 = 28 active lanes).
// Per edge a warp needs its rows of V (5 contiguous segments: m=0, +1, -1, +2, -2) and its slice of the edge's
// Wigner rows.  One elected lane streams them into a per-warp NST-stage shared-memory ring with 1-D bulk async
// copies (cp.async.bulk, completion on an mbarrier: the TMA engine, UBLKCP in SASS), NST edges ahead of the
// math, so HBM latency is hidden without register staging.  The math reads V as 8/16-byte vectors and the
// Wigner rows as float4 broadcasts: 16 FMAs per shared-memory load.  Fixed edge order per destination =>
// bit-deterministic; no atomics.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// one hardware-elected lane of a converged warp: unlike `lane == 0` the compiler knows the guarded bulk-copy issue
// runs in one thread of a uniform warp and emits no per-operand waterfall loops around it
__device__ __forceinline__ bool elect_one_() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void mbar_init_(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx_(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "RB_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra RB_DONE;\n\t"
      "bra RB_WAIT;\n\t"
      "RB_DONE:\n\t"
      "}" ::"r"(smem_addr(bar)), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_addr(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}

__device__ __forceinline__ void load4(const float* p, float* v) {
  const float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void load4(const __nv_bfloat16* p, float* v) {
  const uint2 t = *reinterpret_cast<const uint2*>(p);
  v[0] = __uint_as_float(t.x << 16); v[1] = __uint_as_float(t.x & 0xffff0000u);
  v[2] = __uint_as_float(t.y << 16); v[3] = __uint_as_float(t.y & 0xffff0000u);
}
__device__ __forceinline__ void store4(float* p, float a, float b, float c, float d) {
  *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
}
__device__ __forceinline__ uint32_t bf2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void store4(__nv_bfloat16* p, float a, float b, float c, float d) {
  *reinterpret_cast<uint2*>(p) = make_uint2(bf2(a, b), bf2(c, d));
}
__device__ __forceinline__ void store8(float* p, const float* o) {
  reinterpret_cast<float4*>(p)[0] = make_float4(o[0], o[1], o[2], o[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(o[4], o[5], o[6], o[7]);
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float* o) {
  *reinterpret_cast<uint4*>(p) = make_uint4(bf2(o[0], o[1]), bf2(o[2], o[3]), bf2(o[4], o[5]), bf2(o[6], o[7]));
}

constexpr int RB_NST = 4;  // ring stages (edges in flight per warp)

// Shared-memory stage of one degree group [L_LO, L_HI]: V segments (m=0, +1, -1, +2, -2; degrees ascending inside
// a segment) followed by the group's Wigner slice.
template <int LMAX, int MMAX, int CH, typename TIn, int L_LO, int L_HI>
struct GroupCfg {
  using T = Tab<LMAX, MMAX>;
  static constexpr int ROWB = CH * (int)sizeof(TIn);
  __host__ __device__ static constexpr int la(int am) { return L_LO > am ? L_LO : am; }
  __host__ __device__ static constexpr int nl(int am) { return L_HI >= la(am) ? L_HI - la(am) + 1 : 0; }
  __host__ __device__ static constexpr int seg_off(int am, int neg) {
    int o = 0;
    if (am > 0) o += nl(0) * ROWB;
    for (int k = 1; k < am; ++k) o += 2 * nl(k) * ROWB;
    if (am > 0 && neg) o += nl(am) * ROWB;
    return o;
  }
  static constexpr int VBYTES = seg_off(MMAX + 1, 0);
  static constexpr int W0 = T::woff(L_LO);
  static constexpr int WBYTES = (T::woff(L_HI + 1) - W0) * 4;
  static constexpr int STAGE = VBYTES + WBYTES;
  __host__ __device__ static constexpr int row_off(int l, int mp) {
    return seg_off(mp < 0 ? -mp : mp, mp < 0) + (l - la(mp < 0 ? -mp : mp)) * ROWB;
  }
  static_assert(ROWB % 16 == 0 && WBYTES % 16 == 0, "bulk copies need 16-byte granularity");

  // lane 0: fill one stage with edge (sorted position k, local index el, original id eo)
  __device__ static __forceinline__ void issue(uint8_t* stage, uint64_t* bar, const OpBufs<LMAX>& V,
                                               const float* __restrict__ wig, int el, int eo) {
    mbar_expect_tx_(bar, STAGE);
#pragma unroll
    for (int am = 0; am <= MMAX; ++am) {
      if (nl(am) > 0) {
        const TIn* base = reinterpret_cast<const TIn*>(V.p[am]) + (size_t)el * V.pitch;
        if (am == 0) {
          bulk_g2s(stage + seg_off(0, 0), base + la(0) * CH, nl(0) * ROWB, bar);
        } else {
#pragma unroll
          for (int neg = 0; neg < 2; ++neg)
            bulk_g2s(stage + seg_off(am, neg), base + (neg * (LMAX + 1 - am) + (la(am) - am)) * CH, nl(am) * ROWB,
                     bar);
        }
      }
    }
    bulk_g2s(stage + VBYTES, wig + (size_t)eo * T::WSZ + W0, WBYTES, bar);
  }
};

template <int LMAX, int MMAX, int CH, typename TIn>
struct BackCfg {
  using T = Tab<LMAX, MMAX>;
  template <int G>
  __host__ __device__ static constexpr int stage_bytes() {
    if constexpr (G < T::NG) return GroupCfg<LMAX, MMAX, CH, TIn, T::glo(G), T::ghi(G)>::STAGE;
    else return 0;
  }
  __host__ __device__ static constexpr int stage_of(int g) {
    return g == 0 ? stage_bytes<0>() : g == 1 ? stage_bytes<1>() : g == 2 ? stage_bytes<2>()
         : g == 3 ? stage_bytes<3>() : g == 4 ? stage_bytes<4>() : g == 5 ? stage_bytes<5>() : 0;
  }
  static constexpr int BAR_BYTES = 256;  // NG * RB_NST mbarriers (<= 6 * 4 * 8 bytes), keeps the rings 128-byte aligned
  __host__ __device__ static constexpr int ring_off(int g) {
    int o = BAR_BYTES;
    for (int k = 0; k < g; ++k) o += RB_NST * stage_of(k);
    return o;
  }
  static constexpr int SMEM = ring_off(T::NG);
  // one degree group per (single-warp) CTA: barriers + the largest group's ring
  __host__ __device__ static constexpr int max_stage() {
    int m = 0;
    for (int k = 0; k < T::NG; ++k) m = stage_of(k) > m ? stage_of(k) : m;
    return m;
  }
  static constexpr int SMEM_SPLIT = BAR_BYTES + RB_NST * max_stage();
  static_assert(T::NG <= 6, "at most 6 degree groups");
};

template <int LMAX, int MMAX, int CH, int VC, typename TIn, int L_LO, int L_HI>
__device__ __forceinline__ void rotate_back_fwd_group(const OpBufs<LMAX>& V, const float* __restrict__ alpha,
                                                      int heads, const int* __restrict__ perm,
                                                      const float* __restrict__ wig, float* __restrict__ out, int n,
                                                      int kb, int ke, int e0, uint8_t* ring, uint64_t* bars,
                                                      int lane) {
  using T = Tab<LMAX, MMAX>;
  using G = GroupCfg<LMAX, MMAX, CH, TIn, L_LO, L_HI>;
  constexpr int NC = (L_HI + 1) * (L_HI + 1) - L_LO * L_LO;
  const bool active = lane < CH / 4;
  const int c4 = active ? 4 * lane : 0;
  const bool issuer = elect_one_();
  float acc[NC][4];
#pragma unroll
  for (int i = 0; i < NC; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < RB_NST; ++s) mbar_init_(&bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  // original edge ids of the next 32 edges to issue, one per lane
  int pbase = kb;
  int pv = (pbase + lane < ke) ? __ldg(perm + pbase + lane) : 0;
#pragma unroll
  for (int s = 0; s < RB_NST; ++s) {
    const int j = kb + s;
    const int eo = __shfl_sync(0xffffffffu, pv, s);
    if (issuer && j < ke) G::issue(ring + s * G::STAGE, &bars[s], V, wig, j - e0, eo);
  }
  float a_next = (alpha && kb < ke) ? __ldg(alpha + (size_t)kb * heads + c4 / VC) : 1.f;
  int s = 0;
  uint32_t phase = 0;
  for (int k = kb; k < ke; ++k) {
    const float a = a_next;
    if (alpha && k + 1 < ke) a_next = __ldg(alpha + (size_t)(k + 1) * heads + c4 / VC);
    mbar_wait_(&bars[s], phase);
    const uint8_t* st = ring + s * G::STAGE;
    const float* w = reinterpret_cast<const float*>(st + G::VBYTES);
#pragma unroll
    for (int l = L_LO; l <= L_HI; ++l) {
      const float sc = a * T::rescale(l);
#pragma unroll
      for (int mp = -T::mm(l); mp <= T::mm(l); ++mp) {
        float v[4];
        load4(reinterpret_cast<const TIn*>(st + G::row_off(l, mp)) + c4, v);
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] *= sc;
        const float4* wr = reinterpret_cast<const float4*>(w + (T::wrow(l, mp) - G::W0));
#pragma unroll
        for (int q = 0; q < T::rowlen(l) / 4; ++q) {
          const float4 t = wr[q];
          const float tw[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (4 * q + u < 2 * l + 1) {
              const int i = l * l - L_LO * L_LO + 4 * q + u;
#pragma unroll
              for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(tw[u], v[j], acc[i][j]);
            }
          }
        }
      }
    }
    __syncwarp();  // every lane is done with stage s before it is refilled
    const int jn = k + RB_NST;
    if (jn < ke) {
      if (jn - pbase >= 32) {
        pbase += 32;
        pv = (pbase + lane < ke) ? __ldg(perm + pbase + lane) : 0;
      }
      const int eo = __shfl_sync(0xffffffffu, pv, jn - pbase);
      if (issuer) G::issue(ring + s * G::STAGE, &bars[s], V, wig, jn - e0, eo);
    }
    if (++s == RB_NST) { s = 0; phase ^= 1; }
  }
  if (active) {
    float* o = out + ((size_t)n * T::L + L_LO * L_LO) * CH + c4;
#pragma unroll
    for (int i = 0; i < NC; ++i) store4(o + (size_t)i * CH, acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
  }
}

// SPLIT: one degree group per single-warp CTA (blockIdx.y = group, nodes fastest).  The groups run different fully
// unrolled code (16-18 KB each, 64-72 KB per kernel): with all groups in every CTA the instruction cache thrashed
// (ncu r1k: 29 % / 44 % of the stall samples "no instruction"); split, the CTAs resident on an SM mostly share one
// group's code.
template <int LMAX, int MMAX, int CH, int VC, typename TIn, bool SPLIT>
__global__ void __launch_bounds__(SPLIT ? 32 : 32 * Tab<LMAX, MMAX>::NG) rotate_back_agg_fwd_kernel(
    OpBufs<LMAX> V, const float* __restrict__ alpha /* [E, heads] sorted */, int heads,
    const int* __restrict__ rowptr, const int* __restrict__ perm, const float* __restrict__ wig,
    float* __restrict__ out /* [Nt, L, CH] */, int n0, int e0) {
  using T = Tab<LMAX, MMAX>;
  using B = BackCfg<LMAX, MMAX, CH, TIn>;
  extern __shared__ __align__(128) uint8_t rb_smem[];
  const int g = SPLIT ? (int)blockIdx.y : (int)(threadIdx.x >> 5), lane = threadIdx.x & 31;
  const int n = n0 + blockIdx.x;
  const int kb = rowptr[n], ke = rowptr[n + 1];
  uint64_t* bars = reinterpret_cast<uint64_t*>(rb_smem) + (SPLIT ? 0 : g * RB_NST);
#define GROUP(G)                                                                                                   \
  case G:                                                                                                          \
    if constexpr (G < T::NG)                                                                                       \
      rotate_back_fwd_group<LMAX, MMAX, CH, VC, TIn, T::glo(G), T::ghi(G)>(                                        \
          V, alpha, heads, perm, wig, out, n, kb, ke, e0, rb_smem + (SPLIT ? B::BAR_BYTES : B::ring_off(G)), bars, \
          lane);                                                                                                   \
    break;
  switch (g) {
    GROUP(0) GROUP(1) GROUP(2) GROUP(3) GROUP(4) GROUP(5)
    default: break;
  }
#undef GROUP
}

// backward: d out[n] -> dV (same edge-row layout as V: the SO(2) dgrad/wgrad GEMMs consume it as it is)
// and per-group partials of d alpha[e, h] = sum_{j, c in h} dva[j, c] * V[e, j, c]  (dalpha [NG, Etot, heads];
// the segment-softmax backward sums the NG partials in fixed order).  Same ring as the forward.
template <int LMAX, int MMAX, int CH, int VC, typename TIn, typename TOut, int L_LO, int L_HI>
__device__ __forceinline__ void rotate_back_bwd_group(const OpBufs<LMAX>& V, const OpBufs<LMAX>& dY,
                                                      const float* __restrict__ alpha, int heads,
                                                      const int* __restrict__ perm, const float* __restrict__ wig,
                                                      const float* __restrict__ dout, float* __restrict__ dalpha_g,
                                                      int n, int kb, int ke, int e0, uint8_t* ring, uint64_t* bars,
                                                      int lane) {
  using T = Tab<LMAX, MMAX>;
  using G = GroupCfg<LMAX, MMAX, CH, TIn, L_LO, L_HI>;
  constexpr int NC = (L_HI + 1) * (L_HI + 1) - L_LO * L_LO;
  const bool active = lane < CH / 4;
  const int c4 = active ? 4 * lane : 0;
  const bool issuer = elect_one_();
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < RB_NST; ++s) mbar_init_(&bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  int pbase = kb;
  int pv = (pbase + lane < ke) ? __ldg(perm + pbase + lane) : 0;
#pragma unroll
  for (int s = 0; s < RB_NST; ++s) {
    const int j = kb + s;
    const int eo = __shfl_sync(0xffffffffu, pv, s);
    if (issuer && j < ke) G::issue(ring + s * G::STAGE, &bars[s], V, wig, j - e0, eo);
  }
  float dy[NC][4];
  {
    const float* d = dout + ((size_t)n * T::L + L_LO * L_LO) * CH + c4;
#pragma unroll
    for (int i = 0; i < NC; ++i) load4(d + (size_t)i * CH, dy[i]);
  }
  float a_next = kb < ke ? __ldg(alpha + (size_t)kb * heads + c4 / VC) : 1.f;
  int s = 0;
  uint32_t phase = 0;
  for (int k = kb; k < ke; ++k) {
    const float a = a_next;
    if (k + 1 < ke) a_next = __ldg(alpha + (size_t)(k + 1) * heads + c4 / VC);
    mbar_wait_(&bars[s], phase);
    const uint8_t* st = ring + s * G::STAGE;
    const float* w = reinterpret_cast<const float*>(st + G::VBYTES);
    const int el = k - e0;
    float da = 0.f;
#pragma unroll
    for (int l = L_LO; l <= L_HI; ++l) {
      const float rs = T::rescale(l);
      float dre[2 * MMAX + 1][4];  // index mp + MMAX
#pragma unroll
      for (int mp = -T::mm(l); mp <= T::mm(l); ++mp) {
        float g4[4] = {0.f, 0.f, 0.f, 0.f};
        const float4* wr = reinterpret_cast<const float4*>(w + (T::wrow(l, mp) - G::W0));
#pragma unroll
        for (int q = 0; q < T::rowlen(l) / 4; ++q) {
          const float4 t = wr[q];
          const float tw[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            if (4 * q + u < 2 * l + 1) {
              const int i = l * l - L_LO * L_LO + 4 * q + u;
#pragma unroll
              for (int j = 0; j < 4; ++j) g4[j] = fmaf(tw[u], dy[i][j], g4[j]);
            }
          }
        }
        float v[4];
        load4(reinterpret_cast<const TIn*>(st + G::row_off(l, mp)) + c4, v);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          g4[j] *= rs;
          da = fmaf(g4[j], v[j], da);
          dre[mp + MMAX][j] = g4[j] * a;
        }
      }
      if (active) {
#pragma unroll
        for (int mp = -T::mm(l); mp <= T::mm(l); ++mp) {
          const float* g = dre[mp + MMAX];
          store4(op_ptr<LMAX, TOut>(dY, el, l, mp, CH, c4), g[0], g[1], g[2], g[3]);
        }
      }
    }
    // d alpha partial of this degree group: sum over the VC channels of a head = VC/4 adjacent lanes
    if (!active) da = 0.f;
#pragma unroll
    for (int o = VC / 8; o > 0; o >>= 1) da += __shfl_xor_sync(0xffffffffu, da, o);
    if (active && (lane % (VC / 4)) == 0) dalpha_g[(size_t)k * heads + c4 / VC] = da;
    __syncwarp();
    const int jn = k + RB_NST;
    if (jn < ke) {
      if (jn - pbase >= 32) {
        pbase += 32;
        pv = (pbase + lane < ke) ? __ldg(perm + pbase + lane) : 0;
      }
      const int eo = __shfl_sync(0xffffffffu, pv, jn - pbase);
      if (issuer) G::issue(ring + s * G::STAGE, &bars[s], V, wig, jn - e0, eo);
    }
    if (++s == RB_NST) { s = 0; phase ^= 1; }
  }
}

template <int LMAX, int MMAX, int CH, int VC, typename TIn, typename TOut, bool SPLIT>
__global__ void __launch_bounds__(SPLIT ? 32 : 32 * Tab<LMAX, MMAX>::NG) rotate_back_agg_bwd_kernel(
    OpBufs<LMAX> V, OpBufs<LMAX> dY, const float* __restrict__ alpha, int heads,
    const int* __restrict__ rowptr, const int* __restrict__ perm, const float* __restrict__ wig,
    const float* __restrict__ dout /* [Nt, L, CH] */, float* __restrict__ dalpha /* [NG, Etot, heads] sorted */,
    long long part_stride, int n0, int e0) {
  using T = Tab<LMAX, MMAX>;
  using B = BackCfg<LMAX, MMAX, CH, TIn>;
  extern __shared__ __align__(128) uint8_t rb_smem[];
  const int g = SPLIT ? (int)blockIdx.y : (int)(threadIdx.x >> 5), lane = threadIdx.x & 31;
  const int n = n0 + blockIdx.x;
  const int kb = rowptr[n], ke = rowptr[n + 1];
  uint64_t* bars = reinterpret_cast<uint64_t*>(rb_smem) + (SPLIT ? 0 : g * RB_NST);
  float* dag = dalpha + (size_t)g * part_stride;
#define GROUP(G)                                                                                                 \
  case G:                                                                                                        \
    if constexpr (G < T::NG)                                                                                     \
      rotate_back_bwd_group<LMAX, MMAX, CH, VC, TIn, TOut, T::glo(G), T::ghi(G)>(                                \
          V, dY, alpha, heads, perm, wig, dout, dag, n, kb, ke, e0,                                              \
          rb_smem + (SPLIT ? B::BAR_BYTES : B::ring_off(G)), bars, lane);                                        \
    break;
  switch (g) {
    GROUP(0) GROUP(1) GROUP(2) GROUP(3) GROUP(4) GROUP(5)
    default: break;
  }
#undef GROUP
}

// ------------------------------------------------------------------------------------------
// EdgeDegreeEmbedding tail.  out[n, l*l+l+nn, c] = (1/rf) sum_e D^l[0, nn] rescale_l m0[e, l, c]
// one warp per destination; lanes = 2 edge slots x C channels
// ------------------------------------------------------------------------------------------
template <int LMAX, int MMAX, int C>
__global__ void __launch_bounds__(128) edge_degree_fwd_kernel(
    const float* __restrict__ m0 /* [E, (LMAX+1)*C] original order */, const int* __restrict__ rowptr,
    const int* __restrict__ perm, const float* __restrict__ wig, float inv_rf, float* __restrict__ out, int Nt) {
  using T = Tab<LMAX, MMAX>;
  static_assert(C == 16, "two edge slots x 16 channels per warp");
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * 4 + warp;
  if (n >= Nt) return;
  const int slot = lane >> 4, c = lane & 15;
  float acc[T::L];
#pragma unroll
  for (int i = 0; i < T::L; ++i) acc[i] = 0.f;
  for (int k = rowptr[n] + slot; k < rowptr[n + 1]; k += 2) {
    const int eo = perm[k];
    const float* w = wig + (size_t)eo * T::WSZ;
    const float* r = m0 + (size_t)eo * (LMAX + 1) * C + c;
#pragma unroll
    for (int l = 0; l <= LMAX; ++l) {
      const float v = __ldg(r + l * C) * T::rescale(l);
      const float* wr = w + T::wrow(l, 0);
#pragma unroll
      for (int nn = 0; nn < 2 * l + 1; ++nn) acc[l * l + nn] = fmaf(__ldg(wr + nn), v, acc[l * l + nn]);
    }
  }
#pragma unroll
  for (int i = 0; i < T::L; ++i) {
    const float t = acc[i] + __shfl_xor_sync(0xffffffffu, acc[i], 16);
    if (slot == 0) out[(size_t)n * T::L * C + i * C + c] = t * inv_rf;
  }
}

// d m0[e, l, c] = (1/rf) rescale_l sum_nn D^l[0, nn] dout[dst(e), l*l+nn, c]; thread = (edge, c)
template <int LMAX, int MMAX, int C>
__global__ void edge_degree_bwd_kernel(const float* __restrict__ dout, const int* __restrict__ dst_orig,
                                       const float* __restrict__ wig, float inv_rf, float* __restrict__ dm0, int E) {
  using T = Tab<LMAX, MMAX>;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= E * C) return;
  const int e = t / C, c = t % C;
  const float* w = wig + (size_t)e * T::WSZ;
  const float* d = dout + (size_t)dst_orig[e] * T::L * C + c;
#pragma unroll
  for (int l = 0; l <= LMAX; ++l) {
    const float* wr = w + T::wrow(l, 0);
    float g = 0.f;
#pragma unroll
    for (int nn = 0; nn < 2 * l + 1; ++nn) g = fmaf(__ldg(wr + nn), d[(l * l + nn) * C], g);
    dm0[(size_t)e * (LMAX + 1) * C + l * C + c] = g * T::rescale(l) * inv_rf;
  }
}

}  // namespace efl

// ------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------
using namespace efl;

#define DISPATCH_LM(lmax, mmax, CALL)                                                    \
  if (lmax == 6 && mmax == 2) { CALL(6, 2); }                                            \
  else if (lmax == 8 && mmax == 2) { CALL(8, 2); }                                       \
  else if (lmax == 4 && mmax == 2) { CALL(4, 2); }                                       \
  else { efl_set_error("unsupported (lmax, mmax); built for (6,2), (8,2), (4,2)"); return EFL_ERR_UNSUPPORTED; }

extern "C" int efl_rotate_in_fwd(int lmax, int mmax, int C, int bf16_out, const float* x_src, const float* x_dst,
                                 const int* src, const int* dst, const int* perm, const float* wig,
                                 const float* rad, int rad_ld, void* a0, void* a1, void* a2, long long pitch,
                                 int ones_pad, int e0, int Ec, void* stream) {
  if (Ec <= 0) return EFL_OK;
  EFL_REQUIRE(C == 16, "efl_rotate_in_fwd: sphere_channels must be 16");
  EFL_REQUIRE(ones_pad >= 0 && ones_pad <= 64, "efl_rotate_in_fwd: ones_pad must be in [0, 64]");
  cudaStream_t s = (cudaStream_t)stream;
  dim3 grid((Ec + 3) / 4), block(128);
#define CALL(LM, MM)                                                                                          \
  {                                                                                                           \
    OpBufs<LM> o{{a0, a1, a2}, pitch};                                                                        \
    if (bf16_out)                                                                                             \
      rotate_in_fwd_kernel<LM, MM, 16, __nv_bfloat16><<<grid, block, 0, s>>>(x_src, x_dst, src, dst, perm,    \
                                                                             wig, rad, rad_ld, o, ones_pad,   \
                                                                             e0, Ec);                         \
    else                                                                                                      \
      rotate_in_fwd_kernel<LM, MM, 16, float><<<grid, block, 0, s>>>(x_src, x_dst, src, dst, perm, wig, rad,  \
                                                                     rad_ld, o, ones_pad, e0, Ec);            \
  }
  DISPATCH_LM(lmax, mmax, CALL)
#undef CALL
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

extern "C" int efl_rotate_in_bwd(int lmax, int mmax, int C, int bf16_in, const float* x_src, const float* x_dst,
                                 const int* src, const int* dst, const int* perm, const float* wig,
                                 const float* rad, int rad_ld, const void* da0, const void* da1, const void* da2,
                                 long long pitch, float* drad, float* dxe, int e0, int Ec, void* stream) {
  if (Ec <= 0) return EFL_OK;
  EFL_REQUIRE(C == 16, "efl_rotate_in_bwd: sphere_channels must be 16");
  cudaStream_t s = (cudaStream_t)stream;
  dim3 grid((Ec + 3) / 4), block(128);
#define CALL(LM, MM)                                                                                           \
  {                                                                                                            \
    OpBufs<LM> o{{(void*)da0, (void*)da1, (void*)da2}, pitch};                                                 \
    if (bf16_in)                                                                                               \
      rotate_in_bwd_kernel<LM, MM, 16, __nv_bfloat16><<<grid, block, 0, s>>>(x_src, x_dst, src, dst, perm,     \
                                                                             wig, rad, rad_ld, o, drad, dxe,   \
                                                                             e0, Ec);                          \
    else                                                                                                       \
      rotate_in_bwd_kernel<LM, MM, 16, float><<<grid, block, 0, s>>>(x_src, x_dst, src, dst, perm, wig, rad,   \
                                                                     rad_ld, o, drad, dxe, e0, Ec);            \
  }
  DISPATCH_LM(lmax, mmax, CALL)
#undef CALL
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// Node-owned backward of rotate-in.  side 0: destination half (seg_ptr = CSR rowptr, seg_idx NULL, nodes [n0, n1) of
// the chunk); side 1: source half (seg_ptr/seg_idx = source grouping, all source nodes, accumulate over chunks).
extern "C" int efl_rotate_in_bwd_nodes(int lmax, int mmax, int C, int bf16_in, int side, const float* x_node,
                                       const int* seg_ptr, const int* seg_idx, const int* perm, const float* wig,
                                       const float* rad, int rad_ld, const void* da0, const void* da1,
                                       const void* da2, long long pitch, float* drad, float* dx_out, int e0, int Ec,
                                       int n0, int n1, int accumulate, void* stream) {
  if (n1 <= n0) return EFL_OK;
  EFL_REQUIRE(C == 16, "efl_rotate_in_bwd_nodes: sphere_channels must be 16");
  EFL_REQUIRE(side == 0 || seg_idx != nullptr, "efl_rotate_in_bwd_nodes: the source side needs the source grouping");
  cudaStream_t s = (cudaStream_t)stream;
  // a warp owns a node, one warp per CTA: the hardware backfills warp by warp (16 resident per SM at 128 registers);
  // 2 and 4 warps per CTA measured the same
  constexpr int wpb = 1;
  dim3 grid((n1 - n0 + wpb - 1) / wpb), block(32 * wpb);
  const int coff = side == 0 ? C : 0;
#define CALL_W(LM, MM, TIN, W)                                                                                   \
  rotate_in_bwd_node_kernel<LM, MM, 16, TIN, W><<<grid, block, 0, s>>>(                                          \
      x_node, seg_ptr, seg_idx, perm, wig, rad, rad_ld, o, drad, dx_out, coff, e0, Ec, n0, n1, accumulate);
#define CALL(LM, MM)                                                                                             \
  {                                                                                                              \
    OpBufs<LM> o{{(void*)da0, (void*)da1, (void*)da2}, pitch};                                                   \
    if (bf16_in) CALL_W(LM, MM, __nv_bfloat16, wpb) else CALL_W(LM, MM, float, wpb)                              \
  }
  DISPATCH_LM(lmax, mmax, CALL)
#undef CALL
#undef CALL_W
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

extern "C" int efl_segment_sum_rows(const float* buf, int L, int ldc, int coff, int C, const int* ptr,
                                    const int* idx, float* out, int num_nodes, int accumulate, void* stream) {
  if (num_nodes <= 0) return EFL_OK;
  segment_sum_rows_kernel<<<num_nodes, 256, 0, (cudaStream_t)stream>>>(buf, L * ldc, ldc, coff, C, L, ptr, idx, out,
                                                                        accumulate);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// launch with dynamic shared memory (opt-in above 48 KB, set once per instantiation)
#define RB_LAUNCH(KERNEL, SMEM_BYTES, ...)                                                              \
  {                                                                                                     \
    static unsigned long long attr_done = 0;                                                            \
    if (efl_first_use_on_device(&attr_done)) {                                                          \
      cudaError_t ae = cudaFuncSetAttribute(KERNEL, cudaFuncAttributeMaxDynamicSharedMemorySize,        \
                                            SMEM_BYTES);                                                \
      if (ae != cudaSuccess) {                                                                          \
        efl_set_error(cudaGetErrorString(ae));                                                          \
        return (int)ae;                                                                                 \
      }                                                                                                 \
    }                                                                                                   \
    KERNEL<<<grid, block, SMEM_BYTES, s>>>(__VA_ARGS__);                                                \
  }

// EFL_ROTBACK_SPLIT=0 keeps all degree groups in one CTA (one warp each); default: one group per single-warp CTA
static bool rotback_split() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("EFL_ROTBACK_SPLIT");
    v = e ? atoi(e) : 1;
  }
  return v != 0;
}

extern "C" int efl_rotate_back_parts(int lmax, int mmax) {
  (void)mmax;
  return lmax > 3 ? lmax - 2 : 1;  // Tab::NG: degree groups = partial d-alpha planes of the backward
}

extern "C" int efl_rotate_back_aggregate_fwd(int lmax, int mmax, int heads, int value_ch, int bf16_in,
                                             const void* v0, const void* v1, const void* v2, long long pitch,
                                             const float* alpha, const int* rowptr, const int* perm,
                                             const float* wig, float* out, int n0, int n1, int e0, void* stream) {
  if (n1 <= n0) return EFL_OK;
  EFL_REQUIRE(heads * value_ch == 112 && value_ch == 16, "efl_rotate_back_aggregate_fwd: needs heads*value == 112, value == 16");
  cudaStream_t s = (cudaStream_t)stream;
  const bool split = rotback_split();
#define FWD(LM, MM, TIN, SP, SM)                                                                               \
  RB_LAUNCH((rotate_back_agg_fwd_kernel<LM, MM, 112, 16, TIN, SP>), (BackCfg<LM, MM, 112, TIN>::SM), o, alpha,  \
            heads, rowptr, perm, wig, out, n0, e0)
#define CALL(LM, MM)                                                                                           \
  {                                                                                                            \
    OpBufs<LM> o{{(void*)v0, (void*)v1, (void*)v2}, pitch};                                                    \
    dim3 grid(n1 - n0, split ? Tab<LM, MM>::NG : 1), block(split ? 32 : 32 * Tab<LM, MM>::NG);                 \
    if (bf16_in) {                                                                                             \
      if (split) FWD(LM, MM, __nv_bfloat16, true, SMEM_SPLIT) else FWD(LM, MM, __nv_bfloat16, false, SMEM)     \
    } else {                                                                                                   \
      if (split) FWD(LM, MM, float, true, SMEM_SPLIT) else FWD(LM, MM, float, false, SMEM)                     \
    }                                                                                                          \
  }
  DISPATCH_LM(lmax, mmax, CALL)
#undef CALL
#undef FWD
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

extern "C" int efl_rotate_back_aggregate_bwd(int lmax, int mmax, int heads, int value_ch, int bf16,
                                             const void* v0, const void* v1, const void* v2, long long pitch,
                                             void* dy0, void* dy1, void* dy2, long long dy_pitch, const float* alpha,
                                             const int* rowptr, const int* perm,
                                             const float* wig, const float* dout, float* dalpha,
                                             long long part_stride, int n0, int n1, int e0, void* stream) {
  if (n1 <= n0) return EFL_OK;
  EFL_REQUIRE(heads * value_ch == 112 && value_ch == 16, "efl_rotate_back_aggregate_bwd: needs heads*value == 112, value == 16");
  cudaStream_t s = (cudaStream_t)stream;
  const bool split = rotback_split();
#define BWD(LM, MM, TIN, SP, SM)                                                                               \
  RB_LAUNCH((rotate_back_agg_bwd_kernel<LM, MM, 112, 16, TIN, TIN, SP>), (BackCfg<LM, MM, 112, TIN>::SM), vi,  \
            dyo, alpha, heads, rowptr, perm, wig, dout, dalpha, part_stride, n0, e0)
#define CALL(LM, MM)                                                                                           \
  {                                                                                                            \
    OpBufs<LM> vi{{(void*)v0, (void*)v1, (void*)v2}, pitch};                                                   \
    OpBufs<LM> dyo{{dy0, dy1, dy2}, dy_pitch};                                                                 \
    dim3 grid(n1 - n0, split ? Tab<LM, MM>::NG : 1), block(split ? 32 : 32 * Tab<LM, MM>::NG);                 \
    if (bf16) {                                                                                                \
      if (split) BWD(LM, MM, __nv_bfloat16, true, SMEM_SPLIT) else BWD(LM, MM, __nv_bfloat16, false, SMEM)     \
    } else {                                                                                                   \
      if (split) BWD(LM, MM, float, true, SMEM_SPLIT) else BWD(LM, MM, float, false, SMEM)                     \
    }                                                                                                          \
  }
  DISPATCH_LM(lmax, mmax, CALL)
#undef CALL
#undef BWD
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

extern "C" int efl_edge_degree_fwd(int lmax, int mmax, int C, const float* m0, const int* rowptr, const int* perm,
                                   const float* wig, float rescale_factor, float* out, int Nt, void* stream) {
  if (Nt <= 0) return EFL_OK;
  EFL_REQUIRE(C == 16, "efl_edge_degree_fwd: sphere_channels must be 16");
  cudaStream_t s = (cudaStream_t)stream;
  dim3 grid((Nt + 3) / 4), block(128);
#define CALL(LM, MM) edge_degree_fwd_kernel<LM, MM, 16><<<grid, block, 0, s>>>(m0, rowptr, perm, wig, 1.0f / rescale_factor, out, Nt);
  DISPATCH_LM(lmax, mmax, CALL)
#undef CALL
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

extern "C" int efl_edge_degree_bwd(int lmax, int mmax, int C, const float* dout, const int* dst_orig,
                                   const float* wig, float rescale_factor, float* dm0, int E, void* stream) {
  if (E <= 0) return EFL_OK;
  EFL_REQUIRE(C == 16, "efl_edge_degree_bwd: sphere_channels must be 16");
  cudaStream_t s = (cudaStream_t)stream;
  const int total = E * 16;
  dim3 grid((total + 255) / 256), block(256);
#define CALL(LM, MM) edge_degree_bwd_kernel<LM, MM, 16><<<grid, block, 0, s>>>(dout, dst_orig, wig, 1.0f / rescale_factor, dm0, E);
  DISPATCH_LM(lmax, mmax, CALL)
#undef CALL
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
