// Wigner-D rows from edge frames.
//
// Replaces SO3_Rotation.set_wigner / RotationToWignerDMatrix / wigner_D / _z_rot_mat
// (reference model/EF_layers.py:485-528, 2207-2229).  The reference materialises two
// dense [E, L, L] tensors; only the rows with |m'| <= mmax of each diagonal block are
// ever used (rotate: wigner[:, mask, :], rotate_inv: wigner_inv[:, :, mask]), so this
// kernel emits exactly those rows (235 values for lmax 6, mmax 2), each padded to 16 bytes: WSZ = 296 floats/edge.
//
// D^l = Z(alpha) J_l Z(beta) J_l Z(gamma) with Z(a)[m,m] = cos(m a), Z(a)[m,-m] = -sin(m a).
//
// One THREAD per edge, every (l, m') row fully unrolled.  The J matrices sit in __constant__ memory: with all lanes
// of a warp on the same (l, m', i, k) the operand is a constant-bank immediate of the FFMA (no shared-memory load per
// multiply), and J_l is 3/4 structural zeros (it only couples m-classes (sign, parity of l+|m|) in a fixed pairing), a
// pattern known at compile time, so the skipped products never issue: 585 FFMA per edge for (6,2) instead of the
// 2 253 of the dense row-times-matrix products.  A thread's rows go to HBM through a per-warp shared-memory stage, one
// degree block at a time: written directly, each 16-byte store of a warp lands in 32 different lines (edges are
// 1 184 B apart) and the kernel is bound by L2 write requests (12.7 ms per 12 M edges, 0.18 of HBM); staged, a warp's
// store instruction covers whole contiguous degree blocks (<= 400 B runs).  (The earlier warp-per-edge kernel serialised the seven degrees of a
// warp's rows through a switch and paid one LDS per FFMA: 2.8 ns/edge; see profiles/r1_radius_bench_v1.jsonl.)
#include <cstdlib>

#include "efl_common.cuh"

namespace efl {

constexpr int WIG_LMAX_BUILT = 8;
__host__ __device__ constexpr int wig_joff(int l) { return l * (2 * l - 1) * (2 * l + 1) / 3; }  // sum_{k<l} (2k+1)^2
// J_l blocks l = 0..8, row-major, concatenated: the table for a smaller lmax is a prefix of it
__constant__ float cJ[wig_joff(WIG_LMAX_BUILT + 1)];

// Structural non-zeros of J_l.  Class of order m: (s, q) = (m < 0, parity of l + |m|); J_l couples (1,1) and (0,0) to
// themselves and (1,0) with (0,1).  (A superset of the numerical non-zeros; tests/test_constants.py checks it.)
__host__ __device__ constexpr bool wig_jnz(int l, int i, int k) {
  const int mi = i - l, mk = k - l;
  const int si = mi < 0, sk = mk < 0;
  const int qi = (l + (mi < 0 ? -mi : mi)) & 1, qk = (l + (mk < 0 ? -mk : mk)) & 1;
  return (si == qi) ? (si == sk && qi == qk) : (si != sk && qi != qk);
}

template <int l, int mp, int NA>
__device__ __forceinline__ void wigner_row(const float (&cA)[NA], const float (&sA)[NA], const float (&cB)[NA],
                                           const float (&sB)[NA], const float (&cG)[NA], const float (&sG)[NA],
                                           float* __restrict__ out) {
  constexpr int n = 2 * l + 1, jo = wig_joff(l), am = mp < 0 ? -mp : mp, rl = (n + 3) & ~3;
  float r2[n], r3[n], r4[rl];
  // row m' of Z(alpha) J: cos(m' a) J[m', :] - sin(m' a) J[-m', :]   (cos even, sin odd in m')
  const float ca = cA[am];
  const float sa = mp < 0 ? -sA[am] : sA[am];
#pragma unroll
  for (int k = 0; k < n; ++k) {
    float v = 0.f;
    if (wig_jnz(l, mp + l, k)) v = ca * cJ[jo + (mp + l) * n + k];
    if (mp != 0 && wig_jnz(l, -mp + l, k)) v = fmaf(-sa, cJ[jo + (-mp + l) * n + k], v);
    r2[k] = v;
  }
  // times Z(beta)
#pragma unroll
  for (int k = 0; k < n; ++k) {
    const int nn = k - l, an = nn < 0 ? -nn : nn;
    if (nn == 0) {
      r3[k] = r2[k];
    } else {
      const float sb = nn < 0 ? -sB[an] : sB[an];
      r3[k] = r2[k] * cB[an] + r2[n - 1 - k] * sb;
    }
  }
  // times J (structural zeros skipped)
#pragma unroll
  for (int k = 0; k < n; ++k) {
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < n; ++i)
      if (wig_jnz(l, i, k)) acc = fmaf(r3[i], cJ[jo + i * n + k], acc);
    r4[k] = acc;
  }
  // times Z(gamma), rows padded to 16 bytes
  float o[rl];
#pragma unroll
  for (int k = 0; k < rl; ++k) {
    const int nn = k - l, an = nn < 0 ? -nn : nn;
    if (k >= n) {
      o[k] = 0.f;
    } else if (nn == 0) {
      o[k] = r4[k];
    } else {
      const float sg = nn < 0 ? -sG[an] : sG[an];
      o[k] = r4[k] * cG[an] + r4[n - 1 - k] * sg;
    }
  }
#pragma unroll
  for (int q = 0; q < rl / 4; ++q)
    *reinterpret_cast<float4*>(out + 4 * q) = make_float4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
}

// shared-memory row stride (floats) of a staged degree block of BL floats: an odd number of 16-byte units, so the
// 128-bit stores of 8 consecutive lanes fall into distinct bank groups
__host__ __device__ constexpr int wig_stride(int BL) { return ((BL / 4) & 1) ? BL : BL + 4; }
template <int LMAX, int MMAX>
__host__ __device__ constexpr int wig_stage_floats() {
  int m = 0;
  for (int l = 0; l <= LMAX; ++l) {
    const int s = wig_stride(Tab<LMAX, MMAX>::nrows(l) * Tab<LMAX, MMAX>::rowlen(l));
    m = s > m ? s : m;
  }
  return m;
}

// copy the staged block of degree l of the warp's 32 edges to the Wigner store, 16 bytes per lane and coalesced
template <int LMAX, int MMAX, int l>
__device__ __forceinline__ void wigner_flush(const float* __restrict__ stage, float* __restrict__ wig, int e0, int E,
                                             int lane) {
  using T = Tab<LMAX, MMAX>;
  constexpr int BL = T::nrows(l) * T::rowlen(l), UPB = BL / 4, S = wig_stride(BL);
  __syncwarp();
#pragma unroll 5
  for (int u = lane; u < 32 * UPB; u += 32) {
    const int ed = u / UPB, q = u - ed * UPB;
    if (e0 + ed < E) {
      const float4 v = *reinterpret_cast<const float4*>(stage + ed * S + 4 * q);
      *reinterpret_cast<float4*>(wig + (size_t)(e0 + ed) * T::WSZ + T::woff(l) + 4 * q) = v;
    }
  }
  __syncwarp();
}

// STAGED: out = the warp's stage (each lane writes its edge's rows at lane * stride), flushed after every degree;
// otherwise out = the edge's slot in the Wigner store.
template <int LMAX, int MMAX, bool STAGED, int l, int mp, int NA>
__device__ __forceinline__ void wigner_rows(const float (&cA)[NA], const float (&sA)[NA], const float (&cB)[NA],
                                            const float (&sB)[NA], const float (&cG)[NA], const float (&sG)[NA],
                                            float* __restrict__ out, float* __restrict__ wig, int e0, int E, int lane) {
  using T = Tab<LMAX, MMAX>;
  if constexpr (l <= LMAX) {
    constexpr int S = wig_stride(T::nrows(l) * T::rowlen(l));
    float* row = STAGED ? out + lane * S + (mp + T::mm(l)) * T::rowlen(l) : out + T::wrow(l, mp);
    if constexpr (l == 0) {
      *reinterpret_cast<float4*>(row) = make_float4(1.0f * cJ[0] * cJ[0], 0.f, 0.f, 0.f);
    } else {
      wigner_row<l, mp>(cA, sA, cB, sB, cG, sG, row);
    }
    if constexpr (mp < T::mm(l)) {
      wigner_rows<LMAX, MMAX, STAGED, l, mp + 1>(cA, sA, cB, sB, cG, sG, out, wig, e0, E, lane);
    } else {
      if constexpr (STAGED) wigner_flush<LMAX, MMAX, l>(out, wig, e0, E, lane);
      wigner_rows<LMAX, MMAX, STAGED, l + 1, -T::mm(l + 1)>(cA, sA, cB, sB, cG, sG, out, wig, e0, E, lane);
    }
  }
}

constexpr int WIG_WARPS = 2;

template <int LMAX, int MMAX, bool STAGED>
__global__ void __launch_bounds__(32 * WIG_WARPS) wigner_kernel(const float* __restrict__ rot, float* __restrict__ wig,
                                                                int E) {
  using T = Tab<LMAX, MMAX>;
  static_assert(T::WSZ % 4 == 0, "16-byte rows");
  __shared__ __align__(16) float stage[STAGED ? WIG_WARPS * 32 * wig_stage_floats<LMAX, MMAX>() : 4];
  const int e_raw = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (!STAGED && e_raw >= E) return;
  const int e = e_raw < E ? e_raw : E - 1;  // staged: lanes past the end compute a duplicate and store nothing
  const float* R = rot + (size_t)e * 9;
  const float R0 = R[0], R1 = R[1], R2 = R[2], R4 = R[4], R6 = R[6], R7 = R[7], R8 = R[8];
  // x = R @ (0,1,0): second column (EF_layers.py:509)
  float x0 = R1, x1 = R4, x2 = R7;
  const float nrm = fmaxf(sqrtf(x0 * x0 + x1 * x1 + x2 * x2), 1e-12f);
  x0 = fminf(fmaxf(x0 / nrm, -1.f), 1.f);
  x1 = fminf(fmaxf(x1 / nrm, -1.f), 1.f);
  x2 = fminf(fmaxf(x2 / nrm, -1.f), 1.f);
  const float beta = acosf(x1);
  const float alpha = atan2f(x0, x2);
  // first row of (Ry(alpha) Rx(beta))^T R  ->  gamma (EF_layers.py:511-517)
  const float ca1 = cosf(alpha), sa1 = sinf(alpha);
  const float r00 = ca1 * R0 - sa1 * R6;
  const float r02 = ca1 * R2 - sa1 * R8;
  const float gamma = atan2f(r02, r00);
  float cA[LMAX + 1], sA[LMAX + 1], cB[LMAX + 1], sB[LMAX + 1], cG[LMAX + 1], sG[LMAX + 1];
#pragma unroll
  for (int k = 0; k <= LMAX; ++k) {  // cos / sin of (k * angle), as the reference's _z_rot_mat forms them
    sincosf((float)k * alpha, &sA[k], &cA[k]);
    sincosf((float)k * beta, &sB[k], &cB[k]);
    sincosf((float)k * gamma, &sG[k], &cG[k]);
  }
  float* out = STAGED ? stage + warp * 32 * wig_stage_floats<LMAX, MMAX>() : wig + (size_t)e * T::WSZ;
  wigner_rows<LMAX, MMAX, STAGED, 0, 0>(cA, sA, cB, sB, cG, sG, out, wig, e_raw - lane, E, lane);
}

template <int LMAX, int MMAX>
static void wigner_launch(const float* rot, float* wig, int E, bool staged, cudaStream_t s) {
  const int T = 32 * WIG_WARPS;
  if (staged) wigner_kernel<LMAX, MMAX, true><<<(E + T - 1) / T, T, 0, s>>>(rot, wig, E);
  else wigner_kernel<LMAX, MMAX, false><<<(E + T - 1) / T, T, 0, s>>>(rot, wig, E);
}

}  // namespace efl

extern "C" int efl_wigner_from_rotmat(int lmax, int mmax, const float* rot, const float* jd, float* wig, int E,
                                      void* stream) {
  if (E <= 0) return EFL_OK;
  cudaStream_t s = (cudaStream_t)stream;
  if (!(mmax == 2 && (lmax == 4 || lmax == 6 || lmax == 8))) {
    efl_set_error("efl_wigner_from_rotmat: unsupported (lmax, mmax); built for (6,2), (8,2), (4,2)");
    return EFL_ERR_UNSUPPORTED;
  }
  // J table -> constant bank, stream-ordered (a prefix of the lmax = 8 table; every caller writes the same values)
  cudaError_t err = cudaMemcpyToSymbolAsync(efl::cJ, jd, sizeof(float) * efl::wig_joff(lmax + 1), 0,
                                            cudaMemcpyDeviceToDevice, s);
  if (err != cudaSuccess) {
    efl_set_error(cudaGetErrorString(err));
    return (int)err;
  }
  static const bool staged = [] {  // EFL_WIGNER_DIRECT=1: per-thread stores straight to HBM (A/B switch)
    const char* v = getenv("EFL_WIGNER_DIRECT");
    return !(v && v[0] == '1');
  }();
  if (lmax == 6) efl::wigner_launch<6, 2>(rot, wig, E, staged, s);
  else if (lmax == 8) efl::wigner_launch<8, 2>(rot, wig, E, staged, s);
  else efl::wigner_launch<4, 2>(rot, wig, E, staged, s);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
