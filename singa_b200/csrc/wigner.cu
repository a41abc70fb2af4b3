// Wigner-D rows from edge frames.
//
// Replaces SO3_Rotation.set_wigner / RotationToWignerDMatrix / wigner_D / _z_rot_mat
// (reference model/EF_layers.py:485-528, 2207-2229).  The reference materialises two
// dense [E, L, L] tensors; only the rows with |m'| <= mmax of each diagonal block are
// ever used (rotate: wigner[:, mask, :], rotate_inv: wigner_inv[:, :, mask]), so this
// kernel emits exactly those rows (235 values for lmax 6, mmax 2), each padded to 16 bytes: WSZ = 296 floats/edge.
//
// D^l = Z(alpha) J_l Z(beta) J_l Z(gamma) with Z(a)[m,m] = cos(m a), Z(a)[m,-m] = -sin(m a).
// One warp per edge; lane j owns kept row j.
#include "efl_common.cuh"

namespace efl {

template <int l>
__device__ __forceinline__ void wigner_row(int mp, const float* __restrict__ J, const float (*sc)[2],
                                           const float (*scb)[2], const float (*scg)[2],
                                           float* __restrict__ out) {
  constexpr int n = 2 * l + 1;
  float r2[n], r3[n], r4[n];
  const int am = mp < 0 ? -mp : mp;
  // signed order: cos(m a) is even, sin(m a) odd
  const float ca = sc[am][0];
  const float sa = mp < 0 ? -sc[am][1] : sc[am][1];
#pragma unroll
  for (int k = 0; k < n; ++k) {
    float v = ca * J[(mp + l) * n + k];
    if (mp != 0) v -= sa * J[(-mp + l) * n + k];
    r2[k] = v;
  }
#pragma unroll
  for (int k = 0; k < n; ++k) {
    const int nn = k - l;
    const int an = nn < 0 ? -nn : nn;
    if (nn == 0) {
      r3[k] = r2[k];
    } else {
      const float sb = nn < 0 ? -scb[an][1] : scb[an][1];
      r3[k] = r2[k] * scb[an][0] + r2[n - 1 - k] * sb;
    }
  }
#pragma unroll
  for (int k = 0; k < n; ++k) {
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < n; ++i) acc = fmaf(r3[i], J[i * n + k], acc);
    r4[k] = acc;
  }
#pragma unroll
  for (int k = 0; k < n; ++k) {
    const int nn = k - l;
    const int an = nn < 0 ? -nn : nn;
    if (nn == 0) {
      out[k] = r4[k];
    } else {
      const float sg = nn < 0 ? -scg[an][1] : scg[an][1];
      out[k] = r4[k] * scg[an][0] + r4[n - 1 - k] * sg;
    }
  }
}

template <int LMAX, int MMAX>
__global__ void __launch_bounds__(128) wigner_kernel(const float* __restrict__ rot, const float* __restrict__ jd,
                                                     float* __restrict__ wig, int E) {
  using T = Tab<LMAX, MMAX>;
  constexpr int JSZ = (LMAX + 1) * (2 * LMAX + 1) * (2 * LMAX + 3) / 3;  // sum (2l+1)^2
  __shared__ float sJ[JSZ];
  __shared__ float sSC[4][3][LMAX + 1][2];
  for (int i = threadIdx.x; i < JSZ; i += blockDim.x) sJ[i] = jd[i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int e = blockIdx.x * 4 + warp;
  if (e >= E) return;
  const float* R = rot + (size_t)e * 9;
  // x = R @ (0,1,0): second column (EF_layers.py:509)
  float x0 = R[1], x1 = R[4], x2 = R[7];
  float nrm = fmaxf(sqrtf(x0 * x0 + x1 * x1 + x2 * x2), 1e-12f);
  x0 = fminf(fmaxf(x0 / nrm, -1.f), 1.f);
  x1 = fminf(fmaxf(x1 / nrm, -1.f), 1.f);
  x2 = fminf(fmaxf(x2 / nrm, -1.f), 1.f);
  const float beta = acosf(x1);
  const float alpha = atan2f(x0, x2);
  // first row of (Ry(alpha) Rx(beta))^T R  ->  gamma (EF_layers.py:511-517)
  const float ca = cosf(alpha), sa = sinf(alpha);
  const float r00 = ca * R[0] - sa * R[6];
  const float r02 = ca * R[2] - sa * R[8];
  const float gamma = atan2f(r02, r00);
  if (lane < 3 * (LMAX + 1)) {
    const int which = lane / (LMAX + 1), k = lane % (LMAX + 1);
    const float ang = which == 0 ? alpha : (which == 1 ? beta : gamma);
    const float a = (float)k * ang;
    sSC[warp][which][k][0] = cosf(a);
    sSC[warp][which][k][1] = sinf(a);
  }
  __syncwarp();
  float* out = wig + (size_t)e * T::WSZ;
  for (int j = lane; j < T::M; j += 32) {
    // decode (l, m') of kept row j in l-primary order
    int l = 0, base = 0;
    while (base + T::nrows(l) <= j) {
      base += T::nrows(l);
      ++l;
    }
    const int mp = j - base - T::mm(l);
    int joff = 0;
    for (int k = 0; k < l; ++k) joff += (2 * k + 1) * (2 * k + 1);
    const float* J = sJ + joff;
    float* o = out + T::wrow(l, mp);
    for (int k = 2 * l + 1; k < T::rowlen(l); ++k) o[k] = 0.f;  // row padding (16-byte rows)
    const float(*sA)[2] = sSC[warp][0];
    const float(*sB)[2] = sSC[warp][1];
    const float(*sG)[2] = sSC[warp][2];
    switch (l) {
      case 0: o[0] = 1.0f * J[0] * J[0]; break;
      case 1: wigner_row<1>(mp, J, sA, sB, sG, o); break;
      case 2: wigner_row<2>(mp, J, sA, sB, sG, o); break;
      case 3: if constexpr (LMAX >= 3) wigner_row<3>(mp, J, sA, sB, sG, o); break;
      case 4: if constexpr (LMAX >= 4) wigner_row<4>(mp, J, sA, sB, sG, o); break;
      case 5: if constexpr (LMAX >= 5) wigner_row<5>(mp, J, sA, sB, sG, o); break;
      case 6: if constexpr (LMAX >= 6) wigner_row<6>(mp, J, sA, sB, sG, o); break;
      case 7: if constexpr (LMAX >= 7) wigner_row<7>(mp, J, sA, sB, sG, o); break;
      case 8: if constexpr (LMAX >= 8) wigner_row<8>(mp, J, sA, sB, sG, o); break;
      default: break;
    }
  }
}

}  // namespace efl

extern "C" int efl_wigner_from_rotmat(int lmax, int mmax, const float* rot, const float* jd, float* wig, int E,
                                      void* stream) {
  if (E <= 0) return EFL_OK;
  cudaStream_t s = (cudaStream_t)stream;
  dim3 grid((E + 3) / 4), block(128);
  if (lmax == 6 && mmax == 2) efl::wigner_kernel<6, 2><<<grid, block, 0, s>>>(rot, jd, wig, E);
  else if (lmax == 8 && mmax == 2) efl::wigner_kernel<8, 2><<<grid, block, 0, s>>>(rot, jd, wig, E);
  else if (lmax == 4 && mmax == 2) efl::wigner_kernel<4, 2><<<grid, block, 0, s>>>(rot, jd, wig, E);
  else {
    efl_set_error("efl_wigner_from_rotmat: unsupported (lmax, mmax); built for (6,2), (8,2), (4,2)");
    return EFL_ERR_UNSUPPORTED;
  }
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
