// fp32 SIMT GEMMs for the SO(2) linear maps: the accuracy ("fp32") mode of the path.
//
//   NT : C[i, n] = sum_k A[i, k] * B[n, k]            forward and dgrad (with a transposed weight copy)
//   TN : C[n, k] (+)= sum_i A[i, n] * B[i, k]          wgrad (reduction over edges)
//
// NT epilogues (reference EF_layers.py:825-834, 721-729):
//   0 PLAIN   : C = acc + bias
//   1 SPLIT   : columns [0, nsplit) -> C0 (fp32, "extra" m=0 outputs: alpha inputs and gate),
//               columns [nsplit, N) -> C   (hidden m=0 rows)
//   2 COMPLEX : SO2_m_Convolution combine.  Rows come in pairs (2e: +m "real", 2e+1: -m "imag") and
//               weight rows are interleaved on the host (2q: real weight q, 2q+1: imag weight q):
//                 out[2e,   q] = acc[2e, 2q]   - acc[2e+1, 2q+1]
//                 out[2e+1, q] = acc[2e+1, 2q] + acc[2e,   2q+1]
// 64x64 tile, 16-deep k step, 256 threads, 4x4 micro-tile with even-aligned rows/cols so the
// complex combine is thread-local.
#include "efl_common.cuh"

namespace efl {

constexpr int BM = 64, BN = 64, BK = 16;

template <typename TA, typename TC, int EPI>
__global__ void __launch_bounds__(256) gemm_nt_kernel(const TA* __restrict__ A, long long lda,
                                                      const TA* __restrict__ B, long long ldb,
                                                      const float* __restrict__ bias, TC* __restrict__ C,
                                                      long long ldc, float* __restrict__ C0, long long ldc0,
                                                      int nsplit, int M, int N, int K, long long az = 0,
                                                      long long bz = 0, long long cz = 0) {
  __shared__ float sA[BK][BM + 4];
  __shared__ float sB[BK][BN + 4];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  if (gridDim.z > 1) {
    // batched over coefficient index z (SO3_LinearV2): operand z of A/C, weight block of degree l(z)
    const int z = blockIdx.z;
    int l = 0;
    while ((l + 1) * (l + 1) <= z) ++l;
    A += z * az;
    B += l * bz;
    C += z * cz;
    if (z > 0) bias = nullptr;
  }
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < K; k0 += BK) {
    // 64 rows x 16 k per operand: 1024 elements, 4 per thread
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int idx = threadIdx.x + t * 256;
      const int r = idx / BK, kk = idx % BK;
      const int gm = m0 + r, gn = n0 + r, gk = k0 + kk;
      sA[kk][r] = (gm < M && gk < K) ? to_f<TA>(A[(long long)gm * lda + gk]) : 0.f;
      sB[kk][r] = (gn < N && gk < K) ? to_f<TA>(B[(long long)gn * ldb + gk]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = sA[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = sB[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  const int rm = m0 + ty * 4, cn = n0 + tx * 4;
  if (EPI == 2) {
    // N is the interleaved width (2 * half); rows/cols are even aligned
#pragma unroll
    for (int ip = 0; ip < 2; ++ip) {
#pragma unroll
      for (int jp = 0; jp < 2; ++jp) {
        const int r = rm + 2 * ip, c = cn + 2 * jp;
        if (r + 1 < M && c + 1 < N) {
          const float re = acc[2 * ip][2 * jp] - acc[2 * ip + 1][2 * jp + 1];
          const float im = acc[2 * ip + 1][2 * jp] + acc[2 * ip][2 * jp + 1];
          C[(long long)r * ldc + c / 2] = from_f<TC>(re);
          C[(long long)(r + 1) * ldc + c / 2] = from_f<TC>(im);
        }
      }
    }
    return;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = rm + i;
    if (r >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = cn + j;
      if (c >= N) continue;
      const float v = acc[i][j] + (bias ? bias[c] : 0.f);
      if (EPI == 1 && c < nsplit) C0[(long long)r * ldc0 + c] = v;
      else C[(long long)r * ldc + (c - (EPI == 1 ? nsplit : 0))] = from_f<TC>(v);
    }
  }
}

// C[n, k] (+)= sum_{i in slice} A[i, n] * B[i, k]; deterministic split over i via blockIdx.z partials
template <typename TA>
__global__ void __launch_bounds__(256) gemm_tn_kernel(const TA* __restrict__ A, long long lda,
                                                      const TA* __restrict__ B, long long ldb,
                                                      float* __restrict__ P /* [splits, N1, N2] */, int Mred,
                                                      int N1, int N2, int rows_per_split, long long az = 0,
                                                      long long bz = 0) {
  __shared__ float sA[BK][BM + 4];
  __shared__ float sB[BK][BN + 4];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int a0 = blockIdx.y * BM, b0 = blockIdx.x * BN;
  int i_begin = blockIdx.z * rows_per_split;
  int i_end = min(Mred, i_begin + rows_per_split);
  if (az != 0) {  // batched over z: full reduction per batch entry, partial z goes to P[z]
    A += blockIdx.z * az;
    B += blockIdx.z * bz;
    i_begin = 0;
    i_end = Mred;
  }
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int i0 = i_begin; i0 < i_end; i0 += BK) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int idx = threadIdx.x + t * 256;
      const int kk = idx / BM, r = idx % BM;  // consecutive threads -> consecutive columns (coalesced)
      const int gi = i0 + kk;
      sA[kk][r] = (gi < i_end && a0 + r < N1) ? to_f<TA>(A[(long long)gi * lda + a0 + r]) : 0.f;
      sB[kk][r] = (gi < i_end && b0 + r < N2) ? to_f<TA>(B[(long long)gi * ldb + b0 + r]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = sA[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = sB[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* Pz = P + (size_t)blockIdx.z * N1 * N2;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = a0 + ty * 4 + i;
    if (r >= N1) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = b0 + tx * 4 + j;
      if (c < N2) Pz[(size_t)r * N2 + c] = acc[i][j];
    }
  }
}

// C[idx] = (accumulate ? C[idx] : 0) + sum_z P[z, idx], fixed order
__global__ void splitk_reduce_kernel(const float* __restrict__ P, int splits, long long n, float* __restrict__ C,
                                     long long ldc, int N2, int accumulate) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float acc = 0.f;
  for (int z = 0; z < splits; ++z) acc += P[(size_t)z * n + i];
  const long long r = i / N2, c = i % N2;
  float* o = C + r * ldc + c;
  *o = accumulate ? (*o + acc) : acc;
}

// out[c] (+)= sum_i X[i, c], deterministic two-stage column sum (bias gradients)
template <typename T>
__global__ void colsum_partial_kernel(const T* __restrict__ X, long long ld, int rows, int cols, int rows_per_block,
                                      float* __restrict__ P) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(rows, r0 + rows_per_block);
  float acc = 0.f;
  for (int r = r0; r < r1; ++r) acc += to_f<T>(X[(long long)r * ld + c]);
  P[(size_t)blockIdx.y * cols + c] = acc;
}

// dW[l] = sum_{z in degree l} P[z]
__global__ void so3_wgrad_reduce_kernel(const float* __restrict__ P, float* __restrict__ dW, int lmax, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  for (int l = 0; l <= lmax; ++l) {
    float acc = 0.f;
    for (int z = l * l; z < (l + 1) * (l + 1); ++z) acc += P[(size_t)z * n + i];
    dW[(size_t)l * n + i] = acc;
  }
}

}  // namespace efl

using namespace efl;

// SO3_LinearV2 (reference EF_layers.py:655-671): out[n, i, :] = x[n, i, :] W[l(i)]^T (+ bias on i = 0).
// x [N, L, Cin], W [lmax+1, Cout, Cin], out [N, L, Cout]; fp32.
extern "C" int efl_so3_linear_fwd(const float* x, const float* W, const float* bias, float* out, int N, int lmax,
                                  int Cin, int Cout, void* stream) {
  if (N <= 0) return EFL_OK;
  const int L = (lmax + 1) * (lmax + 1);
  dim3 grid((Cout + BN - 1) / BN, (N + BM - 1) / BM, L), block(256);
  gemm_nt_kernel<float, float, 0><<<grid, block, 0, (cudaStream_t)stream>>>(
      x, (long long)L * Cin, W, Cin, bias, out, (long long)L * Cout, nullptr, 0, 0, N, Cout, Cin, Cin,
      (long long)Cout * Cin, Cout);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// d x[n, i, :] = d out[n, i, :] W[l(i)];  Wt = W transposed per degree: [lmax+1, Cin, Cout]
extern "C" int efl_so3_linear_bwd_data(const float* dout, const float* Wt, float* dx, int N, int lmax, int Cin,
                                       int Cout, void* stream) {
  if (N <= 0) return EFL_OK;
  const int L = (lmax + 1) * (lmax + 1);
  dim3 grid((Cin + BN - 1) / BN, (N + BM - 1) / BM, L), block(256);
  gemm_nt_kernel<float, float, 0><<<grid, block, 0, (cudaStream_t)stream>>>(
      dout, (long long)L * Cout, Wt, Cout, nullptr, dx, (long long)L * Cin, nullptr, 0, 0, N, Cin, Cout, Cout,
      (long long)Cout * Cin, Cin);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// d W[l] = sum_{i in l} sum_n d out[n, i, :]^T x[n, i, :];  workspace: L * Cout * Cin floats
extern "C" int efl_so3_linear_bwd_weight(const float* dout, const float* x, float* dW, float* workspace, int N,
                                         int lmax, int Cin, int Cout, void* stream) {
  const int L = (lmax + 1) * (lmax + 1);
  const long long n = (long long)Cout * Cin;
  cudaStream_t s = (cudaStream_t)stream;
  if (N <= 0) {
    cudaMemsetAsync(dW, 0, sizeof(float) * n * (lmax + 1), s);
    return EFL_OK;
  }
  dim3 grid((Cin + BN - 1) / BN, (Cout + BM - 1) / BM, L), block(256);
  gemm_tn_kernel<float><<<grid, block, 0, s>>>(dout, (long long)L * Cout, x, (long long)L * Cin, workspace, N, Cout,
                                              Cin, N, Cout, Cin);
  EFL_CHECK_LAUNCH();
  so3_wgrad_reduce_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(workspace, dW, lmax, n);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

extern "C" int efl_gemm_nt_f32(int epilogue, const float* A, long long lda, const float* B, long long ldb,
                               const float* bias, float* C, long long ldc, float* C0, long long ldc0, int nsplit,
                               int M, int N, int K, void* stream) {
  if (M <= 0 || N <= 0) return EFL_OK;
  dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM), block(256);
  cudaStream_t s = (cudaStream_t)stream;
  if (epilogue == 0) gemm_nt_kernel<float, float, 0><<<grid, block, 0, s>>>(A, lda, B, ldb, bias, C, ldc, C0, ldc0, nsplit, M, N, K);
  else if (epilogue == 1) gemm_nt_kernel<float, float, 1><<<grid, block, 0, s>>>(A, lda, B, ldb, bias, C, ldc, C0, ldc0, nsplit, M, N, K);
  else if (epilogue == 2) {
    EFL_REQUIRE(M % 2 == 0 && N % 2 == 0, "efl_gemm_nt_f32: complex epilogue needs even M and N");
    gemm_nt_kernel<float, float, 2><<<grid, block, 0, s>>>(A, lda, B, ldb, bias, C, ldc, C0, ldc0, nsplit, M, N, K);
  } else {
    efl_set_error("efl_gemm_nt_f32: unknown epilogue");
    return EFL_ERR_ARG;
  }
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// workspace: splits * N1 * N2 floats, splits = efl_gemm_tn_splits(Mred)
extern "C" int efl_gemm_tn_splits(int Mred) {
  int s = (Mred + 8191) / 8192;
  return s < 1 ? 1 : (s > 64 ? 64 : s);
}

extern "C" int efl_gemm_tn_f32(const float* A, long long lda, const float* B, long long ldb, float* C,
                               long long ldc, int accumulate, float* workspace, int Mred, int N1, int N2,
                               void* stream) {
  if (N1 <= 0 || N2 <= 0) return EFL_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int splits = efl_gemm_tn_splits(Mred);
  int rows = (Mred + splits - 1) / splits;
  rows = ((rows + BK - 1) / BK) * BK;
  if (rows < BK) rows = BK;
  dim3 grid((N2 + BN - 1) / BN, (N1 + BM - 1) / BM, splits), block(256);
  gemm_tn_kernel<float><<<grid, block, 0, s>>>(A, lda, B, ldb, workspace, Mred, N1, N2, rows);
  EFL_CHECK_LAUNCH();
  const long long n = (long long)N1 * N2;
  splitk_reduce_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(workspace, splits, n, C, ldc, N2, accumulate);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// bf16 rows with 16-byte alignment: each lane owns 8 adjacent columns (one 16-byte load per row), the four warps of
// a block take rows r0 + w, r0 + w + 4, ... (four loads in flight per lane) and are combined in warp order, so the
// sum order is fixed.  Same partial layout as colsum_partial_kernel.
__global__ void __launch_bounds__(128)
colsum_partial_bf16x8_kernel(const __nv_bfloat16* __restrict__ X, long long ld, int rows, int cols, int rows_per_block,
                             float* __restrict__ P) {
  __shared__ float part[4][256];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c0 = blockIdx.x * 256 + lane * 8;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(rows, r0 + rows_per_block);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (c0 < cols) {
    const __nv_bfloat16* base = X + c0;
    int r = r0 + w;
    for (; r + 12 < r1; r += 16) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldg(reinterpret_cast<const uint4*>(base + (long long)(r + 4 * u) * ld));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v[u]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 f = __bfloat1622float2(h[j]);
          acc[2 * j] += f.x;
          acc[2 * j + 1] += f.y;
        }
      }
    }
    for (; r < r1; r += 4) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(base + (long long)r * ld));
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = __bfloat1622float2(h[j]);
        acc[2 * j] += f.x;
        acc[2 * j + 1] += f.y;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) part[w][lane * 8 + j] = acc[j];
  __syncthreads();
  for (int c = threadIdx.x; c < 256; c += 128) {
    const int col = blockIdx.x * 256 + c;
    if (col < cols) P[(size_t)blockIdx.y * cols + col] = ((part[0][c] + part[1][c]) + part[2][c]) + part[3][c];
  }
}

// workspace: (ceil(rows / 512) + 1) * cols floats
extern "C" int efl_colsum(int bf16, const void* X, long long ld, int rows, int cols, float* out, int accumulate,
                          float* workspace, void* stream) {
  if (cols <= 0) return EFL_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int rpb = 512;
  const int nb = rows > 0 ? (rows + rpb - 1) / rpb : 1;
  dim3 grid((cols + 127) / 128, nb), block(128);
  if (bf16 && cols % 8 == 0 && ld % 8 == 0 && ((uintptr_t)X & 15) == 0) {
    dim3 g8((cols + 255) / 256, nb);
    colsum_partial_bf16x8_kernel<<<g8, 128, 0, s>>>((const __nv_bfloat16*)X, ld, rows, cols, rpb, workspace);
  } else if (bf16)
    colsum_partial_kernel<__nv_bfloat16><<<grid, block, 0, s>>>((const __nv_bfloat16*)X, ld, rows, cols, rpb, workspace);
  else colsum_partial_kernel<float><<<grid, block, 0, s>>>((const float*)X, ld, rows, cols, rpb, workspace);
  EFL_CHECK_LAUNCH();
  splitk_reduce_kernel<<<(cols + 255) / 256, 256, 0, s>>>(workspace, nb, cols, out, cols, cols, accumulate);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Column sum for callers that cannot allocate device memory.
//
// efl_colsum needs (ceil(rows / 512) + 1) * cols floats of scratch for its per-block partial sums, and the caller has
// to bring them.  Some callers have nowhere to get them from: code running inside a memory pool that is already
// sized and frozen (a serving process after start-up, a plugin that was handed fixed buffers by its host program), or
// code between cudaStreamBeginCapture and cudaStreamEndCapture, where cudaMalloc is an error.  For those the partial
// sums live in the fixed array below, which is part of the library image: every GPU that loads libefl.so gives up
// EFL_COLSUM_STATIC_MAX_FLOATS * 4 bytes (4 MiB) for it from module load until the process exits.
//
// `g_colsum_static` as written in HOST code is only the host-side handle of the symbol.  It is not a device address,
// so it must never be handed to a kernel; the device address differs per GPU and is looked up with
// cudaGetSymbolAddress while that GPU is current (colsum_static_scratch below), once per device.
#include <cstdio>
#include <mutex>

constexpr long long kColsumStaticFloats = 1ll << 20;
static __device__ __align__(256) float g_colsum_static[kColsumStaticFloats];

// Device address of g_colsum_static on the CURRENT device.  The first call on a device asks the runtime for it and
// keeps the answer; later calls only read the table.
static int colsum_static_scratch(float** out) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static float* addr[kMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    efl_set_error(cudaGetErrorString(e));
    return (int)e;
  }
  EFL_REQUIRE(dev >= 0 && dev < kMaxDevices, "efl_colsum_static: device ordinal outside [0, 64)");
  std::lock_guard<std::mutex> lock(mu);
  if (addr[dev] == nullptr) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_colsum_static);
    if (e != cudaSuccess || p == nullptr) {
      efl_set_error(e != cudaSuccess ? cudaGetErrorString(e) : "efl_colsum_static: null address for the scratch symbol");
      return e != cudaSuccess ? (int)e : EFL_ERR_ARG;
    }
    addr[dev] = p;
  }
  *out = addr[dev];
  return EFL_OK;
}

// Same result as efl_colsum (same kernels, same summation order, so bit-identical), without a workspace argument and
// without any allocation: no cudaMalloc / cudaFree / allocating library call on this path.
//   * Limit: (ceil(rows / 512) + 1) * cols <= EFL_COLSUM_STATIC_MAX_FLOATS (2^20); anything larger is refused with
//     EFL_ERR_ARG and a message that names both numbers -- use efl_colsum with your own workspace then.
//   * The first call on each device looks up the array's device address (a runtime query, not an allocation).  Make
//     that first call with the right device current and before any stream capture begins.
//   * `stream` must belong to the current device.
//   * All calls on one GPU share the one array and it keeps whatever the last call left there.  The partial-sum
//     kernel overwrites every slot the reduction reads, so nothing has to be cleared in between; but two calls whose
//     kernels can run at the same time (different streams without an ordering between them) race on it.
extern "C" int efl_colsum_static(int bf16, const void* X, long long ld, int rows, int cols, float* out,
                                 int accumulate, void* stream) {
  EFL_REQUIRE(rows >= 0 && cols >= 0, "efl_colsum_static: negative shape");
  if (cols == 0) return EFL_OK;
  const long long need = ((long long)(rows + 511) / 512 + 1) * (long long)cols;
  if (need > kColsumStaticFloats) {
    char msg[192];
    snprintf(msg, sizeof msg,
             "efl_colsum_static: needs %lld scratch floats, the built-in array holds %lld; call efl_colsum with a "
             "caller-supplied workspace instead",
             need, kColsumStaticFloats);
    efl_set_error(msg);
    return EFL_ERR_ARG;
  }
  float* scratch = nullptr;
  const int rc = colsum_static_scratch(&scratch);
  if (rc != EFL_OK) return rc;
  return efl_colsum(bf16, X, ld, rows, cols, out, accumulate, scratch, stream);
}
