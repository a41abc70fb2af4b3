// Destination-sorted CSR of a caller-supplied (unsorted) edge_index, plus the source-side grouping.
//
// Replaces the host-side bookkeeping the reference leaves to index_add_/index_select on the raw edge list
// (EF_layers.py:326-328 gather by edge_index[0]/[1], :342-351 index_add_ over edge_index[1], :1180 softmax index):
// all kernels of the path consume edges grouped by destination, so the edge list is sorted ONCE per edge_index:
//   perm        [E]   sorted position -> original edge id; stable, i.e. ascending edge id inside a destination, which
//                     is the summation order every deterministic segment reduction of the path uses
//   rowptr      [num_dst+1]  segment offsets over destinations
//   src_sorted, dst_sorted [E]  end points in sorted order;  dst_orig [E] destinations in original order (int32)
//   src_ptr     [num_src+1], src_idx [E]  sorted positions grouped by source (stable): backward of the gather
// HBM-bound integer work: one conversion+histogram pass (integer atomics: order-independent), two in-place exclusive
// scans and two stable LSD radix sorts over only the significant key bits (cub).  No hidden allocation: temp and
// scratch come from the caller.  Out-of-range node ids are reported through *status (device int), never dereferenced.
#include <cub/cub.cuh>

#include "efl_common.cuh"

namespace efl {
namespace csr {

__global__ void __launch_bounds__(256) convert_count_kernel(const long long* __restrict__ src64,
                                                            const long long* __restrict__ dst64, int E, int num_src,
                                                            int num_dst, int* __restrict__ src32,
                                                            int* __restrict__ dst32, int* __restrict__ iota,
                                                            int* __restrict__ dst_cnt, int* __restrict__ src_cnt,
                                                            int* __restrict__ status) {
  const int stride = gridDim.x * blockDim.x;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < E; e += stride) {
    const long long s = src64[e], d = dst64[e];
    const bool ok = s >= 0 && s < num_src && d >= 0 && d < num_dst;
    if (!ok) {
      atomicExch(status, 1);
      src32[e] = 0;
      dst32[e] = 0;
      iota[e] = e;
      continue;
    }
    src32[e] = (int)s;
    dst32[e] = (int)d;
    iota[e] = e;
    atomicAdd(dst_cnt + (int)d, 1);
    atomicAdd(src_cnt + (int)s, 1);
  }
}

__global__ void __launch_bounds__(256) gather_kernel(const int* __restrict__ perm, const int* __restrict__ src32, int E,
                                                     int* __restrict__ src_sorted) {
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < E; i += stride) src_sorted[i] = __ldg(src32 + __ldg(perm + i));
}

static int key_bits(int n) {
  int b = 1;
  while (b < 31 && (1ll << b) < (long long)n) ++b;
  return b;
}

}  // namespace csr
}  // namespace efl

using namespace efl;

#define CSR_CUDA(expr)                                           \
  do {                                                          \
    cudaError_t e__ = (expr);                                   \
    if (e__ != cudaSuccess) {                                   \
      efl_set_error(cudaGetErrorString(e__));                   \
      return (int)e__;                                          \
    }                                                           \
  } while (0)

extern "C" long long efl_csr_temp_bytes(long long E, long long max_nodes) {
  size_t a = 0, b = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, a, (const int*)nullptr, (int*)nullptr, (const int*)nullptr, (int*)nullptr,
                                  (int)(E > 0 ? E : 1));
  cub::DeviceScan::ExclusiveSum(nullptr, b, (const int*)nullptr, (int*)nullptr, (int)(max_nodes + 1));
  return (long long)(a > b ? a : b) + 256;
}

// edge_index rows: src64 = row 0 (sources), dst64 = row 1 (targets); both [E] int64 device arrays.
// scratch: 3*E ints.  status: one device int, set to 1 when an id is out of range (left untouched otherwise).
extern "C" int efl_csr_from_edge_index(const long long* src64, const long long* dst64, int E, int num_src, int num_dst,
                                       int* perm, int* rowptr, int* src_sorted, int* dst_sorted, int* dst_orig,
                                       int* src_ptr, int* src_idx, int* scratch, void* temp, long long temp_bytes,
                                       int* status, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  EFL_REQUIRE(E >= 0 && num_src >= 0 && num_dst >= 0, "efl_csr_from_edge_index: negative size");
  CSR_CUDA(cudaMemsetAsync(rowptr, 0, sizeof(int) * (size_t)(num_dst + 1), s));
  CSR_CUDA(cudaMemsetAsync(src_ptr, 0, sizeof(int) * (size_t)(num_src + 1), s));
  size_t tb = (size_t)temp_bytes;
  if (E > 0) {
    int* src32 = scratch;
    int* iota = scratch + E;
    int* keys_out = scratch + 2 * (size_t)E;
    const int blocks = E < 148 * 8 * 256 ? (E + 255) / 256 : 148 * 8;
    csr::convert_count_kernel<<<blocks, 256, 0, s>>>(src64, dst64, E, num_src, num_dst, src32, dst_orig, iota, rowptr,
                                                     src_ptr, status);
    CSR_CUDA(cub::DeviceRadixSort::SortPairs(temp, tb, (const int*)dst_orig, dst_sorted, (const int*)iota, perm, E, 0,
                                             csr::key_bits(num_dst), s));
    csr::gather_kernel<<<blocks, 256, 0, s>>>(perm, src32, E, src_sorted);
    tb = (size_t)temp_bytes;
    CSR_CUDA(cub::DeviceRadixSort::SortPairs(temp, tb, (const int*)src_sorted, keys_out, (const int*)iota, src_idx, E,
                                             0, csr::key_bits(num_src), s));
  }
  tb = (size_t)temp_bytes;
  CSR_CUDA(cub::DeviceScan::ExclusiveSum(temp, tb, rowptr, rowptr, num_dst + 1, s));
  tb = (size_t)temp_bytes;
  CSR_CUDA(cub::DeviceScan::ExclusiveSum(temp, tb, src_ptr, src_ptr, num_src + 1, s));
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
