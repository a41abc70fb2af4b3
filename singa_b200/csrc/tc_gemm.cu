// bf16 tensor-core GEMMs for the SO(2) linear maps: tcgen05.mma with TMEM accumulators, operands staged by TMA
// (128-byte swizzle) through a 4-stage mbarrier pipeline, persistent warp-specialised CTAs (one per SM):
//   warp 0   TMA producer          warp 1   MMA issuer (one elected lane)
//   warp 2   TMEM allocator        warps 4-7 epilogue (tcgen05.ld -> registers -> fused epilogue -> global)
// Two 256-column TMEM accumulator stages let the epilogue of tile i overlap the MMAs of tile i+1.
//
//   NT : C[i, n] = sum_k A[i, k] B[n, k]     both operands K-major      (forward, dgrad)
//   TN : P[z][n1, n2] = sum_{i in split z} A[i, n1] B[i, n2]   both operands MN-major (wgrad, split over edges,
//        partials reduced in fixed order by splitk_reduce)
// NT epilogues as in gemm_simt.cu (0 plain, 1 split, 2 complex combine) plus 3 = plain with SiLU; outputs bf16 (fp32
// for the split part).
//
// Replaces the cuBLAS mm calls under SO2_Convolution / SO2_m_Convolution (reference EF_layers.py:721-729, 807-875).
#include <string.h>

#include "tc_common.cuh"

namespace efl {
namespace tc {

constexpr int BM = 128;         // UMMA M (rows of the accumulator = TMEM lanes)
constexpr int BK = 64;          // 64 bf16 = 128 bytes = one swizzle span
constexpr int UMMA_K = 16;
constexpr int STAGES = 4;       // ring depth at the largest stage (BN = 256); smaller stages get more, see nstage
constexpr int MAX_STAGES = 8;
constexpr int MAX_BN = 256;
constexpr int A_BYTES = BM * BK * 2;       // 16 KB
constexpr int B_BYTES = MAX_BN * BK * 2;   // 32 KB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int STG_BYTES = 32 * 128;        // per epilogue warp: 32 rows x 128 bytes
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 1024 /*barriers*/ + 8 * STG_BYTES;
static_assert(SMEM_BYTES <= 232448, "shared memory budget (227 KB per CTA)");
constexpr int TMEM_COLS = 512;
constexpr int NUM_THREADS = 384;   // 4 control warps + 8 epilogue warps

struct Params {
  int mode;      // 0 NT, 1 TN
  int M, N, K;   // NT: C is [M, N], reduce K.  TN: M = N1 (rows of C), N = N2, K = reduced rows of this launch
  int BN;        // accumulator columns per tile (multiple of 16, <= 256)
  int epi;
  __nv_bfloat16* C;
  long long ldc;
  float* C0;
  long long ldc0;
  int nsplit;
  const float* bias;
  float* P;      // TN partials [splits, N1, N2]
  int splits;
  int kb_per_split;
  int tma_c, tma_c0;  // output tensor maps valid: full-width blocks leave through TMA stores
  int ring_bytes;     // operand ring size (<= STAGES * STAGE_BYTES): a smaller ring leaves shared memory for the
                      // HBM-bound kernels of another stream to co-reside with the persistent GEMM CTAs
};

// PAIR: the two CTAs of a (2,1,1) cluster work on one 256-row tile with cta_group::2 MMAs: each CTA stages its own
// 128 rows of A and HALF of the B tile, so the L2->SM operand traffic per SM drops by a third at BN = 256 (and the
// shared-memory reads of B halve); accumulators, epilogue and output rows stay per CTA.
template <bool PAIR>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
               const __grid_constant__ CUtensorMap mapC, const __grid_constant__ CUtensorMap mapC0, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.ring_bytes);
  uint64_t* full = bars;                     // [MAX_STAGES]
  uint64_t* empty = bars + MAX_STAGES;       // [MAX_STAGES]
  uint64_t* tfull = bars + 2 * MAX_STAGES;   // [2]
  uint64_t* tempty = bars + 2 * MAX_STAGES + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * MAX_STAGES + 4);

  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;
  const int tn = p.mode == 1;
  constexpr int TILE_M = PAIR ? 2 * BM : BM;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;          // 0 = leader (issues the MMAs)
  const int cta0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;   // first work item of this CTA (pair)
  const int cta_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int m_tiles = (p.M + TILE_M - 1) / TILE_M;
  const int n_tiles = p.N / p.BN;
  const int kb_total = (p.K + BK - 1) / BK;
  const int work_items = m_tiles * n_tiles * (tn ? p.splits : 1);

  if (threadIdx.x == 0) {
    for (int s = 0; s < MAX_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], PAIR ? 512 : 256);  // pair: both CTAs' epilogue threads arrive at the leader
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    if (PAIR) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                   "r"((uint32_t)TMEM_COLS));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                   "r"((uint32_t)TMEM_COLS));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();   // the peer's barriers must be initialised before anything signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // number of 64-wide MN blocks of the B operand in TN mode
  const int bn_cta = PAIR ? p.BN / 2 : p.BN;   // B columns staged by this CTA
  const int b_blocks = (bn_cta + 63) / 64;
  // One stage = the A block + this CTA's part of the B tile (both multiples of 1 KB = the swizzle atom).  The ring
  // is as deep as the 192 KB operand region allows: the operand loads are latency-bound (~2500 cycles from "stage
  // free" to "stage full" under load, ncu r1i), so bytes in flight, not stage count, set the pace.
  const int stage_bytes = tn ? (2 + b_blocks) * 64 * 128 : A_BYTES + bn_cta * 128;
  const int nstage = min(MAX_STAGES, p.ring_bytes / stage_bytes);
  const uint32_t stage_tx = (uint32_t)stage_bytes * (PAIR ? 2u : 1u);

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = cta0; w < work_items; w += cta_step) {
        int z = 0, t = w;
        if (tn) { z = w / (m_tiles * n_tiles); t = w % (m_tiles * n_tiles); }
        const int m_blk = t / n_tiles, n_blk = t % n_tiles;
        const int kb0 = tn ? z * p.kb_per_split : 0;
        const int kb1 = tn ? min(kb_total, kb0 + p.kb_per_split) : kb_total;
        const int a_row = m_blk * TILE_M + (int)rank * BM;        // this CTA's rows of A (NT) / columns (TN)
        const int b_col = n_blk * p.BN + (int)rank * bn_cta;      // this CTA's part of the B tile
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * stage_bytes;
          uint8_t* sb = sa + A_BYTES;
          if (!PAIR) {
            mbar_expect_tx(&full[stage], stage_tx);
            if (!tn) {
              tma_load_2d(sa, &mapA, &full[stage], kb * BK, a_row);
              tma_load_2d(sb, &mapB, &full[stage], kb * BK, b_col);
            } else {
              for (int b = 0; b < 2; ++b) tma_load_2d(sa + b * 8192, &mapA, &full[stage], a_row + b * 64, kb * BK);
              for (int b = 0; b < b_blocks; ++b) tma_load_2d(sb + b * 8192, &mapB, &full[stage], b_col + b * 64, kb * BK);
            }
          } else {
            // the leader's barrier collects the bytes of both CTAs (its expect_tx covers both; the peer's bytes may
            // land first, the phase cannot complete before the leader's arrive)
            if (rank == 0) mbar_expect_tx(&full[stage], stage_tx);
            const uint32_t fb = mapa_u32(&full[stage], 0);
            if (!tn) {
              tma_load_2d_pair(sa, &mapA, fb, kb * BK, a_row);
              tma_load_2d_pair(sb, &mapB, fb, kb * BK, b_col);
            } else {
              for (int b = 0; b < 2; ++b) tma_load_2d_pair(sa + b * 8192, &mapA, fb, a_row + b * 64, kb * BK);
              for (int b = 0; b < b_blocks; ++b) tma_load_2d_pair(sb + b * 8192, &mapB, fb, b_col + b * 64, kb * BK);
            }
          }
          if (++stage == nstage) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1 && rank == 0) {
    // MMA issuer.  The loop is kept as short as possible: one thread issues four MMAs and the commits per K block,
    // and at ~110 instructions per K block (descriptor construction, waterfall loops around every tcgen05
    // instruction) that thread, not the tensor pipe, set the pace (ncu r1i: 750 cycles per K block for every BN,
    // producer spinning on "empty").  Descriptors differ between stages / K steps only in the 14-bit start-address
    // field (bytes >> 4), so they are advanced by adds; control flow is warp-uniform, the issuing lane is elected.
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    // pair: M = 256 (m_dim field = M >> 4 at bit 24)
    const uint32_t idesc = PAIR ? ((make_idesc(p.BN, tn) & ~(31u << 24)) | ((uint32_t)(256 >> 4) << 24)) : make_idesc(p.BN, tn);
    const uint32_t smem_base = smem_u32(smem);
    const uint64_t adesc0 = tn ? make_desc(smem_base, 8192, 1024) : make_desc(smem_base, 16, 1024);
    const uint64_t bdesc0 = tn ? make_desc(smem_base + A_BYTES, 8192, 1024) : make_desc(smem_base + A_BYTES, 16, 1024);
    const uint64_t kstep = tn ? (2048u >> 4) : (32u >> 4);
    const uint64_t sstep = (uint64_t)(stage_bytes >> 4);
    const bool issuer = elect_one();
    for (int w = cta0; w < work_items; w += cta_step) {
      int z = 0;
      if (tn) z = w / (m_tiles * n_tiles);
      const int kb0 = tn ? z * p.kb_per_split : 0;
      const int kb1 = tn ? min(kb_total, kb0 + p.kb_per_split) : kb_total;
      mbar_wait(&tempty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + (uint32_t)(acc * MAX_BN);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        if (issuer) {
          const uint64_t ad = adesc0 + sstep * (uint64_t)stage, bd = bdesc0 + sstep * (uint64_t)stage;
          if (PAIR) {
            tc_mma_pair(tmem_d, ad, bd, idesc, kb > kb0 ? 1u : 0u);
#pragma unroll
            for (int k = 1; k < BK / UMMA_K; ++k) tc_mma_pair(tmem_d, ad + kstep * k, bd + kstep * k, idesc, 1u);
            tc_commit_pair(&empty[stage]);
            if (kb == kb1 - 1) tc_commit_pair(&tfull[acc]);
          } else {
            tc_mma(tmem_d, ad, bd, idesc, kb > kb0 ? 1u : 0u);
#pragma unroll
            for (int k = 1; k < BK / UMMA_K; ++k) tc_mma(tmem_d, ad + kstep * k, bd + kstep * k, idesc, 1u);
            tc_commit(&empty[stage]);
            if (kb == kb1 - 1) tc_commit(&tfull[acc]);
          }
        }
        __syncwarp();
        if (++stage == nstage) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4) {
    // Epilogue: TMEM -> registers -> (bias / complex combine / convert) -> per-warp 32x128-byte staging tile in
    // shared memory (16-byte chunks XOR-swizzled by row, conflict-free both ways) -> coalesced 128-byte row
    // stores (8 lanes per row, 4 rows per instruction).  A direct lane=row store would touch 32 cache lines per
    // instruction and made the LSU, not the tensor pipe, the limiter of the short-K shapes.
    const int q = warp & 3;
    const int eh = (warp - 4) >> 2;
    uint8_t* stg = smem + p.ring_bytes + 1024 + (warp - 4) * STG_BYTES;
    const uint32_t stg_w = smem_u32(stg) + (uint32_t)lane * 128u;   // this lane's row (write phase)
    const int swz = lane & 7;
    // Full-width blocks (128 bytes per row) leave through ONE TMA store of the staging tile (the XOR swizzle above
    // is the tensor map's 128-byte swizzle; rows beyond M are clipped by the hardware); narrower tail blocks take
    // the manual row stores.
    int acc = 0;
    uint32_t acc_phase = 0;
    bool pending = false;  // a bulk store may still be reading this warp's staging tile
    const uint32_t tempty_leader0 = PAIR ? mapa_u32(&tempty[0], 0) : 0u;
    for (int w = cta0; w < work_items; w += cta_step) {
      int z = 0, t = w;
      if (tn) { z = w / (m_tiles * n_tiles); t = w % (m_tiles * n_tiles); }
      const int m_blk = t / n_tiles, n_blk = t % n_tiles;
      const int row0 = m_blk * TILE_M + (int)rank * BM + q * 32;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr0 = tmem_base + (uint32_t)(acc * MAX_BN) + ((uint32_t)(q * 32) << 16);
      int blk = 0;
      for (int c = 0; c < p.BN; ++blk) {
        const int gcol = n_blk * p.BN + c;  // global accumulator column of the block
        // block kind: 0 = bf16, 64 columns; 1 = fp32, 32 columns; 2 = complex combine, 128 accumulator -> 64 bf16
        int kind = 0, limit = p.BN;
        if (tn) kind = 1;
        else if (p.epi == 2) kind = 2;
        else if (p.epi == 1 && gcol < p.nsplit) { kind = 1; limit = min(p.BN, p.nsplit - n_blk * p.BN); }
        const int width = kind == 0 ? 64 : (kind == 1 ? 32 : 128);
        const int wcols = min(width, limit - c);  // multiple of 16
        if ((blk & 1) == eh) {
          const int ngrp = wcols >> 4;
          const bool use_tma = wcols == width && ((kind == 1) ? p.tma_c0 : p.tma_c);
          if (pending) {
            if (lane == 0) tma_store_wait_read();
            __syncwarp();
            pending = false;
          }
          for (int g0 = 0; g0 < ngrp; g0 += 2) {
            uint32_t vv[32];
            const bool two = g0 + 1 < ngrp;
            tc_ld16(taddr0 + (uint32_t)(c + 16 * g0), vv);
            if (two) tc_ld16(taddr0 + (uint32_t)(c + 16 * g0 + 16), vv + 16);
            tc_ld_wait();
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              if (hh == 1 && !two) break;
              const uint32_t* v = vv + 16 * hh;
              const int gi = g0 + hh;
              const int col = gcol + 16 * gi;
              if (kind == 2) {
                // complex combine: partner row = lane ^ 1, column pairs (2j, 2j+1) -> output column j
                float o[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const float mine_e = __uint_as_float(v[2 * j]);
                  const float mine_o = __uint_as_float(v[2 * j + 1]);
                  const float other_o = __shfl_xor_sync(0xffffffffu, mine_o, 1);
                  o[j] = (lane & 1) ? (mine_e + other_o) : (mine_e - other_o);
                }
                st_shared_v4(stg_w + (uint32_t)((gi ^ swz) << 4), pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]),
                             pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
              } else {
                float f[16];
                if (!tn && p.bias) {
#pragma unroll
                  for (int j = 0; j < 16; j += 4) {
                    const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + col + j));
                    f[j] = __uint_as_float(v[j]) + b4.x;
                    f[j + 1] = __uint_as_float(v[j + 1]) + b4.y;
                    f[j + 2] = __uint_as_float(v[j + 2]) + b4.z;
                    f[j + 3] = __uint_as_float(v[j + 3]) + b4.w;
                  }
                } else {
#pragma unroll
                  for (int j = 0; j < 16; ++j) f[j] = __uint_as_float(v[j]);
                }
                if (p.epi == 3) {   // SiLU before the bf16 store (node-level S2 to_grid GEMM, ops._node_s2_gemm_forward)
#pragma unroll
                  for (int j = 0; j < 16; ++j) f[j] = __fdividef(f[j], 1.0f + __expf(-f[j]));
                }
                if (kind == 1) {
#pragma unroll
                  for (int j = 0; j < 4; ++j)
                    st_shared_v4(stg_w + (uint32_t)(((4 * gi + j) ^ swz) << 4), __float_as_uint(f[4 * j]),
                                 __float_as_uint(f[4 * j + 1]), __float_as_uint(f[4 * j + 2]),
                                 __float_as_uint(f[4 * j + 3]));
                } else {
                  st_shared_v4(stg_w + (uint32_t)(((2 * gi) ^ swz) << 4), pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]),
                               pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
                  st_shared_v4(stg_w + (uint32_t)(((2 * gi + 1) ^ swz) << 4), pack_bf16(f[8], f[9]),
                               pack_bf16(f[10], f[11]), pack_bf16(f[12], f[13]), pack_bf16(f[14], f[15]));
                }
              }
            }
          }
          if (use_tma) {
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              if (tn) tma_store_3d(&mapC0, stg, gcol, row0, z);
              else if (kind == 1) tma_store_2d(&mapC0, stg, gcol, row0);
              else if (kind == 2) tma_store_2d(&mapC, stg, gcol / 2, row0);
              else tma_store_2d(&mapC, stg, gcol - (p.epi == 1 ? p.nsplit : 0), row0);
              tma_store_commit();
            }
            pending = true;
            c += wcols;
            continue;
          }
          __syncwarp();
          // coalesced write-out: lane -> (row 4*it + lane/8, 16-byte chunk lane%8)
          const int nchunks = kind == 0 ? (wcols >> 3) : (kind == 1 ? (wcols >> 2) : (wcols >> 4));
          const int ch = lane & 7;
          uint8_t* gbase;
          long long pitch;  // bytes
          if (tn) {
            gbase = reinterpret_cast<uint8_t*>(p.P + ((size_t)z * p.M) * p.N + gcol);
            pitch = (long long)p.N * 4;
          } else if (kind == 1) {
            gbase = reinterpret_cast<uint8_t*>(p.C0 + gcol);
            pitch = p.ldc0 * 4;
          } else if (kind == 2) {
            gbase = reinterpret_cast<uint8_t*>(p.C + gcol / 2);
            pitch = p.ldc * 2;
          } else {
            gbase = reinterpret_cast<uint8_t*>(p.C + (gcol - (p.epi == 1 ? p.nsplit : 0)));
            pitch = p.ldc * 2;
          }
          if (ch < nchunks) {
#pragma unroll
            for (int it = 0; it < 8; ++it) {
              const int r = 4 * it + (lane >> 3);
              const int grow = row0 + r;
              if (grow < p.M) {
                const uint4 val = ld_shared_v4(smem_u32(stg) + (uint32_t)(r * 128 + ((ch ^ (r & 7)) << 4)));
                *reinterpret_cast<uint4*>(gbase + (size_t)grow * pitch + ch * 16) = val;
              }
            }
          }
          __syncwarp();
        }
        c += wcols;
      }
      tc_fence_before();
      if (PAIR) mbar_arrive_cluster(tempty_leader0 + 8u * (uint32_t)acc);
      else mbar_arrive(&tempty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (pending && lane == 0) tma_store_wait_read();
  }

  tc_fence_before();
  if (PAIR) cluster_sync_all();   // neither CTA may leave (or free TMEM) while its partner's MMAs/arrives are in flight
  else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if (PAIR)
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TMEM_COLS));
    else
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TMEM_COLS));
  }
}

// C[c, r] (+)= sum_z P[z][r, c]: fixed-order reduction of the split-K planes of a swapped wgrad product (planes are
// [R, Cc] = C^T), transposed through a 32x33 shared tile so both sides stay coalesced.
__global__ void splitk_reduce_transposed_kernel(const float* __restrict__ P, int splits, int R, int Cc,
                                                float* __restrict__ C, long long ldc, int accumulate) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const size_t plane = (size_t)R * Cc;
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int r = r0 + j, c = c0 + threadIdx.x;
    float acc = 0.f;
    if (r < R && c < Cc)
      for (int z = 0; z < splits; ++z) acc += P[z * plane + (size_t)r * Cc + c];
    tile[j][threadIdx.x] = acc;
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int c = c0 + j, r = r0 + threadIdx.x;
    if (r < R && c < Cc) {
      float* o = C + (size_t)c * ldc + r;
      const float v = tile[threadIdx.x][j];
      *o = accumulate ? (*o + v) : v;
    }
  }
}

}  // namespace tc
}  // namespace efl

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
using namespace efl;
using efl::tc::Params;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// 2-D bf16 row-major [rows, cols] (leading dimension ld elements); box = {64 columns (128 B), box_rows}
int efl_make_map_bf16(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld, int box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    efl_set_error("cuTensorMapEncodeTiled not available from the driver");
    return EFL_ERR_UNSUPPORTED;
  }
  if (((uintptr_t)base & 15) || ((ld * 2) & 15)) {
    efl_set_error("tc_gemm: operand base and row pitch must be 16-byte aligned");
    return EFL_ERR_ARG;
  }
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    efl_set_error("cuTensorMapEncodeTiled failed");
    return EFL_ERR_ARG;
  }
  return EFL_OK;
}

// Output tile map for the epilogue's TMA stores: row-major [planes,] rows x cols, box = 32 rows x 128 bytes, 128-byte
// swizzle.  Returns 0 and sets *ok = 1 when the layout qualifies (16-byte aligned base and pitches); *ok = 0
// otherwise (the epilogue then uses its manual row stores).
static int make_out_map(CUtensorMap* map, int* ok, int is_f32, const void* base, long long planes, long long rows,
                        long long cols, long long ld) {
  *ok = 0;
  memset(map, 0, sizeof(*map));
  EncodeTiledFn enc = get_encode();
  const int esz = is_f32 ? 4 : 2;
  const int bcols = 128 / esz;
  if (!enc || base == nullptr || cols < bcols || rows <= 0 || ((uintptr_t)base & 15) || ((ld * esz) & 15)) return EFL_OK;
  cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)(planes > 0 ? planes : 1)};
  cuuint64_t strides[2] = {(cuuint64_t)ld * esz, (cuuint64_t)ld * esz * (cuuint64_t)rows};
  cuuint32_t box[3] = {(cuuint32_t)bcols, 32u, 1u};
  cuuint32_t estr[3] = {1u, 1u, 1u};
  const int rank = planes > 0 ? 3 : 2;
  CUresult r = enc(map, is_f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank,
                   const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_SUCCESS) *ok = 1;
  return EFL_OK;
}

// largest multiple of 16 that divides n and is <= 256
static int pick_bn(int n) {
  for (int bn = 256; bn >= 16; bn -= 16)
    if (n % bn == 0) return bn;
  return 0;
}

int efl_num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// CTA-pair mode: 0 = single-CTA tiles only, 1 = pairs where they measured faster (default; nt_use_pair /
// tn_use_pair), 2 = pairs for every GEMM (tests).  EFL_GEMM_PAIR presets it; efl_tc_gemm_pair_mode() switches it at
// run time (tests, tools/kernel_bench.py).
static int g_pair_mode = -1;
static int pair_mode() {
  if (g_pair_mode < 0) {
    const char* e = getenv("EFL_GEMM_PAIR");
    g_pair_mode = e ? atoi(e) : 1;
  }
  return g_pair_mode;
}
// NT: pairs pay off when the main loop, not the epilogue, sets the pace (K >= 512) and the half B tile is still a
// decent MMA operand; there must be enough 256-row tiles to fill the pairs (tools/kernel_bench.py, r1 v4).
static bool nt_use_pair(int M, int K, int bn) {
  const int m = pair_mode();
  if (m >= 2) return true;
  return m == 1 && M >= 256 * (efl_num_sms() / 2) && K >= 512 && bn >= 128;
}
static bool tn_use_pair(int n1, int n2, int bn) {
  const int m = pair_mode();
  if (m >= 2) return true;
  return m == 1 && n2 == bn && (n1 + 255) / 256 * 256 == (n1 + 127) / 128 * 128;
}
extern "C" int efl_tc_gemm_pair_mode(int mode) {
  const int prev = pair_mode();
  if (mode >= 0 && mode <= 2) g_pair_mode = mode;
  return prev;
}

// Operand ring budget: EFL_GEMM_RING_KB (default 192 = 4 x 48 KB) or efl_tc_gemm_ring_kb(); multiples of 48 KB keep
// every stage shape whole.  144 leaves ~45 KB of shared memory per SM for co-resident kernels of another stream.
static int g_ring_kb = -1;
static int ring_bytes() {
  if (g_ring_kb < 0) {
    const char* e = getenv("EFL_GEMM_RING_KB");
    g_ring_kb = e ? atoi(e) : 192;
    if (g_ring_kb < 96 || g_ring_kb > 192) g_ring_kb = 192;
  }
  return g_ring_kb * 1024;
}
extern "C" int efl_tc_gemm_ring_kb(int kb) {
  const int prev = ring_bytes() / 1024;
  if (kb >= 96 && kb <= 192) g_ring_kb = kb;
  return prev;
}

static int launch(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, const CUtensorMap& mc0,
                  const Params& p_in, int work_items, bool pair, cudaStream_t s) {
  Params p = p_in;
  p.ring_bytes = ring_bytes();
  const int smem_bytes = p.ring_bytes + 1024 + 1024 + 8 * efl::tc::STG_BYTES;
  static unsigned long long attr_set = 0;
  if (efl_first_use_on_device(&attr_set)) {
    cudaError_t e = cudaFuncSetAttribute(efl::tc::tc_gemm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         efl::tc::SMEM_BYTES);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(efl::tc::tc_gemm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               efl::tc::SMEM_BYTES);
    if (e != cudaSuccess) {
      efl_set_error(cudaGetErrorString(e));
      return (int)e;
    }
  }
  if (!pair) {
    const int grid = work_items < efl_num_sms() ? work_items : efl_num_sms();
    efl::tc::tc_gemm_kernel<false><<<grid, efl::tc::NUM_THREADS, smem_bytes, s>>>(ma, mb, mc, mc0, p);
    EFL_CHECK_LAUNCH();
    return EFL_OK;
  }
  // one CTA pair (cluster of 2 = the two SMs of a TPC) per work item, persistent over at most num_sms / 2 pairs
  int pairs = efl_num_sms() / 2;
  if (work_items < pairs) pairs = work_items;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(2 * pairs));
  cfg.blockDim = dim3(efl::tc::NUM_THREADS);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, efl::tc::tc_gemm_kernel<true>, ma, mb, mc, mc0, p);
  if (e != cudaSuccess) {
    efl_set_error(cudaGetErrorString(e));
    return (int)e;
  }
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// C[M, N(/2 for epilogue 2)] = A[M, K] . B[N, K]^T; bf16 operands, fp32 accumulate.
extern "C" int efl_tc_gemm_nt(int epilogue, const void* A, long long lda, const void* B, long long ldb,
                              const float* bias, void* C, long long ldc, float* C0, long long ldc0, int nsplit, int M,
                              int N, int K, void* stream) {
  if (M <= 0 || N <= 0) return EFL_OK;
  const int bn = pick_bn(N);
  EFL_REQUIRE(bn > 0, "efl_tc_gemm_nt: N must be a multiple of 16");
  EFL_REQUIRE(epilogue >= 0 && epilogue <= 3, "efl_tc_gemm_nt: unknown epilogue");
  EFL_REQUIRE(epilogue != 2 || (M % 2 == 0), "efl_tc_gemm_nt: complex epilogue needs even M");
  EFL_REQUIRE(epilogue != 1 || (nsplit % 16 == 0), "efl_tc_gemm_nt: nsplit must be a multiple of 16");
  EFL_REQUIRE((ldc % 8) == 0 && ((uintptr_t)C & 15) == 0, "efl_tc_gemm_nt: C must be 16-byte aligned with ldc % 8 == 0");
  CUtensorMap ma, mb;
  int rc = efl_make_map_bf16(&ma, A, M, K, lda, efl::tc::BM);
  if (rc) return rc;
  // CTA pairs pay off once there are enough 256-row tiles to fill the 74 pairs
  const bool pair = nt_use_pair(M, K, bn);
  rc = efl_make_map_bf16(&mb, B, N, K, ldb, pair ? bn / 2 : bn);
  if (rc) return rc;
  Params p{};
  p.mode = 0; p.M = M; p.N = N; p.K = K; p.BN = bn; p.epi = epilogue;
  p.C = (__nv_bfloat16*)C; p.ldc = ldc; p.C0 = C0; p.ldc0 = ldc0; p.nsplit = nsplit; p.bias = bias;
  p.P = nullptr; p.splits = 1; p.kb_per_split = 0;
  CUtensorMap mc, mc0;
  const long long ccols = epilogue == 2 ? N / 2 : (epilogue == 1 ? N - nsplit : N);
  make_out_map(&mc, &p.tma_c, 0, ccols > 0 ? C : nullptr, 0, M, ccols, ldc);
  make_out_map(&mc0, &p.tma_c0, 1, epilogue == 1 ? C0 : nullptr, 0, M, nsplit, ldc0);
  const int tile_m = pair ? 2 * efl::tc::BM : efl::tc::BM;
  const int tiles = ((M + tile_m - 1) / tile_m) * (N / bn);
  return launch(ma, mb, mc, mc0, p, tiles, pair, (cudaStream_t)stream);
}

// Plan of one wgrad GEMM  C[N1, N2] = A[Mred, N1]^T B[Mred, N2].
//  * orientation: the kernel's tile is 128 (M side) x BN (N side, a divisor of the N extent); either operand can
//    take the M side (the swapped product is C^T, transposed back by the split-K reduction).  The model below - tile
//    count x max(MMA cycles, operand bytes / 62 B/clk of L2->SM feed per SM) per K block - picked the faster
//    orientation on all six wgrad shapes of the path (tools/kernel_bench.py --only tnx); a swap must win by 10 %.
//  * CTA pairs only where the 256-row tiles add no padding and the N side is a single tile (the conv1 shapes).
//  * split over the reduced (edge) dimension: the output is only N1 x N2, so without splits just (N1/128) x (N2/BN)
//    CTAs would work; the smallest split (>= 8 K blocks each) whose last wave is >= 94 % full, else the best-filled
//    one; at most 4 waves' worth of partial planes.
struct TnPlan {
  bool swap, pair;
  int bn, splits;
};
static double tn_cost(int n1, int n2) {
  const int bn = pick_bn(n2);
  if (bn <= 0) return 1e30;
  const double tiles = (double)((n1 + 127) / 128) * (n2 / bn);
  const double mma = 2.0 * bn;                                   // 4 MMAs of bn/2 cycles per K block
  const double feed = (16384.0 + ((bn + 63) / 64) * 8192.0) / 62.0;
  return tiles * (mma > feed ? mma : feed);
}
static TnPlan tn_plan(int Mred, int N1, int N2) {
  TnPlan pl{};
  pl.swap = pair_mode() != 0 && (N1 % 16 == 0) && tn_cost(N2, N1) < 0.9 * tn_cost(N1, N2);
  const int n1 = pl.swap ? N2 : N1, n2 = pl.swap ? N1 : N2;
  pl.bn = pick_bn(n2);
  if (pl.bn <= 0) { pl.splits = 1; return pl; }
  pl.pair = tn_use_pair(n1, n2, pl.bn);
  const int tile_m = pl.pair ? 2 * efl::tc::BM : efl::tc::BM;
  const int tiles = ((n1 + tile_m - 1) / tile_m) * (n2 / pl.bn);
  const int kb = (Mred + efl::tc::BK - 1) / efl::tc::BK;
  const int sms = pl.pair ? efl_num_sms() / 2 : efl_num_sms();   // scheduling units: CTAs or CTA pairs
  int smax = kb / 8;
  if (smax < 1) smax = 1;
  const int cap = (4 * sms + tiles - 1) / tiles;
  if (smax > cap) smax = cap;
  int best = 1;
  double best_eff = 0.0;
  for (int s = 1; s <= smax; ++s) {
    const int items = tiles * s;
    const int waves = (items + sms - 1) / sms;
    const double eff = (double)items / ((double)waves * sms);
    if (eff > best_eff + 1e-9) { best_eff = eff; best = s; }
    if (eff >= 0.94) { best = s; break; }
  }
  const int per = (kb + best - 1) / best;
  pl.splits = (kb + per - 1) / per;
  return pl;
}

extern "C" int efl_tc_gemm_tn_splits(int Mred, int N1, int N2) { return tn_plan(Mred, N1, N2).splits; }

// defined in gemm_simt.cu
namespace efl {
__global__ void splitk_reduce_kernel(const float* __restrict__ P, int splits, long long n, float* __restrict__ C,
                                     long long ldc, int N2, int accumulate);
}

// C[N1, N2] (+)= A[Mred, N1]^T . B[Mred, N2]; workspace: efl_tc_gemm_tn_splits(...) * N1 * N2 floats
extern "C" int efl_tc_gemm_tn(const void* A, long long lda, const void* B, long long ldb, float* C, long long ldc,
                              int accumulate, float* workspace, int Mred, int N1, int N2, void* stream) {
  if (N1 <= 0 || N2 <= 0) return EFL_OK;
  EFL_REQUIRE(pick_bn(N2) > 0, "efl_tc_gemm_tn: N2 must be a multiple of 16");
  EFL_REQUIRE(Mred > 0, "efl_tc_gemm_tn: empty reduction");
  const TnPlan pl = tn_plan(Mred, N1, N2);
  // the kernel computes P[z][n1, n2]; swapped: (n1, n2) = (N2, N1) with the operands exchanged
  const int n1 = pl.swap ? N2 : N1, n2 = pl.swap ? N1 : N2;
  const void* Am = pl.swap ? B : A;
  const void* Bm = pl.swap ? A : B;
  const long long ldam = pl.swap ? ldb : lda, ldbm = pl.swap ? lda : ldb;
  CUtensorMap ma, mb;
  int rc = efl_make_map_bf16(&ma, Am, Mred, n1, ldam, efl::tc::BK);
  if (rc) return rc;
  rc = efl_make_map_bf16(&mb, Bm, Mred, n2, ldbm, efl::tc::BK);
  if (rc) return rc;
  const int splits = pl.splits;
  const int kb = (Mred + efl::tc::BK - 1) / efl::tc::BK;
  Params p{};
  p.mode = 1; p.M = n1; p.N = n2; p.K = Mred; p.BN = pl.bn; p.epi = 0;
  p.P = workspace; p.splits = splits; p.kb_per_split = (kb + splits - 1) / splits;
  const int tile_m = pl.pair ? 2 * efl::tc::BM : efl::tc::BM;
  const int tiles = ((n1 + tile_m - 1) / tile_m) * (n2 / pl.bn);
  cudaStream_t s = (cudaStream_t)stream;
  CUtensorMap mc, mc0;
  make_out_map(&mc, &p.tma_c, 0, nullptr, 0, 0, 0, 0);
  make_out_map(&mc0, &p.tma_c0, 1, workspace, splits, n1, n2, n2);
  rc = launch(ma, mb, mc, mc0, p, tiles * splits, pl.pair, s);
  if (rc) return rc;
  const long long n = (long long)N1 * N2;
  if (!pl.swap) {
    efl::splitk_reduce_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(workspace, splits, n, C, ldc, N2, accumulate);
  } else {
    dim3 grid((unsigned)((N1 + 31) / 32), (unsigned)((N2 + 31) / 32)), block(32, 8);
    efl::tc::splitk_reduce_transposed_kernel<<<grid, block, 0, s>>>(workspace, splits, N2, N1, C, ldc, accumulate);
  }
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
