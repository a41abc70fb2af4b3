// Rotate-back + aggregate on the tensor cores (bf16 mode).
//
// Replaces alpha*value (reference EF_layers.py:1186-1190), SO3_Rotation.rotate_inv with get_rotate_inv_rescale
// (:500-505, :1530-1549) and SO3_Embedding._reduce_edge (:342-351).  Per destination node n
//     out[n][c, i] = sum_{edges e -> n} sum_j (alpha[e, head(c)] V_e[j, c]) * (rescale_l D_e[j, i])
// is a chain of per-edge MMAs that ACCUMULATE IN TMEM: channels are the UMMA M dimension (112 -> 128 lanes), the 29
// kept m-primary rows the K dimension (-> 32), the 49 l-primary output rows N (-> 64).  The segment sum over a
// node's incoming edges is the accumulate flag of tcgen05.mma: the [E, 49, 112] message tensor never exists and
// there are no atomics; edge order per node is fixed, so the result is deterministic.
//   A operand  V_e as stored by the SO(2) GEMMs (rows of 112 contiguous channels): TMA, MN-major, 128-byte swizzle;
//              the workers scale it in place by alpha (per head) before the MMA
//   B operand  the edge's Wigner block: the workers expand the 296-float Wigner store (bulk-copied into the stage)
//              into a K-major bf16 [64 x 32] tile; its block-diagonal zero pattern is fixed and zeroed once
// Persistent CTAs, node = blockIdx.x + i * gridDim.x.  warp 0: TMA producer (V boxes + Wigner bulk copy, 14-stage
// ring), warp 1: MMA issuer, warp 2: TMEM allocator, warps 4-7: workers (operand preparation, accumulator drain).
#include <stdlib.h>

#include "tc_common.cuh"

namespace efl {
namespace rbtc {

using namespace efl::tc;

constexpr int MAX_NSTAGE = 14;
constexpr int NTHREADS = 256;
constexpr int ALPHA_BATCH = 128;  // edges of alpha staged in shared memory at a time (heads <= 8)

struct Params {
  int n0, n_nodes, e0, heads, vc, CH;
  const int* rowptr;
  const int* perm;
  const float* wig;
  const float* alpha;  // [E, heads] sorted order, or NULL
  float* out;          // [Nt, L, CH]
  int dbg;             // timing experiments only (EFL_RBTC_DBG): 1 no alpha scaling, 2 no B build, 4 no proxy fence, 8 no V loads
};

template <int LMAX, int MMAX>
struct Cfg {
  using T = Tab<LMAX, MMAX>;
  static constexpr int M = T::M, L = T::L;
  static constexpr int KP = (M + 15) / 16 * 16;
  static constexpr int NP = (L + 15) / 16 * 16;
  static constexpr int R0 = LMAX + 1, R1 = MMAX >= 1 ? 2 * LMAX : 0, R2 = MMAX >= 2 ? 2 * (LMAX - 1) : 0;
  static constexpr int F1 = R0, F2 = R0 + R1;
  static_assert(R0 + R1 + R2 == M && MMAX <= 2, "m-primary row bookkeeping");
  static constexpr int BLK = KP * 128;  // one 64-channel half of the A tile
  static constexpr int A_BYTES = 2 * BLK;
  static constexpr int SBO = (KP / 8) * 128;
  static constexpr int B_BYTES = NP * KP * 2;
  static constexpr int W_BYTES = T::WSZ * 4;
  static constexpr int STAGE = (A_BYTES + B_BYTES + W_BYTES + 1023) / 1024 * 1024;
  // ring depth: the per-edge hand-off chain (TMA -> workers -> MMA -> commit -> producer) is ~3.7 us long, so the
  // throughput of the sequential per-node edge loop is depth / latency: as deep as shared memory allows
  static constexpr int NSTAGE = (200 * 1024 / STAGE) < MAX_NSTAGE ? (200 * 1024 / STAGE) : MAX_NSTAGE;
  __host__ __device__ static constexpr int count_entries() {
    int n = 0;
    for (int l = 0; l <= LMAX; ++l) n += T::nrows(l) * (2 * l + 1);
    return n;
  }
  static constexpr int NENT = count_entries();
  static constexpr int TMEM_COLS = 2 * NP <= 128 ? 128 : 256;
  static constexpr int OFF_TAB = NSTAGE * STAGE;             // int tab_w[NENT], int tab_b[NENT], float tab_s[NENT]
  static constexpr int OFF_ALPHA = OFF_TAB + ((3 * NENT * 4 + 127) / 128) * 128;
  static constexpr int OFF_BAR = OFF_ALPHA + ALPHA_BATCH * 8 * 4;
  static constexpr int SMEM = OFF_BAR + 512 + 1024;   // (3 * NSTAGE + 5) mbarriers / slots of 8 bytes <= 512
  static_assert((3 * NSTAGE + 5) * 8 <= 512 && SMEM <= 232448, "shared memory budget");
};

__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

template <int LMAX, int MMAX>
__global__ void __launch_bounds__(NTHREADS, 1)
rotate_back_tc_fwd_kernel(const __grid_constant__ CUtensorMap mapV0, const __grid_constant__ CUtensorMap mapV1,
                          const __grid_constant__ CUtensorMap mapV2, const Params p) {
  using C = Cfg<LMAX, MMAX>;
  using T = typename C::T;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  int* tab_w = reinterpret_cast<int*>(smem + C::OFF_TAB);
  int* tab_b = tab_w + C::NENT;
  float* tab_s = reinterpret_cast<float*>(tab_b + C::NENT);
  float* s_alpha = reinterpret_cast<float*>(smem + C::OFF_ALPHA);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::OFF_BAR);
  uint64_t* full = bars;                     // [C::NSTAGE] TMA landed
  uint64_t* ready = bars + C::NSTAGE;           // [C::NSTAGE] operands prepared by the workers
  uint64_t* empty = bars + 2 * C::NSTAGE;       // [C::NSTAGE] MMAs of the stage retired
  uint64_t* acc_full = bars + 3 * C::NSTAGE;    // [2]
  uint64_t* acc_free = bars + 3 * C::NSTAGE + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * C::NSTAGE + 4);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- setup: zero the operand stages (pad rows / off-block entries stay zero), expansion table, barriers ----
  {
    uint32_t* z = reinterpret_cast<uint32_t*>(smem);
    for (int i = threadIdx.x; i < C::OFF_TAB / 4; i += NTHREADS) z[i] = 0u;
    int idx = 0;
    for (int l = 0; l <= LMAX; ++l) {
      for (int mp = -T::mm(l); mp <= T::mm(l); ++mp) {
        const int j = T::mprim(l, mp);
        for (int q = 0; q < 2 * l + 1; ++q, ++idx) {
          if ((idx % NTHREADS) == (int)threadIdx.x) {
            const int i = l * l + q;
            tab_w[idx] = T::wrow(l, mp) + q;
            tab_b[idx] = (i >> 3) * C::SBO + (j >> 3) * 128 + (i & 7) * 16 + (j & 7) * 2;
            tab_s[idx] = T::rescale(l);
          }
        }
      }
    }
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < C::NSTAGE; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&ready[s], 128);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&acc_full[a], 1);
      mbar_init(&acc_free[a], 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)C::TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---- producer: V boxes + the edge's Wigner store ----
    int stage = 0;
    uint32_t phase = 0;
    for (int ni = blockIdx.x; ni < p.n_nodes; ni += gridDim.x) {
      const int n = p.n0 + ni;
      const int kb = p.rowptr[n], ke = p.rowptr[n + 1];
      for (int k0 = kb; k0 < ke; k0 += 32) {
        const int pv = (k0 + lane < ke) ? __ldg(p.perm + k0 + lane) : 0;  // original edge ids of the next 32 edges
        const int cnt = min(32, ke - k0);
        for (int t = 0; t < cnt; ++t) {
          const int eo = __shfl_sync(0xffffffffu, pv, t);
          const int el = k0 + t - p.e0;
          uint8_t* st = smem + stage * C::STAGE;
          if (lane == 0) {
            mbar_wait(&empty[stage], phase ^ 1);
            mbar_expect_tx(&full[stage], (uint32_t)(((p.dbg & 8) ? 0 : 2 * C::M * 128) + C::W_BYTES));
          }
          if (lane == 0) {  // one thread issues the seven copies (spreading them over lanes measured slower)
            if (!(p.dbg & 8)) {
              for (int hb = 0; hb < 2; ++hb) {
                uint8_t* blk = st + hb * C::BLK;
                tma_load_2d(blk, &mapV0, &full[stage], hb * 64, el * C::R0);
                if (C::R1 > 0) tma_load_2d(blk + C::F1 * 128, &mapV1, &full[stage], hb * 64, el * C::R1);
                if (C::R2 > 0) tma_load_2d(blk + C::F2 * 128, &mapV2, &full[stage], hb * 64, el * C::R2);
              }
            }
            bulk_g2s(st + C::A_BYTES + C::B_BYTES, p.wig + (size_t)eo * T::WSZ, (uint32_t)C::W_BYTES, &full[stage]);
          }
          if (++stage == C::NSTAGE) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ---- MMA issuer ----
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    const uint32_t idesc = make_idesc2(C::NP, 1, 0);  // A = V tile (MN-major), B = Wigner tile (K-major)
    for (int ni = blockIdx.x; ni < p.n_nodes; ni += gridDim.x) {
      const int n = p.n0 + ni;
      const int kb = p.rowptr[n], ke = p.rowptr[n + 1];
      if (ke <= kb) continue;
      mbar_wait(&acc_free[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + (uint32_t)(acc * C::NP);
      for (int k = kb; k < ke; ++k) {
        mbar_wait(&ready[stage], phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t sa = smem_u32(smem + stage * C::STAGE);
          const uint32_t sb = sa + C::A_BYTES;
#pragma unroll
          for (int ks = 0; ks < C::KP / 16; ++ks)
            tc_mma(tmem_d, make_desc(sa + ks * 2048, (uint32_t)C::BLK, 1024),
                   make_desc_noswz(sb + ks * 256, 128, (uint32_t)C::SBO), idesc, (k > kb || ks > 0) ? 1u : 0u);
          tc_commit(&empty[stage]);
          if (k == ke - 1) tc_commit(&acc_full[acc]);
        }
        __syncwarp();
        if (++stage == C::NSTAGE) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4) {
    // ---- workers: thread = channel (TMEM lane) for the drain; 16-byte chunks / table entries for the operands ----
    const int tid = threadIdx.x - 128;
    const int q4 = warp & 3;
    const int NCHK = p.CH >> 3;             // 16-byte channel chunks per row
    const int nchunks = C::M * NCHK;
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    int hd[4], coff[4];  // head and byte offset (inside a stage) of this thread's fixed A chunks; coff < 0: none
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int ci = tid + 128 * u;
      const int j = ci / NCHK, q = ci % NCHK;
      hd[u] = min((q * 8) / p.vc, p.heads - 1);
      coff[u] = ci < nchunks ? (q >> 3) * C::BLK + j * 128 + (((q & 7) ^ (j & 7)) << 4) : -1;
    }
    constexpr int NB = (C::NENT + 127) / 128;  // Wigner-block entries per thread, kept in registers
    int ent_w[NB], ent_b[NB];
    float ent_s[NB];
#pragma unroll
    for (int u = 0; u < NB; ++u) {
      const int idx = tid + 128 * u;
      ent_w[u] = idx < C::NENT ? tab_w[idx] : 0;
      ent_b[u] = idx < C::NENT ? tab_b[idx] : -1;
      ent_s[u] = idx < C::NENT ? tab_s[idx] : 0.f;
    }
    for (int ni = blockIdx.x; ni < p.n_nodes; ni += gridDim.x) {
      const int n = p.n0 + ni;
      const int kb = p.rowptr[n], ke = p.rowptr[n + 1];
      float* o = p.out + (size_t)n * C::L * p.CH + tid;
      if (ke <= kb) {
        if (tid < p.CH)
          for (int i = 0; i < C::L; ++i) o[(size_t)i * p.CH] = 0.f;
        continue;
      }
      for (int k = kb; k < ke; ++k) {
        // alpha of the next ALPHA_BATCH edges of the node: one coalesced load per batch (a per-edge global load
        // would put an L2 round trip on every edge of this sequential loop)
        const int rel = (k - kb) % ALPHA_BATCH;
        if (rel == 0 && p.alpha) {
          asm volatile("bar.sync 1, 128;" ::: "memory");
          const int cnt = min(ALPHA_BATCH, ke - k) * p.heads;
          for (int i = tid; i < cnt; i += 128) s_alpha[i] = __ldg(p.alpha + (size_t)k * p.heads + i);
          asm volatile("bar.sync 1, 128;" ::: "memory");
        }
        float a[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) a[u] = p.alpha ? s_alpha[rel * p.heads + hd[u]] : 1.f;
        mbar_wait(&full[stage], phase);
        uint8_t* st = smem + stage * C::STAGE;
        // all shared-memory loads of the edge first (the compiler cannot move them above the stores below), then the
        // arithmetic, then the stores:
        // (a) A <- alpha * V in place (this thread's fixed 16-byte chunks, 128-byte-swizzled positions)
        // (b) B <- rescale * Wigner rows, K-major bf16 (only the block entries; the rest stays zero)
        const float* w = reinterpret_cast<const float*>(st + C::A_BYTES + C::B_BYTES);
        uint8_t* sb = st + C::A_BYTES;
        uint4 v[4];
        float wv[NB];
        const bool do_a = p.alpha && !(p.dbg & 1), do_b = !(p.dbg & 2);
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (do_a && coff[u] >= 0) v[u] = *reinterpret_cast<const uint4*>(st + coff[u]);
#pragma unroll
        for (int u = 0; u < NB; ++u)
          if (do_b && ent_b[u] >= 0) wv[u] = w[ent_w[u]] * ent_s[u];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (do_a && coff[u] >= 0) {
            const __nv_bfloat162 a2 = __float2bfloat162_rn(a[u]);
            __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v[u]);
#pragma unroll
            for (int t = 0; t < 4; ++t) h[t] = __hmul2(h[t], a2);
            *reinterpret_cast<uint4*>(st + coff[u]) = v[u];
          }
        }
#pragma unroll
        for (int u = 0; u < NB; ++u)
          if (do_b && ent_b[u] >= 0) *reinterpret_cast<__nv_bfloat16*>(sb + ent_b[u]) = __float2bfloat16_rn(wv[u]);
        if (!(p.dbg & 4)) fence_proxy_async();
        mbar_arrive(&ready[stage]);
        if (++stage == C::NSTAGE) { stage = 0; phase ^= 1; }
      }
      // ---- drain the node's accumulator ----
      mbar_wait(&acc_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t)(acc * C::NP) + ((uint32_t)(q4 * 32) << 16);
      uint32_t v[C::NP];
#pragma unroll
      for (int t = 0; t < C::NP / 16; ++t) tc_ld16(taddr + 16u * t, v + 16 * t);
      tc_ld_wait();
      tc_fence_before();
      mbar_arrive(&acc_free[acc]);
      if (tid < p.CH) {
#pragma unroll
        for (int i = 0; i < C::L; ++i) o[(size_t)i * p.CH] = __uint_as_float(v[i]);
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::TMEM_COLS));
  }
}

}  // namespace rbtc
}  // namespace efl

using namespace efl;

int efl_make_map_bf16(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld, int box_rows);
int efl_num_sms();

template <int LMAX, int MMAX>
static int rb_tc_launch(const void* v0, const void* v1, const void* v2, efl::rbtc::Params& p, int Ec, cudaStream_t s) {
  using C = efl::rbtc::Cfg<LMAX, MMAX>;
  CUtensorMap m0, m1, m2;
  int rc = efl_make_map_bf16(&m0, v0, (long long)Ec * C::R0, p.CH, p.CH, C::R0);
  if (rc) return rc;
  rc = efl_make_map_bf16(&m1, C::R1 > 0 ? v1 : v0, (long long)Ec * (C::R1 > 0 ? C::R1 : C::R0), p.CH, p.CH,
                         C::R1 > 0 ? C::R1 : C::R0);
  if (rc) return rc;
  rc = efl_make_map_bf16(&m2, C::R2 > 0 ? v2 : v0, (long long)Ec * (C::R2 > 0 ? C::R2 : C::R0), p.CH, p.CH,
                         C::R2 > 0 ? C::R2 : C::R0);
  if (rc) return rc;
  auto kern = efl::rbtc::rotate_back_tc_fwd_kernel<LMAX, MMAX>;
  static bool attr = false;
  if (!attr) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM);
    if (e != cudaSuccess) {
      efl_set_error(cudaGetErrorString(e));
      return (int)e;
    }
    attr = true;
  }
  const int grid = p.n_nodes < efl_num_sms() ? p.n_nodes : efl_num_sms();
  kern<<<grid, efl::rbtc::NTHREADS, C::SMEM, s>>>(m0, m1, m2, p);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// bf16 V buffers of chunk edges [e0, e0 + Ec), destinations [n0, n1) (whole rows of the CSR inside the chunk)
extern "C" int efl_rotate_back_aggregate_tc_fwd(int lmax, int mmax, int heads, int value_ch, const void* v0,
                                                const void* v1, const void* v2, const float* alpha,
                                                const int* rowptr, const int* perm, const float* wig, float* out,
                                                int n0, int n1, int e0, int Ec, void* stream) {
  if (n1 <= n0) return EFL_OK;
  const int CH = heads * value_ch;
  EFL_REQUIRE(CH <= 128 && CH % 8 == 0 && value_ch % 8 == 0 && heads <= 8,
              "efl_rotate_back_aggregate_tc_fwd: heads*value <= 128, value % 8 == 0, heads <= 8");
  EFL_REQUIRE(Ec > 0, "efl_rotate_back_aggregate_tc_fwd: empty chunk");
  const char* dbg = getenv("EFL_RBTC_DBG");
  efl::rbtc::Params p{n0, n1 - n0, e0, heads, value_ch, CH, rowptr, perm, wig, alpha, out, dbg ? atoi(dbg) : 0};
  cudaStream_t s = (cudaStream_t)stream;
  if (lmax == 6 && mmax == 2) return rb_tc_launch<6, 2>(v0, v1, v2, p, Ec, s);
  if (lmax == 4 && mmax == 2) return rb_tc_launch<4, 2>(v0, v1, v2, p, Ec, s);
  if (lmax == 8 && mmax == 2) return rb_tc_launch<8, 2>(v0, v1, v2, p, Ec, s);
  efl_set_error("efl_rotate_back_aggregate_tc_fwd: unsupported (lmax, mmax)");
  return EFL_ERR_UNSUPPORTED;
}
