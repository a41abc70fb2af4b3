// Separable S2 activation of the edge path on the tensor cores (bf16 mode), forward and backward.
//
// Replaces SeparableS2Activation / S2Activation (reference EF_layers.py:1736-1773) for the per-edge hidden
// features [M=29 coefficients x CH=128 channels].  Per edge, with the channels as the UMMA M dimension:
//   fwd:  D1[c, g] = sum_j H[c, j] T[g, j]          tcgen05.mma, A = H tile (MN-major, TMA, 128B swizzle)
//         S = SiLU(D1)                               TMEM -> registers -> bf16 -> shared (K-major operand)
//         D2[c, j] = sum_g S[c, g] F[g, j]          tcgen05.mma
//         y = D2, y[0] = SiLU(gate)                  TMEM -> registers -> global
//   bwd:  D1 as above, D3[c, g] = sum_{j>0} dY[c, j] F[g, j];  dG = D3 * SiLU'(D1);
//         D2[c, j] = sum_g dG[c, g] T[g, j]  -> d hidden rows (written in the dup layout of the so2_conv_1
//         backward GEMMs) and d gate = dY[0] * SiLU'(gate).
// The [E, G, CH] grid tensor (32 GB per layer at config 2 in the reference) only ever exists in TMEM.
//
// Persistent CTAs, 12 warps: warp 0 TMA loader, warp 1 MMA issuer, warp 2 TMEM allocator, warps 4-7 and 8-11
// two worker groups handling alternate edges (thread = channel = TMEM lane), each with private TMEM accumulators
// and a private S buffer, so the SiLU of one edge overlaps the MMAs / output of the other.
#include "tc_common.cuh"

namespace efl {
namespace s2tc {

using namespace efl::tc;

constexpr int CH = 128;
constexpr int STAGES = 4;
constexpr int MAX_GROUPS = 4;
constexpr int MAX_THREADS = (4 + 4 * MAX_GROUPS) * 32;
constexpr int TMEM_COLS = 512;

struct Params {
  int bwd;
  int M, KP, G, GP;        // coefficients, padded to 16; grid points, padded to 16
  int rows[3];             // rows per edge in each order buffer (ms0, 2*ms1, 2*ms2)
  int first[3];            // first m-primary row of each buffer
  int ms[3];               // degrees per order
  int Ec;
  int NG;                  // worker groups (edges in flight per CTA)
  int gstride;             // TMEM columns per group: [D1 (aliased by D2) | D3 (bwd)]
  const float* T;          // to_grid   [G, M] fp32, m-primary columns
  const float* F;          // from_grid [G, M]
  const float* gate;       // [Ec, gate_ld]
  long long gate_ld;
  __nv_bfloat16* out[3];   // fwd: plain split buffers; bwd: dup-layout buffers (out[0] already offset to hidden cols)
  long long ld0;           // row pitch (elements) of out[0] per edge
  __nv_bfloat16* dgate;    // bwd: [Ec, ld0] view, already offset
  const __nv_bfloat16* dy0;  // bwd: order-0 rows of dY (row 0 feeds d gate)
};

__device__ __forceinline__ float silu_fast(float x) {
  const float h = 0.5f * x;
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
  return fmaf(h, t, h);
}
__device__ __forceinline__ float dsilu_fast(float x) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * x));
  const float s = fmaf(0.5f, t, 0.5f);
  return s * fmaf(x, 1.f - s, 1.f);
}

__device__ __forceinline__ void tc_ld8(uint32_t taddr, uint32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
}

// smem carve-up (bytes), all relative to a 1024-aligned base
struct Layout {
  int stage_bytes, blk_bytes;  // one operand tile = 2 channel blocks of KP*128 bytes
  int off_T, off_F0, off_S, off_B2, off_bar, s_bytes, sbo;
  int total;
};
__host__ __device__ inline Layout make_layout(int KP, int GP, int bwd, int NG) {
  Layout L;
  L.blk_bytes = KP * 128;
  L.stage_bytes = 2 * L.blk_bytes * (bwd ? 2 : 1);
  int o = STAGES * L.stage_bytes;
  L.off_T = o;  o += GP * 128;
  L.off_F0 = o; o += bwd ? GP * 128 : 0;
  L.sbo = (GP / 8) * 128;            // K-major no-swizzle: bytes between 8-row groups
  L.s_bytes = (CH / 8) * L.sbo;
  L.off_S = o;  o += NG * L.s_bytes;
  L.off_B2 = o; o += (KP / 8) * L.sbo;
  L.off_bar = o; o += 256;
  L.total = o + 1024;
  return L;
}

template <int KPC>
__global__ void __launch_bounds__(MAX_THREADS, 1)
s2act_tc_kernel(const __grid_constant__ CUtensorMap mapH0, const __grid_constant__ CUtensorMap mapH1,
                const __grid_constant__ CUtensorMap mapH2, const __grid_constant__ CUtensorMap mapD0,
                const __grid_constant__ CUtensorMap mapD1, const __grid_constant__ CUtensorMap mapD2, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const Layout L = make_layout(p.KP, p.GP, p.bwd, p.NG);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint64_t* h_full = bars;             // [STAGES]
  uint64_t* h_empty = bars + STAGES;   // [STAGES]
  uint64_t* d1_full = bars + 8;        // [MAX_GROUPS]
  uint64_t* s_ready = bars + 12;
  uint64_t* d2_full = bars + 16;
  uint64_t* d2_free = bars + 20;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 24);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- one-time setup: zero operand stages (pad rows must be 0), S buffers; build constant B operands ----
  {
    uint32_t* z = reinterpret_cast<uint32_t*>(smem);
    const int nwords = L.off_bar / 4;
    for (int i = threadIdx.x; i < nwords; i += blockDim.x) z[i] = 0u;
  }
  __syncthreads();
  {
    __nv_bfloat16* sT = reinterpret_cast<__nv_bfloat16*>(smem + L.off_T);
    __nv_bfloat16* sF0 = reinterpret_cast<__nv_bfloat16*>(smem + L.off_F0);
    __nv_bfloat16* sB2 = reinterpret_cast<__nv_bfloat16*>(smem + L.off_B2);
    for (int i = threadIdx.x; i < p.G * p.M; i += blockDim.x) {
      const int g = i / p.M, j = i % p.M;
      const float t = p.T[i], f = p.F[i];
      // K-major, 128B swizzle, rows g (128 B each), K index j < 64
      const int off_sw = g * 64 + ((((j >> 3) ^ (g & 7)) << 3) | (j & 7));
      sT[off_sw] = __float2bfloat16_rn(t);
      if (p.bwd) sF0[off_sw] = __float2bfloat16_rn(j == 0 ? 0.f : f);
      // K-major, no swizzle: rows j, K index g: (j/8)*sbo + (g/8)*128 + (j%8)*16 + (g%8)*2 bytes
      const int off_ns = ((j >> 3) * L.sbo + (g >> 3) * 128 + (j & 7) * 16) / 2 + (g & 7);
      sB2[off_ns] = __float2bfloat16_rn(p.bwd ? t : f);
    }
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&h_full[s], 1); mbar_init(&h_empty[s], 1); }
    for (int g = 0; g < MAX_GROUPS; ++g) {
      mbar_init(&d1_full[g], 1);
      mbar_init(&s_ready[g], 128);
      mbar_init(&d2_full[g], 1);
      mbar_init(&d2_free[g], 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy smem writes -> async proxy (MMA)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_mine = (p.Ec - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // edges of this CTA
  const int row_bytes_total = p.M * 256;  // bytes per edge per operand (M rows x 128 ch x 2 B)

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < n_mine; ++i) {
        const int e = blockIdx.x + i * gridDim.x;
        mbar_wait(&h_empty[stage], phase ^ 1);
        uint8_t* st = smem + stage * L.stage_bytes;
        mbar_expect_tx(&h_full[stage], (uint32_t)(row_bytes_total * (p.bwd ? 2 : 1)));
        for (int hb = 0; hb < 2; ++hb) {
          uint8_t* blk = st + hb * L.blk_bytes;
          tma_load_2d(blk + p.first[0] * 128, &mapH0, &h_full[stage], hb * 64, e * p.rows[0]);
          tma_load_2d(blk + p.first[1] * 128, &mapH1, &h_full[stage], hb * 64, e * p.rows[1]);
          tma_load_2d(blk + p.first[2] * 128, &mapH2, &h_full[stage], hb * 64, e * p.rows[2]);
          if (p.bwd) {
            uint8_t* dblk = st + 2 * L.blk_bytes + hb * L.blk_bytes;
            tma_load_2d(dblk + p.first[0] * 128, &mapD0, &h_full[stage], hb * 64, e * p.rows[0]);
            tma_load_2d(dblk + p.first[1] * 128, &mapD1, &h_full[stage], hb * 64, e * p.rows[1]);
            tma_load_2d(dblk + p.first[2] * 128, &mapD2, &h_full[stage], hb * 64, e * p.rows[2]);
          }
        }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ---- MMA issuer ----
    // iteration i: first GEMM(s) of edge i (group i % NG), then the second GEMM of edge i-(NG-1), whose worker
    // group has had NG-1 iterations to finish its SiLU.  D2 aliases D1, so edge i may only start once its group
    // has drained D2 of edge i-NG (d2_free).
    int stage = 0;
    uint32_t phase = 0;
    const int NG = p.NG, lag = p.NG - 1;
    const uint32_t idesc1 = make_idesc2(p.GP, 1, 0);   // A = operand tile (MN-major), B = T/F0 (K-major)
    const uint32_t idesc2 = make_idesc2(p.KP, 0, 0);   // A = S (K-major), B = B2 (K-major)
    const uint32_t sT = smem_u32(smem + L.off_T), sF0 = smem_u32(smem + L.off_F0), sB2 = smem_u32(smem + L.off_B2);
    for (int i = 0; i < n_mine + lag; ++i) {
      const int j = i - lag;
      if (j >= 0 && j < n_mine) {
        const int g = j % NG;
        const uint32_t k_th = (uint32_t)(j / NG);
        mbar_wait(&s_ready[g], k_th & 1);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t sS = smem_u32(smem + L.off_S + g * L.s_bytes);
          const uint32_t d2 = tmem_base + (uint32_t)(g * p.gstride);
          for (int k = 0; k < p.GP / 16; ++k)
            tc_mma(d2, make_desc_noswz(sS + k * 256, 128, (uint32_t)L.sbo), make_desc_noswz(sB2 + k * 256, 128, (uint32_t)L.sbo),
                   idesc2, k > 0 ? 1u : 0u);
          tc_commit(&d2_full[g]);
        }
        __syncwarp();
      }
      if (i < n_mine) {
        const int g = i % NG;
        const uint32_t k_th = (uint32_t)(i / NG);
        mbar_wait(&h_full[stage], phase);
        mbar_wait(&d2_free[g], (k_th & 1) ^ 1);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t sa = smem_u32(smem + stage * L.stage_bytes);
          const uint32_t d1 = tmem_base + (uint32_t)(g * p.gstride);
          for (int k = 0; k < p.KP / 16; ++k)
            tc_mma(d1, make_desc(sa + k * 2048, (uint32_t)L.blk_bytes, 1024), make_desc(sT + k * 32, 16, 1024), idesc1,
                   k > 0 ? 1u : 0u);
          if (p.bwd) {
            const uint32_t sd = sa + 2 * L.blk_bytes;
            const uint32_t d3 = d1 + (uint32_t)p.GP;
            for (int k = 0; k < p.KP / 16; ++k)
              tc_mma(d3, make_desc(sd + k * 2048, (uint32_t)L.blk_bytes, 1024), make_desc(sF0 + k * 32, 16, 1024),
                     idesc1, k > 0 ? 1u : 0u);
          }
          tc_commit(&h_empty[stage]);
          tc_commit(&d1_full[g]);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp >= 4 && warp < 4 + 4 * p.NG) {
    // ---- worker groups: thread = channel ----
    const int g = (warp - 4) >> 2;
    const int q = warp & 3;
    const int c = q * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    const uint32_t d1 = tmem_base + (uint32_t)(g * p.gstride) + lane_addr;
    const uint32_t d3 = d1 + (uint32_t)p.GP;
    const uint32_t d2 = d1;
    uint8_t* sS = smem + L.off_S + g * L.s_bytes + (c >> 3) * L.sbo + (c & 7) * 16;
    const int f1 = p.first[1], f2 = p.first[2], ms1 = p.ms[1], ms2 = p.ms[2], nM = p.M;
    uint32_t k_th = 0;
    for (int i = g; i < n_mine; i += p.NG, ++k_th) {
      const int e = blockIdx.x + i * gridDim.x;
      // gate path first: its global loads overlap the wait below
      const float gt = p.gate[(size_t)e * p.gate_ld + c];
      float y0;
      if (!p.bwd) y0 = silu_fast(gt);
      else y0 = __bfloat162float(p.dy0[(size_t)e * p.rows[0] * CH + c]) * dsilu_fast(gt);
      mbar_wait(&d1_full[g], k_th & 1);
      tc_fence_after();
      if (!p.bwd) {
        for (int g0 = 0; g0 < p.GP; g0 += 32) {
          uint32_t v[32];
          tc_ld16(d1 + (uint32_t)g0, v);
          if (g0 + 16 < p.GP) tc_ld16(d1 + (uint32_t)(g0 + 16), v + 16);
          tc_ld_wait();
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            if (g0 + 8 * h < p.GP) {
              float s[8];
#pragma unroll
              for (int t = 0; t < 8; ++t) s[t] = silu_fast(__uint_as_float(v[8 * h + t]));
              *reinterpret_cast<uint4*>(sS + ((g0 >> 3) + h) * 128) =
                  make_uint4(pack_bf16(s[0], s[1]), pack_bf16(s[2], s[3]), pack_bf16(s[4], s[5]), pack_bf16(s[6], s[7]));
            }
          }
        }
      } else {
        for (int g0 = 0; g0 < p.GP; g0 += 16) {
          uint32_t v[16], w[16];
          tc_ld16(d1 + (uint32_t)g0, v);
          tc_ld16(d3 + (uint32_t)g0, w);
          tc_ld_wait();
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            float s[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) s[t] = __uint_as_float(w[8 * h + t]) * dsilu_fast(__uint_as_float(v[8 * h + t]));
            *reinterpret_cast<uint4*>(sS + ((g0 >> 3) + h) * 128) =
                make_uint4(pack_bf16(s[0], s[1]), pack_bf16(s[2], s[3]), pack_bf16(s[4], s[5]), pack_bf16(s[6], s[7]));
          }
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      tc_fence_before();
      mbar_arrive(&s_ready[g]);
      mbar_wait(&d2_full[g], k_th & 1);
      tc_fence_after();
      uint32_t yv[16 * KPC];
#pragma unroll
      for (int t = 0; t < KPC; ++t) tc_ld16(d2 + 16u * t, yv + 16 * t);
      tc_ld_wait();
      tc_fence_before();
      mbar_arrive(&d2_free[g]);
      if (!p.bwd) {
        yv[0] = __float_as_uint(y0);
        __nv_bfloat16* o0 = p.out[0] + (size_t)e * p.rows[0] * CH + c;
        __nv_bfloat16* o1 = p.out[1] + (size_t)e * p.rows[1] * CH + c;
        __nv_bfloat16* o2 = p.out[2] + (size_t)e * p.rows[2] * CH + c;
#pragma unroll
        for (int j = 0; j < 16 * KPC; ++j) {
          if (j < nM) {
            __nv_bfloat16* o = j >= f2 ? o2 + (j - f2) * CH : (j >= f1 ? o1 + (j - f1) * CH : o0 + j * CH);
            *o = __float2bfloat16_rn(__uint_as_float(yv[j]));
          }
        }
      } else {
        p.dgate[(size_t)e * p.ld0 + c] = __float2bfloat16_rn(y0);
        __nv_bfloat16* o0 = p.out[0] + (size_t)e * p.ld0 + c;
        __nv_bfloat16* o1 = p.out[1] + (size_t)e * 4 * ms1 * CH + 2 * c;
        __nv_bfloat16* o2 = p.out[2] + (size_t)e * 4 * ms2 * CH + 2 * c;
#pragma unroll
        for (int j = 0; j < 16 * KPC; ++j) {
          if (j < f1) {
            o0[j * CH] = __float2bfloat16_rn(__uint_as_float(yv[j]));
          } else if (j < nM) {
            const bool two = j >= f2;
            const int ms = two ? ms2 : ms1;
            const int t = j - (two ? f2 : f1);
            if (t < ms) {
              // +m ("real") row t; partner -m row is yv[j + ms]
              uint32_t gi_u = 0u;
#pragma unroll
              for (int jj = 0; jj < 16 * KPC; ++jj) gi_u = (jj == j + ms) ? yv[jj] : gi_u;
              const float gr = __uint_as_float(yv[j]), gi = __uint_as_float(gi_u);
              __nv_bfloat16* r0 = (two ? o2 : o1) + (size_t)t * 2 * CH;
              __nv_bfloat16* r1 = r0 + 2 * ms * CH;
              *reinterpret_cast<uint32_t*>(r0) = pack_bf16(gr, gi);
              *reinterpret_cast<uint32_t*>(r1) = pack_bf16(gi, -gr);
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TMEM_COLS));
  }
}

}  // namespace s2tc
}  // namespace efl

// ---------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------
int efl_make_map_bf16(CUtensorMap* map, const void* base, long long rows, long long cols, long long ld, int box_rows);
int efl_num_sms();

// fwd: h0..h2 hidden rows (plain split buffers, bf16) -> out0..out2 (plain split buffers); gate [Ec, gate_ld] fp32.
// bwd: additionally d0..d2 = d(out) rows; out0 = &dY1_0[0][hidden column], out1/out2 dup-layout buffers, ld0 = row
//      pitch of dY1_0, dgate = &dY1_0[0][gate column].  T/F: to/from grid [G, M] fp32 with m-primary columns.
extern "C" int efl_s2act_tc(int bwd, int lmax, int mmax, int G, const void* h0, const void* h1, const void* h2,
                            const void* d0, const void* d1, const void* d2, void* out0, void* out1, void* out2,
                            long long ld0, void* dgate, const float* gate, long long gate_ld, const float* T,
                            const float* F, int Ec, void* stream) {
  using namespace efl::s2tc;
  if (Ec <= 0) return EFL_OK;
  EFL_REQUIRE(mmax == 2, "efl_s2act_tc: built for mmax == 2");
  Params p{};
  p.bwd = bwd;
  p.ms[0] = lmax + 1; p.ms[1] = lmax; p.ms[2] = lmax - 1;
  p.rows[0] = p.ms[0]; p.rows[1] = 2 * p.ms[1]; p.rows[2] = 2 * p.ms[2];
  p.first[0] = 0; p.first[1] = p.rows[0]; p.first[2] = p.rows[0] + p.rows[1];
  p.M = p.rows[0] + p.rows[1] + p.rows[2];
  p.KP = (p.M + 15) / 16 * 16;
  p.G = G; p.GP = (G + 15) / 16 * 16;
  EFL_REQUIRE(p.KP <= 64 && p.GP <= 96, "efl_s2act_tc: coefficient / grid size out of range");
  p.Ec = Ec; p.T = T; p.F = F; p.gate = gate; p.gate_ld = gate_ld;
  p.out[0] = (__nv_bfloat16*)out0; p.out[1] = (__nv_bfloat16*)out1; p.out[2] = (__nv_bfloat16*)out2;
  p.ld0 = ld0; p.dgate = (__nv_bfloat16*)dgate; p.dy0 = (const __nv_bfloat16*)d0;
  CUtensorMap mh[3], md[3];
  const void* hs[3] = {h0, h1, h2};
  const void* ds[3] = {d0, d1, d2};
  for (int s = 0; s < 3; ++s) {
    int rc = efl_make_map_bf16(&mh[s], hs[s], (long long)Ec * p.rows[s], CH, CH, p.rows[s]);
    if (rc) return rc;
    rc = efl_make_map_bf16(&md[s], bwd ? ds[s] : hs[s], (long long)Ec * p.rows[s], CH, CH, p.rows[s]);
    if (rc) return rc;
  }
  p.gstride = bwd ? 2 * p.GP : p.GP;
  p.NG = TMEM_COLS / p.gstride;
  if (p.NG > MAX_GROUPS) p.NG = MAX_GROUPS;
  const Layout L = make_layout(p.KP, p.GP, bwd, p.NG);
  EFL_REQUIRE(L.total <= 200 * 1024, "efl_s2act_tc: shared memory budget exceeded");
  static int attr_done = 0;
  if (attr_done < L.total) {
    cudaError_t e = cudaFuncSetAttribute(s2act_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(s2act_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(s2act_tc_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) { efl_set_error(cudaGetErrorString(e)); return (int)e; }
    attr_done = 200 * 1024;
  }
  const int grid = Ec < efl_num_sms() ? Ec : efl_num_sms();
  const int threads = (4 + 4 * p.NG) * 32;
  cudaStream_t st = (cudaStream_t)stream;
  if (p.KP == 32) s2act_tc_kernel<2><<<grid, threads, L.total, st>>>(mh[0], mh[1], mh[2], md[0], md[1], md[2], p);
  else if (p.KP == 48) s2act_tc_kernel<3><<<grid, threads, L.total, st>>>(mh[0], mh[1], mh[2], md[0], md[1], md[2], p);
  else s2act_tc_kernel<4><<<grid, threads, L.total, st>>>(mh[0], mh[1], mh[2], md[0], md[1], md[2], p);
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}
