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
#include <cuda_fp16.h>
#include <stdlib.h>

#include "tc_common.cuh"

namespace efl {
namespace s2tc {

using namespace efl::tc;

constexpr int CH = 128;
constexpr int MAX_STAGES = 8;
constexpr int MAX_GROUPS = 6;   // forward (6,2): 6 x 80 TMEM columns; the per-edge hand-off chain is latency-bound, so
                                // edges in flight (worker groups) are what buys throughput
constexpr int MAX_THREADS = (4 + 4 * MAX_GROUPS) * 32;
constexpr int TMEM_COLS = 512;

struct Params {
  int bwd;
  int M, KP, G, GP;        // coefficients, padded to 16; grid points, padded to 16
  int rows[3];             // rows per edge in each order buffer (ms0, 2*ms1, 2*ms2)
  int first[3];            // first m-primary row of each buffer
  int ms[3];               // degrees per order
  int Ec;
  int stages;              // operand ring depth (<= MAX_STAGES)
  int ncb, cht;            // channel blocks of 128 per item row, total channels (row pitch)
  int dup;                 // bwd: write order>0 rows in the dup layout (edge path) or plainly (node path)
  int NG;                  // worker groups (edges in flight per CTA)
  int gstride;             // TMEM columns per group: [D1 (aliased by D2) | D3 (bwd)]
  const float* T;          // to_grid   [G, M] fp32, m-primary columns
  const float* F;          // from_grid [G, M]
  const float* gate;       // [Ec, gate_ld]
  long long gate_ld;
  __nv_bfloat16* out[3];   // fwd: plain split buffers; bwd: dup-layout buffers (out[0] already offset to hidden cols)
  long long ld0;           // pitch (elements) between consecutive items of dgate
  long long ld_out;        // pitch (elements) between consecutive items of out[0] (plain output path)
  __nv_bfloat16* dgate;    // bwd: [Ec, ld0] view, already offset
  const __nv_bfloat16* dy0;  // bwd: order-0 rows of dY (row 0 feeds d gate)
  int dbg;                   // timing experiments, always 0 in the shipped entry points (see the launcher)
};

__device__ __forceinline__ float silu_fast(float x) {
  const float h = 0.5f * x;
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
  return fmaf(h, t, h);
}
// SiLU of two fp32 accumulator values GIVEN AS h = x/2 (the to_grid operand in smem is pre-scaled by 1/2, exact in
// bf16), evaluated in packed fp16: x*sigmoid(x) = h + h*tanh(h).  Result = fp16x2 operand word (lo = first value).  fp16 keeps 11 mantissa bits (the bf16 operand it
// replaces kept 8).
__device__ __forceinline__ uint32_t silu_f16x2(uint32_t lo_f32_bits, uint32_t hi_f32_bits) {
  uint32_t r;
  asm("{\n\t"
      ".reg .b32 h, t;\n\t"
      "cvt.rn.f16x2.f32 h, %2, %1;\n\t"
      "tanh.approx.f16x2 t, h;\n\t"
      "fma.rn.f16x2 %0, h, t, h;\n\t"
      "}"
      : "=r"(r)
      : "f"(__uint_as_float(lo_f32_bits)), "f"(__uint_as_float(hi_f32_bits)));
  return r;
}
// w * silu'(x) given h = x/2 and w2 = w/2 (both halvings are folded into the constant MMA operands):
// silu'(x) = (1 + t + h (1 - t^2)) / 2, t = tanh(h)  ->  one MUFU + three FFMA per grid value
__device__ __forceinline__ float dsilu_times_half(float h, float w) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
  const float u = fmaf(-t, t, 1.f);
  const float v = fmaf(h, u, t);
  return fmaf(w, v, w);
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
__host__ __device__ inline Layout make_layout(int KP, int GP, int bwd, int NG, int STAGES) {
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
  L.off_bar = o; o += 512;
  L.total = o + 1024;
  return L;
}

template <int MS0, int MS1, int MS2, int GP, bool BWD, int MAXT>
__global__ void __launch_bounds__(MAXT, 1)
s2act_tc_kernel(const __grid_constant__ CUtensorMap mapH0, const __grid_constant__ CUtensorMap mapH1,
                const __grid_constant__ CUtensorMap mapH2, const __grid_constant__ CUtensorMap mapD0,
                const __grid_constant__ CUtensorMap mapD1, const __grid_constant__ CUtensorMap mapD2, const Params p) {
  constexpr int M = MS0 + 2 * MS1 + 2 * MS2;       // coefficient rows per item
  constexpr int F1 = MS0, F2 = MS0 + 2 * MS1;      // first row of the order-1 / order-2 buffers
  constexpr int R0 = MS0, R1 = 2 * MS1, R2 = 2 * MS2;
  constexpr int KP = (M + 15) / 16 * 16, KPC = KP / 16;
  constexpr bool DUP = BWD && MS1 > 0;             // edge backward writes the dup layout
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const Layout L = make_layout(KP, GP, BWD, p.NG, p.stages);
  const int STAGES = p.stages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint64_t* h_full = bars;             // [STAGES]
  uint64_t* h_empty = bars + MAX_STAGES;
  uint64_t* d1_full = bars + 2 * MAX_STAGES;  // [MAX_GROUPS]
  uint64_t* s_ready = bars + 2 * MAX_STAGES + MAX_GROUPS;
  uint64_t* d2_full = bars + 2 * MAX_STAGES + 2 * MAX_GROUPS;
  uint64_t* d2_free = bars + 2 * MAX_STAGES + 3 * MAX_GROUPS;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * MAX_STAGES + 4 * MAX_GROUPS);
  const int warp = uniform_warp_idx(), lane = threadIdx.x & 31;

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
    for (int i = threadIdx.x; i < p.G * M; i += blockDim.x) {
      const int g = i / M, j = i % M;
      const float t = p.T[i], f = p.F[i];
      // K-major, 128B swizzle, rows g (128 B each), K index j < 64
      const int off_sw = g * 64 + ((((j >> 3) ^ (g & 7)) << 3) | (j & 7));
      sT[off_sw] = __float2bfloat16_rn(0.5f * t);  // D1 = x/2 on the grid: what tanh wants (see silu_f16x2)
      if (BWD) sF0[off_sw] = __float2bfloat16_rn(j == 0 ? 0.f : f);
      // K-major, no swizzle: rows j, K index g: (j/8)*sbo + (g/8)*128 + (j%8)*16 + (g%8)*2 bytes
      const int off_ns = ((j >> 3) * L.sbo + (g >> 3) * 128 + (j & 7) * 16) / 2 + (g & 7);
      // forward: S and from_grid are fp16 operands (SiLU is evaluated in packed f16x2), backward: bf16
      if (BWD) sB2[off_ns] = __float2bfloat16_rn(0.5f * t);  // S holds 2 dG (dsilu_times_half)
      else reinterpret_cast<__half*>(sB2)[off_ns] = __float2half_rn(f);
    }
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < MAX_STAGES; ++s) { mbar_init(&h_full[s], 1); mbar_init(&h_empty[s], BWD ? 2 : 1); }
    for (int g = 0; g < MAX_GROUPS; ++g) {
      mbar_init(&d1_full[g], BWD ? 2 : 1);  // backward: warps 1 and 2 each commit their GEMM
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
  const int n_items = p.Ec * p.ncb;
  const int n_mine = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // items of this CTA
  const int row_bytes_total = M * 256;  // bytes per edge per operand (M rows x 128 ch x 2 B)

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < n_mine; ++i) {
        const int it = blockIdx.x + i * gridDim.x;
        const int e = it / p.ncb, cb = it % p.ncb;
        mbar_wait(&h_empty[stage], phase ^ 1);
        uint8_t* st = smem + stage * L.stage_bytes;
        mbar_expect_tx(&h_full[stage], (uint32_t)(row_bytes_total * (BWD ? 2 : 1)) >> (p.dbg & 1));
        for (int hb = 0; hb < ((p.dbg & 1) ? 1 : 2); ++hb) {
          uint8_t* blk = st + hb * L.blk_bytes;
          tma_load_2d(blk + 0, &mapH0, &h_full[stage], cb * CH + hb * 64, e * R0);
          if (R1 > 0) tma_load_2d(blk + F1 * 128, &mapH1, &h_full[stage], cb * CH + hb * 64, e * R1);
          if (R2 > 0) tma_load_2d(blk + F2 * 128, &mapH2, &h_full[stage], cb * CH + hb * 64, e * R2);
          if (BWD) {
            uint8_t* dblk = st + 2 * L.blk_bytes + hb * L.blk_bytes;
            tma_load_2d(dblk + 0, &mapD0, &h_full[stage], cb * CH + hb * 64, e * R0);
            if (R1 > 0) tma_load_2d(dblk + F1 * 128, &mapD1, &h_full[stage], cb * CH + hb * 64, e * R1);
            if (R2 > 0) tma_load_2d(dblk + F2 * 128, &mapD2, &h_full[stage], cb * CH + hb * 64, e * R2);
          }
        }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ---- MMA issuer 1: first GEMM(s) of every edge (group i % NG).  D2 aliases D1, so edge i may only start once
    // its group has drained D2 of edge i-NG (d2_free).  The second GEMMs are issued by warp 3: one issuing thread
    // for both put ~150 dependent instructions and four mbarrier waits per edge on a single thread and left the
    // worker warps waiting on d1_full/d2_full for 60 % of their time (ncu source view).
    int stage = 0;
    uint32_t phase = 0;
    const int NG = p.NG;
    const uint32_t idesc1 = make_idesc2(GP, 1, 0);   // A = operand tile (MN-major), B = T/F0 (K-major)
    // descriptors differ between stages / K steps only in the start-address field (bytes >> 4): advanced by adds;
    // warp-uniform control flow with an elected issuing lane keeps waterfall loops off every tcgen05 instruction
    const uint64_t a_desc0 = make_desc(smem_u32(smem), (uint32_t)L.blk_bytes, 1024);
    const uint64_t t_desc0 = make_desc(smem_u32(smem + L.off_T), 16, 1024);
    const uint64_t st_step = (uint64_t)(L.stage_bytes >> 4);
    const bool issuer = elect_one();
    for (int i = 0; i < n_mine; ++i) {
      const int g = i % NG;
      const uint32_t k_th = (uint32_t)(i / NG);
      mbar_wait(&h_full[stage], phase);
      mbar_wait(&d2_free[g], (k_th & 1) ^ 1);
      tc_fence_after();
      if (issuer) {
        const uint64_t ad = a_desc0 + st_step * (uint64_t)stage;
        const uint32_t d1 = tmem_base + (uint32_t)(g * p.gstride);
#pragma unroll
        for (int k = 0; k < KP / 16; ++k)
          tc_mma(d1, ad + (uint64_t)(k * (2048 >> 4)), t_desc0 + (uint64_t)(k * (32 >> 4)), idesc1, k > 0 ? 1u : 0u);
        tc_commit(&h_empty[stage]);
        tc_commit(&d1_full[g]);
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (BWD && warp == 2) {
    // ---- backward only, MMA issuer 1b: D3 = dY tile x F0 (to_grid of the upstream gradient), same waits as warp 1 ----
    int stage = 0;
    uint32_t phase = 0;
    const int NG = p.NG;
    const uint32_t idesc1 = make_idesc2(GP, 1, 0);
    const uint64_t d_desc0 = make_desc(smem_u32(smem) + 2 * L.blk_bytes, (uint32_t)L.blk_bytes, 1024);
    const uint64_t f_desc0 = make_desc(smem_u32(smem + L.off_F0), 16, 1024);
    const uint64_t st_step = (uint64_t)(L.stage_bytes >> 4);
    const bool issuer = elect_one();
    for (int i = 0; i < n_mine; ++i) {
      const int g = i % NG;
      const uint32_t k_th = (uint32_t)(i / NG);
      mbar_wait(&h_full[stage], phase);
      mbar_wait(&d2_free[g], (k_th & 1) ^ 1);
      tc_fence_after();
      if (issuer) {
        const uint64_t dd = d_desc0 + st_step * (uint64_t)stage;
        const uint32_t d3 = tmem_base + (uint32_t)(g * p.gstride) + (uint32_t)GP;
#pragma unroll
        for (int k = 0; k < KP / 16; ++k)
          tc_mma(d3, dd + (uint64_t)(k * (2048 >> 4)), f_desc0 + (uint64_t)(k * (32 >> 4)), idesc1, k > 0 ? 1u : 0u);
        tc_commit(&h_empty[stage]);
        tc_commit(&d1_full[g]);
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 3) {
    // ---- MMA issuer 2: second GEMM of every edge, as soon as its worker group has written S (s_ready) ----
    const int NG = p.NG;
    // A = S (K-major), B = B2 (K-major); forward: both fp16 (a_format = b_format = 0)
    const uint32_t idesc2 = BWD ? make_idesc2(KP, 0, 0) : (make_idesc2(KP, 0, 0) & ~((1u << 7) | (1u << 10)));
    const uint64_t s_desc0 = make_desc_noswz(smem_u32(smem + L.off_S), 128, (uint32_t)L.sbo);
    const uint64_t b2_desc0 = make_desc_noswz(smem_u32(smem + L.off_B2), 128, (uint32_t)L.sbo);
    const uint64_t g_step = (uint64_t)(L.s_bytes >> 4);
    const bool issuer = elect_one();
    for (int j = 0; j < n_mine; ++j) {
      const int g = j % NG;
      const uint32_t k_th = (uint32_t)(j / NG);
      mbar_wait(&s_ready[g], k_th & 1);
      tc_fence_after();
      if (issuer) {
        const uint64_t sd = s_desc0 + g_step * (uint64_t)g;
        const uint32_t d2 = tmem_base + (uint32_t)(g * p.gstride);
#pragma unroll
        for (int k = 0; k < GP / 16; ++k)
          tc_mma(d2, sd + (uint64_t)(k * (256 >> 4)), b2_desc0 + (uint64_t)(k * (256 >> 4)), idesc2, k > 0 ? 1u : 0u);
        tc_commit(&d2_full[g]);
      }
      __syncwarp();
    }
  } else if (warp >= 4 && warp < 4 + 4 * p.NG) {
    // ---- worker groups: thread = channel ----
    const int g = (warp - 4) >> 2;
    const int q = warp & 3;
    const int c = q * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    const uint32_t d1 = tmem_base + (uint32_t)(g * p.gstride) + lane_addr;
    const uint32_t d3 = d1 + (uint32_t)GP;
    const uint32_t d2 = d1;
    uint8_t* sS = smem + L.off_S + g * L.s_bytes + (c >> 3) * L.sbo + (c & 7) * 16;
    uint32_t k_th = 0;
    for (int i = g; i < n_mine; i += p.NG, ++k_th) {
      const int it = blockIdx.x + i * gridDim.x;
      const int e = it / p.ncb, cc = (it % p.ncb) * CH + c;
      const size_t cht = (size_t)p.cht;
      // gate path first: its global loads overlap the wait below
      // (the loads are issued here; their first use is after the second GEMM, so their latency hides behind the
      // SiLU phase instead of stalling the warp in front of it)
      const float gt = __ldg(p.gate + (size_t)e * p.gate_ld + cc);
      __nv_bfloat16 dy0v = __float2bfloat16_rn(0.f);
      if (BWD) dy0v = p.dy0[(size_t)e * R0 * cht + cc];
      mbar_wait(&d1_full[g], k_th & 1);
      tc_fence_after();
      if (!BWD) {
#pragma unroll 1
        for (int g0 = 0; g0 < GP; g0 += 32) {
          uint32_t v[32];
          tc_ld16(d1 + (uint32_t)g0, v);
          if (g0 + 16 < GP) tc_ld16(d1 + (uint32_t)(g0 + 16), v + 16);
          tc_ld_wait();
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            if (g0 + 8 * h < GP) {
              uint32_t s2[4];
#pragma unroll
              for (int t = 0; t < 4; ++t) s2[t] = silu_f16x2(v[8 * h + 2 * t], v[8 * h + 2 * t + 1]);
              *reinterpret_cast<uint4*>(sS + ((g0 >> 3) + h) * 128) = make_uint4(s2[0], s2[1], s2[2], s2[3]);
            }
          }
        }
      } else {
        // 16 grid columns per step; the TMEM loads of step k+1 are in flight while step k is evaluated (only three
        // worker warps per scheduler in the backward kernel: the load latency would otherwise be exposed every step)
        constexpr int NCH = GP / 16;
        uint32_t v[2][16], w[2][16];
        tc_ld16(d1, v[0]);
        tc_ld16(d3, w[0]);
        tc_ld_wait();
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
          if (k + 1 < NCH) {
            tc_ld16(d1 + 16u * (k + 1), v[(k + 1) & 1]);
            tc_ld16(d3 + 16u * (k + 1), w[(k + 1) & 1]);
          }
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            float s[8];
#pragma unroll
            for (int t = 0; t < 8; ++t)
              s[t] = dsilu_times_half(__uint_as_float(v[k & 1][8 * h + t]), __uint_as_float(w[k & 1][8 * h + t]));
            *reinterpret_cast<uint4*>(sS + (2 * k + h) * 128) =
                make_uint4(pack_bf16(s[0], s[1]), pack_bf16(s[2], s[3]), pack_bf16(s[4], s[5]), pack_bf16(s[6], s[7]));
          }
          if (k + 1 < NCH) tc_ld_wait();
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
      float y0;
      if (!BWD) y0 = silu_fast(gt);
      else y0 = __bfloat162float(dy0v) * dsilu_fast(gt);
      if (!DUP) {
        if (!BWD) yv[0] = __float_as_uint(y0);
        else p.dgate[(size_t)e * p.ld0 + cc] = __float2bfloat16_rn(y0);
        __nv_bfloat16* o0 = p.out[0] + (size_t)e * (size_t)p.ld_out + cc;
        __nv_bfloat16* o1 = p.out[1] + (size_t)e * R1 * cht + cc;
        __nv_bfloat16* o2 = p.out[2] + (size_t)e * R2 * cht + cc;
        if constexpr (MS1 == 0) {
          // One row block (edge rows / node blocks): adjacent lanes hold adjacent channels, so a lane pair swaps one
          // value per two coefficient rows and each lane stores a packed bf16x2 word - the even lane (row j, channels
          // c, c+1), the odd lane (row j+1, channels c-1, c).  Half the store instructions of the scalar form, 4-byte
          // instead of 2-byte accesses (the scalar stores held this kernel at 4.3 TB/s; the old paired layout ran at 6.1).
          const bool odd = lane & 1;
#pragma unroll
          for (int j = 0; j + 1 < M; j += 2) {
            const float mine_keep = __uint_as_float(odd ? yv[j + 1] : yv[j]);
            const float mine_send = __uint_as_float(odd ? yv[j] : yv[j + 1]);
            const float got = __shfl_xor_sync(0xffffffffu, mine_send, 1);
            __nv_bfloat16* o = o0 + (size_t)(j + (odd ? 1 : 0)) * cht - (odd ? 1 : 0);
            *reinterpret_cast<uint32_t*>(o) = odd ? pack_bf16(got, mine_keep) : pack_bf16(mine_keep, got);
          }
          if (M & 1) o0[(size_t)(M - 1) * cht] = __float2bfloat16_rn(__uint_as_float(yv[M - 1]));
        } else {
#pragma unroll
          for (int j = 0; j < M; ++j) {
            if (p.dbg & 2) break;
            __nv_bfloat16* o = j >= F2 ? o2 + (size_t)(j - F2) * cht : (j >= F1 ? o1 + (size_t)(j - F1) * cht : o0 + (size_t)j * cht);
            *o = __float2bfloat16_rn(__uint_as_float(yv[j]));
          }
        }
      } else {
        p.dgate[(size_t)e * p.ld0 + c] = __float2bfloat16_rn(y0);
        __nv_bfloat16* o0 = p.out[0] + (size_t)e * p.ld0 + c;
#pragma unroll
        for (int j = 0; j < F1; ++j) o0[j * CH] = __float2bfloat16_rn(__uint_as_float(yv[j]));
        // order m > 0: row t of the +m block pairs with row t of the -m block (MS rows later)
        __nv_bfloat16* o1 = p.out[1] + (size_t)e * 4 * MS1 * CH + 2 * c;
#pragma unroll
        for (int t = 0; t < MS1; ++t) {
          const float gr = __uint_as_float(yv[F1 + t]), gi = __uint_as_float(yv[F1 + MS1 + t]);
          *reinterpret_cast<uint32_t*>(o1 + (size_t)t * 2 * CH) = pack_bf16(gr, gi);
          *reinterpret_cast<uint32_t*>(o1 + (size_t)(t + MS1) * 2 * CH) = pack_bf16(gi, -gr);
        }
        __nv_bfloat16* o2 = p.out[2] + (size_t)e * 4 * MS2 * CH + 2 * c;
#pragma unroll
        for (int t = 0; t < MS2; ++t) {
          const float gr = __uint_as_float(yv[F2 + t]), gi = __uint_as_float(yv[F2 + MS2 + t]);
          *reinterpret_cast<uint32_t*>(o2 + (size_t)t * 2 * CH) = pack_bf16(gr, gi);
          *reinterpret_cast<uint32_t*>(o2 + (size_t)(t + MS2) * 2 * CH) = pack_bf16(gi, -gr);
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

static int s2_launch(efl::s2tc::Params& p, const void* const* hs, const void* const* ds, long long n_rows_items,
                     cudaStream_t st) {
  using namespace efl::s2tc;
  p.dbg = 0;  // the partial-load timing experiments (bit 0: first channel half only, bit 1: skip the stores) are compiled in
              // but not reachable from the environment: a library switch must not be able to corrupt results
  p.gstride = p.bwd ? 2 * p.GP : p.GP;
  p.NG = TMEM_COLS / p.gstride;
  // worker groups: forward of the (6,2) edge shape runs 6 (896 threads, <= 73 registers); everything else <= 4
  const int ngcap = (!p.bwd && p.M == 29 && p.GP == 80) ? MAX_GROUPS : 4;
  if (p.NG > ngcap) p.NG = ngcap;
  if (p.KP > 48 && p.NG > 2) p.NG = 2;
  // operand ring: as deep as shared memory allows (<= MAX_STAGES; EFL_S2_STAGES caps it for timing experiments)
  p.stages = 4;
  { const char* d = getenv("EFL_S2_STAGES"); if (d && atoi(d) >= 2 && atoi(d) <= MAX_STAGES) p.stages = atoi(d); }
  Layout L = make_layout(p.KP, p.GP, p.bwd, p.NG, p.stages);
  const int limit = 227 * 1024;
  while (L.total > limit && p.stages > 2) { --p.stages; L = make_layout(p.KP, p.GP, p.bwd, p.NG, p.stages); }
  EFL_REQUIRE(L.total <= limit, "efl_s2act_tc: shared memory budget exceeded");
  CUtensorMap mh[3], md[3];
  for (int s = 0; s < 3; ++s) {
    const int r = p.rows[s] > 0 ? p.rows[s] : 1;
    int rc = efl_make_map_bf16(&mh[s], hs[p.rows[s] > 0 ? s : 0], n_rows_items * r, p.cht, p.cht, r);
    if (rc) return rc;
    rc = efl_make_map_bf16(&md[s], p.bwd ? ds[p.rows[s] > 0 ? s : 0] : hs[p.rows[s] > 0 ? s : 0], n_rows_items * r,
                           p.cht, p.cht, r);
    if (rc) return rc;
  }
  const long long items = (long long)p.Ec * p.ncb;
  const int grid = items < efl_num_sms() ? (int)items : efl_num_sms();
  const int threads = (4 + 4 * p.NG) * 32;
  const size_t smem = (size_t)L.total;
#define S2_LAUNCH(MS0, MS1, MS2, GPV, MAXT, MAXTB)                                                                \
  {                                                                                                              \
    if (p.bwd) {                                                                                                 \
      auto kern = s2act_tc_kernel<MS0, MS1, MS2, GPV, true, MAXTB>;                                              \
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, limit);                            \
      kern<<<grid, threads, smem, st>>>(mh[0], mh[1], mh[2], md[0], md[1], md[2], p);                            \
    } else {                                                                                                     \
      auto kern = s2act_tc_kernel<MS0, MS1, MS2, GPV, false, MAXT>;                                              \
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, limit);                            \
      kern<<<grid, threads, smem, st>>>(mh[0], mh[1], mh[2], md[0], md[1], md[2], p);                            \
    }                                                                                                            \
  }
  if (p.ms[0] == 29 && p.ms[1] == 0 && p.GP == 80) S2_LAUNCH(29, 0, 0, 80, MAX_THREADS, 512)
  else if (p.ms[0] == 39 && p.ms[1] == 0 && p.GP == 96) S2_LAUNCH(39, 0, 0, 96, 640, 640)
  else if (p.ms[0] == 49 && p.ms[1] == 0 && p.GP == 224) S2_LAUNCH(49, 0, 0, 224, 384, 384)
  else {
    efl_set_error("efl_s2act_tc: built for edge rows of (lmax,mmax,G) = (6,2,70), (8,2,90) and node blocks (49 rows, G=210)");
    return EFL_ERR_UNSUPPORTED;
  }
#undef S2_LAUNCH
  EFL_CHECK_LAUNCH();
  return EFL_OK;
}

// Edge path.  h: hidden rows [Ec, M, 128] bf16 (one row of M*128 per edge, m-primary coefficient order, dense);
// fwd: out[e] = &out[e * ld_out], M*128 values; gate [Ec, gate_ld] fp32.
// bwd: additionally d = d(out) rows, dense like h; out = &dY1[0][hidden column], dgate = &dY1[0][gate column],
//      both with item pitch ld_out (the row pitch of dY1).  T/F: to/from grid [G, M] fp32 with m-primary columns.
extern "C" int efl_s2act_tc(int bwd, int lmax, int mmax, int G, const void* h, const void* d, void* out,
                            long long ld_out, void* dgate, const float* gate, long long gate_ld, const float* T,
                            const float* F, int Ec, void* stream) {
  using namespace efl::s2tc;
  if (Ec <= 0) return EFL_OK;
  EFL_REQUIRE(mmax == 2, "efl_s2act_tc: built for mmax == 2");
  Params p{};
  p.bwd = bwd;
  p.M = (lmax + 1) + 2 * lmax + 2 * (lmax - 1);
  p.ms[0] = p.M; p.ms[1] = 0; p.ms[2] = 0;
  p.rows[0] = p.M; p.rows[1] = 0; p.rows[2] = 0;
  p.first[0] = 0; p.first[1] = p.M; p.first[2] = p.M;
  p.KP = (p.M + 15) / 16 * 16;
  p.G = G; p.GP = (G + 15) / 16 * 16;
  EFL_REQUIRE(p.KP <= 64 && p.GP <= 224, "efl_s2act_tc: coefficient / grid size out of range");
  EFL_REQUIRE(ld_out >= (long long)p.M * CH, "efl_s2act_tc: output pitch smaller than one item");
  p.Ec = Ec; p.ncb = 1; p.cht = CH; p.dup = 0;
  p.T = T; p.F = F; p.gate = gate; p.gate_ld = gate_ld;
  p.out[0] = p.out[1] = p.out[2] = (__nv_bfloat16*)out;
  p.ld0 = ld_out; p.ld_out = ld_out; p.dgate = (__nv_bfloat16*)dgate; p.dy0 = (const __nv_bfloat16*)d;
  const void* hs[3] = {h, h, h};
  const void* ds[3] = {d, d, d};
  return s2_launch(p, hs, ds, Ec, (cudaStream_t)stream);
}

// Node-level variant (FeedForwardNetwork, EF_layers.py:234-270): x [N, M, cht] bf16 (l-primary rows, cht a multiple
// of 128), gate [N, gate_ld] fp32, T/F [G, M] fp32.  fwd: out [N, M, cht] bf16.  bwd: dy [N, M, cht] -> out = d x,
// dgate [N, cht] bf16.
extern "C" int efl_s2act_tc_nodes(int bwd, int M, int G, int cht, const void* x, const void* dy, void* out,
                                  void* dgate, const float* gate, long long gate_ld, const float* T, const float* F,
                                  int N, void* stream) {
  using namespace efl::s2tc;
  if (N <= 0) return EFL_OK;
  EFL_REQUIRE(cht % CH == 0, "efl_s2act_tc_nodes: channels must be a multiple of 128");
  Params p{};
  p.bwd = bwd;
  p.ms[0] = M; p.ms[1] = 0; p.ms[2] = 0;
  p.rows[0] = M; p.rows[1] = 0; p.rows[2] = 0;
  p.first[0] = 0; p.first[1] = M; p.first[2] = M;
  p.M = M;
  p.KP = (M + 15) / 16 * 16;
  p.G = G; p.GP = (G + 15) / 16 * 16;
  EFL_REQUIRE(p.KP <= 64 && p.GP <= 224, "efl_s2act_tc_nodes: coefficient / grid size out of range");
  p.Ec = N; p.ncb = cht / CH; p.cht = cht; p.dup = 0;
  p.T = T; p.F = F; p.gate = gate; p.gate_ld = gate_ld;
  p.out[0] = p.out[1] = p.out[2] = (__nv_bfloat16*)out;
  p.ld0 = cht; p.ld_out = (long long)M * cht; p.dgate = (__nv_bfloat16*)dgate; p.dy0 = (const __nv_bfloat16*)dy;
  const void* hs[3] = {x, x, x};
  const void* ds[3] = {dy, dy, dy};
  return s2_launch(p, hs, ds, N, (cudaStream_t)stream);
}
