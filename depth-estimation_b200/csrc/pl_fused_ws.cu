// pl_fused_ws.cu -- mean-field iteration, warp-specialised form (sm_100a).  L in {32, 64}, d = 5.
//
// A thread of the main group owns L/4 consecutive labels of ONE pixel through slice, softmax, operand
// split and epilogue (the TMEM lane <-> row mapping of tcgen05.ld / st is the same one); the two
// shared-memory-bound gather phases of a tile (slice, splat) run on different warp groups.  One CTA per SM, 768 threads
// (the two groups carry about the same number of instructions per warp):
//
//   main group  (warps 0-15, 128 rows x 4 column quarters, a thread owns L/4 labels of one pixel)
//       slice(k+1) -> softmax(k+1) -> X hi/lo into TMEM -> MMAs -> epilogue: P tile -> s_p[k&1]
//   splat group (warps 16-23)
//       P tile of tile k -> HBM (coalesced) and -> lattice (running sums over the vertex-sorted
//       entry ranges, red.global.add.v4)
//
// The P tile is double-buffered in shared memory and handed over with a full / empty mbarrier
// pair, so the splat of tile k overlaps the slice + softmax + MMA of tile k+1.  All global ->
// shared traffic is bulk TMA issued a tile ahead (E0 / P_prev rows, the tile's distinct lattice
// rows, slice record) or two tiles ahead (splat record, double-buffered).
#include <math.h>
#include <stdlib.h>

#include "pl_common.cuh"
#include "pl_tc.cuh"

namespace pl {

namespace {

constexpr int kTile = PL_TILE;
constexpr int kMain = 512;        // main-group threads
constexpr int kSplat = 256;       // splat-group threads
constexpr int kThreads = kMain + kSplat;
constexpr int kD1 = 6;
constexpr int kTE = kTile * kD1;
constexpr int kStage = 120;       // distinct lattice rows staged per tile (more: rows from L2)
constexpr int kRanges = 32;       // splat: entry ranges per tile (24 entries each)

__device__ int g_ws_timeouts = 0;

template <int L>
struct WCfg {
  static constexpr int CPT = L / 4;              // columns per main thread
  static constexpr int KCH = L / 4;
  static constexpr int B_CH = L * 16 + 16;
  static constexpr int B_BYTES = KCH * B_CH;
  static constexpr int PI = L + 4;               // padded row pitch in floats
  static constexpr int P_BYTES = kTile * PI * 4;
  static constexpr int OFF_P = 0;                           // P tile x 2
  static constexpr int OFF_MU = OFF_P + 2 * P_BYTES;
  static constexpr int OFF_IN = OFF_MU + 2 * B_BYTES;       // E0 | P_prev
  static constexpr int OFF_ROWS = OFF_IN + 2 * P_BYTES;
  static constexpr int OFF_RS = OFF_ROWS + kStage * PI * 4;
  static constexpr int OFF_RP = OFF_RS + 6 * kTE;           // splat record x 2
  static constexpr int OFF_RED = OFF_RP + 2 * 9 * kTE;      // max[4][128] | sum[4][128]
  static constexpr int OFF_BAR = OFF_RED + 2 * 4 * kTile * 4;
  static constexpr int SMEM = OFF_BAR + 10 * 8 + 16;
  static constexpr int TMEM_COLS = L <= 32 ? 128 : 256;     // D | X hi | X lo
};

struct WsArgs {
  const float* val_in;
  float* val_out;
  const float* e0;
  const float* p_prev;
  float* p_next;
  const float* mu_canon;
  float scale;
  int64_t n;
  int64_t ntiles;
  float inv_denom;
  const int32_t* vid;
  const uint8_t* rec_slice;
  const uint8_t* rec_splat;
  const int2* useg;
  const int* tile_off;
  const int* tile_U;
  uint4* fill;          // first iteration only: region to fill with 0xFF bytes (the blur's sentinel scratch)
  uint32_t fill_n16;    // its size in 16-byte units
};

__device__ __forceinline__ void red_add_v4(float* addr, float2 a, float2 b) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a.x), "f"(a.y),
               "f"(b.x), "f"(b.y));
}
__device__ __forceinline__ void ffma2(float2& acc, float w, float2 x) {
  unsigned long long a = *reinterpret_cast<unsigned long long*>(&acc);
  float2 ww = make_float2(w, w);
  asm("fma.rn.f32x2 %0, %1, %2, %0;"
      : "+l"(a)
      : "l"(*reinterpret_cast<unsigned long long*>(&ww)),
        "l"(*reinterpret_cast<unsigned long long*>(&x)));
  acc = *reinterpret_cast<float2*>(&a);
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// barrier among the threads of one group (id 1: main, id 2: splat)
__device__ __forceinline__ void group_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc::smem_u32(bar)) : "memory");
}

template <int L, bool FIRST>
__global__ void __launch_bounds__(kThreads, 1) mf_ws_kernel(const WsArgs a) {
  using C = WCfg<L>;
  constexpr int CPT = C::CPT, PI = C::PI;
  extern __shared__ __align__(128) unsigned char smem[];
  float* s_p0 = reinterpret_cast<float*>(smem + C::OFF_P);
  unsigned char* s_mu = smem + C::OFF_MU;
  float* s_e0 = reinterpret_cast<float*>(smem + C::OFF_IN);
  float* s_pp = s_e0 + kTile * PI;
  float* s_rows = reinterpret_cast<float*>(smem + C::OFF_ROWS);
  unsigned char* s_rs = smem + C::OFF_RS;
  unsigned char* s_rp0 = smem + C::OFF_RP;
  float* s_max = reinterpret_cast<float*>(smem + C::OFF_RED);   // [4][128]
  float* s_sum = s_max + 4 * kTile;                             // [4][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::OFF_BAR);
  // bars: [0] Mu landed  [1] MMAs done  [2] tile inputs landed  [3,4] splat record b landed
  //       [5,6] P tile b full (main -> splat)  [7,8] P tile b empty (splat -> main)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);
  const float* s_bary = reinterpret_cast<const float*>(s_rs);
  const uint16_t* s_loc = reinterpret_cast<const uint16_t*>(s_rs + 4 * kTE);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t stride = gridDim.x;
  const int64_t tile0 = blockIdx.x;
  if (FIRST && a.fill_n16 > 0) {   // fire-and-forget: nothing in this kernel reads it
    const uint4 s4 = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
    for (uint32_t i = blockIdx.x * kThreads + tid; i < a.fill_n16; i += gridDim.x * kThreads) a.fill[i] = s4;
  }
  if (tile0 >= a.ntiles) return;
  const uint32_t my_tiles = (uint32_t)((a.ntiles - tile0 + stride - 1) / stride);

  if (warp == 0) {
    if (lane == 0) {
      for (int b = 0; b < 5; b++) tc::mbar_init(&bars[b], 1);
      for (int b = 5; b < 7; b++) tc::mbar_init(&bars[b], kMain);
      for (int b = 7; b < 9; b++) tc::mbar_init(&bars[b], kSplat);
      tc::fence_mbar_init();
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tcgen05_fence_before();
  }
  // up to here nothing of the previous kernel's output has been touched (programmatic dependent launch)
  pdl_wait();
  pdl_launch_dependents();
  if (warp == 0) {
    if (lane == 0) {
      tc::mbar_arrive_expect_tx(&bars[0], 2 * C::B_BYTES);
      tc::tma_load_1d(s_mu, a.mu_canon, 2 * C::B_BYTES, &bars[0]);
      for (uint32_t k = 0; k < 2 && k < my_tiles; k++) {   // splat records of the first two tiles
        tc::mbar_arrive_expect_tx(&bars[3 + k], 9 * kTE);
        tc::tma_load_1d(s_rp0 + k * 9 * kTE, a.rec_splat + (tile0 + k * stride) * (int64_t)(9 * kTE),
                        9 * kTE, &bars[3 + k]);
      }
    }
  }
  __syncthreads();
  tc::tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp < kMain / 32) {
    // =================================== main group =======================================
    const int r = (warp & 3) * 32 + lane;   // row of the tile = TMEM lane
    const int h = warp >> 2;                // column quarter
    const int col0 = h * CPT;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;

    auto load_uvid = [&](int64_t t, int& U, int& uv) {
      U = 0; uv = 0;
      if (!FIRST && t < a.ntiles) {
        U = a.tile_U[t];
        const int u = tid - 2 * kTile;
        if (u >= 0 && u < U && u < kStage) uv = a.useg[a.tile_off[t] + u].x;
      }
    };
    auto issue_inputs = [&](int64_t t, int U, int uv) {
      const int nrows = (int)min((int64_t)kTile, a.n - t * kTile);
      const int Us = FIRST ? 0 : min(U, kStage);   // the first kStage distinct rows; the rest come from L2
      if (tid == 0) {
        uint32_t bytes = (uint32_t)nrows * (L * 4) * (FIRST ? 1 : 2);
        if (!FIRST) bytes += (uint32_t)Us * (L * 4) + 6 * kTE;
        tc::mbar_arrive_expect_tx(&bars[2], bytes);
        if (!FIRST) tc::tma_load_1d(s_rs, a.rec_slice + t * (int64_t)(6 * kTE), 6 * kTE, &bars[2]);
      }
      if (tid < kTile) {
        if (tid < nrows)
          tc::tma_load_1d(s_e0 + tid * PI, a.e0 + (t * kTile + tid) * L, L * 4, &bars[2]);
      } else if (tid < 2 * kTile) {
        const int rr = tid - kTile;
        if (!FIRST && rr < nrows)
          tc::tma_load_1d(s_pp + rr * PI, a.p_prev + (t * kTile + rr) * L, L * 4, &bars[2]);
      } else if (tid < 2 * kTile + kStage) {
        const int u = tid - 2 * kTile;
        if (u < Us) tc::tma_load_1d(s_rows + u * PI, a.val_in + (int64_t)uv * L, L * 4, &bars[2]);
      }
    };

    float e_reg[CPT];
    auto slice_tile = [&](int64_t t, int U) {
      const int nrows = (int)min((int64_t)kTile, a.n - t * kTile);
      if (r >= nrows) {
#pragma unroll
        for (int c = 0; c < CPT; c++) e_reg[c] = 0.f;
        return;
      }
      const float4* e0p = reinterpret_cast<const float4*>(s_e0 + r * PI + col0);
      if (FIRST) {
#pragma unroll
        for (int c4 = 0; c4 < CPT / 4; c4++) {
          const float4 e = e0p[c4];
          e_reg[4 * c4] = e.x; e_reg[4 * c4 + 1] = e.y; e_reg[4 * c4 + 2] = e.z; e_reg[4 * c4 + 3] = e.w;
        }
        return;
      }
      float2 acc[CPT / 2];
#pragma unroll
      for (int c = 0; c < CPT / 2; c++) acc[c] = make_float2(0.f, 0.f);
      const float2* wp = reinterpret_cast<const float2*>(s_bary + r * kD1);
      const uint32_t* lp = reinterpret_cast<const uint32_t*>(s_loc + r * kD1);
      const float2 w01 = wp[0], w23 = wp[1], w45 = wp[2];
      const float w[6] = {w01.x, w01.y, w23.x, w23.y, w45.x, w45.y};
      {
        // Rows of the tile's first kStage distinct vertices are staged in shared memory; entries of a
        // tile with more distinct vertices than that (sharp feature scales: most tiles have 130-180)
        // fetch just their overflow rows from L2 instead of sending the whole tile there.
        const uint32_t l01 = lp[0], l23 = lp[1], l45 = lp[2];
        const uint32_t lo[6] = {l01 & 0xFFFF, l01 >> 16, l23 & 0xFFFF, l23 >> 16, l45 & 0xFFFF, l45 >> 16};
        if (U <= kStage) {      // tile-uniform: every row is staged
#pragma unroll
          for (int j = 0; j < kD1; j++) {
            const float4* rp = reinterpret_cast<const float4*>(s_rows + lo[j] * PI + col0);
#pragma unroll
            for (int c4 = 0; c4 < CPT / 4; c4++) {
              const float4 x = rp[c4];
              ffma2(acc[2 * c4], w[j], make_float2(x.x, x.y));
              ffma2(acc[2 * c4 + 1], w[j], make_float2(x.z, x.w));
            }
          }
        } else {
          const int32_t* vp = a.vid + (t * kTile + r) * kD1;
#pragma unroll
          for (int j = 0; j < kD1; j++) {
            const float4* rp = lo[j] < (uint32_t)kStage
                                   ? reinterpret_cast<const float4*>(s_rows + lo[j] * PI + col0)
                                   : reinterpret_cast<const float4*>(a.val_in + (int64_t)vp[j] * L + col0);
#pragma unroll
            for (int c4 = 0; c4 < CPT / 4; c4++) {
              const float4 x = rp[c4];
              ffma2(acc[2 * c4], w[j], make_float2(x.x, x.y));
              ffma2(acc[2 * c4 + 1], w[j], make_float2(x.z, x.w));
            }
          }
        }
      }
      const float4* ppp = reinterpret_cast<const float4*>(s_pp + r * PI + col0);
#pragma unroll
      for (int c4 = 0; c4 < CPT / 4; c4++) {
        const float4 e = e0p[c4], p = ppp[c4];
        e_reg[4 * c4] = __fadd_rn(e.x, __fmaf_rn(acc[2 * c4].x, a.inv_denom, -p.x));
        e_reg[4 * c4 + 1] = __fadd_rn(e.y, __fmaf_rn(acc[2 * c4].y, a.inv_denom, -p.y));
        e_reg[4 * c4 + 2] = __fadd_rn(e.z, __fmaf_rn(acc[2 * c4 + 1].x, a.inv_denom, -p.z));
        e_reg[4 * c4 + 3] = __fadd_rn(e.w, __fmaf_rn(acc[2 * c4 + 1].y, a.inv_denom, -p.w));
      }
    };

    // prologue: tile 0 sliced with its loads exposed, tile 1 in flight
    int U0, uv0, U1, uv1, U2, uv2;
    load_uvid(tile0, U0, uv0);
    issue_inputs(tile0, U0, uv0);
    load_uvid(tile0 + stride, U1, uv1);
    if (!tc::mbar_wait(&bars[2], 0)) atomicAdd(&g_ws_timeouts, 1);
    slice_tile(tile0, U0);
    group_sync(1, kMain);
    if (tile0 + stride < a.ntiles) issue_inputs(tile0 + stride, U1, uv1);
    load_uvid(tile0 + 2 * stride, U2, uv2);

    constexpr float kLog2e = 1.4426950408889634f;
    for (uint32_t k = 0; k < my_tiles; k++) {
      const int64_t tile = tile0 + (int64_t)k * stride;
      const uint32_t b = k & 1;
      const int64_t next1 = tile + stride, next2 = tile + 2 * stride;

      // ---- softmax numerators -> X hi / lo rows in TMEM ----------------------------------
      {
        float m = -e_reg[0];
#pragma unroll
        for (int c = 1; c < CPT; c++) m = fmaxf(m, -e_reg[c]);
        s_max[h * kTile + r] = m;
        // only the four warps that own the column quarters of the same 32 rows have to meet here
        // (named barriers 3..6, 128 threads each): 8.5 % of the stall samples sat at the 512-thread form
        group_sync(3 + (warp & 3), 4 * 32);
        m = fmaxf(fmaxf(s_max[r], s_max[kTile + r]), fmaxf(s_max[2 * kTile + r], s_max[3 * kTile + r]));
        const float mb = -m * kLog2e;
        float sum = 0.f;
#pragma unroll
        for (int part = 0; part < CPT / 8; part++) {
          uint32_t xh[8], xl[8];
#pragma unroll
          for (int c = 0; c < 8; c++) {
            const float x = ex2_approx(__fmaf_rn(e_reg[8 * part + c], -kLog2e, mb));
            sum += x;
            const float hi = tc::tf32_rna(x);
            xh[c] = __float_as_uint(hi);
            xl[c] = __float_as_uint(tc::tf32_rna(x - hi));
          }
          tc::tmem_st_x8(tmem + lane_base + (uint32_t)(L + col0 + 8 * part), xh);
          tc::tmem_st_x8(tmem + lane_base + (uint32_t)(2 * L + col0 + 8 * part), xl);
        }
        tc::tmem_st_wait();
        s_sum[h * kTile + r] = sum;
      }
      tc::tcgen05_fence_before();
      group_sync(1, kMain);

      if (tid == 0) {
        if (k == 0 && !tc::mbar_wait(&bars[0], 0)) atomicAdd(&g_ws_timeouts, 1);
        tc::tcgen05_fence_after();
        constexpr uint32_t idesc = tc::idesc_tf32(kTile, L);
        const uint32_t b_hi = tc::smem_u32(s_mu), b_lo = b_hi + C::B_BYTES;
#pragma unroll
        for (int pass = 0; pass < 3; pass++) {  // lo*hi, hi*lo, hi*hi
          const uint32_t bb = (pass == 1) ? b_lo : b_hi;
#pragma unroll
          for (int ks = 0; ks < L / 8; ks++) {
            const uint64_t bd = tc::smem_desc_kmajor(bb + ks * 2 * C::B_CH, C::B_CH, 128);
            tc::mma_tf32_ts(tmem, tmem + (uint32_t)((pass == 0 ? 2 * L : L) + ks * 8), bd, idesc,
                            (pass | ks) != 0);
          }
        }
        tc::mma_commit(&bars[1]);
      }

      // ---- while the MMAs run: E of the next tile -----------------------------------------
      if (next1 < a.ntiles) {
        if (!tc::mbar_wait(&bars[2], (k + 1) & 1)) atomicAdd(&g_ws_timeouts, 1);
        slice_tile(next1, U1);
      }
      group_sync(1, kMain);   // input buffers are dead
      if (next2 < a.ntiles) issue_inputs(next2, U2, uv2);
      U1 = U2;
      load_uvid(tile + 3 * stride, U2, uv2);

      // ---- epilogue: TMEM -> P = acc * scale / sum -> P tile b ------------------------------
      if (!tc::mbar_wait(&bars[1], k & 1)) atomicAdd(&g_ws_timeouts, 1);
      tc::tcgen05_fence_after();
      if (k >= 2 && !tc::mbar_wait(&bars[7 + b], ((k >> 1) - 1) & 1)) atomicAdd(&g_ws_timeouts, 1);
      {
        float* s_p = s_p0 + b * (kTile * PI);
        const float inv = __fdividef(a.scale, (s_sum[r] + s_sum[kTile + r]) +
                                                  (s_sum[2 * kTile + r] + s_sum[3 * kTile + r]));
#pragma unroll
        for (int c8 = 0; c8 < CPT / 8; c8++) {
          float v[8];
          tc::tmem_ld_x8(tmem + lane_base + (uint32_t)(col0 + c8 * 8), v);
          float* dst = s_p + r * PI + col0 + c8 * 8;
          *reinterpret_cast<float4*>(dst) = make_float4(v[0] * inv, v[1] * inv, v[2] * inv, v[3] * inv);
          *reinterpret_cast<float4*>(dst + 4) = make_float4(v[4] * inv, v[5] * inv, v[6] * inv, v[7] * inv);
        }
      }
      tc::tcgen05_fence_before();
      mbar_arrive(&bars[5 + b]);   // P tile b is full
    }
  } else {
    // =================================== splat group ======================================
    const int t2 = tid - kMain;
    for (uint32_t k = 0; k < my_tiles; k++) {
      const int64_t tile = tile0 + (int64_t)k * stride;
      const uint32_t b = k & 1, par = (k >> 1) & 1;
      const int64_t p0 = tile * kTile;
      const int rows = (int)min((int64_t)kTile, a.n - p0);
      const int valid = rows * kD1;
      const float* s_p = s_p0 + b * (kTile * PI);
      const unsigned char* s_rp = s_rp0 + b * 9 * kTE;
      const float* s_sw = reinterpret_cast<const float*>(s_rp);
      const int* s_ev = reinterpret_cast<const int*>(s_rp + 4 * kTE);
      const uint8_t* s_spix = s_rp + 8 * kTE;

      if (!tc::mbar_wait(&bars[5 + b], par)) atomicAdd(&g_ws_timeouts, 1);
      // ---- P -> HBM (coalesced rows) -------------------------------------------------------
      {
        constexpr int F4 = L / 4;
#pragma unroll
        for (int i = t2; i < kTile * F4; i += kSplat) {
          const int row = i / F4, c4 = i % F4;
          if (row < rows)
            *reinterpret_cast<float4*>(a.p_next + (p0 + row) * L + 4 * c4) =
                *reinterpret_cast<const float4*>(s_p + row * PI + 4 * c4);
        }
      }
      // ---- splat: 32 ranges of 24 vertex-sorted entries x 8 column groups --------------------
      if (!tc::mbar_wait(&bars[3 + b], par)) atomicAdd(&g_ws_timeouts, 1);
      {
        constexpr int NH = L / 32;
        constexpr int RLEN = kTE / kRanges;   // 24
        const int cg = t2 & 7, k0 = (t2 >> 3) * RLEN;
        if (k0 < valid) {
          int cur = s_ev[k0];
          float2 acc[2 * NH];
#pragma unroll
          for (int q = 0; q < 2 * NH; q++) acc[q] = make_float2(0.f, 0.f);
          float* out = a.val_out + 4 * cg;
#pragma unroll 2
          for (int kk = 0; kk < RLEN; kk += 4) {
            const int4 ev = *reinterpret_cast<const int4*>(s_ev + k0 + kk);
            const float4 ew = *reinterpret_cast<const float4*>(s_sw + k0 + kk);
            const uchar4 ep = *reinterpret_cast<const uchar4*>(s_spix + k0 + kk);
#define PL_WS_SPLAT(VV, WW, PP)                                                        \
  {                                                                                    \
    const float* src = s_p + (int)(PP) * PI + 4 * cg;                                  \
    float4 x[NH];                                                                      \
    _Pragma("unroll") for (int q = 0; q < NH; q++)                                     \
        x[q] = *reinterpret_cast<const float4*>(src + 32 * q);                         \
    if ((VV) != cur) {                                                                 \
      _Pragma("unroll") for (int q = 0; q < NH; q++) {                                 \
        red_add_v4(out + (int64_t)cur * L + 32 * q, acc[2 * q], acc[2 * q + 1]);       \
        acc[2 * q] = make_float2(0.f, 0.f);                                            \
        acc[2 * q + 1] = make_float2(0.f, 0.f);                                        \
      }                                                                                \
      cur = (VV);                                                                      \
    }                                                                                  \
    _Pragma("unroll") for (int q = 0; q < NH; q++) {                                   \
      ffma2(acc[2 * q], (WW), make_float2(x[q].x, x[q].y));                            \
      ffma2(acc[2 * q + 1], (WW), make_float2(x[q].z, x[q].w));                        \
    }                                                                                  \
  }
            PL_WS_SPLAT(ev.x, ew.x, ep.x)
            PL_WS_SPLAT(ev.y, ew.y, ep.y)
            PL_WS_SPLAT(ev.z, ew.z, ep.z)
            PL_WS_SPLAT(ev.w, ew.w, ep.w)
#undef PL_WS_SPLAT
          }
#pragma unroll
          for (int q = 0; q < NH; q++)
            red_add_v4(out + (int64_t)cur * L + 32 * q, acc[2 * q], acc[2 * q + 1]);
        }
      }
      mbar_arrive(&bars[7 + b]);   // P tile b is empty
      group_sync(2, kSplat);       // every splat thread is done with record b
      if (t2 == 0 && k + 2 < my_tiles) {
        tc::mbar_arrive_expect_tx(&bars[3 + b], 9 * kTE);
        tc::tma_load_1d(s_rp0 + b * 9 * kTE, a.rec_splat + (tile + 2 * stride) * (int64_t)(9 * kTE),
                        9 * kTE, &bars[3 + b]);
      }
    }
  }

  tc::tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, C::TMEM_COLS);
}

template <int L, bool FIRST>
int launch_ws(const WsArgs& a, cudaStream_t st) {
  using C = WCfg<L>;
  static PerDeviceOnce attr_set;
  if (attr_set.need()) {
    PL_CUDA(cudaFuncSetAttribute((mf_ws_kernel<L, FIRST>),
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
  }
  const unsigned grid = (unsigned)(a.ntiles < PL_SM_COUNT ? a.ntiles : PL_SM_COUNT);
  if (!FIRST) {   // behind a blur kernel of ours; the first iteration follows memsets
    PL_LAUNCH_PDL((mf_ws_kernel<L, FIRST>), grid, kThreads, C::SMEM, st, a);
    return PL_OK;
  }
  PL_LAUNCH((mf_ws_kernel<L, FIRST>), grid, kThreads, C::SMEM, st, a);
  return PL_OK;
}

}  // namespace

bool mf_ws_supported(const Lattice* L, int labels) {
  return L != nullptr && (labels == 32 || labels == 64) && L->d == 5 && L->ntiles > 0 &&
         L->rec_slice != nullptr && L->rec_splat != nullptr;
}

// One iteration (same contract as mf_tc_iteration; mu_canon from mf_tc_prepare_mu).
int mf_ws_iteration(Lattice* L, bool first, const float* val_in, float* val_out, const float* e0,
                    const float* p_prev, float* p_next, const float* mu_canon, float scale,
                    int labels, cudaStream_t st, void* fill, size_t fill_bytes) {
  WsArgs a = {};
  if (first && fill != nullptr && fill_bytes % 16 == 0 && fill_bytes / 16 < 0xffffffffull) {
    a.fill = reinterpret_cast<uint4*>(fill);
    a.fill_n16 = (uint32_t)(fill_bytes / 16);
  } else if (fill != nullptr && fill_bytes > 0) {
    PL_CUDA(cudaMemsetAsync(fill, 0xFF, fill_bytes, st));
  }
  a.val_in = val_in; a.val_out = val_out; a.e0 = e0; a.p_prev = p_prev; a.p_next = p_next;
  a.mu_canon = mu_canon; a.scale = scale; a.n = L->n; a.ntiles = L->ntiles;
  a.inv_denom = 1.0f / (1 + powf(2, -L->d));
  a.vid = L->vid; a.rec_slice = L->rec_slice; a.rec_splat = L->rec_splat;
  a.useg = L->useg; a.tile_off = L->tile_off; a.tile_U = L->tile_U;
  if (labels == 64) return first ? launch_ws<64, true>(a, st) : launch_ws<64, false>(a, st);
  if (labels == 32) return first ? launch_ws<32, true>(a, st) : launch_ws<32, false>(a, st);
  set_last_error("warp-specialised mean-field kernel: unsupported label count %d", labels);
  return PL_ERR_INVALID;
}

int ws_timeouts() {
  int v = 0;
  cudaMemcpyFromSymbol(&v, g_ws_timeouts, sizeof(int));
  return v;
}

}  // namespace pl
