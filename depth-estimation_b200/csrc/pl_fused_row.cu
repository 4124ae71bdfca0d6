// pl_fused_row.cu -- mean-field iteration, row-per-thread form (sm_100a).  L in {32, 64}, d = 5.
//
// Same arithmetic as pl_fused_tc.cu (crf/crf_module.py:100-102 per tile of 128 pixels):
//     E = E0 + slice(blurred lattice) - P_prev ; X = exp(-E - max) ; P = scale * (X Mu^T) / sum(X)
//     P -> HBM and splat(P) into the zeroed lattice buffer
// but laid out so that the instruction count per tile is about half:
//   * 512 threads = 128 rows x 4 column quarters; a thread owns L/4 consecutive labels of ONE
//     pixel for slice / softmax / operand split / epilogue (the TMEM lane <-> row mapping of
//     tcgen05.ld is the same one, so nothing is transposed).  The per-entry weight / index
//     loads are amortised over L/4 columns, the row max / sum need no shuffles (one
//     shared-memory exchange between the four quarter warps), the FMAs are packed
//     (fma.rn.f32x2 -> FFMA2).
//   * every global -> shared transfer is a bulk TMA copy signalled on an mbarrier, issued a
//     whole tile ahead: E0 / P_prev rows (one 4L-byte copy per row into a padded pitch, so
//     the row-per-thread LDS.128 are bank-conflict free), the tile's distinct lattice rows,
//     the packed slice / splat records.  No thread waits on a global load in steady state.
//   * the A operand never touches shared memory: the softmax threads write X hi / lo rows into
//     TMEM with tcgen05.st (lane = pixel row, column = label) and the MMAs take A from there.
//   * one CTA per SM (188 KB of shared memory); the MMA of tile k overlaps the slice of tile
//     k+1, whose inputs are already resident.
//   * normalisation by the row sum is deferred to the epilogue (P = acc * scale / sum), so
//     the operand split does not wait for the sum exchange.
#include <math.h>
#include <stdlib.h>

#include "pl_common.cuh"
#include "pl_tc.cuh"

namespace pl {

namespace {

constexpr int kTile = PL_TILE;   // 128 pixels = UMMA M
constexpr int kThreads = 512;
constexpr int kD1 = 6;
constexpr int kTE = kTile * kD1;  // entries per tile
constexpr int kStage = 128;       // distinct lattice rows staged per tile (more: rows from L2)
constexpr int kRanges = 64;       // splat: entry ranges per tile (12 entries each)

__device__ int g_row_timeouts = 0;

constexpr int rmax(int a, int b) { return a > b ? a : b; }

template <int L, bool ATM = false>
struct RCfg {
  static constexpr int CPT = L / 4;              // columns per thread
  static constexpr int KCH = L / 4;              // 16-byte K chunks per operand row
  static constexpr int A_CH = kTile * 16 + 16;   // bytes per K chunk of A (padded)
  static constexpr int B_CH = L * 16 + 16;
  static constexpr int A_BYTES = KCH * A_CH;
  static constexpr int B_BYTES = KCH * B_CH;
  static constexpr int PI = L + 4;               // padded row pitch in floats
  // A operand in TMEM (ATM): the region only stages the P tile
  static constexpr int Q_BYTES = ATM ? kTile * PI * 4 : rmax(2 * A_BYTES, kTile * PI * 4);
  static constexpr int IN_BYTES = 2 * kTile * PI * 4;
  static constexpr int ROWS_BYTES = kStage * PI * 4;
  static constexpr int OFF_Q = 0;
  static constexpr int OFF_MU = OFF_Q + Q_BYTES;
  static constexpr int OFF_IN = OFF_MU + 2 * B_BYTES;
  static constexpr int OFF_ROWS = OFF_IN + IN_BYTES;
  static constexpr int OFF_RS = OFF_ROWS + ROWS_BYTES;
  static constexpr int OFF_RP = OFF_RS + 6 * kTE;
  static constexpr int OFF_RED = OFF_RP + 9 * kTE;           // max[4][128] | sum[4][128]
  static constexpr int OFF_BAR = OFF_RED + 2 * 4 * kTile * 4;
  static constexpr int SMEM = OFF_BAR + 64 + 16;
  // ATM: D [0, L) | X hi [L, 2L) | X lo [2L, 3L), rounded up to a power of two
  static constexpr int TMEM_COLS = ATM ? (L <= 32 ? 128 : 256) : (L < 32 ? 32 : L);
};

struct RowArgs {
  const float* val_in;   // blurred lattice of the previous iteration [M+1][L]
  float* val_out;        // zeroed lattice buffer to splat into
  const float* e0;       // [n][L]
  const float* p_prev;   // [n][L]
  float* p_next;         // [n][L]
  const float* mu_canon; // Mu hi / lo planes in the canonical operand layout (mf_tc_prepare_mu)
  float scale;
  int64_t n;
  int64_t ntiles;
  float inv_denom;
  const int32_t* vid;        // [n][6]  (tiles with more than kStage distinct vertices)
  const uint8_t* rec_slice;  // per tile { bary f32[TE] | loc u16[TE] }
  const uint8_t* rec_splat;  // per tile { w f32[TE] | vertex i32[TE] | pixel u8[TE] }
  const int2* useg;
  const int* tile_off;
  const int* tile_U;
};

__device__ __forceinline__ void red_add_v4(float* addr, float2 a, float2 b) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a.x), "f"(a.y),
               "f"(b.x), "f"(b.y));
}

// acc += w * x on both halves: one FFMA2 (the scalar weight is broadcast by the instruction)
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

template <int L, bool FIRST, bool ATM>
__global__ void __launch_bounds__(kThreads, 1) mf_row_kernel(const RowArgs a) {
  using C = RCfg<L, ATM>;
  constexpr int CPT = C::CPT, PI = C::PI;
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* s_q = smem + C::OFF_Q;       // Q hi | lo planes -> P tile (pitch PI)
  unsigned char* s_mu = smem + C::OFF_MU;
  float* s_e0 = reinterpret_cast<float*>(smem + C::OFF_IN);   // [128][PI]
  float* s_pp = s_e0 + kTile * PI;                            // [128][PI]
  float* s_rows = reinterpret_cast<float*>(smem + C::OFF_ROWS);
  unsigned char* s_rs = smem + C::OFF_RS;
  unsigned char* s_rp = smem + C::OFF_RP;
  float* s_max = reinterpret_cast<float*>(smem + C::OFF_RED);  // [4][128]
  float* s_sum = s_max + 4 * kTile;                            // [4][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::OFF_BAR);
  // bars: [0] Mu landed  [1] MMAs done  [2] tile inputs landed  [3] splat record landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const float* s_bary = reinterpret_cast<const float*>(s_rs);
  const uint16_t* s_loc = reinterpret_cast<const uint16_t*>(s_rs + 4 * kTE);
  const float* s_sw = reinterpret_cast<const float*>(s_rp);
  const int* s_ev = reinterpret_cast<const int*>(s_rp + 4 * kTE);
  const uint8_t* s_spix = s_rp + 8 * kTE;
  float* s_p = reinterpret_cast<float*>(s_q);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int r = (warp & 3) * 32 + lane;   // row of the tile = TMEM lane
  const int q = warp >> 2;                // column quarter
  const int col0 = q * CPT;
  const int64_t stride = gridDim.x;
  const int64_t tile0 = blockIdx.x;
  if (tile0 >= a.ntiles) return;

  // ---- setup ----------------------------------------------------------------------------
  if (warp == 0) {
    if (lane == 0) {
#pragma unroll
      for (int b = 0; b < 4; b++) tc::mbar_init(&bars[b], 1);
      tc::fence_mbar_init();
      tc::mbar_arrive_expect_tx(&bars[0], 2 * C::B_BYTES);
      tc::tma_load_1d(s_mu, a.mu_canon, 2 * C::B_BYTES, &bars[0]);
      tc::mbar_arrive_expect_tx(&bars[3], 9 * kTE);
      tc::tma_load_1d(s_rp, a.rec_splat + tile0 * (int64_t)(9 * kTE), 9 * kTE, &bars[3]);
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tcgen05_fence_before();
  }

  // distinct-vertex list of a tile: threads 256..383 hold one vertex id each
  auto load_uvid = [&](int64_t t, int& U, int& uv) {
    U = 0; uv = 0;
    if (!FIRST && t < a.ntiles) {
      U = a.tile_U[t];
      const int u = tid - 256;
      if (u >= 0 && u < U && U <= kStage) uv = a.useg[a.tile_off[t] + u].x;
    }
  };
  // all inputs of tile t -> shared memory (bulk copies, completion on bars[2])
  auto issue_inputs = [&](int64_t t, int U, int uv) {
    const int nrows = (int)min((int64_t)kTile, a.n - t * kTile);
    const int Us = (!FIRST && U <= kStage) ? U : 0;
    if (tid == 0) {
      uint32_t bytes = (uint32_t)nrows * (L * 4) * (FIRST ? 1 : 2);
      if (!FIRST) bytes += (uint32_t)Us * (L * 4) + 6 * kTE;
      tc::mbar_arrive_expect_tx(&bars[2], bytes);
      if (!FIRST)
        tc::tma_load_1d(s_rs, a.rec_slice + t * (int64_t)(6 * kTE), 6 * kTE, &bars[2]);
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
  // E rows of tile t from the landed inputs (slice + E0 - P_prev); rows beyond n get E = 0
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
    if (U <= kStage) {
      const uint32_t l01 = lp[0], l23 = lp[1], l45 = lp[2];
      const uint32_t lo[6] = {l01 & 0xFFFF, l01 >> 16, l23 & 0xFFFF, l23 >> 16, l45 & 0xFFFF, l45 >> 16};
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
    } else {  // too many distinct vertices to stage: rows straight from L2
      const int32_t* vp = a.vid + (t * kTile + r) * kD1;
#pragma unroll
      for (int j = 0; j < kD1; j++) {
        const float4* rp = reinterpret_cast<const float4*>(a.val_in + (int64_t)vp[j] * L + col0);
#pragma unroll
        for (int c4 = 0; c4 < CPT / 4; c4++) {
          const float4 x = rp[c4];
          ffma2(acc[2 * c4], w[j], make_float2(x.x, x.y));
          ffma2(acc[2 * c4 + 1], w[j], make_float2(x.z, x.w));
        }
      }
    }
    const float4* ppp = reinterpret_cast<const float4*>(s_pp + r * PI + col0);
#pragma unroll
    for (int c4 = 0; c4 < CPT / 4; c4++) {
      const float4 e = e0p[c4], p = ppp[c4];
      // E = E0 + (filtered / (1 + 2^-d) - P)   (crf_module.py:101, gaussian_matrix.py:352)
      e_reg[4 * c4] = __fadd_rn(e.x, __fmaf_rn(acc[2 * c4].x, a.inv_denom, -p.x));
      e_reg[4 * c4 + 1] = __fadd_rn(e.y, __fmaf_rn(acc[2 * c4].y, a.inv_denom, -p.y));
      e_reg[4 * c4 + 2] = __fadd_rn(e.z, __fmaf_rn(acc[2 * c4 + 1].x, a.inv_denom, -p.z));
      e_reg[4 * c4 + 3] = __fadd_rn(e.w, __fmaf_rn(acc[2 * c4 + 1].y, a.inv_denom, -p.w));
    }
  };

  // ---- prologue: tile 0 is sliced with its loads exposed, tile 1 is put in flight ---------
  int U0, uv0;
  load_uvid(tile0, U0, uv0);
  __syncthreads();
  tc::tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot;
  issue_inputs(tile0, U0, uv0);
  int U1, uv1;
  load_uvid(tile0 + stride, U1, uv1);
  if (!tc::mbar_wait(&bars[2], 0)) atomicAdd(&g_row_timeouts, 1);
  slice_tile(tile0, U0);
  __syncthreads();
  if (tile0 + stride < a.ntiles) issue_inputs(tile0 + stride, U1, uv1);
  int U2, uv2;
  load_uvid(tile0 + 2 * stride, U2, uv2);

  constexpr float kLog2e = 1.4426950408889634f;
  for (uint32_t k = 0;; k++) {
    const int64_t tile = tile0 + (int64_t)k * stride;
    if (tile >= a.ntiles) break;
    const uint32_t par = k & 1;
    const int64_t p0 = tile * kTile;
    const int rows = (int)min((int64_t)kTile, a.n - p0);
    const int valid = rows * kD1;
    const int64_t next1 = tile + stride, next2 = tile + 2 * stride;

    // ---- softmax numerators X = exp(-E - max) -> hi / lo operand planes ------------------
    {
      float m = -e_reg[0];
#pragma unroll
      for (int c = 1; c < CPT; c++) m = fmaxf(m, -e_reg[c]);
      s_max[q * kTile + r] = m;
      __syncthreads();
      m = fmaxf(fmaxf(s_max[r], s_max[kTile + r]), fmaxf(s_max[2 * kTile + r], s_max[3 * kTile + r]));
      const float mb = -m * kLog2e;
      float sum = 0.f;
      uint32_t xh[CPT], xl[CPT];   // ATM only (dead code otherwise)
#pragma unroll
      for (int c4 = 0; c4 < CPT / 4; c4++) {
        float4 x, hi, lo;
        x.x = ex2_approx(__fmaf_rn(e_reg[4 * c4], -kLog2e, mb));
        x.y = ex2_approx(__fmaf_rn(e_reg[4 * c4 + 1], -kLog2e, mb));
        x.z = ex2_approx(__fmaf_rn(e_reg[4 * c4 + 2], -kLog2e, mb));
        x.w = ex2_approx(__fmaf_rn(e_reg[4 * c4 + 3], -kLog2e, mb));
        sum += (x.x + x.y) + (x.z + x.w);
        hi.x = tc::tf32_rna(x.x); lo.x = tc::tf32_rna(x.x - hi.x);
        hi.y = tc::tf32_rna(x.y); lo.y = tc::tf32_rna(x.y - hi.y);
        hi.z = tc::tf32_rna(x.z); lo.z = tc::tf32_rna(x.z - hi.z);
        hi.w = tc::tf32_rna(x.w); lo.w = tc::tf32_rna(x.w - hi.w);
        if (ATM) {
          xh[4 * c4] = __float_as_uint(hi.x); xh[4 * c4 + 1] = __float_as_uint(hi.y);
          xh[4 * c4 + 2] = __float_as_uint(hi.z); xh[4 * c4 + 3] = __float_as_uint(hi.w);
          xl[4 * c4] = __float_as_uint(lo.x); xl[4 * c4 + 1] = __float_as_uint(lo.y);
          xl[4 * c4 + 2] = __float_as_uint(lo.z); xl[4 * c4 + 3] = __float_as_uint(lo.w);
        } else {
          const int ch = q * (CPT / 4) + c4;   // K chunk of the canonical A layout
          *reinterpret_cast<float4*>(s_q + ch * C::A_CH + r * 16) = hi;
          *reinterpret_cast<float4*>(s_q + C::A_BYTES + ch * C::A_CH + r * 16) = lo;
        }
      }
      if (ATM) {
        // X hi / lo rows straight into TMEM (lane = row, column = label): the A operand of the
        // MMAs, never staged in shared memory
        const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
        if (CPT == 16) {
          tc::tmem_st_x16(tmem + lane_base + (uint32_t)(L + col0), xh);
          tc::tmem_st_x16(tmem + lane_base + (uint32_t)(2 * L + col0), xl);
        } else {
          tc::tmem_st_x8(tmem + lane_base + (uint32_t)(L + col0), xh);
          tc::tmem_st_x8(tmem + lane_base + (uint32_t)(2 * L + col0), xl);
        }
        tc::tmem_st_wait();
      }
      s_sum[q * kTile + r] = sum;
    }
    if (ATM) tc::tcgen05_fence_before();
    else tc::fence_proxy_async();
    __syncthreads();

    // ---- acc = X Mu^T on the tensor cores (3 x TF32 split): one thread issues ------------
    if (tid == 0) {
      if (k == 0 && !tc::mbar_wait(&bars[0], 0)) atomicAdd(&g_row_timeouts, 1);
      tc::tcgen05_fence_after();
      constexpr uint32_t idesc = tc::idesc_tf32(kTile, L);
      const uint32_t a_hi = tc::smem_u32(s_q), a_lo = a_hi + C::A_BYTES;
      const uint32_t b_hi = tc::smem_u32(s_mu), b_lo = b_hi + C::B_BYTES;
#pragma unroll
      for (int pass = 0; pass < 3; pass++) {  // lo*hi, hi*lo, hi*hi
        const uint32_t ab = (pass == 0) ? a_lo : a_hi;
        const uint32_t bb = (pass == 1) ? b_lo : b_hi;
#pragma unroll
        for (int ks = 0; ks < L / 8; ks++) {
          const uint64_t bd = tc::smem_desc_kmajor(bb + ks * 2 * C::B_CH, C::B_CH, 128);
          if (ATM) {   // A: 8 TMEM columns of the hi (or lo) block
            tc::mma_tf32_ts(tmem, tmem + (uint32_t)((pass == 0 ? 2 * L : L) + ks * 8), bd, idesc,
                            (pass | ks) != 0);
          } else {
            const uint64_t ad = tc::smem_desc_kmajor(ab + ks * 2 * C::A_CH, C::A_CH, 128);
            tc::mma_tf32_ss(tmem, ad, bd, idesc, (pass | ks) != 0);
          }
        }
      }
      tc::mma_commit(&bars[1]);
    }

    // ---- while the MMAs run: E of the next tile (its inputs landed during this tile) ------
    if (next1 < a.ntiles) {
      if (!tc::mbar_wait(&bars[2], (k + 1) & 1)) atomicAdd(&g_row_timeouts, 1);
      slice_tile(next1, U1);
    }
    __syncthreads();  // input buffers are dead: refill them for the tile after
    if (next2 < a.ntiles) issue_inputs(next2, U2, uv2);
    U1 = U2;
    load_uvid(tile + 3 * stride, U2, uv2);

    // ---- epilogue: TMEM -> P = acc * scale / sum -> P tile in shared memory ---------------
    if (!tc::mbar_wait(&bars[1], par)) atomicAdd(&g_row_timeouts, 1);
    tc::tcgen05_fence_after();
    {
      const float ssum = (s_sum[r] + s_sum[kTile + r]) + (s_sum[2 * kTile + r] + s_sum[3 * kTile + r]);
      const float inv = __fdividef(a.scale, ssum);
#pragma unroll
      for (int c8 = 0; c8 < CPT / 8; c8++) {
        float v[8];
        tc::tmem_ld_x8(tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(col0 + c8 * 8), v);
        float* dst = s_p + r * PI + col0 + c8 * 8;
        *reinterpret_cast<float4*>(dst) = make_float4(v[0] * inv, v[1] * inv, v[2] * inv, v[3] * inv);
        *reinterpret_cast<float4*>(dst + 4) = make_float4(v[4] * inv, v[5] * inv, v[6] * inv, v[7] * inv);
      }
    }
    tc::tcgen05_fence_before();
    __syncthreads();

    // ---- P -> HBM (coalesced rows) --------------------------------------------------------
    {
      constexpr int F4 = L / 4;
#pragma unroll
      for (int i = tid; i < kTile * F4; i += kThreads) {
        const int row = i / F4, c4 = i % F4;
        if (row < rows)
          *reinterpret_cast<float4*>(a.p_next + (p0 + row) * L + 4 * c4) =
              *reinterpret_cast<const float4*>(s_p + row * PI + 4 * c4);
      }
    }

    // ---- splat P: 64 ranges of 12 vertex-sorted entries x 8 column groups; running sums,
    //      one red.global.add.v4 per (vertex change, 4 columns) -----------------------------
    {
      if (!tc::mbar_wait(&bars[3], par)) atomicAdd(&g_row_timeouts, 1);
      constexpr int NH = L / 32;
      constexpr int RLEN = kTE / kRanges;   // 12
      const int cg = tid & 7, k0 = (tid >> 3) * RLEN;
      if (k0 < valid) {
        int cur = s_ev[k0];
        float2 acc[2 * NH];
#pragma unroll
        for (int h = 0; h < 2 * NH; h++) acc[h] = make_float2(0.f, 0.f);
        float* out = a.val_out + 4 * cg;
#pragma unroll
        for (int kk = 0; kk < RLEN; kk += 4) {
          const int4 ev = *reinterpret_cast<const int4*>(s_ev + k0 + kk);
          const float4 ew = *reinterpret_cast<const float4*>(s_sw + k0 + kk);
          const uchar4 ep = *reinterpret_cast<const uchar4*>(s_spix + k0 + kk);
#define PL_ROW_SPLAT(VV, WW, PP)                                                       \
  {                                                                                    \
    const float* src = s_p + (int)(PP) * PI + 4 * cg;                                  \
    float4 x[NH];                                                                      \
    _Pragma("unroll") for (int h = 0; h < NH; h++)                                     \
        x[h] = *reinterpret_cast<const float4*>(src + 32 * h);                         \
    if ((VV) != cur) {                                                                 \
      _Pragma("unroll") for (int h = 0; h < NH; h++) {                                 \
        red_add_v4(out + (int64_t)cur * L + 32 * h, acc[2 * h], acc[2 * h + 1]);       \
        acc[2 * h] = make_float2(0.f, 0.f);                                            \
        acc[2 * h + 1] = make_float2(0.f, 0.f);                                        \
      }                                                                                \
      cur = (VV);                                                                      \
    }                                                                                  \
    _Pragma("unroll") for (int h = 0; h < NH; h++) {                                   \
      ffma2(acc[2 * h], (WW), make_float2(x[h].x, x[h].y));                            \
      ffma2(acc[2 * h + 1], (WW), make_float2(x[h].z, x[h].w));                        \
    }                                                                                  \
  }
          PL_ROW_SPLAT(ev.x, ew.x, ep.x)
          PL_ROW_SPLAT(ev.y, ew.y, ep.y)
          PL_ROW_SPLAT(ev.z, ew.z, ep.z)
          PL_ROW_SPLAT(ev.w, ew.w, ep.w)
#undef PL_ROW_SPLAT
        }
#pragma unroll
        for (int h = 0; h < NH; h++)
          red_add_v4(out + (int64_t)cur * L + 32 * h, acc[2 * h], acc[2 * h + 1]);
      }
    }
    __syncthreads();  // P tile and splat record are dead
    if (tid == 0 && next1 < a.ntiles) {
      tc::mbar_arrive_expect_tx(&bars[3], 9 * kTE);
      tc::tma_load_1d(s_rp, a.rec_splat + next1 * (int64_t)(9 * kTE), 9 * kTE, &bars[3]);
    }
  }

  tc::tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, C::TMEM_COLS);
}

template <int L, bool FIRST, bool ATM>
int launch_row_impl(const RowArgs& a, cudaStream_t st) {
  using C = RCfg<L, ATM>;
  static PerDeviceOnce attr_set;
  if (attr_set.need()) {
    PL_CUDA(cudaFuncSetAttribute((mf_row_kernel<L, FIRST, ATM>),
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
  }
  const unsigned grid = (unsigned)(a.ntiles < PL_SM_COUNT ? a.ntiles : PL_SM_COUNT);
  PL_LAUNCH((mf_row_kernel<L, FIRST, ATM>), grid, kThreads, C::SMEM, st, a);
  return PL_OK;
}

template <int L, bool FIRST>
int launch_row(const RowArgs& a, cudaStream_t st) {
  // The A operand (X hi / lo) goes registers -> TMEM (tcgen05.st) and the MMAs read it there;
  // PL_MF_ROW_SMEM_A=1 (A/B knob) stages it in shared memory operand planes instead.
  static const bool smem_a = getenv("PL_MF_ROW_SMEM_A") != nullptr;
  return smem_a ? launch_row_impl<L, FIRST, false>(a, st) : launch_row_impl<L, FIRST, true>(a, st);
}

}  // namespace

bool mf_row_supported(const Lattice* L, int labels) {
  return L != nullptr && (labels == 32 || labels == 64) && L->d == 5 && L->ntiles > 0 &&
         L->rec_slice != nullptr && L->rec_splat != nullptr;
}

// One iteration (same contract as mf_tc_iteration; mu_canon from mf_tc_prepare_mu).
int mf_row_iteration(Lattice* L, bool first, const float* val_in, float* val_out, const float* e0,
                     const float* p_prev, float* p_next, const float* mu_canon, float scale,
                     int labels, cudaStream_t st) {
  RowArgs a = {};
  a.val_in = val_in; a.val_out = val_out; a.e0 = e0; a.p_prev = p_prev; a.p_next = p_next;
  a.mu_canon = mu_canon; a.scale = scale; a.n = L->n; a.ntiles = L->ntiles;
  a.inv_denom = 1.0f / (1 + powf(2, -L->d));
  a.vid = L->vid; a.rec_slice = L->rec_slice; a.rec_splat = L->rec_splat;
  a.useg = L->useg; a.tile_off = L->tile_off; a.tile_U = L->tile_U;
  if (labels == 64) return first ? launch_row<64, true>(a, st) : launch_row<64, false>(a, st);
  if (labels == 32) return first ? launch_row<32, true>(a, st) : launch_row<32, false>(a, st);
  set_last_error("row-form mean-field kernel: unsupported label count %d", labels);
  return PL_ERR_INVALID;
}

int row_timeouts() {
  int v = 0;
  cudaMemcpyFromSymbol(&v, g_row_timeouts, sizeof(int));
  return v;
}

}  // namespace pl
