// pl_fused_tc.cu -- one kernel per mean-field iteration with the label-compatibility
// transform on the 5th-generation tensor cores (tcgen05, sm_100a).  L in {16, 32, 64}.
//
// crf/crf_module.py:100-102, pixel-major, per tile of PL_TILE = 128 pixels (= UMMA M):
//     E  = E0 + slice(blurred lattice) - P_prev          (MODE 0 / 2: E = E0)
//     Q  = softmax(-E)                                    registers, V-lane groups
//     P  = scale * Q Mu^T                                 tcgen05.mma kind::tf32, 3xTF32 split
//     P -> HBM and splat(P) into the zeroed lattice buffer (MODE 2: P -> HBM only)
//
// The contraction is 128 x L x L per tile.  fp32 parity (1e-4 of the output scale after five
// iterations) rules out a single TF32 pass, so both operands are split x = hi + lo with
// hi = tf32(x), lo = tf32(x - hi) and three MMAs accumulate lo*hi + hi*lo + hi*hi in the
// fp32 TMEM accumulator (the dropped lo*lo term is 2^-22 relative).  Operand layout in
// shared memory is the K-major no-swizzle canonical form: 16-byte K chunks, each chunk
// holding all rows contiguously (8-row x 16-byte core matrices of 128 B), one 16-byte pad
// per chunk so the softmax threads' float4 stores are bank-conflict free.
//   A = Q hi/lo   [128 rows][L]   written by the softmax phase
//   B = Mu hi/lo  [L rows][L]     pre-split once per call (mu_split_kernel), fetched per
//                                 tile with one bulk TMA copy (cp.async.bulk -> UBLKCP)
//   D = P         TMEM, 128 lanes x L columns fp32, read back with tcgen05.ld
// Shared memory per CTA at L = 64: Q planes 64.5 KB (P is staged over them once the MMAs
// have completed), lattice rows / Mu planes 36 KB (Mu overwrites the staged rows after the
// slice phase), entry lists 9.8 KB -> 110 KB, two CTAs per SM.
#include <math.h>

#include "pl_common.cuh"
#include "pl_tc.cuh"

namespace pl {

namespace {

constexpr int kTile = PL_TILE;   // 128 pixels = UMMA M
constexpr int kThreads = 256;
constexpr int kCtasPerSm = 2;

__device__ int g_tc_timeouts = 0;

constexpr int cmax(int a, int b) { return a > b ? a : b; }

template <int L>
struct Cfg {
  static constexpr int V = L / 4;               // float4 lanes per pixel row (4 / 8 / 16)
  static constexpr int RPP = kThreads / V;      // rows per pass = lane groups per CTA
  static constexpr int RPT = kTile / RPP;       // rows per thread (2 / 4 / 8)
  static constexpr int KCH = L / 4;             // 16-byte K chunks per operand row
  static constexpr int A_CH = kTile * 16 + 16;  // bytes per K chunk of A (padded)
  static constexpr int B_CH = L * 16 + 16;      // bytes per K chunk of B (padded)
  static constexpr int A_BYTES = KCH * A_CH;    // one plane
  static constexpr int B_BYTES = KCH * B_CH;
  static constexpr int PLD = L + 4;             // P row pitch in floats
  static constexpr int QA_BYTES = cmax(2 * A_BYTES, kTile * PLD * 4);
  // lattice rows staged per tile: what fits under the Q planes, at most one per thread
  static constexpr int STAGE_ROWS = QA_BYTES / (L * 4) < kThreads ? QA_BYTES / (L * 4) : kThreads;
  static constexpr int TMEM_COLS = L < 32 ? 32 : L;
};

struct MfArgs {
  const float* val_in;   // blurred lattice of the previous iteration [M+1][L]   (MODE 1)
  float* val_out;        // zeroed lattice buffer to splat into                   (MODE 0, 1)
  const float* e0;       // [n][L]
  const float* p_prev;   // [n][L]                                                (MODE 1)
  float* p_next;         // [n][L]
  float* q_out;          // [n][L] or null                                        (MODE 2)
  const float* mu_canon; // Mu hi / lo planes in the canonical operand layout
  float scale;
  int64_t n;
  int64_t ntiles;
  int d1;
  float inv_denom;
  const int32_t* vid;        // [n][d1]            (fallback when a tile has too many vertices)
  const uint8_t* rec_slice;  // per tile { bary f32[TE] | loc u16[TE] }
  const uint8_t* rec_splat;  // per tile { w f32[TE] | vertex i32[TE] | pixel u8[TE] }
  const int2* useg;
  const int* tile_off;
  const int* tile_U;
};

__device__ __forceinline__ void red_add_v4(float* addr, float4 x) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(x.x), "f"(x.y),
               "f"(x.z), "f"(x.w));
}

template <int W>
__device__ __forceinline__ float group_max(float v) {
#pragma unroll
  for (int o = W / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o, W));
  return v;
}
template <int W>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int o = W / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o, W);
  return v;
}

// MODE 0: first iteration (E = E0, splat P)   MODE 1: later iteration (slice, splat P)
// MODE 2: stand-alone softmax + compatibility (no lattice)
// D1T: d + 1 known at compile time (6 for the 5-D stereo guide) or 0 = run-time.
//
// Persistent: each CTA walks tiles blockIdx.x, +gridDim.x, ...  Per tile
//   [lists land by TMA] -> stage lattice rows -> slice -> softmax -> Q planes -> MMAs ->
//   TMEM -> P tile -> HBM + splat
// while the NEXT tile's E0 / P_prev rows (registers), slice lists (TMA) and distinct-vertex
// ids (registers) are already in flight; Mu is fetched once per CTA.
template <int L, int MODE, int D1T>
__global__ void __launch_bounds__(kThreads, kCtasPerSm) mf_tc_kernel(const MfArgs a) {
  using C = Cfg<L>;
  constexpr int V = C::V, RPP = C::RPP, RPT = C::RPT;
  extern __shared__ __align__(128) unsigned char smem[];
  const int d1 = D1T > 0 ? D1T : a.d1;
  const int TE = (MODE == 2) ? 0 : kTile * d1;
  unsigned char* s_qa = smem;                       // lattice rows -> Q hi | lo planes -> P tile
  unsigned char* s_mu = smem + C::QA_BYTES;         // Mu hi | lo planes (whole kernel)
  unsigned char* s_rs = s_mu + 2 * C::B_BYTES;      // slice record  { bary | loc }
  unsigned char* s_rp = s_rs + 6 * TE;              // splat record  { w | vertex | pixel }
  int* s_uvid = reinterpret_cast<int*>(s_rp + 9 * TE);          // [256] distinct vertex ids
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_uvid + (MODE == 2 ? 0 : kThreads));
  // bars: [0] Mu landed  [1] MMAs done  [2] slice record landed  [3] splat record landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const float* s_bary = reinterpret_cast<const float*>(s_rs);
  const uint16_t* s_loc = reinterpret_cast<const uint16_t*>(s_rs + 4 * TE);
  const float* s_sw = reinterpret_cast<const float*>(s_rp);
  const int* s_ev = reinterpret_cast<const int*>(s_rp + 4 * TE);
  const uint8_t* s_spix = s_rp + 8 * TE;
  const float* s_val = reinterpret_cast<const float*>(s_qa);
  float* s_p = reinterpret_cast<float*>(s_qa);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = tid / V, l = tid % V;
  int64_t tile = blockIdx.x;
  if (tile >= a.ntiles) return;  // (grid never exceeds ntiles; keeps the setup below uniform)

  // ---- one-time setup: barriers, TMEM, Mu and the first tile's lists ---------------------
  if (warp == 0) {
    if (lane == 0) {
#pragma unroll
      for (int b = 0; b < 4; b++) tc::mbar_init(&bars[b], 1);
      tc::fence_mbar_init();
      tc::mbar_arrive_expect_tx(&bars[0], 2 * C::B_BYTES);
      tc::tma_load_1d(s_mu, a.mu_canon, 2 * C::B_BYTES, &bars[0]);
      if (MODE == 1) {
        tc::mbar_arrive_expect_tx(&bars[2], 6 * TE);
        tc::tma_load_1d(s_rs, a.rec_slice + tile * (int64_t)(6 * TE), 6 * TE, &bars[2]);
      }
      if (MODE != 2) {
        tc::mbar_arrive_expect_tx(&bars[3], 9 * TE);
        tc::tma_load_1d(s_rp, a.rec_splat + tile * (int64_t)(9 * TE), 9 * TE, &bars[3]);
      }
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tcgen05_fence_before();
  }

  // first tile's E0 / P_prev rows and distinct-vertex id -> registers
  float4 e_reg[RPT], pp_reg[RPT];
  int U = 0, my_uvid = 0;
  auto prefetch = [&](int64_t t) {
    const int64_t q0 = t * kTile;
    const int nrows = (int)min((int64_t)kTile, a.n - q0);
#pragma unroll
    for (int i = 0; i < RPT; i++) {
      const int r = g + i * RPP;
      e_reg[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      pp_reg[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r < nrows) {
        e_reg[i] = *reinterpret_cast<const float4*>(a.e0 + (q0 + r) * L + 4 * l);
        if (MODE == 1)
          pp_reg[i] = *reinterpret_cast<const float4*>(a.p_prev + (q0 + r) * L + 4 * l);
      }
    }
    if (MODE == 1) {
      U = a.tile_U[t];
      my_uvid = (tid < U) ? a.useg[a.tile_off[t] + tid].x : 0;
    }
  };
  prefetch(tile);
  __syncthreads();
  tc::tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot;

  for (uint32_t it = 0; tile < a.ntiles; tile += gridDim.x, it++) {
    const uint32_t par = it & 1;
    const int64_t p0 = tile * kTile;
    const int rows = (int)min((int64_t)kTile, a.n - p0);
    const int valid = rows * d1;
    const int64_t next = tile + gridDim.x;
    const bool has_next = next < a.ntiles;

    // ---- slice: E = E0 + (slice(blurred lattice) - P_prev) --------------------------------
    if (MODE == 1) {
      const bool staged = U <= C::STAGE_ROWS;
      if (staged) {
        s_uvid[tid] = my_uvid;
        __syncthreads();
        // rows of the tile's distinct vertices: V lanes per row, coalesced 16-byte pieces
        float* sv = reinterpret_cast<float*>(s_qa);
        for (int u = g; u < U; u += RPP)
          *reinterpret_cast<float4*>(sv + u * L + 4 * l) =
              *reinterpret_cast<const float4*>(a.val_in + (int64_t)s_uvid[u] * L + 4 * l);
      }
      if (!tc::mbar_wait(&bars[2], par)) atomicAdd(&g_tc_timeouts, 1);
      __syncthreads();
#define PL_SLICE_FMA(WW, XX)            \
  acc.x = __fmaf_rn((WW), (XX).x, acc.x); \
  acc.y = __fmaf_rn((WW), (XX).y, acc.y); \
  acc.z = __fmaf_rn((WW), (XX).z, acc.z); \
  acc.w = __fmaf_rn((WW), (XX).w, acc.w);
#pragma unroll
      for (int i = 0; i < RPT; i++) {
        const int r = g + i * RPP;
        if (r < rows) {
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
          if (staged) {
            if (D1T == 6) {
              // six (weight, row) pairs of this pixel: 3 x LDS.64 + 3 x LDS.32 (broadcast)
              const float2* wp = reinterpret_cast<const float2*>(s_bary + r * 6);
              const uint32_t* lp = reinterpret_cast<const uint32_t*>(s_loc + r * 6);
              const float2 w01 = wp[0], w23 = wp[1], w45 = wp[2];
              const uint32_t l01 = lp[0], l23 = lp[1], l45 = lp[2];
              const float4 x0 = *reinterpret_cast<const float4*>(s_val + (l01 & 0xFFFF) * L + 4 * l);
              const float4 x1 = *reinterpret_cast<const float4*>(s_val + (l01 >> 16) * L + 4 * l);
              const float4 x2 = *reinterpret_cast<const float4*>(s_val + (l23 & 0xFFFF) * L + 4 * l);
              const float4 x3 = *reinterpret_cast<const float4*>(s_val + (l23 >> 16) * L + 4 * l);
              const float4 x4 = *reinterpret_cast<const float4*>(s_val + (l45 & 0xFFFF) * L + 4 * l);
              const float4 x5 = *reinterpret_cast<const float4*>(s_val + (l45 >> 16) * L + 4 * l);
              PL_SLICE_FMA(w01.x, x0) PL_SLICE_FMA(w01.y, x1) PL_SLICE_FMA(w23.x, x2)
              PL_SLICE_FMA(w23.y, x3) PL_SLICE_FMA(w45.x, x4) PL_SLICE_FMA(w45.y, x5)
            } else {
              for (int j = 0; j < d1; j++) {
                const float w = s_bary[r * d1 + j];
                const float4 x =
                    *reinterpret_cast<const float4*>(s_val + (int)s_loc[r * d1 + j] * L + 4 * l);
                PL_SLICE_FMA(w, x)
              }
            }
          } else {  // too many distinct vertices to stage: rows straight from L2
            for (int j = 0; j < d1; j++) {
              const float w = s_bary[r * d1 + j];
              const float4 x = *reinterpret_cast<const float4*>(
                  a.val_in + (int64_t)a.vid[(p0 + r) * d1 + j] * L + 4 * l);
              PL_SLICE_FMA(w, x)
            }
          }
#undef PL_SLICE_FMA
          // E = E0 + (filtered / (1 + 2^-d) - P)   (crf_module.py:101, gaussian_matrix.py:352)
          e_reg[i].x = __fadd_rn(e_reg[i].x, __fmaf_rn(acc.x, a.inv_denom, -pp_reg[i].x));
          e_reg[i].y = __fadd_rn(e_reg[i].y, __fmaf_rn(acc.y, a.inv_denom, -pp_reg[i].y));
          e_reg[i].z = __fadd_rn(e_reg[i].z, __fmaf_rn(acc.z, a.inv_denom, -pp_reg[i].z));
          e_reg[i].w = __fadd_rn(e_reg[i].w, __fmaf_rn(acc.w, a.inv_denom, -pp_reg[i].w));
        }
      }
      __syncthreads();  // staged rows and the slice record are dead
      if (tid == 0 && has_next) {
        tc::mbar_arrive_expect_tx(&bars[2], 6 * TE);
        tc::tma_load_1d(s_rs, a.rec_slice + next * (int64_t)(6 * TE), 6 * TE, &bars[2]);
      }
    }

    // ---- softmax(-E) -> Q hi / lo operand planes ------------------------------------------
    {
      float m[RPT];
#pragma unroll
      for (int i = 0; i < RPT; i++) {
        const float4 e = e_reg[i];
        m[i] = (g + i * RPP < rows) ? fmaxf(fmaxf(-e.x, -e.y), fmaxf(-e.z, -e.w)) : -INFINITY;
      }
#pragma unroll
      for (int o = V / 2; o > 0; o >>= 1)
#pragma unroll
        for (int i = 0; i < RPT; i++) m[i] = fmaxf(m[i], __shfl_xor_sync(0xffffffffu, m[i], o, V));
      float s[RPT];
#pragma unroll
      for (int i = 0; i < RPT; i++) {
        float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
        if (g + i * RPP < rows) {
          x.x = __expf(-e_reg[i].x - m[i]);
          x.y = __expf(-e_reg[i].y - m[i]);
          x.z = __expf(-e_reg[i].z - m[i]);
          x.w = __expf(-e_reg[i].w - m[i]);
        }
        e_reg[i] = x;
        s[i] = x.x + x.y + x.z + x.w;
      }
#pragma unroll
      for (int o = V / 2; o > 0; o >>= 1)
#pragma unroll
        for (int i = 0; i < RPT; i++) s[i] += __shfl_xor_sync(0xffffffffu, s[i], o, V);
#pragma unroll
      for (int i = 0; i < RPT; i++) {
        const int r = g + i * RPP;
        const float inv = (r < rows) ? __fdividef(1.0f, s[i]) : 0.f;
        float4 x = e_reg[i];
        x.x *= inv; x.y *= inv; x.z *= inv; x.w *= inv;
        if (MODE == 2 && a.q_out != nullptr && r < rows)
          *reinterpret_cast<float4*>(a.q_out + (p0 + r) * L + 4 * l) = x;
        float4 hi, lo;
        hi.x = tc::tf32_rna(x.x); lo.x = tc::tf32_rna(x.x - hi.x);
        hi.y = tc::tf32_rna(x.y); lo.y = tc::tf32_rna(x.y - hi.y);
        hi.z = tc::tf32_rna(x.z); lo.z = tc::tf32_rna(x.z - hi.z);
        hi.w = tc::tf32_rna(x.w); lo.w = tc::tf32_rna(x.w - hi.w);
        // K chunk l, row r of the canonical A layout
        *reinterpret_cast<float4*>(s_qa + l * C::A_CH + r * 16) = hi;
        *reinterpret_cast<float4*>(s_qa + C::A_BYTES + l * C::A_CH + r * 16) = lo;
      }
    }
    tc::fence_proxy_async();  // Q planes (generic proxy) -> visible to the tensor core's reads
    __syncthreads();

    // ---- P = Q Mu^T on the tensor cores: one thread issues --------------------------------
    if (tid == 0) {
      if (it == 0 && !tc::mbar_wait(&bars[0], 0)) atomicAdd(&g_tc_timeouts, 1);
      tc::tcgen05_fence_after();
      constexpr uint32_t idesc = tc::idesc_tf32(kTile, L);
      const uint32_t a_hi = tc::smem_u32(s_qa), a_lo = a_hi + C::A_BYTES;
      const uint32_t b_hi = tc::smem_u32(s_mu), b_lo = b_hi + C::B_BYTES;
#pragma unroll
      for (int pass = 0; pass < 3; pass++) {  // lo*hi, hi*lo, hi*hi
        const uint32_t ab = (pass == 0) ? a_lo : a_hi;
        const uint32_t bb = (pass == 1) ? b_lo : b_hi;
#pragma unroll
        for (int ks = 0; ks < L / 8; ks++) {  // K = 8 per instruction = two 16-byte chunks
          const uint64_t ad = tc::smem_desc_kmajor(ab + ks * 2 * C::A_CH, C::A_CH, 128);
          const uint64_t bd = tc::smem_desc_kmajor(bb + ks * 2 * C::B_CH, C::B_CH, 128);
          tc::mma_tf32_ss(tmem, ad, bd, idesc, (pass | ks) != 0);
        }
      }
      tc::mma_commit(&bars[1]);
    }
    // the next tile's inputs start their trip while the MMAs, the epilogue and the splat run
    if (has_next) prefetch(next);
    __syncwarp();
    if (!tc::mbar_wait(&bars[1], par)) atomicAdd(&g_tc_timeouts, 1);
    __syncwarp();
    tc::tcgen05_fence_after();

    // ---- epilogue: TMEM -> registers -> P tile in shared memory (over the dead Q planes) --
    {
      // TMEM lane quarter q (a warp can only read lanes 32*(warp%4)..+31) x column group
      constexpr int NCG = (kThreads / 128) < (L / 8) ? (kThreads / 128) : (L / 8);
      constexpr int CPW = L / NCG;  // columns per warp, multiple of 8
      const int q = warp & 3, hsel = warp >> 2;
      const int row = q * 32 + lane;
      if (hsel < NCG) {
#pragma unroll
        for (int c8 = 0; c8 < CPW / 8; c8++) {
          const int col = hsel * CPW + c8 * 8;
          float v[8];
          tc::tmem_ld_x8(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)col, v);
          float4 o0 = make_float4(v[0] * a.scale, v[1] * a.scale, v[2] * a.scale, v[3] * a.scale);
          float4 o1 = make_float4(v[4] * a.scale, v[5] * a.scale, v[6] * a.scale, v[7] * a.scale);
          *reinterpret_cast<float4*>(s_p + row * C::PLD + col) = o0;
          *reinterpret_cast<float4*>(s_p + row * C::PLD + col + 4) = o1;
        }
      }
    }
    tc::tcgen05_fence_before();
    __syncthreads();

    // ---- P -> HBM (coalesced rows) --------------------------------------------------------
#pragma unroll
    for (int i = 0; i < RPT; i++) {
      const int r = g + i * RPP;
      if (r < rows)
        *reinterpret_cast<float4*>(a.p_next + (p0 + r) * L + 4 * l) =
            *reinterpret_cast<const float4*>(s_p + r * C::PLD + 4 * l);
    }

    // ---- splat P: every lane group walks a contiguous range of the vertex-sorted entries,
    //      keeps a running sum and flushes one red.global.add.v4 per (vertex, range) ----------
    if (MODE != 2) {
      if (!tc::mbar_wait(&bars[3], par)) atomicAdd(&g_tc_timeouts, 1);
      const int cpg = (((valid + RPP - 1) / RPP) + 3) & ~3;  // entries per group, multiple of 4
      // the record's pads (entries >= valid) carry weight 0 and vertex 0: no bounds check inside
      const int k0 = g * cpg, k1 = min((valid + 3) & ~3, k0 + cpg);
      if (k0 < k1) {
        int cur = s_ev[k0];
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int k = k0; k < k1; k += 4) {  // pads beyond `valid` carry weight 0, vertex 0
          const int4 ev = *reinterpret_cast<const int4*>(s_ev + k);
          const float4 ew = *reinterpret_cast<const float4*>(s_sw + k);
          const uchar4 ep = *reinterpret_cast<const uchar4*>(s_spix + k);
          const float4 x0 = *reinterpret_cast<const float4*>(s_p + (int)ep.x * C::PLD + 4 * l);
          const float4 x1 = *reinterpret_cast<const float4*>(s_p + (int)ep.y * C::PLD + 4 * l);
          const float4 x2 = *reinterpret_cast<const float4*>(s_p + (int)ep.z * C::PLD + 4 * l);
          const float4 x3 = *reinterpret_cast<const float4*>(s_p + (int)ep.w * C::PLD + 4 * l);
#define PL_SPLAT_STEP(VV, WW, XX, KK)                                         \
  {                                                                           \
    if ((VV) != cur) {                                                        \
      red_add_v4(a.val_out + (int64_t)cur * L + 4 * l, acc);                  \
      acc = make_float4(0.f, 0.f, 0.f, 0.f);                                  \
      cur = (VV);                                                             \
    }                                                                         \
    acc.x = __fmaf_rn((WW), (XX).x, acc.x);                                   \
    acc.y = __fmaf_rn((WW), (XX).y, acc.y);                                   \
    acc.z = __fmaf_rn((WW), (XX).z, acc.z);                                   \
    acc.w = __fmaf_rn((WW), (XX).w, acc.w);                                   \
  }
          PL_SPLAT_STEP(ev.x, ew.x, x0, k)
          PL_SPLAT_STEP(ev.y, ew.y, x1, k + 1)
          PL_SPLAT_STEP(ev.z, ew.z, x2, k + 2)
          PL_SPLAT_STEP(ev.w, ew.w, x3, k + 3)
#undef PL_SPLAT_STEP
        }
        red_add_v4(a.val_out + (int64_t)cur * L + 4 * l, acc);
      }
    }
    __syncthreads();  // P tile and splat record are dead: the next tile may overwrite them
    if (MODE != 2 && tid == 0 && has_next) {
      tc::mbar_arrive_expect_tx(&bars[3], 9 * TE);
      tc::tma_load_1d(s_rp, a.rec_splat + next * (int64_t)(9 * TE), 9 * TE, &bars[3]);
    }
  }

  tc::tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, C::TMEM_COLS);
}

// Mu[l][k] = muT[k][l] -> hi / lo planes in the canonical K-major operand layout
// (chunk kc = k / 4: rows l = 0..L-1, 16 bytes each, then one 16-byte pad).
__global__ void __launch_bounds__(256)
mu_split_kernel(const float* __restrict__ muT, int L, float* __restrict__ out) {
  const int chunk_floats = L * 4 + 4;
  const int plane = (L / 4) * chunk_floats;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < plane; i += gridDim.x * blockDim.x) {
    const int kc = i / chunk_floats, rem = i - kc * chunk_floats;
    float hi = 0.f, lo = 0.f;
    if (rem < L * 4) {
      const int row = rem >> 2, k = kc * 4 + (rem & 3);
      const float m = muT[k * L + row];
      hi = tc::tf32_rna(m);
      lo = tc::tf32_rna(m - hi);
    }
    out[i] = hi;
    out[plane + i] = lo;
  }
}

template <int L>
size_t tc_smem(int d1, bool lattice) {
  using C = Cfg<L>;
  const int TE = lattice ? kTile * d1 : 0;
  return (size_t)C::QA_BYTES + 2 * C::B_BYTES + 15 * (size_t)TE + (lattice ? 4 * kThreads : 0) + 64;
}

template <int L, int MODE, int D1T>
int launch_tc(const MfArgs& a, cudaStream_t st) {
  const size_t smem = tc_smem<L>(a.d1, MODE != 2);
  static size_t attr_smem_dev[64] = {};
  size_t& attr_smem = attr_smem_dev[current_device_index()];
  if (smem > attr_smem) {
    PL_CUDA(cudaFuncSetAttribute((mf_tc_kernel<L, MODE, D1T>),
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_smem = smem;
  }
  const int64_t cap = (int64_t)PL_SM_COUNT * (smem <= 113 * 1024 ? kCtasPerSm : 1);
  const unsigned grid = (unsigned)(a.ntiles < cap ? a.ntiles : cap);
  PL_LAUNCH((mf_tc_kernel<L, MODE, D1T>), grid, kThreads, smem, st, a);
  return PL_OK;
}

template <int MODE>
int dispatch_tc(int L, const MfArgs& a, cudaStream_t st) {
#define PL_TC_CASE(LL)                                                   \
  case LL:                                                               \
    if (MODE != 2 && a.d1 == 6) return launch_tc<LL, MODE, 6>(a, st);    \
    return launch_tc<LL, MODE, 0>(a, st);
  switch (L) {
    PL_TC_CASE(16)
    PL_TC_CASE(32)
    PL_TC_CASE(64)
    default:
      set_last_error("tensor-core mean-field kernel: unsupported label count %d", L);
      return PL_ERR_INVALID;
  }
#undef PL_TC_CASE
}

size_t tc_smem_for(int L, int d1, bool lattice) {
  switch (L) {
    case 16: return tc_smem<16>(d1, lattice);
    case 32: return tc_smem<32>(d1, lattice);
    case 64: return tc_smem<64>(d1, lattice);
    default: return (size_t)1 << 30;
  }
}

}  // namespace

bool mf_tc_supported(const Lattice* L, int labels) {
  if (labels != 16 && labels != 32 && labels != 64) return false;
  if (L == nullptr) return true;  // stand-alone softmax + compatibility
  return L->ntiles > 0 && L->rec_slice != nullptr &&
         tc_smem_for(labels, L->d + 1, true) <= 200 * 1024;
}

size_t mf_tc_mu_bytes(int labels) { return 2 * (size_t)(labels / 4) * (labels * 16 + 16); }

int mf_tc_prepare_mu(const float* muT, int labels, float* mu_canon, cudaStream_t st) {
  PL_LAUNCH(mu_split_kernel, 8, 256, 0, st, muT, labels, mu_canon);
  return PL_OK;
}

// One iteration.  first: E = E0 (no slice, p_prev unused).  val_out must be zeroed.
int mf_tc_iteration(Lattice* L, bool first, const float* val_in, float* val_out, const float* e0,
                    const float* p_prev, float* p_next, const float* mu_canon, float scale,
                    int labels, cudaStream_t st) {
  MfArgs a = {};
  a.val_in = val_in; a.val_out = val_out; a.e0 = e0; a.p_prev = p_prev; a.p_next = p_next;
  a.q_out = nullptr; a.mu_canon = mu_canon; a.scale = scale; a.n = L->n; a.ntiles = L->ntiles;
  a.d1 = L->d + 1;
  a.inv_denom = 1.0f / (1 + powf(2, -L->d));
  a.vid = L->vid; a.rec_slice = L->rec_slice; a.rec_splat = L->rec_splat;
  a.useg = L->useg; a.tile_off = L->tile_off; a.tile_U = L->tile_U;
  if (first) return dispatch_tc<0>(labels, a, st);
  return dispatch_tc<1>(labels, a, st);
}

// P = scale * softmax(-E) Mu^T without a lattice (pl_softmax_compat), optional Q output.
int softmax_compat_tc(const float* e, const float* mu_canon, float* p, float* q_out, int64_t n,
                      int labels, float scale, cudaStream_t st) {
  MfArgs a = {};
  a.e0 = e; a.p_next = p; a.q_out = q_out; a.mu_canon = mu_canon; a.scale = scale; a.n = n;
  a.ntiles = (n + kTile - 1) / kTile;
  a.d1 = 0;
  return dispatch_tc<2>(labels, a, st);
}

int tc_timeouts() {
  int v = 0;
  cudaMemcpyFromSymbol(&v, g_tc_timeouts, sizeof(int));
  return v;
}

}  // namespace pl
