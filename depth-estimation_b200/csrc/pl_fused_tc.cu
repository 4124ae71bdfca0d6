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
constexpr int kStageRows = 144;  // lattice rows staged per tile

__device__ int g_tc_timeouts = 0;

constexpr int cmax(int a, int b) { return a > b ? a : b; }

template <int L>
struct Cfg {
  static constexpr int V = L / 4;               // float4 lanes per pixel row (4 / 8 / 16)
  static constexpr int RPP = kThreads / V;      // rows per pass
  static constexpr int RPT = kTile / RPP;       // rows per thread (2 / 4 / 8)
  static constexpr int KCH = L / 4;             // 16-byte K chunks per operand row
  static constexpr int A_CH = kTile * 16 + 16;  // bytes per K chunk of A (padded)
  static constexpr int B_CH = L * 16 + 16;      // bytes per K chunk of B (padded)
  static constexpr int A_BYTES = KCH * A_CH;    // one plane
  static constexpr int B_BYTES = KCH * B_CH;
  static constexpr int PLD = L + 4;             // P row pitch in floats
  static constexpr int QA_BYTES = cmax(2 * A_BYTES, kTile * PLD * 4);
  static constexpr int R_BYTES = cmax(kStageRows * L * 4, 2 * B_BYTES);
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
  int d1;
  float inv_denom;
  const uint16_t* loc;
  const float* bary;
  const int32_t* vid;
  const uint8_t* ent_pix;
  const float* ent_w;
  const int2* useg;
  const int* tile_off;
  const int* tile_U;
};

__device__ __forceinline__ void red_add_v4(float* addr, float4 x) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(x.x), "f"(x.y),
               "f"(x.z), "f"(x.w)
               : "memory");
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
template <int L, int MODE>
__global__ void __launch_bounds__(kThreads, 2) mf_tc_kernel(const MfArgs a) {
  using C = Cfg<L>;
  constexpr int V = C::V, RPP = C::RPP, RPT = C::RPT;
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* s_qa = smem;                          // Q hi | Q lo operand planes; later P
  unsigned char* s_r = smem + C::QA_BYTES;             // staged lattice rows; later Mu hi | lo
  const int d1 = a.d1;
  const int TE = (MODE == 2) ? 0 : kTile * d1;
  float* s_w = reinterpret_cast<float*>(s_r + C::R_BYTES);   // [TE] slice weights (entry order)
  int* s_idx = reinterpret_cast<int*>(s_w + TE);             // [TE] slice row (local or global)
  float* s_sw = reinterpret_cast<float*>(s_idx + TE);        // [TE] splat weights (sorted)
  uint8_t* s_spix = reinterpret_cast<uint8_t*>(s_sw + TE);   // [TE] splat pixel (sorted)
  uint64_t* bars = reinterpret_cast<uint64_t*>(
      smem + ((C::QA_BYTES + C::R_BYTES + 13 * TE + 15) & ~15));  // [0] Mu landed, [1] MMAs done
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t tile = blockIdx.x;
  const int64_t p0 = tile * kTile;
  const int rows = (int)min((int64_t)kTile, a.n - p0);
  const int g = tid / V, l = tid % V;
  const float* s_val = reinterpret_cast<const float*>(s_r);

  // ---- one-time setup: barriers, TMEM, (MODE 0 / 2) Mu fetch ----------------------------
  if (warp == 0) {
    if (lane == 0) {
      tc::mbar_init(&bars[0], 1);
      tc::mbar_init(&bars[1], 1);
      tc::fence_mbar_init();
      if (MODE != 1) {  // region R is free from the start: fetch Mu while the softmax runs
        tc::mbar_arrive_expect_tx(&bars[0], 2 * C::B_BYTES);
        tc::tma_load_1d(s_r, a.mu_canon, 2 * C::B_BYTES, &bars[0]);
      }
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tcgen05_fence_before();
  }

  int U = 0;
  const int2* seg = nullptr;
  bool staged = false;
  if (MODE != 2) {
    U = a.tile_U[tile];
    seg = a.useg + a.tile_off[tile];
    staged = U <= kStageRows;
    const int valid = rows * d1;
    for (int i = tid; i < valid; i += kThreads) {
      s_sw[i] = a.ent_w[tile * TE + i];
      s_spix[i] = a.ent_pix[tile * TE + i];
      if (MODE == 1) {
        s_w[i] = a.bary[tile * TE + i] * a.inv_denom;
        s_idx[i] = staged ? (int)a.loc[tile * TE + i] : a.vid[tile * TE + i];
      }
    }
    if (MODE == 1 && staged) {
      float* sv = reinterpret_cast<float*>(s_r);
      for (int i = tid; i < U * V; i += kThreads) {
        const int u = i / V, v = i - u * V;
        *reinterpret_cast<float4*>(sv + u * L + 4 * v) =
            *reinterpret_cast<const float4*>(a.val_in + (int64_t)seg[u].x * L + 4 * v);
      }
    }
  }
  // this thread's E0 / P_prev vectors, in flight while the staging loads land
  float4 e_reg[RPT], pp_reg[RPT];
#pragma unroll
  for (int i = 0; i < RPT; i++) {
    const int r = g + i * RPP;
    e_reg[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    pp_reg[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < rows) {
      e_reg[i] = *reinterpret_cast<const float4*>(a.e0 + (p0 + r) * L + 4 * l);
      if (MODE == 1) pp_reg[i] = *reinterpret_cast<const float4*>(a.p_prev + (p0 + r) * L + 4 * l);
    }
  }
  __syncthreads();
  tc::tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot;

  // ---- slice + energy + softmax -> Q hi / lo operand planes ------------------------------
#pragma unroll
  for (int i = 0; i < RPT; i++) {
    const int r = g + i * RPP;
    const bool live = r < rows;
    float4 e = e_reg[i];
    if (MODE == 1 && live) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int j = 0; j < d1; j++) {
        const float w = s_w[r * d1 + j];
        const int ix = s_idx[r * d1 + j];
        const float4 x = staged
                             ? *reinterpret_cast<const float4*>(s_val + ix * L + 4 * l)
                             : *reinterpret_cast<const float4*>(a.val_in + (int64_t)ix * L + 4 * l);
        acc.x = __fmaf_rn(w, x.x, acc.x);
        acc.y = __fmaf_rn(w, x.y, acc.y);
        acc.z = __fmaf_rn(w, x.z, acc.z);
        acc.w = __fmaf_rn(w, x.w, acc.w);
      }
      // E = E0 + (filtered - P)   (crf_module.py:101, gaussian_matrix.py:352)
      e.x = __fadd_rn(e.x, __fsub_rn(acc.x, pp_reg[i].x));
      e.y = __fadd_rn(e.y, __fsub_rn(acc.y, pp_reg[i].y));
      e.z = __fadd_rn(e.z, __fsub_rn(acc.z, pp_reg[i].z));
      e.w = __fadd_rn(e.w, __fsub_rn(acc.w, pp_reg[i].w));
    }
    float m = live ? fmaxf(fmaxf(-e.x, -e.y), fmaxf(-e.z, -e.w)) : -INFINITY;
    m = group_max<V>(m);
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (live) {
      x.x = expf(-e.x - m);
      x.y = expf(-e.y - m);
      x.z = expf(-e.z - m);
      x.w = expf(-e.w - m);
    }
    const float s = group_sum<V>(x.x + x.y + x.z + x.w);
    const float inv = live ? 1.0f / s : 0.f;
    x.x *= inv; x.y *= inv; x.z *= inv; x.w *= inv;
    if (MODE == 2 && a.q_out != nullptr && live)
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
  tc::fence_proxy_async();  // Q planes (generic proxy) -> visible to the tensor core's reads
  __syncthreads();          // also: nobody reads the staged lattice rows any more

  // ---- P = Q Mu^T on the tensor cores: one thread issues ---------------------------------
  if (tid == 0) {
    if (MODE == 1) {
      tc::mbar_arrive_expect_tx(&bars[0], 2 * C::B_BYTES);
      tc::tma_load_1d(s_r, a.mu_canon, 2 * C::B_BYTES, &bars[0]);
    }
    if (!tc::mbar_wait(&bars[0], 0)) atomicAdd(&g_tc_timeouts, 1);
    tc::tcgen05_fence_after();
    constexpr uint32_t idesc = tc::idesc_tf32(kTile, L);
    const uint32_t a_hi = tc::smem_u32(s_qa), a_lo = a_hi + C::A_BYTES;
    const uint32_t b_hi = tc::smem_u32(s_r), b_lo = b_hi + C::B_BYTES;
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
  __syncwarp();
  if (!tc::mbar_wait(&bars[1], 0)) atomicAdd(&g_tc_timeouts, 1);
  __syncwarp();
  tc::tcgen05_fence_after();

  // ---- epilogue: TMEM -> registers -> P tile in shared memory (over the dead Q planes) ---
  float* s_p = reinterpret_cast<float*>(s_qa);
  {
    const int q = warp & 3, hsel = warp >> 2;  // TMEM lane quarter, column half
    const int row = q * 32 + lane;
#pragma unroll
    for (int c8 = 0; c8 < (L / 2) / 8; c8++) {
      const int col = hsel * (L / 2) + c8 * 8;
      float v[8];
      tc::tmem_ld_x8(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)col, v);
      float4 o0 = make_float4(v[0] * a.scale, v[1] * a.scale, v[2] * a.scale, v[3] * a.scale);
      float4 o1 = make_float4(v[4] * a.scale, v[5] * a.scale, v[6] * a.scale, v[7] * a.scale);
      *reinterpret_cast<float4*>(s_p + row * C::PLD + col) = o0;
      *reinterpret_cast<float4*>(s_p + row * C::PLD + col + 4) = o1;
    }
  }
  tc::tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, C::TMEM_COLS);

  // ---- P -> HBM (coalesced rows) ----------------------------------------------------------
#pragma unroll
  for (int i = 0; i < RPT; i++) {
    const int r = g + i * RPP;
    if (r < rows)
      *reinterpret_cast<float4*>(a.p_next + (p0 + r) * L + 4 * l) =
          *reinterpret_cast<const float4*>(s_p + r * C::PLD + 4 * l);
  }

  // ---- splat P: segment sums per distinct vertex of the tile -----------------------------
  if (MODE != 2) {
    for (int u = g; u < U; u += RPP) {
      const int2 sg = seg[u];
      const int st = sg.y & 0xFFFF, cnt = sg.y >> 16;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k = st; k < st + cnt; k++) {
        const float w = s_sw[k];
        const float4 x = *reinterpret_cast<const float4*>(s_p + (int)s_spix[k] * C::PLD + 4 * l);
        acc.x = __fmaf_rn(w, x.x, acc.x);
        acc.y = __fmaf_rn(w, x.y, acc.y);
        acc.z = __fmaf_rn(w, x.z, acc.z);
        acc.w = __fmaf_rn(w, x.w, acc.w);
      }
      red_add_v4(a.val_out + (int64_t)sg.x * L + 4 * l, acc);
    }
  }
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
  return (size_t)((C::QA_BYTES + C::R_BYTES + 13 * TE + 15) & ~15) + 32;
}

template <int L, int MODE>
int launch_tc(const MfArgs& a, unsigned grid, cudaStream_t st) {
  const size_t smem = tc_smem<L>(a.d1, MODE != 2);
  static size_t attr_smem = 0;
  if (smem > attr_smem) {
    PL_CUDA(cudaFuncSetAttribute(mf_tc_kernel<L, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
    attr_smem = smem;
  }
  PL_LAUNCH((mf_tc_kernel<L, MODE>), grid, kThreads, smem, st, a);
  return PL_OK;
}

template <int MODE>
int dispatch_tc(int L, const MfArgs& a, unsigned grid, cudaStream_t st) {
  switch (L) {
    case 16: return launch_tc<16, MODE>(a, grid, st);
    case 32: return launch_tc<32, MODE>(a, grid, st);
    case 64: return launch_tc<64, MODE>(a, grid, st);
    default:
      set_last_error("tensor-core mean-field kernel: unsupported label count %d", L);
      return PL_ERR_INVALID;
  }
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
  return L->ntiles > 0 && tc_smem_for(labels, L->d + 1, true) <= 200 * 1024;
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
  MfArgs a;
  a.val_in = val_in; a.val_out = val_out; a.e0 = e0; a.p_prev = p_prev; a.p_next = p_next;
  a.q_out = nullptr; a.mu_canon = mu_canon; a.scale = scale; a.n = L->n; a.d1 = L->d + 1;
  a.inv_denom = 1.0f / (1 + powf(2, -L->d));
  a.loc = L->loc; a.bary = L->bary; a.vid = L->vid; a.ent_pix = L->ent_pix; a.ent_w = L->ent_w;
  a.useg = L->useg; a.tile_off = L->tile_off; a.tile_U = L->tile_U;
  if (first) return dispatch_tc<0>(labels, a, (unsigned)L->ntiles, st);
  return dispatch_tc<1>(labels, a, (unsigned)L->ntiles, st);
}

// P = scale * softmax(-E) Mu^T without a lattice (pl_softmax_compat), optional Q output.
int softmax_compat_tc(const float* e, const float* mu_canon, float* p, float* q_out, int64_t n,
                      int labels, float scale, cudaStream_t st) {
  MfArgs a = {};
  a.e0 = e; a.p_next = p; a.q_out = q_out; a.mu_canon = mu_canon; a.scale = scale; a.n = n;
  a.d1 = 0;
  return dispatch_tc<2>(labels, a, ceil_div_u((uint64_t)n, kTile), st);
}

int tc_timeouts() {
  int v = 0;
  cudaMemcpyFromSymbol(&v, g_tc_timeouts, sizeof(int));
  return v;
}

}  // namespace pl
