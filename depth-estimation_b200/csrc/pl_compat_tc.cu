// pl_compat_tc.cu -- softmax + label-compatibility transform for LARGE label counts on the
// tcgen05 tensor cores (sm_100a):  P[p][j] = scale * sum_k Mu[l0 + j][k] * softmax_k(-E[p][.])
// for j < N output labels (N = L on one GPU, N = L / G when the labels are sharded over G ranks).
//
// crf/crf_module.py:99-102 with charb.forward (:71-75), pixel-major.  L is a multiple of 64
// (128, 256, ...), N a multiple of BN (32 or 64), N <= 256 (TMEM accumulator columns).
//
// Per tile of 128 pixels (= UMMA M), persistent CTAs:
//   pass 1  online softmax statistics (row max, row sum) over all L labels, E from HBM
//   pass 2  for every 64-wide K chunk: re-read the E chunk (L2), Q = exp(-E - max) / sum,
//           split Q = hi + lo (tf32) into the canonical K-major operand planes A, then for every
//           BN-wide block of output labels three MMAs (lo*hi + hi*lo + hi*hi, fp32 parity, see
//           pl_fused_tc.cu) accumulate into that block's TMEM columns.  The Mu blocks (pre-split
//           hi / lo planes, one contiguous record per (K chunk, N block)) stream through a
//           two-deep TMA ring.
//   epilogue TMEM -> registers -> shared memory -> coalesced rows of P
//
// E may be given in the BLOCKED layout an all-gather over label shards produces:
// e[b][p][lw] with b < nblocks, lw < block_w, label = b * block_w + lw  (nblocks = 1: plain [n][L]).
#include <math.h>
#include <stdlib.h>

#include "pl_common.cuh"
#include "pl_tc.cuh"

namespace pl {

namespace {

constexpr int kTile = PL_TILE;  // 128
constexpr int kThreads = 512;
constexpr int kKC = 64;                          // K chunk (labels per A / B record)
constexpr int kA_CH = kTile * 16 + 16;           // bytes per 16-byte K chunk column of A
constexpr int kA_BYTES = (kKC / 4) * kA_CH;      // one plane of A: 33024
constexpr int kPLD = 64 + 4;                     // epilogue staging pitch (floats)
constexpr int kRing = 2;                         // Mu records in flight (TMA ring slots)
constexpr int kSup = 1;                          // 128-pixel tiles per Mu record pass (2 halves the L2 traffic of Mu but measured slower: the kernel is latency-bound, not L2-bound)

__device__ int g_ctc_timeouts = 0;

template <int BN>
struct BCfg {
  static constexpr int B_CH = BN * 16 + 16;
  static constexpr int B_BYTES = (kKC / 4) * B_CH;   // one plane
  static constexpr int REC = 2 * B_BYTES;            // hi | lo
};

constexpr int kMaxBlocks = 16;

struct CArgs {
  const float* e;       // blocked [nblocks][n][block_w]
  // Optional: block b lives at blk[b] ([n][block_w]) instead of e + b * n * block_w.  The
  // pointers may be peer-GPU memory mapped through CUDA IPC (label shards read in place over
  // NVLink, no all-gathered copy); blk[0] == e.
  const float* blk[kMaxBlocks];
  int use_blk;
  int nblocks, block_w;
  const uint8_t* mu_rec;  // [(L/64)][N/BN] records of BCfg<BN>::REC bytes
  float* p;             // [n][N]
  float* q_out;         // [n][L] plain layout, or null
  int64_t n, ntiles;
  int L, N;
  float scale;
  int debug;  // timing experiments only: 1 = no MMAs, 2 = no statistics pass, 4 = no epilogue
};

constexpr int kMaxKC = 8;  // L <= 512

template <int BN>
__global__ void __launch_bounds__(kThreads, 1) compat_tc_kernel(const CArgs a) {
  using B = BCfg<BN>;
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* s_a = smem;                       // A hi | lo planes; epilogue staging
  unsigned char* s_b = smem + kSup * 2 * kA_BYTES;  // kRing B records (ring)
  float* s_m = reinterpret_cast<float*>(s_b + kRing * B::REC);   // [128] row max of -E
  float* s_inv = s_m + kSup * kTile;                         // 1 / row sum
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_inv + kSup * kTile);
  // bars: [0..R) B record landed (ring slot)  [R..2R) ring slot free  [2R] all MMAs on A done
  constexpr int kBarA = 2 * kRing;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + kBarA + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = tid >> 4, l = tid & 15;   // 16 lanes per pixel row
  constexpr int RPP = kThreads / 16;      // rows per pass (32)
  constexpr int RPT = kSup * kTile / RPP; // rows per thread (8), over the kSup tiles
  const int KCH = a.L / kKC;              // K chunks
  const int NB = a.N / BN;                // output blocks
  const int steps_per_tile = KCH * NB;
  const uint32_t tmem_cols = kSup * (a.N <= 32 ? 32 : (a.N <= 64 ? 64 : (a.N <= 128 ? 128 : 256)));

  if (warp == 0) {
    if (lane == 0) {
#pragma unroll
      for (int b = 0; b <= kBarA; b++) tc::mbar_init(&bars[b], 1);
      tc::fence_mbar_init();
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, tmem_cols);
    tc::tcgen05_fence_before();
  }
  __syncthreads();
  tc::tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot;

  // Where this thread's float4 of K chunk kc lives, without the pixel term: the blocked
  // layout puts label (b * block_w + w) of pixel p at e[(b * n + p) * block_w + w].
  int64_t eoff[kMaxKC];
#pragma unroll
  for (int kc = 0; kc < kMaxKC; kc++) {
    const int label = kc * kKC + 4 * l;
    const int b = label / a.block_w;
    eoff[kc] = (a.use_blk ? (int64_t)(a.blk[b < kMaxBlocks ? b : 0] - a.e)
                          : (int64_t)b * a.n * a.block_w) + (label - b * a.block_w);
  }
  const int64_t bw = a.block_w;

  // Every tile consumes the same sequence of Mu records, so the ring simply cycles: global step
  // t of this CTA uses record t % steps_per_tile in slot t % kRing; requests run kRing-1 ahead.
  const uint32_t my_tiles = (uint32_t)((a.ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x);
  const uint32_t total_steps = my_tiles * (uint32_t)steps_per_tile;
  auto request = [&](uint32_t t) {   // tid 0 only
    const uint32_t slot = t % kRing, use = t / kRing;
    if (use > 0 && !tc::mbar_wait(&bars[kRing + slot], (use - 1) & 1)) atomicAdd(&g_ctc_timeouts, 1);
    tc::mbar_arrive_expect_tx(&bars[slot], B::REC);
    tc::tma_load_1d(s_b + slot * B::REC, a.mu_rec + (size_t)(t % steps_per_tile) * B::REC, B::REC,
                    &bars[slot]);
  };
  if (tid == 0)
    for (uint32_t t = 0; t < kRing - 1 && t < total_steps; t++) request(t);

  uint32_t step = 0;      // global MMA-step counter of this CTA
  uint32_t a_uses = 0;    // how many times the A planes have been handed to the tensor core
  const uint32_t ncols1 = tmem_cols / kSup;   // TMEM columns per 128-pixel tile
  for (int64_t st = blockIdx.x; st < a.ntiles; st += gridDim.x) {   // super-tiles of kSup tiles
    const int64_t p0 = st * (kSup * kTile);
    const int rows = (int)min((int64_t)(kSup * kTile), a.n - p0);

    // ---- pass 1: online softmax statistics of -E over all L labels ------------------------
    // RPT independent rows per thread advance together: RPT loads in flight per step
    {
      float m[RPT], sum[RPT];
#pragma unroll
      for (int i = 0; i < RPT; i++) {
        m[i] = -INFINITY;
        sum[i] = 0.f;
      }
#pragma unroll
      for (int kc = 0; kc < kMaxKC; kc++) {
        if (kc >= KCH || (a.debug & 2)) break;
        float4 e[RPT];
#pragma unroll
        for (int i = 0; i < RPT; i++) {
          const int r = g + i * RPP;
          e[i] = (r < rows) ? *reinterpret_cast<const float4*>(a.e + eoff[kc] + (p0 + r) * bw)
                            : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int i = 0; i < RPT; i++) {
          const float mx = fmaxf(fmaxf(-e[i].x, -e[i].y), fmaxf(-e[i].z, -e[i].w));
          const float mn = fmaxf(m[i], mx);
          sum[i] = sum[i] * __expf(m[i] - mn) + __expf(-e[i].x - mn) + __expf(-e[i].y - mn) +
                   __expf(-e[i].z - mn) + __expf(-e[i].w - mn);
          m[i] = mn;
        }
      }
#pragma unroll
      for (int i = 0; i < RPT; i++) {
        const int r = g + i * RPP;
        float mi = m[i], si = sum[i];
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {  // combine the 16 lanes' (max, sum) pairs
          const float m2 = __shfl_xor_sync(0xffffffffu, mi, o, 16);
          const float s2 = __shfl_xor_sync(0xffffffffu, si, o, 16);
          const float mn = fmaxf(mi, m2);
          si = si * __expf(mi - mn) + s2 * __expf(m2 - mn);
          mi = mn;
        }
        if (l == 0) {
          s_m[r] = mi;
          s_inv[r] = (r < rows) ? __fdividef(1.0f, si) : 0.f;
        }
      }
    }
    __syncthreads();

    // ---- pass 2: K chunks -------------------------------------------------------------------
#pragma unroll
    for (int kc = 0; kc < kMaxKC; kc++) {
      if (kc >= KCH) break;
      // this chunk's E values first: the loads overlap the MMAs still reading the previous
      // chunk's A planes
      float4 e[RPT];
#pragma unroll
      for (int i = 0; i < RPT; i++) {
        const int r = g + i * RPP;
        e[i] = (r < rows) ? *reinterpret_cast<const float4*>(a.e + eoff[kc] + (p0 + r) * bw)
                          : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      // the tensor core must be done with the previous contents of the A planes
      if (a_uses > 0 && !tc::mbar_wait(&bars[kBarA], (a_uses - 1) & 1)) atomicAdd(&g_ctc_timeouts, 1);
      tc::tcgen05_fence_after();
#pragma unroll
      for (int i = 0; i < RPT; i++) {
        const int r = g + i * RPP;
        float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < rows) {
          const float m = s_m[r], inv = s_inv[r];
          x.x = __expf(-e[i].x - m) * inv;
          x.y = __expf(-e[i].y - m) * inv;
          x.z = __expf(-e[i].z - m) * inv;
          x.w = __expf(-e[i].w - m) * inv;
          if (a.q_out != nullptr)
            *reinterpret_cast<float4*>(a.q_out + (p0 + r) * a.L + kc * kKC + 4 * l) = x;
        }
        float4 hi, lo;
        hi.x = tc::tf32_rna(x.x); lo.x = tc::tf32_rna(x.x - hi.x);
        hi.y = tc::tf32_rna(x.y); lo.y = tc::tf32_rna(x.y - hi.y);
        hi.z = tc::tf32_rna(x.z); lo.z = tc::tf32_rna(x.z - hi.z);
        hi.w = tc::tf32_rna(x.w); lo.w = tc::tf32_rna(x.w - hi.w);
        // tile t = r / 128 of the super-tile has its own pair of planes
        unsigned char* pa = s_a + (r >> 7) * (2 * kA_BYTES) + l * kA_CH + (r & 127) * 16;
        *reinterpret_cast<float4*>(pa) = hi;
        *reinterpret_cast<float4*>(pa + kA_BYTES) = lo;
      }
      tc::fence_proxy_async();
      __syncthreads();

      if (tid == 0) {
        tc::tcgen05_fence_after();
        const uint32_t idesc = tc::idesc_tf32(kTile, BN);
        for (int nb = 0; nb < NB; nb++, step++) {
          const uint32_t slot = step % kRing;
          if (!tc::mbar_wait(&bars[slot], (step / kRing) & 1)) atomicAdd(&g_ctc_timeouts, 1);
          tc::tcgen05_fence_after();
          const uint32_t b_hi = tc::smem_u32(s_b + slot * B::REC), b_lo = b_hi + B::B_BYTES;
#pragma unroll
          for (int t = 0; t < kSup; t++) {   // one Mu record feeds both tiles of the super-tile
            const uint32_t a_hi = tc::smem_u32(s_a) + t * (2 * kA_BYTES), a_lo = a_hi + kA_BYTES;
#pragma unroll
            for (int pass = 0; pass < 3; pass++) {  // lo*hi, hi*lo, hi*hi
              const uint32_t ab = (pass == 0) ? a_lo : a_hi;
              const uint32_t bb = (pass == 1) ? b_lo : b_hi;
#pragma unroll
              for (int ks = 0; ks < kKC / 8; ks++) {
                const uint64_t ad = tc::smem_desc_kmajor(ab + ks * 2 * kA_CH, kA_CH, 128);
                const uint64_t bd = tc::smem_desc_kmajor(bb + ks * 2 * B::B_CH, B::B_CH, 128);
                if (!(a.debug & 1))
                  tc::mma_tf32_ss(tmem + t * ncols1 + (uint32_t)(nb * BN), ad, bd, idesc,
                                  (kc | pass | ks) != 0);
              }
            }
          }
          tc::mma_commit(&bars[kRing + slot]);   // ring slot free once these MMAs have read it
          if (step + kRing - 1 < total_steps) request(step + kRing - 1);
        }
        tc::mma_commit(&bars[kBarA]);            // A planes free (and, after the last chunk, P ready)
      }
      a_uses++;
      __syncwarp();
    }

    // ---- epilogue: per tile, per 64-column block TMEM -> staging -> coalesced rows of P ------
    if (!tc::mbar_wait(&bars[kBarA], (a_uses - 1) & 1)) atomicAdd(&g_ctc_timeouts, 1);
    __syncwarp();
    tc::tcgen05_fence_after();
    float* s_p = reinterpret_cast<float*>(s_a);
    for (int t = 0; t < kSup; t++) {
      if (t * kTile >= rows) break;
      for (int c0 = 0; c0 < ((a.debug & 4) ? 0 : a.N); c0 += 64) {
        const int cw = min(64, a.N - c0);          // 32 or 64
        const int q = warp & 3, hsel = warp >> 2;  // TMEM lane quarter; column group of the block
        const int row = q * 32 + lane;
        constexpr int NCG = kThreads / 128;        // column groups (4)
        const int cpw = cw / NCG;                  // columns per warp: 8 or 16
        for (int c8 = 0; c8 < cpw / 8; c8++) {
          const int col = hsel * cpw + c8 * 8;
          float v[8];
          tc::tmem_ld_x8(tmem + ((uint32_t)(q * 32) << 16) + t * ncols1 + (uint32_t)(c0 + col), v);
          *reinterpret_cast<float4*>(s_p + row * kPLD + col) =
              make_float4(v[0] * a.scale, v[1] * a.scale, v[2] * a.scale, v[3] * a.scale);
          *reinterpret_cast<float4*>(s_p + row * kPLD + col + 4) =
              make_float4(v[4] * a.scale, v[5] * a.scale, v[6] * a.scale, v[7] * a.scale);
        }
        __syncthreads();
        // 128 rows x cw columns, 16 lanes per row, 32 rows per pass
#pragma unroll
        for (int i = 0; i < kTile / RPP; i++) {
          const int r = g + i * RPP;
          if (t * kTile + r < rows && 4 * l < cw)
            *reinterpret_cast<float4*>(a.p + (p0 + t * kTile + r) * a.N + c0 + 4 * l) =
                *reinterpret_cast<const float4*>(s_p + r * kPLD + 4 * l);
        }
        __syncthreads();
      }
    }
    tc::tcgen05_fence_before();
    __syncthreads();   // TMEM reads done before the next super-tile's first MMA overwrites it
  }

  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, tmem_cols);
}

// Mu[l0 + nb*BN + j][kc*64 + k] = muT[(kc*64 + k) * L + l0 + nb*BN + j]  ->  record (kc, nb):
// hi plane | lo plane, each in the canonical K-major layout (16-byte K chunks of BN rows + pad).
template <int BN>
__global__ void __launch_bounds__(256)
mu_records_kernel(const float* __restrict__ muT, int L, int l0, int N, uint8_t* __restrict__ out) {
  using B = BCfg<BN>;
  const int NB = N / BN;
  const int nrec = (L / kKC) * NB;
  const int plane_floats = B::B_BYTES / 4;
  const int chunk_floats = B::B_CH / 4;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nrec * plane_floats;
       i += gridDim.x * blockDim.x) {
    const int rec = i / plane_floats, w = i - rec * plane_floats;
    const int kc = rec / NB, nb = rec - kc * NB;
    const int ch = w / chunk_floats, rem = w - ch * chunk_floats;
    float hi = 0.f, lo = 0.f;
    if (rem < BN * 4) {
      const int j = rem >> 2, k = kc * kKC + ch * 4 + (rem & 3);
      const float m = muT[(int64_t)k * L + l0 + nb * BN + j];
      hi = tc::tf32_rna(m);
      lo = tc::tf32_rna(m - hi);
    }
    float* base = reinterpret_cast<float*>(out + (size_t)rec * B::REC);
    base[w] = hi;
    base[plane_floats + w] = lo;
  }
}

template <int BN>
size_t ctc_smem() {
  return kSup * 2 * (size_t)kA_BYTES + kRing * (size_t)BCfg<BN>::REC + kSup * 2 * kTile * 4 +
         (2 * kRing + 1) * 8 + 16;
}

template <int BN>
int launch_ctc(const CArgs& a, cudaStream_t st) {
  const size_t smem = ctc_smem<BN>();
  static bool done = false;
  if (!done) {
    PL_CUDA(cudaFuncSetAttribute(compat_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
    done = true;
  }
  const unsigned grid = (unsigned)(a.ntiles < PL_SM_COUNT ? a.ntiles : PL_SM_COUNT);
  PL_LAUNCH((compat_tc_kernel<BN>), grid, kThreads, smem, st, a);
  return PL_OK;
}

}  // namespace

// Output labels l0 .. l0+N of a model with L labels.  Supported: L % 64 == 0, N % 32 == 0,
// 32 <= N <= 256, and the blocked input layout must not split a float4 (block_w % 4 == 0).
bool compat_tc_supported(int L, int N, int block_w) {
  return L >= 64 && L <= kMaxKC * kKC && L % 64 == 0 && N >= 32 && N <= 256 && N % 32 == 0 && block_w % 4 == 0 &&
         L % block_w == 0;
}

size_t compat_tc_mu_bytes(int L, int N) {
  const int BN = (N % 64 == 0) ? 64 : 32;
  const size_t rec = BN == 64 ? BCfg<64>::REC : BCfg<32>::REC;
  return (size_t)(L / kKC) * (N / BN) * rec;
}

int compat_tc_prepare_mu(const float* muT, int L, int l0, int N, uint8_t* mu_rec, cudaStream_t st) {
  if (N % 64 == 0) {
    PL_LAUNCH((mu_records_kernel<64>), 64, 256, 0, st, muT, L, l0, N, mu_rec);
  } else {
    PL_LAUNCH((mu_records_kernel<32>), 64, 256, 0, st, muT, L, l0, N, mu_rec);
  }
  return PL_OK;
}

int compat_tc_run(const float* e, int nblocks, int block_w, const uint8_t* mu_rec, float* p,
                  float* q_out, int64_t n, int L, int N, float scale, cudaStream_t st,
                  const float* const* block_ptrs) {
  CArgs a;
  a.use_blk = 0;
  for (int b = 0; b < kMaxBlocks; b++) a.blk[b] = nullptr;
  if (block_ptrs != nullptr) {
    if (nblocks > kMaxBlocks) {
      set_last_error("at most %d label blocks", kMaxBlocks);
      return PL_ERR_INVALID;
    }
    for (int b = 0; b < nblocks; b++) a.blk[b] = block_ptrs[b];
    a.use_blk = 1;
    e = block_ptrs[0];
  }
  a.e = e; a.nblocks = nblocks; a.block_w = block_w; a.mu_rec = mu_rec; a.p = p; a.q_out = q_out;
  a.n = n; a.ntiles = (n + kSup * kTile - 1) / (kSup * kTile); a.L = L; a.N = N; a.scale = scale;
  a.debug = getenv("PL_CTC_DEBUG") ? atoi(getenv("PL_CTC_DEBUG")) : 0;
  if (N % 64 == 0) return launch_ctc<64>(a, st);
  return launch_ctc<32>(a, st);
}

int ctc_timeouts() {
  int v = 0;
  cudaMemcpyFromSymbol(&v, g_ctc_timeouts, sizeof(int));
  return v;
}

}  // namespace pl
