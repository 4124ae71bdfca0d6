// pl_compat_tc.cu -- softmax + label-compatibility transform for LARGE label counts on the
// tcgen05 tensor cores (sm_100a):  P[p][j] = scale * sum_k Mu[l0 + j][k] * softmax_k(-E[p][.])
// for j < N output labels (N = L on one GPU, N = L / G when the labels are sharded over G ranks).
//
// crf/crf_module.py:99-102 with charb.forward (:71-75), pixel-major.  L is a multiple of 64
// (128, 256, ...), N a multiple of 32, N <= 256 (two TMEM accumulators of N columns).
//
// One persistent CTA per SM over tiles of 128 pixels (= UMMA M), four warp roles, everything
// double-buffered so that the tensor pipe never waits for a phase of another role:
//   warps 0-15  staging: per 32-label K chunk read the E chunk, Q' = exp(-E - rowmax) (the row sum
//               is accumulated on the side and applied in the epilogue: one exp per value), split
//               Q' = hi + lo (tf32) into the canonical K-major operand planes A[chunk & 1].  The row
//               maxima of the NEXT tile are computed in the gaps (one 64-label step after every
//               second chunk, loads issued a chunk ahead).
//   warp 20     one thread issues the MMAs: per 8-label K step three tcgen05.mma kind::tf32 of
//               128 x N x 8 (lo*hi + hi*lo + hi*hi, fp32 parity, see pl_fused_tc.cu) into the
//               128 x N fp32 accumulator acc[tile & 1] in TMEM.
//   warp 21     one thread streams the Mu records (pre-split hi / lo planes of Mu[:, 8 labels], one
//               contiguous record per K step) through a ring of bulk TMA copies, and pulls the next
//               tile's E rows into L2.
//   warps 16-19 epilogue of tile t while the MMAs of tile t+1 run: TMEM -> registers (x scale /
//               row sum) -> shared memory -> coalesced rows of P.
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
constexpr int kSoft = 512;                       // staging threads (warps 0-15)
constexpr int kEpi = 128;                        // epilogue threads (warps 16-19)
constexpr int kThreads = kSoft + kEpi + 64;      // + MMA warp + TMA warp
constexpr int kKC = 32;                          // labels per A chunk
constexpr int kKB = 16;                          // labels per Mu record (two MMA K steps)
constexpr int kA_CH = kTile * 16 + 16;           // bytes per 16-byte K chunk column of A
constexpr int kA_BYTES = (kKC / 4) * kA_CH;      // one plane of A: 16512
constexpr int kA_BUF = 2 * kA_BYTES;             // hi | lo
constexpr int kPLD = 64 + 4;                     // epilogue staging pitch (floats)
constexpr int kStageBytes = kTile * kPLD * 4;    // 34816
constexpr int kMaxRing = 8;                      // Mu records in flight (TMA ring slots), at most
constexpr int kRingBytes = 97 * 1024;            // ring budget: 3 slots at N = 256
constexpr int kMaxL = 512;

__device__ int g_ctc_timeouts = 0;

constexpr int kMaxBlocks = 16;

struct CArgs {
  const float* e;       // blocked [nblocks][n][block_w]
  // Optional: block b lives at blk[b] ([n][block_w]) instead of e + b * n * block_w.  The
  // pointers may be peer-GPU memory mapped through CUDA IPC (label shards read in place over
  // NVLink, no all-gathered copy); blk[0] == e.
  const float* blk[kMaxBlocks];
  int use_blk;
  int nblocks, block_w;
  const uint8_t* mu_rec;  // [L/8] records of rec_bytes(N)
  float* p;             // [n][N]
  // Optional: P is written in the same blocked layout instead, output label block b (pbw labels)
  // to pblk[b] ([n][pbw]); the pointers may be peer-GPU memory (each label owner receives its
  // columns of this rank's pixel rows directly, pl_softmax_compat_shards)
  float* pblk[kMaxBlocks];
  int use_pblk, pbw;
  float* q_out;         // [n][L] plain layout, or null
  int64_t n, ntiles;
  int L, N;
  int ring;             // ring slots
  float scale;
  int debug;  // PL_CTC_DEBUG, timing experiments only: 1 = no MMAs, 2 = no Mu copies, 4 = no statistics loads,
              // 8 = no P stores, 32 = no staging loads, 64 = no L2 prefetch, 128 = print per-role wait cycles
};

__host__ __device__ constexpr int b_ch(int N) { return N * 16 + 16; }       // one K chunk column of B
__host__ __device__ constexpr int rec_bytes(int N) { return 2 * (kKB / 4) * b_ch(N); }  // hi (4 columns) | lo (4 columns)

__device__ __forceinline__ void bar_soft() { asm volatile("bar.sync 1, 512;" ::: "memory"); }
__device__ __forceinline__ void bar_epi() { asm volatile("bar.sync 2, 128;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc::smem_u32(bar)) : "memory");
}
__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float tf32_cut(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }
__device__ __forceinline__ void wait_or_count(uint64_t* bar, uint32_t parity) {
  if (!tc::mbar_wait(bar, parity)) atomicAdd(&g_ctc_timeouts, 1);
}

#define TWAIT(acc, bar, par)                    \
  do {                                          \
    const long long _t0 = clock64();            \
    wait_or_count((bar), (par));                \
    (acc) += clock64() - _t0;                   \
  } while (0)

__global__ void __launch_bounds__(kThreads, 1) compat_tc_kernel(const CArgs a) {
  long long tw0 = 0, tw1 = 0, tw2 = 0;   // PL_CTC_DEBUG=128: cycles each role spends waiting
  const long long t_begin = clock64();
  extern __shared__ __align__(128) unsigned char smem[];
  const int REC = rec_bytes(a.N), BCH = b_ch(a.N);
  unsigned char* s_a = smem;                              // [2] A buffers: hi | lo planes
  float* s_p = reinterpret_cast<float*>(smem + 2 * kA_BUF);   // epilogue staging
  unsigned char* s_b = smem + 2 * kA_BUF + kStageBytes;   // ring of Mu records
  float* s_m = reinterpret_cast<float*>(s_b + a.ring * REC);   // [2][128] row max of -E (tile parity)
  float* s_inv = s_m + 2 * kTile;                              // [2][128] 1 / row sum (tile parity)
  int64_t* s_off = reinterpret_cast<int64_t*>(s_inv + 2 * kTile);   // [L/4] float4 group -> offset
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_off + kMaxL / 4);
  uint64_t* bar_full = bars;                  // [ring] Mu record landed
  uint64_t* bar_free = bars + kMaxRing;       // [ring] MMAs done with the record
  uint64_t* bar_afull = bars + 2 * kMaxRing;  // [2] A buffer written (one arrival per staging warp)
  uint64_t* bar_afree = bar_afull + 2;        // [2] MMAs done with the A buffer
  uint64_t* bar_acc = bar_afull + 4;          // [2] accumulator complete
  uint64_t* bar_accfree = bar_afull + 6;      // [2] epilogue has read the accumulator (128 arrivals)
  uint64_t* bar_inv = bar_afull + 8;          // [2] row sums of the tile written (kSoft arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_afull + 10);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int KCH = a.L / kKC;               // A chunks per tile
  const int steps_per_tile = a.L / kKB;    // Mu records per tile
  const uint32_t acc_cols = a.N <= 32 ? 32 : (a.N <= 64 ? 64 : (a.N <= 128 ? 128 : 256));
  const uint32_t tmem_cols = 2 * acc_cols;

  if (warp == 0) {
    if (lane == 0) {
      for (int b = 0; b < kMaxRing; b++) {
        tc::mbar_init(&bar_full[b], 1);
        tc::mbar_init(&bar_free[b], 1);
      }
      for (int b = 0; b < 2; b++) {
        tc::mbar_init(&bar_afull[b], kSoft / 32);   // one elected arrival per staging warp
        tc::mbar_init(&bar_afree[b], 1);
        tc::mbar_init(&bar_acc[b], 1);
        tc::mbar_init(&bar_accfree[b], kEpi);
        tc::mbar_init(&bar_inv[b], kSoft);
      }
      tc::fence_mbar_init();
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, tmem_cols);
    tc::tcgen05_fence_before();
  }
  // where the float4 group i (labels 4i .. 4i+3) lives, without the pixel term: the blocked
  // layout puts label (b * block_w + w) of pixel p at e[(b * n + p) * block_w + w]
  for (int i = tid; i < a.L / 4; i += kThreads) {
    const int label = 4 * i;
    const int b = label / a.block_w;
    s_off[i] = (a.use_blk ? (int64_t)(a.blk[b < kMaxBlocks ? b : 0] - a.e)
                          : (int64_t)b * a.n * a.block_w) + (label - b * a.block_w);
  }
  __syncthreads();
  tc::tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int64_t bw = a.block_w;
  const uint32_t my_tiles = (uint32_t)((a.ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x);

  if (warp == kSoft / 32 + 5) {
    // ---- TMA producer: every tile consumes the same sequence of Mu records, the ring cycles ----
    if (lane == 0) {
      const uint32_t total = my_tiles * (uint32_t)steps_per_tile;
      uint32_t rec = 0, slot = 0, phase = 1;   // phase: parity of the slot's PREVIOUS use (fresh barrier: 1 passes)
      int64_t next_tile = (int64_t)blockIdx.x + gridDim.x;
      const uint8_t* src_rec = a.mu_rec;
      for (uint32_t t = 0; t < total; t++) {
        if (rec == 0 && !a.use_blk && next_tile < a.ntiles && !(a.debug & 64)) {
          // pull the NEXT tile's E rows into L2: its statistics steps (spread over this tile) then
          // run at L2 latency.  Local memory only: peer blocks bypass the local L2.
          const int64_t q0 = next_tile * kTile;
          const uint32_t bytes = (uint32_t)(min((int64_t)kTile, a.n - q0) * bw * 4);
          for (int b = 0; b < a.nblocks; b++) {
            const char* src = reinterpret_cast<const char*>(a.e + ((int64_t)b * a.n + q0) * bw);
            for (uint32_t o = 0; o < bytes; o += 32768)
              asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src + o),
                           "r"(min(32768u, bytes - o))
                           : "memory");
          }
        }
        TWAIT(tw0, &bar_free[slot], phase);
        if (a.debug & 2) {
          mbar_arrive(&bar_full[slot]);
        } else {
          tc::mbar_arrive_expect_tx(&bar_full[slot], REC);
          tc::tma_load_1d(s_b + slot * REC, src_rec, REC, &bar_full[slot]);
        }
        src_rec += REC;
        if (++rec == (uint32_t)steps_per_tile) {
          rec = 0;
          src_rec = a.mu_rec;
          next_tile += gridDim.x;
        }
        if (++slot == (uint32_t)a.ring) {
          slot = 0;
          phase ^= 1;
        }
      }
      if ((a.debug & 128) && blockIdx.x == 0)
        printf("tma: total %lld wait_free %lld\n", clock64() - t_begin, tw0);
    }
  } else if (warp == kSoft / 32 + 4) {
    // ---- MMA issuer ------------------------------------------------------------------------
    if (lane == 0) {
      const uint32_t idesc = tc::idesc_tf32(kTile, a.N);
      // descriptors are built once; an operand further on is the same descriptor plus (bytes >> 4)
      // in the 14-bit address field (shared memory is < 256 KB: no carry out of the field)
      const uint64_t a_desc0 = tc::smem_desc_kmajor(tc::smem_u32(s_a), kA_CH, 128);
      const uint64_t b_desc0 = tc::smem_desc_kmajor(tc::smem_u32(s_b), BCH, 128);
      const uint64_t a_lo_off = kA_BYTES >> 4, a_buf_off = kA_BUF >> 4, a_k8 = (2 * kA_CH) >> 4;
      const uint64_t b_lo_off = ((kKB / 4) * BCH) >> 4, b_k8 = (2 * BCH) >> 4, b_slot = REC >> 4;
      uint32_t slot = 0, phase = 0, g = 0;    // ring slot + parity of its current use, A chunks consumed
      uint64_t b_cur = b_desc0;
      for (uint32_t t = 0; t < my_tiles; t++) {
        const uint32_t acc = tmem + (t & 1) * acc_cols;
        for (int kc = 0; kc < KCH; kc++, g++) {
          const uint32_t ab = g & 1;
          TWAIT(tw0, &bar_afull[ab], (g >> 1) & 1);
          if (kc == 0 && t >= 2) TWAIT(tw1, &bar_accfree[t & 1], ((t >> 1) - 1) & 1);
          tc::tcgen05_fence_after();
          const uint64_t a_cur = a_desc0 + ab * a_buf_off;
#pragma unroll
          for (int j = 0; j < kKC / kKB; j++) {
            TWAIT(tw2, &bar_full[slot], phase);
            tc::tcgen05_fence_after();
            if (!(a.debug & 1)) {
#pragma unroll
              for (int ks = 0; ks < kKB / 8; ks++) {
                const uint64_t ad = a_cur + (uint64_t)(j * (kKB / 8) + ks) * a_k8;
                const uint64_t bd = b_cur + (uint64_t)ks * b_k8;
                // lo*hi, hi*hi, hi*lo
                tc::mma_tf32_ss(acc, ad + a_lo_off, bd, idesc, (kc | j | ks) != 0);
                tc::mma_tf32_ss(acc, ad, bd, idesc, true);
                tc::mma_tf32_ss(acc, ad, bd + b_lo_off, idesc, true);
              }
            }
            tc::mma_commit(&bar_free[slot]);   // ring slot free once these MMAs have read it
            b_cur += b_slot;
            if (++slot == (uint32_t)a.ring) {
              slot = 0;
              phase ^= 1;
              b_cur = b_desc0;
            }
          }
          tc::mma_commit(&bar_afree[ab]);      // A buffer free
        }
        tc::mma_commit(&bar_acc[t & 1]);       // accumulator of this tile complete
      }
      if ((a.debug & 128) && blockIdx.x == 0)
        printf("mma: total %lld wait_afull %lld wait_accfree %lld wait_bfull %lld tiles %u\n",
               clock64() - t_begin, tw0, tw1, tw2, my_tiles);
    }
  } else if (warp >= kSoft / 32) {
    // ---- warps 16-19: epilogue of tile t under the MMAs of tile t+1 ----------------------------
    const int et = tid - kSoft;               // 0..127 = TMEM lane = pixel row of the tile (warp 16 + q -> lanes 32q..)
    const int g = et >> 4, l = et & 15;       // copy-out: 16 lanes per row, 8 rows per pass
    uint32_t t = 0;
    for (int64_t tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, t++) {
      const int64_t p0 = tile * kTile;
      const int rows = (int)min((int64_t)kTile, a.n - p0);
      const uint32_t par = (t >> 1) & 1;
      TWAIT(tw0, &bar_inv[t & 1], par);
      const float* inv = s_inv + (t & 1) * kTile;
      if (a.q_out != nullptr) {   // Q = exp / sum: rescale the rows written during staging (L2-hot)
        for (int r = g; r < rows; r += kEpi / 16) {
          const float ir = inv[r];
          for (int c = 4 * l; c < a.L; c += 64) {
            float4* qp = reinterpret_cast<float4*>(a.q_out + (p0 + r) * a.L + c);
            float4 x = *qp;
            x.x *= ir; x.y *= ir; x.z *= ir; x.w *= ir;
            *qp = x;
          }
        }
      }
      const float rscale = a.scale * inv[et];
      TWAIT(tw1, &bar_acc[t & 1], par);
      tc::tcgen05_fence_after();
      const uint32_t acc = tmem + (t & 1) * acc_cols + ((uint32_t)((warp & 3) * 32) << 16);
      for (int c0 = 0; c0 < a.N; c0 += 64) {
        const int cw = min(64, a.N - c0);          // 32 or 64
        for (int c16 = 0; c16 < cw; c16 += 16) {
          float v[16];
          tc::tmem_ld_x16(acc + (uint32_t)(c0 + c16), v);
#pragma unroll
          for (int j = 0; j < 4; j++)
            *reinterpret_cast<float4*>(s_p + et * kPLD + c16 + 4 * j) =
                make_float4(v[4 * j] * rscale, v[4 * j + 1] * rscale, v[4 * j + 2] * rscale,
                            v[4 * j + 3] * rscale);
        }
        if (c0 + 64 >= a.N) {      // last TMEM read of this tile: the accumulator may be overwritten
          tc::tcgen05_fence_before();
          mbar_arrive(&bar_accfree[t & 1]);
        }
        bar_epi();
        if (4 * l < cw && !(a.debug & 8)) {
          if (a.use_pblk) {
            const int c = c0 + 4 * l, b = c / a.pbw;
            float* base = a.pblk[b] + (c - b * a.pbw);
#pragma unroll 8
            for (int r = g; r < rows; r += kEpi / 16)
              *reinterpret_cast<float4*>(base + (p0 + r) * a.pbw) =
                  *reinterpret_cast<const float4*>(s_p + r * kPLD + 4 * l);
          } else {
#pragma unroll 8
            for (int r = g; r < rows; r += kEpi / 16)
              *reinterpret_cast<float4*>(a.p + (p0 + r) * a.N + c0 + 4 * l) =
                  *reinterpret_cast<const float4*>(s_p + r * kPLD + 4 * l);
          }
        }
        bar_epi();   // staging free again
      }
    }
    if ((a.debug & 128) && blockIdx.x == 0 && et == 0)
      printf("epi: total %lld wait_inv %lld wait_acc %lld\n", clock64() - t_begin, tw0, tw1);
  } else {
    // ---- warps 0-15: operand staging + the next tile's row maxima -------------------------------
    const int g = tid >> 4, l = tid & 15;     // statistics: 16 lanes per row, 32 rows per pass
    const int g2 = tid >> 3, l2 = tid & 7;    // staging: 8 lanes per row of a 32-label chunk, 64 rows per pass
    constexpr int RPT1 = kTile / (kSoft / 16);   // 8
    constexpr int RPT2 = kTile / (kSoft / 8);    // 4
    const int SS = (a.debug & 4) ? 0 : a.L / 64;   // statistics steps per tile

    auto stat_load = [&](int64_t p0, int rows, int s, float4* e) {
      const float* base = a.e + s_off[s * 16 + l] + p0 * bw;
#pragma unroll
      for (int i = 0; i < RPT1; i++) {
        const int r = g + i * (kSoft / 16);
        e[i] = (r < rows) ? *reinterpret_cast<const float4*>(base + r * bw)
                          : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    };
    auto stat_max = [&](const float4* e, float* m) {
#pragma unroll
      for (int i = 0; i < RPT1; i++)
        m[i] = fmaxf(m[i], fmaxf(fmaxf(-e[i].x, -e[i].y), fmaxf(-e[i].z, -e[i].w)));
    };
    auto stat_store = [&](const float* m, int buf) {
#pragma unroll
      for (int i = 0; i < RPT1; i++) {
        float mi = m[i];
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) mi = fmaxf(mi, __shfl_xor_sync(0xffffffffu, mi, o, 16));
        if (l == 0) s_m[buf * kTile + g + i * (kSoft / 16)] = mi;
      }
    };

    // the first tile's row maxima, up front
    if (blockIdx.x < a.ntiles) {
      const int64_t p0 = (int64_t)blockIdx.x * kTile;
      const int rows = (int)min((int64_t)kTile, a.n - p0);
      float m[RPT1];
#pragma unroll
      for (int i = 0; i < RPT1; i++) m[i] = -INFINITY;
      for (int s = 0; s < SS; s++) {
        float4 e[RPT1];
        stat_load(p0, rows, s, e);
        stat_max(e, m);
      }
      stat_store(m, 0);
    }
    bar_soft();

    float4 ec[RPT2], en[RPT2];   // E chunks kc and kc+1 of the running tile (carried across tiles)
    uint32_t gch = 0, t = 0;     // A chunks produced, tiles done
    for (int64_t tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x, t++) {
      const int64_t p0 = tile * kTile;
      const int rows = (int)min((int64_t)kTile, a.n - p0);
      const int buf = t & 1;
      const int64_t np0 = (tile + gridDim.x) * kTile;
      const bool has_next = tile + gridDim.x < a.ntiles;
      const int nrows = has_next ? (int)min((int64_t)kTile, a.n - np0) : 0;

      float rm[RPT2], rsum[RPT2], mnext[RPT1];
#pragma unroll
      for (int i = 0; i < RPT2; i++) {
        rm[i] = -1.4426950408889634f * s_m[buf * kTile + g2 + i * (kSoft / 8)];   // exp(-e - m) = 2^(-e log2e - m log2e)
        rsum[i] = 0.f;
      }
#pragma unroll
      for (int i = 0; i < RPT1; i++) mnext[i] = -INFINITY;

      float4 es[RPT1];
      // E chunk kc of the tile at q0 (this tile, or the next one near the end of this tile)
      auto load_chunk = [&](int64_t q0, int qrows, int kc, float4* e) {
        const float* base = a.e + s_off[kc * 8 + l2] + q0 * bw;
#pragma unroll
        for (int i = 0; i < RPT2; i++) {
          const int r = g2 + i * (kSoft / 8);
          e[i] = (r < qrows && !(a.debug & 32)) ? *reinterpret_cast<const float4*>(base + r * bw)
                                                 : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      };
      if (t == 0) {   // the first two chunks of the CTA's first tile; later ones are requested below
        load_chunk(p0, rows, 0, ec);
        load_chunk(p0, rows, 1, en);
      }
      for (int kc = 0; kc < KCH; kc++, gch++) {
        const uint32_t ab = gch & 1;
        if (gch >= 2) TWAIT(tw0, &bar_afree[ab], ((gch >> 1) - 1) & 1);
        unsigned char* abuf = s_a + ab * kA_BUF;
#pragma unroll
        for (int i = 0; i < RPT2; i++) {
          const int r = g2 + i * (kSoft / 8);
          float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
          if (r < rows) {
            x.x = ex2_fast(fmaf(ec[i].x, -1.4426950408889634f, rm[i]));
            x.y = ex2_fast(fmaf(ec[i].y, -1.4426950408889634f, rm[i]));
            x.z = ex2_fast(fmaf(ec[i].z, -1.4426950408889634f, rm[i]));
            x.w = ex2_fast(fmaf(ec[i].w, -1.4426950408889634f, rm[i]));
            rsum[i] += (x.x + x.y) + (x.z + x.w);
            if (a.q_out != nullptr)   // unnormalised here, scaled by 1 / sum by the epilogue warps
              *reinterpret_cast<float4*>(a.q_out + (p0 + r) * a.L + kc * kKC + 4 * l2) = x;
          }
          float4 hi, lo;
          // hi = x cut to tf32 (the MMA reads only the top 19 bits anyway), lo = x - hi exactly
          hi.x = tf32_cut(x.x); lo.x = x.x - hi.x;
          hi.y = tf32_cut(x.y); lo.y = x.y - hi.y;
          hi.z = tf32_cut(x.z); lo.z = x.z - hi.z;
          hi.w = tf32_cut(x.w); lo.w = x.w - hi.w;
          unsigned char* pa = abuf + l2 * kA_CH + r * 16;
          *reinterpret_cast<float4*>(pa) = hi;
          *reinterpret_cast<float4*>(pa + kA_BYTES) = lo;
        }
        if ((kc & 1) && has_next && (kc >> 1) < SS) stat_max(es, mnext);   // requested a chunk ago
        // the proxy fence waits for this thread's outstanding loads too: the only ones in flight here
        // are chunk kc+1's, requested a whole chunk ago
        tc::fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_afull[ab]);
#pragma unroll
        for (int i = 0; i < RPT2; i++) ec[i] = en[i];
        // requests: E chunk kc+2 (two chunks of lookahead, maybe the next tile's) and, every second
        // chunk, one 64-label step of the next tile's row maxima
        if (kc + 2 < KCH) {
          load_chunk(p0, rows, kc + 2, en);
        } else if (has_next) {
          load_chunk(np0, nrows, kc + 2 - KCH, en);
        }
        if (has_next && !(kc & 1) && (kc >> 1) < SS) stat_load(np0, nrows, kc >> 1, es);
      }
      // row sums of this tile -> epilogue warps; row maxima of the next tile -> ourselves
#pragma unroll
      for (int i = 0; i < RPT2; i++) {
        float si = rsum[i];
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) si += __shfl_xor_sync(0xffffffffu, si, o, 8);
        if (l2 == 0) s_inv[buf * kTile + g2 + i * (kSoft / 8)] = (g2 + i * (kSoft / 8) < rows) ? __fdividef(1.0f, si) : 0.f;
      }
      mbar_arrive(&bar_inv[buf]);
      if (has_next) stat_store(mnext, buf ^ 1);
      {
        const long long _t0 = clock64();
        bar_soft();
        tw2 += clock64() - _t0;
      }
    }
    if ((a.debug & 128) && blockIdx.x == 0 && tid == 0)
      printf("stage: total %lld wait_afree %lld wait_bar %lld\n", clock64() - t_begin, tw0, tw2);
  }

  __syncthreads();
  if (warp == 0) {
    tc::tcgen05_fence_after();
    tc::tmem_dealloc(tmem, tmem_cols);
  }
}

// Mu[l0 + j][kb*16 + k] = muT[(kb*16 + k) * L + l0 + j]  ->  record kb: hi plane | lo plane, each
// four 16-byte K chunk columns of N rows (+ 16 bytes of pad) in the canonical K-major layout.
__global__ void __launch_bounds__(256)
mu_records_kernel(const float* __restrict__ muT, int L, int l0, int N, uint8_t* __restrict__ out) {
  const int nrec = L / kKB;
  const int chunk_floats = b_ch(N) / 4;
  const int plane_floats = (kKB / 4) * chunk_floats;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nrec * plane_floats;
       i += gridDim.x * blockDim.x) {
    const int kb = i / plane_floats, w = i - kb * plane_floats;
    const int ch = w / chunk_floats, rem = w - ch * chunk_floats;
    float hi = 0.f, lo = 0.f;
    if (rem < N * 4) {
      const int j = rem >> 2, k = kb * kKB + ch * 4 + (rem & 3);
      const float m = muT[(int64_t)k * L + l0 + j];
      hi = tc::tf32_rna(m);
      lo = tc::tf32_rna(m - hi);
    }
    float* base = reinterpret_cast<float*>(out + (size_t)kb * rec_bytes(N));
    base[w] = hi;
    base[plane_floats + w] = lo;
  }
}

int ring_slots(int N) {
  const int r = kRingBytes / rec_bytes(N);
  return r < kMaxRing ? r : kMaxRing;
}

size_t ctc_smem(int N) {
  return 2 * (size_t)kA_BUF + kStageBytes + (size_t)ring_slots(N) * rec_bytes(N) + 4 * kTile * 4 +
         (kMaxL / 4) * 8 + (2 * kMaxRing + 10) * 8 + 16;
}

}  // namespace

// Output labels l0 .. l0+N of a model with L labels.  Supported: L % 64 == 0, N % 32 == 0,
// 32 <= N <= 256, and the blocked input layout must not split a float4 (block_w % 4 == 0).
bool compat_tc_supported(int L, int N, int block_w) {
  return L >= 64 && L <= kMaxL && L % 64 == 0 && N >= 32 && N <= 256 && N % 32 == 0 &&
         block_w % 4 == 0 && L % block_w == 0;
}

size_t compat_tc_mu_bytes(int L, int N) { return (size_t)(L / kKB) * rec_bytes(N); }

int compat_tc_prepare_mu(const float* muT, int L, int l0, int N, uint8_t* mu_rec, cudaStream_t st) {
  PL_LAUNCH(mu_records_kernel, 128, 256, 0, st, muT, L, l0, N, mu_rec);
  return PL_OK;
}

int compat_tc_run(const float* e, int nblocks, int block_w, const uint8_t* mu_rec, float* p,
                  float* q_out, int64_t n, int L, int N, float scale, cudaStream_t st,
                  const float* const* block_ptrs, float* const* out_block_ptrs) {
  CArgs a;
  a.use_blk = 0;
  a.use_pblk = 0;
  a.pbw = block_w;
  for (int b = 0; b < kMaxBlocks; b++) a.pblk[b] = nullptr;
  if (out_block_ptrs != nullptr) {
    if (nblocks > kMaxBlocks || N != L) {
      set_last_error("blocked output needs N == L and at most %d label blocks", kMaxBlocks);
      return PL_ERR_INVALID;
    }
    for (int b = 0; b < nblocks; b++) a.pblk[b] = out_block_ptrs[b];
    a.use_pblk = 1;
  }
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
  a.n = n; a.ntiles = (n + kTile - 1) / kTile; a.L = L; a.N = N; a.scale = scale;
  a.ring = ring_slots(N);
#ifdef PL_DEBUG_KNOBS
  // timing experiments only (builds with -DPL_DEBUG_KNOBS): every bit but 128 skips work and
  // CORRUPTS the result, so the release library does not read the variable at all
  a.debug = getenv("PL_CTC_DEBUG") ? atoi(getenv("PL_CTC_DEBUG")) : 0;
#else
  a.debug = 0;
#endif
  const size_t smem = ctc_smem(N);
  static PerDeviceOnce done;
  if (done.need()) {
    PL_CUDA(cudaFuncSetAttribute(compat_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 210 * 1024));   // one CTA per SM; ctc_smem(N) <= 205 KB for every N
  }
  const int64_t cap = PL_SM_COUNT;
  const unsigned grid = (unsigned)(a.ntiles < cap ? a.ntiles : cap);
  PL_LAUNCH(compat_tc_kernel, grid, kThreads, smem, st, a);
  return PL_OK;
}

int ctc_timeouts() {
  int v = 0;
  cudaMemcpyFromSymbol(&v, g_ctc_timeouts, sizeof(int));
  return v;
}

}  // namespace pl
