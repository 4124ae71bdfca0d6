// pl_tiles.cu -- tile-local view of the lattice and the tiled splat / slice kernels (sm_100a).
//
// Neighbouring pixels share lattice vertices (a 128-pixel run of a natural
// image touches a few dozen distinct vertices, not 768).  The flat kernels in
// pl_filter.cu ignore that: splat sends every (pixel, vertex) contribution to
// L2 as its own atomic, slice gathers every row through L1/L2.  Here the
// lattice build additionally produces, per tile of kTile consecutive points,
//   * its entries (point, remainder) sorted by vertex id  -> ent_pix / ent_w
//   * the list of distinct vertices with their segment in that order -> useg
//   * per entry the index of its vertex in that list                -> loc
// so that
//   splat : stage the tile's src rows in shared memory, sum each vertex's
//           segment in registers (fixed order: deterministic within a tile),
//           ONE red.global.add.v4 per (tile, vertex, float4)   [no smem atomics]
//   slice : stage the tile's distinct vertex rows in shared memory once, then
//           every pixel reads its d+1 rows from there.
// Feature sets without locality (U close to kTile*(d+1)) simply fall back to
// per-entry global access inside the same kernels.
#include <math.h>
#include <stdlib.h>

#include <atomic>
#include <mutex>

#include "pl_common.cuh"

namespace pl {

namespace {

constexpr int kTile = PL_TILE;          // points per tile
constexpr int kThreads = 256;
constexpr unsigned long long kPad = 0xFFFFFFFFFFFFFFFFull;

// ---------------------------------------------------------------------------
// tile build: one CTA per tile
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
tile_build_kernel(const int32_t* __restrict__ vid, const float* __restrict__ bary, int64_t n,
                  int d1, int NS, uint8_t* __restrict__ ent_pix, float* __restrict__ ent_w,
                  uint16_t* __restrict__ loc, int2* __restrict__ useg, int* __restrict__ tile_off,
                  int* __restrict__ tile_U, int* __restrict__ counter,
                  uint8_t* __restrict__ rec_slice, uint8_t* __restrict__ rec_splat) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem_raw);   // [NS]
  unsigned short* ustart = reinterpret_cast<unsigned short*>(keys + NS);        // [TE + 1]
  __shared__ int warp_cnt[kThreads / 32];
  __shared__ int s_off;

  const int tid = threadIdx.x;
  const int TE = kTile * d1;
  const int64_t tile = blockIdx.x;
  const int64_t e0 = tile * TE;
  const int64_t E = n * d1;
  // packed per-tile records (see pl_common.cuh); pads (k >= valid) are written as zeros
  float* rs_bary = reinterpret_cast<float*>(rec_slice + tile * (int64_t)TE * 6);
  uint16_t* rs_loc = reinterpret_cast<uint16_t*>(rs_bary + TE);
  float* rp_w = reinterpret_cast<float*>(rec_splat + tile * (int64_t)TE * 9);
  int32_t* rp_vid = reinterpret_cast<int32_t*>(rp_w + TE);
  uint8_t* rp_pix = reinterpret_cast<uint8_t*>(rp_vid + TE);

  for (int k = tid; k < NS; k += kThreads) {
    unsigned long long key = kPad;
    if (k < TE && e0 + k < E) key = ((unsigned long long)(uint32_t)vid[e0 + k] << 16) | (unsigned)k;
    keys[k] = key;
  }
  __syncthreads();

  // bitonic sort, ascending (pads last; ties impossible: k is part of the key)
  for (int k = 2; k <= NS; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < NS; i += kThreads) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long a = keys[i], b = keys[ixj];
          const bool up = (i & k) == 0;
          if ((a > b) == up) {
            keys[i] = b;
            keys[ixj] = a;
          }
        }
      }
      __syncthreads();
    }
  }

  // head flags + exclusive scan: each thread owns PER consecutive sorted slots
  const int PER = NS / kThreads;  // NS is a power of two >= kThreads
  const int base = tid * PER;
  int heads = 0;
  for (int q = 0; q < PER; q++) {
    const int k = base + q;
    const unsigned long long key = keys[k];
    if (key != kPad && (k == 0 || (keys[k - 1] >> 16) != (key >> 16))) heads++;
  }
  int incl = heads;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, incl, o);
    if ((tid & 31) >= o) incl += y;
  }
  if ((tid & 31) == 31) warp_cnt[tid >> 5] = incl;
  __syncthreads();
  int before = 0, U = 0;
#pragma unroll
  for (int w = 0; w < kThreads / 32; w++) {
    const int c = warp_cnt[w];
    before += (w < (tid >> 5)) ? c : 0;
    U += c;
  }
  int u = before + incl - heads;  // distinct vertices before this thread's slots

  int valid = 0;  // number of real entries in this tile
  {
    const int64_t rem = E - e0;
    valid = rem < TE ? (int)rem : TE;
  }
  for (int i = valid + tid; i < TE; i += kThreads) {
    rs_bary[i] = 0.f;
    rs_loc[i] = 0;
  }
  for (int q = 0; q < PER; q++) {
    const int k = base + q;
    const unsigned long long key = keys[k];
    if (key == kPad) {
      if (k < TE) {
        rp_w[k] = 0.f;
        rp_vid[k] = 0;
        rp_pix[k] = 0;
      }
      continue;
    }
    const bool head = (k == 0 || (keys[k - 1] >> 16) != (key >> 16));
    if (head) {
      ustart[u] = (unsigned short)k;
      u++;
    }
    const int orig = (int)(key & 0xFFFFu);
    const float wgt = bary[e0 + orig];
    ent_pix[e0 + k] = (uint8_t)(orig / d1);
    ent_w[e0 + k] = wgt;
    loc[e0 + orig] = (uint16_t)(u - 1);
    rp_w[k] = wgt;
    rp_vid[k] = (int32_t)(key >> 16);
    rp_pix[k] = (uint8_t)(orig / d1);
    rs_bary[orig] = wgt;
    rs_loc[orig] = (uint16_t)(u - 1);
  }
  if (tid == 0) {
    ustart[U] = (unsigned short)valid;
    s_off = atomicAdd(counter, U);
    atomicMax(counter + 1, U);   // largest tile: sizes the slice kernel's staging buffer
    tile_off[tile] = s_off;
    tile_U[tile] = U;
  }
  __syncthreads();
  const int off = s_off;
  for (int v = tid; v < U; v += kThreads) {
    const int st = ustart[v], en = ustart[v + 1];
    useg[off + v] = make_int2((int)(keys[st] >> 16), st | ((en - st) << 16));
  }
}

// ---------------------------------------------------------------------------
// tile build without a sort (TE <= 1024, i.e. d <= 7): one CTA per tile
//   1. distinct vertices of the tile via a small open-addressing set in shared memory
//   2. local index = rank of the occupied slot
//   3. stable counting sort of the entries by local index: 32-entry chunks, in-chunk rank
//      from match.any, per-(chunk, vertex) counts prefix-summed over the chunks
// Entries of one vertex keep their (pixel, remainder) order, so the tile-internal
// summation order of the splat stays fixed from run to run.
// ---------------------------------------------------------------------------
constexpr int kHashSlots = 2048;
constexpr int kFastMaxTE = 1024;

__global__ void __launch_bounds__(kThreads)
tile_build_fast_kernel(const int32_t* __restrict__ vid, const float* __restrict__ bary, int64_t n,
                       int d1, uint8_t* __restrict__ ent_pix, float* __restrict__ ent_w,
                       uint16_t* __restrict__ loc, int2* __restrict__ useg,
                       int* __restrict__ tile_off, int* __restrict__ tile_U,
                       int* __restrict__ counter, uint8_t* __restrict__ rec_slice,
                       uint8_t* __restrict__ rec_splat) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // per-(32-entry chunk, vertex) counts, <= 32 each: bytes (18 KB instead of 37 KB as 16-bit prefix sums,
  // 5 CTAs per SM instead of 3); an entry adds up the counts of the chunks before its own when it scatters
  unsigned char* s_cnt = smem_raw;                                      // [TE][4 * NW]: a vertex's chunk counts are contiguous
  __shared__ int h_key[kHashSlots];
  __shared__ unsigned short h_lidx[kHashSlots];
  __shared__ int s_uvid[kFastMaxTE];
  __shared__ unsigned short s_start[kFastMaxTE], s_total[kFastMaxTE];
  __shared__ int warp_cnt[kThreads / 32];
  __shared__ int s_off;

  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int TE = kTile * d1;
  const int64_t tile = blockIdx.x;
  const int64_t e0 = tile * TE;
  const int64_t E = n * d1;
  const int valid = (int)((E - e0) < TE ? (E - e0) : TE);
  constexpr int QMAX = kFastMaxTE / kThreads;  // entries per thread
  const int nq = (TE + kThreads - 1) / kThreads;

  float* rs_bary = reinterpret_cast<float*>(rec_slice + tile * (int64_t)TE * 6);
  uint16_t* rs_loc = reinterpret_cast<uint16_t*>(rs_bary + TE);
  float* rp_w = reinterpret_cast<float*>(rec_splat + tile * (int64_t)TE * 9);
  int32_t* rp_vid = reinterpret_cast<int32_t*>(rp_w + TE);
  uint8_t* rp_pix = reinterpret_cast<uint8_t*>(rp_vid + TE);

  for (int i = tid; i < kHashSlots; i += kThreads) h_key[i] = -1;
  int my_vid[QMAX];
  float my_w[QMAX];
#pragma unroll
  for (int q = 0; q < QMAX; q++) {
    const int e = tid + q * kThreads;
    my_vid[q] = (q < nq && e < valid) ? vid[e0 + e] : -1;
    my_w[q] = (q < nq && e < valid) ? bary[e0 + e] : 0.f;
  }
  __syncthreads();

  // 1. insert; remember the slot
  int my_slot[QMAX];
#pragma unroll
  for (int q = 0; q < QMAX; q++) {
    my_slot[q] = 0;
    if (my_vid[q] >= 0) {
      unsigned h = ((unsigned)my_vid[q] * 2654435761u) >> 21;  // top 11 bits
      for (;;) {
        const int old = atomicCAS(&h_key[h], -1, my_vid[q]);
        if (old == -1 || old == my_vid[q]) break;
        h = (h + 1) & (kHashSlots - 1);
      }
      my_slot[q] = (int)h;
    }
  }
  __syncthreads();

  // 2. local index = rank of the occupied slot (8 consecutive slots per thread)
  int U;
  {
    constexpr int SPT = kHashSlots / kThreads;
    int occ = 0;
#pragma unroll
    for (int k = 0; k < SPT; k++) occ += h_key[tid * SPT + k] >= 0 ? 1 : 0;
    int incl = occ;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    if (lane == 31) warp_cnt[wid] = incl;
    __syncthreads();
    int before = 0;
    U = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; w++) {
      const int c = warp_cnt[w];
      before += (w < wid) ? c : 0;
      U += c;
    }
    int u = before + incl - occ;
#pragma unroll
    for (int k = 0; k < SPT; k++) {
      const int key = h_key[tid * SPT + k];
      if (key >= 0) {
        h_lidx[tid * SPT + k] = (unsigned short)u;
        s_uvid[u] = key;
        u++;
      }
    }
  }
  const int nchunks = (TE + 31) / 32;
  const int NW = (nchunks + 3) / 4;            // 32-bit words of counts per vertex (6 for d = 5)
  uint32_t* s_cnt32 = reinterpret_cast<uint32_t*>(s_cnt);
  for (int i = tid; i < U * NW; i += kThreads) s_cnt32[i] = 0u;
  __syncthreads();

  // 3. in-chunk stable rank + per-(chunk, vertex) counts
  int my_u[QMAX], my_r[QMAX];
#pragma unroll
  for (int q = 0; q < QMAX; q++) {
    my_u[q] = 0xFFFF;
    my_r[q] = 0;
    if (q < nq) {  // uniform: whole warps take part in match.any
      const int u = my_vid[q] >= 0 ? (int)h_lidx[my_slot[q]] : 0xFFFF;
      const unsigned peers = __match_any_sync(0xffffffffu, u);
      my_u[q] = u;
      my_r[q] = __popc(peers & ((1u << lane) - 1u));
      const int chunk = wid + q * (kThreads / 32);
      if (u != 0xFFFF && my_r[q] == 0) s_cnt[u * (4 * NW) + chunk] = (unsigned char)__popc(peers);
    }
  }
  __syncthreads();
  // entries per vertex
  for (int u = tid; u < U; u += kThreads) {
    unsigned running = 0;
    for (int wd = 0; wd < NW; wd++) running += __vsadu4(s_cnt32[u * NW + wd], 0u);   // sum of the four bytes
    s_total[u] = (unsigned short)running;
  }
  __syncthreads();
  // exclusive scan of the per-vertex totals (U <= 1024: 4 per thread)
  {
    int t[QMAX], sum = 0;
#pragma unroll
    for (int k = 0; k < QMAX; k++) {
      const int u = tid * QMAX + k;
      t[k] = u < U ? (int)s_total[u] : 0;
      sum += t[k];
    }
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += y;
    }
    __syncthreads();  // warp_cnt reuse
    if (lane == 31) warp_cnt[wid] = incl;
    __syncthreads();
    int before = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; w++) before += (w < wid) ? warp_cnt[w] : 0;
    int run = before + incl - sum;
#pragma unroll
    for (int k = 0; k < QMAX; k++) {
      const int u = tid * QMAX + k;
      if (u < U) s_start[u] = (unsigned short)run;
      run += t[k];
    }
  }
  if (tid == 0) {
    s_off = atomicAdd(counter, U);
    atomicMax(counter + 1, U);   // largest tile: sizes the slice kernel's staging buffer
    tile_off[tile] = s_off;
    tile_U[tile] = U;
  }
  __syncthreads();

  // 4. scatter
#pragma unroll
  for (int q = 0; q < QMAX; q++) {
    const int e = tid + q * kThreads;
    if (q < nq && e < TE) {
      if (my_vid[q] >= 0) {
        const int u = my_u[q];
        const int chunk = wid + q * (kThreads / 32);
        unsigned before = 0;                          // entries of this vertex in earlier chunks
        for (int wd = 0; wd < NW; wd++) {
          const int nb = chunk - 4 * wd;              // bytes of this word that belong to earlier chunks
          if (nb <= 0) break;
          const uint32_t m = nb >= 4 ? 0xFFFFFFFFu : ((1u << (8 * nb)) - 1u);
          before += __vsadu4(s_cnt32[u * NW + wd] & m, 0u);
        }
        const int pos = (int)s_start[u] + (int)before + my_r[q];
        const uint8_t pix = (uint8_t)(e / d1);
        ent_pix[e0 + pos] = pix;
        ent_w[e0 + pos] = my_w[q];
        loc[e0 + e] = (uint16_t)u;
        rp_w[pos] = my_w[q];
        rp_vid[pos] = my_vid[q];
        rp_pix[pos] = pix;
        rs_bary[e] = my_w[q];
        rs_loc[e] = (uint16_t)u;
      } else {  // pads of the last tile: weight 0, vertex 0
        rp_w[e] = 0.f;
        rp_vid[e] = 0;
        rp_pix[e] = 0;
        rs_bary[e] = 0.f;
        rs_loc[e] = 0;
      }
    }
  }
  const int off = s_off;
  for (int u = tid; u < U; u += kThreads)
    useg[off + u] = make_int2(s_uvid[u], (int)s_start[u] | ((int)s_total[u] << 16));
}

int pow2_ceil_i(int x) {
  int p = 1;
  while (p < x) p <<= 1;
  return p;
}

}  // namespace

namespace {

}  // namespace

// pinned two-int slots for small device -> host results: max U of the tile build (read lazily: a slot
// recycled 2048 lattices later can only cost speed, the kernels decide staged / unstaged per tile from
// the tile's own U) and M of the lattice build (read before the build returns)
int* pinned_slot() {
  static int* ring = nullptr;
  static std::once_flag once;
  static std::atomic<unsigned> next{0};
  constexpr unsigned kSlots = 4096;
  std::call_once(once, [] {
    if (cudaHostAlloc((void**)&ring, kSlots * sizeof(int), cudaHostAllocPortable) != cudaSuccess) {
      (void)cudaGetLastError();
      ring = nullptr;
    }
  });
  // two ints per lattice: {sum of U over the tiles, max U}
  return ring == nullptr ? nullptr : ring + (next.fetch_add(2, std::memory_order_relaxed) % kSlots);
}

namespace {

}  // namespace

int readback_small(Lattice* L, void* host_dst, const void* dev_src, size_t bytes, cudaStream_t st) {
  static cudaStream_t side[64] = {};
  static std::mutex mu;
  int dev = 0;
  PL_CUDA(cudaGetDevice(&dev));
  cudaStream_t ss = nullptr;
  if (dev >= 0 && dev < 64) {
    std::lock_guard<std::mutex> lock(mu);
    if (side[dev] == nullptr) PL_CUDA(cudaStreamCreateWithFlags(&side[dev], cudaStreamNonBlocking));
    ss = side[dev];
  }
  if (L->max_U_ev == nullptr) PL_CUDA(cudaEventCreateWithFlags(&L->max_U_ev, cudaEventDisableTiming));
  if (ss == nullptr) {   // no side stream: in the caller's stream
    PL_CUDA(cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, st));
    PL_CUDA(cudaEventRecord(L->max_U_ev, st));
    return PL_OK;
  }
  if (L->aux_ev == nullptr) PL_CUDA(cudaEventCreateWithFlags(&L->aux_ev, cudaEventDisableTiming));
  PL_CUDA(cudaEventRecord(L->aux_ev, st));
  PL_CUDA(cudaStreamWaitEvent(ss, L->aux_ev, 0));
  PL_CUDA(cudaMemcpyAsync(host_dst, dev_src, bytes, cudaMemcpyDeviceToHost, ss));
  PL_CUDA(cudaEventRecord(L->max_U_ev, ss));
  return PL_OK;
}

namespace {
int tiles_max_u(Lattice* L, cudaStream_t st) {
  int* d_max = L->tile_off + 2 * L->ntiles;     // {useg cursor = sum of U, max U}
  L->max_U = -1;
  L->h_max_U = pinned_slot();
  if (L->h_max_U == nullptr) {
    (void)cudaGetLastError();
    L->max_U = 1 << 30;   // unknown: largest staging buffer
    return PL_OK;
  }
  return readback_small(L, L->h_max_U, d_max, 2 * sizeof(int), st);
}
}  // namespace

int lattice_build_tiles(Lattice* L, cudaStream_t st) {
  const int d1 = L->d + 1;
  const int TE = kTile * d1;
  const int64_t ntiles = (L->n + kTile - 1) / kTile;
  int NS = pow2_ceil_i(TE);
  if (NS < kThreads) NS = kThreads;
  const size_t E_pad = (size_t)ntiles * TE;
  PL_TRY(dev_alloc((void**)&L->ent_pix, E_pad, st));
  PL_TRY(dev_alloc((void**)&L->ent_w, sizeof(float) * E_pad, st));
  PL_TRY(dev_alloc((void**)&L->loc, sizeof(uint16_t) * E_pad, st));
  PL_TRY(dev_alloc((void**)&L->useg, sizeof(int2) * E_pad, st));
  PL_TRY(dev_alloc((void**)&L->tile_off, sizeof(int) * (size_t)(2 * ntiles + 2), st));
  L->tile_U = L->tile_off + ntiles;
  int* counter = L->tile_off + 2 * ntiles;
  L->ntiles = ntiles;
  PL_TRY(dev_alloc((void**)&L->rec_slice, E_pad * 6, st));
  PL_TRY(dev_alloc((void**)&L->rec_splat, E_pad * 9, st));
  L->bytes += (int64_t)(E_pad * (1 + 4 + 2 + 8 + 6 + 9) + sizeof(int) * (size_t)(2 * ntiles + 1));
  PL_CUDA(cudaMemsetAsync(counter, 0, 2 * sizeof(int), st));   // useg cursor, max U
  if (TE <= kFastMaxTE) {
    const size_t smem = (size_t)TE * 4 * ((((TE + 31) / 32) + 3) / 4);
    // static (20.5 KB) + dynamic shared memory may exceed the 48 KB default: always opt in
    static PerDeviceOnce fast_attr;
    if (fast_attr.need())
      PL_CUDA(cudaFuncSetAttribute(tile_build_fast_kernel,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    PL_LAUNCH(tile_build_fast_kernel, (unsigned)ntiles, kThreads, smem, st, L->vid, L->bary, L->n,
              d1, L->ent_pix, L->ent_w, L->loc, L->useg, L->tile_off, L->tile_U, counter,
              L->rec_slice, L->rec_splat);
    return tiles_max_u(L, st);
  }
  const size_t smem = sizeof(unsigned long long) * NS + sizeof(unsigned short) * (TE + 2);
  if (smem > 48 * 1024)
    PL_CUDA(cudaFuncSetAttribute(tile_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
  PL_LAUNCH(tile_build_kernel, (unsigned)ntiles, kThreads, smem, st, L->vid, L->bary, L->n, d1, NS,
            L->ent_pix, L->ent_w, L->loc, L->useg, L->tile_off, L->tile_U, counter, L->rec_slice,
            L->rec_splat);
  return tiles_max_u(L, st);
}


namespace {

constexpr int kCW = 64;             // channels per CTA pass (16 float4 lanes per row group)
constexpr int kGroup = 16;          // lanes per row group

__device__ __forceinline__ void red_add_v4(float* addr, float4 x) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(x.x), "f"(x.y),
               "f"(x.z), "f"(x.w)
               : "memory");
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
__device__ __forceinline__ void fma4(float4& acc, float w, const float4 x) {
  float2 lo = make_float2(acc.x, acc.y), hi = make_float2(acc.z, acc.w);
  ffma2(lo, w, make_float2(x.x, x.y));
  ffma2(hi, w, make_float2(x.z, x.w));
  acc = make_float4(lo.x, lo.y, hi.x, hi.y);
}

// ---------------------------------------------------------------------------
// splat, tiled.  grid = (tiles, channel blocks of kCW).   permutohedral.h:452-455
//
// The tile's (d+1) * 128 entries, sorted by vertex, are cut into 32 equal ranges; 8 lanes own a
// range and 8 channels each (float4 indices l and l + 8 of the 64-channel block: every quarter
// warp reads 128 contiguous bytes of one staged src row -> no bank conflicts).  A lane keeps the
// running sum of the current vertex in registers and sends ONE red.global.add.v4 per float4 when
// the vertex changes: balanced work whatever the segment lengths are, at most 31 extra flushes
// per tile.  One 16-byte shared-memory record {weight, src row offset, vertex} per entry.
// ---------------------------------------------------------------------------
constexpr int kRanges = kThreads / 8;   // 32

__global__ void __launch_bounds__(kThreads, 5)
splat_tiled_kernel(const float* __restrict__ src, int64_t src_stride, int Cc, int64_t n, int d1,
                   const uint8_t* __restrict__ rec_splat, float* __restrict__ val,
                   const float* __restrict__ rowscale, int64_t rs_stride) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* s_src = reinterpret_cast<float*>(smem_raw);                 // [kTile][kCW]
  float4* s_rec = reinterpret_cast<float4*>(s_src + kTile * kCW);    // [TE] {w, row offset, vertex, -}

  const int tid = threadIdx.x;
  const int64_t tile = blockIdx.x;
  const int c0 = blockIdx.y * kCW;
  const int cw = min(kCW, Cc - c0);        // multiple of 4
  const int TE = kTile * d1;
  const int64_t p0 = tile * kTile;
  const int rows = (int)min((int64_t)kTile, n - p0);

  // stage src rows (coalesced float4; rows past the end and channels past cw read as zero) ...
  for (int i = tid; i < kTile * (kCW / 4); i += kThreads) {
    const int r = i >> 4, v = i & 15;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < rows && 4 * v < cw) {
      x = *reinterpret_cast<const float4*>(src + (p0 + r) * src_stride + c0 + 4 * v);
      if (rowscale != nullptr) {   // the source is src[p][:] * rowscale[p], never materialised
        const float sc = rowscale[(p0 + r) * rs_stride];
        x.x *= sc; x.y *= sc; x.z *= sc; x.w *= sc;
      }
    }
    *reinterpret_cast<float4*>(s_src + r * kCW + 4 * v) = x;
  }
  // ... and the tile's sorted entry record (pads: weight 0 -> vertex 0, row 0)
  const float* rp_w = reinterpret_cast<const float*>(rec_splat + tile * (int64_t)TE * 9);
  const int32_t* rp_vid = reinterpret_cast<const int32_t*>(rp_w + TE);
  const uint8_t* rp_pix = reinterpret_cast<const uint8_t*>(rp_vid + TE);
  for (int i = tid; i < TE; i += kThreads)
    s_rec[i] = make_float4(rp_w[i], __int_as_float((int)rp_pix[i] * kCW), __int_as_float(rp_vid[i]), 0.f);
  __syncthreads();

  const int rg = tid >> 3, l = tid & 7;
  const int ER = (TE + kRanges - 1) / kRanges;
  const int k0 = rg * ER, k1 = min(TE, k0 + ER);
  const bool on0 = 4 * l < cw, on1 = 32 + 4 * l < cw;
  float* vbase = val + c0 + 4 * l;
  int cur = -1;
  float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
  for (int k = k0; k < k1; k++) {
    const float4 rec = s_rec[k];
    const int v = __float_as_int(rec.z);
    if (v != cur) {
      if (cur >= 0) {
        if (on0) red_add_v4(vbase + (int64_t)cur * Cc, a0);
        if (on1) red_add_v4(vbase + (int64_t)cur * Cc + 32, a1);
      }
      cur = v;
      a0 = make_float4(0.f, 0.f, 0.f, 0.f);
      a1 = a0;
    }
    const float* row = s_src + __float_as_int(rec.y) + 4 * l;
    fma4(a0, rec.x, *reinterpret_cast<const float4*>(row));
    fma4(a1, rec.x, *reinterpret_cast<const float4*>(row + 32));
  }
  if (cur >= 0) {
    if (on0) red_add_v4(vbase + (int64_t)cur * Cc, a0);
    if (on1) red_add_v4(vbase + (int64_t)cur * Cc + 32, a1);
  }
}

// ---------------------------------------------------------------------------
// splat, tiled, d = 5: the same ranges, software-pipelined.  Persistent CTAs (two per SM) walk a
// contiguous run of (tile, channel block) items; while item i is summed, cp.async brings the src
// rows and the raw entry record of item i+1 into the other buffers, so no warp waits on HBM in
// steady state.  A lane keeps 3 of its range's 24 entries {weight, vertex << 7 | row} in registers
// and the 8 lanes of the range hand them round with shuffles: per entry two 128-byte shared-memory
// reads (the src row) and nothing else.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src, bool valid) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
  const int sz = valid ? 16 : 0;   // 0: zero-fill, nothing is read
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

constexpr int kD1p = 6;                       // d + 1 of the pipelined kernel
constexpr int kTEp = kTile * kD1p;            // 768 entries
constexpr int kERp = kTEp / kRanges;          // 24 entries per range
constexpr int kRecRaw = kTEp * 9;             // bytes of a raw splat record

// J > 1: J sources at once, src[p][:] * rowscale[p][j] for j < J, into value rows of J * Cc channels
// ([j][c]): the src tile is staged and gathered once for all of them, only the weights differ.
template <int J>
__global__ void __launch_bounds__(kThreads, 2)
splat_pipe_kernel(const float* __restrict__ src, int64_t src_stride, int Cc, int64_t n, int nblk,
                  int64_t nitems, const uint8_t* __restrict__ rec_splat, float* __restrict__ val,
                  const float* __restrict__ rowscale, int64_t rs_stride) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* s_src = reinterpret_cast<float*>(smem_raw);                    // [2][kTile][kCW]
  unsigned char* s_raw = smem_raw + 2 * kTile * kCW * sizeof(float);    // [2][kRecRaw]

  const int tid = threadIdx.x;
  const int64_t per = (nitems + gridDim.x - 1) / gridDim.x;
  const int64_t i0 = blockIdx.x * per, i1 = min(nitems, i0 + per);
  if (i0 >= i1) return;

  auto request = [&](int64_t item, int sbuf, int rbuf, bool with_rec) {
    const int64_t tile = item / nblk;
    const int c0 = (int)(item - tile * nblk) * kCW;
    const int cw = min(kCW, Cc - c0);
    const int64_t p0 = tile * kTile;
    const int rows = (int)min((int64_t)kTile, n - p0);
    float* dst = s_src + sbuf * (kTile * kCW);
#pragma unroll
    for (int k = 0; k < kTile * (kCW / 4) / kThreads; k++) {
      const int i = tid + k * kThreads;
      const int r = i >> 4, v = i & 15;
      const bool ok = r < rows && 4 * v < cw;
      cp_async16(dst + r * kCW + 4 * v, ok ? src + (p0 + r) * src_stride + c0 + 4 * v : src, ok);
    }
    if (with_rec) {
      const unsigned char* g = rec_splat + tile * (int64_t)kRecRaw;
      for (int i = tid; i < kRecRaw / 16; i += kThreads)
        cp_async16(s_raw + rbuf * kRecRaw + i * 16, g + i * 16, true);
    }
    cp_async_commit();
  };

  const int rg = tid >> 3, l = tid & 7;
  int rbuf = 0;
  int64_t cur_tile = i0 / nblk;
  request(i0, 0, 0, true);
  for (int64_t item = i0; item < i1; item++) {
    const int sbuf = (int)(item - i0) & 1;
    cp_async_wait_all();
    __syncthreads();   // item's buffers landed; everyone is done with the other buffers
    const int64_t tile = item / nblk;
    if (tile != cur_tile) {
      cur_tile = tile;
      rbuf ^= 1;
    }
    if (item + 1 < i1) {
      const bool new_tile = (item + 1) / nblk != tile;
      request(item + 1, sbuf ^ 1, rbuf ^ 1, new_tile);
    }
    const int c0 = (int)(item - tile * nblk) * kCW;
    const int cw = min(kCW, Cc - c0);
    const bool on0 = 4 * l < cw, on1 = 32 + 4 * l < cw;

    // this lane's 3 entries of the range
    const float* rw = reinterpret_cast<const float*>(s_raw + rbuf * kRecRaw);
    const int* rv = reinterpret_cast<const int*>(rw + kTEp);
    const uint8_t* rp = reinterpret_cast<const uint8_t*>(rv + kTEp);
    float w[3];
    int key[3];
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const int k = rg * kERp + j * 8 + l;
      w[j] = rw[k];
      key[j] = (rv[k] << 7) | (int)rp[k];
      // optional per-row factor of the source (src[p][:] * rowscale[p]) folded into the entry weight;
      // pads carry weight 0 and row 0 of the tile, which always exists
      if (J == 1 && rowscale != nullptr) w[j] *= rowscale[(tile * kTile + (int64_t)rp[k]) * rs_stride];
    }
    const float* sb = s_src + sbuf * (kTile * kCW) + 4 * l;
    const int64_t Cv = (int64_t)J * Cc;            // channels of a value row
    float* vbase = val + c0 + 4 * l;
    const float* rs_tile = (J > 1) ? rowscale + tile * kTile * rs_stride : nullptr;
    int cur = -1;
    float4 a0[J], a1[J];
#pragma unroll
    for (int q = 0; q < J; q++) a0[q] = a1[q] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 3; j++) {
#pragma unroll
      for (int i = 0; i < 8; i++) {
        const float wk = __shfl_sync(0xffffffffu, w[j], i, 8);
        const int kk = __shfl_sync(0xffffffffu, key[j], i, 8);
        const int v = kk >> 7;
        if (v != cur) {
          if (cur >= 0) {
#pragma unroll
            for (int q = 0; q < J; q++) {
              if (on0) red_add_v4(vbase + (int64_t)cur * Cv + q * Cc, a0[q]);
              if (on1) red_add_v4(vbase + (int64_t)cur * Cv + q * Cc + 32, a1[q]);
            }
          }
          cur = v;
#pragma unroll
          for (int q = 0; q < J; q++) a0[q] = a1[q] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        const int pix = kk & 127;
        const float* row = sb + pix * kCW;
        const float4 x0 = *reinterpret_cast<const float4*>(row);
        const float4 x1 = *reinterpret_cast<const float4*>(row + 32);
        if (J == 1) {
          fma4(a0[0], wk, x0);
          fma4(a1[0], wk, x1);
        } else {
          const float* rsp = rs_tile + (int64_t)pix * rs_stride;   // L1-resident: 128 x J floats per tile
#pragma unroll
          for (int q = 0; q < J; q++) {
            const float wq = wk * rsp[q];
            fma4(a0[q], wq, x0);
            fma4(a1[q], wq, x1);
          }
        }
      }
    }
#pragma unroll
    for (int q = 0; q < J; q++) {
      if (on0) red_add_v4(vbase + (int64_t)cur * Cv + q * Cc, a0[q]);
      if (on1) red_add_v4(vbase + (int64_t)cur * Cv + q * Cc + 32, a1[q]);
    }
  }
}

// ---------------------------------------------------------------------------
// splat for narrow inputs (1..4 channels), over the tile's vertex-sorted entry record: one thread per
// entry, segmented warp scan over the runs of equal vertex ids (contiguous after the sort, across all
// d+1 remainders), one atomic per run and channel.   permutohedral.h:452-455
// ---------------------------------------------------------------------------
template <int CH>
__global__ void __launch_bounds__(kThreads)
splat_sorted_kernel(const float* __restrict__ src, int64_t src_stride, int64_t n, int d1,
                    const uint8_t* __restrict__ rec_splat, float* __restrict__ val) {
  __shared__ float s_src[kTile * CH];
  const int tid = threadIdx.x, lane = tid & 31;
  const int64_t tile = blockIdx.x;
  const int TE = kTile * d1;
  const int64_t p0 = tile * kTile;
  const int rows = (int)min((int64_t)kTile, n - p0);
  for (int i = tid; i < kTile * CH; i += kThreads) {
    const int r = i / CH, c = i - r * CH;
    s_src[i] = r < rows ? src[(p0 + r) * src_stride + c] : 0.f;
  }
  __syncthreads();
  const float* rp_w = reinterpret_cast<const float*>(rec_splat + tile * (int64_t)TE * 9);
  const int32_t* rp_vid = reinterpret_cast<const int32_t*>(rp_w + TE);
  const uint8_t* rp_pix = reinterpret_cast<const uint8_t*>(rp_vid + TE);
  for (int base = 0; base < TE; base += kThreads) {   // TE is a multiple of 128: warps stay whole
    const int e = base + tid;
    const bool ok = e < TE;
    const float w = ok ? rp_w[e] : 0.f;
    const int v = ok ? rp_vid[e] : -1 - lane;
    const int pix = ok ? (int)rp_pix[e] : 0;
    float x[CH];
#pragma unroll
    for (int c = 0; c < CH; c++) x[c] = w * s_src[pix * CH + c];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int vo = __shfl_up_sync(0xffffffffu, v, o);
      const bool same = lane >= o && vo == v;
#pragma unroll
      for (int c = 0; c < CH; c++) {
        const float y = __shfl_up_sync(0xffffffffu, x[c], o);
        if (same) x[c] += y;
      }
    }
    const int vn = __shfl_down_sync(0xffffffffu, v, 1);
    if (ok && (lane == 31 || vn != v)) {
      if (CH == 4) {
        red_add_v4(val + (int64_t)v * 4, make_float4(x[0], x[1 % CH], x[2 % CH], x[3 % CH]));
      } else {
#pragma unroll
        for (int c = 0; c < CH; c++) atomicAdd(val + (int64_t)v * CH + c, x[c]);
      }
    }
  }
}

// the same for 4, 8 or 16 channels: V float4 lanes per entry, 32 / V entries per warp step
template <int V>
__global__ void __launch_bounds__(kThreads)
splat_sorted_vec_kernel(const float* __restrict__ src, int64_t src_stride, int64_t n, int d1,
                        const uint8_t* __restrict__ rec_splat, float* __restrict__ val) {
  __shared__ __align__(16) float s_src[kTile * 4 * V];
  constexpr int CH = 4 * V;
  const int tid = threadIdx.x, lane = tid & 31;
  const int64_t tile = blockIdx.x;
  const int TE = kTile * d1;
  const int64_t p0 = tile * kTile;
  const int rows = (int)min((int64_t)kTile, n - p0);
  for (int i = tid; i < kTile * V; i += kThreads) {
    const int r = i / V, c = i - r * V;
    reinterpret_cast<float4*>(s_src)[i] =
        r < rows ? *reinterpret_cast<const float4*>(src + (p0 + r) * src_stride + 4 * c)
                 : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __syncthreads();
  const float* rp_w = reinterpret_cast<const float*>(rec_splat + tile * (int64_t)TE * 9);
  const int32_t* rp_vid = reinterpret_cast<const int32_t*>(rp_w + TE);
  const uint8_t* rp_pix = reinterpret_cast<const uint8_t*>(rp_vid + TE);
  const int li = tid % V;
  for (int base = 0; base < TE; base += kThreads / V) {   // (kThreads / V) divides TE: warps stay whole
    const int e = base + tid / V;
    const bool ok = e < TE;
    const float w = ok ? rp_w[e] : 0.f;
    const int v = ok ? rp_vid[e] : -1 - lane;
    const int pix = ok ? (int)rp_pix[e] : 0;
    float4 x = reinterpret_cast<const float4*>(s_src)[pix * V + li];
    x.x *= w; x.y *= w; x.z *= w; x.w *= w;
#pragma unroll
    for (int o = V; o < 32; o <<= 1) {
      const int vo = __shfl_up_sync(0xffffffffu, v, o);
      const bool same = lane >= o && vo == v;
      const float y0 = __shfl_up_sync(0xffffffffu, x.x, o), y1 = __shfl_up_sync(0xffffffffu, x.y, o);
      const float y2 = __shfl_up_sync(0xffffffffu, x.z, o), y3 = __shfl_up_sync(0xffffffffu, x.w, o);
      if (same) { x.x += y0; x.y += y1; x.z += y2; x.w += y3; }
    }
    const int vn = __shfl_down_sync(0xffffffffu, v, V);
    if (ok && (lane + V >= 32 || vn != v)) red_add_v4(val + (int64_t)v * CH + 4 * li, x);
  }
}

// ---------------------------------------------------------------------------
// slice, tiled (+ fused epilogue).  grid = (tiles, channel blocks).  permutohedral.h:473-483
//
// 8 lanes per pixel row, 8 channels each (float4 indices l and l + 8), 4 rows per thread; the
// epilogue operands (sub / add rows) are requested before the gather so their HBM latency hides
// under it.  STAGE = vertex rows the CTA can stage, chosen per lattice from the largest tile.
// ---------------------------------------------------------------------------
template <bool EXACT, int STAGE, bool EPI>
__global__ void __launch_bounds__(kThreads, EPI ? 4 : 6)
slice_tiled_kernel(const float* __restrict__ val, int Cc, int64_t n, int d1,
                   const uint8_t* __restrict__ rec_slice, const int32_t* __restrict__ vid,
                   const int2* __restrict__ useg, const int* __restrict__ tile_off,
                   const int* __restrict__ tile_U, float denom, float* dst, int64_t dst_stride,
                   const float* sub, int64_t sub_stride, const float* add, int64_t add_stride) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* s_val = reinterpret_cast<float*>(smem_raw);                         // [STAGE][kCW]
  float2* s_ent = reinterpret_cast<float2*>(s_val + STAGE * kCW);            // [TE] {weight, row}

  const int tid = threadIdx.x;
  const int64_t tile = blockIdx.x;
  const int c0 = blockIdx.y * kCW;
  const int cw = min(kCW, Cc - c0);
  const int TE = kTile * d1;
  const int64_t p0 = tile * kTile;
  const int rows = (int)min((int64_t)kTile, n - p0);
  const int U = tile_U[tile];
  // The first STAGE distinct vertex rows of the tile are staged; entries whose vertex lies beyond
  // that (tiles of sharp / textured feature sets have more distinct vertices than fit) read their
  // row from L2 -- per entry, not per tile.
  // (staging pays only while most of the tile's rows fit: far beyond STAGE the staged subset is reused
  // too little to cover the copy, measured at 1080p x 128 on textured features)
  const int Us = (U <= STAGE + STAGE / 2) ? min(U, STAGE) : 0;
  const float inv = 1.0f / denom;

  {
    const int2* seg = useg + tile_off[tile];
    for (int i = tid; i < Us * 16; i += kThreads) {
      const int u = i >> 4, v = i & 15;
      if (4 * v < cw)
        *reinterpret_cast<float4*>(s_val + u * kCW + 4 * v) =
            *reinterpret_cast<const float4*>(val + (int64_t)seg[u].x * Cc + c0 + 4 * v);
    }
  }
  const float* rs_bary = reinterpret_cast<const float*>(rec_slice + tile * (int64_t)TE * 6);
  const uint16_t* rs_loc = reinterpret_cast<const uint16_t*>(rs_bary + TE);
  const int valid = rows * d1;
  for (int i = tid; i < TE; i += kThreads) {
    int ix = 0;
    if (i < valid) {
      const int lc = (int)rs_loc[i];
      ix = lc < Us ? lc * kCW : -1 - vid[tile * TE + i];      // >= 0: staged row offset, < 0: vertex id
    }
    s_ent[i] = make_float2(rs_bary[i], __int_as_float(ix));
  }
  __syncthreads();

  const int g = tid >> 3, l = tid & 7;
  const bool on0 = 4 * l < cw, on1 = 32 + 4 * l < cw;
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
  for (int r = g; r < rows; r += kThreads / 8) {
    const int64_t p = p0 + r;
    float4 s0 = zero, s1 = zero, t0 = zero, t1 = zero;
    if (EPI && sub != nullptr) {
      const float* q = sub + p * sub_stride + c0 + 4 * l;
      if (on0) s0 = *reinterpret_cast<const float4*>(q);
      if (on1) s1 = *reinterpret_cast<const float4*>(q + 32);
    }
    if (EPI && add != nullptr) {
      const float* q = add + p * add_stride + c0 + 4 * l;
      if (on0) t0 = *reinterpret_cast<const float4*>(q);
      if (on1) t1 = *reinterpret_cast<const float4*>(q + 32);
    }
    float4 a0 = zero, a1 = zero;
    for (int j = 0; j < d1; j++) {
      const float2 e = s_ent[r * d1 + j];
      const int ix = __float_as_int(e.y);
      float4 x0, x1;
      if (ix >= 0) {
        x0 = *reinterpret_cast<const float4*>(s_val + ix + 4 * l);
        x1 = *reinterpret_cast<const float4*>(s_val + ix + 4 * l + 32);
      } else {
        const float* q = val + (int64_t)(-1 - ix) * Cc + c0 + 4 * l;
        x0 = on0 ? *reinterpret_cast<const float4*>(q) : zero;
        x1 = on1 ? *reinterpret_cast<const float4*>(q + 32) : zero;
      }
      if (EXACT) {
        const float w = e.x;
        a0.x = __fadd_rn(a0.x, __fdiv_rn(__fmul_rn(w, x0.x), denom));
        a0.y = __fadd_rn(a0.y, __fdiv_rn(__fmul_rn(w, x0.y), denom));
        a0.z = __fadd_rn(a0.z, __fdiv_rn(__fmul_rn(w, x0.z), denom));
        a0.w = __fadd_rn(a0.w, __fdiv_rn(__fmul_rn(w, x0.w), denom));
        a1.x = __fadd_rn(a1.x, __fdiv_rn(__fmul_rn(w, x1.x), denom));
        a1.y = __fadd_rn(a1.y, __fdiv_rn(__fmul_rn(w, x1.y), denom));
        a1.z = __fadd_rn(a1.z, __fdiv_rn(__fmul_rn(w, x1.z), denom));
        a1.w = __fadd_rn(a1.w, __fdiv_rn(__fmul_rn(w, x1.w), denom));
      } else {
        const float wi = e.x * inv;
        fma4(a0, wi, x0);
        fma4(a1, wi, x1);
      }
    }
    if (EPI && sub != nullptr) {
      a0.x = __fsub_rn(a0.x, s0.x); a0.y = __fsub_rn(a0.y, s0.y);
      a0.z = __fsub_rn(a0.z, s0.z); a0.w = __fsub_rn(a0.w, s0.w);
      a1.x = __fsub_rn(a1.x, s1.x); a1.y = __fsub_rn(a1.y, s1.y);
      a1.z = __fsub_rn(a1.z, s1.z); a1.w = __fsub_rn(a1.w, s1.w);
    }
    if (EPI && add != nullptr) {
      a0.x = __fadd_rn(t0.x, a0.x); a0.y = __fadd_rn(t0.y, a0.y);
      a0.z = __fadd_rn(t0.z, a0.z); a0.w = __fadd_rn(t0.w, a0.w);
      a1.x = __fadd_rn(t1.x, a1.x); a1.y = __fadd_rn(t1.y, a1.y);
      a1.z = __fadd_rn(t1.z, a1.z); a1.w = __fadd_rn(t1.w, a1.w);
    }
    float* q = dst + p * dst_stride + c0 + 4 * l;
    if (on0) *reinterpret_cast<float4*>(q) = a0;
    if (on1) *reinterpret_cast<float4*>(q + 32) = a1;
  }
}

// ---------------------------------------------------------------------------
// slice contracted on the spot: out[p] += coef * sum_c slice(val)[p][c] * other[p][c].
// The guide gradient (gaussian_matrix.py:450-463) needs the filtered outer products only
// inside such inner products with src / grad rows, so the n x C filter output never exists.
// Same layout as slice_tiled_kernel; the 8 lanes of a row combine with shuffles, channel blocks
// (grid.y) with one atomic per row.
// ---------------------------------------------------------------------------
template <int STAGE>
__global__ void __launch_bounds__(kThreads, 3)
slice_dot_kernel(const float* __restrict__ val, int Cc, int64_t n, int d1,
                 const uint8_t* __restrict__ rec_slice, const int32_t* __restrict__ vid,
                 const int2* __restrict__ useg, const int* __restrict__ tile_off,
                 const int* __restrict__ tile_U, float denom, const float* __restrict__ other,
                 int64_t other_stride, float* out, int64_t out_stride, float coef, int J) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* s_val = reinterpret_cast<float*>(smem_raw);                         // [STAGE][kCW]
  float2* s_ent = reinterpret_cast<float2*>(s_val + STAGE * kCW);            // [TE] {weight, row}

  const int tid = threadIdx.x;
  const int64_t tile = blockIdx.x;
  const int c0 = blockIdx.y * kCW;
  const int cw = min(kCW, Cc - c0);
  const int TE = kTile * d1;
  const int64_t p0 = tile * kTile;
  const int rows = (int)min((int64_t)kTile, n - p0);
  const int U = tile_U[tile];
  const bool staged = U <= STAGE;
  const float inv = coef / denom;

  const float* rs_bary = reinterpret_cast<const float*>(rec_slice + tile * (int64_t)TE * 6);
  const uint16_t* rs_loc = reinterpret_cast<const uint16_t*>(rs_bary + TE);
  const int valid = rows * d1;
  for (int i = tid; i < TE; i += kThreads) {
    int ix = 0;
    if (i < valid) ix = staged ? (int)rs_loc[i] * kCW : vid[tile * TE + i];
    s_ent[i] = make_float2(rs_bary[i], __int_as_float(ix));
  }
  const int g = tid >> 3, l = tid & 7;
  const bool on0 = 4 * l < cw, on1 = 32 + 4 * l < cw;
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  const int64_t Cv = (int64_t)J * Cc;     // channels of a value row: [j][c]
  // the rows of `other` this thread contracts with, once for all J blocks
  float4 o0[kTile * 8 / kThreads], o1[kTile * 8 / kThreads];
#pragma unroll
  for (int i = 0; i < kTile * 8 / kThreads; i++) {
    const int r = g + i * (kThreads / 8);
    o0[i] = o1[i] = zero;
    if (r < rows) {
      const float* q = other + (p0 + r) * other_stride + c0 + 4 * l;
      if (on0) o0[i] = *reinterpret_cast<const float4*>(q);
      if (on1) o1[i] = *reinterpret_cast<const float4*>(q + 32);
    }
  }
  for (int jb = 0; jb < J; jb++) {
    const float* vj = val + (int64_t)jb * Cc + c0;
    __syncthreads();   // the previous block's staged rows are dead (and s_ent is written)
    if (staged) {
      const int2* seg = useg + tile_off[tile];
      for (int i = tid; i < U * 16; i += kThreads) {
        const int u = i >> 4, v = i & 15;
        if (4 * v < cw)
          *reinterpret_cast<float4*>(s_val + u * kCW + 4 * v) =
              *reinterpret_cast<const float4*>(vj + (int64_t)seg[u].x * Cv + 4 * v);
      }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kTile * 8 / kThreads; i++) {   // whole warps stay in the loop for the shuffles
      const int r = g + i * (kThreads / 8);
      const bool ok = r < rows;
      float4 a0 = zero, a1 = zero;
      if (ok) {
        for (int j = 0; j < d1; j++) {
          const float2 e = s_ent[r * d1 + j];
          const int ix = __float_as_int(e.y);
          float4 x0 = zero, x1 = zero;
          if (staged) {
            if (on0) x0 = *reinterpret_cast<const float4*>(s_val + ix + 4 * l);
            if (on1) x1 = *reinterpret_cast<const float4*>(s_val + ix + 4 * l + 32);
          } else {
            const float* q = vj + (int64_t)ix * Cv + 4 * l;
            if (on0) x0 = *reinterpret_cast<const float4*>(q);
            if (on1) x1 = *reinterpret_cast<const float4*>(q + 32);
          }
          fma4(a0, e.x, x0);
          fma4(a1, e.x, x1);
        }
      }
      float dot = a0.x * o0[i].x + a0.y * o0[i].y + a0.z * o0[i].z + a0.w * o0[i].w +
                  a1.x * o1[i].x + a1.y * o1[i].y + a1.z * o1[i].z + a1.w * o1[i].w;
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o, 8);
      if (ok && l == 0) {
        float* dst = out + (p0 + r) * out_stride + jb;
        if (gridDim.y == 1) *dst += inv * dot;
        else atomicAdd(dst, inv * dot);
      }
    }
  }
}

}  // namespace

// val must be zeroed by the caller.  Requires Cc % 4 == 0 and 16-byte aligned rows.
int splat_tiled(Lattice* L, const float* src, int64_t src_stride, int Cc, float* val,
                cudaStream_t st, const float* rowscale, int64_t rs_stride, int J) {
  const int d1 = L->d + 1;
  const size_t smem = sizeof(float) * kTile * kCW + sizeof(float4) * (size_t)kTile * d1;
  static PerDeviceOnce attr_done;
  if (attr_done.need()) {
    PL_CUDA(cudaFuncSetAttribute(splat_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 100 * 1024));
    PL_CUDA(cudaFuncSetAttribute(splat_pipe_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 100 * 1024));
    PL_CUDA(cudaFuncSetAttribute(splat_pipe_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 100 * 1024));
  }
  static const bool no_pipe = getenv("PL_SPLAT_NO_PIPE") != nullptr;
  if (d1 == kD1p && !no_pipe && L->M < (1 << 24)) {   // vertex id and row share one 32-bit key
    const int nblk = (Cc + kCW - 1) / kCW;
    const int64_t nitems = L->ntiles * nblk;
    const size_t smem_p = 2 * sizeof(float) * kTile * kCW + 2 * (size_t)kRecRaw;
    const int64_t cap = 2 * PL_SM_COUNT;
    if (J == 5) {
      PL_LAUNCH(splat_pipe_kernel<5>, (unsigned)(nitems < cap ? nitems : cap), kThreads, smem_p, st, src,
                src_stride, Cc, L->n, nblk, nitems, L->rec_splat, val, rowscale, rs_stride);
      return PL_OK;
    }
    if (J == 1) {
      PL_LAUNCH(splat_pipe_kernel<1>, (unsigned)(nitems < cap ? nitems : cap), kThreads, smem_p, st, src,
                src_stride, Cc, L->n, nblk, nitems, L->rec_splat, val, rowscale, rs_stride);
      return PL_OK;
    }
  }
  if (J != 1) {
    set_last_error("splat_tiled: %d row factors at once needs d = 5 and J = 5", J);
    return PL_ERR_INVALID;
  }
  dim3 grid((unsigned)L->ntiles, (unsigned)((Cc + kCW - 1) / kCW), 1);
  PL_LAUNCH(splat_tiled_kernel, grid, kThreads, smem, st, src, src_stride, Cc, L->n, d1,
            L->rec_splat, val, rowscale, rs_stride);
  return PL_OK;
}

namespace {
template <bool EXACT, int STAGE, bool EPI>
int launch_slice(Lattice* L, const float* val, int Cc, float denom, float* dst, int64_t dst_stride,
                 const float* sub, int64_t sub_stride, const float* add, int64_t add_stride,
                 cudaStream_t st) {
  const int d1 = L->d + 1;
  const size_t smem = sizeof(float) * STAGE * kCW + sizeof(float2) * (size_t)kTile * d1;
  static PerDeviceOnce attr_done;
  if (attr_done.need()) {
    PL_CUDA(cudaFuncSetAttribute((slice_tiled_kernel<EXACT, STAGE, EPI>),
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  }
  dim3 grid((unsigned)L->ntiles, (unsigned)((Cc + kCW - 1) / kCW), 1);
  PL_LAUNCH((slice_tiled_kernel<EXACT, STAGE, EPI>), grid, kThreads, smem, st, val, Cc, L->n, d1,
            L->rec_slice, L->vid, L->useg, L->tile_off, L->tile_U, denom, dst, dst_stride, sub,
            sub_stride, add, add_stride);
  return PL_OK;
}
}  // namespace

int slice_tiled(Lattice* L, const float* val, int Cc, float denom, float* dst, int64_t dst_stride,
                const float* sub, int64_t sub_stride, const float* add, int64_t add_stride,
                bool exact, cudaStream_t st) {
  if (L->max_U < 0) {   // first tiled slice on this lattice: the tile build has long finished
    PL_CUDA(cudaEventSynchronize(L->max_U_ev));
    L->max_U = L->h_max_U[1];
    L->mean_U = L->ntiles > 0 ? (int)((uint32_t)L->h_max_U[0] / (uint64_t)L->ntiles) : 0;
  }
#define PL_SLICE(EX, ST)                                                                         \
  return (sub != nullptr || add != nullptr)                                                      \
             ? launch_slice<EX, ST, true>(L, val, Cc, denom, dst, dst_stride, sub, sub_stride,   \
                                          add, add_stride, st)                                   \
             : launch_slice<EX, ST, false>(L, val, Cc, denom, dst, dst_stride, sub, sub_stride,  \
                                           add, add_stride, st)
  // Feature sets with little locality inside a tile (textured images at sharp scales: 350 distinct
  // vertices per 768 entries) never stage (U > 1.5 x STAGE): the small instance leaves the shared memory
  // to two more CTAs per SM, which is what the L2-latency-bound gathers want (ncu, profiles/r2_summary.md)
  const bool no_locality = L->mean_U > 240;
  if (exact) {
    if (L->max_U <= 48 || no_locality) PL_SLICE(true, 48);
    if (L->max_U <= 96) PL_SLICE(true, 96);
    PL_SLICE(true, 160);
  }
  if (L->max_U <= 48 || no_locality) PL_SLICE(false, 48);
  if (L->max_U <= 96) PL_SLICE(false, 96);
  PL_SLICE(false, 160);
#undef PL_SLICE
}

// narrow inputs: Cc in {1, 2, 3} (scalars) or 4 (one float4 lane per entry, 16-byte aligned rows);
// measured against the per-pixel aggregation: C = 1 1.3-1.6x, C = 4 1.1-1.4x faster, C = 16 (four
// lanes per entry) 1.1x slower, so wider inputs stay with splat_agg_kernel
int splat_sorted(Lattice* L, const float* src, int64_t src_stride, int Cc, float* val, cudaStream_t st) {
  const int d1 = L->d + 1;
  switch (Cc) {
    case 1: PL_LAUNCH(splat_sorted_kernel<1>, (unsigned)L->ntiles, kThreads, 0, st, src, src_stride, L->n, d1, L->rec_splat, val); break;
    case 2: PL_LAUNCH(splat_sorted_kernel<2>, (unsigned)L->ntiles, kThreads, 0, st, src, src_stride, L->n, d1, L->rec_splat, val); break;
    case 3: PL_LAUNCH(splat_sorted_kernel<3>, (unsigned)L->ntiles, kThreads, 0, st, src, src_stride, L->n, d1, L->rec_splat, val); break;
    case 4: PL_LAUNCH(splat_sorted_vec_kernel<1>, (unsigned)L->ntiles, kThreads, 0, st, src, src_stride, L->n, d1, L->rec_splat, val); break;
    default: set_last_error("splat_sorted: C = %d", Cc); return PL_ERR_INVALID;
  }
  return PL_OK;
}

int slice_dot_tiled(Lattice* L, const float* val, int Cc, float denom, const float* other,
                    int64_t other_stride, float* out, int64_t out_stride, float coef, cudaStream_t st,
                    int J) {
  if (L->max_U < 0) {
    PL_CUDA(cudaEventSynchronize(L->max_U_ev));
    L->max_U = L->h_max_U[1];
    L->mean_U = L->ntiles > 0 ? (int)((uint32_t)L->h_max_U[0] / (uint64_t)L->ntiles) : 0;
  }
  const int d1 = L->d + 1;
  static PerDeviceOnce attr_done;
  if (attr_done.need()) {
    PL_CUDA(cudaFuncSetAttribute(slice_dot_kernel<48>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    PL_CUDA(cudaFuncSetAttribute(slice_dot_kernel<96>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    PL_CUDA(cudaFuncSetAttribute(slice_dot_kernel<160>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  }
  dim3 grid((unsigned)L->ntiles, (unsigned)((Cc + kCW - 1) / kCW), 1);
  const size_t ent = sizeof(float2) * (size_t)kTile * d1;
  if (L->max_U <= 48) {
    PL_LAUNCH(slice_dot_kernel<48>, grid, kThreads, sizeof(float) * 48 * kCW + ent, st, val, Cc, L->n, d1,
              L->rec_slice, L->vid, L->useg, L->tile_off, L->tile_U, denom, other, other_stride, out,
              out_stride, coef, J);
  } else if (L->max_U <= 96) {
    PL_LAUNCH(slice_dot_kernel<96>, grid, kThreads, sizeof(float) * 96 * kCW + ent, st, val, Cc, L->n, d1,
              L->rec_slice, L->vid, L->useg, L->tile_off, L->tile_U, denom, other, other_stride, out,
              out_stride, coef, J);
  } else {
    PL_LAUNCH(slice_dot_kernel<160>, grid, kThreads, sizeof(float) * 160 * kCW + ent, st, val, Cc, L->n, d1,
              L->rec_slice, L->vid, L->useg, L->tile_off, L->tile_U, denom, other, other_stride, out,
              out_stride, coef, J);
  }
  return PL_OK;
}

}  // namespace pl
