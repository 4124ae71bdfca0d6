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
  unsigned short* s_cnt = reinterpret_cast<unsigned short*>(smem_raw);  // [nchunks][TE]
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
  for (int i = tid; i < nchunks * U; i += kThreads) s_cnt[(i / U) * TE + (i % U)] = 0;
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
      if (u != 0xFFFF && my_r[q] == 0) s_cnt[chunk * TE + u] = (unsigned short)__popc(peers);
    }
  }
  __syncthreads();
  // prefix over the chunks, per vertex
  for (int u = tid; u < U; u += kThreads) {
    int running = 0;
    for (int c = 0; c < nchunks; c++) {
      const int t = s_cnt[c * TE + u];
      s_cnt[c * TE + u] = (unsigned short)running;
      running += t;
    }
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
        const int pos = (int)s_start[u] + (int)s_cnt[chunk * TE + u] + my_r[q];
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
  PL_TRY(dev_alloc((void**)&L->tile_off, sizeof(int) * (size_t)(2 * ntiles + 1), st));
  L->tile_U = L->tile_off + ntiles;
  int* counter = L->tile_off + 2 * ntiles;
  L->ntiles = ntiles;
  PL_TRY(dev_alloc((void**)&L->rec_slice, E_pad * 6, st));
  PL_TRY(dev_alloc((void**)&L->rec_splat, E_pad * 9, st));
  L->bytes += (int64_t)(E_pad * (1 + 4 + 2 + 8 + 6 + 9) + sizeof(int) * (size_t)(2 * ntiles + 1));
  PL_CUDA(cudaMemsetAsync(counter, 0, sizeof(int), st));
  if (TE <= kFastMaxTE) {
    const size_t smem = sizeof(unsigned short) * (size_t)((TE + 31) / 32) * TE;
    // static (20.5 KB) + dynamic shared memory exceeds the 48 KB default: always opt in
    PL_CUDA(cudaFuncSetAttribute(tile_build_fast_kernel,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    PL_LAUNCH(tile_build_fast_kernel, (unsigned)ntiles, kThreads, smem, st, L->vid, L->bary, L->n,
              d1, L->ent_pix, L->ent_w, L->loc, L->useg, L->tile_off, L->tile_U, counter,
              L->rec_slice, L->rec_splat);
    return PL_OK;
  }
  const size_t smem = sizeof(unsigned long long) * NS + sizeof(unsigned short) * (TE + 2);
  if (smem > 48 * 1024)
    PL_CUDA(cudaFuncSetAttribute(tile_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
  PL_LAUNCH(tile_build_kernel, (unsigned)ntiles, kThreads, smem, st, L->vid, L->bary, L->n, d1, NS,
            L->ent_pix, L->ent_w, L->loc, L->useg, L->tile_off, L->tile_U, counter, L->rec_slice,
            L->rec_splat);
  return PL_OK;
}


namespace {

constexpr int kCW = 64;             // channels per CTA pass (16 float4 lanes per row group)
constexpr int kGroup = 16;          // lanes per row group
constexpr int kGroups = kThreads / kGroup;
constexpr int kStageRows = 160;     // vertex rows a slice CTA can stage (160 x 64 x 4 = 40 KB)

__device__ __forceinline__ void red_add_v4(float* addr, float4 x) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(x.x), "f"(x.y),
               "f"(x.z), "f"(x.w)
               : "memory");
}

// ---------------------------------------------------------------------------
// splat, tiled.  grid = (tiles, channel blocks of kCW).   permutohedral.h:452-455
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
splat_tiled_kernel(const float* __restrict__ src, int64_t src_stride, int Cc, int64_t n, int d1,
                   const uint8_t* __restrict__ ent_pix, const float* __restrict__ ent_w,
                   const int2* __restrict__ useg, const int* __restrict__ tile_off,
                   const int* __restrict__ tile_U, float* __restrict__ val) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* s_src = reinterpret_cast<float*>(smem_raw);                 // [kTile][kCW]
  float* s_w = s_src + kTile * kCW;                                  // [TE]
  uint8_t* s_pix = reinterpret_cast<uint8_t*>(s_w + kTile * d1);     // [TE]

  const int tid = threadIdx.x;
  const int64_t tile = blockIdx.x;
  const int c0 = blockIdx.y * kCW;
  const int cw = min(kCW, Cc - c0);        // multiple of 4
  const int V = cw >> 2;
  const int TE = kTile * d1;
  const int64_t p0 = tile * kTile;
  const int rows = (int)min((int64_t)kTile, n - p0);

  // stage src rows (coalesced float4) and the tile's sorted entry list
  for (int i = tid; i < rows * V; i += kThreads) {
    const int r = i / V, v = i - r * V;
    *reinterpret_cast<float4*>(s_src + r * kCW + 4 * v) =
        *reinterpret_cast<const float4*>(src + (p0 + r) * src_stride + c0 + 4 * v);
  }
  const int valid = rows * d1;
  for (int i = tid; i < valid; i += kThreads) {
    s_w[i] = ent_w[tile * TE + i];
    s_pix[i] = ent_pix[tile * TE + i];
  }
  __syncthreads();

  const int g = tid / kGroup, l = tid % kGroup;
  const int U = tile_U[tile];
  const int2* seg = useg + tile_off[tile];
  if (l < V) {
    for (int u = g; u < U; u += kGroups) {
      const int2 sg = seg[u];
      const int st = sg.y & 0xFFFF, cnt = sg.y >> 16;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int k = st; k < st + cnt; k++) {
        const float w = s_w[k];
        const float4 x = *reinterpret_cast<const float4*>(s_src + (int)s_pix[k] * kCW + 4 * l);
        // product rounded, then added: same two roundings as `val += w*value` (:455)
        acc.x = __fadd_rn(acc.x, __fmul_rn(w, x.x));
        acc.y = __fadd_rn(acc.y, __fmul_rn(w, x.y));
        acc.z = __fadd_rn(acc.z, __fmul_rn(w, x.z));
        acc.w = __fadd_rn(acc.w, __fmul_rn(w, x.w));
      }
      red_add_v4(val + (int64_t)sg.x * Cc + c0 + 4 * l, acc);
    }
  }
}

// ---------------------------------------------------------------------------
// slice, tiled (+ fused epilogue).  grid = (tiles, channel blocks).  permutohedral.h:473-483
// ---------------------------------------------------------------------------
template <bool EXACT>
__global__ void __launch_bounds__(kThreads)
slice_tiled_kernel(const float* __restrict__ val, int Cc, int64_t n, int d1,
                   const uint16_t* __restrict__ loc, const float* __restrict__ bary,
                   const int32_t* __restrict__ vid, const int2* __restrict__ useg,
                   const int* __restrict__ tile_off, const int* __restrict__ tile_U, float denom,
                   float* dst, int64_t dst_stride, const float* sub, int64_t sub_stride,
                   const float* add, int64_t add_stride) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* s_val = reinterpret_cast<float*>(smem_raw);                         // [kStageRows][kCW]
  float* s_w = s_val + kStageRows * kCW;                                     // [TE]
  int* s_idx = reinterpret_cast<int*>(s_w + kTile * d1);                     // [TE] local or global id

  const int tid = threadIdx.x;
  const int64_t tile = blockIdx.x;
  const int c0 = blockIdx.y * kCW;
  const int cw = min(kCW, Cc - c0);
  const int V = cw >> 2;
  const int TE = kTile * d1;
  const int64_t p0 = tile * kTile;
  const int rows = (int)min((int64_t)kTile, n - p0);
  const int U = tile_U[tile];
  const bool staged = U <= kStageRows;   // uniform per CTA
  const float inv = 1.0f / denom;

  if (staged) {
    const int2* seg = useg + tile_off[tile];
    for (int i = tid; i < U * V; i += kThreads) {
      const int u = i / V, v = i - u * V;
      *reinterpret_cast<float4*>(s_val + u * kCW + 4 * v) =
          *reinterpret_cast<const float4*>(val + (int64_t)seg[u].x * Cc + c0 + 4 * v);
    }
  }
  const int valid = rows * d1;
  for (int i = tid; i < valid; i += kThreads) {
    s_w[i] = bary[tile * TE + i];
    s_idx[i] = staged ? (int)loc[tile * TE + i] : vid[tile * TE + i];
  }
  __syncthreads();

  const int g = tid / kGroup, l = tid % kGroup;
  if (l >= V) return;
  for (int r = g; r < rows; r += kGroups) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int j = 0; j < d1; j++) {
      const float w = s_w[r * d1 + j];
      const int ix = s_idx[r * d1 + j];
      const float4 x = staged
                           ? *reinterpret_cast<const float4*>(s_val + ix * kCW + 4 * l)
                           : *reinterpret_cast<const float4*>(val + (int64_t)ix * Cc + c0 + 4 * l);
      if (EXACT) {
        acc.x = __fadd_rn(acc.x, __fdiv_rn(__fmul_rn(w, x.x), denom));
        acc.y = __fadd_rn(acc.y, __fdiv_rn(__fmul_rn(w, x.y), denom));
        acc.z = __fadd_rn(acc.z, __fdiv_rn(__fmul_rn(w, x.z), denom));
        acc.w = __fadd_rn(acc.w, __fdiv_rn(__fmul_rn(w, x.w), denom));
      } else {
        const float wi = w * inv;
        acc.x = __fmaf_rn(wi, x.x, acc.x);
        acc.y = __fmaf_rn(wi, x.y, acc.y);
        acc.z = __fmaf_rn(wi, x.z, acc.z);
        acc.w = __fmaf_rn(wi, x.w, acc.w);
      }
    }
    const int64_t p = p0 + r;
    if (sub != nullptr) {
      const float4 s = *reinterpret_cast<const float4*>(sub + p * sub_stride + c0 + 4 * l);
      acc.x = __fsub_rn(acc.x, s.x); acc.y = __fsub_rn(acc.y, s.y);
      acc.z = __fsub_rn(acc.z, s.z); acc.w = __fsub_rn(acc.w, s.w);
    }
    if (add != nullptr) {
      const float4 s = *reinterpret_cast<const float4*>(add + p * add_stride + c0 + 4 * l);
      acc.x = __fadd_rn(s.x, acc.x); acc.y = __fadd_rn(s.y, acc.y);
      acc.z = __fadd_rn(s.z, acc.z); acc.w = __fadd_rn(s.w, acc.w);
    }
    *reinterpret_cast<float4*>(dst + p * dst_stride + c0 + 4 * l) = acc;
  }
}

}  // namespace

// val must be zeroed by the caller.  Requires Cc % 4 == 0 and 16-byte aligned rows.
int splat_tiled(Lattice* L, const float* src, int64_t src_stride, int Cc, float* val,
                cudaStream_t st) {
  const int d1 = L->d + 1;
  const size_t smem = sizeof(float) * kTile * kCW + (sizeof(float) + 1) * (size_t)kTile * d1 + 16;
  static bool attr_done = false;
  if (!attr_done && smem > 48 * 1024) {
    PL_CUDA(cudaFuncSetAttribute(splat_tiled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
    attr_done = true;
  }
  dim3 grid((unsigned)L->ntiles, (unsigned)((Cc + kCW - 1) / kCW), 1);
  PL_LAUNCH(splat_tiled_kernel, grid, kThreads, smem, st, src, src_stride, Cc, L->n, d1,
            L->ent_pix, L->ent_w, L->useg, L->tile_off, L->tile_U, val);
  return PL_OK;
}

int slice_tiled(Lattice* L, const float* val, int Cc, float denom, float* dst, int64_t dst_stride,
                const float* sub, int64_t sub_stride, const float* add, int64_t add_stride,
                bool exact, cudaStream_t st) {
  const int d1 = L->d + 1;
  const size_t smem = sizeof(float) * kStageRows * kCW + 2 * sizeof(float) * (size_t)kTile * d1;
  static bool attr_done = false;
  if (!attr_done) {
    PL_CUDA(cudaFuncSetAttribute(slice_tiled_kernel<false>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    PL_CUDA(cudaFuncSetAttribute(slice_tiled_kernel<true>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    attr_done = true;
  }
  dim3 grid((unsigned)L->ntiles, (unsigned)((Cc + kCW - 1) / kCW), 1);
  if (exact) {
    PL_LAUNCH((slice_tiled_kernel<true>), grid, kThreads, smem, st, val, Cc, L->n, d1, L->loc,
              L->bary, L->vid, L->useg, L->tile_off, L->tile_U, denom, dst, dst_stride, sub,
              sub_stride, add, add_stride);
  } else {
    PL_LAUNCH((slice_tiled_kernel<false>), grid, kThreads, smem, st, val, Cc, L->n, d1, L->loc,
              L->bary, L->vid, L->useg, L->tile_off, L->tile_U, denom, dst, dst_stride, sub,
              sub_stride, add, add_stride);
  }
  return PL_OK;
}

}  // namespace pl
