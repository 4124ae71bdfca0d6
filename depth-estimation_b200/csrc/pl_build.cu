// pl_build.cu -- lattice construction on the GPU (sm_100a).
//
// Replaces, for a whole feature matrix at once, what the reference does point
// by point inside PermutohedralLattice::splat (crf/lattice/lite/permutohedral.h
// :376-450) and HashTablePermutohedral (:29-169), plus the neighbour lookups of
// blur() (:503-522) which are tabulated once here instead of being re-hashed
// on every axis of every filter call.
//
// Pipeline (all on one stream):
//   1. elevate_kernel     per point: elevate -> greedy remainder-0 point ->
//                         rank -> hyperplane fix-up -> barycentric weights.
//                         Every fp32 operation is spelled with _rn intrinsics
//                         in the reference's order so the INTEGER results
//                         (greedy, rank) are bit-identical to g++ -O2 on x86.
//   2. insert_kernel      per (point, remainder): key -> open-addressing table
//                         of 64-bit words {31-bit fingerprint | entry id}.
//                         Lanes of a warp holding the same key elect one
//                         inserter (match.any + key compare).  A claimed slot
//                         keeps the SMALLEST entry id that maps to it
//                         (atomicMin), i.e. the key's first touch in the
//                         reference's scanline order.
//   3. count/scan/assign  dense vertex ids = rank of the first-touch entry
//                         among all first touches -> exactly the reference's
//                         numbering (hash table `filled++` order, :73-77).
//   4. gather_kernel      vid[p][r] = vertex id.
//   5. neighbour_kernel   per vertex, per axis: the two blur neighbours.
//
// The table capacity is >= 1.25 x entries so it can never fill; a bounded
// probe count plus an overflow flag and a host retry guard the impossible.
#include <math.h>
#include <stdlib.h>

#include <chrono>

#include "pl_common.cuh"

namespace pl {

namespace {

constexpr unsigned long long kEmpty = 0xFFFFFFFFFFFFFFFFull;
constexpr int kProbeLimit = 1 << 14;

struct ScaleFactors {
  float s[PL_MAX_D];
};

// canonical simplex entry (permutohedral.h:346-351): row r, column k
__device__ __forceinline__ int canonical(int r, int k, int D) { return (k <= D - r) ? r : r - (D + 1); }

// 64-bit mix of the d key coordinates; low bits pick the slot, high 31 bits
// are the fingerprint.  Slot positions are not part of parity.
template <int D>
__device__ __forceinline__ uint64_t hash_key(const int16_t* key) {
  uint64_t k = 0;
#pragma unroll
  for (int i = 0; i < D; i++) k = (k + (uint64_t)(int64_t)key[i]) * 0x9E3779B97F4A7C15ull;
  k ^= k >> 29;
  k *= 0xBF58476D1CE4E5B9ull;
  k ^= k >> 32;
  return k;
}

// key of (point p, remainder r) rebuilt from the stored greedy / rank
// (permutohedral.h:446-447, int16 arithmetic like the reference's `short`)
template <int D>
__device__ __forceinline__ void load_key(const int16_t* __restrict__ greedy,
                                         const int8_t* __restrict__ rank, int64_t p, int r,
                                         int16_t* key) {
  const int16_t* g = greedy + p * (D + 1);
  const int8_t* k = rank + p * (D + 1);
#pragma unroll
  for (int i = 0; i < D; i++) key[i] = (int16_t)(g[i] + canonical(r, k[i], D));
}

// ---------------------------------------------------------------------------
// 1. elevate / greedy / rank / barycentric          permutohedral.h:380-441
// ---------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(256)
elevate_kernel(const float* __restrict__ ref, const GuideSrc gs, int64_t n, ScaleFactors sf,
               int16_t* __restrict__ greedy_out, int8_t* __restrict__ rank_out,
               float* __restrict__ bary_out) {
  constexpr int D1 = D + 1;
  const float scale = 1.0f / (D + 1);  // :387
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n;
       p += (int64_t)gridDim.x * blockDim.x) {
    float pos[D];
    if (D == 5 && gs.rgb != nullptr) {
      // ijrgbGuide.forward (crf_module.py:111-114) evaluated in place: ij = mgrid / sqrt(h^2+w^2)
      // in float64, cast to fp32, then fp32 divisions by s_ij / s_rgb -- bit for bit what torch
      // computes, so the [n, 5] feature matrix never exists in HBM.
      const int y = (int)(p / gs.w), x = (int)(p - (int64_t)y * gs.w);
      const int64_t hw = (int64_t)gs.h * gs.w;
      pos[0] = __fdiv_rn(__double2float_rn((double)y / gs.diag), gs.s_ij);
      pos[1 % D] = __fdiv_rn(__double2float_rn((double)x / gs.diag), gs.s_ij);
      pos[2 % D] = __fdiv_rn(gs.rgb[p], gs.s_rgb);
      pos[3 % D] = __fdiv_rn(gs.rgb[hw + p], gs.s_rgb);
      pos[4 % D] = __fdiv_rn(gs.rgb[2 * hw + p], gs.s_rgb);
    } else {
#pragma unroll
      for (int i = 0; i < D; i++) pos[i] = ref[p * D + i];
    }

    // :380-384  (int * float) * float, left to right; a - b + c left to right
    float el[D1];
    el[D] = __fmul_rn(__fmul_rn((float)(-D), pos[D - 1]), sf.s[D - 1]);
#pragma unroll
    for (int i = D - 1; i > 0; i--) {
      float a = __fmul_rn(__fmul_rn((float)i, pos[i - 1]), sf.s[i - 1]);
      float b = __fmul_rn(__fmul_rn((float)(i + 2), pos[i]), sf.s[i]);
      el[i] = __fadd_rn(__fsub_rn(el[i + 1], a), b);
    }
    el[0] = __fadd_rn(el[1], __fmul_rn(__fmul_rn(2.0f, pos[0]), sf.s[0]));

    // :392-403
    int gr[D1];
    int sum = 0;
#pragma unroll
    for (int i = 0; i < D1; i++) {
      float v = __fmul_rn(el[i], scale);
      float up = __fmul_rn(ceilf(v), (float)D1);
      float down = __fmul_rn(floorf(v), (float)D1);
      int g = (__fsub_rn(up, el[i]) < __fsub_rn(el[i], down)) ? (int)up : (int)down;
      gr[i] = (int)(int16_t)g;  // (short) narrowing, :398-399
      sum += gr[i];
    }
    sum = (int)__fmul_rn((float)sum, scale);  // `sum *= scale`, :403

    // :407-411
    int rk[D1];
    float diff[D1];
#pragma unroll
    for (int i = 0; i < D1; i++) {
      rk[i] = 0;
      diff[i] = __fsub_rn(el[i], (float)gr[i]);
    }
#pragma unroll
    for (int i = 0; i < D; i++)
#pragma unroll
      for (int j = i + 1; j < D1; j++) {
        if (diff[i] < diff[j]) rk[i]++;
        else rk[j]++;
      }

    // :413-433
    if (sum > 0) {
#pragma unroll
      for (int i = 0; i < D1; i++) {
        if (rk[i] >= D1 - sum) {
          gr[i] = (int)(int16_t)(gr[i] - D1);
          rk[i] += sum - D1;
        } else {
          rk[i] += sum;
        }
      }
    } else if (sum < 0) {
#pragma unroll
      for (int i = 0; i < D1; i++) {
        if (rk[i] < -sum) {
          gr[i] = (int)(int16_t)(gr[i] + D1);
          rk[i] += D1 + sum;
        } else {
          rk[i] += sum;
        }
      }
    }

    // :436-441  accumulate in i order; slots selected by predicate so the
    // array stays in registers (an out-of-range rank simply matches nothing)
    float b[D + 2];
#pragma unroll
    for (int k = 0; k < D + 2; k++) b[k] = 0.0f;
#pragma unroll
    for (int i = 0; i < D1; i++) {
      float t = __fmul_rn(__fsub_rn(el[i], (float)gr[i]), scale);
#pragma unroll
      for (int k = 0; k < D + 2; k++) {
        if (k == D - rk[i]) b[k] = __fadd_rn(b[k], t);
        if (k == D + 1 - rk[i]) b[k] = __fsub_rn(b[k], t);
      }
    }
    b[0] = __fadd_rn(b[0], __fadd_rn(1.0f, b[D + 1]));

#pragma unroll
    for (int i = 0; i < D1; i++) {
      greedy_out[p * D1 + i] = (int16_t)gr[i];
      rank_out[p * D1 + i] = (int8_t)rk[i];
      bary_out[p * D1 + i] = b[i];
    }
  }
}

// ---------------------------------------------------------------------------
// 2. hash insert                                    permutohedral.h:59-91
// ---------------------------------------------------------------------------
template <int D>
__device__ __forceinline__ bool entry_has_key(uint32_t entry, int r, const int16_t* key,
                                              const int16_t* __restrict__ greedy,
                                              const int8_t* __restrict__ rank) {
  // a key's coordinates are all congruent to its remainder mod d+1, so a
  // different remainder means a different key
  uint32_t p2 = entry / (D + 1);
  int r2 = (int)(entry - p2 * (D + 1));
  if (r2 != r) return false;
  int16_t other[D];
  load_key<D>(greedy, rank, (int64_t)p2, r2, other);
  bool same = true;
#pragma unroll
  for (int i = 0; i < D; i++) same = same && (other[i] == key[i]);
  return same;
}

template <int D>
__device__ __forceinline__ int32_t table_insert(unsigned long long* table, uint32_t mask,
                                                uint64_t k64, const int16_t* key, int r,
                                                uint32_t entry, const int16_t* __restrict__ greedy,
                                                const int8_t* __restrict__ rank, int* overflow) {
  uint32_t h = (uint32_t)k64 & mask;
  const uint32_t fp = (uint32_t)(k64 >> 33);
  const unsigned long long mine = ((unsigned long long)fp << 32) | entry;
  for (int probe = 0; probe < kProbeLimit; probe++) {
    unsigned long long cur = *((volatile unsigned long long*)(table + h));
    if (cur == kEmpty) {
      cur = atomicCAS(table + h, kEmpty, mine);
      if (cur == kEmpty) return (int32_t)h;  // claimed: this key's first insert
    }
    if ((uint32_t)(cur >> 32) == fp &&
        entry_has_key<D>((uint32_t)cur, r, key, greedy, rank)) {
      // same key already present: keep the earliest entry id (first touch)
      if (entry < (uint32_t)cur) atomicMin(table + h, mine);
      return (int32_t)h;
    }
    h = (h + 1) & mask;
  }
  atomicExch(overflow, 1);
  return 0;
}

template <int D>
__global__ void __launch_bounds__(256)
insert_kernel(const int16_t* __restrict__ greedy, const int8_t* __restrict__ rank, int64_t n,
              unsigned long long* table, uint32_t mask, int32_t* __restrict__ slot_out,
              int* overflow) {
  constexpr int D1 = D + 1;
  constexpr int NW = (D + 1) / 2;  // packed key words
  const int lane = threadIdx.x & 31;
  // whole warps iterate together (warp collectives inside)
  for (int64_t base = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) - lane; base < n;
       base += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = base + lane;
    const bool valid = p < n;
    int gr[D1], rk[D1];
#pragma unroll
    for (int i = 0; i < D1; i++) {
      gr[i] = valid ? (int)greedy[p * D1 + i] : 0;
      rk[i] = valid ? (int)rank[p * D1 + i] : 0;
    }
#pragma unroll 1
    for (int r = 0; r < D1; r++) {
      int16_t key[D];
#pragma unroll
      for (int i = 0; i < D; i++) key[i] = (int16_t)(gr[i] + canonical(r, rk[i], D));
      const uint64_t k64 = hash_key<D>(key);
      // warp-level dedup: neighbouring pixels mostly share lattice vertices
      const unsigned peers = __match_any_sync(0xffffffffu, (uint32_t)(k64 >> 32));
      const int leader = __ffs(peers) - 1;
      bool same = true;
#pragma unroll
      for (int w = 0; w < NW; w++) {
        uint32_t lo = (uint16_t)key[2 * w];
        uint32_t hi = (2 * w + 1 < D) ? (uint32_t)(uint16_t)key[2 * w + 1] : 0u;
        uint32_t kw = lo | (hi << 16);
        same = same && (__shfl_sync(0xffffffffu, kw, leader) == kw);
      }
      const bool inserts = valid && (lane == leader || !same);
      int32_t slot = 0;
      if (inserts)
        slot = table_insert<D>(table, mask, k64, key, r, (uint32_t)(p * D1 + r), greedy, rank,
                               overflow);
      __syncwarp();
      const int32_t leader_slot = __shfl_sync(0xffffffffu, slot, leader);
      if (valid) slot_out[p * D1 + r] = inserts ? slot : leader_slot;
    }
  }
}

// ---------------------------------------------------------------------------
// 3. first-touch numbering
// ---------------------------------------------------------------------------
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanChunk = kScanThreads * kScanItems;

__device__ __forceinline__ bool is_first_touch(const unsigned long long* __restrict__ table,
                                               const int32_t* __restrict__ slot, uint64_t e,
                                               uint64_t E) {
  return e < E && (uint32_t)table[slot[e]] == (uint32_t)e;
}

__global__ void __launch_bounds__(kScanThreads)
count_first_kernel(const unsigned long long* __restrict__ table, const int32_t* __restrict__ slot,
                   uint64_t E, uint32_t* __restrict__ chunk_count) {
  const uint64_t base = (uint64_t)blockIdx.x * kScanChunk;
  int c = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; k++)
    c += is_first_touch(table, slot, base + (uint64_t)k * kScanThreads + threadIdx.x, E) ? 1 : 0;
  __shared__ int warp_sum[kScanThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int s = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; w++) s += warp_sum[w];
    chunk_count[blockIdx.x] = (uint32_t)s;
  }
}

// single block: exclusive scan of chunk counts in place; total -> *total_out
__global__ void __launch_bounds__(1024)
scan_chunks_kernel(uint32_t* __restrict__ chunk_count, uint32_t nchunks, uint32_t* total_out) {
  __shared__ uint32_t warp_tot[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (uint32_t base = 0; base < nchunks; base += 1024) {
    uint32_t i = base + threadIdx.x;
    uint32_t v = i < nchunks ? chunk_count[i] : 0u;
    uint32_t x = v;  // inclusive warp scan
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) warp_tot[wid] = x;
    __syncthreads();
    if (wid == 0) {
      uint32_t t = warp_tot[lane];
      uint32_t s = t;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
        if (lane >= o) s += y;
      }
      warp_tot[lane] = s - t;  // exclusive warp offsets
    }
    __syncthreads();
    const uint32_t c = carry;
    if (i < nchunks) chunk_count[i] = c + warp_tot[wid] + (x - v);
    __syncthreads();
    if (threadIdx.x == 1023) carry = c + warp_tot[wid] + x;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total_out = carry;
}

template <int D>
__global__ void __launch_bounds__(kScanThreads)
assign_kernel(const unsigned long long* __restrict__ table, const int32_t* __restrict__ slot,
              uint64_t E, const uint32_t* __restrict__ chunk_offset,
              const int16_t* __restrict__ greedy, const int8_t* __restrict__ rank,
              int32_t* __restrict__ slot_vid, int16_t* __restrict__ vkeys) {
  __shared__ uint32_t warp_cnt[kScanThreads / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint64_t base = (uint64_t)blockIdx.x * kScanChunk;
  uint32_t running = chunk_offset[blockIdx.x];
#pragma unroll 1
  for (int k = 0; k < kScanItems; k++) {
    const uint64_t e = base + (uint64_t)k * kScanThreads + threadIdx.x;
    const bool first = is_first_touch(table, slot, e, E);
    const unsigned bal = __ballot_sync(0xffffffffu, first);
    if (lane == 0) warp_cnt[wid] = __popc(bal);
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; w++) {
      uint32_t c = warp_cnt[w];
      before += (w < wid) ? c : 0u;
      total += c;
    }
    if (first) {
      const uint32_t v = running + before + __popc(bal & ((1u << lane) - 1u));
      slot_vid[slot[e]] = (int32_t)v;
      int16_t key[D];
      const uint64_t p = e / (D + 1);
      load_key<D>(greedy, rank, (int64_t)p, (int)(e - p * (D + 1)), key);
#pragma unroll
      for (int i = 0; i < D; i++) vkeys[(uint64_t)v * D + i] = key[i];
    }
    running += total;
    __syncthreads();
  }
}

// 4. vid[e] = slot_vid[slot[e]] (in place over the slot array)
__global__ void __launch_bounds__(256)
gather_kernel(int32_t* __restrict__ slot_then_vid, const int32_t* __restrict__ slot_vid,
              uint64_t E) {
  for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E;
       e += (uint64_t)gridDim.x * blockDim.x)
    slot_then_vid[e] = slot_vid[slot_then_vid[e]];
}

// ---------------------------------------------------------------------------
// 5. blur neighbours                                permutohedral.h:503-522
// ---------------------------------------------------------------------------
template <int D>
__device__ __forceinline__ int32_t table_find(const unsigned long long* __restrict__ table,
                                              const int32_t* __restrict__ slot_vid,
                                              const int16_t* __restrict__ vkeys, uint32_t mask,
                                              const int16_t* key) {
  const uint64_t k64 = hash_key<D>(key);
  uint32_t h = (uint32_t)k64 & mask;
  const uint32_t fp = (uint32_t)(k64 >> 33);
  for (int probe = 0; probe < kProbeLimit; probe++) {
    const unsigned long long cur = table[h];
    if (cur == kEmpty) return -1;
    if ((uint32_t)(cur >> 32) == fp) {
      const int32_t v = slot_vid[h];
      bool same = true;
#pragma unroll
      for (int i = 0; i < D; i++) same = same && (vkeys[(int64_t)v * D + i] == key[i]);
      if (same) return v;
    }
    h = (h + 1) & mask;
  }
  return -1;
}

template <int D>
__global__ void __launch_bounds__(256)
neighbour_kernel(const unsigned long long* __restrict__ table,
                 const int32_t* __restrict__ slot_vid, const int16_t* __restrict__ vkeys,
                 uint32_t mask, int64_t M, int2* __restrict__ nbr) {
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < M;
       v += (int64_t)gridDim.x * blockDim.x) {
    int16_t key[D];
#pragma unroll
    for (int i = 0; i < D; i++) key[i] = vkeys[v * D + i];
#pragma unroll 1
    for (int j = 0; j <= D; j++) {
      int16_t n1[D], n2[D];
#pragma unroll
      for (int i = 0; i < D; i++) {
        // :505-509 (int16 arithmetic); for j == d the changed coordinate is
        // the implicit last one, which is not part of the stored key
        n1[i] = (int16_t)((i == j) ? key[i] - D : key[i] + 1);
        n2[i] = (int16_t)((i == j) ? key[i] + D : key[i] - 1);
      }
      const int32_t a = table_find<D>(table, slot_vid, vkeys, mask, n1);
      const int32_t b = table_find<D>(table, slot_vid, vkeys, mask, n2);
      // absent neighbours point at row M, which the value buffers keep at zero (:518,522)
      nbr[(int64_t)j * M + v] = make_int2(a < 0 ? (int)M : a, b < 0 ? (int)M : b);
    }
  }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
void reference_scale_factors(int d, float* out) {
  // permutohedral.h:354-371, same expression and types, evaluated on the host
  for (int i = 0; i < d; i++) {
    out[i] = 1.0f / (sqrtf((float)(i + 1) * (i + 2)));
    out[i] *= (d + 1) * sqrtf(2.0 / 3);
  }
}

uint64_t pow2_ceil(uint64_t x) {
  uint64_t p = 1;
  while (p < x) p <<= 1;
  return p;
}

// PL_BUILD_TRACE=1: wall-clock per build phase on stderr (synchronises the stream at each point)
struct BuildTrace {
  bool on;
  cudaStream_t st;
  std::chrono::steady_clock::time_point t0;
  explicit BuildTrace(cudaStream_t s) : on(getenv("PL_BUILD_TRACE") != nullptr), st(s) {
    if (on) { cudaStreamSynchronize(st); t0 = std::chrono::steady_clock::now(); }
  }
  void point(const char* what) {
    if (!on) return;
    cudaStreamSynchronize(st);
    auto t1 = std::chrono::steady_clock::now();
    fprintf(stderr, "[pl build] %-28s %8.3f ms\n", what,
            std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
};

template <int D>
int build_impl(Lattice* L, const float* ref, const GuideSrc& gs, cudaStream_t st) {
  BuildTrace trace(st);
  constexpr int D1 = D + 1;
  const int64_t n = L->n;
  const uint64_t E = (uint64_t)n * D1;
  if (E >= 0xFFFFFFF0ull) {
    set_last_error("n*(d+1) = %llu does not fit the 32-bit entry id", (unsigned long long)E);
    return PL_ERR_INVALID;
  }
  PL_TRY(dev_alloc((void**)&L->greedy, sizeof(int16_t) * E, st));
  PL_TRY(dev_alloc((void**)&L->rank, sizeof(int8_t) * E, st));
  PL_TRY(dev_alloc((void**)&L->bary, sizeof(float) * E, st));
  PL_TRY(dev_alloc((void**)&L->vid, sizeof(int32_t) * E, st));
  L->bytes += (int64_t)((sizeof(int16_t) + sizeof(int8_t) + sizeof(float) + sizeof(int32_t)) * E);

  trace.point("alloc point arrays");
  ScaleFactors sf;
  for (int i = 0; i < PL_MAX_D; i++) sf.s[i] = 0.f;
  reference_scale_factors(D, sf.s);
  PL_LAUNCH((elevate_kernel<D>), grid_for((uint64_t)n, 256, 8), 256, 0, st, ref, gs, n, sf, L->greedy,
            L->rank, L->bary);

  trace.point("elevate");
  const uint32_t nchunks = ceil_div_u(E, kScanChunk);
  uint32_t* chunk_count = nullptr;  // [nchunks] + total + overflow flag
  PL_TRY(dev_alloc((void**)&chunk_count, sizeof(uint32_t) * ((size_t)nchunks + 2), st));
  uint32_t* total_dev = chunk_count + nchunks;
  int* overflow_dev = (int*)(chunk_count + nchunks + 1);

  uint64_t cap = pow2_ceil(E + E / 4 + 1024);
  unsigned long long* table = nullptr;
  uint32_t host_res[2] = {0, 0};
  for (int attempt = 0;; attempt++) {
    if (cap > (1ull << 31)) {
      dev_free(chunk_count, st);
      set_last_error("hash table would need %llu slots", (unsigned long long)cap);
      return PL_ERR_OVERFLOW;
    }
    PL_TRY(dev_alloc((void**)&table, sizeof(unsigned long long) * cap, st));
    PL_CUDA(cudaMemsetAsync(table, 0xFF, sizeof(unsigned long long) * cap, st));
    PL_CUDA(cudaMemsetAsync(total_dev, 0, 2 * sizeof(uint32_t), st));
    trace.point("alloc + clear hash table");
    PL_LAUNCH((insert_kernel<D>), grid_for((uint64_t)n, 256, 8), 256, 0, st, L->greedy, L->rank, n,
              table, (uint32_t)(cap - 1), L->vid, overflow_dev);
    PL_LAUNCH(count_first_kernel, nchunks, kScanThreads, 0, st, table, L->vid, E, chunk_count);
    PL_LAUNCH(scan_chunks_kernel, 1, 1024, 0, st, chunk_count, nchunks, total_dev);
    PL_CUDA(cudaMemcpyAsync(host_res, total_dev, sizeof(host_res), cudaMemcpyDeviceToHost, st));
    PL_CUDA(cudaStreamSynchronize(st));
    if (host_res[1] == 0) break;
    dev_free(table, st);
    table = nullptr;
    cap *= 2;
    if (attempt >= 4) {
      dev_free(chunk_count, st);
      set_last_error("hash insert overflowed %d times", attempt + 1);
      return PL_ERR_OVERFLOW;
    }
  }
  const int64_t M = (int64_t)host_res[0];
  L->M = M;
  trace.point("insert + count + scan + sync");

  int32_t* slot_vid = nullptr;
  PL_TRY(dev_alloc((void**)&slot_vid, sizeof(int32_t) * cap, st));
  PL_TRY(dev_alloc((void**)&L->vkeys, sizeof(int16_t) * (size_t)M * D, st));
  PL_TRY(dev_alloc((void**)&L->nbr, sizeof(int2) * (size_t)M * D1, st));
  L->bytes += (int64_t)(sizeof(int16_t) * (size_t)M * D + sizeof(int2) * (size_t)M * D1);

  PL_LAUNCH((assign_kernel<D>), nchunks, kScanThreads, 0, st, table, L->vid, E, chunk_count,
            L->greedy, L->rank, slot_vid, L->vkeys);
  PL_LAUNCH(gather_kernel, grid_for(E, 256, 8), 256, 0, st, L->vid, slot_vid, E);
  PL_LAUNCH((neighbour_kernel<D>), grid_for((uint64_t)M, 256, 8), 256, 0, st, table, slot_vid,
            L->vkeys, (uint32_t)(cap - 1), M, L->nbr);

  trace.point("assign + gather + neighbours");
  dev_free(slot_vid, st);
  dev_free(table, st);
  dev_free(chunk_count, st);
  const int rt = lattice_build_tiles(L, st);
  trace.point("tile view");
  return rt;
}

}  // namespace

int lattice_build(Lattice* L, const float* ref, cudaStream_t st, const GuideSrc* guide) {
  GuideSrc gs;
  if (guide != nullptr) gs = *guide;
  switch (L->d) {
#define PL_CASE(DD) \
  case DD:          \
    return build_impl<DD>(L, ref, gs, st);
    PL_CASE(1) PL_CASE(2) PL_CASE(3) PL_CASE(4) PL_CASE(5) PL_CASE(6) PL_CASE(7) PL_CASE(8)
    PL_CASE(9) PL_CASE(10) PL_CASE(11) PL_CASE(12) PL_CASE(13) PL_CASE(14) PL_CASE(15) PL_CASE(16)
#undef PL_CASE
    default:
      set_last_error("d = %d outside 1..%d", L->d, PL_MAX_D);
      return PL_ERR_INVALID;
  }
}

}  // namespace pl
