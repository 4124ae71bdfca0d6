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
      const float t = __fmul_rn(__fsub_rn(el[i], (float)gr[i]), scale);
      const int ka = D - rk[i];
      // pure selects (the `if` form became a local-memory array with branches: 73 % of this kernel's
      // stall samples); per slot the same additions happen in the same order
#pragma unroll
      for (int k = 0; k < D + 2; k++) {
        const float up = __fadd_rn(b[k], t), dn = __fsub_rn(b[k], t);
        b[k] = (k == ka) ? up : ((k == ka + 1) ? dn : b[k]);
      }
    }
    b[0] = __fadd_rn(b[0], __fadd_rn(1.0f, b[D + 1]));

    if (D1 % 2 == 0) {
      // packed stores (d = 5: three 4-byte, three 2-byte and three 8-byte stores per point instead of
      // eighteen scalar ones; p * D1 elements keep every pointer naturally aligned for even D1)
      uint32_t* g32 = reinterpret_cast<uint32_t*>(greedy_out + p * D1);
      uint16_t* r16 = reinterpret_cast<uint16_t*>(rank_out + p * D1);
      float2* b2 = reinterpret_cast<float2*>(bary_out + p * D1);
#pragma unroll
      for (int i = 0; i < D1 / 2; i++) {
        g32[i] = (uint32_t)(uint16_t)(int16_t)gr[2 * i] | ((uint32_t)(uint16_t)(int16_t)gr[2 * i + 1] << 16);
        r16[i] = (uint16_t)((uint8_t)(int8_t)rk[2 * i] | ((uint16_t)(uint8_t)(int8_t)rk[2 * i + 1] << 8));
        b2[i] = make_float2(b[2 * i], b[2 * i + 1]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < D1; i++) {
        greedy_out[p * D1 + i] = (int16_t)gr[i];
        rank_out[p * D1 + i] = (int8_t)rk[i];
        bary_out[p * D1 + i] = b[i];
      }
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
              uint64_t E, const uint32_t* __restrict__ chunk_offset, int32_t* __restrict__ slot_vid) {
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

// 4b. per vertex: its key (and, for the growth-quirk replay, where and when it was created).  One
// thread per table slot: an occupied slot holds the entry id of its key's first touch, and
// assign_kernel has given it the vertex id.  This is the only build kernel in front of the neighbour
// table whose OUTPUT is sized by M, so it is the first one that has to wait for M on the host:
// assign and gather are queued before the host blocks.
template <int D>
__global__ void __launch_bounds__(256)
keys_kernel(const unsigned long long* __restrict__ table, const int32_t* __restrict__ slot_vid,
            uint64_t cap, const int16_t* __restrict__ greedy, const int8_t* __restrict__ rank,
            int16_t* __restrict__ vkeys, uint32_t* __restrict__ first_e, int32_t* __restrict__ vslot) {
  const ulonglong2* t2 = reinterpret_cast<const ulonglong2*>(table);
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < cap / 2;
       i += (uint64_t)gridDim.x * blockDim.x) {
    const ulonglong2 pair = t2[i];
#pragma unroll
    for (int k = 0; k < 2; k++) {
      const unsigned long long cur = k == 0 ? pair.x : pair.y;
      if (cur == kEmpty) continue;
      const uint64_t h = 2 * i + k;
      const uint32_t e = (uint32_t)cur;
      const int32_t v = slot_vid[h];
      if (first_e != nullptr) {
        first_e[v] = e;
        vslot[v] = (int32_t)h;
      }
      int16_t key[D];
      const uint32_t p = e / (D + 1);
      load_key<D>(greedy, rank, (int64_t)p, (int)(e - p * (D + 1)), key);
#pragma unroll
      for (int j = 0; j < D; j++) vkeys[(uint64_t)v * D + j] = key[j];
    }
  }
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
      // slot_key: vertex whose stored key describes this slot (the growth-quirk replay may point
      // the slot at a later copy of the same key, or at nothing: v < 0 encodes -(key vertex) - 2)
      const int32_t kv = v >= 0 ? v : -(v + 2);
      bool same = true;
#pragma unroll
      for (int i = 0; i < D; i++) same = same && (vkeys[(int64_t)kv * D + i] == key[i]);
      if (same) return v >= 0 ? v : -1;
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
  // one thread per (vertex, axis, side): a vertex's 2(d+1) probes run side by side instead of one
  // after the other (the kernel is pure probe latency: 23 -> 6 us at M = 8 750)
  constexpr int PER = 2 * (D + 1);
  const int64_t total = M * PER;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t v = t / PER;
    const int q = (int)(t - v * PER);
    const int j = q >> 1, side = q & 1;
    int16_t nk[D];
#pragma unroll
    for (int i = 0; i < D; i++) {
      // :505-509 (int16 arithmetic); for j == d the changed coordinate is the implicit last one,
      // which is not part of the stored key
      const int16_t k = vkeys[v * D + i];
      nk[i] = side == 0 ? (int16_t)((i == j) ? k - D : k + 1) : (int16_t)((i == j) ? k + D : k - 1);
    }
    const int32_t a = table_find<D>(table, slot_vid, vkeys, mask, nk);
    // absent neighbours point at row M, which the value buffers keep at zero (:518,522)
    reinterpret_cast<int*>(nbr + (int64_t)j * M + v)[side] = a < 0 ? (int)M : a;
  }
}

// ---------------------------------------------------------------------------
// 6. the reference's hash-growth quirk (PL_BUILD_REF_GROW_QUIRK)       permutohedral.h:59-62,101-155
//
// HashTablePermutohedral::lookup computes h = hash % capacity and then lookupOffset() may grow()
// the table before it probes from h.  The ONE lookup that triggers each doubling (the first
// lookup after `filled` reaches capacity/2 - 1) therefore probes from a slot computed with the
// old capacity.  If bit log2(old capacity) of its hash is set, it misses an existing key (or
// parks a new key where later lookups cannot find it) and creates an extra vertex with a
// duplicate key.  After the NEXT doubling the rehash walks the old slots in ascending order, so
// the parked copy (lower slot) is re-placed ahead of the regular one and becomes the copy every
// later lookup finds.  Everything else about the table (slot positions, probe lengths) is
// invisible to the filter.
//
// The clean build above yields one vertex per distinct key in first-touch order.  The replay
// below walks the <= 17 doubling events in scanline order, decides for each what the reference
// did, and emits a patch: the creation entries of the extra vertices, and for the handful of
// "special" keys a timeline (from entry x on, lookups return the copy created at entry c).
// Assumptions (both hold unless a probe run is >= capacity/2 long or wraps round the table end,
// probability ~1e-4 per doubling on a table that is at most a quarter full): the stale probe
// reaches an empty slot before a matching key, and slot order of two copies = order of their
// home slots.  Checked against the unmodified reference in tests/ (M, vertex ids, keys,
// neighbours bit-exact).
// ---------------------------------------------------------------------------
constexpr int kQMaxExtra = 64;
constexpr int kQMaxSpecial = 40;
constexpr int kQMaxSegs = 160;
constexpr uint32_t kQNone = 0xFFFFFFFFu;

struct GrowQuirk {
  int n_extras;
  uint32_t extra_entry[kQMaxExtra];   // creation entries of the extra vertices, ascending
  int32_t extra_vertex[kQMaxExtra];   // clean vertex (= key) each one duplicates
  int32_t extra_id[kQMaxExtra];       // final vertex id
  int n_special;
  int32_t special_v[kQMaxSpecial];    // clean vertex ids with more than one copy / a parked copy
  uint32_t special_live[kQMaxSpecial];    // creation entry of the copy lookups find at the end (kQNone: none)
  uint32_t special_parked[kQMaxSpecial];  // copy that becomes live at the next doubling
  int32_t special_live_id[kQMaxSpecial];  // final id of the live copy, -1 none
  int n_segs;
  int32_t seg_v[kQMaxSegs];
  uint32_t seg_start[kQMaxSegs];      // from this entry on ...
  uint32_t seg_copy[kQMaxSegs];       // ... lookups of seg_v return the copy created at this entry
  int32_t seg_id[kQMaxSegs];          // final id of that copy
  int blur_trigger;                   // the last doubling happens at the first lookup of blur()
  unsigned long long blur_cap_old;
  int error;                          // small tables exhausted
};

__device__ __forceinline__ unsigned long long ref_hash(const int16_t* key, int D) {
  // permutohedral.h:109-116: size_t arithmetic, shorts sign-extended
  unsigned long long k = 0;
  for (int i = 0; i < D; i++) {
    k += (unsigned long long)(long long)key[i];
    k *= 2531011ull;
  }
  return k;
}

// number of elements of the ascending array a[0..n) that are < x  (<= x with `incl`)
__device__ __forceinline__ uint32_t count_below(const uint32_t* a, uint32_t n, uint32_t x, bool incl) {
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    const uint32_t v = a[mid];
    if (v < x || (incl && v == x)) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

template <int D>
__global__ void __launch_bounds__(1024)
grow_replay_kernel(const int32_t* __restrict__ vid, const uint32_t* __restrict__ first_e,
                   const int16_t* __restrict__ vkeys, uint32_t E, uint32_t M, GrowQuirk* out) {
  __shared__ GrowQuirk q;
  __shared__ unsigned long long cap;
  __shared__ int phase;             // 0 run, 1 need next touch, 2 done
  __shared__ int32_t touch_v;
  __shared__ uint32_t touch_after, touch_found;
  __shared__ int pend_special;      // special index whose second touch is unconfirmed, -1 none
  __shared__ uint32_t pend_entry;
  __shared__ int cur_special;
  const int tid = threadIdx.x;

  auto find_special = [&](int32_t v) {
    for (int i = 0; i < q.n_special; i++)
      if (q.special_v[i] == v) return i;
    return -1;
  };
  auto add_seg = [&](int32_t v, uint32_t start, uint32_t copy) {
    if (q.n_segs >= kQMaxSegs) { q.error = 1; return; }
    q.seg_v[q.n_segs] = v; q.seg_start[q.n_segs] = start; q.seg_copy[q.n_segs] = copy;
    q.n_segs++;
  };
  auto add_extra = [&](uint32_t e, int32_t v) {
    if (q.n_extras >= kQMaxExtra) { q.error = 1; return; }
    int i = q.n_extras++;
    while (i > 0 && q.extra_entry[i - 1] > e) {
      q.extra_entry[i] = q.extra_entry[i - 1];
      q.extra_vertex[i] = q.extra_vertex[i - 1];
      i--;
    }
    q.extra_entry[i] = e;
    q.extra_vertex[i] = v;
  };
  auto remove_extra = [&](uint32_t e) {
    int i = 0;
    while (i < q.n_extras && q.extra_entry[i] != e) i++;
    if (i == q.n_extras) return;
    for (; i + 1 < q.n_extras; i++) {
      q.extra_entry[i] = q.extra_entry[i + 1];
      q.extra_vertex[i] = q.extra_vertex[i + 1];
    }
    q.n_extras--;
  };
  // entry at which `filled` first reaches T: the T-th smallest of first_e U extras (kQNone: never)
  auto first_reach = [&](unsigned long long T) -> uint32_t {
    if (T == 0) return 0;
    const int m = q.n_extras;
    if ((unsigned long long)M + (unsigned)m < T) return kQNone;
    for (int j = 0; j <= m; j++) {      // j = extras among the T smallest creations
      if ((unsigned long long)j > T) break;
      const unsigned long long c = T - j;   // clean creations among them
      if (c > M) continue;
      const uint32_t last_clean = c > 0 ? first_e[c - 1] : 0u;
      const uint32_t last_extra = j > 0 ? q.extra_entry[j - 1] : 0u;
      if (j < m && c > 0 && !(last_clean < q.extra_entry[j])) continue;   // next extra would be inside
      if (j > 0 && c < M && !(last_extra < first_e[c])) continue;        // next clean would be inside
      if (c == 0 && j == 0) continue;
      return c > 0 && (j == 0 || last_clean > last_extra) ? last_clean : last_extra;
    }
    return kQNone;
  };

  if (tid == 0) {
    q.n_extras = 0; q.n_special = 0; q.n_segs = 0; q.blur_trigger = 0; q.blur_cap_old = 0; q.error = 0;
    cap = 1ull << 15;    // permutohedral.h:36
    pend_special = -1;
    phase = 0;
  }
  __syncthreads();

  for (;;) {
    if (tid == 0) {
      phase = 0;
      const unsigned long long T = cap / 2 - 1;             // :62
      uint32_t ec = first_reach(T);
      if (ec == kQNone) {
        phase = 2;
      } else {
        unsigned long long t = (unsigned long long)ec + 1;  // the lookup that grows the table
        if (pend_special >= 0) {
          if ((unsigned long long)pend_entry >= t) {
            // the doubling comes first: the parked copy is live again before the key is touched,
            // no second copy is created
            remove_extra(pend_entry);
            for (int i = 0; i < q.n_segs; i++)
              if (q.seg_v[i] == q.special_v[pend_special] && q.seg_start[i] == pend_entry &&
                  q.seg_copy[i] == pend_entry) {
                for (int k = i; k + 1 < q.n_segs; k++) {
                  q.seg_v[k] = q.seg_v[k + 1]; q.seg_start[k] = q.seg_start[k + 1];
                  q.seg_copy[k] = q.seg_copy[k + 1];
                }
                q.n_segs--;
                break;
              }
            q.special_live[pend_special] = kQNone;
            ec = first_reach(T);
            t = (unsigned long long)ec + 1;
          }
          pend_special = -1;
        }
        if (ec == kQNone) {
          phase = 2;
        } else if (t >= E) {
          // the table doubles at the first lookup of blur(): parked copies take over there
          q.blur_trigger = 1;
          q.blur_cap_old = cap;
          for (int i = 0; i < q.n_special; i++)
            if (q.special_parked[i] != kQNone) {
              q.special_live[i] = q.special_parked[i];
              q.special_parked[i] = kQNone;
            }
          cap *= 2;
          phase = 2;
        } else {
          const uint32_t te = (uint32_t)t;
          for (int i = 0; i < q.n_special; i++)
            if (q.special_parked[i] != kQNone) {
              q.special_live[i] = q.special_parked[i];
              add_seg(q.special_v[i], te, q.special_parked[i]);
              q.special_parked[i] = kQNone;
            }
          const unsigned long long cap_old = cap;
          cap *= 2;
          const int32_t v = vid[te];
          const unsigned long long H = ref_hash(vkeys + (int64_t)v * D, D);
          int si = find_special(v);
          const uint32_t fe = first_e[v];
          uint32_t live = si >= 0 ? q.special_live[si] : (fe < te ? fe : kQNone);
          if ((H % cap_old) == (H % cap)) {
            // regular lookup; a special key nobody can find gets its regular copy here
            if (si >= 0 && live == kQNone) {
              add_extra(te, v);
              q.special_live[si] = te;
              add_seg(v, te, te);
            }
          } else {
            if (si < 0) {
              if (q.n_special >= kQMaxSpecial) { q.error = 1; phase = 2; }
              else {
                si = q.n_special++;
                q.special_v[si] = v;
                q.special_live[si] = live;
                q.special_parked[si] = kQNone;
                if (live != kQNone) add_seg(v, live, live);
              }
            }
            if (si >= 0) {
              if (fe != te) add_extra(te, v);       // fe == te: this creation is the key's first touch
              q.special_parked[si] = te;
              add_seg(v, te, te);                   // the triggering entry itself uses the parked copy
              if (live != kQNone) {
                add_seg(v, te + 1, live);           // later lookups find the regular copy again
              } else {
                q.special_live[si] = kQNone;
                phase = 1;                          // the next touch of the key creates a regular copy
                touch_v = v;
                touch_after = te;
                cur_special = si;
              }
            }
          }
        }
      }
      if (q.error) phase = 2;
      touch_found = kQNone;
    }
    __syncthreads();
    if (phase == 2) break;
    if (phase == 1) {
      const int32_t v = touch_v;
      // 32 768 entries per trip: eight int4 loads per thread (vid is 16-byte aligned, E a multiple of 2)
      const unsigned long long first = (unsigned long long)touch_after + 1;
      for (unsigned long long base = first & ~3ull; base < E; base += 32768) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
          const unsigned long long e4 = base + ((unsigned long long)k * 1024 + tid) * 4;
          if (e4 + 3 < E) {
            const int4 x = *reinterpret_cast<const int4*>(vid + e4);
            uint32_t hit = kQNone;
            if (x.w == v && e4 + 3 >= first) hit = (uint32_t)(e4 + 3);
            if (x.z == v && e4 + 2 >= first) hit = (uint32_t)(e4 + 2);
            if (x.y == v && e4 + 1 >= first) hit = (uint32_t)(e4 + 1);
            if (x.x == v && e4 >= first) hit = (uint32_t)e4;
            if (hit != kQNone) atomicMin(&touch_found, hit);
          } else {
            for (unsigned long long e = e4 > first ? e4 : first; e < E && e < e4 + 4; e++)
              if (vid[e] == v) atomicMin(&touch_found, (uint32_t)e);
          }
        }
        __syncthreads();
        const uint32_t found = touch_found;
        __syncthreads();   // nobody starts the next chunk's atomics before everybody has read
        if (found != kQNone) break;
      }
      if (tid == 0 && touch_found != kQNone) {
        add_extra(touch_found, v);
        add_seg(v, touch_found, touch_found);
        q.special_live[cur_special] = touch_found;
        pend_special = cur_special;
        pend_entry = touch_found;
      }
    }
    __syncthreads();
  }

  // final ids: id(copy created at entry c) = #clean creations < c + #extras < c
  auto copy_id = [&](uint32_t c) -> int32_t {
    return (int32_t)(count_below(first_e, M, c, false) + count_below(q.extra_entry, q.n_extras, c, false));
  };
  __syncthreads();
  for (int i = tid; i < q.n_segs; i += blockDim.x) q.seg_id[i] = copy_id(q.seg_copy[i]);
  for (int i = tid; i < q.n_extras; i += blockDim.x) q.extra_id[i] = copy_id(q.extra_entry[i]);
  for (int i = tid; i < q.n_special; i += blockDim.x)
    q.special_live_id[i] = q.special_live[i] == kQNone ? -1 : copy_id(q.special_live[i]);
  __syncthreads();
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&q);
    uint32_t* dst = reinterpret_cast<uint32_t*>(out);
    for (int i = tid; i < (int)(sizeof(GrowQuirk) / 4); i += blockDim.x) dst[i] = src[i];
  }
}

// vertex id of entry e after the replay
__device__ __forceinline__ int32_t quirk_vid(const GrowQuirk& q, const uint32_t* __restrict__ first_e,
                                             int32_t v, uint32_t e) {
  for (int i = 0; i < q.n_special; i++) {
    if (q.special_v[i] != v) continue;
    int best = -1;
    for (int k = 0; k < q.n_segs; k++)    // latest segment that has started; ties: the later one
      if (q.seg_v[k] == v && q.seg_start[k] <= e && (best < 0 || q.seg_start[k] >= q.seg_start[best]))
        best = k;
    if (best >= 0) return q.seg_id[best];
  }
  return v + (int32_t)count_below(q.extra_entry, (uint32_t)q.n_extras, first_e[v], false);
}

__global__ void __launch_bounds__(256)
quirk_renumber_kernel(int32_t* __restrict__ vid, const uint32_t* __restrict__ first_e,
                      const GrowQuirk* __restrict__ qg, uint64_t E) {
  __shared__ GrowQuirk q;
  for (int i = threadIdx.x; i < (int)(sizeof(GrowQuirk) / 4); i += blockDim.x)
    reinterpret_cast<uint32_t*>(&q)[i] = reinterpret_cast<const uint32_t*>(qg)[i];
  __syncthreads();
  for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E;
       e += (uint64_t)gridDim.x * blockDim.x)
    vid[e] = quirk_vid(q, first_e, vid[e], (uint32_t)e);
}

// keys in the final numbering, and the hash table's slot -> vertex map pointed at the copy that
// lookups find once splat() is over (what blur()'s neighbour lookups see)
template <int D>
__global__ void __launch_bounds__(256)
quirk_keys_kernel(const int16_t* __restrict__ vkeys_clean, const uint32_t* __restrict__ first_e,
                  const int32_t* __restrict__ vslot, const GrowQuirk* __restrict__ qg, int64_t M,
                  int16_t* __restrict__ vkeys_final, int32_t* __restrict__ slot_vid) {
  __shared__ GrowQuirk q;
  for (int i = threadIdx.x; i < (int)(sizeof(GrowQuirk) / 4); i += blockDim.x)
    reinterpret_cast<uint32_t*>(&q)[i] = reinterpret_cast<const uint32_t*>(qg)[i];
  __syncthreads();
  for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < M + q.n_extras;
       v += (int64_t)gridDim.x * blockDim.x) {
    if (v < M) {
      const int32_t id = (int32_t)v + (int32_t)count_below(q.extra_entry, (uint32_t)q.n_extras, first_e[v], false);
#pragma unroll
      for (int i = 0; i < D; i++) vkeys_final[(int64_t)id * D + i] = vkeys_clean[v * D + i];
      int32_t live = id;
      for (int i = 0; i < q.n_special; i++)
        if (q.special_v[i] == (int32_t)v)
          live = q.special_live_id[i] >= 0 ? q.special_live_id[i] : -(id + 2);   // unfindable: key vertex encoded
      slot_vid[vslot[v]] = live;
    } else {
      const int x = (int)(v - M);
      const int32_t src = q.extra_vertex[x];
#pragma unroll
      for (int i = 0; i < D; i++) vkeys_final[(int64_t)q.extra_id[x] * D + i] = vkeys_clean[(int64_t)src * D + i];
    }
  }
}

// blur()'s very first lookup (axis 0, vertex 0, the "minus" neighbour) is the one that grows the
// table when splat() ended exactly on the threshold: it probes from the stale slot and, if the bit
// is set, reports the neighbour as absent (create = false, permutohedral.h:516)
template <int D>
__global__ void quirk_blur_trigger_kernel(const GrowQuirk* __restrict__ qg,
                                          const int16_t* __restrict__ vkeys, int64_t M,
                                          int2* __restrict__ nbr) {
  if (threadIdx.x != 0 || blockIdx.x != 0 || !qg->blur_trigger || M <= 0) return;
  int16_t n1[D];
#pragma unroll
  for (int i = 0; i < D; i++) n1[i] = (int16_t)((i == 0) ? vkeys[i] - D : vkeys[i] + 1);
  const unsigned long long H = ref_hash(n1, D);
  if ((H % qg->blur_cap_old) != (H % (2 * qg->blur_cap_old))) nbr[0].x = (int)M;
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
  uint32_t* chunk_count = nullptr;  // [nchunks]
  PL_TRY(dev_alloc((void**)&chunk_count, sizeof(uint32_t) * (size_t)nchunks, st));

  uint64_t cap = pow2_ceil(E + E / 4 + 1024);
  unsigned long long* table = nullptr;
  int32_t* slot_vid = nullptr;
  // M and the overflow flag come back through a pinned slot, and the host waits on an event recorded
  // right behind that copy: assign + gather, which need neither, are queued before the host blocks, so
  // the device still has work when the host resumes (a stream synchronise here left it idle for
  // ~25 us per frame, and the launches after it could not catch up before the first filter call)
  uint32_t pageable_res[2] = {0, 0};
  uint32_t* host_res = reinterpret_cast<uint32_t*>(pinned_slot());
  const bool pinned = host_res != nullptr;
  if (!pinned) host_res = pageable_res;
  for (int attempt = 0;; attempt++) {
    if (cap > (1ull << 31)) {
      dev_free(chunk_count, st);
      set_last_error("hash table would need %llu slots", (unsigned long long)cap);
      return PL_ERR_OVERFLOW;
    }
    // one more slot behind the table holds {vertex count, overflow flag}, so that the clear of the table
    // covers them: the count is written by the scan, the flag reads 0xFFFFFFFF until an insert overflows
    PL_TRY(dev_alloc((void**)&table, sizeof(unsigned long long) * (cap + 1), st));
    PL_TRY(dev_alloc((void**)&slot_vid, sizeof(int32_t) * cap, st));
    uint32_t* total_dev = reinterpret_cast<uint32_t*>(table + cap);
    int* overflow_dev = reinterpret_cast<int*>(total_dev + 1);
    PL_CUDA(cudaMemsetAsync(table, 0xFF, sizeof(unsigned long long) * (cap + 1), st));
    trace.point("alloc + clear hash table");
    PL_LAUNCH((insert_kernel<D>), grid_for((uint64_t)n, 256, 8), 256, 0, st, L->greedy, L->rank, n,
              table, (uint32_t)(cap - 1), L->vid, overflow_dev);
    PL_LAUNCH(count_first_kernel, nchunks, kScanThreads, 0, st, table, L->vid, E, chunk_count);
    PL_LAUNCH(scan_chunks_kernel, 1, 1024, 0, st, chunk_count, nchunks, total_dev);
    if (pinned)
      PL_TRY(readback_small(L, host_res, total_dev, 2 * sizeof(uint32_t), st));
    else
      PL_CUDA(cudaMemcpyAsync(host_res, total_dev, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    // after an overflow these two run on an incomplete table: every slot index is still inside the
    // table, their output is discarded, and insert_kernel rewrites L->vid on the retry
    PL_LAUNCH((assign_kernel<D>), nchunks, kScanThreads, 0, st, table, L->vid, E, chunk_count, slot_vid);
    PL_LAUNCH(gather_kernel, grid_for(E, 256, 8), 256, 0, st, L->vid, slot_vid, E);
    if (pinned)
      PL_CUDA(cudaEventSynchronize(L->max_U_ev));
    else
      PL_CUDA(cudaStreamSynchronize(st));
    if (host_res[1] == 0xFFFFFFFFu) break;     // no insert overflowed
    dev_free(table, st);
    dev_free(slot_vid, st);
    table = nullptr;
    slot_vid = nullptr;
    cap *= 2;
    if (attempt >= 4) {
      dev_free(chunk_count, st);
      set_last_error("hash insert overflowed %d times", attempt + 1);
      return PL_ERR_OVERFLOW;
    }
  }
  const int64_t M = (int64_t)host_res[0];
  L->M = M;
  trace.point("insert + count + scan + assign + gather queued, M read back");

  // the reference's table only ever grows once it holds 2^14 - 1 vertices (permutohedral.h:36,62)
  const bool quirk = (L->flags & PL_BUILD_REF_GROW_QUIRK) != 0 && M >= (1 << 14) - 1;
  int16_t* vkeys_clean = nullptr;
  uint32_t* first_e = nullptr;
  int32_t* vslot = nullptr;
  PL_TRY(dev_alloc((void**)&vkeys_clean, sizeof(int16_t) * (size_t)M * D, st));
  if (quirk) {
    PL_TRY(dev_alloc((void**)&first_e, sizeof(uint32_t) * (size_t)M, st));
    PL_TRY(dev_alloc((void**)&vslot, sizeof(int32_t) * (size_t)M, st));
  }
  PL_LAUNCH((keys_kernel<D>), grid_for(cap / 2, 256, 8), 256, 0, st, table, slot_vid, cap, L->greedy, L->rank,
            vkeys_clean, first_e, vslot);
  int64_t Mf = M;
  if (quirk) {
    GrowQuirk* qd = nullptr;
    PL_TRY(dev_alloc((void**)&qd, sizeof(GrowQuirk), st));
    PL_LAUNCH((grow_replay_kernel<D>), 1, 1024, 0, st, L->vid, first_e, vkeys_clean, (uint32_t)E,
              (uint32_t)M, qd);
    int qres[1] = {0};
    int qerr = 0;
    PL_CUDA(cudaMemcpyAsync(qres, &qd->n_extras, sizeof(int), cudaMemcpyDeviceToHost, st));
    PL_CUDA(cudaMemcpyAsync(&qerr, &qd->error, sizeof(int), cudaMemcpyDeviceToHost, st));
    PL_CUDA(cudaStreamSynchronize(st));
    if (qerr != 0) {
      set_last_error("growth-quirk replay ran out of table space (%d extras)", qres[0]);
      return PL_ERR_OVERFLOW;
    }
    Mf = M + qres[0];
    PL_TRY(dev_alloc((void**)&L->vkeys, sizeof(int16_t) * (size_t)Mf * D, st));
    PL_LAUNCH(quirk_renumber_kernel, grid_for(E, 256, 8), 256, 0, st, L->vid, first_e, qd, E);
    PL_LAUNCH((quirk_keys_kernel<D>), grid_for((uint64_t)Mf, 256, 8), 256, 0, st, vkeys_clean, first_e,
              vslot, qd, M, L->vkeys, slot_vid);
    PL_TRY(dev_alloc((void**)&L->nbr, sizeof(int2) * (size_t)Mf * D1, st));
    PL_LAUNCH((neighbour_kernel<D>), grid_for((uint64_t)Mf * 2 * D1, 256, 8), 256, 0, st, table, slot_vid,
              L->vkeys, (uint32_t)(cap - 1), Mf, L->nbr);
    PL_LAUNCH((quirk_blur_trigger_kernel<D>), 1, 32, 0, st, qd, L->vkeys, Mf, L->nbr);
    dev_free(qd, st);
    dev_free(vkeys_clean, st);
    dev_free(first_e, st);
    dev_free(vslot, st);
    L->M = Mf;
  } else {
    L->vkeys = vkeys_clean;
    PL_TRY(dev_alloc((void**)&L->nbr, sizeof(int2) * (size_t)M * D1, st));
    PL_LAUNCH((neighbour_kernel<D>), grid_for((uint64_t)M * 2 * D1, 256, 8), 256, 0, st, table, slot_vid,
              L->vkeys, (uint32_t)(cap - 1), M, L->nbr);
  }
  L->bytes += (int64_t)(sizeof(int16_t) * (size_t)Mf * D + sizeof(int2) * (size_t)Mf * D1);

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
