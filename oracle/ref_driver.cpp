// ref_driver.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Compiles the reference's own CPU lattice (crf/lattice/lite/permutohedral.h)
// IN PLACE from /root/reference -- the header is included by absolute path
// (-DREF_HEADER=...), never copied -- and exposes it through a small C ABI so
// the tests can (a) pin oracle/lattice_oracle.c against the real thing and
// (b) use it as the "reference" CPU baseline.  Output goes to oracle/_ref/
// (git-ignored).  Built by oracle/build_ref.py.
//
// The reference header does not compile against torch 2.11 as committed:
// permutohedral.h:285 calls torch::CPU(kFloat).tensorFromBlob(...), removed
// from ATen long ago.  Instead of patching a copy we give the expression a
// meaning from outside: `CPU` is macro-renamed to a shim declared below in
// namespace torch whose tensorFromBlob forwards to torch::from_blob.  That
// statement is the last line of PermutohedralLattice::filter (wrapping the
// finished output buffer); no arithmetic is touched.
#include <torch/torch.h>
#include <ATen/ATen.h>
#include <cstdint>
#include <cstring>
// everything the reference header includes, so the macros below cannot leak
// into system headers
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>

namespace torch {
struct OracleBlobShim {
  at::Tensor tensorFromBlob(float* data, at::IntArrayRef sizes, void (*deleter)(void*)) {
    // the reference's deleter only prints (permutohedral.h:171-177), i.e. the
    // buffer leaks; free it properly here.
    (void)deleter;
    return torch::from_blob(data, sizes, [](void* p) { delete[] static_cast<float*>(p); },
                            at::TensorOptions().dtype(at::kFloat));
  }
};
inline OracleBlobShim oracle_cpu_shim(at::ScalarType) { return OracleBlobShim(); }
}  // namespace torch

#define CPU oracle_cpu_shim
#define private public   // replay[] (permutohedral.h:558-562) is private; we only read it
#define printf(...) ((void)0)   // the reference narrates every stage on stdout
#include REF_HEADER
#undef printf
#undef private
#undef CPU

extern "C" {

// lattice.cpp:6-10: filter(src[n,C], ref[n,d]) -> [n,C]
void ref_filter(const float* src, const float* ref, int n, int C, int d, float* out) {
  at::Tensor s = torch::from_blob(const_cast<float*>(src), {n, C}, at::kFloat);
  at::Tensor r = torch::from_blob(const_cast<float*>(ref), {n, d}, at::kFloat);
  at::Tensor o = PermutohedralLattice::filter(s, r);
  std::memcpy(out, o.data_ptr<float>(), sizeof(float) * (size_t)n * C);
}

// Runs the reference's splat() point by point and reads back its integer
// state after each call.  Returns M (hashTable.size()).
//   greedy [n*(d+1)] int16, rank [n*(d+1)] int8, weight [n*(d+1)] f32,
//   vid [n*(d+1)] int32 (replay offset / vd), keys [cap*d] int16 (first M rows)
int ref_state(const float* ref, int n, int d, int16_t* greedy, int8_t* rank, float* weight,
              int32_t* vid, int16_t* keys, int cap) {
  const int vd = 1;
  PermutohedralLattice lat(d, vd, n);
  float one = 1.0f;
  for (int p = 0; p < n; p++) {
    lat.splat(const_cast<float*>(ref) + (size_t)p * d, &one);
    for (int i = 0; i <= d; i++) {
      greedy[(size_t)p * (d + 1) + i] = lat.greedy[i];
      rank[(size_t)p * (d + 1) + i] = (int8_t)lat.rank[i];
    }
  }
  for (size_t k = 0; k < (size_t)n * (d + 1); k++) {
    weight[k] = lat.replay[k].weight;
    vid[k] = lat.replay[k].offset / vd;
  }
  int M = lat.hashTable.size();
  int m = M < cap ? M : cap;
  std::memcpy(keys, lat.hashTable.getKeys(), sizeof(int16_t) * (size_t)m * d);
  return M;
}

// Blur neighbours as the reference resolves them (permutohedral.h:503-522):
// nbr[(j*M+v)*2+{0,1}] = vertex id of key-neighbour or -1.
int ref_neighbours(const float* ref, int n, int d, int32_t* nbr, int cap) {
  const int vd = 1;
  PermutohedralLattice lat(d, vd, n);
  float one = 1.0f;
  for (int p = 0; p < n; p++) lat.splat(const_cast<float*>(ref) + (size_t)p * d, &one);
  int M = lat.hashTable.size();
  if (M > cap) return M;
  short* n1 = new short[d + 1];
  short* n2 = new short[d + 1];
  for (int j = 0; j <= d; j++)
    for (int v = 0; v < M; v++) {
      short* key = lat.hashTable.getKeys() + (size_t)v * d;
      for (int k = 0; k < d; k++) { n1[k] = key[k] + 1; n2[k] = key[k] - 1; }
      if (j < d) { n1[j] = key[j] - d; n2[j] = key[j] + d; }
      // lookup() may grow (reallocate) the table once; re-read the base each time
      float* a = lat.hashTable.lookup(n1, false);
      int32_t ia = a ? (int32_t)((a - lat.hashTable.getValues()) / vd) : -1;
      float* b = lat.hashTable.lookup(n2, false);
      int32_t ib = b ? (int32_t)((b - lat.hashTable.getValues()) / vd) : -1;
      nbr[((size_t)j * M + v) * 2 + 0] = ia;
      nbr[((size_t)j * M + v) * 2 + 1] = ib;
    }
  delete[] n1; delete[] n2;
  return M;
}

}  // extern "C"
