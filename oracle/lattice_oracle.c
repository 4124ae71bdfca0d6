/*
 * lattice_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C, single-threaded CPU restatement of the reference's permutohedral
 * lattice filter (reference: crf/lattice/lite/permutohedral.h, the code that
 * crf/lattice/lite/lattice.cpp:6-10 exports as `lattice.filter(src, ref)`).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
 * arm may load this file's shared object; the product path never does.
 *
 * Parity status: PINNED against the reference itself.  oracle/ref_driver.cpp
 * compiles the unmodified reference header in place from /root/reference into
 * oracle/_ref/ and tests/test_oracle_vs_ref.py checks this restatement against
 * it bit for bit (greedy points, ranks, barycentric weights, vertex ids, keys,
 * vertex count, filtered values).  The reference's own tests hold no
 * machine-checkable golden vectors for this path.
 *
 * The restatement is split into "build" (everything that depends only on the
 * feature vectors) and "apply" (splat / blur / slice of one value matrix) so
 * the integer artefacts can be inspected.  The reference interleaves the two
 * (permutohedral.h:236-238) but performs the same float operations on every
 * value in the same order, so results are bit-identical.
 *
 * Build with:  gcc -O2 -ffp-contract=off -fPIC -shared (see oracle/Makefile).
 * -ffp-contract=off keeps gcc from fusing a*b+c, matching the reference as
 * torch's extension builder compiles it (x86-64 baseline, no FMA).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct plo_lattice {
  int n;            /* points */
  int d;            /* feature dims; lattice lives in d+1 dims */
  int M;            /* occupied lattice vertices */
  float *scale;     /* [d]  permutohedral.h:354-371 */
  short *canonical; /* [(d+1)*(d+1)] permutohedral.h:346-351 */
  float *elevated;  /* [n*(d+1)] */
  short *greedy;    /* [n*(d+1)] after the hyperplane fix-up */
  signed char *rank;/* [n*(d+1)] after the hyperplane fix-up */
  float *bary;      /* [n*(d+1)] barycentric weights (slot d+1 dropped) */
  int *vid;         /* [n*(d+1)] vertex id, numbered by first touch */
  short *keys;      /* [cap_keys*d] */
  int cap_keys;
  int *nbr;         /* [(d+1)*M*2]: for axis j, vertex v: {id(key-e_j..), id(key+e_j..)}, -1 if absent */
  /* private open-addressing index over keys (layout is not part of parity) */
  int *slots;       /* entry table: vertex id or -1 */
  size_t nslots;    /* capacity */
  int faithful;     /* 1: reproduce the reference's stale-hash-after-grow quirk */
  int ngrow;        /* table doublings so far */
} plo_lattice;

/* ---- vertex index: key (d shorts) -> vertex id ---------------------------
 *
 * Restates HashTablePermutohedral (permutohedral.h:29-169): open addressing,
 * linear probing, capacity 2^15 doubling when filled >= capacity/2 - 1, keys
 * stored densely in creation order (so vertex id == creation order).
 *
 * REFERENCE QUIRK (reproduced when L->faithful != 0): lookup() computes
 * h = hash % capacity and only then calls lookupOffset(), which may grow()
 * the table before using h (permutohedral.h:102-103 then :62).  The one
 * lookup that triggers each doubling therefore probes from a slot computed
 * with the OLD capacity.  If bit log2(old capacity) of its hash is set it
 * misses an existing key (or parks a new key where later lookups will not
 * find it), so an extra "orphan" vertex with a duplicate key appears: at most
 * one per doubling.  hashTable.size() counts it, the pixel that triggered it
 * slices from it, and after the next doubling the rehash order decides which
 * copy later lookups see.  With L->faithful == 0 the slot is recomputed after
 * growing, i.e. the table behaves as its author intended: one vertex per
 * distinct key, numbered by first touch.  The product implements the latter;
 * tests/ quantify the difference.
 */

static size_t key_hash(const short *key, int d) {
  /* permutohedral.h:109-116; short is sign-extended into size_t */
  size_t k = 0;
  for (int i = 0; i < d; i++) {
    k += (size_t)key[i];
    k *= 2531011;
  }
  return k;
}

static void index_init(plo_lattice *L) {
  L->nslots = (size_t)1 << 15; /* permutohedral.h:36 */
  L->slots = (int *)malloc(sizeof(int) * L->nslots);
  for (size_t i = 0; i < L->nslots; i++) L->slots[i] = -1;
}

/* permutohedral.h:122-155: double, then re-place every entry in old-slot order */
static void index_grow(plo_lattice *L) {
  size_t old_n = L->nslots;
  int *old = L->slots;
  L->nslots = old_n * 2;
  L->slots = (int *)malloc(sizeof(int) * L->nslots);
  for (size_t i = 0; i < L->nslots; i++) L->slots[i] = -1;
  for (size_t i = 0; i < old_n; i++) {
    if (old[i] == -1) continue;
    size_t h = key_hash(L->keys + (size_t)old[i] * L->d, L->d) % L->nslots;
    while (L->slots[h] != -1) { h++; if (h == L->nslots) h = 0; }
    L->slots[h] = old[i];
  }
  free(old);
  L->ngrow++;
}

/* permutohedral.h:59-106 lookup(key, create): returns vertex id or -1. */
static int index_lookup(plo_lattice *L, const short *key, int create) {
  const int d = L->d;
  size_t h = key_hash(key, d) % L->nslots;
  if ((size_t)L->M >= (L->nslots / 2) - 1) {
    index_grow(L);
    if (!L->faithful) h = key_hash(key, d) % L->nslots;
  }
  for (;;) {
    int v = L->slots[h];
    if (v == -1) {
      if (!create) return -1;
      if (L->M == L->cap_keys) {
        L->cap_keys *= 2;
        L->keys = (short *)realloc(L->keys, sizeof(short) * (size_t)L->cap_keys * d);
      }
      memcpy(L->keys + (size_t)L->M * d, key, sizeof(short) * d);
      L->slots[h] = L->M;
      return L->M++;
    }
    if (memcmp(L->keys + (size_t)v * d, key, sizeof(short) * d) == 0) return v;
    h++;
    if (h == L->nslots) h = 0;
  }
}

/* ---- build ---------------------------------------------------------------- */

void plo_scale_factors(int d, float *out) {
  /* permutohedral.h:354-371.  Same expression, same types: 1.0f/sqrtf(float),
     then *= (int)(d+1) * sqrtf((float)(2.0/3)). */
  for (int i = 0; i < d; i++) {
    out[i] = 1.0f / (sqrtf((float)(i + 1) * (i + 2)));
    out[i] *= (d + 1) * sqrtf(2.0 / 3);
  }
}

void plo_free(plo_lattice *L) {
  if (!L) return;
  free(L->scale); free(L->canonical); free(L->elevated); free(L->greedy);
  free(L->rank); free(L->bary); free(L->vid); free(L->keys); free(L->nbr);
  free(L->slots); free(L);
}

plo_lattice *plo_build(const float *ref, int n, int d, int faithful) {
  plo_lattice *L = (plo_lattice *)calloc(1, sizeof(plo_lattice));
  L->faithful = faithful;
  const int d1 = d + 1;
  L->n = n; L->d = d; L->M = 0;
  L->scale = (float *)malloc(sizeof(float) * d);
  L->canonical = (short *)malloc(sizeof(short) * d1 * d1);
  L->elevated = (float *)malloc(sizeof(float) * (size_t)n * d1);
  L->greedy = (short *)malloc(sizeof(short) * (size_t)n * d1);
  L->rank = (signed char *)malloc((size_t)n * d1);
  L->bary = (float *)malloc(sizeof(float) * (size_t)n * d1);
  L->vid = (int *)malloc(sizeof(int) * (size_t)n * d1);
  L->cap_keys = 1 << 12;
  L->keys = (short *)malloc(sizeof(short) * (size_t)L->cap_keys * d);
  index_init(L);
  plo_scale_factors(d, L->scale);

  /* canonical simplex, permutohedral.h:346-351 */
  for (int i = 0; i <= d; i++) {
    for (int j = 0; j <= d - i; j++) L->canonical[i * d1 + j] = (short)i;
    for (int j = d - i + 1; j <= d; j++) L->canonical[i * d1 + j] = (short)(i - d1);
  }

  float *bary = (float *)malloc(sizeof(float) * (d + 2));
  short *key = (short *)malloc(sizeof(short) * d1);
  const float scale = 1.0f / (d + 1); /* permutohedral.h:387 */

  for (int p = 0; p < n; p++) {
    const float *pos = ref + (size_t)p * d;
    float *el = L->elevated + (size_t)p * d1;
    short *gr = L->greedy + (size_t)p * d1;
    signed char *rk = L->rank + (size_t)p * d1;
    const float *sf = L->scale;

    /* elevate, permutohedral.h:380-384 (operation order is part of parity) */
    el[d] = -d * pos[d - 1] * sf[d - 1];
    for (int i = d - 1; i > 0; i--)
      el[i] = (el[i + 1] - i * pos[i - 1] * sf[i - 1] + (i + 2) * pos[i] * sf[i]);
    el[0] = el[1] + 2 * pos[0] * sf[0];

    /* nearest remainder-0 point, permutohedral.h:392-403 */
    int sum = 0;
    for (int i = 0; i <= d; i++) {
      float v = el[i] * scale;
      float up = ceilf(v) * (d + 1);
      float down = floorf(v) * (d + 1);
      if (up - el[i] < el[i] - down) gr[i] = (short)up;
      else gr[i] = (short)down;
      sum += gr[i];
    }
    sum *= scale; /* int <- float truncation quirk, permutohedral.h:403 */

    /* rank, permutohedral.h:407-411 */
    memset(rk, 0, (size_t)d1);
    for (int i = 0; i < d; i++)
      for (int j = i + 1; j <= d; j++)
        if (el[i] - gr[i] < el[j] - gr[j]) rk[i]++;
        else rk[j]++;

    /* hyperplane fix-up, permutohedral.h:413-433 */
    if (sum > 0) {
      for (int i = 0; i <= d; i++) {
        if (rk[i] >= d + 1 - sum) { gr[i] -= d + 1; rk[i] += sum - (d + 1); }
        else rk[i] += sum;
      }
    } else if (sum < 0) {
      for (int i = 0; i <= d; i++) {
        if (rk[i] < -sum) { gr[i] += d + 1; rk[i] += (d + 1) + sum; }
        else rk[i] += sum;
      }
    }

    /* barycentric, permutohedral.h:436-441 */
    memset(bary, 0, sizeof(float) * (d + 2));
    for (int i = 0; i <= d; i++) {
      bary[d - rk[i]] += (el[i] - gr[i]) * scale;
      bary[d + 1 - rk[i]] -= (el[i] - gr[i]) * scale;
    }
    bary[0] += 1.0f + bary[d + 1];

    /* simplex vertices in remainder order, permutohedral.h:444-450,458-460 */
    for (int r = 0; r <= d; r++) {
      for (int i = 0; i < d; i++) key[i] = gr[i] + L->canonical[r * d1 + rk[i]];
      L->vid[(size_t)p * d1 + r] = index_lookup(L, key, 1);
      L->bary[(size_t)p * d1 + r] = bary[r];
    }
  }
  free(bary);

  /* blur neighbours, permutohedral.h:503-522: the reference looks them up per
     axis inside blur(); no vertex is created there, so they are a pure
     function of the final key set and can be tabulated once. */
  const int M = L->M;
  L->nbr = (int *)malloc(sizeof(int) * (size_t)d1 * (M > 0 ? M : 1) * 2);
  short *n1 = (short *)malloc(sizeof(short) * d1);
  short *n2 = (short *)malloc(sizeof(short) * d1);
  for (int j = 0; j <= d; j++) {
    for (int v = 0; v < M; v++) {
      const short *k = L->keys + (size_t)v * d;
      for (int i = 0; i < d; i++) { n1[i] = k[i] + 1; n2[i] = k[i] - 1; }
      /* for j == d the reference writes slot d of a (d+1)-short scratch array
         that the d-coordinate hash never reads (permutohedral.h:508-509) */
      if (j < d) { n1[j] = k[j] - d; n2[j] = k[j] + d; }
      L->nbr[((size_t)j * M + v) * 2 + 0] = index_lookup(L, n1, 0);
      L->nbr[((size_t)j * M + v) * 2 + 1] = index_lookup(L, n2, 0);
    }
  }
  free(n1); free(n2); free(key);
  return L;
}

/* ---- apply: splat, blur, slice ------------------------------------------ */

void plo_apply(const plo_lattice *L, const float *src, int C, float *out) {
  const int n = L->n, d = L->d, d1 = d + 1, M = L->M;
  size_t MC = (size_t)(M > 0 ? M : 1) * C;
  float *val = (float *)calloc(MC, sizeof(float));
  float *tmp = (float *)calloc(MC, sizeof(float));
  float *zero = (float *)calloc(C, sizeof(float));

  /* splat, permutohedral.h:452-455, scanline order */
  for (int p = 0; p < n; p++)
    for (int r = 0; r <= d; r++) {
      float *v = val + (size_t)L->vid[(size_t)p * d1 + r] * C;
      float w = L->bary[(size_t)p * d1 + r];
      const float *s = src + (size_t)p * C;
      for (int c = 0; c < C; c++) v[c] += (w * s[c]);
    }

  /* blur, permutohedral.h:498-534: one Jacobi pass per axis 0..d */
  float *oldv = val, *newv = tmp;
  for (int j = 0; j <= d; j++) {
    for (int v = 0; v < M; v++) {
      int a = L->nbr[((size_t)j * M + v) * 2 + 0];
      int b = L->nbr[((size_t)j * M + v) * 2 + 1];
      const float *vm1 = a >= 0 ? oldv + (size_t)a * C : zero;
      const float *vp1 = b >= 0 ? oldv + (size_t)b * C : zero;
      const float *o = oldv + (size_t)v * C;
      float *q = newv + (size_t)v * C;
      for (int c = 0; c < C; c++) q[c] = 2 * (0.25f * vm1[c] + 0.5f * o[c] + 0.25f * vp1[c]);
    }
    float *t = newv; newv = oldv; oldv = t;
  }

  /* slice, permutohedral.h:473-483: per-term divide by 1 + 2^-d, no
     homogeneous normalisation (permutohedral.h:272-275 commented out) */
  const float denom = 1 + powf(2, -d);
  for (int p = 0; p < n; p++) {
    float *col = out + (size_t)p * C;
    for (int c = 0; c < C; c++) col[c] = 0;
    for (int r = 0; r <= d; r++) {
      const float *b = oldv + (size_t)L->vid[(size_t)p * d1 + r] * C;
      float w = L->bary[(size_t)p * d1 + r];
      for (int c = 0; c < C; c++) col[c] += w * b[c] / denom;
    }
  }
  free(val); free(tmp); free(zero);
}

/* lattice.cpp:6-10 filter(src[n,C], ref[n,d]) -> out[n,C] */
void plo_filter(const float *src, const float *ref, int n, int C, int d, int faithful, float *out) {
  plo_lattice *L = plo_build(ref, n, d, faithful);
  plo_apply(L, src, C, out);
  plo_free(L);
}

/* ---- accessors for the Python side (ctypes) ------------------------------ */
int plo_num_vertices(const plo_lattice *L) { return L->M; }
int plo_num_grows(const plo_lattice *L) { return L->ngrow; }
const float *plo_elevated(const plo_lattice *L) { return L->elevated; }
const short *plo_greedy(const plo_lattice *L) { return L->greedy; }
const signed char *plo_rank(const plo_lattice *L) { return L->rank; }
const float *plo_bary(const plo_lattice *L) { return L->bary; }
const int *plo_vid(const plo_lattice *L) { return L->vid; }
const short *plo_keys(const plo_lattice *L) { return L->keys; }
const int *plo_nbr(const plo_lattice *L) { return L->nbr; }
