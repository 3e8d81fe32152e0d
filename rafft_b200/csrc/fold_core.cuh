/*
 * fold_core.cuh -- the RAFFT folding hot path: one thread block (1..8 warps) per sequence.
 *
 * What the reference does per sequence (rafft/rafft.py:156-216 bfs_pairs, :112-153
 * create_childs, :86-109 find_best_consecutives, :36-83 window_slide; rafft/utils.py:115-132
 * auto_cor, :135-138 eval_one_struct) is re-organised around one observation: every
 * unpaired region the reference calls a Node is exactly the set of unpaired positions of
 * ONE loop of the current structure, and the loop (closing pair + branch helices) fully
 * determines the node's lag ranking, its stems, and the energy change of forming each
 * stem (loop energies are additive).  So the block keeps a table of UNIQUE loops ("nodes")
 * with their sorted candidate stems, and a structure is just (energy, fingerprint, list of
 * node ids).  The `-ms` beam, the itertools.product enumeration, the global `seen` set,
 * the max_branch cut and the stable energy sort then run on integers only and reproduce
 * the reference's order exactly.
 *
 * Two scopes of parallelism:
 *   warp tasks   one node expansion per warp (RF_W* phases, __syncwarp only): bit-parallel
 *                one-hot correlation, bitonic lag ranking, per-lag stem scan over set bits,
 *                loop-energy change of every stem, stable candidate sort;
 *   block phases the beam logic (RF_PAR / RF_SINGLE phases bracketed by block barriers).
 * tests/emu compiles this same file for the host with one thread per scope to check the
 * logic without a GPU; that build is test-only.
 */
#pragma once
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define RF_FN __device__ __forceinline__
#define RF_NOINLINE __device__ __noinline__
#define RF_HD __host__ __device__ inline
#else
#define RF_FN inline
#define RF_NOINLINE inline
#define RF_HD inline
#include <math.h>
#endif

namespace rafft {

/* ------------------------------------------------------------------------------------ */
/* Energy tables on the device (int16 where the values allow; built by the host).        */
struct DevEnergy {
  int32_t hairpin[31], bulge[31], interior[31];
  int16_t stack[36];
  int16_t mmH[96], mmI[96], mm1n[96], mm23[96], mmM[96], mmE[96];
  int16_t d5[24], d3[24];
  int16_t int11[576], int21[2304], int22[9216];
  int32_t ninio, max_ninio, ml_base, ml_intern, ml_closing, terminal_au;
  int32_t n_tri, n_tetra, n_hexa;
  uint32_t tri_key[64], tetra_key[64], hexa_key[64]; /* 2 bits per base, first base highest */
  int32_t tri_e[64], tetra_e[64], hexa_e[64];
  int32_t logterm[4096]; /* (int)(lxc*log(u/30)) for u > 30, host libm (bit-equal to oracle) */
  /* membership bitmaps of the special-hairpin keys (bit = key): most hairpins are not special */
  uint32_t tri_bm[32], tetra_bm[128], hexa_bm[2048];
};

/* Launch-wide constants. */
struct FoldParams {
  int32_t nb_mode, max_stack, max_branch, min_hp, traj;
  double min_nrj;
  double w[4][4];   /* pair weights as the reference's float64 (prep_sequence CENCODING) */
  int32_t wi[4][4]; /* same * 16, exact integers (host checks) */
  int32_t cand_thr; /* keep stems with dE < cand_thr (dcal); exact test per structure later */
  int32_t seen_gen; /* 1..65535: generation tag of this fold's entries of the `seen` table */
};

/* Capacities of one per-block arena (host sizes them; overflow -> status, host regrows). */
struct Caps {
  int32_t L_max;
  int32_t node_cap, pool16_cap, cand_cap, store_cap, nlist_cap;
  int32_t seen_cap, nht_cap; /* powers of two */
  int32_t new_cap;           /* >= max_branch + max_stack + 1 */
  int32_t step_cap;          /* folding steps */
  int32_t todo_cap;
  int32_t chunk_cap;         /* stem combinations examined per block phase */
  int32_t key_cap;           /* per-warp sort keys: pow2 >= max(2*L_max, nb_mode) */
  int32_t mask_words;        /* 32-bit words per one-hot mask (L_max/32 + 2) */
  int32_t wscratch_bytes;    /* scratch slot of one region expansion (TaskQueue::scratch) */
  int64_t arena_bytes;
};

struct NodeRec {
  int32_t ci, cj;         /* closing pair; exterior loop = (-1, L) */
  int32_t pool_off, br_n; /* pool16: br_n*(p,q) then br_n+2 cumulative segment offsets */
  int32_t lv;             /* unpaired positions */
  int32_t e_loop;         /* energy of the loop as it stands */
  int32_t cand_off, cand_n; /* cand_n: -1 not expanded, -2 queued */
  int32_t n_ranked, n_pos;  /* lags scanned / lags with a stem (reference work counters) */
  int64_t w_scan;           /* sum of ceil(d/2) over the ranked lags */
  uint64_t hash;
  int32_t nht_slot, pad_;   /* slot in the unique-loop table (for the clean-up) */
};

/* One stem of a region: exactly one 32-byte sector, 32-byte aligned, so that the product
 * enumeration (one random record per region and combination) costs one sector per look-up.
 * The child loop ids, which only the selected children need, live in Arena::ckid. */
struct Cand {
  int32_t dE;
  uint16_t mip, mjp, nb;      /* innermost pair (local indices of the region) and pairs of the stem */
  uint16_t hh;                /* branches of the region's loop before mip | before mjp << 8 */
  uint32_t geo;               /* original position of the innermost pair: i | j << 12, bit 24 set when
                                 the stem is contiguous in original coordinates (pair k = (i - k, j + k))
                                 and both branch counts fit hh; 0: geometry through the loop record */
  uint64_t fp_lo, fp_hi;      /* sum of pair fingerprints of the stem */
};
#define RF_GEO_OK (1u << 24)
/* geometry + fingerprint of a stem; pos[t]: original position of local index t, brcnt[t]: branches
 * before it */
template <class PosT, class BrT>
RF_FN void cand_geometry(Cand &c, const PosT *pos, const BrT *brcnt, int mip, int mjp, int nb);
static_assert(sizeof(Cand) == 32, "Cand is one sector");
struct CandKid {
  int32_t inner_id, outer_id; /* -1 unresolved, -2 no such node, -3 being created */
};

struct StoreRec {  /* a structure that has been in a beam */
  int32_t e;
  int32_t nl_off, nl_n; /* node ids in the reference's node_list order */
  int32_t done;         /* its whole product has been enumerated (all children are seen) */
  uint64_t fp_lo, fp_hi;
  uint64_t prod;        /* size of its product (saturated), 0 = no candidate anywhere */
  int64_t nact;         /* nodes with candidates */
};

struct NewRec {
  int32_t e, parent_slot;
  uint64_t combo, fp_lo, fp_hi;
};

/* Global output, bump-allocated by finished sequences. */
struct OutBuf {
  int32_t *status;      /* per sequence */
  int64_t *seq_struct;  /* per sequence: first structure index */
  int32_t *seq_nstruct;
  int64_t *seq_db;      /* per sequence: first db byte (rows padded to 16 bytes, 16-byte aligned) */
  int64_t *seq_step;    /* per sequence: first step index */
  int32_t *seq_nstep;
  int64_t *seq_entry;   /* per sequence: first entry index */
  int64_t *counters;    /* per sequence * RAFFT_CNT_N */
  int32_t *energy;      /* structures */
  char *db;
  int32_t *step_n;      /* entries per step */
  int32_t *entries;     /* local structure indices */
  unsigned long long *cursors; /* [0] struct, [1] db, [2] step, [3] entry, [4] queue head */
  int64_t struct_cap, db_cap, step_cap, entry_cap;
};

enum { CNT_STEPS = 0, CNT_NODE_EXP, CNT_W_CORR, CNT_W_SCAN, CNT_N_EVAL, CNT_UNIQ_NODES,
       CNT_UNIQ_W_CORR, CNT_UNIQ_W_SCAN, CNT_COMBOS, CNT_NODE_POS, CNT_T0_NS, CNT_T1_NS, CNT_N_EVAL_EXPAND,
       CNT_TRIVIAL_NODES, CNT_N };

enum { ST_OK = 0, ST_ERR_SEQ = 2, ST_ERR_CAPACITY = 4 };

/* ------------------------------------------------------------------------------------ */
/* Execution policy: CUDA thread block of 1..8 warps, or a single host thread             */
/* (tests/emu only).  Block scope: tid/nthreads/sync.  Warp scope: lane/warp/wsync.       */
#ifdef __CUDACC__
struct Exec {
  RF_FN static int tid() { return threadIdx.x; }
  RF_FN static int nthreads() { return blockDim.x; }
  RF_FN static int stride() { return blockDim.x; } /* of hand-written block-strided loops */
  RF_FN static int lane() { return threadIdx.x & 31; }
  RF_FN static int lanes() { return 32; }
  RF_FN static int warp() { return threadIdx.x >> 5; }
  RF_FN static int nwarps() { return blockDim.x >> 5; }
  RF_FN static void sync() { __syncthreads(); }
  RF_FN static void fence_block() { __threadfence_block(); }
  RF_FN static void wsync() { __syncwarp(); }
  RF_FN static int atomic_add(int *p, int v) { return atomicAdd(p, v); }
  RF_FN static unsigned atomic_or(unsigned *p, unsigned v) { return atomicOr(p, v); }
  RF_FN static int atomic_min(int *p, int v) { return atomicMin(p, v); }
  RF_FN static int atomic_max(int *p, int v) { return atomicMax(p, v); }
  RF_FN static unsigned long long atomic_add64(unsigned long long *p, unsigned long long v) {
    return atomicAdd(p, v);
  }
  RF_FN static int atomic_cas(int *p, int cmp, int v) { return atomicCAS(p, cmp, v); }
  RF_FN static unsigned long long atomic_cas64(unsigned long long *p, unsigned long long cmp,
                                                unsigned long long v) {
    return atomicCAS(p, cmp, v);
  }
  /* sum over the lanes of the calling warp (all 32 lanes must call) */
  RF_FN static long long warp_sum(long long v) {
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
  }
  RF_FN static double dadd(double a, double b) { return __dadd_rn(a, b); }
  RF_FN static double dmul(double a, double b) { return __dmul_rn(a, b); }
  RF_FN static double ddiv(double a, double b) { return __ddiv_rn(a, b); }
  RF_FN static unsigned long long dbits(double a) { return (unsigned long long)__double_as_longlong(a); }
  RF_FN static long long now_ns() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return (long long)t; }
  RF_FN static int popc(unsigned x) { return __popc(x); }
  RF_FN static int ffs(unsigned x) { return __ffs((int)x); }
  /* bits [b, b+32) of the 64-bit value hi:lo, 0 <= b < 32 */
  RF_FN static unsigned funnel_r(unsigned lo, unsigned hi, int b) { return __funnelshift_r(lo, hi, b); }
  /* Sum of v over the block, the same value in every thread, with ONE barrier: per-warp partial sums
   * go to one of two alternating sets of slots (`wsum`: 16 ints; `turn` counts the calls of a run
   * of consecutive calls and must be the same in every thread). */
  RF_FN static int block_sum(int v, int *wsum, int &turn) {
    v = __reduce_add_sync(0xffffffffu, v);
    int *w = wsum + ((turn++ & 1) << 3);
    if ((threadIdx.x & 31) == 0) w[threadIdx.x >> 5] = v;
    __syncthreads();
    int t = 0;
    for (int k = 0; k < (int)(blockDim.x >> 5); k++) t += w[k];
    return t;
  }
  /* Block-wide exclusive prefix sum of 0/1 flags (n <= capacity of flag/pref); returns the
   * total.  Must be called by every thread of the block; `wsum` = nwarps+1 ints of shared
   * memory.  Ends with a barrier, so pref[] and the total are visible to all threads. */
  __device__ __noinline__ static int scan_flags(const int *flag, int *pref, int n, int *wsum) {
    const int T = blockDim.x, t = threadIdx.x;
    const int per = (n + T - 1) / T;
    const int lo = min(n, t * per), hi = min(n, lo + per);
    __syncthreads();
    int c = 0;
    for (int i = lo; i < hi; i++) c += flag[i];
    int inc = c;
    for (int d = 1; d < 32; d <<= 1) {
      int v = __shfl_up_sync(0xffffffffu, inc, d);
      if ((t & 31) >= d) inc += v;
    }
    if ((t & 31) == 31) wsum[t >> 5] = inc;
    __syncthreads();
    int base = 0, total = 0;
    for (int w = 0; w < (T >> 5); w++) {
      int v = wsum[w];
      if (w < (t >> 5)) base += v;
      total += v;
    }
    int run = base + inc - c;
    for (int i = lo; i < hi; i++) { pref[i] = run; run += flag[i]; }
    __syncthreads();
    return total;
  }
};
#define RF_PAR_BEGIN(i, n) \
  Exec::sync();            \
  for (int i = Exec::tid(); i < (n); i += Exec::nthreads()) {
#define RF_PAR_END \
  }                \
  Exec::sync();
#define RF_SINGLE_BEGIN \
  Exec::sync();         \
  if (Exec::tid() == 0) {
#define RF_SINGLE_END \
  }                   \
  Exec::sync();
#define RF_WPAR_BEGIN(i, n) \
  Exec::wsync();            \
  for (int i = Exec::lane(); i < (n); i += 32) {
#define RF_WPAR_END \
  }                 \
  Exec::wsync();
#define RF_WSINGLE_BEGIN \
  Exec::wsync();         \
  if (Exec::lane() == 0) {
#define RF_WSINGLE_END \
  }                    \
  Exec::wsync();
#else
struct Exec {
  static int &order() { static int o = 0; return o; } /* 0 forward, 1 reverse iteration */
  static int tid() { return 0; }
  static int &emu_threads() { static int t = 32; return t; } /* chunk width of the emulation */
  static int nthreads() { return emu_threads(); }
  static int stride() { return 1; }
  static int lane() { return 0; }
  static int lanes() { return 1; }
  static int warp() { return 0; }
  static int nwarps() { return 1; }
  static void sync() {}
  static void fence_block() {}
  static void wsync() {}
  static int atomic_add(int *p, int v) { int o = *p; *p += v; return o; }
  static unsigned atomic_or(unsigned *p, unsigned v) { unsigned o = *p; *p |= v; return o; }
  static int atomic_min(int *p, int v) { int o = *p; if (v < o) *p = v; return o; }
  static int atomic_max(int *p, int v) { int o = *p; if (v > o) *p = v; return o; }
  static unsigned long long atomic_add64(unsigned long long *p, unsigned long long v) {
    unsigned long long o = *p; *p += v; return o;
  }
  static int atomic_cas(int *p, int cmp, int v) { int o = *p; if (o == cmp) *p = v; return o; }
  static unsigned long long atomic_cas64(unsigned long long *p, unsigned long long cmp,
                                         unsigned long long v) {
    unsigned long long o = *p; if (o == cmp) *p = v; return o;
  }
  static long long warp_sum(long long v) { return v; }
  static double dadd(double a, double b) { volatile double r = a + b; return r; }
  static double dmul(double a, double b) { volatile double r = a * b; return r; }
  static double ddiv(double a, double b) { volatile double r = a / b; return r; }
  static unsigned long long dbits(double a) { unsigned long long u; memcpy(&u, &a, 8); return u; }
  static long long now_ns() { return 0; }
  static int popc(unsigned x) { return __builtin_popcount(x); }
  static int ffs(unsigned x) { return __builtin_ffs((int)x); }
  static unsigned funnel_r(unsigned lo, unsigned hi, int b) {
    return b ? (lo >> b) | (hi << (32 - b)) : lo;
  }
  static int block_sum(int v, int *, int &turn) { turn++; return v; }
  static int scan_flags(const int *flag, int *pref, int n, int *) {
    int run = 0;
    for (int i = 0; i < n; i++) { pref[i] = run; run += flag[i]; }
    return run;
  }
};
/* Iterations of a phase must be independent: the emulation runs them forward or reversed. */
#define RF_PAR_BEGIN(i, n)                                   \
  for (int i##_k = 0, i##_n = (n); i##_k < i##_n; i##_k++) { \
    int i = Exec::order() ? i##_n - 1 - i##_k : i##_k;
#define RF_PAR_END }
#define RF_SINGLE_BEGIN {
#define RF_SINGLE_END }
#define RF_WPAR_BEGIN(i, n) RF_PAR_BEGIN(i, n)
#define RF_WPAR_END }
#define RF_WSINGLE_BEGIN {
#define RF_WSINGLE_END }
#endif

/* ------------------------------------------------------------------------------------ */
/* Small helpers                                                                         */
RF_FN uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
/* 128-bit fingerprint of one base pair; a structure's fingerprint is the wrapping sum. */
RF_FN void pair_fp(int i, int j, uint64_t &lo, uint64_t &hi) {
  uint64_t k = ((uint64_t)(uint32_t)i << 32) | (uint32_t)j;
  lo = mix64(k + 0x9E3779B97F4A7C15ULL);
  hi = mix64(k ^ 0xD1B54A32D192ED03ULL) * 0x2545F4914F6CDD1DULL + 0x632BE59BD9B4E019ULL;
}
/* pair type 0..5 = CG GC GU UG AU UA, -1 otherwise (bases 0..3 = A C G U) */
RF_FN int pair_type(int a, int b) {
  const uint64_t T = (5ULL << 12) | (1ULL << 24) | (2ULL << 36) | (3ULL << 44) | (6ULL << 48) |
                     (4ULL << 56);
  return (int)((T >> (4 * (a * 4 + b))) & 15) - 1;
}

template <class PosT, class BrT>
RF_FN void cand_geometry(Cand &c, const PosT *pos, const BrT *brcnt, int mip, int mjp, int nb) {
  const int oi = (int)pos[mip], oj = (int)pos[mjp];
  const int h1 = (int)brcnt[mip], h2 = (int)brcnt[mjp];
  bool plain = h1 < 256 && h2 < 256;
  uint64_t flo = 0, fhi = 0;
  for (int k = 0; k < nb; k++) {
    const int a = (int)pos[mip - k], b = (int)pos[mjp + k];
    plain = plain && a == oi - k && b == oj + k;
    uint64_t l1, h1_;
    pair_fp(a, b, l1, h1_);
    flo += l1; fhi += h1_;
  }
  c.mip = (uint16_t)mip; c.mjp = (uint16_t)mjp; c.nb = (uint16_t)nb;
  c.hh = (uint16_t)((h1 & 255) | ((h2 & 255) << 8));
  c.geo = plain ? ((uint32_t)oi | ((uint32_t)oj << 12) | RF_GEO_OK) : 0u;
  c.fp_lo = flo; c.fp_hi = fhi;
}

/* ------------------------------------------------------------------------------------ */
/* Loop energies (ViennaRNA model restated in SURVEY.md Appendix B; integer dcal/mol).   */
struct EnergyView {
  const DevEnergy *P;
  const uint8_t *s; /* base codes of the sequence */
  int L;
};

RF_FN int e_loglen(const EnergyView &ev, const int32_t *arr, int u) {
  return u <= 30 ? arr[u] : arr[30] + ev.P->logterm[u < 4096 ? u : 4095];
}

RF_NOINLINE int e_hairpin(const EnergyView &ev, int i, int j) {
  const DevEnergy *P = ev.P;
  int u = j - i - 1;
  int e = e_loglen(ev, P->hairpin, u);
  if (u < 3) return e;
  int t = pair_type(ev.s[i], ev.s[j]);
  if (u == 3 || u == 4 || u == 6) {
    /* 2 bits per base, first base highest: independent loads instead of a loop of dependent ones */
    const uint8_t *q = ev.s + i;
    uint32_t key = ((uint32_t)q[0] << 8) | ((uint32_t)q[1] << 6) | ((uint32_t)q[2] << 4) | ((uint32_t)q[3] << 2) | (uint32_t)q[4];
    if (u >= 4) key = (key << 2) | (uint32_t)q[5];
    if (u == 6) key = (key << 4) | ((uint32_t)q[6] << 2) | (uint32_t)q[7];
    if (u == 4) {
      if ((P->tetra_bm[key >> 5] >> (key & 31)) & 1)
        for (int k = 0; k < P->n_tetra; k++)
          if (P->tetra_key[k] == key) return P->tetra_e[k];
    } else if (u == 6) {
      if ((P->hexa_bm[key >> 5] >> (key & 31)) & 1)
        for (int k = 0; k < P->n_hexa; k++)
          if (P->hexa_key[k] == key) return P->hexa_e[k];
    } else {
      if ((P->tri_bm[key >> 5] >> (key & 31)) & 1)
        for (int k = 0; k < P->n_tri; k++)
          if (P->tri_key[k] == key) return P->tri_e[k];
      return e + (t > 1 ? P->terminal_au : 0);
    }
  }
  return e + P->mmH[t * 16 + ev.s[i + 1] * 4 + ev.s[j - 1]];
}

RF_NOINLINE int e_interior(const EnergyView &ev, int i, int j, int p, int q) {
  const DevEnergy *P = ev.P;
  const uint8_t *s = ev.s;
  int n1 = p - i - 1, n2 = j - q - 1;
  int t = pair_type(s[i], s[j]), t2 = pair_type(s[q], s[p]);
  int nl = n1 > n2 ? n1 : n2, ns = n1 > n2 ? n2 : n1;
  if (nl == 0) return P->stack[t * 6 + t2];
  if (ns == 0) {
    int e = e_loglen(ev, P->bulge, nl);
    if (nl == 1) return e + P->stack[t * 6 + t2];
    if (t > 1) e += P->terminal_au;
    if (t2 > 1) e += P->terminal_au;
    return e;
  }
  int si1 = s[i + 1], sj1 = s[j - 1], sp1 = s[p - 1], sq1 = s[q + 1];
  if (ns == 1) {
    if (nl == 1) return P->int11[((t * 6 + t2) * 4 + si1) * 4 + sj1];
    if (nl == 2) {
      if (n1 == 1) return P->int21[(((t * 6 + t2) * 4 + si1) * 4 + sq1) * 4 + sj1];
      return P->int21[(((t2 * 6 + t) * 4 + sq1) * 4 + si1) * 4 + sp1];
    }
    int e = e_loglen(ev, P->interior, nl + 1);
    int a = (nl - ns) * P->ninio;
    e += a < P->max_ninio ? a : P->max_ninio;
    return e + P->mm1n[t * 16 + si1 * 4 + sj1] + P->mm1n[t2 * 16 + sq1 * 4 + sp1];
  }
  if (ns == 2) {
    if (nl == 2) return P->int22[((((t * 6 + t2) * 4 + si1) * 4 + sp1) * 4 + sq1) * 4 + sj1];
    if (nl == 3)
      return P->interior[5] + P->ninio + P->mm23[t * 16 + si1 * 4 + sj1] +
             P->mm23[t2 * 16 + sq1 * 4 + sp1];
  }
  int e = e_loglen(ev, P->interior, n1 + n2);
  int a = (nl - ns) * P->ninio;
  e += a < P->max_ninio ? a : P->max_ninio;
  return e + P->mmI[t * 16 + si1 * 4 + sj1] + P->mmI[t2 * 16 + sq1 * 4 + sp1];
}

/* stem (p,q) seen from a multiloop: 5' neighbour p-1, 3' neighbour q+1 always exist */
RF_FN int e_mlstem(const EnergyView &ev, int p, int q) {
  int t = pair_type(ev.s[p], ev.s[q]);
  int e = ev.P->ml_intern + ev.P->mmM[t * 16 + ev.s[p - 1] * 4 + ev.s[q + 1]];
  return e + (t > 1 ? ev.P->terminal_au : 0);
}
RF_FN int e_extstem(const EnergyView &ev, int p, int q) {
  const DevEnergy *P = ev.P;
  int t = pair_type(ev.s[p], ev.s[q]);
  int e = t > 1 ? P->terminal_au : 0;
  if (p > 0 && q < ev.L - 1) e += P->mmE[t * 16 + ev.s[p - 1] * 4 + ev.s[q + 1]];
  else if (p > 0) e += P->d5[t * 4 + ev.s[p - 1]];
  else if (q < ev.L - 1) e += P->d3[t * 4 + ev.s[q + 1]];
  return e;
}

/* Energy of the loop closed by (ci,cj) (ci < 0: exterior) whose branches are, in order,
 * br[lo1..hi1), then the extra pair (xp,xq) if xp >= 0, then br[lo2..hi2); `unp` unpaired. */
RF_NOINLINE int loop_energy(const EnergyView &ev, int ci, int cj, const int16_t *br, int lo1, int hi1,
                      int xp, int xq, int lo2, int hi2, int unp) {
  int nb = (hi1 - lo1) + (xp >= 0 ? 1 : 0) + (hi2 - lo2);
  if (ci < 0) {
    int e = 0;
    for (int k = lo1; k < hi1; k++) e += e_extstem(ev, br[2 * k], br[2 * k + 1]);
    if (xp >= 0) e += e_extstem(ev, xp, xq);
    for (int k = lo2; k < hi2; k++) e += e_extstem(ev, br[2 * k], br[2 * k + 1]);
    return e;
  }
  if (nb == 0) return e_hairpin(ev, ci, cj);
  if (nb == 1) {
    int p, q;
    if (xp >= 0) { p = xp; q = xq; }
    else if (hi1 > lo1) { p = br[2 * lo1]; q = br[2 * lo1 + 1]; }
    else { p = br[2 * lo2]; q = br[2 * lo2 + 1]; }
    return e_interior(ev, ci, cj, p, q);
  }
  const DevEnergy *P = ev.P;
  int e = P->ml_closing + P->ml_base * unp;
  {
    int tt = pair_type(ev.s[cj], ev.s[ci]);
    e += P->ml_intern + P->mmM[tt * 16 + ev.s[cj - 1] * 4 + ev.s[ci + 1]] +
         (tt > 1 ? P->terminal_au : 0);
  }
  for (int k = lo1; k < hi1; k++) e += e_mlstem(ev, br[2 * k], br[2 * k + 1]);
  if (xp >= 0) e += e_mlstem(ev, xp, xq);
  for (int k = lo2; k < hi2; k++) e += e_mlstem(ev, br[2 * k], br[2 * k + 1]);
  return e;
}
/* ------------------------------------------------------------------------------------ */
/* Per-block working memory, carved from one arena (global memory, L1/L2 resident).       */
/* `seen` entry: 64 + 48 bits of the structure fingerprint and, in the top 16 bits of `hi`, the
 * generation of the fold that wrote it -- entries of earlier folds of the arena read as empty,
 * so the table needs no clean-up pass (the host zeroes it when the generation counter wraps). */
struct alignas(16) SeenEnt { uint64_t lo, hi; };

struct alignas(8) ActRec { int32_t c, off; }; /* candidates passing for the owning structure, first stem record */

struct Arena {
  uint8_t *code;       /* L_max */
  NodeRec *node;       /* node_cap */
  int16_t *pool16;     /* pool16_cap: branches + segment offsets of nodes */
  Cand *cand;          /* cand_cap */
  CandKid *ckid;       /* cand_cap: child loops of the stems */
  StoreRec *store;     /* store_cap */
  char *store_db;      /* store_cap * L */
  int32_t *nlist;      /* nlist_cap: node lists of stored structures */
  int32_t *ncnt;       /* nlist_cap: candidates of that node passing for the owning structure */
  int32_t *ncoff;      /* nlist_cap: first candidate record of that node (copy of NodeRec::cand_off) */
  ActRec *act;         /* nlist_cap: per structure (at its nl_off) the regions that HAVE candidates, compacted in
                          node-list order: what the product enumeration reads (one 32-byte sector = 4 regions) */
  SeenEnt *seen;       /* seen_cap; lo == 0 marks an empty slot (kept clean between runs) */
  unsigned long long *nht;     /* nht_cap: (hash32 << 32) | (node id + 1), 0 = empty */
  NewRec *fresh;       /* new_cap */
  int32_t *sel;        /* max_stack: selected entries of the next beam */
  int32_t *sel_idx;    /* max_stack: the same before ranking */
  int32_t *sel_store;  /* max_stack: selected rank -> store index */
  int32_t *beam;       /* max_stack: store indices */
  int32_t *traj_n;     /* step_cap */
  int32_t *traj;       /* step_cap * max_stack: store indices */
  int32_t *todo;       /* todo_cap * 3 */
  int32_t *todo_id;    /* todo_cap */
  int32_t *flag, *pref; /* sel_cap = max(chunk_cap, new_cap + max_stack) */
  int32_t *sel_e;       /* sel_cap: energies of the entries under selection (the select kernel
                           points this at shared memory when they fit) */
};

/* Scratch of one node expansion (views into the warp's slice of Arena::wscratch). */
struct WarpScratch {
  uint64_t *keys;   /* key_cap: lag keys, later candidate keys */
  uint16_t *pos;    /* L_max: local index -> original position */
  uint16_t *brcnt;  /* L_max: branches of the loop before that position */
  uint32_t *mask;   /* 14 * mask_words: F[A,C,G,U], R[A,C,G,U] (reversed), CL, CLR, bit planes FH, FL, RH, RL */
  int32_t *ranked;  /* nb_mode: lag of rank r */
  int32_t *r_dE;    /* nb_mode */
  uint32_t *r_stem; /* nb_mode: (max_i << 16) | max_j */
  int32_t *r_nb;    /* nb_mode */
  int32_t *tmp;     /* 8 ints */
  uint64_t *sel;    /* RF_SEL_CAP + RF_SEL_SAMPLE keys when key_cap > RF_SEL_MIN: top-n selection (expand_e2) */
};

/* Regions with more than RF_SEL_MIN (padded) lags rank by selection instead of a full sort. */
#define RF_SEL_MIN 2048
#define RF_SEL_CAP 2048
#define RF_SEL_SAMPLE 256

/* Per-sequence state that lives across the phase kernels (global memory; the block-scope
 * kernels work on a shared-memory copy). */
struct SeqState {
  int32_t seq_index, L, status;
  int32_t active;          /* 0 once the fold has converged or failed */
  int32_t stage;           /* 0: a new folding step starts (record the beam), 1: expansion under way */
  int32_t expand_pending;  /* some region of the beam could not be queued this round */
  int32_t n_node, n_pool16, n_cand, n_store, n_nlist, n_seen, n_todo, n_expand;
  int32_t beam_n, n_new, nb_branch, n_steps, n_traj;
  int32_t tmp0, tmp1, tmp2, cur_b, xcut;
  int32_t fail_line;       /* source line of the first capacity failure (diagnostics) */
  int32_t exp_from;        /* loop records from this id on may still need their expansion */
  int32_t qcnt[16], qbase[16]; /* gather: regions per work list (RF_FAST_LISTS fast lists + generic) */
  int32_t emin, emax, cnt_le;
  int32_t wsum[16];
  int64_t seq_off;         /* first letter of the sequence in the batch text */
  unsigned long long cnt[CNT_N];
};
typedef SeqState BlockShared;

/* Region-expansion work list of one round: (sequence slot, node id) + one scratch slot each. */
#define RF_FAST_CLASSES 6
/* every size class (32 << class padded lags) has two work lists, lower and upper half of its lag
 * range: the regions a block expands together are then within 1.5x of each other */
#define RF_FAST_LISTS (2 * RF_FAST_CLASSES)
struct TaskQueue {
  int32_t *n_tasks;   /* device counter; may overshoot cap */
  int32_t cap;
  int32_t *task;      /* 2 ints per task */
  char *scratch;      /* cap * caps.wscratch_bytes */
  int32_t self;       /* slot of the sequence this Fold works on (block scope) */
  /* second list: regions small enough for the shared-memory expansion (expand_fast.cuh);
   * fast_cap == 0 (host emulation) sends everything to the first list */
  int32_t *n_fast;    /* RF_FAST_LISTS counters */
  int32_t fast_cap, fast_lv, fast_br;
  int32_t *fast_task; /* RF_FAST_LISTS lists of fast_cap tasks, 2 ints per task */
  /* third list: regions of up to big_lv positions for the block-per-region expansion
   * (expand_big.cuh); big_cap == 0 sends them to the first list */
  int32_t *n_big;
  int32_t big_cap, big_lv, big_br;
  int32_t *big_task;
};

RF_HD size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
RF_HD int imax(int a, int b) { return a > b ? a : b; }

/* sizes of the per-warp scratch pieces, in the order of WarpScratch */
RF_HD size_t wscratch_carve(WarpScratch &W, char *base, const Caps &c, const FoldParams &fp) {
  size_t off = 0;
  int nrank = fp.nb_mode > 0 ? fp.nb_mode : 1;
#define RF_WCARVE(field, type, count)                \
  off = align_up(off, 16);                           \
  W.field = (type *)(base ? base + off : nullptr);   \
  off += sizeof(type) * (size_t)(count);
  RF_WCARVE(keys, uint64_t, c.key_cap);
  RF_WCARVE(pos, uint16_t, c.L_max);
  RF_WCARVE(brcnt, uint16_t, c.L_max);
  RF_WCARVE(mask, uint32_t, 14 * (size_t)c.mask_words);
  RF_WCARVE(ranked, int32_t, nrank);
  RF_WCARVE(r_dE, int32_t, nrank);
  RF_WCARVE(r_stem, uint32_t, nrank);
  RF_WCARVE(r_nb, int32_t, nrank);
  RF_WCARVE(tmp, int32_t, 8);
  RF_WCARVE(sel, uint64_t, c.key_cap > RF_SEL_MIN ? RF_SEL_CAP + RF_SEL_SAMPLE : 0);
#undef RF_WCARVE
  return align_up(off, 128);
}

/* Carve the arena; with base == nullptr just computes the size.  Returns bytes used. */
RF_HD size_t arena_carve(Arena &A, char *base, const Caps &c, const FoldParams &fp) {
  size_t off = 0;
#define RF_CARVE(field, type, count)                              \
  off = align_up(off, 32);                                        \
  A.field = (type *)(base ? base + off : nullptr);                \
  off += sizeof(type) * (size_t)(count);
  int L = c.L_max;
  int sel_cap = imax(c.chunk_cap, c.new_cap + fp.max_stack);
  /* the two hash tables first: they are the part that must be zero before a run */
  RF_CARVE(seen, SeenEnt, c.seen_cap);
  RF_CARVE(nht, unsigned long long, c.nht_cap);
  RF_CARVE(code, uint8_t, L);
  RF_CARVE(node, NodeRec, c.node_cap);
  RF_CARVE(pool16, int16_t, c.pool16_cap);
  RF_CARVE(cand, Cand, c.cand_cap);
  RF_CARVE(ckid, CandKid, c.cand_cap);
  RF_CARVE(store, StoreRec, c.store_cap);
  RF_CARVE(store_db, char, (size_t)c.store_cap * (size_t)((L + 15) & ~15)); /* rows padded to 16 bytes */
  RF_CARVE(nlist, int32_t, c.nlist_cap);
  RF_CARVE(ncnt, int32_t, c.nlist_cap);
  RF_CARVE(ncoff, int32_t, c.nlist_cap);
  RF_CARVE(act, ActRec, c.nlist_cap);
  RF_CARVE(fresh, NewRec, c.new_cap);
  RF_CARVE(sel, int32_t, fp.max_stack);
  RF_CARVE(sel_idx, int32_t, fp.max_stack);
  RF_CARVE(sel_store, int32_t, fp.max_stack);
  RF_CARVE(beam, int32_t, fp.max_stack);
  RF_CARVE(traj_n, int32_t, c.step_cap);
  RF_CARVE(traj, int32_t, (size_t)c.step_cap * fp.max_stack);
  RF_CARVE(todo, int32_t, (size_t)c.todo_cap * 3);
  RF_CARVE(todo_id, int32_t, c.todo_cap);
  RF_CARVE(flag, int32_t, sel_cap);
  RF_CARVE(pref, int32_t, sel_cap);
  RF_CARVE(sel_e, int32_t, sel_cap);
#undef RF_CARVE
  return align_up(off, 256);
}
/* Byte offsets of the arena pieces (computed once on the host, bound per block/warp). */
struct ArenaOff {
  uint32_t seen, nht, code, node, pool16, cand, ckid, store, store_db, nlist, ncnt, ncoff, act, fresh, sel, sel_idx, sel_store, beam, traj_n, traj, todo, todo_id, flag, pref, sel_e;
};
RF_HD void arena_bind(Arena &A, char *base, const ArenaOff &o) {
  A.seen = (SeenEnt *)(base + o.seen);
  A.nht = (unsigned long long *)(base + o.nht);
  A.code = (uint8_t *)(base + o.code);
  A.node = (NodeRec *)(base + o.node);
  A.pool16 = (int16_t *)(base + o.pool16);
  A.cand = (Cand *)(base + o.cand);
  A.ckid = (CandKid *)(base + o.ckid);
  A.store = (StoreRec *)(base + o.store);
  A.store_db = (char *)(base + o.store_db);
  A.nlist = (int32_t *)(base + o.nlist);
  A.ncnt = (int32_t *)(base + o.ncnt);
  A.ncoff = (int32_t *)(base + o.ncoff);
  A.act = (ActRec *)(base + o.act);
  A.fresh = (NewRec *)(base + o.fresh);
  A.sel = (int32_t *)(base + o.sel);
  A.sel_idx = (int32_t *)(base + o.sel_idx);
  A.sel_store = (int32_t *)(base + o.sel_store);
  A.beam = (int32_t *)(base + o.beam);
  A.traj_n = (int32_t *)(base + o.traj_n);
  A.traj = (int32_t *)(base + o.traj);
  A.todo = (int32_t *)(base + o.todo);
  A.todo_id = (int32_t *)(base + o.todo_id);
  A.flag = (int32_t *)(base + o.flag);
  A.pref = (int32_t *)(base + o.pref);
  A.sel_e = (int32_t *)(base + o.sel_e);
}
#ifndef __CUDACC_RTC__
inline void arena_offsets(ArenaOff &o, const Caps &c, const FoldParams &fp) {
  Arena A;
  char *base = (char *)4096;
  arena_carve(A, base, c, fp);
  o.seen = (uint32_t)((char *)A.seen - base);
  o.nht = (uint32_t)((char *)A.nht - base);
  o.code = (uint32_t)((char *)A.code - base);
  o.node = (uint32_t)((char *)A.node - base);
  o.pool16 = (uint32_t)((char *)A.pool16 - base);
  o.cand = (uint32_t)((char *)A.cand - base);
  o.ckid = (uint32_t)((char *)A.ckid - base);
  o.store = (uint32_t)((char *)A.store - base);
  o.store_db = (uint32_t)((char *)A.store_db - base);
  o.nlist = (uint32_t)((char *)A.nlist - base);
  o.ncnt = (uint32_t)((char *)A.ncnt - base);
  o.ncoff = (uint32_t)((char *)A.ncoff - base);
  o.act = (uint32_t)((char *)A.act - base);
  o.fresh = (uint32_t)((char *)A.fresh - base);
  o.sel = (uint32_t)((char *)A.sel - base);
  o.sel_idx = (uint32_t)((char *)A.sel_idx - base);
  o.sel_store = (uint32_t)((char *)A.sel_store - base);
  o.beam = (uint32_t)((char *)A.beam - base);
  o.traj_n = (uint32_t)((char *)A.traj_n - base);
  o.traj = (uint32_t)((char *)A.traj - base);
  o.todo = (uint32_t)((char *)A.todo - base);
  o.todo_id = (uint32_t)((char *)A.todo_id - base);
  o.flag = (uint32_t)((char *)A.flag - base);
  o.pref = (uint32_t)((char *)A.pref - base);
  o.sel_e = (uint32_t)((char *)A.sel_e - base);
}
#endif
/* bytes at the start of an arena that hold the hash tables (zeroed by the host once) */
RF_HD size_t arena_table_bytes(const Caps &c) {
  return align_up(sizeof(SeenEnt) * (size_t)c.seen_cap, 16) + sizeof(unsigned long long) * (size_t)c.nht_cap;
}

/* Block-shared working set of the combine phase (shared memory on the device): the chunk of
 * stem combinations under examination, the list of beam members still to enumerate with the
 * running sum of their product sizes, and a small hash table that finds duplicates INSIDE a chunk
 * (children of different members can be the same structure; the first one in order wins). */
struct CombineStage {
  uint64_t *lo, *hi;          /* chunk_cap: fingerprints */
  uint64_t *mb_pre;           /* max_stack + 1: first flattened index of the m-th open member */
  uint64_t *mprod, *dprod;    /* max_stack: product of an open member / of an exhausted one */
  int32_t *flag, *pref, *e;   /* chunk_cap */
  int32_t *mem;               /* chunk_cap: open-member number of the item */
  int32_t *mb_idx;            /* max_stack: beam slot of the m-th open member */
  int32_t *htab;              /* hcap: item index + 1, 0 = empty */
  int32_t hcap;               /* power of two >= 2 * chunk_cap */
};
RF_HD size_t stage_carve(CombineStage &S, char *base, int chunk_cap, int max_stack) {
  size_t off = 0;
  int hcap = 64;
  while (hcap < 2 * chunk_cap) hcap <<= 1;
  S.hcap = hcap;
#define RF_SCARVE(field, type, count)              \
  off = align_up(off, 16);                         \
  S.field = (type *)(base ? base + off : nullptr); \
  off += sizeof(type) * (size_t)(count);
  RF_SCARVE(lo, uint64_t, chunk_cap);
  RF_SCARVE(hi, uint64_t, chunk_cap);
  RF_SCARVE(mb_pre, uint64_t, max_stack + 1);
  RF_SCARVE(mprod, uint64_t, max_stack);
  RF_SCARVE(dprod, uint64_t, max_stack);
  RF_SCARVE(flag, int32_t, chunk_cap);
  RF_SCARVE(pref, int32_t, chunk_cap);
  RF_SCARVE(e, int32_t, chunk_cap);
  RF_SCARVE(mem, int32_t, chunk_cap);
  RF_SCARVE(mb_idx, int32_t, max_stack);
  RF_SCARVE(htab, int32_t, hcap);
#undef RF_SCARVE
  return align_up(off, 16);
}

/* ------------------------------------------------------------------------------------ */
struct Fold {
  const DevEnergy *E;
  const FoldParams &prm;
  const Caps &caps;
  Arena A;
  SeqState *sh;
  EnergyView ev;
  TaskQueue Q;
  CombineStage cs; /* bound by the combine kernel only */
  int32_t gen;     /* generation tag of this fold's `seen` entries (1..65535); the launch-wide default
                      can be overridden per fold from device memory, so that a cached launch graph
                      stays valid from one fold to the next */
  RF_HD Fold(const DevEnergy *E_, const FoldParams &p, const Caps &c) : E(E_), prm(p), caps(c), gen(p.seen_gen) {}

  /* --- node geometry ------------------------------------------------------------------ */
  RF_FN const int16_t *node_br(const NodeRec &n) const { return A.pool16 + n.pool_off; }
  RF_FN const int16_t *node_segoff(const NodeRec &n) const {
    return A.pool16 + n.pool_off + 2 * n.br_n;
  }
  RF_FN int seg_start(const NodeRec &n, int k) const {
    return k == 0 ? n.ci + 1 : node_br(n)[2 * (k - 1) + 1] + 1;
  }
  /* segment (= number of branches before it) holding local index t */
  RF_FN int seg_of(const NodeRec &n, int t) const {
    const int16_t *so = node_segoff(n);
    int lo = 0, hi = n.br_n; /* answer in [0, br_n] */
    while (lo < hi) {
      int mid = (lo + hi + 1) >> 1;
      if (so[mid] <= t) lo = mid; else hi = mid - 1;
    }
    /* empty segments share offsets: the LAST segment with so[k] <= t holds t */
    return lo;
  }
  RF_FN int pos_of(const NodeRec &n, int t) const {
    int k = seg_of(n, t);
    return seg_start(n, k) + (t - node_segoff(n)[k]);
  }

  /* kcal test of find_best_consecutives (rafft.py:98-102) with the reference's floats:
   * float32(E_new/100) - float32(E_cur/100) < min_nrj, in double. */
  RF_FN bool passes(int e_cur, int dE) const {
    if (prm.min_nrj == 0.0) return dE < 0;
    double a = (double)(float)Exec::ddiv((double)(e_cur + dE), 100.0);
    double b = (double)(float)Exec::ddiv((double)e_cur, 100.0);
    return Exec::dadd(a, -b) < prm.min_nrj;
  }
  /* candidates of node v are sorted by dE: count the leading ones that pass for energy e */
  RF_FN int count_pass(const NodeRec &n, int e_cur) const {
    int lo = 0, hi = n.cand_n; /* first index that fails */
    if (prm.min_nrj == 0.0) return hi; /* every kept candidate has dE < 0 */
    while (lo < hi) {
      int mid = (lo + hi) >> 1;
      if (passes(e_cur, A.cand[n.cand_off + mid].dE)) lo = mid + 1; else hi = mid;
    }
    return lo;
  }

  /* dot-bracket rows of the structure store are padded to 16 bytes so that they move as vectors */
  struct alignas(16) Vec16 { uint64_t a, b; };
  RF_FN int db_stride() const { return (sh->L + 15) & ~15; }
  RF_FN char *db_row(int store_idx) const { return A.store_db + (size_t)store_idx * (size_t)db_stride(); }

  RF_FN void fail_at(int code, int line) {
    /* called by any thread: status only ever goes 0 -> code */
    sh->status = code;
    sh->fail_line = line;
  }
#define fail(code) fail_at(code, __LINE__)
  RF_FN void count(int which, long long v) {
    Exec::atomic_add64(&sh->cnt[which], (unsigned long long)v);
  }
  /* the same for a per-thread partial sum: one atomic per warp (every thread of the block calls) */
  RF_FN void count_all(int which, long long v) {
    v = Exec::warp_sum(v);
    if (Exec::lane() == 0 && v) count(which, v);
  }

  /* --- node creation --------------------------------------------------------------------
   * Fill node record `id` for the loop closed by (ci,cj) with branches copied from parent
   * ranges.  Returns false on pool overflow. */
  RF_FN bool build_node(int id, int ci, int cj, const NodeRec *par, int lo1, int hi1, int xp, int xq,
                        int lo2, int hi2, int lv) {
    int m = (hi1 - lo1) + (xp >= 0 ? 1 : 0) + (hi2 - lo2);
    int need = 2 * m + m + 2;
    int off = Exec::atomic_add(&sh->n_pool16, need);
    if (off + need > caps.pool16_cap) return false;
    int16_t *br = A.pool16 + off;
    const int16_t *pb = par ? node_br(*par) : nullptr;
    int k = 0;
    for (int b = lo1; b < hi1; b++, k++) { br[2 * k] = pb[2 * b]; br[2 * k + 1] = pb[2 * b + 1]; }
    if (xp >= 0) { br[2 * k] = (int16_t)xp; br[2 * k + 1] = (int16_t)xq; k++; }
    for (int b = lo2; b < hi2; b++, k++) { br[2 * k] = pb[2 * b]; br[2 * k + 1] = pb[2 * b + 1]; }
    int16_t *so = br + 2 * m;
    int acc = 0, start = ci + 1;
    uint64_t h = mix64(((uint64_t)(uint32_t)(ci + 1) << 32) | (uint32_t)cj);
    for (k = 0; k < m; k++) {
      so[k] = (int16_t)acc;
      acc += br[2 * k] - start;
      start = br[2 * k + 1] + 1;
      h = mix64(h ^ (((uint64_t)(uint16_t)br[2 * k] << 16) | (uint16_t)br[2 * k + 1])) + k;
    }
    so[m] = (int16_t)acc;
    acc += cj - start;
    so[m + 1] = (int16_t)acc;
    NodeRec &n = A.node[id];
    n.ci = ci; n.cj = cj; n.pool_off = off; n.br_n = m; n.lv = lv;
    n.cand_off = 0; n.cand_n = -1; n.n_ranked = 0; n.n_pos = 0; n.w_scan = 0; n.hash = h;
    n.nht_slot = -1;
    n.e_loop = loop_energy(ev, ci, cj, br, 0, m, -1, -1, 0, 0, lv);
    /* acc must equal lv: the node holds ALL unpaired positions of its loop */
    if (acc != lv) return false;
    return true;
  }
  RF_FN bool same_loop(const NodeRec &a, const NodeRec &b) const {
    if (a.ci != b.ci || a.cj != b.cj || a.br_n != b.br_n) return false;
    const int16_t *x = node_br(a), *y = node_br(b);
    for (int k = 0; k < 2 * a.br_n; k++)
      if (x[k] != y[k]) return false;
    return true;
  }
  /* Insert node `id` into the unique-loop table or find the identical loop already there. */
  RF_FN int nht_resolve(int id) {
    NodeRec &n = A.node[id];
    unsigned long long tag = ((n.hash >> 32) << 32) | (unsigned)(id + 1);
    unsigned mask = caps.nht_cap - 1;
    unsigned slot = (unsigned)n.hash & mask;
    for (int probe = 0; probe < caps.nht_cap; probe++, slot = (slot + 1) & mask) {
      unsigned long long cur = A.nht[slot];
      if (cur == 0) {
        cur = Exec::atomic_cas64(&A.nht[slot], 0ULL, tag);
        if (cur == 0) { n.nht_slot = (int32_t)slot; return id; }
      }
      if ((cur >> 32) == (tag >> 32)) {
        int other = (int)(cur & 0xffffffffu) - 1;
        if (other == id) return id;
        if (same_loop(A.node[other], n)) return other;
      }
    }
    return -1; /* table full */
  }

  /* --- warp-scope bitonic sort of n (power of two) 64-bit keys ---------------------------- */
  RF_NOINLINE static void wsort(uint64_t *keys, int n, bool descending) {
    for (int k = 2; k <= n; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        RF_WPAR_BEGIN(t, n >> 1)
          int i = ((t & ~(j - 1)) << 1) | (t & (j - 1)); /* lower index of the pair */
          int l = i | j;
          uint64_t a = keys[i], b = keys[l];
          bool up = ((i & k) == 0) != descending; /* ascending run? */
          if ((a > b) == up) { keys[i] = b; keys[l] = a; }
        RF_WPAR_END
      }
    }
  }

  /* --- node expansion: correlation -> top-n lags -> stem scan -> energy change ------------
   * Everything create_childs/find_best_consecutives compute for one unpaired region
   * (rafft.py:112-123, 86-109), done once per unique loop, by ONE warp (warp-scope phases).
   *
   * The region x[0..lv) is held as one-hot bit masks F_X (bit i = [x_i == X]) and their
   * reversals R_X (bit k = [x_{lv-1-k} == X]).  For lag p the partner of i is j = p - i, i.e.
   * reversed index k = i + s with s = lv-1-p, so "x_i = X pairs x_j = Y" for 32 positions at a
   * time is F_X & (R_Y >> s): the correlation is three popcounts per word, the stem scan
   * walks the set bits of the pairable mask. */
  RF_FN unsigned mword(const uint32_t *M, int a, int nw) const { return (a >= 0 && a < nw) ? M[a] : 0u; }
  /* bits [32*wi + s, 32*wi + s + 32) of mask M (nw words), any integer s */
  RF_FN unsigned mshift(const uint32_t *M, int wi, int s, int nw) const {
    int q = 32 * wi + s;
    int a = q >> 5, b = q & 31; /* arithmetic shift = floor */
    return Exec::funnel_r(mword(M, a, nw), mword(M, a + 1, nw), b);
  }

#define RF_EXPAND_VARS                                                                         \
  const int lv = n.lv, m = n.br_n;                                                             \
  const int nlag = 2 * lv - 1;                                                                 \
  const int nw = (lv + 31) >> 5, MW = caps.mask_words;                                         \
  uint32_t *const F = W.mask, *const R = W.mask + 4 * MW, *const CL = W.mask + 8 * MW,         \
                 *const CLR = W.mask + 9 * MW;                                                 \
  const uint8_t *const code = A.code;                                                          \
  int nk = 32;                                                                                 \
  while (nk < nlag) nk <<= 1;                                                                  \
  const int nr = nlag < prm.nb_mode ? nlag : prm.nb_mode;                                      \
  (void)m; (void)nw; (void)F; (void)R; (void)CL; (void)CLR; (void)code; (void)nk; (void)nr;

  /* Warp-scope phases of one region expansion; W is the task's scratch slot. */
  RF_FN void expand_e1(NodeRec &n, WarpScratch &W) {
    RF_EXPAND_VARS
    /* 1. local index -> original position / branch count; one-hot masks; contiguity masks
     *    (CL bit t: position t continues position t-1 in ORIGINAL coordinates) */
    RF_WPAR_BEGIN(x, 10 * MW)
      W.mask[x] = 0;
    RF_WPAR_END
    RF_WPAR_BEGIN(x, 8)
      W.tmp[x] = 0;
    RF_WPAR_END
    RF_WPAR_BEGIN(t, lv)
      int k = seg_of(n, t);
      int first = node_segoff(n)[k];
      int p = seg_start(n, k) + (t - first);
      W.pos[t] = (uint16_t)p;
      W.brcnt[t] = (uint16_t)k;
      int c = code[p], u = lv - 1 - t;
      Exec::atomic_or(&F[c * MW + (t >> 5)], 1u << (t & 31));
      Exec::atomic_or(&R[c * MW + (u >> 5)], 1u << (u & 31));
      if (t > first) {
        Exec::atomic_or(&CL[t >> 5], 1u << (t & 31));
        Exec::atomic_or(&CLR[u >> 5], 1u << (u & 31));
      }
    RF_WPAR_END
    /* 2. pair-weighted correlation of the region against its reverse (utils.py:115-132):
     *    cor[p] = sum_{i+j=p} w(s_i,s_j) = 2 * sum over unordered pairs, exact integers;
     *    c[p] = cor[p] / (min(p, 2lv-2-p) + 1.0) as the reference's float64.
     *    key = order-preserving bits of c[p] with the lag in the low bits, so that sorting
     *    keys descending is the reference's (value desc, lag desc) order (rafft.py:118,92). */
    const int wGC = prm.wi[1][2], wAU = prm.wi[0][3], wGU = prm.wi[2][3];
    /* bit planes of the base codes (A,C,G,U = 0..3): with x = FH ^ RH', y = FL ^ RL' (R' shifted to
     * the lag) the G.C pairs are x&y&(FH^FL), A.U x&y&~(FH^FL), G.U ~x&y&FH -- both orientations
     * at once, so only the first half of the anti-diagonal is visited */
    uint32_t *const FH = W.mask + 10 * MW, *const FL = FH + MW, *const RH = FL + MW, *const RL = RH + MW;
    RF_WPAR_BEGIN(x, MW)
      FH[x] = F[2 * MW + x] | F[3 * MW + x];
      FL[x] = F[1 * MW + x] | F[3 * MW + x];
      RH[x] = R[2 * MW + x] | R[3 * MW + x];
      RL[x] = R[1 * MW + x] | R[3 * MW + x];
    RF_WPAR_END
    RF_WPAR_BEGIN(p, nk)
      uint64_t key = 0;
      if (p < nlag) {
        const int s = lv - 1 - p;
        const int i0 = p < lv ? 0 : p - lv + 1;
        const int d = p < lv ? p + 1 : 2 * lv - 1 - p;
        const int iend = i0 + ((d + 1) >> 1); /* exclusive */
        int acc = 0;
        for (int wi = i0 >> 5; wi <= ((iend - 1) >> 5); wi++) {
          const unsigned fh = FH[wi], fl = FL[wi];
          const unsigned x = fh ^ mshift(RH, wi, s, nw), y = fl ^ mshift(RL, wi, s, nw);
          unsigned range = 0xffffffffu;
          if (32 * wi < i0) range &= 0xffffffffu << (i0 - 32 * wi);
          if (32 * wi + 32 > iend) range &= 0xffffffffu >> (32 * wi + 32 - iend);
          const unsigned xy = x & y & range, t = fh ^ fl;
          acc += wGC * Exec::popc(xy & t) + wAU * Exec::popc(xy & ~t) + wGU * Exec::popc(~x & y & fh & range);
        }
        double cor = Exec::ddiv((double)(2 * (int64_t)acc), 16.0);
        double norm = (double)(p < lv ? p : 2 * lv - 2 - p) + 1.0;
        double c = Exec::ddiv(cor, norm);
        key = (Exec::dbits(c) & ~0x1FFFULL) | (uint64_t)p;
      }
      W.keys[p] = key;
    RF_WPAR_END
  }
  RF_FN void expand_e2(NodeRec &n, WarpScratch &W) {
    RF_EXPAND_VARS
    /* 3. rank: the first nb_mode lags in (value desc, lag desc) order.  Long regions select
     *    instead of sorting everything: a pivot from a sorted sample of the keys is aimed at
     *    ~3 nb_mode keys above it; those are compacted and sorted (the keys are distinct -- the
     *    lag sits in their low bits -- so the result is the same as the full sort's).  A pivot
     *    that leaves fewer than nb_mode or more than RF_SEL_CAP keys falls back to the full sort. */
    if (nk > RF_SEL_MIN && caps.key_cap > RF_SEL_MIN) {
      uint64_t *const smp = W.sel + RF_SEL_CAP;
      const int stride = nk / RF_SEL_SAMPLE;
      RF_WPAR_BEGIN(x, RF_SEL_SAMPLE)
        smp[x] = W.keys[x * stride + (stride >> 1)];
      RF_WPAR_END
      wsort(smp, RF_SEL_SAMPLE, true);
      const int pidx = (int)((3LL * nr * RF_SEL_SAMPLE + nlag - 1) / nlag) + 4;
      if (pidx < RF_SEL_SAMPLE) {
        const uint64_t pivot = smp[pidx];
        RF_WPAR_BEGIN(x, nk)
          const uint64_t k = W.keys[x];
          if (k >= pivot) {
            const int at = Exec::atomic_add(&W.tmp[4], 1);
            if (at < RF_SEL_CAP) W.sel[at] = k;
          }
        RF_WPAR_END
        const int cnt = W.tmp[4];
        if (cnt >= nr && cnt <= RF_SEL_CAP) {
          int np = 32;
          while (np < cnt) np <<= 1;
          RF_WPAR_BEGIN(x, np - cnt)
            W.sel[cnt + x] = 0;
          RF_WPAR_END
          wsort(W.sel, np, true);
          RF_WPAR_BEGIN(r, nr)
            W.ranked[r] = (int)(W.sel[r] & 0x1FFFULL);
          RF_WPAR_END
          return;
        }
      }
    }
    wsort(W.keys, nk, true);
    RF_WPAR_BEGIN(r, nr)
      W.ranked[r] = (int)(W.keys[r] & 0x1FFFULL);
    RF_WPAR_END
  }
  RF_FN void expand_e3(NodeRec &n, WarpScratch &W) {
    RF_EXPAND_VARS
    /* 4. per ranked lag: best-scoring consecutive stem (window_slide, rafft.py:36-83), walking
     *    only the positions whose pair weight is non-zero, and the energy change of forming
     *    it (eval_one_struct difference, rafft.py:98). */
    const double dGC = prm.w[1][2], dAU = prm.w[0][3], dGU = prm.w[2][3];
    RF_WPAR_BEGIN(r, nr)
      const int p = W.ranked[r];
      const int s = lv - 1 - p;
      const int i0 = p < lv ? 0 : p - lv + 1;
      const int d = p < lv ? p + 1 : 2 * lv - 1 - p;
      const int len2 = (d >> 1) + (d & 1);
      const int iend = i0 + len2; /* exclusive */
      int max_nb = 0, tmp_max = 0, max_i = 0, max_j = 0, last_i = -2;
      double max_score = 0.0, prev = 0.0;
      for (int wi = i0 >> 5; wi <= ((iend - 1) >> 5); wi++) {
        unsigned fa = F[0 * MW + wi], fc = F[1 * MW + wi], fg = F[2 * MW + wi], fu = F[3 * MW + wi];
        unsigned ra = mshift(R + 0 * MW, wi, s, nw), rc = mshift(R + 1 * MW, wi, s, nw),
                 rg = mshift(R + 2 * MW, wi, s, nw), ru = mshift(R + 3 * MW, wi, s, nw);
        unsigned gc = (fg & rc) | (fc & rg), au = (fa & ru) | (fu & ra), gu = (fg & ru) | (fu & rg);
        if (dGC == 0.0) gc = 0;
        if (dAU == 0.0) au = 0;
        if (dGU == 0.0) gu = 0;
        unsigned range = 0xffffffffu;
        if (32 * wi < i0) range &= 0xffffffffu << (i0 - 32 * wi);
        if (32 * wi + 32 > iend) range &= 0xffffffffu >> (32 * wi + 32 - iend);
        unsigned P = (gc | au | gu) & range;
        /* in-segment: both arms continue in original coordinates (rafft.py:67), t > 0 */
        unsigned IS = CL[wi] & mshift(CLR, wi, s - 1, nw);
        if ((i0 >> 5) == wi) IS &= ~(1u << (i0 & 31));
        while (P) {
          const int b = Exec::ffs(P) - 1;
          P &= P - 1;
          const int i = 32 * wi + b;
          const double w = ((gc >> b) & 1) ? dGC : ((au >> b) & 1) ? dAU : dGU;
          const bool adjacent = last_i == i - 1;
          double tot = w;
          if ((IS >> b) & 1) tot = Exec::dmul(Exec::dadd(adjacent ? prev : 0.0, w), w);
          tmp_max = adjacent ? tmp_max + 1 : 1;
          if (tot >= max_score && (int)W.pos[p - i] - (int)W.pos[i] > prm.min_hp) {
            max_score = tot; max_nb = tmp_max; max_i = i; max_j = p - i;
          }
          prev = tot;
          last_i = i;
        }
      }
      W.r_stem[r] = ((uint32_t)max_i << 16) | (uint32_t)max_j;
      W.r_nb[r] = max_nb;
      Exec::atomic_add(&W.tmp[0], len2);
    RF_WPAR_END
  }
  RF_FN void expand_e4(NodeRec &n, WarpScratch &W) {
    RF_EXPAND_VARS
    /* the energy change of forming each stem (eval_one_struct difference, rafft.py:98) */
    RF_WPAR_BEGIN(r, nr)
      const int max_nb = W.r_nb[r];
      const int max_i = (int)(W.r_stem[r] >> 16), max_j = (int)(W.r_stem[r] & 0xffff);
      int dE = 0x7fffffff;
      if (max_nb > 0) {
        const int16_t *br = node_br(n);
        int e = 0;
        /* loop closed by the innermost pair: the region's branches between its two arms */
        int a0 = W.pos[max_i], b0 = W.pos[max_j];
        e += loop_energy(ev, a0, b0, br, W.brcnt[max_i], W.brcnt[max_j], -1, -1, 0, 0,
                         max_j - max_i - 1);
        /* loops between consecutive pairs of the stem (stack, or a junction if the stem
         * steps over existing helices: contiguity breaks do not end a stem, rafft.py:71-74) */
        int ak = a0, bk = b0;
        for (int k = 0; k + 1 < max_nb; k++) {
          int an = W.pos[max_i - k - 1], bn = W.pos[max_j + k + 1];
          e += loop_energy(ev, an, bn, br, W.brcnt[max_i - k - 1], W.brcnt[max_i - k], ak, bk,
                           W.brcnt[max_j + k], W.brcnt[max_j + k + 1], 0);
          ak = an; bk = bn;
        }
        /* the region's own loop with the stem as a new branch */
        int io = max_i - max_nb + 1, jo = max_j + max_nb - 1;
        e += loop_energy(ev, n.ci, n.cj, br, 0, W.brcnt[io], ak, bk, W.brcnt[jo], m,
                         io + (lv - 1 - jo));
        dE = e - n.e_loop;
      }
      W.r_dE[r] = dE;
      /* 5. stems below the threshold get a (dE, rank) key: stable order by dE (rafft.py:108) */
      const bool kept = max_nb > 0 && dE < prm.cand_thr;
      W.keys[r] = kept ? (((uint64_t)(uint32_t)(dE ^ 0x80000000)) << 32) | (uint32_t)r : ~0ULL;
      if (max_nb > 0) Exec::atomic_add(&W.tmp[1], 1);
      if (kept) Exec::atomic_add(&W.tmp[2], 1);
    RF_WPAR_END
  }
  RF_FN void expand_e5(NodeRec &n, WarpScratch &W) {
    RF_EXPAND_VARS
    int nrp = 32;
    while (nrp < nr) nrp <<= 1;
    RF_WPAR_BEGIN(r, nrp - nr)
      W.keys[nr + r] = ~0ULL;
    RF_WPAR_END
    const int keep_n = W.tmp[2];
    if (keep_n > 0) wsort(W.keys, nrp, false);
    RF_WSINGLE_BEGIN
      n.n_ranked = nr; n.n_pos = W.tmp[1]; n.w_scan = W.tmp[0];
      int off = Exec::atomic_add(&sh->n_cand, keep_n);
      if (off + keep_n > caps.cand_cap) { fail(ST_ERR_CAPACITY); off = -1; }
      W.tmp[3] = off;
      count(CNT_UNIQ_NODES, 1);
      count(CNT_UNIQ_W_CORR, (long long)lv * (lv + 1) / 2);
      count(CNT_UNIQ_W_SCAN, W.tmp[0]);
    RF_WSINGLE_END
    const int off = W.tmp[3];
    if (off >= 0) {
      RF_WPAR_BEGIN(at, keep_n)
        const int r = (int)(W.keys[at] & 0xffffffffu);
        Cand &c = A.cand[off + at];
        const uint32_t ij = W.r_stem[r];
        const int mip = ij >> 16, mjp = ij & 0xffff, nb = W.r_nb[r];
        c.dE = W.r_dE[r];
        cand_geometry(c, W.pos, W.brcnt, mip, mjp, nb);
        CandKid &ck = A.ckid[off + at];
        ck.inner_id = (mjp - mip > 1) ? -1 : -2;                       /* rafft.py:130 */
        ck.outer_id = (mip - (nb - 1) > 0 || mjp + nb < lv) ? -1 : -2; /* rafft.py:141 */
      RF_WPAR_END
    }
    RF_WSINGLE_BEGIN
      n.cand_off = off < 0 ? 0 : off;
      n.cand_n = off < 0 ? 0 : keep_n;
    RF_WSINGLE_END
  }
  /* all phases in a row (host emulation) */
  RF_FN void expand_node(int v, char *slot) {
    WarpScratch W;
    wscratch_carve(W, slot, caps, prm);
    NodeRec &n = A.node[v];
    expand_e1(n, W); expand_e2(n, W); expand_e3(n, W); expand_e4(n, W); expand_e5(n, W);
  }

  /* --- seen set (global per sequence, rafft.py:196-200) ------------------------------- */
  RF_FN uint64_t seen_tag(uint64_t hi) const {
    return (hi & 0x0000ffffffffffffULL) | ((uint64_t)(uint32_t)gen << 48);
  }
  RF_FN bool seen_has(uint64_t lo, uint64_t hi) const {
    const uint64_t th = seen_tag(hi), gen = (uint64_t)(uint32_t)this->gen;
    unsigned mask = caps.seen_cap - 1;
    unsigned slot = (unsigned)(lo ^ (hi >> 7)) & mask;
    for (;;) {
      const SeenEnt en = A.seen[slot];
      if ((en.hi >> 48) != gen) return false;
      if (en.lo == lo && en.hi == th) return true;
      slot = (slot + 1) & mask;
    }
  }
  /* never concurrent with seen_has (block barriers separate the phases); two inserts of one
   * phase carry different fingerprints (the in-chunk table removed the duplicates) */
  RF_FN void seen_put(uint64_t lo, uint64_t hi, bool count = true) {
    const uint64_t th = seen_tag(hi), gen = (uint64_t)(uint32_t)this->gen;
    unsigned mask = caps.seen_cap - 1;
    unsigned slot = (unsigned)(lo ^ (hi >> 7)) & mask;
    for (;;) {
      const unsigned long long cur = A.seen[slot].hi;
      if ((cur >> 48) != gen &&
          Exec::atomic_cas64((unsigned long long *)&A.seen[slot].hi, cur, th) == cur) {
        A.seen[slot].lo = lo;
        if (count) Exec::atomic_add(&sh->n_seen, 1);
        return;
      }
      slot = (slot + 1) & mask;
    }
  }

  /* decode combination index -> (energy, fingerprint) of the child of store entry S:
   * itertools.product over the nodes that have candidates, last list varies fastest */
  RF_FN void combo_eval(const StoreRec &S, uint64_t combo, int &e, uint64_t &lo, uint64_t &hi) const {
    e = S.e; lo = S.fp_lo; hi = S.fp_hi;
    /* the regions with candidates, compacted by the first pass of phase_combine (A.act); four at a
     * time: the digits are decoded first, then the four stem records are fetched together
     * (independent loads in flight instead of one load latency per region) */
    const ActRec *act = A.act + S.nl_off;
    for (int a0 = (int)S.nact - 1; a0 >= 0; a0 -= 4) {
      int at[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int a = a0 - j;
        at[j] = -1;
        if (a < 0) continue;
        const ActRec ar = act[a];
        const uint32_t c = (uint32_t)ar.c;
        uint32_t dgt = 0;
        if (c != 1) {
          if ((combo >> 32) == 0) { /* the usual case: 32-bit division */
            const uint32_t c32 = (uint32_t)combo, q = c32 / c;
            dgt = c32 - q * c; combo = q;
          } else {
            dgt = (uint32_t)(combo % c); combo /= c;
          }
        }
        at[j] = ar.off + (int)dgt;
      }
      int de[4];
      uint64_t l[4], h[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        de[j] = 0; l[j] = 0; h[j] = 0;
        if (at[j] >= 0) {
          const Cand &cd = A.cand[at[j]];
          de[j] = cd.dE; l[j] = cd.fp_lo; h[j] = cd.fp_hi;
        }
      }
#pragma unroll
      for (int j = 0; j < 4; j++) { e += de[j]; lo += l[j]; hi += h[j]; }
    }
    if (lo == 0) lo = 1; /* 0 marks an empty slot of the seen table */
  }

  /* in-chunk duplicate detection: item i (fingerprint cs.lo/hi[i]) enters the table; items with
   * the same fingerprint share a slot, which ends up holding the smallest item index + 1 */
  RF_FN int chunk_insert(int i) {
    const uint64_t lo = cs.lo[i], hi = cs.hi[i];
    const unsigned mask = (unsigned)cs.hcap - 1;
    unsigned slot = (unsigned)(lo ^ (hi >> 11) ^ (lo >> 29)) & mask;
    for (;;) {
      int v = cs.htab[slot];
      if (v == 0) {
        v = Exec::atomic_cas(&cs.htab[slot], 0, i + 1);
        if (v == 0) return (int)slot;
      }
      if (cs.lo[v - 1] == lo && cs.hi[v - 1] == hi) {
        Exec::atomic_min(&cs.htab[slot], i + 1);
        return (int)slot;
      }
      slot = (slot + 1) & mask;
    }
  }
  /* open member holding flattened combination index gi: last m with mb_pre[m] <= gi */
  RF_FN int member_of(uint64_t gi, int nm) const {
    int lo = 0, hi = nm - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (cs.mb_pre[mid] <= gi) lo = mid; else hi = mid - 1;
    }
    return lo;
  }

  /* ------------------------------------------------------------------------------------ */
  /* ---- phase: start of a fold ---------------------------------------------------------- */
  RF_FN void phase_init(const char *seq, int L, int seq_index) {
    /* ---- reset per-sequence state */
    RF_SINGLE_BEGIN
      sh->seq_index = seq_index; sh->L = L; sh->status = ST_OK;
      sh->n_node = 0; sh->n_pool16 = 0; sh->n_cand = 0; sh->n_store = 0; sh->n_nlist = 0;
      sh->n_seen = 0; sh->n_todo = 0; sh->n_expand = 0; sh->beam_n = 0; sh->n_new = 0;
      sh->nb_branch = 0; sh->n_steps = 0; sh->n_traj = 0;
      sh->active = 1; sh->stage = 0; sh->expand_pending = 0; sh->fail_line = 0; sh->exp_from = 0;
      for (int k = 0; k < CNT_N; k++) sh->cnt[k] = 0;
      sh->cnt[CNT_T0_NS] = (unsigned long long)Exec::now_ns();
      if (L > caps.L_max) sh->status = ST_ERR_CAPACITY;
      else if (L < 1) sh->status = ST_ERR_SEQ; /* the reference cannot fold an empty sequence */
    RF_SINGLE_END
    if (sh->status == ST_OK) {
      RF_PAR_BEGIN(i, L)
        char ch = seq[i];
        int c = ch == 'A' ? 0 : ch == 'C' ? 1 : ch == 'G' ? 2 : ch == 'U' ? 3 : -1;
        if (c < 0) { sh->status = ST_ERR_SEQ; c = 0; }
        A.code[i] = (uint8_t)c;
        A.store_db[i] = '.';
      RF_PAR_END
    }
    if (sh->status == ST_OK) {
      ev.P = E; ev.s = A.code; ev.L = L;
      /* root: exterior loop without branches, completely unfolded structure (rafft.py:226-231) */
      RF_SINGLE_BEGIN
        sh->n_node = 1;
        if (!build_node(0, -1, L, nullptr, 0, 0, -1, -1, 0, 0, L)) fail(ST_ERR_CAPACITY);
        nht_resolve(0);
        StoreRec &S = A.store[0];
        S.e = 0; S.nl_off = 0; S.nl_n = 1; S.done = 0; S.fp_lo = 0; S.fp_hi = 0; S.prod = 0; S.nact = 0;
        A.nlist[0] = 0;
        sh->n_nlist = 1; sh->n_store = 1;
        A.beam[0] = 0; sh->beam_n = 1;
      RF_SINGLE_END
    }
    RF_SINGLE_BEGIN
      if (sh->status != ST_OK) sh->active = 0;
    RF_SINGLE_END
  }

  /* ---- phase: results out, hash tables left empty for the next sequence of this arena ---- */
  RF_FN void phase_finish(OutBuf &out) {
    emit(out);
    /* leave the unique-loop table empty for the next sequence of this arena (the `seen` table
     * is generation-tagged and needs no clean-up) */
    RF_PAR_BEGIN(i, sh->n_node)
      int s = A.node[i].nht_slot;
      if (s >= 0) A.nht[s] = 0;
    RF_PAR_END
  }

  RF_FN void bind_ev() { ev.P = E; ev.s = A.code; ev.L = sh->L; }
  RF_FN void finish_fold() { /* block scope: the fold has converged or failed */
    RF_SINGLE_BEGIN
      sh->active = 0;
      sh->cnt[CNT_T1_NS] = (unsigned long long)Exec::now_ns();
    RF_SINGLE_END
  }

  /* One bfs_pairs call (rafft.py:156-216) is a sequence of phases:
   *   gather -> [region expansions e1..e5, warp tasks] -> combine -> select -> materialise.
   * ---- phase: record the beam (once per step) and queue the regions still to be expanded */
  RF_FN void phase_gather() {
    if (!sh->active) return;
    const int beam_n = sh->beam_n;
    /* glob_traj += [glob_tree] (rafft.py:161) */
    RF_SINGLE_BEGIN
      if (sh->stage == 0) {
        if (sh->n_steps >= caps.step_cap) fail(ST_ERR_CAPACITY);
        else {
          A.traj_n[sh->n_steps] = beam_n;
          for (int b = 0; b < beam_n; b++) A.traj[(size_t)sh->n_steps * prm.max_stack + b] = A.beam[b];
          sh->n_steps++;
          sh->cnt[CNT_STEPS]++;
        }
        sh->stage = 1;
      }
      sh->expand_pending = 0;
      for (int c = 0; c < 16; c++) sh->qcnt[c] = 0;
    RF_SINGLE_END
    if (sh->status != ST_OK) { finish_fold(); return; }
    /* P1: the loops that have not been expanded yet become warp tasks.  They are exactly the loop
     * records created since the last gather that are the first of their kind (every new record
     * belongs to a child that was selected into the beam); one global reservation per work list
     * and block instead of one atomic per region. */
    const int from = sh->exp_from, n_new_nodes = sh->n_node - from;
#define RF_GATHER_CLASS(nd, cls)                                                              \
  int cls = RF_FAST_LISTS;                                                                    \
  if (Q.fast_cap > 0 && (nd).lv <= Q.fast_lv && (nd).br_n <= Q.fast_br) {                     \
    int k2 = 0;                                                                               \
    while ((32 << k2) < 2 * (nd).lv - 1) k2++;                                                \
    cls = 2 * k2 + (2 * (nd).lv - 1 > (24 << k2) ? 1 : 0);                                    \
  } else if (Q.big_cap > 0 && (nd).lv <= Q.big_lv && (nd).br_n <= Q.big_br) {                 \
    cls = RF_FAST_LISTS + 1;                                                                  \
  }
    /* A region that cannot hold a stem needs no expansion: a single position, or the inside of a
     * hairpin loop too short for a pair to span min_hp positions (pos[j] - pos[i] = j - i there,
     * rafft.py:78).  The reference still ranks and scans its lags; those counts have a closed form
     * when every lag is ranked. */
    RF_PAR_BEGIN(x, n_new_nodes)
      NodeRec &nd = A.node[from + x];
      if (nd.nht_slot >= 0 && nd.cand_n == -1) {
        const int nlag0 = 2 * nd.lv - 1;
        if ((nd.lv == 1 || (nd.br_n == 0 && nd.lv <= prm.min_hp + 1)) && nlag0 <= prm.nb_mode) {
          long long ws = 0;
          for (int d = 1; d <= nd.lv; d++) ws += ((d + 1) >> 1) * (d < nd.lv ? 2 : 1);
          nd.cand_off = 0; nd.cand_n = 0; nd.n_ranked = nlag0; nd.n_pos = 0; nd.w_scan = ws;
          count(CNT_TRIVIAL_NODES, 1);
        } else {
          RF_GATHER_CLASS(nd, cls)
          Exec::atomic_add(&sh->qcnt[cls], 1);
        }
      }
    RF_PAR_END
    RF_SINGLE_BEGIN
      for (int c = 0; c <= RF_FAST_LISTS + 1; c++) {
        if (sh->qcnt[c])
          sh->qbase[c] = Exec::atomic_add(c < RF_FAST_LISTS ? Q.n_fast + c : c == RF_FAST_LISTS ? Q.n_tasks : Q.n_big, sh->qcnt[c]);
        sh->qcnt[c] = 0;
      }
    RF_SINGLE_END
    RF_PAR_BEGIN(x, n_new_nodes)
      NodeRec &nd = A.node[from + x];
      if (nd.nht_slot >= 0 && nd.cand_n == -1) {
        RF_GATHER_CLASS(nd, cls)
        const bool fast = cls < RF_FAST_LISTS, big = cls > RF_FAST_LISTS;
        int32_t *list = fast ? Q.fast_task + 2 * (size_t)cls * (size_t)Q.fast_cap : big ? Q.big_task : Q.task;
        const int cap = fast ? Q.fast_cap : big ? Q.big_cap : Q.cap;
        const int at = sh->qbase[cls] + Exec::atomic_add(&sh->qcnt[cls], 1);
        if (at < cap) { list[2 * at] = Q.self; list[2 * at + 1] = from + x; nd.cand_n = -2; }
        else sh->expand_pending = 1; /* next round */
      }
    RF_PAR_END
#undef RF_GATHER_CLASS
    RF_SINGLE_BEGIN
      if (!sh->expand_pending) sh->exp_from = sh->n_node;
    RF_SINGLE_END
  }

  /* ---- phase: combine stems of independent regions (rafft.py:176-203) */
  RF_FN void phase_combine() {
    if (!sh->active || sh->expand_pending) return;
    if (sh->status != ST_OK) { finish_fold(); return; }
    const int beam_n = sh->beam_n;
    /* ---- P2: combine stems of independent regions (rafft.py:176-203) */
    /* P2a: per beam member, how many stems of each of its regions pass (rafft.py:102), the size
     *      of its product, and the reference's work counters (create_childs runs for every
     *      region of every beam member) */
    long long t_exp = 0, t_pos = 0, t_corr = 0, t_scan = 0, t_eval = 0;
    for (int x = Exec::tid(); x < cs.hcap; x += Exec::stride()) cs.htab[x] = 0; /* in-chunk table of the first chunk */
    RF_PAR_BEGIN(b, beam_n)
      StoreRec &S = A.store[A.beam[b]];
      long long c_pos = 0, c_corr = 0, c_scan = 0, c_eval = 0;
      int mact = 0;
      uint64_t prod = 1;
      for (int k = 0; k < S.nl_n; k++) {
        const NodeRec &nd = A.node[A.nlist[S.nl_off + k]];
        c_pos += nd.lv; c_corr += (long long)nd.lv * (nd.lv + 1) / 2; c_scan += nd.w_scan; c_eval += nd.n_pos;
        if (!S.done) {
          int c = count_pass(nd, S.e);
          A.ncnt[S.nl_off + k] = c;
          A.ncoff[S.nl_off + k] = nd.cand_off;
          if (c > 0) {
            ActRec ar; ar.c = c; ar.off = nd.cand_off;
            A.act[S.nl_off + mact] = ar;
            mact++;
            prod = prod > (1ULL << 62) / (uint64_t)c ? (1ULL << 62) : prod * (uint64_t)c;
          }
        }
      }
      t_exp += S.nl_n; t_pos += c_pos; t_corr += c_corr; t_scan += c_scan; t_eval += c_eval;
      if (!S.done) {
        S.nact = mact;
        S.prod = mact ? prod : 0;
        if (mact == 0) S.done = 1;
      }
      cs.mprod[b] = S.done ? 0 : S.prod;
      cs.dprod[b] = S.done ? S.prod : 0;
    RF_PAR_END
    count_all(CNT_NODE_EXP, t_exp); count_all(CNT_NODE_POS, t_pos); count_all(CNT_W_CORR, t_corr);
    count_all(CNT_W_SCAN, t_scan); count_all(CNT_N_EVAL, t_eval); count_all(CNT_N_EVAL_EXPAND, t_eval);
    /* the members still to enumerate ("open"), in beam order, and the running sum of their
     * product sizes: the enumeration below runs over this flattened index space */
    const uint64_t SAT = 1ULL << 62;
    RF_SINGLE_BEGIN
      sh->n_new = 0; sh->nb_branch = 0; sh->cur_b = 0;
      int nm = 0;
      uint64_t pre = 0;
      for (int b = 0; b < beam_n; b++) {
        if (cs.mprod[b]) {
          cs.mb_idx[nm] = b; cs.mb_pre[nm] = pre; nm++;
          pre = pre + cs.mprod[b] > SAT ? SAT : pre + cs.mprod[b];
        }
      }
      cs.mb_pre[nm] = pre;
      sh->tmp0 = nm;
    RF_SINGLE_END
    /* P2b: the open members are enumerated one after the other while nb_branch < max_branch
     *      (rafft.py:180-203), a chunk of the flattened (member, combination) space at a time */
    const int CH = caps.chunk_cap;
    const int nm = sh->tmp0;
    const uint64_t total = cs.mb_pre[nm];
    uint64_t g = 0;
    bool cut = false;
    /* Few barriers per chunk (a sequence walks one or two chunks per step, so the chain of
     * barrier-separated phases IS the latency of this kernel): the running counts live in
     * registers (every thread derives the same values), evaluation and in-chunk insertion are one
     * pass, the in-chunk table is cleared for the next chunk while the children are written, and
     * the `break` position is only looked for when a cut happens. */
    int n_new_r = 0, nb_branch_r = 0; /* sh->n_new, sh->nb_branch (zeroed above) */
    bool failed = false;
    while (prm.max_branch > 0 && g < total && !cut) {
      const int chunk = (int)((total - g) < (uint64_t)CH ? (total - g) : (uint64_t)CH);
      /* evaluate + look up in `seen` (rafft.py:192-196); the same child reached from two members of
       * this chunk: only the first one is new -- unseen items enter the in-chunk table at once (an
       * item's fingerprint is published before its table entry) */
      for (int i = Exec::tid(); i < chunk; i += Exec::stride()) {
        const uint64_t gi = g + (uint64_t)i;
        const int m = member_of(gi, nm);
        const StoreRec &S = A.store[A.beam[cs.mb_idx[m]]];
        int e_; uint64_t lo_, hi_;
        combo_eval(S, gi - cs.mb_pre[m], e_, lo_, hi_);
        cs.e[i] = e_; cs.lo[i] = lo_; cs.hi[i] = hi_; cs.mem[i] = m;
        const int unseen = !seen_has(lo_, hi_);
        cs.flag[i] = unseen;
        if (unseen) { Exec::fence_block(); cs.pref[i] = chunk_insert(i); }
      }
      Exec::sync();
      for (int i = Exec::tid(); i < chunk; i += Exec::stride())
        if (cs.flag[i]) cs.flag[i] = cs.htab[cs.pref[i]] == i + 1;
      const int nnew = Exec::scan_flags(cs.flag, cs.pref, chunk, sh->wsum);
      /* the reference breaks right after the combination that brings nb_branch to
       * max_branch: unseen children number nb0 .. of this chunk are kept up to that one */
      const int base = n_new_r;
      const int room = prm.max_branch - nb_branch_r; /* > 0 */
      const int take = nnew < room ? nnew : room;
      if (base + take > caps.new_cap || 2 * (sh->n_seen + take) + 2 > caps.seen_cap) { failed = true; break; }
      const bool cutting = nnew >= room;
      for (int i = Exec::tid(); i < chunk; i += Exec::stride()) {
        if (cs.flag[i] && cs.pref[i] < take) {
          const int m = cs.mem[i];
          NewRec &r = A.fresh[base + cs.pref[i]];
          r.e = cs.e[i]; r.parent_slot = cs.mb_idx[m]; r.combo = g + (uint64_t)i - cs.mb_pre[m];
          r.fp_lo = cs.lo[i]; r.fp_hi = cs.hi[i];
          seen_put(cs.lo[i], cs.hi[i], false); /* counted once per chunk below: no atomic per child */
          if (cutting && cs.pref[i] == take - 1) sh->xcut = i;
        }
      }
      if (Exec::tid() == 0) sh->n_seen += take; /* nobody reads it before the barrier */
      for (int x = Exec::tid(); x < cs.hcap; x += Exec::stride()) cs.htab[x] = 0; /* for the next chunk */
      Exec::sync(); /* the children are in `seen` before the next chunk looks it up */
      const int upto = cutting ? sh->xcut + 1 : chunk;
      n_new_r += take;
      nb_branch_r += take;
      g += (uint64_t)upto;
      cut = cutting;
    }
    RF_SINGLE_BEGIN
      sh->n_new = n_new_r; sh->nb_branch = nb_branch_r;
      if (failed) fail(ST_ERR_CAPACITY);
    RF_SINGLE_END
    if (sh->status != ST_OK) { finish_fold(); return; }
    /* which members were exhausted, where the enumeration stopped, and the reference's counters:
     * an exhausted member met before the stop is walked again in full by the reference (all its
     * children are in `seen`) */
    RF_SINGLE_BEGIN
      if (prm.max_branch > 0) {
        int mc = nm;       /* open member the `break` happened in */
        uint64_t lp = 0;   /* combinations of it that were examined */
        if (cut) { mc = member_of(g - 1, nm); lp = g - cs.mb_pre[mc]; }
        for (int m = 0; m < mc; m++) A.store[A.beam[cs.mb_idx[m]]].done = 1;
        if (cut && lp == cs.mprod[cs.mb_idx[mc]]) A.store[A.beam[cs.mb_idx[mc]]].done = 1;
        const int bcut = cut ? cs.mb_idx[mc] : beam_n;
        uint64_t walked = g;
        for (int b = 0; b < bcut; b++) walked += cs.dprod[b];
        sh->cnt[CNT_COMBOS] += walked; sh->cnt[CNT_N_EVAL] += walked;
        sh->cur_b = cut ? bcut + 1 : beam_n;
      }
    RF_SINGLE_END
    /* P2c: once nb_branch >= max_branch the `break` only leaves the product loop
     *      (rafft.py:202-203): every later member still tries its first combination */
    {
      const int b0 = sh->cur_b;
      const int rest = beam_n - b0;
      if (rest > 0) {
        RF_PAR_BEGIN(x, cs.hcap)
          cs.htab[x] = 0;
        RF_PAR_END
        RF_PAR_BEGIN(j, rest)
          cs.flag[j] = 0;
          if (cs.mprod[b0 + j]) {
            const StoreRec &S = A.store[A.beam[b0 + j]];
            int e_; uint64_t lo_, hi_;
            combo_eval(S, 0, e_, lo_, hi_);
            cs.e[j] = e_; cs.lo[j] = lo_; cs.hi[j] = hi_;
            cs.flag[j] = !seen_has(lo_, hi_);
          }
        RF_PAR_END
        RF_PAR_BEGIN(j, rest)
          if (cs.flag[j]) cs.pref[j] = chunk_insert(j);
        RF_PAR_END
        RF_PAR_BEGIN(j, rest)
          if (cs.flag[j]) cs.flag[j] = cs.htab[cs.pref[j]] == j + 1;
        RF_PAR_END
        const int nnew = Exec::scan_flags(cs.flag, cs.pref, rest, sh->wsum);
        const int base = sh->n_new;
        if (base + nnew > caps.new_cap || 2 * (sh->n_seen + nnew) + 2 > caps.seen_cap) {
          RF_SINGLE_BEGIN fail(ST_ERR_CAPACITY); RF_SINGLE_END
        } else {
          RF_PAR_BEGIN(j, rest)
            if (cs.flag[j]) {
              NewRec &r = A.fresh[base + cs.pref[j]];
              r.e = cs.e[j]; r.parent_slot = b0 + j; r.combo = 0;
              r.fp_lo = cs.lo[j]; r.fp_hi = cs.hi[j];
              seen_put(cs.lo[j], cs.hi[j]);
            }
            if (cs.mprod[b0 + j] == 1) A.store[A.beam[b0 + j]].done = 1;
          RF_PAR_END
          RF_SINGLE_BEGIN
            int tried = 0;
            for (int j = 0; j < rest; j++) tried += (cs.mprod[b0 + j] | cs.dprod[b0 + j]) != 0;
            sh->cnt[CNT_COMBOS] += tried; sh->cnt[CNT_N_EVAL] += tried;
            sh->n_new += nnew;
            sh->nb_branch += nnew;
          RF_SINGLE_END
        }
      }
    }
    if (sh->status != ST_OK) finish_fold();
  }

  /* ---- phase: the next beam */
  RF_FN void phase_select() {
    if (!sh->active || sh->expand_pending) return;
    const int beam_n = sh->beam_n;
    /* ---- P3: new + old, stable sort by energy, keep max_stack (rafft.py:206-210).
     *      Entry i < n_new is a new child (generation order), then the old beam in order;
     *      the order is (energy, i).  Selection: bisect the energy threshold T of the
     *      keep-th entry, take everything below T and the first entries equal to T. */
    const int n_new = sh->n_new;
    const int total = n_new + beam_n;
    const int keep = total < prm.max_stack ? total : prm.max_stack;
    /* the energies once, side by side (the bisection below reads them ~16 times) */
    RF_PAR_BEGIN(i, total)
      A.sel_e[i] = i < n_new ? A.fresh[i].e : A.store[A.beam[i - n_new]].e;
    RF_PAR_END
#define RF_ENTRY_E(i) (A.sel_e[(i)])
    if (total > keep) {
      RF_SINGLE_BEGIN
        sh->emin = 0x7fffffff; sh->emax = (int)0x80000000;
      RF_SINGLE_END
      {
        int lo_ = 0x7fffffff, hi_ = (int)0x80000000, any = 0;
        for (int i = Exec::tid(); i < total; i += Exec::stride()) {
          int e = RF_ENTRY_E(i);
          lo_ = e < lo_ ? e : lo_; hi_ = e > hi_ ? e : hi_; any = 1;
        }
        if (any) { Exec::atomic_min(&sh->emin, lo_); Exec::atomic_max(&sh->emax, hi_); }
      }
      Exec::sync();
      long long blo = sh->emin, bhi = sh->emax; /* smallest T with count(e <= T) >= keep */
      int turn = 0;
      while (blo < bhi) { /* one barrier per bisection step */
        const long long mid = (blo + bhi) >> 1; /* floor */
        int c = 0;
        for (int i = Exec::tid(); i < total; i += Exec::stride()) c += RF_ENTRY_E(i) <= mid;
        if (Exec::block_sum(c, sh->wsum, turn) >= keep) bhi = mid; else blo = mid + 1;
      }
      const int T_e = (int)blo;
      int n_below;
      {
        int c = 0;
        for (int i = Exec::tid(); i < total; i += Exec::stride()) {
          int e = RF_ENTRY_E(i);
          c += e < T_e;
          A.flag[i] = e == T_e;
        }
        n_below = Exec::block_sum(c, sh->wsum, turn);
      }
      const int need_eq = keep - n_below; /* >= 1 */
      Exec::scan_flags(A.flag, A.pref, total, sh->wsum);
      RF_PAR_BEGIN(i, total)
        int e = RF_ENTRY_E(i);
        A.flag[i] = (e < T_e) || (e == T_e && A.pref[i] < need_eq);
      RF_PAR_END
      Exec::scan_flags(A.flag, A.pref, total, sh->wsum);
      RF_PAR_BEGIN(i, total)
        if (A.flag[i]) A.sel_idx[A.pref[i]] = i;
      RF_PAR_END
    } else {
      RF_PAR_BEGIN(i, total)
        A.sel_idx[i] = i;
      RF_PAR_END
    }
    /* rank the selected entries among themselves */
    RF_PAR_BEGIN(x, keep)
      const int i = A.sel_idx[x];
      const int e = RF_ENTRY_E(i);
      int r = 0;
      for (int y = 0; y < keep; y++) {
        const int q = A.sel_idx[y];
        const int eq = RF_ENTRY_E(q);
        r += (eq < e) || (eq == e && q < i);
      }
      A.sel[r] = i;
    RF_PAR_END
#undef RF_ENTRY_E
    /* same dot-bracket list => stop (rafft.py:213-214): new children are never equal to
     * old members (they were unseen), so the lists match iff no new child was selected */
    RF_SINGLE_BEGIN
      int same = keep == beam_n;
      for (int r = 0; r < keep && same; r++) same = A.sel[r] >= n_new;
      sh->tmp0 = same;
      int nfresh = 0;
      for (int r = 0; r < keep; r++) nfresh += A.sel[r] < n_new;
      sh->tmp1 = nfresh;
      if (sh->n_store + nfresh > caps.store_cap) fail(ST_ERR_CAPACITY);
      sh->n_todo = 0;
    RF_SINGLE_END
    if (sh->status != ST_OK || sh->tmp0) finish_fold();
  }

  /* ---- phase: materialise the selected new children (dot-bracket, node list) */
  RF_FN void phase_materialise() {
    if (!sh->active || sh->expand_pending) return;
    const int L = sh->L;
    const int beam_n = sh->beam_n;
    const int n_new = sh->n_new;
    const int total = n_new + beam_n;
    const int keep = total < prm.max_stack ? total : prm.max_stack;
    /* 4a. which child regions (inner / outer of the chosen stems) are not known yet */
    RF_PAR_BEGIN(r, keep)
      int i = A.sel[r];
      if (i < n_new) {
        const NewRec &nr = A.fresh[i];
        const StoreRec &S = A.store[A.beam[nr.parent_slot]];
        uint64_t combo = nr.combo;
        /* digits are produced last node first */
        int cnt_nodes = 0;
        for (int k = S.nl_n - 1; k >= 0; k--) {
          int c = A.ncnt[S.nl_off + k];
          if (c == 0) continue;
          int v = A.nlist[S.nl_off + k];
          const NodeRec &nd = A.node[v];
          int dgt = (int)(combo % (uint64_t)c);
          combo /= (uint64_t)c;
          CandKid &cd = A.ckid[nd.cand_off + dgt];
          for (int side = 0; side < 2; side++) {
            int *idp = side ? &cd.outer_id : &cd.inner_id;
            if (*idp == -2) continue;
            cnt_nodes++;
            if (*idp == -1 && Exec::atomic_cas(idp, -1, -3) == -1) {
              int at = Exec::atomic_add(&sh->n_todo, 1);
              if (at < caps.todo_cap) {
                A.todo[3 * at] = v; A.todo[3 * at + 1] = dgt; A.todo[3 * at + 2] = side;
              }
            }
          }
        }
        A.sel_store[r] = cnt_nodes; /* node count of the child, reused in 4d */
      }
    RF_PAR_END
    if (sh->n_todo > caps.todo_cap) { RF_SINGLE_BEGIN fail(ST_ERR_CAPACITY); RF_SINGLE_END finish_fold(); return; }
    const int n_todo = sh->n_todo;
    /* 4b. build the loop records */
    RF_SINGLE_BEGIN
      sh->tmp2 = sh->n_node;
      if (sh->n_node + n_todo > caps.node_cap) fail(ST_ERR_CAPACITY);
      else sh->n_node += n_todo;
    RF_SINGLE_END
    if (sh->status != ST_OK) { finish_fold(); return; }
    const int node_base = sh->tmp2;
    RF_PAR_BEGIN(a, n_todo)
      const NodeRec &par = A.node[A.todo[3 * a]];
      const Cand &cd = A.cand[par.cand_off + A.todo[3 * a + 1]];
      int side = A.todo[3 * a + 2];
      int mip = cd.mip, mjp = cd.mjp, nb = cd.nb;
      int ci, cj, lo1, hi1, xp, xq, lo2, hi2, clv;
      const bool plain = (cd.geo & RF_GEO_OK) != 0; /* geometry in the stem record: no walk over the loop's segments */
      const int oi = (int)(cd.geo & 0xfffu), oj = (int)((cd.geo >> 12) & 0xfffu), h1 = cd.hh & 255, h2 = cd.hh >> 8;
      if (side == 0) { /* get_inner_loop (utils.py:148-152): local (mip, mjp) exclusive */
        if (plain) { ci = oi; cj = oj; lo1 = h1; hi1 = h2; }
        else { ci = pos_of(par, mip); cj = pos_of(par, mjp); lo1 = seg_of(par, mip); hi1 = seg_of(par, mjp); }
        xp = -1; xq = -1; lo2 = 0; hi2 = 0; clv = mjp - mip - 1;
      } else { /* get_outer_loop (utils.py:141-145): local [0, mip-nb] ++ [mjp+nb, lv) */
        int io = mip - nb + 1, jo = mjp + nb - 1;
        ci = par.ci; cj = par.cj; lo1 = 0; hi2 = par.br_n; clv = io + (par.lv - 1 - jo);
        if (plain) { hi1 = h1; xp = oi - (nb - 1); xq = oj + (nb - 1); lo2 = h2; }
        else { hi1 = seg_of(par, io); xp = pos_of(par, io); xq = pos_of(par, jo); lo2 = seg_of(par, jo); }
      }
      if (!build_node(node_base + a, ci, cj, &par, lo1, hi1, xp, xq, lo2, hi2, clv)) fail(ST_ERR_CAPACITY);
    RF_PAR_END
    if (sh->status != ST_OK) { finish_fold(); return; }
    /* 4c. unify identical loops reached through different stems / orders */
    RF_PAR_BEGIN(a, n_todo)
      int id = nht_resolve(node_base + a);
      if (id < 0) fail(ST_ERR_CAPACITY);
      CandKid &cd = A.ckid[A.node[A.todo[3 * a]].cand_off + A.todo[3 * a + 1]];
      if (A.todo[3 * a + 2]) cd.outer_id = id; else cd.inner_id = id;
    RF_PAR_END
    if (sh->status != ST_OK) { finish_fold(); return; }
    /* 4d. store records + node lists (in_side then out_side per stem, rafft.py:187-190): every
     *     selected child finds its store index and the start of its node list by counting the
     *     new children ranked before it */
    RF_PAR_BEGIN(r, keep)
      int at = sh->n_store, nl = sh->n_nlist;
      for (int q = 0; q < r; q++)
        if (A.sel[q] < n_new) { at++; nl += A.sel_store[q]; }
      A.sel_idx[r] = at; A.pref[r] = nl;
      if (r == keep - 1) {
        if (A.sel[r] < n_new) { at++; nl += A.sel_store[r]; }
        sh->tmp0 = at; sh->tmp1 = nl;
      }
    RF_PAR_END
    RF_SINGLE_BEGIN
      if (keep > 0) {
        if (sh->tmp1 > caps.nlist_cap) fail(ST_ERR_CAPACITY);
        else { sh->n_nlist = sh->tmp1; sh->n_store = sh->tmp0; }
      }
    RF_SINGLE_END
    if (sh->status != ST_OK) { finish_fold(); return; }
    RF_PAR_BEGIN(r, keep)
      const int i = A.sel[r];
      if (i < n_new) {
        const NewRec &nr = A.fresh[i];
        StoreRec &C = A.store[A.sel_idx[r]];
        C.e = nr.e; C.fp_lo = nr.fp_lo; C.fp_hi = nr.fp_hi; C.done = 0; C.prod = 0; C.nact = 0;
        C.nl_off = A.pref[r]; C.nl_n = A.sel_store[r];
        A.sel_store[r] = A.sel_idx[r]; /* selected rank -> store index */
      }
    RF_PAR_END
    if (sh->status != ST_OK) { finish_fold(); return; }
    /* copy the parent's dot-bracket: one warp per row, 16 bytes per lane */
    Exec::sync();
    {
      const int nvec = db_stride() >> 4;
      for (int r = Exec::warp(); r < keep; r += Exec::nwarps()) {
        const int i = A.sel[r];
        if (i < n_new) {
          const Vec16 *src = (const Vec16 *)db_row(A.beam[A.fresh[i].parent_slot]);
          Vec16 *dst = (Vec16 *)db_row(A.sel_store[r]);
          for (int w = Exec::lane(); w < nvec; w += Exec::lanes()) dst[w] = src[w];
        }
      }
    }
    Exec::sync();
    /* stems + node list */
    RF_PAR_BEGIN(r, keep)
      int i = A.sel[r];
      if (i < n_new) {
        const NewRec &nr = A.fresh[i];
        const StoreRec &S = A.store[A.beam[nr.parent_slot]];
        StoreRec &C = A.store[A.sel_store[r]];
        char *db = db_row(A.sel_store[r]);
        uint64_t combo = nr.combo;
        int w = C.nl_off + C.nl_n; /* filled backwards: nodes are visited last first */
        for (int k = S.nl_n - 1; k >= 0; k--) {
          int c = A.ncnt[S.nl_off + k];
          if (c == 0) continue;
          const NodeRec &nd = A.node[A.nlist[S.nl_off + k]];
          int dgt = (int)(combo % (uint64_t)c);
          combo /= (uint64_t)c;
          const Cand &cd = A.cand[nd.cand_off + dgt];
          const CandKid &kid = A.ckid[nd.cand_off + dgt];
          if (kid.outer_id != -2) A.nlist[--w] = kid.outer_id;
          if (kid.inner_id != -2) A.nlist[--w] = kid.inner_id;
          if (cd.geo & RF_GEO_OK) {
            const int oi = (int)(cd.geo & 0xfffu), oj = (int)((cd.geo >> 12) & 0xfffu);
            for (int q = 0; q < cd.nb; q++) { db[oi - q] = '('; db[oj + q] = ')'; }
          } else for (int q = 0; q < cd.nb; q++) {
            db[pos_of(nd, cd.mip - q)] = '(';
            db[pos_of(nd, cd.mjp + q)] = ')';
          }
        }
      }
    RF_PAR_END
    /* the next beam */
    RF_PAR_BEGIN(r, keep)
      const int i = A.sel[r];
      A.sel[r] = i < n_new ? A.sel_store[r] : A.beam[i - n_new];
    RF_PAR_END
    RF_PAR_BEGIN(r, keep)
      A.beam[r] = A.sel[r];
    RF_PAR_END
    RF_SINGLE_BEGIN
      sh->beam_n = keep;
      sh->stage = 0;
    RF_SINGLE_END
  }

  /* Copy the result of this sequence into the global output (bump allocation). */
  RF_FN void emit(OutBuf &out) {
    const int L = sh->L, si = sh->seq_index;
    RF_SINGLE_BEGIN
      int nstruct = 0, nstep = 0, nentry = 0;
      if (sh->status == ST_OK) {
        if (prm.traj) {
          nstruct = sh->n_store; nstep = sh->n_steps;
          for (int s = 0; s < nstep; s++) nentry += A.traj_n[s];
        } else {
          nstruct = sh->beam_n; nstep = 1; nentry = sh->beam_n;
        }
      }
      unsigned long long s0 = Exec::atomic_add64(&out.cursors[0], (unsigned long long)nstruct);
      const unsigned long long Lp = (unsigned long long)db_stride();
      unsigned long long d0 = Exec::atomic_add64(&out.cursors[1], (unsigned long long)nstruct * Lp);
      unsigned long long t0 = Exec::atomic_add64(&out.cursors[2], (unsigned long long)nstep);
      unsigned long long e0 = Exec::atomic_add64(&out.cursors[3], (unsigned long long)nentry);
      if ((int64_t)(s0 + nstruct) > out.struct_cap || (int64_t)(d0 + (uint64_t)nstruct * Lp) > out.db_cap ||
          (int64_t)(t0 + nstep) > out.step_cap || (int64_t)(e0 + nentry) > out.entry_cap) {
        if (sh->status == ST_OK) sh->status = ST_ERR_CAPACITY;
        nstruct = nstep = nentry = 0;
      }
      out.status[si] = sh->status;
      out.seq_struct[si] = (int64_t)s0; out.seq_nstruct[si] = nstruct;
      out.seq_db[si] = (int64_t)d0;
      out.seq_step[si] = (int64_t)t0; out.seq_nstep[si] = nstep;
      out.seq_entry[si] = (int64_t)e0;
      for (int k = 0; k < CNT_N; k++) out.counters[(size_t)si * CNT_N + k] = (int64_t)sh->cnt[k];
      if (sh->status != ST_OK) out.counters[(size_t)si * CNT_N + CNT_T1_NS] = sh->fail_line; /* diagnostics */
      sh->tmp0 = nstruct;
    RF_SINGLE_END
    const int nstruct = sh->tmp0;
    if (nstruct == 0) return;
    const int64_t s0 = out.seq_struct[si], d0 = out.seq_db[si], t0 = out.seq_step[si],
                  e0 = out.seq_entry[si];
    if (prm.traj) {
      RF_PAR_BEGIN(k, nstruct)
        out.energy[s0 + k] = A.store[k].e;
      RF_PAR_END
      Exec::sync();
      for (int k = Exec::warp(); k < nstruct; k += Exec::nwarps()) {
        const Vec16 *src = (const Vec16 *)db_row(k);
        Vec16 *dst = (Vec16 *)(out.db + d0 + (int64_t)k * db_stride());
        for (int w = Exec::lane(); w < (db_stride() >> 4); w += Exec::lanes()) dst[w] = src[w];
      }
      Exec::sync();
      RF_SINGLE_BEGIN
        int64_t e = e0;
        for (int s = 0; s < sh->n_steps; s++) {
          out.step_n[t0 + s] = A.traj_n[s];
          for (int b = 0; b < A.traj_n[s]; b++) out.entries[e++] = A.traj[(size_t)s * prm.max_stack + b];
        }
      RF_SINGLE_END
    } else {
      RF_PAR_BEGIN(k, nstruct)
        out.energy[s0 + k] = A.store[A.beam[k]].e;
        out.entries[e0 + k] = k;
      RF_PAR_END
      Exec::sync();
      for (int k = Exec::warp(); k < nstruct; k += Exec::nwarps()) {
        const Vec16 *src = (const Vec16 *)db_row(A.beam[k]);
        Vec16 *dst = (Vec16 *)(out.db + d0 + (int64_t)k * db_stride());
        for (int w = Exec::lane(); w < (db_stride() >> 4); w += Exec::lanes()) dst[w] = src[w];
      }
      Exec::sync();
      RF_SINGLE_BEGIN
        out.step_n[t0] = nstruct;
      RF_SINGLE_END
    }
  }
};
#undef fail

}  // namespace rafft
