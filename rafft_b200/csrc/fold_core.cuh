/*
 * fold_core.cuh -- the RAFFT folding hot path as ONE cooperative thread block per sequence.
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
 * The code is written as uniform control flow + `par_for` / `single` phases (each phase is
 * bracketed by block barriers).  tests/emu compiles this same file for the host with
 * blockDim = 1 to check the logic without a GPU; that build is test-only.
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
#define RF_NOINLINE
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
};

/* Launch-wide constants. */
struct FoldParams {
  int32_t nb_mode, max_stack, max_branch, min_hp, traj;
  double min_nrj;
  double w[4][4];   /* pair weights as the reference's float64 (prep_sequence CENCODING) */
  int32_t wi[4][4]; /* same * 16, exact integers (host checks) */
  int32_t cand_thr; /* keep stems with dE < cand_thr (dcal); exact test per structure later */
};

/* Capacities of one per-block arena (host sizes them; overflow -> status, host regrows). */
struct Caps {
  int32_t L_max;
  int32_t node_cap, pool16_cap, cand_cap, store_cap, nlist_cap;
  int32_t seen_cap, nht_cap; /* powers of two */
  int32_t new_cap;           /* >= max_branch + max_stack + 1 */
  int32_t step_cap;          /* folding steps */
  int32_t todo_cap;
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
};

struct Cand {
  int32_t dE;
  uint16_t mip, mjp, nb, rank;
  int32_t inner_id, outer_id; /* -1 unresolved, -2 no such node, -3 being created */
  uint64_t fp_lo, fp_hi;      /* sum of pair fingerprints of the stem */
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
  int64_t *seq_db;      /* per sequence: first db byte */
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
       CNT_UNIQ_W_CORR, CNT_UNIQ_W_SCAN, CNT_COMBOS, CNT_NODE_POS, CNT_T0_NS, CNT_T1_NS, CNT_N };

enum { ST_OK = 0, ST_ERR_SEQ = 2, ST_ERR_CAPACITY = 4 };

/* ------------------------------------------------------------------------------------ */
/* Execution policy: CUDA thread block, or a single host thread (tests/emu only).        */
#ifdef __CUDACC__
struct Exec {
  RF_FN static int tid() { return threadIdx.x; }
  RF_FN static int nthreads() { return blockDim.x; }
  RF_FN static void sync() { __syncthreads(); }
  RF_FN static int atomic_add(int *p, int v) { return atomicAdd(p, v); }
  RF_FN static unsigned long long atomic_add64(unsigned long long *p, unsigned long long v) {
    return atomicAdd(p, v);
  }
  RF_FN static int atomic_cas(int *p, int cmp, int v) { return atomicCAS(p, cmp, v); }
  RF_FN static unsigned long long atomic_cas64(unsigned long long *p, unsigned long long cmp,
                                                unsigned long long v) {
    return atomicCAS(p, cmp, v);
  }
  RF_FN static double dadd(double a, double b) { return __dadd_rn(a, b); }
  RF_FN static double dmul(double a, double b) { return __dmul_rn(a, b); }
  RF_FN static double ddiv(double a, double b) { return __ddiv_rn(a, b); }
  RF_FN static unsigned long long dbits(double a) { return (unsigned long long)__double_as_longlong(a); }
  RF_FN static long long now_ns() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return (long long)t; }
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
#else
struct Exec {
  static int &order() { static int o = 0; return o; } /* 0 forward, 1 reverse iteration */
  static int tid() { return 0; }
  static int &emu_threads() { static int t = 32; return t; } /* chunk width of the emulation */
  static int nthreads() { return emu_threads(); }
  static void sync() {}
  static int atomic_add(int *p, int v) { int o = *p; *p += v; return o; }
  static unsigned long long atomic_add64(unsigned long long *p, unsigned long long v) {
    unsigned long long o = *p; *p += v; return o;
  }
  static int atomic_cas(int *p, int cmp, int v) { int o = *p; if (o == cmp) *p = v; return o; }
  static unsigned long long atomic_cas64(unsigned long long *p, unsigned long long cmp,
                                         unsigned long long v) {
    unsigned long long o = *p; if (o == cmp) *p = v; return o;
  }
  static double dadd(double a, double b) { volatile double r = a + b; return r; }
  static double dmul(double a, double b) { volatile double r = a * b; return r; }
  static double ddiv(double a, double b) { volatile double r = a / b; return r; }
  static unsigned long long dbits(double a) { unsigned long long u; memcpy(&u, &a, 8); return u; }
  static long long now_ns() { return 0; }
};
/* Iterations of a phase must be independent: the emulation runs them forward or reversed. */
#define RF_PAR_BEGIN(i, n)                                   \
  for (int i##_k = 0, i##_n = (n); i##_k < i##_n; i##_k++) { \
    int i = Exec::order() ? i##_n - 1 - i##_k : i##_k;
#define RF_PAR_END }
#define RF_SINGLE_BEGIN {
#define RF_SINGLE_END }
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

RF_FN int e_hairpin(const EnergyView &ev, int i, int j) {
  const DevEnergy *P = ev.P;
  int u = j - i - 1;
  int e = e_loglen(ev, P->hairpin, u);
  if (u < 3) return e;
  int t = pair_type(ev.s[i], ev.s[j]);
  if (u == 3 || u == 4 || u == 6) {
    uint32_t key = 0;
    for (int k = i; k <= j; k++) key = (key << 2) | ev.s[k];
    if (u == 4) {
      for (int k = 0; k < P->n_tetra; k++)
        if (P->tetra_key[k] == key) return P->tetra_e[k];
    } else if (u == 6) {
      for (int k = 0; k < P->n_hexa; k++)
        if (P->hexa_key[k] == key) return P->hexa_e[k];
    } else {
      for (int k = 0; k < P->n_tri; k++)
        if (P->tri_key[k] == key) return P->tri_e[k];
      return e + (t > 1 ? P->terminal_au : 0);
    }
  }
  return e + P->mmH[t * 16 + ev.s[i + 1] * 4 + ev.s[j - 1]];
}

RF_FN int e_interior(const EnergyView &ev, int i, int j, int p, int q) {
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
RF_FN int loop_energy(const EnergyView &ev, int ci, int cj, const int16_t *br, int lo1, int hi1,
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
/* Per-block working memory, carved from one arena.                                      */
struct Arena {
  uint8_t *code;       /* L_max */
  NodeRec *node;       /* node_cap */
  int16_t *pool16;     /* pool16_cap: branches + segment offsets of nodes */
  Cand *cand;          /* cand_cap */
  StoreRec *store;     /* store_cap */
  char *store_db;      /* store_cap * L */
  int32_t *nlist;      /* nlist_cap: node lists of stored structures */
  uint64_t *seen_lo, *seen_hi; /* seen_cap */
  unsigned long long *nht;     /* nht_cap: (hash32 << 32) | (node id + 1), 0 = empty */
  NewRec *fresh;       /* new_cap */
  uint64_t *keys;      /* new_cap + max_stack: sort keys / lag keys (>= 2*L_max) */
  int32_t *sel;        /* max_stack: selected entries of the next beam */
  int32_t *beam;       /* max_stack: store indices */
  int32_t *traj_n;     /* step_cap */
  int32_t *traj;       /* step_cap * max_stack: store indices */
  int32_t *todo;       /* todo_cap * 3 */
  int32_t *todo_id;    /* todo_cap */
  /* node expansion scratch */
  uint16_t *pos;       /* L_max */
  uint16_t *brcnt;     /* L_max + 1 */
  uint8_t *ncode;      /* L_max */
  int32_t *ranked;     /* nb_mode */
  int32_t *tmp_dE;     /* nb_mode */
  uint32_t *tmp_stem;  /* nb_mode * 2: (mip, mjp) and nb */
  uint64_t *tmp_fp;    /* nb_mode * 2 */
  int32_t *act_v, *act_c; /* L_max: nodes of the current structure that have candidates */
  int32_t *flag, *pref, *ch_e; /* per-chunk: unseen flag, exclusive prefix, energy (blockDim) */
  uint64_t *ch_lo, *ch_hi;     /* per-chunk fingerprints */
};

struct BlockShared {
  int32_t seq_index, L, status;
  int32_t n_node, n_pool16, n_cand, n_store, n_nlist, n_seen, n_todo, n_expand;
  int32_t beam_n, n_new, nb_branch, n_steps, n_traj;
  int32_t tmp0, tmp1, tmp2;
  int64_t cnt[CNT_N];
};

RF_HD size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

/* Carve the arena; with base == nullptr just computes the size.  Returns bytes used. */
RF_HD size_t arena_carve(Arena &A, char *base, const Caps &c, const FoldParams &fp, int nthreads) {
  size_t off = 0;
#define RF_CARVE(field, type, count)                              \
  off = align_up(off, 16);                                        \
  A.field = (type *)(base ? base + off : nullptr);                \
  off += sizeof(type) * (size_t)(count);
  int L = c.L_max;
  int nkeys = c.new_cap + fp.max_stack;
  if (nkeys < 2 * L) nkeys = 2 * L;
  int nrank = fp.nb_mode > 0 ? fp.nb_mode : 1;
  RF_CARVE(code, uint8_t, L);
  RF_CARVE(node, NodeRec, c.node_cap);
  RF_CARVE(pool16, int16_t, c.pool16_cap);
  RF_CARVE(cand, Cand, c.cand_cap);
  RF_CARVE(store, StoreRec, c.store_cap);
  RF_CARVE(store_db, char, (size_t)c.store_cap * L);
  RF_CARVE(nlist, int32_t, c.nlist_cap);
  RF_CARVE(seen_lo, uint64_t, c.seen_cap);
  RF_CARVE(seen_hi, uint64_t, c.seen_cap);
  RF_CARVE(nht, unsigned long long, c.nht_cap);
  RF_CARVE(fresh, NewRec, c.new_cap);
  RF_CARVE(keys, uint64_t, nkeys);
  RF_CARVE(sel, int32_t, fp.max_stack);
  RF_CARVE(beam, int32_t, fp.max_stack);
  RF_CARVE(traj_n, int32_t, c.step_cap);
  RF_CARVE(traj, int32_t, (size_t)c.step_cap * fp.max_stack);
  RF_CARVE(todo, int32_t, (size_t)c.todo_cap * 3);
  RF_CARVE(todo_id, int32_t, c.todo_cap);
  RF_CARVE(pos, uint16_t, L);
  RF_CARVE(brcnt, uint16_t, L + 1);
  RF_CARVE(ncode, uint8_t, L);
  RF_CARVE(ranked, int32_t, nrank);
  RF_CARVE(tmp_dE, int32_t, nrank);
  RF_CARVE(tmp_stem, uint32_t, (size_t)nrank * 2);
  RF_CARVE(tmp_fp, uint64_t, (size_t)nrank * 2);
  RF_CARVE(act_v, int32_t, L);
  RF_CARVE(act_c, int32_t, L);
  RF_CARVE(flag, int32_t, nthreads + 1);
  RF_CARVE(pref, int32_t, nthreads + 1);
  RF_CARVE(ch_e, int32_t, nthreads + 1);
  RF_CARVE(ch_lo, uint64_t, nthreads + 1);
  RF_CARVE(ch_hi, uint64_t, nthreads + 1);
#undef RF_CARVE
  return align_up(off, 256);
}

/* ------------------------------------------------------------------------------------ */
struct Fold {
  const DevEnergy *E;
  FoldParams prm;
  Caps caps;
  Arena A;
  BlockShared *sh;
  EnergyView ev;

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
    while (lo < hi) {
      int mid = (lo + hi) >> 1;
      if (passes(e_cur, A.cand[n.cand_off + mid].dE)) lo = mid + 1; else hi = mid;
    }
    return lo;
  }

  RF_FN void fail(int code) {
    /* called inside single phases or by any thread: status only ever goes 0 -> code */
    sh->status = code;
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
    const NodeRec &n = A.node[id];
    unsigned long long tag = ((n.hash >> 32) << 32) | (unsigned)(id + 1);
    unsigned mask = caps.nht_cap - 1;
    unsigned slot = (unsigned)n.hash & mask;
    for (int probe = 0; probe < caps.nht_cap; probe++, slot = (slot + 1) & mask) {
      unsigned long long cur = A.nht[slot];
      if (cur == 0) {
        cur = Exec::atomic_cas64(&A.nht[slot], 0ULL, tag);
        if (cur == 0) return id;
      }
      if ((cur >> 32) == (tag >> 32)) {
        int other = (int)(cur & 0xffffffffu) - 1;
        if (other == id) return id;
        if (same_loop(A.node[other], n)) return other;
      }
    }
    return -1; /* table full */
  }

  /* --- node expansion: correlation -> top-n lags -> stem scan -> energy change ------------
   * Everything create_childs/find_best_consecutives compute for one unpaired region
   * (rafft.py:112-123, 86-109), done once per unique loop. */
  RF_NOINLINE void expand_node(int v) {
    NodeRec &n = A.node[v];
    const int lv = n.lv, m = n.br_n;
    const int nlag = 2 * lv - 1;
    /* 1. local index -> original position / branch count / base code */
    RF_PAR_BEGIN(t, lv)
      int k = seg_of(n, t);
      int p = seg_start(n, k) + (t - node_segoff(n)[k]);
      A.pos[t] = (uint16_t)p;
      A.brcnt[t] = (uint16_t)k;
      A.ncode[t] = A.code[p];
    RF_PAR_END
    /* 2. pair-weighted correlation of the region against its reverse (utils.py:115-132):
     *    cor[p] = sum_{i+j=p} w(s_i,s_j) = 2 * (half anti-diagonal), exact integers;
     *    c[p] = cor[p] / (min(p, 2lv-2-p) + 1.0) as the reference's float64.
     *    key = order-preserving bits of c[p] with the lag in the low bits, so that sorting
     *    keys descending is the reference's (value desc, lag desc) order (rafft.py:118,92). */
    RF_PAR_BEGIN(p, nlag)
      int i0 = p < lv ? 0 : p - lv + 1;
      int d = p < lv ? p + 1 : 2 * lv - 1 - p;
      int half = d >> 1;
      int acc = 0;
      for (int t = 0; t < half; t++) acc += prm.wi[A.ncode[i0 + t]][A.ncode[p - i0 - t]];
      double cor = Exec::ddiv((double)(2 * (int64_t)acc), 16.0);
      double norm = (double)(p < lv ? p : 2 * lv - 2 - p) + 1.0;
      double c = Exec::ddiv(cor, norm);
      A.keys[p] = (Exec::dbits(c) & ~0x1FFFULL) | (uint64_t)p;
    RF_PAR_END
    /* 3. rank by counting; the first nb_mode lags in order */
    const int nr = nlag < prm.nb_mode ? nlag : prm.nb_mode;
    RF_PAR_BEGIN(p, nlag)
      uint64_t key = A.keys[p];
      int r = 0;
      for (int q = 0; q < nlag; q++) r += A.keys[q] > key;
      if (r < nr) A.ranked[r] = p;
    RF_PAR_END
    /* 4. per ranked lag: longest-scoring consecutive stem (window_slide, rafft.py:36-83)
     *    and the energy change of forming it (eval_one_struct difference, rafft.py:98). */
    RF_PAR_BEGIN(r, nr)
      const int p = A.ranked[r];
      const int i0 = p < lv ? 0 : p - lv + 1;
      const int d = p < lv ? p + 1 : 2 * lv - 1 - p;
      const int len2 = (d >> 1) + (d & 1);
      int max_nb = 0, tmp_max = 0, max_i = 0, max_j = 0;
      double max_score = 0.0, prev = 0.0;
      for (int t = 0; t < len2; t++) {
        int ip = i0 + t, jp = p - i0 - t;
        double tot = prm.w[A.ncode[ip]][A.ncode[jp]];
        if (t > 0 && A.pos[ip] - A.pos[ip - 1] == 1 && A.pos[jp + 1] - A.pos[jp] == 1)
          tot = Exec::dmul(Exec::dadd(prev, tot), tot);
        if (tot == 0.0) tmp_max = 0; else tmp_max++;
        if (tot >= max_score && (int)A.pos[jp] - (int)A.pos[ip] > prm.min_hp) {
          max_score = tot; max_nb = tmp_max; max_i = ip; max_j = jp;
        }
        prev = tot;
      }
      int dE = 0x7fffffff;
      uint64_t flo = 0, fhi = 0;
      if (max_nb > 0) {
        const int16_t *br = node_br(n);
        int e = 0;
        /* loop closed by the innermost pair: the region's branches between its two arms */
        int a0 = A.pos[max_i], b0 = A.pos[max_j];
        e += loop_energy(ev, a0, b0, br, A.brcnt[max_i] , A.brcnt[max_j], -1, -1, 0, 0,
                         max_j - max_i - 1);
        /* loops between consecutive pairs of the stem (stack, or a junction if the stem
         * steps over existing helices: contiguity breaks do not end a stem, rafft.py:71-74) */
        int ak = a0, bk = b0;
        for (int k = 0; k < max_nb; k++) {
          uint64_t l1, h1;
          pair_fp(ak, bk, l1, h1);
          flo += l1; fhi += h1;
          if (k + 1 < max_nb) {
            int an = A.pos[max_i - k - 1], bn = A.pos[max_j + k + 1];
            e += loop_energy(ev, an, bn, br, A.brcnt[max_i - k - 1], A.brcnt[max_i - k], ak, bk,
                             A.brcnt[max_j + k], A.brcnt[max_j + k + 1], 0);
            ak = an; bk = bn;
          }
        }
        /* the region's own loop with the stem as a new branch */
        int io = max_i - max_nb + 1, jo = max_j + max_nb - 1;
        e += loop_energy(ev, n.ci, n.cj, br, 0, A.brcnt[io], ak, bk, A.brcnt[jo], m,
                         io + (lv - 1 - jo));
        dE = e - n.e_loop;
      }
      A.tmp_dE[r] = dE;
      A.tmp_stem[2 * r] = ((uint32_t)max_i << 16) | (uint32_t)max_j;
      A.tmp_stem[2 * r + 1] = (uint32_t)max_nb;
      A.tmp_fp[2 * r] = flo;
      A.tmp_fp[2 * r + 1] = fhi;
    RF_PAR_END
    /* 5. keep stems with dE below the threshold, stably sorted by dE (rafft.py:108) */
    RF_SINGLE_BEGIN
      int keep = 0, npos = 0;
      int64_t ws = 0;
      for (int r = 0; r < nr; r++) {
        int p = A.ranked[r];
        int d = p < lv ? p + 1 : 2 * lv - 1 - p;
        ws += (d >> 1) + (d & 1);
        if (A.tmp_stem[2 * r + 1] > 0) npos++;
        if (A.tmp_stem[2 * r + 1] > 0 && A.tmp_dE[r] < prm.cand_thr) keep++;
      }
      n.n_ranked = nr; n.n_pos = npos; n.w_scan = ws;
      int off = sh->n_cand;
      if (off + keep > caps.cand_cap) { fail(ST_ERR_CAPACITY); keep = 0; }
      sh->n_cand = off + keep;
      n.cand_off = off;
      sh->tmp0 = keep;
      sh->cnt[CNT_UNIQ_NODES] += 1;
      sh->cnt[CNT_UNIQ_W_CORR] += (int64_t)lv * (lv + 1) / 2;
      sh->cnt[CNT_UNIQ_W_SCAN] += ws;
    RF_SINGLE_END
    const int keep = sh->tmp0;
    RF_PAR_BEGIN(r, nr)
      if (keep > 0 && A.tmp_stem[2 * r + 1] > 0 && A.tmp_dE[r] < prm.cand_thr) {
        int dE = A.tmp_dE[r], at = 0;
        for (int q = 0; q < nr; q++)
          if (A.tmp_stem[2 * q + 1] > 0 && A.tmp_dE[q] < prm.cand_thr &&
              (A.tmp_dE[q] < dE || (A.tmp_dE[q] == dE && q < r)))
            at++;
        Cand &c = A.cand[n.cand_off + at];
        uint32_t ij = A.tmp_stem[2 * r];
        int mip = ij >> 16, mjp = ij & 0xffff, nb = (int)A.tmp_stem[2 * r + 1];
        c.dE = dE; c.mip = (uint16_t)mip; c.mjp = (uint16_t)mjp; c.nb = (uint16_t)nb;
        c.rank = (uint16_t)r;
        c.inner_id = (mjp - mip > 1) ? -1 : -2;                       /* rafft.py:130 */
        c.outer_id = (mip - (nb - 1) > 0 || mjp + nb < lv) ? -1 : -2; /* rafft.py:141 */
        c.fp_lo = A.tmp_fp[2 * r];
        c.fp_hi = A.tmp_fp[2 * r + 1];
      }
    RF_PAR_END
    RF_SINGLE_BEGIN
      n.cand_n = keep;
    RF_SINGLE_END
  }

  /* --- seen set (global per sequence, rafft.py:196-200) ------------------------------- */
  RF_FN bool seen_has(uint64_t lo, uint64_t hi) const {
    unsigned mask = caps.seen_cap - 1;
    unsigned slot = (unsigned)(lo ^ (hi >> 7)) & mask;
    for (;;) {
      uint64_t l = A.seen_lo[slot];
      if (l == 0) return false;
      if (l == lo && A.seen_hi[slot] == hi) return true;
      slot = (slot + 1) & mask;
    }
  }
  RF_FN void seen_put(uint64_t lo, uint64_t hi) {
    unsigned mask = caps.seen_cap - 1;
    unsigned slot = (unsigned)(lo ^ (hi >> 7)) & mask;
    for (;;) {
      if (A.seen_lo[slot] == 0 &&
          Exec::atomic_cas64((unsigned long long *)&A.seen_lo[slot], 0ULL, lo) == 0ULL) {
        A.seen_hi[slot] = hi;
        return;
      }
      slot = (slot + 1) & mask;
    }
  }

  /* decode combination index -> (energy, fingerprint) of the child of store entry S */
  RF_FN void combo_eval(const StoreRec &S, int m, uint64_t combo, int &e, uint64_t &lo,
                        uint64_t &hi) const {
    e = S.e; lo = S.fp_lo; hi = S.fp_hi;
    for (int k = m - 1; k >= 0; k--) { /* itertools.product: last list varies fastest */
      uint32_t c = (uint32_t)A.act_c[k];
      uint32_t dgt = (uint32_t)(combo % c);
      combo /= c;
      const Cand &cd = A.cand[A.node[A.act_v[k]].cand_off + dgt];
      e += cd.dE; lo += cd.fp_lo; hi += cd.fp_hi;
    }
    if (lo == 0) lo = 1; /* 0 marks an empty slot of the seen table */
  }

  /* ------------------------------------------------------------------------------------ */
  RF_NOINLINE void run(const char *seq, int L, int seq_index, OutBuf &out) {
    /* ---- reset per-sequence state */
    RF_SINGLE_BEGIN
      sh->seq_index = seq_index; sh->L = L; sh->status = ST_OK;
      sh->n_node = 0; sh->n_pool16 = 0; sh->n_cand = 0; sh->n_store = 0; sh->n_nlist = 0;
      sh->n_seen = 0; sh->n_todo = 0; sh->n_expand = 0; sh->beam_n = 0; sh->n_new = 0;
      sh->nb_branch = 0; sh->n_steps = 0; sh->n_traj = 0;
      for (int k = 0; k < CNT_N; k++) sh->cnt[k] = 0;
      sh->cnt[CNT_T0_NS] = Exec::now_ns();
      if (L > caps.L_max) sh->status = ST_ERR_CAPACITY;
      else if (L < 1) sh->status = ST_ERR_SEQ; /* the reference cannot fold an empty sequence */
    RF_SINGLE_END
    if (sh->status == ST_OK) {
      RF_PAR_BEGIN(i, L)
        char ch = seq[i];
        int c = ch == 'A' ? 0 : ch == 'C' ? 1 : ch == 'G' ? 2 : ch == 'U' ? 3 : -1;
        if (c < 0) { sh->status = ST_ERR_SEQ; c = 0; }
        A.code[i] = (uint8_t)c;
      RF_PAR_END
    }
    if (sh->status == ST_OK) {
      RF_PAR_BEGIN(i, caps.seen_cap)
        A.seen_lo[i] = 0;
      RF_PAR_END
      RF_PAR_BEGIN(i, caps.nht_cap)
        A.nht[i] = 0;
      RF_PAR_END
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
      RF_PAR_BEGIN(i, L)
        A.store_db[i] = '.';
      RF_PAR_END
      while (sh->status == ST_OK) {
        if (!step()) break;
      }
    }
    emit(out);
  }

  /* One bfs_pairs call (rafft.py:156-216).  Returns false when the beam has converged. */
  RF_NOINLINE bool step() {
    const int L = sh->L;
    const int beam_n = sh->beam_n;
    /* glob_traj += [glob_tree] (rafft.py:161) */
    RF_SINGLE_BEGIN
      if (sh->n_steps >= caps.step_cap) fail(ST_ERR_CAPACITY);
      else {
        A.traj_n[sh->n_steps] = beam_n;
        for (int b = 0; b < beam_n; b++) A.traj[(size_t)sh->n_steps * prm.max_stack + b] = A.beam[b];
        sh->n_steps++;
        sh->cnt[CNT_STEPS]++;
      }
      sh->n_expand = 0;
    RF_SINGLE_END
    if (sh->status != ST_OK) return false;

    /* ---- P1: expand the loops of the beam that have not been expanded yet */
    for (int b = 0; b < beam_n; b++) {
      const StoreRec &S = A.store[A.beam[b]];
      if (S.done) continue;
      RF_PAR_BEGIN(k, S.nl_n)
        int v = A.nlist[S.nl_off + k];
        if (A.node[v].cand_n == -1 && Exec::atomic_cas(&A.node[v].cand_n, -1, -2) == -1) {
          int at = Exec::atomic_add(&sh->n_expand, 1);
          if (at < caps.todo_cap) A.todo_id[at] = v;
        }
      RF_PAR_END
    }
    if (sh->n_expand > caps.todo_cap) { RF_SINGLE_BEGIN fail(ST_ERR_CAPACITY); RF_SINGLE_END return false; }
    {
      const int ne = sh->n_expand;
      /* deterministic order (candidate pool layout) regardless of thread interleaving */
      RF_SINGLE_BEGIN
        for (int a = 1; a < ne; a++) {
          int x = A.todo_id[a], c = a - 1;
          while (c >= 0 && A.todo_id[c] > x) { A.todo_id[c + 1] = A.todo_id[c]; c--; }
          A.todo_id[c + 1] = x;
        }
      RF_SINGLE_END
      for (int a = 0; a < ne && sh->status == ST_OK; a++) expand_node(A.todo_id[a]);
    }
    if (sh->status != ST_OK) return false;

    /* ---- P2: combine stems of independent regions (rafft.py:176-203) */
    RF_SINGLE_BEGIN
      sh->n_new = 0; sh->nb_branch = 0;
    RF_SINGLE_END
    const int T = Exec::nthreads();
    for (int b = 0; b < beam_n && sh->status == ST_OK; b++) {
      StoreRec &S = A.store[A.beam[b]];
      /* reference work counters: create_childs runs for every node of every beam member */
      RF_SINGLE_BEGIN
        int m = 0;
        uint64_t prod = 1;
        for (int k = 0; k < S.nl_n; k++) {
          const NodeRec &nd = A.node[A.nlist[S.nl_off + k]];
          sh->cnt[CNT_NODE_EXP] += 1;
          sh->cnt[CNT_NODE_POS] += nd.lv;
          sh->cnt[CNT_W_CORR] += (int64_t)nd.lv * (nd.lv + 1) / 2;
          sh->cnt[CNT_W_SCAN] += nd.w_scan;
          sh->cnt[CNT_N_EVAL] += nd.n_pos;
          if (!S.done) {
            int c = count_pass(nd, S.e);
            if (c > 0) {
              A.act_v[m] = A.nlist[S.nl_off + k];
              A.act_c[m] = c;
              m++;
              prod = prod > (1ULL << 62) / (uint64_t)c ? (1ULL << 62) : prod * (uint64_t)c;
            }
          }
        }
        if (!S.done) {
          S.nact = m;
          S.prod = m ? prod : 0;
          if (m == 0) S.done = 1;
        }
        /* a member whose product was exhausted earlier: the reference walks it again
         * (every child is already in `seen`), we only account for it */
        if (S.done && S.prod) {
          uint64_t walked = sh->nb_branch >= prm.max_branch ? 1 : S.prod;
          sh->cnt[CNT_COMBOS] += walked;
          sh->cnt[CNT_N_EVAL] += walked;
        }
      RF_SINGLE_END
      if (S.done) continue;
      const int m = (int)S.nact;
      const uint64_t limit = sh->nb_branch >= prm.max_branch ? 1 : S.prod;
      uint64_t processed = 0;
      bool cut = false;
      while (processed < limit && !cut && sh->status == ST_OK) {
        const int chunk = (int)((limit - processed) < (uint64_t)T ? (limit - processed) : (uint64_t)T);
        /* evaluate + look up (rafft.py:192-196); children of one structure are distinct */
        RF_PAR_BEGIN(i, chunk)
          int e_; uint64_t lo_, hi_;
          combo_eval(S, m, processed + (uint64_t)i, e_, lo_, hi_);
          A.ch_e[i] = e_; A.ch_lo[i] = lo_; A.ch_hi[i] = hi_;
          A.flag[i] = !seen_has(lo_, hi_);
        RF_PAR_END
        RF_SINGLE_BEGIN
          /* exclusive prefix over the chunk + position of the max_branch cut: the reference
           * breaks right after the combination that brings nb_branch to max_branch */
          int run = 0, xcut = -1;
          for (int i = 0; i < chunk && xcut < 0; i++) {
            A.pref[i] = run;
            run += A.flag[i];
            if (sh->nb_branch + run >= prm.max_branch) xcut = i;
          }
          sh->tmp0 = xcut;
          sh->tmp1 = run;               /* new children kept from this chunk */
          sh->tmp2 = sh->n_new;         /* where they go */
          if (sh->n_new + run > caps.new_cap) { fail(ST_ERR_CAPACITY); sh->tmp1 = 0; }
        RF_SINGLE_END
        const int xcut = sh->tmp0, base = sh->tmp2;
        const int upto = xcut >= 0 ? xcut + 1 : chunk;
        if (sh->status == ST_OK) {
          RF_PAR_BEGIN(i, upto)
            if (A.flag[i]) {
              NewRec &r = A.fresh[base + A.pref[i]];
              r.e = A.ch_e[i]; r.parent_slot = b; r.combo = processed + (uint64_t)i;
              r.fp_lo = A.ch_lo[i]; r.fp_hi = A.ch_hi[i];
              seen_put(A.ch_lo[i], A.ch_hi[i]);
            }
          RF_PAR_END
          RF_SINGLE_BEGIN
            sh->n_new += sh->tmp1;
            sh->nb_branch += sh->tmp1;
            sh->n_seen += sh->tmp1;
            sh->cnt[CNT_COMBOS] += upto;
            sh->cnt[CNT_N_EVAL] += upto;
            if (2 * (sh->n_seen + prm.max_branch + prm.max_stack + T) > caps.seen_cap)
              fail(ST_ERR_CAPACITY);
          RF_SINGLE_END
        }
        processed += (uint64_t)upto;
        cut = xcut >= 0;
      }
      if (processed == S.prod) {
        RF_SINGLE_BEGIN
          S.done = 1;
        RF_SINGLE_END
      }
    }
    if (sh->status != ST_OK) return false;

    /* ---- P3: new + old, stable sort by energy, keep max_stack (rafft.py:206-210) */
    const int n_new = sh->n_new;
    const int total = n_new + beam_n;
    RF_PAR_BEGIN(i, total)
      int e = i < n_new ? A.fresh[i].e : A.store[A.beam[i - n_new]].e;
      A.keys[i] = ((uint64_t)(uint32_t)(e ^ 0x80000000)) << 32 | (uint32_t)i;
    RF_PAR_END
    const int keep = total < prm.max_stack ? total : prm.max_stack;
    RF_PAR_BEGIN(i, total)
      uint64_t key = A.keys[i];
      int r = 0;
      for (int q = 0; q < total; q++) r += A.keys[q] < key;
      if (r < keep) A.sel[r] = i;
    RF_PAR_END
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
    if (sh->status != ST_OK || sh->tmp0) return false;

    /* ---- P4: materialise the selected new children (dot-bracket, node list) */
    const int store_base = sh->n_store;
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
          int v = A.nlist[S.nl_off + k];
          const NodeRec &nd = A.node[v];
          int c = count_pass(nd, S.e);
          if (c == 0) continue;
          int dgt = (int)(combo % (uint64_t)c);
          combo /= (uint64_t)c;
          Cand &cd = A.cand[nd.cand_off + dgt];
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
        A.fresh[i].parent_slot |= cnt_nodes << 16; /* node count of the child, reused in 4d */
      }
    RF_PAR_END
    if (sh->n_todo > caps.todo_cap) { RF_SINGLE_BEGIN fail(ST_ERR_CAPACITY); RF_SINGLE_END return false; }
    const int n_todo = sh->n_todo;
    /* 4b. build the loop records */
    RF_SINGLE_BEGIN
      sh->tmp2 = sh->n_node;
      if (sh->n_node + n_todo > caps.node_cap) fail(ST_ERR_CAPACITY);
      else sh->n_node += n_todo;
    RF_SINGLE_END
    if (sh->status != ST_OK) return false;
    const int node_base = sh->tmp2;
    RF_PAR_BEGIN(a, n_todo)
      const NodeRec &par = A.node[A.todo[3 * a]];
      const Cand &cd = A.cand[par.cand_off + A.todo[3 * a + 1]];
      int side = A.todo[3 * a + 2];
      int mip = cd.mip, mjp = cd.mjp, nb = cd.nb;
      bool ok;
      if (side == 0) { /* get_inner_loop (utils.py:148-152): local (mip, mjp) exclusive */
        ok = build_node(node_base + a, pos_of(par, mip), pos_of(par, mjp), &par, seg_of(par, mip),
                        seg_of(par, mjp), -1, -1, 0, 0, mjp - mip - 1);
      } else { /* get_outer_loop (utils.py:141-145): local [0, mip-nb] ++ [mjp+nb, lv) */
        int io = mip - nb + 1, jo = mjp + nb - 1;
        ok = build_node(node_base + a, par.ci, par.cj, &par, 0, seg_of(par, io), pos_of(par, io),
                        pos_of(par, jo), seg_of(par, jo), par.br_n, io + (par.lv - 1 - jo));
      }
      if (!ok) fail(ST_ERR_CAPACITY);
    RF_PAR_END
    if (sh->status != ST_OK) return false;
    /* 4c. unify identical loops reached through different stems / orders */
    RF_PAR_BEGIN(a, n_todo)
      int id = nht_resolve(node_base + a);
      if (id < 0) fail(ST_ERR_CAPACITY);
      Cand &cd = A.cand[A.node[A.todo[3 * a]].cand_off + A.todo[3 * a + 1]];
      if (A.todo[3 * a + 2]) cd.outer_id = id; else cd.inner_id = id;
    RF_PAR_END
    if (sh->status != ST_OK) return false;
    /* 4d. store records + node lists (in_side then out_side per stem, rafft.py:187-190) */
    RF_SINGLE_BEGIN
      int at = sh->n_store, nl = sh->n_nlist;
      for (int r = 0; r < keep; r++) {
        int i = A.sel[r];
        if (i >= n_new) continue;
        NewRec &nr = A.fresh[i];
        int cnt_nodes = nr.parent_slot >> 16;
        nr.parent_slot &= 0xffff;
        StoreRec &C = A.store[at];
        C.e = nr.e; C.fp_lo = nr.fp_lo; C.fp_hi = nr.fp_hi; C.done = 0; C.prod = 0; C.nact = 0;
        C.nl_off = nl; C.nl_n = cnt_nodes;
        nl += cnt_nodes;
        A.keys[r] = (uint64_t)at; /* selected rank -> store index */
        at++;
      }
      if (nl > caps.nlist_cap) fail(ST_ERR_CAPACITY);
      else { sh->n_nlist = nl; sh->n_store = at; }
    RF_SINGLE_END
    if (sh->status != ST_OK) return false;
    /* copy the parent's dot-bracket */
    RF_PAR_BEGIN(x, keep * L)
      int r = x / L, pos = x - r * L;
      int i = A.sel[r];
      if (i < n_new) {
        int parent = A.beam[A.fresh[i].parent_slot];
        A.store_db[(size_t)A.keys[r] * L + pos] = A.store_db[(size_t)parent * L + pos];
      }
    RF_PAR_END
    /* stems + node list */
    RF_PAR_BEGIN(r, keep)
      int i = A.sel[r];
      if (i < n_new) {
        const NewRec &nr = A.fresh[i];
        const StoreRec &S = A.store[A.beam[nr.parent_slot]];
        StoreRec &C = A.store[(int)A.keys[r]];
        char *db = A.store_db + (size_t)A.keys[r] * L;
        uint64_t combo = nr.combo;
        int w = C.nl_off + C.nl_n; /* filled backwards: nodes are visited last first */
        for (int k = S.nl_n - 1; k >= 0; k--) {
          const NodeRec &nd = A.node[A.nlist[S.nl_off + k]];
          int c = count_pass(nd, S.e);
          if (c == 0) continue;
          int dgt = (int)(combo % (uint64_t)c);
          combo /= (uint64_t)c;
          const Cand &cd = A.cand[nd.cand_off + dgt];
          if (cd.outer_id != -2) A.nlist[--w] = cd.outer_id;
          if (cd.inner_id != -2) A.nlist[--w] = cd.inner_id;
          for (int q = 0; q < cd.nb; q++) {
            db[pos_of(nd, cd.mip - q)] = '(';
            db[pos_of(nd, cd.mjp + q)] = ')';
          }
        }
      }
    RF_PAR_END
    /* the next beam */
    RF_SINGLE_BEGIN
      for (int r = 0; r < keep; r++) {
        int i = A.sel[r];
        A.sel[r] = i < n_new ? (int)A.keys[r] : A.beam[i - n_new];
      }
      for (int r = 0; r < keep; r++) A.beam[r] = A.sel[r];
      sh->beam_n = keep;
    RF_SINGLE_END
    return true;
  }

  /* Copy the result of this sequence into the global output (bump allocation). */
  RF_NOINLINE void emit(OutBuf &out) {
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
      unsigned long long d0 = Exec::atomic_add64(&out.cursors[1], (unsigned long long)nstruct * L);
      unsigned long long t0 = Exec::atomic_add64(&out.cursors[2], (unsigned long long)nstep);
      unsigned long long e0 = Exec::atomic_add64(&out.cursors[3], (unsigned long long)nentry);
      if ((int64_t)(s0 + nstruct) > out.struct_cap || (int64_t)(d0 + (uint64_t)nstruct * L) > out.db_cap ||
          (int64_t)(t0 + nstep) > out.step_cap || (int64_t)(e0 + nentry) > out.entry_cap) {
        if (sh->status == ST_OK) sh->status = ST_ERR_CAPACITY;
        nstruct = nstep = nentry = 0;
      }
      out.status[si] = sh->status;
      out.seq_struct[si] = (int64_t)s0; out.seq_nstruct[si] = nstruct;
      out.seq_db[si] = (int64_t)d0;
      out.seq_step[si] = (int64_t)t0; out.seq_nstep[si] = nstep;
      out.seq_entry[si] = (int64_t)e0;
      sh->cnt[CNT_T1_NS] = Exec::now_ns();
      for (int k = 0; k < CNT_N; k++) out.counters[(size_t)si * CNT_N + k] = sh->cnt[k];
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
      RF_PAR_BEGIN(x, nstruct * L)
        out.db[d0 + x] = A.store_db[x];
      RF_PAR_END
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
      RF_PAR_BEGIN(x, nstruct * L)
        int k = x / L;
        out.db[d0 + x] = A.store_db[(size_t)A.beam[k] * L + (x - k * L)];
      RF_PAR_END
      RF_SINGLE_BEGIN
        out.step_n[t0] = nstruct;
      RF_SINGLE_END
    }
  }
};

}  // namespace rafft
