/*
 * expand_fast.cuh -- region expansion for regions of up to 512 unpaired positions (CUDA only):
 * everything create_childs / find_best_consecutives / window_slide compute for one unpaired
 * region (reference rafft/rafft.py:112-123, 86-109, 36-83; rafft/utils.py:115-138), by ONE warp,
 * with the whole working set in shared memory and registers.
 *
 *   1. local index -> original position / branch count / base code      (lanes over positions)
 *   2. one-hot masks F_X, reversed masks R_X, contiguity masks            (warp ballots)
 *   3. correlation of the region against its reverse, one lag per lane: 3 popcounts per 32
 *      positions; exact 32-bit ranking key  floor(acc * 2^18 / norm) << 10 | lag
 *   4. ranking: bitonic network on registers (shuffles across lanes, min/max inside a lane)
 *   5. per ranked lag (one lane each): best consecutive stem, then the energy change of forming
 *      it from per-region prefix sums of the branch terms -- O(stem length)
 *   6. kept stems compacted by ballot, sorted by (dE, rank) on registers, written as Cand records
 *
 * Longer regions, more than 127 branches, -n > 256 or exotic pair weights take the generic
 * kernels k_expand<1..5> (fold_core.cuh), which keep their scratch in global memory.
 */
#pragma once
#include "fold_core.cuh"

namespace rafft {

#define RF_FAST_LMAX 512
#define RF_FAST_BRMAX 127
#define RF_FAST_NRMAX 256
#define RF_FAST_SHIFT 18
#define RF_FAST_LAGBITS 10

/* Launch-wide layout of the fast path (host fills it; lives in the kernel parameter). */
struct FastCfg {
  int32_t enabled;
  int32_t L_cap;   /* regions of up to this many positions (multiple of 32, <= 512) */
  int32_t br_cap;  /* ... and up to this many branches */
  int32_t mws;     /* words per mask incl. one zero word on either side: L_cap/32 + 2 */
  int32_t wgc, wau, wgu; /* pair weights divided by their gcd (small integers) */
  int32_t smem_per_warp;
  int32_t o_mask, o_pos, o_brcnt, o_lcode, o_pre, o_A; /* byte offsets in the warp slice */
  int32_t o_stem, o_ck, o_nb;                                  /* pieces of A after the ranking */
  int32_t o_wst;                                               /* 4 ints of warp state */
  int32_t o_rank; /* rank of every lag (uint16), behind o_nb inside A; dead once the scan order is built */
};

#ifndef __CUDACC_RTC__
inline int32_t fast_pow2(int64_t x) { int64_t p = 32; while (p < x) p <<= 1; return (int32_t)p; }
inline int fast_gcd(int a, int b) { while (b) { int t = a % b; a = b; b = t; } return a; }
/* Returns false when the parameters cannot use the fast path at all. */
inline bool build_fast_cfg(FastCfg &fc, int L_max, const FoldParams &F) {
  memset(&fc, 0, sizeof fc);
  if (F.nb_mode > RF_FAST_NRMAX) return false;
  int g = fast_gcd(fast_gcd(F.wi[1][2], F.wi[0][3]), F.wi[2][3]);
  if (g <= 0) return false;
  fc.wgc = F.wi[1][2] / g; fc.wau = F.wi[0][3] / g; fc.wgu = F.wi[2][3] / g;
  if (fc.wgc > 4 || fc.wau > 4 || fc.wgu > 4) return false; /* key would not fit 32 bits */
  int lc = (L_max + 31) / 32 * 32;
  fc.L_cap = lc > RF_FAST_LMAX ? RF_FAST_LMAX : lc;
  fc.br_cap = RF_FAST_BRMAX;
  fc.mws = fc.L_cap / 32 + 2;
  const int nr = F.nb_mode > 0 ? F.nb_mode : 1;
  const int nk_cap = fast_pow2(2 * (int64_t)fc.L_cap);
  const int ck_cap = fast_pow2(nr);
  size_t off = 0;
  auto carve = [&](size_t bytes) { off = (off + 15) / 16 * 16; size_t o = off; off += bytes; return (int32_t)o; };
  fc.o_mask = carve(4 * 6 * (size_t)fc.mws);
  fc.o_pos = carve(2 * (size_t)fc.L_cap);
  fc.o_brcnt = carve((size_t)fc.L_cap);
  fc.o_lcode = carve((size_t)fc.L_cap);
  fc.o_pre = carve(4 * 2 * (size_t)(fc.br_cap + 1));
  fc.o_wst = carve(16);
  size_t a_keys = 4 * (size_t)nk_cap;
  size_t a_after = 4 * (size_t)nr + 4 * (size_t)ck_cap + 2 * (size_t)nr + 32 + 2 * (size_t)nk_cap;
  fc.o_A = carve(a_keys > a_after ? a_keys : a_after);
  fc.o_stem = fc.o_A;
  fc.o_ck = fc.o_stem + 4 * nr;
  fc.o_ck = (fc.o_ck + 15) / 16 * 16;
  fc.o_nb = fc.o_ck + 4 * ck_cap;
  fc.o_rank = (fc.o_nb + 2 * nr + 15) / 16 * 16;
  fc.smem_per_warp = (int32_t)((off + 127) / 128 * 128);
  fc.enabled = 1;
  return true;
}
#endif

#ifdef __CUDACC__

/* Bitonic sort, descending, of 32*K 32-bit keys held K per lane in BLOCKED layout (key index =
 * K*lane + r): the stages with distance j < K are compare-exchanges between registers of one
 * lane (2 instructions per key), only j >= K needs a shuffle -- half as many shuffle stages as
 * the strided layout. */
template <int K, int J>
__device__ __forceinline__ void fast_inlane(uint32_t (&v)[K], int k, int lane) {
#pragma unroll
  for (int r = 0; r < K; r++) {
    if ((r & J) == 0 && (r | J) < K) {
      const uint32_t a = v[r], b = v[r | J];
      const uint32_t hi = max(a, b), lo = min(a, b);
      const bool desc = ((K * lane + r) & k) == 0;
      v[r] = desc ? hi : lo;
      v[r | J] = desc ? lo : hi;
    }
  }
}

/* keys a[32*r + lane] (any input order will do) -> registers, sorted descending: afterwards
 * rank t is register t % K of lane t / K */
template <int K>
__device__ __forceinline__ void fast_sort_regs(const uint32_t *a, uint32_t (&v)[K], int lane) {
#pragma unroll
  for (int r = 0; r < K; r++) v[r] = a[32 * r + lane];
  for (int k = 2; k <= 32 * K; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j < K) {
        switch (j) {
          case 1: fast_inlane<K, 1>(v, k, lane); break;
          case 2: fast_inlane<K, 2>(v, k, lane); break;
          case 4: fast_inlane<K, 4>(v, k, lane); break;
          case 8: fast_inlane<K, 8>(v, k, lane); break;
          default: fast_inlane<K, 16>(v, k, lane); break;
        }
      } else {
        /* k > j >= K: the direction of the run and the side of the pair depend on the lane only */
        const int jl = j / K;
        const bool keepmax = ((lane & jl) == 0) == (((K * lane) & k) == 0);
#pragma unroll
        for (int r = 0; r < K; r++) {
          const uint32_t o = __shfl_xor_sync(0xffffffffu, v[r], jl);
          v[r] = keepmax ? max(v[r], o) : min(v[r], o);
        }
      }
    }
  }
}

template <int K>
__device__ __noinline__ void fast_sort_desc(uint32_t *a, int lane) {
  uint32_t v[K];
  fast_sort_regs<K>(a, v, lane);
  __syncwarp();
#pragma unroll
  for (int r = 0; r < K; r++) a[K * lane + r] = v[r];
}

__device__ __forceinline__ void fast_sort_dispatch(uint32_t *a, int n /* power of two >= 32 */, int lane) {
  __syncwarp();
  switch (n >> 5) {
    case 1: fast_sort_desc<1>(a, lane); break;
    case 2: fast_sort_desc<2>(a, lane); break;
    case 4: fast_sort_desc<4>(a, lane); break;
    default: fast_sort_desc<8>(a, lane); break;
  }
  __syncwarp();
}

/* bits [32*wi + s, 32*wi + s + 32) of a mask whose words -1 and nw are zero */
__device__ __forceinline__ unsigned fast_shift(const uint32_t *M, int wi, int s) {
  const int q = 32 * wi + s;
  const int a = q >> 5;
  return __funnelshift_r(M[a], M[a + 1], q & 31);
}

struct FastTask {
  const DevEnergy *P;
  const uint8_t *code; /* base codes of the sequence */
  int L;
  const int16_t *br;   /* branches of the region's loop */
  int ci, cj, m, lv, e_loop;
  const uint16_t *pos;
  const uint8_t *brcnt;
  const int32_t *pre_ml, *pre_ext;
};

/* closing pair (i,j) of a multiloop seen from inside */
__device__ __forceinline__ int fast_mlclose(const FastTask &T, int i, int j) {
  const DevEnergy *P = T.P;
  const int tt = pair_type(T.code[j], T.code[i]);
  return P->ml_closing + P->ml_intern + P->mmM[tt * 16 + T.code[j - 1] * 4 + T.code[i + 1]] +
         (tt > 1 ? P->terminal_au : 0);
}

/* Energy change of forming the stem (max_i, max_j, nb) in the region (the eval_one_struct
 * difference of rafft.py:98; same integers as Fold::expand_e4 / loop_energy). */
__device__ __forceinline__ int fast_stem_dE(const FastTask &T, const EnergyView &ev, int max_i, int max_j, int nb) {
  const DevEnergy *P = T.P;
  const int16_t *br = T.br;
  int e = 0;
  int ak = T.pos[max_i], bk = T.pos[max_j];
  { /* loop closed by the innermost pair */
    const int h1 = T.brcnt[max_i], h2 = T.brcnt[max_j];
    const int nbr = h2 - h1;
    if (nbr == 0) e += e_hairpin(ev, ak, bk);
    else if (nbr == 1) e += e_interior(ev, ak, bk, br[2 * h1], br[2 * h1 + 1]);
    else e += fast_mlclose(T, ak, bk) + P->ml_base * (max_j - max_i - 1) + T.pre_ml[h2] - T.pre_ml[h1];
  }
  /* loops between consecutive pairs of the stem */
  for (int k = 1; k < nb; k++) {
    const int an = T.pos[max_i - k], bn = T.pos[max_j + k];
    if (an + 1 == ak && bn - 1 == bk) {
      e += P->stack[pair_type(T.code[an], T.code[bn]) * 6 + pair_type(T.code[bk], T.code[ak])];
    } else {
      const int l0 = T.brcnt[max_i - k], l1 = T.brcnt[max_i - k + 1];
      const int r0 = T.brcnt[max_j + k - 1], r1 = T.brcnt[max_j + k];
      if (l1 == l0 && r1 == r0) e += e_interior(ev, an, bn, ak, bk);
      else
        e += fast_mlclose(T, an, bn) + e_mlstem(ev, ak, bk) + (T.pre_ml[l1] - T.pre_ml[l0]) +
             (T.pre_ml[r1] - T.pre_ml[r0]);
    }
    ak = an; bk = bn;
  }
  /* the region's own loop with the stem as a new branch */
  const int io = max_i - nb + 1, jo = max_j + nb - 1;
  const int h1 = T.brcnt[io], h2 = T.brcnt[jo];
  if (T.ci < 0) {
    e += T.pre_ext[h1] + e_extstem(ev, ak, bk) + (T.pre_ext[T.m] - T.pre_ext[h2]);
  } else if (h1 + (T.m - h2) == 0) {
    e += e_interior(ev, T.ci, T.cj, ak, bk);
  } else {
    e += fast_mlclose(T, T.ci, T.cj) + P->ml_base * (io + (T.lv - 1 - jo)) + e_mlstem(ev, ak, bk) +
         T.pre_ml[h1] + (T.pre_ml[T.m] - T.pre_ml[h2]);
  }
  return e - T.e_loop;
}

/* One region, one warp; `ws` = the warp's slice of shared memory.  The expansion is cut into
 * phases that the kernel separates by block barriers: all warps of an SM then run the same few
 * KB of code (the whole expansion is ~70 KB of SASS, far beyond the instruction caches; a
 * free-running version spent most of its time in no_instruction stalls,
 * profiles/r01_ncu_v4a_fast_unsynced.md). */
#define RF_FAST_BIND                                                                            \
  NodeRec &n = ((NodeRec *)(arena + aoff.node))[v];                                             \
  const uint8_t *code = (const uint8_t *)(arena + aoff.code);                                   \
  const int16_t *br = (const int16_t *)(arena + aoff.pool16) + n.pool_off;                      \
  const int ci = n.ci, cj = n.cj, m = n.br_n, lv = n.lv;                                        \
  const int MWS = fc.mws;                                                                       \
  uint32_t *mask = (uint32_t *)(ws + fc.o_mask) + 1; /* word -1 is the zero pad */              \
  uint32_t *FH = mask, *FL = mask + MWS, *RH = mask + 2 * MWS, *RL = mask + 3 * MWS,            \
           *CL = mask + 4 * MWS, *CLR = mask + 5 * MWS;                                         \
  uint16_t *pos = (uint16_t *)(ws + fc.o_pos);                                                  \
  uint8_t *brcnt = (uint8_t *)(ws + fc.o_brcnt);                                                \
  uint8_t *lcode = (uint8_t *)(ws + fc.o_lcode);                                                \
  int32_t *pre_ml = (int32_t *)(ws + fc.o_pre), *pre_ext = pre_ml + (fc.br_cap + 1);            \
  uint32_t *keys = (uint32_t *)(ws + fc.o_A);                                                   \
  uint32_t *r_stem = (uint32_t *)(ws + fc.o_stem);                                              \
  uint32_t *ck = (uint32_t *)(ws + fc.o_ck);                                                    \
  uint16_t *r_nb = (uint16_t *)(ws + fc.o_nb);                                                  \
  int32_t *wst = (int32_t *)(ws + fc.o_wst); /* keep_n, n_pos, w_scan */                        \
  const int nlag = 2 * lv - 1;                                                                  \
  const int nr = nlag < prm.nb_mode ? nlag : prm.nb_mode;                                       \
  const unsigned FULL = 0xffffffffu;                                                            \
  (void)code; (void)br; (void)ci; (void)cj; (void)m; (void)FH; (void)FL; (void)RH; (void)RL; (void)CL; (void)CLR;     \
  (void)pos; (void)brcnt; (void)lcode; (void)pre_ml; (void)pre_ext; (void)keys;     \
  (void)r_stem; (void)ck; (void)r_nb; (void)wst; (void)nr; (void)FULL;

#define RF_FAST_ARGS                                                                            \
  const DevEnergy *E, const FoldParams &prm, const Caps &caps, const FastCfg &fc, char *arena,  \
      const ArenaOff &aoff, SeqState *st, int v, char *ws, int lane

/* phase 1: local index -> original position, branches before it, base code + contiguity flag;
 *          prefix sums of the branch terms (multiloop view; exterior view for the exterior loop) */
__device__ __forceinline__ void fast_p1(RF_FAST_ARGS) {
  RF_FAST_BIND
  const int16_t *so = br + 2 * m;
  const int L = st->L;
  for (int t = lane; t < lv; t += 32) {
    int lo = 0, hi = m;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (so[mid] <= t) lo = mid; else hi = mid - 1;
    }
    const int first = so[lo];
    const int p = (lo == 0 ? ci + 1 : br[2 * (lo - 1) + 1] + 1) + (t - first);
    pos[t] = (uint16_t)p;
    brcnt[t] = (uint8_t)lo;
    lcode[t] = (uint8_t)(code[p] | (t > first ? 4 : 0));
  }
  EnergyView ev; ev.P = E; ev.s = code; ev.L = L;
  int carry_ml = 0, carry_ext = 0;
  if (lane == 0) { pre_ml[0] = 0; pre_ext[0] = 0; }
  for (int b0 = 0; b0 < m; b0 += 32) {
    const int b = b0 + lane;
    int tm = 0, te = 0;
    if (b < m) {
      const int p = br[2 * b], q = br[2 * b + 1];
      const int t = pair_type(code[p], code[q]);
      const int au = t > 1 ? E->terminal_au : 0;
      const int s5 = code[p > 0 ? p - 1 : 0], s3 = code[q < L - 1 ? q + 1 : L - 1];
      tm = E->ml_intern + E->mmM[t * 16 + s5 * 4 + s3] + au;
      if (ci < 0) te = e_extstem(ev, p, q);
    }
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int a = __shfl_up_sync(FULL, tm, d), b2 = __shfl_up_sync(FULL, te, d);
      if (lane >= d) { tm += a; te += b2; }
    }
    if (b < m) { pre_ml[b + 1] = carry_ml + tm; pre_ext[b + 1] = carry_ext + te; }
    carry_ml += __shfl_sync(FULL, tm, 31);
    carry_ext += __shfl_sync(FULL, te, 31);
  }
}

/* phase 2: bit-plane masks by ballot.  Base codes A,C,G,U = 0..3: FH/FL bit t = high/low bit of
 *          x_t, RH/RL bit u = the same of x_{lv-1-u}; CL bit t = position t continues t-1 in
 *          ORIGINAL coordinates, CLR its reversal.  With x = FH^RH', y = FL^RL' (R' = R shifted
 *          to the lag):  G.C pairs = x&y&(FH^FL), A.U = x&y&~(FH^FL), G.U = ~x&y&FH. */
__device__ __forceinline__ void fast_p2(RF_FAST_ARGS) {
  RF_FAST_BIND
  const int nw = (lv + 31) >> 5;
  for (int w = 0; w < nw; w++) {
    const int t = 32 * w + lane;
    const int lc = t < lv ? lcode[t] : 0, lr = t < lv ? lcode[lv - 1 - t] : 0;
    const unsigned fh = __ballot_sync(FULL, lc & 2), fl = __ballot_sync(FULL, lc & 1),
                   fcl = __ballot_sync(FULL, lc & 4);
    const unsigned rh = __ballot_sync(FULL, lr & 2), rl = __ballot_sync(FULL, lr & 1),
                   rcl = __ballot_sync(FULL, lr & 4);
    if (lane == 0) {
      FH[w] = fh; FL[w] = fl; CL[w] = fcl; RH[w] = rh; RL[w] = rl; CLR[w] = rcl;
    }
  }
  if (lane < 6) { mask[lane * MWS - 1] = 0; mask[lane * MWS + nw] = 0; }
}

/* phase 3: correlation (utils.py:115-132): acc[p] = sum over unordered pairs i+j=p of w(x_i,x_j)
 *    = cor[p]/2 up to the common weight factor (only the first half of the anti-diagonal is
 *    visited); c[p] = cor[p]/(min(p, 2lv-2-p)+1).  Two different quotients with denominators
 *    <= 512 differ by >= 2^-18, so floor(acc * 2^18 / norm) orders (and equates) them exactly as
 *    the reference's floats. */
__device__ __forceinline__ void fast_p3(RF_FAST_ARGS, int nk) {
  RF_FAST_BIND
  const int wGC = fc.wgc, wAU = fc.wau, wGU = fc.wgu;
  for (int p = lane; p < nk; p += 32) {
    uint32_t key = 0;
    if (p < nlag) {
      const int s = lv - 1 - p;
      const int i0 = p < lv ? 0 : p - lv + 1;
      const int d = p < lv ? p + 1 : 2 * lv - 1 - p;
      const int iend = i0 + ((d + 1) >> 1); /* exclusive */
      int acc = 0;
      for (int wi = i0 >> 5; wi <= ((iend - 1) >> 5); wi++) {
        const unsigned fh = FH[wi], fl = FL[wi];
        const unsigned x = fh ^ fast_shift(RH, wi, s), y = fl ^ fast_shift(RL, wi, s);
        unsigned range = FULL;
        if (32 * wi < i0) range &= FULL << (i0 - 32 * wi);
        if (32 * wi + 32 > iend) range &= FULL >> (32 * wi + 32 - iend);
        const unsigned xy = x & y & range, t = fh ^ fl;
        acc += wGC * __popc(xy & t) + wAU * __popc(xy & ~t) + wGU * __popc(~x & y & fh & range);
      }
      key = ((((uint32_t)acc << RF_FAST_SHIFT) / (uint32_t)d) << RF_FAST_LAGBITS) | (uint32_t)p;
    }
    keys[p] = key;
  }
}

/* phase 4: rank: (value desc, lag desc) = the reference's order (rafft.py:118,92) */
template <int K>
__device__ __forceinline__ void fast_p4(RF_FAST_ARGS) {
  RF_FAST_BIND
  uint32_t vreg[K];
  fast_sort_regs<K>(keys, vreg, lane);
  __syncwarp(); /* rank_of lies inside the key buffer */
  /* rank t = register t % K of lane t / K; the padding keys (0) sort behind every real lag */
  uint16_t *rank_of = (uint16_t *)(ws + fc.o_rank);
#pragma unroll
  for (int q = 0; q < K; q++) {
    const int t = K * lane + q;
    if (t < nlag) rank_of[vreg[q] & ((1u << RF_FAST_LAGBITS) - 1)] = (uint16_t)t;
  }
}

/* phase 5: per ranked lag the best consecutive stem (window_slide, rafft.py:36-83) over the set
 *          bits of the pairable mask */
__device__ __forceinline__ void fast_p5(RF_FAST_ARGS) {
  RF_FAST_BIND
  const double dGC = prm.w[1][2], dAU = prm.w[0][3], dGU = prm.w[2][3];
  const unsigned mGC = dGC == 0.0 ? 0u : FULL, mAU = dAU == 0.0 ? 0u : FULL, mGU = dGU == 0.0 ? 0u : FULL;
  /* through a shuffle so that they live in registers: ptxas otherwise re-loads them from the
   * constant bank inside the bit loop */
  const int hGC = __shfl_sync(FULL, __double2hiint(dGC), 0), hAU = __shfl_sync(FULL, __double2hiint(dAU), 0),
            hGU = __shfl_sync(FULL, __double2hiint(dGU), 0);
  /* Scan order: the ranked lags by decreasing length of their anti-diagonal (p = lv-1, lv-2, lv,
   * lv-3, lv+1, ...), compacted by ballot.  Every lane scans one lag per round; lanes of a
   * round then carry nearly the same work and the incomplete last round holds the shortest
   * lags.  Results stay indexed by rank, so the order is invisible outside. */
  uint32_t *perm = ck; /* (lag << 16) | rank; ck is not used before phase 6 */
  {
    const uint16_t *rank_of = (const uint16_t *)(ws + fc.o_rank);
    int n_perm = 0;
    for (int k0 = 0; k0 < nlag && n_perm < nr; k0 += 32) {
      const int k = k0 + lane;
      const int p = (k & 1) ? lv - 1 - ((k + 1) >> 1) : lv - 1 + (k >> 1);
      const int rk = k < nlag ? rank_of[p] : 0xffff;
      const unsigned bal = __ballot_sync(FULL, rk < nr);
      if (rk < nr) perm[n_perm + __popc(bal & ((1u << lane) - 1))] = ((uint32_t)p << 16) | (uint32_t)rk;
      n_perm += __popc(bal);
    }
    __syncwarp();
  }
  int w_scan = 0;
  for (int x = lane; x < nr; x += 32) {
    const uint32_t pr = perm[x];
    const int p = (int)(pr >> 16), r = (int)(pr & 0xffffu);
    const int s = lv - 1 - p;
    const int i0 = p < lv ? 0 : p - lv + 1;
    const int d = p < lv ? p + 1 : 2 * lv - 1 - p;
    const int len2 = (d + 1) >> 1;
    int iend = i0 + len2; /* exclusive */
    /* hairpin test (rafft.py:78): pos[j] - pos[i] decreases along the scan, so the positions
     * that fail it are the innermost ones; they can never become the stem and are clipped */
    while (iend > i0 && !((int)pos[p - (iend - 1)] - (int)pos[iend - 1] > prm.min_hp)) iend--;
    int max_i = 0, max_j = 0;
    double max_score = 0.0, prev = 0.0;
    unsigned carry = 0; /* was the position just before this word pairable (and in range)? */
    unsigned best_bit = 0;
    /* run length so far in the low half, word index in the high half: one register to keep */
    int run_w = (i0 >> 5) << 16, best_rw = 0;
    for (int wi = i0 >> 5; wi <= ((iend - 1) >> 5) && iend > i0; wi++, run_w += 0x10000) {
      const unsigned fh = FH[wi], fl = FL[wi];
      const unsigned x = fh ^ fast_shift(RH, wi, s), y = fl ^ fast_shift(RL, wi, s);
      unsigned range = FULL;
      if (32 * wi < i0) range &= FULL << (i0 - 32 * wi);
      if (32 * wi + 32 > iend) range &= FULL >> (32 * wi + 32 - iend);
      const unsigned xy = x & y & range, t = fh ^ fl;
      const unsigned gc = xy & t & mGC, au = xy & ~t & mAU, gu = ~x & y & fh & range & mGU;
      unsigned Pm = gc | au | gu;
      const unsigned ADJ = (Pm << 1) | carry; /* the previous position is pairable too */
      carry = Pm >> 31;
      /* both arms continue in original coordinates (rafft.py:67), t > 0 */
      unsigned IS = CL[wi] & fast_shift(CLR, wi, s - 1);
      if ((i0 >> 5) == wi) IS &= ~(1u << (i0 & 31));
      const int wbase = wi << 16;
      while (Pm) {
        const unsigned bit = Pm & (0u - Pm);
        Pm ^= bit;
        /* the weights are multiples of 1/16 up to 16: the low words of the doubles are zero */
        const double w = __hiloint2double((gc & bit) ? hGC : (au & bit) ? hAU : hGU, 0);
        const bool adjacent = (ADJ & bit) != 0;
        const double chained = __dmul_rn(__dadd_rn(adjacent ? prev : 0.0, w), w);
        const double tot = (IS & bit) ? chained : w;
        run_w = adjacent ? run_w + 1 : wbase + 1;
        if (tot >= max_score) { max_score = tot; best_rw = run_w; best_bit = bit; }
        prev = tot;
      }
    }
    const int max_nb = best_rw & 0xffff;
    if (best_bit) { max_i = 32 * (best_rw >> 16) + __ffs((int)best_bit) - 1; max_j = p - max_i; }
    w_scan += len2;
    r_stem[r] = ((uint32_t)max_i << 16) | (uint32_t)max_j;
    r_nb[r] = (uint16_t)max_nb;
  }
  w_scan = __reduce_add_sync(FULL, w_scan);
  if (lane == 0) wst[2] = w_scan;
}

/* phase 6: energy change of every stem; stems below the threshold get a (dE, rank) key: stable
 *          order by dE (rafft.py:108); kept keys are compacted by ballot */
__device__ __forceinline__ void fast_p6(RF_FAST_ARGS) {
  RF_FAST_BIND
  EnergyView ev; ev.P = E; ev.s = code; ev.L = st->L;
  FastTask T;
  T.P = E; T.code = code; T.L = ev.L; T.br = br; T.ci = ci; T.cj = cj; T.m = m; T.lv = lv; T.e_loop = n.e_loop;
  T.pos = pos; T.brcnt = brcnt; T.pre_ml = pre_ml; T.pre_ext = pre_ext;
  int keep_n = 0, n_pos = 0;
  for (int r0 = 0; r0 < nr; r0 += 32) {
    const int r = r0 + lane;
    int max_nb = 0, dE = 0x7fffffff;
    if (r < nr) {
      max_nb = r_nb[r];
      if (max_nb > 0) {
        const uint32_t ij = r_stem[r];
        n_pos++;
        dE = fast_stem_dE(T, ev, (int)(ij >> 16), (int)(ij & 0xffff), max_nb);
      }
    }
    const bool kept = max_nb > 0 && dE < prm.cand_thr;
    const unsigned bal = __ballot_sync(FULL, kept);
    if (kept) ck[keep_n + __popc(bal & ((1u << lane) - 1))] = ~((((uint32_t)(dE + 0x800000)) << 8) | (uint32_t)r);
    keep_n += __popc(bal);
  }
  n_pos = __reduce_add_sync(FULL, n_pos);
  int nkc = 32;
  while (nkc < keep_n) nkc <<= 1;
  __syncwarp();
  for (int x = keep_n + lane; x < nkc; x += 32) ck[x] = 0;
  if (lane == 0) { wst[0] = keep_n; wst[1] = n_pos; }
}

/* phase 7: sort the kept stems, allocate and write their records */
__device__ __forceinline__ void fast_p7(RF_FAST_ARGS) {
  RF_FAST_BIND
  const int keep_n = wst[0], n_pos = wst[1], w_scan = wst[2];
  int off = 0;
  if (keep_n > 0) {
    int nkc = 32;
    while (nkc < keep_n) nkc <<= 1;
    fast_sort_dispatch(ck, nkc, lane);
    if (lane == 0) {
      off = atomicAdd(&st->n_cand, keep_n);
      if (off + keep_n > caps.cand_cap) { st->status = ST_ERR_CAPACITY; st->fail_line = -__LINE__; off = -1; }
    }
    off = __shfl_sync(FULL, off, 0);
    if (off >= 0) {
      Cand *cand = (Cand *)(arena + aoff.cand) + off;
      CandKid *ckid = (CandKid *)(arena + aoff.ckid) + off;
      for (int at = lane; at < keep_n; at += 32) {
        const uint32_t key = ~ck[at];
        const int r = (int)(key & 255u);
        const uint32_t ij = r_stem[r];
        const int mip = ij >> 16, mjp = ij & 0xffff, nb = r_nb[r];
        Cand &c = cand[at];
        c.dE = (int32_t)(key >> 8) - 0x800000;
        ckid[at].inner_id = (mjp - mip > 1) ? -1 : -2;                       /* rafft.py:130 */
        ckid[at].outer_id = (mip - (nb - 1) > 0 || mjp + nb < lv) ? -1 : -2; /* rafft.py:141 */
        cand_geometry(c, pos, brcnt, mip, mjp, nb);
      }
    }
  }
  if (lane == 0) {
    n.n_ranked = nr; n.n_pos = n_pos; n.w_scan = w_scan;
    n.cand_off = off < 0 ? 0 : off;
    n.cand_n = off < 0 ? 0 : keep_n;
    atomicAdd(&st->cnt[CNT_UNIQ_NODES], 1ULL);
    atomicAdd(&st->cnt[CNT_UNIQ_W_CORR], (unsigned long long)((long long)lv * (lv + 1) / 2));
    atomicAdd(&st->cnt[CNT_UNIQ_W_SCAN], (unsigned long long)w_scan);
  }
}

#endif /* __CUDACC__ */

}  // namespace rafft
