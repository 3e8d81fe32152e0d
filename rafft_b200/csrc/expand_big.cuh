/*
 * expand_big.cuh -- region expansion for regions of 513 .. 4096 unpaired positions (CUDA only): the
 * same work as expand_fast.cuh (create_childs / find_best_consecutives / window_slide of the
 * reference, rafft/rafft.py:112-123, 86-109, 36-83; rafft/utils.py:115-138), by ONE THREAD BLOCK per
 * region with the whole working set in shared memory.  BASELINE config 4 (3,000-nt inputs) lives
 * here: its regions are too long for the warp-per-region path and used to take the generic
 * kernels, whose scratch is in global memory.
 *
 *   1. local index -> original position / branch count / base code      (threads over positions)
 *   2. bit-plane masks FH/FL/RH/RL and contiguity masks CL/CLR            (one ballot word per warp)
 *   3. correlation, one lag per thread: 3 popcounts per 32 positions; exact 64-bit ranking key
 *      floor(acc * 2^24 / norm) << 13 | lag  (denominators <= 4096: different quotients differ by
 *      >= 2^-24, so the floor orders and equates them exactly as the reference's floats)
 *   4. the nb_mode best lags by MSB-first radix selection over the keys (7 passes of 8 bits, block
 *      histogram in shared memory), ranked among themselves by counting
 *   5. per ranked lag (one thread each) the best consecutive stem        (same bit walk as the fast path)
 *   6. energy change of every stem from the per-region prefix sums        (fast_stem_dE)
 *   7. kept stems sorted by (dE, rank) on registers by one warp, written as Cand records
 *
 * Regions with more than 127 branches or -n > 256 still take the generic kernels.
 */
#pragma once
#include "expand_fast.cuh"

namespace rafft {

#define RF_BIG_LMAX 4096
#define RF_BIG_THREADS 256
#define RF_BIG_SHIFT 24
#define RF_BIG_LAGBITS 13

struct BigCfg {
  int32_t enabled;
  int32_t L_cap;   /* regions of up to this many positions (multiple of 32, <= 4096) */
  int32_t br_cap;
  int32_t mws;     /* words per mask incl. one zero word on either side */
  int32_t wgc, wau, wgu;
  int32_t smem_bytes;
  int32_t o_mask, o_pos, o_brcnt, o_lcode, o_pre, o_keys, o_hist, o_sel, o_stem, o_nb, o_ck, o_st;
};

#ifndef __CUDACC_RTC__
/* Returns false when the parameters cannot use the block-per-region path at all. */
inline bool build_big_cfg(BigCfg &bc, int L_max, const FoldParams &F) {
  memset(&bc, 0, sizeof bc);
  if (F.nb_mode > RF_FAST_NRMAX || F.nb_mode < 1) return false;
  int g = fast_gcd(fast_gcd(F.wi[1][2], F.wi[0][3]), F.wi[2][3]);
  if (g <= 0) return false;
  bc.wgc = F.wi[1][2] / g; bc.wau = F.wi[0][3] / g; bc.wgu = F.wi[2][3] / g;
  if (bc.wgc > 4 || bc.wau > 4 || bc.wgu > 4) return false;
  int lc = (L_max + 31) / 32 * 32;
  bc.L_cap = lc > RF_BIG_LMAX ? RF_BIG_LMAX : lc;
  bc.br_cap = RF_FAST_BRMAX;
  bc.mws = bc.L_cap / 32 + 2;
  const int nr = F.nb_mode;
  size_t off = 0;
  auto carve = [&](size_t bytes) { off = (off + 15) / 16 * 16; size_t o = off; off += bytes; return (int32_t)o; };
  bc.o_mask = carve(4 * 6 * (size_t)bc.mws);
  bc.o_pos = carve(2 * (size_t)bc.L_cap);
  bc.o_brcnt = carve((size_t)bc.L_cap);
  bc.o_lcode = carve((size_t)bc.L_cap);
  bc.o_pre = carve(4 * 2 * (size_t)(bc.br_cap + 1));
  bc.o_keys = carve(8 * 2 * (size_t)bc.L_cap);   /* one 64-bit key per lag (2 L - 1 of them) */
  bc.o_hist = carve(4 * 256);
  bc.o_sel = carve(8 * (size_t)RF_FAST_NRMAX);
  bc.o_stem = carve(4 * (size_t)nr);
  bc.o_nb = carve(2 * (size_t)nr);
  bc.o_ck = carve(4 * (size_t)fast_pow2(nr));
  bc.o_st = carve(64);
  bc.smem_bytes = (int32_t)((off + 127) / 128 * 128);
  bc.enabled = 1;
  return true;
}
#endif

#ifdef __CUDACC__

/* One region, one block of RF_BIG_THREADS threads; `ws` = the block's dynamic shared memory. */
__device__ __forceinline__ void big_expand(const DevEnergy *E, const FoldParams &prm, const Caps &caps,
                                           const BigCfg &bc, char *arena, const ArenaOff &aoff, SeqState *st,
                                           int v, char *ws) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NT = blockDim.x, NW = NT >> 5;
  const unsigned FULL = 0xffffffffu;
  NodeRec &n = ((NodeRec *)(arena + aoff.node))[v];
  const uint8_t *code = (const uint8_t *)(arena + aoff.code);
  const int16_t *br = (const int16_t *)(arena + aoff.pool16) + n.pool_off;
  const int ci = n.ci, cj = n.cj, m = n.br_n, lv = n.lv;
  const int MWS = bc.mws;
  uint32_t *mask = (uint32_t *)(ws + bc.o_mask) + 1; /* word -1 is the zero pad */
  uint32_t *FH = mask, *FL = mask + MWS, *RH = mask + 2 * MWS, *RL = mask + 3 * MWS, *CL = mask + 4 * MWS,
           *CLR = mask + 5 * MWS;
  uint16_t *pos = (uint16_t *)(ws + bc.o_pos);
  uint8_t *brcnt = (uint8_t *)(ws + bc.o_brcnt);
  uint8_t *lcode = (uint8_t *)(ws + bc.o_lcode);
  int32_t *pre_ml = (int32_t *)(ws + bc.o_pre), *pre_ext = pre_ml + (bc.br_cap + 1);
  unsigned long long *keys = (unsigned long long *)(ws + bc.o_keys);
  int32_t *hist = (int32_t *)(ws + bc.o_hist);
  unsigned long long *sel = (unsigned long long *)(ws + bc.o_sel);
  uint32_t *r_stem = (uint32_t *)(ws + bc.o_stem);
  uint16_t *r_nb = (uint16_t *)(ws + bc.o_nb);
  uint32_t *ck = (uint32_t *)(ws + bc.o_ck);
  int32_t *bst = (int32_t *)(ws + bc.o_st); /* 0 keep_n, 1 n_pos, 2 w_scan, 3 n_sel, 4 digit, 5 remaining, 6 cand offset */
  const int nlag = 2 * lv - 1;
  const int nr = nlag < prm.nb_mode ? nlag : prm.nb_mode;
  const int L = st->L;
  EnergyView ev; ev.P = E; ev.s = code; ev.L = L;

  /* ---- 1. positions, branch counts, base codes + contiguity flag; prefix sums of the branch terms */
  {
    const int16_t *so = br + 2 * m;
    for (int t = tid; t < lv; t += NT) {
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
    if (tid < 8) bst[tid] = 0;
    if (warp == 0) {
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
  }
  __syncthreads();
  /* ---- 2. bit planes by ballot (see expand_fast.cuh phase 2), one 32-position word per warp at a time */
  const int nw = (lv + 31) >> 5;
  for (int w = warp; w < nw; w += NW) {
    const int t = 32 * w + lane;
    const int lc = t < lv ? lcode[t] : 0, lr = t < lv ? lcode[lv - 1 - t] : 0;
    const unsigned fh = __ballot_sync(FULL, lc & 2), fl = __ballot_sync(FULL, lc & 1), fcl = __ballot_sync(FULL, lc & 4);
    const unsigned rh = __ballot_sync(FULL, lr & 2), rl = __ballot_sync(FULL, lr & 1), rcl = __ballot_sync(FULL, lr & 4);
    if (lane == 0) { FH[w] = fh; FL[w] = fl; CL[w] = fcl; RH[w] = rh; RL[w] = rl; CLR[w] = rcl; }
  }
  if (tid < 6) { mask[tid * MWS - 1] = 0; mask[tid * MWS + nw] = 0; }
  __syncthreads();
  /* ---- 3. correlation (utils.py:115-132), one lag per thread: acc[p] = sum over the unordered pairs
   *         i + j = p of the gcd-reduced weight (first half of the anti-diagonal), exact 64-bit key */
  {
    const int wGC = bc.wgc, wAU = bc.wau, wGU = bc.wgu;
    for (int p = tid; p < nlag; p += NT) {
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
      keys[p] = ((((unsigned long long)(unsigned)acc << RF_BIG_SHIFT) / (unsigned long long)(unsigned)d) << RF_BIG_LAGBITS) |
                (unsigned long long)(unsigned)p;
    }
  }
  __syncthreads();
  /* ---- 4. rank: (value desc, lag desc) = the reference's order (rafft.py:118,92).  nlag > nr here
   *         (a region this long has thousands of lags): the nr-th largest key by radix selection,
   *         most significant digit first; the keys are distinct (the lag sits in their low bits) */
  {
    unsigned long long prefix = 0; /* the digits found so far */
    int remaining = nr;            /* rank of the wanted key among the keys that share them */
    for (int shift = 48; shift >= 0; shift -= 8) {
      for (int x = tid; x < 256; x += NT) hist[x] = 0;
      __syncthreads();
      for (int p = tid; p < nlag; p += NT) {
        const unsigned long long k = keys[p];
        if ((k >> (shift + 8)) == prefix) atomicAdd(&hist[(int)((k >> shift) & 255u)], 1);
      }
      __syncthreads();
      if (warp == 0) { /* bins from 255 down, 8 per lane: the bin where the running count reaches `remaining` */
        int c = 0;
#pragma unroll
        for (int q = 0; q < 8; q++) c += hist[255 - 8 * lane - q];
        int cum = c;
#pragma unroll
        for (int dd = 1; dd < 32; dd <<= 1) {
          const int a = __shfl_up_sync(FULL, cum, dd);
          if (lane >= dd) cum += a;
        }
        if (cum >= remaining && cum - c < remaining) {
          int above = cum - c;
          for (int q = 0; q < 8; q++) {
            const int b = 255 - 8 * lane - q, h = hist[b];
            if (above + h >= remaining) { bst[4] = b; bst[5] = remaining - above; break; }
            above += h;
          }
        }
      }
      __syncthreads();
      prefix = (prefix << 8) | (unsigned long long)(unsigned)bst[4];
      remaining = bst[5];
    }
    const unsigned long long thr = prefix; /* the nr-th largest key itself */
    for (int p = tid; p < nlag; p += NT) {
      const unsigned long long k = keys[p];
      if (k >= thr) { const int at = atomicAdd(&bst[3], 1); if (at < RF_FAST_NRMAX) sel[at] = k; }
    }
    __syncthreads();
  }
  uint16_t *ranked = (uint16_t *)hist; /* lag of rank r (the histogram is dead) */
  if (tid < nr) {
    const unsigned long long k = sel[tid];
    int r = 0;
    for (int y = 0; y < nr; y++) r += sel[y] > k;
    ranked[r] = (uint16_t)(k & ((1u << RF_BIG_LAGBITS) - 1));
  }
  __syncthreads();
  /* ---- 5. per ranked lag (one thread each) the best consecutive stem (window_slide, rafft.py:36-83)
   *         over the set bits of the pairable mask: the walk of expand_fast.cuh phase 5 */
  if (tid < nr) {
    const double dGC = prm.w[1][2], dAU = prm.w[0][3], dGU = prm.w[2][3];
    const unsigned mGC = dGC == 0.0 ? 0u : FULL, mAU = dAU == 0.0 ? 0u : FULL, mGU = dGU == 0.0 ? 0u : FULL;
    const int hGC = __double2hiint(dGC), hAU = __double2hiint(dAU), hGU = __double2hiint(dGU);
    const int r = tid;
    const int p = ranked[r];
    const int s = lv - 1 - p;
    const int i0 = p < lv ? 0 : p - lv + 1;
    const int d = p < lv ? p + 1 : 2 * lv - 1 - p;
    const int len2 = (d + 1) >> 1;
    int iend = i0 + len2; /* exclusive */
    /* hairpin test (rafft.py:78): the positions that fail it are the innermost ones */
    while (iend > i0 && !((int)pos[p - (iend - 1)] - (int)pos[iend - 1] > prm.min_hp)) iend--;
    int max_i = 0, max_j = 0;
    double max_score = 0.0, prev = 0.0;
    unsigned carry = 0, best_bit = 0;
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
      const unsigned ADJ = (Pm << 1) | carry;
      carry = Pm >> 31;
      unsigned IS = CL[wi] & fast_shift(CLR, wi, s - 1);
      if ((i0 >> 5) == wi) IS &= ~(1u << (i0 & 31));
      const int wbase = wi << 16;
      /* run lengths beyond 65535 cannot occur (L <= 4096) */
      while (Pm) {
        const unsigned bit = Pm & (0u - Pm);
        Pm ^= bit;
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
    r_stem[r] = ((uint32_t)max_i << 16) | (uint32_t)max_j;
    r_nb[r] = (uint16_t)max_nb;
    atomicAdd(&bst[2], len2);
    /* ---- 6. energy change of the stem; stems below the threshold get a (dE, rank) key */
    if (max_nb > 0) {
      FastTask T;
      T.P = E; T.code = code; T.L = L; T.br = br; T.ci = ci; T.cj = cj; T.m = m; T.lv = lv; T.e_loop = n.e_loop;
      T.pos = pos; T.brcnt = brcnt; T.pre_ml = pre_ml; T.pre_ext = pre_ext;
      const int dE = fast_stem_dE(T, ev, max_i, max_j, max_nb);
      atomicAdd(&bst[1], 1);
      if (dE < prm.cand_thr) ck[atomicAdd(&bst[0], 1)] = ~((((uint32_t)(dE + 0x800000)) << 8) | (uint32_t)r);
    }
  }
  __syncthreads();
  /* ---- 7. sort the kept stems (one warp, registers), allocate and write their records */
  const int keep_n = bst[0];
  int nkc = 32;
  while (nkc < keep_n) nkc <<= 1;
  for (int x = keep_n + tid; x < nkc; x += NT) ck[x] = 0;
  __syncthreads();
  if (warp == 0 && keep_n > 0) {
    fast_sort_dispatch(ck, nkc, lane);
    if (lane == 0) {
      int off = atomicAdd(&st->n_cand, keep_n);
      if (off + keep_n > caps.cand_cap) { st->status = ST_ERR_CAPACITY; st->fail_line = -__LINE__; off = -1; }
      bst[6] = off;
    }
  }
  __syncthreads();
  const int off = keep_n > 0 ? bst[6] : 0;
  if (keep_n > 0 && off >= 0) {
    Cand *cand = (Cand *)(arena + aoff.cand) + off;
    CandKid *ckid = (CandKid *)(arena + aoff.ckid) + off;
    for (int at = tid; at < keep_n; at += NT) {
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
  if (tid == 0) {
    n.n_ranked = nr; n.n_pos = bst[1]; n.w_scan = bst[2];
    n.cand_off = off < 0 ? 0 : off;
    n.cand_n = off < 0 ? 0 : keep_n;
    atomicAdd(&st->cnt[CNT_UNIQ_NODES], 1ULL);
    atomicAdd(&st->cnt[CNT_UNIQ_W_CORR], (unsigned long long)((long long)lv * (lv + 1) / 2));
    atomicAdd(&st->cnt[CNT_UNIQ_W_SCAN], (unsigned long long)bst[2]);
  }
  __syncthreads(); /* the block's shared memory is reused by its next region */
}

#endif /* __CUDACC__ */

}  // namespace rafft
