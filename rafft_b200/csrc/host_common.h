/*
 * host_common.h -- host-side preparation shared by the CUDA library (rafft_cuda.cu) and the
 * host-emulation test build (tests/emu): argument validation, device table layout, arena
 * capacities.  Plain C++; no CUDA, no torch.
 */
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <stdlib.h>
#include <string>
#include "../../include/rafft_b200.h"
#include "fold_core.cuh"
#include "turner2004.inc"

namespace rafft {

/* Built-in Turner-2004 tables (reconstructed, tools/gen_tables.py). */
inline void default_energy_params(rafft_energy_params *ep) {
  static const int stack[36] = T04_STACK;
  static const int hairpin[31] = T04_HAIRPIN, bulge[31] = T04_BULGE, interior[31] = T04_INTERIOR;
  static const int mmH[96] = T04_MISMATCH_H, mmI[96] = T04_MISMATCH_I, mm1n[96] = T04_MISMATCH_1NI;
  static const int mm23[96] = T04_MISMATCH_23I, mmM[96] = T04_MISMATCH_M, mmE[96] = T04_MISMATCH_EXT;
  static const int d5[24] = T04_DANGLE5, d3[24] = T04_DANGLE3;
  static const int i11[576] = T04_INT11, i21[2304] = T04_INT21, i22[9216] = T04_INT22;
  static const struct { const char *s; int e; } tri[] = T04_TRILOOPS, tet[] = T04_TETRALOOPS,
                                                hex[] = T04_HEXALOOPS;
  memset(ep, 0, sizeof *ep);
  for (int i = 0; i < 36; i++) (&ep->stack[0][0])[i] = stack[i];
  for (int i = 0; i < 31; i++) {
    ep->hairpin[i] = hairpin[i]; ep->bulge[i] = bulge[i]; ep->interior[i] = interior[i];
  }
  for (int i = 0; i < 96; i++) {
    (&ep->mismatchH[0][0][0])[i] = mmH[i];   (&ep->mismatchI[0][0][0])[i] = mmI[i];
    (&ep->mismatch1nI[0][0][0])[i] = mm1n[i]; (&ep->mismatch23I[0][0][0])[i] = mm23[i];
    (&ep->mismatchM[0][0][0])[i] = mmM[i];   (&ep->mismatchExt[0][0][0])[i] = mmE[i];
  }
  for (int i = 0; i < 24; i++) { (&ep->dangle5[0][0])[i] = d5[i]; (&ep->dangle3[0][0])[i] = d3[i]; }
  for (int i = 0; i < 576; i++) (&ep->int11[0][0][0][0])[i] = i11[i];
  for (int i = 0; i < 2304; i++) (&ep->int21[0][0][0][0][0])[i] = i21[i];
  for (int i = 0; i < 9216; i++) (&ep->int22[0][0][0][0][0][0])[i] = i22[i];
  ep->ninio = T04_NINIO; ep->max_ninio = T04_MAX_NINIO; ep->ml_base = T04_ML_BASE;
  ep->ml_intern = T04_ML_INTERN; ep->ml_closing = T04_ML_CLOSING;
  ep->terminal_au = T04_TERMINAL_AU; ep->lxc = T04_LXC;
  ep->n_tri = T04_N_TRI; ep->n_tetra = T04_N_TETRA; ep->n_hexa = T04_N_HEXA;
  for (int i = 0; i < T04_N_TRI; i++) { strncpy(ep->tri[i], tri[i].s, 7); ep->tri_e[i] = tri[i].e; }
  for (int i = 0; i < T04_N_TETRA; i++) { strncpy(ep->tetra[i], tet[i].s, 7); ep->tetra_e[i] = tet[i].e; }
  for (int i = 0; i < T04_N_HEXA; i++) { strncpy(ep->hexa[i], hex[i].s, 11); ep->hexa_e[i] = hex[i].e; }
}

inline int base_code_host(char c) {
  switch (c) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'U': return 3; }
  return -1;
}

/* API tables -> device layout.  Returns false (err set) for values the int16 layout or the
 * special-loop keys cannot hold. */
inline bool build_dev_energy(const rafft_energy_params *ep, DevEnergy *D, std::string &err) {
  memset(D, 0, sizeof *D);
  auto put16 = [&](int16_t *dst, const int32_t *src, int n, const char *name) {
    for (int i = 0; i < n; i++) {
      if (src[i] > 32767 || src[i] < -32768) { err = std::string("energy table out of int16 range: ") + name; return false; }
      dst[i] = (int16_t)src[i];
    }
    return true;
  };
  for (int i = 0; i < 31; i++) { D->hairpin[i] = ep->hairpin[i]; D->bulge[i] = ep->bulge[i]; D->interior[i] = ep->interior[i]; }
  if (!put16(D->stack, &ep->stack[0][0], 36, "stack")) return false;
  if (!put16(D->mmH, &ep->mismatchH[0][0][0], 96, "mismatch_hairpin")) return false;
  if (!put16(D->mmI, &ep->mismatchI[0][0][0], 96, "mismatch_interior")) return false;
  if (!put16(D->mm1n, &ep->mismatch1nI[0][0][0], 96, "mismatch_interior_1n")) return false;
  if (!put16(D->mm23, &ep->mismatch23I[0][0][0], 96, "mismatch_interior_23")) return false;
  if (!put16(D->mmM, &ep->mismatchM[0][0][0], 96, "mismatch_multi")) return false;
  if (!put16(D->mmE, &ep->mismatchExt[0][0][0], 96, "mismatch_exterior")) return false;
  if (!put16(D->d5, &ep->dangle5[0][0], 24, "dangle5")) return false;
  if (!put16(D->d3, &ep->dangle3[0][0], 24, "dangle3")) return false;
  if (!put16(D->int11, &ep->int11[0][0][0][0], 576, "int11")) return false;
  if (!put16(D->int21, &ep->int21[0][0][0][0][0], 2304, "int21")) return false;
  if (!put16(D->int22, &ep->int22[0][0][0][0][0][0], 9216, "int22")) return false;
  D->ninio = ep->ninio; D->max_ninio = ep->max_ninio; D->ml_base = ep->ml_base;
  D->ml_intern = ep->ml_intern; D->ml_closing = ep->ml_closing; D->terminal_au = ep->terminal_au;
  if (ep->n_tri > 64 || ep->n_tetra > 64 || ep->n_hexa > 64 || ep->n_tri < 0 || ep->n_tetra < 0 || ep->n_hexa < 0) {
    err = "too many special hairpins"; return false;
  }
  auto key = [&](const char *s, int n, uint32_t &k) {
    k = 0;
    for (int i = 0; i < n; i++) {
      int c = base_code_host(s[i]);
      if (c < 0) return false;
      k = (k << 2) | (uint32_t)c;
    }
    return true;
  };
  D->n_tri = ep->n_tri; D->n_tetra = ep->n_tetra; D->n_hexa = ep->n_hexa;
  for (int i = 0; i < ep->n_tri; i++) { if (!key(ep->tri[i], 5, D->tri_key[i])) { err = "bad triloop"; return false; } D->tri_e[i] = ep->tri_e[i]; }
  for (int i = 0; i < ep->n_tetra; i++) { if (!key(ep->tetra[i], 6, D->tetra_key[i])) { err = "bad tetraloop"; return false; } D->tetra_e[i] = ep->tetra_e[i]; }
  for (int i = 0; i < ep->n_hexa; i++) { if (!key(ep->hexa[i], 8, D->hexa_key[i])) { err = "bad hexaloop"; return false; } D->hexa_e[i] = ep->hexa_e[i]; }
  for (int i = 0; i < ep->n_tri; i++) D->tri_bm[D->tri_key[i] >> 5] |= 1u << (D->tri_key[i] & 31);
  for (int i = 0; i < ep->n_tetra; i++) D->tetra_bm[D->tetra_key[i] >> 5] |= 1u << (D->tetra_key[i] & 31);
  for (int i = 0; i < ep->n_hexa; i++) D->hexa_bm[D->hexa_key[i] >> 5] |= 1u << (D->hexa_key[i] & 31);
  for (int u = 0; u < 4096; u++) D->logterm[u] = u > 30 ? (int)(ep->lxc * log((double)u / 30.0)) : 0;
  return true;
}

/* fold() arguments -> launch constants (reference rafft/rafft.py:219-223, bin/rafft:11-27). */
inline bool build_fold_params(const rafft_params *p, FoldParams *F, std::string &err) {
  memset(F, 0, sizeof *F);
  if (p->max_stack < 1 || p->max_stack > 32767) { err = "max_stack must be in [1, 32767]"; return false; }
  if (p->nb_mode < 0 || p->nb_mode > 8191) { err = "nb_mode must be in [0, 8191]"; return false; }
  if (p->min_hp < 0) { err = "min_hp must be >= 0"; return false; }
  if (p->max_branch > 1000000) { err = "max_branch must be <= 1000000"; return false; }
  if (p->temp != 37.0) { err = "only temp = 37.0 is supported (the built-in tables carry no enthalpies)"; return false; }
  if (!(fabs(p->min_nrj) <= 1.0e6)) { err = "min_nrj out of range"; return false; }
  const double ws[3] = {p->gc_wei, p->au_wei, p->gu_wei};
  for (int k = 0; k < 3; k++) {
    double s = ws[k] * 16.0;
    if (!(ws[k] >= 0.0 && ws[k] <= 16.0) || s != floor(s)) {
      err = "pair weights must be multiples of 1/16 in [0, 16] (exact-correlation contract, DESIGN.md section 3)";
      return false;
    }
  }
  F->nb_mode = p->nb_mode; F->max_stack = p->max_stack; F->max_branch = p->max_branch;
  F->min_hp = p->min_hp; F->traj = p->traj ? 1 : 0; F->min_nrj = p->min_nrj;
  /* bases 0..3 = A C G U */
  F->w[0][3] = F->w[3][0] = p->au_wei;
  F->w[1][2] = F->w[2][1] = p->gc_wei;
  F->w[2][3] = F->w[3][2] = p->gu_wei;
  for (int a = 0; a < 4; a++)
    for (int b = 0; b < 4; b++) F->wi[a][b] = (int32_t)(F->w[a][b] * 16.0);
  F->cand_thr = p->min_nrj == 0.0 ? 0 : (int32_t)ceil(p->min_nrj * 100.0) + 2;
  F->seen_gen = 1; /* the caller advances it per fold of an arena */
  return true;
}

inline int32_t pow2_at_least(int64_t x) {
  int64_t p = 16;
  while (p < x) p <<= 1;
  return (int32_t)p;
}

/* Arena capacities for sequences up to L_max, about twice the high-water marks measured on
 * the benchmark mix; `scale` grows on every regrowth after an overflow. */
inline Caps make_caps(int L_max, const FoldParams &F, int nthreads, int scale) {
  Caps c;
  memset(&c, 0, sizeof c);
  int64_t mb = F.max_branch > 0 ? F.max_branch : 0;
  int64_t ms = F.max_stack;
  int64_t mbc = mb < 4096LL * scale ? mb : 4096LL * scale; /* children kept per step */
  c.L_max = L_max < 16 ? 16 : L_max;
  c.chunk_cap = (int32_t)(4 * nthreads > ms ? 4 * nthreads : ms);
  if (const char *e = getenv("RAFFT_CHUNK")) { /* tuning: stem combinations examined per block phase */
    const int v = atoi(e);
    if (v >= 64 && v <= 8192 && v >= ms) c.chunk_cap = v;
  }
  int64_t kmin = 2 * (int64_t)c.L_max > F.nb_mode ? 2 * (int64_t)c.L_max : F.nb_mode;
  c.key_cap = pow2_at_least(kmin < 32 ? 32 : kmin);
  c.mask_words = c.L_max / 32 + 2;
  const int64_t node_base = (256 + 6 * ms + 2 * (int64_t)c.L_max) * scale;
  c.node_cap = (int32_t)(node_base * 3 / 2); /* loop records are small; the stems dominate */
  c.pool16_cap = c.node_cap * 12;
  int per_node = F.nb_mode / 4 < 8 ? 8 : F.nb_mode / 4 > 64 ? 64 : F.nb_mode / 4;
  c.cand_cap = (int32_t)(node_base * per_node);
  /* structures that ever enter a beam, and their region lists: both grow with the number of
   * folding steps, i.e. with the length (3000 nt, -ms 200: 3.5k structures, 18 regions each) */
  const int64_t per_ms = 16 > c.L_max / 64 ? 16 : c.L_max / 64, per_st = 12 > c.L_max / 100 ? 12 : c.L_max / 100;
  c.store_cap = (int32_t)(1 + (ms * per_ms + 48) * scale); /* at least one per folding step (-ms 1) */
  c.nlist_cap = (int32_t)(c.store_cap * per_st);
  c.new_cap = (int32_t)(mbc + ms + 1);
  /* up to max_branch + max_stack new fingerprints per folding step, table kept under half full;
   * the number of steps grows with the length (13 at <= 500 nt, 20-35 at 3000-4000 nt) */
  c.seen_cap = pow2_at_least((16 + c.L_max / 50) * (mbc + ms) * scale + 64);
  c.nht_cap = pow2_at_least(4 * (int64_t)c.node_cap);
  c.step_cap = 48 * scale;
  c.todo_cap = (int32_t)(256 + 8 * ms) * scale;
  WarpScratch W;
  c.wscratch_bytes = (int32_t)wscratch_carve(W, nullptr, c, F);
  Arena A;
  c.arena_bytes = (int64_t)arena_carve(A, nullptr, c, F);
  return c;
}

}  // namespace rafft
