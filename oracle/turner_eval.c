/*
 * ORACLE -- test infrastructure, NOT the product.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may use anything under oracle/.
 *
 * CPU restatement of the energy evaluation the reference delegates to ViennaRNA:
 *   rafft/utils.py:135-138  eval_one_struct -> fold_compound.eval_structure(dot-bracket)
 *   rafft-rs/src/vienna.rs:31-34  vrna_eval_structure_pt(fc, pt) -> int (dcal/mol)
 * Third-party dependency not under /root/reference: ViennaRNA 2.4.18
 * (rafft-rs/.github/build_env.yml:8; README.org:8 ">= 2.4.17").  Its published algorithm
 * (loop decomposition, Turner-2004 parameters, dangles=2, special hairpins, no constraints)
 * is restated here from SURVEY.md Appendix B.
 *
 * Parity pinning: verified against the 11,505 (sequence, structure, energy) triples the
 * reference ships (benchmark_results/*_scores.csv -> tests/golden/energy_kat.csv.gz), the
 * -25.80 kcal/mol known answer of rafft-rs/src/vienna.rs:92-109 and every line of
 * example/rafft.out / rafft_20.out.  Table entries none of these touch are parity-unpinned
 * (DESIGN.md section 4).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "../include/rafft_b200.h"
#include "../rafft_b200/csrc/turner2004.inc"

/* ---- built-in tables (same generated initialisers the product uses) ---- */
void oracle_default_params(rafft_energy_params *ep) {
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
    ep->hairpin[i] = hairpin[i];
    ep->bulge[i] = bulge[i];
    ep->interior[i] = interior[i];
  }
  for (int i = 0; i < 96; i++) {
    (&ep->mismatchH[0][0][0])[i] = mmH[i];
    (&ep->mismatchI[0][0][0])[i] = mmI[i];
    (&ep->mismatch1nI[0][0][0])[i] = mm1n[i];
    (&ep->mismatch23I[0][0][0])[i] = mm23[i];
    (&ep->mismatchM[0][0][0])[i] = mmM[i];
    (&ep->mismatchExt[0][0][0])[i] = mmE[i];
  }
  for (int i = 0; i < 24; i++) {
    (&ep->dangle5[0][0])[i] = d5[i];
    (&ep->dangle3[0][0])[i] = d3[i];
  }
  for (int i = 0; i < 576; i++) (&ep->int11[0][0][0][0])[i] = i11[i];
  for (int i = 0; i < 2304; i++) (&ep->int21[0][0][0][0][0])[i] = i21[i];
  for (int i = 0; i < 9216; i++) (&ep->int22[0][0][0][0][0][0])[i] = i22[i];
  ep->ninio = T04_NINIO;
  ep->max_ninio = T04_MAX_NINIO;
  ep->ml_base = T04_ML_BASE;
  ep->ml_intern = T04_ML_INTERN;
  ep->ml_closing = T04_ML_CLOSING;
  ep->terminal_au = T04_TERMINAL_AU;
  ep->lxc = T04_LXC;
  ep->n_tri = T04_N_TRI;
  ep->n_tetra = T04_N_TETRA;
  ep->n_hexa = T04_N_HEXA;
  for (int i = 0; i < T04_N_TRI; i++) { strncpy(ep->tri[i], tri[i].s, 7); ep->tri_e[i] = tri[i].e; }
  for (int i = 0; i < T04_N_TETRA; i++) { strncpy(ep->tetra[i], tet[i].s, 7); ep->tetra_e[i] = tet[i].e; }
  for (int i = 0; i < T04_N_HEXA; i++) { strncpy(ep->hexa[i], hex[i].s, 11); ep->hexa_e[i] = hex[i].e; }
}

/* base code 0..3 = A C G U, -1 otherwise */
static int base_code(char c) {
  switch (c) {
    case 'A': return 0;
    case 'C': return 1;
    case 'G': return 2;
    case 'U': return 3;
  }
  return -1;
}

/* pair type 0..5 = CG GC GU UG AU UA, -1 = not a canonical pair */
static int pair_type(int a, int b) {
  static const int T[4][4] = {{-1, -1, -1, 4}, {-1, -1, 0, -1}, {-1, 1, -1, 2}, {5, -1, 3, -1}};
  if (a < 0 || b < 0) return -1;
  return T[a][b];
}

static int loglen(const rafft_energy_params *P, const int32_t *arr, int u) {
  if (u <= 30) return arr[u];
  return arr[30] + (int)(P->lxc * log((double)u / 30.0));
}

/* Hairpin closed by (i,j), 0-based (Appendix B "Hairpin"). */
static int e_hairpin(const rafft_energy_params *P, const char *seq, const int *s, int i, int j) {
  int u = j - i - 1, t = pair_type(s[i], s[j]);
  int e = loglen(P, P->hairpin, u);
  if (u < 3) return e;
  if (u == 4)
    for (int k = 0; k < P->n_tetra; k++)
      if (!strncmp(seq + i, P->tetra[k], 6)) return P->tetra_e[k];
  if (u == 6)
    for (int k = 0; k < P->n_hexa; k++)
      if (!strncmp(seq + i, P->hexa[k], 8)) return P->hexa_e[k];
  if (u == 3) {
    for (int k = 0; k < P->n_tri; k++)
      if (!strncmp(seq + i, P->tri[k], 5)) return P->tri_e[k];
    return e + (t > 1 ? P->terminal_au : 0);
  }
  return e + P->mismatchH[t][s[i + 1]][s[j - 1]];
}

/* Interior loop / bulge / stack closed by (i,j) with inner pair (p,q) (Appendix B). */
static int e_interior(const rafft_energy_params *P, const int *s, int i, int j, int p, int q) {
  int n1 = p - i - 1, n2 = j - q - 1;
  int t = pair_type(s[i], s[j]), t2 = pair_type(s[q], s[p]);
  int nl = n1 > n2 ? n1 : n2, ns = n1 > n2 ? n2 : n1;
  if (nl == 0) return P->stack[t][t2];
  if (ns == 0) {
    int e = loglen(P, P->bulge, nl);
    if (nl == 1) return e + P->stack[t][t2];
    if (t > 1) e += P->terminal_au;
    if (t2 > 1) e += P->terminal_au;
    return e;
  }
  int si1 = s[i + 1], sj1 = s[j - 1], sp1 = s[p - 1], sq1 = s[q + 1];
  if (ns == 1) {
    if (nl == 1) return P->int11[t][t2][si1][sj1];
    if (nl == 2) {
      if (n1 == 1) return P->int21[t][t2][si1][sq1][sj1];
      return P->int21[t2][t][sq1][si1][sp1];
    }
    int e = loglen(P, P->interior, nl + 1);
    int a = (nl - ns) * P->ninio;
    e += a < P->max_ninio ? a : P->max_ninio;
    return e + P->mismatch1nI[t][si1][sj1] + P->mismatch1nI[t2][sq1][sp1];
  }
  if (ns == 2) {
    if (nl == 2) return P->int22[t][t2][si1][sp1][sq1][sj1];
    if (nl == 3)
      return P->interior[5] + P->ninio + P->mismatch23I[t][si1][sj1] + P->mismatch23I[t2][sq1][sp1];
  }
  int e = loglen(P, P->interior, n1 + n2);
  int a = (nl - ns) * P->ninio;
  e += a < P->max_ninio ? a : P->max_ninio;
  return e + P->mismatchI[t][si1][sj1] + P->mismatchI[t2][sq1][sp1];
}

/* Stem (p,q) of type t inside a multiloop: 5' neighbour base a, 3' neighbour base b. */
static int e_mlstem(const rafft_energy_params *P, int t, int a, int b) {
  int e = P->ml_intern;
  if (a >= 0 && b >= 0) e += P->mismatchM[t][a][b];
  else if (a >= 0) e += P->dangle5[t][a];
  else if (b >= 0) e += P->dangle3[t][b];
  if (t > 1) e += P->terminal_au;
  return e;
}

static int e_extstem(const rafft_energy_params *P, int t, int a, int b) {
  int e = 0;
  if (a >= 0 && b >= 0) e += P->mismatchExt[t][a][b];
  else if (a >= 0) e += P->dangle5[t][a];
  else if (b >= 0) e += P->dangle3[t][b];
  if (t > 1) e += P->terminal_au;
  return e;
}

/* Energy of the loop closed by (i,j) (i = -1: exterior loop) given pair table pt
 * (pt[k] = partner or -1).  Iterative; no recursion (helices nest ~1000 deep at 3000 nt). */
static long loop_energy(const rafft_energy_params *P, const char *seq, const int *s, const int *pt,
                        int L, int i, int j) {
  int lo = i + 1, hi = (i < 0) ? L : j;
  int nbr = 0, p1 = -1, q1 = -1, unp = 0;
  long e = 0;
  /* first pass: count branches */
  for (int p = lo; p < hi;) {
    if (pt[p] > p) {
      if (!nbr) { p1 = p; q1 = pt[p]; }
      nbr++;
      p = pt[p] + 1;
    } else {
      unp++;
      p++;
    }
  }
  if (i < 0) {
    for (int p = 0; p < L;) {
      if (pt[p] > p) {
        int q = pt[p], t = pair_type(s[p], s[q]);
        e += e_extstem(P, t, p > 0 ? s[p - 1] : -1, q < L - 1 ? s[q + 1] : -1);
        p = q + 1;
      } else p++;
    }
    return e;
  }
  if (nbr == 0) return e_hairpin(P, seq, s, i, j);
  if (nbr == 1) return e_interior(P, s, i, j, p1, q1);
  e = P->ml_closing + (long)P->ml_base * unp;
  e += e_mlstem(P, pair_type(s[j], s[i]), s[j - 1], s[i + 1]);
  for (int p = lo; p < hi;) {
    if (pt[p] > p) {
      int q = pt[p];
      e += e_mlstem(P, pair_type(s[p], s[q]), s[p - 1], s[q + 1]);
      p = q + 1;
    } else p++;
  }
  return e;
}

/* Full free energy in dcal/mol of the structure given as a 0-based pair table (-1 =
 * unpaired).  Returns T04_INF-scale values for structures with impossible loops, like
 * ViennaRNA.  Non-canonical pairs are not supported (return INT32_MIN). */
int oracle_eval_pt(const rafft_energy_params *P, const char *seq, const int *pt, int L) {
  int *s = (int *)malloc(sizeof(int) * (L > 0 ? L : 1));
  long e = 0;
  for (int k = 0; k < L; k++) s[k] = base_code(seq[k]);
  for (int k = 0; k < L; k++)
    if (pt[k] > k && pair_type(s[k], s[pt[k]]) < 0) { free(s); return INT32_MIN; }
  e += loop_energy(P, seq, s, pt, L, -1, L);
  for (int k = 0; k < L; k++)
    if (pt[k] > k) e += loop_energy(P, seq, s, pt, L, k, pt[k]);
  free(s);
  if (e > 2000000000L) e = 2000000000L;
  return (int)e;
}

/* Same from dot-bracket text; returns INT32_MIN on unbalanced input. */
int oracle_eval_db(const rafft_energy_params *P, const char *seq, const char *db, int L) {
  int *pt = (int *)malloc(sizeof(int) * (2 * (L > 0 ? L : 1)));
  int *st = pt + L, sp = 0, ok = 1;
  for (int k = 0; k < L; k++) {
    pt[k] = -1;
    if (db[k] == '(') st[sp++] = k;
    else if (db[k] == ')') {
      if (!sp) { ok = 0; break; }
      int i = st[--sp];
      pt[i] = k;
      pt[k] = i;
    }
  }
  int e = (ok && sp == 0) ? oracle_eval_pt(P, seq, pt, L) : INT32_MIN;
  free(pt);
  return e;
}
