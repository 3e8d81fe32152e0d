/*
 * kin_cuda.cuh -- the folding-graph part of the kinetics (reference rafft/rafft_kin.py):
 * which structures of consecutive folding steps are connected (get_connected_prev, :48-56: every
 * pair of the earlier structure is a pair of the later one) and the non-redundant numbering of
 * the structures (kinetics, :106-115).  The rate matrix and the eigen-decomposition stay on the
 * host (numpy longdouble / LAPACK, rafft_b200/rafft_kin.py), exactly as in the reference.
 *
 * Layout: n structures of one sequence of length L as dot-bracket rows (n * L chars, the folding
 * steps concatenated); pair tables int16[n][L] (partner or -1) are built on the device.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rafft {

/* one warp per structure: lane 0 parses the brackets with the two stacks of paired_positions
 * (reference rafft/utils.py:53-67: "(", "<" / ")", ">" regular, "[" / "]" pseudo-knot pile);
 * the lanes then hash the row.  bad[s] = 1 for an unbalanced or foreign character. */
__global__ void kin_pairtable_kernel(const char *__restrict__ db, int n, int L, int16_t *__restrict__ pt,
                                     int16_t *__restrict__ stack /* n * 2 * L */, uint64_t *__restrict__ hash,
                                     int32_t *__restrict__ bad) {
  const int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (s >= n) return;
  const char *row = db + (size_t)s * L;
  int16_t *p = pt + (size_t)s * L;
  if (lane == 0) {
    int16_t *reg = stack + (size_t)s * 2 * L, *pk = reg + L;
    int nr = 0, nk = 0, err = 0;
    for (int i = 0; i < L; i++) {
      const char c = row[i];
      p[i] = -1;
      if (c == '(' || c == '<') reg[nr++] = (int16_t)i;
      else if (c == '[') pk[nk++] = (int16_t)i;
      else if (c == ')' || c == '>') {
        if (!nr) { err = 1; break; }
        const int j = reg[--nr]; p[j] = (int16_t)i; p[i] = (int16_t)j;
      } else if (c == ']') {
        if (!nk) { err = 1; break; }
        const int j = pk[--nk]; p[j] = (int16_t)i; p[i] = (int16_t)j;
      }
    }
    /* the reference leaves unmatched opening brackets unpaired (paired_positions never pops
     * them); only a closing bracket without partner raises there */
    for (int k = 0; k < nr; k++) p[reg[k]] = -1;
    for (int k = 0; k < nk; k++) p[pk[k]] = -1;
    bad[s] = err;
  }
  uint64_t h = 0x9E3779B97F4A7C15ULL;
  for (int i = lane; i < L; i += 32) {
    uint64_t z = (uint64_t)(unsigned char)row[i] + ((uint64_t)i << 8) + 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    h += z ^ (z >> 31);
  }
  for (int d = 16; d; d >>= 1) h += __shfl_xor_sync(0xffffffffu, h, d);
  if (lane == 0) hash[s] = h;
}

/* uniq[i] = smallest j <= i whose row equals row i (kinetics :106-115 keeps the first occurrence) */
__global__ void kin_unique_kernel(const char *__restrict__ db, const uint64_t *__restrict__ hash, int n, int L,
                                  int32_t *__restrict__ uniq) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t h = hash[i];
  int first = i;
  for (int j = 0; j < i; j++) {
    if (hash[j] != h) continue;
    const char *a = db + (size_t)i * L, *b = db + (size_t)j * L;
    bool same = true;
    for (int k = 0; k < L && same; k++) same = a[k] == b[k];
    if (same) { first = j; break; }
  }
  uniq[i] = first;
}

/* one warp per (current structure c of step t, structure p of the step before): conn = 1 iff
 * pairs(p) is a subset of pairs(c).  Step 0 looks "back" at the LAST step (fast_paths[-1],
 * rafft_kin.py:75: Python's negative index).  conn rows of step t start at conn_off[t];
 * row of c has n_prev entries. */
__global__ void kin_connect_kernel(const int16_t *__restrict__ pt, const int64_t *__restrict__ step_off,
                                   const int64_t *__restrict__ conn_off, int n_steps, int L,
                                   uint8_t *__restrict__ conn, int64_t total_pairs) {
  const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= total_pairs) return;
  /* step holding pair w: last t with conn_off[t] <= w */
  int lo = 0, hi = n_steps - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (conn_off[mid] <= w) lo = mid; else hi = mid - 1;
  }
  const int t = lo, tp = t == 0 ? n_steps - 1 : t - 1;
  const int64_t n_prev = step_off[tp + 1] - step_off[tp];
  const int64_t r = w - conn_off[t];
  const int64_t c = step_off[t] + r / n_prev, p = step_off[tp] + r % n_prev;
  const int16_t *pc = pt + c * L, *pp = pt + p * L;
  int viol = 0;
  for (int i = lane; i < L; i += 32) {
    const int j = pp[i];
    viol |= (j > i) && (pc[i] != j);
  }
  viol = __any_sync(0xffffffffu, viol);
  if (lane == 0) conn[w] = viol ? 0 : 1;
}

}  // namespace rafft
