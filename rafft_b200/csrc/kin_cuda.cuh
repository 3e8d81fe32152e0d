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
 * the lanes then hash the row.  bad[s] = 1 for a closing bracket without partner.  round_only:
 * the ViennaRNA pair table of the distance matrix below (other brackets stay unpaired). */
__global__ void kin_pairtable_kernel(const char *__restrict__ db, int n, int L, int16_t *__restrict__ pt,
                                     int16_t *__restrict__ stack /* n * 2 * L */, uint64_t *__restrict__ hash,
                                     int32_t *__restrict__ bad, int round_only = 0) {
  const int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (s >= n) return;
  const char *row = db + (size_t)s * L;
  int16_t *p = pt + (size_t)s * L;
  if (lane == 0) {
    int16_t *reg = stack + (size_t)s * 2 * L, *pk = reg + L;
    int nr = 0, nk = 0, err = 0;
    for (int i = 0; i < L; i++) {
      char c = row[i];
      if (round_only && c != '(' && c != ')') c = '.'; /* vrna_ptable: only round brackets pair */
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

/* Landscape distance matrix (reference utility/surface.py:19-26, get_distance_matrix): dist[a][b] =
 * RNA.bp_distance = number of pairs present in exactly one of the two structures, i.e. over the
 * pair tables  sum_i [pa[i] != pb[i]] * ([pa[i] > i] + [pb[i] > i]).
 * One block per 32 x 32 tile of the upper triangle; the 64 pair-table rows of the tile stream
 * through shared memory in chunks of KIN_DCHUNK positions (read once per tile instead of once per
 * pair); each of the 256 threads owns 4 pairs (one row a, four columns b).  Rows of the chunk are
 * padded by one 32-bit word so the per-thread row walks hit distinct banks. */
#define KIN_DTILE 32
#define KIN_DCHUNK 128
__global__ void __launch_bounds__(256) kin_bpdist_kernel(const int16_t *__restrict__ pt, int n, int L,
                                                         int32_t *__restrict__ dist) {
  /* tile (ta, tb), tb >= ta, from the linear block index */
  const int nt = (n + KIN_DTILE - 1) / KIN_DTILE;
  int ta = 0, rem = blockIdx.x;
  while (rem >= nt - ta) { rem -= nt - ta; ta++; }
  const int tb = ta + rem;
  constexpr int ROW = KIN_DCHUNK + 2;
  __shared__ int16_t sa[KIN_DTILE][ROW], sb[KIN_DTILE][ROW];
  const int tid = threadIdx.x;
  const int ra = tid >> 3, cb = (tid & 7) * 4; /* row of the a tile, first of 4 columns of the b tile */
  int acc[4] = {0, 0, 0, 0};
  for (int c0 = 0; c0 < L; c0 += KIN_DCHUNK) {
    const int cl = min(KIN_DCHUNK, L - c0);
    for (int x = tid; x < KIN_DTILE * KIN_DCHUNK; x += 256) {
      const int r = x / KIN_DCHUNK, c = x % KIN_DCHUNK;
      const int ga = ta * KIN_DTILE + r, gb = tb * KIN_DTILE + r;
      sa[r][c] = (c < cl && ga < n) ? pt[(size_t)ga * L + c0 + c] : (int16_t)-1;
      sb[r][c] = (c < cl && gb < n) ? pt[(size_t)gb * L + c0 + c] : (int16_t)-1;
    }
    __syncthreads();
    for (int c = 0; c < cl; c++) {
      const int i = c0 + c;
      const int pa = sa[ra][c];
      const int ua = pa > i;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int pb = sb[cb + k][c];
        acc[k] += (pa != pb) ? ua + (pb > i) : 0;
      }
    }
    __syncthreads();
  }
  const int a = ta * KIN_DTILE + ra;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const int b = tb * KIN_DTILE + cb + k;
    if (a < n && b < n) {
      dist[(size_t)a * n + b] = acc[k];
      dist[(size_t)b * n + a] = acc[k];
    }
  }
}

}  // namespace rafft
