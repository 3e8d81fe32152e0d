/*
 * rafft_cuda.cu -- the C-ABI library (include/rafft_b200.h) over the sm_100a kernels.
 *
 * Kernels in this file:
 *   fold_kernel    persistent thread blocks; each block pulls the next sequence from a device
 *                  queue and runs the whole fold (fold_core.cuh) out of its own arena
 *   scan_kernel    exclusive scans that turn per-sequence counts into the ordered layout
 *   pack_kernel    gathers the bump-allocated results into input order
 *   eval_kernel    whole-structure loop-decomposition energies (rafft_eval_structures)
 *
 * There is no CPU fallback: every entry point fails with RAFFT_ERR_CUDA when no device is
 * usable.  Nothing here includes or links anything under oracle/.
 */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <chrono>
#include <string>
#include <vector>
#include "host_common.h"
#include "expand_fast.cuh"
#include "kin_cuda.cuh"

using namespace rafft;

#define RF_BLOCK_THREADS_MAX 256

/* ------------------------------------------------------------------------------------ */
/* kernels                                                                               */

/* Everything the phase kernels of one wave need (kernel parameter, constant bank).
 * A wave = `n` sequences folded together; slot j of the wave works on sequence order[j], owns
 * arena j and state j.  The fold of a sequence is a chain of small kernels per folding
 * step: each kernel's hot code fits the instruction cache and every
 * resident warp runs the same phase. */
struct WaveCtx {
  const DevEnergy *E;
  FoldParams prm;
  Caps caps;
  ArenaOff aoff;
  char *arenas;
  SeqState *states;
  const char *seqs;
  const int64_t *offs;
  const int32_t *order;
  int32_t n;
  TaskQueue Q;
  FastCfg fast;
  int32_t *fast_next; /* dynamic work counter of k_expand_fast */
  int32_t sel_smem;   /* bytes of shared memory for the energies under selection (0: global) */
  int32_t *n_active;
  OutBuf out;
};

#define RF_WAVE_PARAM const __grid_constant__ WaveCtx c

/* block-scope kernels: one block per wave slot, state staged through shared memory */
struct BlockBind {
  __device__ __forceinline__ static bool load(const WaveCtx &c, Fold &f, SeqState *sh, bool need_active) {
    const int j = blockIdx.x;
    SeqState *st = c.states + j;
    if (need_active && !st->active) return false;
    for (int k = threadIdx.x; k < (int)(sizeof(SeqState) / 4); k += blockDim.x) ((int *)sh)[k] = ((const int *)st)[k];
    arena_bind(f.A, c.arenas + (size_t)j * (size_t)c.caps.arena_bytes, c.aoff);
    f.sh = sh;
    f.Q = c.Q;
    f.Q.self = j;
    __syncthreads();
    f.bind_ev();
    return true;
  }
  __device__ __forceinline__ static void store(const WaveCtx &c, SeqState *sh) {
    __syncthreads();
    SeqState *st = c.states + blockIdx.x;
    for (int k = threadIdx.x; k < (int)(sizeof(SeqState) / 4); k += blockDim.x) ((int *)st)[k] = ((const int *)sh)[k];
  }
};

__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_init(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  Fold f(c.E, c.prm, c.caps);
  const int j = blockIdx.x;
  arena_bind(f.A, c.arenas + (size_t)j * (size_t)c.caps.arena_bytes, c.aoff);
  f.sh = &sh;
  const int si = c.order[j];
  const int64_t o = c.offs[si];
  if (threadIdx.x == 0) sh.seq_off = o;
  f.phase_init(c.seqs + o, (int)(c.offs[si + 1] - o), si);
  BlockBind::store(c, &sh);
}

__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_gather(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  Fold f(c.E, c.prm, c.caps);
  if (!BlockBind::load(c, f, &sh, true)) return;
  f.phase_gather();
  BlockBind::store(c, &sh);
}

__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_combine(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  Fold f(c.E, c.prm, c.caps);
  extern __shared__ __align__(16) char combine_smem[];
  if (!BlockBind::load(c, f, &sh, true)) return;
  stage_carve(f.cs, combine_smem, c.caps.chunk_cap, c.prm.max_stack);
  f.phase_combine();
  BlockBind::store(c, &sh);
}

__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_select(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  Fold f(c.E, c.prm, c.caps);
  extern __shared__ __align__(16) char select_smem[];
  if (!BlockBind::load(c, f, &sh, true)) return;
  if (c.sel_smem) f.A.sel_e = (int32_t *)select_smem;
  f.phase_select();
  BlockBind::store(c, &sh);
}

__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_materialise(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  Fold f(c.E, c.prm, c.caps);
  if (!BlockBind::load(c, f, &sh, true)) return;
  f.phase_materialise();
  if (threadIdx.x == 0 && sh.active) atomicAdd(c.n_active, 1);
  BlockBind::store(c, &sh);
}

__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_finish(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  Fold f(c.E, c.prm, c.caps);
  BlockBind::load(c, f, &sh, false);
  OutBuf out = c.out;
  f.phase_finish(out);
}

/* warp-scope kernels: one warp per queued region, grid-stride over the work list */
template <int PHASE>
__global__ void __launch_bounds__(128) k_expand(RF_WAVE_PARAM) {
  const int nt = min(*c.Q.n_tasks, c.Q.cap);
  const int wpb = blockDim.x >> 5;
  for (int t = blockIdx.x * wpb + (threadIdx.x >> 5); t < nt; t += gridDim.x * wpb) {
    const int j = c.Q.task[2 * t], v = c.Q.task[2 * t + 1];
    Fold f(c.E, c.prm, c.caps);
    arena_bind(f.A, c.arenas + (size_t)j * (size_t)c.caps.arena_bytes, c.aoff);
    f.sh = c.states + j;
    f.bind_ev();
    WarpScratch W;
    wscratch_carve(W, c.Q.scratch + (size_t)t * (size_t)c.caps.wscratch_bytes, c.caps, c.prm);
    NodeRec &n = f.A.node[v];
    if (PHASE == 1) f.expand_e1(n, W);
    if (PHASE == 2) f.expand_e2(n, W);
    if (PHASE == 3) f.expand_e3(n, W);
    if (PHASE == 4) f.expand_e4(n, W);
    if (PHASE == 5) f.expand_e5(n, W);
  }
}

/* shared-memory expansion: one warp per region; the warps of a block take RF_FAST_WARPS regions of
 * one size class at a time and walk through the phases together (block barriers), so that an SM
 * executes one phase at a time and its code stays in the instruction caches. */
template <int K>
__device__ __forceinline__ void fast_class(const WaveCtx &c, int cls, char *ws, int *s_base) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int nt = min(c.Q.n_fast[cls], c.Q.fast_cap);
  const int32_t *list = c.Q.fast_task + 2 * (size_t)cls * (size_t)c.Q.fast_cap;
  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) *s_base = atomicAdd(c.fast_next + cls, nwarps);
    __syncthreads();
    const int base = *s_base;
    if (base >= nt) break;
    const int t = base + warp;
    const bool has = t < nt;
    int j = 0, v = 0;
    if (has) { j = list[2 * t]; v = list[2 * t + 1]; }
    char *arena = c.arenas + (size_t)j * (size_t)c.caps.arena_bytes;
    SeqState *st = c.states + j;
#define RF_FAST_CALL c.E, c.prm, c.caps, c.fast, arena, c.aoff, st, v, ws, lane
    /* three code phases per region (set up + rank | scan | energies + records), separated by
     * block barriers; warp barriers order the shared-memory traffic inside a phase */
    if (has) {
      fast_p1(RF_FAST_CALL);
      __syncwarp();
      fast_p2(RF_FAST_CALL);
      __syncwarp();
      fast_p3(RF_FAST_CALL, 32 * K);
      __syncwarp();
      fast_p4<K>(RF_FAST_CALL);
    }
    __syncthreads();
    if (has) fast_p5(RF_FAST_CALL);
    __syncthreads();
    if (has) {
      fast_p6(RF_FAST_CALL);
      __syncwarp();
      fast_p7(RF_FAST_CALL);
    }
#undef RF_FAST_CALL
  }
}

template <int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_expand_fast(RF_WAVE_PARAM) {
  extern __shared__ __align__(16) char fast_smem[];
  __shared__ int s_base;
  char *ws = fast_smem + (size_t)(threadIdx.x >> 5) * (size_t)c.fast.smem_per_warp;
  /* largest regions first */
#pragma unroll 1
  for (int sub = 1; sub >= 0; sub--) fast_class<32>(c, 10 + sub, ws, &s_base);
#pragma unroll 1
  for (int sub = 1; sub >= 0; sub--) fast_class<16>(c, 8 + sub, ws, &s_base);
#pragma unroll 1
  for (int sub = 1; sub >= 0; sub--) fast_class<8>(c, 6 + sub, ws, &s_base);
#pragma unroll 1
  for (int sub = 1; sub >= 0; sub--) fast_class<4>(c, 4 + sub, ws, &s_base);
#pragma unroll 1
  for (int sub = 1; sub >= 0; sub--) fast_class<2>(c, 2 + sub, ws, &s_base);
#pragma unroll 1
  for (int sub = 1; sub >= 0; sub--) fast_class<1>(c, sub, ws, &s_base);
}

/* launch shapes of k_expand_fast: (threads per block, blocks per SM) */
typedef void (*fast_kernel_t)(const WaveCtx);
struct FastShape { int threads, blocks; fast_kernel_t fn; };
static const FastShape g_fast_shapes[] = {
    {512, 1, k_expand_fast<512, 1>}, {256, 2, k_expand_fast<256, 2>}, {256, 3, k_expand_fast<256, 3>},
    {128, 4, k_expand_fast<128, 4>}, {128, 6, k_expand_fast<128, 6>}, {384, 2, k_expand_fast<384, 2>},
    {768, 1, k_expand_fast<768, 1>}, {128, 5, k_expand_fast<128, 5>}, {64, 8, k_expand_fast<64, 8>},
    {64, 12, k_expand_fast<64, 12>}};

/* One block.  Ordered offsets: struct_off/db_off/step_off (n+1 each) and the per-sequence
 * base of its entries, ent_base (n+1); totals[0..3] = structures, db bytes, steps, entries. */
__global__ void scan_kernel(const int64_t *__restrict__ offs, int n, OutBuf out,
                            int64_t *__restrict__ struct_off, int64_t *__restrict__ db_off,
                            int64_t *__restrict__ step_off, int64_t *__restrict__ ent_base,
                            int64_t *__restrict__ totals) {
  __shared__ long long part[4][1024];
  const int T = blockDim.x, t = threadIdx.x;
  const int per = (n + T - 1) / T;
  const int lo = min(n, t * per), hi = min(n, lo + per);
  long long a[4] = {0, 0, 0, 0};
  for (int i = lo; i < hi; i++) {
    long long ns = out.seq_nstruct[i], L = offs[i + 1] - offs[i];
    long long ne = 0;
    for (int s = 0; s < out.seq_nstep[i]; s++) ne += out.step_n[out.seq_step[i] + s];
    a[0] += ns; a[1] += ns * L; a[2] += out.seq_nstep[i]; a[3] += ne;
  }
  for (int k = 0; k < 4; k++) part[k][t] = a[k];
  __syncthreads();
  if (t == 0) {
    for (int k = 0; k < 4; k++) {
      long long run = 0;
      for (int j = 0; j < T; j++) { long long v = part[k][j]; part[k][j] = run; run += v; }
      totals[k] = run;
    }
  }
  __syncthreads();
  for (int k = 0; k < 4; k++) a[k] = part[k][t];
  for (int i = lo; i < hi; i++) {
    long long ns = out.seq_nstruct[i], L = offs[i + 1] - offs[i];
    long long ne = 0;
    for (int s = 0; s < out.seq_nstep[i]; s++) ne += out.step_n[out.seq_step[i] + s];
    struct_off[i] = a[0]; db_off[i] = a[1]; step_off[i] = a[2]; ent_base[i] = a[3];
    a[0] += ns; a[1] += ns * L; a[2] += out.seq_nstep[i]; a[3] += ne;
  }
  /* every thread whose range ends at n holds the totals (empty trailing ranges included) */
  if (hi == n) { struct_off[n] = a[0]; db_off[n] = a[1]; step_off[n] = a[2]; ent_base[n] = a[3]; }
}

/* One block per sequence: copy its structures, dot-brackets, steps and entries into the
 * ordered layout.  entry_off has step_off[n] + 1 elements. */
__global__ void pack_kernel(const int64_t *__restrict__ offs, int n, OutBuf out,
                            const int64_t *__restrict__ struct_off, const int64_t *__restrict__ db_off,
                            const int64_t *__restrict__ step_off, const int64_t *__restrict__ ent_base,
                            int32_t *__restrict__ energy, char *__restrict__ db,
                            int64_t *__restrict__ entry_off, int32_t *__restrict__ entry_struct) {
  const int i = blockIdx.x;
  const int ns = out.seq_nstruct[i], nt = out.seq_nstep[i];
  const int64_t L = offs[i + 1] - offs[i];
  const int64_t s_src = out.seq_struct[i], d_src = out.seq_db[i], t_src = out.seq_step[i],
                e_src = out.seq_entry[i];
  const int64_t s_dst = struct_off[i], d_dst = db_off[i], t_dst = step_off[i], e_dst = ent_base[i];
  for (int k = threadIdx.x; k < ns; k += blockDim.x) energy[s_dst + k] = out.energy[s_src + k];
  const int64_t nbytes = (int64_t)ns * L;
  /* source rows are padded to 16 bytes (fold_core.cuh db_stride), the result is compact */
  const int64_t Lp = (L + 15) & ~(int64_t)15;
  for (int64_t k = threadIdx.x; k < nbytes; k += blockDim.x) {
    const int64_t r = k / L;
    db[d_dst + k] = out.db[d_src + r * Lp + (k - r * L)];
  }
  const int64_t ne = ent_base[i + 1] - e_dst;
  for (int64_t k = threadIdx.x; k < ne; k += blockDim.x) entry_struct[e_dst + k] = out.entries[e_src + k];
  if (threadIdx.x == 0) {
    int64_t run = e_dst;
    for (int s = 0; s < nt; s++) { entry_off[t_dst + s] = run; run += out.step_n[t_src + s]; }
    if (i == n - 1) entry_off[t_dst + nt] = run;
  }
}

/* Whole-structure energy: one warp per structure.  Lane 0 builds the pair table with an
 * explicit stack, then the lanes share the loops (one loop per closing pair + exterior). */
__global__ void eval_kernel(const DevEnergy *__restrict__ E, const char *__restrict__ seqs,
                            const char *__restrict__ dbs, const int64_t *__restrict__ offs, int n,
                            int32_t *__restrict__ scratch /* 2 ints per letter */,
                            uint8_t *__restrict__ codes, int32_t *__restrict__ energy) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n) return;
  const int64_t o = offs[warp];
  const int L = (int)(offs[warp + 1] - o);
  const char *seq = seqs + o, *db = dbs + o;
  int32_t *pt = scratch + 2 * o, *st = pt + L;
  uint8_t *code = codes + o;
  int bad = 0;
  for (int k = lane; k < L; k += 32) {
    char ch = seq[k];
    int c = ch == 'A' ? 0 : ch == 'C' ? 1 : ch == 'G' ? 2 : ch == 'U' ? 3 : -1;
    if (c < 0) { bad = 1; c = 0; }
    code[k] = (uint8_t)c;
  }
  bad = __any_sync(0xffffffffu, bad);
  __syncwarp();
  int ok = 1;
  if (lane == 0) {
    int sp = 0;
    for (int k = 0; k < L && ok; k++) {
      pt[k] = -1;
      if (db[k] == '(') st[sp++] = k;
      else if (db[k] == ')') {
        if (!sp) ok = 0;
        else { int i = st[--sp]; pt[i] = k; pt[k] = i; }
      } else if (db[k] != '.') ok = 0;
    }
    if (sp) ok = 0;
  }
  ok = __shfl_sync(0xffffffffu, ok, 0);
  __syncwarp();
  if (!ok || bad) { if (lane == 0) energy[warp] = INT32_MIN; return; }
  EnergyView ev; ev.P = E; ev.s = code; ev.L = L;
  long long e = 0;
  int nc = 0;
  for (int i = lane - 1; i < L; i += 32) { /* i = -1: exterior loop */
    int j;
    if (i >= 0) { j = pt[i]; if (j <= i) continue; if (pair_type(code[i], code[j]) < 0) { nc = 1; continue; } }
    else j = L;
    int nbr = 0, p1 = -1, q1 = -1, unp = 0;
    for (int p = i + 1; p < j;) {
      if (pt[p] > p) { if (!nbr) { p1 = p; q1 = pt[p]; } nbr++; p = pt[p] + 1; }
      else { unp++; p++; }
    }
    if (i < 0) {
      for (int p = 0; p < L;) {
        if (pt[p] > p) { if (pair_type(code[p], code[pt[p]]) >= 0) e += e_extstem(ev, p, pt[p]); p = pt[p] + 1; }
        else p++;
      }
    } else if (nbr == 0) e += e_hairpin(ev, i, j);
    else if (nbr == 1) { if (pair_type(code[p1], code[q1]) >= 0) e += e_interior(ev, i, j, p1, q1); }
    else {
      int tt = pair_type(code[j], code[i]);
      e += E->ml_closing + (long long)E->ml_base * unp + E->ml_intern +
           E->mmM[tt * 16 + code[j - 1] * 4 + code[i + 1]] + (tt > 1 ? E->terminal_au : 0);
      for (int p = i + 1; p < j;) {
        if (pt[p] > p) { if (pair_type(code[p], code[pt[p]]) >= 0) e += e_mlstem(ev, p, pt[p]); p = pt[p] + 1; }
        else p++;
      }
    }
  }
  for (int d = 16; d; d >>= 1) {
    e += __shfl_xor_sync(0xffffffffu, e, d);
    nc |= __shfl_xor_sync(0xffffffffu, nc, d);
  }
  if (lane == 0) energy[warp] = nc ? INT32_MIN : (e > 2000000000LL ? 2000000000 : (int32_t)e);
}

/* ------------------------------------------------------------------------------------ */
/* host side                                                                             */

struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
};

struct HostBlock {  /* pinned host memory handed out inside a rafft_result */
  void *p = nullptr;
  size_t cap = 0;
};

struct rafft_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, evx0 = nullptr, evx1 = nullptr;
  DevEnergy *d_energy = nullptr;
  DevBuf arenas, seqs, offs, order, idx, big, ordered, scan, evalbuf, states, tasks, ftasks, scratch, small, kinbuf;
  std::vector<HostBlock> pinned_free;
  int64_t *h_small = nullptr; /* pinned: totals + cursors */
  int block_threads = 64;   /* threads per block of the block-scope phase kernels */
  int blocks_per_sm = 16;   /* grid of the warp-task kernels: blocks per SM */
  int64_t max_wave = 0;     /* sequences folded together (0 = as many as memory allows) */
  int64_t max_tasks = 0;    /* capacity of the region work list (0 = automatic) */
  int fast_shape = 2;       /* index into g_fast_shapes (RAFFT_FAST_SHAPE) */
  int use_fast = 1;         /* shared-memory expansion for small regions (RAFFT_FAST=0 disables) */
  uint64_t arena_sig = 0;   /* layout of the arenas whose hash tables are known to be empty */
  bool arena_dirty = true;
  uint32_t seen_gen = 0;    /* generation of the last fold in the arenas (tag of its `seen` entries) */
  size_t arena_clean = 0;   /* leading bytes of the arena buffer known to be clean */
  std::string err;
};

struct ResultImpl {
  rafft_result pub;
  rafft_ctx *ctx;
  HostBlock block;
};

static std::string g_create_err;

#define CK(call)                                                                       \
  do {                                                                                 \
    cudaError_t e_ = (call);                                                           \
    if (e_ != cudaSuccess) {                                                           \
      ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                   \
      return RAFFT_ERR_CUDA;                                                           \
    }                                                                                  \
  } while (0)

static int dev_reserve(rafft_ctx *ctx, DevBuf &b, size_t bytes) {
  if (bytes <= b.cap) return RAFFT_OK;
  if (getenv("RAFFT_DEBUG_ALLOC")) fprintf(stderr, "[rafft] cudaMalloc %zu bytes (had %zu)\n", bytes, b.cap);
  if (b.p) { CK(cudaFree(b.p)); b.p = nullptr; b.cap = 0; }
  size_t want = bytes + bytes / 8 + 256;
  cudaError_t e = cudaMalloc(&b.p, want);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    want = bytes;
    e = cudaMalloc(&b.p, want);
  }
  if (e != cudaSuccess) {
    b.p = nullptr;
    ctx->err = std::string("cudaMalloc(") + std::to_string(bytes) + "): " + cudaGetErrorString(e);
    return RAFFT_ERR_CUDA;
  }
  b.cap = want;
  return RAFFT_OK;
}

static int host_block_get(rafft_ctx *ctx, size_t bytes, HostBlock *hb) {
  int best = -1;
  for (size_t k = 0; k < ctx->pinned_free.size(); k++)
    if (ctx->pinned_free[k].cap >= bytes && (best < 0 || ctx->pinned_free[k].cap < ctx->pinned_free[best].cap))
      best = (int)k;
  if (best >= 0) {
    *hb = ctx->pinned_free[best];
    ctx->pinned_free.erase(ctx->pinned_free.begin() + best);
    return RAFFT_OK;
  }
  /* drop smaller cached blocks before allocating a larger one */
  for (auto &b : ctx->pinned_free) cudaFreeHost(b.p);
  ctx->pinned_free.clear();
  size_t want = bytes + bytes / 4 + 4096;
  if (getenv("RAFFT_DEBUG_ALLOC")) fprintf(stderr, "[rafft] cudaHostAlloc %zu bytes (request %zu)\n", want, bytes);
  CK(cudaHostAlloc(&hb->p, want, cudaHostAllocDefault));
  hb->cap = want;
  return RAFFT_OK;
}

extern "C" {

const char *rafft_version(void) { return "rafft_b200 0.1 sm_100a"; }

void rafft_default_energy_params(rafft_energy_params *ep) {
  if (ep) default_energy_params(ep);
}

const char *rafft_last_error(const rafft_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

void rafft_destroy(rafft_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  DevBuf *bufs[] = {&ctx->arenas, &ctx->seqs, &ctx->offs, &ctx->order, &ctx->idx,
                    &ctx->states, &ctx->tasks, &ctx->ftasks, &ctx->scratch, &ctx->small,
                    &ctx->big, &ctx->ordered, &ctx->scan, &ctx->evalbuf, &ctx->kinbuf};
  for (DevBuf *b : bufs)
    if (b->p) cudaFree(b->p);
  for (auto &b : ctx->pinned_free) cudaFreeHost(b.p);
  if (ctx->h_small) cudaFreeHost(ctx->h_small);
  if (ctx->d_energy) cudaFree(ctx->d_energy);
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  if (ctx->evx0) cudaEventDestroy(ctx->evx0);
  if (ctx->evx1) cudaEventDestroy(ctx->evx1);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

int rafft_create(rafft_ctx **out, int cuda_device, const rafft_energy_params *ep_in) {
  if (!out) { g_create_err = "rafft_create: out is NULL"; return RAFFT_ERR_ARG; }
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    g_create_err = std::string("rafft_create: no usable CUDA device (") +
                   (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0") +
                   "); this library has no CPU fallback";
    (void)cudaGetLastError();
    return RAFFT_ERR_CUDA;
  }
  if (cuda_device < 0 || cuda_device >= ndev) {
    g_create_err = "rafft_create: cuda_device out of range";
    return RAFFT_ERR_ARG;
  }
  rafft_energy_params *ep = (rafft_energy_params *)malloc(sizeof *ep);
  if (ep_in) *ep = *ep_in; else default_energy_params(ep);
  DevEnergy *D = (DevEnergy *)malloc(sizeof *D);
  std::string err;
  if (!build_dev_energy(ep, D, err)) {
    g_create_err = "rafft_create: " + err;
    free(ep); free(D);
    return RAFFT_ERR_ARG;
  }
  free(ep);
  rafft_ctx *ctx = new rafft_ctx;
  ctx->device = cuda_device;
  int rc = [&]() -> int {
    CK(cudaSetDevice(cuda_device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, cuda_device));
    ctx->sm_count = prop.multiProcessorCount;
    CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&ctx->ev0));
    CK(cudaEventCreate(&ctx->ev1));
    CK(cudaEventCreate(&ctx->evx0));
    CK(cudaEventCreate(&ctx->evx1));
    CK(cudaMalloc((void **)&ctx->d_energy, sizeof(DevEnergy)));
    CK(cudaMemcpy(ctx->d_energy, D, sizeof(DevEnergy), cudaMemcpyHostToDevice));
    CK(cudaHostAlloc((void **)&ctx->h_small, 64 * sizeof(int64_t), cudaHostAllocDefault));
    const char *s = getenv("RAFFT_BLOCK_THREADS");
    if (s) { int v = atoi(s); if (v >= 32 && v <= RF_BLOCK_THREADS_MAX && v % 32 == 0) ctx->block_threads = v; }
    s = getenv("RAFFT_SEEN_GEN0"); /* tests: start near the wrap of the `seen` generation counter */
    if (s) { int v = atoi(s); if (v >= 0 && v <= 65535) ctx->seen_gen = (uint32_t)v; }
    s = getenv("RAFFT_BLOCKS_PER_SM");
    if (s) { int v = atoi(s); if (v >= 1 && v <= 32) ctx->blocks_per_sm = v; }
    s = getenv("RAFFT_MAX_WAVE");
    if (s) { long v = atol(s); if (v >= 1) ctx->max_wave = v; }
    s = getenv("RAFFT_FAST_SHAPE");
    if (s) { int v = atoi(s); if (v >= 0 && v < (int)(sizeof g_fast_shapes / sizeof g_fast_shapes[0])) ctx->fast_shape = v; }
    s = getenv("RAFFT_FAST");
    if (s) ctx->use_fast = atoi(s) != 0;
    s = getenv("RAFFT_MAX_TASKS");
    if (s) { long v = atol(s); if (v >= 1) ctx->max_tasks = v; }
    return RAFFT_OK;
  }();
  free(D);
  if (rc != RAFFT_OK) {
    g_create_err = "rafft_create: " + ctx->err;
    rafft_destroy(ctx);
    return rc;
  }
  *out = ctx;
  return RAFFT_OK;
}

void rafft_result_free(rafft_result *res) {
  if (!res) return;
  ResultImpl *r = (ResultImpl *)res;
  if (r->block.p) r->ctx->pinned_free.push_back(r->block);
  delete r;
}

}  // extern "C"

/* Layout of the index part of OutBuf inside ctx->idx for n sequences. */
static size_t idx_layout(OutBuf &o, char *base, int n) {
  size_t off = 0;
  auto carve = [&](size_t bytes) { off = align_up(off, 16); char *p = base ? base + off : nullptr; off += bytes; return p; };
  o.cursors = (unsigned long long *)carve(8 * sizeof(unsigned long long));
  o.status = (int32_t *)carve(sizeof(int32_t) * (size_t)n);
  o.seq_struct = (int64_t *)carve(sizeof(int64_t) * (size_t)n);
  o.seq_nstruct = (int32_t *)carve(sizeof(int32_t) * (size_t)n);
  o.seq_db = (int64_t *)carve(sizeof(int64_t) * (size_t)n);
  o.seq_step = (int64_t *)carve(sizeof(int64_t) * (size_t)n);
  o.seq_nstep = (int32_t *)carve(sizeof(int32_t) * (size_t)n);
  o.seq_entry = (int64_t *)carve(sizeof(int64_t) * (size_t)n);
  o.counters = (int64_t *)carve(sizeof(int64_t) * (size_t)n * CNT_N);
  return align_up(off, 256);
}

static size_t big_layout(OutBuf &o, char *base) {
  size_t off = 0;
  auto carve = [&](size_t bytes) { off = align_up(off, 256); char *p = base ? base + off : nullptr; off += bytes; return p; };
  o.energy = (int32_t *)carve(sizeof(int32_t) * (size_t)o.struct_cap);
  o.db = (char *)carve((size_t)o.db_cap);
  o.step_n = (int32_t *)carve(sizeof(int32_t) * (size_t)o.step_cap);
  o.entries = (int32_t *)carve(sizeof(int32_t) * (size_t)o.entry_cap);
  return align_up(off, 256);
}

/* The fold proper, inputs on the device.  h_offs: host copy of the offsets (n+1). */
static int fold_device(rafft_ctx *ctx, const char *d_seqs, const int64_t *d_offs, const int64_t *h_offs,
                       int32_t n, const rafft_params *params, rafft_result **out_res,
                       int64_t h2d_bytes) {
  std::string err;
  const bool dbg = getenv("RAFFT_DEBUG_ALLOC") != nullptr;
  auto t_start = std::chrono::steady_clock::now();
  auto lap = [&](const char *what) {
    if (dbg) fprintf(stderr, "[rafft] %-18s +%.2f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count());
  };
  FoldParams F;
  if (!build_fold_params(params, &F, err)) { ctx->err = "rafft_fold_batch: " + err; return RAFFT_ERR_ARG; }
  int L_max = 1;
  int64_t total_len = h_offs[n] - h_offs[0];
  for (int i = 0; i < n; i++) {
    int64_t L = h_offs[i + 1] - h_offs[i];
    if (L < 0 || L > 4096) {
      ctx->err = "rafft_fold_batch: sequence length must be in [0, 4096]";
      return RAFFT_ERR_ARG;
    }
    if (L > L_max) L_max = (int)L;
  }
  const int T = ctx->block_threads;
  /* per-kernel overrides of the block size (tuning; same results for any value) */
  auto threads_of = [&](const char *name, int dflt) {
    const char *e = getenv(name);
    const int v = e ? atoi(e) : 0;
    return (v >= 32 && v <= RF_BLOCK_THREADS_MAX && v % 32 == 0) ? v : dflt;
  };
  /* measured on B200 (tools/threads_sweep.sh): the product enumeration wants 4 warps per sequence
   * (more look-ups in flight), selection and materialisation 2 warps up to -ms 100 and 4 beyond */
  const int T_ms = params->max_stack > 100 && T == 64 ? 128 : T;
  const int T_gather = threads_of("RAFFT_GATHER_THREADS", T), T_combine = threads_of("RAFFT_COMBINE_THREADS", T == 64 ? 128 : T);
  const int T_select = threads_of("RAFFT_SELECT_THREADS", T_ms), T_mat = threads_of("RAFFT_MATERIALISE_THREADS", T_ms);
  const int T_finish = threads_of("RAFFT_FINISH_THREADS", RF_BLOCK_THREADS_MAX);
  ResultImpl *R = new ResultImpl;
  memset(&R->pub, 0, sizeof R->pub);
  R->ctx = ctx;
  R->block = HostBlock();
  auto fail = [&](int rc) { rafft_result_free(&R->pub); return rc; };
  int64_t launches = 0, d2h = 0;
  float kernel_ms = 0.f;
  double expand_ms = 0.0;
  int64_t expand_launches = 0;

  /* index part of the output */
  OutBuf O;
  memset(&O, 0, sizeof O);
  size_t idx_bytes = idx_layout(O, nullptr, n);
  if (int rc = dev_reserve(ctx, ctx->idx, idx_bytes)) return fail(rc);
  idx_layout(O, (char *)ctx->idx.p, n);
  if (cudaMemsetAsync(ctx->idx.p, 0, idx_bytes, ctx->stream) != cudaSuccess) { ctx->err = "memset"; return fail(RAFFT_ERR_CUDA); }
  if (int rc = dev_reserve(ctx, ctx->order, sizeof(int32_t) * (size_t)(n + 1))) return fail(rc);

  std::vector<int32_t> todo;          /* sequences still to fold (empty = all) */
  std::vector<int32_t> h_status(n > 0 ? n : 1), h_order;
  int scale = 1;
  int64_t used[4] = {0, 0, 0, 0};      /* cursors after the previous round */
  for (int round = 0; n > 0; round++) {
    const int m = round == 0 ? n : (int)todo.size();
    int Lr = L_max;
    if (round > 0) {
      Lr = 1;
      for (int i : todo) Lr = std::max<int64_t>(Lr, h_offs[i + 1] - h_offs[i]);
    }
    Caps caps = make_caps(Lr, F, T, scale);
    /* output capacities for this round (bump-allocated after what is already used) */
    int64_t per_struct = F.traj ? caps.store_cap : F.max_stack;
    int64_t per_step = F.traj ? caps.step_cap : 1;
    int64_t sum_L = 0;
    if (round == 0) sum_L = total_len;
    else for (int i : todo) sum_L += h_offs[i + 1] - h_offs[i];
    OutBuf N = O;
    N.struct_cap = used[0] + per_struct * m;
    N.db_cap = used[1] + per_struct * (sum_L + 15 * (int64_t)m); /* rows padded to 16 bytes */
    N.step_cap = used[2] + per_step * m;
    N.entry_cap = used[3] + (F.traj ? per_step * F.max_stack : (int64_t)F.max_stack) * m;
    size_t big_bytes = big_layout(N, nullptr);
    if (round == 0) {
      if (int rc = dev_reserve(ctx, ctx->big, big_bytes)) return fail(rc);
      big_layout(N, (char *)ctx->big.p);
    } else {
      DevBuf nb;
      if (int rc = dev_reserve(ctx, nb, big_bytes)) return fail(rc);
      big_layout(N, (char *)nb.p);
      cudaMemcpyAsync(N.energy, O.energy, sizeof(int32_t) * (size_t)used[0], cudaMemcpyDeviceToDevice, ctx->stream);
      cudaMemcpyAsync(N.db, O.db, (size_t)used[1], cudaMemcpyDeviceToDevice, ctx->stream);
      cudaMemcpyAsync(N.step_n, O.step_n, sizeof(int32_t) * (size_t)used[2], cudaMemcpyDeviceToDevice, ctx->stream);
      cudaMemcpyAsync(N.entries, O.entries, sizeof(int32_t) * (size_t)used[3], cudaMemcpyDeviceToDevice, ctx->stream);
      CK(cudaStreamSynchronize(ctx->stream));
      cudaFree(ctx->big.p);
      ctx->big = nb;
    }
    O = N;
    /* waves: as many sequences at once as the memory budget allows (one arena each) */
    size_t free_b = 0, total_b = 0;
    CK(cudaMemGetInfo(&free_b, &total_b));
    int64_t budget = (int64_t)((free_b + ctx->arenas.cap + ctx->scratch.cap) * 0.7);
    int64_t wave = std::min<int64_t>(m, std::max<int64_t>(1, (budget * 3 / 4) / caps.arena_bytes));
    if (ctx->max_wave > 0) wave = std::min<int64_t>(wave, ctx->max_wave);
    int64_t qcap = std::min<int64_t>(wave * 64, std::max<int64_t>(64, (budget / 4) / caps.wscratch_bytes));
    /* regions of up to 512 positions take the shared-memory expansion (expand_fast.cuh); only
     * longer ones need a global scratch slot */
    FastCfg fastcfg;
    int64_t fcap = 0;
    if (ctx->use_fast && build_fast_cfg(fastcfg, Lr, F)) {
      fcap = wave * std::max<int64_t>(128, 4 * (int64_t)F.max_stack);
      qcap = std::min<int64_t>(qcap, wave * 8 + 1024);
    } else {
      memset(&fastcfg, 0, sizeof fastcfg);
    }
    if (ctx->max_tasks > 0) { qcap = std::min<int64_t>(qcap, ctx->max_tasks); fcap = std::min<int64_t>(fcap, ctx->max_tasks); }
    {
      /* The hash tables at the head of every arena must be empty before a run; a run leaves
       * them empty again, so they are zeroed only when the buffer or its layout changes. */
      void *before = ctx->arenas.p;
      if (int rc = dev_reserve(ctx, ctx->arenas, (size_t)(wave * caps.arena_bytes))) return fail(rc);
      uint64_t sig = 1469598103934665603ULL;
      const unsigned char *cb = (const unsigned char *)&caps;
      for (size_t k = 0; k < sizeof caps; k++) sig = (sig ^ cb[k]) * 1099511628211ULL;
      const size_t need = (size_t)(wave * caps.arena_bytes);
      if (before != ctx->arenas.p || sig != ctx->arena_sig || ctx->arena_dirty) {
        CK(cudaMemsetAsync(ctx->arenas.p, 0, need, ctx->stream));
        ctx->arena_sig = sig;
        ctx->arena_dirty = false;
        ctx->arena_clean = need;
      } else if (need > ctx->arena_clean) {
        CK(cudaMemsetAsync((char *)ctx->arenas.p + ctx->arena_clean, 0, need - ctx->arena_clean, ctx->stream));
        ctx->arena_clean = need;
      }
    }
    if (int rc = dev_reserve(ctx, ctx->states, sizeof(SeqState) * (size_t)wave)) return fail(rc);
    if (int rc = dev_reserve(ctx, ctx->tasks, sizeof(int32_t) * 2 * (size_t)qcap)) return fail(rc);
    if (int rc = dev_reserve(ctx, ctx->ftasks, sizeof(int32_t) * 2 * RF_FAST_LISTS * (size_t)(fcap + 1))) return fail(rc);
    if (int rc = dev_reserve(ctx, ctx->scratch, (size_t)qcap * (size_t)caps.wscratch_bytes)) return fail(rc);
    if (int rc = dev_reserve(ctx, ctx->small, 256)) return fail(rc);
    WaveCtx wc;
    memset(&wc, 0, sizeof wc);
    wc.E = ctx->d_energy; wc.prm = F; wc.caps = caps;
    arena_offsets(wc.aoff, caps, F);
    wc.arenas = (char *)ctx->arenas.p; wc.states = (SeqState *)ctx->states.p;
    wc.seqs = d_seqs; wc.offs = d_offs;
    wc.Q.n_tasks = (int32_t *)ctx->small.p; wc.Q.cap = (int32_t)qcap; wc.Q.task = (int32_t *)ctx->tasks.p;
    wc.Q.scratch = (char *)ctx->scratch.p; wc.Q.self = 0;
    wc.n_active = (int32_t *)ctx->small.p + 1;
    wc.Q.n_fast = (int32_t *)ctx->small.p + 2; wc.Q.fast_cap = (int32_t)fcap; wc.Q.fast_task = (int32_t *)ctx->ftasks.p;
    wc.Q.fast_lv = fastcfg.L_cap; wc.Q.fast_br = fastcfg.br_cap;
    wc.fast = fastcfg;
    wc.fast_next = (int32_t *)ctx->small.p + 2 + RF_FAST_LISTS;
    const FastShape &fsh = g_fast_shapes[ctx->fast_shape];
    const size_t fast_smem = (size_t)(fsh.threads / 32) * (size_t)fastcfg.smem_per_warp;
    /* a branch of a loop covers at least min_hp + 2 positions */
    const bool any_generic = fcap == 0 || Lr > fastcfg.L_cap || Lr / (F.min_hp + 2) > fastcfg.br_cap;
    wc.out = O;
    CombineStage cs_probe;
    const size_t stage_bytes = stage_carve(cs_probe, nullptr, caps.chunk_cap, F.max_stack);
    CK(cudaFuncSetAttribute(k_combine, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stage_bytes));
    {
      const size_t sel_bytes = sizeof(int32_t) * (size_t)std::max<int64_t>(caps.chunk_cap, (int64_t)caps.new_cap + F.max_stack);
      wc.sel_smem = sel_bytes <= 40 * 1024 ? (int32_t)sel_bytes : 0;
    }
    const int egrid = ctx->sm_count * ctx->blocks_per_sm;
    if (fcap > 0) CK(cudaFuncSetAttribute(fsh.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fast_smem));
    {
      /* longest-first processing order (ties: input order): counting sort over the lengths on
       * the host, which holds the offsets anyway */
      std::vector<int32_t> start(4098, 0);
      h_order.resize((size_t)m);
      for (int k = 0; k < m; k++) {
        const int si = round == 0 ? k : todo[k];
        start[4096 - (int)(h_offs[si + 1] - h_offs[si]) + 1]++;
      }
      for (int b = 0; b < 4097; b++) start[b + 1] += start[b];
      for (int k = 0; k < m; k++) {
        const int si = round == 0 ? k : todo[k];
        h_order[(size_t)start[4096 - (int)(h_offs[si + 1] - h_offs[si])]++] = si;
      }
      CK(cudaMemcpyAsync(ctx->order.p, h_order.data(), sizeof(int32_t) * (size_t)m, cudaMemcpyHostToDevice, ctx->stream));
      h2d_bytes += sizeof(int32_t) * (int64_t)m;
    }
    CK(cudaEventRecord(ctx->ev0, ctx->stream));
    for (int64_t w0 = 0; w0 < m; w0 += wave) {
      const int nw = (int)std::min<int64_t>(wave, m - w0);
      wc.order = (const int32_t *)ctx->order.p + w0;
      wc.n = nw;
      /* every fold of an arena tags its `seen` entries with a new generation; when the 16-bit
       * counter wraps the tables are zeroed once */
      if (++ctx->seen_gen > 65535u) {
        CK(cudaMemsetAsync(ctx->arenas.p, 0, ctx->arena_clean, ctx->stream));
        ctx->seen_gen = 1;
      }
      wc.prm.seen_gen = (int32_t)ctx->seen_gen;
      k_init<<<nw, T, 0, ctx->stream>>>(wc);
      launches++;
      for (int it = 0;; it++) {
        CK(cudaMemsetAsync(ctx->small.p, 0, 128, ctx->stream)); /* n_tasks, n_active, list counters, list cursors */
        k_gather<<<nw, T_gather, 0, ctx->stream>>>(wc);
        launches++;
        if (fcap > 0) {
          CK(cudaEventRecord(ctx->evx0, ctx->stream));
          fsh.fn<<<ctx->sm_count * fsh.blocks, fsh.threads, fast_smem, ctx->stream>>>(wc);
          CK(cudaEventRecord(ctx->evx1, ctx->stream));
          launches++;
        }
        if (any_generic) {
          k_expand<1><<<egrid, 128, 0, ctx->stream>>>(wc);
          k_expand<2><<<egrid, 128, 0, ctx->stream>>>(wc);
          k_expand<3><<<egrid, 128, 0, ctx->stream>>>(wc);
          k_expand<4><<<egrid, 128, 0, ctx->stream>>>(wc);
          k_expand<5><<<egrid, 128, 0, ctx->stream>>>(wc);
          launches += 5;
        }
        k_combine<<<nw, T_combine, stage_bytes, ctx->stream>>>(wc);
        k_select<<<nw, T_select, (size_t)wc.sel_smem, ctx->stream>>>(wc);
        k_materialise<<<nw, T_mat, 0, ctx->stream>>>(wc);
        launches += 3;
        CK(cudaMemcpyAsync(ctx->h_small + 16, (int32_t *)ctx->small.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        d2h += 8;
        if (fcap > 0) {
          float xms = 0.f;
          CK(cudaEventElapsedTime(&xms, ctx->evx0, ctx->evx1));
          expand_ms += xms;
          expand_launches++;
        }
        const int32_t *hs = (const int32_t *)(ctx->h_small + 16);
        if (hs[1] == 0) break;
        if (it > 100000) { ctx->err = "rafft_fold_batch: no progress"; return fail(RAFFT_ERR_CUDA); }
      }
      k_finish<<<nw, T_finish, 0, ctx->stream>>>(wc);
      launches++;
      if (dbg) { /* high-water marks of the arena pieces against their capacities */
        std::vector<SeqState> hs((size_t)nw);
        cudaMemcpyAsync(hs.data(), ctx->states.p, sizeof(SeqState) * (size_t)nw, cudaMemcpyDeviceToHost, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
        int hw[7] = {0, 0, 0, 0, 0, 0, 0};
        for (const SeqState &q : hs) {
          const int v[7] = {q.n_node, q.n_pool16, q.n_cand, q.n_store, q.n_nlist, q.n_seen, q.n_steps};
          for (int k = 0; k < 7; k++) hw[k] = std::max(hw[k], v[k]);
        }
        fprintf(stderr, "[rafft] high water (cap): node %d (%d) pool16 %d (%d) cand %d (%d) store %d (%d) nlist %d (%d) seen %d (%d) steps %d (%d); arena %.1f MB, scale %d\n",
                hw[0], caps.node_cap, hw[1], caps.pool16_cap, hw[2], caps.cand_cap, hw[3], caps.store_cap, hw[4], caps.nlist_cap,
                hw[5], caps.seen_cap, hw[6], caps.step_cap, caps.arena_bytes / 1048576.0, scale);
      }
    }
    CK(cudaEventRecord(ctx->ev1, ctx->stream));
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(h_status.data(), O.status, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(ctx->h_small, O.cursors, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    d2h += sizeof(int32_t) * (int64_t)n + 64;
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    kernel_ms += ms;
    for (int k = 0; k < 4; k++) used[k] = ctx->h_small[k];
    todo.clear();
    for (int i = 0; i < n; i++)
      if (h_status[i] == ST_ERR_CAPACITY) todo.push_back(i);
    if (dbg && !todo.empty()) {
      std::vector<int64_t> hc((size_t)n * CNT_N);
      cudaMemcpy(hc.data(), O.counters, sizeof(int64_t) * hc.size(), cudaMemcpyDeviceToHost);
      fprintf(stderr, "[rafft] round %d: %zu sequences over capacity (scale %d):", round, todo.size(), scale);
      for (size_t k = 0; k < todo.size() && k < 12; k++)
        fprintf(stderr, " #%d L=%lld line %lld;", todo[k], (long long)(h_offs[todo[k] + 1] - h_offs[todo[k]]), (long long)hc[(size_t)todo[k] * CNT_N + CNT_T1_NS]);
      fprintf(stderr, "\n");
    }
    if (todo.empty() || round >= 5) break;
    scale *= 4;
  }

  lap("fold rounds done");
  /* ordered layout on the device, then one copy per array into pinned host memory */
  int64_t *d_scan = nullptr;
  size_t scan_bytes = sizeof(int64_t) * (4 * (size_t)(n + 1) + 8);
  if (int rc = dev_reserve(ctx, ctx->scan, scan_bytes)) return fail(rc);
  d_scan = (int64_t *)ctx->scan.p;
  int64_t *d_struct_off = d_scan, *d_db_off = d_scan + (n + 1), *d_step_off = d_scan + 2 * (size_t)(n + 1),
          *d_ent_base = d_scan + 3 * (size_t)(n + 1), *d_totals = d_scan + 4 * (size_t)(n + 1);
  CK(cudaEventRecord(ctx->ev0, ctx->stream));
  scan_kernel<<<1, 1024, 0, ctx->stream>>>(d_offs, n, O, d_struct_off, d_db_off, d_step_off, d_ent_base, d_totals);
  launches++;
  CK(cudaMemcpyAsync(ctx->h_small, d_totals, 4 * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  d2h += 32;
  const int64_t nS = ctx->h_small[0], nD = ctx->h_small[1], nT = ctx->h_small[2], nE = ctx->h_small[3];
  /* device ordered buffer: energy | entry_struct | entry_off | db */
  size_t o_energy = 0, o_entry = align_up(o_energy + sizeof(int32_t) * (size_t)nS, 256),
         o_entoff = align_up(o_entry + sizeof(int32_t) * (size_t)nE, 256),
         o_db = align_up(o_entoff + sizeof(int64_t) * (size_t)(nT + 1), 256),
         o_end = align_up(o_db + (size_t)nD, 256);
  if (int rc = dev_reserve(ctx, ctx->ordered, o_end)) return fail(rc);
  char *dord = (char *)ctx->ordered.p;
  if (n > 0) {
    pack_kernel<<<n, 256, 0, ctx->stream>>>(d_offs, n, O, d_struct_off, d_db_off, d_step_off, d_ent_base,
                                           (int32_t *)(dord + o_energy), dord + o_db,
                                           (int64_t *)(dord + o_entoff), (int32_t *)(dord + o_entry));
    launches++;
  } else {
    CK(cudaMemsetAsync(dord + o_entoff, 0, sizeof(int64_t), ctx->stream));
  }
  CK(cudaEventRecord(ctx->ev1, ctx->stream));
  CK(cudaGetLastError());
  /* host block: [ordered part as on the device] status | struct_off | db_off | step_off | counters */
  size_t h_status_o = o_end, h_soff = align_up(h_status_o + sizeof(int32_t) * (size_t)n, 16),
         h_doff = h_soff + sizeof(int64_t) * (size_t)(n + 1), h_toff = h_doff + sizeof(int64_t) * (size_t)(n + 1),
         h_cnt = h_toff + sizeof(int64_t) * (size_t)(n + 1), h_end = h_cnt + sizeof(int64_t) * (size_t)n * CNT_N;
  lap("scan+pack queued");
  if (int rc = host_block_get(ctx, h_end + 64, &R->block)) return fail(rc);
  lap("host block");
  char *hb = (char *)R->block.p;
  CK(cudaMemcpyAsync(hb, dord, o_end, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(hb + h_soff, d_scan, sizeof(int64_t) * 3 * (size_t)(n + 1), cudaMemcpyDeviceToHost, ctx->stream));
  if (n > 0) CK(cudaMemcpyAsync(hb + h_cnt, O.counters, sizeof(int64_t) * (size_t)n * CNT_N, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  d2h += (int64_t)o_end + (int64_t)sizeof(int64_t) * (3 * (n + 1) + (int64_t)n * CNT_N);
  float ms2 = 0.f;
  CK(cudaEventElapsedTime(&ms2, ctx->ev0, ctx->ev1));
  kernel_ms += ms2;
  lap("d2h done");
  memcpy(hb + h_status_o, h_status.data(), sizeof(int32_t) * (size_t)n);
  rafft_result &P = R->pub;
  P.n_seq = n;
  P.status = (const int32_t *)(hb + h_status_o);
  P.struct_off = (const int64_t *)(hb + h_soff);
  P.db_off = (const int64_t *)(hb + h_doff);
  P.step_off = (const int64_t *)(hb + h_toff);
  P.energy_dcal = (const int32_t *)(hb + o_energy);
  P.db = hb + o_db;
  P.entry_off = (const int64_t *)(hb + o_entoff);
  P.entry_struct = (const int32_t *)(hb + o_entry);
  P.counters = (const int64_t *)(hb + h_cnt);
  P.kernel_ms = kernel_ms;
  P.kernel_launches = launches;
  P.h2d_bytes = h2d_bytes;
  P.d2h_bytes = d2h;
  P.expand_ms = expand_ms;
  P.expand_launches = expand_launches;
  *out_res = &R->pub;
  return RAFFT_OK;
}

extern "C" {

int rafft_fold_batch(rafft_ctx *ctx, const char *seqs, const int64_t *offs, int32_t n,
                     const rafft_params *params, rafft_result **out) {
  if (!ctx) return RAFFT_ERR_ARG;
  if (!out || !params || n < 0 || (n > 0 && (!seqs || !offs))) { ctx->err = "rafft_fold_batch: bad argument"; return RAFFT_ERR_ARG; }
  *out = nullptr;
  CK(cudaSetDevice(ctx->device));
  static const int64_t zero_offs[1] = {0};
  if (n == 0) offs = zero_offs;
  for (int i = 0; i < n; i++)
    if (offs[i + 1] < offs[i]) { ctx->err = "rafft_fold_batch: offsets must be non-decreasing"; return RAFFT_ERR_ARG; }
  const int64_t base = offs[0], total = offs[n] - base;
  if (int rc = dev_reserve(ctx, ctx->seqs, (size_t)total + 16)) return rc;
  if (int rc = dev_reserve(ctx, ctx->offs, sizeof(int64_t) * (size_t)(n + 1))) return rc;
  std::vector<int64_t> rel(n + 1);
  for (int i = 0; i <= n; i++) rel[i] = offs[i] - base;
  if (total > 0) CK(cudaMemcpyAsync(ctx->seqs.p, seqs + base, (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(ctx->offs.p, rel.data(), sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return fold_device(ctx, (const char *)ctx->seqs.p, (const int64_t *)ctx->offs.p, rel.data(), n, params, out,
                     total + (int64_t)sizeof(int64_t) * (n + 1));
}

int rafft_fold_batch_device(rafft_ctx *ctx, const void *d_seqs, const void *d_offs, int32_t n,
                            int64_t total_len, int32_t max_len, const rafft_params *params,
                            rafft_result **out) {
  if (!ctx) return RAFFT_ERR_ARG;
  if (!out || !params || n < 0 || (n > 0 && (!d_seqs || !d_offs))) { ctx->err = "rafft_fold_batch_device: bad argument"; return RAFFT_ERR_ARG; }
  (void)total_len; (void)max_len;
  *out = nullptr;
  CK(cudaSetDevice(ctx->device));
  /* the offsets (8(n+1) bytes) are needed on the host to size the arenas and the result */
  std::vector<int64_t> h_offs(n + 1, 0);
  if (n > 0) CK(cudaMemcpy(h_offs.data(), d_offs, sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyDeviceToHost));
  if (n > 0 && h_offs[0] != 0) { ctx->err = "rafft_fold_batch_device: offs[0] must be 0"; return RAFFT_ERR_ARG; }
  for (int i = 0; i < n; i++)
    if (h_offs[i + 1] < h_offs[i]) { ctx->err = "rafft_fold_batch_device: offsets must be non-decreasing"; return RAFFT_ERR_ARG; }
  return fold_device(ctx, (const char *)d_seqs, (const int64_t *)d_offs, h_offs.data(), n, params, out, 0);
}

int rafft_eval_structures(rafft_ctx *ctx, const char *seqs, const char *structs, const int64_t *offs,
                          int32_t n, int32_t *energy_dcal) {
  if (!ctx) return RAFFT_ERR_ARG;
  if (n < 0 || (n > 0 && (!seqs || !structs || !offs || !energy_dcal))) { ctx->err = "rafft_eval_structures: bad argument"; return RAFFT_ERR_ARG; }
  if (n == 0) return RAFFT_OK;
  CK(cudaSetDevice(ctx->device));
  const int64_t base = offs[0], total = offs[n] - base;
  std::vector<int64_t> rel(n + 1);
  for (int i = 0; i <= n; i++) {
    rel[i] = offs[i] - base;
    if (i && rel[i] < rel[i - 1]) { ctx->err = "rafft_eval_structures: offsets must be non-decreasing"; return RAFFT_ERR_ARG; }
    if (i && rel[i] - rel[i - 1] > 4096) { ctx->err = "rafft_eval_structures: sequence length must be <= 4096"; return RAFFT_ERR_ARG; }
  }
  /* seqs | dbs | codes | offs | scratch(2 ints per letter) | energies */
  size_t o_seq = 0, o_db = align_up((size_t)total, 256), o_code = o_db + align_up((size_t)total, 256),
         o_offs = o_code + align_up((size_t)total, 256), o_scr = align_up(o_offs + sizeof(int64_t) * (size_t)(n + 1), 256),
         o_e = align_up(o_scr + sizeof(int32_t) * 2 * (size_t)total, 256), o_end = o_e + sizeof(int32_t) * (size_t)n;
  if (int rc = dev_reserve(ctx, ctx->evalbuf, o_end)) return rc;
  char *d = (char *)ctx->evalbuf.p;
  if (total > 0) {
    CK(cudaMemcpyAsync(d + o_seq, seqs + base, (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d + o_db, structs + base, (size_t)total, cudaMemcpyHostToDevice, ctx->stream));
  }
  CK(cudaMemcpyAsync(d + o_offs, rel.data(), sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyHostToDevice, ctx->stream));
  const int wpb = 4;
  eval_kernel<<<(n + wpb - 1) / wpb, wpb * 32, 0, ctx->stream>>>(ctx->d_energy, d + o_seq, d + o_db,
                                                               (const int64_t *)(d + o_offs), n, (int32_t *)(d + o_scr),
                                                               (uint8_t *)(d + o_code), (int32_t *)(d + o_e));
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(energy_dcal, d + o_e, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  return RAFFT_OK;
}

int rafft_kin_graph(rafft_ctx *ctx, const char *db, int32_t L, const int64_t *step_off, int32_t n_steps,
                    int32_t *uniq, uint8_t *conn, int64_t conn_cap, int64_t *conn_off) {
  if (!ctx) return RAFFT_ERR_ARG;
  if (n_steps < 0 || L < 1 || L > 32767 || !step_off || !conn_off || (n_steps > 0 && (!db || !uniq))) {
    ctx->err = "rafft_kin_graph: bad argument";
    return RAFFT_ERR_ARG;
  }
  for (int t = 0; t < n_steps; t++)
    if (step_off[t + 1] < step_off[t] || step_off[0] != 0) { ctx->err = "rafft_kin_graph: bad step offsets"; return RAFFT_ERR_ARG; }
  const int64_t n = n_steps ? step_off[n_steps] : 0;
  conn_off[0] = 0;
  for (int t = 0; t < n_steps; t++) {
    const int tp = t == 0 ? n_steps - 1 : t - 1;
    conn_off[t + 1] = conn_off[t] + (step_off[t + 1] - step_off[t]) * (step_off[tp + 1] - step_off[tp]);
  }
  const int64_t total = conn_off[n_steps];
  if (n == 0) return RAFFT_OK;
  if (n > 0x7fffffff / 2) { ctx->err = "rafft_kin_graph: too many structures"; return RAFFT_ERR_ARG; }
  if (total > conn_cap || (total > 0 && !conn)) { ctx->err = "rafft_kin_graph: conn buffer too small"; return RAFFT_ERR_ARG; }
  CK(cudaSetDevice(ctx->device));
  /* db | pt | stacks | hash | bad | uniq | step_off | conn_off | conn */
  size_t o_db = 0, o_pt = align_up((size_t)n * L, 256), o_st = align_up(o_pt + sizeof(int16_t) * (size_t)n * L, 256),
         o_hash = align_up(o_st + sizeof(int16_t) * 2 * (size_t)n * L, 256), o_bad = o_hash + sizeof(uint64_t) * (size_t)n,
         o_uniq = o_bad + sizeof(int32_t) * (size_t)n, o_so = align_up(o_uniq + sizeof(int32_t) * (size_t)n, 256),
         o_co = o_so + sizeof(int64_t) * (size_t)(n_steps + 1), o_conn = align_up(o_co + sizeof(int64_t) * (size_t)(n_steps + 1), 256),
         o_end = o_conn + (size_t)total + 16;
  if (int rc = dev_reserve(ctx, ctx->kinbuf, o_end)) return rc;
  char *d = (char *)ctx->kinbuf.p;
  CK(cudaMemcpyAsync(d + o_db, db, (size_t)n * L, cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d + o_so, step_off, sizeof(int64_t) * (size_t)(n_steps + 1), cudaMemcpyHostToDevice, ctx->stream));
  CK(cudaMemcpyAsync(d + o_co, conn_off, sizeof(int64_t) * (size_t)(n_steps + 1), cudaMemcpyHostToDevice, ctx->stream));
  const int wpb = 4;
  kin_pairtable_kernel<<<(unsigned)((n + wpb - 1) / wpb), wpb * 32, 0, ctx->stream>>>(
      d + o_db, (int)n, L, (int16_t *)(d + o_pt), (int16_t *)(d + o_st), (uint64_t *)(d + o_hash), (int32_t *)(d + o_bad));
  kin_unique_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(d + o_db, (const uint64_t *)(d + o_hash), (int)n, L,
                                                                         (int32_t *)(d + o_uniq));
  if (total > 0) {
    const int64_t blocks = (total + wpb - 1) / wpb;
    if (blocks > 0x7fffffff) { ctx->err = "rafft_kin_graph: graph too large"; return RAFFT_ERR_ARG; }
    kin_connect_kernel<<<(unsigned)blocks, wpb * 32, 0, ctx->stream>>>((const int16_t *)(d + o_pt), (const int64_t *)(d + o_so),
                                                                     (const int64_t *)(d + o_co), n_steps, L,
                                                                     (uint8_t *)(d + o_conn), total);
  }
  CK(cudaGetLastError());
  std::vector<int32_t> bad((size_t)n);
  CK(cudaMemcpyAsync(bad.data(), d + o_bad, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(uniq, d + o_uniq, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  if (total > 0) CK(cudaMemcpyAsync(conn, d + o_conn, (size_t)total, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  for (int64_t i = 0; i < n; i++)
    if (bad[(size_t)i]) { ctx->err = "rafft_kin_graph: closing bracket without partner in structure " + std::to_string(i); return RAFFT_ERR_SEQ; }
  return RAFFT_OK;
}

int rafft_bp_distance_matrix(rafft_ctx *ctx, const char *db, int32_t n, int32_t L, int32_t *dist) {
  if (!ctx) return RAFFT_ERR_ARG;
  if (n < 0 || L < 1 || L > 32767 || (n > 0 && (!db || !dist))) { ctx->err = "rafft_bp_distance_matrix: bad argument"; return RAFFT_ERR_ARG; }
  if (n == 0) return RAFFT_OK;
  if (n > 46340) { ctx->err = "rafft_bp_distance_matrix: too many structures"; return RAFFT_ERR_ARG; }
  CK(cudaSetDevice(ctx->device));
  /* db | pt | stacks | hash | bad | dist */
  size_t o_db = 0, o_pt = align_up((size_t)n * L, 256), o_st = align_up(o_pt + sizeof(int16_t) * (size_t)n * L, 256),
         o_hash = align_up(o_st + sizeof(int16_t) * 2 * (size_t)n * L, 256), o_bad = o_hash + sizeof(uint64_t) * (size_t)n,
         o_dist = align_up(o_bad + sizeof(int32_t) * (size_t)n, 256), o_end = o_dist + sizeof(int32_t) * (size_t)n * n;
  if (int rc = dev_reserve(ctx, ctx->kinbuf, o_end)) return rc;
  char *d = (char *)ctx->kinbuf.p;
  CK(cudaMemcpyAsync(d + o_db, db, (size_t)n * L, cudaMemcpyHostToDevice, ctx->stream));
  const int wpb = 4;
  kin_pairtable_kernel<<<(unsigned)((n + wpb - 1) / wpb), wpb * 32, 0, ctx->stream>>>(
      d + o_db, (int)n, L, (int16_t *)(d + o_pt), (int16_t *)(d + o_st), (uint64_t *)(d + o_hash), (int32_t *)(d + o_bad), 1);
  const int nt = (n + KIN_DTILE - 1) / KIN_DTILE;
  kin_bpdist_kernel<<<(unsigned)(nt * (nt + 1) / 2), 256, 0, ctx->stream>>>((const int16_t *)(d + o_pt), n, L, (int32_t *)(d + o_dist));
  CK(cudaGetLastError());
  std::vector<int32_t> bad((size_t)n);
  CK(cudaMemcpyAsync(bad.data(), d + o_bad, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaMemcpyAsync(dist, d + o_dist, sizeof(int32_t) * (size_t)n * n, cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  for (int i = 0; i < n; i++)
    if (bad[(size_t)i]) { ctx->err = "rafft_bp_distance_matrix: closing bracket without partner in structure " + std::to_string(i); return RAFFT_ERR_SEQ; }
  return RAFFT_OK;
}

}  // extern "C"
