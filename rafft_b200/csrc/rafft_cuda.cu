/*
 * rafft_cuda.cu -- the C-ABI library (include/rafft_b200.h) over the sm_100a kernels.
 *
 * Kernels in this file:
 *   k_ctl_reset / k_init / k_gather / k_expand_fast / k_expand_big / k_expand<1..5> / k_combine /
 *   k_select / k_materialise / (k_beam) / k_advance / k_finish
 *                  the level-synchronous phase kernels of the fold (fold_core.cuh, expand_fast.cuh,
 *                  expand_big.cuh).
 *                  One folding iteration of every sequence of a wave is the chain
 *                  gather -> expand -> combine -> select -> materialise -> advance; the chain is the
 *                  body of a CUDA-graph WHILE node whose condition k_advance sets on the device, so a
 *                  whole fold is ONE graph launch with no host round trip.  The
 *                  block-scope kernels are persistent: a fixed grid pulls the still-active sequences
 *                  from a compacted list, so late iterations cost what their few live sequences cost.
 *   scan_kernel    exclusive scans that turn per-sequence counts into the ordered layout
 *   pack_kernel    gathers the bump-allocated results into input order
 *   eval_kernel    whole-structure loop-decomposition energies (rafft_eval_structures)
 *   eval_static_kernel  the same with its scratch in a file-scope device array (rafft_eval_structures_static:
 *                  no allocation inside the call)
 *   int_peak_kernel  IADD3/LOP3 issue-rate microbenchmark (rafft_int_peak: the roofline denominator)
 *
 * There is no CPU fallback: every entry point fails with RAFFT_ERR_CUDA when no device is
 * usable.  Nothing here includes or links anything under oracle/.
 */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include "host_common.h"
#include "expand_fast.cuh"
#include "expand_big.cuh"
#include "kin_cuda.cuh"

using namespace rafft;

#define RF_BLOCK_THREADS_MAX 256

/* ------------------------------------------------------------------------------------ */
/* kernels                                                                               */

/* Device-resident control block of one fold loop (one per lane). */
struct LoopCtl {
  int32_t n_tasks;                   /* generic region work list of this iteration */
  int32_t n_next;                    /* sequences appended to the next active list */
  int32_t n_fast[RF_FAST_LISTS];     /* shared-memory expansion work lists */
  int32_t fast_next[RF_FAST_LISTS];  /* ... and their dynamic cursors */
  int32_t n_cur;                     /* entries of the current active list */
  int32_t parity;                    /* current active list = alist[parity] */
  int32_t iters;                     /* folding iterations run */
  int32_t seen_gen;                  /* generation tag of this fold's `seen` entries */
  int32_t work[4];                   /* dynamic cursors of gather / combine / select / materialise */
  int32_t fast_total[RF_FAST_LISTS]; /* regions per work list over the whole fold (diagnostics) */
  int32_t generic_total;
  int32_t n_big, big_next, big_total; /* work list of the block-per-region expansion */
  int32_t pad_[12];
};

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
  BigCfg big;
  LoopCtl *ctl;
  int32_t *alist;     /* two lists of acap wave slots: the sequences still folding; behind them acap flags
                         (slot still folding after this step?) written by k_materialise, read by k_advance */
  int32_t acap;
  int32_t sel_smem;   /* bytes of shared memory for the energies under selection (0: global) */
  int32_t fuse;       /* 1: select + materialise + the next step's gather run as one kernel (k_beam) */
  int32_t *h_prog;    /* host-mapped: [0] iterations finished, [1] 1 once no sequence is active */
  OutBuf out;
};

#define RF_WAVE_PARAM const __grid_constant__ WaveCtx c

/* block-scope kernels: state of wave slot j staged through shared memory */
struct BlockBind {
  __device__ __forceinline__ static bool load(const WaveCtx &c, Fold &f, SeqState *sh, int j, bool need_active) {
    SeqState *st = c.states + j;
    if (need_active && !st->active) return false;
    for (int k = threadIdx.x; k < (int)(sizeof(SeqState) / 4); k += blockDim.x) ((int *)sh)[k] = ((const int *)st)[k];
    arena_bind(f.A, c.arenas + (size_t)j * (size_t)c.caps.arena_bytes, c.aoff);
    f.sh = sh;
    f.Q = c.Q;
    f.Q.self = j;
    f.gen = c.ctl->seen_gen;
    __syncthreads();
    f.bind_ev();
    return true;
  }
  __device__ __forceinline__ static void store(const WaveCtx &c, SeqState *sh, int j) {
    __syncthreads();
    SeqState *st = c.states + j;
    for (int k = threadIdx.x; k < (int)(sizeof(SeqState) / 4); k += blockDim.x) ((int *)st)[k] = ((const int *)sh)[k];
  }
};

/* Persistent block-scope kernels: every block pulls the next still-active wave slot from the
 * compacted list of this iteration (dynamic cursor `which` of the control block). */
#define RF_FOR_ACTIVE(which)                                                         \
  __shared__ int s_slot_;                                                             \
  const int n_cur_ = c.ctl->n_cur;                                                    \
  const int32_t *alist_ = c.alist + (size_t)c.ctl->parity * (size_t)c.acap;           \
  for (;;) {                                                                          \
    __syncthreads();                                                                  \
    if (threadIdx.x == 0) {                                                           \
      const int k_ = atomicAdd(&c.ctl->work[which], 1);                               \
      s_slot_ = k_ < n_cur_ ? alist_[k_] : -1;                                        \
    }                                                                                 \
    __syncthreads();                                                                  \
    const int j = s_slot_;                                                            \
    if (j < 0) break;
#define RF_END_ACTIVE }

/* start of a fold: the control block (one block of 64 threads; plain launch, its arguments
 * change from fold to fold while the graph of the loop is cached) */
__global__ void k_ctl_reset(LoopCtl *ctl, int nw, int gen) {
  const int t = threadIdx.x;
  if (t < RF_FAST_LISTS) { ctl->n_fast[t] = 0; ctl->fast_next[t] = 0; ctl->fast_total[t] = 0; }
  if (t < 4) ctl->work[t] = 0;
  if (t == 0) {
    ctl->generic_total = 0; ctl->n_big = 0; ctl->big_next = 0; ctl->big_total = 0;
    ctl->n_tasks = 0; ctl->n_next = 0; ctl->n_cur = nw; ctl->parity = 0; ctl->iters = 0; ctl->seen_gen = gen;
  }
}

__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_init(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  Fold f(c.E, c.prm, c.caps);
  const int j = blockIdx.x;
  arena_bind(f.A, c.arenas + (size_t)j * (size_t)c.caps.arena_bytes, c.aoff);
  f.sh = &sh;
  const int si = c.order[j];
  const int64_t o = c.offs[si];
  if (threadIdx.x == 0) { sh.seq_off = o; c.alist[j] = j; }
  f.phase_init(c.seqs + o, (int)(c.offs[si + 1] - o), si);
  if (c.fuse) { /* the first folding step's gather: the root region goes on a work list */
    f.Q = c.Q;
    f.Q.self = j;
    __syncthreads();
    f.phase_gather();
  }
  BlockBind::store(c, &sh, j);
}

__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_gather(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  RF_FOR_ACTIVE(0)
    Fold f(c.E, c.prm, c.caps);
    if (!BlockBind::load(c, f, &sh, j, true)) continue;
    f.phase_gather();
    BlockBind::store(c, &sh, j);
  RF_END_ACTIVE
}

/* the expansion work lists of this step are consumed once the expansion kernels have ended; with
 * the fused chain the NEXT step's gather fills them again inside k_beam, so they are cleared here */
__device__ __forceinline__ void clear_work_lists(LoopCtl *ctl) {
  const int t = threadIdx.x;
  if (t < RF_FAST_LISTS) { ctl->fast_total[t] += ctl->n_fast[t]; ctl->n_fast[t] = 0; ctl->fast_next[t] = 0; }
  if (t == 0) {
    ctl->generic_total += ctl->n_tasks; ctl->n_tasks = 0;
    ctl->big_total += ctl->n_big; ctl->n_big = 0; ctl->big_next = 0;
  }
}

__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_combine(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  extern __shared__ __align__(16) char combine_smem[];
  if (c.fuse && blockIdx.x == 0) clear_work_lists(c.ctl);
  RF_FOR_ACTIVE(1)
    Fold f(c.E, c.prm, c.caps);
    if (!BlockBind::load(c, f, &sh, j, true)) continue;
    stage_carve(f.cs, combine_smem, c.caps.chunk_cap, c.prm.max_stack);
    f.phase_combine();
    BlockBind::store(c, &sh, j);
  RF_END_ACTIVE
}

__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_select(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  extern __shared__ __align__(16) char select_smem[];
  RF_FOR_ACTIVE(2)
    Fold f(c.E, c.prm, c.caps);
    if (!BlockBind::load(c, f, &sh, j, true)) continue;
    if (c.sel_smem) f.A.sel_e = (int32_t *)select_smem;
    f.phase_select();
    BlockBind::store(c, &sh, j);
  RF_END_ACTIVE
}

__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_materialise(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  int32_t *aflag = c.alist + 2 * (size_t)c.acap;
  RF_FOR_ACTIVE(3)
    Fold f(c.E, c.prm, c.caps);
    if (!BlockBind::load(c, f, &sh, j, true)) { if (threadIdx.x == 0) aflag[j] = 0; continue; }
    f.phase_materialise();
    if (threadIdx.x == 0) aflag[j] = sh.active; /* k_advance builds the next list from these */
    BlockBind::store(c, &sh, j);
  RF_END_ACTIVE
}

/* select + materialise of this folding step and the gather of the next one, per sequence: two
 * launches and two round trips of the sequence state less per step than the separate kernels */
__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_beam(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  extern __shared__ __align__(16) char select_smem[];
  int32_t *aflag = c.alist + 2 * (size_t)c.acap;
  RF_FOR_ACTIVE(2)
    Fold f(c.E, c.prm, c.caps);
    if (!BlockBind::load(c, f, &sh, j, true)) { if (threadIdx.x == 0) aflag[j] = 0; continue; }
    if (c.sel_smem) f.A.sel_e = (int32_t *)select_smem;
    f.phase_select();
    __syncthreads();
    f.phase_materialise();
    __syncthreads();
    f.phase_gather();
    if (threadIdx.x == 0) aflag[j] = sh.active;
    BlockBind::store(c, &sh, j);
  RF_END_ACTIVE
}

/* end of a folding iteration (one block of RF_ADVANCE_THREADS threads): the sequences of the current
 * active list that are still folding are compacted IN ORDER into the other list -- the initial order
 * is longest first, and keeping it means the persistent kernels of every later step still start their
 * heaviest sequences first (an atomic append would leave them in completion order: heaviest last) --
 * then the work lists and cursors are cleared and the loop condition is set (graph WHILE node) /
 * published to the host (stream mode) */
#define RF_ADVANCE_THREADS 1024
__global__ void __launch_bounds__(RF_ADVANCE_THREADS) k_advance(RF_WAVE_PARAM, cudaGraphConditionalHandle handle, int use_handle) {
  __shared__ int s_warp[RF_ADVANCE_THREADS / 32];
  LoopCtl *ctl = c.ctl;
  const int t = threadIdx.x, T = blockDim.x;
  const int n_cur = ctl->n_cur, parity = ctl->parity;
  const int32_t *cur = c.alist + (size_t)parity * (size_t)c.acap;
  int32_t *nxt = c.alist + (size_t)(parity ^ 1) * (size_t)c.acap;
  const int per = (n_cur + T - 1) / T;
  const int lo = min(n_cur, t * per), hi = min(n_cur, lo + per);
  const int32_t *aflag = c.alist + 2 * (size_t)c.acap; /* dense, and cur[] is ascending: sequential reads */
  int cnt = 0;
  for (int k = lo; k < hi; k++) cnt += aflag[cur[k]] != 0;
  int inc = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, inc, d);
    if ((t & 31) >= d) inc += v;
  }
  if ((t & 31) == 31) s_warp[t >> 5] = inc;
  __syncthreads();
  int base = 0, total = 0;
  for (int w = 0; w < (T >> 5); w++) {
    const int v = s_warp[w];
    if (w < (t >> 5)) base += v;
    total += v;
  }
  int at = base + inc - cnt;
  for (int k = lo; k < hi; k++) {
    const int j = cur[k];
    if (aflag[j]) nxt[at++] = j;
  }
  if (!c.fuse) clear_work_lists(ctl);
  if (t < 4) ctl->work[t] = 0;
  if (t == 0) {
    int n = total;
    const int it = ++ctl->iters;
    /* no fold needs more iterations than folding steps fit its trajectory plus the rounds a full work list
     * can defer: beyond that the loop is stopped (the device-side WHILE has no host to give up for it) and
     * the host reports the call as failed */
    const bool runaway = n > 0 && it >= 8 * c.caps.step_cap + 4096;
    if (runaway) n = 0;
    ctl->n_cur = n; ctl->n_next = 0; ctl->parity = parity ^ 1;
    if (use_handle) cudaGraphSetConditional(handle, n > 0 ? 1u : 0u);
    if (n == 0) c.h_prog[1] = runaway ? 2 : 1;
    __threadfence_system();
    c.h_prog[0] = it;
  }
}

__global__ void __launch_bounds__(RF_BLOCK_THREADS_MAX) k_finish(RF_WAVE_PARAM) {
  __shared__ SeqState sh;
  Fold f(c.E, c.prm, c.caps);
  BlockBind::load(c, f, &sh, blockIdx.x, false);
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
  if (nt == 0) return; /* an empty class costs nothing (late steps: most are; the count is the same for every thread) */
  const int32_t *list = c.Q.fast_task + 2 * (size_t)cls * (size_t)c.Q.fast_cap;
  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) *s_base = atomicAdd(c.ctl->fast_next + cls, nwarps);
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

/* block-per-region expansion of long regions (expand_big.cuh): persistent blocks pull regions */
__global__ void __launch_bounds__(RF_BIG_THREADS, 2) k_expand_big(RF_WAVE_PARAM) {
  extern __shared__ __align__(16) char big_smem[];
  __shared__ int s_task;
  const int nt = min(c.ctl->n_big, c.Q.big_cap);
  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) s_task = atomicAdd(&c.ctl->big_next, 1);
    __syncthreads();
    const int t = s_task;
    if (t >= nt) break;
    const int j = c.Q.big_task[2 * t], v = c.Q.big_task[2 * t + 1];
    big_expand(c.E, c.prm, c.caps, c.big, c.arenas + (size_t)j * (size_t)c.caps.arena_bytes, c.aoff, c.states + j, v, big_smem);
  }
}

/* launch shapes of k_expand_fast: (threads per block, blocks per SM) */
typedef void (*fast_kernel_t)(const WaveCtx);
struct FastShape { int threads, blocks; fast_kernel_t fn; };
static const FastShape g_fast_shapes[] = {
    {128, 6, k_expand_fast<128, 6>}, {256, 2, k_expand_fast<256, 2>}, {256, 3, k_expand_fast<256, 3>},
    {256, 4, k_expand_fast<256, 4>}};

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
                            int64_t *__restrict__ entry_off, int32_t *__restrict__ entry_struct,
                            int64_t bS, int64_t bD, int64_t bT, int64_t bE, int64_t *__restrict__ g_off) {
  /* bS/bD/bT/bE: what the lanes before this one hold (structures, dot-bracket bytes, steps, entries):
   * the arrays written here are lane-local, the OFFSETS stored in them index the whole result;
   * g_off = struct_off | db_off | step_off (n + 1 each) of this lane's sequences with the bases added */
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
    int64_t run = e_dst + bE;
    for (int s = 0; s < nt; s++) { entry_off[t_dst + s] = run; run += out.step_n[t_src + s]; }
    if (i == n - 1) entry_off[t_dst + nt] = run;
    g_off[i] = s_dst + bS; g_off[(n + 1) + i] = d_dst + bD; g_off[2 * (n + 1) + i] = t_dst + bT;
    if (i == n - 1) {
      g_off[n] = struct_off[n] + bS; g_off[(n + 1) + n] = db_off[n] + bD; g_off[2 * (n + 1) + n] = step_off[n] + bT;
    }
  }
}

/* INT32 issue-rate microbenchmark (rafft_int_peak): 8 independent chains per thread, IADD3 and LOP3
 * alternating; every resident warp can issue every cycle, so the achieved warp-instructions per
 * second are the issue-slot peak the fold kernels are measured against. */
__global__ void __launch_bounds__(256) int_peak_kernel(unsigned *out, int iters, unsigned seed, int mode) {
  unsigned a[8];
#pragma unroll
  for (int j = 0; j < 8; j++) a[j] = seed + threadIdx.x * 8 + j;
  const unsigned b = seed * 3 + 1, c = seed ^ 0x55aa;
  if (mode == 0) {
    for (int i = 0; i < iters; i++) {
#pragma unroll
      for (int u = 0; u < 16; u++) {
#pragma unroll
        for (int j = 0; j < 8; j++) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[j]) : "r"(b));
      }
    }
  } else {
    for (int i = 0; i < iters; i++) {
#pragma unroll
      for (int u = 0; u < 16; u++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
          if (u & 1) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[j]) : "r"(b));
          else asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[j]) : "r"(b), "r"(c));
        }
      }
    }
  }
  unsigned x = 0;
#pragma unroll
  for (int j = 0; j < 8; j++) x ^= a[j];
  if (x == 0x12345678u) out[0] = x;
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

/* ---- rafft_eval_structures_static: the evaluator without any allocation ------------------------
 * Scratch of eval_kernel (pair table + stack: 2 ints per letter, base codes: 1 byte per letter) in a
 * file-scope array of a fixed maximum size instead of a cudaMalloc'ed buffer.  Every GPU that loads
 * this library holds the array (9 MiB) until the process exits.  All calls on one device share it
 * and it keeps the contents of the last call; the kernel writes every element before reading it
 * (pt[k] = -1 as lane 0 walks the structure, st[] pushed before popped, code[] filled first), so
 * nothing has to start at zero -- but two calls that overlap on the GPU race on it.
 * In host code the NAME of the array is a host-side stand-in, not a device address: the kernel gets
 * the address that cudaGetSymbolAddress returned with the context's device current (static_scratch). */
#define RF_STATIC_EVAL_MAX_LETTERS (1 << 20)
static_assert(RF_STATIC_EVAL_MAX_LETTERS == RAFFT_STATIC_EVAL_MAX_LETTERS, "header and library disagree");
static __device__ __align__(256) unsigned char g_static_eval_scratch[(size_t)RF_STATIC_EVAL_MAX_LETTERS * 9 + 256];

/* eval_kernel with its scratch inside `scratch_base` (pt/st at byte 0, codes at 8 * max letters) and
 * every offset checked on the device: a sequence whose letters do not lie inside [0, total_len], or
 * is longer than 4096, gets INT32_MIN and touches nothing. */
__global__ void eval_static_kernel(const DevEnergy *__restrict__ E, const char *__restrict__ seqs,
                                   const char *__restrict__ dbs, const int64_t *__restrict__ offs, int n,
                                   int64_t total_len, unsigned char *__restrict__ scratch_base,
                                   int32_t *__restrict__ energy) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n) return;
  const int64_t o = offs[warp], o1 = offs[warp + 1];
  if (o < 0 || o1 < o || o1 > total_len || o1 > (int64_t)RF_STATIC_EVAL_MAX_LETTERS || o1 - o > 4096) {
    if (lane == 0) energy[warp] = INT32_MIN;
    return;
  }
  const int L = (int)(o1 - o);
  const char *seq = seqs + o, *db = dbs + o;
  int32_t *pt = (int32_t *)scratch_base + 2 * o, *st = pt + L;
  uint8_t *code = scratch_base + (size_t)8 * RF_STATIC_EVAL_MAX_LETTERS + o;
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

#define RF_MAX_LANES 2      /* concurrent sub-batches per device */
#define RF_MAX_DEVICES 16   /* contexts of one rafft_fold_batch_multi call */
#define RF_GRAPH_SLOTS 4
#define RF_PROF_EVENTS 12

enum { FAM_GATHER = 0, FAM_EXPAND_FAST, FAM_EXPAND_GENERIC, FAM_COMBINE, FAM_SELECT, FAM_MATERIALISE,
       FAM_ADVANCE, FAM_OTHER, FAM_N };

/* the launch graph of one wave's fold (init -> WHILE{iteration} -> finish), cached by its kernel
 * arguments: a repeated call with the same shapes re-launches the instantiated graph */
struct GraphSlot {
  cudaGraphExec_t exec = nullptr;
  cudaGraph_t graph = nullptr;
  std::vector<unsigned char> key;
  uint64_t stamp = 0;
};

/* One lane = one stream with its own arenas, work lists, control block and output staging: an
 * independent fold of a contiguous range of the batch.  A context runs up to RF_MAX_LANES lanes
 * concurrently (lane 0 on a high-priority stream: it finishes first and its result copy overlaps
 * the fold of lane 1). */
struct Lane {
  cudaStream_t stream = nullptr;
  cudaEvent_t ev1 = nullptr;
  cudaEvent_t pev[RF_PROF_EVENTS] = {};
  DevBuf arenas, order, idx, big, ordered, scan, states, tasks, ftasks, btasks, scratch, ctl, alist;
  int64_t *h_small = nullptr;  /* pinned + mapped: [0..7] totals/cursors, [8..] loop progress words */
  uint64_t arena_sig = 0;      /* layout of the arenas whose hash tables are known to be empty */
  bool arena_dirty = true;
  uint32_t seen_gen = 0;       /* generation of the last fold in the arenas (tag of its `seen` entries) */
  size_t arena_clean = 0;      /* leading bytes of the arena buffer known to be clean */
  GraphSlot graphs[RF_GRAPH_SLOTS];
  uint64_t graph_clock = 0;
  std::string err;
};

struct rafft_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;   /* = lane[0].stream (evaluator / kinetics entry points) */
  cudaEvent_t ev_job = nullptr;
  DevEnergy *d_energy = nullptr;
  Lane lane[RF_MAX_LANES];
  DevBuf seqs, offs, evalbuf, kinbuf, peakbuf;
  std::vector<HostBlock> pinned_free;
  std::mutex pin_mu;
  int block_threads = 64;   /* threads per block of the block-scope phase kernels */
  int blocks_per_sm = 16;   /* grid of the warp-task kernels: blocks per SM */
  int64_t max_wave = 0;     /* sequences folded together (0 = as many as memory allows) */
  int64_t max_tasks = 0;    /* capacity of the region work list (0 = automatic) */
  int fast_shape = 2;       /* index into g_fast_shapes (RAFFT_FAST_SHAPE) */
  int use_fast = 1;         /* shared-memory expansion for small regions (RAFFT_FAST=0 disables) */
  int use_big = 1;          /* block-per-region shared-memory expansion for long regions (RAFFT_BIG=0 disables) */
  int loop_mode = 0;        /* 0 graph WHILE node, 1 stream launches with bounded look-ahead */
  int lanes = 0;            /* 0 automatic, 1..RF_MAX_LANES forced */
  int profile = 0;          /* per-kernel-family CUDA events (stepwise loop, one lane) */
  int persist = 1;          /* blocks per SM of the persistent block-scope kernels = occupancy * persist */
  int fuse = 0;             /* 1: select + materialise + next gather as one kernel, k_beam (RAFFT_FUSE=1; measured: -2 % at 1,250
                               sequences, +2 % at 10,000: profiles/r02_fuse.md) */
  int lane_split = 50;      /* per cent of the estimated cost that goes to lane 0 when a call runs two lanes */
  int64_t mem_budget = 0;   /* bytes this context may use for arenas + scratch */
  double traj_ratio = 5.0;  /* structures per (sequence x max_stack) of the last traj = 1 call */
  uint32_t gen0 = 0;
  unsigned char *static_scratch = nullptr; /* device address of g_static_eval_scratch on this context's device */
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
/* the same inside a lane: the message goes to the lane (lanes run on their own host threads) */
#define LCK(call)                                                                      \
  do {                                                                                 \
    cudaError_t e_ = (call);                                                           \
    if (e_ != cudaSuccess) {                                                           \
      ln.err = std::string(#call) + ": " + cudaGetErrorString(e_);                     \
      return RAFFT_ERR_CUDA;                                                           \
    }                                                                                  \
  } while (0)

static int dev_reserve_s(std::string &err, DevBuf &b, size_t bytes) {
  if (bytes <= b.cap) return RAFFT_OK;
  if (getenv("RAFFT_DEBUG_ALLOC")) fprintf(stderr, "[rafft] cudaMalloc %zu bytes (had %zu)\n", bytes, b.cap);
  if (b.p) { cudaFree(b.p); b.p = nullptr; b.cap = 0; }
  size_t want = bytes + bytes / 8 + 256;
  cudaError_t e = cudaMalloc(&b.p, want);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    want = bytes;
    e = cudaMalloc(&b.p, want);
  }
  if (e != cudaSuccess) {
    b.p = nullptr;
    err = std::string("cudaMalloc(") + std::to_string(bytes) + "): " + cudaGetErrorString(e);
    return RAFFT_ERR_CUDA;
  }
  b.cap = want;
  return RAFFT_OK;
}
static int dev_reserve(rafft_ctx *ctx, DevBuf &b, size_t bytes) { return dev_reserve_s(ctx->err, b, bytes); }

static int host_block_get(rafft_ctx *ctx, size_t bytes, HostBlock *hb) {
  std::lock_guard<std::mutex> g(ctx->pin_mu);
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
  /* portable: the lanes of other devices copy straight into it (rafft_fold_batch_multi) */
  CK(cudaHostAlloc(&hb->p, want, cudaHostAllocPortable));
  hb->cap = want;
  return RAFFT_OK;
}

static void lane_destroy(Lane &ln) {
  DevBuf *bufs[] = {&ln.arenas, &ln.order, &ln.idx, &ln.big, &ln.ordered, &ln.scan, &ln.states,
                    &ln.tasks, &ln.ftasks, &ln.btasks, &ln.scratch, &ln.ctl, &ln.alist};
  for (DevBuf *b : bufs)
    if (b->p) cudaFree(b->p);
  for (GraphSlot &g : ln.graphs) {
    if (g.exec) cudaGraphExecDestroy(g.exec);
    if (g.graph) cudaGraphDestroy(g.graph);
  }
  if (ln.h_small) cudaFreeHost(ln.h_small);
  if (ln.ev1) cudaEventDestroy(ln.ev1);
  for (cudaEvent_t e : ln.pev)
    if (e) cudaEventDestroy(e);
  if (ln.stream) cudaStreamDestroy(ln.stream);
}

extern "C" {

const char *rafft_version(void) { return "rafft_b200 0.2 sm_100a"; }

int rafft_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
  return n;
}

void rafft_default_energy_params(rafft_energy_params *ep) {
  if (ep) default_energy_params(ep);
}

const char *rafft_last_error(const rafft_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

void rafft_destroy(rafft_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  for (Lane &ln : ctx->lane) lane_destroy(ln);
  DevBuf *bufs[] = {&ctx->seqs, &ctx->offs, &ctx->evalbuf, &ctx->kinbuf, &ctx->peakbuf};
  for (DevBuf *b : bufs)
    if (b->p) cudaFree(b->p);
  for (auto &b : ctx->pinned_free) cudaFreeHost(b.p);
  if (ctx->d_energy) cudaFree(ctx->d_energy);
  if (ctx->ev_job) cudaEventDestroy(ctx->ev_job);
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
    int prio_lo = 0, prio_hi = 0;
    CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi)); /* numerically lower = higher priority */
    for (int k = 0; k < RF_MAX_LANES; k++) {
      Lane &ln = ctx->lane[k];
      CK(cudaStreamCreateWithPriority(&ln.stream, cudaStreamNonBlocking, k == 0 ? prio_hi : prio_lo));
      CK(cudaEventCreate(&ln.ev1));
      for (cudaEvent_t &pe : ln.pev) CK(cudaEventCreate(&pe));
      CK(cudaHostAlloc((void **)&ln.h_small, 64 * sizeof(int64_t), cudaHostAllocMapped | cudaHostAllocPortable));
      memset(ln.h_small, 0, 64 * sizeof(int64_t));
    }
    ctx->stream = ctx->lane[0].stream;
    CK(cudaEventCreate(&ctx->ev_job));
    CK(cudaMalloc((void **)&ctx->d_energy, sizeof(DevEnergy)));
    CK(cudaMemcpy(ctx->d_energy, D, sizeof(DevEnergy), cudaMemcpyHostToDevice));
    size_t free_b = 0, total_b = 0;
    CK(cudaMemGetInfo(&free_b, &total_b));
    ctx->mem_budget = (int64_t)(free_b * 0.7);
    const char *s = getenv("RAFFT_BLOCK_THREADS");
    if (s) { int v = atoi(s); if (v >= 32 && v <= RF_BLOCK_THREADS_MAX && v % 32 == 0) ctx->block_threads = v; }
    s = getenv("RAFFT_SEEN_GEN0"); /* tests: start near the wrap of the `seen` generation counter */
    if (s) { int v = atoi(s); if (v >= 0 && v <= 65535) for (Lane &ln : ctx->lane) ln.seen_gen = (uint32_t)v; }
    s = getenv("RAFFT_BLOCKS_PER_SM");
    if (s) { int v = atoi(s); if (v >= 1 && v <= 32) ctx->blocks_per_sm = v; }
    s = getenv("RAFFT_MAX_WAVE");
    if (s) { long v = atol(s); if (v >= 1) ctx->max_wave = v; }
    s = getenv("RAFFT_FAST_SHAPE");
    if (s) { int v = atoi(s); if (v >= 0 && v < (int)(sizeof g_fast_shapes / sizeof g_fast_shapes[0])) ctx->fast_shape = v; }
    s = getenv("RAFFT_FAST");
    if (s) ctx->use_fast = atoi(s) != 0;
    s = getenv("RAFFT_BIG");
    if (s) ctx->use_big = atoi(s) != 0;
    s = getenv("RAFFT_MAX_TASKS");
    if (s) { long v = atol(s); if (v >= 1) ctx->max_tasks = v; }
    s = getenv("RAFFT_LOOP"); /* graph (default) | stream */
    if (s) ctx->loop_mode = strcmp(s, "stream") == 0 ? 1 : 0;
    s = getenv("RAFFT_LANES");
    if (s) { int v = atoi(s); if (v >= 0 && v <= RF_MAX_LANES) ctx->lanes = v; }
    s = getenv("RAFFT_PROFILE");
    if (s) ctx->profile = atoi(s) != 0;
    s = getenv("RAFFT_FUSE");
    if (s) ctx->fuse = atoi(s) != 0;
    s = getenv("RAFFT_LANE_SPLIT");
    if (s) { int v = atoi(s); if (v >= 5 && v <= 95) ctx->lane_split = v; }
    s = getenv("RAFFT_PERSIST");
    if (s) { int v = atoi(s); if (v >= 1 && v <= 8) ctx->persist = v; }
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

int rafft_set_option(rafft_ctx *ctx, const char *name, int64_t value) {
  if (!ctx || !name) return RAFFT_ERR_ARG;
  std::string k(name);
  if (k == "profile") ctx->profile = value != 0;
  else if (k == "lanes" && value >= 0 && value <= RF_MAX_LANES) ctx->lanes = (int)value;
  else if (k == "loop" && (value == 0 || value == 1)) ctx->loop_mode = (int)value;
  else if (k == "max_wave" && value >= 0) ctx->max_wave = value;
  else if (k == "persist" && value >= 1 && value <= 8) ctx->persist = (int)value;
  else if (k == "lane_split" && value >= 5 && value <= 95) ctx->lane_split = (int)value;
  else if (k == "fuse" && (value == 0 || value == 1)) ctx->fuse = (int)value;
  else { ctx->err = "rafft_set_option: unknown option or value out of range: " + k; return RAFFT_ERR_ARG; }
  return RAFFT_OK;
}

void rafft_result_free(rafft_result *res) {
  if (!res) return;
  ResultImpl *r = (ResultImpl *)res;
  if (r->block.p) {
    std::lock_guard<std::mutex> g(r->ctx->pin_mu);
    r->ctx->pinned_free.push_back(r->block);
  }
  delete r;
}

}  // extern "C"

/* Layout of the index part of OutBuf inside Lane::idx for n sequences. */
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

/* ---- one call = one job; its lanes (streams, possibly on several devices) fold contiguous ranges
 * of the batch and copy their part of the result into ONE pinned host block -------------------- */
struct LaneTask {
  rafft_ctx *ctx = nullptr;
  int lane = 0;          /* lane of ctx */
  int i0 = 0, i1 = 0;    /* its sequences [i0, i1) of the job */
  const char *d_seqs = nullptr;
  const int64_t *d_offs = nullptr; /* n + 1 offsets of the whole job on this lane's device */
  int first_of_ctx = 0;  /* records / owns ev_job of its context */
  /* results */
  int rc = RAFFT_OK;
  int64_t totals[4] = {0, 0, 0, 0};
  int64_t launches = 0, d2h = 0, h2d = 0, iters = 0;
  float span_ms = 0.f;   /* ev_job -> end of this lane's pack */
  double fam_ms[FAM_N] = {0, 0, 0, 0, 0, 0, 0, 0};
  int64_t fam_n[FAM_N] = {0, 0, 0, 0, 0, 0, 0, 0};
  std::vector<int32_t> h_status;
  bool copied = false;   /* its sections are in the job's host block */
  int waves = 0;
  OutBuf O;              /* device output of the fold (between the two halves of the lane) */
};

struct FoldJob {
  int n = 0;
  const int64_t *h_offs = nullptr;
  const rafft_params *params = nullptr;
  FoldParams F;
  std::vector<LaneTask> lanes;
  std::mutex mu;
  std::condition_variable cv;
  int published = 0;                 /* lanes 0 .. published-1 have added their totals to cum */
  int64_t cum[RF_MAX_LANES * RF_MAX_DEVICES + 1][4];
  bool abort = false;
  /* host block: energy | entry_struct | entry_off | db | status | struct_off | db_off | step_off | counters */
  HostBlock block;
  int64_t cap[4] = {0, 0, 0, 0};     /* capacities of the four data sections (structures, db bytes, steps, entries) */
  size_t o_energy = 0, o_entry = 0, o_entoff = 0, o_db = 0, o_status = 0, o_soff = 0, o_doff = 0, o_toff = 0, o_cnt = 0, o_end = 0;
  bool overflow = false;
};

static void job_layout(FoldJob &J) {
  const size_t n = (size_t)J.n;
  J.o_energy = 0;
  J.o_entry = align_up(J.o_energy + sizeof(int32_t) * (size_t)J.cap[0], 256);
  J.o_entoff = align_up(J.o_entry + sizeof(int32_t) * (size_t)J.cap[3], 256);
  J.o_db = align_up(J.o_entoff + sizeof(int64_t) * (size_t)(J.cap[2] + 1), 256);
  J.o_status = align_up(J.o_db + (size_t)J.cap[1], 256);
  J.o_soff = align_up(J.o_status + sizeof(int32_t) * n, 16);
  J.o_doff = J.o_soff + sizeof(int64_t) * (n + 1);
  J.o_toff = J.o_doff + sizeof(int64_t) * (n + 1);
  J.o_cnt = J.o_toff + sizeof(int64_t) * (n + 1);
  J.o_end = J.o_cnt + sizeof(int64_t) * n * CNT_N + 64;
}

/* kernel launch either on a stream or as a node of a graph (same arguments) */
struct Emit {
  cudaStream_t stream = nullptr;
  cudaGraph_t graph = nullptr;
  cudaGraphNode_t prev = nullptr;
  bool have_prev = false;
  cudaError_t err = cudaSuccess;
  void kernel(const void *fn, dim3 grid, dim3 block, size_t smem, void **args) {
    if (err != cudaSuccess) return;
    if (!graph) { err = cudaLaunchKernel(fn, grid, block, args, smem, stream); return; }
    cudaKernelNodeParams kp;
    memset(&kp, 0, sizeof kp);
    kp.func = (void *)fn; kp.gridDim = grid; kp.blockDim = block; kp.sharedMemBytes = (unsigned)smem;
    kp.kernelParams = args; kp.extra = nullptr;
    cudaGraphNode_t node;
    err = cudaGraphAddKernelNode(&node, graph, have_prev ? &prev : nullptr, have_prev ? 1 : 0, &kp);
    prev = node; have_prev = true;
  }
};

/* launch shape of one wave's fold */
struct WavePlan {
  WaveCtx wc;
  int nw = 0;
  int T = 64, T_gather = 64, T_combine = 128, T_select = 64, T_mat = 64, T_finish = 256;
  int g_gather = 0, g_combine = 0, g_select = 0, g_mat = 0, g_fast = 0, egrid = 0;
  size_t stage_bytes = 0, fast_smem = 0;
  const FastShape *fsh = nullptr;
  bool use_fast = false, any_generic = false, use_big = false;
  int g_big = 0;
  size_t big_smem = 0;
};

/* the kernels of one folding iteration; fam_ev != nullptr: a CUDA event before every kernel family
 * (profile mode, stream only) */
static void emit_iteration(Emit &E, WavePlan &P, cudaGraphConditionalHandle handle, int use_handle,
                           cudaEvent_t *fam_ev) {
  WaveCtx &wc = P.wc;
  void *a_wc[] = {&wc};
  int ne = 0;
  auto mark = [&]() { if (fam_ev && E.err == cudaSuccess) E.err = cudaEventRecord(fam_ev[ne++], E.stream); };
  mark();
  if (!wc.fuse) E.kernel((const void *)k_gather, dim3(P.g_gather), dim3(P.T_gather), 0, a_wc);
  mark();
  if (P.use_fast) E.kernel((const void *)P.fsh->fn, dim3(P.g_fast), dim3(P.fsh->threads), P.fast_smem, a_wc);
  if (P.use_big) E.kernel((const void *)k_expand_big, dim3(P.g_big), dim3(RF_BIG_THREADS), P.big_smem, a_wc);
  mark();
  if (P.any_generic) {
    E.kernel((const void *)k_expand<1>, dim3(P.egrid), dim3(128), 0, a_wc);
    E.kernel((const void *)k_expand<2>, dim3(P.egrid), dim3(128), 0, a_wc);
    E.kernel((const void *)k_expand<3>, dim3(P.egrid), dim3(128), 0, a_wc);
    E.kernel((const void *)k_expand<4>, dim3(P.egrid), dim3(128), 0, a_wc);
    E.kernel((const void *)k_expand<5>, dim3(P.egrid), dim3(128), 0, a_wc);
  }
  mark();
  E.kernel((const void *)k_combine, dim3(P.g_combine), dim3(P.T_combine), P.stage_bytes, a_wc);
  mark();
  if (!wc.fuse) E.kernel((const void *)k_select, dim3(P.g_select), dim3(P.T_select), (size_t)wc.sel_smem, a_wc);
  mark();
  if (wc.fuse) E.kernel((const void *)k_beam, dim3(P.g_mat), dim3(P.T_mat), (size_t)wc.sel_smem, a_wc);
  else E.kernel((const void *)k_materialise, dim3(P.g_mat), dim3(P.T_mat), 0, a_wc);
  mark();
  void *a_adv[] = {&wc, &handle, &use_handle};
  E.kernel((const void *)k_advance, dim3(1), dim3(RF_ADVANCE_THREADS), 0, a_adv);
  mark();
}

/* Graph of one wave's fold: k_init -> WHILE { iteration } -> k_finish.  Returns nullptr (and leaves
 * `why`) when the graph cannot be built; the caller then drives the same kernels from the stream. */
static cudaGraphExec_t wave_graph(Lane &ln, WavePlan &P, std::string &why) {
  /* key: every kernel argument and launch dimension */
  std::vector<unsigned char> key(sizeof(WaveCtx) + 16 * sizeof(int) + 4 * sizeof(size_t));
  {
    unsigned char *k = key.data();
    memcpy(k, &P.wc, sizeof(WaveCtx)); k += sizeof(WaveCtx);
    const int dims[16] = {P.nw, P.T, P.T_gather, P.T_combine, P.T_select, P.T_mat, P.T_finish, P.g_gather,
                          P.g_combine, P.g_select, P.g_mat, P.g_fast, P.egrid, P.use_fast, P.any_generic,
                          P.fsh ? P.fsh->threads : 0};
    memcpy(k, dims, sizeof dims); k += sizeof dims;
    memcpy(k, &P.stage_bytes, sizeof(size_t)); k += sizeof(size_t);
    memcpy(k, &P.fast_smem, sizeof(size_t)); k += sizeof(size_t);
    memcpy(k, &P.big_smem, sizeof(size_t)); k += sizeof(size_t);
    const size_t bigdims = (size_t)P.g_big * 2 + (P.use_big ? 1 : 0);
    memcpy(k, &bigdims, sizeof(size_t));
  }
  GraphSlot *victim = &ln.graphs[0];
  for (GraphSlot &g : ln.graphs) {
    if (g.exec && g.key == key) { g.stamp = ++ln.graph_clock; return g.exec; }
    if (g.stamp < victim->stamp) victim = &g;
  }
  if (victim->exec) { cudaGraphExecDestroy(victim->exec); victim->exec = nullptr; }
  if (victim->graph) { cudaGraphDestroy(victim->graph); victim->graph = nullptr; }
  victim->key.clear();
  cudaGraph_t g = nullptr;
  auto fail = [&](const char *what, cudaError_t e) {
    why = std::string(what) + ": " + cudaGetErrorString(e);
    (void)cudaGetLastError();
    if (g) cudaGraphDestroy(g);
    return (cudaGraphExec_t) nullptr;
  };
  cudaError_t e = cudaGraphCreate(&g, 0);
  if (e != cudaSuccess) return fail("cudaGraphCreate", e);
  Emit top;
  top.graph = g;
  void *a_wc[] = {&P.wc};
  top.kernel((const void *)k_init, dim3(P.nw), dim3(P.T), 0, a_wc);
  if (top.err != cudaSuccess) return fail("k_init node", top.err);
  cudaGraphConditionalHandle handle;
  e = cudaGraphConditionalHandleCreate(&handle, g, 1, cudaGraphCondAssignDefault);
  if (e != cudaSuccess) return fail("cudaGraphConditionalHandleCreate", e);
  cudaGraphNodeParams cp = {};
  cp.type = cudaGraphNodeTypeConditional;
  cp.conditional.handle = handle;
  cp.conditional.type = cudaGraphCondTypeWhile;
  cp.conditional.size = 1;
  cudaGraphNode_t wnode;
  e = cudaGraphAddNode(&wnode, g, &top.prev, 1, &cp);
  if (e != cudaSuccess) return fail("conditional WHILE node", e);
  Emit body;
  body.graph = cp.conditional.phGraph_out[0];
  emit_iteration(body, P, handle, 1, nullptr);
  if (body.err != cudaSuccess) return fail("iteration nodes", body.err);
  top.prev = wnode;
  top.kernel((const void *)k_finish, dim3(P.nw), dim3(P.T_finish), 0, a_wc);
  if (top.err != cudaSuccess) return fail("k_finish node", top.err);
  cudaGraphExec_t exec = nullptr;
  e = cudaGraphInstantiate(&exec, g, 0);
  if (e != cudaSuccess) return fail("cudaGraphInstantiate", e);
  victim->exec = exec; victim->graph = g; victim->key = key; victim->stamp = ++ln.graph_clock;
  return exec;
}

/* Fold the sequences [i0, i1) of the job on one lane and copy its part of the result into the
 * job's host block.  Runs on its own host thread when the job has several lanes. */
static int lane_fold(FoldJob &J, int k) {  /* first half: the fold itself, up to the totals of its result */
  LaneTask &T = J.lanes[(size_t)k];
  rafft_ctx *ctx = T.ctx;
  Lane &ln = ctx->lane[T.lane];
  const FoldParams &F = J.F;
  const int m_all = T.i1 - T.i0;
  const int64_t *h_offs = J.h_offs + T.i0;   /* offsets of this lane's sequences */
  const int64_t *d_offs = T.d_offs + T.i0;
  const bool dbg = getenv("RAFFT_DEBUG_ALLOC") != nullptr;
  const bool profile = ctx->profile != 0;
  LCK(cudaSetDevice(ctx->device));
  LCK(cudaStreamWaitEvent(ln.stream, ctx->ev_job, 0));
  int L_max = 1;
  int64_t total_len = h_offs[m_all] - h_offs[0];
  for (int i = 0; i < m_all; i++) L_max = std::max<int64_t>(L_max, h_offs[i + 1] - h_offs[i]);
  const int Tb = ctx->block_threads;
  auto threads_of = [&](const char *name, int dflt) {
    const char *e = getenv(name);
    const int v = e ? atoi(e) : 0;
    return (v >= 32 && v <= RF_BLOCK_THREADS_MAX && v % 32 == 0) ? v : dflt;
  };
  WavePlan P;
  P.T = Tb;
  /* block sizes of the beam kernels: set per round below, once the wave size is known */
  P.T_gather = threads_of("RAFFT_GATHER_THREADS", Tb);
  P.T_finish = threads_of("RAFFT_FINISH_THREADS", RF_BLOCK_THREADS_MAX);

  /* index part of the output */
  OutBuf O;
  memset(&O, 0, sizeof O);
  const size_t idx_bytes = idx_layout(O, nullptr, m_all);
  if (int rc = dev_reserve_s(ln.err, ln.idx, idx_bytes)) return rc;
  idx_layout(O, (char *)ln.idx.p, m_all);
  LCK(cudaMemsetAsync(ln.idx.p, 0, idx_bytes, ln.stream));
  if (int rc = dev_reserve_s(ln.err, ln.order, sizeof(int32_t) * (size_t)(m_all + 1))) return rc;
  if (int rc = dev_reserve_s(ln.err, ln.ctl, sizeof(LoopCtl))) return rc;
  int32_t *h_prog = (int32_t *)(ln.h_small + 8);

  std::vector<int32_t> todo, h_order;
  T.h_status.assign((size_t)(m_all > 0 ? m_all : 1), 0);
  int scale = 1;
  int64_t used[4] = {0, 0, 0, 0};
  const int n_lanes_ctx = std::max(1, (int)std::count_if(J.lanes.begin(), J.lanes.end(), [&](const LaneTask &t) { return t.ctx == ctx; }));
  for (int round = 0; m_all > 0; round++) {
    const int m = round == 0 ? m_all : (int)todo.size();
    int Lr = L_max;
    if (round > 0) {
      Lr = 1;
      for (int i : todo) Lr = std::max<int64_t>(Lr, h_offs[i + 1] - h_offs[i]);
    }
    Caps caps = make_caps(Lr, F, Tb, scale);
    /* output capacities for this round (bump-allocated after what is already used) */
    const int64_t per_struct = F.traj ? caps.store_cap : F.max_stack;
    const int64_t per_step = F.traj ? caps.step_cap : 1;
    int64_t sum_L = 0;
    if (round == 0) sum_L = total_len;
    else for (int i : todo) sum_L += h_offs[i + 1] - h_offs[i];
    OutBuf N = O;
    N.struct_cap = used[0] + per_struct * m;
    N.db_cap = used[1] + per_struct * (sum_L + 15 * (int64_t)m); /* rows padded to 16 bytes */
    N.step_cap = used[2] + per_step * m;
    N.entry_cap = used[3] + (F.traj ? per_step * F.max_stack : (int64_t)F.max_stack) * m;
    const size_t big_bytes = big_layout(N, nullptr);
    if (round == 0) {
      if (int rc = dev_reserve_s(ln.err, ln.big, big_bytes)) return rc;
      big_layout(N, (char *)ln.big.p);
    } else {
      DevBuf nb;
      if (int rc = dev_reserve_s(ln.err, nb, big_bytes)) return rc;
      big_layout(N, (char *)nb.p);
      cudaMemcpyAsync(N.energy, O.energy, sizeof(int32_t) * (size_t)used[0], cudaMemcpyDeviceToDevice, ln.stream);
      cudaMemcpyAsync(N.db, O.db, (size_t)used[1], cudaMemcpyDeviceToDevice, ln.stream);
      cudaMemcpyAsync(N.step_n, O.step_n, sizeof(int32_t) * (size_t)used[2], cudaMemcpyDeviceToDevice, ln.stream);
      cudaMemcpyAsync(N.entries, O.entries, sizeof(int32_t) * (size_t)used[3], cudaMemcpyDeviceToDevice, ln.stream);
      const cudaError_t se = cudaStreamSynchronize(ln.stream);
      cudaFree(ln.big.p);
      ln.big = nb; /* the lane owns the new buffer from here on (freed with the lane) */
      if (se != cudaSuccess) { ln.err = std::string("result regrowth: ") + cudaGetErrorString(se); return RAFFT_ERR_CUDA; }
    }
    O = N;
    /* waves: as many sequences at once as the memory budget allows (one arena each) */
    const int64_t budget = ctx->mem_budget / n_lanes_ctx;
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
    /* regions beyond the warp-per-region path take the block-per-region one (expand_big.cuh) */
    BigCfg bigcfg;
    int64_t bcap = 0;
    if (ctx->use_big && fcap > 0 && Lr > fastcfg.L_cap && build_big_cfg(bigcfg, Lr, F))
      bcap = wave * std::min<int64_t>(2 * (int64_t)F.max_stack + 8, 512);
    else
      memset(&bigcfg, 0, sizeof bigcfg);
    if (ctx->max_tasks > 0) { qcap = std::min<int64_t>(qcap, ctx->max_tasks); fcap = std::min<int64_t>(fcap, ctx->max_tasks); bcap = std::min<int64_t>(bcap, ctx->max_tasks); }
    {
      /* The unique-loop tables at the head of every arena must be empty before a run; a run leaves
       * them empty again, so they are zeroed only when the buffer or its layout changes (or after
       * a failed call: arena_dirty). */
      void *before = ln.arenas.p;
      if (int rc = dev_reserve_s(ln.err, ln.arenas, (size_t)(wave * caps.arena_bytes))) return rc;
      uint64_t sig = 1469598103934665603ULL;
      const unsigned char *cb = (const unsigned char *)&caps;
      for (size_t q = 0; q < sizeof caps; q++) sig = (sig ^ cb[q]) * 1099511628211ULL;
      const size_t need = (size_t)(wave * caps.arena_bytes);
      if (before != ln.arenas.p || sig != ln.arena_sig || ln.arena_dirty) {
        LCK(cudaMemsetAsync(ln.arenas.p, 0, need, ln.stream));
        ln.arena_sig = sig;
        ln.arena_clean = need;
      } else if (need > ln.arena_clean) {
        LCK(cudaMemsetAsync((char *)ln.arenas.p + ln.arena_clean, 0, need - ln.arena_clean, ln.stream));
        ln.arena_clean = need;
      }
      ln.arena_dirty = true; /* until this fold has run to its end */
    }
    {
      /* measured on B200 (tools/threads_sweep.sh, tools/probe/r2_small_threads.sh): with many sequences per
       * resident block the product enumeration wants 4 warps per sequence (more look-ups in flight),
       * selection and materialisation 2 warps up to -ms 100 and 4 beyond; when a wave is so small that a
       * persistent block gets about ONE sequence (a shard of a strong-scaling run), the kernel's time is
       * the time of its heaviest sequence and wider blocks shorten it (-5 % at 1,250 sequences, cfg 2 / 3) */
      const bool small = wave <= (int64_t)ctx->sm_count * 10 && Tb == 64;
      const int T_ms = small ? (J.params->max_stack > 100 ? 256 : 128) : (J.params->max_stack > 100 && Tb == 64 ? 128 : Tb);
      P.T_combine = threads_of("RAFFT_COMBINE_THREADS", small ? 256 : (Tb == 64 ? 128 : Tb));
      P.T_select = threads_of("RAFFT_SELECT_THREADS", T_ms);
      P.T_mat = threads_of("RAFFT_MATERIALISE_THREADS", T_ms);
    }
    if (int rc = dev_reserve_s(ln.err, ln.states, sizeof(SeqState) * (size_t)wave)) return rc;
    if (int rc = dev_reserve_s(ln.err, ln.tasks, sizeof(int32_t) * 2 * (size_t)qcap)) return rc;
    if (int rc = dev_reserve_s(ln.err, ln.ftasks, sizeof(int32_t) * 2 * RF_FAST_LISTS * (size_t)(fcap + 1))) return rc;
    if (int rc = dev_reserve_s(ln.err, ln.scratch, (size_t)qcap * (size_t)caps.wscratch_bytes)) return rc;
    if (int rc = dev_reserve_s(ln.err, ln.btasks, sizeof(int32_t) * 2 * (size_t)(bcap + 1))) return rc;
    if (int rc = dev_reserve_s(ln.err, ln.alist, sizeof(int32_t) * 3 * (size_t)wave)) return rc;
    WaveCtx &wc = P.wc;
    memset(&wc, 0, sizeof wc);
    wc.E = ctx->d_energy; wc.prm = F; wc.caps = caps;
    arena_offsets(wc.aoff, caps, F);
    wc.arenas = (char *)ln.arenas.p; wc.states = (SeqState *)ln.states.p;
    wc.seqs = T.d_seqs; wc.offs = d_offs;
    LoopCtl *ctl = (LoopCtl *)ln.ctl.p;
    wc.ctl = ctl;
    wc.alist = (int32_t *)ln.alist.p; wc.acap = (int32_t)wave;
    wc.h_prog = h_prog;
    wc.fuse = ctx->fuse;
    wc.Q.n_tasks = &ctl->n_tasks; wc.Q.cap = (int32_t)qcap; wc.Q.task = (int32_t *)ln.tasks.p;
    wc.Q.scratch = (char *)ln.scratch.p; wc.Q.self = 0;
    wc.Q.n_fast = ctl->n_fast; wc.Q.fast_cap = (int32_t)fcap; wc.Q.fast_task = (int32_t *)ln.ftasks.p;
    wc.Q.fast_lv = fastcfg.L_cap; wc.Q.fast_br = fastcfg.br_cap;
    wc.fast = fastcfg;
    wc.Q.n_big = &ctl->n_big; wc.Q.big_cap = (int32_t)bcap; wc.Q.big_task = (int32_t *)ln.btasks.p;
    wc.Q.big_lv = bigcfg.L_cap; wc.Q.big_br = bigcfg.br_cap;
    wc.big = bigcfg;
    P.use_big = bcap > 0;
    P.big_smem = (size_t)bigcfg.smem_bytes;
    P.g_big = ctx->sm_count * 2;
    if (P.use_big) LCK(cudaFuncSetAttribute(k_expand_big, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P.big_smem));
    P.fsh = &g_fast_shapes[ctx->fast_shape];
    P.fast_smem = (size_t)(P.fsh->threads / 32) * (size_t)fastcfg.smem_per_warp;
    P.use_fast = fcap > 0;
    /* a branch of a loop covers at least min_hp + 2 positions */
    P.any_generic = fcap == 0 || Lr > (bcap > 0 ? bigcfg.L_cap : fastcfg.L_cap) || Lr / (F.min_hp + 2) > fastcfg.br_cap;
    wc.out = O;
    CombineStage cs_probe;
    P.stage_bytes = stage_carve(cs_probe, nullptr, caps.chunk_cap, F.max_stack);
    LCK(cudaFuncSetAttribute(k_combine, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P.stage_bytes));
    {
      const size_t sel_bytes = sizeof(int32_t) * (size_t)std::max<int64_t>(caps.chunk_cap, (int64_t)caps.new_cap + F.max_stack);
      wc.sel_smem = sel_bytes <= 40 * 1024 ? (int32_t)sel_bytes : 0;
    }
    P.egrid = ctx->sm_count * ctx->blocks_per_sm;
    P.g_fast = ctx->sm_count * P.fsh->blocks;
    if (P.use_fast) LCK(cudaFuncSetAttribute(P.fsh->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)P.fast_smem));
    {
      /* longest-first processing order (ties: input order): counting sort over the lengths on
       * the host, which holds the offsets anyway */
      std::vector<int32_t> start(4098, 0);
      h_order.resize((size_t)m);
      for (int q = 0; q < m; q++) {
        const int si = round == 0 ? q : todo[q];
        start[4096 - (int)(h_offs[si + 1] - h_offs[si]) + 1]++;
      }
      for (int b = 0; b < 4097; b++) start[b + 1] += start[b];
      for (int q = 0; q < m; q++) {
        const int si = round == 0 ? q : todo[q];
        h_order[(size_t)start[4096 - (int)(h_offs[si + 1] - h_offs[si])]++] = si;
      }
      LCK(cudaMemcpyAsync(ln.order.p, h_order.data(), sizeof(int32_t) * (size_t)m, cudaMemcpyHostToDevice, ln.stream));
      LCK(cudaStreamSynchronize(ln.stream)); /* h_order is pageable and reused */
      T.h2d += sizeof(int32_t) * (int64_t)m;
    }
    for (int64_t w0 = 0; w0 < m; w0 += wave) {
      const int nw = (int)std::min<int64_t>(wave, m - w0);
      T.waves++;
      wc.order = (const int32_t *)ln.order.p + w0;
      wc.n = nw;
      P.nw = nw;
      /* persistent block-scope kernels: a grid that fills the GPU once (occupancy x SMs), never
       * more blocks than sequences */
      auto grid_of = [&](const void *fn, int threads, size_t smem) {
        int nb = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fn, threads, smem) != cudaSuccess || nb < 1) { (void)cudaGetLastError(); nb = 1; }
        return (int)std::min<int64_t>(nw, (int64_t)nb * ctx->sm_count * ctx->persist);
      };
      P.g_gather = grid_of((const void *)k_gather, P.T_gather, 0);
      P.g_combine = grid_of((const void *)k_combine, P.T_combine, P.stage_bytes);
      P.g_select = grid_of((const void *)k_select, P.T_select, (size_t)wc.sel_smem);
      P.g_mat = wc.fuse ? grid_of((const void *)k_beam, P.T_mat, (size_t)wc.sel_smem) : grid_of((const void *)k_materialise, P.T_mat, 0);
      /* every fold of an arena tags its `seen` entries with a new generation; when the 16-bit
       * counter wraps the tables are zeroed once */
      if (++ln.seen_gen > 65535u) {
        LCK(cudaMemsetAsync(ln.arenas.p, 0, ln.arena_clean, ln.stream));
        ln.seen_gen = 1;
      }
      wc.prm.seen_gen = 0; /* the kernels read the generation from the control block */
      h_prog[0] = 0; h_prog[1] = 0; /* the lane's stream is idle here: no kernel writes them now */
      k_ctl_reset<<<1, 64, 0, ln.stream>>>(ctl, nw, (int)ln.seen_gen);
      T.launches++;
      const int per_iter = (P.use_fast ? 1 : 0) + (P.use_big ? 1 : 0) + (P.any_generic ? 5 : 0) + (wc.fuse ? 2 : 4) + 1;
      bool done = false;
      if (!profile && ctx->loop_mode == 0) {
        std::string why;
        cudaGraphExec_t exec = wave_graph(ln, P, why);
        if (exec) {
          LCK(cudaGraphLaunch(exec, ln.stream));
          done = true;
        } else {
          if (dbg) fprintf(stderr, "[rafft] launch graph unavailable (%s): stream loop\n", why.c_str());
          ctx->loop_mode = 1;
        }
      }
      if (!done) {
        Emit E;
        E.stream = ln.stream;
        void *a_wc[] = {&wc};
        cudaGraphConditionalHandle nohandle = 0;
        if (profile) LCK(cudaEventRecord(ln.pev[0], ln.stream));
        E.kernel((const void *)k_init, dim3(nw), dim3(P.T), 0, a_wc);
        if (profile) {
          /* stepwise: events around every kernel family, one host synchronisation per iteration */
          LCK(cudaEventRecord(ln.pev[1], ln.stream));
          LCK(cudaStreamSynchronize(ln.stream));
          float ms = 0.f;
          LCK(cudaEventElapsedTime(&ms, ln.pev[0], ln.pev[1]));
          T.fam_ms[FAM_OTHER] += ms; T.fam_n[FAM_OTHER]++;
          for (int it = 0; it < 1000000; it++) {
            emit_iteration(E, P, nohandle, 0, ln.pev);
            if (E.err != cudaSuccess) break;
            LCK(cudaStreamSynchronize(ln.stream));
            const int fams[7] = {FAM_GATHER, FAM_EXPAND_FAST, FAM_EXPAND_GENERIC, FAM_COMBINE, FAM_SELECT, FAM_MATERIALISE, FAM_ADVANCE};
            for (int f = 0; f < 7; f++) {
              LCK(cudaEventElapsedTime(&ms, ln.pev[f], ln.pev[f + 1]));
              const bool ran = !(fams[f] == FAM_EXPAND_FAST && !P.use_fast) && !(fams[f] == FAM_EXPAND_GENERIC && !P.any_generic) &&
                               !(wc.fuse && (fams[f] == FAM_GATHER || fams[f] == FAM_SELECT));
              if (ran) { T.fam_ms[fams[f]] += ms; T.fam_n[fams[f]] += fams[f] == FAM_EXPAND_GENERIC ? 5 : 1; }
              if (dbg) {
                static const char *const nm[7] = {"gather", "fast", "generic", "combine", "select", "materialise", "advance"};
                if (f == 0) fprintf(stderr, "[rafft] it %2d:", it);
                fprintf(stderr, "  %s %.3f", nm[f], ms);
              }
            }
            if (dbg) fprintf(stderr, "  ms\n");
            if (h_prog[1]) break;
          }
        } else {
          /* stream loop: the host keeps at most 3 iterations ahead of the device (progress words in
           * mapped host memory written by k_advance); iterations launched after the last live one
           * find an empty active list and end at once */
          volatile int32_t *hp = h_prog;
          int launched = 0;
          while (E.err == cudaSuccess && !hp[1]) {
            if (launched - hp[0] < 3) { emit_iteration(E, P, nohandle, 0, nullptr); launched++; }
            else std::this_thread::yield();
          }
        }
        if (profile) LCK(cudaEventRecord(ln.pev[0], ln.stream));
        E.kernel((const void *)k_finish, dim3(nw), dim3(P.T_finish), 0, a_wc);
        if (profile) {
          LCK(cudaEventRecord(ln.pev[1], ln.stream));
          LCK(cudaStreamSynchronize(ln.stream));
          float ms = 0.f;
          LCK(cudaEventElapsedTime(&ms, ln.pev[0], ln.pev[1]));
          T.fam_ms[FAM_OTHER] += ms; T.fam_n[FAM_OTHER]++;
        }
        if (E.err != cudaSuccess) { ln.err = std::string("kernel launch: ") + cudaGetErrorString(E.err); return RAFFT_ERR_CUDA; }
      }
      /* iterations of this wave (for the launch count) */
      LCK(cudaMemcpyAsync(ln.h_small + 32, &ctl->iters, sizeof(int32_t), cudaMemcpyDeviceToHost, ln.stream));
      LCK(cudaStreamSynchronize(ln.stream));
      const int iters = *(const int32_t *)(ln.h_small + 32);
      if (((volatile int32_t *)h_prog)[1] == 2) { ln.err = "no progress: the folding loop was stopped after " + std::to_string(iters) + " iterations"; return RAFFT_ERR_CUDA; }
      T.iters += iters;
      T.launches += 2 + (int64_t)iters * per_iter;
      T.d2h += 4;
      if (dbg) { /* regions per size class (work lists of the shared-memory expansion, lower/upper half of each class) */
        LoopCtl hc;
        cudaMemcpyAsync(&hc, ctl, sizeof hc, cudaMemcpyDeviceToHost, ln.stream);
        cudaStreamSynchronize(ln.stream);
        fprintf(stderr, "[rafft] lane %d regions per list (lags <= 24,32,48,64,...):", k);
        for (int q = 0; q < RF_FAST_LISTS; q++) fprintf(stderr, " %d", hc.fast_total[q]);
        fprintf(stderr, "; big %d; generic %d\n", hc.big_total, hc.generic_total);
      }
      if (dbg) { /* high-water marks of the arena pieces against their capacities */
        std::vector<SeqState> hs((size_t)nw);
        cudaMemcpyAsync(hs.data(), ln.states.p, sizeof(SeqState) * (size_t)nw, cudaMemcpyDeviceToHost, ln.stream);
        cudaStreamSynchronize(ln.stream);
        int hw[7] = {0, 0, 0, 0, 0, 0, 0};
        for (const SeqState &q : hs) {
          const int v[7] = {q.n_node, q.n_pool16, q.n_cand, q.n_store, q.n_nlist, q.n_seen, q.n_steps};
          for (int x = 0; x < 7; x++) hw[x] = std::max(hw[x], v[x]);
        }
        fprintf(stderr, "[rafft] lane %d: %d iterations; high water (cap): node %d (%d) pool16 %d (%d) cand %d (%d) store %d (%d) nlist %d (%d) seen %d (%d) steps %d (%d); arena %.1f MB, scale %d\n",
                k, iters, hw[0], caps.node_cap, hw[1], caps.pool16_cap, hw[2], caps.cand_cap, hw[3], caps.store_cap, hw[4], caps.nlist_cap,
                hw[5], caps.seen_cap, hw[6], caps.step_cap, caps.arena_bytes / 1048576.0, scale);
      }
    }
    LCK(cudaGetLastError());
    LCK(cudaMemcpyAsync(T.h_status.data(), O.status, sizeof(int32_t) * (size_t)m_all, cudaMemcpyDeviceToHost, ln.stream));
    LCK(cudaMemcpyAsync(ln.h_small, O.cursors, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ln.stream));
    LCK(cudaStreamSynchronize(ln.stream));
    ln.arena_dirty = false; /* every fold of the round ran to its end: the loop tables are clean */
    T.d2h += sizeof(int32_t) * (int64_t)m_all + 64;
    for (int q = 0; q < 4; q++) used[q] = ln.h_small[q];
    todo.clear();
    for (int i = 0; i < m_all; i++)
      if (T.h_status[(size_t)i] == ST_ERR_CAPACITY) todo.push_back(i);
    if (dbg && !todo.empty())
      fprintf(stderr, "[rafft] lane %d round %d: %zu sequences over capacity (scale %d)\n", k, round, todo.size(), scale);
    if (todo.empty() || round >= 5) break;
    scale *= 4;
  }

  /* ordered layout of this lane on the device */
  const int n = m_all;
  const size_t scan_bytes = sizeof(int64_t) * (7 * (size_t)(n + 1) + 8);
  if (int rc = dev_reserve_s(ln.err, ln.scan, scan_bytes)) return rc;
  int64_t *d_scan = (int64_t *)ln.scan.p;
  int64_t *d_struct_off = d_scan, *d_db_off = d_scan + (n + 1), *d_step_off = d_scan + 2 * (size_t)(n + 1),
          *d_ent_base = d_scan + 3 * (size_t)(n + 1), *d_goff = d_scan + 4 * (size_t)(n + 1),
          *d_totals = d_scan + 7 * (size_t)(n + 1);
  if (n > 0) {
    scan_kernel<<<1, 1024, 0, ln.stream>>>(d_offs, n, O, d_struct_off, d_db_off, d_step_off, d_ent_base, d_totals);
    T.launches++;
    LCK(cudaMemcpyAsync(ln.h_small, d_totals, 4 * sizeof(int64_t), cudaMemcpyDeviceToHost, ln.stream));
    LCK(cudaStreamSynchronize(ln.stream));
    T.d2h += 32;
    for (int q = 0; q < 4; q++) T.totals[q] = ln.h_small[q];
  }
  const int64_t nS = T.totals[0], nD = T.totals[1], nT = T.totals[2], nE = T.totals[3];
  (void)nS; (void)nD; (void)nT; (void)nE;
  T.O = O;
  return RAFFT_OK;
}

/* second half of a lane: its results into the ordered layout (offsets continue after `base`, what
 * the lanes before it hold) and into the job's host block */
static int lane_pack(FoldJob &J, int k, const int64_t *base) {
  LaneTask &T = J.lanes[(size_t)k];
  rafft_ctx *ctx = T.ctx;
  Lane &ln = ctx->lane[T.lane];
  const int n = T.i1 - T.i0;
  const int64_t *d_offs = T.d_offs + T.i0;
  const OutBuf &O = T.O;
  const int64_t nS = T.totals[0], nD = T.totals[1], nT = T.totals[2], nE = T.totals[3];
  int64_t *d_scan = (int64_t *)ln.scan.p;
  int64_t *d_struct_off = d_scan, *d_db_off = d_scan + (n + 1), *d_step_off = d_scan + 2 * (size_t)(n + 1),
          *d_ent_base = d_scan + 3 * (size_t)(n + 1), *d_goff = d_scan + 4 * (size_t)(n + 1);
  /* device ordered buffer: energy | entry_struct | entry_off | db */
  const size_t o_energy = 0, o_entry = align_up(o_energy + sizeof(int32_t) * (size_t)nS, 256),
               o_entoff = align_up(o_entry + sizeof(int32_t) * (size_t)nE, 256),
               o_db = align_up(o_entoff + sizeof(int64_t) * (size_t)(nT + 1), 256),
               o_end = align_up(o_db + (size_t)nD, 256);
  if (int rc = dev_reserve_s(ln.err, ln.ordered, o_end)) return rc;
  char *dord = (char *)ln.ordered.p;
  if (n > 0) {
    pack_kernel<<<n, 256, 0, ln.stream>>>(d_offs, n, O, d_struct_off, d_db_off, d_step_off, d_ent_base,
                                         (int32_t *)(dord + o_energy), dord + o_db,
                                         (int64_t *)(dord + o_entoff), (int32_t *)(dord + o_entry),
                                         base[0], base[1], base[2], base[3], d_goff);
    T.launches++;
  }
  LCK(cudaEventRecord(ln.ev1, ln.stream));
  LCK(cudaGetLastError());
  /* copy into the job's host block (skipped when a section would not fit: the job then gets a
   * larger block and every lane copies again, lane_copy_out) */
  bool fits;
  {
    std::lock_guard<std::mutex> lk(J.mu);
    fits = true;
    for (int q = 0; q < 4; q++) fits = fits && base[q] + T.totals[q] <= J.cap[q];
  }
  if (fits && n > 0) {
    char *hb = (char *)J.block.p;
    if (nS) LCK(cudaMemcpyAsync(hb + J.o_energy + sizeof(int32_t) * (size_t)base[0], dord + o_energy, sizeof(int32_t) * (size_t)nS, cudaMemcpyDeviceToHost, ln.stream));
    if (nE) LCK(cudaMemcpyAsync(hb + J.o_entry + sizeof(int32_t) * (size_t)base[3], dord + o_entry, sizeof(int32_t) * (size_t)nE, cudaMemcpyDeviceToHost, ln.stream));
    LCK(cudaMemcpyAsync(hb + J.o_entoff + sizeof(int64_t) * (size_t)base[2], dord + o_entoff, sizeof(int64_t) * (size_t)(nT + 1), cudaMemcpyDeviceToHost, ln.stream));
    if (nD) LCK(cudaMemcpyAsync(hb + J.o_db + (size_t)base[1], dord + o_db, (size_t)nD, cudaMemcpyDeviceToHost, ln.stream));
    LCK(cudaMemcpyAsync(hb + J.o_soff + sizeof(int64_t) * (size_t)T.i0, d_goff, sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyDeviceToHost, ln.stream));
    LCK(cudaMemcpyAsync(hb + J.o_doff + sizeof(int64_t) * (size_t)T.i0, d_goff + (n + 1), sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyDeviceToHost, ln.stream));
    LCK(cudaMemcpyAsync(hb + J.o_toff + sizeof(int64_t) * (size_t)T.i0, d_goff + 2 * (size_t)(n + 1), sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyDeviceToHost, ln.stream));
    LCK(cudaMemcpyAsync(hb + J.o_cnt + sizeof(int64_t) * (size_t)T.i0 * CNT_N, O.counters, sizeof(int64_t) * (size_t)n * CNT_N, cudaMemcpyDeviceToHost, ln.stream));
    T.copied = true;
    T.d2h += (int64_t)(sizeof(int32_t) * (size_t)(nS + nE) + sizeof(int64_t) * (size_t)(nT + 1) + (size_t)nD) +
             (int64_t)sizeof(int64_t) * (3 * (n + 1) + (int64_t)n * CNT_N);
  }
  LCK(cudaStreamSynchronize(ln.stream));
  LCK(cudaEventElapsedTime(&T.span_ms, ctx->ev_job, ln.ev1));
  if (fits && n > 0) memcpy((char *)J.block.p + J.o_status + sizeof(int32_t) * (size_t)T.i0, T.h_status.data(), sizeof(int32_t) * (size_t)n);
  return RAFFT_OK;
}

/* second chance after a host block that was too small: the lane's ordered buffer is still on the device */
static int lane_copy_out(FoldJob &J, int k) {
  LaneTask &T = J.lanes[(size_t)k];
  rafft_ctx *ctx = T.ctx;
  Lane &ln = ctx->lane[T.lane];
  const int n = T.i1 - T.i0;
  if (n <= 0) return RAFFT_OK;
  LCK(cudaSetDevice(ctx->device));
  const int64_t nS = T.totals[0], nD = T.totals[1], nT = T.totals[2], nE = T.totals[3];
  const int64_t *base = J.cum[k];
  const size_t o_energy = 0, o_entry = align_up(o_energy + sizeof(int32_t) * (size_t)nS, 256),
               o_entoff = align_up(o_entry + sizeof(int32_t) * (size_t)nE, 256),
               o_db = align_up(o_entoff + sizeof(int64_t) * (size_t)(nT + 1), 256);
  char *dord = (char *)ln.ordered.p, *hb = (char *)J.block.p;
  int64_t *d_goff = (int64_t *)ln.scan.p + 4 * (size_t)(n + 1);
  OutBuf O;
  memset(&O, 0, sizeof O);
  idx_layout(O, (char *)ln.idx.p, n);
  if (nS) LCK(cudaMemcpyAsync(hb + J.o_energy + sizeof(int32_t) * (size_t)base[0], dord + o_energy, sizeof(int32_t) * (size_t)nS, cudaMemcpyDeviceToHost, ln.stream));
  if (nE) LCK(cudaMemcpyAsync(hb + J.o_entry + sizeof(int32_t) * (size_t)base[3], dord + o_entry, sizeof(int32_t) * (size_t)nE, cudaMemcpyDeviceToHost, ln.stream));
  LCK(cudaMemcpyAsync(hb + J.o_entoff + sizeof(int64_t) * (size_t)base[2], dord + o_entoff, sizeof(int64_t) * (size_t)(nT + 1), cudaMemcpyDeviceToHost, ln.stream));
  if (nD) LCK(cudaMemcpyAsync(hb + J.o_db + (size_t)base[1], dord + o_db, (size_t)nD, cudaMemcpyDeviceToHost, ln.stream));
  LCK(cudaMemcpyAsync(hb + J.o_soff + sizeof(int64_t) * (size_t)T.i0, d_goff, sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyDeviceToHost, ln.stream));
  LCK(cudaMemcpyAsync(hb + J.o_doff + sizeof(int64_t) * (size_t)T.i0, d_goff + (n + 1), sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyDeviceToHost, ln.stream));
  LCK(cudaMemcpyAsync(hb + J.o_toff + sizeof(int64_t) * (size_t)T.i0, d_goff + 2 * (size_t)(n + 1), sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyDeviceToHost, ln.stream));
  LCK(cudaMemcpyAsync(hb + J.o_cnt + sizeof(int64_t) * (size_t)T.i0 * CNT_N, O.counters, sizeof(int64_t) * (size_t)n * CNT_N, cudaMemcpyDeviceToHost, ln.stream));
  LCK(cudaStreamSynchronize(ln.stream));
  memcpy(hb + J.o_status + sizeof(int32_t) * (size_t)T.i0, T.h_status.data(), sizeof(int32_t) * (size_t)n);
  T.copied = true;
  T.d2h += (int64_t)(sizeof(int32_t) * (size_t)(nS + nE) + sizeof(int64_t) * (size_t)(nT + 1) + (size_t)nD) +
           (int64_t)sizeof(int64_t) * (3 * (n + 1) + (int64_t)n * CNT_N);
  return RAFFT_OK;
}

static double seq_cost(int64_t L, int max_stack) { return (double)L * (double)L * (8.0 + max_stack) + 2000.0; }

/* n sequences (offsets offs[0..n]) dealt to `parts` consumers as contiguous ranges of the input
 * order with equal estimated cost; every range is non-empty when n >= parts.  lo: parts + 1. */
static void plan_ranges(const int64_t *offs, int n, int parts, int max_stack, int *lo) {
  double tot = 0;
  for (int i = 0; i < n; i++) tot += seq_cost(offs[i + 1] - offs[i], max_stack);
  lo[0] = 0;
  double run = 0;
  int d = 1;
  for (int i = 0; i < n && d < parts; i++) {
    run += seq_cost(offs[i + 1] - offs[i], max_stack);
    /* close part d-1 after sequence i once it holds its share -- or when the sequences left are
     * just enough to give every later part one */
    while (d < parts && i + 1 > lo[d - 1] && (run >= tot * d / parts || n - (i + 1) <= parts - d)) lo[d++] = i + 1;
  }
  for (; d < parts; d++) lo[d] = n;
  lo[parts] = n;
}

/* The fold proper.  ctxs[d] folds the sequences [dev_lo[d], dev_lo[d+1]) whose letters and offsets
 * (all n + 1 of them) are at d_seqs[d] / d_offs[d] on its device; h_offs: host copy of the offsets. */
static int fold_job(rafft_ctx *const *ctxs, int n_ctx, const int *dev_lo, const char *const *d_seqs,
                    const int64_t *const *d_offs, const int64_t *h_offs, int32_t n, const rafft_params *params,
                    rafft_result **out_res, int64_t h2d_bytes) {
  rafft_ctx *ctx = ctxs[0]; /* owns the result and reports errors */
  FoldJob J;
  std::string err;
  if (!build_fold_params(params, &J.F, err)) { ctx->err = "rafft_fold_batch: " + err; return RAFFT_ERR_ARG; }
  for (int i = 0; i < n; i++) {
    const int64_t L = h_offs[i + 1] - h_offs[i];
    if (L < 0 || L > 4096) { ctx->err = "rafft_fold_batch: sequence length must be in [0, 4096]"; return RAFFT_ERR_ARG; }
  }
  J.n = n; J.h_offs = h_offs; J.params = params;
  const FoldParams &F = J.F;
  /* lanes: every device splits its range in two when the result copy is worth overlapping
   * (the first half folds on the high-priority stream and is copied out while the second folds) */
  for (int d = 0; d < n_ctx; d++) {
    const int lo = dev_lo[d], hi = dev_lo[d + 1];
    int64_t nt = h_offs[hi] - h_offs[lo];
    int want = ctxs[d]->lanes;
    /* measured on B200 (profiles/r02_lanes.md): two half-batches double the number of kernel tails and
     * lose more than the overlapped result copy gains, so one lane is the default; RAFFT_LANES=2 /
     * rafft_set_option("lanes", 2) turn the overlap on */
    if (want == 0) want = 1;
    (void)nt;
    if (ctxs[d]->profile || hi - lo < 2) want = 1;
    int cut = hi;
    if (want == 2) { /* lane 0 takes lane_split per cent of the estimated cost */
      double tot = 0, run = 0;
      for (int i = lo; i < hi; i++) tot += seq_cost(h_offs[i + 1] - h_offs[i], F.max_stack);
      cut = lo;
      while (cut < hi - 1 && run < 0.01 * ctxs[d]->lane_split * tot) { run += seq_cost(h_offs[cut + 1] - h_offs[cut], F.max_stack); cut++; }
      if (cut <= lo || cut >= hi) { want = 1; cut = hi; }
    }
    for (int l = 0; l < want; l++) {
      LaneTask t;
      t.ctx = ctxs[d]; t.lane = l;
      t.i0 = l == 0 ? lo : cut; t.i1 = (l == 0 && want == 2) ? cut : hi;
      t.d_seqs = d_seqs[d]; t.d_offs = d_offs[d];
      t.first_of_ctx = l == 0;
      J.lanes.push_back(t);
    }
  }
  const int NL = (int)J.lanes.size();
  memset(J.cum, 0, sizeof J.cum);
  /* capacities of the host block's sections: exact bounds without trajectories, an estimate from the
   * last call with them (a section that turns out too small costs a second copy, not the result) */
  {
    const int64_t total_nt = h_offs[n] - h_offs[0];
    const double r = F.traj ? ctx->traj_ratio * 1.25 + 0.5 : 1.0;
    J.cap[0] = (int64_t)((double)n * F.max_stack * r) + 16;
    J.cap[1] = (int64_t)((double)total_nt * F.max_stack * r) + 256;
    J.cap[2] = F.traj ? (int64_t)n * 40 + 16 : (int64_t)n + 1;
    J.cap[3] = F.traj ? (int64_t)((double)n * F.max_stack * 16 * (r / 6.75)) + 64 : (int64_t)n * F.max_stack + 16;
  }
  job_layout(J);
  ResultImpl *R = new ResultImpl;
  memset(&R->pub, 0, sizeof R->pub);
  R->ctx = ctx;
  R->block = HostBlock();
  auto cleanup = [&](int rc) {
    /* one exit for every failure: the arenas of every lane that took part may hold a half-run fold */
    for (LaneTask &t : J.lanes) t.ctx->lane[t.lane].arena_dirty = true;
    R->block = J.block;
    rafft_result_free(&R->pub);
    return rc;
  };
  if (int rc = host_block_get(ctx, J.o_end, &J.block)) return cleanup(rc);
  /* start-of-job event on every device: lanes time their span against it */
  for (LaneTask &t : J.lanes) {
    if (!t.first_of_ctx) continue;
    if (cudaSetDevice(t.ctx->device) != cudaSuccess || cudaEventRecord(t.ctx->ev_job, t.ctx->lane[0].stream) != cudaSuccess) {
      ctx->err = "rafft_fold_batch: cannot record the start event";
      (void)cudaGetLastError();
      return cleanup(RAFFT_ERR_CUDA);
    }
  }
  auto run_lane = [&](int k) {
    LaneTask &t = J.lanes[(size_t)k];
    t.rc = lane_fold(J, k);
    /* publish in lane order, whatever happened: the lanes behind wait for this */
    int64_t base[4];
    bool go;
    {
      std::unique_lock<std::mutex> lk(J.mu);
      J.cv.wait(lk, [&] { return J.published >= k; });
      if (t.rc != RAFFT_OK) J.abort = true;
      for (int q = 0; q < 4; q++) {
        base[q] = J.cum[k][q];
        J.cum[k + 1][q] = base[q] + (t.rc == RAFFT_OK ? t.totals[q] : 0);
        if (J.cum[k + 1][q] > J.cap[q]) J.overflow = true;
      }
      J.published = k + 1;
      go = !J.abort;
      J.cv.notify_all();
    }
    if (go) t.rc = lane_pack(J, k, base);
  };
  if (NL == 1) {
    run_lane(0);
  } else {
    std::vector<std::thread> th;
    for (int k = 1; k < NL; k++) th.emplace_back(run_lane, k);
    run_lane(0);
    for (auto &t : th) t.join();
  }
  cudaSetDevice(ctx->device);
  for (int k = 0; k < NL; k++) {
    LaneTask &t = J.lanes[(size_t)k];
    if (t.rc != RAFFT_OK) {
      ctx->err = "rafft_fold_batch: " + t.ctx->lane[t.lane].err;
      return cleanup(t.rc);
    }
  }
  if (J.abort) { ctx->err = "rafft_fold_batch: a lane failed"; return cleanup(RAFFT_ERR_CUDA); }
  if (J.overflow) {
    /* exact sizes are known now: a block that holds everything, every lane copies (again) */
    HostBlock old = J.block;
    for (int q = 0; q < 4; q++) J.cap[q] = J.cum[NL][q] + 16;
    job_layout(J);
    J.block = HostBlock();
    { std::lock_guard<std::mutex> g(ctx->pin_mu); ctx->pinned_free.push_back(old); }
    if (int rc = host_block_get(ctx, J.o_end, &J.block)) return cleanup(rc);
    for (int k = 0; k < NL; k++) {
      LaneTask &t = J.lanes[(size_t)k];
      t.copied = false;
      if (int rc = lane_copy_out(J, k)) { ctx->err = "rafft_fold_batch: " + t.ctx->lane[t.lane].err; return cleanup(rc); }
    }
    cudaSetDevice(ctx->device);
  }
  if (F.traj && n > 0) ctx->traj_ratio = std::max(1.0, (double)J.cum[NL][0] / ((double)n * F.max_stack));
  char *hb = (char *)J.block.p;
  if (n == 0) { /* no lane wrote anything: the four offset arrays hold their single 0 */
    ((int64_t *)(hb + J.o_entoff))[0] = 0;
    ((int64_t *)(hb + J.o_soff))[0] = 0; ((int64_t *)(hb + J.o_doff))[0] = 0; ((int64_t *)(hb + J.o_toff))[0] = 0;
  }
  R->block = J.block;
  rafft_result &P = R->pub;
  P.n_seq = n;
  P.status = (const int32_t *)(hb + J.o_status);
  P.struct_off = (const int64_t *)(hb + J.o_soff);
  P.db_off = (const int64_t *)(hb + J.o_doff);
  P.step_off = (const int64_t *)(hb + J.o_toff);
  P.energy_dcal = (const int32_t *)(hb + J.o_energy);
  P.db = hb + J.o_db;
  P.entry_off = (const int64_t *)(hb + J.o_entoff);
  P.entry_struct = (const int32_t *)(hb + J.o_entry);
  P.counters = (const int64_t *)(hb + J.o_cnt);
  P.h2d_bytes = h2d_bytes;
  for (LaneTask &t : J.lanes) {
    P.kernel_ms = std::max(P.kernel_ms, (double)t.span_ms);
    P.kernel_launches += t.launches;
    P.h2d_bytes += t.h2d;
    P.d2h_bytes += t.d2h;
    P.iterations = std::max<int64_t>(P.iterations, t.iters);
    for (int f = 0; f < FAM_N; f++) { P.family_ms[f] += t.fam_ms[f]; P.family_launches[f] += t.fam_n[f]; }
  }
  P.expand_ms = P.family_ms[FAM_EXPAND_FAST] + P.family_ms[FAM_EXPAND_GENERIC];
  P.expand_launches = P.family_launches[FAM_EXPAND_FAST] + P.family_launches[FAM_EXPAND_GENERIC];
  P.lanes = NL;
  P.loop_mode = ctx->profile ? 2 : ctx->loop_mode;
  *out_res = &R->pub;
  return RAFFT_OK;
}

extern "C" {

int rafft_fold_batch(rafft_ctx *ctx, const char *seqs, const int64_t *offs, int32_t n,
                     const rafft_params *params, rafft_result **out) {
  if (!ctx) return RAFFT_ERR_ARG;
  return rafft_fold_batch_multi(&ctx, 1, seqs, offs, n, params, out);
}

int rafft_fold_batch_multi(rafft_ctx *const *ctxs, int32_t n_ctx, const char *seqs, const int64_t *offs, int32_t n,
                           const rafft_params *params, rafft_result **out) {
  if (!ctxs || n_ctx < 1 || n_ctx > RF_MAX_DEVICES || !ctxs[0]) return RAFFT_ERR_ARG;
  rafft_ctx *ctx = ctxs[0];
  for (int d = 0; d < n_ctx; d++) {
    if (!ctxs[d]) { ctx->err = "rafft_fold_batch_multi: NULL context"; return RAFFT_ERR_ARG; }
    for (int e = 0; e < d; e++)
      if (ctxs[e] == ctxs[d] || ctxs[e]->device == ctxs[d]->device) { ctx->err = "rafft_fold_batch_multi: one context per device, each once"; return RAFFT_ERR_ARG; }
  }
  if (!out || !params || n < 0 || (n > 0 && (!seqs || !offs))) { ctx->err = "rafft_fold_batch: bad argument"; return RAFFT_ERR_ARG; }
  *out = nullptr;
  static const int64_t zero_offs[1] = {0};
  if (n == 0) offs = zero_offs;
  for (int i = 0; i < n; i++)
    if (offs[i + 1] < offs[i]) { ctx->err = "rafft_fold_batch: offsets must be non-decreasing"; return RAFFT_ERR_ARG; }
  const int64_t base = offs[0], total = offs[n] - base;
  std::vector<int64_t> rel((size_t)n + 1);
  for (int i = 0; i <= n; i++) rel[(size_t)i] = offs[i] - base;
  /* contiguous ranges of the input order with equal estimated cost: every device's part of the
   * result is then one contiguous piece of every output array */
  if (n_ctx > n && n > 0) n_ctx = n;
  if (n == 0) n_ctx = 1;
  std::vector<int> dev_lo((size_t)n_ctx + 1, 0);
  plan_ranges(rel.data(), n, n_ctx, params->max_stack, dev_lo.data());
  std::vector<const char *> d_seqs((size_t)n_ctx);
  std::vector<const int64_t *> d_offs((size_t)n_ctx);
  int64_t h2d = 0;
  for (int d = 0; d < n_ctx; d++) {
    rafft_ctx *cx = ctxs[d];
    cudaError_t e = cudaSetDevice(cx->device);
    if (e != cudaSuccess) { ctx->err = std::string("cudaSetDevice: ") + cudaGetErrorString(e); return RAFFT_ERR_CUDA; }
    if (int rc = dev_reserve(cx, cx->seqs, (size_t)total + 16)) { ctx->err = cx->err; return rc; }
    if (int rc = dev_reserve(cx, cx->offs, sizeof(int64_t) * (size_t)(n + 1))) { ctx->err = cx->err; return rc; }
    /* every device gets the letters of its own range and all offsets */
    const int64_t lo = rel[(size_t)dev_lo[(size_t)d]], hi = rel[(size_t)dev_lo[(size_t)d + 1]];
    if (hi > lo) e = cudaMemcpyAsync((char *)cx->seqs.p + lo, seqs + base + lo, (size_t)(hi - lo), cudaMemcpyHostToDevice, cx->lane[0].stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(cx->offs.p, rel.data(), sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyHostToDevice, cx->lane[0].stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(cx->lane[0].stream);
    if (e != cudaSuccess) { ctx->err = std::string("input copy: ") + cudaGetErrorString(e); return RAFFT_ERR_CUDA; }
    h2d += (hi - lo) + (int64_t)sizeof(int64_t) * (n + 1);
    d_seqs[(size_t)d] = (const char *)cx->seqs.p;
    d_offs[(size_t)d] = (const int64_t *)cx->offs.p;
  }
  cudaSetDevice(ctx->device);
  return fold_job(ctxs, n_ctx, dev_lo.data(), d_seqs.data(), d_offs.data(), rel.data(), n, params, out, h2d);
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
  std::vector<int64_t> h_offs((size_t)n + 1, 0);
  if (n > 0) CK(cudaMemcpy(h_offs.data(), d_offs, sizeof(int64_t) * (size_t)(n + 1), cudaMemcpyDeviceToHost));
  if (n > 0 && h_offs[0] != 0) { ctx->err = "rafft_fold_batch_device: offs[0] must be 0"; return RAFFT_ERR_ARG; }
  for (int i = 0; i < n; i++)
    if (h_offs[(size_t)i + 1] < h_offs[(size_t)i]) { ctx->err = "rafft_fold_batch_device: offsets must be non-decreasing"; return RAFFT_ERR_ARG; }
  const int dev_lo[2] = {0, n};
  const char *ds = (const char *)d_seqs;
  const int64_t *dofs = (const int64_t *)d_offs;
  return fold_job(&ctx, 1, dev_lo, &ds, &dofs, h_offs.data(), n, params, out, 0);
}

int rafft_plan_ranges(const int64_t *offs, int32_t n, int32_t parts, int32_t max_stack, int32_t *lo) {
  if (!offs || !lo || n < 0 || parts < 1) return RAFFT_ERR_ARG;
  for (int i = 0; i < n; i++)
    if (offs[i + 1] < offs[i]) return RAFFT_ERR_ARG;
  plan_ranges(offs, n, parts, max_stack, lo);
  return RAFFT_OK;
}

int rafft_int_peak(rafft_ctx *ctx, double *issue_warp_inst_per_s, double *mix_warp_inst_per_s) {
  if (!ctx || !issue_warp_inst_per_s) return RAFFT_ERR_ARG;
  CK(cudaSetDevice(ctx->device));
  if (int rc = dev_reserve(ctx, ctx->peakbuf, 256)) return rc;
  const int iters = 1500, blocks = ctx->sm_count * 8;
  Lane &ln = ctx->lane[0];
  double best[2] = {0, 0};
  for (int mode = 0; mode < 2; mode++) {
    for (int rep = 0; rep < 4; rep++) {
      CK(cudaEventRecord(ln.pev[0], ln.stream));
      int_peak_kernel<<<blocks, 256, 0, ln.stream>>>((unsigned *)ctx->peakbuf.p, iters, 7u + rep, mode);
      CK(cudaEventRecord(ln.pev[1], ln.stream));
      CK(cudaStreamSynchronize(ln.stream));
      float ms = 0.f;
      CK(cudaEventElapsedTime(&ms, ln.pev[0], ln.pev[1]));
      const double winst = (double)blocks * 8.0 * iters * 16.0 * 8.0;
      if (rep > 0) best[mode] = std::max(best[mode], winst / (ms * 1e-3));
    }
  }
  *issue_warp_inst_per_s = best[0];                        /* IADD3 stream: every scheduler issues every cycle */
  if (mix_warp_inst_per_s) *mix_warp_inst_per_s = best[1];  /* IADD3/LOP3 alternating (LOP3 runs at half rate) */
  return RAFFT_OK;
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

/* The device address of the static scratch array on ctx's device: looked up once per context (a
 * context is bound to one device), with that device current.  No allocation. */
static int static_scratch_of(rafft_ctx *ctx, unsigned char **out) {
  if (!ctx->static_scratch) {
    CK(cudaSetDevice(ctx->device));
    void *p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_static_eval_scratch));
    if (!p) { ctx->err = "rafft_eval_structures_static: no device address for the static scratch"; return RAFFT_ERR_CUDA; }
    ctx->static_scratch = (unsigned char *)p;
  }
  *out = ctx->static_scratch;
  return RAFFT_OK;
}

int rafft_eval_structures_static(rafft_ctx *ctx, const void *d_seqs, const void *d_structs, const void *d_offs,
                                 int32_t n, int64_t total_len, void *d_energy_dcal) {
  if (!ctx) return RAFFT_ERR_ARG;
  if (n < 0 || total_len < 0 || (n > 0 && (!d_offs || !d_energy_dcal)) || (total_len > 0 && (!d_seqs || !d_structs))) {
    ctx->err = "rafft_eval_structures_static: bad argument";
    return RAFFT_ERR_ARG;
  }
  if (total_len > (int64_t)RF_STATIC_EVAL_MAX_LETTERS) {
    ctx->err = "rafft_eval_structures_static: " + std::to_string(total_len) + " letters exceed the static scratch maximum of " +
               std::to_string((long long)RF_STATIC_EVAL_MAX_LETTERS) + " (RAFFT_STATIC_EVAL_MAX_LETTERS); split the batch or use rafft_eval_structures";
    return RAFFT_ERR_ARG;
  }
  if (n > (1 << 22)) { ctx->err = "rafft_eval_structures_static: more than 4194304 structures in one call"; return RAFFT_ERR_ARG; }
  if (n == 0) return RAFFT_OK;
  CK(cudaSetDevice(ctx->device));
  unsigned char *scratch = nullptr;
  if (int rc = static_scratch_of(ctx, &scratch)) return rc;
  const int wpb = 4;
  eval_static_kernel<<<(n + wpb - 1) / wpb, wpb * 32, 0, ctx->stream>>>(ctx->d_energy, (const char *)d_seqs, (const char *)d_structs,
                                                                      (const int64_t *)d_offs, n, total_len, scratch,
                                                                      (int32_t *)d_energy_dcal);
  CK(cudaGetLastError());
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
