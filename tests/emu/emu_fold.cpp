/*
 * TEST-ONLY host emulation of the device code: compiles rafft_b200/csrc/fold_core.cuh with a
 * host execution policy (every `par` phase runs its iterations one after the other, forward
 * or reversed) so that the ordering logic of the CUDA path can be checked against the
 * oracle in the CPU test tier.  This is NOT a CPU fallback: nothing in rafft_b200/ loads
 * it and the product fails loudly without the CUDA library.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string>
#include <vector>
#include "../../rafft_b200/csrc/host_common.h"

using namespace rafft;

static int g_queue_cap = 0;

extern "C" {

void emu_set_queue_cap(int cap) { g_queue_cap = cap; }

void emu_default_energy_params(rafft_energy_params *ep) { default_energy_params(ep); }

/* Same text format as oracle_fold.  hiwater: 8 ints (nodes, pool16, cands, store, nlist,
 * seen, steps, arena KiB).  Returns the sequence status or a negative argument error. */
int emu_fold(const rafft_energy_params *ep_in, const char *seq, int L, const rafft_params *prm,
             int reverse, int nthreads, int scale, char **text, int64_t *cnt, int32_t *hiwater) {
  rafft_energy_params *ep = (rafft_energy_params *)malloc(sizeof *ep);
  if (ep_in) *ep = *ep_in; else default_energy_params(ep);
  DevEnergy *D = (DevEnergy *)malloc(sizeof *D);
  FoldParams F;
  std::string err;
  if (!build_dev_energy(ep, D, err) || !build_fold_params(prm, &F, err)) {
    fprintf(stderr, "emu_fold: %s\n", err.c_str());
    free(ep); free(D);
    return -1;
  }
  Exec::order() = reverse;
  Exec::emu_threads() = nthreads > 0 ? nthreads : 32;
  Caps caps = make_caps(L, F, Exec::nthreads(), scale > 0 ? scale : 1);
  std::vector<char> arena((size_t)caps.arena_bytes, 0);  /* hash tables start empty */
  Fold f(D, F, caps);
  ArenaOff aoff;
  arena_offsets(aoff, caps, F);
  arena_bind(f.A, arena.data(), aoff);
  SeqState sh;
  f.sh = &sh;
  /* region-expansion work list of one round (tiny capacities exercise the deferral path) */
  int qcap = g_queue_cap > 0 ? g_queue_cap : caps.todo_cap;
  int32_t n_tasks = 0;
  std::vector<int32_t> tasks(2 * (size_t)qcap);
  std::vector<char> scratch((size_t)caps.wscratch_bytes);
  f.Q.n_tasks = &n_tasks; f.Q.cap = qcap; f.Q.task = tasks.data(); f.Q.scratch = scratch.data(); f.Q.self = 0;
  f.Q.n_fast = nullptr; f.Q.fast_cap = 0; f.Q.fast_lv = 0; f.Q.fast_br = 0; f.Q.fast_task = nullptr; /* CUDA-only path */
  f.Q.n_big = nullptr; f.Q.big_cap = 0; f.Q.big_lv = 0; f.Q.big_br = 0; f.Q.big_task = nullptr;
  CombineStage probe;
  std::vector<char> stage(stage_carve(probe, nullptr, caps.chunk_cap, F.max_stack) + 64);
  stage_carve(f.cs, (char *)(((uintptr_t)stage.data() + 15) & ~(uintptr_t)15), caps.chunk_cap, F.max_stack);
  /* output for one sequence */
  int64_t scap = (int64_t)caps.store_cap, ecap = (int64_t)caps.step_cap * F.max_stack;
  std::vector<int32_t> status(1), nstruct(1), nstep(1), energy(scap), step_n(caps.step_cap), entries(ecap);
  std::vector<int64_t> sstruct(1), sdb(1), sstep(1), sentry(1), counters(CNT_N);
  std::vector<char> db((size_t)scap * (size_t)(((L > 0 ? L : 1) + 15) & ~15) + 16);
  unsigned long long cursors[8] = {0};
  OutBuf out;
  out.status = status.data(); out.seq_struct = sstruct.data(); out.seq_nstruct = nstruct.data();
  out.seq_db = sdb.data(); out.seq_step = sstep.data(); out.seq_nstep = nstep.data();
  out.seq_entry = sentry.data(); out.counters = counters.data(); out.energy = energy.data();
  out.db = db.data(); out.step_n = step_n.data(); out.entries = entries.data(); out.cursors = cursors;
  out.struct_cap = scap; out.db_cap = (int64_t)db.size(); out.step_cap = caps.step_cap; out.entry_cap = ecap;
  /* the host loop of rafft_cuda.cu, for one sequence: rounds of phase "kernels" */
  /* two folds in the same arena: the second one (the one reported) meets the `seen` entries the
   * first one left behind under another generation tag */
  for (int gen = 1; gen <= 2; gen++) {
  F.seen_gen = gen;
  f.gen = gen;
  n_tasks = 0;
  cursors[0] = cursors[1] = cursors[2] = cursors[3] = cursors[4] = cursors[5] = cursors[6] = cursors[7] = 0;
  f.phase_init(seq, L, 0);
  for (int round = 0; sh.active; round++) {
    n_tasks = 0;
    f.phase_gather();
    int nt = n_tasks < qcap ? n_tasks : qcap;
    for (int t = 0; t < nt; t++) {
      int a = Exec::order() ? nt - 1 - t : t;
      f.expand_node(tasks[2 * a + 1], scratch.data());
    }
    f.phase_combine();
    f.phase_select();
    f.phase_materialise();
    if (round > 100000) { fprintf(stderr, "emu_fold: no progress\n"); return -3; }
  }
  f.phase_finish(out);
  /* a fold must leave the unique-loop table empty for the next sequence of the arena (the
   * `seen` table in front of it is generation-tagged instead) */
  {
    size_t t0 = (size_t)((char *)f.A.nht - arena.data()), tb = sizeof(unsigned long long) * (size_t)caps.nht_cap;
    for (size_t k = t0; k < t0 + tb; k++)
      if (arena[k]) { fprintf(stderr, "emu_fold: unique-loop table not clean after run\n"); return -2; }
  }
  }
  if (status[0] == ST_ERR_CAPACITY && getenv("EMU_VERBOSE")) fprintf(stderr, "emu_fold: capacity exceeded at fold_core.cuh line %d\n", -sh.fail_line < 0 ? sh.fail_line : -sh.fail_line);
  if (cnt) for (int k = 0; k < CNT_N; k++) cnt[k] = counters[k];
  if (hiwater) {
    hiwater[0] = sh.n_node; hiwater[1] = sh.n_pool16; hiwater[2] = sh.n_cand; hiwater[3] = sh.n_store;
    hiwater[4] = sh.n_nlist; hiwater[5] = sh.n_seen; hiwater[6] = sh.n_steps;
    hiwater[7] = (int32_t)(caps.arena_bytes >> 10);
  }
  if (text) {
    std::string o;
    char buf[64];
    int64_t e = sentry[0];
    for (int s = 0; s < nstep[0]; s++) {
      snprintf(buf, sizeof buf, "# %d\n", s);
      o += buf;
      for (int b = 0; b < step_n[sstep[0] + s]; b++, e++) {
        int k = entries[e];
        o.append(db.data() + sdb[0] + (size_t)k * (size_t)((L + 15) & ~15), L);
        snprintf(buf, sizeof buf, " %d\n", energy[sstruct[0] + k]);
        o += buf;
      }
    }
    *text = (char *)malloc(o.size() + 1);
    memcpy(*text, o.c_str(), o.size() + 1);
  }
  free(ep); free(D);
  return status[0];
}

void emu_free(void *p) { free(p); }
}
