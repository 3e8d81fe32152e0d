/*
 * rafft_b200 -- C-ABI of the B200-native RAFFT folding hot path.
 *
 * The reference (lemerleau/RAFFT) has no plugin/FFI registry for this path: its boundary is
 * the Python call  rafft.rafft.fold(sequence, nb_mode, max_stack, max_branch, min_hp,
 * min_nrj, traj, temp, gc_wei, au_wei, gu_wei)   (reference rafft/rafft.py:219-239), and
 * the only C-ABI it crosses is ViennaRNA's energy evaluation
 *     vrna_fold_compound(seq, md, VRNA_OPTION_EVAL_ONLY) / vrna_eval_structure_pt(fc, pt)
 * (reference rafft-rs/src/vienna.rs:20-24,31-34; Python: rafft/utils.py:18-21,135-138).
 * The entry points below are what a binding for the path binds instead; each cites the
 * reference interface it replaces.  Plain pointers and sizes only, integer status codes,
 * no exceptions cross the ABI.  One context per CUDA device; a context is not thread-safe
 * (mirrors `#[pyclass(unsendable)]`, reference rafft-rs/src/bindings.rs:10-11).
 */
#ifndef RAFFT_B200_H
#define RAFFT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RAFFT_OK 0
#define RAFFT_ERR_ARG 1      /* bad argument (NULL, n < 0, unsupported parameter value) */
#define RAFFT_ERR_SEQ 2      /* letter outside A,C,G,U (reference: KeyError, rafft/utils.py:73-80) */
#define RAFFT_ERR_CUDA 3     /* CUDA runtime failure; message in rafft_last_error */
#define RAFFT_ERR_CAPACITY 4 /* a per-sequence device arena overflowed even after regrowth */

/* Turner-2004 style nearest-neighbour parameters in dcal/mol.  Pair types 0..5 = CG GC GU
 * UG AU UA, bases 0..3 = A C G U.  Replaces ViennaRNA's vrna_param_t as read by
 * vrna_params_load (reference rafft-rs/src/vienna.rs:75-81). */
#define RAFFT_MAX_SPECIAL 64
typedef struct rafft_energy_params {
  int32_t stack[6][6];
  int32_t hairpin[31], bulge[31], interior[31];
  int32_t mismatchH[6][4][4], mismatchI[6][4][4], mismatch1nI[6][4][4];
  int32_t mismatch23I[6][4][4], mismatchM[6][4][4], mismatchExt[6][4][4];
  int32_t dangle5[6][4], dangle3[6][4];
  int32_t int11[6][6][4][4];
  int32_t int21[6][6][4][4][4];
  int32_t int22[6][6][4][4][4][4];
  int32_t ninio, max_ninio, ml_base, ml_intern, ml_closing, terminal_au;
  double lxc;
  int32_t n_tri, n_tetra, n_hexa;
  char tri[RAFFT_MAX_SPECIAL][8];    /* 5 letters, NUL padded */
  char tetra[RAFFT_MAX_SPECIAL][8];  /* 6 letters */
  char hexa[RAFFT_MAX_SPECIAL][12];  /* 8 letters */
  int32_t tri_e[RAFFT_MAX_SPECIAL], tetra_e[RAFFT_MAX_SPECIAL], hexa_e[RAFFT_MAX_SPECIAL];
} rafft_energy_params;

/* Arguments of fold() (reference rafft/rafft.py:219-221; CLI flags bin/rafft:11-27). */
typedef struct rafft_params {
  int32_t nb_mode;    /* -n   number of positional lags searched per unpaired region */
  int32_t max_stack;  /* -ms  structures saved per folding step */
  int32_t max_branch; /* --max_branch (fold() default 100, CLI default 1000) */
  int32_t min_hp;     /* -mh  minimum unpaired positions in a hairpin */
  int32_t traj;       /* --traj: 1 = return every step, 0 = only the last one */
  double min_nrj;     /* -mn  a stem is kept iff its energy change is < min_nrj (kcal/mol) */
  double temp;        /* --temp: only 37.0 is supported by the built-in tables */
  double gc_wei, au_wei, gu_wei; /* -gc -au -gu pair weights of the correlation */
} rafft_params;

/* Work counters per sequence (the ALGORITHMIC work of the reference's formulation, used for
 * the roofline figures; definitions in DESIGN.md section 6). */
enum {
  RAFFT_CNT_STEPS = 0,    /* bfs_pairs calls (rafft/rafft.py:156) */
  RAFFT_CNT_NODE_EXP,     /* create_childs calls the reference makes (rafft/rafft.py:168) */
  RAFFT_CNT_W_CORR,       /* sum over those calls of Lv(Lv+1)/2 pair-weight accumulations */
  RAFFT_CNT_W_SCAN,       /* sum over (call, ranked lag) of ceil(d/2) scan positions */
  RAFFT_CNT_N_EVAL,       /* eval_one_struct calls (rafft/rafft.py:98,192) */
  RAFFT_CNT_UNIQ_NODES,   /* unique loop regions this implementation actually expanded */
  RAFFT_CNT_UNIQ_W_CORR,  /* pair-weight accumulations actually performed */
  RAFFT_CNT_UNIQ_W_SCAN,  /* scan positions actually visited */
  RAFFT_CNT_COMBOS,       /* stem combinations enumerated (itertools.product items) */
  RAFFT_CNT_NODE_POS,     /* sum over the reference's create_childs calls of the region length Lv */
  RAFFT_CNT_T0_NS,        /* device globaltimer when the fold of this sequence started ... */
  RAFFT_CNT_T1_NS,        /* ... and ended (diagnostics: schedule / tail analysis) */
  RAFFT_CNT_N_EVAL_EXPAND,/* the part of N_EVAL made by find_best_consecutives (rafft/rafft.py:98) */
  RAFFT_CNT_TRIVIAL_NODES,/* unique loop regions that cannot hold a stem (one position, or a hairpin inside of
                             <= min_hp + 1 positions): settled without an expansion */
  RAFFT_CNT_N
};

/* Result of one rafft_fold_batch call; all arrays are host memory owned by the library
 * until rafft_result_free.  Sequence i has the unique structures
 * [struct_off[i], struct_off[i+1]); structure s of sequence i (length L) has energy
 * energy_dcal[s] (integer dcal/mol; the reference's float is float32(E/100)) and
 * dot-bracket text db[db_off[i] + (s - struct_off[i]) * L .. + L) (not NUL terminated).
 * Its folding steps are [step_off[i], step_off[i+1]); step t lists the structures of its
 * beam in order as entry_struct[entry_off[t] .. entry_off[t+1]) (indices local to the
 * sequence).  With traj = 0 only the final beam is returned (one step per sequence). */
typedef struct rafft_result {
  int32_t n_seq;
  const int32_t *status;        /* n_seq: RAFFT_OK or an error code for that sequence */
  const int64_t *struct_off;    /* n_seq + 1 */
  const int32_t *energy_dcal;   /* struct_off[n_seq] */
  const int64_t *db_off;        /* n_seq + 1 */
  const char *db;               /* db_off[n_seq] */
  const int64_t *step_off;      /* n_seq + 1 */
  const int64_t *entry_off;     /* step_off[n_seq] + 1 */
  const int32_t *entry_struct;  /* entry_off[step_off[n_seq]] */
  const int64_t *counters;      /* n_seq * RAFFT_CNT_N */
  double kernel_ms;             /* device time of the fold kernel(s), CUDA events */
  int64_t kernel_launches;      /* number of kernel launches this call made */
  int64_t h2d_bytes, d2h_bytes; /* bytes copied host->device / device->host by this call */
  double expand_ms;             /* device time of the region-expansion kernel launches alone ... */
  int64_t expand_launches;      /* ... and how many of them there were (profile mode only, else 0) */
  int64_t iterations;           /* folding iterations of the longest-running lane */
  /* per kernel family, CUDA events around every launch (filled in profile mode, rafft_set_option
   * "profile"; zero otherwise): RAFFT_FAM_* */
  double family_ms[8];
  int64_t family_launches[8];
  int32_t lanes;                /* concurrent sub-batches (streams) the call was split into */
  int32_t loop_mode;            /* 0 one CUDA-graph launch per fold (device-side WHILE), 1 stream launches
                                   with bounded look-ahead, 2 stepwise (profile mode) */
} rafft_result;

enum { RAFFT_FAM_GATHER = 0, RAFFT_FAM_EXPAND_FAST, RAFFT_FAM_EXPAND_GENERIC, RAFFT_FAM_COMBINE,
       RAFFT_FAM_SELECT, RAFFT_FAM_MATERIALISE, RAFFT_FAM_ADVANCE, RAFFT_FAM_OTHER };

typedef struct rafft_ctx rafft_ctx;

/* Fill *ep with the built-in Turner-2004 tables (reconstructed). */
void rafft_default_energy_params(rafft_energy_params *ep);

/* Create a context on a CUDA device.  ep == NULL selects the built-in tables.  Replaces
 * RNA.md()/RNA.fold_compound(...) (reference rafft/utils.py:18-21) and
 * vrna_fold_compound (rafft-rs/src/vienna.rs:20-24).  Fails (RAFFT_ERR_CUDA) when no
 * usable CUDA device exists -- there is no CPU fallback. */
int rafft_create(rafft_ctx **out, int cuda_device, const rafft_energy_params *ep);
void rafft_destroy(rafft_ctx *ctx);
/* Message of the last failing call on ctx (ctx == NULL: last rafft_create failure). */
const char *rafft_last_error(const rafft_ctx *ctx);

/* Fold n sequences (ASCII, concatenated in seqs; sequence i = seqs[offs[i] .. offs[i+1])).
 * Replaces n calls of rafft.rafft.fold (reference rafft/rafft.py:219-239).  Synchronous. */
int rafft_fold_batch(rafft_ctx *ctx, const char *seqs, const int64_t *offs, int32_t n,
                     const rafft_params *params, rafft_result **out);
void rafft_result_free(rafft_result *res);

/* The same batch folded by several devices of one box from ONE host process: ctxs[0..n_ctx) are
 * contexts of different CUDA devices; the sequences are dealt to them as contiguous ranges of equal
 * estimated cost (the path has no exchange step: SURVEY.md section 8e), every device folds its range
 * on its own host thread and stream and copies its part straight into the one result, which is in
 * input order and owned by ctxs[0].  Replaces the reference's process pool over the sequences of a
 * set (benchmark_results/bench_fft.py:17-22).  rafft_fold_batch is the n_ctx = 1 case. */
int rafft_fold_batch_multi(rafft_ctx *const *ctxs, int32_t n_ctx, const char *seqs, const int64_t *offs,
                           int32_t n, const rafft_params *params, rafft_result **out);

/* The ranges rafft_fold_batch_multi deals to `parts` devices: lo[p] .. lo[p+1] (parts + 1 values),
 * contiguous in input order, equal estimated cost (~ L^2 (8 + max_stack)).  Needs no device. */
int rafft_plan_ranges(const int64_t *offs, int32_t n, int32_t parts, int32_t max_stack, int32_t *lo);

/* Same, but the inputs already live in device memory (d_seqs: the concatenated letters,
 * d_offs: n+1 int64) and the timed work is the kernel only; results are still returned on
 * the host.  Used by bench.py for the device-resident figure. */
int rafft_fold_batch_device(rafft_ctx *ctx, const void *d_seqs, const void *d_offs, int32_t n,
                            int64_t total_len, int32_t max_len, const rafft_params *params,
                            rafft_result **out);

/* Free energy of whole structures (dot-bracket, same layout as seqs) in dcal/mol, on the
 * device, one structure per sequence.  Replaces vrna_eval_structure_pt
 * (reference rafft-rs/src/vienna.rs:31-34) / fold_compound.eval_structure
 * (rafft/utils.py:135-138). */
int rafft_eval_structures(rafft_ctx *ctx, const char *seqs, const char *structs,
                          const int64_t *offs, int32_t n, int32_t *energy_dcal);

/* rafft_eval_structures for callers that cannot allocate (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same energies as rafft_eval_structures, with everything in device
 * memory the CALLER owns -- d_seqs / d_structs: the concatenated letters and dot-brackets, d_offs:
 * n + 1 int64 offsets (d_offs[0] = 0, d_offs[n] = total_len), d_energy_dcal: n int32 -- and the
 * kernel's scratch (pair tables, base codes: 9 bytes per letter) in a file-scope device array of
 * fixed size instead of a cudaMalloc'ed buffer.  The call makes no cudaMalloc / cudaFree /
 * cudaHostAlloc and no allocating library call; it launches one kernel on the context's stream and
 * returns after that stream has been synchronised (inputs must be complete before the call).
 *   limit     total_len <= RAFFT_STATIC_EVAL_MAX_LETTERS (1,048,576 letters per call); more is
 *             refused with RAFFT_ERR_ARG and a message naming the limit.  Per structure <= 4096 letters;
 *             a structure whose offsets are out of range, longer than that, malformed or holding a
 *             non-canonical pair gets INT32_MIN (as in rafft_eval_structures) and touches nothing.
 *   cost      every GPU on which this library is loaded holds the array (9 MiB) until the process exits.
 *   sharing   all calls on one device share the array, which keeps the last call's contents (the
 *             kernel writes every element before it reads it; nothing must start at zero).  Calls that
 *             overlap on the GPU -- two contexts of one device used from two threads -- RACE on it:
 *             serialise them. */
#define RAFFT_STATIC_EVAL_MAX_LETTERS 1048576
int rafft_eval_structures_static(rafft_ctx *ctx, const void *d_seqs, const void *d_structs, const void *d_offs,
                                 int32_t n, int64_t total_len, void *d_energy_dcal);

/* Folding graph of the kinetics (reference rafft/rafft_kin.py:48-56 get_connected_prev, :106-115
 * the non-redundant numbering in kinetics).  db: n = step_off[n_steps] dot-bracket rows of L
 * characters, the folding steps concatenated (the blocks of a `rafft --traj` output).
 *   uniq[i]   index of the first row equal to row i (struct_map of the reference)
 *   conn      for step t, for every structure c of step t and every structure p of step t-1
 *             (step 0 looks at the LAST step: fast_paths[-1], rafft_kin.py:75), 1 iff every pair of
 *             p is a pair of c; the row of c starts at conn_off[t] + (c - step_off[t]) * n_prev
 *   conn_off  n_steps + 1 offsets into conn (written by the call); conn_cap = capacity of conn
 * A closing bracket without partner is RAFFT_ERR_SEQ (the reference raises IndexError,
 * rafft/utils.py:62-65). */
int rafft_kin_graph(rafft_ctx *ctx, const char *db, int32_t L, const int64_t *step_off, int32_t n_steps,
                    int32_t *uniq, uint8_t *conn, int64_t conn_cap, int64_t *conn_off);

/* Landscape distance matrix (reference utility/surface.py:19-26 get_distance_matrix, which calls
 * RNA.bp_distance for every pair): db = n dot-bracket rows of L characters; dist = n * n int32,
 * dist[a * n + b] = number of base pairs present in exactly one of the structures a and b
 * (symmetric, zero diagonal).  Errors as rafft_kin_graph. */
int rafft_bp_distance_matrix(rafft_ctx *ctx, const char *db, int32_t n, int32_t L, int32_t *dist);

/* Tuning / diagnostics switches of a context: "profile" (1: time every kernel family with CUDA
 * events, stepwise loop; results unchanged), "lanes" (0 automatic, 1, 2), "loop" (0 graph, 1 stream),
 * "max_wave", "persist".  Unknown names fail with RAFFT_ERR_ARG. */
int rafft_set_option(rafft_ctx *ctx, const char *name, int64_t value);

/* Issue-rate microbenchmark on the context's device: warp-instructions per second of a stream of
 * independent IADD3 (every scheduler of every SM issues every cycle: the issue-slot peak the fold
 * kernels are measured against, bench.py roofline) and of an IADD3/LOP3 mix.  ~10 ms. */
int rafft_int_peak(rafft_ctx *ctx, double *issue_warp_inst_per_s, double *mix_warp_inst_per_s);

/* CUDA devices usable by rafft_create (0 without a driver / GPU). */
int rafft_device_count(void);

/* Library/build identification, e.g. "rafft_b200 0.1 sm_100a". */
const char *rafft_version(void);

#ifdef __cplusplus
}
#endif
#endif /* RAFFT_B200_H */
