"""Python entry points of the folding path, drop-in for the reference's rafft/rafft.py:
`fold` keeps the signature, defaults and return types of rafft.rafft.fold (:219-239);
`fold_batch` folds many sequences in one call of the CUDA library.  All of the work
(correlation, lag ranking, stem scan, energy evaluation, beam search) runs on the GPU
through the C-ABI in include/rafft_b200.h; this module only marshals arguments/results."""
import ctypes as C
import os

import numpy as np

from . import _native as N
from .utils import Structure, energy_from_dcal



class FoldContext:
    """One `rafft_ctx` (one CUDA device).  Not thread-safe, like the reference's
    `unsendable` handle (rafft-rs/src/bindings.rs:10-11)."""

    def __init__(self, device=0, energy_params=None):
        self._lib = N.lib()
        self._h = C.c_void_p()
        ep = C.byref(energy_params) if energy_params is not None else None
        rc = self._lib.rafft_create(C.byref(self._h), int(device), ep)
        if rc != N.RAFFT_OK:
            raise N.NativeError(rc, (self._lib.rafft_last_error(None) or b"").decode())
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.rafft_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _err(self, rc):
        return N.NativeError(rc, (self._lib.rafft_last_error(self._h) or b"").decode())

    def set_option(self, name, value):
        "tuning / diagnostics switch of the context (include/rafft_b200.h rafft_set_option)"
        rc = self._lib.rafft_set_option(self._h, name.encode(), int(value))
        if rc != N.RAFFT_OK:
            raise self._err(rc)

    def int_peak(self):
        """(issue-slot peak, IADD3/LOP3 mix) in warp-instructions per second, measured now on this
        context's device by the library's microbenchmark kernel"""
        a, b = C.c_double(), C.c_double()
        rc = self._lib.rafft_int_peak(self._h, C.byref(a), C.byref(b))
        if rc != N.RAFFT_OK:
            raise self._err(rc)
        return a.value, b.value

    # -- raw calls ---------------------------------------------------------------------
    def fold_raw(self, seq_bytes, offs, params):
        """seq_bytes: concatenated ASCII letters; offs: int64 numpy array (n+1).
        Returns a RawResult: numpy VIEWS of the library-owned (pinned) arrays, valid until its
        close() / garbage collection hands the memory back to the library."""
        offs = np.ascontiguousarray(offs, dtype=np.int64)
        n = len(offs) - 1
        res = C.POINTER(N.Result)()
        rc = self._lib.rafft_fold_batch(self._h, seq_bytes, offs.ctypes.data_as(C.POINTER(C.c_int64)),
                                        n, C.byref(params), C.byref(res))
        if rc != N.RAFFT_OK:
            raise self._err(rc)
        return RawResult(res, offs, self._lib, self)

    def fold_raw_device(self, d_seqs_ptr, d_offs_ptr, n, total_len, max_len, params):
        """Inputs already on this context's device (raw device pointers)."""
        res = C.POINTER(N.Result)()
        rc = self._lib.rafft_fold_batch_device(self._h, C.c_void_p(d_seqs_ptr), C.c_void_p(d_offs_ptr),
                                               n, total_len, max_len, C.byref(params), C.byref(res))
        if rc != N.RAFFT_OK:
            raise self._err(rc)
        return RawResult(res, None, self._lib, self)

    def eval_structures(self, seqs, structs):
        "integer energies (dcal/mol) of whole structures; replaces fold_compound.eval_structure"
        assert len(seqs) == len(structs)
        n = len(seqs)
        offs = np.zeros(n + 1, dtype=np.int64)
        for i, (s, d) in enumerate(zip(seqs, structs)):
            if len(s) != len(d):
                raise ValueError("sequence and structure %d differ in length" % i)
            offs[i + 1] = offs[i] + len(s)
        out = np.zeros(max(n, 1), dtype=np.int32)
        rc = self._lib.rafft_eval_structures(self._h, "".join(seqs).encode(), "".join(structs).encode(),
                                             offs.ctypes.data_as(C.POINTER(C.c_int64)), n,
                                             out.ctypes.data_as(C.POINTER(C.c_int32)))
        if rc != N.RAFFT_OK:
            raise self._err(rc)
        return out[:n]


def eval_structures_static(ctx, d_seqs_ptr, d_structs_ptr, d_offs_ptr, n, total_len, d_energy_ptr):
    """rafft_eval_structures_static: the evaluator on caller-owned DEVICE buffers (raw pointers) with the
    kernel's scratch in a fixed static device array -- no allocation inside the call.  Not in the
    reference; see include/rafft_b200.h for the limit (1,048,576 letters) and the sharing rule."""
    rc = ctx._lib.rafft_eval_structures_static(ctx._h, C.c_void_p(d_seqs_ptr), C.c_void_p(d_structs_ptr),
                                               C.c_void_p(d_offs_ptr), int(n), int(total_len), C.c_void_p(d_energy_ptr))
    if rc != N.RAFFT_OK:
        raise ctx._err(rc)


def fold_raw_multi(contexts, seq_bytes, offs, params):
    """One batch folded by several devices from this process (rafft_fold_batch_multi): contiguous
    ranges of the input order, one host thread + stream per device inside the library, one result
    in input order.  contexts: FoldContext objects of different devices; the first owns the result."""
    offs = np.ascontiguousarray(offs, dtype=np.int64)
    n = len(offs) - 1
    res = C.POINTER(N.Result)()
    lib = contexts[0]._lib
    hs = (C.c_void_p * len(contexts))(*[c._h.value for c in contexts])
    rc = lib.rafft_fold_batch_multi(hs, len(contexts), seq_bytes, offs.ctypes.data_as(C.POINTER(C.c_int64)),
                                    n, C.byref(params), C.byref(res))
    if rc != N.RAFFT_OK:
        raise contexts[0]._err(rc)
    return RawResult(res, offs, lib, contexts[0])


def _np(ptr, count, dtype):
    if count == 0:
        return np.zeros(0, dtype=dtype)
    return np.ctypeslib.as_array(ptr, shape=(count,)).view(dtype)


class RawResult:
    """Arrays of one rafft_result (see include/rafft_b200.h), as zero-copy views of the memory the
    library owns; close() (or garbage collection) calls rafft_result_free."""

    def __init__(self, res, offs, lib, owner=None):
        self._res, self._lib = res, lib
        self._owner = owner  # the result hands its pinned block back to this context: keep it alive
        r = res.contents
        n = r.n_seq
        self.n_seq = n
        self.kernel_ms = r.kernel_ms
        self.kernel_launches = r.kernel_launches
        self.expand_ms = r.expand_ms
        self.expand_launches = r.expand_launches
        self.iterations = r.iterations
        self.family_ms = dict(zip(N.FAMILY_NAMES, r.family_ms))
        self.family_launches = dict(zip(N.FAMILY_NAMES, r.family_launches))
        self.lanes, self.loop_mode = r.lanes, r.loop_mode
        self.h2d_bytes, self.d2h_bytes = r.h2d_bytes, r.d2h_bytes
        self.status = _np(r.status, n, np.int32)
        self.struct_off = _np(r.struct_off, n + 1, np.int64)
        self.db_off = _np(r.db_off, n + 1, np.int64)
        self.step_off = _np(r.step_off, n + 1, np.int64)
        self.counters = _np(r.counters, n * N.CNT_N, np.int64).reshape(n, N.CNT_N)
        nS, nD, nT = int(self.struct_off[n]), int(self.db_off[n]), int(self.step_off[n])
        self.n_struct = nS
        self.energy = _np(r.energy_dcal, nS, np.int32)
        self.db = _np(C.cast(r.db, C.POINTER(C.c_uint8)), nD, np.uint8)
        self.entry_off = _np(r.entry_off, nT + 1, np.int64)
        self.entry_struct = _np(r.entry_struct, int(self.entry_off[nT]) if nT else 0, np.int32)
        if offs is not None:
            self.lengths = np.diff(offs)
        else:
            ns = np.diff(self.struct_off)
            self.lengths = np.where(ns > 0, np.diff(self.db_off) // np.maximum(ns, 1), 0)

    def close(self):
        "hand the arrays back to the library; the views must not be used afterwards"
        if self._res is not None:
            # a result outliving its (explicitly closed) context is dropped, not handed back
            if self._owner is None or self._owner._h.value:
                self._lib.rafft_result_free(self._res)
            self._res = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def steps(self, i):
        """sequence i -> list of steps, each a list of (dot-bracket, E_dcal)."""
        L = int(self.lengths[i])
        s0, d0 = int(self.struct_off[i]), int(self.db_off[i])
        out = []
        for t in range(int(self.step_off[i]), int(self.step_off[i + 1])):
            step = []
            for e in range(int(self.entry_off[t]), int(self.entry_off[t + 1])):
                k = int(self.entry_struct[e])
                step.append((self.db[d0 + k * L:d0 + (k + 1) * L].tobytes().decode(), int(self.energy[s0 + k])))
            out.append(step)
        return out


def make_params(nb_mode=100, max_stack=1, max_branch=100, min_hp=3, min_nrj=0.0, traj=False,
                temp=37.0, gc_wei=3.0, au_wei=2.0, gu_wei=1.0):
    return N.Params(int(nb_mode), int(max_stack), int(max_branch), int(min_hp), 1 if traj else 0,
                    float(min_nrj), float(temp), float(gc_wei), float(au_wei), float(gu_wei))


_contexts = {}


def device_count():
    "CUDA devices the library can use"
    return int(N.lib().rafft_device_count())


def get_context(device=0):
    "process-wide context of a device, created on first use"
    ctx = _contexts.get(device)
    if ctx is None:
        ctx = _contexts[device] = FoldContext(device)
    return ctx


def _structure(db, e):
    st = Structure()
    st.str_struct = db
    st.energy_dcal = e
    st.energy = energy_from_dcal(e)
    return st


def _structures(step):
    return [_structure(db, e) for db, e in step]


_LETTER_OK = np.zeros(256, dtype=bool)
_LETTER_OK[[ord(c) for c in "ACGU"]] = True


def _check_letters(seq_bytes):
    """One vectorised pass over the joined letters.  The reference fails with KeyError in
    prep_sequence (rafft/utils.py:73-80) for letters outside its table; its `N` rows are all-zero
    and ViennaRNA treats N specially: not supported here."""
    arr = np.frombuffer(seq_bytes, dtype=np.uint8)
    bad = ~_LETTER_OK[arr]
    if bad.any():
        raise KeyError(chr(int(arr[int(np.argmax(bad))])))


class BatchFolds:
    """What fold_batch returns: a read-only sequence with one `fold`-style result per input
    sequence -- a list of Structure, or (list, trajectory) with traj=True -- built on access from
    zero-copy views of the library's result arrays (nothing is converted for sequences nobody
    looks at).  `.raw` is the RawResult (numpy views: energy, db, offsets) for bulk consumers."""

    def __init__(self, raw, traj):
        self.raw, self._traj = raw, traj
        bad = np.nonzero(raw.status != N.RAFFT_OK)[0]
        if len(bad):
            i = int(bad[0])
            raise N.NativeError(int(raw.status[i]), "sequence %d could not be folded" % i)

    def __len__(self):
        return self.raw.n_seq

    def _beam(self, i, t):
        raw = self.raw
        L = int(raw.lengths[i])
        s0, d0 = int(raw.struct_off[i]), int(raw.db_off[i])
        ks = raw.entry_struct[int(raw.entry_off[t]):int(raw.entry_off[t + 1])]
        if L == 0 or len(ks) == 0:
            return []
        n_s = int(raw.struct_off[i + 1]) - s0
        rows = raw.db[d0:d0 + n_s * L].reshape(n_s, L)[ks]       # one gather for the whole beam
        text = rows.tobytes().decode("ascii")
        es = raw.energy[s0 + ks].tolist()
        return [_structure(text[k * L:(k + 1) * L], e) for k, e in enumerate(es)]

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[k] for k in range(*i.indices(len(self)))]
        if i < 0:
            i += len(self)
        if not 0 <= i < len(self):
            raise IndexError(i)
        t0, t1 = int(self.raw.step_off[i]), int(self.raw.step_off[i + 1])
        final = self._beam(i, t1 - 1)
        if not self._traj:
            return final
        return final, [self._beam(i, t) for t in range(t0, t1 - 1)] + [final]

    def __iter__(self):
        return (self[i] for i in range(len(self)))

    def best(self):
        """(dot-bracket bytes as uint8 rows are ragged, so:) lists of the lowest-energy structure of
        every sequence and its energy in dcal/mol -- the `rafft --bench` first line, vectorised."""
        raw = self.raw
        last = raw.step_off[1:] - 1
        first_entry = raw.entry_struct[raw.entry_off[last]] + raw.struct_off[:-1]
        e = raw.energy[first_entry]
        L = raw.lengths
        start = raw.db_off[:-1] + (first_entry - raw.struct_off[:-1]) * L
        buf = raw.db
        return [buf[int(a):int(a + l)].tobytes().decode("ascii") for a, l in zip(start, L)], e


def fold_batch(sequences, nb_mode=100, max_stack=1, max_branch=100, min_hp=3, min_nrj=0.0,
               traj=False, temp=37.0, gc_wei=3.0, au_wei=2.0, gu_wei=1.0, device=0, context=None,
               devices=None):
    """Fold many sequences in one library call.  Returns a BatchFolds: one `fold`-style result per
    sequence (list of Structure, or (list, trajectory) with traj=True), materialised on access.
    devices=[d0, d1, ...] folds the batch on several GPUs from this one process (contiguous
    ranges of the input, one host thread + stream per device inside the library, results in input
    order) -- the replacement of the reference's process pool over the sequences of a set
    (benchmark_results/bench_fft.py:17-22)."""
    sequences = list(sequences)
    lens = np.fromiter(map(len, sequences), dtype=np.int64, count=len(sequences))
    if len(lens) and int(lens.min()) == 0:
        raise ValueError("empty sequence")
    seq_bytes = "".join(sequences).encode("ascii", "replace")
    _check_letters(seq_bytes)
    offs = np.zeros(len(sequences) + 1, dtype=np.int64)
    np.cumsum(lens, out=offs[1:])
    params = make_params(nb_mode, max_stack, max_branch, min_hp, min_nrj, traj, temp, gc_wei,
                         au_wei, gu_wei)
    if devices is not None and len(devices) > 1:
        raw = fold_raw_multi([get_context(d) for d in devices], seq_bytes, offs, params)
    else:
        ctx = context if context is not None else get_context(devices[0] if devices else device)
        raw = ctx.fold_raw(seq_bytes, offs, params)
    return BatchFolds(raw, traj)


def fold(sequence, nb_mode=100, max_stack=1, max_branch=100, min_hp=3, min_nrj=0.0, traj=False,
         temp=37.0, gc_wei=3.0, au_wei=2.0, gu_wei=1.0):
    "fold a given sequence (same arguments, defaults and return value as the reference)"
    return fold_batch([sequence], nb_mode, max_stack, max_branch, min_hp, min_nrj, traj, temp,
                      gc_wei, au_wei, gu_wei)[0]
