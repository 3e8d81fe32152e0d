"""Kinetics on the fast-folding graph, drop-in for the reference's rafft/rafft_kin.py.

`kinetics(fast_paths, max_time, n_steps, initial_pop=None)` keeps the reference's signature and
return value (rafft/rafft_kin.py:94-150).  The graph part -- which structures of consecutive
folding steps are connected (get_connected_prev, :48-56) and the non-redundant numbering of the
structures (:106-115) -- runs on the GPU through `rafft_kin_graph` (include/rafft_b200.h); the
Metropolis rate matrix (:68-91, numpy longdouble) and its eigen-decomposition
(scipy.linalg.eig / inv, :131-132) stay on the host as in the reference, so populations agree
with it to LAPACK rounding.  Plotting (plot_traj, :18-45) needs matplotlib, which is not
available on this box."""
import ctypes as C
from copy import deepcopy

import numpy as np

from . import _native as N
from .rafft import get_context

KT = 0.61  # rafft/rafft_kin.py:70


def folding_graph(fast_paths, device=0, context=None):
    """fast_paths: list (folding steps) of lists of objects with .str_struct.
    -> (uniq, conn, conn_off, step_off): see rafft_kin_graph in include/rafft_b200.h."""
    ctx = context if context is not None else get_context(device)
    rows = [st.str_struct for step in fast_paths for st in step]
    n_steps = len(fast_paths)
    step_off = np.zeros(n_steps + 1, dtype=np.int64)
    np.cumsum([len(step) for step in fast_paths], out=step_off[1:])
    if not rows:
        return np.zeros(0, np.int32), np.zeros(0, np.uint8), np.zeros(n_steps + 1, np.int64), step_off
    L = len(rows[0])
    if any(len(r) != L for r in rows):
        raise ValueError("structures of one folding graph must have the same length")
    sizes = np.diff(step_off)
    total = int((sizes * np.roll(sizes, 1)).sum())
    uniq = np.zeros(len(rows), dtype=np.int32)
    conn = np.zeros(max(total, 1), dtype=np.uint8)
    conn_off = np.zeros(n_steps + 1, dtype=np.int64)
    rc = ctx._lib.rafft_kin_graph(ctx._h, "".join(rows).encode(), L,
                                  step_off.ctypes.data_as(C.POINTER(C.c_int64)), n_steps,
                                  uniq.ctypes.data_as(C.POINTER(C.c_int32)),
                                  conn.ctypes.data_as(C.POINTER(C.c_uint8)), total,
                                  conn_off.ctypes.data_as(C.POINTER(C.c_int64)))
    if rc == N.RAFFT_ERR_SEQ:
        raise IndexError("pop from empty list")  # what paired_positions raises (rafft/utils.py:62-65)
    if rc != N.RAFFT_OK:
        raise ctx._err(rc)
    return uniq, conn[:total], conn_off, step_off


def get_transition_mat(fast_paths, nb_struct, struct_map, graph=None, device=0):
    """Metropolis rates between connected structures of consecutive steps
    (rafft/rafft_kin.py:68-91); `graph` = folding_graph(fast_paths)."""
    uniq, conn, conn_off, step_off = graph if graph is not None else folding_graph(fast_paths, device)
    transition_mat = np.zeros((nb_struct, nb_struct), dtype=np.longdouble)
    flat = [st for step in fast_paths for st in step]
    n_steps = len(fast_paths)
    for t in range(n_steps):
        tp = n_steps - 1 if t == 0 else t - 1
        n_cur, n_prev = int(step_off[t + 1] - step_off[t]), int(step_off[tp + 1] - step_off[tp])
        if n_cur == 0 or n_prev == 0:
            continue
        block = conn[int(conn_off[t]):int(conn_off[t + 1])].reshape(n_cur, n_prev)
        for ci, pi in zip(*np.nonzero(block)):
            cur, prev = flat[int(step_off[t]) + int(ci)], flat[int(step_off[tp]) + int(pi)]
            map_cur, cur_nrj = struct_map[cur.str_struct]
            map_prev, prev_nrj = struct_map[prev.str_struct]
            delta_nrj = cur_nrj - prev_nrj
            if map_cur != map_prev:
                transition_mat[map_prev, map_cur] = min(1.0, np.exp(-delta_nrj / KT))
                transition_mat[map_cur, map_prev] = min(1.0, np.exp(delta_nrj / KT))
    for si in range(nb_struct):
        transition_mat[si, si] = -transition_mat[si, :].sum()
    return transition_mat


def kinetics(fast_paths, max_time, n_steps, initial_pop=None, device=0):
    """Same inputs and outputs as the reference (rafft/rafft_kin.py:94-150):
    -> trajectory, times, struct_list, str_equi_pop."""
    from scipy.linalg import eig, inv
    graph = folding_graph(fast_paths, device)
    uniq = graph[0]
    flat = [st for step in fast_paths for st in step]
    # the first occurrence of every structure, in order (rafft_kin.py:106-115)
    struct_list = [flat[i] for i in range(len(flat)) if uniq[i] == i]
    struct_map = {st.str_struct: (si, st.energy) for si, st in enumerate(struct_list)}
    nb_struct = len(struct_list)
    transition_mat = get_transition_mat(fast_paths, nb_struct, struct_map, graph)

    if initial_pop is None:
        init_pop = np.array([1.0] + [0.0 for _ in range(nb_struct - 1)], dtype=np.longdouble)
    else:
        init_pop = np.array([0.0 for _ in range(nb_struct)], dtype=np.longdouble)
        for p, w in initial_pop:
            init_pop[p] = w
    trajectory = [deepcopy(init_pop)]

    V, W = eig(transition_mat.T, check_finite=True)
    iW = inv(W)
    time_step = max_time / n_steps
    times = [np.exp(-4)]
    for st in range(n_steps):
        time = np.exp(time_step * st - 4)
        times += [time]
        tmp_pop = W @ np.diag(np.exp(V * time)) @ (iW @ init_pop)
        trajectory += [tmp_pop.real / tmp_pop.real.sum()]
    equi_pop = trajectory[-1]
    str_equi_pop = [(st.str_struct, st.energy, ep, struct_map[st.str_struct][0])
                    for st, ep in zip(struct_list, equi_pop.real)]
    return trajectory, times, struct_list, str_equi_pop


def plot_traj(*args, **kwargs):
    raise RuntimeError("plot_traj needs matplotlib, which is unavailable on this box "
                       "(reference rafft/rafft_kin.py:18-45 is a plotting helper, out of scope)")
