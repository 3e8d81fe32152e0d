"""ctypes binding of the C-ABI in include/rafft_b200.h (librafft_b200.so).

This is the only way into the product path.  It fails loudly when the CUDA library is
missing or no CUDA device is usable; there is no CPU fallback and nothing here touches
oracle/."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "librafft_b200.so")

RAFFT_OK, RAFFT_ERR_ARG, RAFFT_ERR_SEQ, RAFFT_ERR_CUDA, RAFFT_ERR_CAPACITY = 0, 1, 2, 3, 4
MAX_SPECIAL = 64
COUNTER_NAMES = ("steps", "node_exp", "w_corr", "w_scan", "n_eval", "uniq_nodes", "uniq_w_corr",
                 "uniq_w_scan", "combos", "node_pos", "t0_ns", "t1_ns", "n_eval_expand", "trivial_nodes")
CNT_N = len(COUNTER_NAMES)

# every symbol include/rafft_b200.h declares (tests check the library exports all of them)
EXPORTS = ("rafft_version", "rafft_default_energy_params", "rafft_create", "rafft_destroy",
           "rafft_last_error", "rafft_fold_batch", "rafft_fold_batch_device", "rafft_result_free",
           "rafft_eval_structures", "rafft_kin_graph", "rafft_bp_distance_matrix",
           "rafft_fold_batch_multi", "rafft_set_option", "rafft_int_peak", "rafft_plan_ranges", "rafft_device_count",
           "rafft_eval_structures_static")
STATIC_EVAL_MAX_LETTERS = 1 << 20  # RAFFT_STATIC_EVAL_MAX_LETTERS


class EnergyParams(C.Structure):
    """struct rafft_energy_params (pair types 0..5 = CG GC GU UG AU UA, bases A C G U)."""
    _fields_ = [
        ("stack", C.c_int32 * 36),
        ("hairpin", C.c_int32 * 31), ("bulge", C.c_int32 * 31), ("interior", C.c_int32 * 31),
        ("mismatchH", C.c_int32 * 96), ("mismatchI", C.c_int32 * 96), ("mismatch1nI", C.c_int32 * 96),
        ("mismatch23I", C.c_int32 * 96), ("mismatchM", C.c_int32 * 96), ("mismatchExt", C.c_int32 * 96),
        ("dangle5", C.c_int32 * 24), ("dangle3", C.c_int32 * 24),
        ("int11", C.c_int32 * 576), ("int21", C.c_int32 * 2304), ("int22", C.c_int32 * 9216),
        ("ninio", C.c_int32), ("max_ninio", C.c_int32), ("ml_base", C.c_int32),
        ("ml_intern", C.c_int32), ("ml_closing", C.c_int32), ("terminal_au", C.c_int32),
        ("lxc", C.c_double),
        ("n_tri", C.c_int32), ("n_tetra", C.c_int32), ("n_hexa", C.c_int32),
        ("tri", (C.c_char * 8) * MAX_SPECIAL), ("tetra", (C.c_char * 8) * MAX_SPECIAL),
        ("hexa", (C.c_char * 12) * MAX_SPECIAL),
        ("tri_e", C.c_int32 * MAX_SPECIAL), ("tetra_e", C.c_int32 * MAX_SPECIAL),
        ("hexa_e", C.c_int32 * MAX_SPECIAL),
    ]


class Params(C.Structure):
    """struct rafft_params: the arguments of fold() (reference rafft/rafft.py:219-221)."""
    _fields_ = [("nb_mode", C.c_int32), ("max_stack", C.c_int32), ("max_branch", C.c_int32),
                ("min_hp", C.c_int32), ("traj", C.c_int32), ("min_nrj", C.c_double),
                ("temp", C.c_double), ("gc_wei", C.c_double), ("au_wei", C.c_double),
                ("gu_wei", C.c_double)]


class Result(C.Structure):
    """struct rafft_result (host arrays owned by the library until rafft_result_free)."""
    _fields_ = [("n_seq", C.c_int32),
                ("status", C.POINTER(C.c_int32)),
                ("struct_off", C.POINTER(C.c_int64)),
                ("energy_dcal", C.POINTER(C.c_int32)),
                ("db_off", C.POINTER(C.c_int64)),
                ("db", C.POINTER(C.c_char)),
                ("step_off", C.POINTER(C.c_int64)),
                ("entry_off", C.POINTER(C.c_int64)),
                ("entry_struct", C.POINTER(C.c_int32)),
                ("counters", C.POINTER(C.c_int64)),
                ("kernel_ms", C.c_double),
                ("kernel_launches", C.c_int64),
                ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64),
                ("expand_ms", C.c_double), ("expand_launches", C.c_int64),
                ("iterations", C.c_int64),
                ("family_ms", C.c_double * 8), ("family_launches", C.c_int64 * 8),
                ("lanes", C.c_int32), ("loop_mode", C.c_int32)]


FAMILY_NAMES = ("gather", "expand_fast", "expand_generic", "combine", "select", "materialise", "advance", "other")
# kernel names (as ncu prints them) of every family
FAMILY_KERNELS = {"gather": ("k_gather",), "expand_fast": ("k_expand_fast", "k_expand_big"), "expand_generic": ("k_expand",),
                  "combine": ("k_combine",), "select": ("k_select",), "materialise": ("k_materialise", "k_beam"),
                  "advance": ("k_advance",), "other": ("k_init", "k_finish", "k_ctl_reset")}


class NativeError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("rafft_b200 error %d: %s" % (code, msg))
        self.code = code


_lib = None


def lib():
    """Load librafft_b200.so (once).  Raises if it has not been built: no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "rafft_b200: %s is missing. Build it with `python -m rafft_b200.build` (needs nvcc). "
            "There is no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    L.rafft_version.restype = C.c_char_p
    L.rafft_default_energy_params.argtypes = [C.POINTER(EnergyParams)]
    L.rafft_default_energy_params.restype = None
    L.rafft_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(EnergyParams)]
    L.rafft_create.restype = C.c_int
    L.rafft_destroy.argtypes = [C.c_void_p]
    L.rafft_destroy.restype = None
    L.rafft_last_error.argtypes = [C.c_void_p]
    L.rafft_last_error.restype = C.c_char_p
    L.rafft_fold_batch.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_int64), C.c_int32,
                                   C.POINTER(Params), C.POINTER(C.POINTER(Result))]
    L.rafft_fold_batch.restype = C.c_int
    L.rafft_fold_batch_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64,
                                          C.c_int32, C.POINTER(Params), C.POINTER(C.POINTER(Result))]
    L.rafft_fold_batch_device.restype = C.c_int
    L.rafft_fold_batch_multi.argtypes = [C.POINTER(C.c_void_p), C.c_int32, C.c_char_p, C.POINTER(C.c_int64),
                                         C.c_int32, C.POINTER(Params), C.POINTER(C.POINTER(Result))]
    L.rafft_fold_batch_multi.restype = C.c_int
    L.rafft_plan_ranges.argtypes = [C.POINTER(C.c_int64), C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32)]
    L.rafft_plan_ranges.restype = C.c_int
    L.rafft_device_count.restype = C.c_int
    L.rafft_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_int64]
    L.rafft_set_option.restype = C.c_int
    L.rafft_int_peak.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.rafft_int_peak.restype = C.c_int
    L.rafft_result_free.argtypes = [C.POINTER(Result)]
    L.rafft_result_free.restype = None
    L.rafft_eval_structures.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.POINTER(C.c_int64),
                                        C.c_int32, C.POINTER(C.c_int32)]
    L.rafft_eval_structures.restype = C.c_int
    L.rafft_eval_structures_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64,
                                               C.c_void_p]
    L.rafft_eval_structures_static.restype = C.c_int
    L.rafft_kin_graph.argtypes = [C.c_void_p, C.c_char_p, C.c_int32, C.POINTER(C.c_int64), C.c_int32,
                                  C.POINTER(C.c_int32), C.POINTER(C.c_uint8), C.c_int64, C.POINTER(C.c_int64)]
    L.rafft_kin_graph.restype = C.c_int
    L.rafft_bp_distance_matrix.argtypes = [C.c_void_p, C.c_char_p, C.c_int32, C.c_int32, C.POINTER(C.c_int32)]
    L.rafft_bp_distance_matrix.restype = C.c_int
    _lib = L
    return L


def default_energy_params():
    ep = EnergyParams()
    lib().rafft_default_energy_params(C.byref(ep))
    return ep
