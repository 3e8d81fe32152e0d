"""ORACLE shim (test infrastructure, NOT the product): the two ViennaRNA names the reference
imports (rafft/utils.py:7 `from RNA import fold_compound, md`), backed by the C restatement
in oracle/turner_eval.c.  Lets the UNMODIFIED reference Python run in this container:
    PYTHONPATH=oracle/shim:/root/reference python /root/reference/bin/rafft ...
eval_structure returns float32(E_dcal/100) widened to double, like ViennaRNA's
`(float)energy / 100.` seen through SWIG."""
import ctypes
import os
import struct

_here = os.path.dirname(os.path.abspath(__file__))
_lib = ctypes.CDLL(os.path.join(_here, "..", "..", "_ref", "liboracle.so"))
_lib.oracle_eval_db.restype = ctypes.c_int
_P = ctypes.create_string_buffer(400000)
_lib.oracle_default_params(_P)
N_EVAL = [0]


class md:
    def __init__(self):
        self.temperature = 37.0


class fold_compound:
    def __init__(self, sequence, model=None):
        if model is not None and float(model.temperature) != 37.0:
            raise NotImplementedError("oracle shim: only 37.0 C (built-in Turner-2004 tables)")
        self.sequence = sequence
        self._s = sequence.encode()

    def eval_structure(self, structure):
        N_EVAL[0] += 1
        e = _lib.oracle_eval_db(_P, self._s, structure.encode(), len(self._s))
        return struct.unpack("f", struct.pack("f", e / 100.0))[0]
