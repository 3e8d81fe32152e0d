"""rafft_b200 -- B200-native implementation of RAFFT's folding hot path.

`fold` / `fold_batch` are drop-ins for the reference's `rafft.fold` (rafft/__init__.py:1);
everything numeric runs in the sm_100a CUDA library behind include/rafft_b200.h."""
from .rafft import FoldContext, fold, fold_batch, get_context, make_params  # noqa: F401
from .utils import (Structure, dot_bracket, paired_positions, parse_rafft_output,  # noqa: F401
                    read_fasta, format_trajectory, energy_from_dcal)

__version__ = "0.1"
