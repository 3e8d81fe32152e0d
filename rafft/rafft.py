"rafft.rafft of the reference (rafft/rafft.py:219-239 `fold`) -> rafft_b200.rafft"
from rafft_b200.rafft import fold, fold_batch  # noqa: F401
