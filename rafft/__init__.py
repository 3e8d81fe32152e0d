"""`rafft` -- the reference's import names over the B200 implementation.

The reference package exports `from rafft.rafft import fold` and `from rafft.rafft_kin import
kinetics` (rafft/__init__.py:1-2, setup.py:9-10); user code written against it
(README.org:171-180: `from rafft import fold`, `from rafft.utils import parse_rafft_output`)
imports unchanged.  Everything here re-exports rafft_b200; the numeric work runs in the CUDA
library behind include/rafft_b200.h."""
from rafft.rafft import fold  # noqa: F401
from rafft.rafft_kin import kinetics  # noqa: F401
