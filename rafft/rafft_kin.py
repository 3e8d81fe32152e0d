"rafft.rafft_kin of the reference (rafft/rafft_kin.py: kinetics, get_transition_mat, ...) -> rafft_b200.rafft_kin"
from rafft_b200.rafft_kin import *  # noqa: F401,F403
from rafft_b200.rafft_kin import kinetics  # noqa: F401
