"rafft.utils of the reference (Structure, dot_bracket, paired_positions, read_fasta, parse_rafft_output) -> rafft_b200.utils"
from rafft_b200.utils import (Structure, dot_bracket, energy_from_dcal, format_trajectory,  # noqa: F401
                              paired_positions, parse_rafft_output, read_fasta)
