"""ORACLE -- test infrastructure, NOT the product.  Only tests/ may import this.

Plain-Python restatement of the folding-graph part of the reference's kinetics:
  paired_positions     rafft/utils.py:53-67   (two piles: "(" "<" / ")" ">" and "[" / "]")
  get_connected_prev   rafft/rafft_kin.py:48-56  (every pair of the earlier structure is a pair
                       of the later one)
  the step loop of get_transition_mat  rafft/rafft_kin.py:72-75 (step 0 looks at fast_paths[-1])
  the non-redundant numbering of kinetics  rafft/rafft_kin.py:106-115
Parity pinning: tests/test_kin.py checks it against tests/golden/kin_golden.json.gz, produced by
the UNMODIFIED reference (rafft.rafft_kin, rafft.utils) in the build container by
tools/make_golden_kin.py."""


def paired_positions(structure):
    pile_reg, pile_pk, pairs = [], [], []
    for i, c in enumerate(structure):
        if c in "<(":
            pile_reg.append(i)
        elif c == "[":
            pile_pk.append(i)
        elif c in ">)":
            pairs.append((pile_reg.pop(), i))
        elif c == "]":
            pairs.append((pile_pk.pop(), i))
    return pairs


def connected_prev(cur, prev_step):
    cur_pairs = set(paired_positions(cur))
    return [si for si, s in enumerate(prev_step) if len(set(paired_positions(s)) - cur_pairs) == 0]


def folding_graph(steps):
    """steps: list of lists of dot-bracket strings.  -> (uniq, conn): uniq[i] = index of the first
    equal structure in the concatenated order; conn[t][c] = list of connected indices of step t-1
    (step 0: the last step)."""
    flat = [s for step in steps for s in step]
    first = {}
    uniq = []
    for i, s in enumerate(flat):
        first.setdefault(s, i)
        uniq.append(first[s])
    conn = [[connected_prev(s, steps[t - 1]) for s in step] for t, step in enumerate(steps)]
    return uniq, conn
