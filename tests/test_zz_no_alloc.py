"""rafft_eval_structures_static -- the evaluator with its scratch in a static device array, for callers
that cannot allocate (no counterpart in the reference).
Through the C-ABI, against the existing entry point rafft_eval_structures, bit-exact (integer
dcal/mol: the tolerance of every energy test here is 0), on two calls in a row that share the array.

Collected last on purpose; passes alone in a fresh process and
after the other tests (it only needs the session's context)."""
import numpy as np
import pytest

from rafft_b200.synth import config_batch, splitmix_seq

pytestmark = pytest.mark.gpu

I32_MIN = -(2 ** 31)


def _structures(ctx, seqs, nb_mode=100, max_stack=12):
    "(sequence, dot-bracket) pairs: every structure of the trajectories of a small fold on the GPU"
    from rafft_b200 import rafft as R
    offs = np.zeros(len(seqs) + 1, dtype=np.int64)
    np.cumsum([len(s) for s in seqs], out=offs[1:])
    raw = ctx.fold_raw("".join(seqs).encode(), offs, R.make_params(nb_mode, max_stack, 1000, traj=True))
    assert (raw.status == 0).all()
    out = []
    for i, s in enumerate(seqs):
        seen = set()
        for step in raw.steps(i):
            for db, _ in step:
                if db not in seen:
                    seen.add(db)
                    out.append((s, db))
    raw.close()
    return out


def _static_eval(ctx, pairs, torch, total_len=None, out_fill=12345):
    "one call of the new entry point on device buffers the CALLER (this test, through torch) owns"
    from rafft_b200 import rafft as R
    dev = torch.device("cuda", ctx.device)
    seqs = "".join(s for s, _ in pairs).encode()
    dbs = "".join(d for _, d in pairs).encode()
    offs = np.zeros(len(pairs) + 1, dtype=np.int64)
    np.cumsum([len(s) for s, _ in pairs], out=offs[1:])
    d_seqs = torch.frombuffer(bytearray(seqs or b"\0"), dtype=torch.uint8).to(dev)
    d_dbs = torch.frombuffer(bytearray(dbs or b"\0"), dtype=torch.uint8).to(dev)
    d_offs = torch.from_numpy(offs).to(dev)
    d_out = torch.full((max(len(pairs), 1),), out_fill, dtype=torch.int32, device=dev)
    torch.cuda.synchronize(dev)       # inputs complete before the call (the library uses its own stream)
    R.eval_structures_static(ctx, d_seqs.data_ptr(), d_dbs.data_ptr(), d_offs.data_ptr(), len(pairs),
                             int(offs[-1]) if total_len is None else total_len, d_out.data_ptr())
    return d_out.cpu().numpy()[:len(pairs)]


def test_static_eval_matches_existing_entry_point_twice(ctx):
    import torch
    from rafft_b200 import _native as N
    # call 1: ~4,000 structures of 60 sequences (50-500 nt), plus malformed / non-canonical / odd-letter rows
    pairs1 = _structures(ctx, config_batch("cfg2", 60))
    pairs1 += [("GGGAAACCC", "(((...)))"), ("GGGAAACCC", "(((...))."), ("GGGAAACCC", "((.....))"), ("GGGANACCC", "(((...)))"),
               ("AAAAAAAAA", "(((...)))"), ("G", "."), (splitmix_seq(77, 3000), "." * 3000)]
    want1 = ctx.eval_structures([s for s, _ in pairs1], [d for _, d in pairs1])
    got1 = _static_eval(ctx, pairs1, torch)
    assert (got1 == want1).all(), np.nonzero(got1 != want1)[0][:5]
    assert want1[-6] == I32_MIN and want1[-4] == I32_MIN and want1[-3] == I32_MIN    # the error rows really are error rows
    # call 2, right after, on other (fewer, longer, differently aligned) structures: the array still holds
    # call 1's pair tables and base codes, and nothing of them may leak into the result
    pairs2 = _structures(ctx, [splitmix_seq(5000 + k, 700 + 37 * k) for k in range(6)], nb_mode=50, max_stack=5)[::-1]
    want2 = ctx.eval_structures([s for s, _ in pairs2], [d for _, d in pairs2])
    got2 = _static_eval(ctx, pairs2, torch)
    assert (got2 == want2).all(), np.nonzero(got2 != want2)[0][:5]
    # and call 1's input once more
    assert (_static_eval(ctx, pairs1, torch) == want1).all()
    assert len(pairs1) > 1000 and len(pairs2) > 20 and (want1 != I32_MIN).sum() > 1000


def test_static_eval_limits(ctx):
    import torch
    from rafft_b200 import _native as N
    from rafft_b200 import rafft as R
    # beyond the stated maximum: refused with a message that names the limit, nothing launched
    one = [("GGGAAACCC", "(((...)))")]
    with pytest.raises(N.NativeError) as ei:
        _static_eval(ctx, one, torch, total_len=N.STATIC_EVAL_MAX_LETTERS + 1)
    assert ei.value.code == N.RAFFT_ERR_ARG and "1048576" in str(ei.value) and "RAFFT_STATIC_EVAL_MAX_LETTERS" in str(ei.value)
    # exactly at the maximum (256 structures of 4096 letters): accepted, equal to the existing entry point
    big = [(splitmix_seq(9000 + k, 4096), "." * 4096) for k in range(N.STATIC_EVAL_MAX_LETTERS // 4096)]
    big[3] = (big[3][0], "(" + "." * 4094 + ")")
    want = ctx.eval_structures([s for s, _ in big], [d for _, d in big])
    assert (_static_eval(ctx, big, torch) == want).all()
    # offsets that lie about the total: the kernel refuses those rows instead of leaving the array
    two = [("GGGAAACCC", "(((...)))"), ("GGGGAAAACCCC", "((((....))))")]
    got = _static_eval(ctx, two, torch, total_len=9)
    assert got[0] == ctx.eval_structures(["GGGAAACCC"], ["(((...)))"])[0] and got[1] == I32_MIN
    # an empty batch and NULL arguments
    R.eval_structures_static(ctx, 0, 0, 0, 0, 0, 0)
    with pytest.raises(N.NativeError):
        R.eval_structures_static(ctx, 0, 0, 0, 1, 9, 0)
