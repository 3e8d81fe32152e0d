"""How often do folds touch energy-table entries that no reference artefact pins?

The built-in Turner-2004 tables are reconstructed: entries determined by the
reference's 11,505 (sequence, structure, energy) triples are pinned, the rest follow published rules.
This tool folds a stratified sample of a benchmark batch with the ORACLE (CPU; tooling, not the product),
decomposes every structure of every final beam into table keys (tools/loopdecomp.py) and counts the
structures / sequences whose energy involves at least one key that tools/_fit/determined.json does not
list as determined by the data.  usage: unpinned_audit.py cfg2|cfg3 [stride]"""
import collections, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import pyoracle
from loopdecomp import decompose
from rafft_b200.synth import config_batch

wl = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
stride = int(sys.argv[2]) if len(sys.argv) > 2 else 25
n, ms = (100, 50) if wl == "cfg2" else (200, 200)
det = set(json.load(open(os.path.join(ROOT, "tools", "_fit", "determined.json")))["determined"].keys())
seqs = config_batch("cfg2")[::stride]
steps, _ = pyoracle.fold_batch(seqs, n, ms, 1000)
n_struct = n_struct_unp = n_seq_best_unp = n_seq_any_unp = 0
by_table = collections.Counter()
for s, st in zip(seqs, steps):
    any_unp = False
    for k, (db, e) in enumerate(st[-1]):
        keys, _ = decompose(s, db)
        unp = [key for key in keys if str(key) not in det]
        n_struct += 1
        if unp:
            n_struct_unp += 1
            any_unp = True
            if k == 0:
                n_seq_best_unp += 1
            for key in unp:
                by_table[key[0]] += 1
    n_seq_any_unp += any_unp
rec = {"workload": wl, "nb_mode": n, "max_stack": ms, "sequences": len(seqs), "sample": "every %dth sequence of the 10k batch" % stride,
       "final_structures": n_struct, "final_structures_touching_unpinned": n_struct_unp,
       "fraction_structures": n_struct_unp / max(n_struct, 1),
       "sequences_whose_best_structure_touches_unpinned": n_seq_best_unp,
       "sequences_with_any_final_structure_touching_unpinned": n_seq_any_unp,
       "unpinned_key_uses_by_table": dict(by_table.most_common()),
       "note": "pinned = determined by the reference's own (sequence, structure, energy) triples; an unpinned entry follows the "
               "published Turner-2004 construction rule and is unverified against ViennaRNA (absent here)"}
print(json.dumps(rec, indent=1))
