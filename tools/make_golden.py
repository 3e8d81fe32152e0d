"""Generate tests/golden/* from the reference tree (/root/reference; build container only).

  rafft.out, rafft_20.out     copies of example/rafft.out, example/rafft_20.out
  analysis_n50_ms5.txt        the recorded `-n 50 -ms 5` run, analysis.org:723-754
  energy_kat.csv.gz           the 11,505 (sequence, structure, E_dcal) triples of
                              benchmark_results/*_scores.csv (nrj = RNA.energy_of_struct,
                              benchmark_results/scoring.py:125)
  kernel_kat.json             constants of the Rust unit tests that pin the leaf kernels
                              (rafft-rs/src/autocorrelation.rs:101-180, encoding.rs:531-563,
                              vienna.rs:92-109)
  ref_folds.json.gz           trajectories produced HERE by the unmodified reference Python
                              (rafft.rafft.fold through oracle/shim) on seeded inputs
The -m "not gpu" tests check the oracle against all of these; the -m gpu tests check the
CUDA path against the oracle and against these files.  Nothing on the GPU box reads
/root/reference.
"""
import csv, glob, gzip, io, json, os, random, re, shutil, sys, time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, HERE)
from refrun import ref_fold, splitmix_seq  # noqa: E402


def floats(txt):
    return [float(x) for x in re.findall(r"-?\d+\.?\d*(?:e-?\d+)?", txt)]


def kernel_kat():
    ac = open(REF + "/rafft-rs/src/autocorrelation.rs").read()
    vecs = re.findall(r"vec!\[(.*?)\]", ac, re.S)
    # order in file: a, bb, (commented b), ab, _ac
    a, bb, ab, acn = floats(vecs[0]), floats(vecs[1]), floats(vecs[3]), floats(vecs[4])
    assert len(a) == 82 and len(bb) == 82 and len(ab) == 163 and len(acn) == 163
    seq = re.search(r'"([ACGU]{82})"', ac).group(1)
    enc = open(REF + "/rafft-rs/src/encoding.rs").read()
    t = enc[enc.index("fn test_consecutivepairs"):]
    seq14 = re.search(r'let sequence = "([ACGU]+)"', t).group(1)
    split16 = re.search(r'^\s*let tobesplit = "([ACGU]+)"', t, re.M).group(1)  # not the commented-out one
    cp = re.findall(r"(encoded|splitenc)\.consecutive_pairs_at_lag\((\d+), (\d+)\), "
                    r"\((\d+), (\d+), (\d+), ([\d.]+)\)", t)
    whole = [[int(l), int(mh), int(a1), int(b1), int(c1), float(s)]
             for w, l, mh, a1, b1, c1, s in cp if w == "encoded"]
    split = [[int(l), int(mh), int(a1), int(b1), int(c1), float(s)]
             for w, l, mh, a1, b1, c1, s in cp if w == "splitenc"]
    assert len(whole) == 13 and len(split) == 2
    vi = open(REF + "/rafft-rs/src/vienna.rs").read()
    pt = [int(x) for x in re.findall(r"-?\d+", re.search(r"from_vec\(vec!\[(.*?)\]", vi, re.S).group(1))]
    assert pt[0] == 82 and len(pt) == 83
    return {
        "convolution": {"a": a, "b": bb, "ab": ab, "cite": "rafft-rs/src/autocorrelation.rs:101-140"},
        "autocorrelation": {"seq": seq, "weights": {"gc": 3.0, "au": 2.0, "gu": 1.0}, "ac": acn,
                            "cite": "rafft-rs/src/autocorrelation.rs:142-180"},
        "consecutive_pairs": {"seq": seq14, "cases": whole, "cite": "rafft-rs/src/encoding.rs:531-553",
                              "columns": ["lag", "min_hp", "max_nb", "max_i", "max_j", "score"]},
        "consecutive_pairs_split": {"seq": split16, "pos": list(range(5)) + list(range(11, 16)),
                                    "cases": split, "cite": "rafft-rs/src/encoding.rs:555-562"},
        "vrna_eval": {"seq": seq, "pt_1based": pt, "energy_dcal": -2580,
                      "cite": "rafft-rs/src/vienna.rs:92-109"},
    }


def energy_kat():
    tri = {}
    for f in sorted(glob.glob(REF + "/benchmark_results/*_scores.csv")):
        for r in csv.DictReader(open(f)):
            seq, st = r["seq"], r["struct"]
            if len(seq) != len(st) or set(seq) - set("ACGU") or set(st) - set("()."):
                continue
            tri[(seq, st)] = int(round(float(r["nrj"]) * 100))
    buf = io.StringIO()
    w = csv.writer(buf, lineterminator="\n")
    for (seq, st), e in sorted(tri.items(), key=lambda kv: (len(kv[0][0]), kv[0])):
        w.writerow([seq, st, e])
    with gzip.GzipFile(os.path.join(OUT, "energy_kat.csv.gz"), "wb", mtime=0) as g:
        g.write(buf.getvalue().encode())
    return len(tri)


def ref_cases():
    """(name, seq, kwargs) fed to the unmodified reference."""
    fs = open(REF + "/example/rafft.out").readline().strip()
    cases = [
        ("cfg1_L120_seed1_n100_ms50", splitmix_seq(1, 120), dict(nb_mode=100, max_stack=50)),
        ("frameshift_n100_ms50", fs, dict(nb_mode=100, max_stack=50)),
        ("frameshift_fold_defaults_ms20_mb100", fs, dict(nb_mode=100, max_stack=20, max_branch=100)),
        ("L200_seed200_n100_ms20", splitmix_seq(200, 200), dict(nb_mode=100, max_stack=20)),
        ("L300_seed7_n100_ms10", splitmix_seq(7, 300), dict(nb_mode=100, max_stack=10)),
        ("L150_seed9_n200_ms100", splitmix_seq(9, 150), dict(nb_mode=200, max_stack=100)),
        ("L60_mn_neg1", splitmix_seq(11, 60), dict(nb_mode=50, max_stack=10, min_nrj=-1.0)),
        ("L60_mn_pos", splitmix_seq(12, 60), dict(nb_mode=50, max_stack=5, min_nrj=0.75)),
        ("L80_mh5", splitmix_seq(13, 80), dict(nb_mode=100, max_stack=10, min_hp=5)),
        ("L80_weights_111", splitmix_seq(14, 80), dict(nb_mode=100, max_stack=10, gc=1.0, au=1.0, gu=1.0)),
        ("L80_weights_half", splitmix_seq(15, 80), dict(nb_mode=100, max_stack=10, gc=2.5, au=1.5, gu=0.5)),
        ("L1", "G", dict(nb_mode=100, max_stack=5)),
        ("L2", "GC", dict(nb_mode=100, max_stack=5)),
        ("L5", "GGACC", dict(nb_mode=100, max_stack=5)),
        ("L8_hairpin", "GGGAAACC", dict(nb_mode=100, max_stack=5)),
        ("polyA40", "A" * 40, dict(nb_mode=100, max_stack=5)),
        ("GC_repeat", "GC" * 30, dict(nb_mode=100, max_stack=10)),
        ("long_GC_stem_fp_overflow", "G" * 45 + "AAAA" + "C" * 45, dict(nb_mode=20, max_stack=3)),
        ("mb1", splitmix_seq(16, 90), dict(nb_mode=100, max_stack=10, max_branch=1)),
        ("mb3", splitmix_seq(17, 90), dict(nb_mode=100, max_stack=10, max_branch=3)),
        ("n1", splitmix_seq(18, 90), dict(nb_mode=1, max_stack=10)),
        ("n3_ms50", splitmix_seq(19, 90), dict(nb_mode=3, max_stack=50)),
    ]
    rnd = random.Random(20261104)
    for k in range(40):
        L = rnd.choice([12, 20, 28, 35, 50, 64, 82, 100, 120, 150, 200, 250])
        n = rnd.choice([5, 20, 50, 100, 200])
        ms = rnd.choice([1, 2, 5, 10, 20, 50])
        mb = rnd.choice([3, 10, 100, 1000])
        cases.append(("rnd%02d_L%d_n%d_ms%d_mb%d" % (k, L, n, ms, mb), splitmix_seq(5000 + k, L),
                      dict(nb_mode=n, max_stack=ms, max_branch=mb)))
    return cases


def main():
    os.makedirs(OUT, exist_ok=True)
    shutil.copyfile(REF + "/example/rafft.out", OUT + "/rafft.out")
    shutil.copyfile(REF + "/example/rafft_20.out", OUT + "/rafft_20.out")
    lines = open(REF + "/analysis.org").read().split("\n")
    k0 = next(i for i, l in enumerate(lines) if "-n 50 -ms 5" in l)
    k1 = next(i for i in range(k0, len(lines)) if lines[i].startswith("#+begin_example"))
    k2 = next(i for i in range(k1, len(lines)) if lines[i].startswith("#+end_example"))
    blk = lines[k1 + 1:k2]
    assert blk[0].startswith("# ---------0") and blk[-1].startswith("...."), (blk[0], blk[-1])
    open(OUT + "/analysis_n50_ms5.txt", "w").write("\n".join(blk) + "\n")
    json.dump(kernel_kat(), open(OUT + "/kernel_kat.json", "w"), indent=0)
    print("energy triples:", energy_kat())
    out = []
    t0 = time.time()
    for name, seq, kw in ref_cases():
        t = time.time()
        steps = ref_fold(seq, **kw)
        p = dict(nb_mode=100, max_stack=1, max_branch=1000, min_hp=3, min_nrj=0.0, gc=3.0, au=2.0, gu=1.0)
        p.update(kw)
        out.append({"name": name, "seq": seq, "params": p, "steps": steps})
        print("%-40s L=%d steps=%d  %.1fs" % (name, len(seq), len(steps), time.time() - t), flush=True)
    with gzip.GzipFile(OUT + "/ref_folds.json.gz", "wb", mtime=0) as g:
        g.write(json.dumps(out, separators=(",", ":")).encode())
    print("reference folds:", len(out), "in %.0fs" % (time.time() - t0))


if __name__ == "__main__":
    main()
