#!/bin/bash
# profiles/ files of one tools/evidence_r02.sh run:  tools/profile_summary_r02.sh <tag> "<title>"
T=${1:?tag}; TITLE=${2:-}
G=gpurun_out
for c in cfg2 cfg3 cfg4; do [ -s $G/${T}_bench_$c.json ] && cp $G/${T}_bench_$c.json profiles/${T}_bench_$c.json; done
[ -s $G/${T}_bench_cfg2_ref.json ] && cp $G/${T}_bench_cfg2_ref.json profiles/${T}_bench_cfg2_reference_arm.json
for c in cfg2 cfg3 cfg4; do cp $G/${T}_counters_$c.txt profiles/${T}_counters_$c.txt; cp $G/${T}_counters_$c.csv.gz profiles/${T}_launches_$c.csv.gz; done
cp $G/${T}_kernel_counters.json profiles/r02_kernel_counters.json
{
echo "# ncu summary ${T}: ${TITLE}"
echo
echo "Produced by \`tools/evidence_r02.sh ${T}\` on one B200 box: \`pytest -m gpu\` ($(tail -1 $G/${T}_pytest.txt | sed 's/ in .*//')), \`smoke()\`, the bench lines"
echo "\`${T}_bench_cfg{2,3,4}.json\` + \`${T}_bench_cfg2_reference_arm.json\`; then, each only after the same command had exited 0 without ncu:"
echo "per workload one metrics pass over every launch (\`--metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum"
echo "--clock-control none\`, stream loop so that every kernel of the fold is a launch ncu lists: \`${T}_launches_cfg*.csv.gz\`, tables \`${T}_counters_cfg*.txt\`,"
echo "merged into \`r02_kernel_counters.json\`, which \`bench.py\` reads) and \`ncu --set full --clock-control none --import-source on -k regex:<kernel> -s 1 -c 1\`"
echo "(second launch of the kernel = folding step 2, the largest)."
echo
python - <<PY
import json
for c in ("cfg2", "cfg3", "cfg4"):
    try:
        d = json.load(open("$G/${T}_bench_%s.json" % c))
    except Exception:
        continue
    print("* %s: %.2f ms per step, %s seq/s device-resident; e2e %s seq/s (median %.2f ms, p95 %.2f ms); Python API %s seq/s%s" % (
        c, d["ms_per_step"], format(round(d["value"]), ","), format(round(d["e2e"]["value"]), ","), d["e2e"]["median_ms"], d["e2e"]["p95_ms"],
        format(round(d["e2e_python_api"]["value"]), ","),
        "; CPU port %d threads %.1f seq/s" % (d["cpu_baseline"]["cores"], d["cpu_baseline"]["value"]) if "cpu_baseline" in d else ""))
PY
echo
for c in cfg2 cfg3 cfg4; do
echo "## Per kernel family, one fold of $c (ncu metrics pass; per-launch times are serialised and cold, the SHARE is what must agree with bench.py's event times)"
echo
echo '```'
cat $G/${T}_counters_$c.txt
echo '```'
echo
done
echo "## --set full, folding step 2"
echo
python tools/ncu_md.py "k_expand_fast cfg2"=$G/${T}_cfg2_k_expand_fast.raw.csv "k_expand_fast cfg3"=$G/${T}_cfg3_k_expand_fast.raw.csv "k_expand_big cfg4"=$G/${T}_cfg4_k_expand_big.raw.csv k_combine=$G/${T}_cfg2_k_combine.raw.csv k_materialise=$G/${T}_cfg2_k_materialise.raw.csv
echo
echo "## Where k_expand_fast's instructions go (cfg2; source page mapped to lines with tools/ncu_lines.py, bucketed by tools/phase_buckets.py)"
echo
echo '```'
NCU_LINES_INNER=1 python tools/ncu_lines.py $G/${T}_cfg2_k_expand_fast.source.csv.gz rafft_b200/librafft_b200.so k_expand_fastILi256ELi3 4000 2>/dev/null | python tools/phase_buckets.py > /tmp/${T}_buckets.txt; sed -n 1,22p /tmp/${T}_buckets.txt
echo '```'
} > profiles/${T}_ncu_summary.md
wc -l profiles/${T}_ncu_summary.md
