#!/bin/bash
# profiles/ files of one tools/round_check.sh run:  tools/profile_summary.sh <tag> <profile-version> "<title>"
T=${1:?tag}; V=${2:?version}; TITLE=${3:-}
G=gpurun_out
for c in cfg2 cfg3 cfg4; do [ -s $G/${T}_bench_$c.json ] && cp $G/${T}_bench_$c.json profiles/r01_bench_${V}_$c.json; done
[ -s $G/${T}_bench_cfg2_ref.json ] && cp $G/${T}_bench_cfg2_ref.json profiles/r01_bench_${V}_cfg2_reference_arm.json
cp $G/${T}_launches.csv profiles/r01_launches_${V}.csv
python tools/traffic_table.py $G/${T}_dram.csv 1 > profiles/r01_dram_${V}.txt
{
echo "# ncu summary ${V}: ${TITLE}"
echo
echo "Produced by \`tools/round_check.sh ${T}\` on one GPU box: \`pytest -m gpu\`, \`smoke()\`, the bench lines \`r01_bench_${V}_cfg{2,3,4}.json\` +"
echo "\`r01_bench_${V}_cfg2_reference_arm.json\`, then \`tools/evidence.sh\` (every ncu pass after the same command had exited 0 without ncu):"
echo "launch list \`r01_launches_${V}.csv\` (\`--metrics gpu__time_duration.sum --clock-control none\`), per-kernel DRAM traffic \`r01_dram_${V}.txt\`,"
echo "\`ncu --set full --clock-control none --import-source on -k regex:<kernel> -s 1 -c 1\` (second launch = folding iteration 2, the largest)."
echo
python - <<PY
import json
for c in ("cfg2", "cfg3", "cfg4"):
    try:
        d = json.load(open("$G/${T}_bench_%s.json" % c))
    except Exception:
        continue
    print("* %s: %.2f ms per step, %s seq/s device-resident, e2e %s seq/s (steps %s ms)%s" % (
        c, d["ms_per_step"], format(round(d["value"]), ","), format(round(d["e2e"]["value"]), ","), d["e2e"].get("ms_steps_rank0"),
        ", CPU port %d threads %.0f seq/s" % (d["cpu_baseline"]["cores"], d["cpu_baseline"]["value"]) if "cpu_baseline" in d else ""))
PY
echo
echo "## Launch list: share of the step (second fold of the run, per-launch times in us)"
echo
echo '```'
python tools/launch_table.py $G/${T}_launches.csv 1
echo '```'
echo
echo "## --set full, folding iteration 2"
echo
python tools/ncu_md.py k_expand_fast=$G/${T}_k_expand_fast.raw.csv k_combine=$G/${T}_k_combine.raw.csv k_materialise=$G/${T}_k_materialise.raw.csv
} > profiles/r01_ncu_${V}_summary.md
