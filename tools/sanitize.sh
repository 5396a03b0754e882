#!/bin/bash
# compute-sanitizer memcheck / racecheck / synccheck over smoke() (pcg_persistent, pcg_dual, pcg_cluster single + pair on
# 642 rows) and tools/sanitize_cases.py (bicg_persistent; the partitioned kernels as threaded virtual ranks, world 2 and 4).
# usage (on the GPU box): tools/sanitize.sh [outdir]     logs: <outdir>/<tool>_<case>.log, one summary line each on stdout
O=${1:-gpurun_out/sanitizer}
mkdir -p $O
run() {
    tool=$1; name=$2; shift 2
    timeout ${SAN_TIMEOUT:-300} compute-sanitizer --tool $tool --print-limit 20 --log-file $O/${tool}_${name}.log "$@" > $O/${tool}_${name}.out 2>&1
    echo "== $tool $name: exit $? :: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $O/${tool}_${name}.log | tail -1) :: $(tail -1 $O/${tool}_${name}.out)"
}
for tool in memcheck racecheck synccheck; do
    run $tool smoke python -c "import __graft_entry__ as g; g.smoke()"
    run $tool bicg python tools/sanitize_cases.py single
    run $tool ranks python tools/sanitize_cases.py ranks
done
