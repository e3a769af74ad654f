#!/bin/bash
# One profiling pass on the GPU box (run through gpurun): launch list of the default bench, one `ncu --set full` capture per
# workload, summaries exported to CSV/text ON the box (gpurun_out/ is capped at 64 MiB; a full report of every kernel is larger).
#   bash tools/gpu_profile_pass.sh <tag> [configs1] [configs3] [configs2]
# Writes gpurun_out/<tag>_launches.csv, <tag>_ncu_full_summary.csv, <tag>_traffic.json, <tag>_<kernel>_source_lines.txt and keeps
# the report of the last workload named (gpurun_out/<tag>_<workload>.ncu-rep) when it is small enough.
set -u
tag=$1; shift
which=${*:-configs1 configs3 configs2}
mkdir -p gpurun_out
O=gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline > $O/${tag}_ncu_launch.log 2>&1
echo "launch list rc=$?"
specs=""
for w in $which; do
    rep=/tmp/${tag}_${w}.ncu-rep
    timeout 900 ncu --set full --clock-control none --import-source on -o ${rep%.ncu-rep} -f python tools/profile_driver.py $w > $O/${tag}_ncu_${w}.log 2>&1
    echo "ncu full $w rc=$?"
    case $w in configs1) key=configs1@64 ;; configs3) key=configs3@64 ;; configs2) key=configs2@64 ;; esac
    specs="$specs $rep:$key"
    case $w in
        configs1) ks="peaks_stream_kernel index_bands_kernel limbs_stream3_kernel" ;;
        configs3) ks="peaks_kernel limbs_kernel" ;;
        configs2) ks="fuse_multiscale fuse_linear gauss_smooth py_connections_kernel py_humans_kernel" ;;
    esac
    for k in $ks; do
        python tools/ncu_lines.py $rep $k 60 > $O/${tag}_${k}_source_lines.txt 2>&1
    done
    sz=$(stat -c %s $rep 2>/dev/null || echo 0)
    if [ "$sz" -lt 30000000 ]; then cp $rep $O/; fi
done
python tools/ncu_summary.py $O/${tag}_ncu_full_summary.csv $O/${tag}_traffic.json $specs > $O/${tag}_summary.log 2>&1
echo "summary rc=$?"
ls -la $O
