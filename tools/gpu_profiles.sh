#!/bin/sh
# GPU-box script: launch list of the bench command and ncu --set full captures of the dominant kernels, reduced to the
# text summaries that are committed under profiles/ (the .ncu-rep files stay on the box: gpurun_out/ is capped at 64 MiB)
#   gpurun --timeout 2400 -- 'sh tools/gpu_profiles.sh r02e'
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT /tmp/ncu
export SEQM_BENCH_SUB=0
python bench.py --steps 2 --warmup 3 > /dev/null 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches_bench_steps2.csv \
    python bench.py --steps 2 --warmup 3 > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none -k regex:scf_small_kernel -c 1 -s 3 -o /tmp/ncu/scf_small \
    python bench.py --steps 2 --warmup 3 > /dev/null 2>&1
python tools/ncu_summary.py /tmp/ncu/scf_small.ncu-rep > $OUT/${TAG}_scf_small_ncu.txt 2>&1
python tools/exp_c60_ncu.py > /dev/null 2>&1 || exit 2
timeout 900 ncu --set full --clock-control none -k regex:'sp2_step_kernel|xl_extrapolate|pair_grad_xx|pair_fwd_kernel|fock_pair_kernel' -s 12 -c 12 \
    -o /tmp/ncu/c60 python tools/exp_c60_ncu.py > /dev/null 2>&1
python tools/ncu_summary.py /tmp/ncu/c60.ncu-rep > $OUT/${TAG}_c60_ncu.txt 2>&1
timeout 900 ncu --set full --clock-control none -k regex:eigh_cluster_kernel -c 1 -o /tmp/ncu/cluster env NMOL=32 python tools/exp_c60_ncu.py > /dev/null 2>&1
python tools/ncu_summary.py /tmp/ncu/cluster.ncu-rep > $OUT/${TAG}_eigh_cluster_ncu.txt 2>&1
ls -la $OUT | tail -8
