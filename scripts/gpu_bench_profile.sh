#!/bin/bash
# Run on the GPU box (under gpurun): bench both arms, then the ncu launch list and
# full captures of the dominant kernels. Outputs land in gpurun_out/.
set -u
mkdir -p gpurun_out
TAG=${1:-r1}
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_${TAG}.json 2> gpurun_out/bench_${TAG}.err
echo "bench rc=$?"; tail -c 300 gpurun_out/bench_${TAG}.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_${TAG}.json 2> gpurun_out/bench_ref_${TAG}.err
echo "ref rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
$CMD > gpurun_out/plain_${TAG}.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv \
    --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/ncu_launches_${TAG}.log 2>&1
echo "ncu launches rc=$?"
$CMD > gpurun_out/plain2_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:count_small -s 3 -c 2 \
    -o gpurun_out/prof_count_${TAG} -f $CMD > gpurun_out/ncu_full_${TAG}.log 2>&1
echo "ncu full rc=$?"
GCMD="python scripts/gram_probe.py"
$GCMD > gpurun_out/plain3_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gram_dmma -c 1 \
    -o gpurun_out/prof_gram_${TAG} -f $GCMD > gpurun_out/ncu_gram_${TAG}.log 2>&1
echo "ncu gram rc=$?"
ls -la gpurun_out | tail -12
