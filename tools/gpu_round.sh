#!/bin/bash
# One gpurun call: the ncu launch list first (its per-launch instruction counts feed bench.py's
# issue-slot roofline, keyed by the kernel source hash), then bench (both arms), then --set full
# captures of the two hot kernels on a short version of the same command (B200_PROFILING.md recipe).
set -u
TAG=${1:-r1}
OUT=gpurun_out
mkdir -p $OUT
SHORT="python bench.py --textures 8 --steps 1 --warmup 3 --no-cpu --no-e2e"
$SHORT > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
    --log-file $OUT/${TAG}_launches.csv $SHORT > $OUT/${TAG}_ncu1.log 2>&1
echo "launch list rc=$?"
python tools/profile_counts.py $OUT/${TAG}_launches.csv albedo-2048-bc7-normal 8 > $OUT/${TAG}_counts.log 2>&1
cp profiles/bench_counts.json $OUT/${TAG}_bench_counts.json
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; tail -c 3000 $OUT/${TAG}_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_ref.json 2>> $OUT/${TAG}_bench.err; echo "ref rc=$?"; tail -c 600 $OUT/${TAG}_bench_ref.json
# the HBM-bound encoders (north_star kernel 3): same workload shape, normal maps to BC5 and metal-rough maps to BC4
python bench.py --kind normal --format bc5 --no-cpu > $OUT/${TAG}_bench_bc5.json 2>> $OUT/${TAG}_bench.err; echo "bc5 rc=$?"
python bench.py --kind metal_rough --format bc4 --no-cpu > $OUT/${TAG}_bench_bc4.json 2>> $OUT/${TAG}_bench.err; echo "bc4 rc=$?"
# BC5: the fused mip + encode kernel (levels 0..4 are encoded inside mip_fused_kernel<NORMAL, 1>)
ncu --set full --clock-control none --import-source on -k regex:mip_fused -s 3 -c 1 -f -o $OUT/${TAG}_bc5 python bench.py --kind normal --format bc5 --textures 8 --steps 1 --warmup 3 --no-cpu --no-e2e > $OUT/${TAG}_ncu4.log 2>&1
echo "bc5 full rc=$?"
ncu --set full --clock-control none --import-source on -k regex:bc7_encode -s 3 -c 1 -f -o $OUT/${TAG}_bc7 $SHORT > $OUT/${TAG}_ncu2.log 2>&1
echo "bc7 full rc=$?"
ncu --set full --clock-control none --import-source on -k regex:mip_fused -s 3 -c 1 -f -o $OUT/${TAG}_mip $SHORT > $OUT/${TAG}_ncu3.log 2>&1
echo "mip full rc=$?"
ls -la $OUT
