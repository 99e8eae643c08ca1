#!/bin/bash
# Round profile capture (run under gpurun, ONE GPU): every ncu pass only after the same command exited 0 without ncu.
#   1. step timeline (torch.profiler) of the graphed step
#   2. ncu launch list of `bench.py --steps 2 --warmup 3 --no-cpu-baseline`
#   3. ncu --set full of the three MLP kernels (tools/time_train_mlp.py 4096 2) and of the Chamfer sweeps
# Everything lands in gpurun_out/; tools/summarise_profiles.py turns it into the files under profiles/.
set -u
OUT=gpurun_out
TAG=${1:-r1}
mkdir -p $OUT
python tools/step_timeline.py > $OUT/${TAG}_timeline.txt 2>&1 || echo "timeline failed"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_bench_plain.json 2> $OUT/${TAG}_bench_plain.err || { echo "bench failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_launches.log 2>&1 || echo "ncu launch list failed"
python tools/time_train_mlp.py 4096 2 > $OUT/${TAG}_time_train_mlp.json 2>&1 || { echo "time_train_mlp failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'tc_(fwd|bwd|wgrad)_kernel' -s 6 -c 3 -f -o $OUT/${TAG}_mlp \
    python tools/time_train_mlp.py 4096 2 > $OUT/${TAG}_ncu_mlp.log 2>&1 || echo "ncu mlp failed"
python tools/time_chamfer.py 28830 --raster > $OUT/${TAG}_time_chamfer.json 2>&1 || echo "time_chamfer failed"
# raster-order clouds: three sweeps per call (screen of every 8th destination, screen and exact sweep over the stage lists)
ncu --set full --clock-control none --import-source on -k regex:'chamfer_tc_kernel' -s 3 -c 3 -f -o $OUT/${TAG}_chamfer \
    python tools/time_chamfer.py 28830 --raster > $OUT/${TAG}_ncu_chamfer.log 2>&1 || echo "ncu chamfer failed"
ls -la $OUT
