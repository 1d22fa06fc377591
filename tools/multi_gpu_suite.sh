#!/bin/bash
# Runs the three sharded workloads on N GPUs of one box (one process per GPU): BASELINE config 4 (bench.py),
# config 3 (SETP1 KITTI, batch 64 split over the GPUs, and 64 frames per GPU) and config 5 (training step,
# 16 frames per GPU).  Usage: tools/multi_gpu_suite.sh N [tag]   -> gpurun_out/*_<tag>_n<N>.json
set -u
N=${1:-1}; TAG=${2:-run}
OUT=gpurun_out; mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
if [ "$N" = "1" ]; then RUN="python"; fi
PER=$((64 / N))
timeout 600 $RUN bench.py --gpus $N --steps 10 --warmup 3 > $OUT/bench_${TAG}_n$N.json 2> $OUT/bench_${TAG}_n$N.err
timeout 300 $RUN tools/setp1_bench.py --frames $PER > $OUT/setp1_kitti_strong_${TAG}_n$N.json 2> $OUT/setp1_kitti_strong_${TAG}_n$N.err
timeout 300 $RUN tools/setp1_bench.py --frames 64 > $OUT/setp1_kitti_weak_${TAG}_n$N.json 2> $OUT/setp1_kitti_weak_${TAG}_n$N.err
timeout 600 $RUN tools/train_bench.py --frames 16 --steps 5 > $OUT/train_${TAG}_n$N.json 2> $OUT/train_${TAG}_n$N.err
python - <<PY
import json, glob
for f in sorted(glob.glob("$OUT/*_${TAG}_n$N.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split("/")[-1], d.get("metric"), round(d.get("value", 0), 1), d.get("unit"), "ms/step", round(d.get("ms_per_step", 0), 3),
              "e2e", (d.get("e2e") or {}).get("value"))
    except Exception as e:
        print(f, "ERR", e)
PY
