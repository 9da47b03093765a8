#!/bin/bash
# Multi-GPU evidence on ONE box with N GPUs: headline bench, per-phase timeline, the three topology configurations and the
# multi-rank tests. Everything lands in gpurun_out/. usage: scripts/run_multi_gpu_suite.sh N
N=${1:-8}
OUT=gpurun_out
mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $RUN --master-port 29601 bench.py --gpus $N --steps 40 --warmup 5 --also median > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err
tail -c 600 $OUT/bench_n$N.json
timeout 120 $RUN --master-port 29602 scripts/phase_timeline.py --gar krum > $OUT/timeline_n$N.json 2> $OUT/timeline_n$N.err
cat $OUT/timeline_n$N.json
for cfg in cifarnet-ssmw resnet50-msmw vgg16-learn; do
  timeout 240 $RUN --master-port 29603 applications/benchmarks/topology_bench.py --config $cfg --steps 20 2> $OUT/topo_${cfg}_n$N.err | tail -1 | tee $OUT/topo_${cfg}_n$N.json
done
timeout 400 python -m pytest tests/test_multi_gpu.py -x -q 2>&1 | tail -3 | tee $OUT/pytest_mgpu_n$N.log
