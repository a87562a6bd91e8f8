#!/bin/bash
# Sort-first (tiles) strong-scaling runs of C4 on one 8-GPU box: bash tools/mg_scaling.sh <tag>
tag=${1:-mg}
run() { n=$1; shift; name=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 20 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/${tag}_$name.json 2> gpurun_out/${tag}_$name.err; tail -1 gpurun_out/${tag}_$name.json | cut -c1-170; }
run 2 tiles2 --mode tiles --config C4
run 4 tiles4 --mode tiles --config C4
run 8 tiles8 --mode tiles --config C4
REB_SCENE_TRI_ORDER=hierarchical run 8 tiles8_coherent --mode tiles --config C4
