run() { n=$1; shift; tag=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 20 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/s39_$tag.json 2> gpurun_out/s39_$tag.err; tail -1 gpurun_out/s39_$tag.json | cut -c1-200; }
run 2 tiles2_rep --mode tiles --config C4 --shadow-share replicate
run 4 tiles4_rep --mode tiles --config C4 --shadow-share replicate
run 8 tiles8_rep --mode tiles --config C4 --shadow-share replicate
run 4 tiles4_bc --mode tiles --config C4
run 2 views2
