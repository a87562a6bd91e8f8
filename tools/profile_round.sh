#!/bin/bash
# usage: tools/profile_round.sh <tag>   e.g. r1_v10
# One gpurun call: (1) the bench without a profiler (its JSON line is the measurement), (2) the ncu launch list of the
# same command, (3) one `ncu --set full` capture of the seven kernels of one warmed-up C2 frame.  Then summarises
# everything into profiles/<tag>_* with tools/profile_collect.py.
tag=$1
/usr/local/graft/bin/gpurun --timeout 1500 -- "timeout 400 python bench.py --steps 50 --warmup 5 > gpurun_out/${tag}_bench.json 2>gpurun_out/${tag}_bench.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu1.log 2>&1; \
ncu --set full --import-source on --clock-control none -k regex:'k_vertex_both|k_setup_raster|k_tile' --launch-skip 14 --launch-count 7 -o gpurun_out/${tag}_full python tools/prof_frame.py C2 4 > gpurun_out/${tag}_ncu2.log 2>&1; tail -2 gpurun_out/${tag}_ncu2.log; cat gpurun_out/${tag}_bench.json" 2>&1 | tail -8
python tools/profile_collect.py $tag
