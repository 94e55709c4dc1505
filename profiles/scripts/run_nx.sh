#!/bin/bash
N=$1; shift
mkdir -p gpurun_out
run() { tag=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" > gpurun_out/r02_n${N}_$tag.json 2> gpurun_out/r02_n${N}_$tag.err; echo $tag rc=$?; python -c "
import json; d=json.load(open('gpurun_out/r02_n${N}_$tag.json')); print(round(d['value'],1), round(d['ms_per_step'],3), d['config']['parallelism'][:160], d['config']['stage_ms'], round(d['e2e']['value'],1))" || tail -5 gpurun_out/r02_n${N}_$tag.err | cut -c1-300; }
for spec in "$@"; do set -- $spec; run "$@"; done
