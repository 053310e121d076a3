#!/bin/bash
# usage: profiles/scale_run.sh N   (on the GPU box: gpurun --gpus N -- 'bash profiles/scale_run.sh N')
# per-N bench lines of config3 (north star), config4 (4K + environment light) and config5 (training) -> gpurun_out/scale_N<N>.jsonl
N=$1
out=gpurun_out/scale_N${N}.jsonl
: > $out
run() {
  if [ "$N" = 1 ]; then python bench.py --gpus 1 "$@"
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; fi
}
for w in config3 config4 config5; do
  run --workload $w --steps 20 --warmup 5 --no-cpu-baseline 2> gpurun_out/scale_N${N}_${w}.err | tail -1 >> $out
done
cat $out
