#!/bin/bash
# BASELINE config 3, "plus 2/4/8-GPU prefill:decode splits": every TP ratio in {1:1, 1:2, 2:1, 1:4, 4:1} (and 2:2, 4:4)
# of the Llama-3-70B-shaped cache on one 8-GPU box. Jobs that fit side by side run on disjoint GPU sets.
#   gpurun --gpus 8 -- bash tools/reshard_sweep.sh
set -u
out=gpurun_out/reshard_sweep
mkdir -p $out
run() {  # run <gpus csv> <port> <P> <D>
  local n=$(( $3 + $4 ))
  CUDA_VISIBLE_DEVICES=$1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n \
    --master-addr 127.0.0.1 --master-port $2 bench_reshard.py --prefill-tp $3 --decode-tp $4 --steps 10 --warmup 3 \
    > $out/tp$3_tp$4.log 2>&1
  grep '^{' $out/tp$3_tp$4.log | tail -1 > $out/tp$3_tp$4.json
  echo "tp$3->tp$4 rc=$? $(wc -c < $out/tp$3_tp$4.json) bytes"
}
run 0,1 29601 1 1 & run 2,3,4 29602 1 2 & run 5,6,7 29603 2 1 & wait
run 0,1,2,3 29604 2 2 & wait
run 0,1,2,3,4 29605 1 4; run 0,1,2,3,4 29606 4 1
run 0,1,2,3,4,5,6,7 29607 4 4
