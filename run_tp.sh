#!/bin/bash
# Hang-safe multi-GPU launch (used under gpurun --gpus N):  ./run_tp.sh NGPU SECONDS LOGNAME script.py [args...]
# The torchrun tree runs in its own session; after SECONDS the whole process group is killed, then any process
# still holding a GPU context is killed by its exact PID, so a stuck collective cannot hold the box.
mkdir -p gpurun_out
log=gpurun_out/$3.log
setsid python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "${@:4}" > "$log" 2>&1 &
pid=$!
for i in $(seq 1 "$2"); do kill -0 $pid 2>/dev/null || break; sleep 1; done
if kill -0 $pid 2>/dev/null; then
  echo "TIMEOUT after $2 s: killing process group $pid" >> "$log"
  kill -KILL -- -$pid 2>/dev/null
  sleep 2
  for p in $(nvidia-smi --query-compute-apps=pid --format=csv,noheader 2>/dev/null); do kill -KILL "$p" 2>/dev/null; done
fi
wait $pid 2>/dev/null
echo "rc=$?" >> "$log"
tail -n "${TAIL:-40}" "$log"
