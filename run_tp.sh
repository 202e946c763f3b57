#!/bin/bash
# Hang-safe multi-GPU launch (used under gpurun --gpus N):  ./run_tp.sh NGPU SECONDS LOGNAME script.py [args...]
# The torchrun tree runs in its own session; after SECONDS the whole process group is killed, then any process
# of that session still holding a GPU context is killed by its exact PID (other tenants' processes are never touched).
mkdir -p gpurun_out
log=gpurun_out/$3.log
setsid python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "${@:4}" > "$log" 2>&1 &
pid=$!
for i in $(seq 1 "$2"); do kill -0 $pid 2>/dev/null || break; sleep 1; done
if kill -0 $pid 2>/dev/null; then
  echo "TIMEOUT after $2 s: killing process group $pid" >> "$log"
  kill -KILL -- -$pid 2>/dev/null
  sleep 2
  mine=" $(pgrep -s "$pid" | tr '\n' ' ') "   # processes of the torchrun session only
  for p in $(nvidia-smi --query-compute-apps=pid --format=csv,noheader 2>/dev/null); do
    case "$mine" in *" $p "*) kill -KILL "$p" 2>/dev/null ;; esac
  done
fi
wait $pid 2>/dev/null
echo "rc=$?" >> "$log"
tail -n "${TAIL:-40}" "$log"
