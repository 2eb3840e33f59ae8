#!/bin/bash
# Start W ranks of one tutorial part on this machine, one process per GPU (or per CPU "node").
#   bash examples/launch_local.sh part3 4 [extra flags passed to every rank]
# Equivalent of logging into four nodes and starting master/partX and slave/partX by hand.
PART=${1:-part3}
W=${2:-2}
shift 2
DIR=$(cd "$(dirname "$0")" && pwd)
PORT=${PORT:-29544}
pids=()
for ((r = 1; r < W; r++)); do
  python "$DIR/$PART.py" --master-ip 127.0.0.1 --num-nodes "$W" --rank "$r" --port "$PORT" "$@" > "/tmp/b200dp_${PART}_rank$r.log" 2>&1 &
  pids+=($!)
done
python "$DIR/$PART.py" --master-ip 127.0.0.1 --num-nodes "$W" --rank 0 --port "$PORT" "$@"
rc=$?
for p in "${pids[@]}"; do wait "$p" || rc=$?; done
exit $rc
