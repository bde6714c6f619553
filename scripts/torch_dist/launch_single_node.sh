#!/bin/bash
# Launch one node of a (possibly multi-node) job with torchrun (reference: scripts/torch_dist/launch_single_node.sh).
#   usage: launch_single_node.sh NODE_RANK NNODES MASTER_ADDR MASTER_PORT SCRIPT [script args...]
# A per-node lock file prevents two launches from sharing the GPUs; torchrun --max-restarts=0 (no elastic restart:
# a failed rank ends the job, recovery = resume from the last checkpoint).
set -euo pipefail
NODE_RANK=${1:?node rank}; NNODES=${2:?nnodes}; MADDR=${3:?master addr}; MPORT=${4:?master port}; SCRIPT=${5:?script}; shift 5
cd "$(dirname "$0")/../.."
source scripts/torch_dist/set_env.sh
export MASTER_ADDR=$MADDR MASTER_PORT=$MPORT
NPROC=${NPROC_PER_NODE:-$(nvidia-smi -L | wc -l)}
LOCK=/tmp/scaletorch_b200_node.lock
exec 9>"$LOCK"
if ! flock -n 9; then echo "[node $NODE_RANK] another launch holds $LOCK" >&2; exit 3; fi
LOGDIR=${LOGDIR:-logs}; mkdir -p "$LOGDIR"
cleanup() { trap - TERM INT; [ -n "${PID:-}" ] && kill -TERM "$PID" 2>/dev/null && sleep 5 && kill -KILL "$PID" 2>/dev/null || true; }
trap cleanup TERM INT
python -m torch.distributed.run --nnodes="$NNODES" --node-rank="$NODE_RANK" --nproc-per-node="$NPROC" \
  --master-addr "$MADDR" --master-port "$MPORT" --max-restarts=0 "$SCRIPT" "$@" > "$LOGDIR/node_${NODE_RANK}.log" 2>&1 &
PID=$!
wait "$PID"; RC=$?
echo "[node $NODE_RANK] exit code $RC (log: $LOGDIR/node_${NODE_RANK}.log)"
exit $RC
