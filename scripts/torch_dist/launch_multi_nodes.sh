#!/bin/bash
# ssh fan-out over the hosts of node_list.txt (reference: scripts/torch_dist/launch_multi_nodes.sh:1-259): the first
# host is the rendezvous master, every node runs launch_single_node.sh, the launcher waits for all of them, reports
# per-node exit codes, and on SIGINT/SIGTERM (or the first failing node) tears the others down.
#   usage: launch_multi_nodes.sh NODE_LIST SCRIPT [script args...]        (NODE_LIST: one host/IP per line, '#' comments)
set -uo pipefail
LIST=${1:?node list file}; SCRIPT=${2:?script}; shift 2
REPO=$(cd "$(dirname "$0")/../.." && pwd)
mapfile -t NODES < <(grep -v '^\s*#' "$LIST" | awk 'NF {print $1}')
NN=${#NODES[@]}
[ "$NN" -gt 0 ] || { echo "no nodes in $LIST" >&2; exit 2; }
MADDR=${NODES[0]}; MPORT=${MASTER_PORT:-29500}
SSH="ssh -o StrictHostKeyChecking=no -o ConnectTimeout=10"
declare -a PIDS
teardown() {
  trap - TERM INT
  echo "[launcher] tearing down"
  for i in "${!NODES[@]}"; do
    $SSH "${NODES[$i]}" "pid=\$(cat /tmp/scaletorch_b200_launch.pid 2>/dev/null) && kill -TERM -- -\$pid 2>/dev/null; sleep 5; kill -KILL -- -\$pid 2>/dev/null" &
  done
  wait
}
trap 'teardown; exit 130' TERM INT
for i in "${!NODES[@]}"; do
  $SSH "${NODES[$i]}" "cd $REPO && setsid bash -c 'echo \$\$ > /tmp/scaletorch_b200_launch.pid; exec bash scripts/torch_dist/launch_single_node.sh $i $NN $MADDR $MPORT $SCRIPT $*'" &
  PIDS[$i]=$!
  echo "[launcher] node $i -> ${NODES[$i]} (ssh pid ${PIDS[$i]})"
done
FAIL=0
for i in "${!NODES[@]}"; do
  wait "${PIDS[$i]}"; RC=$?
  echo "[launcher] node $i (${NODES[$i]}) exit code $RC"
  if [ $RC -ne 0 ] && [ $FAIL -eq 0 ]; then FAIL=$RC; teardown; fi
done
exit $FAIL
