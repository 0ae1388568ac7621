#!/bin/bash
# Start / stop a job on this box (replaces the ssh fan-out of the reference's examples/mnist/run.sh).
#   ./run.sh start [cluster.conf] [model.conf]     ./run.sh stop
cd "$(dirname "$0")/../.." || exit 1
case "$1" in
  start)
    python -m singa_b200.launch --cluster_conf "${2:-examples/mnist/cluster.conf}" \
        --model_conf "${3:-examples/mnist/mlp.conf}" --workdir /tmp &
    echo $! > /tmp/singa_b200_launch.pid ;;
  stop)
    [ -f /tmp/singa_b200_launch.pid ] && kill -- -"$(ps -o pgid= "$(cat /tmp/singa_b200_launch.pid)" | tr -d ' ')" 2>/dev/null
    rm -f /tmp/singa_b200_launch.pid ;;
  *) echo "usage: $0 start|stop [cluster.conf] [model.conf]" ;;
esac
