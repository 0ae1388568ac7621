#!/bin/bash
# Process control for a single-box job (the reference's script/node.sh fanned ps/ls/scp/kill out over ssh).
#   node.sh ps | kill | gpus
case "$1" in
  ps)   pgrep -af "singa_b200.cli|tools/pm.py" ;;
  kill) for p in $(pgrep -f "singa_b200.cli"); do kill "$p"; done ;;
  gpus) python -m singa_b200.launch --discover ;;
  *)    echo "usage: $0 ps|kill|gpus" ;;
esac
