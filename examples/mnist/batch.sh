#!/bin/bash
# Sweep the Stack-B parameter-manager benchmark over servers x client threads x clients (the reference's
# examples/mnist/batch.sh swept {1,2,4,8,16}); prints one JSON line per configuration.
cd "$(dirname "$0")/../.." || exit 1
port=60000
for ns in 1 2 4; do for nt in 1 2 4; do for nc in 1 4; do
  python tools/pm.py --local --nservers $ns --nclients $nc --threads $nt --rounds "${ROUNDS:-20}" --port $port
  port=$((port + 50))
done; done; done
