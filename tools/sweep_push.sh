#!/bin/bash
# Sweep the fused push knobs at N GPUs: tools/sweep_push.sh <ngpus> <dtype> "<bucket sizes>" "<cta budgets>"
N=$1; D=$2; BUCKETS=${3:-"6000000"}; CTAS=${4:-"296"}; PORT=29600
for b in $BUCKETS; do for c in $CTAS; do
  PORT=$((PORT+1))
  SINGA_B200_PUSH_BUCKET=$b SINGA_B200_PUSH_CTAS=$c timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
    --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 200 --warmup 20 --dtype $D ${EXTRA:-} 2>/dev/null \
    | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=$N $D bucket=$b ctas=$c ${EXTRA:-}', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['e2e']['final_loss_sum_window'])"
done; done
