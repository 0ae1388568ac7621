#!/bin/bash
# Benchmark matrix for BASELINE.md: every config x {ours, torch library arm} x {fp32 (tf32), bf16} on N GPUs, one JSON
# line per run under $OUT (default gpurun_out/matrix).  Usage: tools/run_matrix.sh <ngpus> [configs...]
N=${1:-1}; shift
CONFIGS=${@:-"mnist_mlp lenet cifar10 alexnet wide_mlp"}
OUT=${OUT:-gpurun_out/matrix}; mkdir -p $OUT
PORT=29700
run() {  # name, args...
  local name=$1; shift
  PORT=$((PORT+1))
  if [ "$N" = "1" ]; then
    timeout 400 python bench.py --gpus 1 "$@" > $OUT/$name.json 2> $OUT/$name.err
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $N "$@" 2> $OUT/$name.err | tail -1 > $OUT/$name.json
  fi
  python - "$OUT/$name.json" "$name" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(f"{sys.argv[2]:44s} {d['ms_per_step']:9.4f} ms/step {d['value']:12.0f} img/s  e2e {d['e2e']['ms_per_step']:9.4f} ms  graph={d['config'].get('cuda_graph')} loss={d['e2e'].get('final_loss_sum_window', d['e2e'].get('final_loss'))}")
except Exception as e:
    print(f"{sys.argv[2]:44s} FAILED ({e})")
PY
}
for c in $CONFIGS; do
  steps=100; warm=10
  case $c in alexnet|wide_mlp) steps=20; warm=5;; esac
  for d in ${DTYPES:-fp32 bf16}; do
    if [ "$N" = "1" ]; then
      run ${c}_n1_ours_$d --config $c --dtype $d --steps $steps --warmup $warm
      run ${c}_n1_torch_$d --config $c --dtype $d --impl torch --steps $steps --warmup $warm
    else
      run ${c}_n${N}_nccl_$d --config $c --dtype $d --steps $steps --warmup $warm --push-pull nccl
      # the fused exchange must train to the same loss as the NCCL exchange of the same kernels
      want=$(python -c "import json;print(','.join(str(v) for v in json.loads(open('$OUT/${c}_n${N}_nccl_$d.json').read().strip().splitlines()[-1])['e2e']['final_loss_sum_window']))" 2>/dev/null)
      run ${c}_n${N}_fused_$d --config $c --dtype $d --steps $steps --warmup $warm --push-pull fused ${want:+--expect-loss-window $want}
      run ${c}_n${N}_torch_$d --config $c --dtype $d --impl torch --steps $steps --warmup $warm
      if [ "$c" = "cifar10" ]; then
        for st in 0 2 8; do
          run ${c}_n${N}_async_s${st}_$d --config $c --dtype $d --steps $steps --warmup $warm --push-pull async --staleness $st --sync-frequency 4
        done
      fi
    fi
  done
done
