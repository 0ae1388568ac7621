set -x
mkdir -p gpurun_out/c2
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_conv_gpu.py -x -q -m gpu 2>&1 | tail -15
timeout 600 python -m pytest tests/test_model_gpu.py -x -q -m gpu 2>&1 | tail -8
for c in alexnet cifar10 lenet; do for d in fp32 bf16; do
  steps=100; warm=10; case $c in alexnet) steps=20; warm=5;; esac
  timeout 300 python bench.py --gpus 1 --config $c --dtype $d --steps $steps --warmup $warm > gpurun_out/c2/${c}_n1_ours_$d.json 2> gpurun_out/c2/${c}_n1_ours_$d.err
  python -c "
import json,sys
d=json.loads(open('gpurun_out/c2/${c}_n1_ours_$d.json').read().strip().splitlines()[-1]); print('$c $d', d['ms_per_step'], d['e2e'])"
done; done
SINGA_B200_TRACE_OUT=gpurun_out/c2/tr_alex_fp32 timeout 300 python bench.py --gpus 1 --config alexnet --dtype fp32 --steps 5 --warmup 3 > /dev/null 2> gpurun_out/c2/tr.err
SINGA_B200_TRACE_OUT=gpurun_out/c2/tr_alex_bf16 timeout 300 python bench.py --gpus 1 --config alexnet --dtype bf16 --steps 5 --warmup 3 > /dev/null 2>> gpurun_out/c2/tr.err
ls gpurun_out/c2
