set -x
mkdir -p gpurun_out/m1 gpurun_out/ncu
for c in mnist_mlp lenet cifar10 alexnet wide_mlp; do
  steps=100; warm=10
  case $c in alexnet|wide_mlp) steps=20; warm=5;; esac
  for d in fp32 bf16; do
    timeout 300 python bench.py --gpus 1 --config $c --dtype $d --impl torch --steps $steps --warmup $warm > gpurun_out/m1/${c}_n1_torch_$d.json 2> gpurun_out/m1/${c}_n1_torch_$d.err
    tail -c 400 gpurun_out/m1/${c}_n1_torch_$d.json
  done
done
timeout 120 python tools/conv_case.py fp32 fwd 256 96 27 27 256 5 1 2 > gpurun_out/ncu/conv2_fwd_plain.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_tcgen05 -s 3 -c 1 -o gpurun_out/ncu/conv2_fwd_fp32 -f python tools/conv_case.py fp32 fwd 256 96 27 27 256 5 1 2 > gpurun_out/ncu/conv2_fwd_ncu.log 2>&1
cat gpurun_out/ncu/conv2_fwd_plain.log
timeout 120 python tools/conv_case.py fp32 wgrad 256 96 27 27 256 5 1 2 > gpurun_out/ncu/conv2_wgrad_plain.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_tcgen05 -s 4 -c 1 -o gpurun_out/ncu/conv2_wgrad_fp32 -f python tools/conv_case.py fp32 wgrad 256 96 27 27 256 5 1 2 > gpurun_out/ncu/conv2_wgrad_ncu.log 2>&1
cat gpurun_out/ncu/conv2_wgrad_plain.log
