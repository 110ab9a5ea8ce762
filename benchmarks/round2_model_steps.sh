mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "3xtf32 or attention" 2>&1 | tail -8
for m in bert_base bert_large; do
  timeout 300 python benchmarks/model_step.py $m 14 2>&1 | tail -3 > gpurun_out/step_${m}_flash.txt; tail -1 gpurun_out/step_${m}_flash.txt
  HETU_FLASH_ATTN=0 timeout 300 python benchmarks/model_step.py $m 14 2>&1 | tail -3 > gpurun_out/step_${m}_noflash.txt; tail -1 gpurun_out/step_${m}_noflash.txt
done
HETU_MATMUL_PRECISION=bf16 timeout 300 python benchmarks/model_step.py bert_large 14 2>&1 | tail -2
timeout 300 python benchmarks/model_step.py resnet18 14 2>&1 | tail -1
