# A/B of the MoE step: overlapped exchange (ops/moe_ffn.py) with 2 / 1 capacity parts per expert vs the op chain.
# usage: bash benchmarks/moe_overlap_ab.sh <ngpu> [skip-tests]
N=${1:-2}
mkdir -p gpurun_out
if [ -z "$2" ]; then
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_zw_moe_gate_gpu.py -x -q -m gpu > gpurun_out/moe_kernels.log 2>&1; echo "rc=$?" >> gpurun_out/moe_kernels.log
tail -4 gpurun_out/moe_kernels.log
fi
timeout 900 python -m pytest tests/test_distributed.py -x -q -m gpu -k symm > gpurun_out/moe_dist$N.log 2>&1; echo "rc=$?" >> gpurun_out/moe_dist$N.log
tail -4 gpurun_out/moe_dist$N.log
run() {  # tag, env...
  tag=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --workload moe --steps 20 --warmup 5 > gpurun_out/moe_n${N}_$tag.json 2> gpurun_out/moe_n${N}_$tag.err
  echo "$tag rc=$? $(tail -1 gpurun_out/moe_n${N}_$tag.json | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["e2e"]["value"])')"
}
run chunks2 HETU_MOE_CHUNKS=2
run chunks1 HETU_MOE_CHUNKS=1
run chunks4 HETU_MOE_CHUNKS=4
run chain HETU_MOE_OVERLAP=0
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 benchmarks/moe_profile.py gpurun_out/moe_prof_n${N}.txt > gpurun_out/moe_prof_n${N}.log 2>&1
if [ -z "$2" ]; then
timeout 600 python bench.py --gpus 1 --workload moe --steps 20 --warmup 5 > gpurun_out/moe_n1.json 2> gpurun_out/moe_n1.err; echo "n1 rc=$? $(tail -1 gpurun_out/moe_n1.json | cut -c1-330)"
timeout 250 python benchmarks/moe_profile.py gpurun_out/moe_prof_n1.txt > gpurun_out/moe_prof_n1.log 2>&1
fi
