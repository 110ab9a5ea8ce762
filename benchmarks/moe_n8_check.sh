N=8
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_distributed.py -x -q -m gpu -k symm > gpurun_out/moe_dist$N.log 2>&1; echo "rc=$?" >> gpurun_out/moe_dist$N.log
tail -3 gpurun_out/moe_dist$N.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --workload moe --steps 20 --warmup 5 > gpurun_out/moe_n${N}.json 2> gpurun_out/moe_n${N}.err
echo "rc=$? $(tail -1 gpurun_out/moe_n${N}.json | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["value"], d["ms_per_step"], d["e2e"]["value"])')"
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 benchmarks/moe_profile.py gpurun_out/moe_prof_n${N}.txt > gpurun_out/moe_prof_n${N}.log 2>&1
head -16 gpurun_out/moe_prof_n${N}.txt | cut -c1-120
