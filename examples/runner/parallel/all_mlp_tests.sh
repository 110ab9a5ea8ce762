#!/bin/bash
# Equivalence of every placement with the single-device run (reference all_mlp_tests.sh).
#   bash all_mlp_tests.sh            # GPUs (needs 4), NCCL
#   HETU_CPU=1 bash all_mlp_tests.sh # same scripts on a gloo CPU world
cd "$(dirname "${BASH_SOURCE[0]}")"
set -e
rm -rf results && mkdir -p results
run() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port $((20000 + RANDOM % 20000)) "$@"; }
RTOL=${RTOL:-$([ "$HETU_CPU" = "1" ] && echo 1e-5 || echo 2e-3)}
python test_mlp_base.py
run 2 test_mlp_pp.py --stages 2
run 4 test_mlp_pp.py --stages 4
run 2 test_mlp_pp.py --stages 2 --pipeline gpipe
run 2 test_mlp_pp.py --stages 2 --pipeline pipedream
for split in left right middle; do
  run 2 test_mlp_mp_fixed.py --split $split
  run 4 test_mlp_mp_pp.py --split $split
done
python validate_results.py --rtol "$RTOL"
