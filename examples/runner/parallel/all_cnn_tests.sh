#!/bin/bash
# CNN placements vs the single-device run (reference all_cnn_tests.sh).  HETU_CPU=1 for a gloo CPU world.
cd "$(dirname "${BASH_SOURCE[0]}")"
set -e
mkdir -p results && rm -f results/cnn_*.npy
run() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port $((20000 + RANDOM % 20000)) "$@"; }
RTOL=${RTOL:-$([ "$HETU_CPU" = "1" ] && echo 1e-5 || echo 5e-3)}
python test_model_cnn_base.py
run 2 test_model_cnn.py --stages 2
run 4 test_model_cnn.py --stages 4
run 2 test_model_cnn.py --stages 2 --pipeline gpipe
for s in left right middle; do run 5 test_model_cnn_complex.py --split $s; done
python validate_results.py --rtol "$RTOL" --prefix cnn_
