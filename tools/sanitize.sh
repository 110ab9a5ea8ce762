#!/bin/bash
# Memory / race checking of the CUDA kernels with compute-sanitizer (run on a GPU box; slow: 10-50x).
#   bash tools/sanitize.sh [memcheck|racecheck|initcheck|synccheck] [pytest -k expression]
# Host libraries are covered by tests/test_sanitizers.py (ThreadSanitizer / AddressSanitizer builds).
tool=${1:-memcheck}
expr=${2:-"unary_binary or reduce_topk or softmax_layernorm or embedding or moe_dispatch"}
cd "$(dirname "${BASH_SOURCE[0]}")/.."
exec compute-sanitizer --tool "$tool" --error-exitcode 99 --target-processes all \
     python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k "$expr"
