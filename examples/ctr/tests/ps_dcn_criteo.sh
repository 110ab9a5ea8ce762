#!/bin/bash
# reference: examples/ctr/tests/ps_dcn_criteo.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../run_hetu.py
heturun=${workdir}/../../../bin/heturun
# bash ps_dcn_criteo.sh [workers] : 1 server, N workers, embeddings and dense parameters on the PS
${heturun} -s 1 -w ${1:-4} python ${mainpy} --model dcn_criteo --comm PS --bsp 0 --cache lfuopt --bound 3 --all
