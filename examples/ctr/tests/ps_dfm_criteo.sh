#!/bin/bash
# reference: examples/ctr/tests/ps_dfm_criteo.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../run_hetu.py
heturun=${workdir}/../../../bin/heturun
# bash ps_dfm_criteo.sh [workers] : 1 server, N workers, embeddings and dense parameters on the PS
${heturun} -s 1 -w ${1:-4} python ${mainpy} --model dfm_criteo --comm PS --bsp 0 --cache lfuopt --bound 3 --all
