#!/bin/bash
# reference: examples/ctr/tests/hybrid_dfm_criteo.sh
workdir=$(cd "$(dirname "$0")"; pwd)
mainpy=${workdir}/../run_hetu.py
heturun=${workdir}/../../../bin/heturun
# dense parameters all-reduced, embedding table on the PS behind the HET cache
${heturun} -s 1 -w ${1:-4} python ${mainpy} --model dfm_criteo --comm Hybrid --bsp 0 --cache lfuopt --bound 3 --all
