#!/usr/bin/env bash
# Build the UNMODIFIED reference (a copy of /root/reference under baseline/_ref) for B200.
#
# What is built, following the reference's own CMake files but driving nvcc/g++ directly (the
# reference's top-level CMake insists on MPI, ZeroMQ, protobuf and oneDNN, none of which is in the image):
#   build/lib/libc_runtime_api.so   src/{common,cuda_common,ops}   (src/CMakeLists.txt, GPU branch)
#   build/lib/libmpi_nccl.so? -> libc_communication  (see below)   src/communication/mpi_nccl_communication.cu
# MPI: the reference's communicator needs MPI_Init/Comm_rank/Comm_size/Allgather/Bcast/Send/Recv only;
# baseline/mpi_stub provides those over TCP sockets (rank/size from the launcher's environment), so the
# reference source is compiled as is.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="$HERE/_ref"
SRC_ROOT="${REFERENCE_SRC:-/root/reference}"
JOBS="${JOBS:-8}"
ARCH="${REF_ARCH:--gencode arch=compute_100,code=sm_100}"

if [ ! -d "$REF/src" ]; then
    rm -rf "$REF"; cp -r "$SRC_ROOT" "$REF"; chmod -R u+w "$REF"
fi
OBJ="$REF/build/obj"; LIB="$REF/build/lib"
mkdir -p "$OBJ" "$LIB" "$OBJ/comm"

INC="-I$REF/src/header -I$REF/src/common -I$REF/src/memory_pool -I$REF/src/cuda_common -I/usr/include/x86_64-linux-gnu"
DEFS="-DDEVICE_GPU -DCUDNN8"
NVCC="nvcc -O3 -std=c++14 $ARCH -Xcompiler -fPIC,-fopenmp -w $DEFS $INC"

# --- libc_runtime_api.so -------------------------------------------------------------------------
srcs=()
for f in "$REF"/src/common/*.cc "$REF"/src/cuda_common/*.cc "$REF"/src/ops/*.cu; do
    case "$(basename "$f")" in
        CuSparseCsrmv.cu) continue;;      # removed by the reference's CMake on CUDA >= 11
    esac
    [ -n "${REF_SKIP:-}" ] && [[ " $REF_SKIP " == *" $(basename "$f") "* ]] && continue
    srcs+=("$f")
done
printf '%s\n' "${srcs[@]}" | xargs -P "$JOBS" -I{} bash -c '
    f="$1"; o="'"$OBJ"'/$(basename "$f").o"
    if [ ! -f "$o" ] || [ "$f" -nt "$o" ]; then
        '"$NVCC"' -x cu -c "$f" -o "$o" 2> "$o.log" || { echo "FAILED $f"; cat "$o.log"; rm -f "$o"; exit 0; }
    fi' _ {}
# CPU op library (src/dnnl_ops): the reference's CMake wants oneDNN 1.x; the image only has the oneDNN 3.x *header*
# that ships inside torch (no libdnnl).  The op files that are plain C++/OpenMP (embedding lookup, sparse and dense
# optimizer updates, initialisers, ...) compile against that header and need no dnnl symbol, so they are built as
# they are; the files that call the oneDNN 1.x API (matmul, conv, pooling, ...) do not compile and are left out:
# the Python side then uses its numpy fallback for those ops, exactly as python/hetu/_base.py provides.
TORCH_INC="$(python -c 'import torch,os;print(os.path.join(os.path.dirname(torch.__file__),"include"))')"
CPU_OPS="ArraySet BroadcastTo Dropout EmbeddingLookup Gelu Initializers Optimizers Pad ReduceSumAxisZero Reshape Transpose"
for n in $CPU_OPS; do
    o="$OBJ/cpu_$n.o"
    [ -f "$o" ] || g++ -O3 -fPIC -fopenmp -std=c++17 -w -I"$TORCH_INC" -I"$REF/src/common" -I"$REF/src/header" \
        -c "$REF/src/dnnl_ops/$n.cpp" -o "$o"
done
nvcc -shared -o "$LIB/libc_runtime_api.so" "$OBJ"/*.o \
    -L/usr/lib/x86_64-linux-gnu -lcudart -lcublas -lcusparse -lcurand -lcudnn -lgomp
echo "built $LIB/libc_runtime_api.so"

# --- MPI stand-in + the reference's MPI/NCCL communicator ---------------------------------------
g++ -O2 -fPIC -shared -o "$LIB/libmpi.so" "$HERE/mpi_stub/mpi_stub.cc" -I"$HERE/mpi_stub" -lpthread
$NVCC -I"$HERE/mpi_stub" -x cu -c "$REF/src/communication/mpi_nccl_communication.cu" -o "$OBJ/comm/mpi_nccl_communication.o"
nvcc -shared -o "$LIB/lib_mpi_nccl_runtime_api.so" "$OBJ/comm/mpi_nccl_communication.o" \
    -L"$LIB" -lmpi -L/usr/lib/x86_64-linux-gnu -lnccl -lcudart -Xlinker -rpath -Xlinker '$ORIGIN'
echo "built $LIB/lib_mpi_nccl_runtime_api.so"
