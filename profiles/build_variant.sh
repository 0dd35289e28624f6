#!/bin/bash
# usage: build_variant.sh NAME [GIT_REV|-] [-Dflags...]  -> scratch/lib_NAME.so from the working tree (-) or from a commit
NAME=$1; REV=${2:--}; shift 2
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared"
mkdir -p scratch
if [ "$REV" = "-" ]; then SRC=.; else SRC=/tmp/w/src_$REV; rm -rf $SRC; mkdir -p $SRC; git archive $REV parc_b200/csrc include | tar -x -C $SRC; fi
nvcc $F -I $SRC/include "$@" -Xptxas -v $SRC/parc_b200/csrc/*.cu -o scratch/lib_$NAME.so 2>&1 | grep -A2 "env_step_kernel" | grep -E "registers" | head -1 | sed "s/^/$NAME: /"
