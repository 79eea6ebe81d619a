#!/bin/bash
# A/B helper: builds projectiontrainer_b200/libpt_stage1_<tag>.so from the csrc/ of git revision $1 (default HEAD)
# next to the working-tree library; probes select it with PT_STAGE1_LIB.  usage: tools/build_ab.sh [rev] [tag]
set -e
rev=${1:-HEAD}; tag=${2:-old}
root=$(cd "$(dirname "$0")/.." && pwd)
tmp=$(mktemp -d)
mkdir -p $tmp/projectiontrainer_b200/csrc $tmp/include
git -C $root archive $rev projectiontrainer_b200/csrc include | tar -x -C $tmp
objs=()
for f in $tmp/projectiontrainer_b200/csrc/*.cu; do
  o=$tmp/$(basename $f .cu).o
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --use_fast_math \
       --expt-relaxed-constexpr -I $tmp/include -c $f -o $o &
  objs+=($o)
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $root/projectiontrainer_b200/libpt_stage1_$tag.so "${objs[@]}" -lcudart
rm -rf $tmp
echo $root/projectiontrainer_b200/libpt_stage1_$tag.so
