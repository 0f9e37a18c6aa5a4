#!/bin/bash
# Builds library variants for same-box A/B runs (tools/gpu.sh ab, tools/gpu.sh ab) into tools/ab/ (git-ignored, travels
# with gpurun).  usage: tools/build_ab.sh <tag> [ENV=VALUE ...]    e.g.  tools/build_ab.sh f2 SEVENN_TAB_F2=1
#                                                                        tools/build_ab.sh cpw1 SEVENN_TAB_CPW=1
# The default library is rebuilt afterwards so the tree is left as it was.
set -e
tag=$1; shift
mkdir -p tools/ab
env "$@" python -m sevennet_exp_b200.build | tail -1
cp sevennet_exp_b200/_build/libsevenn_b200.so tools/ab/libsevenn_b200_$tag.so
python -m sevennet_exp_b200.build | tail -1
[ -f tools/ab/libsevenn_b200_ref.so ] || cp sevennet_exp_b200/_build/libsevenn_b200.so tools/ab/libsevenn_b200_ref.so
ls -la tools/ab
