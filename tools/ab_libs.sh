#!/bin/bash
# A/B of side-by-side builds of the library (NRW_LIB, engine.py) on one box: for every lib the score kernel's
# variants on the given mesh; same seeds, so score_crc must agree between libs.
# Usage (on the box): bash tools/ab_libs.sh <tag> <mesh> <variants> lib1.so lib2.so ...
set -u
tag=$1; mesh=$2; variants=$3; shift 3
mkdir -p gpurun_out
for lib in "$@"; do
    NRW_LIB=$PWD/$lib NRW_VARIANTS=$variants timeout 200 python tools/k1s_variants.py $mesh 2>> gpurun_out/${tag}_ab.err | tee -a gpurun_out/${tag}_ab.jsonl | cut -c1-330
done
