#!/bin/bash
# tools/build_variant.sh <name> "<extra nvcc flags>"  ->  build/lib_<name>.so  (A/B of kernel variants: tools/ab.sh)
name=$1; extra=$2
make -s -j8 -C bluesky_gym_hallucination_b200/csrc OUT=../../build/lib_$name.so OBJDIR=../../build/obj_$name EXTRA="$extra"
