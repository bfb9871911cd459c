#!/bin/bash
# compute-sanitizer pass over the small-shape workload (one case per process, each under its own timeout); logs in $1.
OUT=${1:-gpurun_out/sanitizer}; T=${2:-150}; mkdir -p $OUT
CS=/usr/local/cuda/bin/compute-sanitizer
one() { local tool=$1 c=$2; timeout $T $CS --tool $tool --print-limit 20 --log-file $OUT/${tool}_$c.log python bench/sanitizer_target.py $c > $OUT/${tool}_$c.out 2>&1; echo "$tool $c rc=$? $(tail -n 1 $OUT/${tool}_$c.log 2>/dev/null) $(tail -n 1 $OUT/${tool}_$c.out | cut -c1-60)"; }
one memcheck kernels
one memcheck resnet18
one racecheck kernels
one synccheck resnet18
one memcheck lenet
one initcheck lenet
one memcheck vgg11
one racecheck resnet18
one synccheck kernels
