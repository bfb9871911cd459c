#!/bin/bash
# compute-sanitizer pass over the small-shape workload (one case per process, each under its own timeout); logs in $1.
OUT=${1:-gpurun_out/sanitizer}; mkdir -p $OUT
CS=/usr/local/cuda/bin/compute-sanitizer
for c in kernels lenet resnet18 vgg11; do
  timeout 600 $CS --tool memcheck --print-limit 20 --log-file $OUT/memcheck_$c.log python bench/sanitizer_target.py $c > $OUT/memcheck_$c.out 2>&1; echo "memcheck $c rc=$? $(tail -n 1 $OUT/memcheck_$c.log)"
done
for c in kernels resnet18; do
  timeout 600 $CS --tool racecheck --print-limit 20 --log-file $OUT/racecheck_$c.log python bench/sanitizer_target.py $c > $OUT/racecheck_$c.out 2>&1; echo "racecheck $c rc=$? $(tail -n 1 $OUT/racecheck_$c.log)"
  timeout 600 $CS --tool synccheck --print-limit 20 --log-file $OUT/synccheck_$c.log python bench/sanitizer_target.py $c > $OUT/synccheck_$c.out 2>&1; echo "synccheck $c rc=$? $(tail -n 1 $OUT/synccheck_$c.log)"
done
timeout 600 $CS --tool initcheck --print-limit 20 --log-file $OUT/initcheck_lenet.log python bench/sanitizer_target.py lenet > $OUT/initcheck_lenet.out 2>&1; echo "initcheck lenet rc=$? $(tail -n 1 $OUT/initcheck_lenet.log)"
