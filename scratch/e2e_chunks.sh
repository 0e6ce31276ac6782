#!/bin/bash
N=$1
for c in 2 4 16; do
echo "chunks=$c"
TCB_BENCH_CHUNKS=$c bash scratch/e2e_nN.sh $N
done
