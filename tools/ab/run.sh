#!/bin/bash
# interleaved A/B on one box: tools/ab/A.so and tools/ab/B.so are swapped in as hivemind_b200/libhm_engine.so
for rep in 1 2 3; do for v in A B; do cp tools/ab/$v.so hivemind_b200/libhm_engine.so; echo -n "$v: "; timeout 120 python tools/ab/kernels.py "$@" 2>/dev/null | tail -1; done; done
cp tools/ab/B.so hivemind_b200/libhm_engine.so
