#!/bin/bash
# interleaved timing of several builds on one box: tools/ab/multi.sh A dbg1 dbg6 ... (each tools/ab/<name>.so)
keep=$(mktemp); cp hivemind_b200/libhm_engine.so $keep
for rep in 1 2; do for v in "$@"; do cp tools/ab/$v.so hivemind_b200/libhm_engine.so; echo -n "$v: "; timeout 120 python tools/ab/kernels.py 2>/dev/null | tail -1; done; done
cp $keep hivemind_b200/libhm_engine.so
