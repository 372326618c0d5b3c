#!/bin/bash
# Usage (on the GPU box): tools/ab.sh <workload> <variant> [<variant> ...]  -- times each variants/<name>.so on one bench workload.
w=$1; shift
cp scran_b200/libscranb200.so /tmp/keep.so
for v in "$@"; do
  cp variants/$v.so scran_b200/libscranb200.so
  python bench.py --workload $w --steps 3 --warmup 2 --no-e2e --no-cpu > /tmp/ab_$v.json 2>/tmp/ab_$v.err || { echo "$v FAILED"; tail -3 /tmp/ab_$v.err; continue; }
  python -c "
import json,sys;d=json.load(open('/tmp/ab_$v.json'));print('$v','$w',round(d['ms_per_step'],2),{k:round(x['ms'],2) for k,x in d['kernels'].items()})"
done
cp /tmp/keep.so scran_b200/libscranb200.so
