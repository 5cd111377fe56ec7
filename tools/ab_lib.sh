#!/bin/bash
# A/B of two builds of the library on one box: libmoca_b200_base.so (previous commit) vs libmoca_b200.so (working tree)
mkdir -p gpurun_out
cp moca_video_b200/libmoca_b200.so /tmp/new.so
for i in 0 1; do
  cp moca_video_b200/libmoca_b200_base.so moca_video_b200/libmoca_b200.so
  echo "== base"; timeout 300 python tools/profile_forward.py 2 gpurun_out/ops_base.csv 2>&1 | grep -E "forward|gemm |attn_self"
  cp /tmp/new.so moca_video_b200/libmoca_b200.so
  echo "== new"; timeout 300 python tools/profile_forward.py 2 gpurun_out/ops_new.csv 2>&1 | grep -E "forward|gemm |attn_self"
done
python - <<PY
import csv,collections,re
def load(p):
    agg=collections.defaultdict(lambda:[0,0.0])
    for r in csv.DictReader(open(p)):
        k=re.sub(r' tiles=\d+','',r["label"].split(" ",1)[1]); agg[k][0]+=1; agg[k][1]+=float(r["us"])
    return agg
a=load("gpurun_out/ops_base.csv"); b=load("gpurun_out/ops_new.csv")
for k in sorted(set(a)&set(b), key=lambda k:-abs(a[k][1]-b[k][1]))[:25]:
    print(f"{a[k][1]/a[k][0]:7.1f} -> {b[k][1]/b[k][0]:7.1f} x{a[k][0]:3d} ({b[k][1]-a[k][1]:+8.1f})  {k}")
PY
