#!/bin/bash
# tuning aid: per-CTA barrier arrivals of the select kernel on one workload (default: the headline)
W=${1:-resnet18_bior4.4_L5_p50_unclamped}
WT_SELECT_STAMPS_CTAS=1 timeout 100 python tools/select_probe.py $W 2>&1 | grep "cta  *[0-9]*:" | tail -296 > gpurun_out/ctas2.txt
python - <<PY
rows=[l.split() for l in open("gpurun_out/ctas2.txt")]
s0=[float(r[3]) for r in rows]; s1=[float(r[4]) for r in rows]; s2=[float(r[5]) for r in rows]
print("tail", " ".join(f"{a:.0f}" for a in s2[:20]), "end", rows[0][7], rows[0][8])
for i in range(0,len(rows),8): print(i, " ".join(f"{a:4.0f}/{b:4.0f}" for a,b in zip(s0[i:i+8],s1[i:i+8])))
PY
