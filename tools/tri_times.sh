# per-kernel times of a triangle layer (ncu launch list; cold-cache, serialised) + event timing
N=${1:-131072}
python tools/probe.py --n $N --kernels auto --modes 1 2>&1 | grep attn
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/tri_launches.csv python tools/probe.py --n $N --kernels auto --modes 1 > /dev/null 2>&1
python - <<PY
import csv,collections
rows=[r for r in csv.reader(open("gpurun_out/tri_launches.csv")) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value"); gi=hdr.index("Grid Size")
agg=collections.defaultdict(list)
for r in rows[1:]:
    try: agg[(r[ki][:50],r[gi])].append(float(r[vi].replace(",","")))
    except: pass
for k,v in agg.items():
    if "tmx" in k[0]: print(k, len(v), "avg us %.1f"%(sum(v)/len(v)/1e3))
PY
