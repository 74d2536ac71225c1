#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "qnn_mul or concatenate or quantize or float_edges or depthwise or mobilenet or reference_goldens or chained" > gpurun_out/Tn3.log 2>&1; echo "tests rc=$?"
tail -6 gpurun_out/Tn3.log
cat > /tmp/dwprobe.py <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import torch
from tachikoma_b200 import models, runtime as RT, tachikoma as T
name = "mobilenetv2"; batch = 128
part = T.partition_for_tachikoma(models.MODELS[name](batch=batch))
sym, js, names, consts = T.serialize(part)
mod = RT.TachikomaJSONRuntime(sym, js, names)
mod.set_option("conv_impl", "tc"); mod.init(consts)
x = torch.from_numpy(models.synthetic_input(name, batch)).cuda()
out = torch.empty(tuple(part.body.shape), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(3): mod.run(x, out, stream=st, trace_mode=0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): mod.run(x, out, stream=st, trace_mode=0)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
print("mobilenetv2 b128 TKB_DW_OLD=%s: %.3f ms per step = %.0f img/s" % (os.environ.get("TKB_DW_OLD"), ms, batch / ms * 1e3))
PY
TKB_DW_OLD=1 python /tmp/dwprobe.py; python /tmp/dwprobe.py
TKB_MODEL=mobilenetv2 TKB_BATCH=128 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/launches_mbv2.csv python tools/profile_step.py tc 0 2 > gpurun_out/ncu_mbv2.log 2>&1
python - <<'PY'
import csv
rows=[r for r in csv.DictReader([l for l in open('gpurun_out/launches_mbv2.csv') if not l.startswith('==')])]
ids=sorted(set(int(r['ID']) for r in rows)); half=ids[len(ids)//2:]
agg={}
for r in rows:
    if int(r['ID']) not in half: continue
    k=r['Kernel Name'][:48]; a=agg.setdefault(k,[0,0.0,0.0]); v=float(r['Metric Value'].replace(',',''))
    if r['Metric Name']=='gpu__time_duration.sum': a[0]+=1; a[1]+=v/1000 if r['Metric Unit']=='ns' else v
    else: a[2]+= v*{'byte':1,'Kbyte':1e3,'Mbyte':1e6,'Gbyte':1e9}.get(r['Metric Unit'],1)
for k,(n,t,b) in sorted(agg.items(), key=lambda kv:-kv[1][1])[:12]: print("%-50s n=%3d %8.1f us %8.1f MB dram %6.2f TB/s" % (k,n,t,b/1e6,b/t/1e6 if t else 0))
PY
