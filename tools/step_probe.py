"""Step time of one model, compute-only and traced (mode 1), under environment experiments:
python tools/step_probe.py [KEY=VAL ...]   (each argument is one extra environment variant, e.g. TKB_SMS=132)"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import os, sys
sys.path.insert(0, %r)
import torch
from tachikoma_b200 import models, runtime as RT, tachikoma as T
name = os.environ.get("TKB_MODEL", "resnet18"); batch = int(os.environ.get("TKB_BATCH", "64"))
part = T.partition_for_tachikoma(models.MODELS[name](batch=batch))
sym, js, names, consts = T.serialize(part)
mod = RT.TachikomaJSONRuntime(sym, js, names)
mod.set_option("conv_impl", os.environ.get("TKB_IMPL", "tc")); mod.init(consts)
x = torch.from_numpy(models.synthetic_input(name, batch)).cuda()
out = torch.empty(tuple(part.body.shape), dtype=getattr(torch, str(part.body.dtype)), device="cuda")
st = torch.cuda.current_stream().cuda_stream
res = []
for mode in (0, 1):
    for _ in range(5): mod.run(x, out, stream=st, trace_mode=mode)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(100): mod.run(x, out, stream=st, trace_mode=mode)
    e1.record(); torch.cuda.synchronize()
    res.append(e0.elapsed_time(e1) * 10)
print("compute-only %%.1f us   traced %%.1f us   (%%.0f / %%.0f per s)" %% (res[0], res[1], batch * 1e6 / res[0], batch * 1e6 / res[1]))
''' % ROOT
for v in ["base"] + sys.argv[1:]:
    e = dict(os.environ)
    if v != "base":
        for kv in v.split(","):
            k, val = kv.split("=", 1); e[k] = val
    r = subprocess.run([sys.executable, "-c", CHILD], env=e, capture_output=True, text=True, timeout=600)
    print("%-40s %s" % (v, (r.stdout.strip().splitlines() or [r.stderr.strip()[-300:]])[-1]), flush=True)
