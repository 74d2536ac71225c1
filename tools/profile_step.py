"""One ResNet-18 batch-64 step (after 2 warm-up steps) for ncu: python tools/profile_step.py [conv_impl] [trace_mode]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from tachikoma_b200 import models, runtime as RT, tachikoma as T

impl = sys.argv[1] if len(sys.argv) > 1 else "tc"
trace = int(sys.argv[2]) if len(sys.argv) > 2 else 0
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
name = os.environ.get("TKB_MODEL", "resnet18")
batch = int(os.environ.get("TKB_BATCH", "64"))
part = T.partition_for_tachikoma(models.MODELS[name](batch=batch))
sym, js, names, consts = T.serialize(part)
mod = RT.TachikomaJSONRuntime(sym, js, names)
mod.set_option("conv_impl", impl)
mod.init(consts)
mod.trace_enable(trace)
x = torch.from_numpy(models.synthetic_input(name, batch)).cuda()
out = torch.empty(tuple(part.body.shape), dtype=torch.uint8, device="cuda")
for _ in range(steps):
    mod.run(x, out, stream=torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
if os.environ.get("TKB_PLAN"):
    print(mod.plan_summary())
print("launches per step:", mod.launch_count() // steps, "checksum", int(out.sum().item()))
