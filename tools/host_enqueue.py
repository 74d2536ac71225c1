"""Host enqueue time vs device time of a ResNet-18 b64 step (is the step launch-bound?)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tachikoma_b200 import models, runtime as RT, tachikoma as T
part = T.partition_for_tachikoma(models.resnet18(batch=64))
sym, js, names, consts = T.serialize(part)
mod = RT.TachikomaJSONRuntime(sym, js, names)
mod.set_option("conv_impl", "tc")
mod.init(consts)
x = torch.from_numpy(models.synthetic_input("resnet18", 64)).cuda()
out = torch.empty(tuple(part.body.shape), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for trace in (0, 1):
    mod.trace_enable(trace)
    for _ in range(5):
        mod.run(x, out, stream=st)
    torch.cuda.synchronize()
    RT.spin(20000.0, st)   # GPU busy for 20 ms: the loop below measures pure host enqueue time
    t0 = time.perf_counter()
    for _ in range(20):
        mod.run(x, out, stream=st)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"trace={trace}: host enqueue {1e3*(t1-t0)/20:.3f} ms/step, total incl. spin {1e3*(t2-t0):.1f} ms")
