"""Time one float32 convolution: python tools/fconv_probe.py C H O K stride [batch]   (TKB_FCONV_DIRECT=1: direct kernel)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tachikoma_b200 import relay as R, runtime as RT, tachikoma as T

C, H, O, K, S = [int(v) for v in sys.argv[1:6]]
B = int(sys.argv[6]) if len(sys.argv) > 6 else 8
rng = np.random.default_rng(0)
x = R.var("x", (B, C, H, H))
w = R.const((rng.standard_normal((O, C, K, K)) / np.sqrt(C * K * K)).astype("float32"))
f = R.Function([x], R.relu(R.add(R.conv2d(x, w, strides=(S, S), padding=(K // 2, K // 2)), R.const(rng.standard_normal((O, 1, 1)).astype("float32")))))
part = T.partition_for_tachikoma(f)
sym, js, names, consts = T.serialize(part)
mod = RT.TachikomaJSONRuntime(sym, js, names)
mod.init(consts)
xd = torch.randn((B, C, H, H), device="cuda")
out = torch.empty(tuple(part.body.shape), dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    mod.run(xd, out, stream=st)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    mod.run(xd, out, stream=st)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / 20
OH = (H + 2 * (K // 2) - K) // S + 1
fl = 2.0 * B * OH * OH * O * C * K * K
print(f"fconv C={C} H={H} O={O} K={K} s={S} b={B}: {us:.1f} us, {fl / us / 1e6:.2f} TFLOP/s fp32; {mod.plan_summary().splitlines()[1].strip()[:80]}")
