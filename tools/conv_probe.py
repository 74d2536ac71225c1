"""Time one conv layer in isolation: python tools/conv_probe.py C H O K stride [batch]  (TKB_DBG selects experiments)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tachikoma_b200 import relay as R, runtime as RT, tachikoma as T

C, H, O, K, S = [int(v) for v in sys.argv[1:6]]
B = int(sys.argv[6]) if len(sys.argv) > 6 else 64
rng = np.random.default_rng(0)
x = rng.integers(0, 256, (B, C, H, H)).astype("uint8")
w = rng.integers(-127, 128, (O, C, K, K)).astype("int8")
sw = (rng.uniform(0.5, 1.5, O) / (127 * np.sqrt(K * K * C))).astype("float32")
d = R.var("data", x.shape, "uint8")
e = R.qnn_conv2d(d, R.const(w), 0, 0, np.float32(0.02), R.const(sw), (K, K), O, strides=(S, S), padding=(K // 2, K // 2))
e = R.bias_add(e, R.const(rng.integers(-4096, 4096, O).astype("int32")))
e = R.qnn_requantize(e, R.const((np.float32(0.02) * sw).astype("float32")), 0, np.float32(0.024), 0, axis=1, out_dtype="int32")
e = R.cast(R.clip(e, 0.0, 255.0), "uint8")
# NHWC in/out so that no layout conversion is timed: wrap with data_layout NHWC instead
d2 = R.var("data", (B, H, H, C), "uint8")
e2 = R.qnn_conv2d(d2, R.const(np.ascontiguousarray(w.transpose(2, 3, 1, 0))), 0, 0, np.float32(0.02), R.const(sw), (K, K), O,
                  strides=(S, S), padding=(K // 2, K // 2), data_layout="NHWC", kernel_layout="HWIO")
e2 = R.bias_add(e2, R.const(rng.integers(-4096, 4096, O).astype("int32")), axis=3)
e2 = R.qnn_requantize(e2, R.const((np.float32(0.02) * sw).astype("float32")), 0, np.float32(0.024), 0, axis=3, out_dtype="int32")
e2 = R.cast(R.clip(e2, 0.0, 255.0), "uint8")
part = T.partition_for_tachikoma(R.Function([d2], e2))
sym, js, names, consts = T.serialize(part)
mod = RT.TachikomaJSONRuntime(sym, js, names)
mod.set_option("conv_impl", os.environ.get("TKB_IMPL", "tc"))
mod.init(consts)
xd = torch.from_numpy(np.ascontiguousarray(x.transpose(0, 2, 3, 1))).cuda()
out = torch.empty(tuple(part.body.shape), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    mod.run(xd, out, stream=st)
torch.cuda.synchronize()
mod.set_option("time_tc", "1")
n = 10
tot = 0.0
for _ in range(n):
    RT.spin(300.0, st)  # host queues the launch behind the spin: the events see device time only
    mod.run(xd, out, stream=st)
    torch.cuda.synchronize()
    ms, cnt = mod.tc_time()
    tot += ms
us = tot * 1e3 / n
OH = (H + 2 * (K // 2) - K) // S + 1
macs = B * OH * OH * O * C * K * K
print(f"dbg={os.environ.get('TKB_DBG','0')} C={C} H={H} O={O} K={K} s={S}: {us:.1f} us kernel (CUDA events), {2*macs/us/1e6:.0f} TOPS; plan: {mod.plan_summary().splitlines()[1].strip()[:110]}")
