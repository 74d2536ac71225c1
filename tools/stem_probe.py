"""Time the ResNet stem (7x7/2 conv on the NCHW uint8 image) in isolation: python tools/stem_probe.py [batch] [zp]
TKB_DBG selects experiments (1 no epilogue work, 4 no patch loads, 32 no MMAs, 8 zero tiles)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tachikoma_b200 import relay as R, runtime as RT, tachikoma as T

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
ZP = int(sys.argv[2]) if len(sys.argv) > 2 else 128
O = int(os.environ.get("STEM_O", "64"))
rng = np.random.default_rng(0)
x = rng.integers(0, 256, (B, 3, 224, 224)).astype("uint8")
w = rng.integers(-127, 128, (O, 3, 7, 7)).astype("int8")
sw = (rng.uniform(0.5, 1.5, O) / (127 * np.sqrt(147))).astype("float32")
d = R.var("data", x.shape, "uint8")
e = R.qnn_conv2d(d, R.const(w), ZP, 0, np.float32(0.02), R.const(sw), (7, 7), O, strides=(2, 2), padding=(3, 3))
e = R.bias_add(e, R.const(rng.integers(-4096, 4096, O).astype("int32")))
e = R.qnn_requantize(e, R.const((np.float32(0.02) * sw).astype("float32")), 0, np.float32(0.024), 0, axis=1, out_dtype="int32")
e = R.cast(R.clip(e, 0.0, 255.0), "uint8")
e = R.max_pool2d(e, (3, 3), (2, 2), (1, 1))  # keeps the conv output channels-last (no boundary transpose is timed)
part = T.partition_for_tachikoma(R.Function([d], e))
sym, js, names, consts = T.serialize(part)
mod = RT.TachikomaJSONRuntime(sym, js, names)
mod.set_option("conv_impl", "tc")
mod.init(consts)
xd = torch.from_numpy(x).cuda()
out = torch.empty(tuple(part.body.shape), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    mod.run(xd, out, stream=st)
torch.cuda.synchronize()
mod.set_option("time_tc", "1")
tot, n = 0.0, 10
for _ in range(n):
    RT.spin(300.0, st)
    mod.run(xd, out, stream=st)
    torch.cuda.synchronize()
    ms, cnt = mod.tc_time()
    tot += ms
print(f"dbg={os.environ.get('TKB_DBG','0')} B={B} zp={ZP}: stem {tot * 1e3 / n:.1f} us (CUDA events, ~3.7 us of event overhead); "
      + [l.strip() for l in mod.plan_summary().splitlines() if "tcgen05" in l][0][:150])
