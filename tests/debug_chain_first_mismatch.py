"""Locate the first trace record of the small ResNet-18 that differs from the oracle (chained launch debugging)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # repo root (this checker lives under tests/: it uses the oracle)
import numpy as np
from oracle import interp
from tachikoma_b200 import models, runtime as RT, tachikoma as T, trace as TR

batch, image = int(os.environ.get("B", 2)), int(os.environ.get("IMG", 64))
func = models.resnet18(batch=batch, image=image, num_classes=100)
x = np.ascontiguousarray(models.synthetic_input("resnet18", batch)[:, :, :image, :image])
part = T.partition_for_tachikoma(func)
sym, js, names, consts = T.serialize(part)
mod = RT.TachikomaJSONRuntime(sym, js, names)
mod.set_option("conv_impl", "tc"); mod.init(consts)
print(mod.plan_summary()[-1500:])
out = np.empty(part.body.shape, dtype=part.body.dtype)
for it in range(2):
    mod.run(x, out, trace_mode=2)
    got = TR.load_trace(mod.trace_blob())
    _, recs = interp.run_json(js, consts, [x], want_trace=True)
    nbad = 0
    for (gn, ga), (wn, wa) in zip(got, recs):
        wa = np.asarray(wa)
        if not np.array_equal(ga, wa):
            d = np.argwhere(ga != wa)
            print("iter", it, "MISMATCH", gn, ga.shape, "n_diff", len(d), "first", d[:3].tolist(), "chan set", sorted(set(d[:, 1].tolist()))[:12] if ga.ndim == 4 else "")
            nbad += 1
            if nbad >= 3:
                break
    print("iter", it, "bad records:", nbad)
