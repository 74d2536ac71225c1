"""Compute-only step time of ResNet-18 b64 under a set of TKB_DBG experiment bits (chained launch): python tools/chain_probe.py"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import os, sys, time
sys.path.insert(0, %r)
import torch
from tachikoma_b200 import models, runtime as RT, tachikoma as T
name = os.environ.get("TKB_MODEL", "resnet18"); batch = int(os.environ.get("TKB_BATCH", "64"))
part = T.partition_for_tachikoma(models.MODELS[name](batch=batch))
sym, js, names, consts = T.serialize(part)
mod = RT.TachikomaJSONRuntime(sym, js, names)
mod.set_option("conv_impl", os.environ.get("TKB_IMPL", "tc")); mod.init(consts)
x = torch.from_numpy(models.synthetic_input(name, batch)).cuda()
out = torch.empty(tuple(part.body.shape), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(5): mod.run(x, out, stream=st, trace_mode=0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50): mod.run(x, out, stream=st, trace_mode=0)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 50
mod.set_option("time_tc", "1")
RT.spin(3000.0, st); mod.run(x, out, stream=st, trace_mode=0); torch.cuda.synchronize()
tc, n = mod.tc_time()
import ctypes
st = (ctypes.c_uint64 * 8)()
RT.load_library().tkb_debug_chain_stats(st, 1)
mod.run(x, out, stream=st_ if False else st and torch.cuda.current_stream().cuda_stream, trace_mode=0); torch.cuda.synchronize()
RT.load_library().tkb_debug_chain_stats(st, 1)
print("step %%.1f us   tc launches %%d: %%.1f us   stalled items %%d, wait %%.1f us per CTA" %% (ms * 1e3, n, tc * 1e3, st[0], st[1] / 148 / 1965.0)
      + ("; per tile: epilogue %%.2f us, store %%.2f us, publish %%.2f us" %% (st[3] / st[2] / 1965.0, st[4] / st[2] / 1965.0, st[5] / st[2] / 1965.0) if st[2] else ""))
''' % ROOT
for label, env in [("baseline nochain", {"TKB_IMPL": "tc_nochain"}), ("chain: siblings only (default)", {}), ("chain: every layer, split at dependencies", {"TKB_CHAIN_POLICY": "3"}),
                   ("chain: dependent layers merged", {"TKB_CHAIN_POLICY": "2"})] + [(a, dict(b)) for a, b in [tuple(x.split("=", 1)) and (x, dict([x.split("=", 1)])) for x in sys.argv[1:]]]:
    e = dict(os.environ); e.update(env)
    r = subprocess.run([sys.executable, "-c", CHILD], env=e, capture_output=True, text=True, timeout=300)
    print("%-45s %s" % (label, (r.stdout.strip().splitlines() or [r.stderr.strip()[-300:]])[-1]), flush=True)
