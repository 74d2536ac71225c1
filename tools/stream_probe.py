"""Throughput of back-to-back ResNet-18 b64 steps issued round-robin on S streams (one workspace per stream)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tachikoma_b200 import models, runtime as RT, tachikoma as T
name = os.environ.get("TKB_MODEL", "resnet18"); batch = int(os.environ.get("TKB_BATCH", "64"))
part = T.partition_for_tachikoma(models.MODELS[name](batch=batch))
sym, js, names, consts = T.serialize(part)
for impl in (os.environ.get("TKB_IMPLS", "tc_nochain,tc").split(",")):
    mod = RT.TachikomaJSONRuntime(sym, js, names)
    mod.set_option("conv_impl", impl); mod.init(consts)
    for S in (1, 2, 3, 4):
        for mode in (0, 1):
            streams = [torch.cuda.Stream() for _ in range(S)]
            xs = [torch.from_numpy(models.synthetic_input(name, batch, seed=1 + i)).cuda() for i in range(S)]
            outs = [torch.empty(tuple(part.body.shape), dtype=torch.uint8, device="cuda") for _ in range(S)]
            for i in range(3 * S):
                mod.run(xs[i % S], outs[i % S], stream=streams[i % S].cuda_stream, trace_mode=mode)
            torch.cuda.synchronize()
            K = 60
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for st in streams: st.wait_event(e0)
            for i in range(K):
                mod.run(xs[i % S], outs[i % S], stream=streams[i % S].cuda_stream, trace_mode=mode)
            for st in streams: torch.cuda.current_stream().wait_stream(st)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / K
            print("%-11s streams %d trace_mode %d: %.1f us per step = %.0f img/s" % (impl, S, mode, ms * 1e3, batch / ms * 1e3), flush=True)
    del mod
