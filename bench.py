#!/usr/bin/env python
"""bench.py -- traced int8 inferences/s of the tachikoma QNN hot path on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU algorithm (oracle port)

Workload (N=1): BASELINE.json configs[2], int8 ResNet-18 224x224, integer-only, batch 64 per GPU, full
per-op trace, synthetic uint8 images and random-init quantized weights (tachikoma_b200/models.py).
A "step" is one batch of 64 images through partition -> JSON runtime -> kernels.  N>1: one process per
GPU (torchrun), the batch shards by rank (weak scaling, no data-path collective); torch.distributed is
used only for the barrier and the max-over-ranks of the device time.

Legs (all timed with CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks):
  compute_only   input resident in HBM, trace off
  value          input resident in HBM, every operator's tensors captured on the device in their logical layout
                 (trace mode 1: readable afterwards through tkb_trace_blob, verified below)
  e2e            pinned HOST input -> C ABI -> logits + every distinct traced tensor in pinned host memory
                 (trace mode 2: H2D and D2H inside the timed region), two batches in flight on two streams;
                 config.trace_to_host_e2e_one_batch_in_flight is the strictly serial form (run, wait for the trace, run)
  e2e_serialized as e2e, plus the complete NDArray-list blob (repeated records filled in) every step
  sustained      >= 2 s of back-to-back `value` steps (clocks / power under real load)
After the timed legs the run PROVES what it timed: the serialized trace of the first and of the last end-to-end step
and the device-resident trace are compared with the CPU oracle on a subsample of the batch ("bit_exact_trace"); a
mismatch makes the process exit non-zero.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "traced int8 inferences/sec"
UNIT = "inferences/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="resnet18")
    ap.add_argument("--batch", type=int, default=64, help="images per GPU per step")
    ap.add_argument("--conv-impl", default="auto")  # the runtime's default plan (all ResNet-18 layers on tcgen05 either way)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-verify", action="store_true", help="skip the oracle check of the timed traces")
    ap.add_argument("--sustained-s", type=float, default=2.0)
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap,utilization.gpu")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.proc = gpu_index, [], None
        self.marks = {}

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def mark(self, name):
        self.marks[name] = time.perf_counter()

    def _summary(self, rows):
        sm, mx, pw, reasons = [], [], [], set()
        for _, r in rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); pw.append(float(r[3]))
                for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}, None
        time.sleep(0.15)
        self.proc.terminate()
        allrows = self._summary(self.rows)
        sus = None
        if "sustained_begin" in self.marks and "sustained_end" in self.marks:
            a, b = self.marks["sustained_begin"], self.marks["sustained_end"]
            sus = self._summary([r for r in self.rows if a + 0.2 <= r[0] <= b])
        return allrows, sus


SHAPES = {"mlp": "784-128-10 MLP (configs[0])", "lenet5": "LeNet-5 28x28 (configs[1])",
          "resnet18": "resnet18 224x224 (configs[2])", "mobilenetv2": "mobilenetv2 224x224 (configs[3])",
          "bert_encoder": "BERT-base encoder, 1 of 12 layers, seq 128, hidden 768 (configs[4]; batch = sequences)",
          "bert_base": "BERT-base encoder, all 12 layers, seq 128, hidden 768 (configs[4]; batch = sequences)"}


def workload_string(model, batch):
    return (f"int8 {SHAPES.get(model, model)} integer-only, batch {batch} per GPU, full per-op trace "
            "(BASELINE.json configs)")


def build_model(name, batch):
    from tachikoma_b200 import models, tachikoma as T
    func = models.MODELS[name](batch=batch)
    part = T.partition_for_tachikoma(func)
    symbol, graph_json, const_names, consts = T.serialize(part)
    return part, symbol, graph_json, const_names, consts


def cpu_oracle_rate(name, sample, repeats=1):
    """images/s of the oracle (CPU port of the reference algorithm) on `sample` images, all host threads."""
    from oracle import interp
    from tachikoma_b200 import models
    interp.set_threads(os.cpu_count())  # torchrun exports OMP_NUM_THREADS=1
    part, symbol, graph_json, const_names, consts = build_model(name, sample)
    x = models.synthetic_input(name, sample)
    interp.run_json(graph_json, consts, [x])  # warm-up (library load, page faults)
    t0 = time.perf_counter()
    for _ in range(repeats):
        interp.run_json(graph_json, consts, [x], want_trace=True)
    return sample * repeats / (time.perf_counter() - t0)


def run_reference(args, rank):
    """Reference arm: the reference's own CPU implementation is not buildable here (no LLVM/oneDNN), so this
    times the oracle port of its algorithm (a scalar C restatement, NOT oneDNN) on the host cores; rank 0 only."""
    if rank != 0:
        return
    from oracle import interp
    from tachikoma_b200 import models
    cores = interp.set_threads(os.cpu_count())  # torchrun exports OMP_NUM_THREADS=1
    t0 = time.perf_counter()
    probe = cpu_oracle_rate(args.model, 1)
    # bounded sample per step so that (steps + warmup) * sample images finish within ~2 minutes
    sample = int(max(1, min(args.batch, 120.0 * probe / max(1, args.steps + args.warmup))))
    part, symbol, graph_json, const_names, consts = build_model(args.model, sample)
    x = models.synthetic_input(args.model, sample)
    for _ in range(args.warmup):
        interp.run_json(graph_json, consts, [x], want_trace=True)
    t1 = time.perf_counter()
    for _ in range(args.steps):
        interp.run_json(graph_json, consts, [x], want_trace=True)
    dt = time.perf_counter() - t1
    value = sample * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8*s8->s32", "data": "synthetic",
        "config": {"workload": workload_string(args.model, args.batch)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{sample} images/step (bounded sample of the batch of {args.batch}) x {args.steps} steps, scalar C "
                                   f"restatement of the reference lowering (oracle/qnn_oracle.c, not oneDNN), OpenMP on {cores} "
                                   "threads, trace written"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.perf_counter() - t0,
    }
    print(json.dumps(line), file=_JSON_OUT, flush=True)


def _ncu_traffic():
    """DRAM bytes per tcgen05 launch (read + write) from the committed ncu --set full capture of the same step
    (profiles/r02_tc_traffic.json, else round 1's); None when absent."""
    for name in ("r02_tc_traffic.json", "r01_tc_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                return float(json.load(f)["dram_bytes_per_launch"]), name
        except Exception:
            continue
    return None, None


_JSON_OUT = sys.stdout


def _claim_stdout():
    """stdout carries exactly ONE JSON line: keep a private handle to it and point file descriptor 1 at stderr, so
    that nothing a library prints (NCCL's version banner, for one) can land beside the line."""
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def bind_to_gpu_numa_node(local_rank):
    """Pin this rank's host threads (and therefore its first-touch pinned allocations) to the CPUs of the NUMA node its
    GPU hangs off, so that the trace stream's pinned buffers are local to the PCIe root the GPU writes through."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = getattr(torch.cuda.get_device_properties(local_rank), "pci_domain_id", 0)
        dev = torch.cuda.get_device_properties(local_rank).pci_device_id
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0" % (dom, bus, dev)
        node = int(open(path + "/numa_node").read().strip())
        cpulist = open(path + "/local_cpulist").read().strip()
        cpus = set()
        for part in cpulist.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if cpus:
            os.sched_setaffinity(0, cpus)
        return {"numa_node": node, "cpus": len(cpus) if cpus else len(allowed)}
    except Exception as e:  # noqa: BLE001
        return {"numa_node": None, "error": str(e)[:80]}


def verify_against_oracle(model, batch, graph_json, consts, x, logits, blob, sub=4):
    """Compare a GPU-produced (logits, serialized trace) of one batch with the oracle on the first and the last `sub`
    images (inferences are independent: the oracle's records for a sub-batch equal the leading / trailing rows of the
    batch's records).  Returns a list of problems (empty = bit-exact)."""
    from oracle import interp
    from tachikoma_b200 import trace as TR
    recs = TR.load_trace(blob)
    problems = []
    spans = [(0, min(sub, batch))] if batch <= sub else [(0, sub), (batch - sub, batch)]
    for lo, hi in spans:
        n = hi - lo
        part, _, js, _, cs = build_model(model, n)
        rin = x.shape[0] // batch  # input rows per unit of the batch (BERT: 128 token rows per sequence)
        outs, want = interp.run_json(js, cs, [np.ascontiguousarray(x[lo * rin:hi * rin])], want_trace=True)
        k = logits.shape[0] // batch
        if not np.array_equal(np.asarray(outs[0]), logits[lo * k:hi * k]):
            problems.append("logits of images %d..%d differ from the oracle" % (lo, hi))
        if len(want) != len(recs):
            problems.append("trace has %d records, oracle %d" % (len(recs), len(want)))
            continue
        const_names = TR.const_record_names(js)
        for (gn, ga), (wn, wa) in zip(recs, want):
            wa = np.asarray(wa)
            if gn != wn:
                problems.append("record name %s != %s" % (gn, wn))
                break
            if gn in const_names:  # weights, bias, scales, zero points: identical for every batch size
                if ga.shape != wa.shape or not np.array_equal(ga, wa):
                    problems.append("constant record %s differs" % gn)
                continue
            if wa.ndim == 0 or wa.shape[0] % n or ga.shape[0] != batch * (wa.shape[0] // n):
                problems.append("record %s: cannot slice shape %s / %s by batch" % (gn, ga.shape, wa.shape))
                continue
            r = wa.shape[0] // n
            if not np.array_equal(ga[lo * r:hi * r], wa):
                problems.append("record %s differs from the oracle on images %d..%d" % (gn, lo, hi))
        if len(problems) > 8:
            break
    return problems


def main():
    args = parse()
    _claim_stdout()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    from tachikoma_b200 import models, runtime as RT

    torch.cuda.set_device(local_rank)
    numa = bind_to_gpu_numa_node(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world > 1:
            t = torch.tensor([v], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return v

    part, symbol, graph_json, const_names, consts = build_model(args.model, args.batch)
    mod = RT.TachikomaJSONRuntime(symbol, graph_json, const_names)
    mod.set_option("conv_impl", args.conv_impl)
    mod.init(consts)
    macs, hbm_bytes = mod.work()
    # two different batches: the end-to-end leg alternates them, and the first / last steps that get verified differ
    xs_np = [models.synthetic_input(args.model, args.batch, seed=1 + rank + 100 * j) for j in range(2)]
    xs_host = [torch.from_numpy(a).pin_memory() for a in xs_np]
    x_dev = xs_host[0].cuda()
    out_shape = tuple(part.body.shape)
    out_dt = getattr(torch, str(part.body.dtype))
    out_dev = torch.empty(out_shape, dtype=out_dt, device="cuda")
    out_host = torch.empty(out_shape, dtype=out_dt).pin_memory()
    stream = torch.cuda.current_stream().cuda_stream

    def timed(n_warm, n_steps, fn):
        """-> (total ms, launches, per-step ms list).  One event per step boundary: the total is first-to-last event."""
        for i in range(n_warm):
            fn(i - n_warm)
        barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(n_steps + 1)]
        l0 = mod.launch_count()
        ev[0].record()
        for i in range(n_steps):
            fn(i)
            ev[i + 1].record()
        barrier()
        ms = max_over_ranks(ev[0].elapsed_time(ev[-1]))
        per = [ev[i].elapsed_time(ev[i + 1]) for i in range(n_steps)]
        return ms, mod.launch_count() - l0, per

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()

    # (i) compute only: device-resident input, no trace
    ms_compute, _, _ = timed(args.warmup, args.steps, lambda i: mod.run(x_dev, out_dev, stream=stream, trace_mode=0))
    # (ii) headline `value`: device-resident input, every operator's tensors captured on the device (logical layout)
    ms_value, launches, per_value = timed(args.warmup, args.steps, lambda i: mod.run(x_dev, out_dev, stream=stream, trace_mode=1))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    blob_dev = mod.trace_blob() if not args.no_verify else None  # deferred read-out of the device-resident trace
    readout_ms = 1e3 * (time.perf_counter() - t0)
    logits_dev = out_dev.cpu().numpy().copy()

    # (iii) e2e through the C ABI with HOST buffers: H2D of the batch, trace streamed to pinned host, D2H of the logits
    first = {}

    def e2e_step(i):
        j = 0 if i == -args.warmup else (1 if i == args.steps - 1 else i % 2)
        mod.run(xs_host[j], out_host, stream=stream, trace_mode=2)
        mod.trace_sync()
        if i == -args.warmup and not args.no_verify:  # the first end-to-end step of the run (untimed warm-up step)
            first["blob"] = mod.trace_blob()
            first["logits"] = out_host.numpy().copy()

    ms_e2e, _, _ = timed(max(1, args.warmup), args.steps, e2e_step)
    last_blob = mod.trace_blob() if not args.no_verify else None  # the LAST timed step's trace, still in the pinned buffer
    last_logits = out_host.numpy().copy()

    # (iii-b) the same end-to-end work with TWO batches in flight: the caller alternates two streams (each has its own
    # workspace and pinned blob, include/tkb200.h), tkb_run returns once the logits are on the host, and the trace of
    # batch i keeps streaming to the host while batch i+1 is uploaded and computed.  Before a stream is reused the
    # caller waits for that stream's previous trace (tkb_trace_sync_on): every trace is complete in pinned host memory
    # before its buffer is overwritten, which is when a consumer would read it.
    pipe_streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    pipe_out = [torch.empty(out_shape, dtype=out_dt).pin_memory() for _ in range(2)]
    pipe_warm = max(2, args.warmup)

    def pipe_step(i):
        k = i % 2
        sk = pipe_streams[k].cuda_stream
        if i - 2 >= -pipe_warm:
            mod.trace_sync(stream=sk)
        mod.run(xs_host[k], pipe_out[k], stream=sk, trace_mode=2)

    for i in range(-pipe_warm, 0):
        pipe_step(i)
    for st in pipe_streams:
        mod.trace_sync(stream=st.cuda_stream)
    barrier()
    p0 = torch.cuda.Event(enable_timing=True)
    p_end = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    p0.record(pipe_streams[0])
    pipe_streams[1].wait_event(p0)
    for i in range(args.steps):
        pipe_step(i)
    for k in range(2):
        p_end[k].record(pipe_streams[k])  # ordered after the run's trace transfer (the run joins its trace stream)
    barrier()
    ms_e2e_pipe = max_over_ranks(max(p0.elapsed_time(p_end[0]), p0.elapsed_time(p_end[1])))
    k_last = (args.steps - 1) % 2
    pipe_blob = mod.trace_blob(stream=pipe_streams[k_last].cuda_stream) if not args.no_verify else None
    pipe_logits = pipe_out[k_last].numpy().copy()

    # (iv) as (iii) plus the complete serialized blob every step (repeated records filled by the parallel host copy)
    def e2e_blob_step(i):
        mod.run(xs_host[i % 2], out_host, stream=stream, trace_mode=2)
        mod.trace_blob_view()

    ms_e2e_blob, _, _ = timed(1, max(2, args.steps // 2), e2e_blob_step)
    ms_e2e_blob *= args.steps / max(2, args.steps // 2)
    trace_bytes = mod.trace_bytes()
    trace_d2h = mod.trace_d2h_bytes()

    # (v) sustained: >= sustained_s seconds of back-to-back `value` steps
    sus_steps = max(args.steps, int(args.sustained_s * 1e3 / max(ms_value / args.steps, 1e-3)))
    sus_steps = min(sus_steps, 200000)
    barrier()
    sampler.mark("sustained_begin")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(sus_steps):
        mod.run(x_dev, out_dev, stream=stream, trace_mode=1)
    e1.record()
    barrier()
    sampler.mark("sustained_end")
    ms_sus = max_over_ranks(e0.elapsed_time(e1))

    # (vi) PCIe probe: plain pinned-memory copies of the trace's size, all ranks at the same time
    probe_bytes = int(max(trace_d2h, 1 << 26))
    pd = torch.empty(probe_bytes, dtype=torch.uint8, device="cuda")
    ph = torch.empty(probe_bytes, dtype=torch.uint8).pin_memory()

    def copy_rate(dst, src):
        dst.copy_(src, non_blocking=True)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(4):
            dst.copy_(src, non_blocking=True)
        b.record()
        barrier()
        return world * 4 * probe_bytes / (max_over_ranks(a.elapsed_time(b)) * 1e-3) / 1e9  # all ranks together

    d2h_gbs, h2d_gbs = copy_rate(ph, pd), copy_rate(pd, ph)
    del pd, ph
    clocks, clocks_sus = sampler.stop() if rank == 0 else (None, None)

    # roofline of the dominant kernel class (tcgen05 int8 conv/dense): CUDA events around every launch, same stream
    mod.set_option("time_tc", "1")
    tc_ms, tc_n = 0.0, 0
    runs = max(3, args.steps // 4)
    for _ in range(runs):
        RT.spin(3000.0, stream)  # the host queues the whole step while the GPU spins: events see device time only
        mod.run(x_dev, out_dev, stream=stream, trace_mode=0)
        torch.cuda.synchronize()
        a, b = mod.tc_time()
        tc_ms += a
        tc_n += b
    mod.set_option("time_tc", "0")
    peak_tops = RT.probe_int8_peak(4000)
    tc_macs = mod.tc_macs()  # MACs of the tcgen05 launches only (depthwise / few-channel pointwise layers run on CUDA cores)
    achieved_tops = 2.0 * tc_macs * runs / (tc_ms * 1e-3) / 1e12 if tc_ms > 0 else 0.0

    # proof: what was timed is bit-exact (every rank checks its own shard; any failure fails the job)
    problems = []
    if not args.no_verify:
        interp_threads = os.cpu_count()
        from oracle import interp
        interp.set_threads(interp_threads)
        problems += ["first e2e step: " + p for p in
                     verify_against_oracle(args.model, args.batch, graph_json, consts, xs_np[0], first["logits"], first["blob"])]
        problems += ["last e2e step: " + p for p in
                     verify_against_oracle(args.model, args.batch, graph_json, consts, xs_np[1], last_logits, last_blob)]
        problems += ["last pipelined e2e step: " + p for p in
                     verify_against_oracle(args.model, args.batch, graph_json, consts, xs_np[k_last], pipe_logits, pipe_blob)]
        problems += ["device-resident trace: " + p for p in
                     verify_against_oracle(args.model, args.batch, graph_json, consts, xs_np[0], logits_dev, blob_dev)]
    bad = max_over_ranks(float(len(problems) > 0)) > 0

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        if problems:
            print("rank %d: %s" % (rank, problems[:4]), file=sys.stderr)
            sys.exit(3)
        return
    total_images = args.batch * world
    rate = lambda ms: total_images * args.steps / (ms * 1e-3)
    value = rate(ms_value)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    traffic, traffic_src = _ncu_traffic()
    e2e_serial = rate(ms_e2e)
    e2e_value = rate(ms_e2e_pipe)
    d2h_step = int(trace_d2h + np.prod(out_shape) * out_host.element_size())
    e2e_gbs = world * d2h_step * args.steps / (ms_e2e_pipe * 1e-3) / 1e9
    per_sorted = sorted(per_value)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_value / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8*s8->s32", "data": "synthetic",
        "bit_exact_trace": (not problems and not bad) if not args.no_verify else None,
        "config": {
            "workload": workload_string(args.model, args.batch),
            "value_mode": "inputs resident in HBM; every operator's inputs/outputs captured on the device in their logical "
                          "layout (trace mode 1), readable through tkb_trace_blob until the next run; the read-out itself is "
                          "config.trace_readout_ms and the streamed form is e2e",
            "compute_only": rate(ms_compute),
            "trace_to_host_e2e": e2e_value,
            "trace_to_host_e2e_one_batch_in_flight": e2e_serial,
            "e2e_mode": "two batches in flight on two caller streams (double-buffered workspaces and pinned blobs); each "
                        "trace is complete in pinned host memory before its stream is reused",
            "e2e_d2h_gbs_one_batch_in_flight": world * d2h_step * args.steps / (ms_e2e * 1e-3) / 1e9,
            "e2e_serialized_blob": rate(ms_e2e_blob),
            "trace_readout_ms": readout_ms,
            "sustained_value": total_images * sus_steps / (ms_sus * 1e-3), "sustained_steps": sus_steps,
            "sustained_s": ms_sus * 1e-3, "sustained_clocks": clocks_sus,
            "step_ms_min_median_p95": [per_sorted[0], per_sorted[len(per_sorted) // 2], per_sorted[int(0.95 * (len(per_sorted) - 1))]],
            "trace_bytes_per_step": trace_bytes, "trace_d2h_bytes_per_step": trace_d2h,
            "l2": "no explicit flush: each step streams its whole activation arena (> 126 MB L2) through HBM",
            "conv_impl": args.conv_impl, "int8_macs_per_step": macs, "int8_macs_per_step_tcgen05": tc_macs, "numa": numa,
            "verified": "first and last end-to-end step + the device-resident trace of the `value` leg vs the CPU oracle on "
                        "the first 4 and last 4 images of the batch: logits and every serialized trace record; also the last "
                        "step of the two-batches-in-flight e2e leg"
                        if not args.no_verify else "skipped (--no-verify)",
        },
        "e2e": {"value": e2e_value, "unit": UNIT,
                "h2d_bytes_per_step": int(xs_np[0].nbytes), "d2h_bytes_per_step": d2h_step,
                "d2h_gbs": e2e_gbs, "pcie_peak_gbs": {"d2h": d2h_gbs, "h2d": h2d_gbs,
                                                      "how": "cudaMemcpyAsync of %d B pinned<->device, %d rank(s) at once, aggregate" % (probe_bytes, world)},
                "frac_of_pcie": e2e_gbs / d2h_gbs if d2h_gbs else None},
        "gpu_launches": int(launches),
        "roofline": {
            "bound": "tensor", "kernel": "tcgen05.mma kind::i8 conv/dense kernels (tc_halo_kernel / tc_gemm_kernel); all %d launches of a step" % (tc_n // runs if runs else 0),
            "achieved": achieved_tops,
            "peak": peak_tops, "unit": "TFLOP/s", "frac": achieved_tops / peak_tops if peak_tops else None, "traffic": traffic,
            "note": "int8 ops (2*MAC) counted as FLOPs; achieved = 2 * MACs of the step's tcgen05 launches / their summed CUDA-event time "
                    "(events on the launching stream around every run of consecutive tcgen05 launches -- layers that follow one another "
                    "with no other kernel in between share one pair, so event cost and the loss of launch overlap an event between two "
                    "kernels causes are not billed to them; TKB_TC_TIME_PER_LAUNCH=1 restores a pair per launch; the host enqueues "
                    "behind a device spin); peak = this run's "
                    "tcgen05 kind::i8 issue-rate probe (MEASURED_PEAKS.json has no int8 entry; 2x its bf16 burst = %.0f); traffic = DRAM "
                    "bytes per launch from profiles/%s" % (2 * peaks.get("bf16_tflops", 0), traffic_src),
            "avg_launch_us": 1e3 * tc_ms / tc_n if tc_n else None, "launches_per_step": tc_n // runs if runs else None,
            "frac_vs_2x_bf16_burst": achieved_tops / (2 * peaks["bf16_tflops"]) if peaks.get("bf16_tflops") else None,
        },
        "clocks": clocks,
    }
    if problems:
        line["config"]["verify_problems"] = problems[:8]
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count()
        sample = min(8, args.batch)
        r = cpu_oracle_rate(args.model, sample)
        line["cpu_baseline"] = {"value": r, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{sample} images of the same workload, scalar C restatement of the reference lowering "
                                          "(oracle/qnn_oracle.c + OpenMP; not oneDNN) with trace"}
    print(json.dumps(line), file=_JSON_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()
    if problems or bad:
        print("bit-exactness check FAILED: %s" % problems[:4], file=sys.stderr)
        sys.exit(3)


if __name__ == "__main__":
    main()
