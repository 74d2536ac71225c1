"""Per-layer tcgen05 times of one model step: python tools/tc_layers.py launches.csv   (csv = ncu launch list of
`TKB_MODEL=... python tools/profile_step.py tc 0 N`; prints the plan's tcgen05 lines beside the last step's launches)"""
import csv, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tachikoma_b200 import models, runtime as RT, tachikoma as T
name = os.environ.get("TKB_MODEL", "resnet18"); batch = int(os.environ.get("TKB_BATCH", "64"))
part = T.partition_for_tachikoma(models.MODELS[name](batch=batch))
sym, js, names, consts = T.serialize(part)
mod = RT.TachikomaJSONRuntime(sym, js, names); mod.set_option("conv_impl", "tc"); mod.init(consts)
plan = [l.strip() for l in mod.plan_summary().splitlines() if "tcgen05" in l]
rows = list(csv.DictReader([l for l in open(sys.argv[1]) if not l.startswith("==")]))
tc = [r for r in rows if "tc_" in r["Kernel Name"]]
per = len(plan)
last = tc[-per:]
tot = 0.0
for p, r in zip(plan, last):
    t = float(r["Metric Value"].replace(",", "")) / (1000.0 if r["Metric Unit"] in ("ns", "nsecond") else 1.0)
    tot += t
    k = r["Kernel Name"]; k = k[k.find("tc_"):k.find(">") + 1]
    print("%7.1f us  %-34s %-14s %s" % (t, k, r["Grid Size"], p[p.find("->") + 3:][:95]))
print("tcgen05 total %.1f us in %d launches" % (tot, per))
