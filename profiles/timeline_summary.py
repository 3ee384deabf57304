"""Summarise a PTTS_TIMELINE=1 dump (stderr of any run): per kernel, time from its PDL wait to the next kernel's
wait (its cost on the dependent chain) and, for the skinny tcgen05 GEMMs, the in-kernel phases of CTA 0."""
import re
import sys
from collections import defaultdict

rows = []
for line in open(sys.argv[1]):
    m = re.match(r"\[timeline\] (.+?)\s+entry\s+([-\d.]+)\s+wait\s+([-\d.]+)\s+exit\s+([-\d.]+) \|\s+([-\d.]+)\s+([-\d.]+)\s+([-\d.]+)\s+([-\d.]+)", line)
    if m:
        rows.append((m.group(1).strip(),) + tuple(float(x) for x in m.groups()[1:]))
agg = defaultdict(list)
for i, r in enumerate(rows[:-1]):
    name, entry, wait, exit_, ops, ln, acc, slabs = r
    nxt_wait = rows[i + 1][2]
    if wait < 0 or nxt_wait < 0:
        continue
    d = {"chain": nxt_wait - wait, "early": wait - entry}
    if exit_ > 0:
        d["cta0"] = exit_ - wait
    if name.startswith("gemm") and ops > 0 and acc > 0:
        d.update({"w->ops": ops - wait, "ops->acc": acc - ops, "acc->slabs": (slabs - acc) if slabs > 0 else 0.0,
                  "epilogue": exit_ - (slabs if slabs > 0 else acc), "ln_ready": (ln - wait)})
    agg[name].append(d)
keys = ["chain", "early", "cta0", "w->ops", "ops->acc", "acc->slabs", "epilogue", "ln_ready"]
print(f"{'kernel':24s} {'n':>3s} " + " ".join(f"{k:>10s}" for k in keys))
tot = 0.0
for name, ds in agg.items():
    line = f"{name:24s} {len(ds):3d} "
    for k in keys:
        v = [d[k] for d in ds if k in d]
        line += f" {sum(v) / len(v):10.2f}" if v else f" {'':>10s}"
    tot += sum(d["chain"] for d in ds)
    print(line)
print(f"sum of chain costs over the dump: {tot:.1f} us")
