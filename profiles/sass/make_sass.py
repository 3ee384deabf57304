"""SASS evidence from the built objects (run `python -m pocket_tts_b200.build --force` first):
  profiles/sass/summary.txt                      per kernel: instruction count + Blackwell-native mnemonics
  profiles/sass/<unit>.tensor_tma_lines.sass.txt every tcgen05 / TMEM / TMA / mbarrier / mma.sync line of the unit
UTC*MMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM / STTM = tcgen05.ld / st, UTMALDG / UBLKCP = TMA tensor / bulk
copies, SYNCS = mbarrier, HMMA = mma.sync (legacy tensor path), LDGSTS = cp.async.
    python profiles/sass/make_sass.py"""
import collections
import os
import re
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
BUILD = ROOT / "pocket_tts_b200" / "build"
OUT = ROOT / "profiles" / "sass"
UNITS = ["gemm_tc", "seanet_fused", "flow_head_ss", "flow_head_tc", "flow_head", "attention", "flow_mega", "kernels", "mimi_kernels",
         "engine_encoder", "gemm_simt", "static_scratch"]
LISTED = {"gemm_tc", "seanet_fused", "flow_head_tc", "flow_head_ss", "attention"}
KEYS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "HMMA", "LDGSTS", "MUFU", "FFMA", "LDG", "STG",
        "LDS", "STS", "BAR"]
PAT = re.compile(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z0-9_]+)")


def demangle(sym: str) -> str:
    s = subprocess.run(["c++filt", sym], capture_output=True, text=True).stdout.strip()
    s = s.replace("(anonymous namespace)::", "").replace("ptts::", "").replace("void ", "")
    depth, out = 0, []
    for ch in s:                      # cut the argument list: first '(' outside template brackets
        if ch == "<":
            depth += 1
        elif ch == ">":
            depth -= 1
        elif ch == "(" and depth == 0:
            break
        out.append(ch)
    return "".join(out)


def main():
    summary = [__doc__.strip(), ""]
    for unit in UNITS:
        obj = BUILD / f"{unit}.o"
        if not obj.exists():
            continue
        txt = subprocess.run(["cuobjdump", "-sass", str(obj)], capture_output=True, text=True).stdout
        cur, counts, listing = None, collections.OrderedDict(), []
        for line in txt.splitlines():
            m = re.match(r"\s+Function : (\S+)", line)
            if m:
                cur = demangle(m.group(1))
                counts[cur] = collections.Counter()
                listing.append(f"\n=== {cur}")
                continue
            m = PAT.match(line)
            if m and cur:
                op = m.group(1)
                counts[cur][op] += 1
                counts[cur]["_total"] += 1
                if op.startswith(("UTC", "LDTM", "STTM", "UTMA", "UBLKCP", "SYNCS", "HMMA")):
                    listing.append(re.sub(r"\s+/\*[0-9a-f]{16}\*/\s*$", "", line.rstrip()))
        summary.append(f"## {unit}.cu")
        for k, c in counts.items():
            agg = {key: sum(v for op, v in c.items() if op.startswith(key)) for key in KEYS}
            summary.append(f"{k[:120]}: {c['_total']} instr; " + ", ".join(f"{key} {v}" for key, v in agg.items() if v))
        summary.append("")
        if unit in LISTED:
            (OUT / f"{unit}.tensor_tma_lines.sass.txt").write_text(
                f"# every tcgen05 / TMEM / TMA / mbarrier / mma.sync instruction line of {unit}.cu (cuobjdump -sass), per kernel\n"
                + "\n".join(listing) + "\n")
    (OUT / "summary.txt").write_text("\n".join(summary) + "\n")


if __name__ == "__main__":
    main()
