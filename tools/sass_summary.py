"""SASS evidence for profiles/: per kernel of librsb200.so, the instruction count and the mnemonics that show the
sm_100a features in use (UBLKCP = 1-D bulk async copy / TMA engine, SYNCS = mbarrier, FFMA2 / FADD2 / FMUL2 = packed
FP32x2, MUFU.EX2, cluster st.async / mapa, ...).  python tools/sass_summary.py > profiles/r02_sass_summary.txt"""
import collections, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "radiosim_b200", "librsb200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
kern, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = re.sub(r"\(.*", "", kern)
        counts[kern] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        counts[kern][m.group(1)] += 1
WATCH = ["UBLKCP", "SYNCS", "MUFU.EX2", "FFMA2", "FADD2", "FMUL2", "ST.ASYNC", "MAPA", "UCGABAR", "BAR.SYNC", "LDS.128",
         "STS.64", "STG.E.128", "LDGSTS", "DFMA", "ATOMS", "ATOMG", "RED", "REDG", "CCTL", "ERRBAR", "FENCE.VIEW.ASYNC"]
print(f"# {os.path.basename(so)}: cuobjdump -sass, sm_100a")
for k, c in counts.items():
    tot = sum(c.values())
    hit = {w: sum(v for m, v in c.items() if m.startswith(w)) for w in WATCH}
    print(f"{k}: {tot} instructions; " + ", ".join(f"{w} {n}" for w, n in hit.items() if n))
