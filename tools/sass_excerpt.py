"""SASS evidence that the two contraction kernels are tcgen05 / TMEM / TMA code (B200_PROFILING.md mnemonics): counts and the first lines of
UTCHMMA (tcgen05.mma), LDTM (tcgen05.ld), UTCBAR (tcgen05.commit), UBLKCP (cp.async.bulk), UTMALDG (cp.async.bulk.tensor), SYNCS (mbarrier)
per kernel of libb2r.so.   python tools/sass_excerpt.py > profiles/r02_sass_excerpt.md"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "poseestimator_b200", "csrc", "libb2r.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
kernels = {}
cur = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); kernels[cur] = []
    elif cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
        kernels[cur].append(re.sub(r"/\*[0-9a-fx]+\*/", "", line).strip())
MN = ["UTCHMMA", "LDTM", "UTCBAR", "UBLKCP", "UTMALDG", "SYNCS", "FFMA2", "VIMNMX3", "FMNMX3"]
print("# r02 SASS excerpt of libb2r.so (`cuobjdump -sass`, sm_100a)\n")
print("| kernel | instructions | " + " | ".join(MN) + " |\n|---|---|" + "---|" * len(MN))
keep = {}
for name, ins in kernels.items():
    counts = [sum(1 for i in ins if mn in i) for mn in MN]
    if counts[0] or counts[1] or counts[3] or counts[4]:
        short = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
        short = short.replace("(anonymous namespace)::", "").replace("b2r::", "").split("(")[0]
        print(f"| `{short}` | {len(ins)} | " + " | ".join(str(c) for c in counts) + " |")
        keep[short] = ins
print()
for short, ins in keep.items():
    print(f"## `{short}`\n\n```")
    shown = 0
    for mn in ("UTCHMMA", "LDTM", "UTCBAR", "UBLKCP", "UTMALDG"):
        for i in [x for x in ins if mn in x][:3]:
            print(i); shown += 1
    print("```\n")
