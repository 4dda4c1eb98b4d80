"""Instruction mix of the shipped library per kernel, from `cuobjdump -sass` (no GPU needed):
python scripts/sass_counts.py > profiles/r02_sass_counts.md"""
import collections, os, re, subprocess, sys

SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "sfm_b200", "libhsfm_ba.so")
WANT = ["DMMA", "DFMA", "DADD", "DMUL", "LDGSTS", "LDG", "STG", "LDS", "STS", "REDG", "ATOMG", "ATOMS", "MUFU", "BAR", "SHFL", "UTMALDG", "UBLKRED", "UTCMMA"]
txt = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
cur, counts, regs = None, collections.OrderedDict(), {}
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = cur.replace("(anonymous namespace)::", "").replace("hsfm::", "").replace("void ", "")
        cur = re.sub(r"\(.*", "", cur)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1).split(".")[0]
        counts[cur]["total"] += 1
        for w in WANT:
            if op == w or (w in ("REDG", "LDG", "STG") and op == w):
                counts[cur][w] += 1
print("# SASS instruction mix per kernel of `sfm_b200/libhsfm_ba.so` (sm_100a; static counts, `cuobjdump -sass`)\n")
print("FP64 tensor work is `DMMA` (tcgen05 has no FP64 kind); asynchronous staging is `LDGSTS` (cp.async); atomics are `REDG`/`ATOMG`.\n")
print("| kernel | SASS instr | " + " | ".join(WANT) + " |")
print("|---|---:|" + "---:|" * len(WANT))
for k, c in counts.items():
    if c["total"] < 200 or k.startswith("cub::") or "cub::" in k:
        continue
    print(f"| `{k[:70]}` | {c['total']} | " + " | ".join(str(c[w]) if c[w] else "" for w in WANT) + " |")
