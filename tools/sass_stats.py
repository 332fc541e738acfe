"""Static SASS statistics of one kernel in libb200cs.so (dev helper): instruction count by opcode,
and the size of each basic block (split at branch targets / branches)."""
import collections, re, subprocess, sys
so = sys.argv[1]; pat = sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", txt)
for f in funcs[1:]:
    name = f.split("\n", 1)[0]
    if not re.search(pat, name):
        continue
    ins = re.findall(r"/\*([0-9a-f]{4,})\*/\s+(.*?);", f)
    ops = collections.Counter()
    for addr, s in ins:
        m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", s.strip())
        ops[m.group(2).split(".")[0]] += 1
    print(name[:100], "total", len(ins))
    print("  ", dict(ops.most_common(25)))
