"""Static SASS view of one kernel (dev helper, no GPU needed): basic blocks with their opcode mix,
the pipe each opcode issues on (alu = half rate on sm_100: FSEL/FSETP/FMNMX/LOP3/IADD3/...), and
the backward branches that close loops.

    python tools/sass_blocks.py <object or .so> <kernel-name regex> [--min N] [--from ADDR --to ADDR]
"""
import collections
import re
import subprocess
import sys

ALU = {"FSEL", "FSETP", "FMNMX", "FMNMX3", "LOP3", "IADD3", "ISETP", "LEA", "SHF", "SEL", "MOV", "PRMT", "VIADD",
       "PLOP3", "IABS", "IMNMX", "VIMNMX", "VIMNMX3", "FCHK", "P2R", "R2P", "CS2R", "I2FP", "F2FP", "FSET", "ISET", "LOP", "BMSK", "SGXT", "FLO", "POPC"}
FMA = {"FFMA", "FMUL", "FADD", "FFMA2", "FMUL2", "FADD2", "IMAD", "HFMA2", "HADD2", "HMUL2", "FFMA32I", "FMUL32I", "FADD32I"}
LSU = {"LDS", "STS", "LDG", "STG", "LDL", "STL", "LDGSTS", "ATOMS", "ATOMG", "RED", "LDSM", "SYNCS", "UTMALDG", "UBLKCP", "LDC", "LDCU"}


def pipe(op):
    if op in ALU:
        return "alu"
    if op in FMA:
        return "fma"
    if op in LSU:
        return "lsu"
    if op in ("MUFU",):
        return "xu"
    if op in ("BRA", "BSSY", "BSYNC", "EXIT", "WARPSYNC", "BAR", "CALL", "RET", "NANOSLEEP", "YIELD", "BREAK"):
        return "ctl"
    if op in ("SHFL", "VOTE", "MATCH", "REDUX", "CREDUX", "VOTEU"):
        return "warp"
    if op.startswith("U") or op.startswith("R2UR") or op == "S2UR":
        return "uni"
    return "oth"


def main():
    so, pat = sys.argv[1], sys.argv[2]
    args = sys.argv[3:]
    minn = int(args[args.index("--min") + 1]) if "--min" in args else 1
    lo = int(args[args.index("--from") + 1], 16) if "--from" in args else 0
    hi = int(args[args.index("--to") + 1], 16) if "--to" in args else 1 << 30
    txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    for f in re.split(r"\n\s*Function : ", txt)[1:]:
        name = f.split("\n", 1)[0]
        if not re.search(pat, name):
            continue
        ins = []
        for m in re.finditer(r"/\*([0-9a-f]{4,})\*/\s+(.*?);", f):
            addr, s = int(m.group(1), 16), m.group(2).strip()
            mm = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", s)
            op = mm.group(2).split(".")[0]
            tgt = None
            if op in ("BRA", "BSSY", "CALL"):
                t = re.search(r"0x([0-9a-f]+)\s*$", s)
                if t:
                    tgt = int(t.group(1), 16)
            ins.append((addr, op, s, tgt))
        leaders = {ins[0][0]}
        for i, (addr, op, s, tgt) in enumerate(ins):
            if op in ("BRA", "EXIT", "RET", "BSYNC", "CALL", "WARPSYNC") and i + 1 < len(ins):
                leaders.add(ins[i + 1][0])
            if tgt is not None and op != "BSSY":
                leaders.add(tgt)
        print(name[:120], "total", len(ins))
        blocks, cur = [], []
        for it in ins:
            if it[0] in leaders and cur:
                blocks.append(cur)
                cur = []
            cur.append(it)
        if cur:
            blocks.append(cur)
        tot = collections.Counter()
        for b in blocks:
            if len(b) < minn or b[0][0] < lo or b[0][0] > hi:
                continue
            ops = collections.Counter(o for _, o, _, _ in b)
            pipes = collections.Counter(pipe(o) for _, o, _, _ in b)
            tot.update(pipes)
            back = [f"<-{t:04x}" for a, o, s, t in b if o == "BRA" and t is not None and t <= a]
            fwd = [f"->{t:04x}" for a, o, s, t in b if o == "BRA" and t is not None and t > a]
            print(f"  {b[0][0]:05x} n={len(b):4d} {dict(pipes)} {' '.join(back)} {' '.join(fwd)}")
            print(f"        {dict(ops.most_common(14))}")
        print("  pipes in range:", dict(tot))


if __name__ == "__main__":
    main()
