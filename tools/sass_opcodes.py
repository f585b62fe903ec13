"""SASS opcode histogram per kernel of the product library (no GPU needed): which memory / async-copy / tensor opcodes each kernel uses.
    python tools/sass_opcodes.py [lib.so] > profiles/r2_sass_opcodes.txt"""
import collections, re, subprocess, sys
lib = sys.argv[1] if len(sys.argv) > 1 else "lz4-frame-conduit_b200/libb200lz4.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", out)), capture_output=True, text=True).stdout.splitlines()
hist, cur, order = {}, None, []
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = re.sub(r"\(.*", "", names[len(order)]); order.append(cur); hist[cur] = collections.Counter(); continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur:
        hist[cur][m.group(1)] += 1
MEM = ["LDG", "STG", "LD", "ST", "LDS", "STS", "ATOMS", "ATOMG", "ATOM", "RED", "LDL", "STL", "LDGSTS", "UBLKCP", "UTMALDG", "UTMASTG", "HMMA", "UTCMMA", "SHFL", "VOTE", "MEMBAR", "NANOSLEEP", "BAR"]
print(f"# cuobjdump -sass {lib}: static opcode counts per kernel, sm_100a.  LD/ST = generic, LDL/STL = local memory (spills, stack);")
print("# no LDGSTS / UBLKCP / UTMA* / *MMA anywhere: every access on this path is data dependent and there is no GEMM shape (DESIGN.md §3)")
for k in order:
    h = hist[k]
    print(f"\n{k}: {sum(h.values())} instructions")
    print("   memory/sync: " + "  ".join(f"{o}={h[o]}" for o in MEM if h[o]))
    print("   top: " + "  ".join(f"{o}={n}" for o, n in h.most_common(10)))
