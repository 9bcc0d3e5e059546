"""Print the SASS of one kernel of libswarm_b200.so (substring match on the mangled name) and a histogram of its
opcodes:  python tools/sass_fn.py k_bd2d_tiledILb1 [--hist] [--grep REGEX]"""
import re, subprocess, sys, collections, pathlib
lib = pathlib.Path(__file__).resolve().parent.parent / "swarmrl_b200" / "libswarm_b200.so"
pat = sys.argv[1]
txt = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
blocks = re.split(r"\n\s*Function : ", txt)
sel = [b for b in blocks[1:] if pat in b.split("\n", 1)[0]]
assert sel, "no such function"
body = sel[0]
lines = [l for l in body.split("\n") if re.search(r"/\*[0-9a-f]{4,}\*/\s+\S", l)]
ins = []
for l in lines:
    m = re.search(r"/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m: ins.append((int(m.group(1), 16), m.group(2).strip()))
if "--hist" in sys.argv:
    c = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", t).split()[0].split(".")[0] for _, t in ins)
    print(len(ins), "instructions")
    for k, v in c.most_common(40): print(f"{k:12s} {v}")
elif "--grep" in sys.argv:
    rx = re.compile(sys.argv[sys.argv.index("--grep") + 1])
    for k, (a, t) in enumerate(ins):
        if rx.search(t): print(k, hex(a), t)
else:
    for k, (a, t) in enumerate(ins): print(k, hex(a), t)
