"""Instruction mix of the main loop (largest backward branch target span) of the tile kernels in a built library.
  python scripts/sass_mix.py [lib.so]      prints, per tile kernel with G = 8, the loop's size and opcode histogram"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "margin_b200/csrc/libmargin_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], stdout=subprocess.PIPE, text=True).stdout
funcs = re.split(r"\n\s*Function : ", out)[1:]
for f in funcs:
    name = f.split("\n", 1)[0].strip()
    if "tile_kernel" not in name or "ILi8E" not in name:
        continue
    ins = re.findall(r"/\*([0-9a-f]{4,5})\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)[^;]*?(?:(0x[0-9a-f]+)\s*)?;", f)
    best = None
    for addr, op, tgt in ins:
        if op.startswith("BRA") and tgt:
            a, t = int(addr, 16), int(tgt, 16)
            if t < a and (best is None or a - t > best[1] - best[0]):
                best = (t, a)
    if not best:
        continue
    body = [op.split(".")[0] for addr, op, tgt in ins if best[0] <= int(addr, 16) <= best[1]]
    mix = collections.Counter(body).most_common()
    print("%s\n  loop 0x%x..0x%x: %d instructions: %s" % (name[:60], best[0], best[1], len(body), ", ".join("%s %d" % m for m in mix[:16])))
