"""profiles/r02_sass_excerpt.txt: mnemonic counts of every kernel in the built library (cuobjdump -sass) and the SASS of
one PCG iteration of the second half of kernel 5.  CPU only (needs the in-tree libgato_b200.so and cuobjdump)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "gato_python_b200", "libgato_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
ins = re.compile(r"/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)")
keep = ["FFMA", "FFMA2", "DFMA", "DMUL", "DADD", "SHFL", "LDS", "STS", "LDG", "STG", "BAR", "UBLKCP", "SYNCS", "STAS", "UCGABAR_ARV",
        "UCGABAR_WAIT", "MUFU", "LDL", "STL", "ATOMG", "RED"]
out = ["SASS of gato_python_b200/libgato_b200.so (cuobjdump -sass, sm_100a), round 2 final: mnemonic counts per kernel "
       "(scripts/sass_excerpt.py)", ""]
loop = None
for f in txt.split("Function : ")[1:]:
    name = f.split("\n", 1)[0].strip()
    if "ILi14ELi7E" not in name:
        continue
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip().split("(gato::")[0]
    dem = dem.replace("void gato::", "").replace("(int)", "").replace("(bool)", "b")
    lines = [l for l in f.split("\n") if ins.search(l)]
    ops = [ins.search(l).group(1) for l in lines]
    c = collections.Counter(ops)
    out.append(f"{dem:<62} total {len(ops):6d}  " + "  ".join(f"{k} {c[k]}" for k in keep if c[k]))
    if dem.startswith("gato_pcg_scaled<14, 7, 128, false>"):
        bars = [i for i, o in enumerate(ops) if o == "BAR"]
        a, b = bars[4], bars[8]      # one iteration: from the barrier that ends the set-up to the fourth barrier of the loop
        cc = collections.Counter(ops[a:b])
        loop = [f"", f"PCG loop of {dem} (between its 5th and 9th BAR.SYNC: one iteration, {b - a} instructions):",
                "  " + "  ".join(f"{k} {v}" for k, v in cc.most_common(20)), ""] + [l.split("/* 0x")[0].rstrip() for l in lines[a:b + 1]]
out += loop or []
open(os.path.join(ROOT, "profiles", "r02_sass_excerpt.txt"), "w").write("\n".join(out) + "\n")
print("\n".join(out[:24]))
