"""Dump the SASS of one kernel of libb2gs.so (substring of the mangled name) as `offset instruction` lines.
usage: python tools/sass_fn.py <name-substring> <out.txt>   Dev tool (cuobjdump, no GPU needed)."""
import os, subprocess, sys, re
pat, out = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump","-sass",os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "monogs_dev_b200", "libb2gs.so")],capture_output=True,text=True).stdout
on=False; lines=[]
for l in txt.splitlines():
    if "Function :" in l: on = pat in l
    elif on:
        m=re.match(r"\s+/\*([0-9a-f]{4})\*/\s+(.*?)\s*/\*", l)
        if m: lines.append(m.group(1)+" "+m.group(2))
open(out,"w").write("\n".join(lines)+"\n"); print(len(lines))
