#!/usr/bin/env python
"""Which kernels changed between two builds?  python scripts/sass_diff.py old.so new.so  (cuobjdump -sass of both, instruction text compared per kernel).
Used at the end of round 2, when no GPU minutes were left, to show that the commits made after the last GPU-tested one (7a09e94) leave every
kernel but k_rescue bit-identical: profiles/r2z_sass_diff.txt."""
import collections, re, subprocess, sys
def load(p):
    d = collections.defaultdict(list); fn = None
    out = subprocess.run(["cuobjdump", "-sass", p], capture_output=True, text=True, errors="ignore").stdout
    for l in out.splitlines():
        m = re.search(r"Function : (\S+)", l)
        if m: fn = m.group(1); continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", l)
        if m and fn: d[fn].append(m.group(1).strip())
    return d
a = load(sys.argv[1]); b = load(sys.argv[2])
same = [k for k in a if k in b and a[k] == b[k]]
diff = [k for k in a if k in b and a[k] != b[k]]
print("kernels compared:", len([k for k in a if k in b]), " identical SASS:", len(same), " changed:", len(diff))
for k in diff: print("  changed:", k[:100], len(a[k]), "->", len(b[k]), "instructions")
print("only in old:", [k[:70] for k in a if k not in b][:10])
