"""Compare the SASS of two builds function by function.

    python scripts/sass_diff.py /tmp/sass_base/conv_igemm.sass <(cuobjdump -sass build/conv_igemm.cu.o)

Used when a change is meant to leave existing kernels untouched (new template instantiations, new
kernels next to them): every function present in both dumps must have identical instructions.
"""
import re
import sys


def functions(path):
    out, name, body = {}, None, []
    for line in open(path):
        m = re.search(r"Function : (\S+)", line)
        if m:
            if name:
                out[name] = body
            name, body = m.group(1), []
        elif name and "/*" in line:
            body.append(" ".join(line.split()))   # column padding depends on the longest line of the dump
    if name:
        out[name] = body
    return out


a, b = functions(sys.argv[1]), functions(sys.argv[2])
changed = [n for n in a if n in b and a[n] != b[n]]
print(f"{len(a)} functions before, {len(b)} after; removed {sorted(set(a) - set(b))}; "
      f"added {len(set(b) - set(a))}; changed {len(changed)}")
for n in changed:
    print("  CHANGED", n)
sys.exit(1 if changed or set(a) - set(b) else 0)
