"""Write one SASS listing per kernel of libb200dp.so under profiles/sass/ (cuobjdump -sass, sm_100a).

    python scripts/dump_sass.py            # CPU-only; run after a build

Each file is the disassembly of ONE kernel instantiation (address / encoding columns stripped so that a diff
between builds shows instruction changes only), preceded by a header with the demangled name, the resource line
of `cuobjdump -res-usage` and the count of the instructions that prove what the kernel is made of
(UTCHMMA / UTCQMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA tile load, UBLKCP = cp.async.bulk,
LDGMC / multimem = in-switch reduction, REDG = red.global, SYNCS = mbarrier).  profiles/sass/INDEX.md lists them.
"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cs744_pytorch_distributed_tutorial_b200", "lib", "libb200dp.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEY = re.compile(r"^(UTC\w*MMA[\w.]*|UTCBAR[\w.]*|LDTM[\w.]*|STTM[\w.]*|UTMALDG[\w.]*|UTMASTG[\w.]*|UBLKCP[\w.]*|LDGMC[\w.]*|"
                 r"REDG[\w.]*|SYNCS[\w.]*|UCGABAR[\w.]*|HMMA[\w.]*)")


def short_name(demangled: str) -> str:
    n = re.sub(r"^void\s+", "", demangled)
    n = n.replace("b200dp::", "").replace("(anonymous namespace)::", "")
    n = re.sub(r"\(int\)|\(bool\)", "", n)
    n = re.sub(r"\(.*$", "", n)                       # drop the parameter list
    n = re.sub(r"[<>,: ]+", "_", n).strip("_")
    return n


def main() -> None:
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    usage, cur = {}, None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            cur = m.group(1)
        elif cur and "REG:" in line:
            usage[cur] = line.strip()
            cur = None
    funcs, name = collections.OrderedDict(), None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            funcs[name] = []
            continue
        if name is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(.*?)\s*/\* 0x[0-9a-f]+ \*/", line)
        if m:
            funcs[name].append(m.group(1).rstrip(" ;") + " ;")
    names = list(funcs)
    dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    os.makedirs(OUT, exist_ok=True)
    for f in os.listdir(OUT):
        if f.endswith(".sass") or f == "INDEX.md":
            os.remove(os.path.join(OUT, f))
    index = []
    for mangled, demangled in zip(names, dem):
        body = funcs[mangled]
        counts = collections.Counter()
        for ins in body:
            op = re.sub(r"^@!?U?P\d+\s+", "", ins).split()[0] if ins.split() else ""
            m = KEY.match(op)
            if m:
                counts[m.group(1)] += 1
        fn = short_name(demangled) + ".sass"
        with open(os.path.join(OUT, fn), "w") as f:
            f.write(f"// {demangled}\n// {usage.get(mangled, '')}\n")
            f.write("// " + ", ".join(f"{k} x{v}" for k, v in sorted(counts.items())) + "\n")
            f.write("\n".join(body) + "\n")
        index.append((fn, len(body), counts))
    with open(os.path.join(OUT, "INDEX.md"), "w") as f:
        f.write("# SASS listings of libb200dp.so (one file per kernel instantiation; scripts/dump_sass.py)\n\n")
        f.write("| listing | instructions | tcgen05.mma (UTC*MMA) | tcgen05.ld (LDTM) | TMA (UTMALDG) | cp.async.bulk (UBLKCP) | "
                "multimem (LDGMC) | red.global (REDG) | mbarrier (SYNCS) |\n|---|---:|---:|---:|---:|---:|---:|---:|---:|\n")
        for fn, n, c in sorted(index):
            s = lambda pat: sum(v for k, v in c.items() if re.match(pat, k))
            f.write(f"| `{fn}` | {n} | {s(r'UTC\w*MMA')} | {s('LDTM')} | {s('UTMALDG')} | {s('UBLKCP')} | {s('LDGMC')} | "
                    f"{s('REDG')} | {s('SYNCS')} |\n")
    print(f"{len(index)} listings written to {OUT}")


if __name__ == "__main__":
    main()
