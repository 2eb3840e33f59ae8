"""Regenerate profiles/sass_mnemonics.txt and profiles/ptxas_resources.txt from the current sources.

    python scripts/sass_evidence.py

CPU-only (nvcc / cuobjdump cross-compile for sm_100a).  The first file proves which instructions the
kernels contain (tcgen05.mma = UTCHMMA, tcgen05.ld = LDTM, TMA = UTMALDG, multimem.ld_reduce = LDGMC,
system-scope release/acquire = *.STRONG.SYS); the second records registers / spills / stack per kernel.
"""
import collections
import datetime
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cs744_pytorch_distributed_tutorial_b200", "lib", "libb200dp.so")
SRCS = ["csrc/conv_igemm.cu", "csrc/collectives.cu", "csrc/elementwise.cu", "csrc/bn_stream.cu"]
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "--expt-relaxed-constexpr",
         "-I", os.path.join(ROOT, "include"), "-Xcompiler", "-fPIC"]
EXPERIMENTAL = ("wgrad_push", "reduce_staged")      # the optional weight-gradient push path (B200DP_WGRAD_PUSH=1)
today = datetime.date.today().isoformat()


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def sass():
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    total, per, name = collections.Counter(), collections.OrderedDict(), None
    interesting = re.compile(r"UTC|UTMA|UBLKCP|LDTM|STTM|SYNCS|REDG|ATOMG|LDGMC|STGMC|MULTIMEM|MMEM|STRONG\.SYS|MEMBAR|UCGABAR|CCTL")
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            per[name] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Za-z0-9_.]+)", line)
        if m and name and interesting.search(m.group(1)):
            total[m.group(1)] += 1
            per[name][m.group(1)] += 1
    return total, per


def main():
    total, per = sass()
    names = demangle(list(per))
    with open(os.path.join(ROOT, "profiles", "sass_mnemonics.txt"), "w") as f:
        f.write(f"# SASS mnemonic counts of lib/libb200dp.so (cuobjdump -sass, sm_100a), {today}; scripts/sass_evidence.py\n")
        f.write("# tcgen05.mma -> UTCHMMA (.2CTA = cta_group::2), tcgen05.ld -> LDTM, tcgen05.commit -> UTCBAR, TMA -> UTMALDG,\n"
                "# mbarrier -> SYNCS.*, red.global.add.v4.f32 -> REDG...F32x4, multimem.ld_reduce -> LDGMC.E.ADD.F32x4 (multimem.st shows as STG.E.128.STRONG.SYS),\n"
                "# st.release.sys / ld.acquire.sys -> ST/LD ...STRONG.SYS, cluster barrier -> UCGABAR_*, cp.async.bulk -> UBLKCP\n")
        for k, v in total.most_common():
            f.write(f"{v:7d} {k}\n")
        f.write("\n# per kernel ([optional] = only used with B200DP_WGRAD_PUSH=1)\n")
        for n, c in sorted(per.items(), key=lambda kv: names[kv[0]]):
            def cnt(pat):
                return sum(v for k, v in c.items() if re.search(pat, k))
            tags = [("UTCHMMA", cnt(r"^UTCHMMA")), ("UTMALDG", cnt(r"^UTMALDG")), ("LDTM", cnt(r"^LDTM")),
                    ("UTCBAR", cnt(r"^UTCBAR")), ("UBLKCP(cp.async.bulk)", cnt(r"^UBLKCP")), ("REDG.F32x4", cnt(r"REDG.*F32x4")), ("multimem.ld_reduce(LDGMC)", cnt(r"LDGMC|STGMC|MULTIMEM|MMEM")),
                    ("STRONG.SYS", cnt(r"STRONG\.SYS"))]
            tags = " ".join(f"{k}={v}" for k, v in tags if v)
            short = names[n].replace("(anonymous namespace)::", "").replace("void ", "")
            short = re.sub(r"\(.*", "", short)
            exp = "[optional] " if any(e in short for e in EXPERIMENTAL) else ""
            f.write(f"{exp}{short}  {tags or '-'}\n")
    with open(os.path.join(ROOT, "profiles", "ptxas_resources.txt"), "w") as f:
        f.write(f"# ptxas -v resource usage (sm_100a), {today}; scripts/sass_evidence.py\n")
        for src in SRCS:
            r = subprocess.run(["nvcc", *FLAGS, "-Xptxas", "-v", "-c", os.path.join(ROOT, src), "-o", "/dev/null"],
                               capture_output=True, text=True)
            f.write(f"## {src}\n")
            cur, spill = None, ""
            for line in r.stderr.splitlines():
                m = re.search(r"Compiling entry function '(\S+)'", line)
                if m:
                    cur = m.group(1)
                m = re.search(r"(\d+) bytes spill stores, (\d+) bytes spill loads", line)
                if m:
                    spill = "" if m.group(1) == "0" and m.group(2) == "0" else f"  SPILL stores={m.group(1)} loads={m.group(2)}"
                m = re.search(r"Used (\d+) registers(.*)", line)
                if m and cur:
                    short = demangle([cur])[cur].replace("(anonymous namespace)::", "").replace("void ", "")
                    short = re.sub(r"\(.*", "", short)
                    exp = "[optional] " if any(e in short for e in EXPERIMENTAL) else ""
                    f.write(f"{exp}{short:66s} regs={m.group(1)}{m.group(2).rstrip()}{spill}\n")
                    cur = None
    print("wrote profiles/sass_mnemonics.txt and profiles/ptxas_resources.txt")


if __name__ == "__main__":
    main()
