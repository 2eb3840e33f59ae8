"""Hottest source lines / SASS instructions of one result of an .ncu-rep, by warp-stall samples.

    python scripts/ncu_source_hot.py REPORT RESULT_INDEX [N]

Uses the ncu_report module that ships with Nsight Compute.  Prints the N (default 25) SASS instructions with the
most sampled stalls, their source line (needs -lineinfo and --import-source on) and the stall reasons.
"""
import collections
import glob
import sys

sys.path.insert(0, glob.glob("/opt/nvidia/nsight-compute/*/extras/python")[0])
import ncu_report  # noqa: E402

rep, idx = sys.argv[1], int(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
ctx = ncu_report.load_report(rep)
action = ctx.range_by_idx(0).action_by_idx(idx)
print("kernel:", action.name()[:120])
m = action.metric_by_name("smsp__pcsamp_sample_buffer_full")
names = [n for n in action.metric_names() if n.startswith("smsp__pcsamp_warps_issue_stalled_") and "not_issued" not in n]
per_pc = collections.defaultdict(lambda: collections.Counter())
total = collections.Counter()
for n in names:
    met = action.metric_by_name(n)
    if not met.has_correlation_ids():
        continue
    ids = met.correlation_ids()
    for k in range(met.num_instances()):
        v = met.as_uint64(k)
        if v:
            pc = ids.as_uint64(k)
            reason = n.replace("smsp__pcsamp_warps_issue_stalled_", "")
            per_pc[pc][reason] += v
            total[reason] += v
all_samples = sum(total.values())
print("samples:", all_samples, " by reason:", ", ".join(f"{r} {100 * v / all_samples:.0f}%" for r, v in total.most_common(8)))
sass = {}
try:
    for pc in per_pc:
        sass[pc] = action.sass_by_pc(pc)
except Exception as e:  # older interface
    print("no sass lookup:", e)
src = {}
try:
    for pc in per_pc:
        info = action.source_info(pc)
        if info is not None:
            src[pc] = f"{info.file_name().split('/')[-1]}:{info.line()}"
except Exception as e:
    print("no source lookup:", e)
by_line = collections.Counter()
for pc, c in per_pc.items():
    by_line[src.get(pc, "?")] += sum(c.values())
print("\nhottest source lines:")
for line, v in by_line.most_common(12):
    print(f"  {100 * v / all_samples:5.1f}%  {line}")
print("\nhottest instructions:")
for pc, c in sorted(per_pc.items(), key=lambda kv: -sum(kv[1].values()))[:top]:
    v = sum(c.values())
    why = ", ".join(f"{r} {n}" for r, n in c.most_common(3))
    print(f"  {100 * v / all_samples:5.1f}%  {src.get(pc, '?'):28s} {sass.get(pc, '')[:70]:70s} [{why}]")
