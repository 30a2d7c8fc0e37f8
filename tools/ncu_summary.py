"""Summarise an .ncu-rep (raw + source pages) here on the CPU box: key metrics and top stall sites."""
import collections, csv, re, subprocess, sys


def launches(path):
    """Per-kernel launch counts / average duration from an `ncu --metrics gpu__time_duration.sum --csv` log."""
    rows = [r for r in csv.reader(open(path, errors="replace")) if len(r) > 10]
    hdr = rows[0]
    ki, vi, gi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            agg.setdefault((re.sub(r"\(.*", "", r[ki])[:70], r[gi]), []).append(float(r[vi].replace(",", "")))
        except ValueError:
            pass
    total = sum(sum(v) for v in agg.values())
    for (name, grid), v in agg.items():
        print("%-70s grid %-16s n=%-4d avg %9.1f us  share %5.1f %%" % (name, grid, len(v), sum(v) / len(v) / 1e3,
                                                                      100 * sum(v) / max(total, 1)))


if sys.argv[1] == "--launches":
    launches(sys.argv[2])
    sys.exit(0)
rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
d = dict(zip(hdr, zip(units, vals)))
keys = ['Kernel Name', 'gpu__time_duration.sum', 'sm__cycles_elapsed.avg.per_second',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct', 'smsp__inst_executed.sum']
for k in keys:
    if k in d:
        print("| %s | %s %s |" % (k, d[k][1], d[k][0]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
ix = {h: i for i, h in enumerate(rows[hi])}
data = rows[hi + 1:]
tot = sum(int(r[ix['# Samples']] or 0) for r in data)
print("total samples", tot)
names = [h for h in rows[hi] if h.startswith('stall_') and 'Not Issued' not in h]
for r in sorted(data, key=lambda r: -int(r[ix['# Samples']] or 0))[:topn]:
    st = {k[6:]: int(r[ix[k]] or 0) for k in names}
    st = {k: v for k, v in st.items() if v > 0.05 * int(r[ix['# Samples']] or 1)}
    print(r[ix['Address']][-5:], r[ix['# Samples']], r[ix['Instructions Executed']], r[ix['Source']][:64], st)
