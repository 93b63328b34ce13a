"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list: the LAST train step (from the last q_sample
kernel to the end), per kernel: launches, total us, share.  Usage: summarize_launches.py in.csv out.md [first]
(`first`: aggregate the first complete step instead — for lists cut off by `ncu -c N`)."""
import collections
import csv
import sys


def main(src, dst, first_complete=False):
    rows = [r for r in csv.reader(l for l in open(src) if l.startswith('"'))]
    hdr, rows = rows[0], rows[1:]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    names = [r[ki].split("(")[0].replace("b200::", "") for r in rows]
    qs = [i for i, n in enumerate(names) if n == "q_sample_kernel"]
    last = qs[-1]
    end = len(names)
    if first_complete and len(qs) >= 2:  # a truncated list (ncu -c N): the first step that is complete
        last, end = qs[0], qs[1]
    names, rows = names[:end], rows[:end]
    agg = collections.OrderedDict()
    for n, r in zip(names[last:], rows[last:]):
        if n.startswith("void at::"):
            n = "torch:" + n.split("<")[0].replace("void at::", "")
        d = agg.setdefault(n, [0, 0.0])
        d[0] += 1
        d[1] += float(r[vi]) / 1e3
    tot = sum(v[1] for v in agg.values())
    out = [f"# ncu launch list, {'first complete' if first_complete else 'last'} train step of `{src}`", "",
           f"{end - last} launches, {tot:.1f} us summed (cold-cache, serialised: compare SHARES, not absolutes)", "",
           "| kernel | launches | total us | share |", "|---|---:|---:|---:|"]
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append(f"| {n} | {c} | {t:.1f} | {t / tot * 100:.1f}% |")
    open(dst, "w").write("\n".join(out) + "\n")
    print("\n".join(out))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], len(sys.argv) > 3 and sys.argv[3] == "first")
