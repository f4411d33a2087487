"""Copies the round's ncu CSVs from gpurun_out/ into profiles/, prints their summaries and rewrites profiles/traffic.json
(DRAM bytes per fused launch + the sha256 of the csrc/pair_umma.cu they were captured from)."""
import collections
import csv
import hashlib
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
csv.field_size_limit(10 ** 9)
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e3, "us": 1, "ns": 1e-3, "s": 1e6}


def raw(path, want):
    rows = list(csv.reader(open(path)))
    h, u = rows[0], rows[1]
    idx = {n: i for i, n in enumerate(h)}
    out = {"name": [r[idx["Kernel Name"]] for r in rows[2:]]}
    for w in want:
        i = idx[w]
        out[w] = [float(r[i].replace(",", "")) * SCALE.get(u[i], 1) for r in rows[2:]]
    return out


def main():
    for f in ("r02_bench_launches.csv", "r02_pair_umma_full.csv", "r02_stream_umma_full.csv"):
        shutil.copy(os.path.join(ROOT, "gpurun_out", f), os.path.join(ROOT, "profiles", f))
    rows = list(csv.reader(open(os.path.join(ROOT, "profiles", "r02_bench_launches.csv"))))
    hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
    h = rows[hi]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    tot, cnt = collections.Counter(), collections.Counter()
    for r in rows[hi + 1:]:
        if len(r) <= vi:
            continue
        k = "pair_umma_kernel" if "pair_umma" in r[ki] else ("stream_umma_kernel" if "stream_umma" in r[ki] else r[ki][:40])
        tot[k] += float(r[vi].replace(",", ""))
        cnt[k] += 1
    T = sum(tot.values())
    print("bench launch list: %d launches, %.1f ms" % (sum(cnt.values()), T / 1e6))
    for k, v in tot.most_common():
        print("  %-40s %4d launches %10.1f us %5.1f %%" % (k, cnt[k], v / 1e3, 100 * v / T))
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]
    for f in ("r02_pair_umma_full.csv", "r02_stream_umma_full.csv"):
        d = raw(os.path.join(ROOT, "profiles", f), want)
        print(f)
        for i, n in enumerate(d["name"]):
            t, a, b, p = (d[w][i] for w in want)
            print("  %2d %-34s %8.1f us  read %7.1f MB  written %7.1f MB  %5.2f TB/s  tensor pipe active %4.1f %%" %
                  (i, n[-34:], t, a / 1e6, b / 1e6, (a + b) / t / 1e6, p))
        if "pair" in f:
            avg = (sum(d[want[1]]) + sum(d[want[2]])) / len(d["name"])
            sha = hashlib.sha256(open(os.path.join(ROOT, "visual_onoma_to_wave_b200", "csrc", "pair_umma.cu"), "rb").read()).hexdigest()
            json.dump({"pair_umma_kernel_dram_bytes_per_launch": avg, "batch": 128, "pair_umma_cu_sha256": sha,
                       "source": "profiles/r02_pair_umma_full.csv: ncu --set full over the 28 fused launches of one Generator.forward at "
                                 "B=128 x T=172 (dram__bytes_read.sum + dram__bytes_write.sum, mean per launch)"},
                      open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
            print("  traffic.json: %.1f MB per launch, sum of launch times %.1f us" % (avg / 1e6, sum(d[want[0]])))


if __name__ == "__main__":
    sys.exit(main())
