"""Turn the round's ncu captures into the text summaries committed under profiles/.

    python scripts/ncu_summary.py launches gpurun_out/s3_launches.csv  > profiles/r2_launches.txt
    python scripts/ncu_summary.py full gpurun_out/s3_head.ncu-rep      > profiles/r2_ncu_head.txt

`launches`: the --metrics gpu__time_duration.sum launch list of bench.py, grouped per kernel: count,
mean / total time and SHARE of all libavtubes time (cold-cache, serialised times: the share is what
must agree with bench.py's stage_ms, not the absolute).
`full`: one line per profiled launch of an `ncu --set full` report with the counters the roofline
discussion uses (duration, DRAM bytes, tensor pipe, issue slots, registers, L2 throughput).
"""
import csv
import io
import re
import subprocess
import sys

OURS = re.compile(r"avt::|prep_visual|sim_tc|bwd_d|finish_rows|row_grad|audio_|eval_maps|auc_kernel|mtc_|"
                  r"rasterise|spectrogram|video_metrics|ciou_counts|dv_finish|weighted|transpose_audio|sim_fp32|bwd_fp32")


def short(name):
    name = name.replace("avt::", "").replace("void ", "")
    return re.sub(r"\(.*$", "", name)[:60]


def launches(path):
    rows = [r for r in csv.reader(open(path, errors="replace")) if r]
    h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr = rows[h]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg, order = {}, []
    for r in rows[h + 1:]:
        if len(r) <= mv:
            continue
        t = float(r[mv].replace(",", ""))
        t = t / 1e3 if r[mu] in ("ns", "nsecond") else (t * 1e3 if r[mu] in ("ms", "msecond") else t)   # -> us
        k = short(r[kn])
        if k not in agg:
            agg[k] = [0, 0.0, bool(OURS.search(r[kn]))]
            order.append(k)
        agg[k][0] += 1
        agg[k][1] += t
    ours = sum(v[1] for v in agg.values() if v[2])
    allt = sum(v[1] for v in agg.values())
    print(f"# {path}: {sum(v[0] for v in agg.values())} launches, {allt:.0f} us in total, {ours:.0f} us in libavtubes kernels")
    print(f"# {'kernel':60s} {'n':>5s} {'mean us':>9s} {'total us':>10s} {'share of ours':>14s}")
    for k in sorted(agg, key=lambda k: -agg[k][1]):
        n, t, o = agg[k]
        print(f"{k:62s} {n:5d} {t / n:9.1f} {t:10.1f} {('%13.1f%%' % (100 * t / ours)) if o else '   (library)'}")


WANT = [("gpu__time_duration.sum", "dur"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_active", "L1%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "warp inst"),
        ("launch__grid_size", "grid"), ("launch__occupancy_limit_shared_mem", "occ smem")]


def full(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    kn = hdr.index("Kernel Name")
    cols = [(hdr.index(m), lab) for m, lab in WANT if m in hdr]
    print(f"# ncu --set full --clock-control none: {path} ({len(rows) - 2} launches); dram bytes per launch")
    print("# " + "kernel".ljust(44) + " ".join(lab.rjust(12) for _, lab in cols))
    for r in rows[2:]:
        cells = []
        for i, _ in cols:
            u = units[i].replace("byte", "B").replace("second", "s").replace("register/thread", "").replace("%", "")
            cells.append((r[i] + u).rjust(12))
        print(short(r[kn])[:45].ljust(46) + " ".join(cells))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
