"""Dev tool: text summaries of the ncu artefacts gpurun brought back.
  python tools/ncu_summary.py full  gpurun_out/<tag>_mf_full.ncu-rep   > profiles/<tag>_mf_kernel_ncu_summary.txt
  python tools/ncu_summary.py list  gpurun_out/<tag>_launches.csv      > profiles/<tag>_launch_shares.txt"""
import collections, csv, io, re, subprocess, sys

KEEP = re.compile(r"^(dram__bytes_(read|write)\.sum$|gpu__time_duration\.sum|launch__(block_size|grid_size|registers_per_thread|"
                  r"waves_per_multiprocessor|occupancy_limit_(registers|shared_mem))|sm__inst_executed_pipe_(alu|fma|fp64|lsu|xu|uniform)"
                  r"\.avg\.pct_of_peak_sustained_active|sm__pipe_fp64_cycles_active\.avg\.pct|sm__throughput\.avg\.pct|"
                  r"sm__warps_active\.avg\.pct_of_peak_sustained_active|smsp__average_warps_issue_stalled_.*_per_issue_active|"
                  r"smsp__inst_executed\.sum$|smsp__inst_executed_pipe_fp64\.sum$|sm__inst_executed_pipe_fp64\.sum$|smsp__issue_active\.avg\.pct|"
                  r"l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum$|sm__pipe_(fma|fmaheavy|alu)_cycles_active\.avg\.pct_of_peak_sustained_active|sm__inst_executed_pipe_fmaheavy\.avg\.pct_of_peak_sustained_active|sm__cycles_elapsed\.max|lts__t_sector_hit_rate\.pct|smsp__cycles_active\.avg$|sm__inst_executed\.sum$)")

def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    names, units, vals = rows[0], rows[1], rows[2]
    for n, u, v in zip(names, units, vals):
        if n == "Kernel Name":
            print("# kernel:", v)
        if KEEP.search(n):
            print("%-90s %-16s %s" % (n, u, v))

def raw(path):
    """same summary from the CSV of `ncu -i x.ncu-rep --page raw --csv` (what tools/gpu_session.sh brings back: the
    .ncu-rep files themselves are ~15 MB each and gpurun_out is capped at 64 MiB)"""
    rows = list(csv.reader(open(path)))
    names, units = rows[0], rows[1]
    for vals in rows[2:]:
        for n, u, v in zip(names, units, vals):
            if n == "Kernel Name":
                print("# kernel:", v)
            if KEEP.search(n) and not re.search(r"\.(max|min|sum)\.pct|per_second|_allocated", n):
                if "stalled" in n and float(v.replace(",", "")) < 0.3:
                    continue
                print("%-90s %-16s %s" % (n, u, v))


def launches(path):
    txt = open(path).read()
    txt = txt[txt.index('"ID"'):]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(io.StringIO(txt)):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        if row["Metric Unit"] in ("us", "usecond"): v *= 1e3
        if row["Metric Unit"] in ("ms", "msecond"): v *= 1e6
        k = re.sub(r"\(.*", "", row["Kernel Name"])[:70]
        agg[k][0] += 1; agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    print("# kernel, launches, total ns, share")
    for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:25]:
        print("%-72s %6d %14.0f %6.2f%%" % (k, n, t, 100 * t / tot))

def kernels(path, pairs=None):
    """per-kernel table from an `ncu --metrics sm__pipe_fp64_cycles_active...,smsp__inst_executed.sum,
    gpu__time_duration.sum,launch__registers_per_thread --csv` log; pairs = pair-boxes per launch (for instr/pair)"""
    txt = open(path).read()
    txt = txt[txt.index('"ID"'):]
    rows = collections.OrderedDict()
    for row in csv.DictReader(io.StringIO(txt)):
        key = (row["ID"], row["Kernel Name"])
        rows.setdefault(key, {})[row["Metric Name"]] = (float(row["Metric Value"].replace(",", "")), row["Metric Unit"])
    for (_, name), m in rows.items():
        if "sym_kernel" not in name and "mf_kernel" not in name and "st_kernel" not in name:
            continue
        ms, unit = m["gpu__time_duration.sum"]
        ms *= {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}[unit]
        pipe = [v for k, (v, _) in m.items() if k.startswith("sm__pipe_fp64_cycles_active")][0]
        inst = m["smsp__inst_executed.sum"][0]
        line = "%-62s fp64_pipe_active %5.1f%%  regs %3d  ms %8.1f" % (name[:62], pipe, m["launch__registers_per_thread"][0], ms)
        if pairs:
            line += "  warp instr per 32 interactions %6.1f" % (inst * 32 / float(pairs))
        print(line)

if sys.argv[1] == "kernels":
    kernels(sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else None)
else:
    {"full": full, "list": launches, "raw": raw}[sys.argv[1]](sys.argv[2])
