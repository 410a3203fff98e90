"""profiles/ncu_kernels.json from an `ncu --set full` capture of ONE whole step: per kernel, the AVERAGE over the step's
launches of DRAM bytes and executed FP64 flops (bench.py multiplies nothing: it divides by its own average launch time).

usage: ncu_kernels.py out.json raw.csv [source-label]
raw.csv = `ncu -i capture.ncu-rep --page raw --csv` of a capture that holds every launch of one resident step (see
tools/gpu_session_ncu_step.sh).  Executed FP64 flops = 2 * DFMA + DADD + DMUL thread instructions (predicated on), from
smsp__sass_thread_inst_executed_op_d*_pred_on.sum.per_cycle_elapsed times the elapsed SMSP cycles.  Also prints a table for
profiles/*.txt: duration, FP64-pipe / tensor-pipe utilisation, occupancy, DRAM bytes per kernel."""
import collections, csv, json, re, sys


def main():
    dst, raw = sys.argv[1], sys.argv[2]
    label = sys.argv[3] if len(sys.argv) > 3 else raw
    rows = list(csv.reader(open(raw)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    ci = {h: i for i, h in enumerate(hdr)}

    def f(r, name, default=0.0):
        try:
            return float(r[ci[name]].replace(",", ""))
        except (KeyError, ValueError):
            return default

    def scale(name, table):
        return table.get(units[ci[name]].lower(), 1.0) if name in ci else 1.0
    bytes_u = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12}
    time_u = {"ms": 1, "us": 1e-3, "s": 1e3, "ns": 1e-6, "msecond": 1, "usecond": 1e-3, "nsecond": 1e-6, "second": 1e3}
    agg = collections.OrderedDict()
    for r in body:
        name = re.sub(r"^void ", "", r[ci["Kernel Name"]]).split("(")[0].replace("cbs::", "")
        key = name.split("<")[0]
        a = agg.setdefault(key, collections.defaultdict(float))
        cyc = f(r, "smsp__cycles_elapsed.avg") or f(r, "sm__cycles_elapsed.avg")
        ops = {k: f(r, "smsp__sass_thread_inst_executed_op_%s_pred_on.sum.per_cycle_elapsed" % k) * cyc for k in ("dfma", "dadd", "dmul")}
        ms = f(r, "gpu__time_duration.sum") * scale("gpu__time_duration.sum", time_u)
        a["n"] += 1
        a["ms"] += ms
        a["dram"] += f(r, "dram__bytes_read.sum") * scale("dram__bytes_read.sum", bytes_u) + f(r, "dram__bytes_write.sum") * scale("dram__bytes_write.sum", bytes_u)
        a["flop"] += 2 * ops["dfma"] + ops["dadd"] + ops["dmul"]
        for k in ("dfma", "dadd", "dmul"):
            a[k] += ops[k]
        # duration-weighted utilisation figures
        for short, metric in (("fp64_pipe", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
                              ("tensor_pipe", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
                              ("warps_active", "sm__warps_active.avg.pct_of_peak_sustained_active"),
                              ("issue_active", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                              ("dram_pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")):
            a[short] += f(r, metric) * ms
        a["regs"] = f(r, "launch__registers_per_thread")
    table = {}
    print("%-26s %5s %9s %9s %10s %9s %8s %8s %8s %6s" % ("kernel (one C5 step)", "n", "ms total", "GB dram", "GFLOP f64", "fp64pipe%", "tensor%", "warps%", "issue%", "regs"))
    for key, a in agg.items():
        n = a["n"]
        table[key] = {"launches_in_step": int(n), "dram_bytes_per_launch": a["dram"] / n, "fp64_flop_per_launch": a["flop"] / n,
                      "fp64_thread_inst_per_launch": {k: a[k] / n for k in ("dfma", "dadd", "dmul")},
                      "duration_ms_under_ncu": a["ms"] / n, "source": label}
        print("%-26s %5d %9.3f %9.3f %10.1f %9.1f %8.1f %8.1f %8.1f %6d" % (key, n, a["ms"], a["dram"] / 1e9, a["flop"] / 1e9, a["fp64_pipe"] / a["ms"],
                                                                        a["tensor_pipe"] / a["ms"], a["warps_active"] / a["ms"], a["issue_active"] / a["ms"], a["regs"]))
    json.dump(table, open(dst, "w"), indent=1)


if __name__ == "__main__":
    main()
