"""profiles/ncu_kernels.json from `ncu --set full` captures: per kernel launch, DRAM bytes and executed FP64 flops.

usage: ncu_kernels.py out.json key=capture.ncu-rep[:launch_index] ...
`key` is the kernel-name prefix bench.py matches (e.g. k_glm_fit_lanes).  Executed FP64 flops = 2 * DFMA + DADD + DMUL
thread instructions (predicated on), from smsp__sass_thread_inst_executed_op_d*_pred_on.sum.per_cycle_elapsed times the
elapsed SMSP cycles.  Also prints a short summary (tools/ncu_raw_summary.py keys) for profiles/*.txt."""
import csv, io, json, subprocess, sys


def raw_rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    dst = sys.argv[1]
    table = {}
    for arg in sys.argv[2:]:
        key, rep = arg.split("=", 1)
        idx = 0
        if ":" in rep and rep.rsplit(":", 1)[1].isdigit():
            rep, i = rep.rsplit(":", 1); idx = int(i)
        rows = raw_rows(rep)
        hdr, val = rows[0], rows[2 + idx]
        g = {h: v for h, v in zip(hdr, val)}

        def f(name, default=0.0):
            try:
                return float(g[name].replace(",", ""))
            except (KeyError, ValueError):
                return default
        units = {h: u for h, u in zip(hdr, rows[1])}

        def bytes_of(name):
            v = f(name); u = units.get(name, "byte").lower()
            return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12}.get(u, 1)
        cyc = f("smsp__cycles_elapsed.avg") or f("sm__cycles_elapsed.avg")
        ops = {k: f("smsp__sass_thread_inst_executed_op_%s_pred_on.sum.per_cycle_elapsed" % k) * cyc for k in ("dfma", "dadd", "dmul")}
        table[key] = {"kernel": g.get("Kernel Name", "?")[:100],
                      "dram_bytes_per_launch": bytes_of("dram__bytes_read.sum") + bytes_of("dram__bytes_write.sum"),
                      "fp64_flop_per_launch": 2 * ops["dfma"] + ops["dadd"] + ops["dmul"],
                      "fp64_thread_inst": ops, "duration_ms_under_ncu": f("gpu__time_duration.sum") * {"ms": 1, "us": 1e-3, "s": 1e3, "ns": 1e-6}.get(units.get("gpu__time_duration.sum", "ms").lower(), 1),
                      "source": rep}
        print(key, json.dumps(table[key]))
    json.dump(table, open(dst, "w"), indent=1)


if __name__ == "__main__":
    main()
