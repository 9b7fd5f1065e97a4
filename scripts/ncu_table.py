"""The per-kernel summary table of profiles/README.md from an `ncu --set full` report:
   python scripts/ncu_table.py gpurun_out/x.ncu-rep   (runs `ncu -i ... --page raw --csv` and picks the columns)."""
import csv, io, subprocess, sys

report = sys.argv[1]
raw = subprocess.run(["ncu", "-i", report, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
header, units, data = rows[0], rows[1], rows[2:]
column = {name: index for index, name in enumerate(header)}


def value(row, name, scale=1.0):
    if name not in column or row[column[name]] in ("", "n/a"):
        return float("nan")
    number = float(row[column[name]].replace(",", ""))
    unit = units[column[name]]
    factor = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "usecond": 1e-3, "msecond": 1.0, "nsecond": 1e-6, "second": 1e3, "us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3}.get(unit, 1.0)
    return number * factor * scale


print("| kernel | time ms | DRAM read GB | DRAM write GB | IPC | issue active % | warps active % | regs | grid x block | shared wavefronts |")
print("|---|---|---|---|---|---|---|---|---|---|")
for row in data:
    name = row[column["Kernel Name"]].split("(")[0]
    grid = row[column["Grid Size"]].replace(" ", "") if "Grid Size" in column else "?"
    block = row[column["Block Size"]].replace(" ", "") if "Block Size" in column else "?"
    print("| `%s` | %.3f | %.3f | %.3f | %.2f | %.0f | %.0f | %s | %s x %s | %.3g |" % (
        name[-48:], value(row, "gpu__time_duration.sum"), value(row, "dram__bytes_read.sum", 1e-9), value(row, "dram__bytes_write.sum", 1e-9),
        value(row, "sm__inst_executed.avg.per_cycle_active"), value(row, "sm__inst_issued.avg.pct_of_peak_sustained_active"),
        value(row, "sm__warps_active.avg.pct_of_peak_sustained_active"), row[column["launch__registers_per_thread"]] if "launch__registers_per_thread" in column else "?",
        grid, block, value(row, "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")))
