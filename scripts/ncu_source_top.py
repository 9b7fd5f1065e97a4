"""Reads `ncu -i report.ncu-rep --page source --csv [--print-source sass|cuda,sass]` output: the SASS (or CUDA source) lines with
the most warp-stall samples, and the opcode mix.   python scripts/ncu_source_top.py source.csv [lines]"""
import collections, csv, re, sys

rows = list(csv.reader(open(sys.argv[1])))
limit = int(sys.argv[2]) if len(sys.argv) > 2 else 30
header = next(i for i, row in enumerate(rows) if "Source" in row and "# Samples" in row)
index = {name: i for i, name in enumerate(rows[header])}
data = [row for row in rows[header + 1:] if len(row) == len(rows[header])]


def number(row, key):
    try:
        return float(row[index[key]])
    except (ValueError, KeyError):
        return 0.0


samples = sum(number(row, "# Samples") for row in data) or 1.0
instructions = sum(number(row, "Instructions Executed") for row in data) or 1.0
print(f"{len(data)} lines, {instructions:.4g} warp instructions, {samples:.0f} samples")
print("--- lines with the most samples")
for row in sorted(data, key=lambda row: -number(row, "# Samples"))[:limit]:
    print(f"{100 * number(row, '# Samples') / samples:5.1f}%  inst {number(row, 'Instructions Executed'):9.3g}  {row[index['Source']].strip()[:110]}")
groups = collections.defaultdict(lambda: [0.0, 0.0])
for row in data:
    match = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_.]+)", row[index["Source"]].strip())
    name = ".".join(match.group(2).split(".")[:2]) if match else "?"
    groups[name][0] += number(row, "Instructions Executed")
    groups[name][1] += number(row, "# Samples")
print("--- opcodes")
for name, (count, stalled) in sorted(groups.items(), key=lambda item: -item[1][1])[:limit]:
    print(f"{name:20s} inst {100 * count / instructions:5.1f}%  samples {100 * stalled / samples:5.1f}%")
