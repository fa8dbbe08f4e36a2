"""Join an `ncu --page source --csv` dump (per-SASS-instruction counters) with nvdisasm -g line info
and aggregate by source function: executed instructions, stall samples, no-instruction stalls.
    ncu -i rep.ncu-rep --page source --csv > src.csv
    python tools/ncu_hot_functions.py src.csv rosy_b200/csrc/bc7_encode.o rosy_b200/csrc/bc7_encode.cu [section-substring] [--lines N]
An object with several kernels has one .text section per kernel and instruction offsets restart in each: pass a substring of
the mangled name of the profiled kernel (e.g. bc7_encode_kernelILi512E, mip_fused_kernelILb1ELi1E) to pick its section.
"""
import bisect, collections, csv, os, re, subprocess, sys, tempfile

src_csv, obj, cu = sys.argv[1:4]
want_section = sys.argv[4] if len(sys.argv) > 4 and not sys.argv[4].startswith("--") else None
nlines = int(sys.argv[sys.argv.index("--lines") + 1]) if "--lines" in sys.argv else 25
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
line_of, cur, section = {}, None, None
for line in sass.split("\n"):
    m = re.match(r"\s*\.section\s+(\S+)", line)
    if m:
        section = m.group(1)
        continue
    if not section or not section.startswith(".text."):
        continue
    if want_section and want_section not in section:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/", line)
    if m:
        line_of[int(m.group(1), 16)] = cur
text = open(cu).read().split("\n")
starts, names = [], {}
for i, l in enumerate(text, 1):
    if re.match(r"^(__device__|__global__|inline |static |int |size_t )", l):
        txt = " ".join(text[i - 1 : i + 2]).replace("__launch_bounds__(NT)", "")
        m = re.search(r"(\w+)\s*\(", txt)
        if m:
            starts.append(i)
            names[i] = m.group(1)
rows = list(csv.reader(open(src_csv)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
base = None
agg = collections.defaultdict(lambda: collections.Counter())
lines = collections.defaultdict(lambda: collections.Counter())
for r in rows[hi + 1 :]:
    if len(r) < len(hdr) or not r[0].startswith("0x"):
        continue
    a = int(r[0], 16)
    if base is None:
        base = a
    key = line_of.get(a - base)
    fn = "<other>"
    if key and key[0] == os.path.basename(cu):
        j = bisect.bisect_right(starts, key[1]) - 1
        fn = names[starts[j]] if j >= 0 else "?"
    for k in ("Instructions Executed", "# Samples", "stall_no_inst", "stall_wait", "stall_math", "stall_short_sb", "stall_not_selected", "Thread Instructions Executed"):
        v = int(float(r[col[k]] or 0))
        agg[fn][k] += v
        if key:
            lines[key][k] += v
    agg[fn]["static"] += 1
tot = sum(v["Instructions Executed"] for v in agg.values())
tots = sum(v["# Samples"] for v in agg.values())
print(f"{'function':24s} {'static':>6s} {'inst%':>6s} {'samp%':>6s} {'no_inst':>8s} {'wait':>7s} {'math':>7s} {'short_sb':>8s} {'lanes':>6s}")
for fn, v in sorted(agg.items(), key=lambda kv: -kv[1]["Instructions Executed"]):
    ie = v["Instructions Executed"]
    print(f"{fn:24s} {v['static']:6d} {100*ie/tot:6.1f} {100*v['# Samples']/max(tots,1):6.1f} {v['stall_no_inst']:8d} {v['stall_wait']:7d} {v['stall_math']:7d} {v['stall_short_sb']:8d} {v['Thread Instructions Executed']/max(ie,1):6.1f}")
print("hottest source lines (by samples):")
for key, v in sorted(lines.items(), key=lambda kv: -kv[1]["# Samples"])[:nlines]:
    print(f"  {key[0]}:{key[1]:<5d} samples {v['# Samples']:6d} inst {100*v['Instructions Executed']/tot:5.1f}%  no_inst {v['stall_no_inst']:5d}  | {text[key[1]-1].strip()[:90] if key[0]==os.path.basename(cu) else ''}")
