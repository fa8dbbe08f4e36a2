"""Attribute the SASS instructions of a .o/.cubin to the source functions of a .cu file (via -lineinfo).
    python tools/sass_by_function.py rosy_b200/csrc/bc7_encode.o rosy_b200/csrc/bc7_encode.cu
"""
import bisect, collections, os, re, subprocess, sys, tempfile

obj, cu = sys.argv[1], sys.argv[2]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
cnt, cur, section = collections.Counter(), None, None
for line in sass.split("\n"):
    m = re.match(r"\s*\.section\s+(\S+)", line)
    if m:
        section = m.group(1)
        continue
    if not section or not section.startswith(".text."):
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,6}\*/", line):
        cnt[(section[:50], cur)] += 1
src = open(cu).read().split("\n")
starts, names = [], {}
for i, l in enumerate(src, 1):
    if re.match(r"^(__device__|__global__|inline |static |template|int |Bc7Tier |size_t )", l) and "(" in " ".join(src[i - 1 : i + 2]):
        if re.match(r"^template", l):
            continue
        txt = " ".join(src[i - 1 : i + 2]).replace("__launch_bounds__(NT)", "").replace("__launch_bounds__(256)", "")
        m = re.search(r"(\w+)\s*\(", txt)
        if m:
            starts.append(i)
            names[i] = m.group(1)
agg = collections.Counter()
for (sec, key), c in cnt.items():
    if key is None or key[0] != os.path.basename(cu):
        agg[(sec, "<other>")] += c
        continue
    j = bisect.bisect_right(starts, key[1]) - 1
    agg[(sec, names[starts[j]] if j >= 0 else "?")] += c
bysec = collections.Counter()
for (sec, n), c in agg.items():
    bysec[sec] += c
for sec, tot in bysec.items():
    print(f"{sec}: {tot} instr = {tot*16/1024:.1f} KB")
    for (s, n), c in sorted(agg.items(), key=lambda kv: -kv[1]):
        if s == sec:
            print(f"   {n:28s}{c:7d} {c*16/1024:7.1f} KB")
