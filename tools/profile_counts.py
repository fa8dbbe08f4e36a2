"""Turn an ncu launch list (tools/gpu_round.sh, --csv with smsp__inst_executed.sum and dram bytes)
into profiles/bench_counts.json: per-launch warp-instruction and DRAM-byte counts that bench.py
combines with its live CUDA-event durations (roofline.achieved for the issue-bound BC7 kernel).

    python tools/profile_counts.py gpurun_out/r1a_launches.csv albedo-2048-bc7-normal [textures per batched launch]
"""
import collections
import csv
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path, key = sys.argv[1], sys.argv[2]
tex_per_mip_launch = int(sys.argv[3]) if len(sys.argv) > 3 else 1  # textures covered by one mip launch in the profiled command
rows = list(csv.reader(open(path)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[hi]
ki, mi, vi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
agg = collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows[hi + 1 :]:
    if len(r) > vi:
        name = "bc7" if "bc7_encode" in r[ki] else ("mip" if "mip_" in r[ki] else ("bc45" if "bc45" in r[ki] else None))
        if name:
            agg[name][r[mi]].append(float(r[vi].replace(",", "")))


def mean(name, metric):
    v = agg.get(name, {}).get(metric)
    return sum(v) / len(v) if v else None


with open(os.path.join(ROOT, "rosy_b200", "csrc", "bc7_encode.cu"), "rb") as f:
    sha = hashlib.sha256(f.read()).hexdigest()[:16]
out_path = os.path.join(ROOT, "profiles", "bench_counts.json")
doc = json.load(open(out_path)) if os.path.exists(out_path) else {}
doc[key] = {
    "source": os.path.basename(path),
    "bc7_source_sha": sha,
    "launches_averaged": len(agg.get("bc7", {}).get("gpu__time_duration.sum", [])),
    "bc7_textures_per_launch": tex_per_mip_launch,
    "bc7_warp_inst_per_texture": (mean("bc7", "smsp__inst_executed.sum") or 0) / tex_per_mip_launch or None,
    "bc7_dram_bytes_per_texture": ((mean("bc7", "dram__bytes_read.sum") or 0) + (mean("bc7", "dram__bytes_write.sum") or 0)) / tex_per_mip_launch,
    "bc7_ns_per_launch_under_ncu": mean("bc7", "gpu__time_duration.sum"),
    "mip_warp_inst_per_launch": mean("mip", "smsp__inst_executed.sum"),
    "mip_dram_bytes_per_launch": (mean("mip", "dram__bytes_read.sum") or 0) + (mean("mip", "dram__bytes_write.sum") or 0),
    "mip_ns_per_launch_under_ncu": mean("mip", "gpu__time_duration.sum"),
    "mip_textures_per_launch": tex_per_mip_launch,
    "mip_dram_bytes_per_texture": ((mean("mip", "dram__bytes_read.sum") or 0) + (mean("mip", "dram__bytes_write.sum") or 0)) / tex_per_mip_launch,
}
json.dump(doc, open(out_path, "w"), indent=1)
print(json.dumps(doc[key], indent=1))
