#!/bin/bash
# Turn the artefacts of one tools/gpu_round.sh run (gpurun_out/<tag>_*) into the tracked summaries under profiles/:
# ncu --set full summaries per kernel, the BC7 hot-function table, the launch list, the bench lines and the SASS
# mnemonic evidence (UTMALDG / SYNCS for the TMA mip kernel, IDP.4A for BC7, no tensor-core ops).
#   tools/collect_profiles.sh r2c
set -e
TAG=$1
cd "$(dirname "$0")/.."
for k in bc7 mip bc5; do
    if [ -f gpurun_out/${TAG}_$k.ncu-rep ]; then
        ncu -i gpurun_out/${TAG}_$k.ncu-rep --page raw --csv > /tmp/${TAG}_$k.raw.csv 2>/dev/null
        python tools/ncu_pick.py /tmp/${TAG}_$k.raw.csv > profiles/${TAG}_${k}_ncu_full_summary.txt
    fi
done
if [ -f gpurun_out/${TAG}_bc7.ncu-rep ]; then
    ncu -i gpurun_out/${TAG}_bc7.ncu-rep --page source --csv > /tmp/${TAG}_bc7.src.csv 2>/dev/null
    python tools/ncu_hot_functions.py /tmp/${TAG}_bc7.src.csv rosy_b200/csrc/bc7_encode.o rosy_b200/csrc/bc7_encode.cu bc7_encode_kernelILi512E > profiles/${TAG}_bc7_hot_functions.txt || true
fi
for f in launches.csv bench.json bench_ref.json bench_bc5.json bench_bc4.json mapping.json; do
    [ -f gpurun_out/${TAG}_$f ] && cp gpurun_out/${TAG}_$f profiles/${TAG}_$f
done
# the per-launch instruction counts bench.py's issue-slot roofline uses (keyed by the BC7 source hash)
[ -f gpurun_out/${TAG}_bench_counts.json ] && cp gpurun_out/${TAG}_bench_counts.json profiles/bench_counts.json
# SASS evidence per kernel: mnemonic histogram + the lines that prove TMA / mbarrier / dp4a
{
    for o in mip_kernels bc7_encode bc45_encode; do
        cuobjdump -sass rosy_b200/csrc/$o.o 2>/dev/null | python -c "
import re, sys, collections
cur, cnt = None, collections.defaultdict(collections.Counter)
for line in sys.stdin:
    m = re.search(r'Function : (\S+)', line)
    if m:
        cur = m.group(1); continue
    m = re.match(r'\s+/\*[0-9a-f]{4,6}\*/\s+(.*?);', line)
    if m and cur:
        ins = re.sub(r'^@!?U?P\d+\s+', '', m.group(1)).split()[0]
        cnt[cur][ins] += 1
for f, c in cnt.items():
    tot = sum(c.values())
    print('== %s.o  %s  (%d SASS instructions)' % ('$o', f, tot))
    keys = [k for k in c if re.match(r'UTMA|SYNCS|IDP|FADD2|FFMA2|VIMNMX3|MUFU|F2I|I2F|HMMA|IMMA|UTC|LDSM|FENCE|UTMACCTL|ATOM|RED', k)]
    print('   evidence:', ', '.join('%s x%d' % (k, c[k]) for k in sorted(keys)))
    print('   top:', ', '.join('%s x%d' % kv for kv in c.most_common(12)))
"
    done
} > profiles/${TAG}_sass_evidence.txt
ls -la profiles | grep ${TAG}
