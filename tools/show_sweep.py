import json, sys
from collections import defaultdict
rows = json.load(open(sys.argv[1]))
best = defaultdict(list)
for r in rows:
    if r.get('op') == 'broadcast':
        print(r); continue
    best[(r['dtype'], r['bytes'])].append(r)
for key in sorted(best):
    rs = best[key]
    nccl = [r for r in rs if r['impl'] == 'nccl'][0]
    ours = sorted([r for r in rs if r['impl'] == 'b2c'], key=lambda r: r['us_med'])
    o = ours[0]
    tag = lambda r: f"{r['algo'][:3]}{r['variant']}{'z' if r.get('zerocopy') else ''}/{r['max_blocks']}"  # noqa: E731
    print(f"{key[0]:9s} {key[1]:>11d} nccl {nccl['us_med']:8.1f}us {nccl['busbw_GBps']:7.1f} | b2c {o['us_med']:8.1f}us {o['busbw_GBps']:7.1f} GB/s {tag(o)} | " + " ".join(f"{tag(r)}:{r['us_med']:.0f}" for r in ours[1:6]))
