"""Print the per-launch metrics of an `ncu --csv --log-file` launch list as one line per kernel launch."""
import csv, sys
from collections import defaultdict
lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
d = defaultdict(dict)
for r in csv.DictReader(lines):
    d[(int(r['ID']), r['Kernel Name'][:44], r['Grid Size'], r['Block Size'])][r['Metric Name'].split('.')[0][-28:]] = r['Metric Value']
for k, v in sorted(d.items()):
    print(k, v)
