#!/usr/bin/env python3
"""Print a compact per-kernel table from an `ncu --csv --metrics ...` log (tools/ncu_quick.sh)."""
import csv, sys, re, collections
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hdr = rows[0]; ik, im, iv, iid = (hdr.index(x) for x in ('Kernel Name', 'Metric Name', 'Metric Value', 'ID'))
per = collections.OrderedDict()
for r in rows[1:]:
    per.setdefault(r[iid], {'k': re.sub(r'\(.*', '', r[ik])})[r[im]] = r[iv].replace(',', '')
short = lambda m: re.sub(r'smsp__average_warps_issue_stalled_|_per_issue_active.ratio|\.avg\.pct_of_peak_sustained_(active|elapsed)|\.sum', '', m)
for i, d in per.items():
    print('== %s #%s' % (d['k'], i))
    print('   ' + '  '.join('%s=%s' % (short(m), ('%.3g' % float(v)) if re.match(r'^[0-9.eE+-]+$', v) else v) for m, v in d.items() if m != 'k'))
