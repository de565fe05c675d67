#!/usr/bin/env python
"""Print selected metrics of an `ncu --page raw --csv` dump: tools/ncu_pick.py file.csv [substr ...]"""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = sys.argv[2:] or ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct',
                        'l1tex__m_xbar2l1tex_read_bytes.sum', 'sm__pipe_tensor_cycles_active_realtime.avg.pct',
                        'lts__throughput.avg.pct', 'dram__throughput.avg.pct', 'sm__cycles_elapsed.max',
                        'launch__registers_per_thread', 'smsp__inst_executed.sum', 'sm__throughput.avg.pct']
for i, h in enumerate(hdr):
    if any(w in h for w in want):
        print('%-90s %-10s %s' % (h, units[i], ' | '.join(r[i] for r in rows[2:])))
