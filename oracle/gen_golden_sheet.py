#!/usr/bin/env python
"""TEST INFRASTRUCTURE.  Transcribes the reference's only recorded outputs -- the two blocks of
/root/reference/master_car_statistics.csv.xlsx, sheet 'master_car_statistics'  --
into tests/golden/master_car_statistics_sheet.csv, so that the CSV writer, the per-car decision
rules and the comparison tool can be checked against them where /root/reference does not exist.

    python oracle/gen_golden_sheet.py          # run in the build container only

Columns: block ('without_erosion' = V2, 'with_erosion' = V3 / cvs_erosion.py), then the eight
columns the sheet keeps of append_to_master_csv (cvs_erosion.py:243-254).
"""
import csv
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lidar_object_detection_b200 import compare as C  # noqa: E402  (host-only module: zip + XML reader)

SRC = '/root/reference/master_car_statistics.csv.xlsx'
DST = os.path.join(ROOT, 'tests', 'golden', 'master_car_statistics_sheet.csv')


def main():
    blocks = C.sheet_blocks(C.read_xlsx_sheet(SRC, 'master_car_statistics'))
    names = {'Without Erosion': 'without_erosion', 'Erosion': 'with_erosion'}
    with open(DST, 'w', newline='') as fh:
        w = csv.writer(fh, lineterminator='\n')
        w.writerow(('block',) + C.COLUMNS)
        for title, rows in blocks.items():
            for d in rows:
                w.writerow([names[title]] + [d[c] for c in C.COLUMNS])
    print('wrote', DST, {names[t]: len(r) for t, r in blocks.items()})


if __name__ == '__main__':
    main()
