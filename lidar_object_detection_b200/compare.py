"""Comparison of a ``master_car_statistics.csv`` with the reference's recorded results
(the ``.xlsx``-free comparison tool for Appendix B).

The reference keeps its only recorded outputs in ``master_car_statistics.csv.xlsx``
(sheet ``master_car_statistics``): two blocks of eight columns, left = without erosion
(V2), right = with 3x3 erosion (V3 / cvs_erosion.py), the columns of
``append_to_master_csv`` (cvs_erosion.py:243-254) minus ``matched_bbox_id`` and
``timestamp``.  No spreadsheet library is installed, so ``read_xlsx_sheet`` reads the
workbook as what it is: a zip of XML parts.

``check_rows`` verifies the invariants every row of a master CSV must satisfy (they follow
from cvs_erosion.py:214-229, 243-254), ``compare_rows`` joins two tables on (frame, car_id),
and ``erosion_effect`` is the with/without-erosion summary the spreadsheet's charts show.

    python -m lidar_object_detection_b200.compare ours.csv tests/golden/master_car_statistics_sheet.csv
"""
import csv
import re
import zipfile
from xml.etree import ElementTree as ET

COLUMNS = ('frame', 'car_id', 'total_points', 'points_inside_bbox', 'points_outside_bbox',
           'inside_percentage', 'outside_percentage', 'is_matched')
_NS = {'m': 'http://schemas.openxmlformats.org/spreadsheetml/2006/main',
       'r': 'http://schemas.openxmlformats.org/officeDocument/2006/relationships'}
_PKG_REL = '{http://schemas.openxmlformats.org/package/2006/relationships}Relationship'


# --------------------------------------------------------------------------
# xlsx without a spreadsheet library
# --------------------------------------------------------------------------

def _col_index(ref):
    """'A1' -> 0, 'Q74' -> 16, 'AA3' -> 26."""
    n = 0
    for ch in re.match(r'[A-Z]+', ref).group(0):
        n = n * 26 + (ord(ch) - 64)
    return n - 1


def read_xlsx_sheet(path, sheet=None):
    """-> list of rows, each a dict {column index: value}; numbers as float (ints when
    integral), shared/inline strings as str, booleans as bool.  ``sheet``: name, or None
    for the first sheet."""
    with zipfile.ZipFile(path) as z:
        wb = ET.fromstring(z.read('xl/workbook.xml'))
        rels = ET.fromstring(z.read('xl/_rels/workbook.xml.rels'))
        target = {r.get('Id'): r.get('Target') for r in rels.iter(_PKG_REL)}
        part = None
        for s in wb.find('m:sheets', _NS):
            if sheet is None or s.get('name') == sheet:
                part = target[s.get('{%s}id' % _NS['r'])]
                break
        if part is None:
            raise KeyError('no sheet %r in %s' % (sheet, path))
        part = part.lstrip('/')
        if not part.startswith('xl/'):
            part = 'xl/' + part
        shared = []
        if 'xl/sharedStrings.xml' in z.namelist():
            for si in ET.fromstring(z.read('xl/sharedStrings.xml')).findall('m:si', _NS):
                shared.append(''.join(t.text or '' for t in si.iter('{%s}t' % _NS['m'])))
        rows = []
        for row in ET.fromstring(z.read(part)).find('m:sheetData', _NS):
            out = {}
            for c in row.findall('m:c', _NS):
                v = c.find('m:v', _NS)
                t = c.get('t')
                if t == 'inlineStr':
                    val = ''.join(x.text or '' for x in c.iter('{%s}t' % _NS['m']))
                elif v is None or v.text is None:
                    continue
                elif t == 's':
                    val = shared[int(v.text)]
                elif t == 'b':
                    val = v.text == '1'
                elif t == 'str':
                    val = v.text
                else:
                    f = float(v.text)
                    val = int(f) if f == int(f) and 'e' not in v.text.lower() and '.' not in v.text else f
                out[_col_index(c.get('r'))] = val
            rows.append(out)
    return rows


def _as_bool(x):
    if isinstance(x, str):
        return x.strip().lower() in ('true', '1')
    return bool(x)


def sheet_blocks(rows, width=len(COLUMNS)):
    """Split the reference sheet into its blocks.  A block starts at every column of the
    header row whose text is 'frame'; the row above holds the block's title.  ->
    {title: [row dict with COLUMNS keys]}"""
    hdr_i = next(i for i, r in enumerate(rows) if 'frame' in [v for v in r.values() if isinstance(v, str)])
    hdr = rows[hdr_i]
    starts = sorted(c for c, v in hdr.items() if v == 'frame')
    blocks = {}
    for k, c0 in enumerate(starts):
        names = COLUMNS[:width]          # by position: the author renamed the percentage headers by hand
        title = str(rows[hdr_i - 1].get(c0, 'block%d' % k) if hdr_i > 0 else 'block%d' % k).strip()
        out = []
        for r in rows[hdr_i + 1:]:
            if c0 not in r or not isinstance(r[c0], (int, float)):
                continue
            d = {n: r.get(c0 + j) for j, n in enumerate(names) if n}
            d['is_matched'] = _as_bool(d.get('is_matched', False))
            for n in ('frame', 'car_id', 'total_points', 'points_inside_bbox', 'points_outside_bbox'):
                d[n] = int(d[n])
            for n in ('inside_percentage', 'outside_percentage'):
                d[n] = float(d[n])
            out.append(d)
        blocks[str(title)] = out
    return blocks


# --------------------------------------------------------------------------
# tables
# --------------------------------------------------------------------------

def read_master_csv(path):
    """Rows of a ``master_car_statistics.csv`` (the 10 columns of cvs_erosion.py:243-254, or
    the 8 of the sheet, plus an optional leading 'block' column) as dicts with typed values."""
    out = []
    with open(path, newline='') as fh:
        for r in csv.DictReader(fh):
            d = dict(r)
            for n in ('frame', 'car_id', 'total_points', 'points_inside_bbox', 'points_outside_bbox'):
                d[n] = int(float(d[n]))
            if d.get('matched_bbox_id') not in (None, ''):
                d['matched_bbox_id'] = int(float(d['matched_bbox_id']))
            for n in ('inside_percentage', 'outside_percentage'):
                d[n] = float(d[n])
            d['is_matched'] = _as_bool(d['is_matched'])
            out.append(d)
    return out


def check_rows(rows, min_points=10):
    """Invariants of every row (cvs_erosion.py:214-229, 243-254).  -> list of
    (frame, car_id, message); empty when the table is consistent."""
    bad = []
    seen = set()
    for d in rows:
        key = (d['frame'], d['car_id'])
        t, i, o = d['total_points'], d['points_inside_bbox'], d['points_outside_bbox']

        def fail(msg):
            bad.append((key[0], key[1], msg))
        if key in seen:
            fail('duplicate (frame, car_id)')
        seen.add(key)
        if t <= 0:
            fail('a car without points has no row')
        if i + o != t:
            fail('inside + outside != total')
        if d['is_matched']:
            if i < min_points:
                fail('matched with fewer than min_points inside')
            if t > 0 and (round(i / t * 100, 2) != round(d['inside_percentage'], 2)
                          or round(o / t * 100, 2) != round(d['outside_percentage'], 2)):
                fail('percentages do not follow from the counts')
            if 'matched_bbox_id' in d and d['matched_bbox_id'] < 0:
                fail('matched without a box id')
        else:
            if i != 0 or o != t or d['inside_percentage'] != 0.0 or d['outside_percentage'] != 100.0:
                fail('unmatched row must read 0 / total / 0.0 / 100.0')
            if 'matched_bbox_id' in d and d['matched_bbox_id'] != -1:
                fail('unmatched row with a box id')
    return bad


def compare_rows(ours, theirs, columns=COLUMNS[2:]):
    """Join on (frame, car_id).  -> dict with 'only_ours', 'only_theirs' (key lists),
    'differ' ([(key, column, ours, theirs)]) and 'equal' (number of identical rows)."""
    a = {(d['frame'], d['car_id']): d for d in ours}
    b = {(d['frame'], d['car_id']): d for d in theirs}
    differ, equal = [], 0
    for k in sorted(a.keys() & b.keys()):
        same = True
        for c in columns:
            x, y = a[k].get(c), b[k].get(c)
            if isinstance(x, float) or isinstance(y, float):
                ok = abs(float(x) - float(y)) <= 0.005 + 1e-9
            else:
                ok = x == y
            if not ok:
                differ.append((k, c, x, y))
                same = False
        equal += same
    return {'only_ours': sorted(a.keys() - b.keys()), 'only_theirs': sorted(b.keys() - a.keys()),
            'differ': differ, 'equal': equal}


def erosion_effect(before, after):
    """The spreadsheet's own comparison: per (frame, car) present in both tables, the change
    erosion makes.  -> dict of aggregates (means over cars matched in both)."""
    a = {(d['frame'], d['car_id']): d for d in before}
    b = {(d['frame'], d['car_id']): d for d in after}
    keys = sorted(a.keys() & b.keys())
    grown = [k for k in keys if b[k]['total_points'] > a[k]['total_points']]
    both = [k for k in keys if a[k]['is_matched'] and b[k]['is_matched']]
    n = max(len(both), 1)
    return {
        'cars': len(keys), 'matched_in_both': len(both),
        'total_points_grew': grown,                                  # erosion is anti-extensive: must be empty
        'match_flag_changed': [k for k in keys if a[k]['is_matched'] != b[k]['is_matched']],
        'mean_inside_pct_before': sum(a[k]['inside_percentage'] for k in both) / n,
        'mean_inside_pct_after': sum(b[k]['inside_percentage'] for k in both) / n,
        'mean_points_removed': sum(a[k]['total_points'] - b[k]['total_points'] for k in keys) / max(len(keys), 1),
        'mean_outside_removed': sum(a[k]['points_outside_bbox'] - b[k]['points_outside_bbox'] for k in both) / n,
    }


def main(argv=None):
    import argparse
    import json
    ap = argparse.ArgumentParser(description=__doc__.split('\n\n')[0])
    ap.add_argument('ours')
    ap.add_argument('theirs', help='a master CSV, the sheet fixture (with a block column) or the .xlsx itself')
    ap.add_argument('--block', default=None, help="which block of the sheet ('with_erosion' / 'without_erosion')")
    args = ap.parse_args(argv)
    ours = read_master_csv(args.ours)
    if args.theirs.endswith('.xlsx'):
        blocks = sheet_blocks(read_xlsx_sheet(args.theirs))
        theirs = blocks[args.block] if args.block else list(blocks.values())[-1]
    else:
        theirs = read_master_csv(args.theirs)
        if args.block:
            theirs = [d for d in theirs if d.get('block') == args.block]
    rep = compare_rows(ours, theirs)
    rep['inconsistent_rows_ours'] = check_rows(ours)
    print(json.dumps(rep, indent=1, default=str))
    return 0 if not rep['inconsistent_rows_ours'] else 1


if __name__ == '__main__':
    raise SystemExit(main())
