"""The reference's recorded results (master_car_statistics.csv.xlsx) as golden
vectors for the CSV writer and the per-car decision rules, and the comparison tool itself.
CPU only: the fixture was transcribed by oracle/gen_golden_sheet.py from the unmodified workbook."""
import csv
import io
import os
import zipfile

import numpy as np
import pytest

from lidar_object_detection_b200 import compare as C

HERE = os.path.dirname(os.path.abspath(__file__))
SHEET = os.path.join(HERE, 'golden', 'master_car_statistics_sheet.csv')


def blocks():
    rows = C.read_master_csv(SHEET)
    return ([d for d in rows if d['block'] == 'without_erosion'], [d for d in rows if d['block'] == 'with_erosion'])


def test_recorded_rows_satisfy_the_row_invariants():
    before, after = blocks()
    assert len(before) == len(after) == 72
    assert C.check_rows(before) == [] and C.check_rows(after) == []
    frames = sorted({d['frame'] for d in after})
    assert len(frames) == 18 and 570 not in frames and 2717 not in frames      # Appendix B observation


def test_erosion_effect_of_the_recorded_rows():
    before, after = blocks()
    e = C.erosion_effect(before, after)
    assert e['cars'] == 72 and e['total_points_grew'] == [] and e['match_flag_changed'] == []
    assert e['mean_inside_pct_after'] > e['mean_inside_pct_before']             # why the reference erodes at all
    # frame 100, car 0: 2994 -> 2926 points, 85.24 -> 86.33 % inside
    a = {(d['frame'], d['car_id']): d for d in after}[(100, 0)]
    assert (a['total_points'], a['points_inside_bbox'], a['inside_percentage']) == (2926, 2526, 86.33)


def test_csv_writer_reproduces_the_recorded_percentages():
    """lf_format_csv_rows (host C, no device) on the recorded counts must print exactly the recorded
    percentages and match flags: pins the writer (A11) and the unmatched-row convention (A9) to
    results the reference itself produced."""
    from lidar_object_detection_b200 import io as IO
    for rows in blocks():
        table = np.array([[d['frame'], d['car_id'], 0 if d['is_matched'] else -1, d['total_points'],
                           d['points_inside_bbox']] for d in rows], dtype=np.int64)
        text = IO.format_csv_rows(table, 'T', with_header=True).decode()
        got = list(csv.DictReader(io.StringIO(text)))
        assert len(got) == len(rows)
        for g, d in zip(got, rows):
            assert int(g['total_points']) == d['total_points']
            assert int(g['points_outside_bbox']) == d['points_outside_bbox']
            assert float(g['inside_percentage']) == d['inside_percentage'], (d, g)
            assert float(g['outside_percentage']) == d['outside_percentage'], (d, g)
            assert g['is_matched'] == str(d['is_matched'])


def test_decision_rule_on_the_recorded_rows():
    """is_matched <=> best_count >= min_points (cvs_erosion.py:197): every recorded unmatched row has
    inside = 0 and every matched one at least 10 inside; 2098/11 (10 points, none in a box) is the
    sheet's own boundary case."""
    _, after = blocks()
    for d in after:
        assert d['is_matched'] == (d['points_inside_bbox'] >= 10)
    d = {(d['frame'], d['car_id']): d for d in after}[(2098, 11)]
    assert d['total_points'] == 10 and not d['is_matched']


def test_compare_rows_reports_differences():
    before, after = blocks()
    rep = C.compare_rows(after, after)
    assert rep['equal'] == 72 and not rep['differ'] and not rep['only_ours'] and not rep['only_theirs']
    mine = [dict(d) for d in after if d['frame'] != 100]
    mine[0]['total_points'] += 1
    rep = C.compare_rows(mine, after)
    assert [k[0] for k in rep['only_theirs']] == [100] * 5
    assert rep['differ'] == [((mine[0]['frame'], mine[0]['car_id']), 'total_points', mine[0]['total_points'],
                              mine[0]['total_points'] - 1)]
    assert C.check_rows(mine)[0][2] == 'inside + outside != total'


def test_check_rows_flags_every_rule():
    ok = {'frame': 1, 'car_id': 0, 'matched_bbox_id': 2, 'total_points': 40, 'points_inside_bbox': 30,
          'points_outside_bbox': 10, 'inside_percentage': 75.0, 'outside_percentage': 25.0, 'is_matched': True}
    assert C.check_rows([ok]) == []
    for patch, msg in (({'points_inside_bbox': 5, 'points_outside_bbox': 35, 'inside_percentage': 12.5,
                         'outside_percentage': 87.5}, 'fewer than min_points'),
                       ({'inside_percentage': 74.0}, 'percentages'),
                       ({'matched_bbox_id': -1}, 'without a box id'),
                       ({'is_matched': False}, 'unmatched row must read'),
                       ({'total_points': 0, 'points_inside_bbox': 0, 'points_outside_bbox': 0}, 'no row')):
        bad = C.check_rows([dict(ok, **patch)])
        assert bad and any(msg in b[2] for b in bad), (patch, bad)
    assert C.check_rows([ok, ok])[0][2].startswith('duplicate')


def _tiny_xlsx(path):
    sheet = ('<?xml version="1.0"?><worksheet xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main">'
             '<sheetData>'
             '<row r="1"><c r="A1" t="s"><v>1</v></c><c r="J1" t="s"><v>2</v></c></row>'
             '<row r="2">' + ''.join('<c r="%s2" t="s"><v>0</v></c>' % c if c in 'AJ' else
                                     '<c r="%s2" t="inlineStr"><is><t>x</t></is></c>' % c for c in 'ABCDEFGHJKLMNOPQ') +
             '</row>'
             '<row r="3"><c r="A3"><v>7</v></c><c r="B3"><v>0</v></c><c r="C3"><v>40</v></c><c r="D3"><v>30</v></c>'
             '<c r="E3"><v>10</v></c><c r="F3"><v>75</v></c><c r="G3"><v>25</v></c><c r="H3" t="b"><v>1</v></c>'
             '<c r="J3"><v>7</v></c><c r="K3"><v>0</v></c><c r="L3"><v>36</v></c><c r="M3"><v>0</v></c>'
             '<c r="N3"><v>36</v></c><c r="O3"><v>0</v></c><c r="P3"><v>100</v></c><c r="Q3" t="b"><v>0</v></c></row>'
             '</sheetData></worksheet>')
    with zipfile.ZipFile(path, 'w') as z:
        z.writestr('xl/workbook.xml',
                   '<workbook xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main" '
                   'xmlns:r="http://schemas.openxmlformats.org/officeDocument/2006/relationships">'
                   '<sheets><sheet name="other" sheetId="2" r:id="rId2"/>'
                   '<sheet name="master_car_statistics" sheetId="1" r:id="rId1"/></sheets></workbook>')
        z.writestr('xl/_rels/workbook.xml.rels',
                   '<Relationships xmlns="http://schemas.openxmlformats.org/package/2006/relationships">'
                   '<Relationship Id="rId1" Target="worksheets/sheet1.xml"/>'
                   '<Relationship Id="rId2" Target="worksheets/sheet2.xml"/></Relationships>')
        z.writestr('xl/sharedStrings.xml',
                   '<sst xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main">'
                   '<si><t>frame</t></si><si><t>Without Erosion</t></si><si><r><t>Ero</t></r><r><t>sion </t></r></si></sst>')
        z.writestr('xl/worksheets/sheet1.xml', sheet)
        z.writestr('xl/worksheets/sheet2.xml', sheet.replace('<v>40</v>', '<v>41</v>'))


def test_xlsx_reader_without_a_spreadsheet_library(tmp_path):
    p = str(tmp_path / 'book.xlsx')
    _tiny_xlsx(p)
    b = C.sheet_blocks(C.read_xlsx_sheet(p, 'master_car_statistics'))
    assert list(b) == ['Without Erosion', 'Erosion']
    assert b['Without Erosion'] == [{'frame': 7, 'car_id': 0, 'total_points': 40, 'points_inside_bbox': 30,
                                     'points_outside_bbox': 10, 'inside_percentage': 75.0,
                                     'outside_percentage': 25.0, 'is_matched': True}]
    assert b['Erosion'][0]['total_points'] == 36 and b['Erosion'][0]['is_matched'] is False
    assert C.sheet_blocks(C.read_xlsx_sheet(p))['Without Erosion'][0]['total_points'] == 41     # first sheet
    with pytest.raises(KeyError):
        C.read_xlsx_sheet(p, 'nope')


def test_command_line(tmp_path, capsys):
    _, after = blocks()
    ours = tmp_path / 'ours.csv'
    from lidar_object_detection_b200 import io as IO
    table = np.array([[d['frame'], d['car_id'], 0 if d['is_matched'] else -1, d['total_points'],
                       d['points_inside_bbox']] for d in after], dtype=np.int64)
    ours.write_bytes(IO.format_csv_rows(table, '2025-01-01T00:00:00', with_header=True))
    assert C.main([str(ours), SHEET, '--block', 'with_erosion']) == 0
    out = capsys.readouterr().out
    assert '"equal": 72' in out and '"differ": []' in out
