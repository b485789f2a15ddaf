"""Minimal .xlsx writer / reader (one sheet of numbers under a header row) for the two spreadsheets the reference scripts produce
around the hot path: the per-epoch training record (models/CvT(Par).py:492-494: `pd.DataFrame(history.history)` + an `epoch` column,
`to_excel(index=False)`) and the prediction / metric sheet (models/CvT_test(Par).py:440-477) that tools/Compare_r.py:21-24 and
tools/Compare_plot.py:13-26 read back with pandas.  pandas needs openpyxl for both directions; it is not installed here, so the
container (an OOXML zip: workbook + one worksheet with inline strings) is written directly.  Missing values are empty cells, which
pandas reads as NaN - exactly what `pd.concat([df, metrics_df], axis=1).to_excel` leaves below the single metrics row.
"""
from __future__ import annotations

import math
import re
import zipfile
from collections import OrderedDict
from typing import Dict, List, Sequence
from xml.sax.saxutils import escape

_CT = ('<?xml version="1.0" encoding="UTF-8" standalone="yes"?>'
       '<Types xmlns="http://schemas.openxmlformats.org/package/2006/content-types">'
       '<Default Extension="rels" ContentType="application/vnd.openxmlformats-package.relationships+xml"/>'
       '<Default Extension="xml" ContentType="application/xml"/>'
       '<Override PartName="/xl/workbook.xml" ContentType="application/vnd.openxmlformats-officedocument.spreadsheetml.sheet.main+xml"/>'
       '<Override PartName="/xl/worksheets/sheet1.xml" ContentType="application/vnd.openxmlformats-officedocument.spreadsheetml.worksheet+xml"/>'
       '</Types>')
_RELS = ('<?xml version="1.0" encoding="UTF-8" standalone="yes"?>'
         '<Relationships xmlns="http://schemas.openxmlformats.org/package/2006/relationships">'
         '<Relationship Id="rId1" Type="http://schemas.openxmlformats.org/officeDocument/2006/relationships/officeDocument" Target="xl/workbook.xml"/>'
         '</Relationships>')
_WB = ('<?xml version="1.0" encoding="UTF-8" standalone="yes"?>'
       '<workbook xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main" '
       'xmlns:r="http://schemas.openxmlformats.org/officeDocument/2006/relationships">'
       '<sheets><sheet name="Sheet1" sheetId="1" r:id="rId1"/></sheets></workbook>')
_WB_RELS = ('<?xml version="1.0" encoding="UTF-8" standalone="yes"?>'
            '<Relationships xmlns="http://schemas.openxmlformats.org/package/2006/relationships">'
            '<Relationship Id="rId1" Type="http://schemas.openxmlformats.org/officeDocument/2006/relationships/worksheet" Target="worksheets/sheet1.xml"/>'
            '</Relationships>')


def _col(i: int) -> str:
    s = ""
    i += 1
    while i:
        i, r = divmod(i - 1, 26)
        s = chr(65 + r) + s
    return s


def write_xlsx(path: str, columns: "Dict[str, Sequence]"):
    """columns: ordered {header: values}; columns may have different lengths (short ones leave empty cells, read back as NaN)."""
    names = list(columns)
    nrows = max([len(columns[n]) for n in names] + [0])
    rows = ['<row r="1">' + "".join('<c r="%s1" t="inlineStr"><is><t>%s</t></is></c>' % (_col(j), escape(str(n))) for j, n in enumerate(names)) + "</row>"]
    for i in range(nrows):
        cells = []
        for j, n in enumerate(names):
            col = columns[n]
            if i >= len(col) or col[i] is None:
                continue
            v = col[i]
            if isinstance(v, str):
                cells.append('<c r="%s%d" t="inlineStr"><is><t>%s</t></is></c>' % (_col(j), i + 2, escape(v)))
                continue
            v = float(v)
            if math.isnan(v) or math.isinf(v):
                continue
            cells.append('<c r="%s%d"><v>%s</v></c>' % (_col(j), i + 2, repr(int(v)) if v == int(v) and abs(v) < 1e15 else repr(v)))
        rows.append('<row r="%d">%s</row>' % (i + 2, "".join(cells)))
    sheet = ('<?xml version="1.0" encoding="UTF-8" standalone="yes"?>'
             '<worksheet xmlns="http://schemas.openxmlformats.org/spreadsheetml/2006/main"><sheetData>' + "".join(rows) + "</sheetData></worksheet>")
    with zipfile.ZipFile(path, "w", zipfile.ZIP_DEFLATED) as z:
        z.writestr("[Content_Types].xml", _CT)
        z.writestr("_rels/.rels", _RELS)
        z.writestr("xl/workbook.xml", _WB)
        z.writestr("xl/_rels/workbook.xml.rels", _WB_RELS)
        z.writestr("xl/worksheets/sheet1.xml", sheet)


def read_xlsx(path: str) -> "OrderedDict[str, List]":
    """Reads back what write_xlsx wrote (inline strings / plain numbers; shared strings are resolved too): {header: column with None for
    empty cells}."""
    with zipfile.ZipFile(path) as z:
        sheet = z.read("xl/worksheets/sheet1.xml").decode("utf8")
        shared = []
        if "xl/sharedStrings.xml" in z.namelist():
            shared = re.findall(r"<si>.*?<t[^>]*>(.*?)</t>.*?</si>", z.read("xl/sharedStrings.xml").decode("utf8"), re.S)
    table: Dict[int, Dict[int, object]] = {}
    for m in re.finditer(r'<c r="([A-Z]+)(\d+)"([^>]*?)(?:/>|>(.*?)</c>)', sheet, re.S):
        colname, row, attrs, body = m.group(1), int(m.group(2)), m.group(3), m.group(4) or ""
        ci = 0
        for ch in colname:
            ci = ci * 26 + (ord(ch) - 64)
        ci -= 1
        if 't="inlineStr"' in attrs:
            t = re.search(r"<t[^>]*>(.*?)</t>", body, re.S)
            val: object = t.group(1) if t else ""
        else:
            v = re.search(r"<v>(.*?)</v>", body, re.S)
            if not v:
                continue
            val = shared[int(v.group(1))] if 't="s"' in attrs else float(v.group(1))
        table.setdefault(row, {})[ci] = val
    if 1 not in table:
        return OrderedDict()
    ncol = max(table[1]) + 1
    nrow = max(table)
    out: "OrderedDict[str, List]" = OrderedDict()
    for j in range(ncol):
        out[str(table[1].get(j, "Unnamed: %d" % j))] = [table.get(r, {}).get(j) for r in range(2, nrow + 1)]
    return out
