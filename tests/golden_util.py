"""Loads tests/golden/ref_testdata.json (made by tests/golden/make_golden.py from the reference's own
rust/scheduler/testdata fixtures) into oracle-style host columns."""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def load():
    with open(os.path.join(HERE, "golden", "ref_testdata.json")) as f:
        return json.load(f)


def lineitem_columns(part, Utf8):
    """one fixture partition -> dict of numpy / Utf8 columns.  Decimal text -> float64 exactly as CSV parsing does."""
    out = {}
    for k, v in part.items():
        if k in ("l_quantity", "l_extendedprice", "l_discount", "l_tax"):
            out[k] = np.array([float(x) for x in v], dtype=np.float64)
        elif k in ("l_returnflag", "l_linestatus"):
            out[k] = Utf8.from_list(v)
        else:
            out[k] = np.array(v, dtype=np.int32)
    return out


def concat_lineitem(parts, Utf8):
    cols = [lineitem_columns(p, Utf8) for p in parts]
    out = {}
    for k in cols[0]:
        if isinstance(cols[0][k], Utf8):
            out[k] = Utf8.from_list([x for c in cols for x in c[k].to_list()])
        else:
            out[k] = np.concatenate([c[k] for c in cols])
    return out
