"""Sample selection by metadata for `-m/--metadata` + `-f/--filter`.

Host-side mirror of filter_mx() in scripts/partitioning_HC.py:50-170 (off the GPU hot path;
kept so the drop-in CLI accepts the same arguments).  Filter grammar:
'column op value[,value...]' joined by ';', op in == != > >= < <=.  A `date` column is
compared as datetime; `iso_week` (YYYYWww) is translated to the Monday of that ISO week and
compared against `date`; other columns compare as strings for ==/!= and as numbers otherwise.
"""
from datetime import date

import pandas

_CMP = {
    "==": lambda s, v: s == v,
    "!=": lambda s, v: s != v,
    ">": lambda s, v: s > v,
    ">=": lambda s, v: s >= v,
    "<": lambda s, v: s < v,
    "<=": lambda s, v: s <= v,
}


def _add_date_columns(mx):
    if "date" not in mx.columns:
        return mx
    mx = mx.copy()
    pos = mx.columns.get_loc("date")
    mx.insert(pos, "date_original", mx["date"])
    if "year" in mx.columns:
        mx.insert(mx.columns.get_loc("date"), "year_original", mx["year"])
        mx = mx.drop(columns=["year"])
    mx["date"] = pandas.to_datetime(mx["date"], errors="coerce")
    mx.insert(mx.columns.get_loc("date") + 1, "year", mx["date"].dt.year)
    if not {"iso_week_nr", "iso_year", "iso_week"} & set(mx.columns):
        iso = mx["date"].dt.isocalendar()
        year = iso.year.astype(str)
        week = iso.week.astype(str)
        label = year.replace("<NA>", "-") + "W" + week.replace("<NA>", "--").map(lambda x: str(x).zfill(2))
        pos = mx.columns.get_loc("date")
        mx.insert(pos + 2, "iso_year", year)
        mx.insert(pos + 3, "iso_week_nr", week)
        mx.insert(pos + 4, "iso_week", label)
    return mx


def select_samples(meta, filters):
    """meta: string DataFrame (first column = sample id).  Returns ids passing every filter."""
    mx = _add_date_columns(meta)
    sample_column = mx.columns[0]
    for spec in filters.split(";"):
        parts = spec.split(" ")
        col, op = parts[0], parts[1]
        cond = " ".join(parts[2:])
        if "," in cond:
            wanted = cond.split(",")
            if op == "==":
                mx = mx[mx[col].isin(wanted)]
            elif op == "!=":
                mx = mx[~mx[col].isin(wanted)]
        elif col == "date":
            if op in _CMP:
                mx = mx[_CMP[op](mx["date"].astype("datetime64[ns]"), pandas.Timestamp(cond))]
        elif col == "iso_week":
            if "date" not in mx.columns:
                print("\tCannot apply the 'iso_week' filter because column 'date' was not found in the metadata!")
                continue
            year, week = cond.split("W")[0], cond.split("W")[1]
            monday = pandas.to_datetime(date.fromisocalendar(int(year), int(week), 1))
            if op in _CMP:
                mx = mx[_CMP[op](mx["date"].astype("datetime64[ns]"), monday)]
        elif op in ("==", "!="):
            mx = mx[_CMP[op](mx[col], cond)]
        else:
            # scripts/partitioning_HC.py:144-155: the column becomes numeric, the comparison is applied ("<=" is
            # mapped to ">=" there, :152-153; an operator outside the grammar filters nothing), and the column is
            # then rewritten as int strings -- so a later '==' / '!=' / list filter on it sees "5", not "5.0", and a
            # value that is still NaN at that point raises pandas' IntCastingNaNError exactly as in the reference
            mx = mx.copy()
            mx[col] = pandas.to_numeric(mx[col], errors="coerce")
            if op == ">":
                mx = mx[mx[col] > float(cond)]
            elif op == ">=":
                mx = mx[mx[col] >= float(cond)]
            elif op == "<":
                mx = mx[mx[col] < float(cond)]
            elif op == "<=":
                mx = mx[mx[col] >= float(cond)]
            mx = mx.copy()
            mx[col] = mx[col].astype(int).astype(str)
    return mx[sample_column].tolist()
