"""`-m/--metadata` + `-f/--filter` sample selection (reportree_b200/metadata_filter.py) against
expectations derived BY HAND from the reference's own fixture tests/metadata.tsv and the grammar of
filter_mx (scripts/partitioning_HC.py:50-170).  The reference branch itself crashes under the pandas 3
of this image, so it cannot generate the expectations; tests/HC_test.py:16-21
gives the canonical filter string ('country_origin != B;date >= 2023-02-02')."""
import io

import pandas
import pytest

from reportree_b200 import metadata_filter

# /root/reference/tests/metadata.tsv, verbatim (a reproduced reference fixture)
METADATA = """sample\tdate\tcountry origin\tsource isolation
sample1\t2023-03-02\tX Y Z\tA
sample2\t2023-03-03\tA\tB
sample3\t2023-03-04\tX Y Z\tC
sample3_2\t2023-03-05\tA\tD
sample3_3\t2023-03-06\tA\tA
sample4\t2023-03-02\tA\tB
sample5\t2023-03-03\tA\tC
sample6\t2023-03-04\tA\tD
sample7\t2023-03-05\tX Y Z\tA
sample8\t2023-03-06\tA\tB
sample9\t2023-03-02\tA\tC
sample9_2\t2023-03-03\tA\tD
sample10\t2023-03-04\tA\tA
sample11\t2023-03-05\tA\tB
sample12\t2023-03-01\tA\tC
sample13\t2023-03-02\tB\tD
sample14\t2023\tA\tA
sample15\t2023-03-04\tA\tB
sample_reference\t2023-03-05\tA\tC
"""
ALL = [ln.split("\t")[0] for ln in METADATA.splitlines()[1:]]


def _meta():
    """loaded the way the CLI does (scripts/partitioning_HC.py:382-384: dtype=str, blanks -> underscores)"""
    meta = pandas.read_table(io.StringIO(METADATA), dtype=str)
    meta.columns = [c.replace(" ", "_") for c in meta.columns]
    return meta.replace(" ", "_", regex=True)


def _without(*drop):
    return [s for s in ALL if s not in drop]


CASES = [
    # the filter of the reference's own HC test: sample13 is country B; sample14's date "2023" is not a day
    ("country_origin != B;date >= 2023-02-02", _without("sample13", "sample14")),
    ("date == 2023-03-02", ["sample1", "sample4", "sample9", "sample13"]),
    ("date != 2023-03-02", _without("sample1", "sample4", "sample9", "sample13")),   # NaT != x is True: sample14 stays
    ("date < 2023-03-02", ["sample12"]),
    ("date <= 2023-03-02", ["sample1", "sample4", "sample9", "sample12", "sample13"]),
    ("date > 2023-03-05", ["sample3_3", "sample8"]),
    # iso_week is translated to the Monday of that ISO week and compared with `date`: 2023W10 starts on 2023-03-06
    ("iso_week >= 2023W10", ["sample3_3", "sample8"]),
    ("iso_week == 2023W10", ["sample3_3", "sample8"]),
    ("iso_week == 2023W09", []),                                  # Monday 2023-02-27: nobody was sampled that day
    ("iso_week < 2023W10", _without("sample3_3", "sample8", "sample14")),
    ("country_origin == X_Y_Z", ["sample1", "sample3", "sample7"]),
    ("source_isolation == A,B", ["sample1", "sample2", "sample3_3", "sample4", "sample7", "sample8", "sample10", "sample11",
                                 "sample14", "sample15"]),
    ("source_isolation != A,B", ["sample3", "sample3_2", "sample5", "sample6", "sample9", "sample9_2", "sample12", "sample13",
                                 "sample_reference"]),
    # numeric comparison on the derived `year` column; NaN (sample14) compares False
    ("year > 2022", _without("sample14")),
    ("year >= 2024", []),
    ("year < 2024", _without("sample14")),
    ("year <= 2024", []),                                         # the reference maps "<=" to ">=" (:152-153)
    # after a numeric filter the column holds int strings ("2023", not "2023.0"): a later string filter matches them
    ("year > 2022;year == 2023", _without("sample14")),
    ("year > 2022;year != 2023", []),
    ("year > 2022;year == 2023,2024", _without("sample14")),
    ("year > 2022;source_isolation == D;date >= 2023-03-04", ["sample3_2", "sample6"]),
    # operators outside the grammar filter nothing on list / date conditions
    ("source_isolation ~ A,B", ALL),
    ("date ~ 2023-03-02", ALL),
]


@pytest.mark.parametrize("flt,expected", CASES, ids=[c[0] for c in CASES])
def test_select_samples(flt, expected):
    assert metadata_filter.select_samples(_meta(), flt) == expected


def test_numeric_filter_on_text_column_raises_like_the_reference():
    """to_numeric(coerce) leaves NaN, an operator outside the grammar removes nothing, astype(int) then raises
    (pandas IntCastingNaNError, a ValueError) -- not a KeyError from an operator table."""
    with pytest.raises(ValueError):
        metadata_filter.select_samples(_meta(), "source_isolation ~ A")


def test_iso_week_without_date_column_is_skipped(capsys):
    meta = _meta().drop(columns=["date"])
    assert metadata_filter.select_samples(meta, "iso_week >= 2023W10") == ALL
    assert "Cannot apply the 'iso_week' filter" in capsys.readouterr().out


def test_derived_columns_follow_the_reference_layout():
    mx = metadata_filter._add_date_columns(_meta())
    assert list(mx.columns) == ["sample", "date_original", "date", "year", "iso_year", "iso_week_nr", "iso_week",
                                "country_origin", "source_isolation"]
    row = mx[mx["sample"] == "sample8"].iloc[0]
    assert (row["date_original"], row["iso_week"], row["iso_year"], row["iso_week_nr"]) == ("2023-03-06", "2023W10", "2023", "10")
