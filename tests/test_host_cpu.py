"""CPU tests of the C++ host mirror above the C ABI (csrc/host/): the predicate push-down grammar of
src/lance_optimizer.cpp:204-344 evaluated into the prefilter bitmap, the function->metric mapping of :350-371,
and that the library exports every lbh_* symbol include/lance_b200_host.h declares."""
import ctypes
import json
import os
import re

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
FX = json.load(open(os.path.join(HERE, "golden", "reference_sql_fixtures.json")))


def test_host_symbols_exported(pkg, host):
    text = re.sub(r"/\*.*?\*/", "", open(host.HOST_HEADER).read(), flags=re.S)
    names = sorted(set(re.findall(r"\b(lbh_[a-z0-9_]+)\s*\(", text)))
    L = ctypes.CDLL(pkg.LIB_PATH)
    assert not [n for n in names if not hasattr(L, n)]
    assert sorted(host.HOST_SIGNATURES) == names


def fixture_columns(host, case):
    cols = host.Columns()
    for name, vals in case["columns"].items():
        cols.append(name, vals)
    return cols


PY_PREDICATES = {
    "lang = 'en'": lambda c: [v == "en" for v in c["lang"]],
    "lang = 'es'": lambda c: [v == "es" for v in c["lang"]],
    "score > 20": lambda c: [v > 20 for v in c["score"]],
    "lang IS NOT NULL": lambda c: [v is not None for v in c["lang"]],
    "lang IN ('en', 'fr')": lambda c: [v in ("en", "fr") for v in c["lang"]],
    "NOT (lang = 'en')": lambda c: [v != "en" for v in c["lang"]],
}


@pytest.mark.parametrize("case", [c for c in FX["cases"] if "predicate" in c], ids=lambda c: c["name"])
def test_reference_filter_predicates(host, case):
    """The WHERE clauses of test/sql/lance_optimizer_filter.test, in the string form the optimizer pushes down."""
    cols = fixture_columns(host, case)
    n = len(case["ids"])
    got = cols.evaluate(case["predicate"], n)
    assert got.tolist() == PY_PREDICATES[case["predicate"]](case["columns"])


def test_three_valued_logic_and_grammar(host):
    cols = host.Columns()
    lang = ["en", None, "it's", "fr", "en", None]
    score = [10, 20, None, 40, 50, 60]
    w = [0.5, 1.5, 2.5, None, -1.0, 3.0]
    ok = [True, False, None, True, False, True]
    cols.append("lang", lang)
    cols.append("score", score)
    cols.append("w", w)
    cols.append("ok", ok)
    n = 6
    ev = lambda p: cols.evaluate(p, n).tolist()
    T, F = True, False
    assert ev("lang = 'en'") == [T, F, F, F, T, F]
    assert ev("lang != 'en'") == [F, F, T, T, F, F]              # NULL != x is NULL, not TRUE
    assert ev("NOT (lang = 'en')") == [F, F, T, T, F, F]
    assert ev("lang IS NULL") == [F, T, F, F, F, T]
    assert ev("lang = 'it''s'") == [F, F, T, F, F, F]            # FormatConstant's quote escaping (:146-156)
    assert ev("score >= 20 AND score <= 50") == [F, T, F, T, T, F]  # BETWEEN as emitted (:336-341)
    assert ev("20 < score") == [F, F, F, T, T, T]                # constant on the left (:216-232)
    assert ev("(lang = 'en') OR (score > 40)") == [T, F, F, F, T, T]  # OR children parenthesised (:258-262)
    assert ev("(lang = 'zz') OR (score > 40)") == [F, F, F, F, T, T]  # NULL OR FALSE is NULL
    assert ev("lang = 'en' AND score > 20") == [F, F, F, F, T, F]
    assert ev("lang IN ('en', 'fr')") == [T, F, F, T, T, F]
    assert ev("lang NOT IN ('en', 'fr')") == [F, F, T, F, F, F]
    assert ev("lang NOT IN ('en', NULL)") == [F, F, F, F, F, F]  # x NOT IN (.., NULL) is never TRUE
    assert ev("score IN (10, 60)") == [T, F, F, F, F, T]
    assert ev("w > 1.0") == [F, T, T, F, F, T]
    assert ev("w <= -1") == [F, F, F, F, T, F]
    assert ev("score > 19.5") == [F, T, F, T, T, T]
    assert ev("ok = true") == [T, F, F, T, F, T]
    assert ev("ok = false") == [F, T, F, F, T, F]
    assert ev("NOT (ok = true)") == [F, T, F, F, T, F]
    assert ev("score = NULL") == [F] * 6
    # labels past the stored column values (rows appended without extras) read as NULL
    assert cols.evaluate("score IS NULL", 8).tolist() == [F, F, T, F, F, F, T, T]


@pytest.mark.parametrize("bad", ["", "lang =", "lang = 'en", "nosuch = 1", "lang = 'en' garbage", "(lang = 'en'",
                                 "lang IN ()", "1 = 2", "lang IS 3"])
def test_malformed_predicates_raise(host, bad):
    cols = host.Columns()
    cols.append("lang", ["en"])
    with pytest.raises(host.LanceError, match="lance predicate"):
        cols.evaluate(bad, 1)


def test_distance_function_mapping(host):
    """src/lance_optimizer.cpp:350-371 and the metric match of :512-541 ('dot' == 'ip')."""
    assert host.distance_function_metric("array_distance") == "l2"
    assert host.distance_function_metric("array_cosine_distance") == "cosine"
    assert host.distance_function_metric("array_inner_product") == "dot"
    assert host.distance_function_metric("array_negative_inner_product") is None
    assert host.metric_matches_index("dot", "ip") and host.metric_matches_index("l2", "l2")
    assert not host.metric_matches_index("cosine", "l2")  # lance_optimizer.test:77-99: must not rewrite


def test_random_predicates_against_numpy(host):
    rng = np.random.default_rng(3)
    n = 5000
    score = rng.integers(0, 100, n)
    lang = rng.choice(["en", "fr", "es"], n)
    cols = host.Columns()
    cols.append("score", score.tolist())
    cols.append("lang", lang.tolist())
    assert (cols.evaluate("score > 20", n) == (score > 20)).all()
    assert (cols.evaluate("score < 1", n) == (score < 1)).all()
    assert (cols.evaluate("lang IN ('en', 'es') AND score >= 50", n) == (np.isin(lang, ["en", "es"]) & (score >= 50))).all()
    assert (cols.evaluate("(lang = 'fr') OR (NOT (score < 90))", n) == ((lang == "fr") | ~(score < 90))).all()
