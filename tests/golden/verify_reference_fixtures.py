"""Run in the authoring container only: checks tests/golden/reference_sql_fixtures.json against the
reference's own test files under /root/reference (read-only).  For every case the table rows must
appear as INSERT literals in the cited file and the expected ids/distances must appear, in order,
in one result block ('----' ... blank line) inside the cited line range."""
import json
import os
import re
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def fmt_vec(v):
    return "[" + ", ".join(("%.1f" % x) for x in v) + "]"


def result_blocks(lines, lo, hi):
    blocks, cur, inside = [], [], False
    for ln in lines[lo - 1:hi + 12]:
        s = ln.strip()
        if s == "----":
            inside, cur = True, []
        elif inside and s == "":
            blocks.append(cur)
            inside = False
        elif inside:
            cur.append(s.split("\t"))
    if inside:
        blocks.append(cur)
    return blocks


def main():
    fx = json.load(open(os.path.join(HERE, "reference_sql_fixtures.json")))
    bad = 0
    for c in fx["cases"]:
        path, rng = c["cite"].rsplit(":", 1)
        lo, hi = (int(x) for x in rng.split("-"))
        text = open(os.path.join(REF, path)).read()
        lines = text.split("\n")
        for r in c["rows"]:
            if fmt_vec(r) not in text:
                print("MISSING row", r, "in", path)
                bad += 1
        blocks = result_blocks(lines, lo, hi)
        want = None
        if "expect_ids" in c:
            want = [str(i) for i in c["expect_ids"]]
            got_ok = any([row[0] for row in b] == want for b in blocks)
        elif "expect_row_ids" in c:
            want = [[str(i), "%.6f" % d] for i, d in zip(c["expect_row_ids"], c["expect_distances"])]
            got_ok = any(b == want for b in blocks)
        elif "expect_count" in c:
            want = [[str(c["expect_count"])]]
            got_ok = any(b == want for b in blocks)
        else:
            want = ["%.6f" % d for d in c["expect_distances"]]
            got_ok = any([row[-1] for row in b] == want for b in blocks)
        if not got_ok:
            print("MISMATCH", c["name"], "want", want, "blocks", blocks)
            bad += 1
        if "predicate" in c:
            # the WHERE text must be in the cited query (modulo the d. alias)
            pred = c["predicate"]
            region = "\n".join(lines[lo - 1:hi])
            norm = re.sub(r"\bd\.", "", region)
            if pred not in norm:
                print("PREDICATE not found", pred, "in", c["cite"])
                bad += 1
    print("checked", len(fx["cases"]), "cases;", bad, "problems")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
