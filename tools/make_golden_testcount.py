#!/usr/bin/env python3
"""Fixture of the reference's own `testcount` golden (Makefile:317): the inputs of
`historian count -model data/testcount.jukescantor.json -tree data/testcount.nh -recon data/testcount.historian.fa`
and its expected output data/testcount.count.json, flattened to tests/golden/testcount_recon.json.
Run where /root/reference exists."""
import json
import os
import re

REF = "/root/reference/data"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "testcount_recon.json")


def parse_newick(s):
    """-> (names, parent, branch_length) in postorder, root last (tree.cpp:141)."""
    names, parent, bl = [], [], []
    pos = 0

    def node():
        nonlocal pos
        kids = []
        if s[pos] == "(":
            pos += 1
            while True:
                kids.append(node())
                if s[pos] == ",":
                    pos += 1
                    continue
                assert s[pos] == ")"
                pos += 1
                break
        m = re.match(r"([^:,();]*)(?::([0-9.eE+-]+))?", s[pos:])
        pos += m.end()
        names.append(m.group(1)); parent.append(-1); bl.append(float(m.group(2) or 0))
        me = len(names) - 1
        for k in kids:
            parent[k] = me
        return me

    node()
    return names, parent, bl


def main():
    names, parent, bl = parse_newick(open(os.path.join(REF, "testcount.nh")).read().strip())
    seqs, cur = {}, None
    for line in open(os.path.join(REF, "testcount.historian.fa")):
        line = line.strip()
        if line.startswith(">"):
            cur = line[1:].split()[0]; seqs[cur] = ""
        elif cur:
            seqs[cur] += line
    model = json.load(open(os.path.join(REF, "testcount.jukescantor.json")))
    text = open(os.path.join(REF, "testcount.count.json")).read()
    text = re.sub(r'("insTime": [0-9.eE+-]+)\n', r"\1,\n", text)   # the reference's writer omits this comma (model.cpp:946)
    expect = json.loads(text)
    json.dump({"names": names, "parent": parent, "branch_length": bl, "rows": [seqs[n] for n in names], "model": model,
               "expect": expect, "source": "reference data/testcount.{nh,historian.fa,jukescantor.json,count.json}"},
              open(OUT, "w"), indent=1)
    print(OUT)


if __name__ == "__main__":
    main()
