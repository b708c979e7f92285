#!/usr/bin/env python3
"""ORACLE INFRASTRUCTURE: accept oracle/_ref (the unmodified reference compiled
against oracle/gslshim) by running the reference's own golden tests.

This replays the `make test` target list of the reference (Makefile:200-331)
through the same diff-against-golden rule as perl/testexpect.pl, without writing
into /root/reference.  Only runs where /root/reference exists (this container).

    python oracle/run_ref_goldens.py [-v]
"""
import os
import re
import subprocess
import sys

REF = os.environ.get("HISTORIAN_REF", "/root/reference")
BIN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "bin")


def roundfloats(text, n):
    """perl/roundfloats.pl N: round every float to N significant digits."""
    def fix(m):
        x = float(m.group(0))
        if abs(x) < 10.0 ** (-n):
            return "0"
        return "%.*g" % (n, x)
    return re.sub(r"-?\d+\.\d+(?:[eE][-+]?\d+)?", fix, text)


# (program, args, golden, options)
T = []
def t(prog, args, golden, **kw):
    T.append((prog, args.split(), golden, kw))

t("testregex", "/dev/null", "/dev/null")
t("testlogsumexp", "-fast", "data/logsumexp.txt")
t("testlogsumexp", "-slow", "data/logsumexp.txt")
t("testseqio", "data/testaligncount.fa", "data/testaligncount.fa")
t("testseqio", "data/gp120.fa", "data/gp120.fa")
t("testnexus", "data/testnexus.nex", "data/testnexus.nex")
t("teststockholm", "data/cbs.stock", "data/cbs.stock")
t("teststockholm", "data/Lysine.stock", "data/Lysine.stock")
t("testmatexp", "data/testrates.json 1", "data/testrates.probs.json", round=10)
t("testmatexp", "-eigen data/testrates.json 1", "data/testrates.probs.json", round=10)
t("testrateio", "data/testrates.json", "data/testrates.out.json", round=4)
t("testrateio", "data/testrates.out.json", "data/testrates.out.json", round=4)
t("testrateio", "data/testrates.mix2.json", "data/testrates.mix2.out.json", round=4)
t("testrateio", "data/testrates.mix2.out.json", "data/testrates.mix2.out.json", round=4)
t("testmerge", "data/testmerge1.xy.fa data/testmerge1.xz.fa", "data/testmerge1.xyz.fa")
t("testmerge", "data/testmerge1.xy.fa data/testmerge1.ayz.fa", "data/testmerge1.xyaz.fa")
t("testmerge", "data/testmerge1.xz.fa data/testmerge1.ayz.fa", "data/testmerge1.xzay.fa")
t("testmerge", "data/testmerge1.axyz.fa data/testmerge1.xz.fa", "data/testmerge1.axyz.fa")
t("testmerge", "data/testmerge1.xy.fa data/testmerge1.xz.fa data/testmerge1-fail.ayz.fa", "data/empty", fail=True)
t("testseqprofile", "ACGT AAGCT", "data/testseqprofile.aagct.json")
t("testforward", "-all -matrix data/testforward.id100.len2.fa data/testforward.nosub.json 1", "data/testforward.id100.len2.nosub.out")
t("testforward", "-hubs -best data/testforward.len2.fa data/testforward.nosub.json 1", "data/testforward.len2.nosub.best.out")
t("testforward", "-hubs -best data/testforward.len2.fa data/testforward.jukescantor.json 1", "data/testforward.len2.jc.best.out")
t("testforward", "-hubs -best data/testforward.len2-4.fa data/testforward.jukescantor.json .1 .01", "data/testforward.len2-4.xdel.out")
t("testforward", "-hubs -best data/testforward.len2-4.fa data/testforward.jukescantor.json .01 1", "data/testforward.len2-4.yins.out")
t("testforward", "-all 10 data/testforward.len2-4.fa data/testforward.jukescantor.json .1", "data/testforward.len2-4.n10.all.out")
t("testforward", "-hubs 10 data/testforward.len2-4.fa data/testforward.jukescantor.json .1", "data/testforward.len2-4.n10.hubs.out")
t("testnullforward", "data/testforward.nosub.json 1", "data/testnullforward.nosub.out")
t("testbackward", "data/testforward.len2.fa data/testforward.jukescantor.json 1", "data/testbackward.len2.out")
t("testbackward", "data/testforward.len2-4.fa data/testforward.jukescantor.json 1", "data/testbackward.len2-4.out")
t("testtreeio", "data/PF16593.nhx", "data/PF16593.nhx")
t("testtreeio", "data/testnj.out.nh", "data/testnj.out.nh")
t("testtreeio", "data/testreroot.nh C", "data/testreroot.c.nh")
t("testnj", "data/testnj.jukescantor.json data/testnj.fa", "data/testnj.out.nh")
t("testnj", "data/testamino.json data/PF16593.testspan.fa", "data/PF16593.testspan.testnj.nh")
t("testupgma", "data/testnj.jukescantor.json data/testnj.fa", "data/testupgma.out.nh")
t("testupgma", "data/testamino.json data/PF16593.testspan.fa", "data/PF16593.testspan.testupgma.nh")
t("testquickalign", "data/PF16593.pair.fa data/testamino.json 1", "data/testquickalign.out.fa")
t("testsubcount", "data/testrates.json A T 1", "data/testsubcount1.json")
t("testsubcount", "data/testforward.jukescantor.json A T 1", "data/testsubcount2.json")
t("testsubcount", "data/testforward.jukescantor.json A T 1", "data/testsubcount3.json")
t("testsubcount", "data/testrates.mix2.json A T 1", "data/testsubcount.mix2.json")
t("testnumsubcount", "data/testforward.jukescantor.json A T A T .01 4", "data/testnumsubcount1.out")
t("testnumsubcount", "data/testforward.jukescantor.json A T A T 1 4", "data/testnumsubcount2.out")
t("testnumsubcount", "data/testforward.jukescantor.json A T C G 1 4", "data/testnumsubcount3.out")
t("testnumsubcount", "data/testrates.json A T A T 1", "data/testnumsubcount4.out")
t("testaligncount", "-sub data/testnj.jukescantor.json data/testaligncount.fa data/testaligncount.nh", "data/testaligncount.out")
t("testaligncount", "-eigen data/testnj.jukescantor.json data/testaligncount.fa data/testaligncount.nh", "data/testaligncount.out")
t("testaligncount", "-sub data/testcount.jukescantor.json data/testaligncount2.fa data/testcount.nh", "data/testaligncount2.out.json")
t("testsumprod", "data/testnj.jukescantor.json data/testaligncount.fa data/testaligncount.nh", "data/testsumprod.out")
t("testcountio", "data/testcount.count.json", "data/testcount.count.json")
t("historian", "recon -careful -norefine -output fasta -model data/testcount.jukescantor.json -guide data/testcount.fa -tree data/testcount.nh", "data/testcount.historian.fa")
t("historian", "recon -careful -norefine -output fasta -model data/testnj.jukescantor.json -nexus data/testnexus.nex", "data/testnexus.hist.fa")
t("historian", "recon -careful -norefine -output fasta -profsamples 100 -guide data/PF16593.testspan.fa -model data/testamino.json -tree data/PF16593.testspan.testnj.nh -band 10", "data/PF16593.testspan.testnj.historian.fa")
t("historian", "recon -careful -norefine -output fasta -profsamples 100 -guide data/PF16593.testspan.fa -tree data/PF16593.testspan.testnj.nh -model data/testamino.json", "data/PF16593.testspan.testnj.historian.fa")
t("historian", "recon -careful -norefine -output fasta -profsamples 100 -guide data/PF16593.testspan.fa -model data/testamino.json -nj", "data/PF16593.testspan.testnj.historian.fa")
t("historian", "recon -careful -norefine -output fasta -profsamples 100 -seqs data/PF16593.fa -tree data/PF16593.nhx -model data/testamino.json -nj", "data/PF16593.historian.fa")
t("historian", "count -careful -model data/testcount.jukescantor.json -recon data/testcount.fa -tree data/testcount.nh", "data/testcount.out.json")
t("historian", "count -careful -model data/testcount.jukescantor.json -tree data/testcount.nh -recon data/testcount.historian.fa", "data/testcount.count.json")
t("historian", "count -careful -model data/testrates.mix2.json -recon data/testcount.mix2.fa -tree data/testcount.mix2.nh", "data/testcount.mix2.count.json")
t("historian", "sum data/testcount.out.json data/testcount.out.json", "data/testcount.sum.json")
t("historian", "data/zerolen.fa -output fasta", "data/zerolen.aligned.fa")
t("historian", "data/zerolen2.fa -output fasta", "data/zerolen2.aligned.fa")


def run_all(verbose=False):
    results = []
    for prog, args, golden, kw in T:
        cmd = [os.path.join(BIN, prog)] + args
        try:
            p = subprocess.run(cmd, cwd=REF, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
            out, code = p.stdout.decode(errors="replace"), p.returncode
        except subprocess.TimeoutExpired:
            out, code = "", -999
        want = open(os.path.join(REF, golden)).read() if golden != "/dev/null" else ""
        if kw.get("round"):
            out = roundfloats(out, kw["round"])
        if kw.get("fail"):
            ok = code != 0
        else:
            ok = code == 0 and out == want
        results.append((ok, prog, " ".join(args), golden, code))
        if verbose and not ok:
            import difflib
            d = list(difflib.unified_diff(want.splitlines(), out.splitlines(), "golden", "got", lineterm="", n=0))
            print("\n".join(d[:40]))
    return results


if __name__ == "__main__":
    verbose = "-v" in sys.argv
    res = run_all(verbose)
    for ok, prog, args, golden, code in res:
        print("%s  %s %s  -> %s%s" % ("ok    " if ok else "NOT OK", prog, args, golden, "" if ok else "  (exit %d)" % code))
    nbad = sum(1 for r in res if not r[0])
    print("%d/%d reference goldens reproduced by oracle/_ref" % (len(res) - nbad, len(res)))
    sys.exit(1 if nbad else 0)
