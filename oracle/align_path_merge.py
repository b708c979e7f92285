"""ORACLE (test infrastructure, not product code): alignPathMerge restated with the reference's own loops.

Follows reference src/alignpath.cpp:98-207 statement by statement: AlignSeqMap's two maps (:120-139), linkedColumns'
closure over (alignment, column) <-> (row, residue) links (:141-151) and the merge loop that emits, at every step, the
linked columns of the first alignment (lowest number) whose next column is ready (:153-207).  Pure Python, cubic in the
worst case like the reference: for small cases only.  Pinned by tests/test_guide_merge.py against the reference binary's
merged guide alignment (tests/golden/guide_*.npz, made from oracle/_ref/bin/testspan by tools/make_golden_guide.py).

An alignment is {row: [bool per column]} (AlignPath, alignpath.h)."""


def align_path_merge(alignments):
    n_al = len(alignments)
    seq_len, cols = {}, []
    for al in alignments:                                      # alignpath.cpp:101-117
        if not al:
            cols.append(0)
            continue
        c = {len(p) for p in al.values()}
        assert len(c) == 1, "alignment path is not flush"
        cols.append(c.pop())
        for row, path in al.items():
            ln = sum(1 for b in path if b)
            assert seq_len.setdefault(row, ln) == ln, "incompatible number of residues for row %d" % row
    col_row_pos = [dict() for _ in range(n_al)]                # alignColRowToPos[nAlign][col][row] = pos
    row_pos_col = {}                                           # rowPosAlignToCol[row][pos][nAlign] = col
    for k, al in enumerate(alignments):                        # alignpath.cpp:120-138
        pos = {row: 0 for row in al}
        for c in range(cols[k]):
            for row in sorted(al):
                if al[row][c]:
                    p = pos[row]
                    pos[row] += 1
                    col_row_pos[k].setdefault(c, {})[row] = p
                    row_pos_col.setdefault(row, {}).setdefault(p, {})[k] = c
            assert c in col_row_pos[k], "column %d of alignment %d is empty" % (c, k)

    def linked(k, c):                                          # alignpath.cpp:141-151
        ac, queue = {}, {k: c}
        while len(queue) > len(ac):
            for kk in sorted(queue):
                if kk not in ac:
                    ac[kk] = queue[kk]
                    for row, p in col_row_pos[kk][queue[kk]].items():
                        for k2, c2 in row_pos_col[row][p].items():
                            if k2 in ac:
                                assert ac[k2] == c2, "inconsistent alignments"
                            queue.setdefault(k2, c2)
        return ac

    out = {row: [] for row in seq_len}
    nxt = [0] * n_al
    while True:                                                # alignpath.cpp:160-199
        all_done = none_ready = True
        for k in range(n_al):
            if nxt[k] < cols[k]:
                all_done = False
                lc = linked(k, nxt[k])
                if all(nxt[k2] == c2 for k2, c2 in lc.items()):
                    none_ready = False
                    for path in out.values():
                        path.append(False)
                    for k2, c2 in lc.items():
                        for row, path in alignments[k2].items():
                            if path[c2]:
                                out[row][-1] = True
                        nxt[k2] += 1
                    break
        assert not (none_ready and not all_done), "no alignments ready"
        if all_done:
            return out
