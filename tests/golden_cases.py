"""Hand-checked golden graphs G1-G11.  Each case: V, edge rows (src,dst) in
TABLE ORDER (= the stable key), query rows, and the expected answers worked out by hand.
lengths: None = SQL NULL.  paths: None = NULL, list = vertex-only path (mode 0);
paths_e = interleaved [v,e,v,...] with edge id = input row number (mode 1)."""

CASES = {
    # G1 path graph 0->1->2->3
    "G1_path": dict(V=4, edges=[(0, 1), (1, 2), (2, 3)],
                    q=[(0, 3), (0, 1), (3, 0), (1, 3), (2, 2)],
                    lengths=[3, 1, None, 2, 0], reach=[1, 1, 0, 1, 1],
                    paths=[[0, 1, 2, 3], [0, 1], None, [1, 2, 3], [2]],
                    paths_e=[[0, 0, 1, 1, 2, 2, 3], [0, 0, 1], None, [1, 1, 2, 2, 3], [2]]),
    # G2 diamond 0->{1,2}->3: tie-break parent(3) = 1 (smaller frontier vertex id)
    "G2_diamond": dict(V=4, edges=[(0, 1), (0, 2), (1, 3), (2, 3)],
                       q=[(0, 3)], lengths=[2], reach=[1], paths=[[0, 1, 3]], paths_e=[[0, 0, 1, 2, 3]]),
    # G3 diamond, table order reversed: 2's edge row comes first, parent(3) must STILL be 1
    # (frontier vertices are scanned in ascending id, not in discovery or row order)
    "G3_diamond_swapped": dict(V=4, edges=[(2, 3), (0, 2), (0, 1), (1, 3)],
                               q=[(0, 3)], lengths=[2], reach=[1], paths=[[0, 1, 3]],
                               paths_e=[[0, 2, 1, 3, 3]]),
    # G4 multi-edges 0->1 (rows 0,1) and 1->2 (rows 2,3): lowest CSR offset = earliest row wins
    "G4_multiedge": dict(V=3, edges=[(0, 1), (0, 1), (1, 2), (1, 2)],
                         q=[(0, 2)], lengths=[2], reach=[1], paths=[[0, 1, 2]], paths_e=[[0, 0, 1, 2, 2]]),
    # G5 cycle through the source (R4: seen has no source bit; source may be re-discovered)
    "G5_cycle": dict(V=4, edges=[(0, 1), (1, 2), (2, 0), (2, 3)],
                     q=[(0, 3), (1, 0), (0, 2), (3, 0)],
                     lengths=[3, 2, 2, None], reach=[1, 1, 1, 0],
                     paths=[[0, 1, 2, 3], [1, 2, 0], [0, 1, 2], None],
                     paths_e=[[0, 0, 1, 1, 2, 3, 3], [1, 1, 2, 2, 0], [0, 0, 1, 1, 2], None]),
    # G6 self-loop on the source and on an inner vertex
    "G6_selfloop": dict(V=3, edges=[(0, 0), (0, 1), (1, 1), (1, 2)],
                        q=[(0, 2), (0, 0), (1, 0)], lengths=[2, 0, None], reach=[1, 1, 0],
                        paths=[[0, 1, 2], [0], None], paths_e=[[0, 1, 1, 3, 2], [0], None]),
    # G7 unreachable + isolated vertices (4,5 isolated)
    "G7_isolated": dict(V=6, edges=[(0, 1), (2, 3)],
                        q=[(0, 3), (4, 5), (0, 1), (5, 5), (2, 3)],
                        lengths=[None, None, 1, 0, 1], reach=[0, 0, 1, 1, 1],
                        paths=[None, None, [0, 1], [5], [2, 3]], paths_e=[None, None, [0, 0, 1], [5], [2, 1, 3]]),
    # G11 degree-0 vertex at id 0 and at id V-1
    "G11_zero_ends": dict(V=5, edges=[(1, 2), (2, 3), (1, 3)],
                          q=[(1, 3), (0, 4), (0, 1), (3, 4), (2, 3)],
                          lengths=[1, None, None, None, 1], reach=[1, 0, 0, 0, 1],
                          paths=[[1, 3], None, None, None, [2, 3]],
                          paths_e=[[1, 2, 3], None, None, None, [2, 1, 3]]),
}

# expected CSR arrays (reference layout v[V+2], e[E]) for two of the cases, by hand
CSR_EXPECT = {
    "G2_diamond": dict(v=[0, 2, 3, 4, 4, 4], e=[1, 2, 3, 3], eid=[0, 1, 2, 3]),
    "G3_diamond_swapped": dict(v=[0, 2, 3, 4, 4, 4], e=[2, 1, 3, 3], eid=[1, 2, 3, 0]),
    "G4_multiedge": dict(v=[0, 2, 4, 4, 4], e=[1, 1, 2, 2], eid=[0, 1, 2, 3]),
    "G11_zero_ends": dict(v=[0, 0, 2, 3, 3, 3, 3], e=[2, 3, 3], eid=[0, 2, 1]),
}
