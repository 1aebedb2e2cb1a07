"""Shared test helpers: run a query set through an implementation-neutral interface and compare."""
import numpy as np


def unpack_paths(off, ln, valid, vals):
    out = []
    for i in range(len(valid)):
        if not valid[i]:
            out.append(None)
        else:
            out.append([int(x) for x in vals[int(off[i]):int(off[i]) + int(ln[i])]])
    return out


def lengths_list(out, valid):
    return [int(out[i]) if valid[i] else None for i in range(len(valid))]


def assert_paths_equal(a, b, what=""):
    """a, b = (off, len, valid, vals) tuples; byte-for-byte equality incl. offsets."""
    for k, name in enumerate(("offset", "length", "valid", "values")):
        x, y = np.asarray(a[k]), np.asarray(b[k])
        if x.shape != y.shape or not np.array_equal(x, y):
            bad = np.nonzero(x[:min(len(x), len(y))] != y[:min(len(x), len(y))])[0]
            raise AssertionError("%s: %s differ (shapes %s %s, first bad idx %s)" %
                                 (what, name, x.shape, y.shape, bad[:5]))


def check_path_valid(V, v, e, src, dst, path):
    """path is a real walk src->dst in the CSR."""
    assert path[0] == src and path[-1] == dst
    for a, b in zip(path[:-1], path[1:]):
        assert b in e[v[a]:v[a + 1]], "no edge %d->%d" % (a, b)
