"""Comparison rules of the north star: canonical sort, bit-exact integers / counts / row sets,
<= 1e-9 relative on Float64 sums and averages."""
import math

REL_TOL = 1e-9


def canon(rows, keys):
    return sorted(rows, key=lambda r: tuple(r[k] for k in keys))


def assert_rows_match(got, want, keys, exact=(), approx=()):
    got, want = canon(got, keys), canon(want, keys)
    assert len(got) == len(want), f"{len(got)} rows, expected {len(want)}"
    for g, w in zip(got, want):
        for k in list(keys) + list(exact):
            assert g[k] == w[k], (k, g, w)
        for k in approx:
            assert math.isclose(g[k], w[k], rel_tol=REL_TOL, abs_tol=0.0) or g[k] == w[k], (k, g[k], w[k])
