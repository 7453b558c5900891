"""CPU oracle: restatement of the reference algorithms for the hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under `handyrl_b200/` imports this package;
only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs do,
and there only as the checker / the timed CPU baseline, never as the product.
Every function cites the reference file:line it follows; the restatements are
pinned against outputs of the imported reference itself (tests/golden/, made
by tools/make_golden.py) because the reference's own tests hold no golden
vectors for this path.
"""
