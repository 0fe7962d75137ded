"""Build container only (needs /root/reference): the oracle against the LIVE reference
source under oracle/ref_shim.py, on cases that have no committed golden vector."""
import numpy as np
import pyarrow as pa
import pytest

pytestmark = pytest.mark.reference

from oracle import datasets_oracle as O  # noqa: E402


def test_repartition_matches_reference():
    from oracle import ref_shim

    ns = ref_shim.load()
    tabs = [pa.table({"v": np.arange(a, b, dtype=np.int64)}) for a, b in ((0, 10), (10, 13), (13, 13), (13, 50))]
    for nb in (1, 3, 7):
        op = ns.shuffle_and_partition.SimpleShufflePartitionOp(None, random_shuffle=False, random_seed=None)
        out, _ = op.execute(ref_shim.make_block_list(ns, tabs), nb, False)
        want = [b["v"].to_pylist() if b.num_columns else [] for b in ref_shim.blocks_of(out)]
        got = [b["v"].to_pylist() if b.num_columns else [] for b in O.repartition_dataset(tabs, nb)]
        assert got == want


def test_sort_and_shuffle_match_reference_on_fresh_seeds():
    import random

    from oracle import ref_shim

    rng = np.random.default_rng(77)
    keys = rng.integers(-30, 30, 5000, dtype=np.int64)
    tabs = [pa.table({"k": keys[i::3], "r": np.arange(i, 5000, 3, dtype=np.int64)}) for i in range(3)]
    for desc in (False, True):
        random.seed(1)
        ref = pa.concat_tables(ref_shim.ref_sort(tabs, "k", descending=desc, push_based=True))
        assert ref.equals(O.global_stable_sort(tabs, "k", desc))
    ref = ref_shim.ref_random_shuffle(tabs, 9)
    mine = O.random_shuffle_dataset(tabs, 9)
    assert all(a.equals(b) for a, b in zip(ref, mine))
