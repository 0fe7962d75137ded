"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz from the reference.

Runs the reference's own hot-path source (sort_impl, _ShufflePartitionOp,
_GroupbyOp, ArrowBlockAccessor.*) from /root/reference under oracle/ref_shim.py
on small seeded inputs and stores inputs + outputs.  Only runnable in the build
container (the GPU box has no /root/reference); the produced fixtures are
committed and are what the GPU-side tests read.

    python -m oracle.gen_golden
"""
import os
import random

import numpy as np
import pyarrow as pa

from oracle import datasets_oracle as O
from oracle import ref_shim

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _split(n, nblocks):
    cuts = np.linspace(0, n, nblocks + 1).astype(int)
    return list(zip(cuts[:-1], cuts[1:]))


def gen_sort_int64():
    n, nb = 20000, 4
    rng = np.random.default_rng(0)
    out = {}
    for name, keys in (
        ("unique", rng.integers(-(2**63), 2**63 - 1, n, dtype=np.int64)),
        ("ties", rng.integers(-500, 500, n, dtype=np.int64)),
    ):
        tabs = [
            pa.table({"value": np.arange(a, b, dtype=np.int64), "key": keys[a:b]})
            for a, b in _split(n, nb)
        ]
        out[f"{name}_keys"] = keys
        for desc in (False, True):
            res = []
            for push in (False, True):
                random.seed(7)
                blocks = ref_shim.ref_sort(tabs, "key", descending=desc, push_based=push)
                assert len(blocks) == nb
                cat = pa.concat_tables([b for b in blocks if b.num_rows])
                res.append(cat["value"].to_numpy())
                # a key value never spans two output blocks (sort.py:118-119)
                seen = set()
                for b in blocks:
                    ks = set(b["key"].to_pylist()) if b.num_rows else set()
                    assert not (ks & seen)
                    seen |= ks
            assert np.array_equal(res[0], res[1]), "plans disagree"
            out[f"{name}_{'desc' if desc else 'asc'}_value_order"] = res[0]
    out["block_cuts"] = np.array(_split(n, nb))
    np.savez_compressed(os.path.join(OUT, "sort_int64.npz"), **out)


def gen_sort_graysort():
    n, nb = 8000, 4
    out = {}
    for name, entropy in (("uniform", 10), ("ties", 1)):
        t = O.graysort_table(n, seed=11)
        keys = O.fsb_to_numpy(t["key"]).copy()
        if entropy < 10:
            keys[:, : 10 - entropy] = 7  # only the last byte varies -> ~31 rows per key
            keys[:, -1] &= 0x3F
            t = pa.table({"key": O.fsb_from_numpy(keys), "payload": t["payload"]})
        tabs = [t.slice(a, b - a) for a, b in _split(n, nb)]
        out[f"{name}_keys"] = keys
        for desc in (False, True):
            random.seed(3)
            blocks = ref_shim.ref_sort(tabs, "key", descending=desc)
            cat = pa.concat_tables([b for b in blocks if b.num_rows])
            rowid = O.fsb_to_numpy(cat["payload"])[:, :8].copy().view("<u8").ravel()
            out[f"{name}_{'desc' if desc else 'asc'}_rowid_order"] = rowid
    out["block_cuts"] = np.array(_split(n, nb))
    np.savez_compressed(os.path.join(OUT, "sort_graysort.npz"), **out)


def gen_partition():
    ns = ref_shim.load()
    rng = np.random.default_rng(5)
    n = 5000
    keys = rng.integers(0, 100, n, dtype=np.int64)
    t = pa.table({"key": keys, "value": np.arange(n, dtype=np.int64)})
    out = {"keys": keys}
    asc_b = [10, 10, 37, 99, 150]
    desc_b = [150, 99, 37, 10, 10]
    for desc, bounds in ((False, asc_b), (True, desc_b)):
        key = [("key", "descending" if desc else "ascending")]
        parts = ns.BlockAccessor.for_block(t).sort_and_partition(bounds, key, desc)
        tag = "desc" if desc else "asc"
        out[f"{tag}_boundaries"] = np.array(bounds, dtype=np.int64)
        out[f"{tag}_sizes"] = np.array([p.num_rows for p in parts])
        out[f"{tag}_value_order"] = pa.concat_tables(parts)["value"].to_numpy()
    # merge_sorted_blocks on mapper-ordered sorted blocks
    blocks = [
        ns.BlockAccessor.for_block(t.slice(a, b - a)).sort_and_partition(
            [], [("key", "ascending")], False
        )[0]
        for a, b in _split(n, 3)
    ]
    merged, meta = ns.arrow_block.ArrowBlockAccessor.merge_sorted_blocks(
        blocks, [("key", "ascending")], False
    )
    out["merge_value_order"] = merged["value"].to_numpy()
    out["merge_cuts"] = np.array(_split(n, 3))
    assert meta.num_rows == n
    np.savez_compressed(os.path.join(OUT, "partition.npz"), **out)


def gen_boundaries():
    ns = ref_shim.load()
    out = {}
    rng = np.random.default_rng(9)
    # the deterministic half: np.sort + nearest-rank quantiles (sort.py:110-115)
    cases = []
    for nsamp, r in ((80, 8), (10, 4), (7, 3), (40, 5), (1, 2), (2, 3), (33, 16), (100, 1)):
        samples = rng.integers(-1000, 1000, nsamp, dtype=np.int64)
        s = np.sort(samples)
        import warnings

        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            ret = [np.quantile(s, q, interpolation="nearest") for q in np.linspace(0, 1, r)][1:]
        out[f"q_{nsamp}_{r}_samples"] = samples
        out[f"q_{nsamp}_{r}_bounds"] = np.array(ret, dtype=np.int64)
        cases.append((nsamp, r))
    out["q_cases"] = np.array(cases)
    # the sampled half, replayed with random.seed
    n, nb = 4000, 4
    keys = rng.integers(0, 10**6, n, dtype=np.int64)
    tabs = [pa.table({"key": keys[a:b]}) for a, b in _split(n, nb)]
    bl = ref_shim.make_block_list(ns, tabs)
    random.seed(1234)
    b = ns.sort.sample_boundaries(bl.get_blocks(), [("key", "ascending")], nb)
    out["sampled_keys"] = keys
    out["sampled_cuts"] = np.array(_split(n, nb))
    out["sampled_bounds"] = np.array(b, dtype=np.int64)
    # empty dataset -> [None]*(R-1)  (sort.py:100-103)
    ebl = ref_shim.make_block_list(ns, [pa.table({"key": pa.array([], pa.int64())})] * 3)
    eb = ns.sort.sample_boundaries(ebl.get_blocks(), [("key", "ascending")], 3)
    assert eb == [None, None]
    np.savez_compressed(os.path.join(OUT, "boundaries.npz"), **out)


def gen_shuffle():
    n, nb = 4000, 4
    tabs = [pa.table({"value": np.arange(a, b, dtype=np.int64)}) for a, b in _split(n, nb)]
    out = {"block_cuts": np.array(_split(n, nb))}
    for seed in (42, 43):
        blocks = ref_shim.ref_random_shuffle(tabs, seed)
        again = ref_shim.ref_random_shuffle(tabs, seed)
        assert all(a.equals(b) for a, b in zip(blocks, again))
        out[f"seed{seed}_sizes"] = np.array([b.num_rows for b in blocks])
        out[f"seed{seed}_values"] = pa.concat_tables(blocks)["value"].to_numpy()
    # num_blocks != input blocks, ragged (ceil slicing: shuffle_and_partition.py:56-59)
    tabs2 = [pa.table({"value": np.arange(a, b, dtype=np.int64)}) for a, b in ((0, 10), (10, 13), (13, 13), (13, 50))]
    blocks = ref_shim.ref_random_shuffle(tabs2, 5, num_blocks=3)
    out["ragged_cuts"] = np.array([(0, 10), (10, 13), (13, 13), (13, 50)])
    out["ragged_sizes"] = np.array([b.num_rows if b.num_columns else 0 for b in blocks])
    out["ragged_values"] = np.concatenate(
        [b["value"].to_numpy() for b in blocks if b.num_columns and b.num_rows]
    )
    np.savez_compressed(os.path.join(OUT, "shuffle.npz"), **out)


def gen_groupby():
    ns = ref_shim.load()
    A = ns.aggregate
    n, nb, g = 6000, 4, 150
    rng = np.random.default_rng(21)
    k = rng.integers(0, g, n, dtype=np.int64)
    vf = rng.random(n)
    vi = rng.integers(-1000, 1000, n, dtype=np.int64)
    tabs = [pa.table({"k": k[a:b], "vf": vf[a:b], "vi": vi[a:b]}) for a, b in _split(n, nb)]
    out = {"k": k, "vf": vf, "vi": vi, "block_cuts": np.array(_split(n, nb))}
    aggs = (A.Sum("vf"), A.Mean("vf"), A.Count(), A.Sum("vi"), A.Min("vi"), A.Max("vf"), A.Mean("vi"))
    res = []
    for push in (False, True):
        random.seed(99)
        blocks = ref_shim.ref_groupby(tabs, "k", aggs, push_based=push)
        cat = pa.concat_tables([b for b in blocks if b.num_rows])
        res.append(cat)
    assert res[0].schema == res[1].schema
    for name in res[0].column_names:
        a, b = res[0][name].to_numpy(), res[1][name].to_numpy()
        assert np.allclose(a, b, rtol=1e-12), name
        out[f"out::{name}"] = a
    out["out_names"] = np.array(res[0].column_names)
    out["out_types"] = np.array([str(f.type) for f in res[0].schema])
    # name-conflict munging (arrow_block.py:469-471; test_dataset.py:3043-3047)
    random.seed(99)
    blocks = ref_shim.ref_groupby(tabs, "k", (A.Sum("vi"), A.Sum("vi"), A.Count(), A.Count()))
    out["conflict_names"] = np.array(pa.concat_tables(blocks).column_names)
    np.savez_compressed(os.path.join(OUT, "groupby.npz"), **out)


def gen_nulls():
    """Null policy (null_aggregate.py) and nulls-last sort, recorded from the reference."""
    ns = ref_shim.load()
    A = ns.aggregate
    n, nb, g = 3000, 3, 40
    rng = np.random.default_rng(33)
    k = rng.integers(0, g, n, dtype=np.int64)
    vf = rng.random(n)
    vi = rng.integers(-50, 50, n, dtype=np.int64)
    mf = rng.random(n) < 0.15          # nulls sprinkled in
    mi = rng.random(n) < 0.15
    mf[k == 3] = True                  # group 3: all null in vf
    mi[k == 5] = False                 # group 5: no nulls in vi
    tabs = [
        pa.table({"k": k[a:b], "vf": pa.array(vf[a:b], mask=mf[a:b]), "vi": pa.array(vi[a:b], mask=mi[a:b])})
        for a, b in _split(n, nb)
    ]
    out = {"k": k, "vf": vf, "vi": vi, "mf": mf, "mi": mi, "block_cuts": np.array(_split(n, nb))}
    for ignore in (True, False):
        aggs = (A.Sum("vf", ignore), A.Mean("vf", ignore), A.Min("vi", ignore), A.Max("vi", ignore),
                A.Sum("vi", ignore), A.Mean("vi", ignore), A.Count(), A.AbsMax("vi", ignore),
                A.Std("vf", ignore_nulls=ignore), A.Std("vi", ddof=0, ignore_nulls=ignore))
        random.seed(5)
        blocks = ref_shim.ref_groupby(tabs, "k", aggs)
        cat = pa.concat_tables([b for b in blocks if b.num_rows], promote=True)
        tag = "ign" if ignore else "strict"
        out[f"{tag}_names"] = np.array(cat.column_names)
        for name in cat.column_names:
            col = cat[name].combine_chunks()
            valid = np.array([v is not None for v in col.to_pylist()])
            vals = np.array([0 if v is None else v for v in col.to_pylist()], dtype=np.float64)
            out[f"{tag}::{name}::valid"] = valid
            out[f"{tag}::{name}::values"] = vals
    # sort with null keys: nulls last in both directions, input order among them
    ks = pa.array(vi, mask=mi)
    st = [pa.table({"key": ks[a:b], "row": np.arange(a, b, dtype=np.int64)}) for a, b in _split(n, nb)]
    for desc in (False, True):
        random.seed(6)
        blocks = ref_shim.ref_sort(st, "key", descending=desc)
        cat = pa.concat_tables([b for b in blocks if b.num_rows])
        out[f"sort_{'desc' if desc else 'asc'}_row_order"] = cat["row"].to_numpy()
    np.savez_compressed(os.path.join(OUT, "nulls.npz"), **out)


def _pack_rows(rows):
    """list of bytes/None -> (offsets int64, data uint8, valid bool) -- npz holds no objects."""
    lens = [0 if r is None else len(r) for r in rows]
    offs = np.zeros(len(rows) + 1, dtype=np.int64)
    np.cumsum(lens, out=offs[1:])
    data = np.frombuffer(b"".join(b"" if r is None else r for r in rows), dtype=np.uint8)
    return offs, data, np.array([r is not None for r in rows])


def gen_varlen():
    """string / binary keys (variable length) through the reference's sort and groupby:
    lengths 0..19 (so up to three 8-byte chunks), shared prefixes, embedded and trailing NUL
    bytes ("ab" vs "ab\0"), duplicates, and nulls."""
    ns = ref_shim.load()
    A = ns.aggregate
    n, nb = 4000, 4
    rng = np.random.default_rng(77)
    alphabet = np.frombuffer(b"ab\x00\xffz", dtype=np.uint8)
    rows = []
    for i in range(n):
        ln = int(rng.integers(0, 20))
        r = alphabet[rng.integers(0, len(alphabet), ln)].tobytes()
        if rng.random() < 0.3:  # long shared prefix: decided in the 2nd or 3rd chunk
            r = b"prefix--prefix--" + r[:3]
        rows.append(r)
    for i in range(0, n, 7):  # duplicates, to see stability
        rows[i] = rows[(i * 13) % n]
    mask = rng.random(n) < 0.05
    out = {"block_cuts": np.array(_split(n, nb))}
    out["bin_offsets"], out["bin_data"], _ = _pack_rows(rows)
    out["null_mask"] = mask
    v = rng.integers(-100, 100, n, dtype=np.int64)
    out["v"] = v
    # (with null keys the reference's sample_boundaries fails in np.sort -- None < bytes --
    # so only the null-free case can be recorded; the null case is checked against pyarrow's
    # sort_indices, the primitive the reference sorts with, in the GPU tests)
    for tag, with_nulls in (("plain", False),):
        keys = pa.array([None if (with_nulls and m) else r for r, m in zip(rows, mask)], pa.binary())
        tabs = [pa.table({"key": keys[a:b], "row": np.arange(a, b, dtype=np.int64)}) for a, b in _split(n, nb)]
        for desc in (False, True):
            random.seed(8)
            blocks = ref_shim.ref_sort(tabs, "key", descending=desc)
            cat = pa.concat_tables([b for b in blocks if b.num_rows])
            out[f"{tag}_{'desc' if desc else 'asc'}_row_order"] = cat["row"].to_numpy()
    # ASCII strings as pa.string(): same order as their UTF-8 bytes
    words = ["".join(chr(97 + int(c)) for c in rng.integers(0, 4, int(rng.integers(1, 12)))) for _ in range(n)]
    out["str_offsets"], out["str_data"], _ = _pack_rows([w.encode() for w in words])
    tabs = [pa.table({"key": pa.array(words[a:b], pa.string()), "row": np.arange(a, b, dtype=np.int64),
                      "v": v[a:b]}) for a, b in _split(n, nb)]
    random.seed(9)
    blocks = ref_shim.ref_sort(tabs, "key", descending=False)
    out["str_asc_row_order"] = pa.concat_tables([b for b in blocks if b.num_rows])["row"].to_numpy()
    # groupby on the string key
    random.seed(10)
    blocks = ref_shim.ref_groupby(tabs, "key", (A.Count(), A.Sum("v"), A.Min("v"), A.Mean("v")))
    cat = pa.concat_tables([b for b in blocks if b.num_rows])
    out["gb_key_offsets"], out["gb_key_data"], _ = _pack_rows([k.encode() for k in cat["key"].to_pylist()])
    for name in cat.column_names[1:]:
        out[f"gb::{name}"] = cat[name].to_numpy()
    out["gb_names"] = np.array(cat.column_names)
    np.savez_compressed(os.path.join(OUT, "varlen.npz"), **out)


def _ipc_bytes(t: pa.Table) -> np.ndarray:
    sink = pa.BufferOutputStream()
    with pa.ipc.new_stream(sink, t.schema) as w:
        w.write_table(t)
    return np.frombuffer(sink.getvalue().to_pybytes(), dtype=np.uint8).copy()


def gen_partial_agg():
    """The COMBINED (not finalised) blocks of the reference: ArrowBlockAccessor.combine
    (arrow_block.py:396-467) and aggregate_combined_blocks(finalize=False) (:488-587) -- list
    columns [accumulator parts..., has_data] with None for groups whose result is null
    (null_aggregate.py:29-65).  Stored as Arrow IPC streams (npz holds no list columns)."""
    ns = ref_shim.load()
    A = ns.aggregate
    AB = ns.arrow_block.ArrowBlockAccessor
    rng = np.random.default_rng(11)
    n = 3000
    k = np.sort(rng.integers(0, 120, n)).astype(np.int64)
    v = rng.integers(-50, 50, n).astype(np.float64) / 4.0           # exactly representable
    w = rng.integers(-1000, 1000, n).astype(np.int64)
    vmask = rng.random(n) < 0.15
    vmask[k == 7] = True                                             # a group with no valid value
    t = pa.table({"k": k, "v": pa.array(v, mask=vmask), "w": w})
    aggs = (A.Sum("v"), A.Mean("v"), A.Count(), A.Max("w"), A.Min("v"), A.Std("v"),
            A.Sum("v", ignore_nulls=False), A.Mean("w"), A.Sum("w"), A.AbsMax("w"))
    halves = [t.slice(0, 1700), t.slice(1700)]                       # key 0..: groups straddle the cut
    parts = [AB(h).combine("k", aggs) for h in halves]
    merged, _ = AB.aggregate_combined_blocks(parts, "k", aggs, False)
    out = {"input": _ipc_bytes(t), "cut": np.array([1700]), "combined_0": _ipc_bytes(parts[0]),
           "combined_1": _ipc_bytes(parts[1]), "merged_partial": _ipc_bytes(merged)}
    np.savez_compressed(os.path.join(OUT, "partial_agg.npz"), **out)


def main():
    os.makedirs(OUT, exist_ok=True)
    gen_partial_agg()
    gen_varlen()
    gen_nulls()
    gen_sort_int64()
    gen_sort_graysort()
    gen_partition()
    gen_boundaries()
    gen_shuffle()
    gen_groupby()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
