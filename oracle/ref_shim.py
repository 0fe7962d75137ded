"""TEST INFRASTRUCTURE ONLY -- not part of the product path.

Loads the reference's own hot-path Python files from /root/reference under an
in-process stand-in for the ``ray`` task runtime, so that the reference's
``sort_impl`` / ``_ShufflePartitionOp`` / ``_GroupbyOp`` can be executed in
this container (``ray`` itself is not installed and cannot be built: no bazel).

Only ``oracle/gen_golden.py`` and the container-only pinning tests use this
module.  It cannot travel to the GPU box (/root/reference does not exist
there); what travels are the golden vectors it generated (tests/golden/).

What is stubbed: ``ray.remote`` (eager in-process
call, generator results listed, ``num_returns`` split), ``ray.get/put/wait/
nodes``, ``ray.types.ObjectRef``, the runtime-context node id, and the few
names the data package imports from ``ray.util`` / ``ray._private``.
Nothing is copied from the reference: its files are imported where they lie.
"""
import ast
import os
import sys
import types
import warnings

REFERENCE_ROOT = os.environ.get("DS_REFERENCE_ROOT", "/root/reference")
_PY = os.path.join(REFERENCE_ROOT, "python", "ray")


def available() -> bool:
    return os.path.isdir(os.path.join(_PY, "data", "_internal"))


class _Ref:
    """Stand-in for ray.ObjectRef: holds the value directly."""

    __slots__ = ("value",)

    def __init__(self, value):
        self.value = value

    def __class_getitem__(cls, item):
        return cls


def _unwrap(x):
    return x.value if isinstance(x, _Ref) else x


class _RemoteFn:
    def __init__(self, fn, opts):
        self._fn = fn
        self._opts = dict(opts)

    def options(self, **opts):
        merged = dict(self._opts)
        merged.update(opts)
        return _RemoteFn(self._fn, merged)

    def remote(self, *args, **kwargs):
        args = [_unwrap(a) for a in args]
        kwargs = {k: _unwrap(v) for k, v in kwargs.items()}
        out = self._fn(*args, **kwargs)
        if isinstance(out, types.GeneratorType):
            out = list(out)
        n = self._opts.get("num_returns", 1)
        if n == 1:
            return _Ref(out)
        out = list(out)
        assert len(out) == n, (len(out), n)
        return [_Ref(o) for o in out]


def _remote(*args, **kwargs):
    if len(args) == 1 and callable(args[0]) and not kwargs:
        return _RemoteFn(args[0], {})

    def deco(fn):
        return _RemoteFn(fn, kwargs)

    return deco


def _get(refs, **_):
    if isinstance(refs, (list, tuple)):
        return [_unwrap(r) for r in refs]
    return _unwrap(refs)


def _install():
    if "ray" in sys.modules and getattr(sys.modules["ray"], "_ds_oracle_shim", False):
        return sys.modules["ray"]
    if "ray" in sys.modules:
        raise RuntimeError("a real `ray` is already imported; shim not needed")

    def pkg(name, path):
        m = types.ModuleType(name)
        m.__path__ = [path]
        sys.modules[name] = m
        return m

    ray = pkg("ray", _PY)
    ray._ds_oracle_shim = True
    ray.remote = _remote
    ray.get = _get
    ray.put = lambda v: _Ref(v)
    ray.wait = lambda refs, **kw: (list(refs), [])
    ray.nodes = lambda: [
        {"NodeID": "node0", "Alive": True, "Resources": {"CPU": float(os.cpu_count() or 1)}}
    ]
    ray.ObjectRef = _Ref

    rtypes = types.ModuleType("ray.types")
    rtypes.ObjectRef = _Ref
    sys.modules["ray.types"] = rtypes
    ray.types = rtypes

    class _NodeId:
        def hex(self):
            return "node0"

    class _Ctx:
        node_id = _NodeId()

    rc = types.ModuleType("ray.runtime_context")
    rc.get_runtime_context = lambda: _Ctx()
    sys.modules["ray.runtime_context"] = rc
    ray.runtime_context = rc

    actor = types.ModuleType("ray.actor")
    actor.ActorHandle = type("ActorHandle", (), {})
    sys.modules["ray.actor"] = actor
    ray.actor = actor

    priv = pkg("ray._private", os.path.join(_PY, "_private_shim_none"))
    consts = types.ModuleType("ray._private.ray_constants")
    consts.env_integer = lambda k, d: int(os.environ.get(k, d))
    sys.modules["ray._private.ray_constants"] = consts
    priv.ray_constants = consts
    ray._private = priv

    import cloudpickle

    ray.cloudpickle = cloudpickle
    sys.modules["ray.cloudpickle"] = cloudpickle

    util = pkg("ray.util", os.path.join(_PY, "util"))
    ray.util = util
    ann = types.ModuleType("ray.util.annotations")
    for name in ("PublicAPI", "DeveloperAPI", "Deprecated"):

        def _mk():
            def deco(*a, **k):
                if len(a) == 1 and callable(a[0]) and not k:
                    return a[0]
                return lambda f: f

            return deco

        setattr(ann, name, _mk())
    sys.modules["ray.util.annotations"] = ann
    util.annotations = ann

    sched = types.ModuleType("ray.util.scheduling_strategies")

    class NodeAffinitySchedulingStrategy:
        def __init__(self, node_id=None, soft=False):
            self.node_id, self.soft = node_id, soft

    sched.NodeAffinitySchedulingStrategy = NodeAffinitySchedulingStrategy
    sched.SchedulingStrategyT = object
    sys.modules["ray.util.scheduling_strategies"] = sched
    util.scheduling_strategies = sched

    client = types.ModuleType("ray.util.client")

    class _ClientRay:
        @staticmethod
        def is_connected():
            return True

    client.ray = _ClientRay()
    sys.modules["ray.util.client"] = client
    util.client = client

    data = pkg("ray.data", os.path.join(_PY, "data"))
    ray.data = data
    internal = pkg("ray.data._internal", os.path.join(_PY, "data", "_internal"))
    data._internal = internal
    pkg("ray.data._internal.arrow_ops", os.path.join(_PY, "data", "_internal", "arrow_ops"))
    return ray


_LOADED = None


def load():
    """Import the reference hot-path modules; returns a namespace object."""
    global _LOADED
    if _LOADED is not None:
        return _LOADED
    if not available():
        raise RuntimeError(f"reference tree not present at {REFERENCE_ROOT}")
    warnings.filterwarnings("ignore", category=DeprecationWarning)
    warnings.filterwarnings("ignore", category=FutureWarning)
    warnings.filterwarnings("ignore", category=UserWarning)
    _install()
    import importlib

    util = importlib.import_module("ray.data._internal.util")
    util._VERSION_VALIDATED = True
    pb = importlib.import_module("ray.data._internal.progress_bar")
    pb._enabled = False
    ns = types.SimpleNamespace()
    ns.block = importlib.import_module("ray.data.block")
    ns.context = importlib.import_module("ray.data.context")
    ns.arrow_block = importlib.import_module("ray.data._internal.arrow_block")
    ns.sort = importlib.import_module("ray.data._internal.sort")
    ns.shuffle = importlib.import_module("ray.data._internal.shuffle")
    ns.push = importlib.import_module("ray.data._internal.push_based_shuffle")
    ns.shuffle_and_partition = importlib.import_module(
        "ray.data._internal.shuffle_and_partition"
    )
    ns.aggregate = importlib.import_module("ray.data.aggregate")
    ns.block_list = importlib.import_module("ray.data._internal.block_list")
    ns.BlockList = ns.block_list.BlockList
    ns.BlockAccessor = ns.block.BlockAccessor

    # grouped_dataset.py imports ray.data.dataset (the whole product); take only the
    # three op classes (grouped_dataset.py:24-68) out of its AST and exec them against
    # the modules loaded above.
    src_path = os.path.join(_PY, "data", "grouped_dataset.py")
    with open(src_path) as f:
        tree = ast.parse(f.read(), src_path)
    keep = [
        n
        for n in tree.body
        if isinstance(n, ast.ClassDef)
        and n.name in ("_GroupbyOp", "SimpleShuffleGroupbyOp", "PushBasedGroupbyOp")
    ]
    mod = ast.Module(body=keep, type_ignores=[])
    from typing import Any, List, Tuple, Union

    g = {
        "ShuffleOp": ns.shuffle.ShuffleOp,
        "SimpleShufflePlan": ns.shuffle.SimpleShufflePlan,
        "PushBasedShufflePlan": ns.push.PushBasedShufflePlan,
        "Block": ns.block.Block,
        "BlockAccessor": ns.block.BlockAccessor,
        "BlockExecStats": ns.block.BlockExecStats,
        "BlockMetadata": ns.block.BlockMetadata,
        "KeyFn": ns.block.KeyFn,
        "KeyType": ns.block.KeyType,
        "AggregateFn": ns.aggregate.AggregateFn,
        "List": List,
        "Tuple": Tuple,
        "Union": Union,
        "Any": Any,
    }
    exec(compile(mod, src_path, "exec"), g)
    ns.SimpleShuffleGroupbyOp = g["SimpleShuffleGroupbyOp"]
    ns.PushBasedGroupbyOp = g["PushBasedGroupbyOp"]
    _LOADED = ns
    return ns


def make_block_list(ns, tables):
    import ray

    metas = [ns.BlockAccessor.for_block(t).get_metadata(None, None) for t in tables]
    return ns.BlockList([ray.put(t) for t in tables], metas, owned_by_consumer=False)


def blocks_of(block_list):
    import ray

    return ray.get(block_list.get_blocks())


def ref_sort(tables, key, descending=False, push_based=False):
    """Run the reference's sort_impl (sort.py:120-154) on pyarrow tables."""
    ns = load()
    ctx = ns.context.DatasetContext.get_current()
    old = ctx.use_push_based_shuffle
    ctx.use_push_based_shuffle = push_based
    try:
        out, _ = ns.sort.sort_impl(make_block_list(ns, tables), False, key, descending)
        return blocks_of(out)
    finally:
        ctx.use_push_based_shuffle = old


def ref_random_shuffle(tables, seed, num_blocks=None, push_based=False):
    """Run the reference's shuffle op (stage_impl.py:92-119 wiring)."""
    ns = load()
    cls = (
        ns.shuffle_and_partition.PushBasedShufflePartitionOp
        if push_based
        else ns.shuffle_and_partition.SimpleShufflePartitionOp
    )
    op = cls(None, random_shuffle=True, random_seed=seed)
    n = num_blocks or len(tables)
    out, _ = op.execute(make_block_list(ns, tables), n, False)
    return blocks_of(out)


def ref_groupby(tables, key, aggs, push_based=False):
    """Re-creation of do_agg's driver lines (grouped_dataset.py:118-154)."""
    ns = load()
    bl = make_block_list(ns, tables)
    num_reducers = len(tables)
    if key is None:
        num_reducers, boundaries = 1, []
    else:
        boundaries = ns.sort.sample_boundaries(
            bl.get_blocks(), [(key, "ascending")], num_reducers
        )
    cls = ns.PushBasedGroupbyOp if push_based else ns.SimpleShuffleGroupbyOp
    op = cls(map_args=[boundaries, key, aggs], reduce_args=[key, aggs])
    out, _ = op.execute(bl, num_reducers, False)
    return blocks_of(out)
