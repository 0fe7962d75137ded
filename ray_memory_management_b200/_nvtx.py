"""NVTX ranges around the stages and the ds_* kernel families, so that an Nsight Systems
timeline of a job reads like the reference's ``ray timeline`` (one range per map / reduce /
exchange phase, nested ranges per library call).  A no-op when NVTX is unavailable (CPU-only
processes); off unless ``DS_NVTX=1`` or ``enable(True)``, since a push/pop pair per call is
host work on the launch path."""
import contextlib
import os

_on = os.environ.get("DS_NVTX", "0") == "1"
_push = _pop = None


def _load():
    global _push, _pop
    if _push is None:
        try:
            import torch

            torch.cuda.nvtx.range_push("ds")  # raises when the NVTX library is missing
            torch.cuda.nvtx.range_pop()
            _push, _pop = torch.cuda.nvtx.range_push, torch.cuda.nvtx.range_pop
        except Exception:
            _push = _pop = False
    return bool(_push)


def enable(on: bool = True) -> bool:
    """Switch the ranges on or off; returns whether NVTX is usable in this process."""
    global _on
    _on = bool(on)
    return _load() if _on else False


def enabled() -> bool:
    return _on and _load()


@contextlib.contextmanager
def range(name: str):
    if _on and _load():
        _push(name)
        try:
            yield
        finally:
            _pop()
    else:
        yield
