"""NVTX ranges around the host-side entry points (tracing aid), off by default.

``IPOKE_NVTX=1`` (read once at import) wraps the engine / optimizer / MaCow-program entry points in
``torch.cuda.nvtx`` ranges named ``ipoke.<what>``, so a profiler timeline (ncu ``--nvtx``, nsys) groups the kernel
launches of one call.  With the variable unset ``annotate`` returns the function itself: the default path carries no
wrapper and no overhead."""
import functools
import os

ENABLED = os.environ.get("IPOKE_NVTX", "0") not in ("", "0")


def annotate(name):
    def deco(fn):
        if not ENABLED:
            return fn
        import torch

        @functools.wraps(fn)
        def wrapped(*args, **kwargs):
            torch.cuda.nvtx.range_push("ipoke." + name)
            try:
                return fn(*args, **kwargs)
            finally:
                torch.cuda.nvtx.range_pop()
        return wrapped
    return deco
