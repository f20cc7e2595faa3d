"""Optional NVTX ranges around the module calls, named after the reference operators they replace, for nsys / ncu
``--nvtx`` captures (the reference brackets its steps the same way in makani/utils/trainer_profile.py:652-733).
Enabled with MAKANI_B200_NVTX=1 at import time; otherwise the decorator returns the function unchanged (no overhead)."""
import functools
import os

ENABLED = os.environ.get("MAKANI_B200_NVTX", "0") == "1"


def annotate(name: str):
    def deco(fn):
        if not ENABLED:
            return fn
        import torch

        @functools.wraps(fn)
        def wrapped(*args, **kwargs):
            torch.cuda.nvtx.range_push(name)
            try:
                return fn(*args, **kwargs)
            finally:
                torch.cuda.nvtx.range_pop()
        return wrapped
    return deco
