"""Strings both arms of bench.py print, so the driver can divide one by the other.

`bench.py` (our framework) and `baseline/ref_arm.py` (the unmodified reference) must agree on
`metric`, `unit` and `higher_is_better`; everything that differs between the two arms (batch, what the
timed region contains, decomposition) is stated in `config`, never in `metric`.
The metric is BASELINE.json's; it is read from there so the three never drift apart."""
import json
import os

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_FALLBACK = "AlexNet blk1-2 images/s @1/2/4/8 B200 (max-over-ranks) vs ref V4/V5 MPI+CUDA"


def _metric() -> str:
    try:
        with open(os.path.join(_ROOT, "BASELINE.json")) as f:
            return json.load(f).get("metric") or _FALLBACK
    except Exception:
        return _FALLBACK


METRIC = _metric()
UNIT = "images/s"
HIGHER_IS_BETTER = True
SCALING = "strong"          # the global batch is fixed as N grows
MODEL = "AlexNet blocks 1-2 (conv1-relu-pool1-conv2-relu-pool2-lrn2)"
