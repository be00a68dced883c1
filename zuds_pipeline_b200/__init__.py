"""zuds_pipeline_b200 — B200-native Alard-Lupton image subtraction behind the zuds hotpants API.

Only the hot path of zuds-survey/zuds-pipeline lives here: what `zuds.hotpants.prepare_hotpants` +
`subprocess.check_call('hotpants ...')` do for `zuds.subtraction.Subtraction.from_images`
(zuds/hotpants.py:15-95, zuds/subtraction.py:57-226).  The arithmetic is in libzb200.so
(csrc/*.cu, C ABI in include/zb200.h); this package is the Python host side.
"""
from . import fitsio, params, synth  # noqa: F401

__all__ = ["fitsio", "params", "synth"]
