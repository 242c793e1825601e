"""Long skip connections that do not ride through every pipeline stage.

A layer decorated with ``@skippable(stash=['name'])`` hands a tensor over with
``yield stash('name', t)``; a later layer decorated with ``@skippable(pop=['name'])`` takes it
back with ``t = yield pop('name')``. Between the two, the tensor travels straight from the
producing stage's device to the consuming stage's device (``portal.py``), not through the stages
in between. ``Namespace`` objects keep equal names in repeated blocks apart
(``layer.isolate(ns)``), ``verify_skippables`` checks a model's stash/pop pairing up front.

Public names as in reference ``torchgpipe/skip/__init__.py``; the layout and tracker helpers are
exported as well because the multi-process mode and the benchmark drivers use them.
"""
from torchgpipe_b200.skip.layout import SkipLayout, inspect_skip_layout
from torchgpipe_b200.skip.namespace import Namespace
from torchgpipe_b200.skip.skippable import Skippable, pop, skippable, stash, verify_skippables
from torchgpipe_b200.skip.tracker import current_skip_tracker, use_skip_tracker

__all__ = [
    # user API
    'skippable', 'stash', 'pop', 'verify_skippables', 'Namespace',
    # building blocks
    'Skippable', 'SkipLayout', 'inspect_skip_layout', 'current_skip_tracker', 'use_skip_tracker',
]
