"""Drop-in aliasing for code written against the reference package name.

``install()`` registers this package and its sub-modules in ``sys.modules``
under ``torchgpipe`` (and makes ``import torchgpipe_balancing`` raise the same
``ImportError`` as the reference's deprecated shim, ``torchgpipe_balancing.py:1-2``).
"""
import importlib
import sys

_SUBMODULES = ['gpipe', 'microbatch', 'batchnorm', 'checkpoint', 'stream', 'phony', 'dependency',
               'copy', 'pipeline', 'worker', 'balance', 'balance.blockpartition',
               'balance.profile', 'skip', 'skip.skippable', 'skip.namespace', 'skip.layout',
               'skip.tracker', 'skip.portal', 'distributed', 'distributed.context',
               'distributed.gpipe', 'distributed.utils']


def install(name: str = 'torchgpipe') -> None:
    """Makes ``import torchgpipe[.x]`` resolve to ``torchgpipe_b200[.x]``."""
    import torchgpipe_b200
    sys.modules[name] = torchgpipe_b200
    for sub in _SUBMODULES:
        sys.modules[f'{name}.{sub}'] = importlib.import_module(f'torchgpipe_b200.{sub}')
