"""Measured machine constants the runtime reports against."""
import json
import os

__all__ = ['link_bandwidth_gbs']

_DEFAULT_LINK_GBS = 775.0


def link_bandwidth_gbs() -> float:
    """NVLink peer bandwidth per direction in GB/s: ``TGB_LINK_GBS`` if set, else the committed
    measurement of ``benchmarks/p2p_bandwidth.py`` (``profiles/p2p_bandwidth.json``:
    ``unidirectional_gbs``), else 775 (that measurement on 4 x B200 through NVSwitch)."""
    env = os.environ.get('TGB_LINK_GBS')
    if env:
        try:
            return float(env)
        except ValueError:
            pass
    root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    try:
        with open(os.path.join(root, 'profiles', 'p2p_bandwidth.json')) as f:
            return float(json.load(f)['unidirectional_gbs'])
    except (OSError, ValueError, KeyError):
        return _DEFAULT_LINK_GBS
