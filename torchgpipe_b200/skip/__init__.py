"""Long skip connections across pipeline stages (``@skippable`` / ``stash`` / ``pop``)."""
from torchgpipe_b200.skip.namespace import Namespace
from torchgpipe_b200.skip.skippable import pop, skippable, stash, verify_skippables

__all__ = ['skippable', 'stash', 'pop', 'verify_skippables', 'Namespace']
