# Same behaviour as the reference's deprecated module (torchgpipe_balancing.py:1-2).
raise ImportError("import 'torchgpipe_b200.balance' (torchgpipe.balance) instead")
