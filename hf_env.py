"""Environment selector shim.

Every reference script starts with ``import hf_env; hf_env.set_env('202111')`` (ref:resnet_dp.py:1-2),
which on High-Flyer's cluster activates a dated site-packages snapshot.  Off-cluster there is one
environment, so this records the request and does nothing else.
"""
requested_env = None


def set_env(name: str) -> None:
    global requested_env
    requested_env = name
