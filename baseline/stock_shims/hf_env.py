"""Reference-arm shim: `hf_env.set_env` is a no-op off the High-Flyer cluster."""


def set_env(name):
    return None
