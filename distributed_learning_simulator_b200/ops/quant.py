def available() -> bool:
    from . import runtime

    return runtime.available()
