from ..ml_type import ExecutorHookPoint


class Hook:
    def __init__(self, stripable: bool = False) -> None:
        self.enabled = True

    def yield_hooks(self):
        for point in ExecutorHookPoint:
            fn = getattr(self, "_" + point.name.lower(), None)
            if fn is not None:
                yield point, fn
