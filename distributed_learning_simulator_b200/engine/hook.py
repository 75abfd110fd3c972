"""Hook machinery for Trainer / Inferencer.

Re-creates the hook API the reference relies on from ``cyy_torch_toolbox``
(``append_named_hook``/``remove_named_hook``/``append_hook``/``remove_hook``/
``has_hook_obj``/``get_hook``/``disable_hook``; call sites
/root/reference/simulation_lib/worker/worker.py:75-83,
worker/aggregation_worker.py:42-80, worker/gradient_worker.py:33-42).
Hooks are called with keyword arguments; ``executor=`` is always present,
``epoch=`` on epoch-scoped points (method/fed_obd/worker.py:50-51).
"""
from typing import Any, Callable

from .ml_type import ExecutorHookPoint


class Hook:
    """Object hook: methods named ``_<hook point in lower case>`` are bound to that
    point (``_before_execute``, ``_after_epoch``, ``_model_loaded`` ...)."""

    def __init__(self, stripable: bool = False) -> None:
        self._enabled: bool = True
        self.stripable = stripable

    @property
    def enabled(self) -> bool:
        return self._enabled

    def enable(self) -> None:
        self._enabled = True

    def disable(self) -> None:
        self._enabled = False

    def yield_hooks(self):
        for point in ExecutorHookPoint:
            fn = getattr(self, "_" + point.name.lower(), None)
            if fn is not None:
                yield point, fn


class HookCollection:
    def __init__(self) -> None:
        self._hooks: dict[ExecutorHookPoint, list[tuple[str, Callable]]] = {}
        self._hook_objs: dict[str, Hook] = {}
        self._disabled_hooks: set[str] = set()

    # ---- named function hooks -------------------------------------------------
    def append_named_hook(self, hook_point: ExecutorHookPoint, name: str, fun: Callable) -> None:
        entries = self._hooks.setdefault(hook_point, [])
        assert all(n != name for n, _ in entries), f"duplicate hook {name}"
        entries.append((name, fun))

    def prepend_named_hook(self, hook_point: ExecutorHookPoint, name: str, fun: Callable) -> None:
        entries = self._hooks.setdefault(hook_point, [])
        assert all(n != name for n, _ in entries), f"duplicate hook {name}"
        entries.insert(0, (name, fun))

    def remove_named_hook(self, name: str, hook_point: ExecutorHookPoint | None = None) -> None:
        for point, entries in self._hooks.items():
            if hook_point is not None and point != hook_point:
                continue
            entries[:] = [(n, f) for n, f in entries if n != name]

    def has_named_hook(self, name: str) -> bool:
        return any(n == name for entries in self._hooks.values() for n, _ in entries)

    # ---- object hooks ---------------------------------------------------------
    def append_hook(self, hook: Hook, hook_name: str | None = None) -> None:
        if hook_name is None:
            hook_name = type(hook).__name__
        assert hook_name not in self._hook_objs
        self._hook_objs[hook_name] = hook
        for point, fn in hook.yield_hooks():
            self.append_named_hook(point, f"{hook_name}.{point.name}", fn)

    def remove_hook(self, hook_name: str) -> None:
        hook = self._hook_objs.pop(hook_name, None)
        if hook is None:
            return
        for point, _ in hook.yield_hooks():
            self.remove_named_hook(f"{hook_name}.{point.name}", point)

    def has_hook_obj(self, hook_name: str) -> bool:
        return hook_name in self._hook_objs

    def get_hook(self, hook_name: str) -> Hook:
        return self._hook_objs[hook_name]

    def enable_hook(self, hook_name: str) -> None:
        self._disabled_hooks.discard(hook_name)
        if hook_name in self._hook_objs:
            self._hook_objs[hook_name].enable()

    def disable_hook(self, hook_name: str) -> None:
        self._disabled_hooks.add(hook_name)
        if hook_name in self._hook_objs:
            self._hook_objs[hook_name].disable()

    # ---- dispatch -------------------------------------------------------------
    def has_hook(self, hook_point: ExecutorHookPoint) -> bool:
        return any(self._is_live(n) for n, _ in self._hooks.get(hook_point, ()))

    def _is_live(self, name: str) -> bool:
        base = name.split(".", 1)[0]
        return name not in self._disabled_hooks and base not in self._disabled_hooks

    def live_hook_names(self, hook_point: ExecutorHookPoint) -> list[str]:
        return [n for n, _ in self._hooks.get(hook_point, ()) if self._is_live(n)]

    def exec_hooks(self, hook_point: ExecutorHookPoint, **kwargs: Any) -> None:
        for name, fun in list(self._hooks.get(hook_point, ())):
            if self._is_live(name):
                fun(**kwargs)
