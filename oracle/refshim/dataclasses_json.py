"""Minimal stand-in for the `dataclasses_json` package (not installed here).

Gives dataclasses a `to_dict()` / `from_dict()` pair good enough for the reference's
settings tree (nested dataclass + IntEnum).  Test infrastructure only.
"""
import dataclasses
import enum
import typing


def _plain(v):
    if dataclasses.is_dataclass(v) and not isinstance(v, type):
        return {f.name: _plain(getattr(v, f.name)) for f in dataclasses.fields(v)}
    if isinstance(v, enum.Enum):
        return v.value
    if isinstance(v, (list, tuple)):
        return [_plain(x) for x in v]
    if isinstance(v, dict):
        return {k: _plain(x) for k, x in v.items()}
    return v


def _build(tp, v):
    origin = typing.get_origin(tp)
    if origin is typing.Union:
        for a in typing.get_args(tp):
            if a is type(None):
                continue
            return _build(a, v) if v is not None else None
    if isinstance(tp, type) and dataclasses.is_dataclass(tp) and isinstance(v, dict):
        hints = typing.get_type_hints(tp)
        return tp(**{k: _build(hints.get(k, typing.Any), x) for k, x in v.items()})
    if isinstance(tp, type) and issubclass(tp, enum.Enum) and not isinstance(v, enum.Enum):
        return tp(v)
    return v


def dataclass_json(cls):
    cls.to_dict = lambda self: _plain(self)
    cls.from_dict = classmethod(lambda c, d: _build(c, d))
    return cls
