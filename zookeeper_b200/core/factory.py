"""``@factory``: a component whose ``build()`` result stands in for the component wherever it is
used as a field value (reference zookeeper/core/factory.py:11-131)."""
from __future__ import annotations

import functools
import inspect
from typing import Type

from . import utils
from .component import component
from .factory_registry import FACTORY_REGISTRY

_BUILD_SIGNATURE_ERROR = (
    "Classes decorated with @factory must implement a `build()` method taking "
    "precisely one positional argument, `self`."
)


def _build_signature(cls: Type) -> inspect.Signature:
    build = getattr(cls, "build", None)
    if build is None:
        raise TypeError(_BUILD_SIGNATURE_ERROR)
    try:
        sig = inspect.signature(build)
    except (TypeError, ValueError):
        raise TypeError(_BUILD_SIGNATURE_ERROR) from None
    params = list(sig.parameters.values())
    variadic = (inspect.Parameter.VAR_POSITIONAL, inspect.Parameter.VAR_KEYWORD)
    if len(params) != 1 or params[0].name != "self" or params[0].kind in variadic:
        raise TypeError(_BUILD_SIGNATURE_ERROR)
    return sig


def _memoise_build(cls: Type) -> None:
    raw_build = cls.build
    produced = cls.__component_factory_return_type__

    @functools.wraps(raw_build)
    def build(self):
        if self.__component_factory_value__ is utils.missing:
            value = raw_build(self)
            if not utils.type_check(value, produced):
                raise TypeError(
                    f"@factory '{cls}' has a `build()` method annotated with "
                    f"return type {produced}, "
                    f"which is not satisfied by the return value {value}."
                )
            self.__component_factory_value__ = value
        return self.__component_factory_value__

    cls.build = build


def _relabel_printing(cls: Type) -> None:
    for dunder in ("__str__", "__repr__"):
        original = getattr(cls, dunder)

        def printer(self, _original=original):
            return _original(self).replace("<Unconfigured component ", "<Unconfigured factory ")

        setattr(cls, dunder, functools.wraps(original)(printer))


def factory(cls: Type):
    """Class decorator: ``cls`` becomes a component with a cached, type-checked ``build()``;
    it is registered as a producer of its annotated return type."""
    cls = component(cls)
    sig = _build_signature(cls)
    if sig.return_annotation is sig.empty:
        raise TypeError(
            "The `build()` method of a @factory class must have an annotated return "
            "type annotation, e.g.:\n\n"
            "```\n"
            "@factory\n"
            "class MyFactory:\n"
            "    ...\n"
            "    def build(self) -> SomeReturnType:\n"
            "        ...\n"
            "        return some_value\n"
            "```"
        )
    cls.__component_factory_return_type__ = sig.return_annotation
    cls.__component_factory_value__ = utils.missing
    _memoise_build(cls)
    _relabel_printing(cls)
    FACTORY_REGISTRY.setdefault(sig.return_annotation, set()).add(cls)
    return cls
