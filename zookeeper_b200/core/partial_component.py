"""``PartialComponent``: a component class with some field values pre-bound, in the spirit of
``functools.partial`` (reference zookeeper/core/partial_component.py:26-116)."""
from __future__ import annotations

import inspect
from typing import Generic, Type, TypeVar

from . import utils

C = TypeVar("C")

_KWARGS_HELP = (
    "Keyword arguments passed to `PartialComponent` must be either:\n"
    "- An immutable value (int, float, bool, string, or None).\n"
    "- A function or lambda accepting no arguments and returning the value that "
    "should be passed to the component upon instantiation.\n"
    "- An @component class that will be used to instantiate a component instance "
    "for the corresponding field value.\n"
    "- Another `PartialComponent`.\n"
    "Wrapping non-immutable values in a function / lambda allows the values "
    "to be lazily evaluated; they won't be created at all if the partial "
    "component is never instantiated."
)


def _unknown_kwarg(name: str, cls: Type) -> TypeError:
    return TypeError(
        f"Keyword argument '{name}' passed to `PartialComponent` does "
        "not correspond to any field of component class "
        f"'{cls.__name__}'."
    )


class PartialComponent(Generic[C]):
    """``PartialComponent(Cls, a=3)(b=4)`` builds ``Cls(a=3, b=4)``; bound values are produced
    lazily, only when an instance is actually requested."""

    def __init__(self, component_class: Type[C], **kwargs):
        if utils.is_component_instance(component_class):
            raise TypeError(
                "`PartialComponent` must be passed component classes, not component "
                f"instances. Received: {repr(component_class)}"
            )
        if not utils.is_component_class(component_class):
            raise TypeError(
                "The class passed to `PartialComponent` must be a component class. "
                f"Received: {component_class}."
            )
        if not kwargs:
            raise TypeError("`PartialComponent` must receive at least one keyword argument.")

        thunks = {}
        for name, value in kwargs.items():
            if name not in component_class.__component_fields__:
                raise _unknown_kwarg(name, component_class)
            if utils.is_immutable(value):
                thunks[name] = utils.wrap_in_callable(value)
            elif utils.is_component_class(value) or isinstance(value, PartialComponent):
                thunks[name] = value  # calling either yields a fresh component instance
            elif inspect.isfunction(value) and not inspect.signature(value).parameters:
                thunks[name] = value
            else:
                raise TypeError(_KWARGS_HELP)
        self._component_class = component_class
        self._lazy_kwargs = thunks

    def __set_name__(self, owner, name):
        # PEP 487 hook: a PartialComponent assigned straight into a class body is a mistake.
        raise ValueError(
            "`PartialComponent` instances should not be directly assigned to class "
            "bodies. You should instead use `PartialComponent` inside a "
            "`ComponentField`, like so:\n"
            "```\n"
            "@component\n"
            "class ParentComponentClass:\n"
            "    child_component: SomeChildComponentType = ComponentField(\n"
            "        PartialComponent(\n"
            "            SomeDefaultChildComponentClass,\n"
            "            some_arg=some_default_value,\n"
            "            some_other_arg=some_other_default_value,\n"
            "            ...\n"
            "        )\n"
            "    )\n"
            "```"
        )

    def __call__(self, **kwargs) -> C:
        for name in kwargs:
            if name not in self._component_class.__component_fields__:
                raise _unknown_kwarg(name, self._component_class)
        bound = {n: thunk() for n, thunk in self._lazy_kwargs.items() if n not in kwargs}
        bound.update(kwargs)
        return self._component_class(**bound)
