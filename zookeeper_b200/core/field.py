"""Declarative fields of components (reference zookeeper/core/field.py:13-281).

``Field`` holds a plain configurable value, ``ComponentField`` a nested sub-component.  Both are
class-level descriptors-by-convention: the component's attribute lookup (component.py) consults
them; they never bind to instances themselves.
"""
from __future__ import annotations

import inspect
from typing import Any, Callable, Generic, Type, TypeVar, Union

from . import utils
from .partial_component import PartialComponent

C = TypeVar("C")
F = TypeVar("F")

_EMPTY = inspect.Signature.empty


def _annotation_for(owner: Type, name: str) -> Any:
    """First annotation for ``name`` along the MRO of ``owner`` (or ``missing``)."""
    for klass in inspect.getmro(owner):
        ann = getattr(klass, "__annotations__", {})
        if name in ann:
            return ann[name]
    return utils.missing


class _FieldNameError(ValueError, RuntimeError):
    """The reference raises ``ValueError`` from ``__set_name__`` (zookeeper/core/field.py:81) and its
    test expects the ``RuntimeError`` CPython < 3.12 wrapped such exceptions in (field_test.py:116-125);
    CPython >= 3.12 re-raises the original with a note instead.  Being both keeps the caller-visible
    behaviour on either side of that interpreter change."""


class Field(Generic[C, F]):
    """A typed, configurable value.  ``default`` may be an immutable value or a function of zero
    arguments / of ``self`` that produces the default lazily (evaluated at most once per
    instance, the component caches it)."""

    def __init__(self, default: Union[utils.Missing, F, Callable[[], F], Callable[[C], F]] = utils.missing,
                 *, allow_missing: bool = False):
        self.name = None
        self.type = None
        self.host_component_class = None
        self.allow_missing = allow_missing
        self._registered = False
        self._return_annotation = _EMPTY

        if allow_missing and default is not utils.missing:
            raise ValueError("If a `Field` has `allow_missing=True`, no default can be provided.")

        if default is utils.missing or utils.is_immutable(default):
            self._default = default
            return
        if inspect.isfunction(default):
            sig = inspect.signature(default)
            params = list(sig.parameters.values())
            variadic = (inspect.Parameter.VAR_POSITIONAL, inspect.Parameter.VAR_KEYWORD)
            if len(params) <= 1 and not (params and params[0].kind in variadic):
                self._default = default
                self._return_annotation = sig.return_annotation
                return
        raise TypeError(
            "If `default` is passed to `Field`, it must be either:\n"
            "- An immutable value (int, float, bool, string, or None).\n"
            "- A function or lambda accepting no arguments or a single\n"
            "  argument (`self`), and returning the default value.\n"
            f"Received: {default}."
        )

    # PEP 487: called when the class body that contains this field has been executed.
    def __set_name__(self, owner: Type[C], name: str):
        self._claim(name)
        declared = _annotation_for(owner, name)
        returned = self._return_annotation
        if declared is utils.missing:
            if returned is _EMPTY:
                raise TypeError(
                    "Fields must be defined inside the component class definition, "
                    "with a type annotation in one of the following ways:\n\n"
                    "```\n"
                    "@component\n"
                    "class ComponentClass:\n"
                    "    ...\n"
                    "    # Like this\n"
                    "    name_1: type_1 = Field(default_1)\n"
                    "    ...\n"
                    "    # Or like this\n"
                    "    @Field\n"
                    "    def name_2(self) -> type_2:\n"
                    "        ...\n"
                    "        return default_2\n"
                    "```\n\n"
                    f"Unable to find a type annotation for field '{name}' on class "
                    f"'{owner.__name__}'."
                )
            declared = returned
        elif returned is not _EMPTY and returned != declared:
            raise TypeError(
                f"Two non-equal type annotations found for field '{name}': "
                f"{declared} and {returned}."
            )
        self._bind(owner, name, declared)

    def _claim(self, name: str) -> None:
        if self._registered:
            raise ValueError("This field has already been registered to a component.")
        if name.startswith("_"):
            raise _FieldNameError("Field names cannot start with underscores.")

    def _bind(self, owner: Type, name: str, annotation: Any) -> None:
        self.name, self.host_component_class, self.type = name, owner, annotation
        self._registered = True

    def _require_registered(self) -> None:
        if not self._registered:
            raise ValueError("This field has not been registered to a component.")

    def __repr__(self) -> str:
        if not self._registered:
            return "<Unregistered Field>"
        return f"<Field '{self.name}' of {self.host_component_class.__name__} with type {self.type}>"

    @property
    def has_default(self) -> bool:
        self._require_registered()
        return self._default is not utils.missing

    _KIND = "Field"
    _NO_DEFAULT = "has no default or configured value."

    def _check_owner(self, instance: Any) -> None:
        self._require_registered()
        if not self.has_default:
            raise AttributeError(f"{self._KIND} '{self.name}' {self._NO_DEFAULT}")
        if not isinstance(instance, self.host_component_class):
            owner = self.host_component_class.__name__
            raise TypeError(
                f"{self._KIND} '{self.name}' belongs to component '{owner}'; `get_default` must be "
                f"called with an instance of '{owner}'. Received: {repr(instance)}."
            )

    def get_default(self, instance: C) -> F:
        self._check_owner(instance)
        default = self._default
        if not inspect.isfunction(default):
            return default
        value = default(instance) if inspect.signature(default).parameters else default()
        if utils.is_component_instance(value):
            raise TypeError(
                f"Field '{self.name}' of component '{instance.__component_name__}' "
                "is returning a component instance as its default value. To use "
                "components in fields, use `ComponentField` rather than `Field`."
            )
        return value


class ComponentField(Field, Generic[C, F]):
    """A field whose value is a nested component.  The annotation names the base type every
    candidate must inherit from; ``default`` may be a component class or a ``PartialComponent``,
    and extra keyword arguments pre-set fields of that default."""

    _KIND = "ComponentField"
    _NO_DEFAULT = "has no default or configured component class."

    def __init__(self, default: Union[utils.Missing, F, PartialComponent[F]] = utils.missing, *,
                 allow_missing: bool = False, **kwargs):
        if allow_missing and default is not utils.missing:
            raise ValueError("If a `Field` has `allow_missing=True`, no default can be provided.")
        if default is utils.missing:
            if kwargs:
                raise TypeError(
                    "Keyword arguments can only be passed to `ComponentField` if "
                    "a default component class is also passed."
                )
        elif isinstance(default, PartialComponent) or utils.is_component_class(default):
            if kwargs:
                default = PartialComponent(default, **kwargs)
        elif utils.is_component_instance(default):
            raise TypeError(
                "The `default` passed to `ComponentField` must be a component class, "
                f"not a component instance. Received: {repr(default)}."
            )
        else:
            raise TypeError(
                "The `default` passed to `ComponentField` must be either a component "
                "class or a `PartialComponent`."
            )
        self.name = utils.missing
        self.type = utils.missing
        self.host_component_class = utils.missing
        self.allow_missing = allow_missing
        self._registered = False
        self._default = default

    def __set_name__(self, owner: Type[C], name: str):
        self._claim(name)
        declared = _annotation_for(owner, name)
        if declared is utils.missing:
            raise TypeError(
                "ComponentFields must be defined inside the component class definition, "
                "with a type-annotation as follows:\n\n"
                "```\n"
                "@component\n"
                "class ParentComponentClass:\n"
                "    field_name: SomeChildComponentType = ComponentField(\n"
                "        SomeDefaultChildComponentClass\n"
                "    )\n"
                "```\n"
                "\nUnlike `Field`, `ComponentField` cannot be used as a decorator.\n\n"
                f"Unable to find a type annotation for field '{name}' on class "
                f"'{owner.__name__}'."
            )
        self._bind(owner, name, declared)

    def get_default(self, component_instance: C) -> F:
        self._check_owner(component_instance)
        # A fresh child without arguments: as a nested sub-component it picks up whatever it is
        # missing from its parent.
        return self._default()
