"""Shared helpers of the configuration runtime: sentinels, predicates, type checking, string
parsing and interactive prompts.  Behavioural contract: reference zookeeper/core/utils.py:11-195
(the error / warning texts asserted by the reference's tests are kept verbatim)."""
from __future__ import annotations

import ast
import inspect
import re
import typing
from typing import Any, Callable, Iterator, Sequence, Type, TypeVar

import click


class Missing:
    """Type of the ``missing`` sentinel (a field without default / value)."""

    _instance = None

    def __new__(cls):
        if cls._instance is None:
            cls._instance = super().__new__(cls)
        return cls._instance

    def __repr__(self):
        return "<missing>"


missing = Missing()


def warn(message: str) -> None:
    click.secho(f"WARNING: {message}", fg="yellow", err=True)


# ----------------------------------------------------------------------------- predicates
def is_component_class(cls: Any) -> bool:
    # The marker has to sit in the class' own namespace: a plain subclass of a component is not
    # a component until it is decorated itself.
    return inspect.isclass(cls) and "__component_name__" in vars(cls)


def is_component_instance(obj: Any) -> bool:
    return is_component_class(type(obj))


def is_factory_class(cls: Any) -> bool:
    return is_component_class(cls) and hasattr(cls, "__component_factory_return_type__")


def is_factory_instance(obj: Any) -> bool:
    return is_factory_class(type(obj))


def generate_subclasses(cls: Any) -> Iterator[Type]:
    """``cls`` followed by every transitive subclass, depth first."""
    if not inspect.isclass(cls):
        return
    stack = [cls]
    while stack:
        current = stack.pop()
        yield current
        stack.extend(reversed(current.__subclasses__()))


def generate_component_subclasses(cls: Any) -> Iterator[Type]:
    for sub in generate_subclasses(cls):
        if is_component_class(sub) and not inspect.isabstract(sub):
            yield sub


def generate_component_ancestors_with_field(instance: Any, field_name: str,
                                            include_instance: bool = False) -> Iterator[Any]:
    """Walk up the parent chain, nearest first, yielding components that declare ``field_name``."""
    node = instance if include_instance else instance.__component_parent__
    while node is not None:
        if field_name in node.__component_fields__:
            yield node
        node = node.__component_parent__


# ----------------------------------------------------------------------------- type checking
def _value_matches(value: Any, expected: Any) -> bool:
    try:
        import typeguard
    except ImportError:  # pragma: no cover - typeguard ships with the image
        return _structural_check(value, expected)
    try:
        params = list(inspect.signature(typeguard.check_type).parameters)
        if params and params[0] == "argname":      # typeguard 2.x: check_type(argname, value, type)
            typeguard.check_type("", value, expected)
        else:                                        # typeguard >= 3: check_type(value, type)
            typeguard.check_type(value, expected)
    except Exception as exc:
        if exc.__class__.__name__ in ("TypeCheckError", "TypeError"):
            return False
        return _structural_check(value, expected)
    return True


def _structural_check(value: Any, expected: Any) -> bool:
    """Small stand-in used only when typeguard cannot decide."""
    if expected is Any or expected is inspect.Signature.empty:
        return True
    origin = typing.get_origin(expected)
    args = typing.get_args(expected)
    if origin is typing.Union:
        return any(_structural_check(value, a) for a in args)
    if origin in (list, set, frozenset) or origin is typing.Sequence:
        if not isinstance(value, origin if inspect.isclass(origin) else (list, tuple)):
            return False
        return not args or all(_structural_check(v, args[0]) for v in value)
    if origin is tuple:
        if not isinstance(value, tuple):
            return False
        if not args or (len(args) == 2 and args[1] is Ellipsis):
            return not args or all(_structural_check(v, args[0]) for v in value)
        return len(args) == len(value) and all(_structural_check(v, a) for v, a in zip(value, args))
    if origin is dict:
        if not isinstance(value, dict):
            return False
        return not args or all(_structural_check(k, args[0]) and _structural_check(v, args[1])
                               for k, v in value.items())
    if expected is float:
        return isinstance(value, (int, float)) and not isinstance(value, bool)
    if inspect.isclass(expected):
        return isinstance(value, expected)
    if origin is not None and inspect.isclass(origin):
        return isinstance(value, origin)
    return True


def type_check(value: Any, expected_type: Any) -> bool:
    """Does ``value`` satisfy ``expected_type``?  For a @factory instance the question is whether
    what ``build()`` returns is a sub-type; if that cannot be decided a warning is printed and
    the value is accepted."""
    if is_factory_instance(value):
        produced = value.__component_factory_return_type__
        try:
            return issubclass(produced, expected_type)
        except TypeError:
            warn(f"Unable to check that {produced} is a sub-type of {expected_type}.")
            return True
    return _value_matches(value, expected_type)


# ----------------------------------------------------------------------------- misc helpers
T = TypeVar("T")


def wrap_in_callable(value: T) -> Callable[[], T]:
    return lambda: value


_SCALARS = (int, float, bool, str, frozenset)


def is_immutable(value: Any) -> bool:
    """None, scalars, strings, frozensets, and tuples made only of those (one level deep)."""
    if value is None or isinstance(value, _SCALARS):
        return True
    return isinstance(value, tuple) and all(v is None or isinstance(v, _SCALARS) for v in value)


def type_name_str(tp: Any) -> str:
    try:
        for attr in ("__qualname__", "__name__"):
            if hasattr(tp, attr):
                return str(getattr(tp, attr))
        return str(tp)
    except Exception:
        return "<unknown type>"


def convert_to_snake_case(name: str) -> str:
    step1 = re.sub(r"(.)([A-Z][a-z]+)", r"\1_\2", name)
    step2 = re.sub(r"([a-z0-9])([A-Z])", r"\1_\2", step1)
    return re.sub(r"__+", "_", step2).lower()


def parse_value_from_string(string: str) -> Any:
    """Python literal if it parses as one, the raw string otherwise."""
    try:
        return ast.literal_eval(string)
    except (ValueError, SyntaxError):
        return str(string)
    except Exception:
        raise ValueError(f"Could not parse '{string}'.")


def prompt_for_value(field_name: str, field_type: Any) -> Any:
    return click.prompt(
        f"\nNo value found for field '{field_name}' of type '{field_type}'. ",
        prompt_suffix="Please enter a value for this parameter:\n> ",
        value_proc=parse_value_from_string,
    )


def prompt_for_component_subclass(component_name: str, classes: Sequence[Type[T]]) -> Type[T]:
    by_name = {cls.__qualname__: cls for cls in classes}
    names = sorted(by_name)

    def to_index(text: str) -> int:
        try:
            idx = int(text) - 1
        except ValueError:
            idx = -1
        if not 0 <= idx < len(names):
            raise click.UsageError(f"Please enter a number between 1 and {len(names)}.")
        return idx

    menu = "\n".join(f"{i + 1})  {n}" for i, n in enumerate(names))
    choice = click.prompt(
        f"\nNo instance found for nested component '{component_name}'. Please choose "
        "from one of the following component subclasses to instantiate:\n" + menu,
        prompt_suffix="\n> ",
        value_proc=to_index,
    )
    return by_name[names[choice]]
