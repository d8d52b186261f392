"""The ``@component`` decorator and ``configure()``.

Behavioural contract: reference zookeeper/core/component.py:102-695.  A component is an ordinary
class whose class-level ``Field`` / ``ComponentField`` objects describe configurable values.  The
decorator swaps in attribute hooks so that reading a field resolves it lazily from, in order of
priority,
  1. a value given to ``configure()`` on this instance or on the nearest ancestor that declares a
     field of the same name (scoped inheritance),
  2. a keyword argument given to ``__init__`` (or assigned before configuration),
  3. the field's default, else the value of the nearest ancestor declaring the field,
type-checks it against the annotation, and caches it in the instance ``__dict__``.
A field holding a @factory instance reads as the result of its ``build()``.
"""
from __future__ import annotations

import inspect
from typing import Any, Dict, Iterator, List, Optional, Type

from . import utils
from .factory_registry import FACTORY_REGISTRY
from .field import ComponentField, Field

INDENT = " " * 4


# --------------------------------------------------------------------------------------------
# field resolution
# --------------------------------------------------------------------------------------------
def _accept(instance, field: Field, value: Any) -> Any:
    """Type-check ``value`` for ``field`` and cache it on the instance."""
    if not utils.type_check(value, field.type):
        raise TypeError(
            f"Field '{field.name}' of component '{instance.__component_name__}' is "
            f"annotated with type '{field.type}', which is not satisfied by "
            f"value {repr(value)}."
        )
    object.__setattr__(instance, field.name, value)
    return value


def _resolve_field(instance, name: str) -> Any:
    """Raw value of field ``name`` (a @factory instance is returned as such, not built)."""
    cache = instance.__dict__
    if name in cache:
        return cache[name]
    field = instance.__component_fields__[name]
    given_at_init = instance.__component_instantiated_field_values__

    # 1) configured here or on the closest ancestor declaring the same field
    for holder in utils.generate_component_ancestors_with_field(instance, name, include_instance=True):
        configured = holder.__component_configured_field_values__
        if name in configured:
            value = _accept(instance, field, configured[name])
            given_at_init.pop(name, None)
            return value

    # 2) passed to __init__ / assigned before configuration
    if name in given_at_init:
        return _accept(instance, field, given_at_init[name])

    # 3) own default, else whatever the closest ancestor with such a field resolves to
    try:
        value = field.get_default(instance)
    except AttributeError as no_default:
        ancestor = next(utils.generate_component_ancestors_with_field(instance, name), None)
        if ancestor is None:
            raise no_default from None
        try:
            value = _resolve_field(ancestor, name)
        except AttributeError:
            raise no_default from None   # blame this component, not the ancestor
    return _accept(instance, field, value)


def base_getattr(instance, name: str):
    """``getattr`` that does not build factories (used while wiring sub-components)."""
    if utils.is_component_instance(instance) and name in type(instance).__component_fields__:
        return _resolve_field(instance, name)
    return getattr(instance, name)


def _install_attribute_hooks(cls: Type) -> None:
    plain_get = cls.__getattribute__
    plain_set = cls.__setattr__
    plain_del = cls.__delattr__
    plain_dir = cls.__dir__

    def __getattribute__(self, name):
        fields = plain_get(self, "__component_fields__")
        if name not in fields:
            return plain_get(self, name)
        value = _resolve_field(self, name)
        return value.build() if utils.is_factory_instance(value) else value

    def __setattr__(self, name, value):
        if isinstance(getattr(cls, name, None), Field):
            if self.__component_configured__:
                raise ValueError(
                    "Setting already configured component field values directly is "
                    "prohibited. Use Zookeeper component configuration to set field"
                    " values."
                )
            self.__component_instantiated_field_values__[name] = value
            return
        plain_set(self, name, value)

    def __delattr__(self, name):
        if name in self.__component_fields__:
            raise ValueError("Deleting component field values is prohibited.")
        plain_del(self, name)

    def __dir__(self) -> List[str]:
        return sorted(set(plain_dir(self)) | set(self.__component_fields__))

    cls.__getattribute__ = __getattribute__
    cls.__base_getattribute__ = lambda self, name: base_getattr(self, name)
    cls.__setattr__ = __setattr__
    cls.__delattr__ = __delattr__
    cls.__dir__ = __dir__


# --------------------------------------------------------------------------------------------
# printing
# --------------------------------------------------------------------------------------------
def _render_fields(instance, single_line: bool) -> Iterator[str]:
    for name, field in instance.__component_fields__.items():
        try:
            value = base_getattr(instance, name)
        except AttributeError:
            if not field.allow_missing:
                raise
            value = utils.missing
        ancestor = next(utils.generate_component_ancestors_with_field(instance, name), None)
        inherited = (value is not utils.missing and ancestor is not None
                     and base_getattr(ancestor, name) is value)
        if utils.is_component_instance(value):
            if inherited:
                text = "<inherited component instance>"
            elif single_line:
                text = repr(value)
            else:
                text = f"\n{INDENT}".join(str(value).split("\n"))
        elif inherited:
            text = "<inherited value>"
        elif callable(value):
            text = "<callable>"
        elif isinstance(value, str):
            text = f'"{value}"'
        else:
            text = str(value)
        yield f"{name}={text}"


def _unconfigured(instance) -> str:
    return f"<Unconfigured component '{instance.__component_name__}' instance>"


def __component_repr__(self) -> str:
    if not self.__component_configured__:
        return _unconfigured(self)
    return f"{type(self).__name__}({', '.join(_render_fields(self, single_line=True))})"


def __component_str__(self) -> str:
    if not self.__component_configured__:
        return _unconfigured(self)
    body = f",\n{INDENT}".join(_render_fields(self, single_line=False))
    return f"{type(self).__name__}(\n{INDENT}{body}\n)"


def __component_init__(instance, **kwargs):
    """Keyword arguments must name fields; they act as instance-level defaults."""
    for name in kwargs:
        if name not in instance.__component_fields__:
            raise TypeError(
                "Keyword arguments passed to component `__init__` must correspond to "
                f"component fields. Received non-matching argument '{name}'."
            )
    instance.__component_instantiated_field_values__ = dict(kwargs)
    instance.__component_configured_field_values__ = {}
    # names of fields for which some value is reachable from this instance (grows in configure)
    instance.__component_fields_with_values_in_scope__ = {
        f.name for f in instance.__component_fields__.values() if f.has_default
    } | set(kwargs)


# --------------------------------------------------------------------------------------------
# @component
# --------------------------------------------------------------------------------------------
def _check_post_configure(cls: Type) -> None:
    hook = getattr(cls, "__post_configure__", None)
    if hook is None and not hasattr(cls, "__post_configure__"):
        return
    if not callable(hook):
        raise TypeError("The `__post_configure__` attribute of a @component class must be a method.")
    params = inspect.signature(hook).parameters
    if len(params) > 1 or (len(params) == 1 and "self" not in params):
        raise TypeError(
            "The `__post_configure__` method of a @component class must take no "
            f"arguments except `self`, but `{cls.__name__}.__post_configure__` "
            f"accepts arguments {tuple(params)}."
        )


def component(cls: Type):
    """Class decorator turning ``cls`` into a configurable component."""
    if not inspect.isclass(cls):
        raise TypeError("Only classes can be decorated with @component.")
    if inspect.isabstract(cls):
        raise TypeError("Abstract classes cannot be decorated with @component.")
    if utils.is_component_class(cls):
        raise TypeError(
            f"The class {cls.__name__} is already a component; the @component decorator "
            "cannot be applied again."
        )
    if cls.__init__ not in (object.__init__, __component_init__):
        raise TypeError("Component classes must not define a custom `__init__` method.")
    cls.__init__ = __component_init__
    _check_post_configure(cls)

    # fields of the whole MRO, base classes first so that subclasses override
    fields: Dict[str, Field] = {}
    for klass in reversed(inspect.getmro(cls)):
        fields.update({n: v for n, v in vars(klass).items() if isinstance(v, Field)})
    if not fields:
        utils.warn(f"Component {cls.__name__} has no defined fields.")
    for name, field in fields.items():
        current = getattr(cls, name)
        if not isinstance(current, Field):
            owner = field.host_component_class
            raise ValueError(
                f"Field '{name}' is defined on super-class {owner.__name__}. "
                f"In subclass {cls.__name__}, '{name}' has been overriden with value: "
                f"{current}.\n\n"
                f"If you wish to change the default value of field '{name}' in a "
                f"subclass of {owner.__name__}, please wrap the new default "
                "value in a new `Field` instance."
            )
    cls.__component_fields__ = fields

    _install_attribute_hooks(cls)
    cls.__str__ = __component_str__
    cls.__repr__ = __component_repr__
    cls.__component_name__ = cls.__name__
    cls.__component_parent__ = None
    cls.__component_configured__ = False
    return cls


# --------------------------------------------------------------------------------------------
# configure()
# --------------------------------------------------------------------------------------------
def _candidates_for(field: ComponentField) -> List[Type]:
    """Component classes and @factory classes able to fill ``field``."""
    found = list(utils.generate_component_subclasses(field.type))
    for sub in utils.generate_subclasses(field.type):
        found.extend(FACTORY_REGISTRY.get(sub, []))
    return found


def _class_named(text: str, candidates: List[Type]) -> Optional[Type]:
    snake = utils.convert_to_snake_case(text)
    for cand in candidates:
        if text in (cand.__name__, cand.__qualname__) or snake == utils.convert_to_snake_case(cand.__name__):
            return cand
    return None


def _type_label(tp: Any) -> str:
    return tp.__name__ if inspect.isclass(tp) else str(tp)


def _no_class_error(full_name: str, label: str, candidates: List[Type]) -> ValueError:
    if candidates:
        listing = "\n    ".join([""] + [utils.type_name_str(c) for c in candidates])
        return ValueError(
            f"Component field '{full_name}' of type '{label}' "
            f"has no default or configured class. Please configure "
            f"'{full_name}' with one of the following @component or "
            "@factory classes:" + listing
        )
    return ValueError(
        f"Component field '{full_name}' of type '{label}' "
        f"has no default or configured class. No defined @component "
        "or @factory class satisfies this type. Please define an "
        f"@component class subclassing '{label}', or an "
        "@factory class with a `build()` method returning a "
        f"'{label}' instance. This class must be imported "
        "before invoking `configure()`."
    )


def _unknown_key_error(key: str, component_name: str) -> ValueError:
    return ValueError(
        f"Key '{key}' does not correspond to any field of component "
        f"'{component_name}'."
        "\n\n"
        "If you have nested components as follows:\n\n"
        "```\n"
        "@component\n"
        "class ChildComponent:\n"
        "    a: int = Field(0)\n"
        "\n"
        "@task\n"
        "class SomeTask:\n"
        "    child: ChildComponent = ComponentField(ChildComponent)\n"
        "    def run(self):\n"
        "        print(self.child.a)\n"
        "```\n\n"
        "then trying to configure `a=<SOME_VALUE>` will fail. You instead need to "
        "fully qualify the key name, and configure the value with "
        "`child.a=<SOME_VALUE>`."
    )


def _settle_field(instance, field: Field, conf: Dict[str, Any], interactive: bool) -> None:
    """Decide where the value of one field comes from; consume its key from ``conf``."""
    full_name = f"{instance.__component_name__}.{field.name}"
    label = _type_label(field.type)
    nested = isinstance(field, ComponentField)
    candidates = _candidates_for(field) if nested else []
    configured = instance.__component_configured_field_values__

    if field.name in conf:
        value = conf.pop(field.name)
        if nested and candidates and isinstance(value, str):
            chosen = _class_named(value, candidates)
            if chosen is not None:
                value = chosen()
        configured[field.name] = value
    elif field.name in instance.__component_fields_with_values_in_scope__ or field.allow_missing:
        return_early = field.name not in instance.__component_fields_with_values_in_scope__
        if return_early:
            return
    elif nested and len(candidates) == 1:
        only = candidates[0]
        utils.warn(
            f"'{utils.type_name_str(only)}' is the only concrete component "
            f"class that satisfies the type of the annotated field '{full_name}'. "
            "Using an instance of this class by default.",
        )
        configured[field.name] = only()
    elif interactive:
        if nested:
            if not candidates:
                raise ValueError(
                    "No component or factory class is defined which satisfies the "
                    f"type {label} of field {full_name}. If such a class "
                    "has been defined, it must be imported before calling "
                    "`configure`."
                )
            configured[field.name] = utils.prompt_for_component_subclass(full_name, candidates)()
        else:
            configured[field.name] = utils.prompt_for_value(full_name, field.type)
    elif nested:
        raise _no_class_error(full_name, label, candidates)
    else:
        raise ValueError(
            "No configuration value found for annotated field "
            f"'{full_name}' of type '{label}'."
        )
    instance.__component_fields_with_values_in_scope__.add(field.name)


def configure(instance, conf: Dict[str, Any], name: Optional[str] = None, interactive: bool = False):
    """Inject ``conf`` into ``instance`` and, recursively, into its sub-components.

    ``conf`` maps (possibly dotted) field names to values; values from ``conf`` beat ``__init__``
    arguments and defaults.  Consumed keys are removed from ``conf``.  A component can be
    configured once; afterwards its fields are read-only."""
    if not utils.is_component_instance(instance):
        raise TypeError(
            "Only @component, @factory, and @task instances can be configured. "
            f"Received: {instance}."
        )
    if instance.__component_configured__:
        raise ValueError(f"Component '{instance.__component_name__}' has already been configured.")
    if name is not None:
        instance.__component_name__ = name

    fields = instance.__component_fields__
    for field in fields.values():
        _settle_field(instance, field, conf, interactive)

    # whatever is left must be addressed to a sub-component ("child.key")
    for key in conf:
        head = key.split(".")[0] if "." in key else None
        ok = head is not None and isinstance(fields.get(head), ComponentField)
        if not ok:
            raise _unknown_key_error(key, instance.__component_name__)

    for field in fields.values():
        if not isinstance(field, ComponentField):
            continue
        try:
            child = base_getattr(instance, field.name)
        except AttributeError:
            if field.allow_missing:
                continue
            raise
        if not utils.is_component_instance(child) or child.__component_configured__:
            continue
        child.__component_parent__ = instance
        child.__component_fields_with_values_in_scope__ |= instance.__component_fields_with_values_in_scope__
        prefix = field.name + "."
        scoped = {k[len(prefix):]: v for k, v in conf.items() if k.startswith(prefix)}
        configure(child, scoped, name=f"{instance.__component_name__}.{field.name}", interactive=interactive)

    instance.__component_configured__ = True
    if hasattr(type(instance), "__post_configure__"):
        instance.__post_configure__()
