"""Behavioural spec of the configuration runtime (@component / configure), restating what the
reference's zookeeper/core/component_test.py checks (cited per test), against zookeeper_b200.core."""
import abc
from typing import List, Tuple
from unittest.mock import patch

import click
import pytest

from zookeeper_b200.core import ComponentField, Field, component, configure, factory


def make_ab():
    @component
    class A:
        a: int = Field()
        b: str = Field("foo")

    return A


# component_test.py:23-66 ------------------------------------------------------------------
def test_decorator_rejects_functions_abstract_classes_and_custom_init():
    with pytest.raises(TypeError, match="Only classes can be decorated with @component."):
        component(lambda: None)

    class Abstract(abc.ABC):
        @abc.abstractmethod
        def foo(self):
            pass

    with pytest.raises(TypeError, match="Abstract classes cannot be decorated with @component."):
        component(Abstract)

    class WithInit:
        def __init__(self, a, b=5):
            self.a, self.b = a, b

    with pytest.raises(TypeError, match="Component classes must not define a custom `__init__` method."):
        component(WithInit)

    A = make_ab()
    with pytest.raises(TypeError, match="is already a component"):
        component(A)


# component_test.py:68-98 ------------------------------------------------------------------
def test_generated_init_takes_field_kwargs_only():
    A = make_ab()
    x = A(a=2)
    assert (x.a, x.b) == (2, "foo")
    x = A(a=0, b="bar")
    assert (x.a, x.b) == (0, "bar")
    with pytest.raises(TypeError, match=r"takes 1 positional argument but 2 were given"):
        A("foobar")
    with pytest.raises(TypeError, match="Keyword arguments passed to component `__init__` must correspond to "
                                        "component fields. Received non-matching argument 'some_other_field_name'"):
        A(some_other_field_name=0)


# component_test.py:101-156 ----------------------------------------------------------------
def test_configure_overrides_and_scoping():
    A = make_ab()
    x = A()
    configure(x, {"a": 0, "b": "bar"})
    assert (x.a, x.b) == (0, "bar")

    @component
    class Child:
        a: int = Field()
        b: str = Field()
        c: List[float] = Field()

    @component
    class Parent:
        b: str = Field("bar")
        child: Child = ComponentField(Child)

    @component
    class GrandParent:
        a: int = Field()
        b: str = Field()
        parent: Parent = ComponentField(Parent)

    g = GrandParent()
    conf = {"a": 10, "parent.child.a": 15, "b": "foo", "parent.child.b": "baz", "parent.child.c": [0, 4.2]}
    configure(g, conf)
    assert g.a == 10 and g.parent.child.a == 15
    assert g.b == "foo" and g.parent.b == "foo" and g.parent.child.b == "baz"
    assert g.parent.child.c == [0, 4.2]
    assert "a" not in conf and "b" not in conf       # consumed keys are removed (component.py:528)
    with pytest.raises(ValueError, match="has already been configured"):
        configure(g, {})


# component_test.py:159-196 ----------------------------------------------------------------
def test_single_concrete_subclass_is_instantiated_automatically(capsys):
    class AbstractChild:
        pass

    @component
    class Child1(AbstractChild):
        pass

    @component
    class Parent:
        child: AbstractChild = ComponentField()

    p = Parent()
    configure(p, {})
    assert isinstance(p.child, Child1)
    assert "is the only concrete component class that satisfies the type" in capsys.readouterr().err

    @component
    class Child2(AbstractChild):
        pass

    with pytest.raises(ValueError, match=r"^Component field 'Parent.child' of type 'AbstractChild' has no default or configured class."):
        configure(Parent(), {})


# component_test.py:199-277 ----------------------------------------------------------------
def test_missing_value_errors_and_interactive_prompts():
    A = make_ab()
    with pytest.raises(ValueError, match=r"^No configuration value found for annotated field 'FAKE_NAME.a' of type 'int'."):
        configure(A(), {"b": "bar"}, name="FAKE_NAME")

    x = A()
    with patch("click.prompt", return_value=42) as prompt:
        configure(x, {"b": "bar"}, name="FAKE_NAME", interactive=True)
    assert (x.a, x.b) == (42, "bar")
    prompt.assert_called_once()

    class AbstractChild:
        pass

    @component
    class Child1(AbstractChild):
        pass

    @component
    class Child2(AbstractChild):
        pass

    class Child3_Abstract(AbstractChild):
        pass

    @component
    class Child3A(Child3_Abstract):
        pass

    @component
    class Child3B(Child3_Abstract):
        pass

    @component
    class Parent:
        child: AbstractChild = ComponentField()

    for i, expected in enumerate([Child1, Child2, Child3A, Child3B]):  # alphabetical, components only
        p = Parent()
        with patch("click.prompt", return_value=i) as prompt:
            configure(p, {}, interactive=True)
        assert isinstance(p.child, expected)
        prompt.assert_called_once()


# component_test.py:280-331 ----------------------------------------------------------------
def test_str_and_repr_golden_strings():
    @component
    class Child1:
        a: int = Field()

    @component
    class Child2:
        a: int = Field()
        b: str = Field()
        c: List[float] = Field()
        d: int = Field(allow_missing=True)
        child_1: Child1 = ComponentField()

    @component
    class Parent:
        b: str = Field("bar")
        child_1: Child1 = ComponentField(Child1)
        child_2: Child2 = ComponentField(Child2)

    p = Parent()
    assert repr(p) == "<Unconfigured component 'Parent' instance>"
    configure(p, {"child_1.a": 5, "child_2.a": 10, "b": "foo", "child_2.c": [1.5, -1.2]}, name="parent")
    assert click.unstyle(repr(p)) == (
        'Parent(b="foo", child_1=Child1(a=5), child_2=Child2(a=10, b=<inherited value>, '
        "c=[1.5, -1.2], d=<missing>, child_1=<inherited component instance>))"
    )
    assert click.unstyle(str(p)) == """Parent(
    b="foo",
    child_1=Child1(
        a=5
    ),
    child_2=Child2(
        a=10,
        b=<inherited value>,
        c=[1.5, -1.2],
        d=<missing>,
        child_1=<inherited component instance>
    )
)"""


# component_test.py:334-358 ----------------------------------------------------------------
def test_type_check_on_read_and_override_guard():
    A = make_ab()
    x = A()
    configure(x, {"a": 4.5}, name="x")
    with pytest.raises(TypeError, match="Field 'a' of component 'x' is annotated with type '<class 'int'>', which is not satisfied by value 4.5."):
        x.a

    @component
    class Super:
        foo: str = Field("bar")

    with pytest.raises(ValueError, match="Field 'foo' is defined on super-class"):

        @component
        class Sub(Super):
            foo = 1


# component_test.py:361-460 ----------------------------------------------------------------
def test_factory_values_type_check_and_inheritance(capsys):
    class Base:
        pass

    class Concrete(Base):
        pass

    @factory
    class F1:
        def build(self) -> Base:
            return Concrete()

    @factory
    class F2:
        def build(self) -> Concrete:
            return Concrete()

    @factory
    class F3:
        def build(self) -> Tuple[int, int, int]:
            return (1, 2, 3)

    @component
    class A1:
        base: Base = ComponentField(F1)

    @component
    class A2:
        base: Base = ComponentField(F2)

    @component
    class A3:
        base: Tuple[float, float, float] = ComponentField(F3)

    assert isinstance(A1().base, Concrete) and isinstance(A2().base, Concrete)
    capsys.readouterr()
    assert A3().base == (1, 2, 3)
    assert capsys.readouterr().err == (
        "WARNING: Unable to check that typing.Tuple[int, int, int] is a sub-type of "
        "typing.Tuple[float, float, float].\n"
    )

    # a child field resolved through a @factory parent (larq/zookeeper issue 121)
    @component
    class Child:
        x: int = Field()

    @factory
    class Factory:
        child: Child = ComponentField(Child)
        x: int = Field(5)

        def build(self) -> int:
            return self.child.x

    f = Factory()
    configure(f, {})
    assert f.child.x == 5 and f.build() == 5

    # inherited factory value (issue 123)
    @factory
    class IntFactory:
        def build(self) -> int:
            return 5

    @component
    class Kid:
        x: int = ComponentField()

    @component
    class Mum:
        child: Kid = ComponentField(Kid)
        x: int = ComponentField(IntFactory)

    p = Mum()
    configure(p, {})
    assert p.x == 5 and p.child.x == 5
    p = Mum()
    configure(p, {"child.x": 7})
    assert p.x == 5 and p.child.x == 7


# component_test.py:463-496 ----------------------------------------------------------------
def test_post_configure_hook():
    with pytest.raises(TypeError, match="The `__post_configure__` attribute of a @component class must be a method."):

        @component
        class A:
            __post_configure__ = 3.14

    with pytest.raises(TypeError, match="The `__post_configure__` method of a @component class must take no arguments except `self`"):

        @component
        class B:
            def __post_configure__(self, x):
                pass

    @component
    class C:
        a: int = Field(0)
        b: float = Field(3.14)

        def __post_configure__(self):
            self.c = self.a + self.b

    c = C()
    configure(c, {"a": 1, "b": -3.14})
    assert c.c == 1 - 3.14


# component_test.py:499-572 ----------------------------------------------------------------
def test_unknown_keys_and_non_component_targets():
    @component
    class Child:
        a: int = Field(1)

    @component
    class Parent:
        b: str = Field("foo")
        child: Child = ComponentField(Child)

    @component
    class GrandParent:
        c: float = Field(3.14)
        parent: Parent = ComponentField(Parent)

    with pytest.raises(ValueError, match="Key 'd' does not correspond to any field of component 'GrandParent'."):
        configure(GrandParent(), {"d": "bar"}, name="GrandParent")
    with pytest.raises(ValueError, match="Key 'b' does not correspond to any field of component 'GrandParent'."):
        configure(GrandParent(), {"b": "bar"})
    g = GrandParent()
    configure(g, {"parent.b": "bar"})
    assert g.parent.b == "bar"
    with pytest.raises(ValueError, match="Key 'non_existent_field' does not correspond to any field of component 'GrandParent.parent'."):
        configure(GrandParent(), {"parent.non_existent_field": "bar"})

    class Plain:
        a: int = Field()

    msg = "Only @component, @factory, and @task instances can be configured."
    with pytest.raises(TypeError, match=msg):
        configure(Plain(), conf={"a": 5})

    @component
    class B:
        b: int = Field()

    with pytest.raises(TypeError, match=msg):
        configure(B, conf={"b": 3})          # a class, not an instance

    class SubOfComponent(B):
        c: int = Field()

    with pytest.raises(TypeError, match=msg):
        configure(SubOfComponent(), conf={"b": 3, "c": 42})   # undecorated subclass


# component_test.py:575-646 ----------------------------------------------------------------
def test_allow_missing_fields():
    @component
    class A:
        a: int = Field()
        b: float = Field(allow_missing=True)

        @Field
        def c(self) -> float:
            if hasattr(self, "b"):
                return self.b
            return self.a

    with pytest.raises(ValueError, match="No configuration value found for annotated field 'A.a' of type 'int'."):
        configure(A(), {"b": 3.14})
    x = A()
    configure(x, {"a": 0})
    assert x.c == 0
    x = A()
    configure(x, {"a": 0, "b": 3.14})
    assert x.c == 3.14

    class Base:
        a: int = Field()

    @component
    class Child1(Base):
        a = Field(5)

    @component
    class Child2(Base):
        a = Field(5)

    @component
    class Parent:
        child: Base = ComponentField()
        child_allow_missing: Base = ComponentField(allow_missing=True)

    with pytest.raises(ValueError, match="Component field 'Parent.child' of type 'Base' has no default or configured class."):
        configure(Parent(), {"child_allow_missing": "Child2"})
    p = Parent()
    configure(p, {"child": "Child1"})
    assert not hasattr(p, "child_allow_missing")

    @component
    class Kid:
        a: int = Field(allow_missing=True)

    @component
    class Mum:
        a: int = Field(5)
        child: Kid = ComponentField(Kid)

    m = Mum()
    configure(m, {})
    assert m.child.a == 5


# component_test.py:649-670 ----------------------------------------------------------------
def test_setattr_delattr_rules():
    @component
    class A:
        a: int = Field(6)
        b: float = Field(allow_missing=True)

    x = A()
    x.a = 3
    x.b = 5.0
    configure(x, {"a": 0})
    assert (x.a, x.b) == (0, 5.0)
    with pytest.raises(ValueError, match="Setting already configured component field values directly is "
                                         "prohibited. Use Zookeeper component configuration to set field values."):
        x.a = 5
    with pytest.raises(ValueError, match="Deleting component field values is prohibited."):
        del x.a
    assert {"a", "b"} <= set(dir(x))
