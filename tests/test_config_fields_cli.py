"""Field / ComponentField / PartialComponent / @factory / @task / CLI behaviour, restating the
reference's field_test.py, partial_component_test.py, factory_test.py, task_test.py, cli_test.py and
utils_test.py against zookeeper_b200.core."""
from typing import List, Optional

import pytest
from click import testing

from zookeeper_b200.core import ComponentField, Field, PartialComponent, cli, component, configure, factory, task
from zookeeper_b200.core import utils
from zookeeper_b200.core.factory_registry import FACTORY_REGISTRY


# field_test.py:10-140 ---------------------------------------------------------------------
def test_field_default_validation():
    for bad in ([1, 2, 3], lambda x, y, z: x + y + z, lambda *a: 0):
        with pytest.raises(TypeError, match="If `default` is passed to `Field`, it must be either:"):
            Field(bad)
    assert Field(allow_missing=True).allow_missing
    with pytest.raises(ValueError):
        Field(3.14, allow_missing=True)


def test_field_defaults_and_registration():
    class A:
        foo: int = Field()
        five: int = Field(5)
        none: Optional[int] = Field(None)
        lazy: int = Field(lambda: 7)
        uses_self: List[int] = Field(lambda instance: [instance.baz])
        baz = 2

        @Field
        def decorated() -> float:
            return 3.14

    a = A()
    assert not A.foo.has_default
    with pytest.raises(AttributeError, match="Field 'foo' has no default or configured value"):
        A.foo.get_default(a)
    assert A.five.get_default(a) == 5 and A.none.get_default(a) is None
    assert A.lazy.get_default(a) == 7 and A.uses_self.get_default(a) == [2]
    assert A.decorated.get_default(a) == 3.14 and A.decorated.type is float
    assert repr(A.five) == "<Field 'five' of A with type <class 'int'>>"
    with pytest.raises(TypeError, match="`get_default` must be called with an instance of 'A'"):
        A.five.get_default(object())

    loose = Field(5)
    assert repr(loose) == "<Unregistered Field>"
    for probe in (lambda: loose.has_default, lambda: loose.get_default(object())):
        with pytest.raises(ValueError, match="This field has not been registered to a component"):
            probe()
    # Python <= 3.11 wraps __set_name__ errors in RuntimeError, 3.12 re-raises the ValueError
    with pytest.raises((RuntimeError, ValueError)):

        class B:
            _field = Field()

    with pytest.raises((RuntimeError, TypeError)):

        class C:
            untyped = Field(1)


# field_test.py:143-206 --------------------------------------------------------------------
class AbstractClass:
    a: int = Field()


@component
class ConcreteComponent:
    a: int = Field(2)


def test_component_field_defaults():
    with pytest.raises(TypeError, match="The `default` passed to `ComponentField` must be a component class, not a component instance."):
        ComponentField(ConcreteComponent())
    with pytest.raises(TypeError, match="must be either a component class or a `PartialComponent`"):
        ComponentField(3.14)
    with pytest.raises(TypeError, match="Keyword arguments can only be passed"):
        ComponentField(a=1, b=2, c=3)
    assert ComponentField(allow_missing=True).allow_missing
    with pytest.raises(ValueError):
        ComponentField(3.14, allow_missing=True)

    class A:
        none: AbstractClass = ComponentField()
        cls: AbstractClass = ComponentField(ConcreteComponent)
        partial: AbstractClass = ComponentField(PartialComponent(ConcreteComponent, a=5))
        kwargs: AbstractClass = ComponentField(ConcreteComponent, a=7)

    a = A()
    assert not A.none.has_default
    with pytest.raises(AttributeError, match="ComponentField 'none' has no default or configured component class."):
        A.none.get_default(a)
    assert isinstance(A.cls.get_default(a), ConcreteComponent)
    assert A.partial.get_default(a).a == 5
    assert A.kwargs.get_default(a).a == 7


# partial_component_test.py:30-99 ----------------------------------------------------------
def test_partial_component():
    msg = "The class passed to `PartialComponent` must be a component class."
    for bad in (2.71, (lambda x: x * 2), type("Plain", (), {})):
        with pytest.raises(TypeError, match=msg):
            PartialComponent(bad, a=3)
    with pytest.raises(TypeError, match="`PartialComponent` must be passed component classes, not component instances."):
        PartialComponent(ConcreteComponent(), a=3)
    with pytest.raises(TypeError, match="`PartialComponent` must receive at least one keyword argument."):
        PartialComponent(ConcreteComponent)
    with pytest.raises(TypeError, match="does not correspond to any field"):
        PartialComponent(ConcreteComponent, zzz=3)
    with pytest.raises(TypeError, match="Keyword arguments passed to `PartialComponent` must be either"):
        PartialComponent(ConcreteComponent, a=[1, 2])

    class AbstractChild:
        pass

    @component
    class Child1(AbstractChild):
        a: int = Field()
        b: str = Field("foo")

    @component
    class Parent:
        a: int = Field(10)
        child: AbstractChild = ComponentField()

    p = PartialComponent(Parent, child=Child1)()
    configure(p, {})
    assert isinstance(p.child, Child1) and p.child.b == "foo" and p.child.a == 10
    p = PartialComponent(Parent, child=PartialComponent(Child1, a=5))()
    configure(p, {})
    assert p.child.a == 5
    assert PartialComponent(Child1, a=3)(b="x").b == "x"
    with pytest.raises((RuntimeError, ValueError)):

        class Holder:
            oops = PartialComponent(Child1, a=3)


# utils_test.py:6-69 -----------------------------------------------------------------------
def test_predicates():
    class Plain:
        def build(self) -> object:
            pass

    @component
    class Comp:
        pass

    @factory
    class Fact:
        def build(self) -> object:
            pass

    assert not utils.is_component_class(5) and not utils.is_component_class(lambda: "foo")
    assert not utils.is_component_class(Plain) and utils.is_component_class(Comp)
    assert not utils.is_component_class(Comp()) and utils.is_component_instance(Comp())
    assert not utils.is_component_instance(Comp) and not utils.is_component_instance(Plain())
    assert not utils.is_factory_class(Plain) and utils.is_factory_class(Fact) and not utils.is_factory_class(Fact())
    assert utils.is_factory_instance(Fact()) and not utils.is_factory_instance(Fact) and not utils.is_factory_instance(Plain())
    assert utils.convert_to_snake_case("BinaryNetMnist") == "binary_net_mnist"
    assert utils.convert_to_snake_case("Test_Task") == "test_task" and utils.convert_to_snake_case("testTask") == "test_task"
    assert utils.parse_value_from_string("[1, 2.5]") == [1, 2.5] and utils.parse_value_from_string("a/b@c") == "a/b@c"
    assert utils.is_immutable((1, "a", None)) and not utils.is_immutable([1]) and not utils.is_immutable((1, [2]))


# factory_test.py:6-50 ---------------------------------------------------------------------
def test_factory_validation_and_registry():
    sig_msg = r"Classes decorated with @factory must implement a `build\(\)` method taking precisely one positional argument"
    with pytest.raises(TypeError, match=sig_msg):

        @factory
        class NoBuild:
            pass

    with pytest.raises(TypeError, match=sig_msg):

        @factory
        class TooMany:
            def build(self, a, b, c):
                pass

    with pytest.raises(TypeError, match=r"The `build\(\)` method of a @factory class must have an annotated return type annotation"):

        @factory
        class NoAnnotation:
            def build(self):
                pass

    class AbstractType:
        pass

    calls = []

    @factory
    class Good:
        def build(self) -> AbstractType:
            calls.append(1)
            return AbstractType()

    assert Good in FACTORY_REGISTRY.get(AbstractType, set())
    g = Good()
    assert repr(g) == "<Unconfigured factory 'Good' instance>"
    assert g.build() is g.build() and len(calls) == 1   # built once, cached

    @factory
    class Liar:
        def build(self) -> int:
            return "not an int"

    with pytest.raises(TypeError, match="which is not satisfied by the return value"):
        Liar().build()


# task_test.py:6-53, cli_test.py:28-139 ----------------------------------------------------
@pytest.fixture
def runner():
    saved = dict(cli.commands)
    cli.commands = dict()

    @task
    class TestTask:
        a: int = Field()
        b: str = Field("foo")
        c: bool = Field(False)

        def run(self):
            print(self.a, self.b, self.c)

    yield testing.CliRunner()
    cli.commands = saved


def test_task_validation(runner):
    @task
    class T1:
        def run(self):
            pass

    @task
    class T2:
        @classmethod
        def run(cls):
            pass

    @task
    class T3:
        @staticmethod
        def run():
            pass

    with pytest.raises(TypeError, match="Classes decorated with @task must define a `run` method."):

        @task
        class NoRun:
            pass

    with pytest.raises(TypeError, match=r"^A @task class must define a `run` method taking no arguments except `self`"):

        @task
        class BadRun:
            def run(a, b):
                pass

    with pytest.raises(ValueError, match="Task naming conflict"):

        @task
        class test_task:
            def run(self):
                pass


def test_cli_aliases_values_and_flags(runner):
    for alias in ("test_task", "TestTask", "Test_Task", "testTask"):
        assert runner.invoke(cli, [alias, "a=5"]).exit_code == 0
    assert runner.invoke(cli, ["NotTestTask", "a=5"]).exit_code != 0
    r = runner.invoke(cli, ["test_task", "a=5"])
    assert r.exit_code == 0 and r.output == "5 foo False\n"
    r = runner.invoke(cli, ["test_task", "a=5", "b=bar", "c=True"])
    assert r.output == "5 bar True\n"
    r = runner.invoke(cli, ["test_task", "a=5", "b=https://some-path/foo/bar@somewhere"])
    assert r.exit_code == 0 and r.output == "5 https://some-path/foo/bar@somewhere False\n"
    assert runner.invoke(cli, ["test_task", "a=5", "--c"]).output == "5 foo True\n"
    assert runner.invoke(cli, ["test_task", "a=5", "--no-c"]).output == "5 foo False\n"
    for bad in ("x-y=1.0", "x@y=1.0"):
        assert runner.invoke(cli, ["test_task", "a=5", bad]).exit_code == 2
    assert runner.invoke(cli, ["test_task"]).exit_code == 1          # missing value for 'a'


def test_cli_nested_keys_and_component_selection(runner):
    @component
    class Child:
        x_Y_z: float = Field(0.0)

    @task
    class ParentTask:
        a: int = Field(2)
        child: Child = ComponentField(Child)

        def run(self):
            print(self.a, self.child.x_Y_z)

    r = runner.invoke(cli, ["ParentTask", "a=5", "child.x_Y_z=1.0"])
    assert r.exit_code == 0 and r.output == "5 1.0\n"

    class Base:
        name = "abstract_base"

    @component
    class BaseComponent(Base):
        name = "component_base"

    @factory
    class BaseFactory:
        def build(self) -> Base:
            value = Base()
            value.name = "factory_base"
            return value

    @task
    class PickTask:
        base: Base = ComponentField()

        def run(self):
            print(self.base.name)

    assert runner.invoke(cli, ["pick_task"]).exit_code == 1
    r = runner.invoke(cli, ["pick_task", "base=BaseComponent"])
    assert r.exit_code == 0 and r.output == "component_base\n"
    r = runner.invoke(cli, ["pick_task", "base=base_factory"])
    assert r.exit_code == 0 and r.output == "factory_base\n"
