"""``@task``: a component with an argument-less ``run()`` that is exposed as a CLI sub-command
(reference zookeeper/core/task.py:10-61)."""
from __future__ import annotations

import inspect

import click

from .cli import ConfigParam, cli
from .component import component, configure
from .utils import convert_to_snake_case


def task(cls):
    cls = component(cls)
    run = getattr(cls, "run", None)
    if not callable(run):
        raise TypeError("Classes decorated with @task must define a `run` method.")
    params = inspect.signature(run).parameters
    if len(params) > 1 or (len(params) == 1 and "self" not in params):
        raise TypeError(
            "A @task class must define a `run` method taking no arguments except "
            f"`self`, which runs the task, but `{cls.__name__}.run` accepts arguments "
            f"{tuple(params)}."
        )
    taken = {convert_to_snake_case(existing) for existing in cli.commands}
    if convert_to_snake_case(cls.__name__) in taken:
        raise ValueError(
            f"Task naming conflict. Task with name '{cls.__name__}' (or similar) "
            "already registered. Note that the task name is the name of the class that "
            "the @task decorator is applied to."
        )

    @cli.command(cls.__name__, context_settings=dict(ignore_unknown_options=True))
    @click.option("-i", "--interactive", is_flag=True, default=False, help="Interactively configure task.")
    @click.argument("config", type=ConfigParam(), nargs=-1)
    def command(config, interactive):
        instance = cls()
        configure(instance, dict(config), interactive=interactive)
        instance.run()

    return cls
