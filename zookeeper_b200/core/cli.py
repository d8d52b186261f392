"""Command-line front end: one click sub-command per @task, ``key=value`` / ``--flag`` /
``--no-flag`` arguments (reference zookeeper/core/cli.py:8-58)."""
from __future__ import annotations

import re

import click

from .utils import convert_to_snake_case, parse_value_from_string

_KEY = r"[\w.]+"


class ConfigParam(click.ParamType):
    """Turns one CLI token into a ``(key, value)`` pair."""

    name = "config"

    def convert(self, value, param, ctx):
        if re.fullmatch(f"--no-{_KEY}", value):
            return value[len("--no-"):], False
        if re.fullmatch(f"--{_KEY}", value):
            return value[2:], True
        key, sep, raw = value.partition("=")
        if not sep or "=" in raw or not re.fullmatch(_KEY, key):
            self.fail(
                "configuration parameters must be of the form 'key=value', where "
                "the key contains only alpha-numeric characters, '_', and '.', "
                f"and the value doesn't contain '='. Received '{value}'.",
                param,
                ctx,
            )
        try:
            parsed = parse_value_from_string(raw)
        except Exception:
            self.fail(
                f"unable to parse value of configuration parameter {value}. The "
                "only supported types are `int`, `float`, `str`, `None`, and "
                "lists/tuples of the above.",
                param,
                ctx,
            )
        return key, parsed


class CamelCaseGroup(click.Group):
    """A command answers to every spelling that snake-cases to the same name."""

    def get_command(self, ctx, cmd_name):
        wanted = convert_to_snake_case(cmd_name)
        for registered in self.list_commands(ctx):
            if convert_to_snake_case(registered) == wanted:
                return super().get_command(ctx, registered)
        return None


@click.group(cls=CamelCaseGroup)
def cli():
    pass
