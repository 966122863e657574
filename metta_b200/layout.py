"""Reads include/mg_config.h so Python and the kernels share one definition of the lowered layout."""

from __future__ import annotations

import os
import re
from types import SimpleNamespace

_HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "mg_config.h")


def _parse(path: str) -> SimpleNamespace:
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    env: dict[str, int] = {}
    pending = []
    for m in re.finditer(r"^#define[ \t]+(\w+)(\([^)]*\))?[ \t]+(.+?)[ \t]*$", text, flags=re.M):
        name, args, expr = m.group(1), m.group(2), m.group(3)
        if args or name.endswith("_H_"):
            continue  # function-like macros are restated below
        pending.append((name, expr))
    for name, expr in pending:
        env[name] = int(eval(expr, {"__builtins__": {}}, env))  # noqa: S307 - our own header
    return SimpleNamespace(**env)


L = _parse(_HEADER)


def list_item(i: int) -> int:
    return 1 + 2 * i


def list_val(i: int) -> int:
    return 2 + 2 * i
