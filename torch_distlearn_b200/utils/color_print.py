"""Coloured protocol tracing (reference: lua/colorPrint.lua — global ``printServer`` (red) and
``printClient(node, s)`` (blue)).  Plain functions here, not globals, and the client printer has
the arity the library actually calls it with (the reference's example override gets it wrong,
SURVEY §2.9)."""
from __future__ import annotations

import sys

_RED, _BLUE, _RESET = "\033[31m", "\033[34m", "\033[0m"


def _colour() -> bool:
    return sys.stdout.isatty()


def print_server(s: str) -> None:
    """lua/colorPrint.lua:3-9"""
    msg = f"[server] {s}"
    print(f"{_RED}{msg}{_RESET}" if _colour() else msg, flush=True)


def print_client(node: int, s: str) -> None:
    """lua/colorPrint.lua:11-17"""
    msg = f"[client {node}] {s}"
    print(f"{_BLUE}{msg}{_RESET}" if _colour() else msg, flush=True)


printServer, printClient = print_server, print_client
