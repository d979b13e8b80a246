"""``tf.app.flags`` / ``tf.app.run`` equivalents (templates/00_mnist_replica.py:48-86,274-275;
run_mnist_distributed.py:182): ``DEFINE_string/integer/float/boolean`` register flags on a global
``FLAGS`` object that parses ``sys.argv`` lazily on first attribute access; ``app.run(main, argv)``
parses known flags and calls ``main(leftover_argv)``. Booleans accept ``--flag``, ``--noflag``,
``--flag=true/false``."""
from __future__ import annotations

import argparse
import sys
from typing import Any, Callable, List, Optional


class _FlagValues:
    def __init__(self):
        self.__dict__["_parser"] = argparse.ArgumentParser(allow_abbrev=False, add_help=True)
        self.__dict__["_parsed"] = None
        self.__dict__["_defs"] = {}
        self.__dict__["_unparsed"] = []

    def _define(self, name: str, default: Any, help_: str, kind: str) -> None:
        if name in self._defs:
            raise ValueError("Duplicate flag %s" % name)
        self._defs[name] = (default, kind)
        if kind == "bool":
            self._parser.add_argument("--" + name, nargs="?", const=True, default=default,
                                      type=lambda v: str(v).lower() in ("1", "true", "t", "yes", "y"),
                                      help=help_)
            self._parser.add_argument("--no" + name, dest=name, action="store_false", help=argparse.SUPPRESS)
        else:
            typ = {"string": str, "integer": int, "float": float}[kind]
            self._parser.add_argument("--" + name, type=typ, default=default, help=help_)
        self.__dict__["_parsed"] = None

    def _parse(self, argv: Optional[List[str]] = None) -> List[str]:
        args = sys.argv[1:] if argv is None else list(argv)
        ns, unparsed = self._parser.parse_known_args(args)
        self.__dict__["_parsed"] = ns
        self.__dict__["_unparsed"] = unparsed
        return unparsed

    def __getattr__(self, name: str):
        if name.startswith("_"):
            raise AttributeError(name)
        if self._parsed is None:
            self._parse()
        try:
            return getattr(self._parsed, name)
        except AttributeError:
            raise AttributeError("Unknown command line flag %r" % name)

    def __setattr__(self, name: str, value) -> None:
        if self._parsed is None:
            self._parse()
        setattr(self._parsed, name, value)

    def flag_values_dict(self) -> dict:
        if self._parsed is None:
            self._parse()
        return dict(vars(self._parsed))

    def _reset(self) -> None:
        self.__init__()


FLAGS = _FlagValues()


def DEFINE_string(name, default, help):  # noqa: A002
    FLAGS._define(name, default, help, "string")


def DEFINE_integer(name, default, help):  # noqa: A002
    FLAGS._define(name, default, help, "integer")


def DEFINE_float(name, default, help):  # noqa: A002
    FLAGS._define(name, default, help, "float")


def DEFINE_boolean(name, default, help):  # noqa: A002
    FLAGS._define(name, default, help, "bool")


DEFINE_bool = DEFINE_boolean


def app_run(main: Optional[Callable] = None, argv: Optional[List[str]] = None) -> None:
    """Parses known flags, then calls ``main(argv_with_leftovers)`` and exits with its result."""
    args = sys.argv if argv is None else argv
    unparsed = FLAGS._parse(args[1:])
    main = main or sys.modules["__main__"].main
    sys.exit(main(args[:1] + unparsed))
