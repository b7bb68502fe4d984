"""Configuration tokenizer: ordered ``name = value`` pairs.

Grammar (behaviour of reference ``src/utils/config.h:20-189``): whitespace separated
tokens, ``#`` starts a comment to end of line, ``"..."`` is a single-line string and
``'...'`` a multi-line string, both with ``\\`` escapes, ``=`` is its own token.  Order is
preserved and duplicates are kept (consumers apply them in order, so later wins).
"""
from __future__ import annotations

from typing import Iterator, List, Tuple

ConfigPairs = List[Tuple[str, str]]

_EOF = ""


class ConfigError(ValueError):
    pass


class _Reader:
    def __init__(self, text: str):
        self._t = text
        self._i = 0
        self._ch = self._get()

    def _get(self) -> str:
        if self._i >= len(self._t):
            return _EOF
        c = self._t[self._i]
        self._i += 1
        return c

    def _skip_line(self) -> None:
        while True:
            self._ch = self._get()
            if self._ch in (_EOF, "\n", "\r"):
                return

    def _parse_str(self, quote: str, multiline: bool) -> str:
        out = []
        while True:
            self._ch = self._get()
            if self._ch == _EOF:
                raise ConfigError("ConfigReader: unterminated string")
            if self._ch == "\\":
                out.append(self._get())
            elif self._ch == quote:
                return "".join(out)
            elif self._ch in "\r\n" and not multiline:
                raise ConfigError("ConfigReader: unterminated string")
            else:
                out.append(self._ch)

    def next_token(self):
        """Return (token, saw_newline_before_token); token is None at end of input."""
        tok: List[str] = []
        new_line = False
        while self._ch != _EOF:
            c = self._ch
            if c == "#":
                self._skip_line()
                new_line = True
            elif c in "\"'":
                if tok:
                    raise ConfigError("ConfigReader: token followed directly by string")
                s = self._parse_str(c, multiline=(c == "'"))
                self._ch = self._get()
                return s, new_line
            elif c == "=":
                if not tok:
                    self._ch = self._get()
                    return "=", new_line
                return "".join(tok), new_line
            elif c in "\r\n\t ":
                if c in "\r\n" and not tok:
                    new_line = True
                self._ch = self._get()
                if tok:
                    return "".join(tok), new_line
            else:
                tok.append(c)
                self._ch = self._get()
        if tok:
            return "".join(tok), new_line
        return None, True


def iter_config(text: str) -> Iterator[Tuple[str, str]]:
    rd = _Reader(text)
    while True:
        name, _ = rd.next_token()
        if name is None or name == "=":
            return
        eq, nl = rd.next_token()
        if eq is None or nl or eq != "=":
            return
        val, nl = rd.next_token()
        if val is None or nl or val == "=":
            return
        yield name, val


def parse_config_string(text: str) -> ConfigPairs:
    return list(iter_config(text))


def parse_config_file(path: str) -> ConfigPairs:
    try:
        with open(path, "r") as f:
            return parse_config_string(f.read())
    except OSError as e:
        raise ConfigError("cannot open file %s" % path) from e


def parse_cli_overrides(args) -> ConfigPairs:
    """``name=value`` command-line arguments (reference ``src/cxxnet_main.cpp:67-72``)."""
    out: ConfigPairs = []
    for a in args:
        if "=" in a:
            k, v = a.split("=", 1)
            v = v.split()[0] if v.split() else ""
            if k and v:
                out.append((k, v))
    return out
