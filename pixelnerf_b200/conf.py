"""HOCON-subset reader standing in for pyhocon (absent from the image).

Covers exactly the syntax the reference's conf/*.conf and expconf.conf use
(SURVEY.md section 5): nested ``name { }`` blocks, ``k = v`` / ``k=v``, ``#``
comments, ``include required("relative/path")``, ``[]`` lists, True/False/true/false,
ints, floats, bare and quoted strings, and deep merge of repeated blocks.  The
getters mirror the pyhocon calls the reference makes (conf.get_int / get_float /
get_bool / get_string / get_list, dotted keys, optional defaults, ``conf["a.b"]``).
"""
import os
import re

_MISSING = object()


class ConfigTree(dict):
    def _lookup(self, key, default=_MISSING):
        cur = self
        for part in str(key).split("."):
            if not isinstance(cur, dict) or not dict.__contains__(cur, part):
                if default is _MISSING:
                    raise KeyError(key)
                return default
            cur = dict.__getitem__(cur, part)
        return cur

    def __getitem__(self, key):
        return self._lookup(key)

    def __contains__(self, key):
        return self._lookup(key, None) is not None or dict.__contains__(self, key)

    def get(self, key, default=None):
        return self._lookup(key, default)

    def get_int(self, key, default=_MISSING):
        v = self._lookup(key, default)
        return v if v is None else int(v)

    def get_float(self, key, default=_MISSING):
        v = self._lookup(key, default)
        return v if v is None else float(v)

    def get_bool(self, key, default=_MISSING):
        v = self._lookup(key, default)
        if isinstance(v, str):
            return v.lower() in ("true", "yes", "on", "1")
        return v if v is None else bool(v)

    def get_string(self, key, default=_MISSING):
        v = self._lookup(key, default)
        return v if v is None else str(v)

    def get_list(self, key, default=_MISSING):
        v = self._lookup(key, default)
        return v if v is None else list(v)

    def get_config(self, key, default=_MISSING):
        return self._lookup(key, default)

    def put(self, key, value):
        cur = self
        parts = str(key).split(".")
        for part in parts[:-1]:
            if not isinstance(dict.get(cur, part), ConfigTree):
                dict.__setitem__(cur, part, ConfigTree())
            cur = dict.__getitem__(cur, part)
        dict.__setitem__(cur, parts[-1], value)

    @staticmethod
    def from_dict(d):
        t = ConfigTree()
        for k, v in d.items():
            dict.__setitem__(t, k, ConfigTree.from_dict(v) if isinstance(v, dict) else v)
        return t


def _merge(dst, src):
    for k, v in src.items():
        if isinstance(v, dict) and isinstance(dict.get(dst, k), dict):
            _merge(dict.__getitem__(dst, k), v)
        else:
            dict.__setitem__(dst, k, v)


_TOKEN = re.compile(r"""
    (?P<ws>[ \t\r]+) | (?P<nl>\n) | (?P<comment>(\#|//)[^\n]*) |
    (?P<str>"(?:[^"\\]|\\.)*") | (?P<punct>[{}\[\]=:,()]) |
    (?P<bare>[^\s{}\[\]=:,()"\#]+)
""", re.X)


def _scalar(tok):
    if tok in ("True", "true"):
        return True
    if tok in ("False", "false"):
        return False
    if tok in ("null", "None"):
        return None
    try:
        return int(tok)
    except ValueError:
        pass
    try:
        return float(tok)
    except ValueError:
        return tok


class _Parser:
    def __init__(self, text, basedir):
        self.toks = [(m.lastgroup, m.group()) for m in _TOKEN.finditer(text)
                     if m.lastgroup not in ("ws", "comment")]
        self.i = 0
        self.basedir = basedir

    def peek(self):
        return self.toks[self.i] if self.i < len(self.toks) else ("eof", "")

    def next(self):
        t = self.peek()
        self.i += 1
        return t

    def skip_sep(self):
        while self.peek()[0] == "nl" or self.peek()[1] == ",":
            self.i += 1

    def parse_object(self, until):
        tree = ConfigTree()
        while True:
            self.skip_sep()
            kind, tok = self.peek()
            if tok == until or kind == "eof":
                self.i += 1
                return tree
            self.next()
            if tok == "include":
                _merge(tree, self.parse_include())
                continue
            key = tok[1:-1] if kind == "str" else tok
            kind2, tok2 = self.peek()
            if tok2 == "{":
                self.next()
                value = self.parse_object("}")
            elif tok2 in ("=", ":"):
                self.next()
                value = self.parse_value()
            else:
                raise ValueError(f"conf: expected '=' or '{{' after {key!r}, got {tok2!r}")
            sub = ConfigTree()
            sub.put(key, value)
            _merge(tree, sub)

    def parse_value(self):
        kind, tok = self.next()
        if tok == "{":
            return self.parse_object("}")
        if tok == "[":
            items = []
            while True:
                self.skip_sep()
                if self.peek()[1] == "]":
                    self.next()
                    return items
                items.append(self.parse_value())
        if kind == "str":
            return tok[1:-1]
        # bare tokens up to end of line form one value ("a b" -> "a b")
        parts = [tok]
        while self.peek()[0] == "bare":
            parts.append(self.next()[1])
        return _scalar(parts[0]) if len(parts) == 1 else " ".join(parts)

    def parse_include(self):
        # include required("path") | include "path"
        kind, tok = self.next()
        if tok == "required":
            assert self.next()[1] == "("
            kind, tok = self.next()
            assert self.next()[1] == ")"
        path = os.path.join(self.basedir, tok[1:-1])
        return ConfigFactory.parse_file(path)


class ConfigFactory:
    @staticmethod
    def parse_string(text, basedir="."):
        return _Parser(text, basedir).parse_object(until="\0")

    @staticmethod
    def parse_file(path):
        with open(path, "r") as f:
            return ConfigFactory.parse_string(f.read(), os.path.dirname(os.path.abspath(path)))

    @staticmethod
    def from_dict(d):
        return ConfigTree.from_dict(d)
