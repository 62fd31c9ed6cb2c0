"""GBNF grammars for constrained decoding.

Parity: `llama.cpp/src/llama-grammar.cpp` (parser `:200-520`, stack machine `llama_grammar_advance_stack` /
`llama_grammar_accept`), the `grammar` sampler (`llama-sampling.cpp:2625`) and `json-schema-to-grammar.cpp` for the JSON
subset (`json_schema_to_gbnf` below covers object / array / string / number / integer / boolean / null / enum / const).

Representation: every rule is a list of alternatives, every alternative a list of elements; an element is
`("char", ranges, negated)` (code-point ranges), `("any",)` or `("ref", rule_name)`.  Repetitions and groups are rewritten
into generated helper rules at parse time, as the reference does.  Matching keeps a *set of stacks*; a stack is a tuple of
pending element lists (innermost first).  A token is acceptable when feeding its characters leaves at least one stack alive;
the grammar is complete when some stack is empty.
"""
from __future__ import annotations

import json
from functools import lru_cache

Element = tuple
Stack = tuple  # tuple of tuples of Elements (innermost frame first)


class GrammarError(ValueError):
    pass


class _Parser:
    def __init__(self, text: str):
        self.s = text
        self.i = 0
        self.rules: dict[str, list[list[Element]]] = {}
        self.n_gen = 0

    # -- lexing helpers --
    def ws(self, newlines: bool = True):
        s = self.s
        while self.i < len(s):
            c = s[self.i]
            if c == "#":
                while self.i < len(s) and s[self.i] != "\n":
                    self.i += 1
            elif c in " \t" or (newlines and c in "\r\n"):
                self.i += 1
            else:
                break

    def name(self) -> str:
        j = self.i
        while j < len(self.s) and (self.s[j].isalnum() or self.s[j] in "-_"):
            j += 1
        if j == self.i:
            raise GrammarError(f"expected a rule name at offset {self.i}")
        n = self.s[self.i:j]
        self.i = j
        return n

    def char(self) -> int:
        s = self.s
        c = s[self.i]
        if c != "\\":
            self.i += 1
            return ord(c)
        e = s[self.i + 1]
        self.i += 2
        if e == "x":
            v = int(s[self.i:self.i + 2], 16); self.i += 2; return v
        if e == "u":
            v = int(s[self.i:self.i + 4], 16); self.i += 4; return v
        if e == "U":
            v = int(s[self.i:self.i + 8], 16); self.i += 8; return v
        return ord({"n": "\n", "r": "\r", "t": "\t"}.get(e, e))

    def gen(self, base: str) -> str:
        self.n_gen += 1
        return f"{base}_{self.n_gen}"

    # -- grammar --
    def parse(self) -> dict[str, list[list[Element]]]:
        self.ws()
        while self.i < len(self.s):
            n = self.name()
            self.ws()
            if self.s[self.i:self.i + 3] != "::=":
                raise GrammarError(f"expected '::=' after rule {n!r}")
            self.i += 3
            self.ws()
            self.rules[n] = self.alternates(n, nested=False)
            self.ws()
        if "root" not in self.rules:
            raise GrammarError("grammar has no 'root' rule")
        for alts in self.rules.values():
            for alt in alts:
                for el in alt:
                    if el[0] == "ref" and el[1] not in self.rules:
                        raise GrammarError(f"undefined rule {el[1]!r}")
        return self.rules

    def alternates(self, rule: str, nested: bool) -> list[list[Element]]:
        alts = [self.sequence(rule, nested)]
        while self.i < len(self.s) and self.s[self.i] == "|":
            self.i += 1
            self.ws()
            alts.append(self.sequence(rule, nested))
        return alts

    def sequence(self, rule: str, nested: bool) -> list[Element]:
        s = self.s
        seq: list[Element] = []
        last_start = 0          # index in seq where the last item (possibly several literal chars) begins
        while self.i < len(s):
            c = s[self.i]
            if c == '"':
                self.i += 1
                last_start = len(seq)
                while s[self.i] != '"':
                    cp = self.char()
                    seq.append(("char", ((cp, cp),), False))
                self.i += 1
                self.ws(nested)
            elif c == "[":
                self.i += 1
                neg = s[self.i] == "^"
                if neg:
                    self.i += 1
                ranges = []
                while s[self.i] != "]":
                    lo = self.char()
                    hi = lo
                    if s[self.i] == "-" and s[self.i + 1] != "]":
                        self.i += 1
                        hi = self.char()
                    ranges.append((lo, hi))
                self.i += 1
                last_start = len(seq)
                seq.append(("char", tuple(ranges), neg))
                self.ws(nested)
            elif c == ".":
                self.i += 1
                last_start = len(seq)
                seq.append(("any",))
                self.ws(nested)
            elif c == "(":
                self.i += 1
                self.ws()
                sub = self.gen(rule)
                self.rules[sub] = self.alternates(rule, nested=True)
                if s[self.i] != ")":
                    raise GrammarError(f"expected ')' at offset {self.i}")
                self.i += 1
                last_start = len(seq)
                seq.append(("ref", sub))
                self.ws(nested)
            elif c.isalnum() or c in "-_":
                save = self.i
                n = self.name()
                self.ws(nested)
                if s[self.i:self.i + 3] == "::=":     # start of the next rule (only reachable when newlines are skipped)
                    self.i = save
                    break
                last_start = len(seq)
                seq.append(("ref", n))
            elif c in "*+?{":
                if last_start >= len(seq):
                    raise GrammarError(f"repetition with nothing to repeat at offset {self.i}")
                lo, hi = {"*": (0, None), "+": (1, None), "?": (0, 1)}.get(c, (None, None))
                if c == "{":
                    j = s.index("}", self.i)
                    body = s[self.i + 1:j]
                    self.i = j + 1
                    if "," in body:
                        a, b = body.split(",", 1)
                        lo, hi = int(a or 0), (int(b) if b.strip() else None)
                    else:
                        lo = hi = int(body)
                else:
                    self.i += 1
                self.ws(nested)
                item = seq[last_start:]
                del seq[last_start:]
                seq.extend(self.repeat(rule, item, lo, hi))
                last_start = len(seq)
            else:
                break
        return seq

    def repeat(self, rule: str, item: list[Element], lo: int, hi: int | None) -> list[Element]:
        out: list[Element] = []
        for _ in range(lo):
            out.extend(item)
        if hi is None:           # item*  ==>  star ::= item star | <empty>
            star = self.gen(rule)
            self.rules[star] = [list(item) + [("ref", star)], []]
            out.append(("ref", star))
        else:                    # up to (hi - lo) optional copies, nested so that the match stays a prefix
            prev = None
            for _ in range(hi - lo):
                opt = self.gen(rule)
                self.rules[opt] = [list(item) + ([("ref", prev)] if prev else []), []]
                prev = opt
            if prev:
                out.append(("ref", prev))
        return out


def _char_matches(el: Element, cp: int) -> bool:
    if el[0] == "any":
        return True
    hit = any(lo <= cp <= hi for lo, hi in el[1])
    return hit != el[2]


class Grammar:
    """Compiled GBNF grammar + matcher state."""

    def __init__(self, text: str, _rules=None, _stacks=None):
        self.rules = _rules if _rules is not None else _Parser(text).parse()
        self.text = text
        if _stacks is None:
            _stacks = set()
            for alt in self.rules["root"]:
                self._expand((tuple(alt),), _stacks, set())
        self.stacks: set[Stack] = _stacks

    def clone(self) -> "Grammar":
        return Grammar(self.text, self.rules, set(self.stacks))

    # expand the innermost frame until it starts with a terminal (or the stack is empty = grammar complete)
    def _expand(self, stack: Stack, out: set, seen: set):
        if stack in seen:
            return
        seen.add(stack)
        while stack and not stack[0]:
            stack = stack[1:]
        if not stack:
            out.add(())
            return
        head = stack[0][0]
        if head[0] != "ref":
            out.add(stack)
            return
        rest = (stack[0][1:],) + stack[1:]
        for alt in self.rules[head[1]]:
            self._expand((tuple(alt),) + rest, out, seen)

    @staticmethod
    def _step(rules_expand, stacks: set, cp: int) -> set:
        nxt: set = set()
        seen: set = set()
        for st in stacks:
            if not st:
                continue
            if _char_matches(st[0][0], cp):
                rules_expand((st[0][1:],) + st[1:], nxt, seen)
        return nxt

    def _feed(self, stacks: set, text: str) -> set:
        for ch in text:
            stacks = self._step(self._expand, stacks, ord(ch))
            if not stacks:
                break
        return stacks

    # -- public --
    def accepts(self, text: str) -> bool:
        """Would feeding `text` keep the grammar alive?"""
        return bool(text) and bool(self._feed(self.stacks, text))

    def accept(self, text: str):
        st = self._feed(self.stacks, text)
        if not st:
            raise GrammarError(f"text {text!r} is not accepted by the grammar")
        self.stacks = st

    @property
    def complete(self) -> bool:
        """True when the input consumed so far is a full sentence of the grammar (EOS allowed)."""
        return () in self.stacks

    @property
    def dead_end(self) -> bool:
        return self.stacks == {()}

    def matches(self, text: str) -> bool:
        """Full-string match from the *initial* state (helper for tests)."""
        g = Grammar(self.text, self.rules)
        st = g._feed(g.stacks, text)
        return () in st


class GrammarConstraint:
    """Binds a Grammar to a vocabulary: answers "is token t allowed now?" and produces full-vocabulary masks."""

    def __init__(self, grammar: Grammar | str, token_text: list[str | None], eos_ids: tuple[int, ...] = ()):
        self.g = grammar if isinstance(grammar, Grammar) else Grammar(grammar)
        self.token_text = token_text      # None = control / byte-fragment token: never allowed under a grammar
        self.eos_ids = set(eos_ids)

    def allowed(self, tok: int) -> bool:
        if tok in self.eos_ids:
            return self.g.complete
        t = self.token_text[tok]
        return t is not None and self.g.accepts(t)

    def mask(self):
        """Boolean list over the vocabulary (True = allowed).  First-character prefilter keeps this cheap."""
        firsts: dict[int, bool] = {}
        out = []
        for tid, t in enumerate(self.token_text):
            if tid in self.eos_ids:
                out.append(self.g.complete)
                continue
            if not t:
                out.append(False)
                continue
            c0 = ord(t[0])
            ok = firsts.get(c0)
            if ok is None:
                ok = firsts[c0] = bool(self.g._step(self.g._expand, self.g.stacks, c0))
            out.append(ok and (len(t) == 1 or self.g.accepts(t)))
        return out

    def accept(self, tok: int):
        if tok in self.eos_ids:
            return
        self.g.accept(self.token_text[tok])


# ---- JSON schema -> GBNF (subset) ---------------------------------------------------------------------
_JSON_PRIMS = r'''
ws ::= [ \t\n]{0,8}
string ::= "\"" ( [^"\\\x00-\x1f] | "\\" ( ["\\/bfnrt] | "u" [0-9a-fA-F]{4} ) )* "\""
integer ::= "-"? ( "0" | [1-9] [0-9]{0,15} )
number ::= integer ( "." [0-9]{1,16} )? ( [eE] [-+]? [0-9]{1,3} )?
boolean ::= "true" | "false"
null ::= "null"
value ::= object | array | string | number | boolean | null
object ::= "{" ws ( string ws ":" ws value ( ws "," ws string ws ":" ws value )* )? ws "}"
array ::= "[" ws ( value ( ws "," ws value )* )? ws "]"
'''


def _digits_between(lo: str, hi: str) -> str:
    """GBNF expression for the decimal strings of len(lo) == len(hi) digits between lo and hi inclusive (leading zeros allowed)."""
    n = len(lo)
    if n == 0:
        return ""
    if lo == "0" * n and hi == "9" * n:
        return "[0-9]" if n == 1 else f"[0-9]{{{n}}}"
    if lo[0] == hi[0]:
        rest = _digits_between(lo[1:], hi[1:])
        return f'"{lo[0]}"' + (f" {rest}" if rest else "")
    alts = []
    a, b = int(lo[0]), int(hi[0])
    low_full, high_full = lo[1:] == "0" * (n - 1), hi[1:] == "9" * (n - 1)
    if not low_full:                                     # the lowest leading digit with a constrained tail
        alts.append(f'"{lo[0]}" {_digits_between(lo[1:], "9" * (n - 1))}')
        a += 1
    if not high_full:
        b -= 1
    if a <= b:                                           # leading digits whose tail is free
        head = f'"{a}"' if a == b else f"[{a}-{b}]"
        alts.append(head + ("" if n == 1 else " [0-9]" if n == 2 else f" [0-9]{{{n - 1}}}"))
    if not high_full:
        alts.append(f'"{hi[0]}" {_digits_between("0" * (n - 1), hi[1:])}')
    return "( " + " | ".join(alts) + " )" if len(alts) > 1 else alts[0]


def _nonneg_range(lo: int, hi: int | None) -> list[str]:
    """Alternatives covering the integers lo..hi (hi None = unbounded, up to 16 digits) written without leading zeros."""
    alts = []
    top = hi if hi is not None else 10 ** 16 - 1
    for n in range(len(str(lo)), len(str(top)) + 1):
        a = max(lo, 0 if n == 1 else 10 ** (n - 1))
        b = min(top, 10 ** n - 1)
        if a <= b:
            alts.append(_digits_between(str(a).zfill(n), str(b).zfill(n)))
    return alts


def int_range_gbnf(lo: int | None, hi: int | None) -> str:
    """GBNF expression for the integers in [lo, hi] (either bound may be None)."""
    if lo is None and hi is None:
        return "integer"
    alts = []
    if lo is None or lo < 0:                              # negative part: "-" followed by a magnitude in [max(1, -hi), -lo]
        mag_lo = 1 if (hi is None or hi >= 0) else -hi
        mag = _nonneg_range(mag_lo, None if lo is None else -lo)
        if mag:
            alts.append('"-" ( ' + " | ".join(mag) + " )")
    if hi is None or hi >= 0:
        alts += _nonneg_range(max(lo or 0, 0) if lo is not None else 0, hi)
    return "( " + " | ".join(alts) + " )"


_Q = '"\\""'      # the GBNF literal for one double-quote character
_FORMATS = {
    "date": f"{_Q} date {_Q}", "time": f"{_Q} time {_Q}", "date-time": f'{_Q} date "T" time {_Q}',
    "uuid": f'{_Q} hex{{8}} "-" hex{{4}} "-" hex{{4}} "-" hex{{4}} "-" hex{{12}} {_Q}',
}
_FORMAT_PRIMS = r'''
date ::= [0-9]{4} "-" ( "0" [1-9] | "1" [0-2] ) "-" ( "0" [1-9] | [1-2] [0-9] | "3" [0-1] )
time ::= ( [01] [0-9] | "2" [0-3] ) ":" [0-5] [0-9] ":" [0-5] [0-9] ( "." [0-9]{3} )? ( "Z" | ( "+" | "-" ) ( [01] [0-9] | "2" [0-3] ) ":" [0-5] [0-9] )
hex ::= [0-9a-fA-F]
char ::= [^"\\\x00-\x1f] | "\\" ( ["\\/bfnrt] | "u" [0-9a-fA-F]{4} )
'''


def json_schema_to_gbnf(schema: dict | None) -> str:
    """GBNF for JSON documents matching `schema` (None / {} = any JSON value).

    Covers what the reference's converter does for the schemas LLM tooling sends (`common/json-schema-to-grammar.cpp`): `$ref` into
    `$defs` / `definitions`, `allOf` (objects merged), `anyOf` / `oneOf`, `const` / `enum`, type lists, objects with required and
    optional properties in declaration order and `additionalProperties`, arrays with `items` / `prefixItems` / `minItems` / `maxItems`,
    strings with `minLength` / `maxLength` and the `date`, `time`, `date-time`, `uuid` formats, integers with `minimum` / `maximum` /
    `exclusiveMinimum` / `exclusiveMaximum`.  `pattern` is not translated (the string stays unconstrained).  Every repetition in the
    generated grammar is bounded, so a constrained generation always terminates."""
    root_schema = schema or {}
    rules: dict[str, str] = {}
    uses_format = [False]

    def lit(v) -> str:
        return '"' + json.dumps(v).replace("\\", "\\\\").replace('"', '\\"') + '"'

    def rule_name(base: str) -> str:
        base = "".join(c if (c.isalnum() or c == "-") else "-" for c in base) or "r"
        n, i = base, 1
        while n in rules:
            i += 1
            n = f"{base}{i}"
        rules[n] = ""
        return n

    def resolve(sc: dict) -> dict:
        seen = 0
        while isinstance(sc, dict) and "$ref" in sc and seen < 32:
            ref = sc["$ref"]
            if not ref.startswith("#/"):
                raise GrammarError(f"unsupported $ref {ref!r} (only references into this schema)")
            node = root_schema
            for part in ref[2:].split("/"):
                node = node[part.replace("~1", "/").replace("~0", "~")]
            sc, seen = {**node, **{k: v for k, v in sc.items() if k != "$ref"}}, seen + 1
        return sc

    def merge_all_of(sc: dict) -> dict:
        out = {k: v for k, v in sc.items() if k != "allOf"}
        for sub in sc["allOf"]:
            sub = resolve(sub)
            if "allOf" in sub:
                sub = merge_all_of(sub)
            for k, v in sub.items():
                if k == "properties":
                    out["properties"] = {**out.get("properties", {}), **v}
                elif k == "required":
                    out["required"] = list(dict.fromkeys(list(out.get("required", [])) + list(v)))
                else:
                    out.setdefault(k, v)
        return out

    def obj(sc: dict, name: str) -> str:
        props = sc.get("properties", {})
        req = set(sc.get("required", []))
        extra = sc.get("additionalProperties", False if props else True)
        kvs = []
        for k in props:
            sub = visit(props[k], f"{name}-{k}")
            kvs.append((k in req, f'{lit(k)} ws ":" ws {sub}'))
        extra_kv = None
        if extra is not False:
            ev = visit(extra if isinstance(extra, dict) else {}, f"{name}-extra")
            extra_kv = f'string ws ":" ws {ev}'
        n = len(kvs)
        memo: dict[tuple[int, bool], str] = {}

        def rest(i: int, first: bool) -> str:
            """Rule for the properties from index i on; `first`: nothing has been written yet, so no comma goes in front."""
            key = (i, first)
            if key in memo:
                return memo[key]
            sep = "" if first else 'ws "," ws '
            if i == n:
                if extra_kv is None:
                    memo[key] = ""
                    return ""
                rn = rule_name(f"{name}-more")
                rules[rn] = f'( {sep}{extra_kv} ( ws "," ws {extra_kv} ){{0,15}} )?'
                memo[key] = rn
                return rn
            required, kv = kvs[i]
            after = rest(i + 1, False)
            taken = f"{sep}{kv}" + (f" {after}" if after else "")
            if required:
                body = taken
            else:
                skipped = rest(i + 1, first)
                body = f"( {taken} | {skipped} )" if skipped else f"( {taken} )?"
            rn = rule_name(f"{name}-p{i}{'a' if first else 'b'}")
            rules[rn] = body
            memo[key] = rn
            return rn
        inner = rest(0, True)
        rn = rule_name(name)
        rules[rn] = '"{" ws ' + (f"{inner} " if inner else "") + 'ws "}"'
        return rn

    def arr(sc: dict, name: str) -> str:
        rn = rule_name(name)
        if "prefixItems" in sc:
            items = [visit(x, f"{name}-t{i}") for i, x in enumerate(sc["prefixItems"])]
            rules[rn] = '"[" ws ' + ' ws "," ws '.join(items) + ' ws "]"'
            return rn
        item = visit(sc.get("items", {}), f"{name}-item")
        lo, hi = int(sc.get("minItems", 0)), sc.get("maxItems")
        hi = int(hi) if hi is not None else max(lo, 1) + 31
        if hi <= 0:
            rules[rn] = '"[" ws "]"'
            return rn
        more = f'( ws "," ws {item} ){{{max(0, lo - 1)},{hi - 1}}}'
        inner = f"{item} {more}"
        rules[rn] = '"[" ws ' + (f"( {inner} )?" if lo == 0 else inner) + ' ws "]"'
        return rn

    def visit(sc, name: str) -> str:
        if sc is True or sc is None or sc == {}:
            return "value"
        sc = resolve(sc)
        if "allOf" in sc:
            sc = merge_all_of(sc)
        if "const" in sc:
            return lit(sc["const"])
        if "enum" in sc:
            rn = rule_name(name)
            rules[rn] = " | ".join(lit(v) for v in sc["enum"])
            return rn
        if "anyOf" in sc or "oneOf" in sc:
            alts = sc.get("anyOf") or sc.get("oneOf")
            rn = rule_name(name)
            rules[rn] = " | ".join(visit(sub, f"{name}-{i}") for i, sub in enumerate(alts))
            return rn
        t = sc.get("type")
        if isinstance(t, list):
            rn = rule_name(name)
            rules[rn] = " | ".join(visit({**sc, "type": x}, f"{name}-{x}") for x in t)
            return rn
        if t == "object" or (t is None and ("properties" in sc or "additionalProperties" in sc)):
            if "properties" not in sc and sc.get("additionalProperties", True) is True:
                return "object"
            return obj(sc, name)
        if t == "array" or (t is None and ("items" in sc or "prefixItems" in sc)):
            if "items" not in sc and "prefixItems" not in sc and "minItems" not in sc and "maxItems" not in sc:
                return "array"
            return arr(sc, name)
        if t == "string":
            if sc.get("format") in _FORMATS:
                uses_format[0] = True
                rn = rule_name(name)
                rules[rn] = _FORMATS[sc["format"]]
                return rn
            if "minLength" in sc or "maxLength" in sc:
                uses_format[0] = True
                lo, hi = int(sc.get("minLength", 0)), sc.get("maxLength")
                rn = rule_name(name)
                rules[rn] = '"\\"" char' + (f"{{{lo},{int(hi)}}}" if hi is not None else f"{{{lo},{lo + 256}}}") + ' "\\""'
                return rn
            return "string"
        if t == "integer":
            lo, hi = sc.get("minimum"), sc.get("maximum")
            if "exclusiveMinimum" in sc:
                lo = int(sc["exclusiveMinimum"]) + 1
            if "exclusiveMaximum" in sc:
                hi = int(sc["exclusiveMaximum"]) - 1
            if lo is None and hi is None:
                return "integer"
            rn = rule_name(name)
            rules[rn] = int_range_gbnf(None if lo is None else int(lo), None if hi is None else int(hi))
            return rn
        if t in ("number", "boolean", "null"):
            return t
        return "value"

    root = visit(root_schema, "root-v")
    body = "\n".join(f"{k} ::= {v}" for k, v in rules.items())
    return f"root ::= {root}\n{body}\n" + _JSON_PRIMS + (_FORMAT_PRIMS if uses_format[0] else "")
