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


def json_schema_to_gbnf(schema: dict | None) -> str:
    """GBNF for JSON documents matching `schema` (None / {} = any JSON value)."""
    rules: list[str] = []
    counter = [0]

    def lit(v) -> str:
        return '"' + json.dumps(v).replace("\\", "\\\\").replace('"', '\\"') + '"'

    def visit(sc, name: str) -> str:
        if not sc:
            return "value"
        if "const" in sc:
            return lit(sc["const"])
        if "enum" in sc:
            rules.append(f"{name} ::= " + " | ".join(lit(v) for v in sc["enum"]))
            return name
        if "anyOf" in sc or "oneOf" in sc:
            alts = sc.get("anyOf") or sc.get("oneOf")
            parts = []
            for i, sub in enumerate(alts):
                parts.append(visit(sub, f"{name}-{i}"))
            rules.append(f"{name} ::= " + " | ".join(parts))
            return name
        t = sc.get("type")
        if isinstance(t, list):
            rules.append(f"{name} ::= " + " | ".join(visit({**sc, "type": x}, f"{name}-{x}") for x in t))
            return name
        if t == "object" and "properties" in sc:
            props = sc["properties"]
            req = sc.get("required", list(props))
            parts = []
            for k in props:                      # emitted in declaration order; optional ones may be skipped
                counter[0] += 1
                sub = visit(props[k], f"{name}-{k}".replace("_", "-"))
                parts.append((k in req, f'{lit(k)} ws ":" ws {sub}'))
            body, first = "", True
            for required, p in parts:
                sep = "" if first else 'ws "," ws '
                if required:
                    body += f" {sep}{p}"
                    first = False
                else:
                    body += f" ( {sep}{p} )?" if not first else f" ( {p} ws \",\" ws )?"
            rules.append(f'{name} ::= "{{" ws{body} ws "}}"')
            return name
        if t == "array" and "items" in sc:
            item = visit(sc["items"], f"{name}-item")
            lo, hi = sc.get("minItems", 0), sc.get("maxItems")
            rep = f'( ws "," ws {item} )' + ("*" if hi is None else f"{{0,{max(0, hi - 1)}}}")
            inner = f"{item} {rep}"
            rules.append(f'{name} ::= "[" ws ' + (f"( {inner} )?" if lo == 0 else inner) + ' ws "]"')
            return name
        if t in ("string", "integer", "number", "boolean", "null", "object", "array"):
            return t
        return "value"

    root = visit(schema or {}, "root-v")
    return f"root ::= {root}\n" + "\n".join(rules) + "\n" + _JSON_PRIMS
