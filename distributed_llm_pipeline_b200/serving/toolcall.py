"""Tool calling for the OpenAI-compatible chat route: announce the tools to the model, constrain the answer, parse it back.

Reference counterpart: `llama.cpp/common/chat.cpp` (one handler per model family + the "generic" one) and
`common/chat-parser.cpp`.  Two halves, same as there:

* **generation side** (`prepare`): the tool list is written into the system message, and - the reference's generic handler,
  `common_chat_params_init_generic` - the answer is constrained by a grammar built from a JSON schema that allows either
  `{"tool_call": {"name": <one of the tools>, "arguments": <that tool's parameter schema>}}` (`tool_calls: [...]` when parallel
  calls are allowed) or `{"response": "<text>"}`; `tool_choice: "required"` removes the second branch, `"none"` the first.
* **parsing side** (`parse`): recognises the generic envelope and the native syntaxes of the common model families, so a model
  whose template already teaches it a tool syntax works without the grammar:
    - Hermes / Qwen:   `<tool_call>{"name": ..., "arguments": {...}}</tool_call>`
    - Llama 3.x:       `{"name": ..., "parameters": {...}}`  (optionally after `<|python_tag|>`)
    - Mistral Nemo:    `[TOOL_CALLS][{"name": ..., "arguments": {...}}, ...]`
    - Functionary:     `<function=name>{...}</function>`
  Everything outside the recognised calls is returned as the message content.
"""
from __future__ import annotations

import json
import re
import uuid
from dataclasses import dataclass, field


@dataclass
class ToolSetup:
    messages: list[dict]
    json_schema: dict | None = None
    active: bool = False
    parallel: bool = False


@dataclass
class ParsedMessage:
    content: str = ""
    tool_calls: list[dict] = field(default_factory=list)

    def to_openai(self) -> dict:
        m = {"role": "assistant", "content": self.content if (self.content or not self.tool_calls) else None}
        if self.tool_calls:
            m["tool_calls"] = self.tool_calls
        return m


def _functions(tools: list[dict]) -> list[dict]:
    out = []
    for t in tools or []:
        f = t.get("function", t) if isinstance(t, dict) else None
        if isinstance(f, dict) and f.get("name"):
            out.append({"name": str(f["name"]), "description": str(f.get("description", "")),
                        "parameters": f.get("parameters") or {"type": "object", "properties": {}}})
    return out


def prepare(messages: list[dict], tools: list[dict] | None, tool_choice="auto", parallel_tool_calls: bool = False) -> ToolSetup:
    """Returns the messages to render (tool list folded into the system message, earlier tool traffic flattened to text) and the
    JSON schema that constrains the answer."""
    fns = _functions(tools or [])
    msgs = [_flatten(m) for m in messages]
    choice = tool_choice if isinstance(tool_choice, str) else "required" if isinstance(tool_choice, dict) else "auto"
    if isinstance(tool_choice, dict):      # {"type": "function", "function": {"name": ...}}: only that tool may be called
        want = (tool_choice.get("function") or {}).get("name")
        fns = [f for f in fns if f["name"] == want] or fns
    if not fns or choice == "none":
        return ToolSetup(msgs)
    lines = ["You can call tools. Available tools:"]
    for f in fns:
        lines.append(json.dumps({"name": f["name"], "description": f["description"], "parameters": f["parameters"]}))
    lines.append('Answer with JSON only: {"tool_call": {"name": ..., "arguments": {...}}} to call a tool'
                 + (' (or {"tool_calls": [...]} for several)' if parallel_tool_calls else "")
                 + ('.' if choice == "required" else ', or {"response": "..."} to answer the user directly.'))
    note = "\n".join(lines)
    if msgs and msgs[0].get("role") == "system":
        msgs[0] = {"role": "system", "content": f"{msgs[0].get('content', '')}\n\n{note}"}
    else:
        msgs.insert(0, {"role": "system", "content": note})
    call = {"anyOf": [{"type": "object", "properties": {"name": {"const": f["name"]}, "arguments": f["parameters"]},
                       "required": ["name", "arguments"]} for f in fns]} if len(fns) > 1 else \
        {"type": "object", "properties": {"name": {"const": fns[0]["name"]}, "arguments": fns[0]["parameters"]}, "required": ["name", "arguments"]}
    if parallel_tool_calls:
        branch = {"type": "object", "properties": {"tool_calls": {"type": "array", "items": call, "minItems": 1}}, "required": ["tool_calls"]}
    else:
        branch = {"type": "object", "properties": {"tool_call": call}, "required": ["tool_call"]}
    if choice == "required":
        schema = branch
    else:
        schema = {"anyOf": [branch, {"type": "object", "properties": {"response": {"type": "string"}}, "required": ["response"]}]}
    return ToolSetup(msgs, schema, True, parallel_tool_calls)


def _flatten(m: dict) -> dict:
    """Earlier assistant tool calls and tool results become plain text the built-in templates can render."""
    role, content = m.get("role", "user"), m.get("content")
    if isinstance(content, list):       # OpenAI content parts
        content = "".join(p.get("text", "") for p in content if isinstance(p, dict))
    content = content or ""
    if role == "assistant" and m.get("tool_calls"):
        calls = [{"name": c.get("function", {}).get("name"), "arguments": _maybe_json(c.get("function", {}).get("arguments"))} for c in m["tool_calls"]]
        content = (content + "\n" if content else "") + json.dumps({"tool_calls": calls})
    if role == "tool":
        return {"role": "user", "content": f"Tool result ({m.get('name') or m.get('tool_call_id') or 'tool'}): {content}"}
    return {"role": role, "content": content}


def _maybe_json(v):
    if isinstance(v, str):
        try:
            return json.loads(v)
        except json.JSONDecodeError:
            return v
    return v


def _call(name, arguments) -> dict:
    args = arguments if isinstance(arguments, str) else json.dumps(arguments if arguments is not None else {})
    return {"id": "call_" + uuid.uuid4().hex[:24], "type": "function", "function": {"name": str(name), "arguments": args}}


def _json_prefix(s: str):
    """Decode one JSON value at the start of `s` -> (value, chars consumed) or (None, 0)."""
    s2 = s.lstrip()
    try:
        v, end = json.JSONDecoder().raw_decode(s2)
        return v, len(s) - len(s2) + end
    except json.JSONDecodeError:
        return None, 0


_HERMES = re.compile(r"<tool_call>\s*(.*?)\s*</tool_call>", re.S)
_FUNCTIONARY = re.compile(r"<function=([A-Za-z0-9_.\-]+)>\s*(.*?)\s*</function>", re.S)


def parse(text: str, tools: list[dict] | None = None) -> ParsedMessage:
    known = {f["name"] for f in _functions(tools or [])}
    ok = (lambda n: isinstance(n, str) and (not known or n in known))
    t = text.strip()
    # generic envelope (what `prepare` asks for)
    v, used = _json_prefix(t)
    if isinstance(v, dict) and used == len(t):
        if isinstance(v.get("tool_call"), dict) and ok(v["tool_call"].get("name")):
            return ParsedMessage("", [_call(v["tool_call"]["name"], v["tool_call"].get("arguments", {}))])
        if isinstance(v.get("tool_calls"), list) and v["tool_calls"] and all(isinstance(c, dict) and ok(c.get("name")) for c in v["tool_calls"]):
            return ParsedMessage("", [_call(c["name"], c.get("arguments", {})) for c in v["tool_calls"]])
        if isinstance(v.get("response"), str) and len(v) == 1:
            return ParsedMessage(v["response"])
    # Hermes / Qwen
    calls, rest = [], t
    for m in _HERMES.finditer(t):
        c, _ = _json_prefix(m.group(1))
        if isinstance(c, dict) and ok(c.get("name")):
            calls.append(_call(c["name"], c.get("arguments", c.get("parameters", {}))))
            rest = rest.replace(m.group(0), "", 1)
    if calls:
        return ParsedMessage(rest.strip(), calls)
    # Functionary
    for m in _FUNCTIONARY.finditer(t):
        if ok(m.group(1)):
            c, _ = _json_prefix(m.group(2))
            calls.append(_call(m.group(1), c if c is not None else m.group(2)))
            rest = rest.replace(m.group(0), "", 1)
    if calls:
        return ParsedMessage(rest.strip(), calls)
    # Mistral
    if "[TOOL_CALLS]" in t:
        head, _, tail = t.partition("[TOOL_CALLS]")
        arr, used = _json_prefix(tail)
        if isinstance(arr, list) and arr and all(isinstance(c, dict) and ok(c.get("name")) for c in arr):
            return ParsedMessage((head + tail[used:]).strip(), [_call(c["name"], c.get("arguments", {})) for c in arr])
    # Llama 3.x: a bare JSON object with name + parameters (possibly several separated by ';' after <|python_tag|>)
    body = t.split("<|python_tag|>", 1)[1] if "<|python_tag|>" in t else t
    pos, any_json = 0, False
    while pos < len(body):
        c, used = _json_prefix(body[pos:])
        if not (isinstance(c, dict) and ok(c.get("name")) and ("parameters" in c or "arguments" in c)):
            break
        any_json = True
        calls.append(_call(c["name"], c.get("parameters", c.get("arguments", {}))))
        pos += used
        while pos < len(body) and body[pos] in "; \n\t":
            pos += 1
    if any_json and pos >= len(body):
        return ParsedMessage(t.split("<|python_tag|>", 1)[0].strip() if "<|python_tag|>" in t else "", calls)
    return ParsedMessage(text)
