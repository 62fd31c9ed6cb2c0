"""HTTP / SSE orchestrator, wire-compatible with the reference's Rust front-end.

Parity (`orchestrator/src/main.rs`): `POST /chat` with `{"prompt": string}` answers `text/event-stream`; every event is
`data: {"msg_type":"log"|"token","content":string}\\n\\n` (`:18-27,63-95`), a keep-alive comment is sent every second
(`:97`), CORS is permissive and everything else falls back to static files (`:100-111`); default port 3005.  What differs on
purpose: the engine is persistent (the reference spawns `llama-cli` and reloads the model for every request), requests
queue for slots instead of racing, `token` events carry only generated text (the reference forwards raw stdout incl. banners),
multi-byte UTF-8 is never split across events, a disconnecting client cancels its request, and failures arrive as a `log`
event instead of a panicked handler.  Extra routes mirror `llama-server` (`tools/server/server.cpp:164-197`): `/health`,
`/metrics`, `/slots`, `/props`, `/completion`, `/tokenize`, `/detokenize`, `/v1/models`, `/v1/completions`,
`/v1/chat/completions`.
"""
from __future__ import annotations

import argparse
import asyncio
import json
import mimetypes
import os
import sys
import time
import uuid

from ..utils.args import lora_from_args
from .sampler import SamplingParams
from .scheduler import Request, Scheduler

STATIC_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "static")
CORS = [("access-control-allow-origin", "*"), ("vary", "origin, access-control-request-method, access-control-request-headers")]


def sse_event(msg_type: str, content: str) -> bytes:
    """One server-sent event in the reference's wire format (serde field order: msg_type, content)."""
    return b"data: " + json.dumps({"msg_type": msg_type, "content": content}, ensure_ascii=False, separators=(",", ":")).encode() + b"\n\n"


KEEPALIVE = b":\n\n"


def _render_jinja(tmpl: str, messages: list[dict], meta: dict, add_generation_prompt: bool = True) -> str | None:
    """Render the model's own Jinja chat template (what `common_chat_templates_apply` does through minja) in a sandbox; None when
    jinja2 is missing or the template does not render (the marker-based fall-backs below take over)."""
    try:
        from jinja2.sandbox import ImmutableSandboxedEnvironment
    except ImportError:
        return None
    try:
        env = ImmutableSandboxedEnvironment(trim_blocks=True, lstrip_blocks=True)

        def raise_exception(msg):
            raise ValueError(msg)
        env.globals["raise_exception"] = raise_exception
        return env.from_string(tmpl).render(messages=messages, add_generation_prompt=add_generation_prompt, bos_token=str(meta.get("bos_token", "")),
                                            eos_token=str(meta.get("eos_token", "")))
    except Exception:      # noqa: BLE001 - any template problem falls back to the built-in formats
        return None


def apply_chat_template(meta: dict, messages: list[dict], add_generation_prompt: bool = True) -> str:
    """Chat formatting, in the reference's order of preference (`common/chat.cpp`, `llama-chat.cpp`): a built-in format forced by name
    (`--chat-template NAME` -> `meta["dlp.chat_template_name"]`), the model's own Jinja template rendered in a sandbox, the built-in
    format its template text is recognised as (`serving/chat_templates.py`: 46 formats, each checked against the reference's output), and
    for checkpoints without any template ChatML, which is what the reference server falls back to (`common_chat_templates_init`)."""
    from . import chat_templates
    name = meta.get("dlp.chat_template_name")
    if name:
        return chat_templates.apply(str(name), messages, add_generation_prompt)
    tmpl = str(meta.get("tokenizer.chat_template", ""))
    rendered = _render_jinja(tmpl, messages, meta, add_generation_prompt) if "{%" in tmpl or "{{" in tmpl else None
    if rendered is not None:
        return rendered
    detected = chat_templates.detect(tmpl) if tmpl else None
    if detected:
        return chat_templates.apply(detected, messages, add_generation_prompt)
    return chat_templates.apply("chatml", messages, add_generation_prompt)


FINGERPRINT = "dlp_b200"


def _as(kind, value, default):
    """`value` as `kind`, or `default` when it is absent or of the wrong type (llama-server's json_value falls back the same way)."""
    if value is None or isinstance(value, bool) and kind is not bool:
        return default
    try:
        return kind(value)
    except (TypeError, ValueError):
        return default


def _logit_bias(v) -> dict[int, float]:
    """`logit_bias` as {token id: bias} from either wire form ({"id": bias} or [[id, bias], ...]); `false` as a bias means "never" (-inf);
    entries that are not numbers are skipped."""
    pairs = v.items() if isinstance(v, dict) else [x for x in (v or []) if isinstance(x, (list, tuple)) and len(x) == 2] if isinstance(v, list) else []
    out = {}
    for k_, b in pairs:
        try:
            out[int(k_)] = float("-inf") if b is False else float(b)
        except (TypeError, ValueError):
            continue
    return out


def _logprob_items(items: list[dict]) -> list[dict]:
    """OpenAI `logprobs.content` entries from the scheduler's per-token records (post-sampling records carry probabilities: take their logs)."""
    import math

    def one(p):
        lp = p["logprob"] if "logprob" in p else (math.log(p["prob"]) if p.get("prob", 0) > 0 else -9999.0)
        return {"id": p["id"], "token": p["token"], "bytes": p.get("bytes", []), "logprob": lp}
    return [{**one(p), "top_logprobs": [one(t) for t in p.get("top_logprobs", p.get("top_probs", []))]} for p in items]


class Orchestrator:
    def __init__(self, scheduler: Scheduler, model_info: dict | None = None, static_dir: str = STATIC_DIR, n_predict: int = 200,
                 startup_logs: list[str] | None = None, api_key: str | None = None, max_queue: int = 0,
                 slot_save_path: str | None = None):
        self.api_key = api_key            # --api-key: Bearer token required on every API route except /health (server-http.cpp)
        self.max_queue = max_queue        # > 0: refuse new work with 503 once this many requests wait (request queueing, PDF p.6)
        self.slot_save_path = slot_save_path
        self.sched = scheduler
        self.info = model_info or {}
        self.static_dir = static_dir
        self.n_predict = n_predict
        self.startup_logs = startup_logs or []
        self.t_start = time.time()
        self._query: dict[str, str] = {}
        self._server: asyncio.AbstractServer | None = None

    # ---- tiny HTTP/1.1 layer ------------------------------------------------------------------------
    async def _read_request(self, reader: asyncio.StreamReader):
        line = await reader.readline()
        if not line:
            return None
        try:
            method, target, _ = line.decode("latin1").split(" ", 2)
        except ValueError:
            return None
        headers = {}
        while True:
            h = await reader.readline()
            if h in (b"\r\n", b"\n", b""):
                break
            k, _, v = h.decode("latin1").partition(":")
            headers[k.strip().lower()] = v.strip()
        body = b""
        n = int(headers.get("content-length", "0") or 0)
        if n > 64 << 20:
            raise ValueError("body too large")
        if n:
            body = await reader.readexactly(n)
        return method.upper(), target, headers, body

    @staticmethod
    def _head(status: str, headers: list[tuple[str, str]]) -> bytes:
        return (f"HTTP/1.1 {status}\r\n" + "".join(f"{k}: {v}\r\n" for k, v in headers) + "\r\n").encode("latin1")

    async def _send(self, writer, status: str, body: bytes, ctype: str = "application/json", extra=()):
        writer.write(self._head(status, [("content-type", ctype), ("content-length", str(len(body)))] + CORS + list(extra)) + body)
        await writer.drain()

    async def _json(self, writer, obj, status="200 OK"):
        await self._send(writer, status, json.dumps(obj).encode())

    async def handle(self, reader: asyncio.StreamReader, writer: asyncio.StreamWriter):
        try:
            while True:
                req = await self._read_request(reader)
                if req is None:
                    break
                method, target, headers, body = req
                path, _, query = target.partition("?")
                self._query = dict(kv.partition("=")[::2] for kv in query.split("&") if kv)
                keep = await self.route(method, path, headers, body, reader, writer)
                if not keep or headers.get("connection", "").lower() == "close":
                    break
        except (ConnectionError, asyncio.IncompleteReadError, ValueError):
            pass
        finally:
            try:
                writer.close()
            except Exception:
                pass

    async def route(self, method, path, headers, body, reader, writer) -> bool:
        if method == "OPTIONS":  # permissive CORS pre-flight (tower-http CorsLayer::permissive)
            writer.write(self._head("200 OK", CORS + [("access-control-allow-methods", "*"), ("access-control-allow-headers", "*"),
                                                       ("content-length", "0")]))
            await writer.drain()
            return True
        is_api = method == "POST" or path in ("/metrics", "/slots", "/props", "/v1/models", "/models", "/api/tags", "/lora-adapters") or path.startswith("/slots/")
        if self.api_key and is_api and path != "/health":
            auth = headers.get("authorization", "")
            if auth != f"Bearer {self.api_key}" and headers.get("x-api-key", "") != self.api_key:
                await self._json(writer, {"error": {"code": 401, "message": "Invalid API Key", "type": "authentication_error"}}, "401 Unauthorized")
                return True
        if self.max_queue > 0 and method == "POST" and len(self.sched.queue) >= self.max_queue and path not in ("/tokenize", "/detokenize"):
            await self._json(writer, {"error": {"code": 503, "message": "request queue is full", "type": "unavailable_error"}}, "503 Service Unavailable")
            return True
        try:
            if method == "POST" and path == "/chat":
                await self.chat(body, writer)
                return False
            if method == "POST" and path.startswith("/slots/"):
                return await self.slot_action(path, json.loads(body or b"{}"), writer)
            if method == "POST" and path == "/apply-template":
                d = json.loads(body or b"{}")
                await self._json(writer, {"prompt": apply_chat_template(self.info.get("metadata", {}), d.get("messages", []),
                                                                        bool(d.get("add_generation_prompt", True)))})
                return True
            if method == "POST" and path == "/infill":
                d = json.loads(body or b"{}")
                fim = self.sched.tok.fim_ids()
                if fim is None:
                    await self._json(writer, {"error": {"code": 501, "message": "Infill is not supported by this model", "type": "not_supported_error"}},
                                     "501 Not Implemented")
                    return True
                t = self.sched.tok
                ids = [fim["pre"]] + t.encode(d.get("input_prefix", ""), add_bos=False) + [fim["suf"]] + \
                    t.encode(d.get("input_suffix", ""), add_bos=False) + [fim["mid"]]
                d["prompt"] = ([t.bos_id] if t.add_bos else []) + ids
                d.setdefault("infill_sampler", True)      # the reference's /infill runs the infill sampler at the end of its chain
                return await self.completion(d, writer)
            if method == "POST" and path in ("/completion", "/completions", "/v1/completions"):
                return await self.completion(json.loads(body or b"{}"), writer, openai=path.startswith("/v1"))
            if method == "POST" and path == "/v1/messages/count_tokens":     # the prompt the messages render to, tokenised
                prompt = apply_chat_template(self.info.get("metadata", {}), self._anthropic_to_messages(json.loads(body or b"{}")))
                await self._json(writer, {"input_tokens": len(self.sched.tok.encode(prompt))})
                return True
            if method == "POST" and path == "/api/show":                      # Ollama: what is loaded
                md = self.info.get("metadata", {})
                await self._json(writer, {"model_info": {f"{self.info.get('arch', 'llama')}.context_length": self.sched.eng.n_ctx}, "modelfile": "", "parameters": "",
                                          "template": md.get("tokenizer.chat_template", ""), "details": self._ollama_details(), "capabilities": ["completion"]})
                return True
            if path == "/lora-adapters" and method in ("GET", "POST"):
                cur = self.info.get("lora", [])
                if method == "GET":
                    await self._json(writer, cur)
                    return True
                want = json.loads(body or b"[]")
                asked = {int(x.get("id", -1)): float(x.get("scale", 0.0)) for x in want if isinstance(x, dict)} if isinstance(want, list) else None
                if asked is None or any(i not in {a["id"] for a in cur} for i in asked):
                    await self._json(writer, *self._bad_request("expected an array of {id, scale} naming loaded adapters"))
                elif any(abs(asked.get(a["id"], 0.0) - a["scale"]) > 1e-9 for a in cur):
                    # adapters are folded into the weights when the model is loaded (models/weights.py): another scale needs a reload
                    await self._json(writer, {"error": {"code": 501, "message": "adapter scales are applied at load time here: restart with --lora path:scale",
                                                        "type": "not_supported_error"}}, "501 Not Implemented")
                else:
                    await self._json(writer, {"success": True})
                return True
            if method == "POST" and path == "/props":
                await self._json(writer, {"error": {"code": 501, "message": "this server's properties are fixed at start-up", "type": "not_supported_error"}},
                                 "501 Not Implemented")
                return True
            if method == "POST" and path in ("/v1/chat/completions", "/chat/completions"):
                return await self.chat_completions(json.loads(body or b"{}"), writer)
            if method == "POST" and path in ("/api/generate", "/api/chat"):
                d = json.loads(body or b"{}")
                # the reference registers /api/chat on its OpenAI chat handler (server.cpp: `post_chat_completions`), so a body without any of
                # Ollama's own fields is answered in that format; one that carries `options` / `format` / `keep_alive` gets Ollama's NDJSON
                if path == "/api/chat" and isinstance(d, dict) and not any(k in d for k in ("options", "format", "keep_alive")):
                    return await self.chat_completions(d, writer)
                return await self.ollama(path, d, writer)
            if method == "GET" and path == "/api/tags":
                await self._json(writer, {"models": [{"name": self.info.get("name", "model"), "model": self.info.get("name", "model"),
                                                      "details": {"format": "gguf", "family": self.info.get("arch", "llama")}}]})
                return True
            if method == "POST" and path == "/v1/messages":
                return await self.anthropic_messages(json.loads(body or b"{}"), writer)
            if method == "POST" and path == "/tokenize":
                d = json.loads(body or b"{}")
                t = self.sched.tok
                ids = t.encode(d.get("content", ""), add_bos=bool(d.get("add_special", False)))
                if d.get("with_pieces"):      # a piece that is not valid UTF-8 on its own is given as its bytes
                    def piece(i):
                        b = t.token_bytes(i) if t.types[i] != 3 else t.tokens[i].encode("utf-8")
                        try:
                            return b.decode("utf-8")
                        except UnicodeDecodeError:
                            return list(b)
                    await self._json(writer, {"tokens": [{"id": i, "piece": piece(i)} for i in ids]})
                else:
                    await self._json(writer, {"tokens": ids})
                return True
            if method == "POST" and path == "/detokenize":
                # control tokens are written out by name, as the reference's /detokenize does (tokens_to_str with special = true)
                await self._json(writer, {"content": self.sched.tok.decode_with_specials([int(t) for t in json.loads(body or b"{}").get("tokens", [])])})
                return True
            if method == "POST" and path in ("/embedding", "/embeddings", "/v1/embeddings"):
                d = json.loads(body or b"{}")
                inp = d.get("input", d.get("content", ""))
                items = inp if isinstance(inp, list) and (not inp or not isinstance(inp[0], int)) else [inp]
                loop = asyncio.get_running_loop()
                pooling = str(d.get("pooling", self.info.get("pooling", "mean")))
                vecs = [await loop.run_in_executor(None, self.sched.embed, it, pooling) for it in items]
                if path.startswith("/v1"):
                    n_tok = sum(len(self.sched.tok.encode(it)) if isinstance(it, str) else len(it) for it in items)
                    await self._json(writer, {"model": self.info.get("name", "model"), "object": "list", "usage": {"prompt_tokens": n_tok, "total_tokens": n_tok},
                                              "data": [{"embedding": v, "index": i, "object": "embedding"} for i, v in enumerate(vecs)]})
                else:
                    await self._json(writer, [{"index": i, "embedding": v} for i, v in enumerate(vecs)] if isinstance(inp, list) and items is inp
                                     else {"embedding": vecs[0]})
                return True
            if method == "POST" and path in ("/rerank", "/reranking", "/v1/rerank", "/v1/reranking"):
                return await self.rerank(json.loads(body or b"{}"), writer)
            if method == "GET" and path in ("/health", "/v1/health"):
                await self._json(writer, {"status": "ok"})
                return True
            if method == "GET" and path == "/metrics":
                await self._send(writer, "200 OK", self.sched.metrics.prometheus().encode(), "text/plain; version=0.0.4")
                return True
            if method == "GET" and path == "/slots":
                await self._json(writer, self.sched.slots_info())
                return True
            if method == "GET" and path == "/props":
                from .sampler import SamplingParams
                dp = SamplingParams()       # what a request gets when it names nothing (llama-server: `default_generation_settings`)
                md = self.info.get("metadata", {})
                await self._json(writer, {
                    "default_generation_settings": {"n_ctx": self.sched.eng.n_ctx, "params": {
                        **self._generation_settings(Request(prompt="", n_predict=self.n_predict, sampling=dp), {}), "n_predict": self.n_predict,
                        "max_tokens": self.n_predict}},
                    "total_slots": self.sched.n_slots, "model_alias": self.info.get("name", "model"), "model_path": self.info.get("path", self.info.get("name", "")),
                    "modalities": {"vision": False, "audio": False}, "endpoint_slots": True, "endpoint_props": False, "endpoint_metrics": True,
                    "chat_template": md.get("tokenizer.chat_template", ""), "bos_token": md.get("bos_token", ""), "eos_token": md.get("eos_token", ""),
                    "build_info": "dlp_b200", "is_sleeping": False, "webui": False, "webui_settings": {},
                    # this server's own fields
                    "model": self.info, "n_ctx": self.sched.eng.n_ctx, "uptime_s": round(time.time() - self.t_start, 1)})
                return True
            if method == "GET" and path in ("/v1/models", "/models"):          # the OpenAI list and the Ollama list in one answer, as the reference gives it
                name = self.info.get("name", "model")
                await self._json(writer, {"models": [{"name": name, "model": name, "modified_at": "", "size": "", "digest": "", "type": "model", "description": "",
                                                      "tags": [""], "capabilities": ["completion"], "parameters": "", "details": self._ollama_details()}],
                                          "object": "list", "data": [{"id": name, "object": "model", "created": int(self.t_start), "owned_by": "dlp_b200",
                                                                      "meta": {"n_vocab": self.info.get("n_vocab"), "n_ctx_train": self.info.get("n_ctx_train"),
                                                                               "n_embd": self.info.get("n_embd"), "n_params": self.info.get("n_params")}}]})
                return True
            if method in ("GET", "HEAD"):
                return await self.static(path, writer, head=(method == "HEAD"))
            await self._json(writer, {"error": {"code": 404, "message": "File Not Found", "type": "not_found_error"}}, "404 Not Found")
            return True
        except json.JSONDecodeError as e:
            await self._json(writer, *self._bad_request(f"invalid JSON: {e}"))
            return True
        except (ConnectionError, asyncio.CancelledError):
            raise
        except (KeyError, TypeError, ValueError, AttributeError) as e:      # a well-formed body of the wrong shape: answer, never drop the connection
            await self._json(writer, *self._bad_request(f"{type(e).__name__}: {e}"))
            return True

    async def slot_action(self, path: str, d: dict, writer) -> bool:
        """`POST /slots/{id}?action=save|restore|erase` (`tools/server/server.cpp:196-197`): KV state of an idle slot to / from a
        session file under --slot-save-path (utils/session.py)."""
        from ..utils import session
        try:
            idx = int(path.split("/")[2])
            slot = self.sched.slots[idx]
            if idx < 0:
                raise IndexError
        except (ValueError, IndexError):
            await self._json(writer, *self._bad_request("Invalid slot ID"))
            return True
        action = self._query.get("action", "")
        if action not in ("save", "restore", "erase"):
            await self._json(writer, *self._bad_request("Invalid action"))
            return True
        if action in ("save", "restore") and not self.slot_save_path:
            await self._json(writer, {"error": {"code": 501, "message": "This server does not support slots action. Start it with `--slot-save-path`",
                                                "type": "not_supported_error"}}, "501 Not Implemented")
            return True
        fname = os.path.basename(str(d.get("filename", f"slot{idx}.bin")))
        fpath = os.path.join(self.slot_save_path or "", fname)

        def job(eng):
            t0 = time.time()
            if slot.state.name != "IDLE":
                raise RuntimeError("slot is busy")
            if action == "erase":
                n = len(slot.cache_tokens)
                if slot.sid is not None:
                    eng.truncate_seq(slot.sid, 0)
                slot.cache_tokens, slot.n_past = [], 0
                return {"id_slot": idx, "n_erased": n}
            if action == "save":
                if slot.sid is None:
                    raise RuntimeError("slot is empty")
                nbytes = session.save_sequence(eng, slot.sid, fpath, slot.cache_tokens)
                return {"id_slot": idx, "filename": fname, "n_saved": len(slot.cache_tokens), "n_written": nbytes,
                        "timings": {"save_ms": round((time.time() - t0) * 1e3, 3)}}
            if action == "restore":
                if slot.sid is None:
                    slot.sid = eng.alloc_seq()
                try:
                    hdr = session.load_sequence(eng, slot.sid, fpath)
                except (OSError, ValueError, KeyError) as e:
                    raise RuntimeError(f"Unable to restore slot, no available space in KV cache or invalid slot save file ({e})")
                slot.cache_tokens = list(hdr.get("tokens") or [])[: hdr["n_tokens"]]
                slot.n_past = len(slot.cache_tokens)
                if slot.n_past != hdr["n_tokens"]:          # no token ids in the file: the cache cannot be matched to prompts
                    eng.truncate_seq(slot.sid, 0)
                    slot.cache_tokens, slot.n_past = [], 0
                return {"id_slot": idx, "filename": fname, "n_restored": slot.n_past, "n_read": os.path.getsize(fpath),
                        "timings": {"restore_ms": round((time.time() - t0) * 1e3, 3)}}
            raise ValueError("Invalid action")
        try:
            res = await asyncio.get_running_loop().run_in_executor(None, self.sched.run_on_engine, job)
        except (RuntimeError, ValueError, OSError) as e:
            await self._json(writer, *self._bad_request(str(e)))
            return True
        await self._json(writer, res)
        return True

    # ---- static files (ServeDir fallback) -----------------------------------------------------------------
    async def static(self, path: str, writer, head=False) -> bool:
        rel = path.lstrip("/") or "index.html"
        full = os.path.realpath(os.path.join(self.static_dir, rel))
        root = os.path.realpath(self.static_dir)
        if os.path.isdir(full):
            full = os.path.join(full, "index.html")
        if not (full == root or full.startswith(root + os.sep)) or not os.path.isfile(full):
            await self._json(writer, {"error": {"code": 404, "message": "File Not Found", "type": "not_found_error"}}, "404 Not Found")
            return True
        data = open(full, "rb").read()
        ctype = mimetypes.guess_type(full)[0] or "application/octet-stream"
        if ctype.startswith("text/") or ctype in ("application/javascript", "application/json"):
            ctype += "; charset=utf-8" if "charset" not in ctype else ""
        await self._send(writer, "200 OK", b"" if head else data, ctype)
        return True

    # ---- the reference route ----------------------------------------------------------------------------------
    async def _stream_request(self, req: Request, writer, fmt):
        """Submit `req`, forward its events formatted by `fmt(kind, content) -> bytes | None`, 1 s keep-alives in between."""
        loop = asyncio.get_running_loop()
        q: asyncio.Queue = asyncio.Queue()
        req.on_event = lambda kind, content: loop.call_soon_threadsafe(q.put_nowait, (kind, content))
        self.sched.submit(req)
        try:
            while True:
                try:
                    kind, content = await asyncio.wait_for(q.get(), timeout=1.0)
                except asyncio.TimeoutError:
                    writer.write(KEEPALIVE)
                    await writer.drain()
                    continue
                data = fmt(kind, content)
                if data:
                    writer.write(data)
                    await writer.drain()
                if kind == "done":
                    break
        except (ConnectionError, asyncio.CancelledError):
            self.sched.cancel(req)  # client went away: free the slot (the reference leaves the child running)
            raise

    async def chat(self, body: bytes, writer):
        try:
            payload = json.loads(body or b"{}")
            prompt = payload["prompt"]
            if not isinstance(prompt, str):
                raise KeyError
        except (json.JSONDecodeError, KeyError, TypeError):
            # axum's Json extractor rejects malformed bodies with 422 / 400
            await self._send(writer, "422 Unprocessable Entity", b"Failed to deserialize the JSON body into the target type", "text/plain; charset=utf-8")
            return
        writer.write(self._head("200 OK", [("content-type", "text/event-stream"), ("cache-control", "no-cache")] + CORS +
                                [("connection", "close")]))
        for line in self.startup_logs:  # the distributed-placement "proof" lines the UI highlights (RPC / offloaded)
            writer.write(sse_event("log", line))
        await writer.drain()
        req = Request(prompt=prompt, n_predict=self.n_predict, sampling=SamplingParams())

        def fmt(kind, content):
            if kind == "token":
                return sse_event("token", content)
            if kind == "log":
                return sse_event("log", content)
            if kind == "error":
                return sse_event("log", f"error: {content}\n")
            return None

        await self._stream_request(req, writer, fmt)

    # ---- llama-server style / OpenAI style completions -----------------------------------------------------------
    def _sampling_from(self, d: dict) -> SamplingParams:
        schema = d.get("json_schema")
        rf = d.get("response_format") if isinstance(d.get("response_format"), dict) else {}
        if schema is None and rf.get("type") == "json_object":
            schema = rf.get("schema") or {}
        if schema is None and rf.get("type") == "json_schema":
            schema = (rf.get("json_schema") or {}).get("schema") or {}
        trig = d.get("grammar_triggers") or []       # [{"type": "word" | "pattern" | 0..3, "value": "..."}] (server-common.cpp)
        words = tuple(str(t.get("value", "")) for t in trig if isinstance(t, dict) and t.get("type") in ("word", 1, None))
        pats = tuple(str(t.get("value", "")) for t in trig if isinstance(t, dict) and t.get("type") in ("pattern", "pattern_full", 2, 3))
        return SamplingParams(grammar=str(d.get("grammar", "") or ""), json_schema=schema,
                              grammar_lazy=bool(d.get("grammar_lazy", False)), grammar_trigger_words=words, grammar_trigger_patterns=pats,
                              infill=bool(d.get("infill_sampler", False)) or "infill" in (d.get("samplers") or []),
                              top_n_sigma=_as(float, d.get("top_n_sigma"), -1.0), dynatemp_range=_as(float, d.get("dynatemp_range"), 0.0),
                              dynatemp_exponent=_as(float, d.get("dynatemp_exponent"), 1.0), xtc_probability=_as(float, d.get("xtc_probability"), 0.0),
                              xtc_threshold=_as(float, d.get("xtc_threshold"), 0.1), mirostat=_as(int, d.get("mirostat"), 0),
                              mirostat_tau=_as(float, d.get("mirostat_tau"), 5.0), mirostat_eta=_as(float, d.get("mirostat_eta"), 0.1),
                              dry_multiplier=_as(float, d.get("dry_multiplier"), 0.0), dry_base=_as(float, d.get("dry_base"), 1.75),
                              dry_allowed_length=_as(int, d.get("dry_allowed_length"), 2), dry_penalty_last_n=_as(int, d.get("dry_penalty_last_n"), -1),
                              dry_breakers=tuple(t for b in (d.get("dry_sequence_breakers") if d.get("dry_sequence_breakers") is not None else ["\n", ":", "\"", "*"])
                                                 for t in self.sched.tok.encode(str(b), add_bos=False)[-1:]),
                              repeat_last_n=_as(int, d.get("repeat_last_n"), 64),
                              logit_bias=_logit_bias(d.get("logit_bias")),
                              temperature=_as(float, d.get("temperature"), 0.8), top_k=_as(int, d.get("top_k"), 40), top_p=_as(float, d.get("top_p"), 0.95),
                              min_p=_as(float, d.get("min_p"), 0.05), typical_p=_as(float, d.get("typical_p"), 1.0),
                              repeat_penalty=_as(float, d.get("repeat_penalty"), 1.0), frequency_penalty=_as(float, d.get("frequency_penalty"), 0.0),
                              presence_penalty=_as(float, d.get("presence_penalty"), 0.0), seed=max(0, _as(int, d.get("seed"), 0)))

    def _ollama_details(self) -> dict:
        return {"parent_model": "", "format": "gguf", "family": self.info.get("arch", ""), "families": [self.info.get("arch", "")], "parameter_size": "",
                "quantization_level": ""}

    def _bad_request(self, message: str) -> tuple[dict, str]:
        return {"error": {"code": 400, "message": message, "type": "invalid_request_error"}}, "400 Bad Request"

    def _request_from(self, d: dict, prompt, openai: bool) -> Request:
        stop = d.get("stop") or []
        stop = [stop] if isinstance(stop, str) else [str(x) for x in stop]
        try:
            n_predict = int(d.get("n_predict", d.get("max_tokens", self.n_predict)))
        except (TypeError, ValueError):
            n_predict = self.n_predict
        # -1 = until EOS or a full context; any other value below 1 still yields the token sampled from the prompt (the reference checks
        # its budget only after a token has been produced, server-context.cpp process_token)
        limit = self.sched.eng.n_ctx if n_predict == -1 else max(1, n_predict)
        req = Request(prompt=prompt, n_predict=limit, sampling=self._sampling_from(d), stop=stop, cache_prompt=bool(d.get("cache_prompt", True)),
                      ignore_eos=bool(d.get("ignore_eos", False)), n_keep=_as(int, d.get("n_keep"), 0),
                      n_draft=_as(int, d.get("speculative.n_max", d.get("n_draft")), 0),
                      n_probs=_as(int, d.get("n_probs"), 0) or (_as(int, d.get("logprobs"), 0) if openai else 0),
                      post_sampling_probs=bool(d.get("post_sampling_probs", False)))
        req.n_predict_asked = n_predict
        req.want_slot = _as(int, d.get("id_slot"), -1)
        return req

    def _generation_settings(self, req: Request, d: dict) -> dict:
        """What the request ran with, under llama-server's names (`server_task_result_cmpl_final::to_json`, server-task.cpp)."""
        sp = req.sampling
        return {"seed": sp.seed, "temperature": sp.temperature, "dynatemp_range": sp.dynatemp_range, "dynatemp_exponent": sp.dynatemp_exponent,
                "top_k": sp.top_k, "top_p": sp.top_p, "min_p": sp.min_p, "top_n_sigma": sp.top_n_sigma, "xtc_probability": sp.xtc_probability,
                "xtc_threshold": sp.xtc_threshold, "typical_p": sp.typical_p, "repeat_last_n": sp.repeat_last_n, "repeat_penalty": sp.repeat_penalty,
                "presence_penalty": sp.presence_penalty, "frequency_penalty": sp.frequency_penalty, "dry_multiplier": sp.dry_multiplier,
                "dry_base": sp.dry_base, "dry_allowed_length": sp.dry_allowed_length, "dry_penalty_last_n": sp.dry_penalty_last_n,
                "dry_sequence_breakers": d.get("dry_sequence_breakers") if d.get("dry_sequence_breakers") is not None else ["\n", ":", "\"", "*"],
                "mirostat": sp.mirostat, "mirostat_tau": sp.mirostat_tau, "mirostat_eta": sp.mirostat_eta, "stop": list(req.stop),
                "max_tokens": getattr(req, "n_predict_asked", req.n_predict), "n_predict": getattr(req, "n_predict_asked", req.n_predict),
                "n_keep": req.n_keep, "n_discard": 0, "ignore_eos": req.ignore_eos, "stream": bool(d.get("stream", False)),
                "logit_bias": [[k, v] for k, v in sp.logit_bias.items()], "n_probs": req.n_probs, "min_keep": 0, "grammar": sp.grammar,
                "grammar_lazy": sp.grammar_lazy, "grammar_triggers": list(d.get("grammar_triggers") or []), "preserved_tokens": [],
                "chat_format": "Content-only", "reasoning_format": "none", "reasoning_in_content": False, "thinking_forced_open": False,
                "samplers": ["penalties", "dry", "top_n_sigma", "top_k", "typ_p", "top_p", "min_p", "xtc", "temperature"],
                "speculative.n_max": req.n_draft or getattr(self.sched, "n_draft_default", 0), "speculative.n_min": 0, "speculative.p_min": 0.0,
                "timings_per_token": bool(d.get("timings_per_token", False)), "post_sampling_probs": req.post_sampling_probs,
                "backend_sampling": False, "lora": []}

    def _native_final(self, req: Request, d: dict, index: int = 0, content: str | None = None) -> dict:
        """The closing record of `/completion` (also the last stream event): every field of the reference's, in its order."""
        return {"index": index, "content": req.text if content is None else content,
                "tokens": list(req.output_ids) if (d.get("return_tokens") and content is None) else [], "id_slot": req.id_slot, "stop": True,
                "model": self.info.get("name", "model"), "tokens_predicted": len(req.output_ids), "tokens_evaluated": len(req.prompt_ids),
                "generation_settings": self._generation_settings(req, d), "prompt": self.sched.tok.decode_with_specials(req.prompt_ids),
                "has_new_line": "\n" in req.text, "truncated": req.truncated, "stop_type": req.stop_type, "stopping_word": req.stopping_word,
                "tokens_cached": req.tokens_cached, "timings": req.timings(),
                **({"completion_probabilities": req.probs} if req.n_probs > 0 and content is None else {})}

    def _openai_text(self, req: Request, rid: str) -> dict:
        return {"choices": [{"text": req.text, "index": 0, "logprobs": {"content": _logprob_items(req.probs)} if req.n_probs > 0 else None,
                             "finish_reason": req.finish_reason}],
                "created": int(time.time()), "model": self.info.get("name", "model"), "system_fingerprint": FINGERPRINT, "object": "text_completion",
                "usage": {"completion_tokens": len(req.output_ids), "prompt_tokens": len(req.prompt_ids),
                          "total_tokens": len(req.prompt_ids) + len(req.output_ids)}, "id": rid, "timings": req.timings()}

    async def completion(self, d, writer, openai=False) -> bool:
        if not isinstance(d, dict):
            await self._json(writer, *self._bad_request("the request body must be a JSON object"))
            return True
        if "prompt" not in d:
            await self._json(writer, *self._bad_request("key 'prompt' not found"))
            return True
        prompt = d["prompt"]
        # an array of prompts (strings or token arrays; a flat array of integers is ONE prompt) runs as one request each and answers with an array
        many = isinstance(prompt, list) and any(not isinstance(x, int) for x in prompt) and not d.get("stream")
        if many:
            reqs = [self._request_from(d, x, openai) for x in prompt]
            await asyncio.gather(*(self._run_blocking(r) for r in reqs))
            if openai:
                out = self._openai_text(reqs[0], "cmpl-" + uuid.uuid4().hex[:24])
                out["choices"] = [{**self._openai_text(r, "")["choices"][0], "index": k} for k, r in enumerate(reqs)]
                await self._json(writer, out)
            else:
                await self._json(writer, [self._native_final(r, d, k) for k, r in enumerate(reqs)])
            return True
        req = self._request_from(d, prompt, openai)
        rid = "cmpl-" + uuid.uuid4().hex[:24]
        sent, sent_tok = [0], [0]
        if d.get("stream"):
            writer.write(self._head("200 OK", [("content-type", "text/event-stream"), ("cache-control", "no-cache")] + CORS + [("connection", "close")]))

            def fmt(kind, content):
                if kind == "token":
                    ids, sent_tok[0] = req.output_ids[sent_tok[0]:], len(req.output_ids)
                    obj = ({"id": rid, "object": "text_completion", "created": int(time.time()), "model": self.info.get("name", "model"),
                            "system_fingerprint": FINGERPRINT, "choices": [{"index": 0, "text": content, "logprobs": None, "finish_reason": None}]}
                           if openai else {"index": 0, "content": content, "tokens": list(ids), "stop": False, "id_slot": -1,
                                           "tokens_predicted": len(req.output_ids), "tokens_evaluated": len(req.prompt_ids)})
                    if req.n_probs > 0:           # the entries recorded since the last event (a multi-byte piece may span tokens)
                        new, sent[0] = req.probs[sent[0]:], len(req.probs)
                        if openai:
                            obj["choices"][0]["logprobs"] = {"content": _logprob_items(new)}
                        else:
                            obj["completion_probabilities"] = new
                    return b"data: " + json.dumps(obj).encode() + b"\n\n"
                if kind == "done":
                    if openai:
                        obj = self._openai_text(req, rid)
                        obj["choices"][0].update(text="", logprobs=None, finish_reason=content)
                    else:
                        obj = self._native_final(req, d, content="")
                    return b"data: " + json.dumps(obj).encode() + b"\n\n" + (b"data: [DONE]\n\n" if openai else b"")
                return None

            await self._stream_request(req, writer, fmt)
            return False
        await self._run_blocking(req)
        out = self._openai_text(req, rid) if openai else self._native_final(req, d)
        fields = d.get("response_fields")
        if not openai and isinstance(fields, list) and fields:      # only these fields, nested ones by path ("timings/prompt_n"), keyed by the path
            picked = {}
            for path in fields:
                cur, ok = out, True
                for k in str(path).split("/"):
                    if isinstance(cur, dict) and k in cur:
                        cur = cur[k]
                    else:
                        ok = False
                        break
                if ok:
                    picked[str(path)] = cur
            out = picked
        await self._json(writer, out)
        return True

    async def rerank(self, d: dict, writer) -> bool:
        """`/v1/rerank` (tools/server/server.cpp rerank handler: `{"query", "documents", "top_n"}` -> results sorted by
        `relevance_score`).  The reference scores with a classification-head reranker (pooling type RANK); the decoder families
        served here have no such head, so the score is the cosine similarity of the pooled embeddings of query and document -
        the same engine path as `/v1/embeddings` - which keeps the route and its wire format usable with any loaded model."""
        query, docs = d.get("query"), d.get("documents") or d.get("texts")
        if not isinstance(query, str) or not isinstance(docs, list) or not docs:
            await self._json(writer, {"error": {"message": '"query" (string) and "documents" (non-empty array) are required', "type": "invalid_request_error"}},
                             "400 Bad Request")
            return True
        docs = [x if isinstance(x, str) else str(x.get("text", "")) if isinstance(x, dict) else str(x) for x in docs]
        pooling = str(d.get("pooling", "mean"))
        loop = asyncio.get_running_loop()
        qv = await loop.run_in_executor(None, self.sched.embed, query, pooling)
        n_tok = len(self.sched.tok.encode(query))
        scored = []
        for i, doc in enumerate(docs):
            dv = await loop.run_in_executor(None, self.sched.embed, doc, pooling)
            n_tok += len(self.sched.tok.encode(doc))
            scored.append({"index": i, "relevance_score": float(sum(a * b for a, b in zip(qv, dv)))})     # both vectors are unit length
        scored.sort(key=lambda r: -r["relevance_score"])
        top_n = int(d.get("top_n", len(scored)) or len(scored))
        results = scored[:max(1, top_n)]
        if d.get("return_documents") or d.get("return_text"):
            for r in results:
                r["document"] = {"text": docs[r["index"]]}
        await self._json(writer, {"model": str(d.get("model") or self.info.get("name", "model")), "object": "list",
                                  "usage": {"prompt_tokens": n_tok, "total_tokens": n_tok}, "results": results})
        return True

    async def chat_completions(self, d, writer) -> bool:
        from . import toolcall
        if not isinstance(d, dict) or not isinstance(d.get("messages"), list):
            await self._json(writer, *self._bad_request("Expected 'messages' to be an array"))
            return True
        tools = d.get("tools") or []
        setup = toolcall.prepare(d.get("messages", []), tools, d.get("tool_choice", "auto"), bool(d.get("parallel_tool_calls", False)))
        prompt = apply_chat_template(self.info.get("metadata", {}), setup.messages, bool(d.get("add_generation_prompt", True)))
        n_predict = _as(int, d.get("max_tokens", d.get("max_completion_tokens")), self.n_predict)
        n_predict = self.sched.eng.n_ctx if n_predict == -1 else max(1, n_predict)
        stop = d.get("stop") or []
        sampling = self._sampling_from(d)
        if setup.active and sampling.json_schema is None and not sampling.grammar:
            sampling.json_schema = setup.json_schema         # the generic tool-call envelope (common/chat.cpp, generic handler)
        req = Request(prompt=prompt, n_predict=n_predict, sampling=sampling, stop=[stop] if isinstance(stop, str) else list(stop),
                      n_probs=max(1, _as(int, d.get("top_logprobs"), 0)) if d.get("logprobs") else 0)
        rid = "chatcmpl-" + uuid.uuid4().hex[:24]

        def chat_logprobs(items):     # OpenAI chat `logprobs.content` (llama-server adds the token id to each entry)
            return {"content": _logprob_items(items)}
        model = self.info.get("name", "model")
        if d.get("stream"):
            writer.write(self._head("200 OK", [("content-type", "text/event-stream"), ("cache-control", "no-cache")] + CORS + [("connection", "close")]))
            held: list[str] = []

            opened = [False]

            def chunk(delta, finish, **extra):
                base = {"choices": [{"finish_reason": finish, "index": 0, "delta": delta}], "created": int(time.time()), "id": rid, "model": model,
                        "system_fingerprint": FINGERPRINT, "object": "chat.completion.chunk", **extra}
                out = b"data: " + json.dumps(base).encode() + b"\n\n"
                if not opened[0]:             # the stream opens with the role, as the reference's does
                    opened[0] = True
                    if "role" not in delta:
                        out = chunk({"role": "assistant", "content": None}, None) + out
                return out

            def fmt(kind, content):
                if kind == "token":
                    if tools:                     # a tool call is only known once the answer is complete: hold the text back
                        held.append(content)
                        return None
                    return chunk({"content": content}, None)
                if kind == "done":
                    out = b""
                    reason = content
                    if tools:
                        msg = toolcall.parse("".join(held), tools)
                        if msg.tool_calls:
                            out += chunk({"role": "assistant", "tool_calls": [{"index": i, **c} for i, c in enumerate(msg.tool_calls)]}, None)
                            reason = "tool_calls"
                        if msg.content:
                            out += chunk({"content": msg.content}, None)
                    usage = {"completion_tokens": len(req.output_ids), "prompt_tokens": len(req.prompt_ids),
                             "total_tokens": len(req.prompt_ids) + len(req.output_ids)}
                    if isinstance(d.get("stream_options"), dict) and d["stream_options"].get("include_usage"):
                        # OpenAI `stream_options.include_usage`: the closing chunk stays bare, one more chunk without choices carries the usage
                        tail = {"choices": [], "created": int(time.time()), "id": rid, "model": model, "system_fingerprint": FINGERPRINT,
                                "object": "chat.completion.chunk", "usage": usage, "timings": req.timings()}
                        return out + chunk({}, reason) + b"data: " + json.dumps(tail).encode() + b"\n\n" + b"data: [DONE]\n\n"
                    return out + chunk({}, reason, timings=req.timings()) + b"data: [DONE]\n\n"
                return None

            await self._stream_request(req, writer, fmt)
            return False
        await self._run_blocking(req)
        message, reason = {"role": "assistant", "content": req.text}, req.finish_reason
        if tools:
            msg = toolcall.parse(req.text, tools)
            message = msg.to_openai()
            if msg.tool_calls:
                reason = "tool_calls"
        await self._json(writer, {"id": rid, "object": "chat.completion", "created": int(time.time()), "model": model, "system_fingerprint": FINGERPRINT,
                                  "choices": [{"index": 0, "message": message, "finish_reason": reason,
                                               **({"logprobs": chat_logprobs(req.probs)} if req.n_probs > 0 else {})}],
                                  "usage": {"prompt_tokens": len(req.prompt_ids), "completion_tokens": len(req.output_ids),
                                            "total_tokens": len(req.prompt_ids) + len(req.output_ids)}, "timings": req.timings()})
        return True

    # ---- Ollama- and Anthropic-style routes (the reference server speaks both dialects, server.cpp:164-197) ------------
    async def ollama(self, path: str, d: dict, writer) -> bool:
        chat = path == "/api/chat"
        prompt = apply_chat_template(self.info.get("metadata", {}), d.get("messages", [])) if chat else str(d.get("prompt", ""))
        opts = dict(d.get("options") or {})
        opts.setdefault("temperature", 0.8)
        n_predict = int(opts.get("num_predict", self.n_predict))
        stop = opts.get("stop") or []
        req = Request(prompt=prompt, n_predict=n_predict if n_predict > 0 else self.sched.eng.n_ctx, sampling=self._sampling_from(opts),
                      stop=[stop] if isinstance(stop, str) else list(stop))
        model = str(d.get("model") or self.info.get("name", "model"))

        def obj(text, done, reason=None):
            o = {"model": model, "created_at": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()), "done": done}
            if chat:
                o["message"] = {"role": "assistant", "content": text}
            else:
                o["response"] = text
            if done:
                o.update(done_reason=reason, prompt_eval_count=len(req.prompt_ids), eval_count=len(req.output_ids))
            return o
        if d.get("stream", True):                    # Ollama streams newline-delimited JSON by default
            writer.write(self._head("200 OK", [("content-type", "application/x-ndjson")] + CORS + [("connection", "close")]))

            def fmt(kind, content):
                if kind == "token":
                    return json.dumps(obj(content, False)).encode() + b"\n"
                if kind == "done":
                    return json.dumps(obj("", True, content)).encode() + b"\n"
                return None
            await self._stream_request(req, writer, fmt)
            return False
        await self._run_blocking(req)
        await self._json(writer, obj(req.text, True, req.finish_reason))
        return True

    @staticmethod
    def _anthropic_to_messages(d: dict) -> list[dict]:
        norm = []
        if d.get("system"):
            sys_txt = d["system"] if isinstance(d["system"], str) else "".join(b.get("text", "") for b in d["system"])
            norm.append({"role": "system", "content": sys_txt})
        for m in list(d.get("messages", [])):
            c = m.get("content", "")
            norm.append({"role": m.get("role", "user"), "content": c if isinstance(c, str) else "".join(b.get("text", "") for b in c)})
        return norm

    async def anthropic_messages(self, d: dict, writer) -> bool:
        norm = self._anthropic_to_messages(d)
        stop = d.get("stop_sequences") or []
        req = Request(prompt=apply_chat_template(self.info.get("metadata", {}), norm), n_predict=int(d.get("max_tokens", self.n_predict)),
                      sampling=self._sampling_from(d), stop=list(stop))
        rid = "msg_" + uuid.uuid4().hex[:24]
        model = str(d.get("model") or self.info.get("name", "model"))
        if d.get("stream"):
            writer.write(self._head("200 OK", [("content-type", "text/event-stream"), ("cache-control", "no-cache")] + CORS + [("connection", "close")]))

            def ev(name, o):
                return f"event: {name}\ndata: {json.dumps(o)}\n\n".encode()
            n_in = len(self.sched.tok.encode(req.prompt)) if isinstance(req.prompt, str) else len(req.prompt)
            writer.write(ev("message_start", {"type": "message_start", "message": {"id": rid, "type": "message", "role": "assistant", "content": [], "model": model,
                                                                                   "stop_reason": None, "stop_sequence": None,
                                                                                   "usage": {"input_tokens": n_in, "output_tokens": 0}}}))
            writer.write(ev("content_block_start", {"type": "content_block_start", "index": 0, "content_block": {"type": "text", "text": ""}}))

            def fmt(kind, content):
                if kind == "token":
                    return ev("content_block_delta", {"type": "content_block_delta", "index": 0, "delta": {"type": "text_delta", "text": content}})
                if kind == "done":
                    reason = "max_tokens" if content == "length" else "stop_sequence" if req.stop_type == "word" else "end_turn"
                    return (ev("content_block_stop", {"type": "content_block_stop", "index": 0}) +
                            ev("message_delta", {"type": "message_delta", "delta": {"stop_reason": reason, "stop_sequence": req.stopping_word or None},
                                                 "usage": {"output_tokens": len(req.output_ids)}}) +
                            ev("message_stop", {"type": "message_stop"}))
                return None
            await self._stream_request(req, writer, fmt)
            return False
        await self._run_blocking(req)
        await self._json(writer, {"id": rid, "type": "message", "role": "assistant", "model": model,
                                  "content": [{"type": "text", "text": req.text}],
                                  "stop_reason": "max_tokens" if req.finish_reason == "length" else "stop_sequence" if req.stop_type == "word" else "end_turn",
                                  "stop_sequence": req.stopping_word or None,
                                  "usage": {"input_tokens": len(req.prompt_ids), "output_tokens": len(req.output_ids)}})
        return True

    async def _run_blocking(self, req: Request):
        loop = asyncio.get_running_loop()
        fut = loop.create_future()
        req.on_event = lambda kind, content: (kind == "done") and loop.call_soon_threadsafe(lambda: fut.done() or fut.set_result(True))
        self.sched.submit(req)
        await fut

    # ---- lifecycle ----------------------------------------------------------------------------------------
    async def serve(self, host: str = "0.0.0.0", port: int = 3005):
        self._server = await asyncio.start_server(self.handle, host, port, reuse_address=True)
        async with self._server:
            await self._server.serve_forever()


class ModelRouter(Orchestrator):
    """Several models behind one port: requests name a model (`"model"` field of the JSON body, default = the first one), models
    are loaded on first use, at most `max_loaded` stay resident (least recently used idle model is unloaded first) and models
    idle for `sleep_idle_s` seconds are unloaded to free the GPU.  Each loaded model is a full Orchestrator (own scheduler
    thread, slots, KV cache), so every route works per model.

    Parity: llama-server's router mode (`tools/server/server-models.cpp:415-600`: spawn / monitor child servers, LRU unload)
    and `--sleep-idle-seconds` (`server-context.cpp:651-665`); here the models share the process instead of being child
    processes (the native front-end `runtime/orchestrator.cpp` supervises the process as a whole).

    `replicas` > 1 loads every model that many times (data parallelism: replica i on `cuda:(i mod device count)` when the device is
    "cuda", each with its own scheduler, slots and KV cache) and sends each request to the replica with the fewest requests in flight."""

    def __init__(self, models: dict[str, str], build_kw: dict | None = None, max_loaded: int = 1, sleep_idle_s: float = 0.0,
                 static_dir: str = STATIC_DIR, n_predict: int = 200, replicas: int = 1, **orch_kw):
        super().__init__(None, {"name": "router"}, static_dir, n_predict, api_key=orch_kw.get("api_key"))
        self.models = dict(models)
        self.build_kw = build_kw or {}
        self.max_loaded = max(1, max_loaded)
        self.sleep_idle_s = sleep_idle_s
        self.orch_kw = orch_kw
        self.replicas = max(1, int(replicas))
        self.loaded: dict[str, list[Orchestrator]] = {}      # model name -> its replicas
        self.last_used: dict[str, float] = {}
        self.n_loads = 0
        self._rr = 0
        self._load_lock = asyncio.Lock()

    @staticmethod
    def _in_flight(o: Orchestrator) -> int:
        return len(o.sched.queue) + sum(s.state.name != "IDLE" for s in o.sched.slots)

    @classmethod
    def _busy(cls, group: list[Orchestrator]) -> bool:
        return any(cls._in_flight(o) for o in group)

    def _pick(self, group: list[Orchestrator]) -> Orchestrator:
        """Least requests in flight; ties go round so that an idle group is used evenly."""
        self._rr += 1
        return min(enumerate(group), key=lambda io: (self._in_flight(io[1]), (io[0] - self._rr) % len(group)))[1]

    def _unload(self, name: str):
        group = self.loaded.pop(name, None)
        if group is not None:
            for o in group:
                o.sched.stop()
            del group
            import gc
            import torch
            gc.collect()
            if torch.cuda.is_available():
                torch.cuda.empty_cache()

    def _load(self, name: str) -> list[Orchestrator]:
        import torch
        group = []
        for i in range(self.replicas):
            kw = dict(self.build_kw)
            if self.replicas > 1 and str(kw.get("device", "cuda")) == "cuda" and torch.cuda.is_available():
                kw["device"] = f"cuda:{i % torch.cuda.device_count()}"
            sched, info, logs = build_service(self.models[name], **kw)
            sched.start()
            group.append(Orchestrator(sched, info, self.static_dir, self.n_predict, logs, **self.orch_kw))
        self.n_loads += 1
        return group

    async def acquire(self, name: str) -> Orchestrator:
        async with self._load_lock:
            if name not in self.loaded:
                while len(self.loaded) >= self.max_loaded:
                    idle = sorted((n for n, o in self.loaded.items() if not self._busy(o)), key=lambda n: self.last_used.get(n, 0.0))
                    if not idle:
                        await asyncio.sleep(0.05)      # every resident model is serving: wait for one to drain
                        continue
                    self._unload(idle[0])
                self.loaded[name] = await asyncio.get_running_loop().run_in_executor(None, self._load, name)
            self.last_used[name] = time.time()
            return self._pick(self.loaded[name])

    async def reap_idle(self):
        """Background task: unload models nobody has used for sleep_idle_s seconds."""
        while self.sleep_idle_s > 0:
            await asyncio.sleep(min(1.0, self.sleep_idle_s / 2))
            async with self._load_lock:
                now = time.time()
                for n in [n for n, o in self.loaded.items() if now - self.last_used.get(n, now) > self.sleep_idle_s and not self._busy(o)]:
                    self._unload(n)

    async def route(self, method, path, headers, body, reader, writer) -> bool:
        if method == "POST" and path in ("/models/load", "/models/unload"):      # llama-server router mode (server-models.cpp): explicit load / unload
            try:
                name = str(json.loads(body or b"{}").get("model", ""))
            except (json.JSONDecodeError, AttributeError):
                name = ""
            if name not in self.models:
                await self._json(writer, {"error": {"code": 404, "message": f"model {name!r} not found", "type": "not_found_error"}}, "404 Not Found")
            elif path == "/models/load":
                await self.acquire(name)
                await self._json(writer, {"success": True})
            else:
                async with self._load_lock:
                    if name in self.loaded and self._busy(self.loaded[name]):
                        await self._json(writer, {"error": {"code": 409, "message": f"model {name!r} is serving requests", "type": "unavailable_error"}}, "409 Conflict")
                        return True
                    self._unload(name)
                await self._json(writer, {"success": True})
            return True
        if method == "GET" and path in ("/v1/models", "/models"):
            await self._json(writer, {"object": "list", "data": [
                {"id": n, "object": "model", "owned_by": "dlp_b200", "path": p,
                 "status": "loaded" if n in self.loaded else "unloaded"} for n, p in self.models.items()]})
            return True
        if method == "GET" and path in ("/health", "/v1/health"):
            await self._json(writer, {"status": "ok", "models_loaded": sorted(self.loaded), "loads": self.n_loads, "replicas": self.replicas,
                                      "in_flight": {n: [self._in_flight(o) for o in g] for n, g in self.loaded.items()}})
            return True
        if method in ("GET", "HEAD", "OPTIONS") and path not in ("/metrics", "/slots", "/props", "/lora-adapters"):
            return await super().route(method, path, headers, body, reader, writer)
        name = next(iter(self.models))
        if body:
            try:
                name = str(json.loads(body).get("model") or name)
            except (json.JSONDecodeError, AttributeError):
                pass
        name = self._query.get("model", name)
        if name not in self.models:
            await self._json(writer, {"error": {"code": 404, "message": f"model {name!r} not found", "type": "not_found_error"}}, "404 Not Found")
            return True
        orch = await self.acquire(name)
        orch._query = self._query
        try:
            return await orch.route(method, path, headers, body, reader, writer)
        finally:
            self.last_used[name] = time.time()

    async def serve(self, host: str = "0.0.0.0", port: int = 3005):
        reaper = asyncio.ensure_future(self.reap_idle()) if self.sleep_idle_s > 0 else None
        try:
            await super().serve(host, port)
        finally:
            if reaper:
                reaper.cancel()
            for n in list(self.loaded):
                self._unload(n)


def add_parallel_args(ap):
    """Flags that put a distributed engine behind the front-end (the reference: `llama-cli --rpc a,b -ngl 99` per request,
    `orchestrator/src/main.rs:47-48`; `llama-server -sm layer|row -ts ...`, `common/arg.cpp:2310-2358`)."""
    ap.add_argument("--tp", type=int, default=int(os.environ.get("DLP_ARG_TP", "1")), help="tensor-parallel ranks per pipeline stage")
    ap.add_argument("--pp", type=int, default=int(os.environ.get("DLP_ARG_PP", "1")), help="pipeline stages (layer split)")
    ap.add_argument("--tp-impl", default=os.environ.get("DLP_TP_IMPL", "fused"), choices=["fused", "nccl"],
                    help="tensor-parallel collectives: fused into the kernels (NVLink stores) or NCCL")
    ap.add_argument("-ts", "--tensor-split", default=None, help="relative layer share of each pipeline stage, e.g. 3,1")
    ap.add_argument("--rpc", default=os.environ.get("DLP_ARG_RPC"), help="comma separated host:port list of dlp-worker stages (CPU layer split)")
    ap.add_argument("--workers", type=int, default=0, help="launch this many local dlp-worker stages and use them (implies --rpc)")
    ap.add_argument("--worker-base-port", type=int, default=50052)
    ap.add_argument("--no-warmup", action="store_true", help="skip the start-up warm-up (kernel set-up + decode-graph capture per batch size)")
    ap.add_argument("--mlock", action="store_true", help="pin the model file's pages in RAM")
    ap.add_argument("--no-mmap", action="store_true", help="read the model file into memory instead of mapping it")
    ap.add_argument("--direct-io", action="store_true", help="read the model file with O_DIRECT (implies --no-mmap)")
    ap.add_argument("-md", "--model-draft", default=os.environ.get("DLP_ARG_MODEL_DRAFT"),
                    help="draft model (same vocabulary) for speculative decoding of greedy requests inside the server")
    ap.add_argument("--draft-max", type=int, default=16, help="tokens drafted per verification block (0 = off unless the request asks)")


def build_service(model: str, device: str = "cuda", n_ctx: int = 2048, n_slots: int = 4, n_batch: int = 2048, page_size: int = 64,
                  lora=None, pp: int = 1, tp: int = 1, tp_impl: str = "fused", tensor_split=None, rpc: str | None = None,
                  workers: int = 0, worker_base_port: int = 50052, model_draft: str | None = None, draft_max: int = 16):
    """Load a GGUF checkpoint and return (scheduler, model_info, startup_logs).

    model_draft: a second, smaller checkpoint with the same vocabulary, loaded on the front-end process's device; greedy requests
    are then decoded by draft-and-verify blocks (`llama-server -md`, tools/server/server-context.cpp:2029-2068).

    pp * tp > 1: must run under torchrun (one process per rank); only the leader rank returns - followers replay the leader's
    engine calls inside this function until the service shuts down and then exit (parallel/spmd.py).
    rpc / workers: the engine is a chain of dlp-worker CPU stages driven over TCP (parallel/rpc.py)."""
    import torch

    from ..gguf import open_gguf
    from ..models.llama import LlamaEngine
    from ..models.weights import load_weights
    from ..utils.tokenizer import Tokenizer
    dev = device if (device != "cuda" or torch.cuda.is_available()) else "cpu"
    f = open_gguf(model)
    tok = Tokenizer(f.metadata)
    t0 = time.time()
    info_extra = {}
    if workers or rpc:
        from ..parallel.rpc import LocalWorkerPool, RpcEngine
        pool = None
        if workers:
            pool = LocalWorkerPool(model, workers, base_port=worker_base_port, n_ctx=n_ctx)
            rpc = ",".join(f"127.0.0.1:{worker_base_port + i}" for i in range(workers))
        eng = RpcEngine(rpc.split(","), n_ctx=n_ctx, max_seqs=n_slots, recover=pool is not None)
        eng._pool = pool
        from ..models.config import config_from_gguf
        cfg = config_from_gguf(f.metadata, f.tensors)
        logs = [f"load: {os.path.basename(model)} | arch {cfg.arch} | {cfg.n_layer} layers | n_embd {cfg.n_embd} | vocab {cfg.vocab_size}\n"]
        logs += eng.describe()
        logs.append(f"RPC pipeline: {len(eng.workers)} stages, n_ctx {eng.n_ctx}, {n_slots} slots, ready in {time.time() - t0:.2f} s\n")
    elif pp * tp > 1:
        from ..parallel import spmd
        from ..parallel.grid import init_grid, shutdown
        grid = init_grid(pp=pp, tp=tp, device=dev)
        split = [float(x) for x in tensor_split.split(",")] if isinstance(tensor_split, str) and tensor_split else tensor_split
        eng_local, stage, comm, load_s, ranges = spmd.build_engine(f, grid, n_ctx, n_slots + 1, n_batch, page_size if dev != "cpu" else 16,
                                                                   tp_impl, lora, split)
        cfg = eng_local.cfg
        mine = (f"load_tensors: stage {grid.stage} tp-rank {grid.tp_rank}: offloaded layers [{ranges[grid.stage].start}, {ranges[grid.stage].stop}) of "
                f"{cfg.n_layer} to GPU {grid.device} ({eng_local.w.nbytes() / 2**20:.1f} MiB in {load_s:.2f} s"
                f"{', +embedding' if grid.first_stage else ''}{', +lm_head' if grid.last_stage else ''})\n")
        lines = [None] * grid.world
        torch.distributed.all_gather_object(lines, mine)
        leader = spmd.leader_rank(pp, tp)
        ch = spmd.make_channel(grid, leader)
        if grid.rank != leader:
            spmd.follower_loop(eng_local, grid, stage, ch, log=sys.stderr.write)
            shutdown(0)
        eng = spmd.DistributedEngine(eng_local, grid, stage, ch)
        impl = "fused in-kernel collectives" if getattr(comm, "fused", False) else ("NCCL / gloo collectives" if tp > 1 else "none")
        logs = [f"load: {os.path.basename(model)} | arch {cfg.arch} | {cfg.n_layer} layers | n_embd {cfg.n_embd} | vocab {cfg.vocab_size}\n",
                f"grid: {pp} pipeline stage(s) x {tp} tensor-parallel rank(s) = {grid.world} processes | TP: {impl}"
                f"{' (' + comm.symm_backend + ')' if getattr(comm, 'symm_backend', None) else ''} | control: {os.environ.get('DLP_CTRL', 'shm')}\n"]
        logs += lines
        logs.append(f"kv cache: {eng_local.kv.n_pages} pages x {eng_local.page_size} tokens per rank ({eng_local.kv.bytes() / 2**20:.1f} MiB), {n_slots} slots, n_ctx {n_ctx}\n")
        info_extra = {"pp": pp, "tp": tp}
    else:
        w = load_weights(f, dev, lora=lora)
        eng = LlamaEngine(w, n_ctx=n_ctx, max_seqs=n_slots + 1, page_size=page_size if dev != "cpu" else 16, max_batch_tokens=n_batch)
        cfg = w.cfg
        logs = [
            f"load: {os.path.basename(model)} | arch {cfg.arch} | {cfg.n_layer} layers | n_embd {cfg.n_embd} | vocab {cfg.vocab_size}\n",
            f"load_tensors: offloaded {cfg.n_layer + 1}/{cfg.n_layer + 1} layers to GPU ({dev}), {w.nbytes() / 2**20:.1f} MiB in {time.time() - t0:.2f} s\n",
            f"kv cache: {eng.kv.n_pages} pages x {eng.page_size} tokens ({eng.kv.bytes() / 2**20:.1f} MiB), {n_slots} slots, n_ctx {n_ctx}\n",
        ]
    info = {"name": str(f.metadata.get("general.name", os.path.basename(model))), "n_params": cfg.n_params(), "n_ctx_train": cfg.n_ctx_train,
            "metadata": {"tokenizer.chat_template": f.metadata.get("tokenizer.chat_template", ""),
                         "bos_token": tok.tokens[tok.bos_id] if 0 <= tok.bos_id < tok.n_vocab else "",
                         "eos_token": tok.tokens[tok.eos_id] if 0 <= tok.eos_id < tok.n_vocab else ""}, **info_extra,
            "arch": cfg.arch, "n_vocab": tok.n_vocab, "n_embd": cfg.n_embd, "path": model,
            "lora": [{"id": i, "path": getattr(a, "path", str(a)), "scale": float(getattr(a, "scale", 1.0))} for i, a in enumerate(lora or [])]}
    draft_eng = None
    if model_draft:
        fd = open_gguf(model_draft)
        tok_d = Tokenizer(fd.metadata)
        if tok_d.n_vocab != tok.n_vocab or tok_d.tokens[:256] != tok.tokens[:256]:     # common/speculative.cpp: vocabularies must match
            raise ValueError(f"draft model vocabulary ({tok_d.n_vocab} tokens) does not match the target's ({tok.n_vocab})")
        wd = load_weights(fd, dev)
        draft_eng = LlamaEngine(wd, n_ctx=n_ctx, max_seqs=n_slots + 1, page_size=page_size if dev != "cpu" else 16,
                                max_batch_tokens=min(n_batch, 512))
        logs.append(f"draft model: {os.path.basename(model_draft)} | {wd.cfg.n_layer} layers | n_embd {wd.cfg.n_embd} | up to {draft_max} tokens per block\n")
        info["draft_model"] = os.path.basename(model_draft)
    sched = Scheduler(eng, tok, n_batch=n_batch, n_slots=n_slots, draft_engine=draft_eng, n_draft_default=draft_max if draft_eng else 0)
    return sched, info, logs


def service_kwargs(a) -> dict:
    from ..utils.args import apply_loader_flags
    apply_loader_flags(a)
    return dict(pp=a.pp, tp=a.tp, tp_impl=a.tp_impl, tensor_split=a.tensor_split, rpc=a.rpc, workers=a.workers,
                worker_base_port=a.worker_base_port, model_draft=getattr(a, "model_draft", None), draft_max=getattr(a, "draft_max", 16))


def maybe_reexec(a, module: str, argv):
    """`--tp/--pp` outside torchrun: start the same command once per rank."""
    if a.pp * a.tp > 1 and "WORLD_SIZE" not in os.environ and not (a.rpc or a.workers):
        from ..parallel.spmd import reexec_under_torchrun
        reexec_under_torchrun(a.pp * a.tp, list(sys.argv[1:] if argv is None else argv), module=module)


def stop_service(sched):
    sched.stop()
    eng = sched.eng
    if hasattr(eng, "shutdown"):          # distributed engine: release the followers
        eng.shutdown()
        from ..parallel.grid import shutdown
        shutdown(0)
    pool = getattr(eng, "_pool", None)
    if pool is not None:
        pool.stop()


def main(argv=None):
    ap = argparse.ArgumentParser(description="dlp_b200 orchestrator (wire-compatible with the reference's /chat SSE API)")
    ap.add_argument("-m", "--model", default=None)
    ap.add_argument("--models", default=None, help="router mode: comma separated name=path.gguf list (or a directory of .gguf files)")
    ap.add_argument("--models-max", type=int, default=1, help="router mode: models kept resident at once")
    ap.add_argument("--replicas", type=int, default=1, help="router mode: copies of each model (one per GPU in turn); requests go to the least busy copy")
    ap.add_argument("--sleep-idle-seconds", type=float, default=0.0, help="router mode: unload a model after this many idle seconds")
    ap.add_argument("--host", default=os.environ.get("DLP_HOST", "0.0.0.0"))
    ap.add_argument("--port", type=int, default=int(os.environ.get("DLP_PORT", "3005")))
    ap.add_argument("-c", "--ctx-size", type=int, default=2048)
    ap.add_argument("-n", "--predict", type=int, default=200)
    ap.add_argument("-np", "--parallel", type=int, default=4)
    ap.add_argument("-b", "--batch-size", type=int, default=2048)
    ap.add_argument("--device", default="cuda")
    ap.add_argument("--static", default=STATIC_DIR)
    ap.add_argument("--lora", default=os.environ.get("DLP_ARG_LORA"), help="comma separated LoRA adapter GGUFs (path[:scale])")
    ap.add_argument("--api-key", default=os.environ.get("DLP_API_KEY"), help="require 'Authorization: Bearer <key>' on API routes")
    ap.add_argument("--max-queue", type=int, default=0, help="answer 503 once this many requests are waiting (0 = unbounded)")
    ap.add_argument("--slot-save-path", default=None, help="directory for /slots/{id}?action=save|restore files")
    ap.add_argument("--context-shift", action="store_true", help="slide the context window instead of stopping at n_ctx")
    ap.add_argument("--chat-template", default=None, help="built-in chat format by name instead of the model's own template (chatml, llama3, mistral-v7, gemma, ...)")
    ap.add_argument("-a", "--alias", default=None, help="model name reported by the API")
    ap.add_argument("--api-key-file", default=None, help="file with one API key per line (the first is used)")
    ap.add_argument("--chat-template-file", default=None, help="Jinja chat template read from a file (replaces the model's own)")
    ap.add_argument("--no-context-shift", action="store_true", help="stop at a full context (the default here)")
    ap.add_argument("-to", "--timeout", type=int, default=600, help="accepted: socket time-outs are per connection here")
    ap.add_argument("--models-dir", default=None, help="router mode: directory of .gguf files (same as --models DIR)")
    ap.add_argument("--pooling", default=None, choices=("none", "mean", "cls", "last", "rank"), help="default pooling of /embedding (a request may name its own)")
    add_parallel_args(ap)
    from ..utils.args import add_compat, finish_compat
    add_compat(ap)
    a = finish_compat(ap, ap.parse_args(argv))
    if a.api_key_file:
        a.api_key = next((ln.strip() for ln in open(a.api_key_file) if ln.strip()), a.api_key)
    a.models = a.models or a.models_dir
    maybe_reexec(a, "distributed_llm_pipeline_b200.serving.orchestrator", argv)
    if a.models:
        if os.path.isdir(a.models):
            models = {os.path.splitext(f)[0]: os.path.join(a.models, f) for f in sorted(os.listdir(a.models)) if f.endswith(".gguf")}
        else:
            models = dict(item.split("=", 1) if "=" in item else (os.path.splitext(os.path.basename(item))[0], item) for item in a.models.split(","))
        router = ModelRouter(models, dict(device=a.device, n_ctx=a.ctx_size, n_slots=a.parallel, n_batch=a.batch_size), a.models_max,
                             a.sleep_idle_seconds, a.static, a.predict, replicas=a.replicas, api_key=a.api_key, max_queue=a.max_queue, slot_save_path=a.slot_save_path)
        print(f"dlp_b200 router ready: http://localhost:{a.port} ({len(models)} models)", flush=True)
        try:
            asyncio.run(router.serve(a.host, a.port))
        except KeyboardInterrupt:
            pass
        return
    if not a.model:
        ap.error("-m/--model or --models is required")
    sched, info, logs = build_service(a.model, a.device, a.ctx_size, a.parallel, a.batch_size, lora=lora_from_args(a), **service_kwargs(a))
    if not a.no_warmup:
        logs.append(f"warm-up: decode graphs for 1..{sched.n_slots} sequences captured in {sched.warmup():.2f} s\n")
    for l in logs:
        sys.stderr.write(l)
    sched.context_shift = a.context_shift
    if a.chat_template:
        from . import chat_templates
        if a.chat_template not in chat_templates.names():
            ap.error(f"--chat-template: unknown name {a.chat_template!r}; built-in: {', '.join(chat_templates.names())}")
        info.setdefault("metadata", {})["dlp.chat_template_name"] = a.chat_template
    if a.chat_template_file:
        info.setdefault("metadata", {})["tokenizer.chat_template"] = open(a.chat_template_file).read()
    if a.alias:
        info["name"] = a.alias
    if a.pooling:
        info["pooling"] = a.pooling
    sched.start()
    orch = Orchestrator(sched, info, a.static, a.predict, logs, api_key=a.api_key, max_queue=a.max_queue, slot_save_path=a.slot_save_path)
    print(f"dlp_b200 orchestrator ready: http://localhost:{a.port}", flush=True)
    try:
        asyncio.run(orch.serve(a.host, a.port))
    except KeyboardInterrupt:
        pass
    finally:
        stop_service(sched)


if __name__ == "__main__":
    main()
