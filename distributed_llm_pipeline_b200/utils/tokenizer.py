"""Tokenizers driven by GGUF metadata: SentencePiece-style (`tokenizer.ggml.model = "llama"`) and byte-level BPE
(`"gpt2"`, incl. the Llama-3 pre-tokenizer).

Parity: `llm_tokenizer_spm` / `llm_tokenizer_bpe` (`llama.cpp/src/llama-vocab.cpp:110,279`), token-type handling and
byte fallback as in `llama_vocab`; detokenisation mirrors `token_to_piece`.  Own implementation (heap-based bigram
merging for SPM, rank-based merging for BPE).
"""
from __future__ import annotations

import heapq
from functools import lru_cache

try:
    import regex as re
except ImportError:  # pragma: no cover
    import re

from ..gguf.constants import TokenType

SPM_SPACE = "▁"

GPT2_PAT = r"""'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+"""
LLAMA3_PAT = (r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|"
              r"\s+(?!\S)|\s+")


# Pre-tokenizers by `tokenizer.ggml.pre` (`llama.cpp/src/llama-vocab.cpp:279-480` picks the same expressions per name; they are the split
# rules of each family's own tokenizer.json).  Every expression of a list is applied in turn to the pieces produced by the one before;
# matches and the text between matches are both kept as pieces (`unicode_regex_split`).
_CONTR = r"(?:'[sS]|'[tT]|'[rR][eE]|'[vV][eE]|'[mM]|'[lL][lL]|'[dD])"
_GPT2_CORE = r"""'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)"""
_CJK = "[\u4e00-\u9fa5\u3040-\u309f\u30a0-\u30ff]+"
PRE_TOKENIZERS: dict[str, list[str]] = {
    "default": [GPT2_PAT],
    "llama-bpe": [LLAMA3_PAT],
    "gpt-2": [_GPT2_CORE],
    "qwen2": [_CONTR + r"|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+"],
    "chatglm-bpe": [_CONTR + r"|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+"],
    "seed-coder": [_CONTR + r"|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1}| ?[^\s\p{L}\p{N}\r\n]+|\s*[\r\n]+|\s+(?!\S)|\s+"],
    "starcoder": [r"\p{N}", _GPT2_CORE],
    "falcon": [r"[\p{P}\$\+<=>\^~\|`]+", _GPT2_CORE, "[0-9][0-9][0-9]"],
    "deepseek-coder": ["[\r\n]", r"\s?\p{L}+", r"\s?\p{P}+", "[\u4e00-\u9fa5\u0800-\u4e00\uac00-\ud7ff]+", r"\p{N}"],
    "deepseek-v3": [r"\p{N}{1,3}", _CJK,
                    r"[!\"#$%&'()*+,\-./:;<=>?@\[\\\]^_`{|}~][A-Za-z]+|[^\r\n\p{L}\p{P}\p{S}]?[\p{L}\p{M}]+| ?[\p{P}\p{S}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+"],
    "bloom": [r" ?[^(\s|.,!?…。，、।۔،)]+"],
    "viking": [r" ?[^(\s|.,!?…。，、।۔،)]+", r"\p{N}"],
    "tekken": [r"[^\r\n\p{L}\p{N}]?((?=[\p{L}])([^a-z]))*((?=[\p{L}])([^A-Z]))+|[^\r\n\p{L}\p{N}]?((?=[\p{L}])([^a-z]))+((?=[\p{L}])([^A-Z]))*|\p{N}| ?[^\s\p{L}\p{N}]+[\r\n/]*|\s*[\r\n]+|\s+(?!\S)|\s+"],
    "gpt-4o": [r"[^\r\n\p{L}\p{N}]?((?=[\p{L}])([^a-z]))*((?=[\p{L}])([^A-Z]))+" + _CONTR + r"?|[^\r\n\p{L}\p{N}]?((?=[\p{L}])([^a-z]))+((?=[\p{L}])([^A-Z]))*" + _CONTR
               + r"?|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n/]*|\s*[\r\n]+|\s+(?!\S)|\s+"],
    "bailingmoe": [r"'(?:[sSdDmMtT]|[lL][lL]|[vV][eE]|[rR][eE])|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]|\s+(?!\S)|\s+"],
    "exaone-moe": [_CONTR + r"|[^\r\n\p{L}\p{N}]?(?:\p{L}\p{M}*(?: \p{L}\p{M}*)*)+|\p{N}| ?[^\s\p{L}\p{N}]+[\r\n/]?|\s*[\r\n]|\s+(?!\S)|\s+"],
    "chameleon": ["<sentinel:[0-9]+>", "(IMGIMG)((A|B|C|D|E|F|G|H|I){1,4})Z", r"([\t\n]|    |  )", r"\p{N}", r"[\p{P}!-/:-@\[-`{-~]", _GPT2_CORE],
}
for _name, _same in (("llama3", "llama-bpe"), ("falcon3", "llama-bpe"), ("falcon-h1", "llama-bpe"), ("pixtral", "llama-bpe"), ("midm-2.0", "llama-bpe"), ("lfm2", "llama-bpe"),
                     ("jina-es", "gpt-2"), ("jina-de", "gpt-2"), ("jina-v1-en", "gpt-2"), ("jina-v2-es", "gpt-2"), ("jina-v2-de", "gpt-2"), ("jina-v2-code", "gpt-2"),
                     ("a.x-4.0", "gpt-2"), ("mellum", "gpt-2"), ("modern-bert", "gpt-2"), ("trillion", "gpt-2"), ("granite-docling", "gpt-2"),
                     ("kormo", "qwen2"), ("solar-open", "qwen2"), ("bailingmoe2", "bailingmoe"), ("llada-moe", "bailingmoe"), ("llama-v3", "llama-bpe"), ("dbrx", "llama-bpe"), ("smaug-bpe", "llama-bpe"),
                     ("mpt", "gpt-2"), ("olmo", "gpt-2"), ("jais", "gpt-2"), ("phi-2", "gpt-2"), ("exaone4", "gpt-2"), ("roberta-bpe", "gpt-2"), ("gigachat", "gpt-2"),
                     ("stablelm2", "qwen2"), ("deepseek-r1-qwen", "qwen2"), ("hunyuan", "qwen2"), ("megrez", "qwen2"), ("grok-2", "qwen2"),
                     ("glm4", "chatglm-bpe"), ("refact", "starcoder"), ("command-r", "starcoder"), ("smollm", "starcoder"), ("codeshell", "starcoder"),
                     ("exaone", "starcoder"), ("minerva-7b", "starcoder"), ("poro-chat", "bloom"), ("gpt3-finnish", "bloom"), ("llama4", "gpt-4o"),
                     ("minimax-m2", "gpt-4o"), ("deepseek3-llm", "deepseek-v3"), ("hunyuan-dense", "deepseek-v3")):
    PRE_TOKENIZERS[_name] = PRE_TOKENIZERS[_same]
IGNORE_MERGES = {"llama-bpe", "llama3", "llama-v3", "tekken", "falcon3", "falcon-h1", "pixtral", "midm-2.0", "lfm2"}       # a word that is itself a token is taken whole (llama-vocab.cpp `ignore_merges`)


def pre_split(text: str, exprs) -> list[str]:
    pieces = [text]
    for rx in exprs:
        nxt: list[str] = []
        for p in pieces:
            last = 0
            for m in rx.finditer(p):
                if m.end() == m.start():
                    continue
                if m.start() > last:
                    nxt.append(p[last:m.start()])
                nxt.append(m.group(0))
                last = m.end()
            if last < len(p):
                nxt.append(p[last:])
        pieces = nxt
    return pieces


@lru_cache(maxsize=1)
def _bytes_to_unicode() -> dict[int, str]:
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAC + 1)) + list(range(0xAE, 0xFF + 1))
    cs = bs[:]
    n = 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return {b: chr(c) for b, c in zip(bs, cs)}


class Tokenizer:
    def __init__(self, meta: dict):
        self.model = meta.get("tokenizer.ggml.model", "llama")
        self.tokens: list[str] = list(meta["tokenizer.ggml.tokens"])
        self.n_vocab = len(self.tokens)
        scores = meta.get("tokenizer.ggml.scores")
        self.scores = [float(x) for x in scores] if scores is not None else [0.0] * self.n_vocab
        types = meta.get("tokenizer.ggml.token_type")
        self.types = [int(x) for x in types] if types is not None else [int(TokenType.NORMAL)] * self.n_vocab
        self.bos_id = int(meta.get("tokenizer.ggml.bos_token_id", 1))
        self.eos_id = int(meta.get("tokenizer.ggml.eos_token_id", 2))
        self.unk_id = int(meta.get("tokenizer.ggml.unknown_token_id", 0))
        self.eot_id = meta.get("tokenizer.ggml.eot_token_id")
        self.meta_ids = {k: meta.get(f"tokenizer.ggml.fim_{k}_token_id") for k in ("pre", "suf", "mid")}
        self.add_bos = bool(meta.get("tokenizer.ggml.add_bos_token", self.model == "llama"))
        self.add_space_prefix = bool(meta.get("tokenizer.ggml.add_space_prefix", True))
        self.tok2id = {t: i for i, t in enumerate(self.tokens)}
        self.byte_ids: dict[int, int] = {}
        self.codec = None       # WordPiece / Unigram / RWKV vocabularies (utils/tokenizer_extra.py)
        self.add_eos = bool(meta.get("tokenizer.ggml.add_eos_token", self.model in ("bert", "t5")))
        if self.model in ("bert", "t5", "rwkv"):
            from .tokenizer_extra import RwkvCodec, UgmCodec, WpmCodec
            if self.model == "bert":
                self.codec = WpmCodec(self.tokens, self.unk_id)
                self.bos_id = int(meta.get("tokenizer.ggml.cls_token_id", meta.get("tokenizer.ggml.bos_token_id", 101)))
                self.add_bos = bool(meta.get("tokenizer.ggml.add_bos_token", True))      # [CLS] ... [SEP]
                self.sep_id = int(meta.get("tokenizer.ggml.seperator_token_id", meta.get("tokenizer.ggml.eos_token_id", self.eos_id)))
            elif self.model == "t5":
                cm = meta.get("tokenizer.ggml.precompiled_charsmap")
                self.codec = UgmCodec(self.tokens, self.scores, self.types, self.unk_id, self.add_space_prefix,
                                      bool(meta.get("tokenizer.ggml.remove_extra_whitespaces", True)),
                                      charsmap=bytes(bytearray(int(x) & 0xFF for x in cm)) if cm is not None and len(cm) else None)
                self.add_bos = bool(meta.get("tokenizer.ggml.add_bos_token", False))
            else:
                self.codec = RwkvCodec(self.tokens, self.unk_id)
                self.add_bos = bool(meta.get("tokenizer.ggml.add_bos_token", False))
                self.add_eos = bool(meta.get("tokenizer.ggml.add_eos_token", False))
        elif self.model == "llama":
            for b in range(256):
                tid = self.tok2id.get(f"<0x{b:02X}>")
                if tid is not None:
                    self.byte_ids[b] = tid
        else:
            merges = meta.get("tokenizer.ggml.merges", [])
            self.ranks = {tuple(m.split(" ", 1)): i for i, m in enumerate(merges)}
            pre = str(meta.get("tokenizer.ggml.pre", "default"))
            self.pre = pre if pre in PRE_TOKENIZERS else "default"
            self.pre_exprs = [re.compile(e) for e in PRE_TOKENIZERS[self.pre]]
            self.ignore_merges = pre in IGNORE_MERGES
            self.b2u = _bytes_to_unicode()
            self.u2b = {u: b for b, u in self.b2u.items()}
        self.special = {t: i for i, t in enumerate(self.tokens) if self.types[i] in (TokenType.CONTROL, TokenType.USER_DEFINED) and len(t) > 1}
        self.stop_ids = {self.eos_id} | ({int(self.eot_id)} if self.eot_id is not None else set())

    @classmethod
    def from_gguf(cls, src):
        from ..gguf import open_gguf
        f = open_gguf(src) if isinstance(src, str) else src
        return cls(f.metadata)

    # ---- encode -----------------------------------------------------------------------------------
    def _spm_encode(self, text: str) -> list[int]:
        if not text:
            return []
        syms = list(text)
        n = len(syms)
        prev = list(range(-1, n - 1))
        nxt = list(range(1, n + 1))
        nxt[-1] = -1
        alive = [True] * n
        heap: list = []

        def push(i):
            j = nxt[i]
            if i < 0 or j < 0:
                return
            piece = syms[i] + syms[j]
            tid = self.tok2id.get(piece)
            if tid is not None:
                heapq.heappush(heap, (-self.scores[tid], i, piece))

        for i in range(n - 1):
            push(i)
        while heap:
            _, i, piece = heapq.heappop(heap)
            j = nxt[i] if alive[i] else -1
            if j < 0 or not alive[j] or syms[i] + syms[j] != piece:
                continue
            syms[i] = piece
            alive[j] = False
            nxt[i] = nxt[j]
            if nxt[j] >= 0:
                prev[nxt[j]] = i
            if prev[i] >= 0:
                push(prev[i])
            push(i)
        out: list[int] = []
        i = 0
        while i >= 0:
            s = syms[i]
            tid = self.tok2id.get(s)
            if tid is not None:
                out.append(tid)
            else:
                for b in s.encode("utf-8"):
                    out.append(self.byte_ids.get(b, self.unk_id))
            i = nxt[i]
        return out

    def _bpe_word(self, word: str) -> list[str]:
        parts = list(word)
        while len(parts) > 1:
            best, bi = None, -1
            for i in range(len(parts) - 1):
                r = self.ranks.get((parts[i], parts[i + 1]))
                if r is not None and (best is None or r < best):
                    best, bi = r, i
            if best is None:
                break
            parts[bi:bi + 2] = [parts[bi] + parts[bi + 1]]
        return parts

    def _encode_plain(self, text: str) -> list[int]:
        if self.codec is not None:
            return self.codec.encode(text)
        if self.model == "llama":
            t = text.replace(" ", SPM_SPACE)
            return self._spm_encode(t)
        out: list[int] = []
        for m in pre_split(text, self.pre_exprs):
            word = "".join(self.b2u[b] for b in m.encode("utf-8"))
            if self.ignore_merges and word in self.tok2id:
                out.append(self.tok2id[word])
                continue
            for piece in self._bpe_word(word):
                tid = self.tok2id.get(piece)
                if tid is not None:
                    out.append(tid)
                else:
                    out.extend(self.tok2id.get(ch, self.unk_id) for ch in piece)
        return out

    def encode(self, text: str, add_bos: bool | None = None, parse_special: bool = True) -> list[int]:
        ids: list[int] = []
        if add_bos if add_bos is not None else self.add_bos:
            ids.append(self.bos_id)
        segments: list[tuple[str, bool]] = [(text, False)]
        if parse_special and self.special:
            pat = re.compile("|".join(re.escape(t) for t in sorted(self.special, key=len, reverse=True)))
            segments, last = [], 0
            for m in pat.finditer(text):
                if m.start() > last:
                    segments.append((text[last:m.start()], False))
                segments.append((m.group(0), True))
                last = m.end()
            if last < len(text):
                segments.append((text[last:], False))
        prev_special = True        # SPM: a text fragment gets the space prefix at the start AND behind every special token (llama-vocab.cpp SPM case)
        for seg, is_special in segments:
            if is_special:
                ids.append(self.special[seg])
            else:
                if self.model == "llama" and prev_special and self.add_space_prefix:
                    seg = " " + seg
                ids.extend(self._encode_plain(seg))
            prev_special = is_special
        if getattr(self, "add_eos", False) and (add_bos if add_bos is not None else True):
            ids.append(getattr(self, "sep_id", self.eos_id))          # bert: [SEP], t5: </s>
        return ids

    # ---- decode -----------------------------------------------------------------------------------
    def token_bytes(self, tid: int) -> bytes:
        if tid < 0 or tid >= self.n_vocab:
            return b""
        t, typ = self.tokens[tid], self.types[tid]
        if self.codec is not None:
            return self.codec.token_bytes(tid, typ)
        if typ == TokenType.CONTROL:
            return b""
        if self.model == "llama":
            if typ == TokenType.BYTE:
                return bytes([int(t[3:5], 16)])
            return t.replace(SPM_SPACE, " ").encode("utf-8")
        return bytes(self.u2b.get(ch, ord("?")) for ch in t)

    def fim_ids(self) -> dict | None:
        """Fill-in-the-middle token ids (`tokenizer.ggml.fim_{pre,suf,mid}_token_id`, or the usual spellings in the vocab)."""
        out = {}
        for key, names in (("pre", ("<|fim_prefix|>", "<fim_prefix>", "<PRE>", "▁<PRE>")), ("suf", ("<|fim_suffix|>", "<fim_suffix>", "<SUF>", "▁<SUF>")),
                           ("mid", ("<|fim_middle|>", "<fim_middle>", "<MID>", "▁<MID>"))):
            tid = self.meta_ids.get(key)
            if tid is None:
                tid = next((self.special[n] for n in names if n in self.special), None)
            if tid is None:
                return None
            out[key] = int(tid)
        return out

    def token_texts(self) -> list[str | None]:
        """Text of every token for grammar matching: None for control tokens and for byte fragments that are not valid
        UTF-8 on their own (the grammar sampler never picks those)."""
        if getattr(self, "_texts", None) is None:
            out: list[str | None] = []
            for tid in range(self.n_vocab):
                if self.types[tid] in (TokenType.CONTROL, TokenType.UNKNOWN, TokenType.UNUSED):
                    out.append(None)
                    continue
                try:
                    out.append(self.token_bytes(tid).decode("utf-8") or None)
                except UnicodeDecodeError:
                    out.append(None)
            self._texts = out
        return self._texts

    def decode_with_specials(self, ids: list[int]) -> str:
        """Text with control tokens written out (`<s> one two`), nothing stripped: what llama-server reports as the `prompt` it ran."""
        out = bytearray()
        for i in ids:
            if 0 <= i < self.n_vocab and self.types[i] == TokenType.CONTROL:
                out += self.tokens[i].encode("utf-8")
            else:
                out += self.token_bytes(i)
        return out.decode("utf-8", errors="replace")

    def decode(self, ids: list[int], strip_leading_space: bool = True) -> str:
        data = b"".join(self.token_bytes(i) for i in ids)
        s = data.decode("utf-8", errors="replace")
        if self.model in ("llama", "bert", "t5") and strip_leading_space and self.add_space_prefix and s.startswith(" "):
            s = s[1:]
        return s


class StreamDecoder:
    """Incremental detokeniser: holds back incomplete UTF-8 sequences so a multi-byte character is never split
    across stream events (the reference's 64-byte stdout chunks + from_utf8_lossy produce U+FFFD there,
    `orchestrator/src/main.rs:86-90`)."""

    def __init__(self, tok: Tokenizer):
        self.tok = tok
        self.pending = b""
        self.first = True

    def push(self, tid: int) -> str:
        self.pending += self.tok.token_bytes(tid)
        cut = len(self.pending)
        # find the longest prefix that is valid UTF-8
        for back in range(1, min(4, len(self.pending)) + 1):
            b = self.pending[-back]
            if b & 0xC0 == 0x80:
                continue
            need = 1 if b < 0x80 else 2 if b >> 5 == 0b110 else 3 if b >> 4 == 0b1110 else 4 if b >> 3 == 0b11110 else 1
            if need > back:
                cut = len(self.pending) - back
            break
        out, self.pending = self.pending[:cut], self.pending[cut:]
        # (a generated piece keeps its leading space - "one" + " two" - exactly as the reference's token_to_piece gives it to llama-cli
        # and llama-server; only detokenising a whole sequence, `Tokenizer.decode`, drops the space the SPM prefix rule added)
        return out.decode("utf-8", errors="replace")

    def flush(self) -> str:
        s = self.pending.decode("utf-8", errors="replace")
        self.pending = b""
        return s
