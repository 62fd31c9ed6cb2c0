"""The other GGUF vocabulary kinds: WordPiece ("bert"), Unigram ("t5") and the RWKV byte trie ("rwkv").

Reference: `llama.cpp/src/llama-vocab.cpp` - `llm_tokenizer_wpm_session` (:674-780), `llm_tokenizer_ugm_session` (:784-1090),
`llm_tokenizer_rwkv_session` (:1180-1260).  Each codec turns one plain text segment into token ids (`encode`) and a token id back
into bytes (`token_bytes`); special-token splitting, BOS / EOS handling and streaming decode stay in `utils/tokenizer.py`.
"""
from __future__ import annotations

import unicodedata

from ..gguf.constants import TokenType

SPACE = "▁"      # the "lower one eighth block" word-boundary marker


def _is_cjk(cp: int) -> bool:
    return (0x4E00 <= cp <= 0x9FFF or 0x3400 <= cp <= 0x4DBF or 0x20000 <= cp <= 0x2A6DF or 0x2A700 <= cp <= 0x2B73F
            or 0x2B740 <= cp <= 0x2B81F or 0x2B920 <= cp <= 0x2CEAF or 0xF900 <= cp <= 0xFAFF or 0x2F800 <= cp <= 0x2FA1F)


class WpmCodec:
    """WordPiece as GGUF stores it: word-initial pieces carry a leading U+2581, continuation pieces are bare (the converter
    strips "##").  Text is NFD-normalised and lower-cased, split at white space with every punctuation mark / ASCII symbol /
    CJK ideograph a word of its own; each word is covered greedily by the longest vocabulary entries, or becomes UNK as a whole."""

    def __init__(self, tokens: list[str], unk_id: int):
        self.tok2id = {t: i for i, t in enumerate(tokens)}
        self.tokens = tokens
        self.unk_id = unk_id
        self.max_len = max((len(t) for t in tokens), default=1)

    @staticmethod
    def words(text: str) -> list[str]:
        out = [""]
        for ch in unicodedata.normalize("NFD", text):
            cp, cat = ord(ch), unicodedata.category(ch)
            if ch.isspace() or cat == "Zs":
                if out[-1]:
                    out.append("")
                continue
            if cp == 0 or cp == 0xFFFD or cat in ("Cc", "Cf"):
                continue
            low = ch.lower()
            if cat.startswith("P") or (cp < 0x7F and cat.startswith("S")) or _is_cjk(cp):
                if out[-1]:
                    out.append("")
                out[-1] = low
                out.append("")
            else:
                out[-1] += low
        if not out[-1]:
            out.pop()
        return out

    def encode(self, text: str) -> list[int]:
        ids: list[int] = []
        for word in self.words(text):
            w = SPACE + word
            n, i, piece_ids = len(w), 0, []
            while i < n:
                for j in range(min(n, i + self.max_len), i, -1):
                    tid = self.tok2id.get(w[i:j])
                    if tid is not None:
                        piece_ids.append(tid)
                        i = j
                        break
                else:
                    piece_ids = []
                    break
            ids.extend(piece_ids or [self.unk_id])
        return ids

    def token_bytes(self, tid: int, typ: int) -> bytes:
        if typ == TokenType.CONTROL:
            return b""
        return self.tokens[tid].replace(SPACE, " ").encode("utf-8")


class UgmCodec:
    """Unigram language-model segmentation (SentencePiece / T5): the split of the normalised text that maximises the sum of token
    scores, found by a forward Viterbi pass over a prefix trie.  Characters no token covers cost `min_score - 10` and map to UNK;
    user-defined tokens match with score 0.  Normalisation: NFKC stands in for the SentencePiece `nmt_nfkc` character map (the
    reference reads the model's `precompiled_charsmap`; for text without compatibility characters both are the identity), then
    runs of spaces collapse (`remove_extra_whitespaces`), a space is prefixed (`add_space_prefix`) and spaces become U+2581."""

    def __init__(self, tokens: list[str], scores: list[float], types: list[int], unk_id: int, add_space_prefix: bool = True,
                 remove_extra_whitespaces: bool = True):
        self.tokens, self.types, self.unk_id = tokens, types, unk_id
        self.add_space_prefix, self.remove_extra_ws = add_space_prefix, remove_extra_whitespaces
        self.trie: dict = {}
        normal = [s for s, t in zip(scores, types) if t == TokenType.NORMAL]
        self.min_score = min(normal) if normal else 0.0
        self.unk_score = self.min_score - 10.0
        for tid, (t, typ) in enumerate(zip(tokens, types)):
            if typ not in (TokenType.NORMAL, TokenType.USER_DEFINED, TokenType.UNUSED) or not t:
                continue
            node = self.trie
            for ch in t:
                node = node.setdefault(ch, {})
            node[None] = (tid, 0.0 if typ == TokenType.USER_DEFINED else float(scores[tid]))

    def normalize(self, text: str) -> str:
        s = unicodedata.normalize("NFKC", text)
        if self.remove_extra_ws:
            s = " ".join(p for p in s.split(" ") if p) if s.strip(" ") else ""
        if self.add_space_prefix and s:
            s = " " + s
        return s.replace(" ", SPACE)

    def encode(self, text: str) -> list[int]:
        s = self.normalize(text)
        n = len(s)
        if n == 0:
            return []
        best = [(-float("inf"), -1, self.unk_id)] * (n + 1)      # (score, start, token) of the best path ending at each position
        best[0] = (0.0, -1, -1)
        for i in range(n):
            base = best[i][0]
            if base == -float("inf"):
                continue
            node, j, single = self.trie, i, False
            while j < n and s[j] in node:
                node = node[s[j]]
                j += 1
                if None in node:
                    tid, sc = node[None]
                    if base + sc > best[j][0]:
                        best[j] = (base + sc, i, tid)
                    single = single or j == i + 1
            if not single and base + self.unk_score > best[i + 1][0]:
                best[i + 1] = (base + self.unk_score, i, self.unk_id)
        out, pos = [], n
        while pos > 0:
            _, start, tid = best[pos]
            if not (out and tid == self.unk_id and out[-1] == self.unk_id):      # a run of unknown characters is one UNK
                out.append(tid)
            pos = start
        return out[::-1]

    def token_bytes(self, tid: int, typ: int) -> bytes:
        if typ in (TokenType.CONTROL, TokenType.UNKNOWN):
            return b""
        if typ == TokenType.BYTE:
            return bytes([int(self.tokens[tid][3:5], 16)])
        return self.tokens[tid].replace(SPACE, " ").encode("utf-8")


def unescape_rwkv(tok: str) -> bytes:
    """RWKV vocabularies are stored as escaped Python-style byte strings (`\\n`, `\\t`, `\\xNN`, `\\\\`)."""
    out, i, n = bytearray(), 0, len(tok)
    simple = {"n": 0x0A, "t": 0x09, "r": 0x0D, "\\": 0x5C, "'": 0x27, '"': 0x22}
    while i < n:
        ch = tok[i]
        if ch == "\\" and i + 1 < n:
            nx = tok[i + 1]
            if nx == "x" and i + 3 < n:
                try:
                    out.append(int(tok[i + 2:i + 4], 16))
                    i += 4
                    continue
                except ValueError:
                    pass
            if nx in simple:
                out.append(simple[nx])
                i += 2
                continue
        out.extend(ch.encode("utf-8"))
        i += 1
    return bytes(out)


class RwkvCodec:
    """Greedy longest match over raw bytes with a trie of the (unescaped) vocabulary."""

    def __init__(self, tokens: list[str], unk_id: int):
        self.unk_id = unk_id
        self.raw = [unescape_rwkv(t) for t in tokens]
        self.trie: dict = {}
        for tid, b in enumerate(self.raw):
            if not b:
                continue
            node = self.trie
            for byte in b:
                node = node.setdefault(byte, {})
            node.setdefault(None, tid)

    def encode(self, text: str) -> list[int]:
        data, out, i = text.encode("utf-8"), [], 0
        while i < len(data):
            node, j, hit, hit_end = self.trie, i, None, i
            while j < len(data) and data[j] in node:
                node = node[data[j]]
                j += 1
                if None in node:
                    hit, hit_end = node[None], j
            if hit is None:
                out.append(self.unk_id)
                i += 1
            else:
                out.append(hit)
                i = hit_end
        return out

    def token_bytes(self, tid: int, typ: int) -> bytes:
        return b"" if typ == TokenType.CONTROL else self.raw[tid]
