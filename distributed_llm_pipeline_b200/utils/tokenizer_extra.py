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


class CharsMap:
    """SentencePiece's precompiled normalisation table (`tokenizer.ggml.precompiled_charsmap`): a 4-byte length, a double-array trie of that
    many bytes whose keys are UTF-8 byte strings, then the replacement strings, NUL-terminated.  Each trie unit is a packed uint32 -
    child check byte in bits 0-7, "has a value" in bit 8, the XOR offset to the children in bits 10-30 (shifted by 8 more when bit 9 is set);
    the value unit behind a key holds the byte offset of its replacement (llama-vocab.cpp:1010-1050, :1069-1114).  `apply` rewrites a text by
    always taking the longest key that matches at the current byte; bytes that are not valid UTF-8 become U+FFFD."""

    def __init__(self, blob: bytes):
        import struct
        n = struct.unpack_from("<I", blob, 0)[0]
        if 4 + n >= len(blob) or n % 4:
            raise ValueError("precompiled_charsmap: the trie does not fit the blob")
        self.units = struct.unpack_from(f"<{n // 4}I", blob, 4)
        self.repl = blob[4 + n:]

    @staticmethod
    def _offset(u: int) -> int:
        return (u >> 10) << ((u & 0x200) >> 6)

    def longest(self, data: bytes, pos: int) -> tuple[int, bytes | None]:
        """(bytes consumed, replacement) for the longest key matching data[pos:], or (0, None)."""
        units, node = self.units, 0
        node ^= self._offset(units[0])
        best_len, best_val = 0, -1
        for i in range(pos, len(data)):
            c = data[i]
            if c == 0:
                break
            node ^= c
            if node >= len(units) or (units[node] & 0x800000FF) != c:
                break
            u = units[node]
            node ^= self._offset(u)
            if (u >> 8) & 1:
                best_len, best_val = i - pos + 1, units[node] & 0x7FFFFFFF
        if best_len == 0:
            return 0, None
        end = self.repl.index(b"\0", best_val)
        return best_len, self.repl[best_val:end]

    def apply(self, text: str, keep: "dict | None" = None) -> str:
        data, out, pos = text.encode("utf-8"), bytearray(), 0
        while pos < len(data):
            if keep is not None:            # user-defined tokens pass through untouched (trie of their UTF-8 bytes)
                node, j, hit = keep, pos, 0
                while j < len(data) and data[j] in node:
                    node = node[data[j]]
                    j += 1
                    if None in node:
                        hit = j - pos
                if hit:
                    out += data[pos:pos + hit]
                    pos += hit
                    continue
            n, rep = self.longest(data, pos)
            if n:
                out += rep
                pos += n
                continue
            b0 = data[pos]
            width = 1 if b0 < 0x80 else 2 if 0xC0 <= b0 < 0xE0 else 3 if 0xE0 <= b0 < 0xF0 else 4 if 0xF0 <= b0 < 0xF8 else 0
            chunk = data[pos:pos + width]
            try:
                if width == 0 or len(chunk) < width:
                    raise UnicodeDecodeError("utf-8", bytes(chunk), 0, 1, "truncated")
                chunk.decode("utf-8")
                out += chunk
                pos += width
            except UnicodeDecodeError:
                out += b"\xef\xbf\xbd"
                pos += 1
        return out.decode("utf-8", errors="replace")


class UgmCodec:
    """Unigram language-model segmentation (SentencePiece / T5): the split of the normalised text that maximises the sum of token
    scores, found by a forward Viterbi pass over a prefix trie.  Characters no token covers cost `min_score - 10` and map to UNK;
    user-defined tokens match with score 0.  Normalisation: the model's own `precompiled_charsmap` (`CharsMap`) when the file carries one,
    else NFKC (for text without compatibility characters both are the identity), then
    runs of spaces collapse (`remove_extra_whitespaces`), a space is prefixed (`add_space_prefix`) and spaces become U+2581."""

    def __init__(self, tokens: list[str], scores: list[float], types: list[int], unk_id: int, add_space_prefix: bool = True,
                 remove_extra_whitespaces: bool = True, charsmap: bytes | None = None):
        self.tokens, self.types, self.unk_id = tokens, types, unk_id
        self.charsmap = CharsMap(charsmap) if charsmap else None
        self.user_trie: dict = {}
        for t, typ in zip(tokens, types):
            if typ == TokenType.USER_DEFINED and t:
                node = self.user_trie
                for byte in t.encode("utf-8"):
                    node = node.setdefault(byte, {})
                node[None] = True
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
        s = self.charsmap.apply(text, self.user_trie or None) if self.charsmap is not None else unicodedata.normalize("NFKC", text)
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
