"""Sequence state save / restore (inference-time checkpointing).

Parity: llama.cpp session files and per-sequence state (`llama_state_seq_save_file` / `llama_state_save_file`,
`llama.cpp/src/llama-context.cpp:2311-2509`; KV part `llama-kv-cache.cpp:1490-1980`) and the server's
`/slots/:id?action=save|restore|erase` (`tools/server/server.cpp:196-197`).  File layout (little endian):
  magic "DLPS" | u32 version | u32 json_len | json header | K/V payload: for each local layer, K then V,
  [n_tokens, Hkv, D] fp16 (token-major, i.e. gathered out of the pages - the file does not depend on the page layout)
The header records the model fingerprint (arch, dims, layer range, TP shard) and the token ids; restore refuses a
mismatching model and works into a fragmented page pool (pages are allocated wherever the free list has them).
"""
from __future__ import annotations

import json
import struct

import numpy as np
import torch

MAGIC = b"DLPS"
VERSION = 1


def _fingerprint(eng) -> dict:
    c = eng.cfg
    return {"arch": c.arch, "n_embd": c.n_embd, "n_layer": c.n_layer, "n_head_kv_local": eng.Hkv, "head_dim": c.head_dim,
            "layers": [eng.w.layer_range.start, eng.w.layer_range.stop], "tp_rank": eng.w.tp_rank, "tp_size": eng.w.tp_size,
            "vocab": c.vocab_size}


def save_sequence(eng, sid: int, path: str, tokens: list[int] | None = None) -> int:
    """Write the KV state (and token ids) of sequence `sid`; returns bytes written."""
    seq = eng.seqs[sid]
    n, ps = seq.length, eng.page_size
    pages = torch.tensor(seq.pages[: (n + ps - 1) // ps], dtype=torch.long, device=eng.device)
    hdr = dict(_fingerprint(eng), n_tokens=n, tokens=list(tokens) if tokens is not None else None, page_size=ps)
    hj = json.dumps(hdr).encode()
    with open(path, "wb") as f:
        f.write(MAGIC + struct.pack("<II", VERSION, len(hj)) + hj)
        for il in eng.w.layer_range:
            for cache in (eng.kv.k[il], eng.kv.v[il]):
                # [pages, Hkv, page, D] -> [tokens, Hkv, D]
                t = cache[pages].permute(0, 2, 1, 3).reshape(-1, eng.Hkv, eng.D)[:n].contiguous()
                f.write(t.cpu().numpy().astype("<f2", copy=False).tobytes())
        return f.tell()


def load_sequence(eng, sid: int, path: str) -> dict:
    """Restore a saved state into (empty or to-be-overwritten) sequence `sid`; returns the header."""
    with open(path, "rb") as f:
        if f.read(4) != MAGIC:
            raise ValueError(f"{path}: not a dlp session file")
        ver, hl = struct.unpack("<II", f.read(8))
        if ver != VERSION:
            raise ValueError(f"{path}: unsupported session version {ver}")
        hdr = json.loads(f.read(hl))
        fp = _fingerprint(eng)
        for k, v in fp.items():
            if hdr.get(k) != v:
                raise ValueError(f"{path}: session was saved for a different model / shard ({k}: {hdr.get(k)} != {v})")
        n = int(hdr["n_tokens"])
        eng.truncate_seq(sid, 0)
        seq = eng.seqs[sid]
        eng.kv.ensure(seq, n)
        ps = eng.page_size
        npg = (n + ps - 1) // ps
        pages = torch.tensor(seq.pages[:npg], dtype=torch.long, device=eng.device)
        per = n * eng.Hkv * eng.D
        for il in eng.w.layer_range:
            for cache in (eng.kv.k[il], eng.kv.v[il]):
                raw = np.frombuffer(f.read(per * 2), dtype="<f2")
                if raw.size != per:
                    raise ValueError(f"{path}: truncated session file")
                t = torch.from_numpy(raw.copy()).view(n, eng.Hkv, eng.D)
                pad = torch.zeros(npg * ps, eng.Hkv, eng.D, dtype=torch.float16)
                pad[:n] = t
                cache[pages] = pad.view(npg, ps, eng.Hkv, eng.D).permute(0, 2, 1, 3).to(eng.device)
        seq.length = n
        eng.seq_len_dev[sid] = n
    return hdr
