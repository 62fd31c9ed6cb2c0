#!/usr/bin/env python
"""dlp-gguf: inspect and edit GGUF files - dump, hash, set-metadata, new-metadata.

Parity: the gguf-py command-line scripts (`llama.cpp/gguf-py/gguf/scripts/gguf_dump.py`, `gguf_hash.py`, `gguf_set_metadata.py`,
`gguf_new_metadata.py`) and the `examples/gguf` / `examples/gguf-hash` programs, on this repository's own reader / writer.

  dump          key-value pairs and the tensor table (`--json` for machines, `--no-tensors`)
  hash          sha256 per tensor and over all tensor data in file order: two files with the same weights hash the same whatever
                their metadata says (`--check other.gguf` compares)
  set-metadata  overwrite one existing scalar / string key IN PLACE when the new value has the same encoded size (as the
                reference script does: no data is moved)
  new-metadata  copy to a new file while adding / replacing / removing keys (`--set key=value`, `--remove key`, types follow the
                existing key or `key:type=value`), tensor data carried over byte for byte
"""
import argparse
import hashlib
import json
import os
import struct
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

from distributed_llm_pipeline_b200.gguf import GGMLType, GGUFWriter, open_gguf  # noqa: E402
from distributed_llm_pipeline_b200.gguf.constants import ValueType  # noqa: E402


def _short(v, limit=6):
    if isinstance(v, (list, np.ndarray)):
        items = list(v[:limit])
        return f"[{', '.join(repr(x) if isinstance(x, str) else str(x) for x in items)}{', ...' if len(v) > limit else ''}] ({len(v)} items)"
    return repr(v) if isinstance(v, str) else str(v)


def cmd_dump(a):
    f = open_gguf(a.file)
    f0 = f.files[0]
    if a.json:
        out = {"metadata": {k: (v if not isinstance(v, (list, np.ndarray)) or len(v) <= 64 else {"n_items": len(v), "head": [x if isinstance(x, str) else float(x) for x in v[:8]]})
                            for k, v in f.metadata.items()},
               "tensors": {} if a.no_tensors else {n: {"shape": list(t.shape), "type": GGMLType(t.ggml_type).name, "nbytes": int(t.nbytes)} for n, t in f.tensors.items()}}
        print(json.dumps(out, indent=1, default=lambda o: o.tolist() if isinstance(o, np.ndarray) else float(o) if isinstance(o, np.floating) else int(o)))
        return 0
    print(f"* {a.file}: {len(f.metadata)} key-value pairs, {len(f.tensors)} tensors, {len(f.files)} file(s)")
    for i, (k, v) in enumerate(f.metadata.items(), 1):
        print(f"{i:6d}: {ValueType(f0.metadata_types[k]).name:8s} {k} = {_short(v)}")
    if not a.no_tensors:
        total = 0
        for i, (n, t) in enumerate(f.tensors.items(), 1):
            total += int(t.nbytes)
            print(f"{i:6d}: {int(np.prod(t.shape)):12d} | {' x '.join(str(d) for d in t.shape):24s} | {GGMLType(t.ggml_type).name:6s} | {n}")
        print(f"* tensor data: {total / 2**20:.2f} MiB")
    return 0


def tensor_hashes(path: str):
    f = open_gguf(path)
    whole = hashlib.sha256()
    per = {}
    for n in f.tensors:
        raw = np.ascontiguousarray(f.raw(n)).reshape(-1).view(np.uint8)
        per[n] = hashlib.sha256(raw).hexdigest()
        whole.update(raw)
    return per, whole.hexdigest()


def cmd_hash(a):
    per, whole = tensor_hashes(a.file)
    if not a.quiet:
        for n, h in per.items():
            print(f"sha256  {h}  {a.file}:{n}")
    print(f"sha256  {whole}  {a.file}")
    if a.check:
        per2, whole2 = tensor_hashes(a.check)
        bad = [n for n in per if per2.get(n) != per[n]] + [n for n in per2 if n not in per]
        print("same tensor data" if whole == whole2 and not bad else f"DIFFERENT: {len(bad)} tensor(s), e.g. {bad[:3]}")
        return 0 if whole == whole2 and not bad else 1
    return 0


_SCALAR_FMT = {ValueType.UINT8: "<B", ValueType.INT8: "<b", ValueType.UINT16: "<H", ValueType.INT16: "<h", ValueType.UINT32: "<I", ValueType.INT32: "<i",
               ValueType.FLOAT32: "<f", ValueType.BOOL: "<?", ValueType.UINT64: "<Q", ValueType.INT64: "<q", ValueType.FLOAT64: "<d"}


def _parse(value: str, vt: ValueType):
    if vt == ValueType.STRING:
        return value
    if vt == ValueType.BOOL:
        return value.lower() in ("1", "true", "yes")
    if vt in (ValueType.FLOAT32, ValueType.FLOAT64):
        return float(value)
    return int(value, 0)


def cmd_set(a):
    f = open_gguf(a.file)
    f0 = f.files[0]
    if a.key not in f.metadata:
        raise SystemExit(f"no such key {a.key!r} (new-metadata adds keys)")
    vt = ValueType(f0.metadata_types[a.key])
    if vt == ValueType.ARRAY:
        raise SystemExit("arrays cannot be changed in place; use new-metadata")
    old = f.metadata[a.key]
    new = _parse(a.value, vt)
    enc_old = old.encode("utf-8") if vt == ValueType.STRING else struct.pack(_SCALAR_FMT[vt], old)
    enc_new = new.encode("utf-8") if vt == ValueType.STRING else struct.pack(_SCALAR_FMT[vt], new)
    if len(enc_old) != len(enc_new):
        raise SystemExit(f"in-place edit needs the same encoded size ({len(enc_old)} bytes); use new-metadata")
    key_enc = struct.pack("<Q", len(a.key.encode())) + a.key.encode() + struct.pack("<I", int(vt))
    f.close()
    with open(a.file, "r+b") as fh:
        head = fh.read(64 << 20)
        pos = head.find(key_enc)
        if pos < 0 or head.find(key_enc, pos + 1) >= 0:
            raise SystemExit("could not locate the key unambiguously")
        off = pos + len(key_enc) + (8 if vt == ValueType.STRING else 0)
        if head[off:off + len(enc_old)] != enc_old:
            raise SystemExit("file layout does not match the parsed value; refusing to write")
        if a.dry_run:
            print(f"would set {a.key}: {old!r} -> {new!r} at byte {off}")
            return 0
        fh.seek(off)
        fh.write(enc_new)
    print(f"{a.key}: {old!r} -> {new!r}")
    return 0


def cmd_new(a):
    f = open_gguf(a.file)
    f0 = f.files[0]
    sets = {}
    for item in a.set or []:
        k, _, v = item.partition("=")
        k, _, tname = k.partition(":")
        if tname:
            vt = ValueType[tname.upper()]
        elif k in f.metadata:
            vt = ValueType(f0.metadata_types[k])
        else:
            vt = ValueType.STRING
        if vt == ValueType.ARRAY:
            raise SystemExit(f"{k}: array values cannot be given on the command line")
        sets[k] = (_parse(v, vt), vt)
    w = GGUFWriter(a.output)
    for k, v in f.metadata.items():
        if k in (a.remove or []) or k in sets:
            continue
        w.add(k, v, f0.metadata_types[k]) if not isinstance(v, (list, np.ndarray)) else w.add(k, v)
    for k, (v, vt) in sets.items():
        w.add(k, v, vt)
    for n, t in f.tensors.items():
        w.add_tensor(n, t.shape, t.ggml_type, lambda n=n: np.ascontiguousarray(f.raw(n)).reshape(-1))
    w.write()
    print(f"{a.file} -> {a.output}: {len(sets)} key(s) set, {len([k for k in (a.remove or []) if k in f.metadata])} removed")
    return 0


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    sub = ap.add_subparsers(dest="cmd", required=True)
    p = sub.add_parser("dump"); p.add_argument("file"); p.add_argument("--json", action="store_true"); p.add_argument("--no-tensors", action="store_true"); p.set_defaults(fn=cmd_dump)
    p = sub.add_parser("hash"); p.add_argument("file"); p.add_argument("--check"); p.add_argument("-q", "--quiet", action="store_true"); p.set_defaults(fn=cmd_hash)
    p = sub.add_parser("set-metadata"); p.add_argument("file"); p.add_argument("key"); p.add_argument("value"); p.add_argument("--dry-run", action="store_true"); p.set_defaults(fn=cmd_set)
    p = sub.add_parser("new-metadata"); p.add_argument("file"); p.add_argument("output"); p.add_argument("--set", action="append"); p.add_argument("--remove", action="append"); p.set_defaults(fn=cmd_new)
    a = ap.parse_args(argv)
    return a.fn(a)


if __name__ == "__main__":
    sys.exit(main())
