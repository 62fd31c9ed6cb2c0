#!/usr/bin/env python
"""serve-bench: concurrent requests against an HTTP server, aggregate tokens/s and time-to-first-token percentiles.

Counterpart of the reference's `scripts/server-bench.py` / `tools/server/bench` (k6): the same client drives
  * this repository's orchestrator (`/chat`, the reference orchestrator's wire format; or `/completion` streaming), and
  * the reference's `llama-server` (`/completion` streaming, `tools/server/server.cpp:164-197`),
so both arms of BASELINE config 5 (concurrent serving through the front-end) are measured by one instrument.

Prompts are random token ids of a fixed length sent as token arrays where the route accepts them (`/completion`), else as text
of about that many tokens - valid with random-init weights.  Every request is its own TCP connection; TTFT is measured from
send to the first generated-token event on the client.
"""
import argparse
import json
import random
import socket
import statistics
import threading
import time


def _request(host, port, path, body: bytes, on_data):
    s = socket.create_connection((host, port), timeout=600)
    s.sendall((f"POST {path} HTTP/1.1\r\nHost: x\r\nContent-Type: application/json\r\nContent-Length: {len(body)}\r\nConnection: close\r\n\r\n").encode() + body)
    buf, head_done, chunked = b"", False, False
    while True:
        d = s.recv(65536)
        if not d:
            break
        buf += d
        if not head_done and b"\r\n\r\n" in buf:
            head, _, buf = buf.partition(b"\r\n\r\n")
            head_done = True
            chunked = b"transfer-encoding: chunked" in head.lower()
        if head_done:
            while b"\n\n" in buf or b"\r\n\r\n" in buf:      # SSE events end with a blank line
                sep = b"\r\n\r\n" if (b"\r\n\r\n" in buf and (b"\n\n" not in buf or buf.index(b"\r\n\r\n") < buf.index(b"\n\n"))) else b"\n\n"
                ev, _, buf = buf.partition(sep)
                on_data(ev, chunked)
    s.close()


def run_one(a, idx, out, seed_off=0):
    rnd = random.Random(1000 + idx + seed_off)
    toks = [rnd.randrange(3, a.vocab) for _ in range(a.prompt_tokens)]
    t0 = time.perf_counter()
    st = {"first": None, "n": 0, "last": t0}

    def on_data(ev: bytes, chunked):
        for line in ev.split(b"\n"):
            line = line.strip()
            if not line.startswith(b"data:"):
                continue
            try:
                d = json.loads(line[5:].strip())
            except json.JSONDecodeError:
                continue
            if a.route == "/chat":
                if d.get("msg_type") != "token" or not d.get("content"):
                    continue
            else:
                if not d.get("content") and not d.get("tokens"):
                    continue
            now = time.perf_counter()
            if st["first"] is None:
                st["first"] = now - t0
            st["n"] += 1
            st["last"] = now

    if a.route == "/chat":
        body = {"prompt": " ".join(f"w{t % 997}" for t in toks[: max(1, a.prompt_tokens // 2)])}
    else:
        body = {"prompt": toks, "n_predict": a.n_predict, "stream": True, "temperature": 0.0, "ignore_eos": True, "cache_prompt": False}
    try:
        _request(a.host, a.port, a.route, json.dumps(body).encode(), on_data)
    except OSError as e:
        out[idx] = {"error": str(e)}
        return
    out[idx] = {"ttft_s": st["first"], "events": st["n"], "total_s": st["last"] - t0}


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=3005)
    ap.add_argument("--route", default="/completion", choices=["/completion", "/chat"])
    ap.add_argument("-c", "--concurrency", default="1,4,16", help="comma separated numbers of simultaneous requests")
    ap.add_argument("--prompt-tokens", type=int, default=128)
    ap.add_argument("-n", "--n-predict", type=int, default=64)
    ap.add_argument("--vocab", type=int, default=32000)
    ap.add_argument("--warmup-rounds", type=int, default=1, help="unmeasured rounds at each concurrency level before the timed one")
    ap.add_argument("--tag", default="")
    ap.add_argument("-o", "--output", default="")
    a = ap.parse_args()
    rows = []
    for conc in [int(x) for x in a.concurrency.split(",")]:
        for _ in range(a.warmup_rounds):      # both arms get the same treatment: first-use costs stay out of the timed round
            out = [None] * conc
            ts = [threading.Thread(target=run_one, args=(a, i, out, 10000)) for i in range(conc)]
            [t.start() for t in ts]
            [t.join() for t in ts]
        out = [None] * conc
        ts = [threading.Thread(target=run_one, args=(a, i, out)) for i in range(conc)]
        t0 = time.perf_counter()
        [t.start() for t in ts]
        [t.join() for t in ts]
        wall = time.perf_counter() - t0
        ok = [o for o in out if o and "error" not in o and o["ttft_s"] is not None]
        ev = sum(o["events"] for o in ok)
        # one event per generated token on both servers (the orchestrator may merge a multi-byte piece: tokens >= events)
        gen_tokens = ev if a.route == "/chat" else conc * a.n_predict if len(ok) == conc else ev
        row = {"tag": a.tag, "route": a.route, "concurrency": conc, "ok": len(ok), "prompt_tokens": a.prompt_tokens, "n_predict": a.n_predict,
               "wall_s": round(wall, 3), "agg_tok_s": round(gen_tokens / wall, 2) if wall > 0 else None,
               "ttft_ms_p50": round(statistics.median(o["ttft_s"] for o in ok) * 1e3, 2) if ok else None,
               "ttft_ms_max": round(max(o["ttft_s"] for o in ok) * 1e3, 2) if ok else None, "events": ev}
        rows.append(row)
        print(json.dumps(row), flush=True)
    if a.output:
        json.dump(rows, open(a.output, "w"), indent=1)


if __name__ == "__main__":
    main()
