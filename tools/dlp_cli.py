#!/usr/bin/env python
"""dlp-cli: single-process chat / completion front-end (the role `llama-cli` plays behind the reference orchestrator).

Generated text goes to stdout (unbuffered), everything else to stderr - the split the reference orchestrator relies on
(`orchestrator/src/main.rs:62-95`).  `--rpc a:p,b:p` runs the model on dlp-worker CPU stages, like the reference's
`llama-cli --rpc 127.0.0.1:50052,127.0.0.1:50053` (`main.rs:47-48`).
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from distributed_llm_pipeline_b200.utils.args import COMMON, GENERATION, SAMPLING, SPECULATIVE, apply_loader_flags, build_parser, lora_from_args, sampling_from_args  # noqa: E402
from distributed_llm_pipeline_b200.utils.log import get_logger  # noqa: E402


def main(argv=None):
    a = build_parser("dlp-cli - generate text from a GGUF model", [COMMON, SAMPLING, GENERATION, SPECULATIVE]).parse_args(argv)
    apply_loader_flags(a)
    if a.list_devices:
        from distributed_llm_pipeline_b200.utils.fit import list_devices
        print("Available devices:\n  " + "\n  ".join(list_devices()))
        return
    if not a.model:
        sys.exit("dlp-cli: -m model.gguf is required")
    log = get_logger()
    log.verbosity = 4 if a.verbose else 3
    if a.log_file:
        log.add_file(a.log_file)
    import torch

    from distributed_llm_pipeline_b200.serving.scheduler import Request, Scheduler
    from distributed_llm_pipeline_b200.utils.tokenizer import Tokenizer
    t0 = time.time()
    pp, tp = a.pp or 1, a.tp or 1
    sched = None
    if pp * tp > 1 and not a.rpc:
        # a pp x tp grid of processes: this command re-starts itself under torchrun; the leader rank runs the conversation, the
        # other ranks replay its engine calls (parallel/spmd.py) and leave inside build_service when the leader is done
        if "WORLD_SIZE" not in os.environ:
            from distributed_llm_pipeline_b200.parallel.spmd import reexec_under_torchrun
            reexec_under_torchrun(pp * tp, list(sys.argv[1:] if argv is None else argv), script=os.path.abspath(__file__))
        from distributed_llm_pipeline_b200.serving.orchestrator import build_service, stop_service
        sched, _info, logs = build_service(a.model, a.device, a.ctx_size, a.parallel, min(a.batch_size, a.ubatch_size), lora=lora_from_args(a),
                                           pp=pp, tp=tp, tp_impl=a.tp_impl, tensor_split=a.tensor_split)
        for line in logs:
            log.info(line.rstrip("\n"))
        eng, tok = sched.eng, sched.tok
    elif a.rpc:
        from distributed_llm_pipeline_b200.parallel.rpc import RpcEngine
        eng = RpcEngine(a.rpc.split(","), n_ctx=a.ctx_size, max_seqs=a.parallel)
        for line in eng.describe():
            log.info(line)
        tok = Tokenizer.from_gguf(a.model)
    else:
        from distributed_llm_pipeline_b200.gguf import open_gguf
        from distributed_llm_pipeline_b200.models.llama import LlamaEngine
        from distributed_llm_pipeline_b200.models.weights import load_weights
        dev = a.device if (a.device != "cuda" or torch.cuda.is_available()) else "cpu"
        f = open_gguf(a.model)
        if a.fit and dev != "cpu":
            from distributed_llm_pipeline_b200.models.config import config_from_gguf
            from distributed_llm_pipeline_b200.utils.fit import fit_params
            fr = fit_params(f, config_from_gguf(f.metadata, f.tensors), a.ctx_size, a.parallel, torch.cuda.mem_get_info()[0])
            log.info(f"fit: {fr.note}")
            if not fr.fits:
                sys.exit(f"dlp-cli: {fr.note}")
            a.ctx_size, a.parallel = fr.n_ctx, fr.n_slots
        w = load_weights(f, dev, lora=lora_from_args(a))
        eng = LlamaEngine(w, n_ctx=a.ctx_size, max_seqs=a.parallel, page_size=64 if dev != "cpu" else 16,
                          max_batch_tokens=min(a.batch_size, a.ubatch_size))
        tok = Tokenizer(f.metadata)
        log.info(f"load_tensors: offloaded {w.cfg.n_layer + 1}/{w.cfg.n_layer + 1} layers to GPU ({dev}), {w.nbytes() / 2**20:.1f} MiB")
    log.info(f"model loaded in {time.time() - t0:.2f} s | n_ctx {a.ctx_size} | slots {a.parallel}")
    prompt = open(a.file).read() if a.file else a.prompt
    if a.escape and not a.no_escape and not a.file:      # -e (the reference's default): \n \t \' \" \\ in a prompt given on the command line
        for esc, ch in (("\\n", "\n"), ("\\t", "\t"), ("\\'", "'"), ('\\"', '"'), ("\\\\", "\\")):
            prompt = prompt.replace(esc, ch)
    system = open(a.system_prompt_file).read() if a.system_prompt_file else a.system_prompt
    chat_mode = bool(system) or a.conversation or a.interactive or a.interactive_first
    if a.no_conversation:
        a.single_turn, chat_mode = True, False
    messages: list[dict] = [{"role": "system", "content": system}] if system else []
    if chat_mode:       # the conversation goes through the model's chat template (llama-cli's conversation mode)
        from distributed_llm_pipeline_b200.gguf import open_gguf as _open
        from distributed_llm_pipeline_b200.serving.orchestrator import apply_chat_template
        md = _open(a.model).metadata
        chat_meta = {"tokenizer.chat_template": md.get("tokenizer.chat_template", ""), "bos_token": "", "eos_token": ""}
        if a.interactive_first and not prompt and sys.stdin.isatty():
            prompt = input("> ")
        messages.append({"role": "user", "content": a.in_prefix + prompt + a.in_suffix})
        prompt = apply_chat_template(chat_meta, messages)
    cvecs = [(p_, 1.0) for p_ in (a.control_vector or [])] + [(p_, float(sc)) for p_, sc in (a.control_vector_scaled or [])]
    if cvecs:
        if len(cvecs) > 1 or not hasattr(eng, "set_control_vector"):
            sys.exit("dlp-cli: one control vector at a time, on a single-process engine (merge several with tools/dlp_cvector.py)")
        rng = tuple(int(x) for x in a.control_vector_layer_range) if a.control_vector_layer_range else None
        eng.set_control_vector(cvecs[0][0], cvecs[0][1], rng)
        log.info(f"control vector {cvecs[0][0]} x {cvecs[0][1]}" + (f" on layers {rng[0]}..{rng[1]}" if rng else ""))
    if a.verbose_prompt:
        ids = tok.encode(prompt)
        log.info("prompt tokens: " + " ".join(f"{i}:{tok.tokens[i]!r}" for i in ids))
    if a.display_prompt:
        sys.stdout.write(prompt)
    if (a.model_draft or a.lookup) and not a.rpc and sched is None:
        # draft-and-verify generation (llama-speculative / llama-lookup): greedy, single sequence
        from distributed_llm_pipeline_b200.serving.speculative import ModelDrafter, NgramDrafter, SpecStats, generate_speculative
        if a.model_draft:
            dw = load_weights(a.model_draft, dev)
            drafter = ModelDrafter(LlamaEngine(dw, n_ctx=a.ctx_size, max_seqs=1, page_size=eng.page_size, max_batch_tokens=eng.max_tokens))
        else:
            drafter = NgramDrafter()
        st = SpecStats()
        t1 = time.time()
        ids = generate_speculative(eng, tok.encode(prompt), a.predict, drafter, a.draft_max, () if a.ignore_eos else tuple(tok.stop_ids), st)
        sys.stdout.write(tok.decode(ids) + "\n")
        log.info(f"speculative: {st.n_generated} tokens in {st.n_target_passes} target passes ({st.tokens_per_pass:.2f} tok/pass), "
                 f"drafted {st.n_drafted}, accepted {st.n_accepted} ({100 * st.acceptance:.1f}%), {st.n_generated / (time.time() - t1):.1f} tok/s")
        log.flush()
        return
    if sched is None:
        sched = Scheduler(eng, tok, n_batch=a.batch_size)
    sched.context_shift = a.context_shift

    def on_event(kind, content):
        if kind == "token":
            sys.stdout.write(content)
            sys.stdout.flush()
        elif kind in ("log", "error"):
            log.info(content.rstrip("\n"))

    history, last_prompt, last_reply = "", prompt, ""
    while True:
        req = sched.generate(prompt, n_predict=a.predict, sampling=sampling_from_args(a), ignore_eos=a.ignore_eos, on_event=on_event, n_keep=a.keep,
                             stop=list(a.reverse_prompt or []))   # (--lookup / -md take the stand-alone speculative path above)
        sys.stdout.write("\n")
        if not history:
            history = prompt
        last_reply = req.text + "\n"
        t = req.timings()
        log.info(f"prompt eval: {t['prompt_n']} tokens, {t['prompt_per_second']} tok/s | eval: {t['predicted_n']} tokens, {t['predicted_per_second']} tok/s")
        if a.single_turn or not sys.stdin.isatty():
            break
        try:
            prompt = input("> ")
        except EOFError:
            break
        cmd = prompt.strip()
        if cmd in ("/exit", "/quit"):
            break
        if cmd == "/clear":                      # forget the conversation (and its cached prefix)
            history = ""
            sys.stderr.write("(conversation cleared)\n")
            continue
        if cmd == "/regen":                      # sample the last reply again
            prompt = last_prompt
        elif cmd.startswith("/read "):           # use a file's contents as the next message
            try:
                prompt = open(cmd[6:].strip()).read()
            except OSError as e:
                sys.stderr.write(f"cannot read file: {e}\n")
                continue
        if chat_mode:
            if cmd == "/regen":
                messages = messages[:-1] if messages and messages[-1]["role"] == "assistant" else messages
            else:
                messages += [{"role": "assistant", "content": req.text}, {"role": "user", "content": a.in_prefix + prompt + a.in_suffix}]
            prompt = apply_chat_template(chat_meta, messages)
        elif cmd != "/regen":
            history = history + last_reply       # multi-turn: the reply becomes part of the next prompt (prefix stays cached)
            prompt = history + a.in_prefix + prompt + a.in_suffix
        last_prompt = prompt
    log.flush()
    if pp * tp > 1 and not a.rpc:
        stop_service(sched)


if __name__ == "__main__":
    main()
