"""OpenAI-compatible model server: the container contract of the reference's vLLM chart.

    python -m enterprise-inference_b200.server --model <hf id> --served-model-name <name> --port 2080 \
        --block-size 128 --dtype bfloat16 --max-model-len 8192 --tensor-parallel-size 1 ...

What the chart passes (must accept; may no-op): /root/reference/core/helm-charts/vllm/xeon-values.yaml:70-89
and templates/deployment.yaml:66-87.  HTTP surface: /root/reference/docs/api-spec.yaml (POST
/v1/completions, /v1/chat/completions; 422 on a bad body, additionalProperties: false), GET /health
(values.yaml:135-141), GET /metrics with the vllm:* series the Grafana dashboard queries
(observability/xeon-exporter/grafana-xeon-dashboard.json), GET /v1/models (docs/getting-started-example.md:79-81).
"temperature": 0 (what the reference's clients send) is the device's argmax; temperature / top_p / top_k /
min_p / seed sample on the device (csrc/sample.cu) with vLLM's semantics and defaults (the model's generation_config.json,
else temperature 1), `logprobs` / `top_logprobs` come from a log-softmax kernel over the same logits; frequency_penalty,
presence_penalty, repetition_penalty, logit_bias and guided decoding are NOT implemented: accepted and ignored
or rejected (best_of != n, beam search); n parallel samples run as n engine requests.  /v1/embeddings is served by TEI in the reference, not here.
"""
from __future__ import annotations

import argparse
import asyncio
import contextlib
import json
import os
import queue
import signal
import threading
import time
import uuid
from typing import Any, Dict, List, Optional, Union

from fastapi import FastAPI, HTTPException, Request as HttpRequest
from fastapi.responses import JSONResponse, PlainTextResponse, StreamingResponse
from prometheus_client import CollectorRegistry, Counter, Gauge, Histogram, generate_latest
from pydantic import BaseModel, ConfigDict, Field

from .engine import LLMEngine, Request


# ------------------------------------------------------------------------------------------------
# CLI: every flag the chart emits
# ------------------------------------------------------------------------------------------------
def make_arg_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser("enterprise-inference_b200.server", allow_abbrev=False)
    ap.add_argument("--model", required=True)
    ap.add_argument("--served-model-name", "--served_model_name", dest="served_model_name", default=None)
    ap.add_argument("--port", type=int, default=2080)
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--block-size", "--block_size", dest="block_size", type=int, default=128)
    ap.add_argument("--dtype", default="bfloat16")
    ap.add_argument("--max-model-len", "--max_model_len", dest="max_model_len", type=int, default=8192)
    ap.add_argument("--max-num-batched-tokens", "--max_num_batched_tokens", dest="max_num_batched_tokens", type=int, default=2048)
    ap.add_argument("--max-num-seqs", "--max_num_seqs", dest="max_num_seqs", type=int, default=256)
    ap.add_argument("--tensor-parallel-size", "--tensor_parallel_size", dest="tensor_parallel_size", type=int, default=1)
    ap.add_argument("--pipeline-parallel-size", "--pipeline_parallel_size", dest="pipeline_parallel_size", type=int, default=1)
    ap.add_argument("--distributed_executor_backend", "--distributed-executor-backend", dest="distributed_executor_backend", default="mp")
    ap.add_argument("--enable_chunked_prefill", "--enable-chunked-prefill", dest="enable_chunked_prefill", action="store_true")
    ap.add_argument("--enforce-eager", "--enforce_eager", dest="enforce_eager", action="store_true")
    ap.add_argument("--disable-log-requests", "--disable_log_requests", dest="disable_log_requests", action="store_true")
    ap.add_argument("--enable-auto-tool-choice", dest="enable_auto_tool_choice", action="store_true")
    ap.add_argument("--tool-call-parser", dest="tool_call_parser", default=None)
    ap.add_argument("--chat-template", dest="chat_template", default=None)
    # Gaudi-variant flags of the same chart (gaudi-values.yaml:160): accepted, ignored
    ap.add_argument("--gpu-memory-util", "--gpu-memory-utilization", dest="gpu_memory_utilization", type=float, default=0.9)
    ap.add_argument("--max-num-prefill-seqs", dest="max_num_prefill_seqs", type=int, default=None)
    ap.add_argument("--num_scheduler_steps", "--num-scheduler-steps", dest="num_scheduler_steps", type=int, default=1)
    ap.add_argument("--use-padding-aware-scheduling", dest="use_padding_aware_scheduling", action="store_true")
    # ours
    ap.add_argument("--kv-cache-gib", type=float, default=None, help="KV cache size (default: half of free HBM)")
    ap.add_argument("--random-weights", action="store_true", help="random-init weights of the model's shape (no checkpoint)")
    ap.add_argument("--tokenizer", default="auto", help="'auto' (HF tokenizer from the model dir) or 'byte'")
    return ap


def validate_args(args) -> None:
    if args.dtype not in ("bfloat16", "bf16", "auto"):
        raise SystemExit(f"--dtype {args.dtype}: this backend computes in bfloat16 only")
    if args.pipeline_parallel_size != 1:
        raise SystemExit("--pipeline-parallel-size > 1 is not supported (the reference never sets it, values.yaml:188)")
    if args.block_size % 64:
        raise SystemExit("--block-size must be a multiple of 64 (the chart passes 128)")
    if args.tensor_parallel_size not in (1, 2, 4, 8):
        raise SystemExit("--tensor-parallel-size must be 1, 2, 4 or 8")


# ------------------------------------------------------------------------------------------------
# tokenizers
# ------------------------------------------------------------------------------------------------
class ByteTokenizer:
    """Ids are UTF-8 bytes (+3 specials); for random-weight smoke runs and tests."""
    bos_token_id, eos_token_id, pad_token_id = 256, 257, 258

    def __init__(self, vocab: int):
        self.vocab = vocab

    def encode(self, text: str, add_special_tokens: bool = True) -> List[int]:
        ids = [b % self.vocab for b in text.encode("utf-8")]
        return ([self.bos_token_id % self.vocab] if add_special_tokens else []) + ids

    def decode(self, ids: List[int], skip_special_tokens: bool = True) -> str:
        return bytes(i for i in ids if i < 256).decode("utf-8", errors="replace")

    def apply_chat_template(self, messages, add_generation_prompt=True, **kw) -> str:
        out = "".join(f"<|{m['role']}|>\n{_content_text(m.get('content'))}\n" for m in messages)
        return out + ("<|assistant|>\n" if add_generation_prompt else "")


def _content_text(c) -> str:
    if c is None:
        return ""
    if isinstance(c, str):
        return c
    return "".join(p.get("text", "") for p in c if isinstance(p, dict))


def load_tokenizer(kind: str, model_path: Optional[str], vocab: int):
    if kind == "byte" or model_path is None:
        return ByteTokenizer(vocab)
    from transformers import AutoTokenizer
    return AutoTokenizer.from_pretrained(model_path, local_files_only=True)


# ------------------------------------------------------------------------------------------------
# request / response schema (docs/api-spec.yaml; extra fields are rejected with 422)
# ------------------------------------------------------------------------------------------------
class TokenizeRequest(BaseModel):
    model_config = ConfigDict(extra="ignore")
    model: Optional[str] = None
    prompt: str
    add_special_tokens: Optional[bool] = True


class DetokenizeRequest(BaseModel):
    model_config = ConfigDict(extra="ignore")
    model: Optional[str] = None
    tokens: List[int]


class StreamOptions(BaseModel):
    model_config = ConfigDict(extra="forbid")
    include_usage: Optional[bool] = True
    continuous_usage_stats: Optional[bool] = False


class _Sampling(BaseModel):
    model_config = ConfigDict(extra="forbid")
    model: str
    max_tokens: Optional[int] = 16
    temperature: Optional[float] = None      # None: the model's generation_config.json, else 1.0 (vLLM's default)
    top_p: Optional[float] = None
    top_k: Optional[int] = None
    min_p: Optional[float] = 0.0
    n: Optional[int] = 1
    best_of: Optional[int] = None
    seed: Optional[int] = None
    stop: Optional[Union[str, List[str]]] = None
    stream: Optional[bool] = False
    stream_options: Optional[StreamOptions] = None
    logprobs: Optional[Union[int, bool]] = None
    logit_bias: Optional[Dict[str, float]] = None
    frequency_penalty: Optional[float] = 0.0
    presence_penalty: Optional[float] = 0.0
    repetition_penalty: Optional[float] = 1.0
    length_penalty: Optional[float] = 1.0
    use_beam_search: Optional[bool] = False
    early_stopping: Optional[bool] = False
    user: Optional[str] = None
    stop_token_ids: Optional[List[int]] = None
    include_stop_str_in_output: Optional[bool] = False
    ignore_eos: Optional[bool] = False
    min_tokens: Optional[int] = 0
    skip_special_tokens: Optional[bool] = True
    spaces_between_special_tokens: Optional[bool] = True
    truncate_prompt_tokens: Optional[int] = None
    add_special_tokens: Optional[bool] = None
    echo: Optional[bool] = False
    response_format: Optional[Dict[str, Any]] = None
    guided_json: Optional[Any] = None
    guided_regex: Optional[str] = None
    guided_choice: Optional[List[str]] = None
    guided_grammar: Optional[str] = None
    guided_decoding_backend: Optional[str] = None
    guided_whitespace_pattern: Optional[str] = None


class CompletionRequest(_Sampling):
    prompt: Union[str, List[str], List[int], List[List[int]]]
    suffix: Optional[str] = None
    allowed_token_ids: Optional[List[int]] = None


class ChatCompletionRequest(_Sampling):
    messages: List[Dict[str, Any]] = Field(min_length=1)
    top_logprobs: Optional[int] = None
    tools: Optional[List[Dict[str, Any]]] = None
    tool_choice: Optional[Union[str, Dict[str, Any]]] = "auto"
    add_generation_prompt: Optional[bool] = True
    documents: Optional[List[Dict[str, str]]] = None
    chat_template: Optional[str] = None
    chat_template_kwargs: Optional[Dict[str, Any]] = None


# ------------------------------------------------------------------------------------------------
# metrics: the series the reference's Grafana dashboard and ServiceMonitor expect
# ------------------------------------------------------------------------------------------------
class Metrics:
    def __init__(self, model_name: str):
        self.registry = CollectorRegistry()
        L = ["model_name"]
        buckets_t = (0.001, 0.005, 0.01, 0.02, 0.04, 0.06, 0.08, 0.1, 0.25, 0.5, 0.75, 1.0, 2.5, 5.0, 7.5, 10.0)
        buckets_l = (0.3, 0.5, 0.8, 1.0, 1.5, 2.0, 2.5, 5.0, 10.0, 15.0, 20.0, 30.0, 40.0, 50.0, 60.0)
        buckets_n = (1, 2, 5, 10, 20, 50, 100, 200, 500, 1000, 2000, 5000, 10000)
        r = self.registry
        self.e2e = Histogram("vllm:e2e_request_latency_seconds", "request latency", L, buckets=buckets_l, registry=r)
        self.ttft = Histogram("vllm:time_to_first_token_seconds", "TTFT", L, buckets=buckets_t, registry=r)
        self.tpot = Histogram("vllm:time_per_output_token_seconds", "TPOT", L, buckets=buckets_t, registry=r)
        self.prompt_tokens = Counter("vllm:prompt_tokens", "prompt tokens", L, registry=r)
        self.generation_tokens = Counter("vllm:generation_tokens", "generated tokens", L, registry=r)
        self.req_prompt = Histogram("vllm:request_prompt_tokens", "prompt tokens per request", L, buckets=buckets_n, registry=r)
        self.req_gen = Histogram("vllm:request_generation_tokens", "generated tokens per request", L, buckets=buckets_n, registry=r)
        self.success = Counter("vllm:request_success", "finished requests", L + ["finished_reason"], registry=r)
        self.running = Gauge("vllm:num_requests_running", "running", L, registry=r)
        self.waiting = Gauge("vllm:num_requests_waiting", "waiting", L, registry=r)
        self.swapped = Gauge("vllm:num_requests_swapped", "swapped", L, registry=r)
        self.gpu_cache = Gauge("vllm:gpu_cache_usage_perc", "KV cache usage", L, registry=r)
        self.cpu_cache = Gauge("vllm:cpu_cache_usage_perc", "CPU KV cache usage (always 0)", L, registry=r)
        self.m = model_name
        for g in (self.running, self.waiting, self.swapped, self.gpu_cache, self.cpu_cache):
            g.labels(self.m).set(0)

    def finished(self, req: Request):
        m = self.m
        n_out = len(req.output_ids)
        self.e2e.labels(m).observe(req.finish_time - req.arrival)
        if req.first_token_time is not None:
            self.ttft.labels(m).observe(req.first_token_time - req.arrival)
        for a, b in zip(req.token_times, req.token_times[1:]):
            self.tpot.labels(m).observe(b - a)
        self.prompt_tokens.labels(m).inc(len(req.prompt_ids))
        self.generation_tokens.labels(m).inc(n_out)
        self.req_prompt.labels(m).observe(len(req.prompt_ids))
        self.req_gen.labels(m).observe(n_out)
        self.success.labels(m, req.finish_reason or "stop").inc()


# ------------------------------------------------------------------------------------------------
# engine thread: the single thread that drives the ctx (the C-ABI is not re-entrant per ctx)
# ------------------------------------------------------------------------------------------------
class EngineLoop:
    def __init__(self, engine: LLMEngine, metrics: Metrics):
        self.engine, self.metrics = engine, metrics
        self.inbox: "queue.Queue" = queue.Queue()
        self.listeners: Dict[str, Any] = {}      # request_id -> (loop, asyncio.Queue)
        self._sent: Dict[str, int] = {}          # request_id -> output tokens already emitted
        self.alive = True
        self.error: Optional[str] = None
        self.draining = False
        self.thread = threading.Thread(target=self._run, name="b200-engine", daemon=True)

    def start(self):
        self.thread.start()

    def submit(self, req: Request, loop, q: "asyncio.Queue"):
        if self.draining:
            raise HTTPException(503, "server is draining")
        self.listeners[req.request_id] = (loop, q)
        self.inbox.put(req)

    def abort(self, request_id: str):
        """From any thread: the client disconnected, the SSE generator was cancelled, or a stop string was hit.
        The engine thread drops the request, returns its KV blocks and sequence slot, and counts it as aborted."""
        self.inbox.put(("abort", request_id))

    def _emit(self, rid: str, item):
        tgt = self.listeners.get(rid)
        if tgt:
            loop, q = tgt
            loop.call_soon_threadsafe(q.put_nowait, item)

    def _run(self):
        eng = self.engine
        try:
            while self.alive:
                try:
                    while True:
                        req = self.inbox.get(block=not eng.has_work(), timeout=0.05)
                        if isinstance(req, tuple):          # ("abort", request_id)
                            gone = eng.abort_request(req[1])
                            self._sent.pop(req[1], None)
                            if gone is not None:
                                self.metrics.finished(gone)
                                self._emit(gone.request_id, ("done", "abort"))
                            self.listeners.pop(req[1], None)
                            continue
                        try:
                            eng.add_request(req)
                        except (ValueError, OverflowError) as e:
                            self._emit(req.request_id, ("error", str(e)))
                            self.listeners.pop(req.request_id, None)
                except queue.Empty:
                    pass
                if not eng.has_work():
                    continue
                done = eng.step()
                m = self.metrics
                m.running.labels(m.m).set(len(eng.scheduler.running))
                m.waiting.labels(m.m).set(len(eng.scheduler.waiting))
                m.gpu_cache.labels(m.m).set(eng.blocks.usage)
                # tokens not yet handed to the listener (a preempted request can receive the token of the step that was
                # in flight while it sits in `waiting`, so the count is kept per request, not per step)
                preempted = [r for r in eng.scheduler.waiting if r.output_ids]
                for r in list(eng.scheduler.running) + preempted + done:
                    n0 = self._sent.get(r.request_id, 0)
                    if len(r.output_ids) > n0:
                        self._emit(r.request_id, ("tokens", r.output_ids[n0:]))
                        self._sent[r.request_id] = len(r.output_ids)
                for r in done:
                    self._sent.pop(r.request_id, None)
                    m.finished(r)
                    if r.error is not None:             # the scheduler gave up on this one request only
                        self._emit(r.request_id, ("error", r.error))
                    else:
                        self._emit(r.request_id, ("done", r.finish_reason))
                    self.listeners.pop(r.request_id, None)
        except Exception as e:  # the engine died: /health goes non-200 and the pod is restarted
            self.error = f"{type(e).__name__}: {e}"
            for rid in list(self.listeners):
                self._emit(rid, ("error", self.error))

    def healthy(self) -> bool:
        return self.error is None and self.thread.is_alive()


# ------------------------------------------------------------------------------------------------
# app
# ------------------------------------------------------------------------------------------------
def create_app(engine: LLMEngine, tokenizer, served_model_name: str, max_model_len: int, tool_parser=None,
               default_sampling: Optional[Dict[str, Any]] = None) -> FastAPI:
    metrics = Metrics(served_model_name)
    loop_ = EngineLoop(engine, metrics)

    @contextlib.asynccontextmanager
    async def lifespan(_app):
        loop_.start()
        yield
        loop_.draining = True       # SIGTERM: finish what is running (terminationGracePeriodSeconds: 120)
        t0 = time.monotonic()
        while engine.has_work() and loop_.healthy() and time.monotonic() - t0 < 120:
            await asyncio.sleep(0.05)
        loop_.alive = False

    app = FastAPI(title="enterprise-inference b200 decode backend", lifespan=lifespan)
    app.state.engine_loop = loop_
    app.state.metrics = metrics
    eos = getattr(tokenizer, "eos_token_id", None)
    eos_ids = tuple([eos] if isinstance(eos, int) else (eos or []))

    @app.get("/health")
    async def health():
        if not loop_.healthy():
            return PlainTextResponse(loop_.error or "engine thread stopped", status_code=503)
        return PlainTextResponse("", status_code=200)

    # the small utility routes of the image's server (health probes of other stacks, tokenizer access)
    @app.get("/ping")
    @app.post("/ping")
    async def ping():
        return await health()

    @app.get("/version")
    async def version():
        return {"version": "b200-decode 0.1 (vLLM-compatible API surface of the v0.10.2 image)"}

    @app.post("/tokenize")
    async def tokenize(body: TokenizeRequest):
        ids = tokenizer.encode(body.prompt, add_special_tokens=body.add_special_tokens is not False)
        return {"count": len(ids), "max_model_len": max_model_len, "tokens": ids}

    @app.post("/detokenize")
    async def detokenize(body: DetokenizeRequest):
        return {"prompt": tokenizer.decode(list(body.tokens), skip_special_tokens=False)}

    @app.get("/metrics")
    async def metrics_ep():
        return PlainTextResponse(generate_latest(metrics.registry).decode(), media_type="text/plain; version=0.0.4")

    @app.get("/v1/models")
    async def models():
        return {"object": "list", "data": [{"id": served_model_name, "object": "model", "created": int(time.time()),
                                            "owned_by": "enterprise-inference-b200", "root": served_model_name,
                                            "max_model_len": max_model_len}]}

    def error(status: int, msg: str, typ: str = "BadRequestError"):
        return JSONResponse({"object": "error", "message": msg, "type": typ, "param": None, "code": status}, status_code=status)

    def check_model(name: str):
        if name != served_model_name:
            return error(404, f"The model `{name}` does not exist.", "NotFoundError")
        return None

    def unsupported(body: _Sampling) -> Optional[str]:
        """Fields of docs/api-spec.yaml this backend parses but does not implement: a non-default value gets a 400,
        never a 200 whose output silently ignores it."""
        bad = []
        if body.length_penalty not in (None, 1.0):
            bad.append("length_penalty")
        if getattr(body, "allowed_token_ids", None):
            bad.append("allowed_token_ids")
        if getattr(body, "suffix", None):
            bad.append("suffix")
        for f in ("guided_json", "guided_regex", "guided_choice", "guided_grammar"):
            if getattr(body, f) is not None:
                bad.append(f)
        if body.response_format is not None and body.response_format.get("type", "text") != "text":
            bad.append("response_format")
        if bad:
            return ("not supported by this backend: " + ", ".join(bad) +
                    " (guided decoding, allowed_token_ids, suffix and length_penalty are not implemented; leave them unset)")
        for name, lo, hi in (("frequency_penalty", -2.0, 2.0), ("presence_penalty", -2.0, 2.0)):
            v = getattr(body, name)
            if v is not None and not lo <= v <= hi:
                return f"{name} must be in [{lo}, {hi}]"
        if body.repetition_penalty is not None and not 0.0 < body.repetition_penalty <= 100.0:
            return "repetition_penalty must be in (0, 100]"
        if body.logit_bias:
            if len(body.logit_bias) > 1024:
                return "logit_bias accepts at most 1024 entries"
            vocab = engine.runner.cfg.vocab
            for k, v in body.logit_bias.items():
                if not str(k).lstrip("-").isdigit() or not 0 <= int(k) < vocab:
                    return f"logit_bias key {k!r} is not a token id of this model (0..{vocab - 1})"
                if not -100.0 <= float(v) <= 100.0:
                    return "logit_bias values must be in [-100, 100]"
        adjusts = bool(body.logit_bias) or bool(body.frequency_penalty) or bool(body.presence_penalty) or \
            body.repetition_penalty not in (None, 1.0)
        if adjusts and body.logprobs not in (None, False):
            return ("logprobs together with logit_bias or a penalty is not supported (the image reports log-probabilities of "
                    "the raw logits; this backend adjusts the logits in place)")
        if body.seed is not None and not -2 ** 63 <= body.seed < 2 ** 64:
            return "seed must fit 64 bits"
        if body.top_k is not None and not -1 <= body.top_k < 2 ** 31:
            return "top_k must be -1 (off), 0 (off) or a positive 32-bit integer"
        if body.max_tokens is not None and body.max_tokens < 1:
            return "max_tokens must be at least 1"
        return None

    def make_request(body: _Sampling, prompt_ids: List[int]) -> Request:
        if body.truncate_prompt_tokens:
            prompt_ids = prompt_ids[-body.truncate_prompt_tokens:]
        max_tokens = body.max_tokens if body.max_tokens is not None else max(max_model_len - len(prompt_ids), 1)
        d = default_sampling or {}
        temperature = body.temperature if body.temperature is not None else float(d.get("temperature", 1.0))
        top_p = body.top_p if body.top_p is not None else float(d.get("top_p", 1.0))
        top_k = body.top_k if body.top_k is not None else int(d.get("top_k", 0))
        return Request(request_id="cmpl-" + uuid.uuid4().hex, prompt_ids=prompt_ids, max_tokens=max_tokens,
                       ignore_eos=bool(body.ignore_eos), eos_token_ids=eos_ids,
                       stop_token_ids=tuple(body.stop_token_ids or ()), min_tokens=body.min_tokens or 0,
                       temperature=max(float(temperature), 0.0), top_k=max(int(top_k), 0),
                       top_p=min(max(float(top_p), 0.0), 1.0), min_p=min(max(float(body.min_p or 0.0), 0.0), 1.0),
                       seed=body.seed, logprobs=want_logprobs(body),
                       repetition_penalty=float(body.repetition_penalty if body.repetition_penalty is not None else 1.0),
                       frequency_penalty=float(body.frequency_penalty or 0.0), presence_penalty=float(body.presence_penalty or 0.0),
                       logit_bias={int(k): float(v) for k, v in body.logit_bias.items()} if body.logit_bias else None)

    def want_logprobs(body: _Sampling) -> Optional[int]:
        """None: off.  Completions: `logprobs: N`; chat: `logprobs: true` + `top_logprobs: N`."""
        lp = body.logprobs
        if lp is None or lp is False:
            return None
        if lp is True:
            return int(getattr(body, "top_logprobs", None) or 0)
        return int(lp)

    def tok_str(t: int) -> str:
        return tokenizer.decode([int(t)], skip_special_tokens=False)

    def format_logprobs(entries, token_ids, chat: bool, offset0: int = 0):
        """OpenAI shapes (docs/api-spec.yaml): completions = parallel arrays, chat = one object per token."""
        if chat:
            content = []
            for tid, (lp, top) in zip(token_ids, entries):
                ts = tok_str(tid)
                content.append({"token": ts, "logprob": lp, "bytes": list(ts.encode("utf-8")),
                                "top_logprobs": [{"token": tok_str(a), "logprob": b, "bytes": list(tok_str(a).encode("utf-8"))}
                                                 for a, b in top]})
            return {"content": content}
        out = {"text_offset": [], "token_logprobs": [], "tokens": [], "top_logprobs": []}
        off = offset0
        for tid, (lp, top) in zip(token_ids, entries):
            ts = tok_str(tid)
            out["text_offset"].append(off)
            off += len(ts)
            out["tokens"].append(ts)
            out["token_logprobs"].append(lp)
            alts = {tok_str(a): b for a, b in top}
            alts.setdefault(ts, lp)      # the chosen token is always listed (vLLM does the same)
            out["top_logprobs"].append(alts)
        return out

    class Detok:
        """Incremental detokenisation + stop strings."""
        def __init__(self, body: _Sampling):
            self.ids: List[int] = []
            self.sent = 0
            self.stops = [body.stop] if isinstance(body.stop, str) else list(body.stop or [])
            self.include_stop = bool(body.include_stop_str_in_output)
            self.skip_special = body.skip_special_tokens is not False
            self.stopped = False

        def push(self, new_ids: List[int]) -> str:
            self.ids += new_ids
            text = tokenizer.decode(self.ids, skip_special_tokens=self.skip_special)
            for s in self.stops:
                k = text.find(s, max(self.sent - len(s), 0))
                if k >= 0:
                    text = text[:k + (len(s) if self.include_stop else 0)]
                    self.stopped = True
            if not self.stopped and text.endswith("�"):     # incomplete UTF-8 sequence: hold it back
                return ""
            delta, self.sent = text[self.sent:], max(len(text), self.sent)
            return delta

        def flush(self) -> str:
            """At the end of the request: whatever was held back as a possibly incomplete UTF-8 sequence."""
            if self.stopped:
                return ""
            text = tokenizer.decode(self.ids, skip_special_tokens=self.skip_special)
            delta, self.sent = text[self.sent:], max(len(text), self.sent)
            return delta

    async def run(body: _Sampling, prompt_ids: List[int], chat: bool, prompt_text: str = ""):
        if len(prompt_ids) == 0:
            return error(400, "the prompt is empty")
        why = unsupported(body)
        if why:
            return error(400, why)
        n = body.n or 1
        if (body.best_of not in (None, n)) or body.use_beam_search:
            return error(400, "best_of != n and beam search are not supported by this backend")
        if not 1 <= n <= 16:
            return error(400, "n must be between 1 and 16")
        nlp = want_logprobs(body)
        if nlp is not None and not 0 <= nlp <= 20:
            return error(400, "logprobs / top_logprobs must be between 0 and 20")
        # n parallel samples = n engine requests over the same prompt (vLLM: child i samples with seed + i)
        reqs = []
        for i in range(n):
            r_ = make_request(body, prompt_ids)
            if body.seed is not None:
                r_.seed = (body.seed + i) & (2 ** 64 - 1)
            reqs.append(r_)
        req0 = reqs[0]
        if len(prompt_ids) + req0.max_tokens > max_model_len:
            return error(400, f"This model's maximum context length is {max_model_len} tokens. However, you requested "
                              f"{len(prompt_ids) + req0.max_tokens} tokens ({len(prompt_ids)} in the prompt, {req0.max_tokens} "
                              "in the completion).")
        queues = [asyncio.Queue() for _ in reqs]
        for r_, q_ in zip(reqs, queues):
            loop_.submit(r_, asyncio.get_running_loop(), q_)
        created = int(time.time())
        rid = req0.request_id if not chat else "chat" + req0.request_id
        obj = "chat.completion" if chat else "text_completion"
        detoks = [Detok(body) for _ in reqs]
        lp_seen = [0] * n

        def pending_logprobs(i: int):
            """Log-probabilities of request i's tokens not yet sent (None when the request did not ask)."""
            r_ = reqs[i]
            if r_.logprobs is None:
                return None
            a, b_ = lp_seen[i], min(len(r_.logprob_data), len(r_.output_ids))   # the engine thread appends both
            lp_seen[i] = b_
            off = sum(len(tok_str(t)) for t in r_.output_ids[:a]) if not chat else 0
            return format_logprobs(r_.logprob_data[a:b_], r_.output_ids[a:b_], chat, off)

        def usage():
            done = sum(len(r_.output_ids) for r_ in reqs)
            return {"prompt_tokens": len(prompt_ids), "completion_tokens": done, "total_tokens": len(prompt_ids) + done}

        async def events(i: int):
            q, detok, req = queues[i], detoks[i], reqs[i]
            while True:
                kind, val = await q.get()
                if kind == "tokens":
                    d = detok.push(val)
                    if detok.stopped:           # stop string hit: the engine drops the request now
                        loop_.abort(req.request_id)
                    yield "delta", d
                    if detok.stopped:
                        yield "done", "stop"
                        return
                elif kind == "done":
                    rest = detok.flush()
                    if rest:
                        yield "delta", rest
                    yield "done", val
                    return
                else:
                    yield "error", val
                    return

        if body.stream:
            async def sse():
                include_usage = bool(body.stream_options and body.stream_options.include_usage)
                cont = bool(body.stream_options and body.stream_options.continuous_usage_stats)
                merged: asyncio.Queue = asyncio.Queue()

                async def pump(i: int):
                    async for kind, val in events(i):
                        await merged.put((i, kind, val))

                tasks = [asyncio.ensure_future(pump(i)) for i in range(n)]
                first = [True] * n
                live = n
                try:
                    while live:
                        i, kind, val = await merged.get()
                        if kind == "error":
                            yield f"data: {json.dumps({'error': {'message': val, 'type': 'InternalServerError', 'code': 500}})}\n\n"
                            break
                        finish = val if kind == "done" else None
                        if kind == "done":
                            live -= 1
                        text = val if kind == "delta" else ""
                        if kind == "delta" and not text and not first[i]:
                            continue
                        if chat:
                            delta = {"content": text}
                            if first[i]:
                                delta["role"] = "assistant"
                            choice = {"index": i, "delta": delta if kind == "delta" else {}, "logprobs": pending_logprobs(i),
                                      "finish_reason": finish}
                            chunk = {"id": rid, "object": "chat.completion.chunk", "created": created,
                                     "model": served_model_name, "choices": [choice]}
                        else:
                            chunk = {"id": rid, "object": "text_completion", "created": created, "model": served_model_name,
                                     "choices": [{"index": i, "text": text, "logprobs": pending_logprobs(i),
                                                  "finish_reason": finish, "stop_reason": None}]}
                        if include_usage:
                            chunk["usage"] = usage() if cont else None
                        first[i] = False
                        yield f"data: {json.dumps(chunk)}\n\n"
                finally:
                    for t in tasks:
                        t.cancel()
                    for r_ in reqs:            # client gone (generator closed / cancelled) or an error: free the engine side
                        if not r_.finished:
                            loop_.abort(r_.request_id)
                if include_usage:
                    yield f"data: {json.dumps({'id': rid, 'object': 'chat.completion.chunk' if chat else 'text_completion', 'created': created, 'model': served_model_name, 'choices': [], 'usage': usage()})}\n\n"
                yield "data: [DONE]\n\n"
            return StreamingResponse(sse(), media_type="text/event-stream")

        choices = []
        try:
            results = []
            for i in range(n):
                text, finish = "", None
                async for kind, val in events(i):
                    if kind == "delta":
                        text += val
                    elif kind == "done":
                        finish = val
                    else:
                        return error(500, val, "InternalServerError")
                results.append((text, finish))
        finally:
            for r_ in reqs:                    # cancelled (client disconnected) or failed: free the engine side
                if not r_.finished:
                    loop_.abort(r_.request_id)
        for i, (text, finish) in enumerate(results):
            if chat:
                content, tool_calls = text, []
                if tool_parser is not None and getattr(body, "tools", None) and body.tool_choice != "none":
                    content, tool_calls = tool_parser(text, body.tools)
                    if tool_calls:
                        finish = "tool_calls"
                choices.append({"index": i, "message": {"role": "assistant", "content": content, "tool_calls": tool_calls},
                                "logprobs": pending_logprobs(i), "finish_reason": finish, "stop_reason": None})
            else:
                choices.append({"index": i, "text": (prompt_text + text) if body.echo else text, "logprobs": pending_logprobs(i),
                                "finish_reason": finish, "stop_reason": None})
        return {"id": rid, "object": obj, "created": created, "model": served_model_name, "choices": choices, "usage": usage()}

    @app.post("/v1/completions")
    async def completions(body: CompletionRequest):
        bad = check_model(body.model)
        if bad:
            return bad
        p = body.prompt
        if isinstance(p, list) and p and isinstance(p[0], (list, str)):
            if len(p) != 1:
                return error(400, "batched prompts are not supported: send one request per prompt")
            p = p[0]
        if isinstance(p, str):
            ids = tokenizer.encode(p) if body.add_special_tokens is not False else tokenizer.encode(p, add_special_tokens=False)
            text = p
        else:
            ids, text = list(p), ""
            if any((not isinstance(i, int)) or i < 0 or i >= engine.runner.cfg.vocab for i in ids):
                return error(400, "prompt token ids out of range")
        return await run(body, list(ids), chat=False, prompt_text=text)

    @app.post("/v1/chat/completions")
    async def chat_completions(body: ChatCompletionRequest):
        bad = check_model(body.model)
        if bad:
            return bad
        for m in body.messages:
            if "role" not in m:
                return error(400, "every message needs a role")
        try:
            kw = dict(body.chat_template_kwargs or {})
            if body.tools:
                kw["tools"] = body.tools
            if body.chat_template:
                kw["chat_template"] = body.chat_template
            prompt = tokenizer.apply_chat_template(body.messages, add_generation_prompt=bool(body.add_generation_prompt),
                                                   tokenize=False, **kw)
        except Exception as e:
            return error(400, f"chat template failed: {e}")
        ids = tokenizer.encode(prompt, add_special_tokens=False)
        return await run(body, list(ids), chat=True)

    return app


class _RunnerFactory:
    """Picklable recipe for one rank's ModelRunner (used in this process and in spawned TP workers)."""

    def __init__(self, args):
        self.args = args

    def resolve(self):
        from .config import from_hf_config, get_config
        from .weights import resolve_hf_snapshot
        a = self.args
        try:
            path = resolve_hf_snapshot(a.model)
            return path, from_hf_config(path)
        except FileNotFoundError:
            if not a.random_weights:
                raise
            return None, get_config(a.model)

    def __call__(self, rank: int, world: int):
        import torch
        from .runner import ModelRunner
        from .weights import load_safetensors_dir
        a = self.args
        path, cfg = self.resolve()
        torch.cuda.set_device(rank)
        runner = ModelRunner(cfg, max_tokens=a.max_num_batched_tokens, max_model_len=a.max_model_len,
                             block_size=a.block_size, kv_cache_gib=a.kv_cache_gib, n_blocks=getattr(a, "n_blocks", None),
                             tp_rank=rank, tp_world=world, graphs=True)
        if a.random_weights or path is None:
            runner.init_random_weights()
        else:
            runner.load_weights(load_safetensors_dir(path))
        return runner


def build_engine(args):
    """Runner(s) + engine from the chart's flags.  tensor_parallel_size > 1: this process is rank 0 and
    spawns one worker process per further GPU (the chart's `--distributed_executor_backend mp`)."""
    factory = _RunnerFactory(args)
    path, cfg = factory.resolve()
    world = args.tensor_parallel_size
    workers = []
    if world > 1:
        from .tp_exec import launch_workers
        runner, workers = launch_workers(world, factory)
    else:
        runner = factory(0, 1)
    engine = LLMEngine(runner, max_num_batched_tokens=args.max_num_batched_tokens, max_num_seqs=args.max_num_seqs,
                       max_model_len=args.max_model_len)
    tok = load_tokenizer(args.tokenizer if path else "byte", path, cfg.vocab)
    return engine, tok, workers


def load_default_sampling(model_path: Optional[str]) -> Dict[str, Any]:
    """Default sampling parameters as vLLM takes them (`--generation-config auto`): the model's
    generation_config.json next to its weights, when there is one."""
    if not model_path:
        return {}
    f = os.path.join(model_path, "generation_config.json")
    if not os.path.isfile(f):
        return {}
    try:
        with open(f) as fh:
            g = json.load(fh)
    except (OSError, ValueError):
        return {}
    out = {k: g[k] for k in ("temperature", "top_p", "top_k") if isinstance(g.get(k), (int, float))}
    if g.get("do_sample") is False:
        out["temperature"] = 0.0
    return out


def main(argv=None):
    args, unknown = make_arg_parser().parse_known_args(argv)
    if unknown:
        print(f"[b200] ignoring flags this backend does not know: {unknown}", flush=True)
    validate_args(args)
    import uvicorn
    engine, tok, workers = build_engine(args)
    from .tool_parsers import get_parser
    parser = get_parser(args.tool_call_parser) if args.enable_auto_tool_choice else None
    app = create_app(engine, tok, args.served_model_name or args.model, args.max_model_len, tool_parser=parser,
                     default_sampling=load_default_sampling(_RunnerFactory(args).resolve()[0]))
    signal.signal(signal.SIGTERM, lambda *_: (_ for _ in ()).throw(KeyboardInterrupt()))
    try:
        uvicorn.run(app, host=args.host, port=args.port, log_level="warning" if args.disable_log_requests else "info")
    finally:
        if workers:
            from .tp_exec import shutdown
            shutdown(engine.runner, workers)


if __name__ == "__main__":
    main()
