"""CPU: the OpenAI surface and the chart's container contract, with a fake runner behind the engine.
Shapes of the reference's smoke checks: temperature-0 completions return 200 with
choices[0].text; /health 200; /metrics exposes the vllm:* series; bad bodies give 422."""
import json
import types

import numpy as np
import pytest
from fastapi.testclient import TestClient

from conftest import pkg

S = pkg("server")
E = pkg("engine")


class FakeRunner:
    """next token = (sum of context) % 251 — deterministic, depends on the whole sequence."""
    def __init__(self):
        self.n_blocks, self.block_size, self.max_tokens, self.max_model_len = 64, 16, 64, 256
        self.cfg = types.SimpleNamespace(vocab=512)
        self.kv = {}

    def step(self, b):
        for i in range(b.n_tokens):
            self.kv[int(b.slots[i])] = int(b.ids[i])
        out = []
        for r in b.logit_rows:
            bt = b.block_tables[int(b.row_to_seq[r])]
            out.append(sum(self.kv[int(bt[p // 16]) * 16 + p % 16] for p in range(int(b.ctx_lens[r]))) % 251)
        self.last = np.array(out, dtype=np.int32)
        return self.last

    def sample_last(self, temperature, top_k, top_p, seeds, draws, min_p=None):
        """Records the call; a 'sampled' token = greedy token + 1 for rows with temperature > 0."""
        self.sample_calls = getattr(self, "sample_calls", []) + [dict(temperature=list(temperature), top_k=list(top_k),
                                                                         top_p=list(top_p), seeds=list(seeds), draws=list(draws),
                                                                         min_p=list(min_p) if min_p is not None else None)]
        return np.array([t + 1 if temperature[i] > 0 else t for i, t in enumerate(self.last)], dtype=np.int32)


    def adjust_last(self, rows):
        """Records the rows the engine wants adjusted (None = untouched)."""
        snap = [None if r is None else dict(r, prompt=list(r["prompt"]), output=list(r["output"])) for r in rows]
        self.adjust_calls = getattr(self, "adjust_calls", []) + [snap]

    def logprobs_last(self, token_ids, n_top=0):
        """logprob of the chosen token = -0.25; alternatives = the next token ids with -1, -2, ..."""
        n = len(token_ids)
        lp = np.full(n, -0.25, dtype=np.float32)
        ids = np.array([[(int(t) + j) % 251 for j in range(n_top)] for t in token_ids], dtype=np.int32).reshape(n, n_top)
        tlp = np.array([[-0.25 - j for j in range(n_top)] for _ in token_ids], dtype=np.float32).reshape(n, n_top)
        return lp, ids, tlp


class OverlapRunner(FakeRunner):
    """The same model behind the runner's overlap interface (step_async / step_result): ids of -(j+1) are resolved from the
    tokens of the step enqueued just before, as the device does.  LLMEngine then runs in overlap mode."""
    def step_async(self, b):
        prev = getattr(self, "last", np.zeros(0, dtype=np.int32))
        b.ids = [int(prev[-int(t) - 1]) if int(t) < 0 else int(t) for t in b.ids]
        self.n_async = getattr(self, "n_async", 0) + 1
        return FakeRunner.step(self, b).copy()

    def step_result(self, handle):
        return handle


def expect_ids(prompt_ids, n):
    ids = list(prompt_ids)
    for _ in range(n):
        ids.append(sum(ids) % 251)
    return ids[len(prompt_ids):]


@pytest.fixture(params=["synchronous", "overlap"])
def client(request):
    """Every test of the OpenAI surface runs against the synchronous engine and against the overlapped one."""
    eng = E.LLMEngine(FakeRunner() if request.param == "synchronous" else OverlapRunner(), max_num_batched_tokens=64,
                      max_num_seqs=8, max_model_len=256)
    assert eng.overlap == (request.param == "overlap")
    tok = S.ByteTokenizer(512)
    app = S.create_app(eng, tok, "meta-llama/Llama-3.1-8B-Instruct", 256)
    with TestClient(app) as c:
        c.tok = tok
        yield c


MODEL = "meta-llama/Llama-3.1-8B-Instruct"


def test_health_models_metrics(client):
    assert client.get("/health").status_code == 200
    m = client.get("/v1/models").json()
    assert m["data"][0]["id"] == MODEL
    r = client.post("/v1/completions", json={"model": MODEL, "prompt": "What is Deep Learning?", "max_tokens": 5, "temperature": 0})
    assert r.status_code == 200
    text = client.get("/metrics").text
    for series in ("vllm:e2e_request_latency_seconds_bucket", "vllm:prompt_tokens_total", "vllm:generation_tokens_total",
                   "vllm:time_per_output_token_seconds_bucket", "vllm:time_to_first_token_seconds_bucket",
                   "vllm:num_requests_running", "vllm:num_requests_swapped", "vllm:num_requests_waiting",
                   "vllm:gpu_cache_usage_perc", "vllm:cpu_cache_usage_perc", "vllm:request_prompt_tokens_bucket",
                   "vllm:request_generation_tokens_bucket", "vllm:request_success_total"):
        assert series in text, series
    assert f'model_name="{MODEL}"' in text
    assert 'finished_reason="length"' in text


def test_completion_matches_engine_and_usage(client):
    prompt = "What is Deep Learning?"
    ids = client.tok.encode(prompt)
    r = client.post("/v1/completions", json={"model": MODEL, "prompt": prompt, "max_tokens": 25, "temperature": 0}).json()
    want = client.tok.decode(expect_ids(ids, 25))
    assert r["object"] == "text_completion" and r["choices"][0]["text"] == want
    assert r["choices"][0]["finish_reason"] == "length"
    assert r["usage"] == {"prompt_tokens": len(ids), "completion_tokens": 25, "total_tokens": len(ids) + 25}
    # token-id prompt + echo
    r2 = client.post("/v1/completions", json={"model": MODEL, "temperature": 0, "prompt": ids, "max_tokens": 25}).json()
    assert r2["choices"][0]["text"] == want


def test_streaming_sse_with_usage(client):
    with client.stream("POST", "/v1/completions", json={"model": MODEL, "temperature": 0, "prompt": "abc", "max_tokens": 8, "stream": True,
                                                        "stream_options": {"include_usage": True}}) as r:
        assert r.status_code == 200
        lines = [l for l in r.iter_lines() if l.startswith("data: ")]
    assert lines[-1] == "data: [DONE]"
    chunks = [json.loads(l[6:]) for l in lines[:-1]]
    text = "".join(c["choices"][0]["text"] for c in chunks if c["choices"])
    assert text == client.tok.decode(expect_ids(client.tok.encode("abc"), 8))
    assert chunks[-1]["choices"] == [] and chunks[-1]["usage"]["completion_tokens"] == 8
    assert any(c["choices"] and c["choices"][0]["finish_reason"] == "length" for c in chunks)


def test_chat_completion_and_stream(client):
    body = {"model": MODEL, "messages": [{"role": "system", "content": "You are helpful."}, {"role": "user", "content": "Hi"}],
            "max_tokens": 6, "temperature": 0}
    r = client.post("/v1/chat/completions", json=body).json()
    assert r["object"] == "chat.completion" and r["choices"][0]["message"]["role"] == "assistant"
    assert r["usage"]["completion_tokens"] == 6
    with client.stream("POST", "/v1/chat/completions", json={**body, "stream": True}) as s:
        chunks = [json.loads(l[6:]) for l in s.iter_lines() if l.startswith("data: ") and l != "data: [DONE]"]
    assert chunks[0]["choices"][0]["delta"].get("role") == "assistant"
    assert "".join(c["choices"][0]["delta"].get("content", "") for c in chunks if c["choices"]) == r["choices"][0]["message"]["content"]


def test_validation_and_errors(client):
    assert client.post("/v1/completions", json={"model": MODEL}).status_code == 422                     # prompt missing
    assert client.post("/v1/completions", json={"model": MODEL, "temperature": 0, "prompt": "x", "bogus_field": 1}).status_code == 422
    assert client.post("/v1/chat/completions", json={"model": MODEL, "temperature": 0, "messages": []}).status_code == 422
    r = client.post("/v1/completions", json={"model": "other", "prompt": "x"})
    assert r.status_code == 404 and r.json()["object"] == "error"
    r = client.post("/v1/completions", json={"model": MODEL, "temperature": 0, "prompt": "x" * 300, "max_tokens": 10})
    assert r.status_code == 400 and "maximum context length" in r.json()["message"]
    assert client.post("/v1/completions", json={"model": MODEL, "temperature": 0, "prompt": "x", "n": 64}).status_code == 400
    assert client.post("/v1/completions", json={"model": MODEL, "temperature": 0, "prompt": "x", "n": 2, "best_of": 4}).status_code == 400
    assert client.post("/v1/completions", json={"model": MODEL, "temperature": 0, "prompt": "x", "use_beam_search": True}).status_code == 400


@pytest.mark.parametrize("field,value", [("guided_json", {"type": "object"}),
                                         ("guided_regex", "a+"), ("guided_choice", ["a", "b"]), ("guided_grammar", "root ::= 'a'"),
                                         ("response_format", {"type": "json_object"}), ("length_penalty", 2.0)])
def test_fields_the_backend_does_not_implement_get_a_400_not_a_silent_200(client, field, value):
    """docs/api-spec.yaml:275-311,397,479-518 defines them; a 200 whose text ignores them would be a wrong answer."""
    for path, extra in (("/v1/completions", {"prompt": "x"}), ("/v1/chat/completions", {"messages": [{"role": "user", "content": "x"}]})):
        r = client.post(path, json={"model": MODEL, "temperature": 0, "max_tokens": 2, field: value, **extra})
        assert r.status_code == 400 and r.json()["object"] == "error" and field in r.json()["message"], (path, r.text)
    # the defaults of the same fields are fine
    ok = {"frequency_penalty": 0.0, "presence_penalty": 0, "repetition_penalty": 1.0, "logit_bias": {}, "length_penalty": 1.0,
          "response_format": {"type": "text"}, "guided_json": None}
    r = client.post("/v1/completions", json={"model": MODEL, "temperature": 0, "max_tokens": 2, "prompt": "x", **ok})
    assert r.status_code == 200


def test_penalties_and_logit_bias_reach_the_device_and_are_validated(client):
    """docs/api-spec.yaml:275-311,397: frequency_penalty, presence_penalty, logit_bias (and vLLM's repetition_penalty) are
    applied on the device to the step's logits before the token is chosen: the engine hands the runner each request's
    own history and table once per step, then takes every row's token from the sampler (also greedy rows: the step's
    own argmax predates the adjustment).  Out-of-range values and the combination with logprobs are 400s."""
    fr = client.app.state.engine_loop.engine.runner
    base = {"model": MODEL, "prompt": "hello", "max_tokens": 3, "temperature": 0}
    r = client.post("/v1/completions", json={**base, "frequency_penalty": 0.5, "presence_penalty": -1.0, "repetition_penalty": 1.2,
                                             "logit_bias": {"17": 5.0, "300": -100}})
    assert r.status_code == 200, r.text
    assert len(fr.adjust_calls) == 3 and len(fr.sample_calls) == 3          # once per generated token
    row = fr.adjust_calls[-1][0]
    ids = client.tok.encode("hello")
    assert row["prompt"] == ids and len(row["output"]) == 2 and row["bias"] == {17: 5.0, 300: -100.0}
    assert (row["rep"], row["freq"], row["pres"]) == (1.2, 0.5, -1.0)
    assert fr.sample_calls[-1]["temperature"] == [0.0]
    n = len(fr.adjust_calls)
    assert client.post("/v1/completions", json=base).status_code == 200
    assert len(fr.adjust_calls) == n, "a request without penalties must not pay for the adjustment"
    for bad in ({"frequency_penalty": 2.5}, {"presence_penalty": -3}, {"repetition_penalty": 0}, {"logit_bias": {"abc": 1}},
                {"logit_bias": {"99999": 1}}, {"logit_bias": {"5": 101}}, {"logit_bias": {"5": 1}, "logprobs": 2}):
        r = client.post("/v1/completions", json={**base, **bad})
        assert r.status_code == 400 and r.json()["object"] == "error", (bad, r.text)
    r = client.post("/v1/chat/completions", json={"model": MODEL, "messages": [{"role": "user", "content": "x"}], "max_tokens": 2,
                                                  "temperature": 0, "presence_penalty": 0.5})
    assert r.status_code == 200


def test_seed_and_top_k_out_of_range_cannot_kill_the_engine(client):
    """a negative or 65-bit seed / a 2^31 top_k used to raise inside the engine thread and take every
    request down with it.  Negative int64 seeds are accepted (as vLLM does), absurd values get a 400, /health stays 200."""
    base = {"model": MODEL, "prompt": "hello", "max_tokens": 3, "temperature": 1.0}
    assert client.post("/v1/completions", json={**base, "seed": -7}).status_code == 200
    assert client.post("/v1/completions", json={**base, "seed": 2 ** 64 - 1, "n": 2}).status_code == 200     # seed + i wraps
    assert client.post("/v1/completions", json={**base, "seed": 2 ** 70}).status_code == 400
    assert client.post("/v1/completions", json={**base, "top_k": 2 ** 31}).status_code == 400
    assert client.post("/v1/completions", json={**base, "top_k": 2 ** 31 - 1}).status_code == 200              # clamped to the vocab
    assert client.post("/v1/completions", json={**base, "max_tokens": 0}).status_code == 400
    assert client.get("/health").status_code == 200


def test_stop_string_aborts_the_engine_request(client):
    """A stop string ends the request in the engine as well (no decoding on to max_tokens, blocks returned)."""
    loop = client.app.state.engine_loop
    ids = client.tok.encode("hello")
    text = client.tok.decode(expect_ids(ids, 12))
    stop = text[3:5]
    r = client.post("/v1/completions", json={"model": MODEL, "temperature": 0, "prompt": "hello", "max_tokens": 200, "stop": stop})
    assert r.status_code == 200 and r.json()["choices"][0]["finish_reason"] == "stop"
    assert stop not in r.json()["choices"][0]["text"] and text.startswith(r.json()["choices"][0]["text"][:2])
    import time
    for _ in range(100):
        if not loop.engine.has_work() and len(loop.engine.blocks.free) == loop.engine.blocks.n_blocks:
            break
        time.sleep(0.02)
    assert not loop.engine.has_work() and len(loop.engine.blocks.free) == loop.engine.blocks.n_blocks
    assert not loop.listeners       # (how many tokens the engine got through before the abort landed is a race with a fake runner)


def test_client_disconnect_aborts_a_stream(client):
    """Closing an SSE stream early frees the request's KV blocks and sequence slot (vLLM aborts on disconnect)."""
    import time
    loop = client.app.state.engine_loop
    with client.stream("POST", "/v1/completions", json={"model": MODEL, "temperature": 0, "prompt": "hello", "max_tokens": 240,
                                                        "stream": True, "ignore_eos": True}) as s:
        it = s.iter_lines()
        got = 0
        for line in it:
            if line.startswith("data: "):
                got += 1
            if got >= 3:
                break
    for _ in range(200):
        if not loop.engine.has_work():
            break
        time.sleep(0.02)
    assert not loop.engine.has_work() and len(loop.engine.blocks.free) == loop.engine.blocks.n_blocks
    assert client.get("/health").status_code == 200


def test_stop_and_ignore_eos(client):
    ids = client.tok.encode("hello")
    first = expect_ids(ids, 1)[0]
    r = client.post("/v1/completions", json={"model": MODEL, "temperature": 0, "prompt": "hello", "max_tokens": 9, "stop_token_ids": [first]}).json()
    assert r["usage"]["completion_tokens"] == 1 and r["choices"][0]["finish_reason"] == "stop"
    r = client.post("/v1/completions", json={"model": MODEL, "temperature": 0, "prompt": "hello", "max_tokens": 9, "ignore_eos": True}).json()
    assert r["usage"]["completion_tokens"] == 9


def test_chart_flags_are_accepted():
    """Every arg the chart emits (xeon-values.yaml:70-89 + deployment.yaml:72-87, and the Gaudi variant)."""
    argv = ["--block-size", "128", "--dtype", "bfloat16", "--max-model-len", "8192", "--distributed_executor_backend", "mp",
            "--enable_chunked_prefill", "--enforce-eager", "--max-num-batched-tokens", "2048", "--max-num-seqs", "256",
            "--disable-log-requests", "--enable-auto-tool-choice", "--tool-call-parser", "llama3_json",
            "--tensor-parallel-size", "1", "--pipeline-parallel-size", "1", "--model", MODEL, "--served-model-name", MODEL,
            "--port", "2080", "--gpu-memory-util", "0.98", "--max-num-prefill-seqs", "16", "--num_scheduler_steps", "1",
            "--use-padding-aware-scheduling"]
    args, unknown = S.make_arg_parser().parse_known_args(argv)
    assert unknown == [] and args.block_size == 128 and args.port == 2080 and args.max_num_seqs == 256
    S.validate_args(args)
    bad, _ = S.make_arg_parser().parse_known_args(argv + ["--pipeline-parallel-size", "2"])
    with pytest.raises(SystemExit):
        S.validate_args(bad)


def test_health_goes_red_when_the_engine_dies(client):
    loop = client.app.state.engine_loop

    def boom(_):
        raise RuntimeError("device lost")
    loop.engine.runner.step = boom
    if loop.engine.overlap:
        loop.engine.runner.step_async = boom
    r = client.post("/v1/completions", json={"model": MODEL, "temperature": 0, "prompt": "x", "max_tokens": 2})
    assert r.status_code == 500
    assert client.get("/health").status_code == 503


def test_tool_call_parsers():
    T = pkg("tool_parsers")
    tools = [{"type": "function", "function": {"name": "get_weather", "parameters": {}}}]
    c, calls = T.parse_llama3_json('{"name": "get_weather", "parameters": {"city": "Paris"}}', tools)
    assert c is None and calls[0]["function"]["name"] == "get_weather" and json.loads(calls[0]["function"]["arguments"]) == {"city": "Paris"}
    c, calls = T.parse_llama3_json('<|python_tag|>{"name": "get_weather", "parameters": {}}; {"name": "get_weather", "arguments": {"city": "Rome"}}', tools)
    assert len(calls) == 2
    assert T.parse_llama3_json("The weather is nice.", tools) == ("The weather is nice.", [])
    assert T.parse_llama3_json('{"name": "unknown_fn", "parameters": {}}', tools)[1] == []
    c, calls = T.parse_hermes('Let me check.\n<tool_call>\n{"name": "get_weather", "arguments": {"city": "Oslo"}}\n</tool_call>', tools)
    assert c == "Let me check." and calls[0]["type"] == "function" and calls[0]["id"].startswith("chatcmpl-tool-")
    assert T.parse_hermes("<tool_call>not json</tool_call>", tools) == ("<tool_call>not json</tool_call>", [])
    assert T.get_parser(None) is None and T.get_parser("hermes") is T.parse_hermes
    with pytest.raises(SystemExit):
        T.get_parser("mistral")


def test_chat_tool_calls_through_the_server():
    """With --enable-auto-tool-choice the parser turns a JSON assistant turn into tool_calls."""
    class Scripted(FakeRunner):
        script = list(b'{"name": "get_weather", "parameters": {"city": "Paris"}}')

        def step(self, b):
            out = super().step(b)
            for j, r in enumerate(b.logit_rows):
                n_prompt = self.prompt_len
                k = int(b.ctx_lens[r]) - n_prompt
                out[j] = self.script[k] if 0 <= k < len(self.script) else 257
            return out
    fr = Scripted()
    eng = E.LLMEngine(fr, max_num_batched_tokens=64, max_num_seqs=8, max_model_len=256)
    tok = S.ByteTokenizer(512)
    app = S.create_app(eng, tok, MODEL, 256, tool_parser=pkg("tool_parsers").parse_llama3_json)
    msgs = [{"role": "user", "content": "Weather in Paris?"}]
    fr.prompt_len = len(tok.encode(tok.apply_chat_template(msgs), add_special_tokens=False))
    tools = [{"type": "function", "function": {"name": "get_weather", "parameters": {"type": "object"}}}]
    with TestClient(app) as c:
        r = c.post("/v1/chat/completions", json={"model": MODEL, "temperature": 0, "messages": msgs, "tools": tools, "max_tokens": 80}).json()
    ch = r["choices"][0]
    assert ch["finish_reason"] == "tool_calls" and ch["message"]["content"] is None
    assert ch["message"]["tool_calls"][0]["function"]["name"] == "get_weather"


def test_sampling_fields_reach_the_device_sampler():
    """temperature / top_p / top_k / seed of the OpenAI body (docs/api-spec.yaml) travel to the runner's
    sampler with the draw number = tokens generated so far; temperature 0 never calls it; an omitted
    temperature takes the server default (vLLM: generation_config.json, else 1.0)."""
    runner = FakeRunner()
    engine = E.LLMEngine(runner, max_num_batched_tokens=64, max_num_seqs=8, max_model_len=256)
    app = S.create_app(engine, S.ByteTokenizer(512), MODEL, 256, default_sampling={"temperature": 0.6, "top_p": 0.9})
    with TestClient(app) as client:
        r = client.post("/v1/completions", json={"model": MODEL, "prompt": "abc", "max_tokens": 3, "temperature": 0})
        assert r.status_code == 200 and not getattr(runner, "sample_calls", [])
        r = client.post("/v1/completions", json={"model": MODEL, "prompt": "abc", "max_tokens": 3, "temperature": 0.8,
                                                 "top_p": 0.5, "top_k": 7, "seed": 1234, "min_p": 0.05})
        assert r.status_code == 200
        calls = runner.sample_calls
        assert len(calls) == 3
        assert all(abs(c["min_p"][0] - 0.05) < 1e-6 for c in calls)
        assert [c["draws"] for c in calls] == [[0], [1], [2]]
        assert all(c["seeds"] == [1234] and c["top_k"] == [7] for c in calls)
        assert all(abs(c["temperature"][0] - 0.8) < 1e-6 and abs(c["top_p"][0] - 0.5) < 1e-6 for c in calls)
        runner.sample_calls = []
        r = client.post("/v1/completions", json={"model": MODEL, "prompt": "abc", "max_tokens": 2})
        assert r.status_code == 200
        c = runner.sample_calls[0]
        assert abs(c["temperature"][0] - 0.6) < 1e-6 and abs(c["top_p"][0] - 0.9) < 1e-6 and c["top_k"] == [0]
        assert c["seeds"][0] != 0          # no seed in the body: one is drawn per request


def test_n_parallel_samples(client):
    """`n` (docs/api-spec.yaml): n choices with indices 0..n-1 from n engine requests over the same prompt;
    usage counts all of them; a body seed gives child i the seed + i; streaming interleaves the choices."""
    ids = client.tok.encode("abc")
    r = client.post("/v1/completions", json={"model": MODEL, "prompt": "abc", "max_tokens": 5, "temperature": 0, "n": 3}).json()
    assert [c["index"] for c in r["choices"]] == [0, 1, 2]
    want = client.tok.decode(expect_ids(ids, 5))
    assert all(c["text"] == want and c["finish_reason"] == "length" for c in r["choices"])      # greedy: identical
    assert r["usage"] == {"prompt_tokens": len(ids), "completion_tokens": 15, "total_tokens": len(ids) + 15}
    r = client.post("/v1/chat/completions", json={"model": MODEL, "messages": [{"role": "user", "content": "hi"}],
                                                  "max_tokens": 4, "temperature": 0, "n": 2}).json()
    assert [c["index"] for c in r["choices"]] == [0, 1] and r["choices"][0]["message"] == r["choices"][1]["message"]
    with client.stream("POST", "/v1/completions", json={"model": MODEL, "prompt": "abc", "max_tokens": 4, "temperature": 0,
                                                        "n": 2, "stream": True}) as resp:
        lines = [l for l in resp.iter_lines() if l.startswith("data: ")]
    assert lines[-1] == "data: [DONE]"
    chunks = [json.loads(l[6:]) for l in lines[:-1]]
    for i in (0, 1):
        text = "".join(c["choices"][0]["text"] for c in chunks if c["choices"] and c["choices"][0]["index"] == i)
        assert text == client.tok.decode(expect_ids(ids, 4))
        assert [c["choices"][0]["finish_reason"] for c in chunks if c["choices"] and c["choices"][0]["index"] == i][-1] == "length"


def test_utility_routes(client):
    assert client.get("/ping").status_code == 200 and client.post("/ping").status_code == 200
    assert "version" in client.get("/version").json()
    t = client.post("/tokenize", json={"model": MODEL, "prompt": "What is Deep Learning?"}).json()
    assert t["count"] == len(t["tokens"]) > 0 and t["tokens"] == client.tok.encode("What is Deep Learning?")
    d = client.post("/detokenize", json={"model": MODEL, "tokens": t["tokens"]}).json()
    assert d["prompt"].endswith("What is Deep Learning?")


def test_logprobs_shapes(client):
    """`logprobs` (completions: N) and `logprobs` + `top_logprobs` (chat) in the OpenAI shapes of docs/api-spec.yaml."""
    ids = client.tok.encode("abc")
    toks = expect_ids(ids, 4)
    r = client.post("/v1/completions", json={"model": MODEL, "prompt": "abc", "max_tokens": 4, "temperature": 0, "logprobs": 2}).json()
    lp = r["choices"][0]["logprobs"]
    assert lp["tokens"] == [client.tok.decode([t], skip_special_tokens=False) for t in toks]
    assert lp["token_logprobs"] == [-0.25] * 4 and len(lp["top_logprobs"]) == 4 and lp["text_offset"][0] == 0
    assert all(len(d) in (1, 2) and min(d.values()) >= -1.25 for d in lp["top_logprobs"])
    r = client.post("/v1/chat/completions", json={"model": MODEL, "messages": [{"role": "user", "content": "hi"}], "max_tokens": 3,
                                                  "temperature": 0, "logprobs": True, "top_logprobs": 3}).json()
    content = r["choices"][0]["logprobs"]["content"]
    assert len(content) == 3 and all(len(c["top_logprobs"]) == 3 and c["logprob"] == -0.25 and "bytes" in c for c in content)
    r = client.post("/v1/completions", json={"model": MODEL, "prompt": "abc", "max_tokens": 2, "temperature": 0}).json()
    assert r["choices"][0]["logprobs"] is None
    assert client.post("/v1/completions", json={"model": MODEL, "prompt": "abc", "temperature": 0, "logprobs": 50}).status_code == 400
    with client.stream("POST", "/v1/completions", json={"model": MODEL, "prompt": "abc", "max_tokens": 4, "temperature": 0,
                                                        "logprobs": 1, "stream": True}) as resp:
        chunks = [json.loads(l[6:]) for l in resp.iter_lines() if l.startswith("data: ") and l != "data: [DONE]"]
    got = [t for c in chunks if c["choices"] and c["choices"][0]["logprobs"] for t in c["choices"][0]["logprobs"]["tokens"]]
    assert got == lp["tokens"]
