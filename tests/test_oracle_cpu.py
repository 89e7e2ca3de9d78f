"""CPU: the oracle against its frozen golden vectors (and through them the HF stand-in), the
product-side config/rope restatement against the oracle's, and size-independent properties."""
import ast
import os

import numpy as np
import pytest
import torch

from conftest import pkg
from oracle import decode_oracle as O


@pytest.mark.parametrize("name", ["tiny", "tiny-qwen", "tiny-g8", "tiny-g8x8"])
def test_oracle_reproduces_golden(name, golden_dir):
    g = np.load(f"{golden_dir}/{name}.npz")
    cfg = O.CONFIGS[name]
    frozen = ast.literal_eval(str(g["cfg"]))
    assert frozen == {k: v for k, v in cfg.to_dict().items() if k in frozen}
    w = O.make_weights(cfg, seed=int(g["seed"]), norm_jitter=float(g["norm_jitter"]))
    assert O.weights_digest(w) == str(g["digest"]), "torch CPU RNG stream drifted: regenerate fixtures"
    om = O.OracleModel(cfg, w, capacity=256)
    prompt = g["prompt"].tolist()
    toks, logits = om.generate(prompt, len(g["tokens"]), keep_logits=True)
    # Not bit-equal across hosts: R2 leaves the fp32 summation order free, the host BLAS picks it by ISA
    # (AVX2 / AVX-512 / AMX boxes differ), and one bf16 that rounds the other way in a K/V row moves every
    # later row by an ulp.  So: the free-running tokens agree up to the first near-tie, and the logits up to
    # there (same inputs) agree to bf16-ulp level; the teacher-forced logits (same inputs on every row) too.
    want = g["tokens"].tolist()
    gl = torch.from_numpy(g["gen_logits"]).float()
    n = len(want)
    for i, (a, b) in enumerate(zip(toks, want)):
        if a != b:
            m = float(O.top2_margin(gl[i]))
            assert m < 0.08, f"{name}: greedy token {i} differs from the fixture at an oracle margin of {m:.3f}"
            n = i + 1
            break
    assert_logits_close(logits[:n].float(), gl[:n], name)
    st = om.new_seq()
    tf = om.forward(st, prompt + want[:-1]).float()
    assert_logits_close(tf, torch.from_numpy(g["tf_logits"]).float(), name)


def assert_logits_close(a, b, what):
    """<= 2 bf16 ulp of the row's largest logit per element (an ulp of a small logit is far below the noise of its
    own sum), softmax max-abs error <= 1e-3 (north_star bar: 2e-2)."""
    ulp = b.abs().amax(-1, keepdim=True) * 2.0 ** -7
    worst = ((a - b).abs() / ulp).max().item()
    assert worst <= 2.0, f"{what}: logits differ from the fixture by {worst:.2f} bf16 ulp of the row maximum"
    err = (torch.softmax(a, -1) - torch.softmax(b, -1)).abs().max().item()
    assert err <= 1e-3, f"{what}: softmax differs from the fixture by {err:.2e}"


@pytest.mark.parametrize("name", ["tiny", "tiny-qwen", "tiny-g8", "tiny-g8x8"])
def test_golden_within_north_star_tolerance_of_hf(name, golden_dir):
    """The frozen HF (transformers 5.5.0, CPU bf16) logits bound the oracle: softmax max abs
    error <= 2e-2 (north_star); greedy picks agree wherever the oracle's margin is not tiny."""
    g = np.load(f"{golden_dir}/{name}.npz")
    a, b = torch.from_numpy(g["tf_logits"]), torch.from_numpy(g["hf_logits"])
    assert (torch.softmax(a, -1) - torch.softmax(b, -1)).abs().max() <= 2e-2
    margin = O.top2_margin(a)
    dis = a.argmax(-1) != b.argmax(-1)
    assert (margin[dis] < 0.07).all(), "greedy disagreement outside near-ties"


def test_chunked_prefill_equals_token_at_a_time():
    cfg = O.CONFIGS["tiny"]
    om = O.OracleModel(cfg, O.make_weights(cfg), capacity=64)
    ids = O.make_prompt(cfg, 3, 20)
    a = om.forward(om.new_seq(), ids)
    st = om.new_seq()
    b = torch.cat([om.forward(st, ids[:7]), om.forward(st, ids[7:8]), om.forward(st, ids[8:])])
    # not bit-equal: the fp32 GEMM's summation order depends on the row count (as it does in the
    # reference image); the difference must stay at bf16-ulp level
    assert (a.float() - b.float()).abs().max() <= 2 ** -5
    assert (torch.softmax(a.float(), -1) - torch.softmax(b.float(), -1)).abs().max() <= 1e-3


def test_product_config_and_rope_match_oracle():
    C = pkg("config")
    for oname, pname in (("llama-3.1-8b", "llama-3.1-8b"), ("qwen2-32b", "qwen2-32b"), ("llama-3.1-70b", "llama-3.1-70b"),
                         ("tiny", "tiny"), ("tiny-g8x8", "tiny-g8x8")):
        o, p = O.CONFIGS[oname].to_dict(), C.get_config(pname).to_dict()
        o.pop("name"), p.pop("name")
        assert o == p
    # SURVEY §8 shape table
    assert C.get_config("llama-3.1-8b").decode_params() == 7_504_658_432
    assert C.get_config("qwen2-32b").decode_params() == 31_984_189_440
    assert C.get_config("llama-3.1-70b").decode_params() == 69_501_714_432
    assert C.get_config("llama-3.1-8b").kv_bytes_per_token() == 131072
    for name in ("tiny", "qwen2-32b"):
        assert torch.equal(C.rope_cos_sin(C.get_config(name), 4096), O.cos_sin_cache(O.CONFIGS[name], 4096))


def test_rounding_points():
    """R1/R5: the bf16 intermediate roundings are really there (a 'better' fp32 path differs)."""
    torch.manual_seed(0)
    x = torch.randn(4, 512).to(torch.bfloat16)
    w = (1 + 0.1 * torch.randn(512)).to(torch.bfloat16)
    y = O.rms_norm(x, w, 1e-5)
    xf = x.float()
    better = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5) * w.float()).to(torch.bfloat16)
    assert not torch.equal(y, better)
    y2, r2 = O.add_rms_norm(x, x, w, 1e-5)
    assert torch.equal(r2, (x.float() * 2).to(torch.bfloat16))
    assert O.greedy(torch.tensor([[1.0, 3.0, 3.0, 2.0]]).to(torch.bfloat16)).item() == 1  # ties -> lowest


def test_batched_step_matches_per_sequence_forward():
    cfg = O.CONFIGS["tiny"]
    om = O.OracleModel(cfg, O.make_weights(cfg, norm_jitter=0.1), capacity=64)
    prompts = [O.make_prompt(cfg, i, n) for i, n in enumerate((3, 9, 17))]
    sts, ref = [], []
    for p in prompts:
        st = om.new_seq(); om.forward(st, p[:-1]); sts.append(st)
        st2 = om.new_seq(); ref.append(om.forward(st2, p)[-1])
    for fast in (False, True):
        import copy
        got = om.step_batch(copy.deepcopy(sts), [p[-1] for p in prompts], fast_linear=fast)
        assert (torch.softmax(got.float(), -1) - torch.softmax(torch.stack(ref).float(), -1)).abs().max() <= 1e-3


def test_sampling_probs_match_vllm_masks():
    """oracle.sampling_probs against keep-masks produced by vLLM's own apply_top_k_top_p_pytorch
    (tools/make_sampling_golden.py): the same tokens survive top-k / top-p, probabilities sum to 1."""
    import numpy as np
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "sampling_masks.npz"))
    seeds = sorted(int(k.split("_")[1]) for k in z.files if k.startswith("logits_"))
    assert len(seeds) >= 8
    for sd in seeds:
        logits = torch.from_numpy(z[f"logits_{sd}"].copy()).view(torch.bfloat16)
        T, k, p = z[f"params_{sd}"]
        probs = O.sampling_probs(logits, float(T), int(k), float(p))
        keep = torch.from_numpy(z[f"keep_{sd}"])
        assert torch.equal(probs > 0, keep), f"case {sd}: kept sets differ"
        assert abs(float(probs.sum()) - 1.0) < 1e-5
        # the kept probabilities are the softmax of logits / T restricted to the kept set
        ref = (logits.float() / float(T)).masked_fill(~keep, float("-inf")).softmax(-1)
        assert torch.allclose(probs, ref, atol=1e-6)


def _bf(a):
    import numpy as np
    return torch.from_numpy(np.ascontiguousarray(a)).view(torch.bfloat16)


def test_oracle_ops_equal_vllm_native_ops_bit_for_bit():
    """The oracle's rounding points against vLLM's OWN native-path operators (RMSNorm / fused add+RMSNorm,
    SiluAndMul, Llama3 and plain rotary embedding: cache and application), outputs captured from the `vllm`
    package by tools/make_vllm_op_golden.py.  Bit equality pins R1-R3 and R5 at the operator level."""
    import numpy as np
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "vllm_native_ops.npz"))
    for h in (256, 4096):
        x, res, w = _bf(z[f"rms_x_{h}"]), _bf(z[f"rms_res_{h}"]), _bf(z[f"rms_w_{h}"])
        assert torch.equal(O.rms_norm(x, w, 1e-5), _bf(z[f"rms_y_{h}"]))
        y, r = O.add_rms_norm(x, res, w, 1e-5)
        assert torch.equal(y, _bf(z[f"addrms_y_{h}"])) and torch.equal(r, _bf(z[f"addrms_res_{h}"]))
    assert torch.equal(O.silu_mul(_bf(z["silu_gu"])), _bf(z["silu_y"]))
    pos = torch.from_numpy(z["rope_pos"])
    rows = torch.from_numpy(z["rope_cache_rows"])
    q, k = _bf(z["rope_q"]), _bf(z["rope_k"])
    for name, cfg in (("llama3", O.CONFIGS["llama-3.1-8b"]), ("plain", O.CONFIGS["qwen2-32b"])):
        cache = O.cos_sin_cache(cfg, 8192)
        assert torch.equal(cache[rows], _bf(z[f"rope_cache_{name}"])), f"{name}: cos/sin cache differs"
        cs = cache[pos]
        q2 = O.rope_apply(q.view(len(pos), 4, 128), cs[:, :64], cs[:, 64:]).reshape(len(pos), -1)
        k2 = O.rope_apply(k.view(len(pos), 2, 128), cs[:, :64], cs[:, 64:]).reshape(len(pos), -1)
        assert torch.equal(q2, _bf(z[f"rope_q_{name}"])) and torch.equal(k2, _bf(z[f"rope_k_{name}"]))


def test_sampling_probs_min_p_matches_vllm_masks():
    """min_p: oracle.sampling_probs against the keep-masks of vLLM's own MinPLogitsProcessor.apply followed by its
    top-k/top-p function (captured by tools/make_sampling_golden.py in Sampler.sample's order)."""
    import numpy as np
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "sampling_masks.npz"))
    seeds = sorted(int(k.split("_")[2]) for k in z.files if k.startswith("mp_logits_"))
    assert len(seeds) >= 7
    for sd in seeds:
        logits = torch.from_numpy(z[f"mp_logits_{sd}"].copy()).view(torch.bfloat16)
        T, mp, k, p = z[f"mp_params_{sd}"]
        probs = O.sampling_probs(logits, float(T), int(k), float(p), min_p=float(mp))
        keep = torch.from_numpy(z[f"mp_keep_{sd}"])
        assert torch.equal(probs > 0, keep), f"min_p case {sd}: kept sets differ ({int((probs > 0).sum())} vs {int(keep.sum())})"
        assert abs(float(probs.sum()) - 1.0) < 1e-5


@pytest.mark.parametrize("hq,hkv,n", [(4, 1, 1), (32, 8, 129), (40, 8, 700), (64, 8, 2049)])
def test_oracle_attention_against_torch_sdpa(hq, hkv, n):
    """R4 has no host-runnable vLLM implementation in this image (its CPU attention is compiled code the CUDA wheel
    lacks), so it stays a restatement; this narrows it against the independent implementation vLLM-CPU's SDPA backend
    itself calls for prefill rows: torch.nn.functional.scaled_dot_product_attention on the host, GQA, causal last row.
    fp32 SDPA: the bf16 results agree to 1 ulp on almost every element (SDPA scales q and k by sqrt(scale) each, the
    oracle scales the scores: different roundings, same arithmetic).  bf16 SDPA (P rounded to bf16 before P.V, which
    R4 does NOT do): within 2 ulp."""
    g = torch.Generator().manual_seed(hq * 1000 + n)
    q = torch.randn(hq, 128, generator=g).to(torch.bfloat16)
    K = torch.randn(n, hkv, 128, generator=g).to(torch.bfloat16)
    V = torch.randn(n, hkv, 128, generator=g).to(torch.bfloat16)
    want = O.attend(q, K, V, n, 128 ** -0.5).float()
    F = torch.nn.functional
    for dt, max_ulp, max_frac in ((torch.float32, 1, 0.02), (torch.bfloat16, 2, 0.6)):
        o = F.scaled_dot_product_attention(q.to(dt)[None, :, None, :], K.to(dt).permute(1, 0, 2)[None], V.to(dt).permute(1, 0, 2)[None],
                                           scale=128 ** -0.5, enable_gqa=True)[0, :, 0].to(torch.bfloat16).float()
        d = (o - want).abs()
        ulp = want.abs().clamp_min(2.0 ** -126).log2().floor().exp2() * 2.0 ** -7
        assert (d <= torch.maximum(max_ulp * ulp, torch.tensor(2.0 ** -9))).all(), f"{dt}: max diff {float(d.max()):.3e}"
        assert float((d > 0).float().mean()) <= max_frac, f"{dt}: {float((d > 0).float().mean()):.3f} of elements differ"


def test_adjust_logits_equals_vllm_penalties_and_logit_bias():
    """oracle.adjust_logits against rows processed by vLLM's own apply_penalties (+ the logit_bias add of its
    LogitBiasLogitsProcessor), captured by tools/make_penalty_golden.py: bit for bit in fp32."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "penalties.npz"))
    seeds = sorted(int(k.split("_")[1]) for k in z.files if k.startswith("logits_"))
    assert len(seeds) >= 8
    for sd in seeds:
        logits = torch.from_numpy(z[f"logits_{sd}"].copy()).view(torch.bfloat16)
        rep, freq, pres = (float(v) for v in z[f"params_{sd}"])
        bias = {int(i): float(b) for i, b in zip(z[f"bias_ids_{sd}"], z[f"bias_{sd}"])}
        got = O.adjust_logits(logits, z[f"prompt_{sd}"].tolist(), z[f"output_{sd}"].tolist(), rep, freq, pres, bias)
        want = torch.from_numpy(z[f"want_{sd}"])
        assert torch.equal(got, want), f"case {sd}: {(got != want).sum()} logits differ, max {(got - want).abs().max():.3e}"
        changed = (want != logits.float())
        assert changed.any() and changed.sum() <= len(z[f"prompt_{sd}"]) + len(z[f"output_{sd}"]) + len(bias)


def test_oracle_mlp_and_norm_against_vllm_modules():
    """One level above the operator pins: vLLM's own LlamaMLP and RMSNorm MODULES run on the host
    (tools/make_vllm_module_golden.py) against the oracle's layer pieces on the same weights.  The norm is bit-exact; the
    MLP may differ by the summation order of the host's bf16 GEMM from the oracle's fp32 GEMM (R2 leaves it free):
    <= 1 bf16 ulp on <= 1 % of the outputs (bit-equal on the box that made the fixture)."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "vllm_modules.npz"))
    for tag, h, inter, T in (("a", 512, 1024, 5), ("b", 1024, 2048, 17), ("c", 640, 1536, 3)):
        g = torch.Generator().manual_seed(h + T)      # the generator's stream: gate_up, down, x, norm weight, residual
        gu = (0.02 * torch.randn(2 * inter, h, generator=g)).to(torch.bfloat16)
        down = (0.02 * torch.randn(h, inter, generator=g)).to(torch.bfloat16)
        x = torch.randn(T, h, generator=g).to(torch.bfloat16)
        nw = (1 + 0.1 * torch.randn(h, generator=g)).to(torch.bfloat16)
        res = torch.randn(T, h, generator=g).to(torch.bfloat16)
        assert int(gu.view(torch.int16).long().sum() + x.view(torch.int16).long().sum()) == int(z[f"mlp_{tag}_digest"]), "RNG stream drifted"
        got, want = O.linear(O.silu_mul(O.linear(x, gu)), down).float(), _bf(z[f"mlp_{tag}_y"]).float()
        d = (got - want).abs()
        ulp = want.abs().clamp_min(2.0 ** -126).log2().floor().exp2() * 2.0 ** -7
        assert (d <= torch.maximum(ulp, torch.tensor(2.0 ** -12))).all() and float((d > 0).float().mean()) <= 0.01, tag
        y, r = O.add_rms_norm(x, res, nw, 1e-5)
        assert torch.equal(y, _bf(z[f"norm_{tag}_y"])) and torch.equal(r, _bf(z[f"norm_{tag}_r"])), tag
