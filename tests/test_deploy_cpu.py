"""CPU: the deploy/ files against the reference chart (skipped where /root/reference is absent, e.g. on
the GPU box): b200-values.yaml keeps the schema of xeon-values.yaml, and the server's CLI accepts every
argument any of the reference's values files can put on the command line."""
import glob
import os
import shutil

import pytest
import yaml

from conftest import ROOT, pkg

REF = "/root/reference/core/helm-charts/vllm"
needs_ref = pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present")


def load(path):
    with open(path) as f:
        return yaml.safe_load(f)


@needs_ref
def test_b200_values_keep_the_reference_schema():
    ours = load(os.path.join(ROOT, "deploy", "helm-charts", "vllm", "b200-values.yaml"))
    ref = load(os.path.join(REF, "xeon-values.yaml"))
    for key in ("accelDevice", "image", "modelConfigs", "defaultModelConfigs", "LLM_MODEL_ID"):
        assert key in ours, key
    assert set(ours["image"]) >= {"repository", "tag", "pullPolicy"}
    ref_model = next(iter(ref["modelConfigs"].values()))
    for name, mc in ours["modelConfigs"].items():
        assert set(mc) <= set(ref_model), f"{name}: keys outside the reference schema: {set(mc) - set(ref_model)}"
        assert {"extraCmdArgs", "tensor_parallel_size"} <= set(mc)
    assert set(ours["defaultModelConfigs"]) <= set(ref["defaultModelConfigs"])
    assert "meta-llama/Llama-3.1-8B-Instruct" in ours["modelConfigs"]      # the reference's default CPU model


@needs_ref
@pytest.mark.parametrize("values", ["xeon-values.yaml", "gaudi-values.yaml", "gaudi3-values.yaml"])
def test_every_reference_cmd_arg_is_accepted(values):
    S = pkg("server")
    ref = load(os.path.join(REF, values))
    arg_lists = [mc.get("extraCmdArgs") or [] for mc in (ref.get("modelConfigs") or {}).values()]
    arg_lists.append((ref.get("defaultModelConfigs") or {}).get("extraCmdArgs") or [])
    seen = 0
    for argv in arg_lists:
        argv = [str(a) for a in argv]
        args, unknown = S.make_arg_parser().parse_known_args(argv + ["--model", "m", "--port", "2080"])
        # flags this backend has no use for may be unknown, but they must be ignorable (never an error),
        # and the ones that shape the decode path must be understood
        assert args.block_size == 128 or "--block-size" not in argv
        for must in ("--block-size", "--dtype", "--max-model-len", "--max-num-seqs", "--max-num-batched-tokens"):
            assert must not in unknown
        seen += 1
    assert seen >= 2


def test_our_values_args_parse_and_validate():
    S = pkg("server")
    ours = load(os.path.join(ROOT, "deploy", "helm-charts", "vllm", "b200-values.yaml"))
    for name, mc in ours["modelConfigs"].items():
        argv = [str(a) for a in mc["extraCmdArgs"]] + ["--tensor-parallel-size", str(mc["tensor_parallel_size"]),
                                                       "--model", name, "--served-model-name", name, "--port", "2080"]
        args, unknown = S.make_arg_parser().parse_known_args(argv)
        assert unknown == []
        S.validate_args(args)


def test_playbook_models_exist_in_the_values_file():
    """Every model the B200 task file can deploy has a modelConfigs entry, and its default TP degree
    matches the values file (the helm --set only overrides it when the operator asks)."""
    tasks = load(os.path.join(ROOT, "deploy", "playbooks", "deploy-b200-models.yml"))
    table = next(t["ansible.builtin.set_fact"]["b200_models"] for t in tasks if "ansible.builtin.set_fact" in t)
    values = load(os.path.join(ROOT, "deploy", "helm-charts", "vllm", "b200-values.yaml"))
    assert len(table) >= 3
    for m in table:
        assert m["hf"] in values["modelConfigs"], m["hf"]
        assert int(values["modelConfigs"][m["hf"]]["tensor_parallel_size"]) == int(m["tp"]), m
        assert m["key"].startswith("b200-") and m["release"].endswith("-b200")
    tags = {tag for t in tasks for tag in (t.get("tags") if isinstance(t.get("tags"), list) else [])}
    for m in table:
        assert f"install-{m['key']}" in tags
    vars_patch = open(os.path.join(ROOT, "deploy", "patches", "0004-vars-b200-values-file.patch")).read()
    assert '+b200_values_file: "{{ remote_helm_charts_base }}/vllm/b200-values.yaml"' in vars_patch


REF_ROOT = "/root/reference"


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF_ROOT, "core")) or shutil.which("patch") is None,
                    reason="needs the reference tree and patch(1)")
def test_patches_apply_to_the_reference_and_the_patched_menu_works(tmp_path):
    """deploy/patches/*.patch are real unified diffs against the reference: they apply cleanly to a copy of it, the
    patched shell still parses, the menu maps id 31 / the name b200-llama-70b to the playbook's model keys and refuses
    a B200 id on a CPU deployment, and the chart template carries the nvidia.com/gpu branch and the 120 s grace period."""
    import subprocess
    work = tmp_path / "ref"
    shutil.copytree(os.path.join(REF_ROOT, "core"), work / "core")
    patches = sorted(glob.glob(os.path.join(ROOT, "deploy", "patches", "*.patch")))
    assert len(patches) == 5
    for p in patches:
        r = subprocess.run(["patch", "-p1", "--forward", "-i", p], cwd=work, capture_output=True, text=True)
        assert r.returncode == 0, f"{os.path.basename(p)}: {r.stdout}{r.stderr}"
    sel = work / "core/lib/models/model-selection.sh"
    for f in (sel, work / "core/lib/user-menu/parse-user-prompts.sh", work / "core/lib/system/setup-env.sh"):
        assert subprocess.run(["bash", "-n", str(f)]).returncode == 0

    def names(device, models):
        r = subprocess.run(["bash", "-c", f'source "{sel}"; cpu_or_gpu={device}; models="{models}"; get_model_names'],
                           capture_output=True, text=True)
        return r.returncode, r.stdout.strip(), r.stderr
    assert names("b", "31") == (0, "b200-llama-8b", "")
    assert names("b", "33,b200-deepseek-r1-distill-qwen-32b")[1].replace(",", " ").split() == ["b200-llama-70b", "b200-deepseek-r1-distill-qwen-32b"]
    rc, out, err = names("c", "31")
    assert rc != 0 and "B200 model identifier" in err
    assert names("c", "21")[:2] == (0, "cpu-llama-8b")                      # the existing ids are untouched
    tasks = load(os.path.join(ROOT, "deploy", "playbooks", "deploy-b200-models.yml"))
    keys = {m["key"] for t in tasks if "ansible.builtin.set_fact" in t for m in t["ansible.builtin.set_fact"]["b200_models"]}
    assert {"b200-llama-8b", "b200-deepseek-r1-distill-qwen-32b", "b200-llama-70b"} <= keys
    tpl = (work / "core/helm-charts/vllm/templates/deployment.yaml").read_text()
    assert 'nvidia.com/gpu:' in tpl and 'if or (not .Values.accelDevice) (eq .Values.accelDevice "b200")' in tpl
    assert "b200_values_file" in (work / "core/inventory/metadata/vars/inference_llm_models.yml").read_text()
