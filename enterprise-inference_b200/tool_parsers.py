"""Tool-call parsers for /v1/chat/completions, selected by the chart's `--enable-auto-tool-choice
--tool-call-parser llama3_json|hermes` (/root/reference/core/helm-charts/vllm/xeon-values.yaml:87-88,118-119,209-210).

llama3_json: the assistant turn is a JSON object (or several separated by ';')
             {"name": "<fn>", "parameters": {...}}   (also accepts "arguments")
hermes:      <tool_call>{"name": "<fn>", "arguments": {...}}</tool_call>, any number of them, free text around.
Both return (content, tool_calls) with tool_calls in the OpenAI wire format; text that does not parse is
returned unchanged as content with no tool calls.
"""
from __future__ import annotations

import json
import re
import uuid
from typing import Any, Dict, List, Optional, Tuple


def _call(name: str, args: Any) -> Dict[str, Any]:
    if not isinstance(args, str):
        args = json.dumps(args if args is not None else {}, ensure_ascii=False)
    return {"id": "chatcmpl-tool-" + uuid.uuid4().hex[:24], "type": "function", "function": {"name": name, "arguments": args}}


def _known(name: str, tools: Optional[List[Dict[str, Any]]]) -> bool:
    if not tools:
        return True
    return any((t.get("function") or {}).get("name") == name for t in tools)


def parse_llama3_json(text: str, tools=None) -> Tuple[Optional[str], List[Dict[str, Any]]]:
    body = text.strip()
    if body.startswith("<|python_tag|>"):
        body = body[len("<|python_tag|>"):].strip()
    if not body.startswith("{"):
        return text, []
    calls, dec, i = [], json.JSONDecoder(), 0
    try:
        while i < len(body):
            obj, end = dec.raw_decode(body, i)
            if not isinstance(obj, dict) or "name" not in obj or not _known(obj["name"], tools):
                return text, []
            calls.append(_call(obj["name"], obj.get("parameters", obj.get("arguments", {}))))
            i = end
            while i < len(body) and body[i] in " \n\t;":
                i += 1
    except json.JSONDecodeError:
        return text, []
    return (None, calls) if calls else (text, [])


_HERMES = re.compile(r"<tool_call>\s*(.*?)\s*</tool_call>", re.S)


def parse_hermes(text: str, tools=None) -> Tuple[Optional[str], List[Dict[str, Any]]]:
    calls = []
    for m in _HERMES.finditer(text):
        try:
            obj = json.loads(m.group(1))
        except json.JSONDecodeError:
            return text, []
        if not isinstance(obj, dict) or "name" not in obj or not _known(obj["name"], tools):
            return text, []
        calls.append(_call(obj["name"], obj.get("arguments", obj.get("parameters", {}))))
    if not calls:
        return text, []
    content = _HERMES.sub("", text).strip()
    return (content or None), calls


PARSERS = {"llama3_json": parse_llama3_json, "hermes": parse_hermes}


def get_parser(name: Optional[str]):
    if name is None:
        return None
    if name not in PARSERS:
        raise SystemExit(f"--tool-call-parser {name}: supported parsers are {sorted(PARSERS)}")
    return PARSERS[name]
