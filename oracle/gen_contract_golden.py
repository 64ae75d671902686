"""Extract the machine-checkable HTTP / metrics contract from the reference's own server module
(run in the BUILD container only; /root/reference is not on the GPU box).

    python -m oracle.gen_contract_golden     # writes tests/golden/serve_contract.json
"""
from __future__ import annotations

import json
import os
import re
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference")


def main() -> None:
    os.environ.setdefault("OTEL_SDK_DISABLED", "true")
    sys.path.insert(0, str(REF))
    import llm.serve_llm as ref  # the reference module (vLLM import may fail softly; that is handled there)
    from prometheus_client import REGISTRY

    app = ref.create_app()
    routes = sorted({(r.method, r.resource.canonical) for r in app.router.routes() if r.method != "HEAD"})
    fams = {}
    for fam in REGISTRY.collect():
        if not fam.name.startswith("llm_"):
            continue
        buckets = sorted({float(s.labels["le"]) for s in fam.samples if s.name.endswith("_bucket")} - {float("inf")})
        fams[fam.name] = {"type": fam.type, "help": fam.documentation, "buckets": buckets}

    class _B(ref.AsyncVLLMBackend):
        def __init__(self):  # no engine: only the template code path is exercised
            self._model_name = "x"
            self._tokenizer = None
            self._tokenizer_ready = True

    b = _B()
    case = {"prompt": "What is 2+2?\nExplain.", "system_prompt": "You are terse."}
    case["formatted"] = b.apply_chat_template(case["prompt"], case["system_prompt"])
    case["formatted_default_system"] = b.apply_chat_template(case["prompt"], None)

    sys.path.insert(0, str(REF / "tools" / "mcp_universe"))
    import openai_proxy as proxy

    msgs = [{"role": "system", "content": "be brief"}, {"role": "user", "content": [{"type": "text", "text": "hi"}]}]
    flat = {"messages": msgs, "flattened": proxy._flatten_messages(msgs)}

    dash = (REF / "infra/monitoring/grafana").rglob("*.json")
    series = set()
    for p in dash:
        series |= set(re.findall(r"\b(llm_[a-z_]+)\b", p.read_text()))
    series = sorted(s for s in series if not s.startswith("llm_backend"))

    src = (REF / "llm/serve_llm.py").read_text()
    a = src.index("meta: Dict[str, Any] = {\n")
    meta_keys = re.findall(r'^\s+"(\w+)":', src[a:src.index("return web.json_response({\"output\"", a)], re.M)
    out = {
        "routes": routes,
        "metric_families": fams,
        "metrics_content_type": ref.CONTENT_TYPE_LATEST,
        "template_case": case,
        "flatten_case": flat,
        "grafana_series": series,
        "meta_keys": meta_keys,
        "span_names": ["llm.handle_request", "llm.time_to_first_token", "llm.generate"],
        "defaults": {"LLM_MAX_TOKENS": ref.LLM_MAX_TOKENS, "margin": ref.LLM_PROMPT_SAFETY_MARGIN_TOKENS},
    }
    dst = ROOT / "tests" / "golden" / "serve_contract.json"
    dst.write_text(json.dumps(out, indent=1, sort_keys=True))
    print(f"wrote {dst}: {len(routes)} routes, {len(fams)} metric families, {len(series)} grafana series")


if __name__ == "__main__":
    main()
