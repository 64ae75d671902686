"""OpenTelemetry wiring for the llm-backend service.

Contract kept from the reference's llm/tracing.py:14-33: `init_tracer(service_name)` installs a
global TracerProvider whose spans go through a BatchSpanProcessor to the OTLP/HTTP endpoint in
OTEL_EXPORTER_OTLP_ENDPOINT (default http://jaeger:4318/v1/traces); `get_tracer(name)` returns a
tracer after making sure that happened. Export is asynchronous, so an absent collector never
blocks a request. LLM_OTEL_EXPORT=0 keeps spans in-process only (ids still appear in `meta.otel`);
the standard OTEL_SDK_DISABLED=true turns the SDK off altogether.
"""
from __future__ import annotations

import os
import threading

from opentelemetry import trace

_DEFAULT_ENDPOINT = "http://jaeger:4318/v1/traces"
_lock = threading.Lock()
_installed_for: str | None = None


def init_tracer(service_name: str) -> None:
    global _installed_for
    with _lock:
        if _installed_for is not None:
            return
        from opentelemetry.sdk.resources import SERVICE_NAME, Resource
        from opentelemetry.sdk.trace import TracerProvider

        provider = TracerProvider(resource=Resource(attributes={SERVICE_NAME: service_name}))
        if os.environ.get("LLM_OTEL_EXPORT", "1").lower() not in ("0", "false", "no", "off"):
            try:
                from opentelemetry.exporter.otlp.proto.http.trace_exporter import OTLPSpanExporter
                from opentelemetry.sdk.trace.export import BatchSpanProcessor

                endpoint = os.environ.get("OTEL_EXPORTER_OTLP_ENDPOINT", _DEFAULT_ENDPOINT)
                provider.add_span_processor(BatchSpanProcessor(OTLPSpanExporter(endpoint=endpoint)))
            except Exception as exc:  # exporter package missing: keep tracing local, say so once
                print(f"[llm-tracing] OTLP exporter unavailable ({exc}); spans stay in-process")
        trace.set_tracer_provider(provider)
        _installed_for = service_name


def get_tracer(service_name: str):
    init_tracer(service_name)
    return trace.get_tracer(service_name)
