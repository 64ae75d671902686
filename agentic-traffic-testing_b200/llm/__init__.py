"""Drop-in `llm` package: `python3 -m llm.serve_llm` starts the B200-native backend with the CLI,
env, HTTP, metrics and tracing contract of the reference's llm/serve_llm.py."""
