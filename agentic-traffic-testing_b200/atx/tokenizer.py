"""Tokenizers behind the backend's `encode / decode / apply_chat_template` surface
(the three calls llm/serve_llm.py makes on its tokenizer: :633, :662-667, :825-831).

* HFTokenizer  — the checkpoint's own tokenizer (tokenizer.json in the model directory), including
                 its chat template, when a real checkpoint is mounted.
* ByteTokenizer — deterministic offline stand-in (no tokenizer files exist in this environment):
                 UTF-8 bytes are ids 0..255, the Llama-3 special tokens keep their public ids
                 (SURVEY.md Appendix A), every other id decodes to one printable character so that
                 random-weight models still yield valid JSON-serialisable text.
"""
from __future__ import annotations

import re
from pathlib import Path
from typing import Dict, List, Optional, Sequence

LLAMA3_SPECIALS: Dict[str, int] = {
    "<|begin_of_text|>": 128000,
    "<|end_of_text|>": 128001,
    "<|start_header_id|>": 128006,
    "<|end_header_id|>": 128007,
    "<|eom_id|>": 128008,
    "<|eot_id|>": 128009,
}


def llama3_chat_string(messages: Sequence[dict], add_generation_prompt: bool = True) -> str:
    """The Llama-3 instruct layout the reference falls back to (llm/serve_llm.py:671-678)."""
    out = ["<|begin_of_text|>"]
    for m in messages:
        out.append("<|start_header_id|>" + str(m["role"]) + "<|end_header_id|>\n\n" + str(m["content"]) + "<|eot_id|>")
    if add_generation_prompt:
        out.append("<|start_header_id|>assistant<|end_header_id|>\n\n")
    return "".join(out)


class ByteTokenizer:
    name = "byte-stub"

    def __init__(self, vocab_size: int = 128256):
        self.vocab_size = vocab_size
        if vocab_size >= 128256:
            self.specials = dict(LLAMA3_SPECIALS)
        else:  # tiny test vocabularies: park the specials at the top of the range
            self.specials = {s: vocab_size - 1 - i for i, s in enumerate(LLAMA3_SPECIALS)}
        self._by_id = {v: k for k, v in self.specials.items()}
        self._split = re.compile("(" + "|".join(re.escape(s) for s in self.specials) + ")")
        self.bos_token_id = self.specials["<|begin_of_text|>"]
        self.eos_token_id = self.specials["<|eot_id|>"]

    def encode(self, text: str, add_special_tokens: bool = False) -> List[int]:
        ids: List[int] = [self.bos_token_id] if add_special_tokens else []
        for part in self._split.split(text):
            if not part:
                continue
            sid = self.specials.get(part)
            if sid is not None:
                ids.append(sid)
            else:
                ids.extend(part.encode("utf-8"))
        return ids

    def decode(self, ids: Sequence[int], skip_special_tokens: bool = False) -> str:
        out: List[str] = []
        run = bytearray()

        def flush() -> None:
            if run:
                out.append(run.decode("utf-8", errors="replace"))
                run.clear()

        for i in ids:
            i = int(i)
            if 0 <= i < 256:
                run.append(i)
                continue
            flush()
            sp = self._by_id.get(i)
            if sp is not None:
                if not skip_special_tokens:
                    out.append(sp)
            else:
                out.append(chr(0x4E00 + i % 20000))
        flush()
        return "".join(out)

    def apply_chat_template(self, messages: Sequence[dict], tokenize: bool = False,
                            add_generation_prompt: bool = True):
        s = llama3_chat_string(messages, add_generation_prompt)
        return self.encode(s) if tokenize else s


class HFTokenizer:
    def __init__(self, path: Path):
        from transformers import AutoTokenizer

        self.tok = AutoTokenizer.from_pretrained(str(path), local_files_only=True)
        self.name = f"hf:{path}"
        self.vocab_size = len(self.tok)
        self.bos_token_id = self.tok.bos_token_id
        self.eos_token_id = self.tok.eos_token_id

    def encode(self, text: str, add_special_tokens: bool = False) -> List[int]:
        return self.tok.encode(text, add_special_tokens=add_special_tokens)

    def decode(self, ids: Sequence[int], skip_special_tokens: bool = False) -> str:
        return self.tok.decode(list(ids), skip_special_tokens=skip_special_tokens)

    def apply_chat_template(self, messages, tokenize: bool = False, add_generation_prompt: bool = True):
        return self.tok.apply_chat_template(messages, tokenize=tokenize, add_generation_prompt=add_generation_prompt)


def load_tokenizer(model_path: Optional[Path], vocab_size: int):
    if model_path is not None and any((Path(model_path) / f).exists() for f in ("tokenizer.json", "tokenizer.model")):
        return HFTokenizer(Path(model_path))
    return ByteTokenizer(vocab_size)
