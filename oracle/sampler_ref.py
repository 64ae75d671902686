"""CPU restatement of the sampler (TEST INFRASTRUCTURE — see oracle/llama_ref.py header).

Reference behaviour being restated: vLLM `SamplingParams(temperature=0.2)` (llm/serve_llm.py:379,
521-525; top_p 1.0) and HF `do_sample=True, temperature=0.7` (llm/hf_cpu_server.py:91) draw from
softmax(logits / T); greedy parity uses torch.argmax. The draw is restated with the Gumbel-max
identity so that a seed reproduces the same token id on CPU and GPU; the HF / vLLM top-p rule
(`TopPLogitsWarper`: keep the smallest prefix of the sorted distribution whose mass reaches top_p)
is restated as {i : sum_{p_j > p_i} p_j < top_p}. Parity is unpinned by the reference (it has no
sampler tests); the distributional test in tests/ pins it against torch.multinomial statistics.
"""
from __future__ import annotations

import numpy as np

_M32 = np.uint64(0xFFFFFFFF)


def lowbias32(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64) & _M32
    x ^= x >> np.uint64(16)
    x = (x * np.uint64(0x7FEB352D)) & _M32
    x ^= x >> np.uint64(15)
    x = (x * np.uint64(0x846CA68B)) & _M32
    x ^= x >> np.uint64(16)
    return x


def stream_key(seed: int, step: int) -> int:
    lo = (seed & 0xFFFFFFFF)
    hi = (seed >> 32) & 0xFFFFFFFF
    a = int(lowbias32(np.array([(lo + 0x9E3779B9 * step) & 0xFFFFFFFF]))[0])
    b = int(lowbias32(np.array([hi ^ 0x85EBCA6B]))[0])
    return a ^ b


def gumbel(seed: int, step: int, n: int) -> np.ndarray:
    key = np.uint64(stream_key(seed, step))
    x = lowbias32(np.arange(n, dtype=np.uint64) ^ key)
    u = ((x >> np.uint64(8)).astype(np.float32) + np.float32(0.5)) * np.float32(2.0 ** -24)
    return -np.log(-np.log(u.astype(np.float32))).astype(np.float32)


def nucleus_mask(logits: np.ndarray, temperature: float, top_p: float) -> np.ndarray:
    if top_p >= 1.0:
        return np.ones_like(logits, dtype=bool)
    e = np.exp((logits.astype(np.float64) - logits.max()) / temperature)
    p = e / e.sum()
    order = np.argsort(-p, kind="stable")
    ps = p[order]
    mass_above = np.concatenate(([0.0], np.cumsum(ps)[:-1]))  # strictly-greater mass, ties broken by order
    # tokens with equal probability share the same strictly-greater mass
    first_of_run = np.concatenate(([True], ps[1:] != ps[:-1]))
    run_start = np.maximum.accumulate(np.where(first_of_run, np.arange(len(ps)), 0))
    strict = mass_above[run_start]
    keep = np.zeros(len(p), dtype=bool)
    keep[order] = strict < top_p
    return keep


def sample_scores(logits: np.ndarray, temperature: float, seed: int, step: int) -> np.ndarray:
    return (logits.astype(np.float32) * np.float32(1.0 / np.float32(temperature))) + gumbel(seed, step, len(logits))


def sample(logits: np.ndarray, temperature: float, top_p: float, seed: int, step: int) -> int:
    if not temperature > 0:
        return int(np.argmax(logits))  # first maximum, as torch.argmax
    sc = sample_scores(logits, temperature, seed, step)
    sc = np.where(nucleus_mask(logits, temperature, top_p), sc, -np.inf)
    return int(np.argmax(sc))
