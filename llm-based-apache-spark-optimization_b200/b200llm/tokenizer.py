"""Stand-in tokenizer.  No tokenizer asset exists on this box (no
tokenizer.json / tokenizer.model / GGUF anywhere, no network), so text <-> ids is a
deterministic byte-level map that lets `ollama.generate(prompt=str) -> .response: str`
run end to end; a real tokenizer plugs in through the same two methods.
"""
from __future__ import annotations

from typing import List, Sequence

BOS, EOS, OFFSET = 1, 2, 3


class ByteTokenizer:
    def __init__(self, vocab: int):
        self.vocab = vocab

    def encode(self, text: str, bos: bool = True) -> List[int]:
        ids = [BOS] if bos else []
        ids.extend(min(b + OFFSET, self.vocab - 1) for b in text.encode("utf-8"))
        return ids

    def decode(self, ids: Sequence[int]) -> str:
        out = bytearray()
        for t in ids:
            if t in (BOS, EOS) or t < OFFSET:
                continue
            b = t - OFFSET
            # ids beyond the byte range (a random-init model emits any id) map to printable ASCII
            out.append(b if b < 256 else 32 + (t % 95))
        return out.decode("utf-8", "replace")

    @property
    def stop_ids(self):
        return (EOS,)


def render_prompt(system: str, prompt: str, template: str = "", raw: bool = False) -> str:
    """Stand-in for the Modelfile TEMPLATE Ollama renders from .System / .Prompt
    (SURVEY.md 3.5); the model-specific chat templates are a later row (8f-3)."""
    if raw:
        return prompt
    if template:
        return template.replace("{{ .System }}", system).replace("{{ .Prompt }}", prompt)
    return f"{system}\n\n{prompt}" if system else prompt
