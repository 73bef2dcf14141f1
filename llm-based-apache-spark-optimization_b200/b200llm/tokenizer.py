"""Text <-> token ids for `ollama.generate(prompt=str) -> .response: str`, and the prompt templates.

The reference never tokenizes itself: Ollama does it inside the `ollama.generate` call
(Flask/app.py:102-107,160-166; FastAPI/app.py:85-90,105-111) with the tokenizer tables stored in
the model's GGUF blob and the TEMPLATE of its Modelfile.  Both target checkpoints are covered:

  * `SpmTokenizer`   `tokenizer.ggml.model = "llama"`: the SentencePiece BPE of duckdb-nsql (Llama-2 /
                     CodeLlama vocabulary: pieces with scores, byte fallback pieces <0xXX>, "▁" for a space);
  * `BpeTokenizer`   `tokenizer.ggml.model = "gpt2"`: the byte-level BPE of llama3.2 (ranked merges over
                     the GPT-2 byte alphabet, llama3 pre-tokenizer split, whole-word lookup before merging);
  * `ByteTokenizer`  stand-in for the seeded random-init models (no tokenizer asset exists offline).

`from_gguf_metadata` builds the right one from the metadata dict gguf_loader.py parses.  The
algorithms restate the published ones (SentencePiece BPE; GPT-2 BPE as configured for Llama 3);
tests/test_tokenizer.py checks them against the `sentencepiece` and `tokenizers` libraries on
vocabularies trained in the test (checkers only, never imported here).
"""
from __future__ import annotations

import heapq
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

BOS, EOS, OFFSET = 1, 2, 3

# GGUF tokenizer.ggml.token_type values
TT_NORMAL, TT_UNKNOWN, TT_CONTROL, TT_USER_DEFINED, TT_UNUSED, TT_BYTE = 1, 2, 3, 4, 5, 6


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


class _SpecialSplitter:
    """Cuts a text at the special tokens (control / user-defined pieces such as <|eot_id|>) so that they
    map to their single id and the text between them goes through the tokenizer proper."""

    def __init__(self, specials: Dict[str, int]):
        self.specials = {s: i for s, i in specials.items() if s}
        self._by_len = sorted(self.specials, key=len, reverse=True)
        self._firsts = {s[0] for s in self._by_len}

    def split(self, text: str) -> List[Tuple[Optional[int], str]]:
        """[(special id or None, fragment)] in order."""
        if not self._by_len:
            return [(None, text)] if text else []
        out: List[Tuple[Optional[int], str]] = []
        start = i = 0
        n = len(text)
        while i < n:
            if text[i] in self._firsts:
                hit = next((s for s in self._by_len if text.startswith(s, i)), None)   # longest first
                if hit is not None:
                    if i > start:
                        out.append((None, text[start:i]))
                    out.append((self.specials[hit], hit))
                    i += len(hit)
                    start = i
                    continue
            i += 1
        if start < n:
            out.append((None, text[start:]))
        return out


class SpmTokenizer:
    """SentencePiece BPE as stored in a GGUF blob (`tokenizer.ggml.model = "llama"`): pieces with scores.
    A fragment is cut into UTF-8 characters; the adjacent pair whose concatenation is the piece with the highest
    score (leftmost on ties) is merged until no pair is a piece; characters that are no piece fall back to the
    byte pieces <0xXX>.  A space is the piece character U+2581 and one is put in front of the text."""

    SPACE = "▁"

    def __init__(self, tokens: Sequence[str], scores: Sequence[float], token_types: Optional[Sequence[int]] = None,
                 bos_id: int = 1, eos_id: int = 2, unk_id: int = 0, add_bos: bool = True, add_space_prefix: bool = True,
                 extra_stop: Iterable[int] = ()):
        self.tokens = list(tokens)
        self.scores = list(scores)
        if token_types is not None:
            self.types = list(token_types)
        else:                                              # no type table: recognise the byte pieces by their spelling
            self.types = [TT_BYTE if len(t) == 6 and t.startswith("<0x") and t.endswith(">") else TT_NORMAL
                          for t in self.tokens]
        self.vocab = len(self.tokens)
        self.bos_id, self.eos_id, self.unk_id = bos_id, eos_id, unk_id
        self.add_bos, self.add_space_prefix = add_bos, add_space_prefix
        self._id: Dict[str, int] = {}
        for i, t in enumerate(self.tokens):
            self._id.setdefault(t, i)
        self._byte_id = [self._id.get("<0x%02X>" % b, unk_id) for b in range(256)]
        self._byte_of = {i: b for b, i in enumerate(self._byte_id) if i != unk_id and self.types[i] == TT_BYTE}
        self._special = _SpecialSplitter({t: i for i, t in enumerate(self.tokens)
                                          if self.types[i] in (TT_CONTROL, TT_USER_DEFINED)})
        self._stop = tuple(dict.fromkeys([eos_id, *extra_stop]))

    # ---- encode ----
    def _merge(self, text: str) -> List[int]:
        sym = list(text)                                   # piece text per symbol, "" once merged away
        n = len(sym)
        prev = list(range(-1, n - 1))
        nxt = list(range(1, n + 1))
        nxt[-1] = -1
        heap: List[Tuple[float, int, int, int]] = []       # (-score, left index, right index, merged length)

        def push(l: int, r: int) -> None:
            if l < 0 or r < 0:
                return
            i = self._id.get(sym[l] + sym[r])
            if i is not None:
                heapq.heappush(heap, (-self.scores[i], l, r, len(sym[l]) + len(sym[r])))

        for i in range(n - 1):
            push(i, i + 1)
        while heap:
            _, l, r, size = heapq.heappop(heap)
            if not sym[l] or not sym[r] or nxt[l] != r or len(sym[l]) + len(sym[r]) != size:
                continue                                   # one side has changed since the pair was queued
            sym[l] += sym[r]
            sym[r] = ""
            nxt[l] = nxt[r]
            if nxt[r] >= 0:
                prev[nxt[r]] = l
            push(prev[l], l)
            push(l, nxt[l])
        out: List[int] = []
        i = 0 if n else -1
        while i >= 0:
            tid = self._id.get(sym[i])
            if tid is not None:
                out.append(tid)
            else:
                out.extend(self._byte_id[b] for b in sym[i].encode("utf-8"))
            i = nxt[i]
        return out

    def encode(self, text: str, bos: bool = True, parse_special: bool = True) -> List[int]:
        ids = [self.bos_id] if (bos and self.add_bos) else []
        parts = self._special.split(text) if parse_special else ([(None, text)] if text else [])
        prev_special = True                                # the text start counts as one (BOS)
        for sid, frag in parts:
            if sid is not None:
                ids.append(sid)
                prev_special = True
                continue
            if self.add_space_prefix and prev_special:
                frag = " " + frag
            ids.extend(self._merge(frag.replace(" ", self.SPACE)))
            prev_special = False
        return ids

    # ---- decode ----
    def decode(self, ids: Sequence[int]) -> str:
        out = bytearray()
        for t in ids:
            if t < 0 or t >= self.vocab:
                continue
            ty = self.types[t]
            if ty in (TT_CONTROL, TT_UNKNOWN, TT_UNUSED):
                continue
            if t in self._byte_of:
                out.append(self._byte_of[t])
            else:
                out.extend(self.tokens[t].replace(self.SPACE, " ").encode("utf-8"))
        text = out.decode("utf-8", "replace")
        return text[1:] if self.add_space_prefix and text.startswith(" ") else text

    @property
    def stop_ids(self):
        return self._stop


def _byte_alphabet() -> Dict[int, str]:
    """GPT-2's byte -> printable unicode character map (the alphabet byte-level BPE vocabularies are spelled in)."""
    keep = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAD)) + list(range(0xAE, 0x100))
    table = {b: chr(b) for b in keep}
    extra = 0
    for b in range(256):
        if b not in table:
            table[b] = chr(256 + extra)
            extra += 1
    return table


# pre-tokenizer splits by `tokenizer.ggml.pre`
_PRE_PATTERNS = {
    # Llama 3 family ("llama-bpe", "llama3")
    "llama-bpe": r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}| ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+",
    # GPT-2
    "default": r"'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+",
}
_PRE_ALIASES = {"llama3": "llama-bpe", "llama-v3": "llama-bpe", "gpt-2": "default", "gpt2": "default"}


class BpeTokenizer:
    """Byte-level BPE as stored in a GGUF blob (`tokenizer.ggml.model = "gpt2"`): the text is split by the
    pre-tokenizer pattern, every piece spelled in the byte alphabet; a piece that is a token is taken whole
    (Llama 3's ignore_merges), otherwise its adjacent pair with the lowest merge rank (leftmost on ties) is
    merged until no pair has a rank."""

    def __init__(self, tokens: Sequence[str], merges: Sequence[str], token_types: Optional[Sequence[int]] = None,
                 bos_id: int = -1, eos_id: int = -1, add_bos: bool = True, pre: str = "llama-bpe",
                 ignore_merges: Optional[bool] = None, extra_stop: Iterable[int] = ()):
        import regex                                      # \p{L} classes and possessive-free lookahead as published
        self.tokens = list(tokens)
        self.types = list(token_types) if token_types is not None else [TT_NORMAL] * len(self.tokens)
        self.vocab = len(self.tokens)
        self.bos_id, self.eos_id, self.add_bos = bos_id, eos_id, add_bos and bos_id >= 0
        pre = _PRE_ALIASES.get(pre, pre)
        if pre not in _PRE_PATTERNS:
            raise ValueError(f"tokenizer: unknown pre-tokenizer {pre!r} (known: {sorted(_PRE_PATTERNS)})")
        self._split = regex.compile(_PRE_PATTERNS[pre])
        self.ignore_merges = (pre == "llama-bpe") if ignore_merges is None else ignore_merges
        self._id: Dict[str, int] = {}
        for i, t in enumerate(self.tokens):
            self._id.setdefault(t, i)
        self._rank: Dict[Tuple[str, str], int] = {}
        for r, m in enumerate(merges):
            a, _, b = m.partition(" ")
            self._rank.setdefault((a, b), r)
        self._b2u = _byte_alphabet()
        self._u2b = {u: b for b, u in self._b2u.items()}
        self._special = _SpecialSplitter({t: i for i, t in enumerate(self.tokens)
                                          if self.types[i] in (TT_CONTROL, TT_USER_DEFINED)})
        self._stop = tuple(dict.fromkeys([i for i in (eos_id, *extra_stop) if i >= 0]))
        self._cache: Dict[str, List[int]] = {}

    def _word(self, w: str) -> List[int]:
        got = self._cache.get(w)
        if got is not None:
            return got
        if self.ignore_merges and w in self._id:
            ids = [self._id[w]]
        else:
            sym = list(w)
            while len(sym) > 1:
                best, at = None, -1
                for i in range(len(sym) - 1):
                    r = self._rank.get((sym[i], sym[i + 1]))
                    if r is not None and (best is None or r < best):
                        best, at = r, i
                if best is None:
                    break
                sym[at:at + 2] = [sym[at] + sym[at + 1]]
            ids = []
            for s in sym:
                tid = self._id.get(s)
                if tid is not None:
                    ids.append(tid)
                else:                                      # not a token (vocabulary without the full alphabet): per byte
                    ids.extend(self._id[c] for c in s if c in self._id)
        if len(self._cache) < 65536:
            self._cache[w] = ids
        return ids

    def encode(self, text: str, bos: bool = True, parse_special: bool = True) -> List[int]:
        ids = [self.bos_id] if (bos and self.add_bos) else []
        parts = self._special.split(text) if parse_special else ([(None, text)] if text else [])
        for sid, frag in parts:
            if sid is not None:
                ids.append(sid)
                continue
            for piece in self._split.findall(frag):
                ids.extend(self._word("".join(self._b2u[b] for b in piece.encode("utf-8"))))
        return ids

    def decode(self, ids: Sequence[int]) -> str:
        out = bytearray()
        for t in ids:
            if t < 0 or t >= self.vocab or self.types[t] in (TT_CONTROL, TT_UNKNOWN, TT_UNUSED):
                continue
            if self.types[t] == TT_USER_DEFINED:
                out.extend(self.tokens[t].encode("utf-8"))
                continue
            for ch in self.tokens[t]:
                b = self._u2b.get(ch)
                if b is not None:
                    out.append(b)
                else:
                    out.extend(ch.encode("utf-8"))
        return out.decode("utf-8", "replace")

    @property
    def stop_ids(self):
        return self._stop


def from_gguf_metadata(md: Dict[str, object]):
    """The tokenizer a GGUF blob describes (gguf_loader.GGUFFile.metadata): SentencePiece BPE for
    `tokenizer.ggml.model = "llama"`, byte-level BPE for "gpt2".  Generation stops at the end-of-sequence id
    and at the end-of-turn pieces the vocabulary has (what Ollama treats as end of generation)."""
    model = md.get("tokenizer.ggml.model")
    tokens = md.get("tokenizer.ggml.tokens")
    if model is None or tokens is None:
        raise ValueError("GGUF: no tokenizer tables (tokenizer.ggml.model / tokenizer.ggml.tokens)")
    types = md.get("tokenizer.ggml.token_type")
    bos = int(md.get("tokenizer.ggml.bos_token_id", 1 if model == "llama" else -1))
    eos = int(md.get("tokenizer.ggml.eos_token_id", 2 if model == "llama" else -1))
    add_bos = bool(md.get("tokenizer.ggml.add_bos_token", True))
    index = {t: i for i, t in enumerate(tokens)}
    extra = [int(md[k]) for k in ("tokenizer.ggml.eot_token_id", "tokenizer.ggml.eom_token_id") if k in md]
    extra += [index[t] for t in ("<|eot_id|>", "<|end_of_text|>", "<|eom_id|>", "<|im_end|>", "<|end|>") if t in index]
    if model == "llama":
        scores = md.get("tokenizer.ggml.scores")
        if scores is None:
            raise ValueError("GGUF: SentencePiece vocabulary without tokenizer.ggml.scores")
        return SpmTokenizer(tokens, scores, types, bos, eos, int(md.get("tokenizer.ggml.unknown_token_id", 0)), add_bos,
                            bool(md.get("tokenizer.ggml.add_space_prefix", True)), extra)
    if model == "gpt2":
        merges = md.get("tokenizer.ggml.merges")
        if merges is None:
            raise ValueError("GGUF: BPE vocabulary without tokenizer.ggml.merges")
        return BpeTokenizer(tokens, merges, types, bos, eos, add_bos, str(md.get("tokenizer.ggml.pre", "default")),
                            extra_stop=extra)
    raise ValueError(f"GGUF: tokenizer model {model!r} is not supported (llama, gpt2)")


# ---- prompt templates ------------------------------------------------------------------------------
# Ollama renders the Modelfile TEMPLATE (Go text/template) of the model with .System and .Prompt; the
# TEMPLATE lives in Ollama's model manifest, not in the GGUF blob, so the two target models' templates
# are restated here as published in the Ollama library (no manifest exists offline to check them
# against); `template=` of the call or register_template() overrides them.
TEMPLATES: Dict[str, str] = {
    "llama3.2": ("{{ if .System }}<|start_header_id|>system<|end_header_id|>\n\n{{ .System }}<|eot_id|>{{ end }}"
                 "{{ if .Prompt }}<|start_header_id|>user<|end_header_id|>\n\n{{ .Prompt }}<|eot_id|>{{ end }}"
                 "<|start_header_id|>assistant<|end_header_id|>\n\n{{ .Response }}<|eot_id|>"),
    "duckdb-nsql": ("### Instruction:\nYour task is to generate valid duckdb SQL to answer the following question"
                    "{{ if .System }}, given a duckdb database schema.{{ end }}\n\n### Input:\n"
                    "{{ if .System }}Here is the database schema that the SQL query will run on:\n{{ .System }}{{ end }}\n\n"
                    "### Question:\n{{ .Prompt }}\n\n### Response (use duckdb shorthand if possible):\n"),
}
TEMPLATES["llama3.2:1b"] = TEMPLATES["llama3.2"]


def register_template(model: str, template: str) -> None:
    TEMPLATES[model] = template


def _tpl_tokens(template: str):
    """('text', s) / ('action', s) pieces of a Go template, with the {{- and -}} whitespace trimming applied."""
    out, i = [], 0
    trim_next = False
    while True:
        j = template.find("{{", i)
        text = template[i:] if j < 0 else template[i:j]
        if trim_next:
            text = text.lstrip()
        if j < 0:
            if text:
                out.append(("text", text))
            return out
        k = template.find("}}", j)
        if k < 0:
            raise ValueError("template: unclosed {{")
        action = template[j + 2:k]
        if action.startswith("-"):
            text, action = text.rstrip(), action[1:]
        trim_next = action.endswith("-")
        if trim_next:
            action = action[:-1]
        if text:
            out.append(("text", text))
        out.append(("action", action.strip()))
        i = k + 2


def render_template(template: str, fields: Dict[str, str]) -> str:
    """The subset of Go text/template Ollama's generate templates use: {{ .Field }}, {{ if .Field }} ..
    {{ else }} .. {{ end }}, {{- / -}} trimming.  Rendering stops at {{ .Response }} (the model continues there)."""
    toks = _tpl_tokens(template)
    out: List[str] = []
    stack: List[bool] = []        # is the branch at each nesting level live
    for kind, val in toks:
        live = all(stack)
        if kind == "text":
            if live:
                out.append(val)
            continue
        if val.startswith("if "):
            name = val[3:].strip()
            if not name.startswith("."):
                raise ValueError(f"template: unsupported condition {val!r}")
            stack.append(bool(fields.get(name[1:], "")))
        elif val == "else":
            if not stack:
                raise ValueError("template: {{ else }} without {{ if }}")
            stack[-1] = not stack[-1]
        elif val == "end":
            if not stack:
                raise ValueError("template: {{ end }} without {{ if }}")
            stack.pop()
        elif val == ".Response":
            if live:
                break
        elif val.startswith("."):
            if live:
                out.append(fields.get(val[1:], ""))
        else:
            raise ValueError(f"template: unsupported action {{{{ {val} }}}}")
    return "".join(out)


def render_prompt(system: str, prompt: str, template: str = "", raw: bool = False, model: str = "") -> str:
    """What Ollama feeds the tokenizer for generate(model=, system=, prompt=): the prompt as is
    with raw=True, else the call's `template`, else the model's registered TEMPLATE, else system + blank line +
    prompt (the seeded random-init models have no template)."""
    if raw:
        return prompt
    tpl = template or TEMPLATES.get(model, "")
    if tpl:
        return render_template(tpl, {"System": system, "Prompt": prompt})
    return f"{system}\n\n{prompt}" if system else prompt
