"""Tokenizers and prompt templates (b200llm/tokenizer.py) against the libraries whose algorithms they restate.

Ollama tokenizes inside the reference's `ollama.generate` call (Flask/app.py:102-107,160-166) with the tables
of the GGUF blob.  The blobs are not on this box, so the vocabularies are TRAINED here, in the test, with the
`sentencepiece` and `tokenizers` libraries (checkers only); the encoders must then agree id for id.
"""
import json
import os
import random
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "llm-based-apache-spark-optimization_b200"))
from b200llm import tokenizer as T   # noqa: E402

WORDS = ("select from where group by order having count sum avg min max distinct join left inner on as and or not "
         "null table users orders price amount total customer name city country date year month the of to in is "
         "show me all rows with highest lowest average number per each for that have been which query returns "
         "error column missing syntax near unexpected token Spark DataFrame schema string integer double").split()
EXTRA = ["naïve café", "日本語のテキスト", "emoji 🙂 ok", "tabs\tand\nnewlines\r\n", "  leading  double  spaces ", "x=1;y=22;z=333",
         "it's we're I'll DON'T", "SELECT COUNT(*) FROM t WHERE a >= 10.5;", "12345678 9", "a_b-c/d\\e", "ÀÉÎõü ßøæ", " nbsp"]


def _corpus(n_lines=3000, seed=7):
    rnd = random.Random(seed)
    lines = []
    for _ in range(n_lines):
        k = rnd.randint(3, 14)
        ws = [rnd.choice(WORDS) for _ in range(k)]
        if rnd.random() < 0.3:
            ws[rnd.randrange(k)] = str(rnd.randint(0, 99999))
        if rnd.random() < 0.2:
            ws[rnd.randrange(k)] = ws[rnd.randrange(k)].upper()
        lines.append(" ".join(ws) + rnd.choice([".", "?", ";", "", ","]))
    return lines + EXTRA * 3


def _samples(seed=11):
    rnd = random.Random(seed)
    out = list(EXTRA) + ["", " ", "a", "\n", "Show me the total price per customer, ordered by the highest amount."]
    for _ in range(120):
        out.append(" ".join(rnd.choice(WORDS + ["zzqx", "Ünï", "42", "SELECT", "(", ")", "'s"]) for _ in range(rnd.randint(1, 20))))
    return out


@pytest.fixture(scope="module")
def spm(tmp_path_factory):
    import sentencepiece as sp
    d = tmp_path_factory.mktemp("spm")
    corpus = d / "corpus.txt"
    corpus.write_text("\n".join(_corpus()), encoding="utf-8")
    sp.SentencePieceTrainer.train(input=str(corpus), model_prefix=str(d / "m"), vocab_size=420, model_type="bpe",
                                  byte_fallback=True, normalization_rule_name="identity", remove_extra_whitespaces=False,
                                  add_dummy_prefix=True, character_coverage=0.995, split_digits=True,
                                  allow_whitespace_only_pieces=True, minloglevel=2)
    proc = sp.SentencePieceProcessor(model_file=str(d / "m.model"))
    n = proc.GetPieceSize()
    types = []
    for i in range(n):
        types.append(T.TT_UNKNOWN if proc.IsUnknown(i) else T.TT_CONTROL if proc.IsControl(i) else
                     T.TT_BYTE if proc.IsByte(i) else T.TT_UNUSED if proc.IsUnused(i) else T.TT_NORMAL)
    md = {"tokenizer.ggml.model": "llama", "tokenizer.ggml.tokens": [proc.IdToPiece(i) for i in range(n)],
          "tokenizer.ggml.scores": [proc.GetScore(i) for i in range(n)], "tokenizer.ggml.token_type": types,
          "tokenizer.ggml.bos_token_id": proc.bos_id(), "tokenizer.ggml.eos_token_id": proc.eos_id(),
          "tokenizer.ggml.unknown_token_id": proc.unk_id()}
    return proc, T.from_gguf_metadata(md)


def test_sentencepiece_bpe_ids_equal_the_library(spm):
    proc, tok = spm
    assert isinstance(tok, T.SpmTokenizer)
    for text in _samples():
        want = proc.EncodeAsIds(text)
        assert tok.encode(text, bos=False) == want, text
        assert tok.encode(text)[0] == proc.bos_id()
        assert tok.decode(want) == proc.DecodeIds(want), text
    assert tok.stop_ids == (proc.eos_id(),)


def test_sentencepiece_special_pieces_and_byte_fallback(spm):
    proc, tok = spm
    ids = tok.encode("a</s>b", bos=False)
    assert proc.eos_id() in ids                                    # the control piece is one id
    assert ids == tok.encode("a", bos=False) + [proc.eos_id()] + tok.encode("b", bos=False)
    assert proc.eos_id() not in tok.encode("a</s>b", bos=False, parse_special=False)
    snow = tok.encode("☃", bos=False)                              # not in the vocabulary: its three UTF-8 bytes
    assert [tok.tokens[i] for i in snow[-3:]] == ["<0xE2>", "<0x98>", "<0x83>"]
    assert tok.decode(snow) == "☃"


def train_bpe():
    """(library tokenizer, ours from the same tables, vocabulary, the tables as GGUF metadata)"""
    from tokenizers import Regex, Tokenizer, decoders, models, pre_tokenizers, trainers
    pat = T._PRE_PATTERNS["llama-bpe"]
    tk = Tokenizer(models.BPE())
    tk.pre_tokenizer = pre_tokenizers.Sequence([pre_tokenizers.Split(Regex(pat), behavior="isolated", invert=False),
                                                pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=False)])
    tk.decoder = decoders.ByteLevel()
    specials = ["<|begin_of_text|>", "<|end_of_text|>", "<|start_header_id|>", "<|end_header_id|>", "<|eot_id|>"]
    trainer = trainers.BpeTrainer(vocab_size=700, special_tokens=specials, initial_alphabet=pre_tokenizers.ByteLevel.alphabet(),
                                  show_progress=False)
    tk.train_from_iterator(_corpus(), trainer)
    spec = json.loads(tk.to_str())
    spec["model"]["ignore_merges"] = True                          # as in Llama 3's tokenizer.json
    tk = Tokenizer.from_str(json.dumps(spec))
    vocab = tk.get_vocab()
    tokens = [None] * len(vocab)
    for t, i in vocab.items():
        tokens[i] = t
    merges = [m if isinstance(m, str) else " ".join(m) for m in spec["model"]["merges"]]
    types = [T.TT_CONTROL if t in specials else T.TT_NORMAL for t in tokens]
    md = {"tokenizer.ggml.model": "gpt2", "tokenizer.ggml.pre": "llama-bpe", "tokenizer.ggml.tokens": tokens,
          "tokenizer.ggml.merges": merges, "tokenizer.ggml.token_type": types,
          "tokenizer.ggml.bos_token_id": vocab["<|begin_of_text|>"], "tokenizer.ggml.eos_token_id": vocab["<|end_of_text|>"]}
    return tk, T.from_gguf_metadata(md), vocab, md


@pytest.fixture(scope="module")
def bpe():
    return train_bpe()[:3]


def test_byte_level_bpe_ids_equal_the_library(bpe):
    tk, tok, vocab = bpe
    assert isinstance(tok, T.BpeTokenizer) and tok.ignore_merges
    for text in _samples():
        want = tk.encode(text, add_special_tokens=False).ids
        assert tok.encode(text, bos=False) == want, text
        assert tok.decode(want) == tk.decode(want, skip_special_tokens=True), text
        assert tok.decode(want) == text
    assert tok.encode("x")[0] == vocab["<|begin_of_text|>"]
    assert set(tok.stop_ids) == {vocab["<|end_of_text|>"], vocab["<|eot_id|>"]}


def test_rendered_llama3_prompt_tokenizes_with_its_special_tokens(bpe):
    tk, tok, vocab = bpe
    text = T.render_prompt("You explain errors.", "Why did it fail?", model="llama3.2")
    assert text.endswith("<|start_header_id|>assistant<|end_header_id|>\n\n")
    ids = tok.encode(text)
    want = tk.encode(text, add_special_tokens=False).ids            # the library parses its special tokens too
    assert ids[1:] == want
    assert ids.count(vocab["<|eot_id|>"]) == 2 and ids.count(vocab["<|start_header_id|>"]) == 3


def test_templates():
    r = T.render_prompt
    assert r("", "p") == "p" and r("s", "p") == "s\n\np" and r("s", "p", raw=True) == "p"
    assert r("s", "p", "[{{ .System }}|{{ .Prompt }}]") == "[s|p]"
    tpl = "A {{- if .System }} <{{ .System }}> {{- else }} none{{ end }} B{{ if .Prompt }} {{ .Prompt }}{{ end }}{{ .Response }} tail"
    assert r("s", "p", tpl) == "A <s> B p"
    assert r("", "", tpl) == "A none B"
    nsql = r("CREATE TABLE t(a INT)", "count rows", model="duckdb-nsql")
    assert "Here is the database schema that the SQL query will run on:\nCREATE TABLE t(a INT)\n\n### Question:\ncount rows\n\n### Response" in nsql
    assert "schema" not in r("", "count rows", model="duckdb-nsql").split("### Question")[0].replace("### Input", "")
    with pytest.raises(ValueError):
        r("s", "p", "{{ range .Messages }}x{{ end }}")
    with pytest.raises(ValueError):
        r("s", "p", "{{ end }}")


def test_from_gguf_metadata_errors():
    with pytest.raises(ValueError):
        T.from_gguf_metadata({})
    with pytest.raises(ValueError):
        T.from_gguf_metadata({"tokenizer.ggml.model": "t5", "tokenizer.ggml.tokens": ["a"]})
    with pytest.raises(ValueError):
        T.from_gguf_metadata({"tokenizer.ggml.model": "llama", "tokenizer.ggml.tokens": ["a"]})
    with pytest.raises(ValueError):
        T.from_gguf_metadata({"tokenizer.ggml.model": "gpt2", "tokenizer.ggml.tokens": ["a"]})


def _write_tokenizer_gguf(path, md):
    """The tokenizer.ggml.* tables as llama.cpp's converter writes them (gguf-py is the writer, our own reader reads)."""
    import gguf
    w = gguf.GGUFWriter(str(path), "llama")
    w.add_tokenizer_model(md["tokenizer.ggml.model"])
    if "tokenizer.ggml.pre" in md:
        w.add_tokenizer_pre(md["tokenizer.ggml.pre"])
    w.add_token_list(md["tokenizer.ggml.tokens"])
    w.add_token_types(md["tokenizer.ggml.token_type"])
    if "tokenizer.ggml.scores" in md:
        w.add_token_scores(md["tokenizer.ggml.scores"])
    if "tokenizer.ggml.merges" in md:
        w.add_token_merges(md["tokenizer.ggml.merges"])
    w.add_bos_token_id(md["tokenizer.ggml.bos_token_id"])
    w.add_eos_token_id(md["tokenizer.ggml.eos_token_id"])
    w.add_tensor("token_embd.weight", np.zeros((4, 4), np.float32))
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()


def test_tokenizers_survive_the_gguf_file(spm, bpe, tmp_path):
    from b200llm import gguf_loader
    for kind, tok in (("spm", spm[1]), ("bpe", bpe[1])):
        md = {"tokenizer.ggml.model": "llama" if kind == "spm" else "gpt2", "tokenizer.ggml.tokens": tok.tokens,
              "tokenizer.ggml.token_type": tok.types, "tokenizer.ggml.bos_token_id": tok.bos_id,
              "tokenizer.ggml.eos_token_id": tok.eos_id}
        if kind == "spm":
            md["tokenizer.ggml.scores"] = tok.scores
        else:
            md["tokenizer.ggml.pre"] = "llama-bpe"
            md["tokenizer.ggml.merges"] = [f"{a} {b}" for (a, b), _ in sorted(tok._rank.items(), key=lambda kv: kv[1])]
        path = tmp_path / f"{kind}.gguf"
        _write_tokenizer_gguf(path, md)
        got = gguf_loader.load_tokenizer(str(path))
        assert type(got) is type(tok) and got.stop_ids == tok.stop_ids
        for text in _samples()[:40]:
            assert got.encode(text) == tok.encode(text), text
    # a weights-only file has no tokenizer model: the caller keeps its stand-in
    import gguf
    w = gguf.GGUFWriter(str(tmp_path / "plain.gguf"), "llama")
    w.add_token_list(["a", "b"])
    w.add_tensor("token_embd.weight", np.zeros((4, 4), np.float32))
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    assert gguf_loader.load_tokenizer(str(tmp_path / "plain.gguf")) is None
