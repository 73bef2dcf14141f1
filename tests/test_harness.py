"""The benchmark harness that takes the place of the reference's Model_Evaluation_&_Comparision.py (reference lines
19-66: schema + question -> ollama.generate -> exact match / Levenshtein distance / latency; 109-144: the four-query
loop).  CPU: the text columns against a stand-in `ollama` module.  GPU: the engine table on the tiny geometry."""
import importlib.util
import os
import sys
import types

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")


def _load():
    spec = importlib.util.spec_from_file_location("b200_eval_harness", os.path.join(PKG, "Model_Evaluation_&_Comparision.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_edit_distance_is_levenshtein():
    h = _load()
    assert h.edit_distance("", "") == 0
    assert h.edit_distance("kitten", "sitting") == 3
    assert h.edit_distance("SELECT * FROM taxi;", "SELECT * FROM taxi") == 1
    assert h.edit_distance("abc", "") == 3 and h.edit_distance("", "abc") == 3
    assert h.edit_distance("flaw", "lawn") == h.edit_distance("lawn", "flaw") == 2


def test_text_columns_call_generate_the_way_the_reference_does(monkeypatch):
    """Every case goes through ollama.generate(model=, system=<schema sentence>, prompt=<question>) and .response is
    scored: a backend that answers the expected SQL gets 100 % exact match and distance 0; one that answers nothing gets
    the length of the expected text as distance."""
    h = _load()
    calls = []
    expected = dict(h.CASES)

    class Res:
        def __init__(self, text):
            self.response = text
            self.load_duration, self.prompt_eval_duration, self.eval_count, self.eval_duration = 0, 2_000_000, 7, 1_000_000

    fake = types.ModuleType("ollama")
    fake.ResponseError = type("ResponseError", (Exception,), {})

    def generate(model, system=None, prompt=None, options=None, **kw):
        calls.append((model, system, prompt, options))
        return Res(" " + expected[prompt] + "\n" if model == "good" else "")
    fake.generate = generate
    monkeypatch.setitem(sys.modules, "ollama", fake)
    good = h.text_metrics("good", greedy=True)
    assert good["exact_match_rate"] == 100.0 and good["avg_edit_distance"] == 0.0 and len(good["queries"]) == len(h.CASES)
    assert all(c[1].startswith("Here is the database schema that the SQL query will run on: CREATE TABLE taxi") for c in calls)
    assert [c[2] for c in calls] == [q for q, _ in h.CASES]
    assert all(c[3]["temperature"] == 0 for c in calls)
    assert good["queries"][0]["ttft_ms"] == 2.0 and good["queries"][0]["tok_s"] == 7000.0
    bad = h.text_metrics("bad", greedy=False)
    assert bad["exact_match_rate"] == 0.0
    assert bad["avg_edit_distance"] == sum(len(e) for _, e in h.CASES) / len(h.CASES)
    assert "temperature" not in calls[-1][3]              # the reference's own call: Ollama's default sampling

    def boom(**kw):
        raise fake.ResponseError("model 'x' not found")
    fake.generate = boom
    assert "not found" in h.text_metrics("x", greedy=True)["error"]


@pytest.mark.gpu
def test_engine_table_on_the_tiny_geometry():
    h = _load()
    out = h.engine_bench("tiny", [1, 8], 40, 8, 2, 0)
    assert out["model"] == "tiny" and out["ttft_ms_p50"] > 0
    assert [d["bs"] for d in out["decode"]] == [1, 8]
    for d in out["decode"]:
        assert d["ms_per_step"] > 0 and abs(d["tok_s"] - d["bs"] * 1000.0 / d["ms_per_step"]) < 1e-6 and d["bytes_per_step"] > 0
