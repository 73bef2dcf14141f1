"""Ollama REST wire format of b200llm.server (first next row).  The CPU tests drive
the HTTP layer with a stand-in generate function; the GPU test serves the real engine and checks
that the wire answer equals what the in-process shim returns for the same request."""
import json
import os
import sys
import urllib.error
import urllib.request

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "llm-based-apache-spark-optimization_b200")]

from b200llm import server  # noqa: E402


class _Err(Exception):
    def __init__(self, error, status_code):
        super().__init__(error)
        self.error, self.status_code = error, status_code


def _fake_generate(**kw):
    if kw["model"] == "nope":
        raise _Err("model 'nope' not found", 404)
    if kw["model"] == "boom":
        raise RuntimeError("engine exploded")
    text = "SELECT * FROM t WHERE " + "x = 1 AND " * 20 + "y = 2;"
    return {"model": kw["model"], "created_at": "2025-01-01T00:00:00+00:00", "response": text, "done": True,
            "done_reason": "stop", "context": [1, 2, 3], "total_duration": 10, "load_duration": 1,
            "prompt_eval_count": 7, "prompt_eval_duration": 2, "eval_count": 50, "eval_duration": 5,
            "_seen": {k: kw.get(k) for k in ("system", "prompt", "options")}}


@pytest.fixture()
def base_url():
    srv = server.make_server("127.0.0.1", 0, generate_fn=_fake_generate)
    server.serve_in_thread(srv)
    yield f"http://127.0.0.1:{srv.server_address[1]}"
    srv.shutdown()
    srv.server_close()


def _post(url, obj):
    req = urllib.request.Request(url, data=json.dumps(obj).encode(), headers={"Content-Type": "application/json"})
    return urllib.request.urlopen(req, timeout=30)


def test_root_version_tags_show(base_url):
    assert urllib.request.urlopen(base_url + "/", timeout=10).read() == b"Ollama is running"
    assert "version" in json.load(urllib.request.urlopen(base_url + "/api/version", timeout=10))
    tags = json.load(urllib.request.urlopen(base_url + "/api/tags", timeout=10))
    names = {m["name"] for m in tags["models"]}
    assert {"llama3.2", "duckdb-nsql"} <= names              # the two models the reference's apps name
    assert all(m["details"]["family"] == "llama" for m in tags["models"])
    show = json.load(_post(base_url + "/api/show", {"model": "llama3.2:latest"}))
    assert show["model_info"]["llama.block_count"] == 28 and show["model_info"]["llama.embedding_length"] == 3072
    with pytest.raises(urllib.error.HTTPError) as ei:
        _post(base_url + "/api/show", {"model": "mistral"})
    assert ei.value.code == 404 and "not found" in json.load(ei.value)["error"]


def test_generate_non_stream_passes_arguments_through(base_url):
    r = _post(base_url + "/api/generate", {"model": "duckdb-nsql", "system": "schema", "prompt": "all rows",
                                           "stream": False, "options": {"temperature": 0}})
    assert r.headers["Content-Type"].startswith("application/json")
    out = json.load(r)
    assert out["done"] is True and out["response"].startswith("SELECT") and out["eval_count"] == 50
    assert out["_seen"] == {"system": "schema", "prompt": "all rows", "options": {"temperature": 0}}


def test_generate_streams_ndjson_by_default(base_url):
    r = _post(base_url + "/api/generate", {"model": "llama3.2", "prompt": "why did the job fail"})
    assert r.headers["Content-Type"] == "application/x-ndjson"
    lines = [json.loads(ln) for ln in r.read().decode().splitlines() if ln]
    assert len(lines) >= 3 and all(not ln["done"] for ln in lines[:-1]) and lines[-1]["done"] is True
    assert lines[-1]["response"] == "" and lines[-1]["eval_count"] == 50 and lines[-1]["done_reason"] == "stop"
    whole = "".join(ln["response"] for ln in lines)
    assert whole == _fake_generate(model="llama3.2")["response"]       # the way the upstream client reassembles it


def test_generate_errors_use_the_wire_shape(base_url):
    for body, code in (({"prompt": "x"}, 400), ({"model": "nope", "prompt": "x"}, 404),
                       ({"model": "boom", "prompt": "x"}, 500)):
        with pytest.raises(urllib.error.HTTPError) as ei:
            _post(base_url + "/api/generate", body)
        assert ei.value.code == code and isinstance(json.load(ei.value)["error"], str)
    req = urllib.request.Request(base_url + "/api/generate", data=b"{not json", headers={"Content-Type": "application/json"})
    with pytest.raises(urllib.error.HTTPError) as ei:
        urllib.request.urlopen(req, timeout=10)
    assert ei.value.code == 400
    with pytest.raises(urllib.error.HTTPError) as ei:
        _post(base_url + "/api/chat", {"model": "llama3.2"})
    assert ei.value.code == 404


@pytest.mark.gpu
def test_rest_answer_equals_the_shim(monkeypatch):
    import ollama
    srv = server.make_server("127.0.0.1", 0)
    server.serve_in_thread(srv)
    try:
        url = f"http://127.0.0.1:{srv.server_address[1]}/api/generate"
        req = {"model": "tiny", "system": "You are a SQL assistant.", "prompt": "count the rows of sales",
               "options": {"temperature": 0, "num_predict": 24}}
        direct = ollama.generate(**req)
        one = json.load(_post(url, dict(req, stream=False)))
        assert one["response"] == direct.response and one["eval_count"] == direct.eval_count
        assert one["prompt_eval_count"] == direct.prompt_eval_count and one["done_reason"] == direct.done_reason
        lines = [json.loads(ln) for ln in _post(url, req).read().decode().splitlines() if ln]
        assert "".join(ln["response"] for ln in lines) == direct.response and lines[-1]["done"] is True
    finally:
        srv.shutdown()
        srv.server_close()
