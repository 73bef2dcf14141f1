"""Ollama REST wire compatibility for the one route the reference's client library uses.

The reference never speaks HTTP itself: `ollama.generate(...)` (Flask/app.py:102-107,160-166;
FastAPI/app.py:85-90,105-111) is the upstream Python client, which POSTs to
`$OLLAMA_HOST/api/generate` and reads back JSON (NDJSON when streaming).  A deployment that keeps
the upstream client installed can therefore switch engines by pointing OLLAMA_HOST at this server
instead of shadowing the `ollama` module (first "next" row).

Routes (shapes as in Ollama's docs/api.md):
    GET  /               -> "Ollama is running"
    GET  /api/version    -> {"version": ...}
    GET  /api/tags       -> {"models": [...]}          the geometry registry
    POST /api/show       -> details + model_info of one model, 404 if unknown
    POST /api/generate   -> the generate reply; "stream" defaults to true (application/x-ndjson:
                            response chunks with done=false, then one done=true record with the
                            counters), false gives one JSON object
Errors are `{"error": "..."}` with the status code the upstream server uses (400 bad request,
404 unknown model, 500 engine failure).  Everything goes through the same in-process engine and
scheduler as the shim (`ollama.generate` in this repo); there is no second code path.

    python -m b200llm.server --host 127.0.0.1 --port 11434
"""
from __future__ import annotations

import argparse
import datetime
import json
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Any, Callable, Dict, Optional

VERSION = "0.5.0-b200llm"
STREAM_CHUNK_CHARS = 64      # the scheduler returns whole completions; they are cut into chunks on the wire


def _default_generate(**kw: Any):
    import ollama            # this repo's shim (the package directory is on sys.path)
    return ollama.generate(**kw)


def _registry() -> Dict[str, Any]:
    from . import geometry
    return geometry.REGISTRY


def _resolve(name: str):
    from . import geometry
    return geometry.resolve(name)


def _template_of(name: str) -> str:
    """The TEMPLATE generate renders for this model: the registered Modelfile template when the model is served
    from a checkpoint with tokenizer tables, else the stand-in rendering (system, blank line, prompt)."""
    from . import runtime, tokenizer
    if name in runtime._checkpoints:
        tpl = tokenizer.TEMPLATES.get(name) or tokenizer.TEMPLATES.get(name.split(":")[0], "")
        if tpl:
            return tpl
    return "{{ if .System }}{{ .System }}\n\n{{ end }}{{ .Prompt }}"


def _details(geo) -> Dict[str, Any]:
    return {"parent_model": "", "format": "b200llm-bf16", "family": "llama", "families": ["llama"],
            "parameter_size": f"{geo.params / 1e9:.1f}B", "quantization_level": "BF16"}


def _model_entry(name: str, geo) -> Dict[str, Any]:
    return {"name": name, "model": name, "modified_at": "1970-01-01T00:00:00Z", "size": 2 * geo.params,
            "digest": f"b200llm-{geo.name}", "details": _details(geo)}


class OllamaHandler(BaseHTTPRequestHandler):
    protocol_version = "HTTP/1.1"
    server_version = "b200llm"
    generate_fn: Callable[..., Any] = staticmethod(_default_generate)

    # ---- plumbing ----
    def log_message(self, fmt: str, *args: Any) -> None:   # quiet by default
        if getattr(self.server, "verbose", False):
            super().log_message(fmt, *args)

    def _send_json(self, status: int, obj: Any) -> None:
        body = json.dumps(obj).encode()
        self.send_response(status)
        self.send_header("Content-Type", "application/json; charset=utf-8")
        self.send_header("Content-Length", str(len(body)))
        self.end_headers()
        self.wfile.write(body)

    def _send_text(self, status: int, text: str) -> None:
        body = text.encode()
        self.send_response(status)
        self.send_header("Content-Type", "text/plain; charset=utf-8")
        self.send_header("Content-Length", str(len(body)))
        self.end_headers()
        self.wfile.write(body)

    def _read_json(self) -> Optional[Dict[str, Any]]:
        try:
            n = int(self.headers.get("Content-Length") or 0)
            raw = self.rfile.read(n) if n > 0 else b"{}"
            obj = json.loads(raw or b"{}")
        except (ValueError, json.JSONDecodeError):
            self._send_json(400, {"error": "invalid JSON body"})
            return None
        if not isinstance(obj, dict):
            self._send_json(400, {"error": "request body must be a JSON object"})
            return None
        return obj

    # ---- routes ----
    def do_GET(self) -> None:  # noqa: N802 - http.server naming
        path = self.path.split("?")[0].rstrip("/") or "/"
        if path == "/":
            self._send_text(200, "Ollama is running")
        elif path == "/api/version":
            self._send_json(200, {"version": VERSION})
        elif path == "/api/tags":
            self._send_json(200, {"models": [_model_entry(n, g) for n, g in _registry().items()]})
        else:
            self._send_json(404, {"error": f"unknown route {path}"})

    do_HEAD = do_GET  # noqa: N815

    def do_POST(self) -> None:  # noqa: N802
        path = self.path.split("?")[0].rstrip("/")
        body = self._read_json()
        if body is None:
            return
        if path == "/api/show":
            self._show(body)
        elif path == "/api/generate":
            self._generate(body)
        else:
            self._send_json(404, {"error": f"unknown route {path}"})

    def _show(self, body: Dict[str, Any]) -> None:
        name = body.get("model") or body.get("name") or ""
        try:
            geo = _resolve(name)
        except KeyError:
            self._send_json(404, {"error": f"model '{name}' not found"})
            return
        info = {"general.architecture": "llama", "general.parameter_count": geo.params,
                "llama.block_count": geo.n_layers, "llama.embedding_length": geo.d,
                "llama.attention.head_count": geo.n_q, "llama.attention.head_count_kv": geo.n_kv,
                "llama.attention.key_length": geo.head_dim, "llama.feed_forward_length": geo.ffn,
                "llama.vocab_size": geo.vocab, "llama.rope.freq_base": geo.rope_theta,
                "llama.attention.layer_norm_rms_epsilon": geo.rms_eps}
        self._send_json(200, {"modelfile": f"FROM {name}\n", "parameters": "", "template": _template_of(name),
                              "details": _details(geo), "model_info": info})

    def _generate(self, body: Dict[str, Any]) -> None:
        model = body.get("model") or ""
        if not model:
            self._send_json(400, {"error": "model is required"})
            return
        stream = body.get("stream", True)
        kw = {k: body[k] for k in ("prompt", "suffix", "system", "template", "context", "raw", "format", "images",
                                   "options", "keep_alive") if k in body and body[k] is not None}
        try:
            res = type(self).generate_fn(model=model, stream=False, **kw)
        except Exception as e:  # noqa: BLE001 - mapped to the wire error shape
            status = getattr(e, "status_code", 500)
            if not isinstance(status, int) or status < 400:
                status = 500
            self._send_json(status, {"error": getattr(e, "error", None) or str(e)})
            return
        final = res.model_dump() if hasattr(res, "model_dump") else dict(res)
        if not body.get("prompt") and not body.get("system"):
            final["done_reason"] = final.get("done_reason") or "load"     # Ollama's "load the model" request
        if not stream:
            self._send_json(200, final)
            return
        # NDJSON: the text in chunks (done=false), then the record with the counters and an empty response
        self.send_response(200)
        self.send_header("Content-Type", "application/x-ndjson")
        self.send_header("Transfer-Encoding", "chunked")
        self.end_headers()

        def chunk(obj: Dict[str, Any]) -> None:
            data = (json.dumps(obj) + "\n").encode()
            self.wfile.write(f"{len(data):x}\r\n".encode() + data + b"\r\n")

        text = final.get("response") or ""
        for i in range(0, len(text), STREAM_CHUNK_CHARS):
            chunk({"model": model, "created_at": datetime.datetime.now(datetime.timezone.utc).isoformat(),
                   "response": text[i:i + STREAM_CHUNK_CHARS], "done": False})
        last = dict(final)
        last["response"] = ""
        chunk(last)
        self.wfile.write(b"0\r\n\r\n")


def make_server(host: str = "127.0.0.1", port: int = 11434, generate_fn: Optional[Callable[..., Any]] = None,
                verbose: bool = False) -> ThreadingHTTPServer:
    """Server object (not yet serving).  `generate_fn(**kwargs)` defaults to this repo's ollama.generate; tests
    inject a stand-in to exercise the wire format without a GPU."""
    handler = type("Handler", (OllamaHandler,), {})
    if generate_fn is not None:
        handler.generate_fn = staticmethod(generate_fn)
    srv = ThreadingHTTPServer((host, port), handler)
    srv.daemon_threads = True
    srv.verbose = verbose
    return srv


def serve_in_thread(srv: ThreadingHTTPServer) -> threading.Thread:
    t = threading.Thread(target=srv.serve_forever, name="b200llm-http", daemon=True)
    t.start()
    return t


def main() -> None:
    ap = argparse.ArgumentParser(description="Ollama-compatible /api/generate server on the b200llm engine")
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=11434)
    ap.add_argument("--verbose", action="store_true")
    args = ap.parse_args()
    srv = make_server(args.host, args.port, verbose=args.verbose)
    print(f"b200llm: serving the Ollama API on http://{args.host}:{srv.server_address[1]}", flush=True)
    try:
        srv.serve_forever()
    except KeyboardInterrupt:
        pass
    finally:
        srv.server_close()


if __name__ == "__main__":
    main()
