"""Web inference app (GROUP03.pdf p.26-27).

The reference serves ``gr.Interface(fn=predict, inputs=gr.Image(type="pil"),
outputs=gr.Label(num_top_classes=3)).launch(server_name="0.0.0.0", server_port=7861)``.
``gradio`` is not installable here (no network), so:

* if ``gradio`` imports, exactly that interface is launched around ``Predictor.predict``;
* otherwise a dependency-free ``http.server`` app offers the same contract on the same port: a
  browser form to upload an image (``GET /``), ``POST /predict`` with the image bytes (raw body or
  multipart ``file`` field) answering ``{"label": top1, "confidences": [{"label", "confidence"} x3]}``
  -- the JSON shape of a Gradio ``Label`` -- and ``GET /health``.
"""
from __future__ import annotations

import io
import json
import re
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Optional, Sequence

from .inference import Predictor

_PAGE = """<!doctype html><html><head><meta charset="utf-8"><title>CIFAR-10 MobileNetV2 Classifier</title>
<style>body{font-family:sans-serif;max-width:40em;margin:2em auto}pre{background:#eee;padding:1em}</style></head>
<body><h2>CIFAR-10 MobileNetV2 Classifier</h2>
<p>Upload an image; the top-3 classes and probabilities are shown.</p>
<input type="file" id="f" accept="image/*"><button onclick="go()">Submit</button><pre id="out"></pre>
<script>async function go(){const f=document.getElementById('f').files[0];if(!f)return;
const r=await fetch('/predict',{method:'POST',body:await f.arrayBuffer()});
const j=await r.json();document.getElementById('out').textContent=
j.confidences?j.confidences.map(c=>c.label+': '+(100*c.confidence).toFixed(1)+'%').join('\\n'):JSON.stringify(j);}
</script></body></html>"""


def _extract_upload(body: bytes, content_type: str) -> bytes:
    m = re.search(r"boundary=([^;]+)", content_type or "")
    if not m:
        return body
    boundary = ("--" + m.group(1).strip('"')).encode()
    for part in body.split(boundary):
        if b"\r\n\r\n" in part and b"filename=" in part.split(b"\r\n\r\n", 1)[0]:
            return part.split(b"\r\n\r\n", 1)[1].rsplit(b"\r\n", 1)[0]
    return body


def label_json(pred: dict) -> dict:
    items = sorted(pred.items(), key=lambda kv: -kv[1])
    return {"label": items[0][0], "confidences": [{"label": k, "confidence": v} for k, v in items]}


def make_handler(predictor: Predictor):
    class Handler(BaseHTTPRequestHandler):
        def log_message(self, fmt, *args):  # keep stdout for the app's own lines
            pass

        def _send(self, code: int, body: bytes, ctype: str) -> None:
            self.send_response(code)
            self.send_header("Content-Type", ctype)
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

        def do_GET(self):
            if self.path.startswith("/health"):
                self._send(200, b'{"status": "ok"}', "application/json")
            else:
                self._send(200, _PAGE.encode(), "text/html; charset=utf-8")

        def do_POST(self):
            if not self.path.startswith("/predict"):
                self._send(404, b'{"error": "not found"}', "application/json")
                return
            try:
                from PIL import Image
                n = int(self.headers.get("Content-Length", "0"))
                data = _extract_upload(self.rfile.read(n), self.headers.get("Content-Type", ""))
                img = Image.open(io.BytesIO(data))
                out = label_json(predictor.predict(img))
                self._send(200, json.dumps(out).encode(), "application/json")
            except Exception as e:
                self._send(400, json.dumps({"error": str(e)}).encode(), "application/json")
    return Handler


def build_server(predictor: Predictor, host: str = "0.0.0.0", port: int = 7861) -> ThreadingHTTPServer:
    return ThreadingHTTPServer((host, port), make_handler(predictor))


def serve(checkpoint: Optional[str], host: str = "0.0.0.0", port: int = 7861, device: Optional[str] = None,
          force_builtin: bool = False) -> None:
    predictor = Predictor(checkpoint, device)
    if not force_builtin:
        try:
            import gradio as gr
            gr.Interface(fn=predictor.predict, inputs=gr.Image(type="pil"),
                         outputs=gr.Label(num_top_classes=3), title="CIFAR-10 MobileNetV2 Classifier",
                         description="Upload an image; the model returns the top-3 CIFAR-10 classes."
                         ).launch(server_name=host, server_port=port)
            return
        except ImportError:
            pass
    srv = build_server(predictor, host, port)
    print(f"Running on http://{host}:{port}", flush=True)
    try:
        srv.serve_forever()
    finally:
        srv.server_close()


def main(argv: Optional[Sequence[str]] = None) -> int:
    import argparse
    ap = argparse.ArgumentParser(prog="b200dp.serve", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--checkpoint", default="best_mobilenetv2_cifar10_224_mpi.pth")
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=7861)
    ap.add_argument("--device", default=None)
    ap.add_argument("--builtin", action="store_true", help="use the built-in HTTP app even if gradio is installed")
    ns = ap.parse_args(argv)
    serve(ns.checkpoint, ns.host, ns.port, ns.device, ns.builtin)
    return 0


if __name__ == "__main__":
    import sys
    sys.exit(main())
