"""Built-in web app keeps the Gradio contract: image in -> top-3 {label, confidence} (GROUP03.pdf p.26-27)."""
import io
import json
import threading
import urllib.request

from PIL import Image

from b200dp.data.datasets import SyntheticCIFAR10
from b200dp.inference import Predictor
from b200dp.serve import build_server, label_json


def test_http_predict_roundtrip():
    pred = Predictor(None, device="cpu")
    srv = build_server(pred, "127.0.0.1", 0)
    port = srv.server_address[1]
    th = threading.Thread(target=srv.serve_forever, daemon=True)
    th.start()
    try:
        buf = io.BytesIO()
        Image.fromarray(SyntheticCIFAR10(1, False).images[0].numpy()).save(buf, format="PNG")
        req = urllib.request.Request(f"http://127.0.0.1:{port}/predict", data=buf.getvalue(), method="POST")
        out = json.loads(urllib.request.urlopen(req, timeout=30).read())
        assert len(out["confidences"]) == 3 and out["label"] == out["confidences"][0]["label"]
        page = urllib.request.urlopen(f"http://127.0.0.1:{port}/", timeout=30).read().decode()
        assert "CIFAR-10 MobileNetV2 Classifier" in page
        assert json.loads(urllib.request.urlopen(f"http://127.0.0.1:{port}/health", timeout=30).read())["status"] == "ok"
    finally:
        srv.shutdown(); srv.server_close()


def test_label_json_sorted():
    out = label_json({"cat": 0.2, "dog": 0.7, "ship": 0.1})
    assert out["label"] == "dog" and [c["label"] for c in out["confidences"]] == ["dog", "cat", "ship"]
