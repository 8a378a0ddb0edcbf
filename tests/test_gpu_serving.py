"""The serving host on a real GPU: `create_app_from_env` (the analogue of `uvicorn run-sd:app`, app/run-sd.sh:14) is
built from the reference's environment contract, runs the warm-up inference before it answers (app/run-sd.py:144-146),
and serves the reference's routes (app/run-sd.py:153-203).  B200SD_MODEL_SIZE=tiny keeps the model small; the kernels,
CUDA graph, batcher, PNG pool and (on a 2-GPU box) the Dispatcher are the real ones."""
import base64
import io
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ENV = {"APP": "sd21", "POD_NAME": "pod-0", "NODEPOOL": "b200", "MODEL_ID": "stabilityai/stable-diffusion-2-1",
       "DEVICE": "b200", "COMPILED_MODEL_ID": "none", "NUM_OF_RUNS_INF": "4", "B200SD_MODEL_SIZE": "tiny",
       "B200SD_SEED_FROM_PROMPT": "1", "B200SD_BATCH_WAIT_MS": "50"}


def _png(b64: str) -> np.ndarray:
    from PIL import Image
    return np.asarray(Image.open(io.BytesIO(base64.b64decode(b64))).convert("RGB"))


def _psnr(a: np.ndarray, b: np.ndarray) -> float:
    mse = float(((a.astype(np.float64) - b.astype(np.float64)) ** 2).mean())
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)


@pytest.fixture()
def app_env(monkeypatch, cuda, lib_built):
    for k, v in ENV.items():
        monkeypatch.setenv(k, v)
    return monkeypatch


def test_app_from_env_warms_up_and_serves(app_env):
    from fastapi.testclient import TestClient

    from b200sd import serving
    app = serving.create_app_from_env()  # returns only after the warm-up inference (readiness gate)
    c = TestClient(app)
    assert c.get("/readiness").json() == {"message": "pod-0is ready"}
    assert c.get("/health").json() == {"message": "pod-0is healthy"}
    assert "try /load/{n_runs}/infer/{n_inf}" in c.get("/").json()["message"]

    r1 = c.post("/genimage", json={"prompt": "a red fox in the snow"}).json()
    r2 = c.post("/genimage", json={"prompt": "a lighthouse at dusk"}).json()
    assert r1["prompt"] == "a red fox in the snow" and float(r1["latency"]) > 0
    a, b = _png(r1["response"]), _png(r2["response"])
    assert a.shape == b.shape == (128, 128, 3) and a.dtype == np.uint8
    assert a.std() > 5 and _psnr(a, b) < 30, "two prompts must give two different, non-constant images"
    again = _png(c.post("/genimage", json={"prompt": "a red fox in the snow"}).json()["response"])
    assert np.array_equal(a, again), "same prompt, same seed, same batch shape: same bits"

    g = c.post("/generate", json={"prompt": "a red fox in the snow", "num_inference_steps": 3}).json()
    assert _png(g["image"]).shape == (128, 128, 3) and g["execution_time"] > 0

    rep = c.get("/load/2/infer/3").json()["message"]
    assert rep.startswith("benchmark report:RESULT FOR stable_diffusion_512 on pod-0: Latency P0=") and "P99=" in rep


def _fan_out(engine_like, prompts):
    out, errs = [None] * len(prompts), []

    def one(i):
        try:
            out[i] = np.asarray(engine_like.text2img(prompts[i])[0])
        except Exception as e:  # noqa: BLE001
            errs.append(e)
    ts = [threading.Thread(target=one, args=(i,)) for i in range(len(prompts))]
    [t.start() for t in ts]
    [t.join(timeout=300) for t in ts]
    assert not errs, errs
    return out


def test_concurrent_requests_are_batched_and_each_caller_gets_its_own_image(app_env):
    from b200sd import serving
    cfg = serving.read_env()
    eng = serving.BatchingEngine(serving.build_tiny_pipeline(cfg["MODEL_ID"], "cuda:0"), 4, max_batch=8, max_wait_ms=200)
    try:
        prompts = [f"prompt number {i}" for i in range(8)]
        alone = [np.asarray(eng.text2img(p)[0]) for p in prompts[:3]]
        n_before = len(eng.batches)
        together = _fan_out(eng, prompts)
        assert max(eng.batches[n_before:]) > 1, f"requests were not batched: {eng.batches[n_before:]}"
        for i in range(3):  # other batch size, other tilings: same image up to 16-bit rounding noise
            assert _psnr(alone[i], together[i]) > 35, (i, _psnr(alone[i], together[i]))
        assert _psnr(together[0], together[1]) < 30
    finally:
        eng.close()


def test_dispatcher_over_two_gpus(app_env):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    from b200sd import serving
    app_env.setenv("B200SD_GPUS", "2")
    app_env.setenv("B200SD_MAX_BATCH", "8")
    app = serving.create_app_from_env()
    disp = None
    for r in app.routes:  # the Dispatcher is the engine the handlers closed over
        fn = getattr(r, "endpoint", None)
        for cell in (getattr(fn, "__closure__", None) or ()):
            if isinstance(cell.cell_contents, serving.Dispatcher):
                disp = cell.cell_contents
    assert disp is not None and len(disp.engines) == 2
    assert [str(e.pipe.device) for e in disp.engines] == ["cuda:0", "cuda:1"]
    prompts = [f"request {i}" for i in range(32)]
    got = _fan_out(disp, prompts)
    assert all(n > 0 for n in disp.routed), disp.routed
    assert all(len(e.batches) > 1 for e in disp.engines)  # warm-up + at least one served batch on each GPU
    ref = [np.asarray(disp.engines[0].text2img(p)[0]) for p in prompts[:4]]
    for i in range(4):
        assert _psnr(ref[i], got[i]) > 35, (i, _psnr(ref[i], got[i]))
    assert _psnr(got[0], got[1]) < 30
