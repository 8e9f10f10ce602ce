"""bench.py's reference arm runs without a GPU (the reference's own module staged in oracle/_ref when present, else the port
oracle/torch_port.py, on the host cores) and prints the contract's JSON line; the GPU arm must refuse to run without a GPU
instead of falling back."""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def test_reference_arm_prints_the_contract_line():
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["unit"] == "clips/s" and line["value"] > 0
    assert line["higher_is_better"] is True and line["n_gpus"] == 1 and line["steps"] == 1
    assert "workload" in line["config"] and "model" not in line["config"]
    staged = (ROOT / "oracle" / "_ref" / "src" / "models" / "fusion.py").is_file()
    assert line["cpu_baseline"]["kind"] == ("reference" if staged else "port")
    assert line["cpu_baseline"]["cores"] >= 1 and line["cpu_baseline"]["value"] == line["value"]
    assert "B=4096" in line["cpu_baseline"]["sample"] and line["cpu_baseline"]["b8"]["value"] > 0   # both sizes
    assert line["e2e"] == {"value": line["value"], "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert line["vs_baseline"] is None


def test_reference_arm_train_workload_and_staged_reference_equals_port():
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--workload", "train", "--batch", "32",
                          "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["value"] > 0 and "configs[3]" in line["config"]["workload"]
    sys.path.insert(0, str(ROOT))
    from oracle import ref_runner, torch_port
    if not ref_runner.available():
        return
    import torch
    from multimodalemotionrecognition_b200 import XAttnFusionHead

    torch.manual_seed(0)
    sd = {k: v.detach().clone() for k, v in XAttnFusionHead(num_classes=8, audio_n_mels=768, xattn_use_emotion_prior=True).state_dict().items()}
    ref = ref_runner.build_reference_module(sd).eval()
    g = torch.Generator().manual_seed(1)
    xv, xa = torch.randn(4, 16, 512, generator=g), torch.randn(4, 149, 768, generator=g)
    with torch.inference_mode():   # the staged, unmodified reference module and the port are the same computation
        assert (ref(xv[..., None, None], xa) - torch_port.xattn_forward(sd, xv, xa, num_heads=4)).abs().max().item() <= 1e-6


def test_gpu_arm_has_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        return
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--steps", "1", "--warmup", "0"], capture_output=True, text=True,
                         timeout=600, cwd=ROOT)
    assert out.returncode != 0
    assert not [l for l in out.stdout.splitlines() if l.startswith("{")]
