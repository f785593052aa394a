import os
import subprocess
import sys

import pytest
import torch

from pytorch_distributed_b200 import harness, launcher, suspend
from pytorch_distributed_b200.suspend import SUSPEND_EXIT_CODE

from . import _mp_helpers as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TINY = ["--device", "cpu", "--epochs", "1", "--batch-size", "4", "--steps", "3", "--val-steps", "1",
        "--image-size", "32", "--num-classes", "10", "--train-samples", "64", "--val-samples", "16",
        "--width", "8", "--workers", "0"]


def _run(script, extra, env=None, timeout=600):
    e = dict(os.environ, OMP_NUM_THREADS="1", **(env or {}))
    return subprocess.run([sys.executable, os.path.join(ROOT, script), *TINY, *extra], cwd=ROOT, env=e,
                          capture_output=True, text=True, timeout=timeout)


def test_shims_import():
    import hf_env
    import hfai
    import hfai.nccl.distributed as hdist
    from hfai.nn.parallel import DistributedDataParallel
    hf_env.set_env("202111")
    assert hf_env.requested_env == "202111"
    assert hfai.client.receive_suspend_command() is False
    assert callable(hfai.go_suspend) and callable(hfai.client.go_suspend)      # quirk Q2
    assert callable(hfai.multiprocessing.spawn) and callable(hdist.init_process_group)
    assert DistributedDataParallel.__name__ == "DistributedDataParallel"
    ds = hfai.datasets.ImageNet(split="val", transform=None, num_samples=8, image_size=8)
    assert hasattr(ds, "loader") and len(ds) == 8


@pytest.mark.timeout(900)
@pytest.mark.parametrize("script,extra", [
    ("resnet_single_gpu.py", []),
    ("resnet_dp.py", ["--gpus", "2"]),
    ("restnet_ddp.py", ["--gpus", "2"]),
    ("resnet_ddp_apex.py", ["--gpus", "2"]),
])
def test_entry_points_run(script, extra, tmp_path, free_port):
    log = tmp_path / "steps.jsonl"
    if script == "restnet_ddp.py":
        extra = [*extra, "--profile", "--log-jsonl", str(log)]
    r = _run(script, [*extra, "--save-path", str(tmp_path)], env={"MASTER_PORT": str(free_port)})
    assert r.returncode == 0, r.stderr[-2000:]
    assert "cost time per epoch:" in r.stdout and "Acc1:" in r.stdout
    assert r.stdout.count("cost time per epoch:") == 1          # rank-0-only print in the DDP modes
    if script == "restnet_ddp.py":
        # --profile: per-step time (max over ranks) and one record per gradient bucket
        import json
        recs = [json.loads(l) for l in log.read_text().splitlines()]
        steps = [x for x in recs if x.get("event") == "step_time"]
        assert steps and all(x["ms"] > 0 and x["images_per_s"] > 0 for x in steps)
        assert all(x["buckets"] and all(b["bytes"] > 0 and b["us"] > 0 for b in x["buckets"]) for x in steps)


def test_suspend_writes_latest_once_and_resume_continues(tmp_path):
    cfg = H.tiny_cfg("single", save_path=str(tmp_path), max_steps=6)
    sentinel = tmp_path / "suspend.flag"
    os.environ[suspend.SENTINEL_ENV] = str(sentinel)
    try:
        sentinel.write_text("x")
        with pytest.raises(SystemExit) as ei:
            harness.main(cfg, 0)
        assert ei.value.code == SUSPEND_EXIT_CODE and not sentinel.exists()
        ck = torch.load(tmp_path / "latest.pt", map_location="cpu", weights_only=False)
        assert (ck["epoch"], ck["step"]) == (0, 1)              # 'step' is the next step to run
        steps_seen = []
        orig = harness.to_device
        harness.to_device = lambda b, d: (steps_seen.append(1), orig(b, d))[1]
        try:
            harness.main(cfg, 0)
        finally:
            harness.to_device = orig
        assert len(steps_seen) == 5 + 1                         # steps 1..5 of the epoch + 1 val batch
    finally:
        os.environ.pop(suspend.SENTINEL_ENV, None)
        suspend.clear()


@pytest.mark.timeout(600)
def test_ddp_suspend_only_rank0_saves(tmp_path, free_port):
    sentinel = tmp_path / "s.flag"
    sentinel.write_text("x")
    env = {"MASTER_PORT": str(free_port), suspend.SENTINEL_ENV: str(sentinel)}
    r = _run("restnet_ddp.py", ["--gpus", "2", "--save-path", str(tmp_path / "out")], env=env)
    assert r.returncode == 0, r.stderr[-2000:]                  # suspend tears the group down cleanly
    assert (tmp_path / "out" / "latest.pt").exists()
    assert "cost time per epoch" not in r.stdout


def test_config_defaults_are_the_reference_literals():
    """No flags == the hyper-parameter block at the top of each reference main() (ref:restnet_ddp.py:76-84 etc.)."""
    from pytorch_distributed_b200 import config
    want_dirs = {"single": "output/resnet_single", "dp": "output/resnet_dp", "ddp": "output/resnet_ddp",
                 "ddp_amp": "output/resnet_ddp_amp"}
    for mode in config.MODES:
        c = config.from_args(mode, [])
        assert (c.epochs, c.batch_size, c.num_workers, c.lr, c.momentum, c.weight_decay) == (100, 400, 4, 0.1, 0.9, 1e-4)
        assert (c.step_size, c.gamma) == (30, 0.1)
        assert str(c.save_dir).endswith(want_dirs[mode])
        assert (c.image_size, c.num_classes, c.train_samples, c.val_samples) == (224, 1000, 1_281_167, 50_000)
    assert config.from_args("ddp", ["--lr", "0.2", "--batch-size", "64"]).lr == 0.2


@pytest.mark.timeout(600)
def test_reference_style_script_runs_on_the_shims(tmp_path, free_port):
    """A script written the way the reference's are (hf_env, hfai.multiprocessing.spawn(..., bind_numa=True),
    hfai.nccl.distributed, hfai.nn.parallel.DistributedDataParallel, hfai.datasets with a torchvision pipeline, stock
    torch.optim / torch DistributedSampler) trains through this framework unchanged."""
    import hfai
    hfai.multiprocessing.spawn(H.reference_style_main, args=(free_port, str(tmp_path)), nprocs=2, bind_numa=True)
    ck = torch.load(tmp_path / "latest.pt")
    assert len(ck["model"]) == len(ResNetKeys()) and ck["acc"][3] == 8.0          # 4 steps per rank, summed over 2 ranks
    assert ck["optimizer"]["param_groups"][0]["lr"] == 0.01 and ck["scheduler"]["last_epoch"] == 1


def ResNetKeys():
    from pytorch_distributed_b200.models import ResNet
    return ResNet((1, 1, 1, 1), num_classes=10, width=8).state_dict().keys()
