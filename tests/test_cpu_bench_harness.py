"""CPU tests of the benchmark plumbing: both arms degrade cleanly without a GPU, and the reference
harness loads the unmodified tutorial sources (hash-checked) with only the documented parameterisation."""
import json
import os
import subprocess
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(__file__))
from helpers import REPO  # noqa: E402

sys.path.insert(0, os.path.join(REPO, "baseline"))

needs_ref = pytest.mark.skipif(not os.path.exists(os.path.join(REPO, "baseline", "_ref", "MANIFEST.json")),
                               reason="baseline/_ref not installed (run baseline/install_reference.py)")


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
@pytest.mark.parametrize("impl", ["native", "reference"])
def test_bench_without_gpu_prints_one_json_line_and_exits_zero(impl):
    r = subprocess.run([sys.executable, "bench.py", "--impl", impl, "--steps", "2", "--warmup", "1"], cwd=REPO,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-1000:]
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["impl"] == impl and "unavailable" in out


@needs_ref
def test_reference_sources_are_byte_identical_to_the_manifest():
    import hashlib
    import ref_harness
    manifest = json.load(open(os.path.join(REPO, "baseline", "_ref", "MANIFEST.json")))["sha256"]
    assert len(manifest) == 17
    for rel, digest in manifest.items():
        data = open(os.path.join(REPO, "baseline", "_ref", rel), "rb").read()
        assert hashlib.sha256(data).hexdigest() == digest, rel
        if os.path.exists(os.path.join("/root/reference", rel)):
            assert data == open(os.path.join("/root/reference", rel), "rb").read(), rel
    assert ref_harness.available() is None


@needs_ref
def test_reference_module_loads_and_only_world_literals_are_parameterised():
    import ref_harness
    src = open(os.path.join(REPO, "baseline", "_ref", "master", "part2b", "part2b.py")).read()
    assert ref_harness._parameterise_world(src, 4, "part2b.py") == src          # world 4: untouched
    p8 = ref_harness._parameterise_world(src, 8, "part2b.py")
    assert "dist.new_group(list(range(8)))" in p8 and "p.grad / 8" in p8 and "[0, 1, 2, 3]" not in p8
    # everything else is character-for-character the reference
    import difflib
    changed = [l for l in difflib.unified_diff(src.splitlines(), p8.splitlines(), lineterm="", n=0)
               if l[:1] in "+-" and l[:3] not in ("+++", "---")]
    assert len(changed) == 4, changed
    # part2a_extra: the master's hand-unrolled peers 1..3 become a loop over range(1, world); the training loop
    # around the sync block and the slave script stay character-for-character the reference
    ex = (ref_harness.REF_ROOT / "master" / "part2a" / "part2a_extra.py").read_text()
    assert ref_harness._parameterise_world(ex, 4, "part2a_extra.py") == ex
    e8 = ref_harness._parameterise_world(ex, 8, "part2a_extra.py")
    compile(e8, "part2a_extra", "exec")
    assert "for peer in range(1, 8):" in e8 and "grad_mean = grad_mean / 8" in e8 and "grad_3" not in e8
    head, tail = ex.index("            grad_1 = torch.zeros_like(p.grad)"), ex.index("        optimizer.step()")
    assert e8.startswith(ex[:head]) and e8.endswith(ex[tail:])
    sl = (ref_harness.REF_ROOT / "slave" / "part2a" / "part2a_extra.py").read_text()
    assert ref_harness._parameterise_world(sl, 8, "part2a_extra.py") == sl
    mod = ref_harness.load_reference("part3", rank=0, world=1, device="cpu")
    assert mod.device == "cpu" and callable(mod.train_model) and hasattr(mod, "DDP")
    m = mod.mdl.VGG11()
    assert sum(p.numel() for p in m.parameters()) == 9_231_114
    slave = ref_harness.load_reference("part2b", rank=1, world=4, device="cpu")
    assert slave.__file__.endswith(os.path.join("slave", "part2b", "part2b.py"))


@needs_ref
def test_reference_train_model_runs_on_cpu_through_the_harness_loader():
    """The reference's own loop, driven with two synthetic batches (part1: no process group needed)."""
    import contextlib
    import io
    import ref_harness
    mod = ref_harness.load_reference("part1", rank=0, world=1, device="cpu")
    torch.manual_seed(5000)
    model = mod.mdl.VGG11()
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4)
    crit = torch.nn.CrossEntropyLoss()
    g = torch.Generator().manual_seed(0)
    batches = [(torch.randn(4, 3, 32, 32, generator=g), torch.randint(0, 10, (4,), generator=g)) for _ in range(2)]
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        mod.train_model(model, batches, opt, crit, 0)
    assert buf.getvalue().startswith("0 loss: ")


def test_reference_input_pipeline_is_torchvisions_dataset_class_with_the_tutorials_loader_settings():
    """``e2e_with_loader`` of the reference arm: torchvision's CIFAR10 ``__getitem__`` (uint8 array -> PIL ->
    transform chain of the tutorial) behind ``DataLoader(num_workers=2, pin_memory=True)``; only the pixels are
    synthetic (ref: master/part1/part1.py:65-85, master/part3/part3.py:94-102)."""
    pytest.importorskip("torchvision")
    import ref_harness
    from torchvision import datasets, transforms
    dl = ref_harness.reference_dataloader(per_rank_batch=16, rank=0, world=1, n_images=64)
    assert isinstance(dl.dataset, datasets.CIFAR10) and dl.num_workers == 2 and dl.batch_size == 16
    kinds = [type(t) for t in dl.dataset.transform.transforms]
    assert kinds == [transforms.RandomCrop, transforms.RandomHorizontalFlip, transforms.ToTensor, transforms.Normalize]
    x, y = dl.dataset[0]                                      # the class's own __getitem__, in this process
    assert x.shape == (3, 32, 32) and x.dtype == torch.float32 and 0 <= int(y) < 10
    assert len(dl) == 4
    # the sharded flavour: DistributedSampler over the same dataset, a quarter of the images per rank
    d4 = ref_harness.reference_dataloader(per_rank_batch=16, rank=1, world=4, n_images=64 * 4)
    assert type(d4.sampler).__name__ == "DistributedSampler" and len(d4) == 4
