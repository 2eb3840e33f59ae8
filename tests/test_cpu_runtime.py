"""CPU tests of the native runtime pieces that need no GPU (rendezvous store), checkpoints, CLI."""
import json
import os
import subprocess
import sys
import threading

import pytest
import torch

sys.path.insert(0, os.path.dirname(__file__))
from helpers import REPO, free_port, reference_model_module  # noqa: E402


def test_library_builds_and_exports_the_c_abi():
    from cs744_pytorch_distributed_tutorial_b200.ops import _lib
    lib = _lib.lib()
    assert "sm_100a" in _lib.version()
    header = open(os.path.join(REPO, "include", "b200dp.h")).read()
    import re
    declared = set(re.findall(r"\b(b200dp_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 30
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in include/b200dp.h but not exported"


def test_store_set_get_add_barrier_across_three_ranks():
    from cs744_pytorch_distributed_tutorial_b200.parallel.comm import Store
    port, world = free_port(), 3
    results, errors = {}, []

    def rank_main(rank):
        try:
            s = Store("127.0.0.1", port, rank, world, timeout_s=20)
            s.set(f"hello/{rank}", bytes([rank]) * (rank + 1))
            got = [s.get(f"hello/{r}") for r in range(world)]   # blocks until each peer has written
            total = s.add("counter", rank + 1)
            s.barrier("t")
            final = s.add("counter", 0)
            s.barrier("t")
            results[rank] = (got, total, final)
            s.close()
        except Exception as e:  # pragma: no cover
            errors.append((rank, repr(e)))

    threads = [threading.Thread(target=rank_main, args=(r,)) for r in range(world)]
    [t.start() for t in threads]
    [t.join(60) for t in threads]
    assert not errors, errors
    for rank in range(world):
        got, _, final = results[rank]
        assert got == [bytes([r]) * (r + 1) for r in range(world)]
        assert final == 1 + 2 + 3


def test_store_get_times_out_with_a_message_naming_the_key():
    from cs744_pytorch_distributed_tutorial_b200.ops._lib import NativeError
    from cs744_pytorch_distributed_tutorial_b200.parallel.comm import Store
    s = Store("127.0.0.1", free_port(), 0, 1, timeout_s=0.3)
    with pytest.raises(NativeError, match="never/written"):
        s.get("never/written")
    s.close()


def test_store_survives_a_bogus_request_and_sweeps_finished_barriers():
    """The store port is reachable by anything that can reach the host: a request that claims a 3 GiB key must
    cost the server nothing but that connection, and finished barrier rounds must not pile up in the map."""
    import socket
    import struct
    from cs744_pytorch_distributed_tutorial_b200.ops._lib import NativeError
    from cs744_pytorch_distributed_tutorial_b200.parallel import comm as C
    port = free_port()
    s = C.Store("127.0.0.1", port, 0, 1, timeout_s=0.5)
    bogus = socket.create_connection(("127.0.0.1", port))
    bogus.sendall(struct.pack("<BI", 1, 3 << 30))
    bogus.settimeout(5)
    assert bogus.recv(16) == b""            # dropped, nothing allocated
    bogus.close()
    s.set("still/alive", b"yes")
    assert s.get("still/alive") == b"yes"
    for _ in range(6):
        s.barrier("x")
    with pytest.raises(NativeError):        # round 0's keys were erased when round 2 started
        s.get("barrier/0/go")
    s.set("red/1/0", b"a")
    s.set("red/10/0", b"b")
    s.delete_prefix("red/1/")
    assert s.get("red/10/0") == b"b"
    with pytest.raises(NativeError):
        s.get("red/1/0")
    s.close()


def test_checkpoint_roundtrip_and_reference_compat(tmp_path):
    from cs744_pytorch_distributed_tutorial_b200.checkpoint import load_checkpoint, save_checkpoint
    from cs744_pytorch_distributed_tutorial_b200.models.vgg import VGG11
    torch.manual_seed(1)
    m = VGG11()
    opt = torch.optim.SGD(m.parameters(), lr=0.1, momentum=0.9)
    path = str(tmp_path / "ck.pt")
    save_checkpoint(path, m, opt.state_dict(), step=7)
    ck = torch.load(path, weights_only=False)
    assert ck["step"] == 7 and len(ck["model"]) == 58
    assert all(not k.startswith("module.") and v.is_contiguous() for k, v in ck["model"].items())
    m2 = VGG11()
    out = load_checkpoint(path, m2)
    assert out["step"] == 7
    for (k, a), (_, b) in zip(m.state_dict().items(), m2.state_dict().items()):
        assert torch.equal(a, b), k
    ref_mod = reference_model_module()
    if ref_mod is not None:
        ref = ref_mod.VGG11()
        ref.load_state_dict(ck["model"], strict=True)          # ours -> reference
        bare = str(tmp_path / "bare.pt")
        torch.save({"module." + k: v for k, v in ref.state_dict().items()}, bare)   # raw DDP state_dict
        load_checkpoint(bare, m2)                               # reference -> ours


def test_cli_part1_prints_reference_lines_and_json():
    cmd = [sys.executable, "-m", "cs744_pytorch_distributed_tutorial_b200.cli", "--mode", "part1", "--device", "cpu",
           "--global-batch", "16", "--train-size", "64", "--test-size", "32", "--iters", "2"]
    r = subprocess.run(cmd, cwd=REPO, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = r.stdout.strip().splitlines()
    assert lines[0].startswith("0 loss: ")
    assert any(l.startswith("Test set: Average loss:") for l in lines)
    metrics = json.loads(lines[-1])
    assert metrics["mode"] == "part1" and metrics["iters"] == 2 and metrics["samples"] == 32


def test_cli_two_process_part2a_with_reference_flags():
    """The reference's launch recipe: one process per node with --master-ip/--num-nodes/--rank."""
    port = free_port()
    procs = []
    for rank in range(2):
        cmd = [sys.executable, "-m", "cs744_pytorch_distributed_tutorial_b200.cli", "--mode", "part2a",
               "--master-ip", "127.0.0.1", "--num-nodes", "2", "--rank", str(rank), "--port", str(port),
               "--device", "cpu", "--global-batch", "16", "--train-size", "64", "--test-size", "16", "--iters", "2"]
        procs.append(subprocess.Popen(cmd, cwd=REPO, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=600) for p in procs]
    for p, (so, se) in zip(procs, outs):
        assert p.returncode == 0, se[-2000:]
    assert outs[0][0].splitlines()[0].startswith("0 loss: ")
    m = json.loads(outs[0][0].strip().splitlines()[-1])
    assert m["world_size"] == 2 and m["global_batch"] == 16


@pytest.mark.parametrize("part", ["part1", "part2a", "part2a_extra", "part2b", "part3"])
def test_example_scripts_keep_the_reference_command_line(part):
    """examples/partX.py accept exactly the tutorial's flags (ref: master/part2a/part2a.py:136-143) and
    preset the mode; `--help` must not need a GPU or a rendezvous."""
    script = os.path.join(REPO, "examples", part + ".py")
    r = subprocess.run([sys.executable, script, "--help"], cwd=REPO, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    for flag in ("--master-ip", "--num-nodes", "--rank"):
        assert flag in r.stdout
    src = open(script).read()
    assert f'"--mode", "{part}"' in src


def test_example_launcher_runs_two_ranks_of_part3_on_cpu():
    env = dict(os.environ, PORT=str(free_port()))
    cmd = ["bash", os.path.join(REPO, "examples", "launch_local.sh"), "part3", "2", "--device", "cpu",
           "--global-batch", "16", "--train-size", "64", "--test-size", "16", "--iters", "2"]
    r = subprocess.run(cmd, cwd=REPO, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    m = json.loads(r.stdout.strip().splitlines()[-1])
    assert m["mode"] == "part3" and m["world_size"] == 2


def test_descriptor_handoff_over_abstract_unix_sockets():
    """csrc/fdpass.cpp: how ranks exchange the POSIX handles of the VMM heap and the multicast object.
    A pipe's write end travels through SCM_RIGHTS; what the receiver writes into its copy must come
    out of the sender's read end.  Also: a receive with nobody connecting times out instead of hanging."""
    import ctypes
    import uuid
    from cs744_pytorch_distributed_tutorial_b200.ops import _lib
    lib = _lib.lib()
    lib.b200dp_fd_listen.argtypes = [ctypes.c_char_p]
    lib.b200dp_fd_send.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    lib.b200dp_fd_recv.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.c_int]
    name = f"b200dp-test-{uuid.uuid4().hex}".encode()
    lfd = lib.b200dp_fd_listen(name)
    assert lfd >= 0
    try:
        tag = ctypes.c_int(-1)
        assert lib.b200dp_fd_recv(lfd, ctypes.byref(tag), 50) == -2          # nobody there: bounded wait
        r, w = os.pipe()
        sent = []
        t = threading.Thread(target=lambda: sent.append(lib.b200dp_fd_send(name, w, 42, 5000)))
        t.start()
        got = lib.b200dp_fd_recv(lfd, ctypes.byref(tag), 5000)
        t.join()
        assert sent == [0] and got >= 0 and got != w and tag.value == 42
        os.write(got, b"through the copy")
        assert os.read(r, 64) == b"through the copy"
        for fd in (got, r, w):
            os.close(fd)
        assert lib.b200dp_fd_send(b"b200dp-nobody-listens-here", 0, 1, 100) == -1   # connect timeout
    finally:
        os.close(lfd)


@pytest.mark.parametrize("fn,args,needle", [
    ("b200dp_conv_igemm", (None, None, None, None, None, 4, 8, 8, 48, 64, 9, 0, None), "unsupported shape"),
    ("b200dp_conv_igemm", (None, None, None, None, None, 4, 8, 8, 64, 64, 5, 0, None), "taps=5"),
    ("b200dp_conv_igemm", (None, None, None, None, None, 4, 6, 6, 64, 64, 9, 0, None), "cannot be tiled"),
    ("b200dp_conv_wgrad", (None, None, None, 4, 8, 8, 64, 96, 9, 64, 0, 0, None), "Cout=96"),
    ("b200dp_bn_stats", (None, None, 100, 12, None), "C=12"),
    ("b200dp_linear_fwd", (None, None, None, None, 4, 13, 10, None), "multiple of 8"),
    ("b200dp_linear_fwd", (None, None, None, None, 4, 16, 40, None), "O=40"),
    ("b200dp_head_fwd_bwd", (None,) * 11 + (4, 4096, 10, 1.0, None), "F=4096"),
    ("b200dp_conv_dgrad_bn_reduce", (None,) * 8 + (4, 8, 8, 64, 64, 9, 0, 1, 0, None), "are required"),
    ("b200dp_bn_relu_pool_bwd_reduce_from_out", (None, None, None, None, None, 0, 64, None), "unsupported shape"),
])
def test_entry_points_reject_bad_arguments_before_touching_the_gpu(fn, args, needle):
    """Fail loudly: shape / argument errors come back as a code plus a message naming the problem
    (b200dp_last_error), on any host, without a CUDA call having been made."""
    import cs744_pytorch_distributed_tutorial_b200.parallel.comm  # noqa: F401  (declares the comm prototypes)
    from cs744_pytorch_distributed_tutorial_b200.ops._lib import NativeError, call
    with pytest.raises(NativeError) as e:
        call(fn, *args)
    assert needle in str(e.value) and "code -22" in str(e.value)


def test_cli_runs_with_the_native_loader():
    cmd = [sys.executable, "-m", "cs744_pytorch_distributed_tutorial_b200.cli", "--mode", "part1", "--device", "cpu",
           "--global-batch", "16", "--train-size", "72", "--test-size", "40", "--iters", "4", "--loader", "native",
           "--loader-threads", "2"]
    r = subprocess.run(cmd, cwd=REPO, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    m = json.loads(r.stdout.strip().splitlines()[-1])
    assert m["iters"] == 4 and m["samples"] == 40
