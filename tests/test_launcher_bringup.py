"""The reference launcher's worker bring-up, UNMODIFIED, on the installed vLLM with the shim and the
plugin: launch.CustomExecutor spawns a local worker over its own RPC (launch.py:153-190, rpc.py,
rpc_reader.py), WorkerWrapper clobbers local_rank (launch.py:510-520), init_worker / init_device /
load_model are driven by name (launch.py:290-292), collective_rpc and execute_model return from the
output rank (launch.py:322-384), and the b200c plugin is registered inside the worker process.
CPU-only: the worker class is a stub (tests/stub_worker.py); the real Worker needs a GPU.
Runs only where the reference checkout is mounted."""
import os
import signal
import subprocess
import sys
import tempfile

import pytest

REF_SRC = "/root/reference/src"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TESTS = os.path.dirname(os.path.abspath(__file__))

SCRIPT = r'''
import json, os, socket, sys, tempfile
ROOT, REF, TESTS = %r, %r, %r
sys.path[:0] = [ROOT, REF, TESTS]
with socket.socket() as s:
    s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
os.environ["VLLM_SERVER_PORT"] = str(port)
from vllm_distributed_b200 import launch_shim
launch_shim.install()
import launch                                   # the reference's own module, unmodified
from vllm.config import DeviceConfig, ModelConfig, ParallelConfig, VllmConfig
d = tempfile.mkdtemp()
json.dump(dict(architectures=["LlamaForCausalLM"], model_type="llama", hidden_size=256,
               intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
               num_key_value_heads=4, vocab_size=1024, rms_norm_eps=1e-5,
               max_position_embeddings=512, torch_dtype="bfloat16"), open(d + "/config.json", "w"))
cfg = VllmConfig(model_config=ModelConfig(model=d, skip_tokenizer_init=True, dtype="bfloat16", max_model_len=128),
                 parallel_config=ParallelConfig(worker_cls="stub_worker.StubWorker"),
                 device_config=DeviceConfig(device="cpu"))
ex = launch.CustomExecutor(cfg)
assert ex.world_size == 1 and ex.output_rank == 0
(rank, local_rank, echo, calls), = ex.collective_rpc("ping", args=(7,))
assert (rank, echo) == (0, 7) and local_rank == 0        # -1 from the head was clobbered
assert calls == ["__init__", "init_device", "load_model"], calls
assert ex.execute_model("sched-out") == {"rank": 0, "echo": "sched-out"}
ex.check_health()
print("BRINGUP OK", flush=True)
os._exit(0)
'''


# (A single-worker bring-up ran here as a test of its own until round 2; the two tests below
# start local workers the same way and check the same calls, so SCRIPT is kept as the smallest
# recipe for trying the launcher by hand: python -c "SCRIPT % (ROOT, REF_SRC, TESTS)".)


SCRIPT_REMOTE = r'''
import json, os, signal, socket, subprocess, sys, tempfile
ROOT, REF, TESTS = %r, %r, %r
sys.path[:0] = [ROOT, REF, TESTS]
with socket.socket() as s:
    s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
os.environ["VLLM_SERVER_PORT"] = str(port)
os.environ["PYTHONPATH"] = os.pathsep.join([ROOT, TESTS])
from vllm_distributed_b200 import launch_shim
launch = launch_shim.load_launcher(os.path.join(REF, "launch.py"))   # shim + plugin + 0.22 executor protocol
from vllm.config import DeviceConfig, ModelConfig, ParallelConfig, VllmConfig
from vllm.v1.executor.abstract import Executor
assert launch.CustomExecutor.supports_pp and issubclass(launch.CustomExecutor, Executor)
d = tempfile.mkdtemp()
json.dump(dict(architectures=["LlamaForCausalLM"], model_type="llama", hidden_size=256,
               intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
               num_key_value_heads=4, vocab_size=1024, rms_norm_eps=1e-5,
               max_position_embeddings=512, torch_dtype="bfloat16"), open(d + "/config.json", "w"))
cfg = VllmConfig(model_config=ModelConfig(model=d, skip_tokenizer_init=True, dtype="bfloat16", max_model_len=128),
                 parallel_config=ParallelConfig(worker_cls="stub_worker.StubWorker", pipeline_parallel_size=2,
                                                distributed_executor_backend="external_launcher"),
                 device_config=DeviceConfig(device="cpu"))
# the second "host": the reference's own `launch.py remote <ip>` entry point, through the shim
remote = subprocess.Popen([sys.executable, "-u", "-m", "vllm_distributed_b200.launch_shim",
                           os.path.join(REF, "launch.py"), "remote", "127.0.0.1"], env=dict(os.environ))
ex = launch.CustomExecutor(cfg)          # stage 0 is filled locally, stage 1 from the remote node
assert ex.world_size == 2 and ex.output_rank == 1      # first TP worker of the LAST pipeline stage
res = ex.collective_rpc("ping", args=(7,))
assert [r[0] for r in res] == [0, 1] and all(r[3] == ["__init__", "init_device", "load_model"] for r in res), res
assert ex.execute_model("sched-out") == {"rank": 1, "echo": "sched-out"}
# vLLM 0.22 samples in a second RPC: it must be answered by the last stage too
assert ex.sample_tokens("grammar") == {"rank": 1, "sampled": "grammar"}
fut = ex.sample_tokens("g2", non_block=True)
assert fut.result(timeout=30) == {"rank": 1, "sampled": "g2"}
# `timeout` is honoured under the shim (the launcher itself ignores it, launch.py:351-384)
import time
t0 = time.monotonic()
try:
    ex.collective_rpc("nap", timeout=0.3, args=(2.0,))
    raise SystemExit("expected a TimeoutError")
except TimeoutError as e:
    assert time.monotonic() - t0 < 1.5 and "nap" in str(e), e
assert ex.collective_rpc("nap", timeout=30, args=(0.01,)) == ["rested", "rested"]   # workers are fine afterwards
ex.check_health()                                                                  # timeout=10 now means something
print("BRINGUP REMOTE OK", flush=True)
os._exit(0)
'''


@pytest.mark.skipif(not os.path.isdir(REF_SRC), reason="reference checkout not mounted")
def test_reference_executor_fills_a_pipeline_stage_from_a_remote_node():
    """PP=2: stage 0 from a local worker, stage 1 from a node started with the launcher's own
    `launch.py remote <ip>` (launch.py:543-632) over TCP + cloudpickle RPC."""
    with tempfile.TemporaryDirectory() as td:
        log = os.path.join(td, "run.log")
        with open(log, "w") as f:
            p = subprocess.Popen([sys.executable, "-u", "-c", SCRIPT_REMOTE % (ROOT, REF_SRC, TESTS)],
                                 stdout=f, stderr=subprocess.STDOUT, start_new_session=True,
                                 env=dict(os.environ, PYTHONPATH=os.pathsep.join([ROOT, TESTS])))
            try:
                rc = p.wait(timeout=300)
            finally:
                try:
                    os.killpg(p.pid, signal.SIGKILL)   # head, remote node and their workers
                except ProcessLookupError:
                    pass
        text = open(log).read()
    assert rc == 0, text[-3000:]
    assert "BRINGUP REMOTE OK" in text
    assert text.count("b200c plugin registered") >= 2     # in the local AND the remote worker
    assert "WorkerWrapper created for rpc_rank 1, local_rank 0" in text


SCRIPT_TP2_PP2 = r'''
import json, os, socket, subprocess, sys, tempfile
ROOT, REF, TESTS = %r, %r, %r
sys.path[:0] = [ROOT, REF, TESTS]
with socket.socket() as s:
    s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
os.environ["VLLM_SERVER_PORT"] = str(port)
os.environ["PYTHONPATH"] = os.pathsep.join([ROOT, TESTS])
import fake_gpu_launcher
launch = fake_gpu_launcher.load(REF, 2)                  # the head "has" two devices
from vllm.config import DeviceConfig, ModelConfig, ParallelConfig, VllmConfig
d = tempfile.mkdtemp()
json.dump(dict(architectures=["LlamaForCausalLM"], model_type="llama", hidden_size=256,
               intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
               num_key_value_heads=4, vocab_size=1024, rms_norm_eps=1e-5,
               max_position_embeddings=512, torch_dtype="bfloat16"), open(d + "/config.json", "w"))
cfg = VllmConfig(model_config=ModelConfig(model=d, skip_tokenizer_init=True, dtype="bfloat16", max_model_len=128),
                 parallel_config=ParallelConfig(worker_cls="stub_worker.StubWorker", pipeline_parallel_size=2,
                                                tensor_parallel_size=2,
                                                distributed_executor_backend="external_launcher"),
                 device_config=DeviceConfig(device="cpu"))
# the second host: a node with two devices, started through the launcher's own remote_main
remote = subprocess.Popen([sys.executable, "-u", os.path.join(TESTS, "fake_gpu_launcher.py"), REF, "2", "127.0.0.1"],
                          env=dict(os.environ))
ex = launch.CustomExecutor(cfg)
# launch.py:211-247: rank = pp_rank * tp + tp_rank; stage 0 local, stage 1 from ONE remote node
assert ex.world_size == 4 and ex.output_rank == 2        # first TP worker of the LAST stage
res = ex.collective_rpc("ping", args=(7,))
assert [r[0] for r in res] == [0, 1, 2, 3], res
assert [r[1] for r in res[:2]] == [0, 1]                  # node-local GPU index on the head ...
assert sorted(r[1] for r in res[2:]) == [0, 1]            # ... and on the remote node (-1 was clobbered)
assert all(r[3] == ["__init__", "init_device", "load_model"] for r in res), res
assert ex.execute_model("sched-out") == {"rank": 2, "echo": "sched-out"}
assert ex.sample_tokens("grammar") == {"rank": 2, "sampled": "grammar"}
print("BRINGUP TP2xPP2 OK", flush=True)
os._exit(0)
'''


@pytest.mark.skipif(not os.path.isdir(REF_SRC), reason="reference checkout not mounted")
def test_reference_executor_places_tp2_x_pp2_over_two_nodes():
    """The BASELINE topology in small: TP=2 x PP=2, stage 0 on the head's two devices, stage 1 on
    a remote node's two devices (a TP group never spans hosts, launch.py:213-240); ranks, node-local
    device indices and the output rank come out as the communicator layer expects them."""
    with tempfile.TemporaryDirectory() as td:
        log = os.path.join(td, "run.log")
        with open(log, "w") as f:
            p = subprocess.Popen([sys.executable, "-u", "-c", SCRIPT_TP2_PP2 % (ROOT, REF_SRC, TESTS)],
                                 stdout=f, stderr=subprocess.STDOUT, start_new_session=True,
                                 env=dict(os.environ, PYTHONPATH=os.pathsep.join([ROOT, TESTS])))
            try:
                rc = p.wait(timeout=300)
            finally:
                try:
                    os.killpg(p.pid, signal.SIGKILL)
                except ProcessLookupError:
                    pass
        text = open(log).read()
    assert rc == 0, text[-3000:]
    assert "BRINGUP TP2xPP2 OK" in text
    assert text.count("b200c plugin registered") >= 2
