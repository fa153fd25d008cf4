"""The reference launcher's modules import on the installed vLLM once the shim is installed.
Runs only where the reference checkout is mounted (the build container); the GPU box has no
/root/reference and nothing on the GPU path reads it."""
import importlib
import os
import subprocess
import sys

import pytest

REF_SRC = "/root/reference/src"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(os.path.isdir(REF_SRC), reason="covered by the launcher import test below when the reference is mounted")
def test_shim_reports_every_moved_name():
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "from vllm_distributed_b200 import launch_shim\n"
        "rep = launch_shim.install()\n"
        "import vllm.utils as u\n"
        "assert callable(u.get_open_port) and callable(u.run_method) and callable(u.get_ip)\n"
        "assert u.FlexibleArgumentParser is not None\n"
        "from vllm.worker.worker_base import WorkerWrapperBase\n"
        "from vllm.entrypoints.openai.tool_parsers import ToolParserManager\n"
        "print(sorted(rep.items()))\n" % ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-3000:]
    assert "WorkerWrapperBase" in out.stdout


@pytest.mark.skipif(not os.path.isdir(REF_SRC), reason="reference checkout not mounted")
def test_reference_launcher_imports_under_shim():
    code = (
        "import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from vllm_distributed_b200 import launch_shim\n"
        "rep = launch_shim.install()\n"
        "import vllm.utils as u\n"
        "assert callable(u.get_open_port) and callable(u.run_method) and callable(u.get_ip)\n"
        "assert u.FlexibleArgumentParser is not None and 'vllm.worker.worker_base.WorkerWrapperBase' in rep\n"
        "from vllm.entrypoints.openai.tool_parsers import ToolParserManager\n"
        "import rpc, rpc_reader, launch\n"
        "from vllm.v1.executor.abstract import Executor\n"
        "assert issubclass(launch.CustomExecutor, Executor)\n"
        "w = launch.WorkerWrapper(None, 3, 1)\n"
        "assert w.local_rank == 1 and w.rpc_rank == 3\n"
        "assert callable(launch.set_peer_run_worker) and callable(launch.remote_main)\n"
        "print('launcher imported')\n" % (ROOT, REF_SRC))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-3000:]
    assert "launcher imported" in out.stdout
