"""The vLLM plugin registers on a CPU-only box: entry point discoverable from the checkout, fused op
registered with the reference's in-place contract, layer and PP patches applied, platform untouched
when it is not CUDA."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_plugin_registers_without_a_gpu():
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "from importlib.metadata import entry_points\n"
        "eps = {e.name: e.value for e in entry_points(group='vllm.general_plugins')}\n"
        "assert eps.get('b200c') == 'vllm_distributed_b200.plugin:register', eps\n"
        "import torch\n"
        "from vllm_distributed_b200 import plugin\n"
        "plugin.register(); plugin.register()   # idempotent\n"
        "op = torch.ops.vllm.b200c_allreduce_add_rmsnorm.default\n"
        "schema = str(op._schema)\n"
        "assert 'Tensor(a0!) allreduce_in' in schema and 'Tensor(a1!) residual' in schema, schema\n"
        "from vllm.model_executor.models import llama, qwen2\n"
        "assert llama.LlamaDecoderLayer.forward.__name__ == '_layer_forward'\n"
        "assert qwen2.Qwen2DecoderLayer.forward.__name__ == '_layer_forward'\n"
        "from vllm.distributed.parallel_state import GroupCoordinator\n"
        "assert GroupCoordinator.isend_tensor_dict.__module__ == 'vllm_distributed_b200.pp_channel'\n"
        "from vllm.platforms import current_platform\n"
        "assert not getattr(type(current_platform), '_b200c_patched', False)  # not CUDA here\n"
        "print('plugin ok')\n" % ROOT)
    env = dict(os.environ, PYTHONPATH=ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300,
                         cwd=ROOT, env=env)
    assert out.returncode == 0, out.stderr[-3000:]
    assert "plugin ok" in out.stdout
