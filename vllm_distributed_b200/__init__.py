"""B200-native TP communication hot path behind vLLM's device-communicator API.

Package layout (only what the path needs):
  csrc/            hand-written sm_100a CUDA kernels + the C ABI (include/b200c.h)
  _cabi.py         ctypes binding of libb200c.so (no fallback: missing library = error)
  comm.py          per-rank host object (arenas, IPC exchange, torch tensor front-end)
  emu.py           N ranks on one GPU in one cooperative launch (bring-up / 1-GPU tests)
  communicator.py  B200Communicator(CudaCommunicator): the drop-in for vLLM
  plugin.py        vllm.general_plugins entry point (selection seam)
  pp_channel.py    send/recv_tensor_dict over the P2P mailboxes (metadata over a /dev/shm ring)
  model_patch.py   layer-level fused op for the Llama / Qwen2 decoder classes (eager path)
  fusion_pass.py   model-agnostic all_reduce -> fused_add_rms_norm rewrite under torch.compile
  norm_provider.py libb200c's norm kernels as a vLLM IR-op provider (the TP=1 seam)
  launch_shim.py   lets the reference's launch.py / rpc.py bring-up run on vLLM 0.22
"""
__version__ = "0.1.0"


def __getattr__(name: str):
    # `CompilationConfig(inductor_passes={"post_grad_custom_post_pass":
    # "vllm_distributed_b200.b200c_fusion_pass"})` resolves the pass as an attribute of this
    # package (vllm/config/compilation.py:935-940); built lazily so that importing the package
    # never imports torch or vLLM.
    if name == "b200c_fusion_pass":
        from .fusion_pass import _LazyPass
        obj = _LazyPass()
        globals()[name] = obj
        return obj
    raise AttributeError(name)
