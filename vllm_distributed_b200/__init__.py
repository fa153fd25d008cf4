"""B200-native TP communication hot path behind vLLM's device-communicator API.

Package layout (only what the path needs):
  csrc/            hand-written sm_100a CUDA kernels + the C ABI (include/b200c.h)
  _cabi.py         ctypes binding of libb200c.so (no fallback: missing library = error)
  comm.py          per-rank host object (arenas, IPC exchange, torch tensor front-end)
  emu.py           N ranks on one GPU in one cooperative launch (bring-up / 1-GPU tests)
  communicator.py  B200Communicator(CudaCommunicator): the drop-in for vLLM
  plugin.py        vllm.general_plugins entry point (selection seam)
  pp_channel.py    send/recv_tensor_dict over the P2P mailboxes
  launch_shim.py   lets the reference's launch.py / rpc.py bring-up run on vLLM 0.22
"""
__version__ = "0.1.0"
