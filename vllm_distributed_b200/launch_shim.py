"""Compatibility shim: lets the reference launcher's bring-up run on the installed vLLM (0.22).

The reference (`launch.py` / `rpc.py` / `rpc_reader.py` of koush/vllm-distributed) targets a
vLLM ~0.10 whose import paths have since moved.  This module does not
contain or modify the launcher: it installs the moved names under their old import paths and
adapts the two signatures that changed, then runs the launcher file the user points it at:

    python -m vllm_distributed_b200.launch_shim /path/to/vllm-distributed/src/launch.py serve ...

What is mapped (old name used by launch.py -> where it lives in vLLM 0.22):
  vllm.utils.FlexibleArgumentParser              -> vllm.utils.argparse_utils
  vllm.utils.get_ip / get_open_port / get_distributed_init_method -> vllm.utils.network_utils
  vllm.utils.run_method                          -> vllm.v1.serial_utils
  vllm.utils.cuda_device_count_stateless         -> torch / vllm.platforms.cuda
  vllm.worker.worker_base.WorkerWrapperBase      -> vllm.v1.worker.worker_base (ctor adapter)
  vllm.entrypoints.openai.tool_parsers.ToolParserManager -> vllm.tool_parsers
  vllm.entrypoints.openai.api_server.load_log_config -> ...openai.server_utils
  vllm.entrypoints.openai.api_server.maybe_register_tokenizer_info_endpoint -> removed (no-op)
It also registers the b200c plugin in the launcher process, so every worker it spawns (local
fork or remote node) selects B200Communicator.
"""
from __future__ import annotations

import importlib
import os
import sys
import types


def _ensure_module(name: str) -> types.ModuleType:
    try:
        return importlib.import_module(name)
    except Exception:
        mod = types.ModuleType(name)
        sys.modules[name] = mod
        parent, _, child = name.rpartition(".")
        if parent:
            setattr(_ensure_module(parent), child, mod)
        return mod


def _first_attr(candidates):
    """candidates: iterable of (module, attr); returns the first that imports."""
    for mod, attr in candidates:
        try:
            return getattr(importlib.import_module(mod), attr)
        except Exception:
            continue
    return None


def install() -> dict[str, str]:
    """Install the aliases; returns {old dotted name: 'mapped' | 'stub' | 'present'}."""
    report: dict[str, str] = {}

    def alias(old_mod: str, name: str, candidates, stub=None):
        mod = _ensure_module(old_mod)
        key = f"{old_mod}.{name}"
        if hasattr(mod, name):
            report[key] = "present"
            return
        obj = _first_attr(candidates)
        if obj is None:
            obj, report[key] = stub, "stub"
        else:
            report[key] = "mapped"
        setattr(mod, name, obj)

    alias("vllm.utils", "FlexibleArgumentParser",
          [("vllm.utils.argparse_utils", "FlexibleArgumentParser")])
    for fn in ("get_ip", "get_open_port", "get_distributed_init_method"):
        alias("vllm.utils", fn, [("vllm.utils.network_utils", fn)])
    alias("vllm.utils", "run_method", [("vllm.v1.serial_utils", "run_method"),
                                       ("vllm.utils.func_utils", "run_method")])

    def _device_count() -> int:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    alias("vllm.utils", "cuda_device_count_stateless",
          [("vllm.utils.torch_utils", "cuda_device_count_stateless")], stub=_device_count)

    alias("vllm.entrypoints.openai.tool_parsers", "ToolParserManager",
          [("vllm.tool_parsers", "ToolParserManager"),
           ("vllm.tool_parsers.abstract_tool_parser", "ToolParserManager")])
    alias("vllm.entrypoints.openai.api_server", "load_log_config",
          [("vllm.entrypoints.openai.server_utils", "load_log_config")],
          stub=lambda *_a, **_k: None)
    alias("vllm.entrypoints.openai.api_server", "maybe_register_tokenizer_info_endpoint",
          [], stub=lambda *_a, **_k: None)

    # WorkerWrapperBase: the launcher calls WorkerWrapperBase.__init__(vllm_config, rpc_rank);
    # 0.22's constructor is (rpc_rank, global_rank=None).
    wb = _ensure_module("vllm.worker.worker_base")
    _ensure_module("vllm.worker")
    if not hasattr(wb, "WorkerWrapperBase"):
        new_base = _first_attr([("vllm.v1.worker.worker_base", "WorkerWrapperBase")])
        if new_base is not None:
            class WorkerWrapperBase(new_base):  # type: ignore[misc, valid-type]
                def __init__(self, vllm_config, rpc_rank: int = 0, *args, **kwargs):
                    try:
                        super().__init__(rpc_rank, *args, **kwargs)
                    except TypeError:
                        super().__init__(vllm_config, rpc_rank, *args, **kwargs)
                    self.vllm_config = vllm_config
            wb.WorkerWrapperBase = WorkerWrapperBase
            report["vllm.worker.worker_base.WorkerWrapperBase"] = "mapped"
        else:
            report["vllm.worker.worker_base.WorkerWrapperBase"] = "stub"
    else:
        report["vllm.worker.worker_base.WorkerWrapperBase"] = "present"
    return report


def adapt_executor(executor_cls) -> None:
    """Bring the launcher's CustomExecutor up to the executor protocol of vLLM 0.22
    (vllm/v1/executor/abstract.py:44-45, 221-247) without touching its source:
      * supports_pp: the engine refuses pipeline parallelism otherwise (arg_utils.py:2265-2268);
      * sample_tokens: 0.22 samples in a second RPC; like execute_model (launch.py:322-336) the
        answer must come from the first TP worker of the LAST pipeline stage (output_rank), not
        from worker 0 as the base class would take it;
      * timeout: the launcher's collective_rpc accepts `timeout` and ignores it (launch.py:351-384:
        `f.result()` without a bound -- its own check_health passes timeout=10 to no effect).
        The wrapper below takes the launcher's futures (its `non_block=True` path) and waits for
        them with the deadline, raising TimeoutError like vLLM's own executors do.
    """
    executor_cls.supports_pp = True

    orig_rpc = executor_cls.__dict__.get("collective_rpc")
    if orig_rpc is not None and not getattr(orig_rpc, "_b200c_timeout", False):
        import concurrent.futures
        import time

        def collective_rpc(self, method, timeout=None, args=(), kwargs=None, non_block=False,
                           unique_reply_rank=None):
            if timeout is None or non_block:
                return orig_rpc(self, method, timeout, args, kwargs, non_block, unique_reply_rank)
            futures = orig_rpc(self, method, timeout, args, kwargs, True, unique_reply_rank)
            deadline = time.monotonic() + timeout
            try:
                return [f.result(timeout=max(0.0, deadline - time.monotonic())) for f in futures]
            except concurrent.futures.TimeoutError as e:
                name = method if isinstance(method, str) else getattr(method, "__name__", "callable")
                raise TimeoutError(f"RPC call to {name} timed out after {timeout} s") from e
        collective_rpc._b200c_timeout = True
        executor_cls.collective_rpc = collective_rpc

    def sample_tokens(self, grammar_output, non_block: bool = False):
        (output,) = self.collective_rpc("sample_tokens", args=(grammar_output,),
                                        unique_reply_rank=self.output_rank, non_block=non_block)
        return output

    if "sample_tokens" not in executor_cls.__dict__:
        executor_cls.sample_tokens = sample_tokens


def load_launcher(script: str):
    """Install the shim, register the plugin, import the launcher file as module `launch` and
    adapt its executor.  Returns the module: `mod.CustomExecutor` can be handed to vLLM as
    `distributed_executor_backend`, `mod.main()` is the launcher's own CLI."""
    import importlib.util
    script = os.path.abspath(script)
    install()
    from . import plugin
    plugin.register()
    src_dir = os.path.dirname(script)
    if src_dir not in sys.path:
        sys.path.insert(0, src_dir)  # launch.py imports its siblings rpc / rpc_reader
    name = os.path.splitext(os.path.basename(script))[0]
    if name in sys.modules and getattr(sys.modules[name], "__file__", None) == script:
        mod = sys.modules[name]
    else:
        spec = importlib.util.spec_from_file_location(name, script)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod  # multiprocessing pickles worker_main by module name
        spec.loader.exec_module(mod)
    if hasattr(mod, "CustomExecutor"):
        adapt_executor(mod.CustomExecutor)
    return mod


def main(argv: list[str] | None = None) -> None:
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv:
        raise SystemExit("usage: python -m vllm_distributed_b200.launch_shim <launch.py> [args...]")
    script = os.path.abspath(argv[0])
    mod = load_launcher(script)
    sys.argv = [script] + argv[1:]
    mod.main()


if __name__ == "__main__":
    main()
