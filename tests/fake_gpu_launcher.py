"""Test scaffolding (not product code): run the reference launcher as if this host had N CUDA
devices, so that its TP x PP stage placement can be exercised on a CPU-only box.  Only the two
names the launcher asks about devices are replaced in ITS module namespace -- `current_platform`
(is_cuda() -> True) and `cuda_device_count_stateless` (-> N); the launcher's own code is untouched
and the workers are the CPU stub (tests/stub_worker.py).

    python fake_gpu_launcher.py <reference src dir> <N> <server ip>      # a "remote node" with N devices
"""
import os
import sys


def load(ref_src: str, n_devices: int):
    from vllm_distributed_b200 import launch_shim
    launch = launch_shim.load_launcher(os.path.join(ref_src, "launch.py"))

    class AsIfCuda:
        def __init__(self, real):
            self._real = real

        def is_cuda(self):
            return True

        def __getattr__(self, name):
            return getattr(self._real, name)

    launch.current_platform = AsIfCuda(launch.current_platform)
    launch.cuda_device_count_stateless = lambda: n_devices
    return launch


if __name__ == "__main__":
    load(sys.argv[1], int(sys.argv[2])).remote_main(sys.argv[3])
