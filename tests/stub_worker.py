class StubWorker:
    """Stands in for vllm.v1.worker.gpu_worker.Worker on a CPU-only box."""
    def __init__(self, vllm_config=None, local_rank=-1, rank=-1, distributed_init_method=None,
                 is_driver_worker=False, **kwargs):
        self.vllm_config, self.local_rank, self.rank = vllm_config, local_rank, rank
        self.distributed_init_method = distributed_init_method
        self.is_driver_worker = is_driver_worker
        self.calls = ["__init__"]

    def init_device(self):
        self.calls.append("init_device")

    def load_model(self):
        self.calls.append("load_model")

    def check_health(self):
        return None

    def ping(self, x):
        return (self.rank, self.local_rank, x, list(self.calls))

    def execute_model(self, scheduler_output):
        return {"rank": self.rank, "echo": scheduler_output}

    def sample_tokens(self, grammar_output):
        return {"rank": self.rank, "sampled": grammar_output}

    def nap(self, seconds):
        import time
        time.sleep(seconds)
        return "rested"

    def shutdown(self):
        pass
