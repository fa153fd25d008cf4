#!/usr/bin/env python3
"""Generate golden vectors by running the REFERENCE's own dependency code here.

TEST INFRASTRUCTURE -- not part of the product path.

The reference (koush/vllm-distributed, /root/reference/src/launch.py:290-292,
322-349, 524-538) only launches vLLM workers; the arithmetic of the hot path
lives in the un-vendored `vllm` package (0.22.0 in this image).  This script
drives THAT code, unmodified, over gloo on CPU with N processes:

  * vllm.distributed.communication_op.tensor_model_parallel_all_reduce
        (communication_op.py:12-14 -> parallel_state.py:514-541)
  * vllm.ir.ops.fused_add_rms_norm / rms_norm   (ir/ops/layernorm.py:10-58)
  * tensor_model_parallel_all_gather(dim=-1)
        (communication_op.py:17-21 -> base_device_communicator.py:184-207)
  * GroupCoordinator.send_tensor_dict / recv_tensor_dict with
        all_gather_group=tp   (parallel_state.py:833-1050)

and stores inputs + outputs (bf16 as raw uint16 bit patterns) under
tests/golden/*.npz.  It cannot travel to the GPU box (it needs nothing from
/root/reference at run time, but it is only meaningful as a one-off
generator); the fixtures it writes are committed.

Run:  python oracle/gen_golden.py            (writes tests/golden/)
"""
import os
import sys
import tempfile

import numpy as np
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden")

# (name, world, tokens, hidden, eps)
AR_CASES = [
    ("qwen05b_tp2_b1", 2, 1, 896, 1e-6),
    ("qwen05b_tp2_b5", 2, 5, 896, 1e-6),
    ("llama8b_tp2_b3", 2, 3, 4096, 1e-5),
    ("llama70b_tp4_b2", 4, 2, 8192, 1e-5),
    ("llama70b_tp8_b3", 8, 3, 8192, 1e-5),
    ("ragged_tp4_b7_h1000", 4, 7, 1000, 1e-5),
]
# (name, world, rows, cols_per_rank)
AG_CASES = [
    ("ag_tp2", 2, 3, 520),
    ("ag_tp4", 4, 2, 264),
    ("ag_tp8", 8, 5, 96),
]
# (name, tp, pp, tokens, hidden)
PP_CASES = [
    ("pp2_tp1", 1, 2, 3, 896),
    ("pp2_tp2", 2, 2, 4, 512),
]


def bits(t: torch.Tensor) -> np.ndarray:
    assert t.dtype == torch.bfloat16
    return t.contiguous().view(torch.int16).numpy().view(np.uint16).copy()


def make_inputs(seed_base, world, tokens, hidden):
    """per-rank x_r = randn(seed 1234+r).bf16, etc."""
    xs = []
    for r in range(world):
        g = torch.Generator().manual_seed(seed_base + r)
        xs.append(torch.randn(tokens, hidden, generator=g).bfloat16())
    g = torch.Generator().manual_seed(7)
    res = torch.randn(tokens, hidden, generator=g).bfloat16()
    g = torch.Generator().manual_seed(8)
    gamma = torch.randn(hidden, generator=g).bfloat16()
    return xs, res, gamma


def _worker(rank, world, tp, pp, port, tmpdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    from vllm.config import DeviceConfig, VllmConfig, set_current_vllm_config
    from vllm.distributed import parallel_state as ps
    from vllm.distributed.communication_op import (
        tensor_model_parallel_all_gather, tensor_model_parallel_all_reduce)
    import vllm.ir.ops as irops

    cfg = VllmConfig(device_config=DeviceConfig(device="cpu"))
    out = {}
    with set_current_vllm_config(cfg):
        ps.init_distributed_environment(
            world_size=world, rank=rank, local_rank=rank,
            distributed_init_method=f"tcp://127.0.0.1:{port}", backend="gloo")
        ps.initialize_model_parallel(tp, pp)
        tp_rank = ps.get_tp_group().rank_in_group
        pp_rank = ps.get_pp_group().rank_in_group

        if pp == 1:
            for name, w, tokens, hidden, eps in AR_CASES:
                if w != world:
                    continue
                xs, res, gamma = make_inputs(1234, world, tokens, hidden)
                ar = tensor_model_parallel_all_reduce(xs[tp_rank].clone())
                y, r2 = irops.fused_add_rms_norm(ar.clone(), res.clone(), gamma, eps)
                y0 = irops.rms_norm(ar.clone(), gamma, eps)
                out[f"{name}/ar"] = bits(ar)
                out[f"{name}/norm_out"] = bits(y)
                out[f"{name}/residual_out"] = bits(r2)
                out[f"{name}/rms_only"] = bits(y0)
                if rank == 0:
                    for r in range(world):
                        out[f"{name}/x{r}"] = bits(xs[r])
                    out[f"{name}/residual"] = bits(res)
                    out[f"{name}/gamma"] = bits(gamma)
                    out[f"{name}/meta"] = np.array([world, tokens, hidden], dtype=np.int64)
                    out[f"{name}/eps"] = np.array([eps], dtype=np.float64)
            for name, w, rows, cols in AG_CASES:
                if w != world:
                    continue
                g = torch.Generator().manual_seed(4321 + tp_rank)
                slab = torch.randn(rows, cols, generator=g).bfloat16()
                full = tensor_model_parallel_all_gather(slab, dim=-1)
                out[f"{name}/slab{tp_rank}"] = bits(slab)
                out[f"{name}/full"] = bits(full.contiguous())
                if rank == 0:
                    out[f"{name}/meta"] = np.array([world, rows, cols], dtype=np.int64)
        else:
            for name, ctp, cpp, tokens, hidden in PP_CASES:
                if (ctp, cpp) != (tp, pp):
                    continue
                g = torch.Generator().manual_seed(99)
                hs = torch.randn(tokens, hidden, generator=g).bfloat16()
                rs = torch.randn(tokens, hidden, generator=g).bfloat16()
                ppg = ps.get_pp_group()
                tpg = ps.get_tp_group() if tp > 1 else None
                if pp_rank == 0:
                    ppg.send_tensor_dict({"hidden_states": hs, "residual": rs, "step": 5},
                                         all_gather_group=tpg)
                    if rank == 0:
                        out[f"{name}/hidden_states_in"] = bits(hs)
                        out[f"{name}/residual_in"] = bits(rs)
                        out[f"{name}/meta"] = np.array([tp, pp, tokens, hidden], dtype=np.int64)
                else:
                    got = ppg.recv_tensor_dict(all_gather_group=tpg)
                    assert got["step"] == 5
                    out[f"{name}/hidden_states_out_tp{tp_rank}"] = bits(got["hidden_states"])
                    out[f"{name}/residual_out_tp{tp_rank}"] = bits(got["residual"])
        torch.distributed.barrier()
    np.savez(os.path.join(tmpdir, f"rank{rank}.npz"), **out)
    ps.destroy_model_parallel()
    ps.destroy_distributed_environment()


def run(world, tp, pp, port):
    with tempfile.TemporaryDirectory() as td:
        mp.spawn(_worker, args=(world, tp, pp, port, td), nprocs=world, join=True)
        per_rank = [dict(np.load(os.path.join(td, f"rank{r}.npz"))) for r in range(world)]
    return per_rank


def main():
    os.makedirs(OUT, exist_ok=True)
    port = 29650
    # --- TP groups: every rank must hold the same all-reduce / gather result.
    for world in (2, 4, 8):
        per_rank = run(world, world, 1, port); port += 1
        merged = dict(per_rank[0])
        for r in range(1, world):
            for k, v in per_rank[r].items():
                if k in merged:
                    assert np.array_equal(merged[k], v), f"rank {r} disagrees on {k}"
                else:
                    merged[k] = v
        by_case = {}
        for k, v in merged.items():
            case, field = k.split("/")
            by_case.setdefault(case, {})[field] = v
        for case, d in by_case.items():
            np.savez_compressed(os.path.join(OUT, f"{case}.npz"), **d)
            print("wrote", case, sorted(d))
    # --- PP channel
    for (name, tp, pp, _, _) in PP_CASES:
        per_rank = run(tp * pp, tp, pp, port); port += 1
        d = {}
        for pr in per_rank:
            for k, v in pr.items():
                case, field = k.split("/")
                d[field] = v
        np.savez_compressed(os.path.join(OUT, f"{name}.npz"), **d)
        print("wrote", name, sorted(d))


if __name__ == "__main__":
    sys.exit(main())
