"""Diagnostic for the one-shot cluster kernel: run the smoke() shape (4 emulated ranks,
[8, 4096] bf16) and, on a mismatch against the oracle, print WHERE the residual stream
differs (row, 16-byte vector index) so that a wrong cluster shape can be told from an
ordering bug.  Run it plainly and under `ncu --metrics gpu__time_duration.sum`.
"""
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from oracle import oracle  # noqa: E402
from vllm_distributed_b200 import _cabi as C  # noqa: E402
from vllm_distributed_b200.emu import EmuGroup  # noqa: E402


def main():
    world, t, h = 4, 8, 4096
    g = EmuGroup(world, "cuda:0", max_bytes=4 << 20, ag_bytes=4 << 20, p2p_bytes=1 << 20)
    g.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 3000)
    gen = torch.Generator().manual_seed(0)
    xs = [torch.randn(t, h, generator=gen).bfloat16() for _ in range(world)]
    res = torch.randn(t, h, generator=gen).bfloat16()
    gamma = torch.randn(h, generator=gen).bfloat16()
    y_ref, r_ref, ar_ref = oracle.all_reduce_add_rms_norm(xs, res, gamma, 1e-5)
    bad = 0
    for it in range(3):
        for split in (0, 1, 2, 4):
            g.set_tuning(C.TUNE_ONESHOT_SPLIT, split)
            res_outs = [torch.full((t, h), 7.0, dtype=torch.bfloat16, device="cuda") for _ in xs]
            ys, rs = g.all_reduce_add_rmsnorm([x.cuda() for x in xs], [res.cuda() for _ in xs],
                                              gamma.cuda(), 1e-5, variant=C.ONESHOT)
            torch.cuda.synchronize()
            for r in range(world):
                d = (rs[r].cpu().view(torch.int16) != r_ref.view(torch.int16))
                if d.any():
                    bad += 1
                    vec = d.view(t, h // 8, 8).any(-1)  # [rows, vectors]
                    rows = vec.any(-1).nonzero().flatten().tolist()
                    cols = vec.any(0).nonzero().flatten()
                    print(f"it {it} split {split} rank {r}: {int(d.sum())} elements differ; rows {rows}; "
                          f"vector columns {int(cols.min())}..{int(cols.max())} ({cols.numel()} of {h // 8})",
                          flush=True)
                yd = (ys[r].cpu().float() - y_ref.float()).abs().max().item()
                if yd > 0.1:
                    print(f"it {it} split {split} rank {r}: norm max abs diff {yd}", flush=True)
    try:
        g.check_health()
    except Exception as e:  # noqa: BLE001
        print("health:", e)
    print("diag done, mismatching (iteration, split, rank) cases:", bad, flush=True)


if __name__ == "__main__":
    main()
