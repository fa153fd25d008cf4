"""Protocol stress on one GPU (emulated ranks): hundreds of back-to-back calls with random sizes,
variants and group sizes -- region rotation, re-arming, ticket bookkeeping and the flag protocols
must never leak state from one call into the next.  The checker is a rank-order fp32 sum computed
with torch on the GPU (the same arithmetic as the CPU oracle, tests/test_gpu_parity.py pins that)."""
import random

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref_sum(xs):
    acc = xs[0].float()
    for x in xs[1:]:
        acc = acc + x.float()
    return acc.to(xs[0].dtype)


@pytest.mark.parametrize("world", [2, 3, 5, 6, 7, 8])
def test_random_call_sequence(world):
    assert torch.cuda.is_available()
    from vllm_distributed_b200 import _cabi as C
    from vllm_distributed_b200.emu import EmuGroup
    g = EmuGroup(world, "cuda:0", max_bytes=2 << 20, ag_bytes=2 << 20, p2p_bytes=256 << 10)
    g.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 3000)
    rng = random.Random(1000 + world)
    gen = torch.Generator(device="cuda").manual_seed(world)
    hidden_choices = [256, 896, 1000, 2048, 4096, 8192]
    bad = 0
    for it in range(150):
        kind = rng.choice(["ar", "ar", "fused", "fused", "gather", "p2p"])
        if kind == "ar":
            numel = 8 * rng.randint(1, 30000)
            xs = [torch.randn(numel, device="cuda", generator=gen).bfloat16() for _ in range(world)]
            outs = g.all_reduce(xs, variant=rng.choice([0, 1, 2]))
            ref = _ref_sum(xs)
            bad += sum(int(not torch.equal(o, ref)) for o in outs)
        elif kind == "fused":
            h = rng.choice(hidden_choices)
            t = rng.randint(1, max(1, min(96, (2 << 20) // (h * 2 * world))))
            xs = [torch.randn(t, h, device="cuda", generator=gen).bfloat16() for _ in range(world)]
            res = torch.randn(t, h, device="cuda", generator=gen).bfloat16()
            gamma = torch.randn(h, device="cuda", generator=gen).bfloat16()
            ys, rs = g.all_reduce_add_rmsnorm(xs, [res.clone() for _ in range(world)], gamma, 1e-5,
                                              variant=rng.choice([0, 1, 2]))
            x32 = _ref_sum(xs).float() + res.float()
            bad += sum(int(not torch.equal(r, x32.bfloat16())) for r in rs)          # bit-exact
            y = ((x32 * torch.rsqrt(x32.pow(2).mean(-1, keepdim=True) + 1e-5)).bfloat16() * gamma)
            bad += sum(int(not torch.allclose(o.float(), y.float(), atol=1e-3, rtol=1.6e-2)) for o in ys)
            bad += sum(int(not torch.equal(o, ys[0])) for o in ys)                     # ranks agree
        elif kind == "gather":
            rows, cols = rng.randint(1, 40), rng.choice([2, 8, 264, 1002])
            slabs = [torch.randn(rows, cols, device="cuda", generator=gen).bfloat16() for _ in range(world)]
            ref = torch.cat(slabs, dim=-1)
            bad += sum(int(not torch.equal(o, ref)) for o in g.all_gather_lastdim(slabs))
        else:
            src = rng.randrange(world)
            dst = rng.choice([r for r in range(world) if r != src])
            t = torch.randn(8 * rng.randint(1, 40000), device="cuda", generator=gen).bfloat16()
            bad += int(not torch.equal(g.sendrecv(src, dst, t), t))
        assert bad == 0, f"first mismatch at call {it} ({kind}), world {world}"
    g.check_health()
    g.close()


def test_call_counter_wraps_without_breaking_the_region_rotation():
    """The 32-bit call counter wraps after 2^32 calls (days of decode).  0xffffffff is skipped so
    that region = counter % 3 keeps cycling 1, 2, 0 across the wrap."""
    from vllm_distributed_b200 import _cabi as C
    from vllm_distributed_b200.emu import EmuGroup
    world = 4
    g = EmuGroup(world, "cuda:0", max_bytes=1 << 20, ag_bytes=1 << 20, p2p_bytes=64 << 10)
    g.set_tuning(C.TUNE_SPIN_TIMEOUT_MS, 3000)
    gen = torch.Generator(device="cuda").manual_seed(3)

    def one_call(i):
        xs = [torch.randn(4 + i % 5, 4096, device="cuda", generator=gen).bfloat16() for _ in range(world)]
        res = torch.randn_like(xs[0])
        gamma = torch.randn(4096, device="cuda", generator=gen).bfloat16()
        ys, rs = g.all_reduce_add_rmsnorm(xs, [res.clone() for _ in range(world)], gamma, 1e-5,
                                          variant=1 + i % 2)
        x32 = _ref_sum(xs).float() + res.float()
        assert all(torch.equal(r, x32.bfloat16()) for r in rs), f"call {i}"
        assert all(torch.equal(y, ys[0]) for y in ys)

    for i in range(3):                       # counter is now 3: region 0 is next
        one_call(i)
    assert g.ranks[0].get_tuning(C.TUNE_DEBUG_EPOCH) == 3
    start = 0xFFFFFFF0 - (0xFFFFFFF0 % 3)    # a multiple of 3 just below the wrap: same region next
    g.set_tuning(C.TUNE_DEBUG_EPOCH, start)
    for i in range(40):                      # crosses 2^32
        one_call(i)
    after = g.ranks[0].get_tuning(C.TUNE_DEBUG_EPOCH)
    assert after < 64, hex(after)            # wrapped
    assert after == (start + 40) - 0xFFFFFFFF   # exactly one value (0xffffffff) was skipped
    g.check_health()
    g.close()
